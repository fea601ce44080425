#!/usr/bin/env python
"""bench.py - headline benchmark of the B200 emulator backend (contract: task prompt (4)).

Metric (BASELINE.json): time-evolved tasks / second.  A task = one EmulatorProgram integrated
0 -> duration with dop853 at the run() code defaults (atol=1e-7, rtol=1e-14), normalised and
Born-sampled (100 shots) - what `BloqadeTask.run` does (task/bloqade.py:43-88).

Workload (N=1): BASELINE.json configs[2], the 4096-task parameter sweep on a 12-atom chain (D = 4096,
64 final detunings x 64 sweep durations).  It is the configuration the tasks/sec metric and the 1/2/4/8 GPU
scaling are quoted on.  N>1: every rank runs its own 4096-task block of a 64N x 64 grid (weak scaling, no
data-path collective).

   python bench.py [--gpus N] [--steps K] [--warmup W] [--tasks T] [--impl reference]

One JSON line on stdout (rank 0).
  value          tasks/s with inputs resident in HBM (device-timed, successful tasks only)
  e2e            the same sweep through the call a user makes - program.bloqade.python().run(100) on the B200
                 routine + batch.report() - lowering, H2D, GPU, D2H and packaging inside the timed region
  e2e_c_abi      bqa_run_batch with pinned HOST buffers (the C-ABI level of the same step)
  roofline       the dominant kernel of the step (the CTA-per-task evolve kernel)
  roofline_hpsi  the standalone fused H(t) psi on a BASELINE configs[3]-sized state (5x5 lattice, 2^25 amplitudes)
  cpu_baseline   the UNMODIFIED reference (baseline/_ref) on the host cores, bounded sample of the same sweep
  strong         (N > 1) ONE 4096-task sweep split over the ranks
  sharded        (N > 1) a (24 + log2 N)-atom state sharded over the ranks vs the unsharded run; N = 8: configs[4]
  small_configs  configs[0] / configs[1] (2a, 2b) through .run(); configs[3] end to end through .run(interaction_picture=True)
`--impl reference` runs the unmodified reference through its own API and parallel path on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "time-evolved tasks/sec"
UNIT = "tasks/s"
SHOTS = 100
SOLVER = dict(solver_name="dop853", atol=1e-7, rtol=1e-14)


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi, B200_PROFILING.md "clocks DURING the timed region")
# ------------------------------------------------------------------------------------------------------
class Clocks:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                pw.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(self.NAMES, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------------
# the reference itself (baseline/_ref on the GPU box, /root/reference in the build container)
# ------------------------------------------------------------------------------------------------------
def load_reference():
    """import the UNMODIFIED reference package through baseline/refshim.py; None when it is not present"""
    base = os.path.join(ROOT, "baseline")
    if base not in sys.path:
        sys.path.insert(0, base)
    try:
        import refshim

        if not refshim.reference_available():
            return None
        import warnings

        warnings.filterwarnings("ignore")
        refshim.install()
        import bloqade.analog as ba

        return ba
    except Exception as e:  # noqa: BLE001
        log("reference package unavailable:", repr(e))
        return None


def sweep_grid(n_tasks, rank=0):
    """(final_detuning, total_duration) of the C3 sweep block of `rank`: 64 detunings x 64 durations (task =
    i_det * 64 + i_dur); other ranks get the neighbouring block of final detunings (same spread of durations =>
    same work per rank)."""
    side = 64
    dets = np.linspace(4.0, 32.0, side) + 0.2 * rank
    durs = np.linspace(0.5, 4.0, side)
    grid = [(float(round(d, 6)), float(round(t, 6))) for d in dets for t in durs]
    if n_tasks < len(grid):  # sub-sampled sweep keeps the spread of durations
        pick = np.linspace(0, len(grid) - 1, n_tasks).round().astype(int)
        grid = [grid[i] for i in pick]
    while len(grid) < n_tasks:
        grid = grid + grid[: n_tasks - len(grid)]
    return grid


def sweep_programs(n_tasks, rank=0):
    from bloqade_analog_b200 import programs

    return [programs.config3_task(d, t) for d, t in sweep_grid(n_tasks, rank)]


def builder_sweep(grid):
    """the same sweep written with the reference's builder (what a user writes; README.md:88-108 waveforms):
    Chain(12).rydberg...piecewise_linear([0.1, 0.8, 0.1] * total, ...).batch_assign(final_detuning=..., total=...)"""
    from bloqade.analog import cast
    from bloqade.analog.atom_arrangement import Chain

    total = cast("total")
    durations = [0.1 * total, 0.8 * total, 0.1 * total]
    return (Chain(12, lattice_spacing=5.72).rydberg.rabi.amplitude.uniform.piecewise_linear(durations, [0, 15.8, 15.8, 0])
            .detuning.uniform.piecewise_linear(durations, [-16.33, -16.33, "final_detuning", "final_detuning"])
            .batch_assign(final_detuning=[g[0] for g in grid], total=[g[1] for g in grid]))


REF_SAMPLE = 32  # tasks of one reference step: the same list on every box


def reference_sample_grid(n=REF_SAMPLE):
    """n tasks spread over the sweep so that their mean duration equals the sweep's mean (2.25 us)"""
    side = 64
    dets = np.linspace(4.0, 32.0, side)
    durs = np.linspace(0.5, 4.0, side)
    idx = np.linspace(0, side - 1, n).round().astype(int)
    return [(float(round(dets[(7 * k) % side], 6)), float(round(durs[i], 6))) for k, i in enumerate(idx)]


def reference_run(ba, grid, workers):
    """one pass of the UNMODIFIED reference over `grid` through its own public API and parallel path:
    program.bloqade.python().run(100, multiprocessing=True, num_workers=workers)
    (src/bloqade/analog/ir/routine/bloqade.py:233-311, task/batch.py:235-244)"""
    routine = builder_sweep(grid).bloqade.python()
    assert type(routine).__name__ == "BloqadePythonRoutine", "the reference arm must run the reference's own routine"
    t0 = time.perf_counter()
    if workers > 1:
        batch = routine.run(SHOTS, multiprocessing=True, num_workers=workers, **SOLVER)
    else:
        batch = routine.run(SHOTS, **SOLVER)
    dt = time.perf_counter() - t0
    assert len(batch.tasks) == len(grid)
    return dt


def reference_versions():
    import numba
    import scipy

    return f"scipy {scipy.__version__}, numba {numba.__version__}, numpy {np.__version__}"


def cpu_baseline(n_sample, workers):
    """bounded sample of the workload on the host cores: the reference itself when it is present, else the
    oracle port on one core"""
    ba = load_reference()
    if ba is not None:
        reference_run(ba, reference_sample_grid(2), 1)  # numba JIT + imports, untimed
        grid = sorted(reference_sample_grid(n_sample), key=lambda g: -g[1])
        dt = reference_run(ba, grid, workers)
        return {"value": n_sample / dt, "unit": UNIT, "cores": workers, "kind": "reference",
                "sample": f"{n_sample} of the 4096 sweep tasks (mean duration {np.mean([g[1] for g in grid]):.2f} us) through "
                          f"the unmodified reference: program.bloqade.python().run(100, multiprocessing=True, "
                          f"num_workers={workers}), {dt:.1f} s; {reference_versions()}; os.cpu_count()={os.cpu_count()}"}
    from bloqade_analog_b200 import programs
    from oracle import emulator as oe

    oe.run(programs.config3_task(8.0, 0.02), 4, uniforms=np.full(4, 0.5), **SOLVER)  # numba JIT
    grid = reference_sample_grid(min(n_sample, 3))
    t0 = time.perf_counter()
    for d, t in grid:
        oe.run(programs.config3_task(d, t), SHOTS, uniforms=np.random.RandomState(1).random_sample(SHOTS), **SOLVER)
    dt = time.perf_counter() - t0
    return {"value": len(grid) / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{len(grid)} sweep tasks on oracle/emulator.py (reference package not present), {dt:.1f} s"}


WORKLOAD = ("configs[2]: 4096-task detuning/duration sweep on a 12-atom chain (D=4096 full space), dop853 atol=1e-7 "
            "rtol=1e-14, 100 shots/task")


def reference_arm(args):
    """`--impl reference`: the UNMODIFIED reference (baseline/_ref) on the box's host cores through its own public
    API and its own parallel path, on a bounded sample of the same workload (REF_SAMPLE tasks per step, the same
    list on every box).  Rank 0 alone runs it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ba = load_reference()
    if ba is None:
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref (copy of the reference package) not present"}),
              flush=True)
        return
    workers = max(1, min(os.cpu_count() or 1, REF_SAMPLE))
    grid = sorted(reference_sample_grid(REF_SAMPLE), key=lambda g: -g[1])
    reference_run(ba, reference_sample_grid(2), 1)  # numba JIT in this process; forked workers inherit it
    for _ in range(max(0, min(args.warmup, 1))):
        reference_run(ba, grid[-workers:], workers)
    dts = [reference_run(ba, grid, workers) for _ in range(args.steps)]
    dt = float(sum(dts))
    value = REF_SAMPLE * args.steps / dt
    sample = (f"{REF_SAMPLE} of the 4096 sweep tasks per step (spread over the sweep, mean duration "
              f"{np.mean([g[1] for g in grid]):.2f} us = the sweep's mean), program.bloqade.python().run(100, "
              f"multiprocessing=True, num_workers={workers}); {reference_versions()}; os.cpu_count()={os.cpu_count()}")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_tasks_per_step": REF_SAMPLE, "api": "program.bloqade.python().run"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------
def profile_record(name):
    """a number measured under ncu and committed under profiles/ (bench.py never runs under a profiler)"""
    path = os.path.join(ROOT, "profiles", name)
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f)
    return None


def hpsi_roofline(torch, device_index, peaks, peak_kind, n_atoms=25, reps=20):
    """standalone fused H(t) psi on the C4-sized full space: algorithmic bytes 32*D per apply"""
    from bloqade_analog_b200 import _lib, ir, programs
    from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen

    l1, l2 = {25: (5, 5), 24: (6, 4), 20: (5, 4)}.get(n_atoms, (n_atoms, 1))
    prog = programs.adiabatic_program(ir.square_sites(l1, l2, lattice_spacing=5.0))
    h = RydbergHamiltonianCodeGen(device=device_index).emit(prog)
    dim = h.space.size
    stream = torch.cuda.current_stream()
    h.plan.set_stream(stream.cuda_stream)
    psi = torch.randn(dim, 2, dtype=torch.float64, device="cuda")
    psi /= psi.norm()
    out = torch.empty_like(psi)
    D = _lib.MEM_DEVICE
    for _ in range(3):
        h.plan.apply(h.record, 1.7, psi.data_ptr(), out.data_ptr(), mem=D)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    torch.cuda.synchronize()
    for a, b in evs:  # 1 GiB in + out per apply (> L2), alternate directions so nothing stays cached
        a.record(stream)
        h.plan.apply(h.record, 1.7, psi.data_ptr(), out.data_ptr(), mem=D)
        b.record(stream)
    torch.cuda.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in evs])
    alg = 32.0 * dim
    ach = alg / (np.median(ms) * 1e-3) / 1e9
    tiled = n_atoms > 13 and not os.environ.get("BQA_NO_TILED")
    rec = profile_record("hpsi_traffic.json") or {}
    traffic = rec.get("tiled" if tiled else "one_pass", {}).get(str(n_atoms))
    res = {"kernel": "%s (fused -i H(t) psi, 2^%d full space, uniform drive + detuning + vdW diagonal)"
                     % ("apply_tiled_kernel passes" if tiled else "apply_kernel", n_atoms),
           "bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
           "frac_of_8TBps_nominal": ach / 8000.0, "peak_kind": peak_kind + " burst copy", "algorithmic_bytes_per_launch": alg,
           "ms_per_launch_median": float(np.median(ms)), "ms_per_launch_min": float(ms.min()),
           "traffic": traffic["bytes"] if traffic else None, "traffic_source": traffic["source"] if traffic else None}
    # kernels (3) and (4) on the same state: norm (16*D read) and Born sampling of 1000 shots (16*D read; the
    # two-level form for large states writes no CDF array)
    shots = 1000
    uni = torch.rand(shots, dtype=torch.float64, device="cuda")
    sh = torch.empty(shots, n_atoms, dtype=torch.uint8, device="cuda")
    nrm = torch.empty(1, dtype=torch.float64, device="cuda")

    def t_med(fn, reps=10):
        fn()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        torch.cuda.synchronize()
        for a, b in ev:
            a.record(stream)
            fn()
            b.record(stream)
        torch.cuda.synchronize()
        return float(np.median([a.elapsed_time(b) for a, b in ev]))

    ms_n = t_med(lambda: h.plan.norm_noalloc(psi.data_ptr(), 1, nrm.data_ptr()))
    ms_s = t_med(lambda: h.plan.sample(psi.data_ptr(), uni.data_ptr(), mem=D, n_states=1, n_shots=shots, shots_out=sh.data_ptr()))
    res["reductions"] = {
        "norm": {"ms": ms_n, "algorithmic_bytes": 16.0 * dim, "achieved": 16.0 * dim / ms_n / 1e6, "unit": "GB/s",
                 "frac": 16.0 * dim / ms_n / 1e6 / peaks["hbm_gbs"]},
        "sample_1000_shots": {"ms": ms_s, "algorithmic_bytes": 16.0 * dim, "achieved": 16.0 * dim / ms_s / 1e6, "unit": "GB/s",
                              "frac": 16.0 * dim / ms_s / 1e6 / peaks["hbm_gbs"]}}
    h.plan.close()
    del psi, out
    torch.cuda.empty_cache()
    return res


def small_config_lines(torch, ba, device, with_config4=True):
    """BASELINE configs[0] (C1, 6 tasks, 3x3 lattice) and configs[1] (C2, Chain(16) blockade subspace, D=2584)
    end to end through program.bloqade.python().run(100) on the B200 routine: seconds per call"""
    from bloqade.analog.atom_arrangement import Chain, Square

    from bloqade_analog_b200 import routine

    om, de, du = [0, 15.8, 15.8, 0], [-16.33, -16.33, 16.33, 16.33], [0.4, 3.2, 0.4]
    c1 = (Square(3, lattice_spacing="a").rydberg.rabi.amplitude.uniform.piecewise_linear(du, om)
          .detuning.uniform.piecewise_linear(du, de).batch_assign(a=list(np.arange(4.0, 7.0, 0.5))))
    c2 = (Chain(16, lattice_spacing=5.72).rydberg.rabi.amplitude.uniform.piecewise_linear(du, om)
          .detuning.uniform.piecewise_linear(du, de))
    c2b = Chain(16, lattice_spacing=5.72).rydberg.detuning.uniform.piecewise_linear(du, de)
    for i in range(16):  # sixteen local drives = sixteen IndexMapping terms in the reference (north_star's case)
        c2b = c2b.amplitude.location(i).piecewise_linear(du, om)
    # configs[3]: 5x5 lattice, full 2^25 space, ONE task, 1000 shots - in the interaction picture (the reference's own
    # run() option): 6.6 k RHS evaluations instead of 315 k, seconds instead of minutes (profiles/r02_summary.md)
    c4 = (Square(5, lattice_spacing=5.0).rydberg.rabi.amplitude.uniform.piecewise_linear(du, om)
          .detuning.uniform.piecewise_linear(du, de))
    out = {}
    cases = [("config1", c1, {}, SHOTS, 3), ("config2a", c2, {"blockade_radius": 6.0}, SHOTS, 3),
             ("config2b", c2b, {"blockade_radius": 6.0}, SHOTS, 3)]
    if with_config4:
        cases.append(("config4_interaction_picture", c4, {"interaction_picture": True}, 1000, 1))
    for name, prog, kw, shots, reps in cases:
        r = routine.b200_routine(prog)
        if reps > 1:
            r.run(shots, devices=[device], **kw, **SOLVER)
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            batch = r.run(shots, devices=[device], **kw, **SOLVER)
            ts.append(time.perf_counter() - t0)
        n = len(batch.tasks)
        st = [t.integrator_stats for t in batch.tasks.values()]
        out[name] = {"tasks": n, "shots": shots, "s_per_run_call": float(np.median(ts)), "tasks_per_s": n / float(np.median(ts)),
                     "rhs_evals": int(sum(int(s["nfcn"]) for s in st)), "lowering": routine.LAST_LOWERING}
        if name.startswith("config4"):
            bits = np.array(batch.report().bitstrings()[0])  # 0 = Rydberg in the report convention
            out[name]["mean_rydberg_density"] = float(1.0 - bits.mean())
    return out


def sharded_leg(torch, dist, rank, world, local_rank):
    """SURVEY.md 8(e) row 3 on the N GPUs of this run: a (24 + log2 N)-atom full-space state sharded over the ranks
    against the unsharded run on rank 0 (H psi, dop853 evolution, sampling) + timings; at N = 8 also the
    30-atom BASELINE configs[4] H psi."""
    from bloqade_analog_b200 import _lib, ir, programs
    from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen
    from bloqade_analog_b200.sharded import ShardedHamiltonian
    from bloqade_analog_b200.state_vector import AnalogGate

    g = int(np.log2(world))
    n_atoms = 24 + g
    dev = torch.device("cuda", local_rank)
    d = 0.02
    prog = programs.adiabatic_program(ir.chain_sites(n_atoms, lattice_spacing=5.5), durations=[0.1 * d, 0.8 * d, 0.1 * d])
    sh = ShardedHamiltonian(prog)
    ldim = sh.local_dim
    res = {"n_atoms": n_atoms, "world": world, "local_dim": ldim, "duration_us": d}
    full = RydbergHamiltonianCodeGen(device=local_rank).emit(prog) if rank == 0 else None
    gen = torch.Generator(device="cpu").manual_seed(11)
    psi_full = torch.randn(1 << n_atoms, 2, dtype=torch.float64, generator=gen)
    psi_full /= psi_full.norm()
    psi_loc = torch.view_as_complex(psi_full[rank * ldim:(rank + 1) * ldim].contiguous()).to(dev)
    out = sh.apply(psi_loc, 0.6 * d, raw=True)
    out_r = torch.view_as_real(out)
    gathered = [torch.empty_like(out_r) for _ in range(world)] if rank == 0 else None
    dist.gather(out_r, gathered, dst=0)
    if rank == 0:
        pf = psi_full.to(dev)
        ref = torch.empty_like(pf)
        full.plan.apply(full.record, 0.6 * d, pf.data_ptr(), ref.data_ptr(), raw=True, mem=_lib.MEM_DEVICE)
        torch.cuda.synchronize()
        got = torch.cat(gathered)
        res["apply_max_rel_err"] = float((got - ref).abs().max() / ref.abs().max())
        del pf, ref, got, gathered
    del psi_full
    partners = [sh.desc.ys[rank ^ (1 << b)] for b in range(sh.g)]

    def time_apply(shh, n_total, reps=10):
        part = [shh.desc.ys[rank ^ (1 << b)] for b in range(shh.g)]
        o = torch.empty(shh.local_dim, dtype=torch.complex128, device=dev)
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            shh.plan.apply_sharded(shh.record, 0.6 * d, n_total, rank, part, shh._ys, o.data_ptr(), raw=True)
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    res["apply_ms"] = time_apply(sh, n_atoms)
    res["apply_alg_GBps_per_gpu"] = 32.0 * (1 << n_atoms) / (res["apply_ms"] * 1e-3) / 1e9 / world
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    states, stats = sh.evolve()
    torch.cuda.synchronize()
    res["evolve_s"] = time.perf_counter() - t0
    res["stats"] = {k: int(stats[k][0]) for k in ("status", "nfcn", "nstep", "naccpt", "nrejct")}
    res["ms_per_rhs"] = 1e3 * res["evolve_s"] / max(1, res["stats"]["nfcn"])
    res["norm"] = sh.norm(states[0])
    final_r = torch.view_as_real(states[0])
    gathered = [torch.empty_like(final_r) for _ in range(world)] if rank == 0 else None
    dist.gather(final_r, gathered, dst=0)
    np.random.seed(9123892)
    shots = sh.sample(states[0], 200)
    if rank == 0:
        (ref_state,) = AnalogGate(full).apply(full.space.zero_state(np.complex128), engine="streamed")
        got = np.ascontiguousarray(torch.cat(gathered).cpu().numpy()).view(np.complex128).reshape(-1)
        ov = np.vdot(ref_state.data, got)
        res["infidelity"] = float(abs(1.0 - abs(ov) ** 2 / (np.vdot(got, got).real * np.vdot(ref_state.data, ref_state.data).real)))
        np.random.seed(9123892)
        ref_shots = full.plan.sample(got, np.random.random_sample(200))[0]
        res["shots_rows_differing"] = int((ref_shots != shots).any(axis=1).sum())
        full.plan.close()
    sh.close()
    del states, gathered
    torch.cuda.empty_cache()
    if world == 8:  # BASELINE configs[4]: 6x5 lattice, 2^30 amplitudes, 2 GiB per vector per GPU
        prog5 = programs.config5()
        sh5 = ShardedHamiltonian(prog5)
        ms = time_apply(sh5, 30, reps=5)
        res["config5_hpsi_ms"] = ms
        res["config5_alg_GBps_per_gpu"] = 32.0 * (1 << 30) / (ms * 1e-3) / 1e9 / 8
        res["config5_nvlink_floor_ms"] = 3 * (16.0 * (1 << 27)) / 900e9 * 1e3
        sh5.close()
    dist.barrier()
    return res


def gpu_arm(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        # NCCL writes its version banner to stdout at NCCL_DEBUG=VERSION/INFO; stdout carries the JSON line only
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "INFO", "TRACE") and not os.environ.get("BQA_KEEP_NCCL_DEBUG"):
            os.environ["NCCL_DEBUG"] = "WARN"
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from bloqade_analog_b200 import _lib
    from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
    from bloqade_analog_b200.lowering import concat_records

    _lib.require_device()  # no CPU fallback: fails loudly without the CUDA library / a device
    peaks, peak_kind = measured_peaks()

    # ---- host-side lowering of the device-resident leg (outside its timed region) -----------------------
    t0 = time.perf_counter()
    progs = sweep_programs(args.tasks, rank)
    codegen = RydbergHamiltonianCodeGen(CompileCache(), device=local_rank)
    hams = [codegen.emit(p) for p in progs]
    plan = hams[0].plan
    assert all(h.plan is plan for h in hams)
    blob, offsets = concat_records([h.record for h in hams])
    times = np.array([h.emulator_ir.duration for h in hams], dtype=np.float64)
    rs = np.random.RandomState(9123892 + rank)
    uniforms = rs.random_sample((args.tasks, SHOTS))
    lower_s = time.perf_counter() - t0
    dim, n_atoms = plan.dim, plan.n_atoms
    opts = _lib.make_opts(**SOLVER, engine=args.engine)

    stream = torch.cuda.Stream()
    plan.set_stream(stream.cuda_stream)
    plan.enable_timing(True)

    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()

    # device-resident inputs / outputs
    d_blob = torch.from_numpy(blob).cuda()
    d_off = torch.from_numpy(offsets).cuda()
    d_times = torch.from_numpy(times).cuda()
    d_uni = torch.from_numpy(uniforms).cuda()
    d_shots = torch.empty((args.tasks, SHOTS, n_atoms), dtype=torch.uint8, device="cuda")
    # pinned host inputs / outputs for the C-ABI host-buffer leg
    keep = [pinned(blob), pinned(offsets), pinned(times), pinned(uniforms)]
    h_blob, h_off, h_times, h_uni = (k[1] for k in keep)
    h_shots_t = torch.empty((args.tasks, SHOTS, n_atoms), dtype=torch.uint8).pin_memory()
    h_shots = h_shots_t.numpy()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    D = _lib.MEM_DEVICE
    last_stats = {}

    def step_device():
        _, st = plan.run_batch(d_blob.data_ptr(), d_off.data_ptr(), d_times.data_ptr(), d_uni.data_ptr(), opts,
                               shots_out=d_shots.data_ptr(), mem=D, n_tasks=args.tasks, n_shots=SHOTS)
        last_stats["st"] = st

    def step_host():
        _, st = plan.run_batch(h_blob, h_off, h_times, h_uni, opts, shots_out=h_shots)
        last_stats["st"] = st

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn()
        barrier()
        plan.enable_timing(True)  # reset the kernel-time accumulator
        launches0 = _lib.launch_count()
        total_ms = 0.0
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.fill_(1)  # L2 flush between timed iterations (outside the event bracket)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            step_fn()
            b.record(stream)
            b.synchronize()
            total_ms += a.elapsed_time(b)
        barrier()
        return max_over_ranks(total_ms), _lib.launch_count() - launches0

    clocks = Clocks(local_rank) if rank == 0 else None
    dev_ms, launches = timed(step_device, args.steps, args.warmup)
    kern_ms, kern_n = plan.get_timing()
    st = last_stats["st"]
    ok_tasks = int((st["status"] == 1).sum())
    bad = args.tasks - ok_tasks
    bad_ids = [int(i) for i in np.nonzero(st["status"] != 1)[0][:8]]
    nfcn_total = int(st["nfcn"].sum())
    clock_rec = clocks.stop() if clocks else None
    abi_ms, _ = timed(step_host, max(1, min(args.steps, 2)), 1)
    abi_steps = max(1, min(args.steps, 2))

    # ---- end to end through the call a user makes: program.bloqade.python().run(100) on the B200 routine ----
    # timed region = everything .run() does for the rank's whole sweep: lowering of the symbolic program, drawing
    # the sampling uniforms, H2D of records/uniforms, the GPU work, D2H of shots/stats, packaging into the
    # reference's LocalBatch - plus batch.report() (the device->host result the user reads)
    ba = load_reference()
    e2e = None
    if ba is not None:
        from bloqade_analog_b200 import routine as b200_routine

        grid = sweep_grid(args.tasks, rank)
        routine = b200_routine.b200_routine(builder_sweep(grid))
        parts = {}

        def step_run():
            t0 = time.perf_counter()
            np.random.seed(9123892 + rank)
            batch = routine.run(SHOTS, devices=[local_rank], raise_on_error=False, **SOLVER)
            t1 = time.perf_counter()
            rep = batch.report()
            t2 = time.perf_counter()
            parts["run_s"], parts["report_s"] = t1 - t0, t2 - t1
            parts["rows"] = len(rep.dataframe)
            parts["ok"] = sum(int(t.integrator_stats["status"]) == 1 for t in batch.tasks.values())
            return batch

        step_run()  # warm-up (imports, kernels)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_run()
        torch.cuda.synchronize()
        run_s = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": parts["ok"] * world * args.steps / run_s, "unit": UNIT,
               "h2d_bytes_per_step": int(blob.nbytes + offsets.nbytes + times.nbytes + uniforms.nbytes),
               "d2h_bytes_per_step": int(h_shots.nbytes + st.nbytes), "ms_per_step": 1e3 * run_s / args.steps,
               "api": "program.bloqade.python().run(100, ...) on BloqadeB200Routine + batch.report()",
               "lowering": b200_routine.LAST_LOWERING, "run_call_s": parts["run_s"], "report_s": parts["report_s"],
               "report_rows": parts["rows"], "timing": "host wall clock around the call, max over ranks"}
    ok_h = int((last_stats["st"]["status"] == 1).sum())
    abi = {"value": ok_h * world * abi_steps / (abi_ms * 1e-3), "unit": UNIT, "ms_per_step": abi_ms / abi_steps,
           "api": "bqa_run_batch with pinned host buffers (C ABI)"}
    if e2e is None:
        e2e = dict(abi, h2d_bytes_per_step=int(blob.nbytes + offsets.nbytes + times.nbytes + uniforms.nbytes),
                   d2h_bytes_per_step=int(h_shots.nbytes + st.nbytes), note="reference package absent: C-ABI level only")

    # ---- strong scaling: ONE 4096-task sweep split over the ranks (distributed.run_programs_sharded) ---------
    strong = None
    if dist is not None and not args.no_strong:
        from bloqade_analog_b200.distributed import run_programs_sharded

        all_progs = sweep_programs(4096, 0)
        u = np.random.RandomState(7).random_sample((4096, SHOTS))
        run_programs_sharded(all_progs[: 64 * world], SHOTS, uniforms=u[: 64 * world], device=local_rank, **SOLVER)
        barrier()
        t0 = time.perf_counter()
        res = run_programs_sharded(all_progs, SHOTS, uniforms=u, device=local_rank, **SOLVER)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        strong = {"tasks_total": 4096, "value": sum(r.status == 1 for r in res) / dt, "unit": UNIT, "s": dt,
                  "scaling": "strong", "note": "contiguous blocks of 4096/N tasks per rank incl. host lowering of the "
                  "rank's block and the all_gather of shots"}

    # ---- sharded full-space state over the N GPUs ---------------------------------------------------------
    sharded = None
    if dist is not None and not args.no_sharded:
        del d_blob, d_uni, d_shots, flush
        torch.cuda.empty_cache()
        try:
            sharded = sharded_leg(torch, dist, rank, world, local_rank)
        except Exception as e:  # noqa: BLE001
            sharded = {"error": repr(e)}

    total_tasks = ok_tasks * world
    value = total_tasks * args.steps / (dev_ms * 1e-3)

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel of the step: evolve_small_kernel (one launch per step) --------
    alg_bytes = 32.0 * dim * nfcn_total  # 32*D per H psi x RHS evaluations of the launch
    kms = kern_ms / max(1, kern_n)
    ach = alg_bytes / (kms * 1e-3) / 1e9
    evo = profile_record("evolve_kernel_ncu.json") or {}
    roofline = {
        "kernel": "evolve_small_kernel (CTA-per-task DOP853: all stage combines, fused H(t) psi and error norms)",
        "bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
        "peak_kind": peak_kind + " copy bandwidth",
        # DRAM bytes per RHS evaluation from the committed ncu capture of this kernel x the RHS evaluations of the launch
        "traffic": (evo["dram_bytes_per_rhs"] * nfcn_total) if "dram_bytes_per_rhs" in evo else None,
        "traffic_source": evo.get("source"),
        "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": kms, "share_of_step": kern_ms / dev_ms,
        "rhs_evals_per_launch": nfcn_total,
        "binding_resource": evo.get("binding_resource"),
        "note": "the state vectors of a task live in shared memory / L2, so HBM is not the binding resource of this "
                "kernel: 'achieved' is the algorithmic H psi rate (32*D per RHS) it sustains, 'binding_resource' the "
                "pipe utilisation ncu measured (profiles/)",
    }
    hpsi = None
    if not args.no_hpsi:
        try:
            hpsi = hpsi_roofline(torch, local_rank, peaks, peak_kind, args.hpsi_atoms)
        except Exception as e:  # noqa: BLE001
            hpsi = {"error": repr(e)}
    small = None
    if ba is not None and not args.no_small:
        try:
            small = small_config_lines(torch, ba, local_rank, with_config4=not args.no_c4)
        except Exception as e:  # noqa: BLE001
            small = {"error": repr(e)}
    cpu = None
    try:
        cores = max(1, min(os.cpu_count() or 1, 32))
        cpu = cpu_baseline(args.cpu_tasks if args.cpu_tasks > 0 else 2 * cores, cores)
    except Exception as e:  # noqa: BLE001
        cpu = {"error": repr(e)}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "tasks_per_gpu": args.tasks, "engine": args.engine, "dim": dim, "shots": SHOTS, "parallelism": f"task-sharded x{world}, no collective",
                   "l2": "flushed between timed steps (256 MiB write)", "host_lowering_s": round(lower_s, 3),
                   "failed_tasks": bad, "failed_task_ids": bad_ids, "value_counts": "successful tasks only",
                   "mean_rhs_per_task": nfcn_total / args.tasks},
        "e2e": e2e,
        "e2e_c_abi": abi,
        "gpu_launches": int(launches),
        "roofline": roofline,
        "roofline_hpsi": hpsi,
        "cpu_baseline": cpu,
        "clocks": clock_rec,
    }
    if strong is not None:
        line["strong"] = strong
    if sharded is not None:
        line["sharded"] = sharded
    if small is not None:
        line["small_configs"] = small
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tasks", type=int, default=4096, help="sweep tasks per GPU")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-tasks", type=int, default=0, help="tasks of the bounded CPU-baseline sample (0: 2 per core)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (N > 1)")
    ap.add_argument("--no-sharded", action="store_true", help="skip the sharded-state leg (N > 1)")
    ap.add_argument("--no-small", action="store_true", help="skip the configs[0]/configs[1] lines")
    ap.add_argument("--no-c4", action="store_true", help="skip the configs[3] end-to-end run (interaction picture, ~25 s)")
    ap.add_argument("--hpsi-atoms", type=int, default=25)
    ap.add_argument("--no-hpsi", action="store_true")
    ap.add_argument("--engine", default="auto", choices=["auto", "cluster", "cta", "streamed"])
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
