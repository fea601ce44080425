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

One JSON line on stdout (rank 0).  `value` = tasks/s with inputs resident in HBM; `e2e` = the same through
the C ABI with pinned HOST buffers (H2D of records/uniforms and D2H of shots/stats inside the timed region);
`roofline` = the dominant kernel of the step (the CTA-per-task evolve kernel); `roofline_hpsi` = the
standalone fused H(t) psi apply on a BASELINE configs[3]-sized state (5x5 lattice, 2^25 amplitudes);
`cpu_baseline` = the oracle's NumPy/SciPy/Numba port of the reference path on the host cores.
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


def sweep_programs(n_tasks, rank=0):
    """C3 sweep block of `rank`: 64 detunings x 64 durations (task = i_det * 64 + i_dur); other ranks get the
    neighbouring block of final detunings (same spread of durations => same work per rank)."""
    from bloqade_analog_b200 import programs

    side = 64
    dets = np.linspace(4.0, 32.0, side) + 0.2 * rank
    durs = np.linspace(0.5, 4.0, side)
    grid = [(float(round(d, 6)), float(round(t, 6))) for d in dets for t in durs]
    if n_tasks < len(grid):  # sub-sampled sweep keeps the spread of durations
        pick = np.linspace(0, len(grid) - 1, n_tasks).round().astype(int)
        grid = [grid[i] for i in pick]
    while len(grid) < n_tasks:
        grid = grid + grid[: n_tasks - len(grid)]
    return [programs.config3_task(d, t) for d, t in grid]


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
# CPU arm: the oracle port of the reference path (scipy.integrate.ode + per-term NumPy/Numba passes)
# ------------------------------------------------------------------------------------------------------
def _cpu_task(args):
    d, t, seed = args
    from bloqade_analog_b200 import programs
    from oracle import emulator as oe

    prog = programs.config3_task(d, t)
    rs = np.random.RandomState(seed)
    t0 = time.perf_counter()
    oe.run(prog, SHOTS, uniforms=rs.random_sample(SHOTS), **SOLVER)
    return time.perf_counter() - t0


def _cpu_sample_grid(n):
    """n tasks spread over the sweep so that their mean duration equals the sweep's mean (2.25 us)"""
    side = 64
    dets = np.linspace(4.0, 32.0, side)
    durs = np.linspace(0.5, 4.0, side)
    idx = np.linspace(0, side - 1, n).round().astype(int)
    return [(float(round(dets[(7 * k) % side], 6)), float(round(durs[i], 6)), k) for k, i in enumerate(idx)]


def cpu_warmup():
    from bloqade_analog_b200 import programs
    from oracle import emulator as oe

    oe.run(programs.config3_task(8.0, 0.02), 4, uniforms=np.full(4, 0.5), **SOLVER)  # numba JIT


def cpu_baseline_single(n_sample=3):
    """rank-0 bounded sample on ONE core: tasks/s of the oracle port"""
    cpu_warmup()
    grid = _cpu_sample_grid(n_sample)
    t0 = time.perf_counter()
    for g in grid:
        _cpu_task(g)
    dt = time.perf_counter() - t0
    return {"value": n_sample / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{n_sample} of the 4096 sweep tasks (durations {[g[1] for g in grid]} us), oracle/emulator.py "
                      f"(scipy ode dop853 + numba term passes), {dt:.1f} s"}


def reference_arm(args):
    """`--impl reference`: the reference's CPU implementation of the path = the oracle port (the reference
    itself is pure Python that cannot travel to the GPU box), on all host cores through a
    process pool like the reference's own `multiprocessing=True` fan-out (task/batch.py:235-244)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp

    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 64))
    per_step = 2 * workers  # bounded sample: two tasks per worker per step, longest first (~10 s steps)
    grid = sorted(_cpu_sample_grid(per_step), key=lambda g: -g[1])
    ctx = mp.get_context("fork")
    cpu_warmup()
    with ctx.Pool(workers) as pool:
        for _ in range(max(1, min(args.warmup, 1))):
            list(pool.imap_unordered(_cpu_task, grid[-workers:], chunksize=1))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            list(pool.imap_unordered(_cpu_task, grid, chunksize=1))
        dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[2]: 4096-task detuning/duration sweep, 12-atom chain (D=4096), dop853 atol=1e-7 rtol=1e-14, 100 shots",
                   "sample_tasks_per_step": per_step},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port",
                         "sample": f"{per_step} sweep tasks per step over {workers} processes (oracle/emulator.py)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------
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
    # ncu --set full, 2^25: dram read+write of the two tiled passes (profiles/r01_summary.md: 1.39 GB + 1.67 GB);
    # one-pass gather kernel 4.54 GB
    traffic = (3.07e9 if tiled else 4.54e9) if n_atoms == 25 else None
    res = {"kernel": "%s (fused -i H(t) psi, 2^%d full space, uniform drive + detuning + vdW diagonal)"
                     % ("apply_tiled_kernel x2 passes" if tiled else "apply_kernel", n_atoms),
           "bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
           "frac_of_8TBps_nominal": ach / 8000.0, "peak_kind": peak_kind + " burst copy", "algorithmic_bytes_per_launch": alg,
           "ms_per_launch_median": float(np.median(ms)), "ms_per_launch_min": float(ms.min()), "traffic": traffic}
    h.plan.close()
    del psi, out
    torch.cuda.empty_cache()
    return res


def gpu_arm(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from bloqade_analog_b200 import _lib
    from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
    from bloqade_analog_b200.lowering import concat_records

    _lib.require_device()  # no CPU fallback: fails loudly without the CUDA library / a device
    peaks, peak_kind = measured_peaks()

    # ---- host-side lowering (outside the timed region: done once per sweep, like _compile RB:204-231) ----
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
    # pinned host inputs / outputs for the end-to-end leg
    keep = [pinned(blob), pinned(offsets), pinned(times), pinned(uniforms)]
    h_blob, h_off, h_times, h_uni = (k[1] for k in keep)
    h_shots_t = torch.empty((args.tasks, SHOTS, n_atoms), dtype=torch.uint8).pin_memory()
    h_shots = h_shots_t.numpy()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    D, H = _lib.MEM_DEVICE, _lib.MEM_HOST
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
        t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), _lib.launch_count() - launches0

    clocks = Clocks(local_rank) if rank == 0 else None
    dev_ms, launches = timed(step_device, args.steps, args.warmup)
    kern_ms, kern_n = plan.get_timing()
    st = last_stats["st"]
    bad = int((st["status"] != 1).sum())
    nfcn_total = int(st["nfcn"].sum())
    clock_rec = clocks.stop() if clocks else None
    e2e_ms, _ = timed(step_host, args.steps, 1)
    st_h = last_stats["st"]
    bad += int((st_h["status"] != 1).sum())

    total_tasks = args.tasks * world
    value = total_tasks * args.steps / (dev_ms * 1e-3)
    e2e = total_tasks * args.steps / (e2e_ms * 1e-3)

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel of the step: evolve_small_kernel (one launch per step) --------
    alg_bytes = 32.0 * dim * nfcn_total  # 32*D per H psi x RHS evaluations of the launch
    kms = kern_ms / max(1, kern_n)
    ach = alg_bytes / (kms * 1e-3) / 1e9
    roofline = {
        "kernel": "evolve_small_kernel (CTA-per-task DOP853: all stage combines, fused H(t) psi and error norms)",
        "bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
        "peak_kind": peak_kind + " copy bandwidth", "traffic": None,
        "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": kms, "share_of_step": kern_ms / dev_ms,
        "rhs_evals_per_launch": nfcn_total,
        "note": "state vectors live in shared memory / L2, so the effective rate may exceed the HBM roofline",
    }
    hpsi = None
    if not args.no_hpsi:
        try:
            hpsi = hpsi_roofline(torch, local_rank, peaks, peak_kind, args.hpsi_atoms)
        except Exception as e:  # noqa: BLE001
            hpsi = {"error": repr(e)}
    cpu = None
    if world == 1 or True:
        try:
            cpu = cpu_baseline_single(args.cpu_tasks)
        except Exception as e:  # noqa: BLE001
            cpu = {"error": repr(e)}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[2]: 4096-task detuning/duration sweep on a 12-atom chain (D=4096 full space), "
                               "dop853 atol=1e-7 rtol=1e-14, 100 shots/task",
                   "tasks_per_gpu": args.tasks, "engine": args.engine, "dim": dim, "shots": SHOTS, "parallelism": f"task-sharded x{world}, no collective",
                   "l2": "flushed between timed steps (256 MiB write)", "host_lowering_s": round(lower_s, 3),
                   "failed_tasks": bad, "mean_rhs_per_task": nfcn_total / args.tasks},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(blob.nbytes + offsets.nbytes + times.nbytes + uniforms.nbytes),
                "d2h_bytes_per_step": int(h_shots.nbytes + st_h.nbytes), "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "roofline_hpsi": hpsi,
        "cpu_baseline": cpu,
        "clocks": clock_rec,
    }
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
    ap.add_argument("--cpu-tasks", type=int, default=3, help="tasks of the bounded CPU-baseline sample")
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
