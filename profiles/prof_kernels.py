"""Driver for ncu / timing of the kernels that had no profile yet (VERDICT r1 "What's weak" 8):
   python profiles/prof_kernels.py reductions [n_atoms=25]   norm (kernel 3) + Born sampling (kernel 4) on a 2^n state
   python profiles/prof_kernels.py streamed  [n_atoms=25]   a few dop853 steps of the streamed engine (combine / error kernels)
   python profiles/prof_kernels.py c1                       BASELINE configs[0]: 6 tasks of the 3x3 lattice (cluster engine)
   python profiles/prof_kernels.py c2                       BASELINE configs[1]: Chain(16) blockade subspace, D = 2584
   python profiles/prof_kernels.py three_level [n=8] [tasks=148]   three-level (hyperfine) atoms: Chain(n), D = 3^n, a sweep batch
Prints one JSON line with CUDA-event timings (torch events on the plan's stream)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from bloqade_analog_b200 import _lib, ir, programs
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from bloqade_analog_b200.lowering import concat_records

what = sys.argv[1]
D = _lib.MEM_DEVICE


def timed(fn, stream, reps):
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    torch.cuda.synchronize()
    for a, b in evs:
        a.record(stream)
        fn()
        b.record(stream)
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in evs]))


if what == "reductions":
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    l1, l2 = {25: (5, 5), 24: (6, 4), 20: (5, 4)}.get(n, (n, 1))
    h = RydbergHamiltonianCodeGen().emit(programs.adiabatic_program(ir.square_sites(l1, l2, lattice_spacing=5.0)))
    dim = h.space.size
    stream = torch.cuda.current_stream()
    h.plan.set_stream(stream.cuda_stream)
    psi = torch.randn(dim, 2, dtype=torch.float64, device="cuda")
    psi /= psi.norm()
    shots = 1000
    uni = torch.rand(shots, dtype=torch.float64, device="cuda")
    out = torch.empty(shots, n, dtype=torch.uint8, device="cuda")
    nrm = torch.empty(1, dtype=torch.float64, device="cuda")
    f_norm = lambda: h.plan.norm_noalloc(psi.data_ptr(), 1, nrm.data_ptr())
    f_samp = lambda: h.plan.sample(psi.data_ptr(), uni.data_ptr(), mem=D, n_states=1, n_shots=shots, shots_out=out.data_ptr())
    for _ in range(2):
        f_norm(); f_samp()
    ms_norm, ms_samp = timed(f_norm, stream, 10), timed(f_samp, stream, 10)
    print(json.dumps({"what": what, "n_atoms": n, "norm_ms": ms_norm, "norm_GBps": 16.0 * dim / ms_norm / 1e6,
                      "sample_1000_shots_ms": ms_samp, "sample_GBps_alg": 16.0 * dim / ms_samp / 1e6,
                      "norm_value": float(nrm.item())}))
elif what == "streamed":
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    l1, l2 = {25: (5, 5), 24: (6, 4), 20: (5, 4)}.get(n, (n, 1))
    # README protocol shape; only the first 2 ns are integrated (a few dozen dop853 steps at the stability limit)
    prog = programs.adiabatic_program(ir.square_sites(l1, l2, lattice_spacing=5.0))
    h = RydbergHamiltonianCodeGen().emit(prog)
    opts = _lib.make_opts("dop853", 1e-7, 1e-14, engine="streamed")
    offsets = np.array([0, h.record.size], dtype=np.int64)
    t_end = float(sys.argv[3]) if len(sys.argv) > 3 else 0.002
    t0 = time.perf_counter()
    out, st = h.plan.evolve_batch(h.record, offsets, np.array([[t_end]]), opts)
    wall = time.perf_counter() - t0
    s = st[0] if st.ndim == 1 else st[0][0]
    print(json.dumps({"what": what, "n_atoms": n, "t_end_us": t_end, "wall_s": wall, "status": int(s["status"]), "nfcn": int(s["nfcn"]),
                      "nstep": int(s["nstep"]), "ms_per_rhs_incl_setup": 1e3 * wall / max(1, int(s["nfcn"]))}))
elif what in ("c1", "c2"):
    if what == "c1":
        progs = [programs.adiabatic_program(ir.square_sites(3, 3, lattice_spacing=a)) for a in np.arange(4.0, 7.0, 0.5)]
    else:
        progs = [programs.adiabatic_program(ir.chain_sites(16, lattice_spacing=5.72), blockade_radius=6.0)]
    cg = RydbergHamiltonianCodeGen(CompileCache())
    hams = [cg.emit(p) for p in progs]
    res = {"what": what, "dim": int(hams[0].space.size), "tasks": len(hams)}
    opts = _lib.make_opts("dop853", 1e-7, 1e-14, engine=(sys.argv[2] if len(sys.argv) > 2 else "auto"))
    for rep in range(3):
        per = []
        t0 = time.perf_counter()
        nf = 0
        for h in hams:  # C1: every lattice spacing has its own van-der-Waals diagonal = its own plan
            h.plan.enable_timing(True)
            out, st = h.plan.evolve_batch(h.record, np.array([0, h.record.size], dtype=np.int64), np.array([[h.emulator_ir.duration]]), opts)
            ms, k = h.plan.get_timing()
            per.append(ms)
            nf += int(np.asarray(st["nfcn"]).sum())
            if rep == 0:
                import zlib
                res.setdefault("state_crc", []).append(zlib.crc32(np.ascontiguousarray(out).tobytes()))
        res["wall_s"] = time.perf_counter() - t0
        res["kernel_ms_per_task"] = per
        res["rhs_total"] = nf
        res["us_per_rhs"] = 1e3 * sum(per) / nf
    print(json.dumps(res))
elif what == "three_level":
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    n_tasks = int(sys.argv[3]) if len(sys.argv) > 3 else 148
    progs = [programs.adiabatic_program(ir.chain_sites(n, lattice_spacing=5.72), durations=[0.05, 0.4, 0.05],
                                        delta=[-16.33, -16.33, 4.0 + 0.1 * k, 4.0 + 0.1 * k], atom_type=ir.ThreeLevelAtom)
             for k in range(n_tasks)]
    cg = RydbergHamiltonianCodeGen(CompileCache())
    hams = [cg.emit(p) for p in progs]
    plan = hams[0].plan
    blob, off = concat_records([h.record for h in hams])
    times = np.array([[h.emulator_ir.duration] for h in hams])
    opts = _lib.make_opts("dop853", 1e-7, 1e-14)
    plan.enable_timing(True)
    out, st = plan.evolve_batch(blob, off, times, opts)
    ms, k = plan.get_timing()
    print(json.dumps({"what": what, "n_atoms": n, "dim": int(hams[0].space.size), "tasks": n_tasks, "kernel_ms": ms,
                      "tasks_per_s": 1e3 * n_tasks / ms, "mean_rhs_per_task": float(st["nfcn"].mean()),
                      "failed": int((st["status"].reshape(-1) != 1).sum()),
                      "alg_GBps": 32.0 * hams[0].space.size * float(st["nfcn"].sum()) / ms / 1e6}))
