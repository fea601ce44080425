"""CTA engine at dim 4096: the register-carry DOP853 step (default) against the plain step (BQA_NO_FUSED_STEP=1):
   final states, step counters and kernel time on N short C3-sweep tasks.
   python profiles/check_fused_step.py [n_tasks] [duration_us] [n_atoms]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bloqade_analog_b200 import _lib, programs
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from bloqade_analog_b200.lowering import concat_records

n = int(sys.argv[1]) if len(sys.argv) > 1 else 148
dur = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
n_atoms = int(sys.argv[3]) if len(sys.argv) > 3 else 12
progs = [programs.config3_task(4.0 + 0.1 * k, dur, n_atoms) for k in range(n)]
cg = RydbergHamiltonianCodeGen(CompileCache())
hams = [cg.emit(p) for p in progs]
plan = hams[0].plan
blob, off = concat_records([h.record for h in hams])
times = np.array([[0.3 * h.emulator_ir.duration, h.emulator_ir.duration] for h in hams])
opts = _lib.make_opts("dop853", 1e-7, 1e-14, engine="cta")
res = {}
for mode in ("fused", "plain"):
    if mode == "plain":
        os.environ["BQA_NO_FUSED_STEP"] = "1"
    else:
        os.environ.pop("BQA_NO_FUSED_STEP", None)
    plan.evolve_batch(blob, off, times, opts)
    plan.enable_timing(True)
    out, st = plan.evolve_batch(blob, off, times, opts)
    ms, k = plan.get_timing()
    res[mode] = (out, st, ms)
    print(mode, "kernel ms", round(ms, 3), "rhs/task", st["nfcn"].sum(axis=1).mean(), "status ok", int((st["status"] == 1).sum()), "of", st.size)
a, b = res["fused"], res["plain"]
print("states bit-identical:", bool(np.array_equal(a[0], b[0])), " max |diff|", float(np.abs(a[0] - b[0]).max()))
print("counters identical:", all(np.array_equal(a[1][f], b[1][f]) for f in ("status", "nfcn", "nstep", "naccpt", "nrejct")))
print("speed-up", round(b[2] / a[2], 3))
