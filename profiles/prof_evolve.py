"""Small driver for ncu: N short C3-sweep tasks through one evolve-kernel launch.
   python profiles/prof_evolve.py <engine> [n_tasks] [duration_us]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bloqade_analog_b200 import _lib, programs
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from bloqade_analog_b200.lowering import concat_records

engine = sys.argv[1] if len(sys.argv) > 1 else "auto"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 148
dur = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
n_atoms = int(sys.argv[4]) if len(sys.argv) > 4 else 12
progs = [programs.config3_task(4.0 + 0.1 * k, dur, n_atoms) for k in range(n)]
cg = RydbergHamiltonianCodeGen(CompileCache())
hams = [cg.emit(p) for p in progs]
plan = hams[0].plan
blob, off = concat_records([h.record for h in hams])
times = np.array([[h.emulator_ir.duration] for h in hams])
opts = _lib.make_opts("dop853", 1e-7, 1e-14, engine=engine)
plan.enable_timing(True)
out, st = plan.evolve_batch(blob, off, times, opts)
ms, k = plan.get_timing()
print(engine, "atoms", n_atoms, "tasks", n, "kernel ms", ms, "rhs/task", st["nfcn"].mean(), "us per rhs per task-slot", 1e3 * ms / st["nfcn"].mean())
