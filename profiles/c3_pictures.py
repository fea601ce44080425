"""BASELINE configs[2] sweep tasks in the Schroedinger and in the interaction picture (run()'s interaction_picture flag):
RHS evaluations, kernel time and the distance between the final states of the two pictures.
   python profiles/c3_pictures.py [n_tasks=512]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

import bench
from bloqade_analog_b200 import _lib
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from bloqade_analog_b200.lowering import concat_records

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
progs = bench.sweep_programs(4096, 0)[:: 4096 // n]
cg = RydbergHamiltonianCodeGen(CompileCache())
hams = [cg.emit(p) for p in progs]
plan = hams[0].plan
blob, off = concat_records([h.record for h in hams])
times = np.array([[h.emulator_ir.duration] for h in hams])
res = {"tasks": len(hams)}
states = {}
for pic in ("schroedinger", "interaction"):
    opts = _lib.make_opts("dop853", 1e-7, 1e-14, interaction_picture=(pic == "interaction"))
    plan.enable_timing(True)
    out, st = plan.evolve_batch(blob, off, times, opts)
    ms, k = plan.get_timing()
    plan.enable_timing(False)
    ok = st["status"].reshape(-1) == 1
    states[pic] = out.reshape(len(hams), -1)
    res[pic] = {"kernel_ms": ms, "tasks_per_s": 1e3 * int(ok.sum()) / ms, "failed": int((~ok).sum()),
                "mean_rhs_per_task": float(st["nfcn"].mean())}
a, b = states["schroedinger"], states["interaction"]
ov = np.abs(np.einsum("ij,ij->i", a.conj(), b)) ** 2 / (np.einsum("ij,ij->i", a.conj(), a).real * np.einsum("ij,ij->i", b.conj(), b).real)
res["infidelity_between_pictures_max"] = float(np.nanmax(np.abs(1 - ov)))
res["infidelity_between_pictures_median"] = float(np.nanmedian(np.abs(1 - ov)))
print(json.dumps(res))
