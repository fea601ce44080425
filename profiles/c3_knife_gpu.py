"""GPU counterpart of tests/golden/make_c3_knife_edge.py: the C3 tasks whose status differs between SciPy and the
GPU, re-run on the GPU with the absolute tolerance changed by a relative 1e-12 ... 1e-9.
   python profiles/c3_knife_gpu.py 440 951 1783"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bloqade_analog_b200 import _lib, programs
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from bloqade_analog_b200.lowering import concat_records

EPS = [0.0, 1e-12, -1e-12, 1e-11, -1e-11, 1e-10, -1e-10, 1e-9, -1e-9]
tasks = [int(a) for a in sys.argv[1:]] or [440, 951, 1783]
grid = programs.config3_grid(64, 64)
cg = RydbergHamiltonianCodeGen(CompileCache())
hams = [cg.emit(programs.config3_task(*grid[k])) for k in tasks]
plan = hams[0].plan
blob, off = concat_records([h.record for h in hams])
times = np.array([[h.emulator_ir.duration] for h in hams])
doc = {}
for eps in EPS:
    opts = _lib.make_opts("dop853", 1e-7 * (1.0 + eps), 1e-14, engine="cta")
    _, st = plan.evolve_batch(blob, off, times, opts)
    for k, s in zip(tasks, st[:, 0]):
        doc.setdefault(str(k), {"eps": [], "status": [], "nfcn": [], "naccpt": [], "nrejct": []})
        d = doc[str(k)]
        d["eps"].append(eps); d["status"].append(int(s["status"])); d["nfcn"].append(int(s["nfcn"]))
        d["naccpt"].append(int(s["naccpt"])); d["nrejct"].append(int(s["nrejct"]))
for k, d in doc.items():
    print(k, list(zip(d["eps"], d["status"], d["nfcn"])))
print(json.dumps(doc))
