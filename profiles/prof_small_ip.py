"""Engine choice for SMALL states outside the fused Schroedinger dop853 path (interaction picture / dopri5):
cluster engine (generic gather through DSMEM) against the CTA engine (fused RHS routines).
   python profiles/prof_small_ip.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from bloqade_analog_b200 import _lib, programs
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from bloqade_analog_b200.lowering import concat_records

res = []
for n_atoms in (8, 9, 10, 11):
    progs = [programs.config3_task(4.0 + 0.1 * k, 0.5, n_atoms) for k in range(296)]
    cg = RydbergHamiltonianCodeGen(CompileCache())
    hams = [cg.emit(p) for p in progs]
    plan = hams[0].plan
    blob, off = concat_records([h.record for h in hams])
    times = np.array([[h.emulator_ir.duration] for h in hams])
    for solver, ip in (("dop853", True), ("dopri5", False)):
        row = {"atoms": n_atoms, "solver": solver, "interaction_picture": ip}
        for engine in ("cluster", "cta"):
            opts = _lib.make_opts(solver, 1e-7, 1e-14, interaction_picture=ip, engine=engine)
            plan.enable_timing(True)
            out, st = plan.evolve_batch(blob, off, times, opts)
            ms, k = plan.get_timing()
            plan.enable_timing(False)
            row[engine + "_ms"] = ms
            row["rhs_per_task"] = float(st["nfcn"].mean())
        res.append(row)
        print(json.dumps(row), flush=True)
