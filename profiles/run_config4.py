"""BASELINE configs[3] end to end on ONE B200: Square(5, lattice_spacing=5.0) (25 atoms, full 2^25 space),
README adiabatic protocol (0.4 / 3.2 / 0.4 us), dop853 at the code defaults (atol 1e-7, rtol 1e-14),
|0..0> -> evolve -> normalise -> 1000 shots in one library call (streamed engine, tiled H psi).
   python profiles/run_config4.py [scale_of_durations=1.0] [shots=1000] [picture=s|ip] [atol=1e-7] [rtol=1e-14]
(config 4 is additionally timed with atol=1e-10, rtol=1e-7 and with interaction_picture=True)
Prints one JSON line.  The reference cannot run this size in reasonable time (CSR needs ~10 GB and
a solve takes days), so what is checked here is the norm before normalisation, the step statistics and the physics
of the samples (Rydberg density, nearest-neighbour blockade violations)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

from bloqade_analog_b200 import _lib, ir, programs
from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
shots = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
durations = [round(0.4 * scale, 6), round(3.2 * scale, 6), round(0.4 * scale, 6)]
prog = programs.adiabatic_program(ir.square_sites(5, 5, lattice_spacing=5.0), durations=durations)
h = RydbergHamiltonianCodeGen().emit(prog)
picture = sys.argv[3] if len(sys.argv) > 3 else "s"
atol = float(sys.argv[4]) if len(sys.argv) > 4 else 1e-7
rtol = float(sys.argv[5]) if len(sys.argv) > 5 else 1e-14
opts = _lib.make_opts("dop853", atol, rtol, interaction_picture=(picture == "ip"), engine="streamed")
offsets = np.array([0, h.record.size], dtype=np.int64)
times = np.array([[h.emulator_ir.duration]], dtype=np.float64)
np.random.seed(20261105)
uniforms = np.random.random_sample(shots).reshape(1, -1)
n0 = _lib.launch_count()
t0 = time.perf_counter()
shots_arr, stats = h.plan.run_batch(h.record, offsets, times.reshape(-1), uniforms, opts, True)
wall = time.perf_counter() - t0
s = shots_arr[0].astype(np.int64).reshape(shots, 5, 5)
nn_viol = ((s[:, 1:, :] & s[:, :-1, :]).any(axis=(1, 2)) | (s[:, :, 1:] & s[:, :, :-1]).any(axis=(1, 2))).mean()
st = stats[0]
res = {"config": "configs[3]: Square(5, 5.0), 25 atoms, D=2^25, README adiabatic protocol x%g, dop853 %g/%g, %s picture"
                 % (scale, atol, rtol, "interaction" if picture == "ip" else "Schroedinger"),
       "duration_us": float(h.emulator_ir.duration), "wall_s": wall, "status": int(st["status"]), "nfcn": int(st["nfcn"]),
       "nstep": int(st["nstep"]), "naccpt": int(st["naccpt"]), "nrejct": int(st["nrejct"]),
       "ms_per_rhs": 1e3 * wall / max(1, int(st["nfcn"])), "kernel_launches": int(_lib.launch_count() - n0),
       "shots": shots, "mean_rydberg_density": float(s.mean()), "shots_with_nn_blockade_violation": float(nn_viol),
       "rydberg_count_histogram": np.bincount(s.sum(axis=(1, 2)), minlength=14).tolist()}
print(json.dumps(res))

# picture = "both": the same protocol in the other picture too, final states compared (the two pictures are different
# numerical paths: gather RHS with the exp(-i V t) rotation against the tiled RHS)
if len(sys.argv) > 6 and sys.argv[6] == "compare":
    outs = {}
    for pic in ("s", "ip"):
        o = _lib.make_opts("dop853", atol, rtol, interaction_picture=(pic == "ip"), engine="streamed")
        t0 = time.perf_counter()
        psi, stt = h.plan.evolve_batch(h.record, offsets, times, o)
        outs[pic] = (psi.reshape(-1), int(stt.reshape(-1)[0]["nfcn"]), time.perf_counter() - t0)
    a, b = outs["s"][0], outs["ip"][0]
    ov = np.vdot(a, b)
    infid = abs(1.0 - abs(ov) ** 2 / (np.vdot(a, a).real * np.vdot(b, b).real))
    print(json.dumps({"compare": "Schroedinger vs interaction picture, same protocol and tolerances", "infidelity": float(infid),
                      "max_abs_diff": float(np.abs(a - b).max()), "nfcn_s": outs["s"][1], "nfcn_ip": outs["ip"][1],
                      "evolve_s_s": outs["s"][2], "evolve_s_ip": outs["ip"][2]}))
