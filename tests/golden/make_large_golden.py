"""Full-protocol golden for the STREAMED engine + tiled H psi path (the code path of BASELINE configs[3], 5x5 lattice,
2^25 amplitudes) at a size the reference finishes in minutes: 4x4 square lattice, a = 5.0 um (same spacing and
blockade regime as configs[3]), README adiabatic protocol of 4 us, full 2^16 space, dop853 at the run() defaults.
The reference itself (imported through baseline/refshim.py) integrates it; stored: the final state as complex64
(fidelity against it is exact to ~1e-14), per-atom Rydberg densities / norm / <H>, Var H in float64, seeded shots.
tests/test_gpu_parity.py::test_streamed_full_protocol_16_atoms_matches_reference evolves the same program on the GPU
with engine="streamed" (tiled RHS) and compares.

   python tests/golden/make_large_golden.py      (~10-20 min on one core)
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
sys.path.insert(0, ROOT)

import refshim  # noqa: E402

refshim.install()

from bloqade.analog.atom_arrangement import Square  # noqa: E402
from bloqade.analog.emulate.ir.state_vector import StateVector  # noqa: E402

from bloqade_analog_b200 import adapter, lowering, serialize  # noqa: E402

DUR = [0.4, 3.2, 0.4]
OMEGA = [0, 15.8, 15.8, 0]
DELTA = [-16.33, -16.33, 16.33, 16.33]


def main():
    prog = (Square(4, lattice_spacing=5.0).rydberg.rabi.amplitude.uniform.piecewise_linear(DUR, OMEGA)
            .detuning.uniform.piecewise_linear(DUR, DELTA))
    (em,) = prog.bloqade.python().hamiltonian()
    h = em.hamiltonian
    t0 = time.time()
    (final,) = em.evolve(solver_name="dop853", atol=1e-7, rtol=1e-14)
    dt = time.time() - t0
    psi = final.data.copy()
    mine = adapter.convert_program(h.emulator_ir)
    p = np.abs(psi) ** 2
    n = h.space.n_atoms
    dens = np.array([p[(h.space.configurations >> a) & 1 == 1].sum() for a in range(n)])
    avg, var = h.average_and_variance(StateVector(psi.copy(), h.space))
    norm = float(np.linalg.norm(psi))
    sv = StateVector(psi.copy(), h.space)
    sv.normalize()
    np.random.seed(4242)
    uniforms = np.random.random_sample(200)
    np.random.seed(4242)
    shots = sv.sample(200)
    np.savez_compressed(os.path.join(HERE, "large_square4_a5.npz"), final_c64=psi.astype(np.complex64), density=dens,
                        norm=np.array(norm), avg_var=np.array([avg, var]), sample_uniforms=uniforms, sample_shots=shots,
                        record=lowering.pack_task_record(mine, 0), reference_seconds=np.array(dt))
    import json

    with open(os.path.join(HERE, "large_square4_a5.json"), "w") as f:
        json.dump(serialize.program_to_dict(mine), f)
    print("done", dt, "s norm", norm, "density", dens[:4])


if __name__ == "__main__":
    main()
