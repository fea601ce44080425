"""Generate tests/golden/*.npz + *.json by running the REFERENCE itself (imported read-only from
/root/reference through refshim) in the build container.  Commit the outputs; the GPU box has no
reference.  Run:  python tests/golden/make_golden.py

For every case it stores: the program in backend-IR JSON (adapter-converted from the reference's own
EmulatorProgram), the reference Space / van-der-Waals diagonal / dense-free operator fingerprints, RHS
values of RydbergHamiltonian._ode_complex_kernel(_int) on seeded random states, final states of
AnalogGate.apply for dop853 and dopri5 (this image's SciPy 1.18.1), observables and seeded samples.
"""
import json
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
sys.path.insert(0, ROOT)

import refshim  # noqa: E402

refshim.install()

from bloqade.analog import cast, piecewise_constant, piecewise_linear, start, var  # noqa: E402,F401
from bloqade.analog.atom_arrangement import Chain, Square  # noqa: E402
from bloqade.analog.emulate.ir.state_vector import AnalogGate, StateVector  # noqa: E402

from bloqade_analog_b200 import adapter, lowering, programs, serialize  # noqa: E402

DUR = [0.4, 3.2, 0.4]
OMEGA = [0, 15.8, 15.8, 0]
DELTA = [-16.33, -16.33, 16.33, 16.33]


def readme(lattice, durations=DUR, omega=OMEGA, delta=DELTA):
    return (lattice.rydberg.rabi.amplitude.uniform.piecewise_linear(durations, omega)
            .detuning.uniform.piecewise_linear(durations, delta))


def local_drive(lattice, n):
    p = lattice.rydberg.detuning.uniform.piecewise_linear(DUR, DELTA)
    for i in range(n):
        om = [round(v + (0.01 * i if v != 0 else 0.0), 6) for v in OMEGA]
        p = p.amplitude.location(i).piecewise_linear(DUR, om)
    return p


def cases():
    out = {}
    out["c1_square3_a5"] = dict(prog=readme(Square(3, lattice_spacing=5.0)), kw={})
    out["c1_square3_a6p5_short"] = dict(
        prog=readme(Square(3, lattice_spacing=6.5), [0.1, 0.6, 0.1]), kw={}, times=[0.25, 0.5, 0.8])
    out["c2a_chain16_blockade"] = dict(prog=readme(Chain(16, lattice_spacing=5.72)), kw=dict(blockade_radius=6.0))
    out["c2b_chain16_blockade_local"] = dict(prog=local_drive(Chain(16, lattice_spacing=5.72), 16), kw=dict(blockade_radius=6.0))
    grid = programs.config3_grid(64, 64)
    for tag, k in (("c3_task0000", 0), ("c3_task2047", 2047), ("c3_task4095", 4095)):
        d, t = grid[k]
        durs = [round(0.1 * t, 6), round(0.8 * t, 6), round(0.1 * t, 6)]
        out[tag] = dict(prog=readme(Chain(12, lattice_spacing=5.72), durs, OMEGA, [-16.33, -16.33, d, d]), kw={},
                        solvers=("dop853",))
    # phases, local detuning, slices/adds: the asymmetric (transition + transpose) path
    p = (Chain(4, lattice_spacing=6.1).rydberg.rabi.amplitude.uniform.piecewise_linear([0.2, 0.8, 0.2], [0, 9.0, 9.0, 0])
         .phase.uniform.piecewise_constant([0.5, 0.7], [0.3, -1.1])
         .detuning.uniform.constant(4.0, 1.2)
         .detuning.scale([0.1, 0.5, -0.3, 1.0]).piecewise_linear([0.6, 0.6], [0.0, 7.0, -2.0]))
    out["phase_chain4"] = dict(prog=p, kw={}, times=[0.4, 1.2])
    p = (Chain(5, lattice_spacing=5.5).rydberg.rabi.amplitude.location(1).constant(6.0, 1.0)
         .amplitude.location(3, 0.5).piecewise_linear([0.5, 0.5], [0, 8.0, 0])
         .phase.location(1).constant(0.7, 1.0).phase.location(3).poly([0.1, 0.9, -0.4], 1.0)
         .detuning.uniform.poly([1.0, -3.0, 2.0], 1.0))
    out["local_phase_chain5_blockade"] = dict(prog=p, kw=dict(blockade_radius=5.6))
    # three-level atoms: hyperfine + rydberg couplings
    p = (Chain(3, lattice_spacing=6.0).rydberg.rabi.amplitude.uniform.piecewise_linear([0.2, 0.6, 0.2], [0, 7.0, 7.0, 0])
         .phase.uniform.constant(0.4, 1.0)
         .detuning.uniform.piecewise_linear([0.5, 0.5], [-5.0, 5.0, 2.0])
         .hyperfine.rabi.amplitude.uniform.constant(5.0, 1.0)
         .detuning.location(1).constant(1.5, 1.0))
    out["three_level_chain3"] = dict(prog=p, kw=dict(use_hyperfine=True))
    p = (Chain(4, lattice_spacing=5.0).rydberg.rabi.amplitude.uniform.constant(6.0, 0.8)
         .detuning.uniform.constant(2.0, 0.8).hyperfine.rabi.amplitude.location(2).constant(4.0, 0.8)
         .phase.location(2).constant(-0.6, 0.8))
    out["three_level_chain4_blockade"] = dict(prog=p, kw=dict(use_hyperfine=True, blockade_radius=5.1))
    # vacant site: filling scatter + site-index mapping
    p = readme(start.add_position([(0, 0), (0, 5.3), (5.3, 0), (5.3, 5.3), (9.0, 9.0)], [True, False, True, True, True]),
               [0.2, 0.8, 0.2])
    out["vacancy_5sites"] = dict(prog=p, kw={})
    return out


def seed_of(name):
    return zlib.crc32(name.encode()) & 0x7FFFFFFF


def main():
    index = {}
    for name, case in cases().items():
        kw = dict(case["kw"])
        routine = case["prog"].bloqade.python()
        (em,) = routine.hamiltonian(**kw)
        h = em.hamiltonian
        dim = h.space.size
        mine = adapter.convert_program(h.emulator_ir)
        record = lowering.pack_task_record(mine, 0)
        rng = np.random.default_rng(seed_of(name))
        data = dict(
            configurations=h.space.configurations,
            is_full=np.array(h.space.space_type.value == "full_space"),
            rydberg=h.rydberg,
            record=record,
        )
        # waveform values (reference interpreter) on a grid including the breakpoints
        T = h.emulator_ir.duration
        grid_t = np.unique(np.concatenate([np.linspace(0, T, 41), rng.uniform(0, T, 16)]))
        wf_vals = []
        for fields in h.emulator_ir.pulses.values():
            for term in fields.detuning:
                wf_vals.append([term.amplitude.emit()(t) for t in grid_t])
            for term in fields.rabi:
                wf_vals.append([term.amplitude.emit()(t) for t in grid_t])
                if term.phase is not None:
                    wf_vals.append([term.phase.emit()(t) for t in grid_t])
        data["wf_times"] = grid_t
        data["wf_values"] = np.array(wf_vals, dtype=np.float64)
        # RHS parity vectors
        psi = rng.normal(size=dim) + 1j * rng.normal(size=dim)
        psi /= np.linalg.norm(psi)
        ts = np.concatenate([[0.0, T], rng.uniform(0, T, 4)])
        data["rhs_psi"] = psi
        data["rhs_times"] = ts
        data["rhs_out"] = np.array([h._ode_complex_kernel(t, psi, np.zeros(dim, complex)).copy() for t in ts])
        data["rhs_out_int"] = np.array([h._ode_complex_kernel_int(t, psi, np.zeros(dim, complex)).copy() for t in ts])
        data["happly_out"] = h._apply(psi, ts[-1]).copy()
        # evolution
        times = case.get("times", [])
        for solver in case.get("solvers", ("dop853", "dopri5")):
            states = [s.data.copy() for s in em.evolve(times=times, solver_name=solver)]
            data[f"final_{solver}"] = np.array(states)
        states = [s.data.copy() for s in em.evolve(times=times, interaction_picture=True)]
        data["final_dop853_int"] = np.array(states)
        data["times"] = np.array(times if times else [T], dtype=np.float64)
        final = StateVector(data["final_dop853"][-1].copy(), h.space)
        data["avg_var"] = np.array(h.average_and_variance(final))
        data["avg_var_t"] = np.array(h.average_and_variance(final, 0.37 * T))
        # local traces (atom indices = filled-site order)
        nl = h.space.atom_type.n_level
        op1 = rng.normal(size=(nl, nl)) + 1j * rng.normal(size=(nl, nl))
        op2 = rng.normal(size=(nl * nl, nl * nl)) + 1j * rng.normal(size=(nl * nl, nl * nl))
        filled = [i for i, f in enumerate(h.space.program_register.geometry.filling) if f]
        sa, sb = filled[0], filled[-1]
        data["lt_op1"], data["lt_op2"] = op1, op2
        data["lt_sites"] = np.array([sa, sb])
        data["lt_1"] = np.array([final.local_trace(op1, sa), final.local_trace(op1, sb)])
        data["lt_2"] = np.array([final.local_trace(op2, (sa, sb)), final.local_trace(op2, (sb, sa))]) if sa != sb else np.zeros(2, complex)
        # seeded sampling: record the uniforms np.random.choice consumes and the reference's shots
        final.normalize()
        np.random.seed(seed_of(name))
        uniforms = np.random.random_sample(200)
        np.random.seed(seed_of(name))
        data["sample_uniforms"] = uniforms
        data["sample_shots"] = final.sample(200)
        data["sample_state"] = final.data.copy()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **data)
        index[name] = dict(program=serialize.program_to_dict(mine), kwargs={k: (v if not isinstance(v, bool) else bool(v)) for k, v in kw.items()}, dim=int(dim))
        print(name, dim, {k: v.shape for k, v in data.items() if k.startswith("final")})
    with open(os.path.join(HERE, "index.json"), "w") as f:
        json.dump(index, f, indent=1)
    # cross-check the standalone program builders against the reference-derived records
    ck = {
        "c1_square3_a5": programs.adiabatic_program(__import__("bloqade_analog_b200").ir.square_sites(3, lattice_spacing=5.0)),
        "c2a_chain16_blockade": programs.config2("a"),
        # (c2b is not compared here: the reference emits its sixteen per-atom terms in set order; the
        #  builder's record is checked against golden RHS values in tests/test_hostsim_golden.py instead)
        "c3_task2047": programs.config3()[2047],
    }
    for name, prog in ck.items():
        ref = np.load(os.path.join(HERE, name + ".npz"))["record"]
        got = lowering.pack_task_record(prog, 0)
        assert np.array_equal(ref, got), f"builder mismatch for {name}"
    print("standalone builders reproduce the reference-derived task records")


if __name__ == "__main__":
    main()
