"""The drop-in end to end on the GPU, beside the UNMODIFIED reference in the same process:
`routine.install()` -> `program.bloqade.python().run(...)` / `.run_callback(...)` / `.hamiltonian()` on the B200
routine, compared with the reference's own `BloqadePythonRoutine` (src/bloqade/analog/ir/routine/bloqade.py:233-503)
on BASELINE configs 1, 2a, 2b and a slice of config 3.  The reference package comes from baseline/_ref on the GPU
box (baseline/install_ref.py) and from /root/reference in the build container."""
import numpy as np
import pytest
from conftest import cuda_available, infidelity, load_reference, reference_available

pytestmark = [
    pytest.mark.gpu,
    pytest.mark.skipif(not cuda_available(), reason="needs a CUDA device"),
    pytest.mark.skipif(not reference_available(), reason="reference package not present (baseline/_ref)"),
]

OMEGA = [0, 15.8, 15.8, 0]
DELTA = [-16.33, -16.33, 16.33, 16.33]


@pytest.fixture(scope="module")
def ref():
    load_reference()
    import bloqade.analog as ba

    return ba


@pytest.fixture()
def b200():
    from bloqade_analog_b200 import routine

    routine.install()
    yield routine
    routine.uninstall()


def adiabatic(geometry, durations, delta=DELTA):
    return (geometry.rydberg.rabi.amplitude.uniform.piecewise_linear(durations, OMEGA)
            .detuning.uniform.piecewise_linear(durations, delta))


def config1(ref, durations=(0.1, 0.4, 0.1), spacings=np.arange(4.0, 7.0, 0.5)):
    """BASELINE configs[0] with a shortened protocol (the reference needs 8 s per task at the full 4 us)"""
    from bloqade.analog.atom_arrangement import Square

    return adiabatic(Square(3, lattice_spacing="a"), list(durations)).batch_assign(a=list(spacings))


def config2(ref, variant, durations=(0.1, 0.5, 0.1)):
    from bloqade.analog.atom_arrangement import Chain

    chain = Chain(16, lattice_spacing=5.72)
    if variant == "a":
        return adiabatic(chain, list(durations))
    prog = chain.rydberg.detuning.uniform.piecewise_linear(list(durations), DELTA)
    prog = prog.rabi.amplitude
    for i in range(16):  # distinct waveforms -> one IndexMapping term per atom in the reference
        om = [0, round(15.8 + 0.01 * i, 6), round(15.8 + 0.01 * i, 6), 0]
        prog = prog.location(i).piecewise_linear(list(durations), om)
    return prog


def config3_slice(ref, n_det, n_dur, t_max=1.0):
    from bloqade.analog import cast
    from bloqade.analog.atom_arrangement import Chain

    total = cast("total")
    durations = [0.1 * total, 0.8 * total, 0.1 * total]
    prog = (Chain(12, lattice_spacing=5.72).rydberg.rabi.amplitude.uniform.piecewise_linear(durations, OMEGA)
            .detuning.uniform.piecewise_linear(durations, [-16.33, -16.33, "final_detuning", "final_detuning"]))
    dets = np.linspace(4.0, 32.0, n_det)
    durs = np.linspace(0.3, t_max, n_dur)
    grid = [(float(round(d, 6)), float(round(t, 6))) for d in dets for t in durs]
    return prog.batch_assign(final_detuning=[g[0] for g in grid], total=[g[1] for g in grid])


def run_both(program, shots, **kw):
    """the same seeded call on the reference routine and (after install()) on the B200 routine"""
    from bloqade_analog_b200 import routine

    routine.uninstall()
    r_ref = program.bloqade.python()
    assert type(r_ref).__name__ == "BloqadePythonRoutine"
    np.random.seed(9123892)
    ref_batch = r_ref.run(shots, **kw)
    routine.install()
    try:
        r_new = program.bloqade.python()
        assert type(r_new).__name__ == "BloqadeB200Routine"
        np.random.seed(9123892)
        new_batch = r_new.run(shots, **kw)
    finally:
        routine.uninstall()
    return ref_batch, new_batch


def assert_same_batches(ref_batch, new_batch, max_diff_shots_per_task=1):
    """same LocalBatch / Report types and the seeded shots of the reference (a CDF value within one ulp of a
    uniform draw may move a single shot)"""
    assert type(new_batch).__mro__[-1] is object and isinstance(new_batch, type(ref_batch))
    assert list(new_batch.tasks.keys()) == list(ref_batch.tasks.keys())
    rr, rn = ref_batch.report(), new_batch.report()
    assert isinstance(rn, type(rr))
    br, bn = rr.bitstrings(), rn.bitstrings()
    assert len(br) == len(bn)
    for a, b in zip(br, bn):
        assert a.shape == b.shape and a.dtype == b.dtype
        assert (a != b).any(axis=1).sum() <= max_diff_shots_per_task
    assert rn.dataframe.index.names == rr.dataframe.index.names
    assert list(rn.dataframe.index[:5]) == list(rr.dataframe.index[:5])
    for (kr, tr), (kn, tn) in zip(ref_batch.tasks.items(), new_batch.tasks.items()):
        assert tn.metadata == tr.metadata
        assert tn.geometry.filling == tr.geometry.filling
        assert len(tn.result().shot_outputs) == len(tr.result().shot_outputs)
        assert tn.result().shot_outputs[0].pre_sequence == tr.result().shot_outputs[0].pre_sequence


def test_config1_run_matches_reference(ref):
    ref_batch, new_batch = run_both(config1(ref), 100)
    assert len(new_batch.tasks) == 6
    assert_same_batches(ref_batch, new_batch)


@pytest.mark.parametrize("variant", ["a", "b"])
def test_config2_blockade_subspace_run_matches_reference(ref, variant):
    ref_batch, new_batch = run_both(config2(ref, variant), 100, blockade_radius=6.0)
    assert_same_batches(ref_batch, new_batch)
    # no shot violates the nearest-neighbour blockade (1 = ground, 0 = Rydberg)
    (bits,) = new_batch.report().bitstrings()
    assert not ((bits[:, :-1] == 0) & (bits[:, 1:] == 0)).any()


def test_config3_slice_run_matches_reference(ref, b200):
    """64 sweep tasks on the B200 routine; the reference runs every 8th of them serially (same seeded draws on
    that sub-batch)"""
    full = config3_slice(ref, 8, 8)
    np.random.seed(3)
    batch = full.bloqade.python().run(100)
    assert len(batch.tasks) == 64
    bits = batch.report().bitstrings()
    assert len(bits) == 64 and all(b.shape == (100, 12) for b in bits)
    sub = config3_slice(ref, 2, 4)
    ref_batch, new_batch = run_both(sub, 100)
    assert_same_batches(ref_batch, new_batch)


def state_callback(register, metadata, hamiltonian):
    return register.data, hamiltonian.average(register), metadata


@pytest.mark.parametrize("interaction_picture", [False, True])
def test_run_callback_matches_reference(ref, interaction_picture):
    """final states of run_callback: infidelity < 1e-8 against the reference's scipy integration (north_star)"""
    from bloqade_analog_b200 import routine

    program = config1(ref, durations=(0.1, 0.6, 0.1), spacings=[4.0, 5.0, 6.5])
    kw = dict(interaction_picture=interaction_picture, atol=1e-10, rtol=1e-10)
    ref_out = program.bloqade.python().run_callback(state_callback, **kw)
    routine.install()
    try:
        new_out = program.bloqade.python().run_callback(state_callback, **kw)
    finally:
        routine.uninstall()
    assert len(new_out) == len(ref_out) == 3
    for (s_ref, avg_ref, meta_ref), (s_new, avg_new, meta_new) in zip(ref_out, new_out):
        assert infidelity(s_ref, s_new) < 1e-8
        assert abs(avg_ref - avg_new) <= 1e-7 * max(1.0, abs(avg_ref))
        assert meta_new == meta_ref


def test_run_callback_equals_direct_gate_apply(ref, b200):
    """restates tests/test_run_callback.py:14-35 of the reference on the B200 routine: run_callback returns
    exactly what a direct AnalogGate.apply of the same Hamiltonian returns"""
    from bloqade.analog import start

    from bloqade_analog_b200 import adapter
    from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen
    from bloqade_analog_b200.state_vector import AnalogGate

    program = (start.add_position((0, 0)).add_position((0, 6.1)).rydberg.detuning.uniform.constant(15, 0.1)
               .rabi.amplitude.uniform.constant(15, 0.1))
    (task_data,) = list(program.bloqade.python()._generate_ir((), 0.0, "interpret", False))
    h = RydbergHamiltonianCodeGen().emit(adapter.convert_program(task_data.emulator_ir))
    (expected,) = AnalogGate(h).apply(h.space.zero_state(np.complex128))
    (single,) = program.bloqade.python().run_callback(lambda reg, *_: reg.data)
    (multi,) = program.bloqade.python().run_callback(lambda reg, *_: reg.data, multiprocessing=True, num_workers=1)
    np.testing.assert_equal(expected.data, single)
    np.testing.assert_equal(expected.data, multi)


def test_run_callback_exceptions_like_reference(ref, b200):
    """tests/test_run_callback.py:42-70 of the reference"""
    from bloqade.analog import start

    def boom(*args):
        raise ValueError

    program = (start.add_position((0, 0)).add_position((0, 6.1)).rydberg.detuning.uniform.constant(15, 1)
               .rabi.amplitude.uniform.constant(15, 1))
    with pytest.raises(RuntimeError):
        program.bloqade.python().run_callback(boom)
    with pytest.raises(RuntimeError):
        program.bloqade.python().run_callback(boom, multiprocessing=True, num_workers=1)
    (res,) = program.bloqade.python().run_callback(boom, ignore_exceptions=True)
    assert isinstance(res, ValueError)


@pytest.mark.parametrize("n_atoms", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("phi", [0.0, 2 * np.pi / 3, 4 * np.pi / 3])
@pytest.mark.parametrize("interaction", [True, False])
def test_solution_against_exact_diagonalisation(ref, b200, n_atoms, phi, interaction):
    """restates tests/test_emulator_interface.py:42-98 of the reference on the B200 routine: evolved state,
    average and variance at 101 times against exact diagonalisation of the constant Hamiltonian, both pictures"""
    from math import isclose

    from bloqade.analog.atom_arrangement import Chain

    program = Chain(n_atoms, lattice_spacing=6).rydberg.rabi.amplitude.uniform.constant(2 * np.pi, 4.0)
    if phi != 0:
        program = program.phase.uniform.constant(phi, 4)
    (emu,) = program.bloqade.python().hamiltonian()
    times = np.linspace(0, 4, 101)
    states = list(emu.evolve(times=times, interaction_picture=interaction, atol=1e-10, rtol=1e-14))
    h = emu.hamiltonian.tocsr(0)
    e, v = np.linalg.eigh(h.toarray())
    psi0 = emu.zero_state().data
    assert len(states) == 101
    for state, time in zip(states, times):
        assert str(state)
        expected = v @ (np.diag(np.exp(-1j * e * time)) @ (v.T.conj() @ psi0))
        exp_avg = np.vdot(expected, h.dot(expected)).real
        exp_var = (np.vdot(expected, (h @ h).dot(expected)) - exp_avg**2).real
        avg = emu.hamiltonian.average(state, time=time)
        var = emu.hamiltonian.variance(state, time=time)
        avg2, var2 = emu.hamiltonian.average_and_variance(state, time=time)
        assert isclose(abs(np.vdot(state.data, expected)), 1.0, abs_tol=1e-7), f"state at {time}"
        assert abs(np.vdot(state.data, expected) - 1.0) < 1e-7, f"overlap phase at {time}"
        assert isclose(avg, exp_avg, rel_tol=1e-7, abs_tol=1e-7) and isclose(avg2, exp_avg, rel_tol=1e-7, abs_tol=1e-7)
        assert isclose(var, exp_var, rel_tol=1e-7, abs_tol=1e-7) and isclose(var2, exp_var, rel_tol=1e-7, abs_tol=1e-7)


def test_zero_and_fock_states_like_reference(ref, b200):
    """tests/test_emulator_interface.py:11-39 of the reference"""
    from itertools import product

    from bloqade.analog import start

    (emu,) = (start.add_position([(0, 0), (0, 5), (0, 10)]).rydberg.rabi.amplitude.uniform.constant(15.0, 4.0)
              .detuning.uniform.constant(1.0, 4.0).bloqade.python().hamiltonian())
    data = np.zeros(8)
    data[0] = 1
    assert np.array_equal(emu.zero_state().data, data)
    for idx, bitstring in enumerate(product("gr", repeat=3)):
        sv = emu.fock_state("".join(bitstring[::-1]))
        data = np.zeros(8)
        data[idx] = 1
        assert np.array_equal(sv.data, data)


def test_call_signature_and_errors_like_reference(ref, b200):
    import inspect

    from bloqade.analog.ir.routine.bloqade import BloqadePythonRoutine

    r = config1(ref).bloqade.python()
    for name in ("run", "__call__", "run_callback", "hamiltonian", "_compile"):
        want = [p for p in inspect.signature(getattr(BloqadePythonRoutine, name)).parameters.values()]
        got = [p for p in inspect.signature(getattr(type(r), name)).parameters.values()]
        got_ref_part = [(p.name, p.kind, p.default) for p in got[: len(want)]]
        assert got_ref_part == [(p.name, p.kind, p.default) for p in want], name
        assert all(p.kind is p.KEYWORD_ONLY and p.default is not p.empty for p in got[len(want):]), name
    with pytest.raises(ValueError):
        r.run(10, multiprocessing=True, cache_matrices=True)
    with pytest.raises(NotImplementedError):
        r.run(10, solver_name="lsoda")


def _fn_programs():
    """the `.fn(...)` programs of the reference's tests/test_python_codegen.py:270-286 and tests/test_python_emulator.py:85,
    and a Gaussian-smoothed drive"""
    from bloqade.analog import piecewise_linear, start
    from bloqade.analog.ir.control.waveform import GaussianKernel

    def phase_function(t):
        return np.sin(t)

    p1 = (start.add_position([(0, 0), (0, 6)])
          .rydberg.detuning.uniform.piecewise_linear([0.1, 0.8, 0.1], [-10, -10, 20, 20])
          .amplitude.uniform.piecewise_linear([0.1, 0.8, 0.1], [0, 15.7, 15.7, 0])
          .phase.location(0, 1.0).fn(phase_function, 1.0)
          .detuning.scale([0.1, 1.0]).piecewise_constant([0.1, 0.3, 0.1], [0, 1, 2]).poly([2, -1, 3, -4], 0.5))
    p2 = (start.add_position((0, 0)).add_position((0, 6.0))
          .rydberg.detuning.uniform.piecewise_linear([0.1, 1.0, 0.1], [-20, -20, 20, 20])
          .amplitude.uniform.piecewise_linear([0.1, 1.0, 0.1], [0, 10, 10, 0])
          .amplitude.location(1).fn(lambda t: 4 * np.sin(3 * t), 1.2))
    smooth = piecewise_linear([0.1, 0.4, 0.1], [0, 12, 12, 0]).smooth(0.05, GaussianKernel)
    p3 = (start.add_position([(0, 0), (0, 6.2), (6.2, 0)])
          .rydberg.rabi.amplitude.uniform.apply(smooth)
          .detuning.uniform.piecewise_linear([0.1, 0.4, 0.1], [-12, -12, 14, 14]))
    return [("codegen_fn_phase", p1, 1e-8), ("emulator_fn_amplitude", p2, 1e-8), ("gaussian_smooth", p3, 1e-5)]


@pytest.mark.parametrize("which", [0, 1, 2])
def test_python_fn_and_smooth_waveforms_match_reference(ref, which):
    """PythonFn / Smooth waveforms (waveform.py:405-468,571-642) are lowered to a cubic-Hermite table for the device
    (tests/test_pyfn_lowering.py holds the table-vs-interpreter bounds); final states of run_callback against the
    reference's own integration: infidelity < 1e-8 for `.fn(...)` drives.  A Smooth waveform is evaluated by the
    reference with quad(epsabs=1e-4): its H(t) jitters by ~2e-3 rad/us from one RHS call to the next, the table
    interpolates between its knots, so the two answers are only comparable to ~(dOmega T)^2: bound 1e-5."""
    from bloqade_analog_b200 import routine

    name, program, bound = _fn_programs()[which]

    def collect(state, metadata, hamiltonian):
        return state.data.copy()

    (old,) = program.bloqade.python().run_callback(collect, atol=1e-10, rtol=1e-10)
    routine.install()
    try:
        r_new = program.bloqade.python()
        assert type(r_new).__name__ == "BloqadeB200Routine"
        (new,) = r_new.run_callback(collect, atol=1e-10, rtol=1e-10)
    finally:
        routine.uninstall()
    assert infidelity(old, new) < bound, name
