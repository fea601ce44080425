"""`-m gpu` parity tests: the CUDA library, called through its C ABI (ctypes, include/bqa_b200.h),
against (a) golden vectors produced by the reference itself (tests/golden/*.npz, see make_golden.py) and
(b) the CPU oracle (oracle/) on seeded inputs.  Bars: subspace / diagonal / shots bit-exact; single
H psi <= 1e-13 relative; final states fidelity deviation < 1e-8 (north_star); observables 1e-7.
"""
import os
import numpy as np
import pytest
from conftest import GOLDEN_CASES, SMALL_CASES, cuda_available, golden_case, infidelity

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not cuda_available(), reason="needs a CUDA device")]

from bloqade_analog_b200 import _lib, ir, lowering, programs  # noqa: E402
from bloqade_analog_b200.batch import evolve_programs, run_programs  # noqa: E402
from bloqade_analog_b200.hamiltonian import CompileCache, RydbergHamiltonianCodeGen  # noqa: E402
from bloqade_analog_b200.state_vector import AnalogGate, StateVector, make_plan  # noqa: E402

_HAMS = {}


def ham(name):
    if name not in _HAMS:
        c = golden_case(name)
        _HAMS[name] = RydbergHamiltonianCodeGen().emit(c.program)
    return _HAMS[name]


# ---- kernel (1): fused H(t) psi ------------------------------------------------------------------------
@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_apply_matches_reference(name):
    c, h = golden_case(name), ham(name)
    psi = c.data["rhs_psi"]
    for t, ref, ref_int in zip(c.data["rhs_times"], c.data["rhs_out"], c.data["rhs_out_int"]):
        scale = np.abs(ref).max()
        got = h._ode_complex_kernel(float(t), psi, np.zeros_like(psi))
        assert np.abs(got - ref).max() <= 1e-13 * scale
        got = h._ode_complex_kernel_int(float(t), psi, np.zeros_like(psi))
        assert np.abs(got - ref_int).max() <= 1e-12 * scale
    got = h._apply(psi, float(c.data["rhs_times"][-1]))
    assert np.abs(got - c.data["happly_out"]).max() <= 1e-13 * np.abs(c.data["happly_out"]).max()


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_device_rydberg_diagonal_bit_exact(name):
    c = golden_case(name)
    plan = make_plan(c.space)
    did = plan.build_rydberg(lowering.pair_interactions(c.program.register))
    assert np.array_equal(plan.get_diagonal(did), c.data["rydberg"])
    plan.close()


def test_csr_fallback_term_equals_bitflip_form():
    """SparseMatrixCSR / CSC fallback (SO:15-34): the same Rabi term fed as a generic CSR matrix
    (+ its transpose for the conjugate part) gives the same H psi as the recognised bit-flip form"""
    for name in ("phase_chain4", "local_phase_chain5_blockade", "c1_square3_a5"):
        c, h = golden_case(name), ham(name)
        prog = c.program
        (coupling, fields), = prog.pulses.items()
        term = fields.rabi[0]
        rest = ir.Fields(fields.detuning, fields.rabi[1:])
        prog2 = ir.EmulatorProgram(prog.register, prog.duration, {coupling: rest})
        plan = make_plan(c.space)
        did = plan.add_diagonal(c.data["rydberg"])
        # record of the remaining terms + the removed term's waveforms appended as extra waveforms
        table = lowering._WaveformTable()
        rec2 = lowering.pack_task_record(prog2, did, extra_waveforms=[term.amplitude] + ([term.phase] if term.phase is not None else []))
        n_wf = int(rec2[2])
        amp_wf = n_wf - (2 if term.phase is not None else 1)
        phase_wf = n_wf - 1 if term.phase is not None else -1
        from bloqade_analog_b200.state_vector import RabiOperator

        op = RabiOperator(dict(term.operator_data.target_atoms), 0, term.operator_data.operator_type, c.space,
                          term.amplitude, term.phase).op_tocsr()
        op.sort_indices()
        plan.add_csr(amp_wf, phase_wf, op.data, op.indices, op.indptr)
        psi = c.data["rhs_psi"]
        for t, ref in zip(c.data["rhs_times"], c.data["rhs_out"]):
            got = plan.apply(rec2, float(t), psi)
            assert np.abs(got - ref).max() <= 1e-13 * np.abs(ref).max(), name
        plan.close()


def test_apply_vs_oracle_18_atoms():
    """full 2^18 space, beyond anything the reference's tests reach: CUDA apply == oracle restatement"""
    from oracle import emulator as oe

    prog = programs.adiabatic_program(ir.square_sites(6, 3, lattice_spacing=5.3))
    h = RydbergHamiltonianCodeGen(device_diagonal_min_dim=1 << 10).emit(prog)
    oh = oe.Hamiltonian(prog)
    assert np.array_equal(h.rydberg, oh.rydberg)
    rng = np.random.default_rng(7)
    psi = rng.normal(size=h.space.size) + 1j * rng.normal(size=h.space.size)
    psi /= np.linalg.norm(psi)
    for t in (0.0, 0.2, 1.7, 3.9, 4.0):
        ref = oh.rhs(t, psi)
        got = h._ode_complex_kernel(t, psi, np.zeros_like(psi))
        assert np.abs(got - ref).max() <= 1e-13 * np.abs(ref).max()
        ref = oh.rhs_interaction(t, psi)
        got = h._ode_complex_kernel_int(t, psi, np.zeros_like(psi))
        assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max()


# ---- kernels (2)+(5): adaptive evolution -------------------------------------------------------------------
@pytest.mark.parametrize("engine", ["cluster", "cta", "streamed"])
@pytest.mark.parametrize("solver", ["dop853", "dopri5"])
@pytest.mark.parametrize("name", SMALL_CASES)
def test_evolve_small_cases(name, solver, engine):
    c, h = golden_case(name), ham(name)
    gate = AnalogGate(h)
    times = list(c.data["times"])
    # Hairer's published rejection rule (SciPy <= 1.16): different step sequence, same answer within the bar
    out = list(gate.apply(h.space.zero_state(np.complex128), solver_name=solver, times=times, engine=engine, reject_mode=0))
    assert len(out) == len(times)
    for a, b in zip(out, c.data[f"final_{solver}"]):
        assert infidelity(a.data, b) < 1e-8
    # default: same controller as this image's SciPy: states agree far below the solver's own error
    out = list(gate.apply(h.space.zero_state(np.complex128), solver_name=solver, times=times, engine=engine))
    for a, b in zip(out, c.data[f"final_{solver}"]):
        assert np.abs(a.data - b).max() < 2e-8
        assert infidelity(a.data, b) < 1e-13


@pytest.mark.parametrize("engine", ["cluster", "cta", "streamed"])
@pytest.mark.parametrize("name", SMALL_CASES)
def test_evolve_interaction_picture(name, engine):
    c, h = golden_case(name), ham(name)
    out = list(AnalogGate(h).apply(h.space.zero_state(np.complex128), times=list(c.data["times"]),
                                   interaction_picture=True, engine=engine))
    for a, b in zip(out, c.data["final_dop853_int"]):
        assert infidelity(a.data, b) < 1e-8


@pytest.mark.parametrize("name", ["c2a_chain16_blockade", "c2b_chain16_blockade_local", "c3_task0000", "c3_task2047",
                                  "c3_task4095"])
def test_evolve_config_cases(name):
    """BASELINE.json configs 2 (blockade subspace, uniform + sixteen local drives) and 3 (sweep tasks)"""
    c, h = golden_case(name), ham(name)
    (out,) = AnalogGate(h).apply(h.space.zero_state(np.complex128))
    assert infidelity(out.data, c.data["final_dop853"][-1]) < 1e-8
    (out,) = AnalogGate(h).apply(h.space.zero_state(np.complex128), interaction_picture=True)
    assert infidelity(out.data, c.data["final_dop853_int"][-1]) < 1e-8


@pytest.mark.parametrize("engine", ["streamed", "cta", "cluster"])
def test_all_engines_config2_and_3(engine):
    """every engine on the blockade subspace (dim 2584: cluster of 4 with a non-power-of-two chunk) and on a
    C3 task (dim 4096: cluster of 2)"""
    for name in ("c2a_chain16_blockade", "c3_task2047"):
        c, h = golden_case(name), ham(name)
        (out,) = AnalogGate(h).apply(h.space.zero_state(np.complex128), engine=engine)
        assert infidelity(out.data, c.data["final_dop853"][-1]) < 1e-8
        # the controller sits at err ~ 1 on these stability-limited problems: a different summation order of
        # the error norm can flip one borderline accept/reject and move the state within the solver's own
        # tolerance (atol = 1e-7), far inside the fidelity bar above
        assert np.abs(out.data - c.data["final_dop853"][-1]).max() < 1e-5


@pytest.mark.parametrize("name", ["c1_square3_a5", "c3_task0000", "c3_task4095"])
def test_cluster_fast_path_equals_generic_path(name, monkeypatch):
    """the fused DOP853 stage loop (bqa_cluster_fast.cuh) against the generic TaskEvolver on the same engine:
    same arithmetic in the same order, only the error-norm summation order differs"""
    c, h = golden_case(name), ham(name)
    z = h.space.zero_state(np.complex128)
    (fast,) = AnalogGate(h).apply(z, engine="cluster")
    monkeypatch.setenv("BQA_NO_FAST", "1")
    (gen,) = AnalogGate(h).apply(z, engine="cluster")
    monkeypatch.delenv("BQA_NO_FAST")
    assert infidelity(fast.data, gen.data) < 1e-8  # borderline accept/reject flips allowed, see above
    # amplitudes at the solver's noise floor (the all-Rydberg corner, |psi| ~ 1e-5 at atol = 1e-7) move with
    # every borderline accept/reject flip; the bar is the fidelity
    assert np.abs(fast.data - gen.data).max() < 5e-4
    assert infidelity(fast.data, c.data["final_dop853"][-1]) < 1e-8
    monkeypatch.setenv("BQA_FAST_K4", "1")  # alternative tiling (4 elements per thread) at chunk 1024
    (alt,) = AnalogGate(h).apply(z, engine="cluster")
    assert infidelity(alt.data, c.data["final_dop853"][-1]) < 1e-8
    monkeypatch.delenv("BQA_FAST_K4")


def test_cluster_fast_path_multiple_times_and_psi0():
    c, h = golden_case("c1_square3_a6p5_short"), ham("c1_square3_a6p5_short")
    out = list(AnalogGate(h).apply(h.space.zero_state(np.complex128), times=list(c.data["times"]), engine="cluster"))
    for a, b in zip(out, c.data["final_dop853"]):
        assert np.abs(a.data - b).max() < 2e-8


def test_cluster_engine_dim8192_cluster_of_4():
    """13-atom chain: 12 vectors x 128 KB -> cluster of 4; compared with the streamed engine and the oracle"""
    from oracle import emulator as oe

    prog = programs.adiabatic_program(ir.chain_sites(13, 5.9), [0.05, 0.3, 0.05])
    h = RydbergHamiltonianCodeGen().emit(prog)
    z = h.space.zero_state(np.complex128)
    (a,) = AnalogGate(h).apply(z, engine="cluster")
    (b,) = AnalogGate(h).apply(z, engine="streamed")
    ref = oe.evolve(prog)[-1]
    assert infidelity(a.data, ref) < 1e-8 and infidelity(b.data, ref) < 1e-8
    assert np.abs(a.data - b.data).max() < 1e-9


def test_evolve_from_nonzero_state_and_no_alias():
    c, h = golden_case("c1_square3_a6p5_short"), ham("c1_square3_a6p5_short")
    from oracle import emulator as oe

    rng = np.random.default_rng(3)
    psi0 = rng.normal(size=512) + 1j * rng.normal(size=512)
    psi0 /= np.linalg.norm(psi0)
    keep = psi0.copy()
    (out,) = AnalogGate(h).apply(StateVector(psi0, h.space), atol=1e-10, rtol=1e-10)
    assert np.array_equal(psi0, keep)  # the caller's initial state is never mutated
    ref = oe.evolve(c.program, psi0=keep, atol=1e-10, rtol=1e-10)[-1]
    assert infidelity(out.data, ref) < 1e-12


def test_error_codes():
    c, h = golden_case("c1_square3_a6p5_short"), ham("c1_square3_a6p5_short")
    with pytest.raises(RuntimeError, match="Larger nsteps is needed"):
        list(AnalogGate(h).apply(h.space.zero_state(np.complex128), nsteps=5))
    with pytest.raises(RuntimeError, match="Larger nsteps is needed"):
        list(AnalogGate(h).apply(h.space.zero_state(np.complex128), nsteps=5, engine="streamed"))
    with pytest.raises(ValueError):
        list(AnalogGate(h).apply(h.space.zero_state(np.complex128), times=[h.emulator_ir.duration * 2]))
    with pytest.raises(ValueError):
        list(AnalogGate(h).apply(h.space.zero_state(np.complex128), solver_name="rk4"))
    with pytest.raises(NotImplementedError):
        list(AnalogGate(h).apply(h.space.zero_state(np.complex128), solver_name="lsoda"))


# ---- kernel (3): reductions ----------------------------------------------------------------------------------
@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_observables(name):
    c, h = golden_case(name), ham(name)
    final = StateVector(c.data["final_dop853"][-1].copy(), h.space)
    assert abs(final.norm() - np.linalg.norm(final.data)) < 1e-14
    avg, var = h.average_and_variance(final)
    np.testing.assert_allclose([avg, var], c.data["avg_var"], rtol=1e-9, atol=1e-9)
    T = h.emulator_ir.duration
    avg, var = h.average_and_variance(final, 0.37 * T)
    np.testing.assert_allclose([avg, var], c.data["avg_var_t"], rtol=1e-9, atol=1e-9)
    sa, sb = (int(x) for x in c.data["lt_sites"])
    op1, op2 = c.data["lt_op1"], c.data["lt_op2"]
    got = np.array([final.local_trace(op1, sa), final.local_trace(op1, sb)])
    np.testing.assert_allclose(got, c.data["lt_1"], rtol=1e-12, atol=1e-13)
    if sa != sb:
        got = np.array([final.local_trace(op2, (sa, sb)), final.local_trace(op2, (sb, sa))])
        np.testing.assert_allclose(got, c.data["lt_2"], rtol=1e-12, atol=1e-13)
    psi = c.data["rhs_psi"].copy() * 3.7
    sv = StateVector(psi, h.space)
    sv.normalize()
    assert abs(np.linalg.norm(sv.data) - 1.0) < 1e-14


def test_local_trace_errors():
    h = ham("vacancy_5sites")
    sv = StateVector(golden_case("vacancy_5sites").data["final_dop853"][-1].copy(), h.space)
    with pytest.raises(ValueError):
        sv.local_trace(np.eye(3), 0)
    with pytest.raises(ValueError):
        sv.local_trace(np.eye(2), 7)
    with pytest.raises(ValueError, match="empty"):
        sv.local_trace(np.eye(2), 1)


# ---- kernel (4): Born sampling ---------------------------------------------------------------------------------
@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_sampling_bit_exact_with_reference(name):
    """np.random.choice semantics (SP:211-229): same uniforms -> the reference's shots, bit for bit"""
    c, h = golden_case(name), ham(name)
    plan = h.plan
    got = plan.sample(c.data["sample_state"], c.data["sample_uniforms"])[0]
    assert got.dtype == np.uint8
    assert np.array_equal(got, c.data["sample_shots"])
    # the public route draws the same uniforms from NumPy's global RandomState
    import zlib

    np.random.seed(zlib.crc32(name.encode()) & 0x7FFFFFFF)
    got = StateVector(c.data["sample_state"].copy(), h.space).sample(200)
    assert np.array_equal(got, c.data["sample_shots"])


@pytest.mark.parametrize("n,shots", [(20, 5000), (22, 512)])
def test_sampling_large_cdf_vs_numpy(n, shots):
    """(20, 5000): CDF array + binary search; (22, 512): the two-level form for large states with few shots (tile
    offsets searched, one tile re-evaluated per shot, no CDF array)"""
    dim = 1 << n
    rng = np.random.default_rng(11)
    psi = (rng.normal(size=dim) + 1j * rng.normal(size=dim)) * rng.random(dim) ** 4
    psi /= np.linalg.norm(psi)
    u = rng.random(shots)
    plan = _lib.Plan(n, 2, dim)
    got = plan.sample(psi, u)[0]
    cdf = np.cumsum(np.abs(psi) ** 2)
    cdf /= cdf[-1]
    idx = np.searchsorted(cdf, u, side="right")
    ref = ((idx[:, None] >> np.arange(n)[None, :]) & 1).astype(np.uint8)
    # the device scan sums in a different order: a shot may move to the neighbouring state only when the
    # uniform falls within rounding of a CDF edge
    bad = np.nonzero((got != ref).any(axis=1))[0]
    for s in bad:
        assert np.min(np.abs(cdf - u[s])) < 1e-12
    assert bad.size <= 2
    plan.close()


def test_sampling_more_states_than_one_grid_dimension():
    """70 000 small states in one call: the segment dimension of the batched kernels is launched in chunks of 65 535"""
    n, n_states, shots = 4, 70000, 3
    dim = 1 << n
    rng = np.random.default_rng(5)
    psi = rng.normal(size=(n_states, dim)) + 1j * rng.normal(size=(n_states, dim))
    u = rng.random((n_states, shots))
    plan = _lib.Plan(n, 2, dim)
    got = plan.sample(psi, u)
    cdf = np.cumsum(np.abs(psi) ** 2, axis=1)
    cdf /= cdf[:, -1:]
    idx = (cdf[:, None, :] <= u[:, :, None]).sum(axis=2)
    ref = ((idx[:, :, None] >> np.arange(n)[None, None, :]) & 1).astype(np.uint8)
    assert (got != ref).any(axis=2).sum() <= 2
    assert np.array_equal(got[-1], ref[-1]) and np.array_equal(got[65535], ref[65535])
    plan.close()


# ---- AnalogGate.run / batched run -----------------------------------------------------------------------------
def test_run_batch_config3_slice():
    """a slice of the C3 sweep through the batched entry point == per-task reference final states,
    shots == oracle sampling of the reference state with the same uniforms"""
    from oracle import space as ospace

    names = ["c3_task0000", "c3_task2047", "c3_task4095"]
    progs = [golden_case(n).program for n in names]
    states, stats = evolve_programs(progs)
    for n, s, st in zip(names, states, stats):
        assert st["status"] == 1
        assert infidelity(s, golden_case(n).data["final_dop853"][-1]) < 1e-8
    np.random.seed(5)
    uniforms = np.stack([np.random.random_sample(100) for _ in names])
    np.random.seed(5)
    res = run_programs(progs, 100)
    for k, (n, r) in enumerate(zip(names, res)):
        assert r.status == 1 and r.shots.shape == (100, 12)
        psi = states[k] / np.linalg.norm(states[k])
        ref = ospace.sample_state_vector(golden_case(n).space.configurations, psi, uniforms[k], 12, 2)
        assert (r.shots != ref).any(axis=1).sum() <= 1


def test_config3_statuses_match_scipy():
    """ALL 4096 tasks of BASELINE configs[2] through the batched entry point against SciPy's own outcome
    (tests/golden/c3_status.json, make_c3_status.py): integrator status per task and the amount of work.
    dop853's stiffness exit needs 15 consecutive probes above 6.1 and on these stability-limited problems the probe
    hovers at 6.1, so a task can sit on a knife edge: task 1783 fails in SciPy only at atol = 1e-7 exactly and passes
    for atol * (1 +- 1e-12 ... 1e-9) (tests/golden/c3_knife_edge.json), the GPU passes it; task 951 is the same
    situation the other way round (profiles/r02_c3_knife_gpu.json: -4 at 1e-7 exactly, 1 for every perturbed atol;
    SciPy: 1).  Task 440 fails on both for every perturbation.  Any other mismatch fails this test.
    The controller takes SciPy's decisions step for step on short runs (test_hostsim_golden, the small goldens); over
    the ~1500 steps of a sweep task a last-bit difference in one error norm (tree sum here, serial sum in SciPy)
    changes the step sequence - SciPy's own RHS count moves by up to 30 % under an atol change of 1e-12 - so the
    per-task counts are compared in aggregate: SciPy re-evaluates the RHS once after every accepted and every
    rejected step, hence calls - 1 <-> nfcn + naccpt + nrejct."""
    import json

    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "golden", "c3_status.json")) as f:
        gold = json.load(f)
    progs = [programs.config3_task(d, t) for d, t in programs.config3_grid(64, 64)]
    np.random.seed(3)
    res = run_programs(progs, 4, raise_on_error=False)
    status = np.array([r.status for r in res])
    ref = np.array(gold["status"])
    mismatch = sorted(int(k) for k in np.nonzero(status != ref)[0])
    assert set(mismatch) <= {951, 1783}, mismatch
    assert status[440] == -4 and ref[440] == -4
    ok = (status == 1) & (ref == 1)
    nf = np.array([r.nfcn + r.naccpt + r.nrejct for r in res], dtype=np.float64)
    calls = np.array(gold["calls"], dtype=np.float64) - 1
    assert abs(nf[ok].sum() / calls[ok].sum() - 1.0) < 0.05
    assert np.median(np.abs(nf[ok] / calls[ok] - 1.0)) < 0.1


def test_failed_task_does_not_poison_batch():
    progs = [golden_case("c1_square3_a6p5_short").program, golden_case("c1_square3_a5").program]
    res = run_programs(progs, 10, nsteps=2000, raise_on_error=False)
    assert res[0].status == 1 and res[1].status == -2
    with pytest.raises(RuntimeError, match="Larger nsteps"):
        run_programs(progs, 10, nsteps=2000)


@pytest.mark.parametrize("engine", ["cta", "cluster", "streamed"])
def test_output_times_after_a_failure_are_marked(engine):
    """three output times, nsteps (400) too small for the second: the first entry succeeds, the second carries the
    integrator's status (-2) and every later one is -1 (not "success" with a stale state) on every engine"""
    h = ham("c1_square3_a5")
    d = h.emulator_ir.duration
    opts = _lib.make_opts("dop853", 1e-7, 1e-14, 400, False, 1, engine)
    blob, offsets = lowering.concat_records([h.record])
    _, stats = h.plan.evolve_batch(blob, offsets, np.array([[0.02 * d, 0.7 * d, d]]), opts)
    st = np.asarray(stats).reshape(-1)
    assert [int(x) for x in st["status"]] == [1, -2, -1], st["status"]


@pytest.mark.parametrize("engine", ["cta", "cluster", "streamed"])
def test_resume_from_an_output_time_is_bit_identical(engine):
    """bqa_solver_opts::t_start: evolve to t1, take the state, resume from (t1, state) to T == evolve with
    times = [t1, T] (SciPy's ode restarts at every output time, state_vector.py:589-600) - the mid-run checkpoint
    of the streamed / sharded engines rests on this"""
    h = ham("c1_square3_a6p5_short")
    d = h.emulator_ir.duration
    blob, offsets = lowering.concat_records([h.record])
    o = _lib.make_opts("dop853", 1e-7, 1e-14, engine=engine)
    both, st = h.plan.evolve_batch(blob, offsets, np.array([[0.4 * d, d]]), o)
    first, _ = h.plan.evolve_batch(blob, offsets, np.array([[0.4 * d]]), o)
    assert np.array_equal(first[0, 0], both[0, 0])
    o2 = _lib.make_opts("dop853", 1e-7, 1e-14, engine=engine, t_start=0.4 * d)
    rest, st2 = h.plan.evolve_batch(blob, offsets, np.array([[d]]), o2, psi0=first[0, 0])
    assert np.array_equal(rest[0, 0], both[0, 1])
    assert int(st2["nfcn"][0, 0]) == int(st["nfcn"][0, 1]) and int(st2["status"][0, 0]) == 1


def test_gate_run_seeded_matches_reference_shots():
    c, h = golden_case("c1_square3_a6p5_short"), ham("c1_square3_a6p5_short")
    import zlib

    # golden shots were drawn from the state at the last requested time of this case's program run to `times[-1]`;
    # run() goes to `duration`, so compare with the oracle instead
    from oracle import emulator as oe

    np.random.seed(17)
    u = np.random.random_sample(64)
    np.random.seed(17)
    got = AnalogGate(h).run(64, atol=1e-7, rtol=1e-14)
    ref = oe.run(c.program, 64, uniforms=u)
    assert (got != ref).any(axis=1).sum() <= 1


# ---- large states: size-independent properties --------------------------------------------------------------
@pytest.mark.parametrize("n_atoms", [22, 25])
def test_large_apply_properties(n_atoms):
    """full 2^22 / 2^25 (BASELINE config 4 size): Hermiticity <phi|H psi> = conj <psi|H phi>, linearity,
    and agreement of a slice of rows with the oracle formula evaluated directly"""
    import torch

    l1, l2 = (5, 5) if n_atoms == 25 else (11, 2)
    prog = programs.adiabatic_program(ir.square_sites(l1, l2, lattice_spacing=5.0))
    h = RydbergHamiltonianCodeGen().emit(prog)
    dim = h.space.size
    g = torch.Generator(device="cuda").manual_seed(1)
    psi = torch.randn(dim, 2, dtype=torch.float64, device="cuda", generator=g)
    phi = torch.randn(dim, 2, dtype=torch.float64, device="cuda", generator=g)
    psi /= psi.norm()
    phi /= phi.norm()
    out1 = torch.empty_like(psi)
    out2 = torch.empty_like(psi)
    t = 1.3
    D = _lib.MEM_DEVICE
    h.plan.apply(h.record, t, psi.data_ptr(), out1.data_ptr(), raw=True, mem=D)
    h.plan.apply(h.record, t, phi.data_ptr(), out2.data_ptr(), raw=True, mem=D)
    torch.cuda.synchronize()
    vp, vh = torch.view_as_complex(phi), torch.view_as_complex(psi)
    a = torch.vdot(vp, torch.view_as_complex(out1))
    b = torch.vdot(vh, torch.view_as_complex(out2)).conj()
    scale = float(torch.view_as_complex(out1).abs().max())
    assert abs(complex(a) - complex(b)) < 1e-10 * max(1.0, abs(complex(a)))
    # linearity
    comb = (0.3 * psi - 1.7 * phi).contiguous()
    out3 = torch.empty_like(psi)
    h.plan.apply(h.record, t, comb.data_ptr(), out3.data_ptr(), raw=True, mem=D)
    torch.cuda.synchronize()
    assert float((out3 - (0.3 * out1 - 1.7 * out2)).abs().max()) < 1e-12 * scale
    # direct formula on a few rows
    V = h.plan.get_diagonal(h.diag_id)
    psi_h = torch.view_as_complex(psi).cpu().numpy()
    out_h = torch.view_as_complex(out1).cpu().numpy()
    omega = prog.pulses[ir.LevelCoupling.Rydberg].rabi[0].amplitude(t)
    delta = prog.pulses[ir.LevelCoupling.Rydberg].detuning[0].amplitude(t)
    rng = np.random.default_rng(0)
    for i in list(rng.integers(0, dim, 64)) + [0, dim - 1]:
        i = int(i)
        ref = (V[i] - delta * bin(i).count("1")) * psi_h[i]
        for a_ in range(n_atoms):
            ref += 0.5 * omega * psi_h[i ^ (1 << a_)]
        assert abs(out_h[i] - ref) <= 1e-12 * max(1.0, abs(ref))


@pytest.mark.parametrize("n_atoms", [14, 15, 16, 19, 24, 25, 26, 27])
def test_tiled_apply_equals_one_pass(n_atoms, monkeypatch):
    """the shared-memory-tiled H psi (bqa_apply_tiled.cuh: 1, 2 or 3 passes depending on n) against the one-pass
    gather kernel on the same inputs: per-atom Rabi amplitudes with a phase (complex coefficients) and a local
    detuning pattern, raw and -i forms.  Only the summation order differs."""
    import torch
    from decimal import Decimal

    phase = ir.piecewise_linear([1.0, 3.0], [0.0, 1.1, 2.3])
    for ph in (phase, None):
        prog = programs.adiabatic_program(ir.chain_sites(n_atoms, lattice_spacing=5.1), durations=[1.0, 2.0, 1.0],
                                          local_drive=ph is not None, phase=ph)
        fields = prog.pulses[ir.LevelCoupling.Rydberg]
        pattern = {a: Decimal(str(0.1 + 0.07 * ((a * 5) % 11))) for a in range(n_atoms)}
        fields.detuning.append(ir.DetuningTerm(ir.DetuningOperatorData(pattern), ir.piecewise_linear([4.0], [-3.0, 8.0])))
        h = RydbergHamiltonianCodeGen().emit(prog)
        dim = h.space.size
        g = torch.Generator(device="cuda").manual_seed(n_atoms)
        psi = torch.randn(dim, 2, dtype=torch.float64, device="cuda", generator=g)
        psi /= psi.norm()
        D = _lib.MEM_DEVICE
        for raw in (True, False):
            tiled = torch.full_like(psi, float("nan"))
            plain = torch.full_like(psi, float("nan"))
            n0 = _lib.launch_count()
            h.plan.apply(h.record, 1.7, psi.data_ptr(), tiled.data_ptr(), raw=raw, mem=D)
            n_tiled = _lib.launch_count() - n0
            monkeypatch.setenv("BQA_NO_TILED", "1")
            h.plan.apply(h.record, 1.7, psi.data_ptr(), plain.data_ptr(), raw=raw, mem=D)
            monkeypatch.delenv("BQA_NO_TILED")
            torch.cuda.synchronize()
            assert n_tiled == (1 if n_atoms <= 14 else 2 if n_atoms <= 26 else 3)
            scale = float(plain.abs().max())
            assert float((tiled - plain).abs().max()) <= 1e-13 * scale
        del psi, tiled, plain
        h.plan.close()


# ---- sharded state: shard kernels against the unsharded path on ONE device ----------------------
@pytest.mark.parametrize("world,n_atoms", [(2, 15), (2, 18), (4, 17), (8, 18), (2, 26)])
def test_sharded_apply_equals_unsharded(world, n_atoms):
    """every rank's (H psi)_local from bqa_apply_sharded - partner shards passed as plain device pointers, the role
    NVLink peer mappings play on a multi-GPU box - equals the slice of the unsharded H psi; per-atom complex drive,
    local detuning pattern, van-der-Waals diagonal built per shard"""
    import torch
    from decimal import Decimal
    from bloqade_analog_b200.lowering import pack_task_record, pair_interactions

    g = world.bit_length() - 1
    n_loc = n_atoms - g
    phase = ir.piecewise_linear([1.0, 3.0], [0.0, 1.1, 2.3])
    prog = programs.adiabatic_program(ir.chain_sites(n_atoms, lattice_spacing=5.1), durations=[1.0, 2.0, 1.0],
                                      local_drive=True, phase=phase)
    pattern = {a: Decimal(str(0.1 + 0.07 * ((a * 5) % 11))) for a in range(n_atoms)}
    prog.pulses[ir.LevelCoupling.Rydberg].detuning.append(
        ir.DetuningTerm(ir.DetuningOperatorData(pattern), ir.piecewise_linear([4.0], [-3.0, 8.0])))
    h = RydbergHamiltonianCodeGen().emit(prog)
    dim, ldim = h.space.size, 1 << n_loc
    gen = torch.Generator(device="cuda").manual_seed(5)
    psi = torch.randn(dim, 2, dtype=torch.float64, device="cuda", generator=gen)
    psi /= psi.norm()
    D = _lib.MEM_DEVICE
    for raw in (True, False):
        ref = torch.empty_like(psi)
        h.plan.apply(h.record, 2.3, psi.data_ptr(), ref.data_ptr(), raw=raw, mem=D)
        torch.cuda.synchronize()
        scale = float(ref.abs().max())
        for rank in range(world):
            plan = _lib.Plan(n_loc, 2, ldim)
            did = plan.build_rydberg_shard(pair_interactions(prog.register), rank)
            rec = pack_task_record(prog, did)
            out = torch.full((ldim, 2), float("nan"), dtype=torch.float64, device="cuda")
            partners = [psi.data_ptr() + 16 * ldim * (rank ^ (1 << b)) for b in range(g)]
            plan.apply_sharded(rec, 2.3, n_atoms, rank, partners, psi.data_ptr() + 16 * ldim * rank, out.data_ptr(), raw=raw)
            torch.cuda.synchronize()
            assert float((out - ref[rank * ldim:(rank + 1) * ldim]).abs().max()) <= 1e-13 * scale, (rank, raw)
            if raw and rank in (0, world - 1) and n_atoms <= 18:  # shard diagonal == slice of the full diagonal, bit for bit
                assert np.array_equal(plan.get_diagonal(did), h.plan.get_diagonal(h.diag_id)[rank * ldim:(rank + 1) * ldim])
            plan.close()
    h.plan.close()


def test_sharded_barrier_is_bounded_when_a_peer_never_arrives(monkeypatch):
    """rank 0 of a 2-rank sharded evolve whose peer never shows up (its buffers are plain allocations of this
    process): the device-initiated barrier gives up after BQA_BARRIER_TIMEOUT_MS, sets the abort word of every rank
    and the call returns status -6 instead of hanging; after bqa_shard_reset the sync block is clean again"""
    import time
    import torch

    monkeypatch.setenv("BQA_BARRIER_TIMEOUT_MS", "300")
    n_atoms, ldim = 15, 1 << 14
    from bloqade_analog_b200.lowering import pack_task_record, pair_interactions

    prog = programs.adiabatic_program(ir.chain_sites(n_atoms, lattice_spacing=5.5), durations=[0.01, 0.02, 0.01])
    plan = _lib.Plan(n_atoms - 1, 2, ldim)
    did = plan.build_rydberg_shard(pair_interactions(prog.register), 0)
    rec = pack_task_record(prog, did)
    bufs = [_lib.shard_alloc(16 * ldim)[0] for _ in range(2)]
    syncs = [_lib.shard_alloc(_lib.shard_sync_bytes())[0] for _ in range(2)]
    desc = _lib.ShardDesc(2, 0, n_atoms, 0)
    for r in range(2):
        desc.ys[r], desc.sync[r] = bufs[r], syncs[r]
    out = torch.empty((1, ldim), dtype=torch.complex128, device="cuda")
    opts = _lib.make_opts("dop853", 1e-7, 1e-14, engine="streamed")
    t0 = time.perf_counter()
    stats = plan.evolve_sharded(rec, opts, desc, np.array([prog.duration]), out.data_ptr(), None)
    assert time.perf_counter() - t0 < 30.0
    assert int(stats["status"][0]) == -6

    def flags(ptr):
        class H:
            __cuda_array_interface__ = {"shape": (8,), "typestr": "<u8", "data": (ptr, False), "version": 2, "strides": None}
        return torch.as_tensor(H(), device="cuda").cpu().numpy().copy()

    assert flags(syncs[0])[2] == 1 and flags(syncs[1])[2] == 1  # abort word set on both ranks
    _lib.shard_reset(syncs[0])
    assert not flags(syncs[0]).any()
    for b in bufs + syncs:
        _lib.shard_free(b)
    plan.close()


def test_streamed_engine_tiled_rhs_equals_gather_rhs(monkeypatch):
    """the streamed engine with the tiled RHS (default for two-level full spaces) and with the gather RHS
    (BQA_NO_TILED) integrate the same 2^16 state to the same answer"""
    prog = programs.adiabatic_program(ir.square_sites(4, 4, lattice_spacing=6.1), durations=[0.1, 0.3, 0.1])
    h = RydbergHamiltonianCodeGen().emit(prog)
    gate = AnalogGate(h)
    (a,) = gate.apply(h.space.zero_state(np.complex128), engine="streamed")
    monkeypatch.setenv("BQA_NO_TILED", "1")
    (b,) = gate.apply(h.space.zero_state(np.complex128), engine="streamed")
    assert infidelity(a.data, b.data) < 1e-12
    assert np.abs(a.data - b.data).max() < 1e-7


def test_streamed_engine_interaction_picture_tiled_equals_gather(monkeypatch):
    """interaction picture on the streamed engine: tiled passes without the static diagonal + the conj(u) rotation pass
    against the gather RHS (BQA_NO_TILED), 2^16 state"""
    prog = programs.adiabatic_program(ir.square_sites(4, 4, lattice_spacing=5.0), durations=[0.05, 0.3, 0.05])
    h = RydbergHamiltonianCodeGen().emit(prog)
    gate = AnalogGate(h)
    (a,) = gate.apply(h.space.zero_state(np.complex128), engine="streamed", interaction_picture=True)
    monkeypatch.setenv("BQA_NO_TILED", "1")
    (b,) = gate.apply(h.space.zero_state(np.complex128), engine="streamed", interaction_picture=True)
    assert infidelity(a.data, b.data) < 1e-12
    assert np.abs(a.data - b.data).max() < 1e-7


def test_large_streamed_evolve_norm_and_sharding_free_consistency():
    """2^20 state, short evolution on the streamed engine vs the CTA-free oracle (SciPy), fidelity < 1e-8"""
    from oracle import emulator as oe

    prog = programs.adiabatic_program(ir.square_sites(5, 4, lattice_spacing=6.0), [0.02, 0.06, 0.02])
    h = RydbergHamiltonianCodeGen().emit(prog)
    (out,) = AnalogGate(h).apply(h.space.zero_state(np.complex128), atol=1e-9, rtol=1e-9)
    ref = oe.evolve(prog, atol=1e-9, rtol=1e-9)[-1]
    assert infidelity(out.data, ref) < 1e-8
    assert abs(np.linalg.norm(out.data) - 1.0) < 1e-6


def test_large_streamed_interaction_picture_vs_oracle():
    """2^18 state, streamed engine in the INTERACTION picture (gather RHS with the exp(-i V t) rotation; at config-4-like
    parameters it needs ~16x fewer RHS evaluations than the Schroedinger picture) against the oracle's restatement of
    _ode_complex_kernel_int driven by SciPy (SV:342-359, 604-641) at the run() defaults"""
    from oracle import emulator as oe

    prog = programs.adiabatic_program(ir.square_sites(6, 3, lattice_spacing=5.0), [0.008, 0.064, 0.008])
    h = RydbergHamiltonianCodeGen().emit(prog)
    (out,) = AnalogGate(h).apply(h.space.zero_state(np.complex128), interaction_picture=True, engine="streamed")
    ref = oe.evolve(prog, interaction_picture=True)[-1]
    assert infidelity(out.data, ref) < 1e-8
    assert np.abs(out.data - ref).max() < 1e-6


def test_streamed_full_protocol_16_atoms_matches_reference():
    """The code path of BASELINE configs[3] (streamed engine + tiled H psi: both passes, ~18 000 dop853 steps at the
    stability limit) over the FULL 4 us README protocol at the size the reference finishes in minutes: Square(4, 5.0),
    2^16 amplitudes, run() defaults.  Golden = the reference's own final state (tests/golden/make_large_golden.py)."""
    import json

    from bloqade_analog_b200 import serialize

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "large_square4_a5.npz"))
    with open(os.path.join(os.path.dirname(__file__), "golden", "large_square4_a5.json")) as f:
        prog = serialize.program_from_dict(json.load(f))
    h = RydbergHamiltonianCodeGen().emit(prog)
    gate = AnalogGate(h)
    (out,) = gate.apply(h.space.zero_state(np.complex128), engine="streamed")
    ref = g["final_c64"].astype(np.complex128)
    assert infidelity(out.data, ref) < 1e-8
    assert abs(out.norm() - float(g["norm"])) < 1e-6
    p = np.abs(out.data) ** 2
    dens = np.array([p[(h.space.configurations >> a) & 1 == 1].sum() for a in range(16)])
    np.testing.assert_allclose(dens, g["density"], rtol=0, atol=1e-6)
    avg, var = h.average_and_variance(out)
    ravg, rvar = g["avg_var"]
    assert abs(avg - ravg) < 1e-5 * max(1.0, abs(ravg)) and abs(var - rvar) < 1e-5 * max(1.0, abs(rvar))
    # seeded shots of the normalised state: the reference's own draws (np.random.choice consumes random_sample)
    out.normalize()
    shots = h.plan.sample(out.data, g["sample_uniforms"])
    assert (shots.reshape(g["sample_shots"].shape) != g["sample_shots"]).any(axis=1).sum() <= 1
    # the same evolution on the CTA-free gather RHS and on the tiled one is covered by
    # test_streamed_engine_tiled_rhs_equals_gather_rhs


def test_launch_counter_and_no_cpu_fallback():
    before = _lib.launch_count()
    h = ham("c1_square3_a5")
    h._apply(golden_case("c1_square3_a5").data["rhs_psi"], 1.0)
    assert _lib.launch_count() > before


def test_sharded_evolve_two_gpus_matches_unsharded():
    """the real multi-GPU path (CUDA IPC peer mappings, device-initiated barriers, cross-rank error norms) under
    torchrun on 2 GPUs of this box; skipped on a single-GPU box (profiles/r01_shard_check_*.json hold the runs
    made with gpurun --gpus 2/4/8)"""
    import json
    import subprocess
    import sys
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(root, "profiles", "shard_check.py"), "16", "0.2"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert lines, out.stderr[-2000:]
    res = {}
    for l in lines:
        res.update(json.loads(l))
    assert res["stats"]["status"] == 1
    assert res["apply_max_rel_err"] < 1e-13
    assert res["infidelity"] < 1e-8
    assert res["shots_rows_differing"] <= 1
    assert res["checkpoint_resume_bit_identical"] and res["segments_resumed"] == 2
