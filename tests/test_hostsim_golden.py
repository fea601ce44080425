"""CPU checks of the DEVICE ALGORITHM (the same .cuh sources the kernels compile, run serially by
tests/hostsim) against golden vectors produced by the reference itself.  These pin the host logic
(space, lowering, task record) and the integrator / RHS / waveform arithmetic without a GPU; the
`-m gpu` tests repeat them through the CUDA library."""
import numpy as np
import pytest
from conftest import GOLDEN_CASES, SMALL_CASES, golden_case, infidelity

from bloqade_analog_b200 import lowering, programs
from bloqade_analog_b200.space import SpaceType


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_space_and_diagonal_bit_exact(name):
    c = golden_case(name)
    assert np.array_equal(c.space.configurations, c.data["configurations"])
    assert c.space.configurations.dtype == c.data["configurations"].dtype
    assert (c.space.space_type is SpaceType.FullSpace) == bool(c.data["is_full"])
    assert np.array_equal(lowering.rydberg_diagonal(c.space), c.data["rydberg"])


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_record_matches_reference_derived(name):
    c = golden_case(name)
    assert np.array_equal(lowering.pack_task_record(c.program, 0), c.data["record"])


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_device_waveform_eval(name, hostsim):
    """on-device piece-table evaluation == the reference's Decimal interpreter (WF:84-85)"""
    c = golden_case(name)
    rec = c.data["record"]
    ts, vals = c.data["wf_times"], c.data["wf_values"]
    k = 0
    rows = []
    for fields in c.program.pulses.values():
        for term in fields.detuning:
            rows.append(term.amplitude)
        for term in fields.rabi:
            rows.append(term.amplitude)
            if term.phase is not None:
                rows.append(term.phase)
    # map each waveform to its id in the record: re-pack one at a time in the same order
    table = lowering._WaveformTable()
    ids = [table.add(w) for w in rows]
    for wid, ref in zip(ids, vals):
        got = np.array([hostsim.wf_eval(rec, wid, float(t)) for t in ts])
        np.testing.assert_allclose(got, ref, rtol=1e-13, atol=1e-13)
        k += 1
    assert k == len(vals)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_rhs_parity(name, hostsim):
    c = golden_case(name)
    V, rec, psi = c.data["rydberg"], c.data["record"], c.data["rhs_psi"]
    for t, ref, ref_int in zip(c.data["rhs_times"], c.data["rhs_out"], c.data["rhs_out_int"]):
        scale = np.abs(ref).max()
        got = hostsim.apply(c.space, V, rec, float(t), psi)
        assert np.abs(got - ref).max() <= 1e-13 * scale
        got = hostsim.apply(c.space, V, rec, float(t), psi, interaction=True)
        assert np.abs(got - ref_int).max() <= 1e-12 * scale
    got = hostsim.apply(c.space, V, rec, float(c.data["rhs_times"][-1]), psi, raw=True)
    assert np.abs(got - c.data["happly_out"]).max() <= 1e-13 * np.abs(c.data["happly_out"]).max()


def test_c2b_builder_rhs(hostsim):
    """the standalone C2b builder orders its sixteen terms differently from the reference: same physics"""
    c = golden_case("c2b_chain16_blockade_local")
    rec = lowering.pack_task_record(programs.config2("b"), 0)
    for t, ref in zip(c.data["rhs_times"], c.data["rhs_out"]):
        got = hostsim.apply(c.space, c.data["rydberg"], rec, float(t), c.data["rhs_psi"])
        assert np.abs(got - ref).max() <= 1e-13 * np.abs(ref).max()


@pytest.mark.parametrize("name", SMALL_CASES)
@pytest.mark.parametrize("solver", ["dop853", "dopri5"])
def test_evolve_matches_reference_scipy(name, solver, hostsim):
    """same controller as this image's SciPy (dop853: reject_mode 1): identical (x, h) step sequence
    (SciPy 1.18.1 only adds redundant RHS re-evaluations: one more at x+h after an accepted step, one at
    x after a rejected one), so the states agree up to amplified rounding - far below the solver's own
    global error (~1e-6 at atol=1e-7)"""
    c = golden_case(name)
    out, stats, rc = hostsim.evolve(c.space, c.data["rydberg"], c.data["record"], c.data["times"], solver, reject_mode=1)
    assert rc == 1
    ref = c.data[f"final_{solver}"]
    for a, b in zip(out, ref):
        assert np.abs(a - b).max() < 2e-8
        assert infidelity(a, b) < 1e-13


@pytest.mark.parametrize("name", SMALL_CASES)
def test_evolve_default_controller_fidelity(name, hostsim):
    """default (Hairer) rejection rule: fewer steps, fidelity deviation < 1e-8 (north_star bar)"""
    c = golden_case(name)
    out, stats, rc = hostsim.evolve(c.space, c.data["rydberg"], c.data["record"], c.data["times"], "dop853")
    assert rc == 1
    for a, b in zip(out, c.data["final_dop853"]):
        assert infidelity(a, b) < 1e-8


@pytest.mark.parametrize("name", SMALL_CASES)
def test_evolve_interaction_picture(name, hostsim):
    c = golden_case(name)
    out, stats, rc = hostsim.evolve(c.space, c.data["rydberg"], c.data["record"], c.data["times"], "dop853",
                                    interaction=True, reject_mode=1)
    assert rc == 1
    for a, b in zip(out, c.data["final_dop853_int"]):
        assert infidelity(a, b) < 1e-10


def test_nsteps_exceeded_status(hostsim):
    c = golden_case("c1_square3_a6p5_short")
    out, stats, rc = hostsim.evolve(c.space, c.data["rydberg"], c.data["record"], c.data["times"][-1:], nsteps=5)
    assert rc == -2
