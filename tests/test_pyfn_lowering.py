"""PythonFn / Smooth waveforms (ir/control/waveform.py:405-468,571-642) cannot run on the device: the backend lowers
them at plan time to a cubic-Hermite table (2048 uniform knots, ir._hermite_pieces) that the device evaluates like
any other polynomial piece.  Here the table - evaluated by the DEVICE code compiled for the host (tests/hostsim) - is
compared with the reference's own interpreter on the programs of the reference's tests/test_python_codegen.py:270-286
and tests/test_python_emulator.py:65,85 (`.fn(...)`) and on a Gaussian-smoothed waveform.

Stated bounds (relative to max |f|):
  PythonFn, smooth f:  1e-9   (interpolation error h^4/384 max|f''''| with h = T/2048 is below 1e-14 for these;
                               what is left is the finite-difference derivative at the knots)
  Smooth:              3e-4   (the reference evaluates it by scipy.integrate.quad with epsabs = epsrel = 1e-4,
                               waveform.py:632-638: measured here, its value jitters by up to 2.2e-3 absolute = 1.8e-4
                               of the 12 rad/us plateau around the true convolution from one t to the next; the table
                               goes through the reference's values at its knots and cannot do better in between)
The GPU end of it (final states of run_callback against the reference, infidelity < 1e-8) is
tests/test_dropin_gpu.py::test_python_fn_and_smooth_waveforms_match_reference."""
import numpy as np
import pytest
from conftest import load_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="reference package not present")


@pytest.fixture(scope="module")
def ba():
    return load_reference()


def programs(ba):
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
    return {"codegen_fn_phase": (p1, 1e-9), "emulator_fn_amplitude": (p2, 1e-9), "gaussian_smooth": (p3, 3e-4)}


@pytest.mark.parametrize("name", ["codegen_fn_phase", "emulator_fn_amplitude", "gaussian_smooth"])
def test_lowered_table_against_the_reference_interpreter(ba, name, hostsim):
    from bloqade_analog_b200 import adapter, lowering

    prog, bound = programs(ba)[name]
    (em,) = prog.bloqade.python().hamiltonian()
    ir_ref = em.hamiltonian.emulator_ir
    mine = adapter.convert_program(ir_ref)
    rec = lowering.pack_task_record(mine, 0)
    T = float(ir_ref.duration)
    rng = np.random.default_rng(7)
    ts = np.unique(np.concatenate([np.linspace(0.0, T, 97), rng.uniform(0.0, T, 400)]))
    ref_waves, my_waves = [], []
    for fields_ref, fields_mine in zip(ir_ref.pulses.values(), mine.pulses.values()):
        for a, b in zip(fields_ref.detuning, fields_mine.detuning):
            ref_waves.append(a.amplitude)
            my_waves.append(b.amplitude)
        for a, b in zip(fields_ref.rabi, fields_mine.rabi):
            ref_waves.append(a.amplitude)
            my_waves.append(b.amplitude)
            if a.phase is not None:
                ref_waves.append(a.phase)
                my_waves.append(b.phase)
    table = lowering._WaveformTable()  # ids in the record: re-pack one at a time in the same order
    ids = [table.add(w) for w in my_waves]
    worst = 0.0
    for wid, wf in zip(ids, ref_waves):
        f = wf.emit()
        want = np.array([float(f(float(t))) for t in ts])
        got = np.array([hostsim.wf_eval(rec, wid, float(t)) for t in ts])
        scale = max(1.0, np.abs(want).max())
        worst = max(worst, float(np.abs(got - want).max() / scale))
    assert len(ids) >= 2
    print(f"{name}: worst relative deviation {worst:.3e} (bound {bound:g})")
    assert worst <= bound, worst
