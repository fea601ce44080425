"""Drop-in boundary: `program.bloqade.python().run(...)` routed to the B200 backend.
Needs the reference tree (build container only); on a box without a GPU the run must fail loudly."""
import os
import sys

import numpy as np
import pytest
from conftest import GOLDEN, cuda_available

pytestmark = pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref():
    sys.path.insert(0, GOLDEN)
    import refshim

    refshim.install()
    from bloqade.analog.atom_arrangement import Square

    return Square


def program(Square):
    return (Square(3, lattice_spacing="a").rydberg.rabi.amplitude.uniform.piecewise_linear([0.1, 0.3, 0.1], [0, 15.8, 15.8, 0])
            .detuning.uniform.piecewise_linear([0.1, 0.3, 0.1], [-16.33, -16.33, 16.33, 16.33])
            .batch_assign(a=[5.0, 6.0]))


def test_install_routes_python_to_b200_and_compiles(ref):
    from bloqade_analog_b200 import adapter, lowering, routine

    routine.install()
    try:
        r = program(ref).bloqade.python()
        assert type(r).__name__ == "BloqadeB200Routine"
        batch = r._compile(shots=10)
        assert len(batch.tasks) == 2
        recs = [lowering.pack_task_record(adapter.convert_program(t.emulator_ir), 0) for t in batch.tasks.values()]
        assert recs[0].size == recs[1].size and recs[0][0] == 0.5
        ems = r.hamiltonian()
        assert len(ems) == 2
        with pytest.raises(ValueError):
            r.run(10, multiprocessing=True, cache_matrices=True)
        if not cuda_available():
            from bloqade_analog_b200._lib import BackendError

            with pytest.raises(BackendError):  # no CPU fallback
                r.run(10)
    finally:
        routine.uninstall()
    assert type(program(ref).bloqade.python()).__name__ == "BloqadePythonRoutine"


def test_result_packaging_matches_reference_convention(ref):
    from bloqade_analog_b200 import routine
    from bloqade_analog_b200.batch import TaskResult

    r = routine.b200_routine(program(ref))
    batch = r._compile(shots=3)
    task = list(batch.tasks.values())[0]
    shots = np.array([[1, 0, 0, 0, 0, 0, 0, 0, 1]] * 3, dtype=np.uint8)
    routine._package(task, TaskResult(shots, 1, 0, 0, 0, 0))
    out = task.result().shot_outputs
    assert len(out) == 3 and out[0].pre_sequence == [1] * 9
    assert out[0].post_sequence == [0, 1, 1, 1, 1, 1, 1, 1, 0]
    rep = batch.tasks  # same LocalBatch/BloqadeTask types as the reference
    assert type(task).__name__ == "BloqadeTask"
