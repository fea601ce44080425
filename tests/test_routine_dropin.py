"""Drop-in boundary on the CPU: `program.bloqade.python()` routed to the B200 routine, the
result types, the array-built packaging / report against the reference's per-shot loops, and the loud failure
without a GPU.  Needs the reference package (/root/reference here, baseline/_ref on the GPU box)."""
import inspect
import pickle
from collections import OrderedDict

import numpy as np
import pytest
from conftest import cuda_available, load_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="reference package not present")


@pytest.fixture(scope="module")
def ref():
    load_reference()
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
        with pytest.raises(ValueError, match="num_workers"):
            r.run(10, num_workers=2)
        with pytest.raises(ValueError, match="not supported"):
            r.run(10, solver_name="rk4")
        if not cuda_available():
            from bloqade_analog_b200._lib import BackendError

            with pytest.raises(BackendError):  # no CPU fallback
                r.run(10)
            with pytest.raises(BackendError):
                r.run_callback(lambda *a: None)
    finally:
        routine.uninstall()
    assert type(program(ref).bloqade.python()).__name__ == "BloqadePythonRoutine"


def test_signatures_match_the_reference_routine(ref):
    from bloqade.analog.ir.routine.bloqade import BloqadePythonRoutine

    from bloqade_analog_b200 import routine

    cls = routine.routine_class()
    assert issubclass(cls, BloqadePythonRoutine)
    for name in ("run", "__call__", "run_callback", "hamiltonian", "_compile"):
        want = list(inspect.signature(getattr(BloqadePythonRoutine, name)).parameters.values())
        got = list(inspect.signature(getattr(cls, name)).parameters.values())
        assert [(p.name, p.kind, p.default) for p in got[: len(want)]] == [(p.name, p.kind, p.default) for p in want], name
        # anything new is keyword-only with a default
        assert all(p.kind is p.KEYWORD_ONLY and p.default is not p.empty for p in got[len(want):]), name


def _fake_run(routine, r, shots, rng):
    """lower the sweep on the CPU and package made-up measurement digits (no GPU needed)"""
    from bloqade_analog_b200 import _lib

    low = routine._lower(r, (), 0.0, "interpret", False)
    digits = (rng.random((low.n_tasks, shots, low.n_atoms)) < 0.3).astype(np.uint8)
    stats = np.zeros(low.n_tasks, dtype=_lib.STATS_DTYPE)
    return low, digits, r._package(low, shots, digits, stats, None)


def test_packaging_and_report_match_the_reference_loops(ref):
    """task/bloqade.py:66-86 and task/batch.py:121-198 done with arrays == the reference's per-shot loops"""
    import pandas as pd
    from bloqade.analog.submission.ir.task_results import (
        QuEraShotResult,
        QuEraShotStatusCode,
        QuEraTaskResults,
        QuEraTaskStatusCode,
    )
    from bloqade.analog.task.batch import LocalBatch
    from bloqade.analog.task.bloqade import BloqadeTask

    from bloqade_analog_b200 import routine

    r = routine.b200_routine(program(ref))
    low, digits, batch = _fake_run(routine, r, 7, np.random.default_rng(5))
    assert isinstance(batch, LocalBatch) and all(isinstance(t, BloqadeTask) for t in batch.tasks.values())
    # the reference's own construction from the same digits
    ref_tasks = OrderedDict()
    for k, td in enumerate(r._generate_ir((), 0.0, "interpret", False)):
        filling = np.asarray(td.emulator_ir.register.geometry.filling, dtype=int)
        full = np.zeros_like(filling, dtype=int)
        outs = []
        for shot in digits[k]:
            full[filling == 1] = 1 - shot
            outs.append(QuEraShotResult(shot_status=QuEraShotStatusCode.Completed, pre_sequence=filling.tolist(),
                                        post_sequence=full.tolist()))
        res = QuEraTaskResults(task_status=QuEraTaskStatusCode.Completed, shot_outputs=outs)
        ref_tasks[k] = BloqadeTask(7, td.emulator_ir, td.metadata_dict, None, res)
    ref_batch = LocalBatch(None, ref_tasks, None)
    for k in ref_tasks:
        assert batch.tasks[k].result() == ref_tasks[k].result()
        assert batch.tasks[k].metadata == ref_tasks[k].metadata
        assert batch.tasks[k].geometry == ref_tasks[k].geometry
        assert batch.tasks[k].nshots == 7
    rep_new, rep_ref = batch.report(), ref_batch.report()
    pd.testing.assert_frame_equal(rep_new.dataframe, rep_ref.dataframe)
    assert rep_new.dataframe.index.equals(rep_ref.dataframe.index)
    assert rep_new.metas == rep_ref.metas and rep_new.geos == rep_ref.geos
    from bloqade.analog.task.base import Report

    assert isinstance(rep_new, Report)
    for got in (rep_new.bitstrings(), rep_new.bitstrings(clusters=(0, 0)), Report.bitstrings(rep_new)):
        assert len(got) == len(rep_ref.bitstrings())
        for a, b in zip(got, rep_ref.bitstrings()):
            assert np.array_equal(a, b) and a.dtype == b.dtype
    assert all(x.shape[0] == 0 for x in rep_new.bitstrings(clusters=[(1, 0)]))
    assert rep_new.counts() == rep_ref.counts()
    assert rep_new.list_param("a") == [5.0, 6.0]
    pd.testing.assert_frame_equal(rep_new.rydberg_densities(), rep_ref.rydberg_densities())
    # the reference's own (slow) report on the B200 batch gives the same frame as well
    pd.testing.assert_frame_equal(LocalBatch.report(batch).dataframe, rep_ref.dataframe)
    # filter_metadata keeps the batch type and the fast report
    sub = batch.filter_metadata(a=[6.0])
    assert type(sub) is type(batch) and list(sub.tasks) == [1]
    pd.testing.assert_frame_equal(sub.report().dataframe, ref_batch.filter_metadata(a=[6.0]).report().dataframe)


def test_tasks_are_lazy_and_pickle_as_reference_tasks(ref):
    from bloqade.analog.emulate.ir.emulator import EmulatorProgram
    from bloqade.analog.task.bloqade import BloqadeTask

    from bloqade_analog_b200 import routine

    r = routine.b200_routine(program(ref))
    _, _, batch = _fake_run(routine, r, 3, np.random.default_rng(1))
    task = batch.tasks[1]
    assert task._emulator_ir is None and task._result is None  # nothing built yet
    ir = task.emulator_ir  # compiled by the reference pipeline on demand
    assert isinstance(ir, EmulatorProgram) and ir.duration == 0.5
    assert [float(x) for x in ir.register.sites[1]] == [0.0, 6.0]
    clone = pickle.loads(pickle.dumps(task))
    assert type(clone) is BloqadeTask
    assert clone.result() == task.result() and clone.metadata == task.metadata and clone.shots == 3


def test_resolve_devices_maps_the_reference_options(ref, monkeypatch):
    from bloqade_analog_b200 import _lib, routine

    monkeypatch.setattr(_lib, "device_count", lambda: 8)
    monkeypatch.setattr(_lib, "current_device", lambda: 3)
    assert routine._resolve_devices(None, False, None) == [3]
    assert routine._resolve_devices(None, True, None) == list(range(8))
    assert routine._resolve_devices(None, True, 4) == [0, 1, 2, 3]
    assert routine._resolve_devices(None, True, 64) == list(range(8))
    assert routine._resolve_devices(2, False, None) == [0, 1]
    assert routine._resolve_devices([5, 1], False, None) == [5, 1]
    with pytest.raises(ValueError):
        routine._resolve_devices([9], False, None)
    parts = routine._split(list(range(10)), np.arange(10, dtype=float), 3)
    assert sorted(sum(parts, [])) == list(range(10)) and max(map(len, parts)) - min(map(len, parts)) <= 1
