"""One-pass lowering of a batch_assign sweep (bloqade_analog_b200/sweep.py, routine._lower) against the
reference's own compile pipeline run task by task (ir/routine/bloqade.py:184-202 + adapter.convert_program):
every task record must be BIT-identical, the metadata equal, and programs the vector pass cannot handle must
fall back to the task-by-task path.  CPU only; needs the reference package (/root/reference or baseline/_ref)."""
import numpy as np
import pytest
from conftest import load_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="reference package not present")

OMEGA = [0, 15.8, 15.8, 0]
DELTA = [-16.33, -16.33, 16.33, 16.33]


@pytest.fixture(scope="module")
def ba():
    return load_reference()


def both(program, args=(), blockade_radius=0.0, use_hyperfine=False):
    from bloqade_analog_b200 import routine

    r = routine.b200_routine(program)
    fast = routine._lower(r, args, blockade_radius, "interpret", use_hyperfine)
    routine.FAST_LOWERING = False
    try:
        slow = routine._lower(r, args, blockade_radius, "interpret", use_hyperfine)
    finally:
        routine.FAST_LOWERING = True
    return fast, slow


def assert_identical(fast, slow, expect_fast=True):
    assert slow.how.startswith("task-by-task")
    if expect_fast:
        assert fast.how == "vectorised", fast.how
    assert fast.n_tasks == slow.n_tasks
    for k, (a, b) in enumerate(zip(fast.records, slow.records)):
        assert a.shape == b.shape, k
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), k
    assert np.array_equal(fast.durations, slow.durations)
    assert fast.metadata == slow.metadata
    assert fast.filling == slow.filling and fast.n_atoms == slow.n_atoms
    assert [g.task_idx for g in fast.groups] == [g.task_idx for g in slow.groups]
    assert [len(g.site_registers) for g in fast.groups] == [len(g.site_registers) for g in slow.groups]
    for k in range(fast.n_tasks):
        assert [tuple(map(float, s)) for s in fast.sites[k]] == [tuple(map(float, s)) for s in slow.sites[k]]


def test_config1_lattice_spacing_sweep(ba):
    from bloqade.analog.atom_arrangement import Square

    prog = (Square(3, lattice_spacing="a").rydberg.rabi.amplitude.uniform.piecewise_linear([0.4, 3.2, 0.4], OMEGA)
            .detuning.uniform.piecewise_linear([0.4, 3.2, 0.4], DELTA).batch_assign(a=np.arange(4.0, 7.0, 0.5)))
    fast, slow = both(prog)
    assert_identical(fast, slow)
    assert len(fast.groups) == 1 and len(fast.groups[0].site_registers) == 6  # one diagonal per spacing


def test_config3_detuning_duration_sweep(ba):
    from bloqade.analog import cast
    from bloqade.analog.atom_arrangement import Chain

    total = cast("total")
    durations = [0.1 * total, 0.8 * total, 0.1 * total]
    prog = (Chain(12, lattice_spacing=5.72).rydberg.rabi.amplitude.uniform.piecewise_linear(durations, OMEGA)
            .detuning.uniform.piecewise_linear(durations, [-16.33, -16.33, "final", "final"]))
    dets = np.linspace(4.0, 32.0, 12)
    durs = np.linspace(0.5, 4.0, 12)
    grid = [(float(round(d, 6)), float(round(t, 6))) for d in dets for t in durs]
    fast, slow = both(prog.batch_assign(final=[g[0] for g in grid], total=[g[1] for g in grid]))
    assert_identical(fast, slow)
    # and the records equal the ones of the stand-alone program builder that bench.py and the GPU tests use
    from bloqade_analog_b200 import lowering, programs

    for k in (0, 77, 143):
        want = lowering.pack_task_record(programs.config3_task(*grid[k]), 0)
        assert np.array_equal(want.view(np.uint64), fast.records[k].view(np.uint64))


def test_blockade_subspace_varying_positions_split_into_spaces(ba):
    from bloqade.analog.atom_arrangement import Chain

    prog = (Chain(6, lattice_spacing="a").rydberg.rabi.amplitude.uniform.constant(5.0, "t")
            .detuning.uniform.constant(2.0, "t").batch_assign(a=[4.0, 7.0, 4.0], t=[0.3, 0.4, 0.5]))
    fast, slow = both(prog, blockade_radius=5.0)
    assert_identical(fast, slow)
    assert [g.task_idx for g in fast.groups] == [[0, 2], [1]]


def test_local_drives_scaled_locations_and_masks(ba):
    from bloqade.analog import start

    geo = start.add_position([(0, 0), (0, 6.0), (5.5, 0), (5.5, 6.0)])
    prog = (geo.rydberg.detuning.uniform.piecewise_linear([0.2, "t"], [-10, -10, "d"])
            .scale([0.5, 0.0, 1.0, 0.25]).constant("loc_det", 0.3)
            .rydberg.rabi.amplitude.location([0, 2], ["s0", 0.5]).piecewise_linear([0.1, "t", 0.1], [0, "om", "om", 0])
            .location(1).constant(3.0, 0.2)
            .batch_assign(t=[0.3, 0.5, 0.9], d=[5.0, 8.0, 12.0], om=[4.0, 6.0, 9.0], s0=[1.0, 0.7, 0.3],
                          loc_det=[1.0, 2.0, 3.0]))
    fast, slow = both(prog)
    assert_identical(fast, slow)


def test_uniform_phase_and_per_atom_phase(ba):
    from bloqade.analog.atom_arrangement import Chain

    base = Chain(3, lattice_spacing=6.1).rydberg.rabi.amplitude.uniform.constant("om", 0.5)
    uniform_phase = base.phase.uniform.piecewise_constant([0.2, 0.3], [0.0, "phi"]).batch_assign(
        om=[3.0, 5.0], phi=[0.4, 1.2])
    assert_identical(*both(uniform_phase))
    local_phase = (base.phase.location(0).constant("phi", 0.5).location(2).constant(0.3, 0.5)
                   .batch_assign(om=[3.0, 5.0], phi=[0.4, 1.2]))
    assert_identical(*both(local_phase))


def test_more_drives_than_atoms_take_the_per_atom_sum_path(ba):
    from bloqade.analog.atom_arrangement import Chain

    prog = (Chain(2, lattice_spacing=7.0).rydberg.detuning.uniform.constant("a", 0.4)
            .location(0).constant("b", 0.4).location(1).linear(0.0, "c", 0.4)
            .rydberg.rabi.amplitude.uniform.constant(2.0, 0.4).location(0).constant("b", 0.4).location(1).constant("c", 0.4)
            .batch_assign(a=[1.0, 1.5], b=[3.0, 4.0], c=[5.0, 7.0]))
    assert_identical(*both(prog))


def test_hyperfine_and_rydberg_couplings(ba):
    from bloqade.analog.atom_arrangement import Chain

    prog = (Chain(3, lattice_spacing=6.5).hyperfine.rabi.amplitude.uniform.constant("oh", 0.3)
            .hyperfine.detuning.uniform.constant(1.5, 0.3)
            .rydberg.rabi.amplitude.uniform.piecewise_linear([0.1, 0.1, 0.1], [0, "om", "om", 0])
            .batch_assign(oh=[2.0, 3.0], om=[5.0, 6.0]))
    fast, slow = both(prog)
    assert_identical(fast, slow)
    assert fast.groups[0].register.atom_type.n_level == 3


def test_slices_poly_negative_append_and_args(ba):
    from bloqade.analog import piecewise_linear, var
    from bloqade.analog.atom_arrangement import Chain

    t = var("t")
    wf = piecewise_linear([0.2, 0.6, 0.2], [0, 10, 10, 0])
    prog = (Chain(2, lattice_spacing=6.0).rydberg.rabi.amplitude.uniform.apply(wf[0.1:t])
            .detuning.uniform.poly([1.0, "c1", -2.0], t).apply(-wf[:0.4])
            .batch_assign(t=[0.5, 0.6, 0.7]).args(["c1"]))
    assert_identical(*both(prog, args=(0.75,)))


def test_vacancies_keep_the_filling(ba):
    from bloqade.analog import start

    prog = (start.add_position((0, 0)).add_position((0, 5.0), filling=False).add_position((0, 10.0))
            .rydberg.rabi.amplitude.uniform.constant("om", 0.4).detuning.scale([1.0, 5.0, 0.5]).constant(2.5, 0.4)
            .batch_assign(om=[1.0, 2.0, 3.0]))
    fast, slow = both(prog)
    assert_identical(fast, slow)
    assert fast.filling == [1, 0, 1] and fast.n_atoms == 2


def test_waveforms_that_coincide_for_one_task_only_fall_back(ba):
    """task 1 has om == the detuning constant: the reference de-duplicates the two waveform tables there"""
    from bloqade.analog.atom_arrangement import Chain

    prog = (Chain(2, lattice_spacing=6.0).rydberg.rabi.amplitude.uniform.constant("om", 0.4)
            .detuning.uniform.constant(2.0, 0.4).batch_assign(om=[1.0, 2.0, 3.0]))
    fast, slow = both(prog)
    assert fast.how.startswith("task-by-task")
    assert_identical(fast, slow, expect_fast=False)


def test_python_fn_waveforms_fall_back(ba):
    from bloqade.analog.atom_arrangement import Chain

    def ramp(time, slope):
        return slope * time

    prog = (Chain(2, lattice_spacing=6.0).rydberg.rabi.amplitude.uniform.fn(ramp, 0.5)
            .batch_assign(slope=[1.0, 2.0]))
    fast, slow = both(prog)
    assert fast.how.startswith("task-by-task")
    assert_identical(fast, slow, expect_fast=False)


def test_branches_that_differ_between_tasks_fall_back(ba):
    """a segment of zero duration is dropped by the canonicaliser (canonicalize.py:210-212) for one task only"""
    from bloqade.analog.atom_arrangement import Chain

    prog = (Chain(2, lattice_spacing=6.0).rydberg.rabi.amplitude.uniform.piecewise_linear(["t0", 0.3], [1.0, 2.0, 0.0])
            .batch_assign(t0=[0.0, 0.2]))
    fast, slow = both(prog)
    assert fast.how.startswith("task-by-task")
    assert_identical(fast, slow, expect_fast=False)


def test_single_task_programs_and_static_assign(ba):
    from bloqade.analog.atom_arrangement import Square

    prog = (Square(2, lattice_spacing="a").rydberg.rabi.amplitude.uniform.piecewise_linear([0.1, "t", 0.1], OMEGA)
            .assign(a=5.5, t=0.8))
    fast, slow = both(prog)
    assert_identical(fast, slow)
    assert fast.n_tasks == 1 and fast.metadata[0].keys() == {"a", "t"}
