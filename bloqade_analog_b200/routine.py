"""Drop-in routine: `program.bloqade.python().run(...)` on the B200 backend.

Needs a real `bloqade.analog` installation (it subclasses the reference's `BloqadePythonRoutine`,
`src/bloqade/analog/ir/routine/bloqade.py:148-550`, and returns the reference's own `LocalBatch` /
`BloqadeTask` / `QuEraTaskResults` / `Report` types).  What is replaced:

* the per-task compile loop (`_generate_ir`, RB:184-202) by ONE vectorised lowering of the whole sweep
  (`sweep.py`) whenever the tasks differ only in scalar values - cross-checked bit for bit against the reference
  pipeline on sampled tasks, with the task-by-task reference pipeline as the fallback;
* the per-task numerical work (`BloqadeTask.run`, task/bloqade.py:43-88: operator generation + `AnalogGate.run`)
  by one batched GPU call per Hilbert space and device (`bqa_run_batch`), sharded over GPUs with `devices=` /
  `multiprocessing=True, num_workers=N` (independent tasks, no communication);
* the per-shot packaging loop (task/bloqade.py:66-86) and the per-shot loop of `LocalBatch.report()`
  (task/batch.py:121-198) by array operations; `QuEraTaskResults` objects are built lazily, per task, on
  first access.
"""
from __future__ import annotations

import threading
from collections import OrderedDict, namedtuple
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

from . import _lib, adapter, ir
from .hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from .lowering import pack_task_record
from .state_vector import AnalogGate, StateVector
from .vecnum import NonUniformSweep

_CLASSES = None
_ORIGINAL_PYTHON = None
#: set to False to force the task-by-task reference pipeline (tests compare the two paths)
FAST_LOWERING = True
#: how the last sweep was lowered: "vectorised" or "task-by-task: <why>"
LAST_LOWERING = ""


# -----------------------------------------------------------------------------------------------------------
# lowering of a sweep: device-independent task records grouped by Hilbert space
# -----------------------------------------------------------------------------------------------------------
@dataclass
class SpaceGroup:
    """tasks that share a Hilbert space (same `Register` in the sense of emulator.py:197-226)"""

    register: ir.Register
    task_idx: List[int] = field(default_factory=list)
    site_registers: List[ir.Register] = field(default_factory=list)  # distinct geometries -> diag ids 0, 1, ...
    site_keys: Dict[Tuple, int] = field(default_factory=dict)

    def diag_index(self, register: ir.Register) -> int:
        key = tuple(register.sites)
        if key not in self.site_keys:
            self.site_keys[key] = len(self.site_registers)
            self.site_registers.append(register)
        return self.site_keys[key]


@dataclass
class LoweredSweep:
    n_tasks: int
    metadata: List[Dict[str, Any]]
    records: List[np.ndarray]          # one float64 record per task (rows of one 2-D array on the fast path)
    durations: np.ndarray
    groups: List[SpaceGroup]
    filling: List[int]
    sites: List[Any]                   # per task: list of (x, y) Decimals of ALL sites (shared when equal)
    n_atoms: int
    how: str
    make_ir: Callable[[int], Any]      # task number -> the reference's EmulatorProgram (compiled on demand)


def _group_for(groups: List[SpaceGroup], register: ir.Register) -> SpaceGroup:
    for g in groups:
        if g.register == register:
            return g
    groups.append(SpaceGroup(register))
    return groups[-1]


def _lower_task_by_task(task_data_list, why: str, make_ir) -> LoweredSweep:
    """the reference's own compile pipeline per task (RB:184-202) + adapter.convert_program"""
    groups: List[SpaceGroup] = []
    records, durations, metadata, sites = [], [], [], []
    filling, n_atoms = [], 0
    for k, td in enumerate(task_data_list):
        program = adapter.convert_program(td.emulator_ir)
        g = _group_for(groups, program.register)
        g.task_idx.append(k)
        records.append(pack_task_record(program, g.diag_index(program.register)))
        durations.append(program.duration)
        metadata.append(td.metadata_dict)
        geometry = td.emulator_ir.register.geometry
        sites.append(geometry.sites)
        filling = [int(f) for f in geometry.filling]
        n_atoms = len(program.register)
    return LoweredSweep(len(records), metadata, records, np.asarray(durations, dtype=np.float64), groups, filling, sites,
                        n_atoms, "task-by-task: " + why, make_ir)


def _lower(routine, args, blockade_radius, waveform_runtime, use_hyperfine, verify: bool = True) -> LoweredSweep:
    """circuit + parameters -> task records.  Fast path: one vectorised pass (sweep.py), verified on the first,
    middle and last task against the reference pipeline; otherwise the reference pipeline per task."""
    global LAST_LOWERING
    from bloqade.analog.compiler.passes.emulator import assign, flatten, generate_emulator_ir

    from .sweep import SweepProgram

    params = routine.params
    circuit = flatten(routine.circuit)
    assignments = [{**params.static_params, **bp} for bp in params.batch_assignments(*args)]

    def make_ir(k: int):
        _, final_circuit = assign(assignments[k], circuit)
        return generate_emulator_ir(final_circuit, blockade_radius, waveform_runtime, use_hyperfine)

    def slow(why: str) -> LoweredSweep:
        global LAST_LOWERING
        low = _lower_task_by_task(list(routine._generate_ir(args, blockade_radius, waveform_runtime, use_hyperfine)),
                                  why, make_ir)
        LAST_LOWERING = low.how
        return low

    if not FAST_LOWERING:
        return slow("fast lowering disabled")
    try:
        sp = SweepProgram(circuit, assignments, blockade_radius, use_hyperfine)
        n = sp.n_tasks
        groups: List[SpaceGroup] = []
        diag_ids = np.zeros(n, dtype=np.float64)
        if sp.positions_vary:
            sites = []
            for k in range(n):
                reg = sp.register(k)
                g = _group_for(groups, reg)
                g.task_idx.append(k)
                diag_ids[k] = g.diag_index(reg)
                sites.append(sp.sites_decimal(k))
        else:
            reg = sp.register(0)
            g = _group_for(groups, reg)
            g.task_idx = list(range(n))
            g.diag_index(reg)
            sites = [sp.sites_decimal(0)] * n
        recs = pack_task_record(sp.program(groups[0].register), diag_ids, n_tasks=n)
        durations = recs[:, 0].copy()
    except NonUniformSweep as e:
        return slow(str(e))
    if verify:
        # the reference pipeline on sampled tasks must give the same record, bit for bit
        for k in sorted({0, n // 2, n - 1}):
            program = adapter.convert_program(make_ir(k))
            want = pack_task_record(program, int(diag_ids[k]))
            got = recs[k]
            if want.shape != got.shape or not np.array_equal(want.view(np.uint64), got.view(np.uint64)):
                return slow(f"vectorised record of task {k} differs from the reference pipeline's")
    metadata = assignments
    low = LoweredSweep(n, metadata, list(recs), durations, groups, sp.filling, sites, sp.n_atoms, "vectorised", make_ir)
    LAST_LOWERING = low.how
    return low


# -----------------------------------------------------------------------------------------------------------
# execution: one bqa_run_batch / bqa_evolve_batch per (Hilbert space, device)
# -----------------------------------------------------------------------------------------------------------
def _resolve_devices(devices, multiprocessing: bool, num_workers: Optional[int]) -> List[int]:
    """`devices=` (new, keyword-only) or the reference's `multiprocessing / num_workers` (task/batch.py:235-244:
    N worker processes) mapped onto N GPUs"""
    count = _lib.device_count()
    if count < 1:
        raise _lib.BackendError("no CUDA device available: the B200 backend has no CPU fallback")
    if devices is not None:
        devs = list(range(count))[: int(devices)] if isinstance(devices, (int, np.integer)) else [int(d) for d in devices]
        if not devs or any(d < 0 or d >= count for d in devs):
            raise ValueError(f"devices={devices!r}: this machine has {count} CUDA device(s)")
        return devs
    if multiprocessing:
        n = count if num_workers is None else max(1, min(int(num_workers), count))
        return list(range(n))
    return [_lib.current_device()]


class _DeviceContext:
    """plans + van-der-Waals diagonals of the sweep's Hilbert spaces on one device"""

    def __init__(self, device: int, spaces: Dict):
        self.codegen = RydbergHamiltonianCodeGen(CompileCache(shared_spaces=spaces), device=device)

    def plan(self, group: SpaceGroup):
        space, plan = self.codegen._space_and_plan(group.register)
        for i, reg in enumerate(group.site_registers):  # diag ids in first-appearance order, as in the records
            diag_id, _ = self.codegen._diagonal(reg, space, plan)
            assert diag_id == i, "diagonal ids must follow the order of SpaceGroup.site_registers"
        return plan


def _split(idx: List[int], durations: np.ndarray, n_dev: int) -> List[List[int]]:
    """independent tasks over devices: longest first, dealt round-robin (near-equal work without communication)"""
    if n_dev == 1:
        return [list(idx)]
    order = sorted(idx, key=lambda k: -durations[k])
    return [sorted(order[d::n_dev]) for d in range(n_dev)]


def _execute(low: LoweredSweep, devices: List[int], job: Callable):
    """job(plan, task_numbers) runs on one device thread per device; exceptions are re-raised in the caller"""
    spaces: Dict = {}
    # Hilbert spaces are built once on the host and shared by the per-device plans
    for g in low.groups:
        RydbergHamiltonianCodeGen(CompileCache(shared_spaces=spaces)).space_only(g.register)
    work: Dict[int, List[Tuple[SpaceGroup, List[int]]]] = {d: [] for d in devices}
    for g in low.groups:
        for d, part in zip(devices, _split(g.task_idx, low.durations, len(devices))):
            if part:
                work[d].append((g, part))
    errors: List[BaseException] = []

    def run_device(device: int):
        try:
            ctx = _DeviceContext(device, spaces)
            for g, part in work[device]:
                job(ctx.plan(g), part)
        except BaseException as e:  # noqa: BLE001
            errors.append(e)

    if len(devices) == 1:
        run_device(devices[0])
    else:
        threads = [threading.Thread(target=run_device, args=(d,)) for d in devices]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    if errors:
        raise errors[0]


def _concat(low: LoweredSweep, part: List[int]):
    recs = [low.records[k] for k in part]
    offsets = np.zeros(len(recs) + 1, dtype=np.int64)
    np.cumsum([r.size for r in recs], out=offsets[1:])
    return np.ascontiguousarray(np.concatenate(recs)), offsets


def run_lowered(low: LoweredSweep, shots: int, opts, devices: List[int], uniforms: np.ndarray):
    """evolve + normalise + sample every task -> (digits uint8[n_tasks, shots, n_atoms], stats[n_tasks])"""
    out = np.empty((low.n_tasks, int(shots), low.n_atoms), dtype=np.uint8)
    stats = np.zeros(low.n_tasks, dtype=_lib.STATS_DTYPE)

    def job(plan, part):
        blob, offsets = _concat(low, part)
        s, st = plan.run_batch(blob, offsets, low.durations[part], uniforms[part], opts, True)
        out[part] = s
        stats[part] = st

    _execute(low, devices, job)
    return out, stats


def evolve_lowered(low: LoweredSweep, opts, devices: List[int], max_bytes: int = 1 << 30):
    """final state of every task from |0...0> (host complex128 rows), batched per Hilbert space and device"""
    states: List[Optional[np.ndarray]] = [None] * low.n_tasks
    stats = np.zeros(low.n_tasks, dtype=_lib.STATS_DTYPE)

    def job(plan, part):
        chunk = max(1, int(max_bytes // (16 * plan.dim)))
        for i in range(0, len(part), chunk):
            sub = part[i:i + chunk]
            blob, offsets = _concat(low, sub)
            psi, st = plan.evolve_batch(blob, offsets, low.durations[sub].reshape(-1, 1), opts)
            for j, k in enumerate(sub):
                states[k] = psi[j, 0]
                stats[k] = st[j, 0]

    _execute(low, devices, job)
    return states, stats


# -----------------------------------------------------------------------------------------------------------
# result types: subclasses of the reference's BloqadeTask / LocalBatch with array-backed, lazily built parts
# -----------------------------------------------------------------------------------------------------------
def _classes():
    global _CLASSES
    if _CLASSES is not None:
        return _CLASSES
    import pandas as pd
    from bloqade.analog.ir.routine.bloqade import BloqadePythonRoutine, TaskData
    from bloqade.analog.serialize import Serializer
    from bloqade.analog.submission.ir.task_results import (
        QuEraShotResult,
        QuEraShotStatusCode,
        QuEraTaskResults,
        QuEraTaskStatusCode,
    )
    from bloqade.analog.task.base import Geometry, Report
    from bloqade.analog.task.batch import LocalBatch
    from bloqade.analog.task.bloqade import BloqadeTask

    class B200Task(BloqadeTask):
        """`BloqadeTask` (task/bloqade.py:24-88) whose shots live in ONE uint8 array per batch: `post` is this
        task's view [shots, n_sites] in the QuEra convention (1 = ground / present, 0 = Rydberg / empty),
        `task_result_ir` and `emulator_ir` are built on first access."""

        def __init__(self, shots, metadata, geometry, post=None, make_ir=None, emulator_ir=None,
                     compile_cache=None, task_result_ir=None):
            self.shots = shots
            self.metadata = metadata
            self.compile_cache = compile_cache
            self._geom = geometry
            self._post = post
            self._make_ir = make_ir
            self._emulator_ir = emulator_ir
            self._result = task_result_ir

        @property
        def emulator_ir(self):
            if self._emulator_ir is None and self._make_ir is not None:
                self._emulator_ir = self._make_ir()
            return self._emulator_ir

        @emulator_ir.setter
        def emulator_ir(self, value):
            self._emulator_ir = value

        @property
        def task_result_ir(self):
            if self._result is None and self._post is not None:
                pre = [int(f) for f in self._geom.filling]
                done = QuEraShotStatusCode.Completed
                outputs = [QuEraShotResult.construct(shot_status=done, pre_sequence=pre, post_sequence=row)
                           for row in self._post.tolist()]
                self._result = QuEraTaskResults.construct(task_status=QuEraTaskStatusCode.Completed, shot_outputs=outputs)
            return self._result

        @task_result_ir.setter
        def task_result_ir(self, value):
            self._result = value
            self._post = None

        def _geometry(self):
            return self._geom

        def run(self, *args, **kwargs):
            """re-running one task goes through the reference's own CPU code path (task/bloqade.py:43-88)"""
            self._post = None
            return BloqadeTask.run(self, *args, **kwargs)

        def __reduce__(self):
            return (_rebuild_task, (self.shots, self.emulator_ir, self.metadata, self.task_result_ir))

    def _task_to_dict(obj):
        return {"shots": obj.shots, "emulator_ir": obj.emulator_ir, "metadata": obj.metadata,
                "task_result_ir": obj.task_result_ir.dict() if obj.task_result_ir else None}

    # saved batches are plain reference objects: a stock bloqade-analog can load them
    Serializer.serializers[B200Task] = _task_to_dict
    Serializer.type_to_str[B200Task] = Serializer.type_to_str[BloqadeTask]

    class B200Report(Report):
        """`Report` (task/base.py:186-424) of an emulator batch: every task has one cluster and no sorting step,
        so `bitstrings()` is a slice of the row block of each task instead of one mask over ALL rows per task
        (the reference's loop is quadratic in the number of tasks: 47 s for 4096 x 100 shots)"""

        def __init__(self, data, metas, geos, name, task_rows):
            Report.__init__(self, data, metas, geos, name)
            self._task_rows = task_rows  # task_number -> (first row, last row + 1), tasks in dataframe order

        def bitstrings(self, filter_perfect_filling: bool = True, clusters=[]):
            wanted = [clusters] if isinstance(clusters, tuple) else list(clusters)
            if any(tuple(c) != (0, 0) for c in wanted):
                return Report.bitstrings(self, filter_perfect_filling, clusters)
            values = self.dataframe.to_numpy()
            return [values[a:b].copy() for a, b in self._task_rows]

    class B200LocalBatch(LocalBatch):
        """`LocalBatch` (task/batch.py:114-254) with an array-built `report()`"""

        def report(self):
            fast = self._fast_report()
            return fast if fast is not None else LocalBatch.report(self)

        def _fast_report(self):
            tasks = list(self.tasks.items())
            if not tasks or any(not isinstance(t, B200Task) or t._post is None or t._geom.parallel_decoder
                                for _, t in tasks):
                return None
            n_sites = tasks[0][1]._post.shape[1]
            if any(t._post.shape[1] != n_sites for _, t in tasks):
                return None
            data = np.concatenate([t._post for _, t in tasks]).astype(np.int8)
            counts = np.array([t._post.shape[0] for _, t in tasks])
            numbers = np.array([k for k, _ in tasks])
            fillings = ["".join(map(str, (int(f) for f in t._geom.filling))) for _, t in tasks]
            levels_fill, codes_fill = np.unique(np.array(fillings, dtype=object), return_inverse=True)
            levels_num, codes_num = np.unique(numbers, return_inverse=True)
            cluster = np.empty(1, dtype=object)
            cluster[0] = (0, 0)
            index = pd.MultiIndex(
                levels=[levels_num, cluster, levels_fill, levels_fill],
                codes=[np.repeat(codes_num, counts), np.zeros(len(data), dtype=np.int8), np.repeat(codes_fill, counts),
                       np.repeat(codes_fill, counts)],
                names=["task_number", "cluster", "perfect_sorting", "pre_sequence"], verify_integrity=False)
            df = pd.DataFrame(data, index=index)
            metas = [t.metadata for _, t in tasks]
            geos = [t.geometry for _, t in tasks]
            ends = np.cumsum(counts)
            rows = list(zip((ends - counts).tolist(), ends.tolist()))
            return B200Report(df, metas, geos, "Local" if self.name is None else self.name, rows)

    def _batch_to_dict(obj):
        return {"source": None, "tasks": [(k, v) for k, v in obj.tasks.items()], "name": obj.name}

    Serializer.serializers[B200LocalBatch] = _batch_to_dict
    Serializer.type_to_str[B200LocalBatch] = Serializer.type_to_str[LocalBatch]

    class LazyTaskData:
        """`TaskData` (RB:46-56) whose `emulator_ir` is compiled by the reference pipeline on first access"""

        def __init__(self, task_id, metadata_dict, make_ir):
            self.task_id = task_id
            self.metadata_dict = metadata_dict
            self._make_ir = make_ir
            self._ir = None

        @property
        def emulator_ir(self):
            if self._ir is None:
                self._ir = self._make_ir()
            return self._ir

        @property
        def metadata(self):
            return TaskData.metadata.fget(self)

    class BloqadeB200Routine(BloqadePythonRoutine):
        """Same signatures, defaults and return types as `BloqadePythonRoutine` (RB:233-249,314-329,348-365,
        505-512); new options are keyword-only with defaults."""

        def run(self, shots: int, args: Tuple = (), name: Optional[str] = None, blockade_radius: float = 0.0,
                waveform_runtime: str = "interpret", interaction_picture: bool = False, cache_matrices: bool = False,
                multiprocessing: bool = False, num_workers: Optional[int] = None, solver_name: str = "dop853",
                atol: float = 1e-7, rtol: float = 1e-14, nsteps: int = 2_147_483_647, *,
                devices: Union[None, int, Sequence[int]] = None, raise_on_error: bool = True):
            if multiprocessing and cache_matrices:
                raise ValueError("Cannot use multiprocessing and cache_matrices at the same time.")
            if num_workers is not None and not multiprocessing:
                raise ValueError("num_workers is only used when multiprocessing is enabled.")
            if solver_name not in AnalogGate.SUPPORTED_SOLVERS:
                raise ValueError(f"'{solver_name}' not supported.")
            opts = _lib.make_opts(solver_name, atol, rtol, nsteps, interaction_picture)
            devs = _resolve_devices(devices, multiprocessing, num_workers)
            low = _lower(self, args, blockade_radius, waveform_runtime, False)
            # NumPy's global RandomState, task by task in task order: the draws of the reference's serial loop
            # (space.py:217), so a seeded run reproduces the reference's shots
            uniforms = np.random.random_sample((low.n_tasks, int(shots)))
            digits, stats = run_lowered(low, shots, opts, devs, uniforms)
            if raise_on_error:
                for k in range(low.n_tasks):
                    AnalogGate._error_check(solver_name, int(stats["status"][k]))
            return self._package(low, shots, digits, stats, name)

        def _package(self, low: LoweredSweep, shots, digits, stats, name):
            """task/bloqade.py:66-86 for the whole batch at once: post = 1 - digit scattered onto the filled sites"""
            filling = np.asarray(low.filling, dtype=int)
            post = np.zeros((low.n_tasks, int(shots), filling.size), dtype=np.uint8)
            post[:, :, filling == 1] = 1 - digits
            tasks = OrderedDict()
            geometry_cache: Dict[int, Any] = {}
            for k in range(low.n_tasks):
                sites = low.sites[k]
                geom = geometry_cache.get(id(sites))
                if geom is None:
                    geom = geometry_cache[id(sites)] = Geometry(sites, [bool(f) for f in low.filling])
                task = B200Task(shots, low.metadata[k], geom, post[k], make_ir=lambda k=k: low.make_ir(k))
                task.integrator_stats = stats[k]
                tasks[k] = task
            return B200LocalBatch(self.source, tasks, name)

        def __call__(self, *args, shots: int = 1, name: Optional[str] = None, blockade_radius: float = 0.0,
                     waveform_runtime: str = "interpret", interaction_picture: bool = False,
                     multiprocessing: bool = False, num_workers: Optional[int] = None, cache_matrices: bool = False,
                     solver_name: str = "dop853", atol: float = 1e-7, rtol: float = 1e-14,
                     nsteps: int = 2_147_483_647):
            return self.run(shots=shots, args=args, name=name, blockade_radius=blockade_radius,
                            waveform_runtime=waveform_runtime, multiprocessing=multiprocessing, num_workers=num_workers,
                            cache_matrices=cache_matrices, solver_name=solver_name, atol=atol, rtol=rtol, nsteps=nsteps,
                            interaction_picture=interaction_picture)

        def run_callback(self, callback: Callable, program_args: Tuple = (), callback_args: Tuple = (),
                         ignore_exceptions: bool = False, blockade_radius: float = 0.0,
                         waveform_runtime: str = "interpret", interaction_picture: bool = False,
                         cache_matrices: bool = False, multiprocessing: bool = False,
                         num_workers: Optional[int] = None, solver_name: str = "dop853", atol: float = 1e-7,
                         rtol: float = 1e-14, nsteps: int = 2_147_483_647, use_hyperfine: bool = False, *,
                         devices: Union[None, int, Sequence[int]] = None) -> List[Any]:
            """all tasks are evolved in batched GPU calls; the callbacks then run in this process, in task order
            (so they need not be picklable), each on a host copy of its state"""
            if multiprocessing and cache_matrices:
                raise ValueError("Cannot use multiprocessing and cache_matrices at the same time.")
            if solver_name not in AnalogGate.SUPPORTED_SOLVERS:
                raise ValueError(f"'{solver_name}' not supported.")
            opts = _lib.make_opts(solver_name, atol, rtol, nsteps, interaction_picture)
            devs = _resolve_devices(devices, multiprocessing, num_workers)
            low = _lower(self, program_args, blockade_radius, waveform_runtime, use_hyperfine)
            states, stats = evolve_lowered(low, opts, devs)
            cache = CompileCache()
            results: List[Any] = []
            first_error = None
            for k in range(low.n_tasks):
                try:
                    AnalogGate._error_check(solver_name, int(stats["status"][k]))
                    em = B200Emulation(LazyTaskData(k, low.metadata[k], lambda k=k: low.make_ir(k)), cache)
                    state = StateVector(states[k], _LazySpace(em))
                    results.append(callback(state, em.metadata, _LazyHamiltonian(em), *callback_args))
                except BaseException as e:  # same policy as RB:150-158,488-501
                    results.append(e)
                    if first_error is None:
                        first_error = (k, e)
            if first_error is not None and not ignore_exceptions:
                raise RuntimeError(f"{first_error[1].__class__.__name__} occured during child process running for task "
                                   f"number {first_error[0]}:\n{first_error[1]}") from first_error[1]
            return results

        def hamiltonian(self, *args, blockade_radius: float = 0.0, use_hyperfine: bool = False,
                        waveform_runtime: str = "interpret", cache_matrices: bool = False) -> List["B200Emulation"]:
            cache = CompileCache() if cache_matrices else None
            return [B200Emulation(td, cache)
                    for td in self._generate_ir(args, blockade_radius, waveform_runtime, use_hyperfine)]

    _CLASSES = dict(routine=BloqadeB200Routine, task=B200Task, batch=B200LocalBatch, task_data=LazyTaskData)
    return _CLASSES


def _rebuild_task(shots, emulator_ir, metadata, task_result_ir):
    """pickled B200 tasks come back as plain reference BloqadeTask objects"""
    from bloqade.analog.task.bloqade import BloqadeTask

    return BloqadeTask(shots, emulator_ir, metadata, None, task_result_ir)


class _LazyProxy:
    """stands for an object that is only built when a callback touches it"""

    __slots__ = ("_build", "_obj")

    def __init__(self, build):
        object.__setattr__(self, "_build", build)
        object.__setattr__(self, "_obj", None)

    def _get(self):
        if object.__getattribute__(self, "_obj") is None:
            object.__setattr__(self, "_obj", object.__getattribute__(self, "_build")())
        return object.__getattribute__(self, "_obj")

    def __getattr__(self, name):
        return getattr(self._get(), name)

    def __eq__(self, other):
        other = other._get() if isinstance(other, _LazyProxy) else other
        return self._get() == other

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash(self._get())

    def __repr__(self):
        return repr(self._get())


class _LazyHamiltonian(_LazyProxy):
    def __init__(self, emulation):
        super().__init__(lambda: emulation.hamiltonian)


class _LazySpace(_LazyProxy):
    def __init__(self, emulation):
        super().__init__(lambda: emulation.hamiltonian.space)


class B200Emulation:
    """Counterpart of `BloqadeEmulation` (RB:58-145): lazily generated GPU Hamiltonian + evolve()."""

    def __init__(self, task_data, compile_cache: Optional[CompileCache] = None):
        self.task_data = task_data
        self.compile_cache = compile_cache
        self._hamiltonian = None

    @property
    def metadata(self):
        return self.task_data.metadata

    @property
    def hamiltonian(self):
        if self._hamiltonian is None:
            program = adapter.convert_program(self.task_data.emulator_ir)
            self._hamiltonian = RydbergHamiltonianCodeGen(self.compile_cache).emit(program)
        return self._hamiltonian

    def zero_state(self, dtype=np.float64) -> StateVector:
        return self.hamiltonian.space.zero_state(dtype)

    def fock_state(self, fock_state_str: str, dtype=np.float64) -> StateVector:
        space = self.hamiltonian.space
        state = StateVector(np.zeros(space.size, dtype=dtype), space)
        state.data[space.fock_state_to_index(fock_state_str)] = 1
        return state

    def evolve(self, state: Optional[StateVector] = None, solver_name: str = "dop853", atol: float = 1e-7,
               rtol: float = 1e-14, nsteps: int = 2_147_483_647, times=(), interaction_picture: bool = False):
        state = self.zero_state(np.complex128) if state is None else state
        return AnalogGate(self.hamiltonian).apply(state, solver_name, atol, rtol, nsteps, times, interaction_picture)


def routine_class():
    """`BloqadeB200Routine`, created on first use (importing bloqade.analog is deferred until then)."""
    return _classes()["routine"]


def b200_routine(program):
    """`program` is a builder or a Routine; returns the B200 routine for it (no global patching)."""
    routine = program.parse() if hasattr(program, "parse") else program
    return routine_class()(routine.source, routine.circuit, routine.params)


def install():
    """Opt-in: make the literal `program.bloqade.python()` return the B200 routine."""
    global _ORIGINAL_PYTHON
    from bloqade.analog.ir.routine.bloqade import BloqadeServiceOptions

    cls = routine_class()
    if _ORIGINAL_PYTHON is None:
        _ORIGINAL_PYTHON = BloqadeServiceOptions.python
    BloqadeServiceOptions.python = lambda self: cls(self.source, self.circuit, self.params)
    BloqadeServiceOptions.b200 = BloqadeServiceOptions.python


def uninstall():
    global _ORIGINAL_PYTHON
    if _ORIGINAL_PYTHON is not None:
        from bloqade.analog.ir.routine.bloqade import BloqadeServiceOptions

        BloqadeServiceOptions.python = _ORIGINAL_PYTHON
        _ORIGINAL_PYTHON = None
