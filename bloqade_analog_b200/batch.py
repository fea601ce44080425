"""Batched task execution: the B200 replacement of `LocalBatch._run`'s per-task loop
(`src/bloqade/analog/task/batch.py:216-254`) and of `BloqadeTask.run`
(`src/bloqade/analog/task/bloqade.py:43-88`).

All tasks that share a Hilbert space (same `Register` in the reference's sense, emulator.py:197-226)
are evolved, normalised and sampled by ONE `bqa_run_batch` call: one CTA per task, the task list
sharded over GPUs/ranks by contiguous blocks with no communication.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from .ir import EmulatorProgram
from .lowering import concat_records
from .state_vector import AnalogGate


@dataclass
class TaskResult:
    """Outcome of one task: sampled digits (uint8[shots, n_atoms], 1 = Rydberg) + integrator stats."""

    shots: np.ndarray
    status: int
    nfcn: int
    nstep: int
    naccpt: int
    nrejct: int

    def post_sequence(self, filling: Sequence[int]) -> np.ndarray:
        """QuEra convention of task/bloqade.py:66-82: 1 = ground / atom present, 0 = Rydberg or empty
        site; shots are scattered onto the filled sites."""
        filling = np.asarray(filling, dtype=int)
        full = np.zeros((self.shots.shape[0], filling.size), dtype=int)
        full[:, filling == 1] = 1 - self.shots.astype(int)
        return full


def shard_range(n: int, rank: int, world: int) -> range:
    """contiguous block of task numbers owned by `rank` (ceil(n/world) each)"""
    per = -(-n // world)
    return range(min(n, rank * per), min(n, (rank + 1) * per))


def run_programs(
    programs: Sequence[EmulatorProgram],
    shots: int,
    solver_name: str = "dop853",
    atol: float = 1e-7,
    rtol: float = 1e-14,
    nsteps: int = 2_147_483_647,
    interaction_picture: bool = False,
    project_hyperfine: bool = True,
    compile_cache: Optional[CompileCache] = None,
    device: Optional[int] = None,
    reject_mode: int = 1,
    engine: str = "auto",
    uniforms: Optional[np.ndarray] = None,
    raise_on_error: bool = True,
) -> List[TaskResult]:
    """Evolve + normalise + sample every program; results in input order.

    Sampling uniforms are drawn from NumPy's global RandomState task by task, in task order, exactly
    where the reference's serial loop would draw them (space.py:217), unless `uniforms[n_tasks, shots]`
    is given."""
    if solver_name not in AnalogGate.SUPPORTED_SOLVERS:
        raise ValueError(f"'{solver_name}' not supported.")
    opts = _lib.make_opts(solver_name, atol, rtol, nsteps, interaction_picture, reject_mode, engine)
    codegen = RydbergHamiltonianCodeGen(compile_cache, device=device)
    n = len(programs)
    if uniforms is None:
        uniforms = np.stack([np.random.random_sample(int(shots)) for _ in range(n)]) if n else np.zeros((0, shots))
    uniforms = np.asarray(uniforms, dtype=np.float64).reshape(n, int(shots))
    hams = [codegen.emit(p) for p in programs]
    groups: Dict[int, List[int]] = {}
    for i, h in enumerate(hams):
        groups.setdefault(id(h.plan), []).append(i)
    results: List[Optional[TaskResult]] = [None] * n
    for idx in groups.values():
        plan = hams[idx[0]].plan
        blob, offsets = concat_records([hams[i].record for i in idx])
        times = np.array([hams[i].emulator_ir.duration for i in idx], dtype=np.float64)
        shots_arr, stats = plan.run_batch(blob, offsets, times, uniforms[idx], opts, project_hyperfine)
        for j, i in enumerate(idx):
            s = stats[j]
            results[i] = TaskResult(shots_arr[j], int(s["status"]), int(s["nfcn"]), int(s["nstep"]),
                                    int(s["naccpt"]), int(s["nrejct"]))
    if raise_on_error:
        for r in results:
            AnalogGate._error_check(solver_name, r.status)
    return results  # type: ignore[return-value]


def evolve_programs(
    programs: Sequence[EmulatorProgram],
    solver_name: str = "dop853",
    atol: float = 1e-7,
    rtol: float = 1e-14,
    nsteps: int = 2_147_483_647,
    interaction_picture: bool = False,
    compile_cache: Optional[CompileCache] = None,
    device: Optional[int] = None,
    reject_mode: int = 1,
    engine: str = "auto",
) -> Tuple[List[np.ndarray], np.ndarray]:
    """Final state vectors (host complex128) of every program from |0...0>, batched per space."""
    opts = _lib.make_opts(solver_name, atol, rtol, nsteps, interaction_picture, reject_mode, engine)
    codegen = RydbergHamiltonianCodeGen(compile_cache, device=device)
    hams = [codegen.emit(p) for p in programs]
    groups: Dict[int, List[int]] = {}
    for i, h in enumerate(hams):
        groups.setdefault(id(h.plan), []).append(i)
    states: List[Optional[np.ndarray]] = [None] * len(programs)
    all_stats = np.zeros(len(programs), dtype=_lib.STATS_DTYPE)
    for idx in groups.values():
        plan = hams[idx[0]].plan
        blob, offsets = concat_records([hams[i].record for i in idx])
        times = np.array([[hams[i].emulator_ir.duration] for i in idx], dtype=np.float64)
        out, stats = plan.evolve_batch(blob, offsets, times, opts)
        for j, i in enumerate(idx):
            states[i] = out[j, 0]
            all_stats[i] = stats[j, 0]
    return states, all_stats  # type: ignore[return-value]
