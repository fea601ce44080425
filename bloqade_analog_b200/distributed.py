"""Multi-GPU execution of parameter sweeps: one process per GPU, tasks sharded by contiguous blocks, NO
data-path collective (every `BloqadeTask` is self-contained, ir/routine/bloqade.py:225-229;
the reference fans them out over processes, task/batch.py:235-244).  `torch.distributed` is only the
plumbing that gathers the small per-task results (uint8 shots + integrator stats) on every rank.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import numpy as np

from .batch import TaskResult, run_programs, shard_range


def sharded_uniforms(n_tasks: int, shots: int) -> np.ndarray:
    """The uniforms the reference's serial loop would draw (task by task from NumPy's global RandomState,
    space.py:217) - drawn identically on every rank (same seed => same stream), so sharding does not change
    any shot."""
    return np.stack([np.random.random_sample(int(shots)) for _ in range(n_tasks)]) if n_tasks else np.zeros((0, shots))


def run_programs_sharded(programs: Sequence, shots: int, *, runner: Optional[Callable] = None, group=None,
                         uniforms: Optional[np.ndarray] = None, **solver_kwargs) -> List[TaskResult]:
    """Evolve + sample `programs` over all ranks of the default (or given) process group; every rank
    returns the complete, ordered result list.  `runner` defaults to `batch.run_programs` on this rank's
    CUDA device (LOCAL_RANK); tests inject a CPU stand-in to exercise the sharding logic under gloo."""
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n = len(programs)
    if uniforms is None:
        uniforms = sharded_uniforms(n, shots)
    uniforms = np.asarray(uniforms, dtype=np.float64).reshape(n, int(shots))
    mine = shard_range(n, rank, world)
    runner = run_programs if runner is None else runner
    local: List[TaskResult] = []
    if len(mine):
        local = runner([programs[i] for i in mine], shots, uniforms=uniforms[mine.start:mine.stop],
                       raise_on_error=False, **solver_kwargs)
    if world == 1:
        return local
    gathered: List[Optional[list]] = [None] * world
    dist.all_gather_object(gathered, [(r.shots, r.status, r.nfcn, r.nstep, r.naccpt, r.nrejct) for r in local], group=group)
    out: List[TaskResult] = []
    for part in gathered:
        out.extend(TaskResult(*p) for p in part)
    assert len(out) == n
    return out
