"""world_size-2 gloo test of the sweep sharding: contiguous blocks, no data-path collective,
results identical to the single-rank order.  The GPU runner is replaced by a deterministic CPU stand-in: the
point here is the host-side partition / gather logic."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_exactly():
    from bloqade_analog_b200.batch import shard_range

    for n in (0, 1, 5, 8, 4096, 4097):
        for world in (1, 2, 3, 8):
            blocks = [shard_range(n, r, world) for r in range(world)]
            flat = [i for b in blocks for i in b]
            assert flat == list(range(n))
            assert max(len(b) for b in blocks) <= -(-n // world) if n else True


def _fake_runner(programs, shots, uniforms=None, raise_on_error=False, **kw):
    from bloqade_analog_b200.batch import TaskResult

    out = []
    for p, u in zip(programs, uniforms):
        bits = (u[:, None] * (np.arange(3)[None, :] + 1 + p) % 1.0 > 0.5).astype(np.uint8)
        out.append(TaskResult(bits, 1, int(100 * p), 0, 0, 0))
    return out


def _worker(rank, world, port, q):
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bloqade_analog_b200.distributed import run_programs_sharded

    np.random.seed(77)
    res = run_programs_sharded(list(range(7)), 16, runner=_fake_runner)
    q.put((rank, [r.shots.tobytes() for r in res], [r.nfcn for r in res]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_matches_single_rank():
    import torch.multiprocessing as mp

    sys.path.insert(0, ROOT)
    from bloqade_analog_b200.distributed import run_programs_sharded

    np.random.seed(77)
    ref = run_programs_sharded(list(range(7)), 16, runner=_fake_runner)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, shots, nfcn in got:
        assert shots == [r.shots.tobytes() for r in ref]
        assert nfcn == [r.nfcn for r in ref]
