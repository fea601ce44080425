"""Host logic of the sharded state (bloqade_analog_b200/sharded.py): partner schedule and shard-wise Born sampling.
No GPU: the kernels are covered by tests/test_gpu_parity.py (single-device emulation) and profiles/shard_check.py
(two GPUs)."""
import numpy as np
import pytest

from bloqade_analog_b200.sharded import partner_ranks, split_sample_uniforms


def test_partner_ranks_are_xor_involutions():
    for world in (2, 4, 8):
        for r in range(world):
            ps = partner_ranks(r, world)
            assert len(ps) == world.bit_length() - 1
            for b, q in enumerate(ps):
                assert q == r ^ (1 << b) and partner_ranks(q, world)[b] == r
    with pytest.raises(ValueError):
        partner_ranks(0, 3)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_shardwise_sampling_equals_global_cdf_sampling(world):
    """owner/rescaled-uniform split followed by a local CDF search == one search in the global CDF
    (what np.random.choice does, emulate/ir/space.py:217)"""
    rng = np.random.default_rng(world)
    dim = 1 << 12
    p = rng.random(dim) ** 3
    p[rng.integers(0, dim, 100)] = 0.0
    ldim = dim // world
    weights = [p[r * ldim:(r + 1) * ldim].sum() for r in range(world)]
    u = rng.random(5000)
    owner, lu = split_sample_uniforms(u, weights)
    cdf = np.cumsum(p)
    ref = np.searchsorted(cdf, u * cdf[-1], side="right")
    got = np.empty_like(ref)
    for j in range(len(u)):
        loc = p[owner[j] * ldim:(owner[j] + 1) * ldim]
        c = np.cumsum(loc)
        got[j] = owner[j] * ldim + min(np.searchsorted(c, lu[j] * c[-1], side="right"), ldim - 1)
    assert (got != ref).mean() < 2e-3  # only uniforms within rounding of a bin edge may differ
    assert np.all(p[got] > 0)


def test_checkpoint_grid_has_no_one_ulp_segment():
    """0.02 + 0.16 + 0.02 = 0.20000000000000004: three checkpoints of a third each must not leave a fourth segment
    of one ulp (the integrator would end it with status -3)"""
    from bloqade_analog_b200.sharded import checkpoint_grid

    d = 0.02 + 0.16 + 0.02
    g = checkpoint_grid(d, 0.2 / 3.0)
    assert len(g) == 3 and g[-1] == d and g[0] < g[1] < g[2]
    assert checkpoint_grid(1.0, 0.4) == [0.4, 0.8, 1.0]
    assert checkpoint_grid(1.0, 2.0) == [1.0]
