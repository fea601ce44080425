"""Host logic of the shared-memory-tiled H psi (csrc/bqa_apply_tiled.cuh, tiled_plan in bqa_api_ops.cuh), checked
without a GPU through bqa_tiled_plan_describe: every atom of the register is flipped by exactly one pass, the
tile groups of a pass do not overlap, and the tile walk of a pass (tile_seq / tile_base, restated here) visits
every tile exactly once with partner tiles back to back."""
import pytest

from bloqade_analog_b200 import _lib

TILE_BITS = 13


def tile_base(s, tile):
    mid = s["hb"] - s["lo_bits"]
    return ((tile & ((1 << mid) - 1)) << s["lo_bits"]) | ((tile >> mid) << (s["hb"] + s["tb"]))


def tile_seq(s, block, grid, it, n_tiles):
    if s["pair"]:
        q = block + (it >> 1) * grid
        if q >= n_tiles // 2:
            return -1
        b = s["pair_b"]
        t = ((q >> b) << (b + 1)) | (q & ((1 << b) - 1))
        return t | ((it & 1) << b)
    t = block + it * grid
    return t if t < n_tiles else -1


def atoms_of_pass(s):
    """atoms flipped inside the tile by this pass (tile bit p >= p0 -> atom)"""
    out = []
    for p in range(s["p0"], TILE_BITS):
        out.append(p if p < s["lo_bits"] else s["hb"] + (p - s["lo_bits"]))
    return out


@pytest.mark.parametrize("g", [0, 1, 2, 3])
@pytest.mark.parametrize("n_local", list(range(14, 31)))
def test_every_atom_is_flipped_exactly_once(n_local, g):
    passes = _lib.tiled_plan_describe(n_local, n_local + g)
    assert passes, "shapes with more than 13 local atoms are tiled"
    seen = []
    for i, s in enumerate(passes):
        if i == 0:
            assert (s["lo_bits"], s["tb"], s["p0"]) == (TILE_BITS, 0, 0)
        else:
            assert s["lo_bits"] + s["tb"] == TILE_BITS and 2 <= s["lo_bits"] <= 9
            assert s["hb"] >= s["lo_bits"] and s["hb"] + s["tb"] <= n_local
        seen += atoms_of_pass(s)
        if s["n_x"]:
            seen.append(s["xbit"])
            # the leftover atom is never a tile bit of its own pass
            assert s["xbit"] not in range(0, s["lo_bits"]) and s["xbit"] not in range(s["hb"], s["hb"] + s["tb"])
            # rank bits are read from the peer shard, local leftovers walk partner pairs
            assert s["pair"] == (1 if s["xbit"] < n_local else 0)
    assert sorted(seen) == list(range(n_local + g))
    # the rank bits come first, one per pass
    assert [s["xbit"] for s in passes[:g]] == list(range(n_local, n_local + g))


def test_config_shapes():
    """BASELINE configs 4 and 5: 25 atoms in 2 passes (10 fresh atoms with 128-byte lines + one leftover each),
    27 local atoms of the 30-atom register on 8 GPUs in 3 passes (one rank bit each)"""
    c4 = _lib.tiled_plan_describe(25)
    assert len(c4) == 2
    assert (c4[0]["n_x"], c4[0]["xbit"], c4[0]["pair"], c4[0]["pair_b"]) == (1, 13, 1, 0)
    assert (c4[1]["lo_bits"], c4[1]["hb"], c4[1]["tb"], c4[1]["p0"], c4[1]["xbit"]) == (3, 14, 10, 3, 24)
    assert len(_lib.tiled_plan_describe(26)) == 2 and len(_lib.tiled_plan_describe(27)) == 3
    c5 = _lib.tiled_plan_describe(27, 30)
    assert len(c5) == 3 and [s["xbit"] for s in c5] == [27, 28, 29] and not any(s["pair"] for s in c5)


@pytest.mark.parametrize("n_local,grid", [(14, 1), (16, 148), (19, 148), (21, 148), (22, 7)])
def test_tile_walk_visits_every_tile_once_and_pairs_partners(n_local, grid):
    n_tiles = 1 << (n_local - TILE_BITS)
    for s in _lib.tiled_plan_describe(n_local):
        visited, bases = [], set()
        for block in range(min(grid, n_tiles)):
            it = 0
            while True:
                t = tile_seq(s, block, min(grid, n_tiles), it, n_tiles)
                if t < 0:
                    break
                visited.append(t)
                if s["pair"] and it & 1:
                    # the second tile of a pair is the partner of the first under the leftover atom
                    assert tile_base(s, t) == tile_base(s, visited[-2]) ^ (1 << s["xbit"])
                bases.add(tile_base(s, t))
                it += 1
        assert sorted(visited) == list(range(n_tiles))
        # tile bases are distinct and leave the tile bits clear
        assert len(bases) == n_tiles
        tile_mask = ((1 << s["lo_bits"]) - 1) | (((1 << s["tb"]) - 1) << s["hb"])
        assert all(b & tile_mask == 0 for b in bases)
