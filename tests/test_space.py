"""bloqade_analog_b200.space.Space against the reference's own Space (emulate/ir/space.py:39-229) on 21 registers:
the positions / radii of the reference's tests/test_space.py plus 2-D lattices with a blockade radius, knife-edge
radii and random positions; two- and three-level atoms, full spaces and blockade subspaces.  The fixture
(tests/golden/space_cases.*) was produced by the reference itself (tests/golden/make_space_golden.py).  Bit-exact:
configurations, index / state dtypes, every swap_state_at / transition_state_at (row, col) pair, fock strings.
CPU only - subspace construction is host logic (north_star: "subspace index construction must be bit-exact")."""
import json
import os

import numpy as np
import pytest

from bloqade_analog_b200 import ir
from bloqade_analog_b200.space import Space

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "space_cases.json")) as _f:
    META = json.load(_f)
_NPZ = None


def arrays():
    global _NPZ
    if _NPZ is None:
        _NPZ = np.load(os.path.join(HERE, "golden", "space_cases.npz"))
    return _NPZ


def make_space(m):
    atom = ir.TwoLevelAtom if m["n_level"] == 2 else ir.ThreeLevelAtom
    return Space.create(ir.Register(atom, [tuple(p) for p in m["positions"]], m["radius"]))


@pytest.mark.parametrize("name", sorted(META))
def test_configurations_and_dtypes(name):
    m, a = META[name], arrays()
    space = make_space(m)
    ref = a[f"{name}/configurations"]
    assert space.size == m["size"] and space.n_atoms == m["n_atoms"]
    assert space.configurations.dtype == ref.dtype
    assert np.array_equal(space.configurations, ref)
    assert str(np.dtype(space.index_type)) == m["index_dtype"].replace("<class 'numpy.", "").replace("'>", "")
    assert str(np.dtype(space.state_type)) == m["state_dtype"].replace("<class 'numpy.", "").replace("'>", "")
    assert str(space) == m["str"]
    for i, fock in m["fock"].items():
        assert space.index_to_fock_state(int(i)) == fock
        assert space.fock_state_to_index(fock) == int(i)


@pytest.mark.parametrize("name", sorted(META))
def test_swap_and_transition_index_vectors(name):
    m, a = META[name], arrays()
    space = make_space(m)
    pairs = [(0, 1)] if m["n_level"] == 2 else [(0, 1), (1, 2), (0, 2)]
    for atom in range(m["n_atoms"]):
        for s1, s2 in pairs:
            for tag, fn in (("swap", space.swap_state_at), ("trans", space.transition_state_at)):
                for x, y in ((s1, s2), (s2, s1)):
                    key = f"{name}/{tag}_{atom}_{x}_{y}"
                    rows, cols = fn(atom, x, y)
                    assert isinstance(rows, slice) == bool(a[key + "/rows_is_slice"]), key
                    rows = np.arange(space.size) if isinstance(rows, slice) else np.asarray(rows)
                    assert np.array_equal(rows, a[key + "/rows"]), key
                    assert np.array_equal(np.asarray(cols), a[key + "/cols"]), key
                    assert np.asarray(cols).dtype == a[key + "/cols"].dtype, key


def test_partner_table_is_the_gather_form_of_swap_state_at():
    """the int32 table the CUDA kernels read equals the (row, col) pairs above (col where a partner exists, -1 else)"""
    for name in ("square4_r1p1", "chain5_3level_r1p1", "random7_3level_r1p6", "square3_full"):
        m = META[name]
        space = make_space(m)
        table = space.partner_table()
        for atom in range(m["n_atoms"]):
            for s in range(m["n_level"] - 1):
                rows, cols = space.swap_state_at(atom, s, s + 1)
                rows = np.arange(space.size) if isinstance(rows, slice) else np.asarray(rows)
                expect = np.full(space.size, -1, dtype=np.int64)
                if space.space_type.name == "FullSpace":
                    expect[rows] = np.searchsorted(space.configurations, cols)
                else:
                    expect[rows] = cols
                assert np.array_equal(table[atom, s].astype(np.int64), expect), (name, atom, s)
