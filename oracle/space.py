"""ORACLE (test infrastructure only - never imported by the product path).

CPU restatement of the reference's Hilbert-space construction and index maps:
  * `Space.create`                      src/bloqade/analog/emulate/ir/space.py:39-103
  * `swap_state_at / transition_state_at`  space.py:134-178, emulate/ir/atom_type.py:80-122,146-161
  * `sample_state_vector`               space.py:211-229
It follows the reference's algorithm (grow the configuration list atom by atom with Kronecker sums,
blockaded configurations lose the Rydberg option, final sort) independently of the product's
`bloqade_analog_b200.space`, which builds the same list without a sort.

Pinned by tests/test_oracle_golden.py against the literal vectors of the reference's tests/test_space.py
and against arrays dumped from the reference itself (tests/golden/*.npz).
"""
from __future__ import annotations

from decimal import Decimal

import numpy as np


def _dec(x):
    return x if isinstance(x, Decimal) else Decimal(str(x))


def is_state_at(configs, index, state, n_level):
    if n_level == 2:
        return ((configs >> index) & 1) == state
    return (configs // 3**index) % 3 == state


def create_space(sites, n_level: int, blockade_radius):
    """returns (is_full, configurations)  - space.py:39-103"""
    sites = [(_dec(x), _dec(y)) for x, y in sites]
    radius = _dec(blockade_radius)
    n_atom = len(sites)
    total = n_level**n_atom
    check_atoms = []
    for i in range(1, n_atom):
        near = []
        for j in range(i):
            dx, dy = sites[i][0] - sites[j][0], sites[i][1] - sites[j][1]
            if (dx * dx + dy * dy).sqrt() <= radius:
                near.append(j)
        check_atoms.append(near)
    config_type = np.result_type(np.min_scalar_type(total - 1), np.uint32)
    if all(len(c) == 0 for c in check_atoms):
        return True, np.arange(total, dtype=config_type)
    states = np.arange(n_level, dtype=config_type)
    configs = states
    ryd = n_level - 1
    for i, near in enumerate(check_atoms, 1):
        free = np.ones(configs.shape, dtype=bool)
        for j in near:
            free &= ~is_state_at(configs, j, ryd, n_level)
        open_cfg = configs[free]
        blocked_cfg = configs[~free]
        shift = states * config_type.type(n_level**i)
        grown_open = np.kron(open_cfg, np.ones_like(states)) + np.kron(np.ones_like(open_cfg), shift)
        grown_blocked = np.kron(blocked_cfg, np.ones_like(states[:-1])) + np.kron(np.ones_like(blocked_cfg), shift[:-1])
        configs = np.hstack((grown_blocked, grown_open))
    configs.sort()
    return False, configs


def _locate(configs, is_full, rows, targets):
    if is_full:
        return rows, targets.astype(np.int64)
    pos = np.searchsorted(configs, targets)
    ok = pos < configs.size
    ok[ok] = configs[pos[ok]] == targets[ok]
    return rows[ok], pos[ok]


def transition_pairs(configs, is_full, n_level, index, fro, to):
    """(row_indices, col_indices): rows have atom `index` in `fro`, cols the same state with it in `to`"""
    rows = np.flatnonzero(is_state_at(configs, index, fro, n_level))
    src = configs[rows].astype(np.int64)
    tgt = src + (to - fro) * n_level**index
    return _locate(configs, is_full, rows, tgt.astype(configs.dtype))


def swap_pairs(configs, is_full, n_level, index, s1, s2):
    """symmetric exchange s1 <-> s2 on atom `index` as (row_indices, col_indices)"""
    r1, c1 = transition_pairs(configs, is_full, n_level, index, s1, s2)
    r2, c2 = transition_pairs(configs, is_full, n_level, index, s2, s1)
    rows = np.concatenate((r1, r2))
    cols = np.concatenate((c1, c2))
    order = np.argsort(rows, kind="stable")
    return rows[order], cols[order]


def decode_samples(sampled_configs, n_atoms, n_level, project_hyperfine=True):
    """space.py:219-229: digits base n_level, atom 0 least significant; 3-level projection h->0, r->1"""
    sampled = np.array(sampled_configs, copy=True)
    out = np.empty((sampled.size, n_atoms), dtype=np.uint8)
    for a in range(n_atoms):
        out[:, a] = sampled % n_level
        sampled //= n_level
    if project_hyperfine and n_level == 3:
        out[out == 1] = 0
        out[out == 2] = 1
    return out


def sample_state_vector(configs, psi, uniforms, n_atoms, n_level, project_hyperfine=True):
    """np.random.choice(configs, p=|psi|^2) with the uniforms made explicit (numpy legacy RandomState.choice:
    cdf = p.cumsum(); cdf /= cdf[-1]; idx = cdf.searchsorted(u, side='right'))"""
    p = np.abs(psi) ** 2
    cdf = p.cumsum()
    cdf /= cdf[-1]
    idx = cdf.searchsorted(np.asarray(uniforms), side="right")
    idx = np.minimum(idx, configs.size - 1)
    return decode_samples(configs[idx], n_atoms, n_level, project_hyperfine)
