"""Hilbert-space bookkeeping for the B200 emulator: full space and Rydberg-blockade subspace.

Mirror of the reference's `Space` (`src/bloqade/analog/emulate/ir/space.py:21-229`) with the same
attribute and method names.  Index construction must be BIT-EXACT with the reference
(north_star), which is pinned by `tests/test_space.py` against the reference's own literal
vectors (`tests/test_space.py:12-615` upstream) and against golden vectors generated from it.

Construction differs from the reference on purpose: the subspace is grown atom by atom with the
new atom's digit as the most significant one, so the list stays sorted without a final sort, and
the partner tables the CUDA kernels consume (`partner_table`) are produced directly in gather
form (one int32 per (atom, direction, basis state); -1 = partner outside the subspace).
"""
from __future__ import annotations

from dataclasses import dataclass
from decimal import Decimal
from enum import Enum
from typing import Any, Tuple

import numpy as np

from .ir import AtomType, Register, ThreeLevelAtomType


class SpaceType(str, Enum):
    FullSpace = "full_space"
    SubSpace = "sub_space"


def _within(site_1, site_2, radius: Decimal) -> bool:
    # space.py:54 evaluates |r1 - r2| <= R_b on Decimal coordinates (exact decimal arithmetic)
    dx = site_1[0] - site_2[0]
    dy = site_1[1] - site_2[1]
    return (dx * dx + dy * dy).sqrt() <= radius


@dataclass(frozen=True)
class Space:
    space_type: SpaceType
    atom_type: AtomType
    program_register: Register
    configurations: np.ndarray

    def __eq__(self, other: Any) -> bool:
        return isinstance(other, Space) and self.program_register == other.program_register

    def __hash__(self):
        return hash(self.program_register)

    # ---- construction (space.py:39-103) ------------------------------------------------
    @classmethod
    def create(cls, register: Register) -> "Space":
        sites = register.sites
        n_atom = len(sites)
        atom_type = register.atom_type
        n_level = atom_type.n_level
        radius = register.blockade_radius
        total = n_level**n_atom

        # configuration integer type: at least uint32 (space.py:59-60)
        config_type = np.result_type(np.min_scalar_type(total - 1), np.uint32)

        blockers = []  # for atom k>0: lower-index atoms within the blockade radius
        for k in range(1, n_atom):
            blockers.append([j for j in range(k) if _within(sites[k], sites[j], radius)])

        if all(len(b) == 0 for b in blockers):
            configurations = np.arange(total, dtype=config_type)
            return cls(SpaceType.FullSpace, atom_type, register, configurations)

        ryd = atom_type.rydberg_digit
        configs = np.arange(n_level, dtype=config_type)
        for k, neigh in enumerate(blockers, 1):
            weight = config_type.type(n_level**k)
            blocked = np.zeros(configs.shape, dtype=bool)
            for j in neigh:
                blocked |= _digit(configs, j, n_level) == ryd
            free = configs[~blocked]
            parts = []
            for d in range(n_level):
                base = free if d == ryd else configs
                parts.append(base + config_type.type(d) * weight)
            configs = np.concatenate(parts)
            if ryd != n_level - 1:  # pragma: no cover - both atom types keep |r> as the top digit
                configs.sort()
        return cls(SpaceType.SubSpace, atom_type, register, configs)

    # ---- simple properties ---------------------------------------------------------------
    @property
    def index_type(self):
        return np.int32 if self.size < np.iinfo(np.int32).max else np.int64

    @property
    def size(self) -> int:
        return int(self.configurations.size)

    @property
    def n_atoms(self) -> int:
        return len(self.program_register)

    @property
    def n_sites(self) -> int:
        return len(self.program_register.geometry.sites)

    @property
    def state_type(self):
        return self.configurations.dtype

    # ---- digit predicates (atom_type.py:73-78,136-144) -------------------------------------
    def is_state_at(self, index: int, state: int) -> np.ndarray:
        return _digit(self.configurations, index, self.atom_type.n_level) == int(state)

    def is_rydberg_at(self, index: int) -> np.ndarray:
        return self.is_state_at(index, self.atom_type.rydberg_digit)

    # ---- index maps (space.py:134-178, atom_type.py:80-122,146-161) -------------------------
    def _lookup(self, target_configs: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """index of each target configuration in this space, and a validity mask"""
        if self.space_type is SpaceType.FullSpace:
            return target_configs.astype(np.int64), np.ones(target_configs.shape, dtype=bool)
        pos = np.searchsorted(self.configurations, target_configs)
        ok = pos < self.size
        ok[ok] = self.configurations[pos[ok]] == target_configs[ok]
        return pos, ok

    def transition_state_at(self, index: int, fro: int, to: int):
        """(row_indices, col_indices): rows = basis states whose atom `index` is in `fro`;
        cols = the same configuration with that atom moved to `to` (when it is in the space).
        In the full space the second element is the target CONFIGURATION array, which equals
        the index there - same convention as the reference (space.py:165-166)."""
        n_level = self.atom_type.n_level
        rows = np.flatnonzero(self.is_state_at(index, fro))
        cfg = self.configurations[rows]
        weight = cfg.dtype.type(n_level**index)
        if to >= fro:
            tgt = cfg + cfg.dtype.type(to - fro) * weight
        else:
            tgt = cfg - cfg.dtype.type(fro - to) * weight
        if self.space_type is SpaceType.FullSpace:
            return rows, tgt
        pos, ok = self._lookup(tgt)
        return rows[ok], pos[ok]

    def swap_state_at(self, index: int, state_1: int, state_2: int):
        """(row_indices, col_indices) of the symmetric exchange state_1 <-> state_2 on atom
        `index`.  Two-level full space returns (slice(None), configurations ^ bit) like the
        reference (atom_type.py:146-151, space.py:142-143)."""
        n_level = self.atom_type.n_level
        d = _digit(self.configurations, index, n_level)
        weight = n_level**index
        lo, hi = (state_1, state_2) if state_1 <= state_2 else (state_2, state_1)
        delta = self.configurations.dtype.type((hi - lo) * weight)
        tgt = self.configurations.copy()
        m_lo = d == lo
        m_hi = d == hi
        tgt[m_lo] += delta
        tgt[m_hi] -= delta
        involved = m_lo | m_hi
        if n_level == 2:
            rows: Any = slice(None)
            tgt_sel = tgt
            row_idx = None
        else:
            row_idx = np.flatnonzero(involved)
            rows = row_idx
            tgt_sel = tgt[involved]
        if self.space_type is SpaceType.FullSpace:
            return rows, tgt_sel
        pos, ok = self._lookup(tgt_sel)
        if np.all(ok):
            return rows, pos
        if row_idx is None:
            row_idx = np.arange(self.size)
        return row_idx[ok], pos[ok]

    def partner_table(self) -> np.ndarray:
        """Gather-form partner indices for the CUDA kernels.

        Returns int32[n_atoms, n_level-1, size]: entry [a, s, i] is the index of the basis state
        obtained from state i by exchanging digit s <-> s+1 on atom a (s=0: g<->r for two-level
        atoms; s=0: g<->h, s=1: h<->r for three-level atoms), or -1 when atom a is in neither
        of the two levels or the partner lies outside the blockade subspace.  Because an
        exchange is an involution, the scatter loops of the reference (`IndexMapping`,
        sparse_operator.py:115-128; CSC, :15-21) become pure gathers."""
        n_level = self.atom_type.n_level
        if self.size >= 2**31:
            raise ValueError("partner tables are int32; dim too large")
        table = np.full((self.n_atoms, n_level - 1, self.size), -1, dtype=np.int32)
        for a in range(self.n_atoms):
            for s in range(n_level - 1):
                rows, cols = self.swap_state_at(a, s, s + 1)
                if isinstance(rows, slice):
                    rows = np.arange(self.size)
                table[a, s, rows] = cols
        return table

    # ---- fock states (space.py:180-209) -----------------------------------------------------
    def string_to_integer(self, fock_state: str) -> int:
        s = fock_state.replace("|", "").replace(">", "")
        value, shift = 0, 1
        for ch in s:
            value += shift * self.atom_type.str_to_int[ch]
            shift *= self.atom_type.n_level
        return value

    def integer_to_string(self, state_int: int) -> str:
        n_level = self.atom_type.n_level
        state_int = int(state_int)
        chars = []
        for _ in range(self.n_atoms):
            chars.append(self.atom_type.int_to_str[state_int % n_level])
            state_int //= n_level
        return "|" + "".join(chars) + ">"

    def fock_state_to_index(self, fock_state: str) -> int:
        state_int = self.string_to_integer(fock_state)
        if self.space_type is SpaceType.FullSpace:
            return state_int
        index = int(np.searchsorted(self.configurations, state_int))
        if index >= self.size or state_int != int(self.configurations[index]):
            raise ValueError(f"state: {fock_state} not in rydberg blockade subspace.")
        return index

    def index_to_fock_state(self, index: int) -> str:
        if index < 0 or index >= self.size:
            raise ValueError(f"index: {index} out of bounds.")
        return self.integer_to_string(self.configurations[index])

    def zero_state(self, dtype=np.float64):
        from .state_vector import StateVector

        state = np.zeros(self.size, dtype=dtype)
        state[0] = 1.0
        return StateVector(state, self)

    # ---- Born sampling (space.py:211-229) -> CUDA kernel (4) --------------------------------
    def sample_state_vector(self, state_vector: np.ndarray, n_samples: int, project_hyperfine: bool = True):
        """Born-rule sampling on the GPU: |psi|^2 -> inclusive prefix sum (CDF) -> binary search
        (side="right") -> base-n_level digit decode, atom 0 = least significant.

        The uniforms are drawn on the host with `np.random.random_sample(n_samples)`, exactly the
        draw `np.random.choice(p=...)` makes inside the reference (space.py:217), so a run seeded
        with `np.random.seed` reproduces the reference's shots."""
        from . import _lib

        uniforms = np.random.random_sample(int(n_samples))
        project = bool(project_hyperfine) and isinstance(self.atom_type, ThreeLevelAtomType)
        return _lib.sample_host(
            np.ascontiguousarray(state_vector, dtype=np.complex128),
            None if self.space_type is SpaceType.FullSpace and self.atom_type.n_level == 2 else self.configurations,
            self.n_atoms,
            self.atom_type.n_level,
            uniforms,
            project,
        )

    def __str__(self):
        lines = []
        n_digits = len(str(self.size - 1))
        for index in range(min(self.size, 30)):
            lines.append(f"{index: >{n_digits}d}. {self.index_to_fock_state(index)}")
        if self.size > 30:
            lines.append("...")
        return "\n".join(lines) + "\n"


def _digit(configs: np.ndarray, index: int, n_level: int) -> np.ndarray:
    if n_level == 2:
        return (configs >> configs.dtype.type(index)) & configs.dtype.type(1)
    return (configs // configs.dtype.type(n_level**index)) % configs.dtype.type(n_level)
