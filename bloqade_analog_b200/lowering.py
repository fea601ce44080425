"""Lowering of an `EmulatorProgram` to the flat descriptors the CUDA library consumes.

This is the device-plan counterpart of the reference's operator generation
(`emulate/codegen/hamiltonian.py:63-203`, `RydbergHamiltonianCodeGen`):

* `rydberg_diagonal` restates `visit_register` (HC:63-91): V[i] = sum_{a<b} C6/|r_a-r_b|^6 over the
  pairs that are both Rydberg in basis state i, same pair order and same `<= eps` cut, so the
  float64 values are bit-identical to the reference's.
* `pack_task_record` replaces `visit_detuning_term` / `visit_rabi_term` (HC:188-203).  Instead of
  materialising one f64[dim] diagonal per detuning term (HC:98-118) and one IndexMapping / CSR per
  Rabi term (HC:120-186), it keeps what generated them - `target_atoms` scales and the waveform -
  because on the device every such term collapses onto per-atom coefficients
  (csrc/bqa_core.cuh: Coefs).  The record layout is documented in include/bqa_b200.h.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .ir import (
    EmulatorProgram,
    LevelCoupling,
    RabiOperatorType,
    Register,
    ThreeLevelAtomType,
    Waveform,
    flatten_waveform,
)
from .space import Space
from .vecnum import column, tofloat as _flt

RB_C6 = 2 * np.pi * 862690  # rad um^6 / us   (reference constants.py:3)

TASK_HEADER = 8
PIECE_STRIDE = 6
MAX_ATOMS = 32


def rydberg_diagonal(space: Space, register: Optional[Register] = None) -> np.ndarray:
    """van-der-Waals diagonal, value-identical to HC:73-89."""
    register = space.program_register if register is None else register
    sites = [np.asarray([float(x), float(y)]) for x, y in register.sites]
    rydberg = np.zeros(space.size, dtype=np.float64)
    eps = np.finfo(np.float64).eps
    is_ryd = [space.is_rydberg_at(a) for a in range(len(sites))] if len(sites) <= 16 else None
    for a in range(len(sites)):
        ra = is_ryd[a] if is_ryd is not None else space.is_rydberg_at(a)
        for b in range(a + 1, len(sites)):
            distance = np.linalg.norm(sites[a] - sites[b])
            interaction = RB_C6 / (distance**6)
            if interaction <= eps:
                continue
            rb = is_ryd[b] if is_ryd is not None else space.is_rydberg_at(b)
            rydberg[np.logical_and(ra, rb)] += interaction
    return rydberg


def pair_interactions(register: Register) -> np.ndarray:
    """C6/|r_a-r_b|^6 for a<b as an upper-triangular [n,n] f64 matrix (0 where cut by the eps test);
    input of the on-device diagonal builder `bqa_build_rydberg` (same pair order as HC:76-89)."""
    sites = [np.asarray([float(x), float(y)]) for x, y in register.sites]
    n = len(sites)
    out = np.zeros((n, n), dtype=np.float64)
    eps = np.finfo(np.float64).eps
    for a in range(n):
        for b in range(a + 1, n):
            distance = np.linalg.norm(sites[a] - sites[b])
            interaction = RB_C6 / (distance**6)
            if interaction > eps:
                out[a, b] = interaction
    return out


def coupling_slot(n_level: int, coupling: LevelCoupling) -> int:
    """exchange slot s (digits s <-> s+1; detuning acts on digit s+1).
    two-level: g<->r is slot 0 (HC:104-105,126-128); three-level: hyperfine g<->h slot 0,
    rydberg h<->r slot 1 (HC:106-110,130-136)."""
    if n_level == 2:
        return 0
    return 1 if LevelCoupling(coupling) is LevelCoupling.Rydberg else 0


class _WaveformTable:
    def __init__(self):
        self.wf_start: List[int] = [0]
        self.pieces: List[List[float]] = []
        self.coefs: List[float] = []
        self._keys: List[Tuple[tuple, int]] = []

    def add(self, wf: Waveform, dedup: bool = True) -> int:
        flat = flatten_waveform(wf)
        key = tuple(flat)  # waveforms with identical piece tables share one entry
        if dedup:
            # compared entry by entry (not hashed): with the per-task vector numbers of a whole sweep (vecnum.py)
            # tables that coincide for SOME tasks only must be noticed - the comparison raises NonUniformSweep
            for other, wid in self._keys:
                if len(other) == len(key) and other == key:
                    return wid
        for p in flat:
            self.pieces.append([p.lo, p.hi, p.offset, p.scale, float(len(self.coefs)), float(len(p.coeffs))])
            self.coefs.extend(p.coeffs)
        self.wf_start.append(len(self.pieces))
        wid = len(self.wf_start) - 2
        if dedup:
            self._keys.append((key, wid))
        return wid


def pack_task_record(program: EmulatorProgram, diag_id=0, extra_waveforms: Sequence[Waveform] = (),
                     n_tasks: Optional[int] = None) -> np.ndarray:
    """Flatten the time-dependent part of one task into the float64 task record.
    `extra_waveforms` are appended to the waveform table (after the program's own, in order, without
    de-duplication) for generic CSR terms registered with `bqa_plan_add_csr`.

    With `n_tasks`, `program` is a whole sweep whose numbers are per-task vectors (vecnum.py, built by
    sweep.py) and `diag_id` may be an int array [n_tasks]: the result is records[n_tasks, L], row k
    bit-identical to the record of task k lowered on its own."""
    reg = program.register
    n = len(reg)
    if n > MAX_ATOMS:
        raise ValueError(f"at most {MAX_ATOMS} atoms are supported (uint32 configurations)")
    n_level = reg.atom_type.n_level
    table = _WaveformTable()
    det_rows: List[List[float]] = []
    rabi_rows: List[List[float]] = []
    for coupling, fields in program.pulses.items():
        slot = coupling_slot(n_level, coupling)
        for term in fields.detuning:
            w = table.add(term.amplitude)
            scale = [0.0] * n
            for atom, value in term.operator_data.target_atoms.items():
                scale[atom] = scale[atom] + _flt(value)
            det_rows.append([float(w), float(slot)] + scale)
        for term in fields.rabi:
            w = table.add(term.amplitude)
            pw = -1
            if term.phase is not None:
                pw = table.add(term.phase)
            elif term.operator_data.operator_type is RabiOperatorType.RabiAsymmetric:
                raise ValueError("RabiAsymmetric term without a phase waveform")
            scale = [0.0] * n
            for atom, value in term.operator_data.target_atoms.items():
                scale[atom] = scale[atom] + _flt(value)
            rabi_rows.append([float(w), float(pw), float(slot)] + scale)
    for wf in extra_waveforms:
        table.add(wf, dedup=False)
    header = [
        program.duration if n_tasks is not None else float(program.duration),
        diag_id if n_tasks is not None else float(diag_id),
        float(len(table.wf_start) - 1),
        float(len(table.pieces)),
        float(len(det_rows)),
        float(len(rabi_rows)),
        float(len(table.coefs)),
        float(n),
    ]
    flat: List[float] = header + [float(x) for x in table.wf_start]
    for p in table.pieces:
        flat.extend(p)
    flat.extend(table.coefs)
    for r in det_rows:
        flat.extend(r)
    for r in rabi_rows:
        flat.extend(r)
    if n_tasks is None:
        return np.asarray(flat, dtype=np.float64)
    out = np.empty((n_tasks, len(flat)), dtype=np.float64)
    for j, x in enumerate(flat):
        out[:, j] = x if isinstance(x, np.ndarray) else column(x, n_tasks)
    return out


def concat_records(records: Sequence[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """(blob, offsets[n+1]) for a batch of task records."""
    offsets = np.zeros(len(records) + 1, dtype=np.int64)
    np.cumsum([r.size for r in records], out=offsets[1:])
    blob = np.concatenate(records) if records else np.zeros(0, dtype=np.float64)
    return np.ascontiguousarray(blob), offsets
