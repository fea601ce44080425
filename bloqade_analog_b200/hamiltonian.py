"""Operator generation for the B200 backend: EmulatorProgram -> RydbergHamiltonian on the GPU.

Mirror of `RydbergHamiltonianCodeGen` / `CompileCache`
(`src/bloqade/analog/emulate/codegen/hamiltonian.py:34-214`).  The reference materialises one
f64[dim] diagonal per detuning term and one IndexMapping / CSR per Rabi term; here `emit` builds

* the device plan of the register's `Space` (partner tables in gather form),
* the van-der-Waals diagonal - on the host for small spaces (bit-exact restatement of HC:73-89), on
  the device (`bqa_plan_build_rydberg`, same pair order) from `device_diagonal_min_dim` upwards,
* the flat task record (waveform pieces + per-atom scales) that the kernels evaluate on-device.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

from . import _lib
from .ir import EmulatorProgram, LevelCoupling, Register
from .lowering import coupling_slot, pack_task_record, pair_interactions, rydberg_diagonal
from .space import Space
from .state_vector import DetuningOperator, RabiOperator, RydbergHamiltonian, make_plan


@dataclass
class CompileCache:
    """Reuse of spaces / device plans / diagonals across the tasks of a sweep (HC:34-41)."""

    space_cache: Dict[Register, Tuple[Space, "_lib.Plan"]] = field(default_factory=dict)
    diagonal_cache: Dict[Tuple, int] = field(default_factory=dict)
    # host-side Spaces, shareable between the caches of several devices (a Space has no device state)
    shared_spaces: Dict[Register, Space] = field(default_factory=dict)


class RydbergHamiltonianCodeGen:
    def __init__(self, compile_cache: Optional[CompileCache] = None, device: Optional[int] = None,
                 device_diagonal_min_dim: int = 1 << 20):
        self.compile_cache = CompileCache() if compile_cache is None else compile_cache
        self.device = device
        self.device_diagonal_min_dim = device_diagonal_min_dim

    def space_only(self, register: Register) -> Space:
        """the Space of `register`, built once per CompileCache.shared_spaces (no device work)"""
        spaces = self.compile_cache.shared_spaces
        if register not in spaces:
            spaces[register] = Space.create(register)
        return spaces[register]

    def _space_and_plan(self, register: Register):
        cache = self.compile_cache.space_cache
        if register in cache:
            return cache[register]
        space = self.space_only(register)
        plan = make_plan(space, self.device)
        cache[register] = (space, plan)
        return space, plan

    def _diagonal(self, register: Register, space: Space, plan):
        key = (id(plan), tuple(register.sites))
        cache = self.compile_cache.diagonal_cache
        if key in cache:
            return cache[key], None
        host = None
        if space.size >= self.device_diagonal_min_dim:
            diag_id = plan.build_rydberg(pair_interactions(register))
        else:
            host = rydberg_diagonal(space, register)
            diag_id = plan.add_diagonal(host)
        cache[key] = diag_id
        return diag_id, host

    def emit(self, emulator_program: EmulatorProgram) -> RydbergHamiltonian:
        register = emulator_program.register
        space, plan = self._space_and_plan(register)
        diag_id, host_diag = self._diagonal(register, space, plan)
        n_level = register.atom_type.n_level
        det_ops, rabi_ops = [], []
        for coupling, fields in emulator_program.pulses.items():
            slot = coupling_slot(n_level, coupling)
            for term in fields.detuning:
                det_ops.append(DetuningOperator(dict(term.operator_data.target_atoms), slot, space, term.amplitude))
            for term in fields.rabi:
                rabi_ops.append(
                    RabiOperator(dict(term.operator_data.target_atoms), slot, term.operator_data.operator_type, space,
                                 term.amplitude, term.phase)
                )
        record = pack_task_record(emulator_program, diag_id)
        return RydbergHamiltonian(
            emulator_ir=emulator_program,
            space=space,
            detuning_ops=det_ops,
            rabi_ops=rabi_ops,
            plan=plan,
            diag_id=diag_id,
            record=record,
            _rydberg=host_diag,
        )
