"""Synthetic lattice programs of BASELINE.json `configs`, built directly in the
B200 backend IR so that tests and bench.py run where the reference is absent (the GPU box).

Each builder states the reference builder chain that produces the same `EmulatorProgram`
(README.md:88-108 waveforms); `tests/golden/make_golden.py` checks, in the build container, that the
task records packed from these builders equal the ones packed from the reference's own IR.
"""
from __future__ import annotations

from decimal import Decimal
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import ir

README_DURATIONS = [0.4, 3.2, 0.4]
README_OMEGA = [0, 15.8, 15.8, 0]
README_DELTA = [-16.33, -16.33, 16.33, 16.33]


def _uniform(n: int) -> Dict[int, Decimal]:
    return {a: Decimal("1.0") for a in range(n)}


def adiabatic_program(
    sites,
    durations: Sequence = README_DURATIONS,
    omega: Sequence = README_OMEGA,
    delta: Sequence = README_DELTA,
    blockade_radius=0.0,
    local_drive: bool = False,
    phase: Optional[ir.Waveform] = None,
    atom_type: ir.AtomType = ir.TwoLevelAtom,
) -> ir.EmulatorProgram:
    """`lattice.rydberg.rabi.amplitude.uniform.piecewise_linear(durations, omega)
              .detuning.uniform.piecewise_linear(durations, delta)`  (README.md:88-108).
    local_drive=True is the `.amplitude.location(i)` per-atom variant (config 2b)."""
    n = len(sites)
    reg = ir.Register(atom_type, sites, blockade_radius)
    amp = ir.piecewise_linear(durations, omega)
    det = ir.piecewise_linear(durations, delta)
    op_type = ir.RabiOperatorType.RabiSymmetric if phase is None else ir.RabiOperatorType.RabiAsymmetric
    if local_drive:
        # distinct waveform per atom (peak Omega + 0.01*a): the reference merges locations that share a
        # waveform into ONE term, so only distinct waveforms give one IndexMapping term per atom
        rabi = []
        for a in range(n):
            om = [round(float(v) + (0.01 * a if float(v) != 0.0 else 0.0), 6) for v in omega]
            rabi.append(ir.RabiTerm(ir.RabiOperatorData(op_type, {a: Decimal("1")}), ir.piecewise_linear(durations, om), phase))
    else:
        rabi = [ir.RabiTerm(ir.RabiOperatorData(op_type, _uniform(n)), amp, phase)]
    fields = ir.Fields([ir.DetuningTerm(ir.DetuningOperatorData(_uniform(n)), det)], rabi)
    duration = float(sum(ir._dec(d) for d in durations))
    return ir.EmulatorProgram(reg, duration, {ir.LevelCoupling.Rydberg: fields})


def config1(spacings: Sequence[float] = tuple(np.arange(4.0, 7.0, 0.5))) -> List[ir.EmulatorProgram]:
    """C1: `Square(3, lattice_spacing=s)` README sweep, batch_assign(lattice_spacing=arange(4,7,.5)), 6 tasks"""
    return [adiabatic_program(ir.square_sites(3, lattice_spacing=float(s))) for s in spacings]


def config2(variant: str = "a") -> ir.EmulatorProgram:
    """C2: `Chain(16, lattice_spacing=5.72)`, blockade_radius=6.0 (dim 2584); a = uniform drive,
    b = sixteen `.amplitude.location(i)` drives (IndexMapping terms in the reference)"""
    return adiabatic_program(ir.chain_sites(16, 5.72), blockade_radius=6.0, local_drive=(variant == "b"))


def config3_grid(n_detuning: int = 64, n_duration: int = 64):
    """(final_detuning, total_duration) pairs of the C3 sweep, task number = i_det * n_duration + i_dur"""
    dets = np.linspace(4.0, 32.0, n_detuning)
    durs = np.linspace(0.5, 4.0, n_duration)
    return [(float(round(d, 6)), float(round(t, 6))) for d in dets for t in durs]


def config3_task(final_detuning: float, total: float, n_atoms: int = 12) -> ir.EmulatorProgram:
    """one C3 task: `Chain(12, lattice_spacing=5.72)`, Omega ramps 0 -> 15.8 -> 0 over
    [0.1, 0.8, 0.1] * total, detuning -16.33 -> final_detuning.  The durations are the Decimal products the
    reference builder forms for `[0.1 * total, 0.8 * total, 0.1 * total]` with `total = cast("total")`
    (ir/scalar.py:528-529), so the record equals the one `program.bloqade.python()` lowers for the same
    `batch_assign(final_detuning=..., total=...)` point bit for bit (tests/test_sweep_lowering.py)."""
    t = ir._dec(total)
    durations = [Decimal("0.1") * t, Decimal("0.8") * t, Decimal("0.1") * t]
    delta = [-16.33, -16.33, final_detuning, final_detuning]
    return adiabatic_program(ir.chain_sites(n_atoms, 5.72), durations, README_OMEGA, delta)


def config3(n_tasks: int = 4096, n_atoms: int = 12) -> List[ir.EmulatorProgram]:
    side = int(round(np.sqrt(n_tasks)))
    grid = config3_grid(side, side) if side * side == n_tasks else config3_grid(64, 64)[:n_tasks]
    return [config3_task(d, t, n_atoms) for d, t in grid]


def config4(l: int = 5) -> ir.EmulatorProgram:
    """C4: `Square(5, lattice_spacing=5.0)` full 2^25 space, README protocol"""
    return adiabatic_program(ir.square_sites(l, lattice_spacing=5.0))


def config5(l1: int = 6, l2: int = 5) -> ir.EmulatorProgram:
    """C5: `Square(6, 5, lattice_spacing=5.0)` full 2^30 space (sharded)"""
    return adiabatic_program(ir.square_sites(l1, l2, lattice_spacing=5.0))
