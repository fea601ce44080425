"""Emulator IR of the hot path: the input contract of the B200 backend.

Host-side mirror of the reference's emulator IR, with the same class names, field names and
meaning, so programs (and parity tests) read like the reference's own:

* `Register / Fields / DetuningTerm / RabiTerm / DetuningOperatorData / RabiOperatorData /
  RabiOperatorType / LevelCoupling / EmulatorProgram`
  mirror `src/bloqade/analog/emulate/ir/emulator.py:99-246`.
* `TwoLevelAtomType / ThreeLevelAtomType` mirror `emulate/ir/atom_type.py:61-165`
  (digit conventions: atom 0 is the least-significant base-`n_level` digit).
* the waveform nodes `Linear / Constant / Poly / Append / Slice / Add / Scale / Negative /
  PythonFn` mirror the evaluation semantics of `ir/control/waveform.py:255-800` after
  assignment and canonicalisation (the only form the emulator ever sees, emulator.py:32-63).

What differs by design: waveforms here hold plain numbers (no symbolic scalars - the reference has
already assigned them when it hands an `EmulatorProgram` to the hot path), and every waveform can be
*flattened* into a table of polynomial pieces (`flatten_waveform`) that the CUDA kernels evaluate
on-device (north_star: "per-term waveform coefficients evaluated on-device").

`bloqade_analog_b200.adapter` converts a real `bloqade.analog` `EmulatorProgram` into this IR.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from decimal import Decimal
from enum import Enum
from typing import Callable, Dict, List, Optional, Sequence, Tuple, Union

from .vecnum import DecVec, NonUniformSweep, tofloat as _flt

Number = Union[int, float, Decimal, DecVec]


def _dec(x: Number) -> Decimal:
    """Same literal handling as the reference's scalar `cast` (Decimal(str(x))).  A `DecVec` (one Decimal
    per task of a sweep, vecnum.py) passes through: the lowering below is written so that it runs
    unchanged on such vector numbers."""
    if isinstance(x, (Decimal, DecVec)):
        return x
    return Decimal(str(x))


# --------------------------------------------------------------------------------------
# atom types (atom_type.py:61-165)
# --------------------------------------------------------------------------------------
@dataclass(frozen=True)
class AtomType:
    n_level: int = 0
    rydberg_digit: int = 0

    def __hash__(self):
        return hash(self.__class__)


@dataclass(frozen=True)
class TwoLevelAtomType(AtomType):
    """digits: 0 = ground |g>, 1 = Rydberg |r>  (atom_type.py:127-135)"""

    n_level: int = 2
    rydberg_digit: int = 1
    int_to_str = ["g", "r"]
    str_to_int = {"g": 0, "r": 1}

    __hash__ = AtomType.__hash__


@dataclass(frozen=True)
class ThreeLevelAtomType(AtomType):
    """digits: 0 = ground |g>, 1 = hyperfine |h>, 2 = Rydberg |r>  (atom_type.py:63-71)"""

    n_level: int = 3
    rydberg_digit: int = 2
    int_to_str = ["g", "h", "r"]
    str_to_int = {"g": 0, "h": 1, "r": 2}

    __hash__ = AtomType.__hash__


TwoLevelAtom = TwoLevelAtomType()
ThreeLevelAtom = ThreeLevelAtomType()


class LevelCoupling(str, Enum):
    Rydberg = "rydberg"
    Hyperfine = "hyperfine"


class RabiOperatorType(int, Enum):
    RabiAsymmetric = 0
    RabiSymmetric = 1


# --------------------------------------------------------------------------------------
# waveforms (ir/control/waveform.py) - numbers already assigned
# --------------------------------------------------------------------------------------
class Waveform:
    """Base class.  `duration` is a Decimal; `__call__` follows waveform.py:84-85."""

    duration: Decimal

    def __call__(self, clock_s: float) -> float:
        return float(self.eval_decimal(Decimal(str(clock_s))))

    def eval_decimal(self, clock_s: Decimal) -> Decimal:  # pragma: no cover - abstract
        raise NotImplementedError


@dataclass(frozen=True)
class Linear(Waveform):
    """f(t) = start + (stop-start)/duration * t for t <= duration else 0 (waveform.py:282-296)"""

    start: Decimal
    stop: Decimal
    duration: Decimal

    def __init__(self, start: Number, stop: Number, duration: Number):
        object.__setattr__(self, "start", _dec(start))
        object.__setattr__(self, "stop", _dec(stop))
        object.__setattr__(self, "duration", _dec(duration))

    def eval_decimal(self, clock_s):
        if clock_s > self.duration:
            return Decimal(0)
        if self.duration.is_zero():
            raise ValueError("Duration of linear waveform is zero. Cannot divide by zero.")
        return ((self.stop - self.start) / self.duration) * clock_s + self.start


@dataclass(frozen=True)
class Constant(Waveform):
    """f(t) = value for t <= duration else 0 (waveform.py:330-335)"""

    value: Decimal
    duration: Decimal

    def __init__(self, value: Number, duration: Number):
        object.__setattr__(self, "value", _dec(value))
        object.__setattr__(self, "duration", _dec(duration))

    def eval_decimal(self, clock_s):
        if clock_s > self.duration:
            return Decimal(0)
        return self.value


@dataclass(frozen=True)
class Poly(Waveform):
    """f(t) = c[0] + c[1] t + ... for t <= duration else 0 (waveform.py:369-382)"""

    coeffs: Tuple[Decimal, ...]
    duration: Decimal

    def __init__(self, coeffs: Sequence[Number], duration: Number):
        object.__setattr__(self, "coeffs", tuple(_dec(c) for c in coeffs))
        object.__setattr__(self, "duration", _dec(duration))

    def eval_decimal(self, clock_s):
        if clock_s > self.duration:
            return Decimal(0)
        value = Decimal(0)
        power = Decimal(1)
        for c in self.coeffs:
            value += c * power
            power *= clock_s
        return value


@dataclass(frozen=True)
class PythonFn(Waveform):
    """Arbitrary host callable f(t) (waveform.py:452-468).

    It cannot run on the device.  `flatten_waveform` lowers it at plan time to a piecewise-cubic
    Hermite table with `n_samples` uniform knots; the interpolation error
    is O(h^4 f'''') and is reported by `sampled_error_bound`."""

    fn: Callable[[float], float]
    duration: Decimal
    n_samples: int = 2048

    def __init__(self, fn, duration: Number, n_samples: int = 2048):
        object.__setattr__(self, "fn", fn)
        object.__setattr__(self, "duration", _dec(duration))
        object.__setattr__(self, "n_samples", int(n_samples))

    def eval_decimal(self, clock_s):
        if clock_s > self.duration:
            return Decimal(0)
        return Decimal(str(self.fn(float(clock_s))))


@dataclass(frozen=True)
class Append(Waveform):
    """Concatenation; the FIRST segment with t <= append_time + duration wins, evaluated in its
    own local time (waveform.py:698-708)."""

    waveforms: Tuple[Waveform, ...]

    def __init__(self, waveforms: Sequence[Waveform]):
        object.__setattr__(self, "waveforms", tuple(waveforms))

    @property
    def duration(self) -> Decimal:
        total = Decimal(0)
        for w in self.waveforms:
            total += w.duration
        return total

    def eval_decimal(self, clock_s):
        append_time = Decimal(0)
        for w in self.waveforms:
            d = w.duration
            if clock_s <= append_time + d:
                return w.eval_decimal(clock_s - append_time)
            append_time += d
        return Decimal(0)


@dataclass(frozen=True)
class Slice(Waveform):
    """waveform[start:stop]: shift by `start`, zero beyond the slice duration (waveform.py:668-673)"""

    waveform: Waveform
    start: Decimal
    stop: Decimal

    def __init__(self, waveform: Waveform, start: Optional[Number] = None, stop: Optional[Number] = None):
        object.__setattr__(self, "waveform", waveform)
        object.__setattr__(self, "start", _dec(0 if start is None else start))
        object.__setattr__(self, "stop", waveform.duration if stop is None else _dec(stop))

    @property
    def duration(self) -> Decimal:
        return self.stop - self.start

    def eval_decimal(self, clock_s):
        if clock_s > self.duration:
            return Decimal(0)
        return self.waveform.eval_decimal(clock_s + self.start)


@dataclass(frozen=True)
class Negative(Waveform):
    waveform: Waveform

    @property
    def duration(self) -> Decimal:
        return self.waveform.duration

    def eval_decimal(self, clock_s):
        return -self.waveform.eval_decimal(clock_s)


@dataclass(frozen=True)
class Scale(Waveform):
    scalar: Decimal
    waveform: Waveform

    def __init__(self, scalar: Number, waveform: Waveform):
        object.__setattr__(self, "scalar", _dec(scalar))
        object.__setattr__(self, "waveform", waveform)

    @property
    def duration(self) -> Decimal:
        return self.waveform.duration

    def eval_decimal(self, clock_s):
        return self.scalar * self.waveform.eval_decimal(clock_s)


@dataclass(frozen=True)
class Add(Waveform):
    """left + right; the shorter operand contributes 0 after its own duration.
    Note the reference goes through `__call__` (float round trip) for each operand
    (waveform.py:791-792)."""

    left: Waveform
    right: Waveform

    @property
    def duration(self) -> Decimal:
        return max(self.left.duration, self.right.duration)

    def eval_decimal(self, clock_s):
        return Decimal(str(self.left(clock_s))) + Decimal(str(self.right(clock_s)))

    def __call__(self, clock_s):
        return self.left(clock_s) + self.right(clock_s)


def piecewise_linear(durations: Sequence[Number], values: Sequence[Number]) -> Waveform:
    """Builder helper with the semantics of `.piecewise_linear(durations, values)`
    (builder/waveform.py; produces Append of Linear, cf. README.md:88-108)."""
    if len(values) != len(durations) + 1:
        raise ValueError("piecewise_linear needs len(values) == len(durations) + 1")
    segs = [Linear(values[i], values[i + 1], durations[i]) for i in range(len(durations))]
    return segs[0] if len(segs) == 1 else Append(segs)


def piecewise_constant(durations: Sequence[Number], values: Sequence[Number]) -> Waveform:
    if len(values) != len(durations):
        raise ValueError("piecewise_constant needs len(values) == len(durations)")
    segs = [Constant(values[i], durations[i]) for i in range(len(durations))]
    return segs[0] if len(segs) == 1 else Append(segs)


# ---- flattening to device pieces ------------------------------------------------------
@dataclass(frozen=True)
class Piece:
    """value(t) = scale * sum_k coeffs[k] * (t - offset)^k   if  lo < t <= hi  (lo = -inf: no lower
    bound), else 0.  All numbers float64."""

    lo: float
    hi: float
    offset: float
    scale: float
    coeffs: Tuple[float, ...]


def _hermite_pieces(fn, duration: float, n: int):
    """Uniform-knot cubic Hermite table of fn on [0, duration] (derivatives by centred differences
    of a 4x finer sampling).  Returns list of (t0, t1, c0..c3) in local time u = t - t0."""
    import numpy as np

    fine = 4
    m = n * fine
    ts = np.linspace(0.0, duration, m + 1)
    vs = np.array([float(fn(float(t))) for t in ts])
    h = duration / m
    dv = np.empty_like(vs)
    dv[2:-2] = (-vs[4:] + 8 * vs[3:-1] - 8 * vs[1:-3] + vs[:-4]) / (12 * h)
    dv[0] = (-25 * vs[0] + 48 * vs[1] - 36 * vs[2] + 16 * vs[3] - 3 * vs[4]) / (12 * h)
    dv[1] = (-3 * vs[0] - 10 * vs[1] + 18 * vs[2] - 6 * vs[3] + vs[4]) / (12 * h)
    dv[-1] = (25 * vs[-1] - 48 * vs[-2] + 36 * vs[-3] - 16 * vs[-4] + 3 * vs[-5]) / (12 * h)
    dv[-2] = (3 * vs[-1] + 10 * vs[-2] - 18 * vs[-3] + 6 * vs[-4] - vs[-5]) / (12 * h)
    out = []
    H = duration / n
    for k in range(n):
        i0, i1 = k * fine, (k + 1) * fine
        p0, p1, m0, m1 = vs[i0], vs[i1], dv[i0], dv[i1]
        c0 = p0
        c1 = m0
        c2 = (3 * (p1 - p0) / H - 2 * m0 - m1) / H
        c3 = (2 * (p0 - p1) / H + m0 + m1) / (H * H)
        out.append((float(ts[i0]), float(ts[i1]), (float(c0), float(c1), float(c2), float(c3))))
    return out


def flatten_waveform(wf: Waveform) -> List[Piece]:
    """Lower a waveform tree to a flat list of polynomial pieces in ABSOLUTE time.

    The windows reproduce the interpreter's comparisons exactly (waveform.py:286,332,371,668-708):
    every guard is either `t <= b` (upper, inclusive) or `t > a` (the previous Append segment did
    not match), so each leaf is active on (a, b] - or on (-inf, b] for a leaf that is never preceded
    by another Append segment.  Breakpoints are accumulated in Decimal like the reference and only
    then rounded to float64."""
    pieces: List[Piece] = []

    def walk(node: Waveform, shift: Decimal, lo: Optional[Decimal], hi: Optional[Decimal], scale: float):
        # local time of `node` = t_abs - shift ; active window (lo, hi] in absolute time
        def clip_hi(new_hi: Decimal):
            return new_hi if hi is None else min(hi, new_hi)

        if isinstance(node, (Linear, Constant, Poly, PythonFn)):
            h = clip_hi(shift + node.duration)
            flo = -math.inf if lo is None else _flt(lo)
            if lo is not None and h <= lo:
                return
            if isinstance(node, Linear):
                if node.duration.is_zero():
                    raise ValueError("Duration of linear waveform is zero. Cannot divide by zero.")
                coeffs = (_flt(node.start), _flt((node.stop - node.start) / node.duration))
                pieces.append(Piece(flo, _flt(h), _flt(shift), scale, coeffs))
            elif isinstance(node, Constant):
                pieces.append(Piece(flo, _flt(h), _flt(shift), scale, (_flt(node.value),)))
            elif isinstance(node, Poly):
                coeffs = tuple(_flt(c) for c in node.coeffs) or (0.0,)
                pieces.append(Piece(flo, _flt(h), _flt(shift), scale, coeffs))
            else:
                if isinstance(shift, DecVec) or isinstance(h, DecVec) or isinstance(node.duration, DecVec):
                    raise NonUniformSweep("PythonFn waveforms are tabulated per task")
                fshift = float(shift)
                for (t0, t1, c) in _hermite_pieces(node.fn, float(node.duration), node.n_samples):
                    a = fshift + t0
                    b = fshift + t1
                    plo = max(flo, a) if t0 > 0.0 else flo
                    phi = min(float(h), b)
                    if phi > plo or (plo == -math.inf):
                        pieces.append(Piece(plo, phi, a, scale, c))
        elif hasattr(node, "pieces"):  # serialize.PieceTable: an already flattened table
            if isinstance(shift, DecVec) or isinstance(hi, DecVec) or isinstance(lo, DecVec):
                raise NonUniformSweep("piece tables are shifted per task")
            fshift = float(shift)
            flo = -math.inf if lo is None else _flt(lo)
            for p in node.pieces:
                plo = max(flo, p.lo + fshift) if p.lo != -math.inf else flo
                phi = p.hi + fshift if hi is None else min(float(hi), p.hi + fshift)
                if phi > plo:
                    pieces.append(Piece(plo, phi, p.offset + fshift, scale * p.scale, p.coeffs))
        elif isinstance(node, Append):
            append_time = Decimal(0)
            first = True
            for w in node.waveforms:
                d = w.duration
                seg_hi = clip_hi(shift + append_time + d)
                seg_lo = lo if first else (shift + append_time if lo is None else max(lo, shift + append_time))
                walk(w, shift + append_time, seg_lo, seg_hi, scale)
                append_time += d
                first = False
        elif isinstance(node, Slice):
            walk(node.waveform, shift - node.start, lo, clip_hi(shift + node.duration), scale)
        elif isinstance(node, Negative):
            walk(node.waveform, shift, lo, hi, -scale)
        elif isinstance(node, Scale):
            walk(node.waveform, shift, lo, hi, scale * _flt(node.scalar))
        elif isinstance(node, Add):
            walk(node.left, shift, lo, hi, scale)
            walk(node.right, shift, lo, hi, scale)
        else:
            raise TypeError(f"cannot lower waveform node {type(node).__name__} to device pieces")

    walk(wf, Decimal(0), None, None, 1.0)
    return pieces


def eval_pieces(pieces: Sequence[Piece], t: float) -> float:
    """Host evaluation of a piece table - the exact arithmetic the device performs
    (csrc/bqa_core.cuh: wf_eval).  Used by host-side checks only."""
    total = 0.0
    for p in pieces:
        if t > p.lo and t <= p.hi:
            u = t - p.offset
            acc = 0.0
            for c in reversed(p.coeffs):
                acc = acc * u + c
            total += p.scale * acc
    return total


# --------------------------------------------------------------------------------------
# emulator IR (emulator.py:99-246)
# --------------------------------------------------------------------------------------
@dataclass(frozen=True)
class RabiOperatorData:
    operator_type: RabiOperatorType
    target_atoms: Dict[int, Decimal]

    def __init__(self, operator_type: RabiOperatorType, target_atoms: Dict[int, Number]):
        object.__setattr__(self, "operator_type", RabiOperatorType(operator_type))
        object.__setattr__(self, "target_atoms", {int(k): _dec(v) for k, v in target_atoms.items()})

    def __hash__(self):
        return hash(self.operator_type) ^ hash(frozenset(self.target_atoms.items()))


@dataclass(frozen=True)
class DetuningOperatorData:
    target_atoms: Dict[int, Decimal]

    def __init__(self, target_atoms: Dict[int, Number]):
        object.__setattr__(self, "target_atoms", {int(k): _dec(v) for k, v in target_atoms.items()})

    def __hash__(self):
        return hash(frozenset(self.target_atoms.items()))


@dataclass
class RabiTerm:
    operator_data: RabiOperatorData
    amplitude: Waveform
    phase: Optional[Waveform] = None


@dataclass
class DetuningTerm:
    operator_data: DetuningOperatorData
    amplitude: Waveform


@dataclass
class Fields:
    detuning: List[DetuningTerm]
    rabi: List[RabiTerm]


@dataclass(frozen=True)
class Geometry:
    """task/base.py Geometry: all trap sites and their filling (1 = atom present)."""

    sites: List[Tuple[float, float]]
    filling: List[int]


@dataclass(frozen=True)
class Register:
    """emulator.py:157-226.  `sites` are the FILLED sites (atoms), in atom order."""

    atom_type: AtomType
    sites: List[Tuple[Decimal, Decimal]]
    blockade_radius: Decimal
    geometry: Optional[Geometry] = None
    full_index_to_index: Dict[int, int] = field(init=False, default_factory=dict, compare=False)

    def __init__(self, atom_type, sites, blockade_radius: Number = 0, geometry: Optional[Geometry] = None):
        object.__setattr__(self, "atom_type", atom_type)
        object.__setattr__(self, "sites", [(_dec(x), _dec(y)) for x, y in sites])
        object.__setattr__(self, "blockade_radius", _dec(blockade_radius))
        if geometry is None:
            geometry = Geometry([(float(x), float(y)) for x, y in self.sites], len(self.sites) * [1])
        object.__setattr__(self, "geometry", geometry)
        mapping: Dict[int, int] = {}
        for i, filled in enumerate(geometry.filling):
            if filled:
                mapping[i] = len(mapping)
        object.__setattr__(self, "full_index_to_index", mapping)

    def __len__(self):
        return len(self.sites)

    def same(self, other) -> bool:
        return (
            isinstance(other, Register)
            and self.atom_type == other.atom_type
            and self.blockade_radius == other.blockade_radius
            and set(self.sites) == set(other.sites)
        )

    def __eq__(self, other):
        # emulator.py:197-210: with zero blockade radius only the atom count matters
        if not isinstance(other, Register):
            return False
        if self.blockade_radius == Decimal("0") and other.blockade_radius == Decimal("0"):
            return self.atom_type == other.atom_type and len(self.sites) == len(other.sites)
        return self.same(other)

    def __hash__(self):
        if self.blockade_radius == Decimal("0"):
            return hash(self.__class__) ^ hash(self.atom_type) ^ hash(len(self.sites))
        return (
            hash(self.__class__)
            ^ hash(self.atom_type)
            ^ hash(self.blockade_radius)
            ^ hash(frozenset(self.sites))
        )


@dataclass
class EmulatorProgram:
    register: Register
    duration: float
    pulses: Dict[LevelCoupling, Fields]


# --------------------------------------------------------------------------------------
# lattices (ir/location/bravais.py:124-128,159,213): site order itertools.product(range(L1), range(L2))
# --------------------------------------------------------------------------------------
def chain_sites(n: int, lattice_spacing: Number = 1.0) -> List[Tuple[Decimal, Decimal]]:
    a = _dec(lattice_spacing)
    return [(a * i, Decimal(0) * a) for i in range(n)]


def square_sites(l1: int, l2: Optional[int] = None, lattice_spacing: Number = 1.0):
    l2 = l1 if l2 is None else l2
    a = _dec(lattice_spacing)
    return [(a * i, a * j) for i in range(l1) for j in range(l2)]
