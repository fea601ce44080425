"""Numbers that stand for ONE VALUE PER TASK of a parameter sweep.

`program.batch_assign(...)` produces tasks that differ only in the values of scalar variables.  The
reference compiles every task on its own (`ir/routine/bloqade.py:184-202`: assign -> canonicalize ->
EmulatorProgramCodeGen, ~8 ms of Python per task); the B200 routine lowers the whole sweep in ONE pass by
running the same lowering code on these vector numbers:

* `DecVec`   - a vector of `decimal.Decimal` (the reference evaluates every scalar expression in Decimal,
               `ir/scalar.py:284-580`; keeping Decimal keeps the float64 results bit-identical),
* `FloatVec` - a vector of float64 (what `float(Decimal)` gives, per task).

Arithmetic is elementwise.  Every COMPARISON must have the same outcome for all tasks - then it returns a
plain `bool` and the lowering takes the same branch for the whole sweep; otherwise it raises
`NonUniformSweep` and the caller falls back to the task-by-task path.  Hashes use the first element, which
is consistent with that notion of equality.
"""
from __future__ import annotations

import operator
from decimal import Decimal
from typing import Sequence, Union

import numpy as np


class NonUniformSweep(Exception):
    """the tasks of a sweep would take different branches of the lowering (or use an unsupported node)"""


def _uniform(mask) -> bool:
    mask = np.asarray(mask, dtype=bool)
    if mask.all():
        return True
    if not mask.any():
        return False
    raise NonUniformSweep("a comparison has different outcomes for different tasks of the sweep")


class _Vec:
    __slots__ = ("v",)

    def __init__(self, v: np.ndarray):
        self.v = v

    def __len__(self):
        return self.v.shape[0]

    def _other(self, other):
        return other.v if isinstance(other, _Vec) else other

    def _cmp(self, other, op) -> bool:
        return _uniform(op(self.v, self._other(other)))

    def __lt__(self, other):
        return self._cmp(other, operator.lt)

    def __le__(self, other):
        return self._cmp(other, operator.le)

    def __gt__(self, other):
        return self._cmp(other, operator.gt)

    def __ge__(self, other):
        return self._cmp(other, operator.ge)

    def __eq__(self, other):
        if other is None:
            return False
        if isinstance(other, _Vec) and other.v is self.v:
            return True
        return self._cmp(other, operator.eq)

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash(self.v[0])

    def __bool__(self):
        return not self.__eq__(0)

    def __repr__(self):
        return f"{type(self).__name__}({self.v[0]!r}, ... x{len(self)})"

    def _new(self, v):
        return type(self)(v)

    def __add__(self, other):
        return self._new(self.v + self._other(other))

    def __radd__(self, other):
        return self._new(other + self.v)

    def __sub__(self, other):
        return self._new(self.v - self._other(other))

    def __rsub__(self, other):
        return self._new(other - self.v)

    def __mul__(self, other):
        return self._new(self.v * self._other(other))

    def __rmul__(self, other):
        return self._new(other * self.v)

    def __truediv__(self, other):
        return self._new(self.v / self._other(other))

    def __rtruediv__(self, other):
        return self._new(other / self.v)

    def __neg__(self):
        return self._new(-self.v)


class DecVec(_Vec):
    """object array of Decimals, one per task"""

    __slots__ = ()

    def is_zero(self) -> bool:
        return _uniform([x.is_zero() for x in self.v])

    def tofloat(self) -> "FloatVec":
        return FloatVec(self.v.astype(np.float64))


class FloatVec(_Vec):
    """float64 array, one per task"""

    __slots__ = ()

    def tofloat(self) -> "FloatVec":
        return self


Dec = Union[Decimal, DecVec]


def dec_vector(values: Sequence) -> Dec:
    """`Decimal(str(x))` of every value (`ir/scalar.py:311-315`); a plain Decimal when all tasks agree"""
    decs = [v if isinstance(v, Decimal) else Decimal(str(v)) for v in values]
    first = decs[0]
    if all(d == first for d in decs):
        return first
    out = np.empty(len(decs), dtype=object)
    out[:] = decs
    return DecVec(out)


def tofloat(x):
    """float64 of a Decimal / DecVec (correctly rounded, per task)"""
    return x.tofloat() if isinstance(x, _Vec) else float(x)


def vmax(a, b):
    """elementwise max of float-likes (no uniformity requirement: it is a value, not a branch)"""
    if isinstance(a, _Vec) or isinstance(b, _Vec):
        av = a.v if isinstance(a, _Vec) else a
        bv = b.v if isinstance(b, _Vec) else b
        return FloatVec(np.maximum(av, bv).astype(np.float64))
    return max(a, b)


def task_value(x, k: int):
    """the value of task k"""
    return x.v[k] if isinstance(x, _Vec) else x


def column(x, n: int) -> np.ndarray:
    """float64[n] column of a float-like"""
    if isinstance(x, _Vec):
        return np.asarray(x.v, dtype=np.float64)
    return np.full(n, float(x), dtype=np.float64)
