"""ORACLE (test infrastructure only - never imported by the product path).

NumPy restatement of the two explicit Runge-Kutta cores that the reference reaches through
`scipy.integrate.ode(...).set_integrator("dop853" | "dopri5")`
(reference call site: src/bloqade/analog/emulate/ir/state_vector.py:589-599).

The algorithm lives in a third-party dependency that is NOT vendored in /root/reference:
SciPy's compiled `scipy.integrate._dop` (uv.lock pins scipy 1.15.3 / 1.17.1; this image has
1.18.1), which is a translation of E. Hairer's DOPRI5 / DOP853 (Hairer, Norsett, Wanner, "Solving
Ordinary Differential Equations I", 2nd ed., 1993).  This file restates that published algorithm
with SciPy's wrapper settings (scipy/integrate/_ode.py: safety 0.9, beta 0.0 - which DOPRI5 reads as
its default 0.04 -, dfactor/ifactor 0.2/10 resp. 0.3/6, max_step = first_step = 0, i.e. automatic
initial step, and a fresh start on every integrate() call).

`reject_mode="hairer"` is the published rule for the step size after a rejected DOP853 step,
hnew = h / min(1/fac1, err^(1/8)/safety); `reject_mode="scipy118"` is hnew = h*fac1, which is what
this image's SciPy 1.18.1 C translation does (found by tracing its RHS call times; with it the
restatement reproduces SciPy's nfcn / nstep / naccpt / nrejct exactly and its states to ~1e-13).
DOPRI5 follows Hairer in both.

Pinning: `tests/test_oracle_golden.py::test_dop_restatement_matches_scipy` checks the restatement against
SciPy itself (identical counters, same final state) on golden programs.
"""
from __future__ import annotations

import numpy as np
from scipy.integrate._ivp import dop853_coefficients as _d8
from scipy.integrate._ivp.rk import RK45 as _RK45

UROUND = 2.3e-16


class DopResult:
    def __init__(self):
        self.status = 0
        self.nfcn = 0
        self.nstep = 0
        self.naccpt = 0
        self.nrejct = 0
        self.h = 0.0
        self.trace = []  # (x, h, err, accepted)


def _hinit(f, x, y, posneg, f0, iord, hmax, atol, rtol):
    sk = atol + rtol * np.abs(y)
    dnf = float(np.sum((f0 / sk) ** 2))
    dny = float(np.sum((y / sk) ** 2))
    if dnf <= 1e-10 or dny <= 1e-10:
        h = 1e-6
    else:
        h = np.sqrt(dny / dnf) * 0.01
    h = min(h, hmax)
    h = np.copysign(h, posneg)
    y1 = y + h * f0
    f1 = f(x + h, y1)
    der2 = float(np.sum(((f1 - f0) / sk) ** 2))
    der2 = np.sqrt(der2) / h
    der12 = max(abs(der2), np.sqrt(dnf))
    if der12 <= 1e-15:
        h1 = max(1e-6, abs(h) * 1e-3)
    else:
        h1 = (0.01 / der12) ** (1.0 / iord)
    h = min(100 * abs(h), h1, hmax)
    return np.copysign(h, posneg)


def integrate(f, x0, y0, xend, solver="dop853", atol=1e-7, rtol=1e-14, nsteps=2**31 - 1, trace=False,
              reject_mode="hairer"):
    """One `ode.integrate(xend)` call.  `f(t, y) -> dy` on real vectors.  Returns (y, DopResult)."""
    res = DopResult()
    y = np.array(y0, dtype=np.float64, copy=True)
    n = y.size
    d5 = solver == "dopri5"
    safe = 0.9
    if d5:
        beta, fac1, fac2, iord = 0.04, 0.2, 10.0, 5
        expo1 = 0.2 - beta * 0.75
    else:
        beta, fac1, fac2, iord = 0.0, 0.3, 6.0, 8
        expo1 = 1.0 / 8.0 - beta * 0.2
    facc1, facc2 = 1.0 / fac1, 1.0 / fac2
    nstiff = 1000
    x = float(x0)
    posneg = 1.0 if xend - x >= 0 else -1.0
    hmax = abs(xend - x)
    facold = 1e-4
    hlamb = 0.0
    iasti = nonsti = 0
    last = reject = False
    k1 = f(x, y)
    h = _hinit(f, x, y, posneg, k1, iord, hmax, atol, rtol)
    res.nfcn += 2
    A = _d8.A
    C = _d8.C
    while True:
        if res.nstep > nsteps:
            res.status = -2
            break
        if 0.1 * abs(h) <= abs(x) * UROUND:
            res.status = -3
            break
        if (x + 1.01 * h - xend) * posneg > 0.0:
            h = xend - x
            last = True
        res.nstep += 1
        if d5:
            a = _RK45.A
            c = _RK45.C
            ks = [k1]
            for s in range(1, 6):
                ys = y + h * sum(a[s, j] * ks[j] for j in range(s) if a[s, j] != 0.0)
                ks.append(f(x + c[s] * h, ys))
            ysti = ys
            ynew = y + h * sum(_RK45.B[j] * ks[j] for j in range(6) if _RK45.B[j] != 0.0)
            k7 = f(x + h, ynew)
            ks.append(k7)
            res.nfcn += 6
            errv = h * sum(_RK45.E[j] * ks[j] for j in range(7) if _RK45.E[j] != 0.0)
            sk = atol + rtol * np.maximum(np.abs(y), np.abs(ynew))
            err = np.sqrt(float(np.sum((errv / sk) ** 2)) / n)
        else:
            ks = [k1]
            for s in range(1, 12):
                ys = y + h * sum(A[s, j] * ks[j] for j in range(s) if A[s, j] != 0.0)
                ts = x + h if s == 11 else x + C[s] * h
                ks.append(f(ts, ys))
            y1_12 = ys
            res.nfcn += 11
            ynew = y + h * sum(_d8.B[j] * ks[j] for j in range(12) if _d8.B[j] != 0.0)
            sk = atol + rtol * np.maximum(np.abs(y), np.abs(ynew))
            e5 = sum(_d8.E5[j] * ks[j] for j in range(12) if _d8.E5[j] != 0.0)
            e3 = sum(_d8.E3[j] * ks[j] for j in range(12) if _d8.E3[j] != 0.0)
            err5 = float(np.sum((e5 / sk) ** 2))
            err3 = float(np.sum((e3 / sk) ** 2))
            deno = err5 + 0.01 * err3
            if deno <= 0.0:
                deno = 1.0
            err = abs(h) * err5 * np.sqrt(1.0 / (n * deno))
        fac11 = err**expo1
        fac = fac11 / facold**beta
        fac = max(facc2, min(facc1, fac / safe))
        hnew = h / fac
        if trace:
            res.trace.append((x, h, err, err <= 1.0))
        if err <= 1.0:
            facold = max(err, 1e-4)
            res.naccpt += 1
            if d5:
                knew = k7
            else:
                knew = f(x + h, ynew)
                res.nfcn += 1
            if res.naccpt % nstiff == 0 or iasti > 0:
                if d5:
                    stnum = float(np.sum((k7 - ks[5]) ** 2))
                    stden = float(np.sum((ynew - ysti) ** 2))
                    if stden > 0:
                        hlamb = h * np.sqrt(stnum / stden)
                    lim = 3.25
                else:
                    stnum = float(np.sum((knew - ks[11]) ** 2))
                    stden = float(np.sum((ynew - y1_12) ** 2))
                    if stden > 0:
                        hlamb = abs(h) * np.sqrt(stnum / stden)
                    lim = 6.1
                if hlamb > lim:
                    nonsti = 0
                    iasti += 1
                    if iasti == 15:
                        res.status = -4
                        break
                else:
                    nonsti += 1
                    if nonsti == 6:
                        iasti = 0
            k1 = knew
            y = ynew
            x = x + h
            if last:
                h = hnew
                res.status = 1
                break
            if abs(hnew) > hmax:
                hnew = posneg * hmax
            if reject:
                hnew = posneg * min(abs(hnew), abs(h))
            reject = False
        else:
            if reject_mode == "hairer" or d5:
                hnew = h / min(facc1, fac11 / safe)  # Hairer's dop853.f / dopri5.f (SciPy <= 1.16)
            else:
                hnew = h / facc1  # "scipy118": what SciPy 1.18.1's C dop853 does (measured)
            reject = True
            if res.naccpt >= 1:
                res.nrejct += 1
            last = False
        h = hnew
    res.h = h
    return y, res
