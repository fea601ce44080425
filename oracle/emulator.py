"""ORACLE (test infrastructure only - never imported by the product path; bench.py may time it as the
CPU baseline, `cpu_baseline.kind = "port"`).

CPU restatement (NumPy + SciPy, with Numba loops like the reference) of bloqade-analog's Python
emulator hot path, operating on the program description in `bloqade_analog_b200.ir`:

  operator generation   RydbergHamiltonianCodeGen          emulate/codegen/hamiltonian.py:63-203
  RHS                   _ode_complex_kernel(_int)          emulate/ir/state_vector.py:323-359
  Rabi term             RabiOperator.dot                   state_vector.py:279-288
  matvecs               IndexMapping / CSR / CSC loops     emulate/sparse_operator.py:15-34,99-128
  driver                AnalogGate._apply(_interaction_picture), run   state_vector.py:574-699
  expectation values    average / variance / local_trace   state_vector.py:30-80,401-463

Like the reference it keeps one f64[dim] diagonal per detuning term, one IndexMapping (single target
atom) or CSR (several) per Rabi term, evaluates every term in its own pass, and lets
`scipy.integrate.ode` (dop853 / dopri5) call back into Python for every RHS.  The integrator itself
is third-party SciPy (see oracle/dop.py for its restatement and the version note).

Pinned against the reference itself: tests/golden/*.npz hold operators, RHS values, final states and
observables produced by importing /root/reference in the build container
(tests/golden/make_golden.py); tests/test_oracle_golden.py compares this file with them.
"""
from __future__ import annotations

import numpy as np
from scipy.integrate import ode
from scipy.sparse import csr_matrix

from . import space as ospace

try:  # the reference's loops are numba-jitted (sparse_operator.py:15-34,99-128)
    from numba import njit
except Exception:  # pragma: no cover
    def njit(*a, **k):
        def wrap(f):
            return f
        return wrap if not (a and callable(a[0])) else a[0]

RB_C6 = 2 * np.pi * 862690


@njit(cache=True)
def _csr_matvec(nrow, data, indices, indptr, scale, x, out):
    for i in range(nrow):
        acc = 0.0j
        for j in range(indptr[i], indptr[i + 1]):
            acc += data[j] * x[indices[j]]
        out[i] += scale * acc
    return out


@njit(cache=True)
def _csc_matvec(ncol, data, indices, indptr, scale, x, out):
    for i in range(ncol):
        for j in range(indptr[i], indptr[i + 1]):
            out[indices[j]] += scale * data[j] * x[i]
    return out


@njit(cache=True)
def _map_matvec(rows, cols, scale, x, out):
    for k in range(rows.size):
        out[rows[k]] += scale * x[cols[k]]
    return out


class _RabiOp:
    """IndexMapping (one target atom) or CSR (several), as the reference chooses (HC:150-177)"""

    def __init__(self, dim, pairs_per_atom, scales):
        self.dim = dim
        if len(pairs_per_atom) == 1:
            self.kind = "map"
            self.rows = pairs_per_atom[0][0].astype(np.int64)
            self.cols = pairs_per_atom[0][1].astype(np.int64)
        else:
            self.kind = "csr"
            rows = np.concatenate([p[0] for p in pairs_per_atom])
            cols = np.concatenate([p[1] for p in pairs_per_atom])
            vals = np.concatenate([np.full(p[0].size, s) for p, s in zip(pairs_per_atom, scales)])
            m = csr_matrix((vals, (rows, cols)), shape=(dim, dim))
            m.sort_indices()
            self.data, self.indices, self.indptr = m.data.astype(np.float64), m.indices, m.indptr

    def matvec(self, x, out, scale):
        if self.kind == "map":
            return _map_matvec(self.rows, self.cols, scale, x, out)
        return _csr_matvec(self.dim, self.data, self.indices, self.indptr, scale, x, out)

    def rmatvec(self, x, out, scale):
        """out += scale * Op^T x  (the reference's `.T`: IndexMapping swaps rows/cols, CSR becomes CSC)"""
        if self.kind == "map":
            return _map_matvec(self.cols, self.rows, scale, x, out)
        return _csc_matvec(self.dim, self.data, self.indices, self.indptr, scale, x, out)

    def tocsr(self):
        if self.kind == "map":
            return csr_matrix((np.ones(self.rows.size), (self.rows, self.cols)), shape=(self.dim, self.dim))
        return csr_matrix((self.data, self.indices, self.indptr), shape=(self.dim, self.dim))


class Hamiltonian:
    def __init__(self, program):
        reg = program.register
        self.program = program
        self.n_level = reg.atom_type.n_level
        self.n_atoms = len(reg.sites)
        self.is_full, self.configs = ospace.create_space(reg.sites, self.n_level, reg.blockade_radius)
        self.dim = self.configs.size
        self.duration = float(program.duration)
        nl, ryd = self.n_level, self.n_level - 1
        # van der Waals diagonal (HC:73-89)
        pos = [np.asarray([float(x), float(y)]) for x, y in reg.sites]
        self.rydberg = np.zeros(self.dim, dtype=np.float64)
        eps = np.finfo(np.float64).eps
        for a in range(self.n_atoms):
            ra = ospace.is_state_at(self.configs, a, ryd, nl)
            for b in range(a + 1, self.n_atoms):
                v = RB_C6 / (np.linalg.norm(pos[a] - pos[b]) ** 6)
                if v <= eps:
                    continue
                self.rydberg[np.logical_and(ra, ospace.is_state_at(self.configs, b, ryd, nl))] += v
        self.detuning = []  # (diagonal, amplitude)
        self.rabi = []  # (op, amplitude, phase)
        for coupling, fields in program.pulses.items():
            if nl == 2:
                upper, lower = 1, 0
            elif coupling.value == "rydberg":
                upper, lower = 2, 1
            else:
                upper, lower = 1, 0
            for term in fields.detuning:  # HC:98-118
                d = np.zeros(self.dim, dtype=np.float64)
                for atom, value in term.operator_data.target_atoms.items():
                    d[ospace.is_state_at(self.configs, atom, upper, nl)] -= float(value)
                self.detuning.append((d, term.amplitude))
            for term in fields.rabi:  # HC:120-186: fro = upper level, to = lower level
                pairs, scales = [], []
                symmetric = int(term.operator_data.operator_type) == 1
                for atom, value in term.operator_data.target_atoms.items():
                    if symmetric:
                        pairs.append(ospace.swap_pairs(self.configs, self.is_full, nl, atom, upper, lower))
                    else:
                        pairs.append(ospace.transition_pairs(self.configs, self.is_full, nl, atom, upper, lower))
                    scales.append(float(value))
                self.rabi.append((_RabiOp(self.dim, pairs, scales), term.amplitude, term.phase))

    # ---- SV:323-334 --------------------------------------------------------------------------
    def _diag(self, t, with_rydberg=True):
        diag = self.rydberg.copy() if with_rydberg else np.zeros(self.dim)
        for d, amp in self.detuning:
            diag = diag + d * amp(t)
        return diag

    def _rabi_dot(self, x, out, t):
        for op, amp, phase in self.rabi:  # SV:279-288
            a = amp(t) / 2
            if phase is None:
                op.matvec(x, out, a)
            else:
                a = a * np.exp(-1j * phase(t))
                op.matvec(x, out, a)
                op.rmatvec(x, out, np.conj(a))
        return out

    def apply(self, psi, t=None):
        """H(t) psi  (SV:375-398)"""
        t = self.duration if t is None else t
        out = (self._diag(t) * psi).astype(np.complex128)
        return self._rabi_dot(np.ascontiguousarray(psi, dtype=np.complex128), out, t)

    def rhs(self, t, psi):
        return -1j * self.apply(psi, t)

    def rhs_interaction(self, t, psi):  # SV:342-359
        u = np.exp(-1j * self.rydberg * t)
        phi = u * psi
        out = (self._diag(t, with_rydberg=False) * phi).astype(np.complex128)
        self._rabi_dot(phi, out, t)
        return -1j * np.conj(u) * out

    def tocsr(self, t):
        from scipy.sparse import diags

        h = diags(self._diag(t)).tocsr().astype(np.complex128)
        for op, amp, phase in self.rabi:
            a = amp(t) / 2
            if phase is None:
                h = h + op.tocsr() * a
            else:
                m = op.tocsr() * (a * np.exp(-1j * phase(t)))
                h = h + m + m.T.conj()
        return h

    def average_and_variance(self, psi, t=None):  # SV:419-442
        hp = self.apply(psi, t)
        avg = np.vdot(psi, hp).real
        return avg, np.vdot(hp, hp).real - avg**2


def evolve(program, psi0=None, times=(), solver_name="dop853", atol=1e-7, rtol=1e-14, nsteps=2**31 - 1,
           interaction_picture=False, hamiltonian=None, return_counts=False):
    """AnalogGate._apply / _apply_interaction_picture (SV:574-641): list of states at `times`
    (default: [duration]) using scipy.integrate.ode exactly like the reference."""
    h = Hamiltonian(program) if hamiltonian is None else hamiltonian
    times = [h.duration] if len(times) == 0 else list(times)
    if psi0 is None:
        psi0 = np.zeros(h.dim, dtype=np.complex128)
        psi0[0] = 1.0
    psi0 = np.asarray(psi0, dtype=np.complex128)
    fn = h.rhs_interaction if interaction_picture else h.rhs
    calls = [0]

    def f(t, y):
        calls[0] += 1
        return fn(t, y.view(np.complex128)).view(np.float64)

    solver = ode(f)
    solver.set_initial_value(psi0.view(np.float64).copy())
    solver.set_integrator(solver_name, atol=atol, rtol=rtol, nsteps=nsteps)
    out = []
    last = psi0.copy()
    for t in times:
        if solver.t != t:
            solver.integrate(t)
            code = solver.get_return_code()
            if code not in (1, 2):
                raise RuntimeError(f"DOP853/DOPRI5 status {code}")
            y = solver.y.view(np.complex128).copy()
            last = np.exp(-1j * t * h.rydberg) * y if interaction_picture else y
        out.append(last.copy())
    if return_counts:
        return out, calls[0]
    return out


def run(program, shots, uniforms=None, **kw):
    """AnalogGate.run (SV:674-699): evolve, normalise, sample"""
    h = Hamiltonian(program)
    psi = evolve(program, hamiltonian=h, **kw)[-1]
    psi = psi / np.linalg.norm(psi)
    if uniforms is None:
        uniforms = np.random.random_sample(shots)
    return ospace.sample_state_vector(h.configs, psi, uniforms, h.n_atoms, h.n_level)


def local_trace(h: Hamiltonian, psi, op, sites):
    """_expt_one_body_op / _expt_two_body_op (SV:30-80) as plain loops; raw sum (no norm division)"""
    nl = h.n_level
    cfgs = h.configs.astype(np.int64)
    res = 0.0j
    if len(sites) == 1:
        div = nl ** sites[0]
        for i, c in enumerate(cfgs):
            col = (c // div) % nl
            for row in range(nl):
                ele = op[row, col]
                if ele == 0:
                    continue
                new = c - col * div + row * div
                j = np.searchsorted(cfgs, new)
                if j < cfgs.size and cfgs[j] == new:
                    res += ele * psi[i] * np.conj(psi[j])
        return res
    d1, d2 = nl ** sites[1], nl ** sites[0]
    for i, c in enumerate(cfgs):
        c1, c2 = (c // d1) % nl, (c // d2) % nl
        col = c1 + c2 * nl
        for row in range(nl * nl):
            ele = op[row, col]
            if ele == 0:
                continue
            r2, r1 = divmod(row, nl)
            new = c - c1 * d1 + r1 * d1 - c2 * d2 + r2 * d2
            j = np.searchsorted(cfgs, new)
            if j < cfgs.size and cfgs[j] == new:
                res += ele * psi[i] * np.conj(psi[j])
    return res
