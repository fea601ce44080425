"""StateVector / RydbergHamiltonian / AnalogGate of the B200 backend.

Same names, signatures, defaults and error behaviour as the reference's
`src/bloqade/analog/emulate/ir/state_vector.py` (StateVector :83-217, DetuningOperator :261-270,
RabiOperator :273-298, RydbergHamiltonian :301-484, AnalogGate :487-699) - but every numerical
method runs on the GPU through libbqa_b200.so (kernels (1)-(5)); there is no NumPy/SciPy compute
path here.  `StateVector.data` stays a host `np.ndarray[complex128]`.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np

from . import _lib
from .ir import EmulatorProgram, RabiOperatorType, Waveform
from .lowering import pack_task_record, pair_interactions
from .space import Space, SpaceType


@dataclass(frozen=True)
class StateVector:
    data: np.ndarray
    space: Space

    def local_trace(self, matrix: np.ndarray, site_index: Union[int, Tuple[int, int]]) -> complex:
        """Trace of a one- or two-site operator over the state (state_vector.py:88-202).
        One-body results are divided by norm() (SV:175), two-body results are not (SV:133)."""
        n_level = self.space.atom_type.n_level
        reg = self.space.program_register
        if isinstance(site_index, (int, np.integer)):
            shape = (n_level, n_level)
            if matrix.shape != shape:
                raise ValueError(f"expecting operator to be size {shape}, got {matrix.shape}")
            if site_index < 0 or site_index >= self.space.n_sites:
                raise ValueError(f"site_index: {site_index} out of bounds with {self.space.n_sites} sites")
            if not reg.geometry.filling[site_index]:
                raise ValueError(f"Trying to measure site {site_index} which is empty.")
            atom = reg.full_index_to_index[int(site_index)]
            value = _plan_for_space(self.space).local_trace(self.data, matrix, (atom,))
            return value / self.norm()
        site_indices = tuple(site_index)
        shape = (n_level**2, n_level**2)
        if matrix.shape != shape:
            raise ValueError(f"expecting operator to be size {shape}, got site {matrix.shape}")
        for s in site_indices:
            if s < 0 or s >= self.space.n_sites:
                raise ValueError(f"site_index: {s} out of bounds with {self.space.n_sites} sites")
        for s in site_indices:
            if not reg.geometry.filling[s]:
                raise ValueError(f"Trying to measure site {s} which is empty.")
        atoms = tuple(reg.full_index_to_index[int(s)] for s in site_indices)
        return _plan_for_space(self.space).local_trace(self.data, matrix, atoms)

    def sample(self, shots: int, project_hyperfine: bool = True) -> np.ndarray:
        return self.space.sample_state_vector(self.data, shots, project_hyperfine=project_hyperfine)

    def normalize(self) -> None:
        data = self.data
        if data.dtype != np.complex128 or not data.flags.c_contiguous:
            raise TypeError("normalize() needs a contiguous complex128 state")
        _plan_for_space(self.space).normalize(data)

    def norm(self) -> float:
        return _plan_for_space(self.space).norm(self.data)


_SPACE_PLANS = {}


def _plan_for_space(space: Space) -> "_lib.Plan":
    """structure-only plan (no diagonals) cached per Space object for the reduction kernels"""
    key = id(space)
    hit = _SPACE_PLANS.get(key)
    if hit is not None and hit[0] is space:
        return hit[1]
    plan = make_plan(space)
    if len(_SPACE_PLANS) > 8:
        _SPACE_PLANS.clear()
    _SPACE_PLANS[key] = (space, plan)
    return plan


def make_plan(space: Space, device: Optional[int] = None) -> "_lib.Plan":
    n, nl = space.n_atoms, space.atom_type.n_level
    if space.space_type is SpaceType.FullSpace:
        return _lib.Plan(n, nl, space.size, device=device)
    return _lib.Plan(n, nl, space.size, space.configurations.astype(np.uint32), space.partner_table(), device=device)


@dataclass(frozen=True)
class DetuningOperator:
    """One detuning term: -(sum_a scale_a n_a) * amplitude(t)  (state_vector.py:261-270).
    `diagonal` is materialised on the host only when asked for (API parity / tocsr)."""

    target_atoms: dict
    slot: int
    space: Space
    amplitude: Optional[Callable[[float], float]] = None

    @property
    def diagonal(self) -> np.ndarray:
        d = np.zeros(self.space.size, dtype=np.float64)
        for atom, value in self.target_atoms.items():
            d[self.space.is_state_at(atom, self.slot + 1)] -= float(value)
        return d

    def get_diagonal(self, time: float):
        return self.diagonal * self.amplitude(time) if self.amplitude else self.diagonal


@dataclass(frozen=True)
class RabiOperator:
    """One Rabi term (state_vector.py:273-298); `tocsr` builds the host matrix for inspection."""

    target_atoms: dict
    slot: int
    operator_type: RabiOperatorType
    space: Space
    amplitude: Callable[[float], float]
    phase: Optional[Callable[[float], float]] = None

    def op_tocsr(self):
        from scipy.sparse import coo_matrix

        size = self.space.size
        rows_all, cols_all, vals = [], [], []
        for atom, value in self.target_atoms.items():
            if self.operator_type is RabiOperatorType.RabiSymmetric:
                rows, cols = self.space.swap_state_at(atom, self.slot + 1, self.slot)
            else:
                rows, cols = self.space.transition_state_at(atom, self.slot + 1, self.slot)
            if isinstance(rows, slice):
                rows = np.arange(size)
            rows_all.append(np.asarray(rows, dtype=np.int64))
            cols_all.append(np.asarray(cols, dtype=np.int64))
            vals.append(np.full(len(cols), float(value)))
        return coo_matrix(
            (np.concatenate(vals), (np.concatenate(rows_all), np.concatenate(cols_all))), shape=(size, size)
        ).tocsr()

    def tocsr(self, time: float):
        amplitude = self.amplitude(time) / 2
        if self.phase is None:
            return self.op_tocsr() * amplitude
        amplitude = amplitude * np.exp(-1j * self.phase(time))
        mat = self.op_tocsr() * amplitude
        return mat + mat.T.conj()


@dataclass
class RydbergHamiltonian:
    """Hamiltonian of one task (state_vector.py:301-484), resident on the GPU.

    Attributes mirror the reference (`emulator_ir`, `space`, `rydberg`, `detuning_ops`, `rabi_ops`).
    `plan` is the device plan (space + van-der-Waals diagonal), `record` the flat task record
    holding the waveforms and per-atom scales (include/bqa_b200.h)."""

    emulator_ir: EmulatorProgram
    space: Space
    detuning_ops: List[DetuningOperator] = field(default_factory=list)
    rabi_ops: List[RabiOperator] = field(default_factory=list)
    plan: Optional["_lib.Plan"] = None
    diag_id: int = -1
    record: Optional[np.ndarray] = None
    _rydberg: Optional[np.ndarray] = None

    @property
    def rydberg(self) -> np.ndarray:
        if self._rydberg is None:
            self._rydberg = self.plan.get_diagonal(self.diag_id)
        return self._rydberg

    def _check_register(self, register: np.ndarray):
        register_shape = (self.space.size,)
        if register.shape != register_shape:
            raise ValueError(f"Expecting `register` to have  shape {register_shape}, got shape {register.shape}")

    def _apply(self, register: np.ndarray, time: Optional[float] = None, output: Optional[np.ndarray] = None):
        """H(time) @ register on the GPU (kernel (1), raw form; SV:375-398)"""
        self._check_register(register)
        if time is None:
            time = self.emulator_ir.duration
        if output is None:
            output = np.zeros_like(register, dtype=np.complex128)
        return self.plan.apply(self.record, float(time), register, output, raw=True)

    def _ode_complex_kernel(self, time: float, register: np.ndarray, output: np.ndarray):
        """-i H(time) register (SV:323-334)"""
        return self.plan.apply(self.record, float(time), register, output)

    def _ode_complex_kernel_int(self, time: float, register: np.ndarray, output: np.ndarray):
        """interaction-picture right-hand side (SV:342-359)"""
        return self.plan.apply(self.record, float(time), register, output, interaction_picture=True)

    def average(self, register: StateVector, time: Optional[float] = None) -> float:
        return self.average_and_variance(register, time)[0]

    def average_and_variance(self, register: StateVector, time: Optional[float] = None) -> Tuple[float, float]:
        if not isinstance(register, StateVector):
            raise TypeError("register must be a StateVector")
        self._check_register(register.data)
        if time is None:
            time = self.emulator_ir.duration
        return self.plan.expectation(self.record, float(time), register.data)

    def variance(self, register: StateVector, time: Optional[float] = None) -> float:
        return self.average_and_variance(register, time)[1]

    def tocsr(self, time: float):
        """host scipy CSR of H(time) - inspection / testing aid (SV:465-484)"""
        from scipy.sparse import diags

        diagonal = self.rydberg.copy()
        for d in self.detuning_ops:
            diagonal = diagonal + d.get_diagonal(time)
        hamiltonian = diags(diagonal).tocsr()
        for r in self.rabi_ops:
            hamiltonian = hamiltonian + r.tocsr(time)
        return hamiltonian


_DOP_ERRORS = {
    -1: "DOP853/DOPRI5: Input is not consistent.",
    -2: "DOP853/DOPRI5: Larger nsteps is needed.",
    -3: "DOP853/DOPRI5: Step size becomes too small.",
    -4: "DOP853/DOPRI5: Problem is probably stiff (interrupted).",
    -6: "sharded state: a peer rank failed or a cross-rank barrier timed out.",
}


@dataclass(frozen=True)
class AnalogGate:
    SUPPORTED_SOLVERS = ["lsoda", "dop853", "dopri5"]

    hamiltonian: RydbergHamiltonian

    @staticmethod
    def _error_check(solver_name: str, status_code: int):
        """same messages as state_vector.py:494-508"""
        if status_code in (1, 2):
            return
        if status_code in _DOP_ERRORS:
            raise RuntimeError(_DOP_ERRORS[status_code])
        raise RuntimeError(f"DOP853/DOPRI5: unhandled status code {status_code}")

    def _check_args(self, state_vec, solver_name, atol, rtol, nsteps, times):
        duration = self.hamiltonian.emulator_ir.duration
        times = [duration] if len(times) == 0 else times
        if state_vec is None:
            state_vec = self.hamiltonian.space.zero_state(np.complex128)
        if state_vec.space != self.hamiltonian.space:
            raise ValueError("State vector not in the same space as the Hamiltonian.")
        if solver_name not in AnalogGate.SUPPORTED_SOLVERS:
            raise ValueError(f"'{solver_name}' not supported.")
        if any(time > duration or time < 0.0 for time in times):
            raise ValueError(f"Times must be between 0 and duration {duration}. found {times}")
        return state_vec, solver_name, atol, rtol, nsteps, times

    def apply(
        self,
        state: StateVector,
        solver_name: str = "dop853",
        atol: float = 1e-7,
        rtol: float = 1e-14,
        nsteps: int = 2_147_483_647,
        times: Union[Sequence[float], np.ndarray] = (),
        interaction_picture: bool = False,
        *,
        reject_mode: int = 1,
        engine: str = "auto",
    ) -> Iterator[StateVector]:
        """Evolve `state` and yield one StateVector per requested time (SV:574-671).  The whole
        evolution (all times) is ONE library call; like SciPy's `ode`, the integrator restarts at
        every output time."""
        state, solver_name, atol, rtol, nsteps, times = self._check_args(state, solver_name, atol, rtol, nsteps, times)
        opts = _lib.make_opts(solver_name, atol, rtol, nsteps, interaction_picture, reject_mode, engine)
        h = self.hamiltonian
        psi0 = np.ascontiguousarray(np.asarray(state.data).astype(np.complex128, copy=True))
        times_arr = np.asarray([float(t) for t in times], dtype=np.float64).reshape(1, -1)
        offsets = np.array([0, h.record.size], dtype=np.int64)
        out, stats = h.plan.evolve_batch(h.record, offsets, times_arr, opts, psi0=psi0, psi0_shared=True)
        return self._yield_states(out[0], stats[0], solver_name)

    def _yield_states(self, states, stats, solver_name):
        for k in range(states.shape[0]):
            AnalogGate._error_check(solver_name, int(stats[k]["status"]))
            yield StateVector(states[k].copy(), self.hamiltonian.space)

    def run(
        self,
        shots: int = 1,
        solver_name: str = "dop853",
        atol: float = 1e-14,
        rtol: float = 1e-7,
        nsteps: int = 2_147_483_647,
        interaction_picture: bool = False,
        project_hyperfine: bool = True,
        *,
        reject_mode: int = 1,
        engine: str = "auto",
    ) -> np.ndarray:
        """zero state -> evolve -> normalise -> sample, fused in one library call (SV:674-699)"""
        if solver_name not in AnalogGate.SUPPORTED_SOLVERS:
            raise ValueError(f"'{solver_name}' not supported.")
        opts = _lib.make_opts(solver_name, atol, rtol, nsteps, interaction_picture, reject_mode, engine)
        h = self.hamiltonian
        offsets = np.array([0, h.record.size], dtype=np.int64)
        times = np.array([h.emulator_ir.duration], dtype=np.float64)
        uniforms = np.random.random_sample(int(shots)).reshape(1, -1)
        shots_arr, stats = h.plan.run_batch(h.record, offsets, times, uniforms, opts, project_hyperfine)
        AnalogGate._error_check(solver_name, int(stats[0]["status"]))
        return shots_arr[0]
