"""Drop-in routine: `program.bloqade.python().run(...)` on the B200 backend.

Needs a real `bloqade.analog` installation (it subclasses the reference's `BloqadePythonRoutine`,
`src/bloqade/analog/ir/routine/bloqade.py:148-550`, and returns the reference's own `LocalBatch` /
`BloqadeTask` / `QuEraTaskResults` objects).  The reference's compile pipeline (`_generate_ir`, RB:184-202)
is reused untouched; only the per-task numerical work (`BloqadeTask.run`, task/bloqade.py:43-88, i.e.
operator generation + `AnalogGate.run`) is replaced by ONE batched GPU call per Hilbert space.
"""
from __future__ import annotations

from typing import Any, Callable, List, Optional, Tuple

import numpy as np

from . import adapter
from .batch import run_programs
from .hamiltonian import CompileCache, RydbergHamiltonianCodeGen
from .state_vector import AnalogGate, StateVector

_ROUTINE_CLASS = None
_ORIGINAL_PYTHON = None


def _package(task, result):
    """shots -> QuEraTaskResults exactly like task/bloqade.py:66-86 (1 = ground / present, 0 = Rydberg / empty)"""
    from bloqade.analog.submission.ir.task_results import (
        QuEraShotResult,
        QuEraShotStatusCode,
        QuEraTaskResults,
        QuEraTaskStatusCode,
    )

    filling = np.asarray(task.emulator_ir.register.geometry.filling, dtype=int)
    post = result.post_sequence(filling)
    pre = filling.tolist()
    shot_outputs = [
        QuEraShotResult.construct(shot_status=QuEraShotStatusCode.Completed, pre_sequence=pre, post_sequence=row)
        for row in post.tolist()
    ]
    task.task_result_ir = QuEraTaskResults.construct(task_status=QuEraTaskStatusCode.Completed, shot_outputs=shot_outputs)
    return task


class B200Emulation:
    """Counterpart of `BloqadeEmulation` (RB:58-145): lazily generated GPU Hamiltonian + evolve()."""

    def __init__(self, task_data, compile_cache: Optional[CompileCache] = None):
        self.task_data = task_data
        self.compile_cache = compile_cache
        self._hamiltonian = None

    @property
    def metadata(self):
        return self.task_data.metadata

    @property
    def hamiltonian(self):
        if self._hamiltonian is None:
            program = adapter.convert_program(self.task_data.emulator_ir)
            self._hamiltonian = RydbergHamiltonianCodeGen(self.compile_cache).emit(program)
        return self._hamiltonian

    def zero_state(self, dtype=np.float64) -> StateVector:
        return self.hamiltonian.space.zero_state(dtype)

    def fock_state(self, fock_state_str: str, dtype=np.float64) -> StateVector:
        space = self.hamiltonian.space
        state = StateVector(np.zeros(space.size, dtype=dtype), space)
        state.data[space.fock_state_to_index(fock_state_str)] = 1
        return state

    def evolve(self, state: Optional[StateVector] = None, solver_name: str = "dop853", atol: float = 1e-7,
               rtol: float = 1e-14, nsteps: int = 2_147_483_647, times=(), interaction_picture: bool = False):
        state = self.zero_state(np.complex128) if state is None else state
        return AnalogGate(self.hamiltonian).apply(state, solver_name, atol, rtol, nsteps, times, interaction_picture)


def routine_class():
    """`BloqadeB200Routine`, created on first use (importing bloqade.analog is deferred until then)."""
    global _ROUTINE_CLASS
    if _ROUTINE_CLASS is not None:
        return _ROUTINE_CLASS
    from bloqade.analog.ir.routine.bloqade import BloqadePythonRoutine

    class BloqadeB200Routine(BloqadePythonRoutine):
        """Same signatures, defaults and return types as `BloqadePythonRoutine` (RB:233-249,348-365,505-512)."""

        def run(self, shots: int, args: Tuple = (), name: Optional[str] = None, blockade_radius: float = 0.0,
                waveform_runtime: str = "interpret", interaction_picture: bool = False, cache_matrices: bool = False,
                multiprocessing: bool = False, num_workers: Optional[int] = None, solver_name: str = "dop853",
                atol: float = 1e-7, rtol: float = 1e-14, nsteps: int = 2_147_483_647, *, device: Optional[int] = None):
            if multiprocessing and cache_matrices:
                raise ValueError("Cannot use multiprocessing and cache_matrices at the same time.")
            batch = self._compile(shots=shots, args=args, name=name, blockade_radius=blockade_radius,
                                  cache_matrices=cache_matrices, waveform_runtime=waveform_runtime)
            tasks = list(batch.tasks.values())
            programs = [adapter.convert_program(t.emulator_ir) for t in tasks]
            # one GPU batch per Hilbert space; multiprocessing / num_workers have no meaning here (accepted, ignored)
            results = run_programs(programs, shots, solver_name, atol, rtol, nsteps, interaction_picture,
                                   project_hyperfine=True, compile_cache=CompileCache(), device=device)
            for task, res in zip(tasks, results):
                _package(task, res)
            return batch

        def __call__(self, *args, shots: int = 1, **kwargs):
            return self.run(shots, args=args, **kwargs)

        def run_callback(self, callback: Callable, program_args: Tuple = (), callback_args: Tuple = (),
                         ignore_exceptions: bool = False, blockade_radius: float = 0.0,
                         waveform_runtime: str = "interpret", interaction_picture: bool = False,
                         cache_matrices: bool = False, multiprocessing: bool = False,
                         num_workers: Optional[int] = None, solver_name: str = "dop853", atol: float = 1e-7,
                         rtol: float = 1e-14, nsteps: int = 2_147_483_647, use_hyperfine: bool = False) -> List[Any]:
            if multiprocessing and cache_matrices:
                raise ValueError("Cannot use multiprocessing and cache_matrices at the same time.")
            cache = CompileCache()
            results: List[Any] = []
            first_error = None
            for task_data in self._generate_ir(program_args, blockade_radius, waveform_runtime, use_hyperfine):
                try:
                    em = B200Emulation(task_data, cache)
                    (state,) = em.evolve(solver_name=solver_name, atol=atol, rtol=rtol, nsteps=nsteps,
                                         interaction_picture=interaction_picture)
                    results.append(callback(state, task_data.metadata, em.hamiltonian, *callback_args))
                except BaseException as e:  # same policy as RB:150-158,488-501
                    results.append(e)
                    if first_error is None:
                        first_error = (task_data.task_id, e)
            if first_error is not None and not ignore_exceptions:
                raise RuntimeError(f"{first_error[1].__class__.__name__} occured during child process running for task "
                                   f"number {first_error[0]}:\n{first_error[1]}") from first_error[1]
            return results

        def hamiltonian(self, *args, blockade_radius: float = 0.0, use_hyperfine: bool = False,
                        waveform_runtime: str = "interpret", cache_matrices: bool = False) -> List[B200Emulation]:
            cache = CompileCache() if cache_matrices else None
            return [B200Emulation(td, cache)
                    for td in self._generate_ir(args, blockade_radius, waveform_runtime, use_hyperfine)]

    _ROUTINE_CLASS = BloqadeB200Routine
    return _ROUTINE_CLASS


def b200_routine(program):
    """`program` is a builder or a Routine; returns the B200 routine for it (no global patching)."""
    routine = program.parse() if hasattr(program, "parse") else program
    return routine_class()(routine.source, routine.circuit, routine.params)


def install():
    """Opt-in: make the literal `program.bloqade.python()` return the B200 routine."""
    global _ORIGINAL_PYTHON
    from bloqade.analog.ir.routine.bloqade import BloqadeServiceOptions

    cls = routine_class()
    if _ORIGINAL_PYTHON is None:
        _ORIGINAL_PYTHON = BloqadeServiceOptions.python
    BloqadeServiceOptions.python = lambda self: cls(self.source, self.circuit, self.params)
    BloqadeServiceOptions.b200 = BloqadeServiceOptions.python


def uninstall():
    global _ORIGINAL_PYTHON
    if _ORIGINAL_PYTHON is not None:
        from bloqade.analog.ir.routine.bloqade import BloqadeServiceOptions

        BloqadeServiceOptions.python = _ORIGINAL_PYTHON
        _ORIGINAL_PYTHON = None
