"""CPU checks of the drop-in boundary: libbqa_b200.so loads without a GPU, exports every symbol that
include/bqa_b200.h declares, its structs match the ctypes mirrors, and compute entry points fail
loudly (no CPU fallback) when no CUDA device is present."""
import ctypes
import os
import re

import numpy as np
import pytest
from conftest import ROOT, cuda_available

from bloqade_analog_b200 import _lib

HEADER = os.path.join(ROOT, "include", "bqa_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bqa_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_functions()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/bqa_b200.h but not exported"
    assert sorted(_lib.EXPORTS) == names, "ctypes EXPORTS list and header disagree"
    assert lib.bqa_version() >= 100


def test_struct_layouts_match_header():
    assert ctypes.sizeof(_lib.SpaceDesc) == 32
    assert ctypes.sizeof(_lib.SolverOpts) == 48 and _lib.SolverOpts.t_start.offset == 40
    assert ctypes.sizeof(_lib.TaskStats) == 56
    assert _lib.STATS_DTYPE.itemsize == ctypes.sizeof(_lib.TaskStats)


def test_error_channel_without_crossing_exceptions():
    lib = _lib.load()
    rc = lib.bqa_plan_create(None, None)
    assert rc != 0
    assert b"null" in lib.bqa_last_error()


def test_unsupported_solver_names():
    with pytest.raises(NotImplementedError):
        _lib.make_opts("lsoda")
    with pytest.raises(ValueError):
        _lib.make_opts("rk4")
    o = _lib.make_opts("dopri5", 1e-9, 1e-8, 77, True)
    assert (o.solver, o.interaction_picture, o.atol, o.rtol, o.nsteps) == (5, 1, 1e-9, 1e-8, 77)


@pytest.mark.skipif(cuda_available(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback():
    with pytest.raises(_lib.BackendError):
        _lib.Plan(3, 2, 8)
    from bloqade_analog_b200 import ir, programs
    from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen

    with pytest.raises(_lib.BackendError):
        RydbergHamiltonianCodeGen().emit(programs.adiabatic_program(ir.chain_sites(3, 6.0)))
    from bloqade_analog_b200.space import Space

    sp = Space.create(programs.adiabatic_program(ir.chain_sites(3, 6.0)).register)
    with pytest.raises(_lib.BackendError):
        sp.sample_state_vector(np.ones(8, dtype=np.complex128) / np.sqrt(8), 4)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "bloqade_analog_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "hostsim" not in src or f.endswith(".cuh"), f
