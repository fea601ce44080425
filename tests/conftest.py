import ctypes
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


with open(os.path.join(GOLDEN, "index.json")) as _f:
    GOLDEN_INDEX = json.load(_f)
GOLDEN_CASES = sorted(GOLDEN_INDEX)
SMALL_CASES = [c for c in GOLDEN_CASES if GOLDEN_INDEX[c]["dim"] <= 600]


class GoldenCase:
    def __init__(self, name):
        from bloqade_analog_b200 import serialize
        from bloqade_analog_b200.space import Space

        self.name = name
        self.data = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.program = serialize.program_from_dict(GOLDEN_INDEX[name]["program"])
        self.space = Space.create(self.program.register)
        self.dim = self.space.size
        self.n_atoms = self.space.n_atoms
        self.n_level = self.space.atom_type.n_level


_CASES = {}


def golden_case(name) -> GoldenCase:
    if name not in _CASES:
        _CASES[name] = GoldenCase(name)
    return _CASES[name]


def infidelity(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    return abs(1.0 - abs(np.vdot(a, b)) ** 2 / (np.vdot(a, a).real * np.vdot(b, b).real))


# ---- test-only host simulator of the device algorithm ------------------------------------------------
class HostSim:
    def __init__(self):
        src = os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")
        lib = os.path.join(ROOT, "tests", "hostsim", "libbqa_hostsim.so")
        csrc = os.path.join(ROOT, "bloqade_analog_b200", "csrc")
        deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cuh", ".h"))]
        if not os.path.exists(lib) or os.path.getmtime(lib) < max(os.path.getmtime(d) for d in deps):
            subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", csrc, src, "-o", lib], check=True)
        self.lib = ctypes.CDLL(lib)
        self.lib.hostsim_wf_eval.restype = ctypes.c_double
        self.lib.hostsim_record_size.restype = ctypes.c_int64

    @staticmethod
    def _p(a):
        return None if a is None else a.ctypes.data_as(ctypes.c_void_p)

    def _space_args(self, space):
        from bloqade_analog_b200.space import SpaceType

        if space.space_type is SpaceType.FullSpace:
            return None, None
        return (np.ascontiguousarray(space.configurations.astype(np.uint32)), np.ascontiguousarray(space.partner_table()))

    def wf_eval(self, record, w, t):
        return self.lib.hostsim_wf_eval(self._p(record), ctypes.c_int(w), ctypes.c_double(t))

    def apply(self, space, V, record, t, psi, interaction=False, raw=False):
        cfg, part = self._space_args(space)
        psi = np.ascontiguousarray(psi, dtype=np.complex128)
        out = np.zeros(space.size, dtype=np.complex128)
        self.lib.hostsim_apply(space.n_atoms, space.atom_type.n_level, ctypes.c_int64(space.size), self._p(cfg), self._p(part),
                               self._p(V), self._p(record), ctypes.c_double(t), int(interaction), int(raw), self._p(psi),
                               self._p(out))
        return out

    def evolve(self, space, V, record, times, solver="dop853", atol=1e-7, rtol=1e-14, interaction=False, reject_mode=0,
               psi0=None, nsteps=2**31 - 1):
        cfg, part = self._space_args(space)
        dim = space.size
        if psi0 is None:
            psi0 = np.zeros(dim, dtype=np.complex128)
            psi0[0] = 1.0
        psi0 = np.ascontiguousarray(psi0, dtype=np.complex128)
        times = np.ascontiguousarray(times, dtype=np.float64)
        out = np.zeros((times.size, dim), dtype=np.complex128)
        stats = np.zeros((times.size, 6))
        rc = self.lib.hostsim_evolve(space.n_atoms, space.atom_type.n_level, ctypes.c_int64(dim), self._p(cfg), self._p(part),
                                     self._p(V), self._p(record), {"dop853": 8, "dopri5": 5}[solver], ctypes.c_double(atol),
                                     ctypes.c_double(rtol), ctypes.c_int64(nsteps), int(interaction), int(reject_mode),
                                     self._p(psi0), self._p(times), int(times.size), self._p(out), self._p(stats))
        return out, stats, rc


@pytest.fixture(scope="session")
def hostsim():
    return HostSim()


def reference_available() -> bool:
    """the UNMODIFIED reference package: /root/reference/src here, baseline/_ref on the GPU box"""
    base = os.path.join(ROOT, "baseline")
    if base not in sys.path:
        sys.path.insert(0, base)
    import refshim

    return refshim.reference_available()


def load_reference():
    """import bloqade.analog (the reference) through baseline/refshim.py"""
    assert reference_available()
    import refshim

    refshim.install()
    import bloqade.analog

    return bloqade.analog


def cuda_available():
    try:
        from bloqade_analog_b200 import _lib

        return _lib.device_count() > 0
    except Exception:
        return False
