"""ctypes binding of the C ABI in include/bqa_b200.h (libbqa_b200.so, built in-tree by `build()`).

There is NO CPU fallback: if the shared library is missing or no CUDA device is usable, every compute
entry point raises (north_star: "no CPU fallback").  Loading the library and resolving its symbols
works without a GPU (it only links libcudart), which is what the CPU-side tests check.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, byref, c_char_p, c_double, c_int32, c_int64, c_uint8, c_void_p
from typing import Optional, Sequence

import numpy as np

MEM_HOST = 1
MEM_DEVICE = 2
SOLVERS = {"dopri5": 5, "dop853": 8}
ENGINES = {"auto": 0, "cta": 1, "streamed": 2, "cluster": 3}

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbqa_b200.so")

EXPORTS = [
    "bqa_version", "bqa_last_error", "bqa_device_count", "bqa_set_device", "bqa_device_info",
    "bqa_launch_count", "bqa_tiled_plan_describe", "bqa_plan_create", "bqa_plan_destroy", "bqa_plan_set_stream",
    "bqa_plan_add_diagonal", "bqa_plan_build_rydberg", "bqa_plan_get_diagonal", "bqa_plan_add_csr",
    "bqa_apply", "bqa_evolve_batch", "bqa_run_batch", "bqa_norm", "bqa_normalize", "bqa_expectation",
    "bqa_local_trace", "bqa_sample", "bqa_plan_enable_timing", "bqa_plan_get_timing",
    "bqa_shard_sync_bytes", "bqa_shard_alloc", "bqa_shard_open", "bqa_shard_close", "bqa_shard_free", "bqa_shard_reset",
    "bqa_plan_build_rydberg_shard", "bqa_apply_sharded", "bqa_evolve_sharded", "bqa_norm_noalloc", "bqa_current_device",
]


class SpaceDesc(Structure):
    _fields_ = [("n_atoms", c_int32), ("n_level", c_int32), ("dim", c_int64), ("configs", c_void_p),
                ("partners", c_void_p)]


class SolverOpts(Structure):
    _fields_ = [("solver", c_int32), ("interaction_picture", c_int32), ("atol", c_double), ("rtol", c_double),
                ("nsteps", c_int64), ("reject_mode", c_int32), ("engine", c_int32), ("t_start", c_double)]


class ShardDesc(Structure):
    """bqa_shard_desc (include/bqa_b200.h)"""
    _fields_ = [("world", c_int32), ("rank", c_int32), ("n_atoms_total", c_int32), ("pad", c_int32),
                ("ys", c_void_p * 8), ("sync", c_void_p * 8)]


class TaskStats(Structure):
    _fields_ = [("status", c_int32), ("pad", c_int32), ("nfcn", c_int64), ("nstep", c_int64), ("naccpt", c_int64),
                ("nrejct", c_int64), ("h_last", c_double), ("t_last", c_double)]


STATS_DTYPE = np.dtype([("status", np.int32), ("pad", np.int32), ("nfcn", np.int64), ("nstep", np.int64),
                        ("naccpt", np.int64), ("nrejct", np.int64), ("h_last", np.float64), ("t_last", np.float64)])

_lib = None


class BackendError(RuntimeError):
    pass


def load():
    """Load libbqa_b200.so (once).  Raises BackendError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise BackendError(
            f"{LIB_PATH} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; "
            "g.build()').  The B200 backend has no CPU fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    lib.bqa_last_error.restype = c_char_p
    lib.bqa_launch_count.restype = c_int64
    lib.bqa_shard_sync_bytes.restype = c_int64
    lib.bqa_plan_destroy.restype = None
    _lib = lib
    return lib


def _check(rc: int):
    if rc != 0:
        raise BackendError(load().bqa_last_error().decode())


def device_count() -> int:
    n = c_int32(0)
    try:
        _check(load().bqa_device_count(byref(n)))
    except BackendError:
        return 0
    return n.value


def current_device() -> int:
    """the CUDA device current in this thread"""
    n = c_int32(0)
    _check(load().bqa_current_device(byref(n)))
    return n.value


def require_device():
    if device_count() < 1:
        raise BackendError("no CUDA device available: the B200 backend has no CPU fallback")


def launch_count() -> int:
    return int(load().bqa_launch_count())


def tiled_plan_describe(n_local: int, n_total: int = None):
    """pass list of the tiled H psi: dicts with lo_bits, hb, tb, p0, n_x, xbit, pair, pair_b (host arithmetic only)"""
    import numpy as np

    out = np.zeros((8, 8), dtype=np.int32)
    n = load().bqa_tiled_plan_describe(c_int32(n_local), c_int32(n_local if n_total is None else n_total),
                                       out.ctypes.data_as(c_void_p), c_int32(8))
    if n < 0:
        raise BackendError("bqa_tiled_plan_describe: bad shape")
    keys = ("lo_bits", "hb", "tb", "p0", "n_x", "xbit", "pair", "pair_b")
    return [dict(zip(keys, (int(v) for v in out[i]))) for i in range(n)]


def _ptr(a) -> c_void_p:
    """host numpy array or integer device pointer -> void*"""
    if a is None:
        return c_void_p(None)
    if isinstance(a, (int, np.integer)):
        return c_void_p(int(a))
    return c_void_p(a.ctypes.data)


def make_opts(solver_name="dop853", atol=1e-7, rtol=1e-14, nsteps=2_147_483_647, interaction_picture=False,
              reject_mode=1, engine="auto", t_start=0.0) -> SolverOpts:
    if solver_name == "lsoda":
        raise NotImplementedError(
            "solver 'lsoda' has no GPU implementation in the B200 backend (and there is no CPU fallback); "
            "use 'dop853' or 'dopri5'"
        )
    if solver_name not in SOLVERS:
        raise ValueError(f"'{solver_name}' not supported.")
    return SolverOpts(SOLVERS[solver_name], int(bool(interaction_picture)), float(atol), float(rtol), int(nsteps),
                      int(reject_mode), ENGINES[engine] if isinstance(engine, str) else int(engine), float(t_start))


class Plan:
    """Owner of a `bqa_plan*`: one Hilbert space + its static diagonals on the current device."""

    def __init__(self, n_atoms: int, n_level: int, dim: int, configs: Optional[np.ndarray] = None,
                 partners: Optional[np.ndarray] = None, device: Optional[int] = None):
        lib = load()
        require_device()
        if device is not None:
            _check(lib.bqa_set_device(int(device)))
        self.n_atoms, self.n_level, self.dim = int(n_atoms), int(n_level), int(dim)
        self._configs = None if configs is None else np.ascontiguousarray(configs, dtype=np.uint32)
        self._partners = None if partners is None else np.ascontiguousarray(partners, dtype=np.int32)
        desc = SpaceDesc(self.n_atoms, self.n_level, self.dim, _ptr(self._configs), _ptr(self._partners))
        h = c_void_p()
        _check(lib.bqa_plan_create(byref(desc), byref(h)))
        self._h = h
        self.n_diags = 0

    def close(self):
        if getattr(self, "_h", None):
            load().bqa_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, stream: int):
        _check(load().bqa_plan_set_stream(self._h, c_void_p(int(stream))))

    def enable_timing(self, on: bool = True):
        _check(load().bqa_plan_enable_timing(self._h, c_int32(int(on))))

    def get_timing(self):
        """(cumulative device ms of the bracketed evolve kernels, number of such launches)"""
        ms, n = c_double(0.0), c_int64(0)
        _check(load().bqa_plan_get_timing(self._h, byref(ms), byref(n)))
        return ms.value, n.value

    def add_diagonal(self, v: np.ndarray) -> int:
        v = np.ascontiguousarray(v, dtype=np.float64)
        assert v.size == self.dim
        out = c_int32(-1)
        _check(load().bqa_plan_add_diagonal(self._h, _ptr(v), byref(out)))
        self.n_diags += 1
        return out.value

    def build_rydberg(self, pair: np.ndarray) -> int:
        pair = np.ascontiguousarray(pair, dtype=np.float64)
        assert pair.shape == (self.n_atoms, self.n_atoms)
        out = c_int32(-1)
        _check(load().bqa_plan_build_rydberg(self._h, _ptr(pair), byref(out)))
        self.n_diags += 1
        return out.value

    def build_rydberg_shard(self, pair: np.ndarray, rank: int) -> int:
        """diagonal of rank `rank`'s shard of a register of pair.shape[0] atoms (this plan = the local atoms)"""
        pair = np.ascontiguousarray(pair, dtype=np.float64)
        out = c_int32(-1)
        _check(load().bqa_plan_build_rydberg_shard(self._h, c_int32(pair.shape[0]), _ptr(pair), c_int32(rank), byref(out)))
        self.n_diags += 1
        return out.value

    def apply_sharded(self, record: np.ndarray, t: float, n_atoms_total: int, rank: int, partners, psi_in: int,
                      psi_out: int, raw: bool = False):
        """device pointers only; partners[b] = shard of rank ^ (1 << b)"""
        record = np.ascontiguousarray(record, dtype=np.float64)
        arr = (c_void_p * len(partners))(*[int(x) for x in partners])
        _check(load().bqa_apply_sharded(self._h, _ptr(record), c_int64(record.size), c_double(t), c_int32(int(raw)),
                                        c_int32(n_atoms_total), c_int32(rank), arr, c_void_p(int(psi_in)),
                                        c_void_p(int(psi_out))))

    def evolve_sharded(self, record: np.ndarray, opts: "SolverOpts", desc: "ShardDesc", times, psi_out: int,
                       psi0: Optional[int] = None) -> np.ndarray:
        record = np.ascontiguousarray(record, dtype=np.float64)
        times = np.ascontiguousarray(times, dtype=np.float64).reshape(-1)
        stats = np.zeros(times.size, dtype=STATS_DTYPE)
        _check(load().bqa_evolve_sharded(self._h, _ptr(record), c_int64(record.size), byref(opts), byref(desc),
                                         c_void_p(psi0), _ptr(times), c_int32(times.size), c_void_p(int(psi_out)),
                                         _ptr(stats)))
        return stats

    def get_diagonal(self, diag_id: int) -> np.ndarray:
        v = np.empty(self.dim, dtype=np.float64)
        _check(load().bqa_plan_get_diagonal(self._h, c_int32(diag_id), _ptr(v)))
        return v

    def add_csr(self, amp_wf: int, phase_wf: int, data, indices, indptr):
        data = np.ascontiguousarray(data, dtype=np.float64)
        indices = np.ascontiguousarray(indices, dtype=np.int32)
        indptr = np.ascontiguousarray(indptr, dtype=np.int64)
        _check(load().bqa_plan_add_csr(self._h, c_int32(amp_wf), c_int32(phase_wf), c_int64(data.size), _ptr(data),
                                      _ptr(indices), _ptr(indptr)))

    # ---- kernel (1) --------------------------------------------------------------------------
    def apply(self, record: np.ndarray, t: float, psi, out=None, interaction_picture=False, raw=False, mem=MEM_HOST):
        record = np.ascontiguousarray(record, dtype=np.float64)
        if mem == MEM_HOST:
            psi = np.ascontiguousarray(psi, dtype=np.complex128)
            if out is None:
                out = np.empty(self.dim, dtype=np.complex128)
        _check(load().bqa_apply(self._h, _ptr(record), c_int64(record.size), c_double(t),
                               c_int32(int(interaction_picture)), c_int32(int(raw)), _ptr(psi), _ptr(out), c_int32(mem)))
        return out

    # ---- kernels (2)+(5) ---------------------------------------------------------------------
    def evolve_batch(self, blob, offsets, times, opts: SolverOpts, psi0=None, psi0_shared=False, psi_out=None,
                     mem=MEM_HOST, n_tasks=None, n_times=None):
        if mem == MEM_HOST:
            blob = np.ascontiguousarray(blob, dtype=np.float64)
            offsets = np.ascontiguousarray(offsets, dtype=np.int64)
            times = np.ascontiguousarray(times, dtype=np.float64)
            n_tasks = offsets.size - 1
            times = times.reshape(n_tasks, -1)
            n_times = times.shape[1]
            if psi0 is not None:
                psi0 = np.ascontiguousarray(psi0, dtype=np.complex128)
            if psi_out is None:
                psi_out = np.empty((n_tasks, n_times, self.dim), dtype=np.complex128)
        stats = np.zeros((n_tasks, n_times), dtype=STATS_DTYPE)
        _check(load().bqa_evolve_batch(self._h, _ptr(blob), _ptr(offsets), c_int32(n_tasks), byref(opts), _ptr(psi0),
                                      c_int32(int(psi0_shared)), _ptr(times), c_int32(n_times), _ptr(psi_out),
                                      _ptr(stats), c_int32(mem)))
        return psi_out, stats

    def run_batch(self, blob, offsets, times, uniforms, opts: SolverOpts, project_hyperfine=True, shots_out=None,
                  psi_out=None, mem=MEM_HOST, n_tasks=None, n_shots=None):
        if mem == MEM_HOST:
            blob = np.ascontiguousarray(blob, dtype=np.float64)
            offsets = np.ascontiguousarray(offsets, dtype=np.int64)
            times = np.ascontiguousarray(times, dtype=np.float64).ravel()
            n_tasks = offsets.size - 1
            uniforms = np.ascontiguousarray(uniforms, dtype=np.float64).reshape(n_tasks, -1)
            n_shots = uniforms.shape[1]
            if shots_out is None:
                shots_out = np.empty((n_tasks, n_shots, self.n_atoms), dtype=np.uint8)
        stats = np.zeros(n_tasks, dtype=STATS_DTYPE)
        _check(load().bqa_run_batch(self._h, _ptr(blob), _ptr(offsets), c_int32(n_tasks), byref(opts), _ptr(times),
                                   _ptr(uniforms), c_int32(n_shots), c_int32(int(project_hyperfine)), _ptr(shots_out),
                                   _ptr(psi_out), _ptr(stats), c_int32(mem)))
        return shots_out, stats

    # ---- kernel (3) --------------------------------------------------------------------------
    def norm(self, psi, mem=MEM_HOST) -> float:
        if mem == MEM_HOST:
            psi = np.ascontiguousarray(psi, dtype=np.complex128)
        out = c_double(0.0)
        _check(load().bqa_norm(self._h, _ptr(psi), c_int32(mem), byref(out)))
        return out.value

    def norm_noalloc(self, psi_dev: int, n_states: int, norm_dev: int):
        """bqa_norm_noalloc: device pointers only, no allocation, no synchronisation (enqueued on the plan's stream)"""
        _check(load().bqa_norm_noalloc(self._h, c_void_p(int(psi_dev)), c_int32(int(n_states)), c_void_p(int(norm_dev))))

    def normalize(self, psi, mem=MEM_HOST):
        _check(load().bqa_normalize(self._h, _ptr(psi), c_int32(mem)))

    def expectation(self, record, t: float, psi, mem=MEM_HOST):
        record = np.ascontiguousarray(record, dtype=np.float64)
        if mem == MEM_HOST:
            psi = np.ascontiguousarray(psi, dtype=np.complex128)
        avg, var = c_double(0.0), c_double(0.0)
        _check(load().bqa_expectation(self._h, _ptr(record), c_int64(record.size), c_double(t), _ptr(psi), c_int32(mem),
                                     byref(avg), byref(var)))
        return avg.value, var.value

    def local_trace(self, psi, op: np.ndarray, sites: Sequence[int], mem=MEM_HOST) -> complex:
        op = np.ascontiguousarray(op, dtype=np.complex128)
        if mem == MEM_HOST:
            psi = np.ascontiguousarray(psi, dtype=np.complex128)
        res = np.zeros(2, dtype=np.float64)
        sa = int(sites[0])
        sb = int(sites[1]) if len(sites) > 1 else 0
        _check(load().bqa_local_trace(self._h, _ptr(psi), c_int32(mem), c_int32(len(sites)), c_int32(sa), c_int32(sb),
                                     _ptr(op), _ptr(res)))
        return complex(res[0], res[1])

    # ---- kernel (4) --------------------------------------------------------------------------
    def sample(self, psi, uniforms, project_hyperfine=True, mem=MEM_HOST, n_states=None, n_shots=None, shots_out=None):
        if mem == MEM_HOST:
            psi = np.ascontiguousarray(psi, dtype=np.complex128).reshape(-1, self.dim)
            n_states = psi.shape[0]
            uniforms = np.ascontiguousarray(uniforms, dtype=np.float64).reshape(n_states, -1)
            n_shots = uniforms.shape[1]
            shots_out = np.empty((n_states, n_shots, self.n_atoms), dtype=np.uint8)
        _check(load().bqa_sample(self._h, _ptr(psi), c_int32(n_states), _ptr(uniforms), c_int32(n_shots),
                                c_int32(int(project_hyperfine)), _ptr(shots_out), c_int32(mem)))
        return shots_out


def sample_host(state_vector, configurations, n_atoms, n_level, uniforms, project):
    """one-shot helper behind Space.sample_state_vector"""
    dim = state_vector.size
    if configurations is None:
        plan = Plan(n_atoms, n_level, dim)
    else:
        # sampling needs only the configurations; the partner table is irrelevant here
        partners = np.full((n_atoms, n_level - 1, dim), -1, dtype=np.int32)
        plan = Plan(n_atoms, n_level, dim, configurations, partners)
    try:
        return plan.sample(state_vector, uniforms, project)[0]
    finally:
        plan.close()


# ---- peer-visible device buffers of a sharded state (CUDA IPC) ------------------------------------------------
def shard_alloc(nbytes: int):
    """(device pointer, 64-byte IPC handle) of a zero-filled buffer other ranks can map"""
    ptr = c_void_p()
    handle = ctypes.create_string_buffer(64)
    _check(load().bqa_shard_alloc(c_int64(int(nbytes)), byref(ptr), handle))
    return int(ptr.value), handle.raw


def shard_open(handle: bytes) -> int:
    ptr = c_void_p()
    _check(load().bqa_shard_open(ctypes.create_string_buffer(handle, 64), byref(ptr)))
    return int(ptr.value)


def shard_close(ptr: int):
    _check(load().bqa_shard_close(c_void_p(int(ptr))))


def shard_free(ptr: int):
    _check(load().bqa_shard_free(c_void_p(int(ptr))))


def shard_reset(sync_ptr: int):
    """zero this rank's barrier counters / abort word; a host barrier must follow before bqa_evolve_sharded"""
    _check(load().bqa_shard_reset(c_void_p(int(sync_ptr))))


def shard_sync_bytes() -> int:
    return int(load().bqa_shard_sync_bytes())
