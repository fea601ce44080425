"""One full-space state sharded over the GPUs of a node (BASELINE config 5).

No reference counterpart: the reference keeps the state in one NumPy array (`emulate/ir/state_vector.py:88-90`)
and its CSR operators would need ~515 GB at 30 atoms (`emulate/codegen/hamiltonian.py:161`).  Layout: atom k is
index bit k (`emulate/ir/atom_type.py:139,151`); rank r of `world` = 2^g processes (one per GPU) owns the
amplitudes whose top g index bits equal r.  Everything on the data path is device initiated: the tiled H psi
kernel reads the partner rank's stage input in place over NVLink peer access (CUDA IPC mapping), and the
integrator's cross-rank sums and barriers are kernels too (`csrc/bqa_large.cuh`).  `torch.distributed` only
carries the 64-byte IPC handles at set-up and a few scalars (norms, sampling prefix sums).
"""
from __future__ import annotations

import json
import os
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from .ir import EmulatorProgram
from .lowering import pack_task_record, pair_interactions


def _log2(world: int) -> int:
    g = int(world).bit_length() - 1
    if world < 2 or (1 << g) != world or world > 8:
        raise ValueError("world must be 2, 4 or 8")
    return g


def partner_ranks(rank: int, world: int) -> List[int]:
    """partner of `rank` for every rank bit b: the shard that a flip of atom n_local + b reads"""
    return [rank ^ (1 << b) for b in range(_log2(world))]


def split_sample_uniforms(uniforms: np.ndarray, shard_weights: Sequence[float]):
    """Born sampling over shards: shot j with uniform u_j lands on the rank whose cumulative weight interval
    contains u_j * total, and is resolved there with the rescaled uniform.  Returns (owner[j], local_u[j])."""
    w = np.asarray(shard_weights, dtype=np.float64)
    edges = np.concatenate([[0.0], np.cumsum(w)])
    x = np.asarray(uniforms, dtype=np.float64) * edges[-1]
    owner = np.clip(np.searchsorted(edges, x, side="right") - 1, 0, len(w) - 1)
    local = (x - edges[owner]) / np.where(w[owner] > 0, w[owner], 1.0)
    return owner.astype(np.int64), np.clip(local, 0.0, np.nextafter(1.0, 0.0))


def checkpoint_grid(duration: float, every: float):
    """output times of evolve_checkpointed: every, 2 every, ... and the program duration itself as the last one
    (a duration that is a whole multiple of `every` up to rounding must not leave a last segment of one ulp)"""
    n = max(1, int(np.ceil(duration / every - 1e-9)))
    return [every * k for k in range(1, n)] + [duration]


class ShardedHamiltonian:
    """This rank's shard of `RydbergHamiltonian` for a two-level full-space program."""

    def __init__(self, program: EmulatorProgram, world: Optional[int] = None, rank: Optional[int] = None,
                 device: Optional[int] = None, group=None):
        import torch
        import torch.distributed as dist

        self._torch, self._dist, self.group = torch, dist, group
        self.world = dist.get_world_size(group) if world is None else int(world)
        self.rank = dist.get_rank(group) if rank is None else int(rank)
        self.g = _log2(self.world)
        register = program.register
        if register.atom_type.n_level != 2 or float(register.blockade_radius) != 0.0:
            raise ValueError("sharded states: two-level atoms, full space (blockade_radius = 0) only")
        self.program = program
        self.n_atoms = len(register.sites)
        self.n_local = self.n_atoms - self.g
        if self.n_local < 14:
            raise ValueError("sharded states need at least 14 local atoms per rank")
        self.local_dim = 1 << self.n_local
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.plan = _lib.Plan(self.n_local, 2, self.local_dim, device=self.device)
        self.diag_id = self.plan.build_rydberg_shard(pair_interactions(register), self.rank)
        self.record = pack_task_record(program, self.diag_id)
        # peer-visible buffers: stage input ys and the sync block (barrier counters + reduction partials)
        self._ys, ys_handle = _lib.shard_alloc(16 * self.local_dim)
        self._sync, sync_handle = _lib.shard_alloc(_lib.shard_sync_bytes())
        handles: List = [None] * self.world
        dist.all_gather_object(handles, (ys_handle, sync_handle), group=group)
        self._opened: List[int] = []
        self.desc = _lib.ShardDesc(self.world, self.rank, self.n_atoms, 0)
        for r in range(self.world):
            if r == self.rank:
                ys_ptr, sync_ptr = self._ys, self._sync
            else:
                ys_ptr, sync_ptr = _lib.shard_open(handles[r][0]), _lib.shard_open(handles[r][1])
                self._opened += [ys_ptr, sync_ptr]
            self.desc.ys[r] = ys_ptr
            self.desc.sync[r] = sync_ptr
        dist.barrier(group=group)

    # ---- helpers -----------------------------------------------------------------------------------------------
    def _ys_tensor(self):
        """this rank's peer-visible stage-input buffer as a torch tensor (zero copy)"""
        torch = self._torch

        class _Holder:
            pass

        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (self.local_dim, 2), "typestr": "<f8", "data": (self._ys, False),
                                      "version": 2, "strides": None}
        self._ys_holder = h
        return torch.as_tensor(h, device=f"cuda:{self.device}")

    def zero_state(self):
        torch = self._torch
        psi = torch.zeros(self.local_dim, dtype=torch.complex128, device=f"cuda:{self.device}")
        if self.rank == 0:
            psi[0] = 1.0
        return psi

    # ---- RydbergHamiltonian._apply (state_vector.py:375-398) on the shard -----------------------------------------
    def apply(self, psi_local, t: float, raw: bool = True):
        """(H(t) psi)_local (raw) or (-i H(t) psi)_local; collective: every rank calls it"""
        torch, dist = self._torch, self._dist
        ys = self._ys_tensor()
        ys.copy_(torch.view_as_real(psi_local.contiguous()))
        torch.cuda.synchronize()
        dist.barrier(group=self.group)  # every shard is in its peer-visible buffer
        out = torch.empty_like(psi_local)
        partners = [self.desc.ys[r] for r in partner_ranks(self.rank, self.world)]
        self.plan.apply_sharded(self.record, float(t), self.n_atoms, self.rank, partners, self._ys, out.data_ptr(), raw=raw)
        dist.barrier(group=self.group)  # peers are done reading this rank's buffer
        return out

    # ---- AnalogGate._apply (state_vector.py:574-602) ---------------------------------------------------------------
    def evolve(self, times: Sequence[float] = (), psi0_local=None, solver_name: str = "dop853", atol: float = 1e-7,
               rtol: float = 1e-14, nsteps: int = 2_147_483_647, reject_mode: int = 1, t_start: float = 0.0):
        """returns (states [n_times, local_dim] complex128 on this GPU, stats); collective.  `t_start` is the time of
        `psi0_local` (resume from a checkpoint, see `evolve_checkpointed`)"""
        torch, dist = self._torch, self._dist
        times = np.asarray(list(times) if len(times) else [self.program.duration], dtype=np.float64)
        opts = _lib.make_opts(solver_name, atol, rtol, nsteps, False, reject_mode, "streamed", t_start)
        out = torch.empty((times.size, self.local_dim), dtype=torch.complex128, device=f"cuda:{self.device}")
        p0 = None if psi0_local is None else int(psi0_local.contiguous().data_ptr())
        _lib.shard_reset(self._sync)  # barrier counters / abort word back to zero, ranks in step again
        dist.barrier(group=self.group)
        stats = self.plan.evolve_sharded(self.record, opts, self.desc, times, out.data_ptr(), p0)
        dist.barrier(group=self.group)
        return out, stats

    # ---- mid-run checkpoints (the reference only serialises finished batches, serialize.py:115-213) --
    def checkpoint_path(self, directory: str) -> str:
        return os.path.join(directory, f"shard_{self.rank:02d}_of_{self.world:02d}.npy")

    def save_checkpoint(self, directory: str, t: float, psi_local) -> None:
        """this rank's shard -> `directory`/shard_RR_of_WW.npy, then (rank 0, after every rank has written) meta.json;
        a resume only trusts a directory whose meta.json exists, so a crash mid-write leaves the previous one valid"""
        os.makedirs(directory, exist_ok=True)
        tmp = self.checkpoint_path(directory) + ".tmp.npy"
        np.save(tmp, psi_local.detach().cpu().numpy())
        os.replace(tmp, self.checkpoint_path(directory))
        self._dist.barrier(group=self.group)
        if self.rank == 0:
            meta = {"t": float(t), "world": self.world, "n_atoms": self.n_atoms, "local_dim": self.local_dim,
                    "duration": float(self.program.duration)}
            with open(os.path.join(directory, "meta.json.tmp"), "w") as f:
                json.dump(meta, f)
            os.replace(os.path.join(directory, "meta.json.tmp"), os.path.join(directory, "meta.json"))
        self._dist.barrier(group=self.group)

    def load_checkpoint(self, directory: str):
        """(t, psi_local on this GPU) of the checkpoint in `directory`, or None when there is no complete one"""
        meta_file = os.path.join(directory, "meta.json")
        if not os.path.exists(meta_file):
            return None
        with open(meta_file) as f:
            meta = json.load(f)
        if meta["world"] != self.world or meta["n_atoms"] != self.n_atoms:
            raise ValueError("checkpoint was written for another register / number of GPUs")
        psi = self._torch.from_numpy(np.load(self.checkpoint_path(directory))).to(f"cuda:{self.device}")
        return float(meta["t"]), psi

    def checkpoint_grid(self, every: float):
        return checkpoint_grid(float(self.program.duration), every)

    def evolve_checkpointed(self, directory: str, every: float, **solver):
        """AnalogGate._apply to the end of the program with a checkpoint every `every` microseconds of program time;
        resumes from `directory` when it holds a complete checkpoint.  SciPy's ode restarts the integrator at every
        output time (state_vector.py:589-600), so the segments ARE the reference's flow for times = the checkpoint
        grid: an interrupted and resumed run ends bit for bit like an uninterrupted one with the same grid.
        Returns (final shard, list of per-segment stats)."""
        grid = self.checkpoint_grid(every)
        t, psi, all_stats = 0.0, None, []
        loaded = self.load_checkpoint(directory)
        if loaded is not None:
            t, psi = loaded
        for t_next in grid:
            if t_next <= t:
                continue
            states, stats = self.evolve([t_next], psi0_local=psi, t_start=t, **solver)
            all_stats.append(stats)
            if int(stats["status"][0]) != 1:
                raise RuntimeError(f"sharded evolve failed with status {int(stats['status'][0])} in ({t}, {t_next}]")
            t, psi = t_next, states[0]
            self.save_checkpoint(directory, t, psi)
        return psi, all_stats

    # ---- StateVector.norm / sample (state_vector.py:210-217, space.py:211-229) ----------------------------------------
    def norm2_shards(self, psi_local) -> np.ndarray:
        """|psi_r|^2 of every rank's shard (same array on every rank)"""
        mine = float(self.plan.norm(int(psi_local.data_ptr()), mem=_lib.MEM_DEVICE)) ** 2
        parts: List = [None] * self.world
        self._dist.all_gather_object(parts, mine, group=self.group)
        return np.asarray(parts, dtype=np.float64)

    def norm(self, psi_local) -> float:
        return float(np.sqrt(self.norm2_shards(psi_local).sum()))

    def sample(self, psi_local, shots: int, uniforms: Optional[np.ndarray] = None) -> np.ndarray:
        """[shots, n_atoms] uint8 Rydberg occupations (1 = Rydberg), same on every rank; uniforms as drawn by
        `np.random.random_sample(shots)` on every rank with the same seed"""
        if uniforms is None:
            uniforms = np.random.random_sample(int(shots))
        owner, local_u = split_sample_uniforms(uniforms, self.norm2_shards(psi_local))
        mine = np.nonzero(owner == self.rank)[0]
        rows = np.zeros((len(mine), self.n_atoms), dtype=np.uint8)
        if len(mine):
            torch = self._torch
            dev = f"cuda:{self.device}"
            u_dev = torch.as_tensor(np.ascontiguousarray(local_u[mine]), device=dev)
            shots_dev = torch.empty((len(mine), self.n_local), dtype=torch.uint8, device=dev)
            self.plan.sample(int(psi_local.data_ptr()), int(u_dev.data_ptr()), mem=_lib.MEM_DEVICE, n_states=1,
                             n_shots=len(mine), shots_out=int(shots_dev.data_ptr()))
            rows[:, : self.n_local] = shots_dev.cpu().numpy()
            for b in range(self.g):
                rows[:, self.n_local + b] = (self.rank >> b) & 1
        parts: List = [None] * self.world
        self._dist.all_gather_object(parts, (mine, rows), group=self.group)
        out = np.zeros((int(shots), self.n_atoms), dtype=np.uint8)
        for idx, r in parts:
            out[idx] = r
        return out

    def close(self):
        for p in self._opened:
            _lib.shard_close(p)
        self._opened = []
        if getattr(self, "_ys", None):
            _lib.shard_free(self._ys)
            _lib.shard_free(self._sync)
            self._ys = self._sync = None
        self.plan.close()
