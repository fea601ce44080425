// bqa_api_shard.cuh - C ABI of the SHARDED full-space state (BASELINE config 5): a
// register of n atoms whose 2^n amplitudes do not fit one GPU.  Rank r of `world` = 2^g ranks holds the
// amplitudes whose top g index bits equal r.  Diagonal terms and flips of the n - g local atoms never leave
// the GPU; a flip of a "rank bit" atom reads the partner rank's shard IN PLACE over NVLink peer access
// (CUDA IPC mapping) inside the same tiled H psi kernel, and the integrator's cross-rank reductions and
// barriers are device initiated too (bqa_large.cuh: lg_barrier, lg_total3).  No reference counterpart: the
// reference holds one NumPy array (emulate/ir/state_vector.py:88-90) and cannot run this size.
#pragma once

static const int64_t kShardSyncBytes = 64 * 1024;  // [flags: 8 x u64][partial sums: f64 ...]
static const int64_t kShardFlagBytes = 64;

static int shard_check(const bqa_plan* p, const bqa_shard_desc* d) {
  if (!p || !d) return fail("shard: null argument");
  if (d->world < 2 || d->world > 8 || (d->world & (d->world - 1))) return fail("shard: world must be 2, 4 or 8");
  if (d->rank < 0 || d->rank >= d->world) return fail("shard: bad rank");
  int g = 0;
  while ((1 << g) < d->world) ++g;
  if (d->n_atoms_total != p->P.n_atoms + g) return fail("shard: plan must hold n_atoms_total - log2(world) atoms");
  if (!p->P.full_space || p->P.n_level != 2) return fail("shard: two-level full space only");
  for (int r = 0; r < d->world; ++r)
    if (!d->ys[r] || !d->sync[r]) return fail("shard: missing peer mapping of rank %d", r);
  return 0;
}

// van-der-Waals diagonal of this rank's shard: index bits above the local atoms come from `rank`
__global__ void build_rydberg_shard_kernel(int64_t dim, int n_total, uint32_t hi, const double* __restrict__ pair, double* V) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t mask = (uint32_t)i | hi;
    double v = 0.0;
    for (int a = 0; a < n_total; ++a) {
      if (!((mask >> a) & 1u)) continue;
      for (int b = a + 1; b < n_total; ++b) {
        const double c = pair[a * n_total + b];
        if (((mask >> b) & 1u) && c != 0.0) v = __dadd_rn(v, c);
      }
    }
    V[i] = v;
  }
}

extern "C" {

int64_t bqa_shard_sync_bytes(void) { return kShardSyncBytes; }

int bqa_shard_alloc(int64_t bytes, void** dev_ptr, void* handle64) {
  BQA_ENTER();
  if (!dev_ptr || !handle64 || bytes <= 0) return fail("bqa_shard_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* d = nullptr;
  CK(cudaMalloc(&d, (size_t)bytes));
  CK(cudaMemset(d, 0, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, d);
  if (e != cudaSuccess) {
    cudaFree(d);
    return fail("cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  memcpy(handle64, &h, 64);
  *dev_ptr = d;
  return 0;
}

int bqa_shard_open(const void* handle64, void** peer_ptr) {
  BQA_ENTER();
  if (!handle64 || !peer_ptr) return fail("bqa_shard_open: null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  CK(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int bqa_shard_close(void* peer_ptr) {
  BQA_ENTER();
  CK(cudaIpcCloseMemHandle(peer_ptr));
  return 0;
}

int bqa_shard_free(void* dev_ptr) {
  BQA_ENTER();
  CK(cudaFree(dev_ptr));
  return 0;
}

int bqa_shard_reset(void* sync_own) {
  BQA_ENTER();
  if (!sync_own) return fail("bqa_shard_reset: null argument");
  CK(cudaMemset(sync_own, 0, (size_t)kShardFlagBytes));  // barrier counter, barrier count, abort word
  CK(cudaDeviceSynchronize());
  return 0;
}

int bqa_plan_build_rydberg_shard(bqa_plan* p, int32_t n_total, const double* pair_host, int32_t rank, int32_t* diag_id) {
  BQA_ENTER_PLAN(p);
  if (!p || !pair_host || !diag_id) return fail("bqa_plan_build_rydberg_shard: null argument");
  if (!p->P.full_space || p->P.n_level != 2 || n_total < p->P.n_atoms || n_total > 31)
    return fail("bqa_plan_build_rydberg_shard: two-level full space, n_total <= 31");
  double *d_pair = nullptr, *V = nullptr;
  CK(cudaMalloc(&d_pair, sizeof(double) * n_total * n_total));
  CK(cudaMemcpy(d_pair, pair_host, sizeof(double) * n_total * n_total, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&V, sizeof(double) * p->P.dim));
  build_rydberg_shard_kernel<<<grid_for(p, p->P.dim, 256), 256, 0, p->stream>>>(
      p->P.dim, n_total, (uint32_t)rank << p->P.n_atoms, d_pair, V);
  LAUNCHED();
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(p->stream));
  cudaFree(d_pair);
  p->diags.push_back(V);
  *diag_id = (int32_t)p->diags.size() - 1;
  return refresh_diag_table(p);
}

// out_local = [-i] (H(t) psi)_local.  partners[b] = device pointer of the shard of rank ^ (1 << b) (peer
// mapped, or any device buffer: the single-GPU tests pass the other shard directly).  No barrier inside:
// the caller guarantees the partner shards are complete.  Device memory only.
int bqa_apply_sharded(bqa_plan* p, const double* record, int64_t record_len, double t, int32_t raw,
                      int32_t n_atoms_total, int32_t rank, const void* const* partners, const void* psi_in,
                      void* psi_out) {
  BQA_ENTER_PLAN(p);
  if (!p || !record || !partners || !psi_in || !psi_out) return fail("bqa_apply_sharded: null argument");
  const int g = n_atoms_total - p->P.n_atoms;
  if (g < 1 || g > 3 || rank < 0 || rank >= (1 << g)) return fail("bqa_apply_sharded: bad shard geometry");
  if (!p->P.full_space || p->P.n_level != 2 || p->P.n_csr != 0) return fail("bqa_apply_sharded: two-level full space only");
  Staged rec;
  if (rec.in(record, sizeof(double) * record_len, BQA_MEM_HOST, p->stream)) return 1;
  const int diag_id = (int)record[1];
  const double* V = (diag_id >= 0 && diag_id < (int)p->diags.size()) ? p->diags[diag_id] : nullptr;
  const cplx* pp[3] = {nullptr, nullptr, nullptr};
  for (int b = 0; b < g; ++b) pp[b] = (const cplx*)partners[b];
  std::vector<TileSpec> passes;
  if (!tiled_plan(p->P.n_atoms, n_atoms_total, rank, pp, raw ? 1 : 0, passes))
    return fail("bqa_apply_sharded: needs at least 14 local atoms");
  if (apply_tiled_launch(p, passes, record, (const double*)rec.dev, V, t, (const cplx*)psi_in, (cplx*)psi_out)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

// Adaptive evolution of ONE sharded state (dop853 / dopri5, Schroedinger picture): every rank calls this
// with the same record, options and times.  psi0_local: device pointer of this rank's shard of the initial
// state, or null for |0...0>.  psi_out_local: device, [n_times][local dim].  stats: host, [n_times].
int bqa_evolve_sharded(bqa_plan* p, const double* record, int64_t record_len, const bqa_solver_opts* o,
                       const bqa_shard_desc* d, const void* psi0_local, const double* times, int32_t n_times,
                       void* psi_out_local, bqa_task_stats* stats) {
  BQA_ENTER_PLAN(p);
  if (!record || !times || n_times < 1) return fail("bqa_evolve_sharded: bad argument");
  if (check_opts(o)) return 1;
  if (shard_check(p, d)) return 1;
  if (o->interaction_picture) return fail("bqa_evolve_sharded: interaction picture not supported");
  int g = 0;
  while ((1 << g) < d->world) ++g;
  LgShard sh;
  memset(&sh, 0, sizeof(sh));
  sh.world = d->world;
  sh.rank = d->rank;
  sh.n_total = d->n_atoms_total;
  sh.ys = (cplx*)d->ys[d->rank];
  for (int b = 0; b < g; ++b) sh.partners[b] = (const cplx*)d->ys[d->rank ^ (1 << b)];
  for (int r = 0; r < d->world; ++r) {
    sh.peer_flag[r] = (unsigned long long*)d->sync[r];
    sh.peer_partial[r] = (double*)((char*)d->sync[r] + kShardFlagBytes);
  }
  sh.partial = sh.peer_partial[d->rank];
  if ((int64_t)sizeof(double) * 4 * p->sm_count * 4 + kShardFlagBytes > kShardSyncBytes) return fail("sync block too small");
  Staged rec, st;
  if (rec.in(record, sizeof(double) * record_len, BQA_MEM_HOST, p->stream)) return 1;
  if (st.out(stats, sizeof(TaskStats) * (size_t)n_times, BQA_MEM_HOST)) return 1;
  std::vector<int64_t> h_off = {0, record_len};
  std::vector<double> h_times(times, times + n_times);
  if (evolve_large_device(p, (const double*)rec.dev, h_off, record, 1, o, (const cplx*)psi0_local, psi0_local ? 1 : 0,
                          h_times.data(), n_times, (cplx*)psi_out_local, (TaskStats*)st.dev, &sh))
    return 1;
  if (st.flush(p->stream)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

}  // extern "C"
