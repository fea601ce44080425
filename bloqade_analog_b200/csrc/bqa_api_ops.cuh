// bqa_api_ops.cuh - C ABI entry points that launch kernels (included at the end of bqa_api.cu)
#pragma once

static int grid_for(const bqa_plan* p, int64_t n, int threads, int per_sm = 8) {
  int64_t want = (n + threads - 1) / threads;
  int64_t cap = (int64_t)p->sm_count * per_sm;  // a multiple of the SM count
  return (int)std::max<int64_t>(1, std::min(want, cap));
}

// gridDim.y is limited to 65535: batched segment kernels are launched in chunks of this many segments
constexpr int kMaxSegPerLaunch = 65535;

// segment-wise |psi|^2 sums -> d_out[n_seg] (two deterministic passes)
static int norm2_segments(bqa_plan* p, const cplx* d_psi, int n_seg, double* d_partial, int blocks, double* d_out) {
  const int64_t dim = p->P.dim;
  for (int s0 = 0; s0 < n_seg; s0 += kMaxSegPerLaunch) {
    const int ns = std::min(kMaxSegPerLaunch, n_seg - s0);
    dim3 g(blocks, ns);
    norm2_partial_kernel<<<g, 256, 0, p->stream>>>(d_psi + (size_t)s0 * dim, dim, dim, d_partial + (size_t)s0 * blocks);
    LAUNCHED();
  }
  finish_sum_kernel<<<n_seg, 256, 0, p->stream>>>(d_partial, blocks, 1, d_out);
  LAUNCHED();
  CK(cudaGetLastError());
  return 0;
}

// Born sampling of n_seg device states; all pointers device
static int sample_device(bqa_plan* p, const cplx* d_psi, int n_seg, const double* d_uniforms, int n_shots,
                         int project, uint8_t* d_shots) {
  const int64_t dim = p->P.dim;
  const int n_tiles = (int)((dim + kScanTile - 1) / kScanTile);
  // large states with few shots: no CDF array, each shot re-evaluates one tile (sample_search_tiles_kernel)
  const bool two_level = dim >= ((int64_t)1 << 18) && (int64_t)n_shots * kScanTile * 2 <= dim;
  void* sc = nullptr;
  size_t need = sizeof(double) * ((size_t)n_seg * n_tiles + n_seg + (two_level ? 0 : (size_t)n_seg * dim));
  if (plan_scratch(p, need, &sc)) return 1;
  double* tile_sums = (double*)sc;
  double* totals = tile_sums + (size_t)n_seg * n_tiles;
  double* cdf = totals + n_seg;
  const int project_flag = (project && p->P.n_level == 3) ? 1 : 0;
  for (int s0 = 0; s0 < n_seg; s0 += kMaxSegPerLaunch) {
    const int ns = std::min(kMaxSegPerLaunch, n_seg - s0);
    dim3 g(n_tiles, ns);
    scan_tile_sums_kernel<<<g, kScanThreads, 0, p->stream>>>(d_psi + (size_t)s0 * dim, dim, dim, n_tiles,
                                                            tile_sums + (size_t)s0 * n_tiles);
    LAUNCHED();
  }
  scan_tile_offsets_kernel<<<n_seg, 1024, 0, p->stream>>>(tile_sums, n_tiles, totals);
  LAUNCHED();
  for (int s0 = 0; s0 < n_seg; s0 += kMaxSegPerLaunch) {
    const int ns = std::min(kMaxSegPerLaunch, n_seg - s0);
    const cplx* psi0 = d_psi + (size_t)s0 * dim;
    const double* uni0 = d_uniforms + (size_t)s0 * n_shots;
    uint8_t* shots0 = d_shots + (size_t)s0 * n_shots * p->P.n_atoms;
    if (two_level) {
      sample_search_tiles_kernel<<<dim3(n_shots, ns), kScanThreads, 0, p->stream>>>(
          psi0, dim, dim, n_tiles, tile_sums + (size_t)s0 * n_tiles, totals + s0, p->P.configs, p->P.n_atoms, p->P.n_level,
          project_flag, uni0, n_shots, shots0);
      LAUNCHED();
      continue;
    }
    dim3 g(n_tiles, ns);
    scan_write_cdf_kernel<<<g, kScanThreads, 0, p->stream>>>(psi0, dim, dim, n_tiles, tile_sums + (size_t)s0 * n_tiles,
                                                            totals + s0, cdf + (size_t)s0 * dim);
    LAUNCHED();
    dim3 gs((n_shots + 127) / 128, ns);
    sample_search_kernel<<<gs, 128, 0, p->stream>>>(cdf + (size_t)s0 * dim, dim, p->P.configs, p->P.n_atoms, p->P.n_level,
                                                    project_flag, uni0, n_shots, shots0);
    LAUNCHED();
  }
  CK(cudaGetLastError());
  return 0;
}

// Shared-memory-tiled H(t) psi for large two-level full spaces (bqa_apply_tiled.cuh): pass list.
// n_loc atoms are local to this device; atoms n_loc .. n_total-1 are the rank bits of a sharded state whose
// partner shards are `partners[b]` (rank ^ (1 << b)); n_total == n_loc when not sharded.
static int tiled_pass_count(int n_loc, int g, int maxfresh) {
  int k = std::max(1, g);
  while (kTileBits + maxfresh * (k - 1) + (k - g) < n_loc) ++k;
  return k;
}

static bool tiled_plan(int n_loc, int n_total, int64_t rank, const cplx* const* partners, int raw,
                       std::vector<TileSpec>& passes) {
  if (n_loc <= kTileBits) return false;
  const int g = n_total - n_loc;
  // k passes cover 13 + maxfresh (k - 1) tile bits plus ONE leftover atom each (partner amplitudes gathered
  // from global memory: the peer shards first, then local atoms).  A later pass takes 11 fresh atoms with
  // 64-byte global segments (lo_bits = 2) or 10 with full 128-byte lines (lo_bits = 3): the latter when it
  // does not cost a pass.
  int maxfresh = kTileBits - 2;
  if (!getenv("BQA_TILED_LO2") && tiled_pass_count(n_loc, g, maxfresh - 1) == tiled_pass_count(n_loc, g, maxfresh))
    --maxfresh;
  const int k = tiled_pass_count(n_loc, g, maxfresh);
  const int local_extras = std::max(0, n_loc - (kTileBits + maxfresh * (k - 1)));
  // unsharded: the leftover atom of the first pass is atom 13, i.e. the neighbouring tile; the remaining
  // leftovers are the top local atoms.  Every local leftover is walked in partner pairs (TileSpec::pair).
  const bool adj = g == 0 && local_extras > 0;
  const int top_extras = local_extras - (adj ? 1 : 0);
  const int n_tiled = n_loc - top_extras;
  const bool pairs = !getenv("BQA_TILED_NO_PAIR");
  int next = 0;
  for (int i = 0; i < k; ++i) {
    TileSpec s{};
    if (i == 0) {
      s.lo_bits = kTileBits;
      s.hb = kTileBits;
      s.first = 1;
      next = kTileBits + (adj ? 1 : 0);
    } else {
      // `fresh` new atoms; the high group is widened downwards to at least 4 bits (lo_bits <= 9) over atoms
      // already done, which p0 then skips
      // (sharing the remaining atoms evenly between the later passes of a sharded state - 13 + 7 + 7 instead of
      // 13 + 10 + 4 at 27 local atoms, 1 KB rows instead of 128-byte ones over NVLink - was measured SLOWER on 8 GPUs:
      // 13.8 against 12.9 ms at 30 atoms, profiles/r02_summary.md)
      const int fresh = std::min(n_tiled - next, maxfresh);
      const int tb = std::max(fresh, 4);
      s.lo_bits = kTileBits - tb;
      s.hb = next + fresh - tb;
      s.tb = tb;
      s.p0 = kTileBits - fresh;
      next += fresh;
    }
    if (i < g) {
      s.n_x = 1;
      s.xbit = n_loc + i;
      s.x_ptr = partners[i];
    } else if (adj && i == 0) {
      s.n_x = 1;
      s.xbit = kTileBits;
    } else if (i - g - (adj ? 1 : 0) < top_extras) {
      s.n_x = 1;
      s.xbit = n_loc - 1 - (i - g - (adj ? 1 : 0));
    }
    if (s.n_x && !s.x_ptr && pairs) {
      // tile-index bit of the leftover atom: the free index bits are [lo_bits, hb) and [hb + tb, n_loc)
      s.pair = 1;
      s.pair_b = s.xbit < s.hb ? s.xbit - s.lo_bits : (s.hb - s.lo_bits) + (s.xbit - (s.hb + s.tb));
    }
    // L2 eviction hints (TileSpec::hint): on by default, measured 0.772 -> 0.760 ms and 3.30 -> 3.07 GB at 2^25
    s.hint = getenv("BQA_TILED_HINT") ? atoi(getenv("BQA_TILED_HINT")) : 3;
    s.n_total = n_total;
    s.idx_hi = (long long)(rank << n_loc);
    s.raw = raw;
    passes.push_back(s);
  }
  passes.back().last = 1;
  return true;
}

static bool tiled_passes(const bqa_plan* p, int raw, std::vector<TileSpec>& passes) {
  const PlanView& P = p->P;
  if (!P.full_space || P.n_level != 2 || P.n_csr != 0) return false;
  if (getenv("BQA_NO_TILED")) return false;
  return tiled_plan(P.n_atoms, P.n_atoms, 0, nullptr, raw, passes);
}

typedef void (*TiledKernel)(const PlanView, const double*, const double*, double, const TileSpec, const cplx*, cplx*);

// kernel variant of one pass: real / complex coefficients, first / later pass, with / without a leftover atom
static TiledKernel tiled_kernel_for(bool real, const TileSpec& s) {
#define BQA_TILED_VARIANTS(K, R)                                                                          \
  K<R, false, false, 0>, K<R, false, true, 0>, K<R, false, false, 2>, K<R, false, true, 2>, K<R, false, false, 3>, \
      K<R, false, true, 3>, K<R, true, false, 0>, K<R, true, true, 0>
#define BQA_TILED_REMOTE(K, R) K<R, false, true, 0, true>, K<R, false, true, 2, true>, K<R, false, true, 3, true>, K<R, true, true, 0, true>
  static const TiledKernel tab[24] = {BQA_TILED_VARIANTS(apply_tiled_kernel, false),
                                      BQA_TILED_VARIANTS(apply_tiled_kernel, true),
                                      BQA_TILED_REMOTE(apply_tiled_kernel, false), BQA_TILED_REMOTE(apply_tiled_kernel, true)};
  static std::atomic<uint64_t> attr_set{0};
  if (!once_per_device(attr_set)) {
    for (TiledKernel k : tab)
      if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileDynSmem) != cudaSuccess) return nullptr;
    once_per_device(attr_set, true);
  }
  return tab[tiled_variant(real, s)];
}

// out = [-i] H(t) psi by tiled passes; `in` and `out` must not alias.  rec_host: the same record on the host
static int apply_tiled_launch(bqa_plan* p, const std::vector<TileSpec>& passes, const double* rec_host,
                              const double* d_rec, const double* V, double t, const cplx* in, cplx* out) {
  const bool real = !record_has_phase(rec_host);
  const int64_t n_tiles = p->P.dim >> kTileBits;
  const int grid = (int)std::min<int64_t>(n_tiles, p->sm_count);
  for (const TileSpec& s : passes) {
    TiledKernel kern = tiled_kernel_for(real, s);
    if (!kern) return fail("apply_tiled: cannot raise the dynamic shared memory limit");
    kern<<<grid, kTileThreads, kTileDynSmem, p->stream>>>(p->P, d_rec, V, t, s, in, out);
    LAUNCHED();
  }
  CK(cudaGetLastError());
  return 0;
}

extern "C" {

int bqa_tiled_plan_describe(int32_t n_local, int32_t n_total, int32_t* out, int32_t max_passes) {
  std::vector<TileSpec> passes;
  const cplx* const peer = reinterpret_cast<const cplx*>(16);  // stands for "a peer shard": never dereferenced
  const cplx* none[3] = {peer, peer, peer};
  if (!out || n_total < n_local || n_total - n_local > 3) return -1;
  if (!tiled_plan(n_local, n_total, 0, none, 0, passes)) return 0;
  if ((int)passes.size() > max_passes) return -1;
  for (size_t i = 0; i < passes.size(); ++i) {
    const TileSpec& s = passes[i];
    const int32_t row[8] = {s.lo_bits, s.hb, s.tb, s.first ? 0 : s.p0, s.n_x, s.xbit, s.pair, s.pair_b};
    memcpy(out + 8 * i, row, sizeof(row));
  }
  return (int)passes.size();
}

int bqa_plan_add_csr(bqa_plan* p, int32_t amp_wf, int32_t phase_wf, int64_t nnz, const double* data,
                     const int32_t* indices, const int64_t* indptr) {
  BQA_ENTER_PLAN(p);
  if (!p || !data || !indices || !indptr) return fail("bqa_plan_add_csr: null argument");
  if (p->csr.size() >= 8) return fail("bqa_plan_add_csr: at most 8 generic terms");
  const int64_t dim = p->P.dim;
  // transpose on the host (CSC of M read as CSR of M^T) so the device only ever gathers
  std::vector<int64_t> t_indptr(dim + 1, 0);
  for (int64_t j = 0; j < nnz; ++j) t_indptr[indices[j] + 1]++;
  for (int64_t i = 0; i < dim; ++i) t_indptr[i + 1] += t_indptr[i];
  std::vector<int32_t> t_indices(nnz);
  std::vector<double> t_data(nnz);
  std::vector<int64_t> fill(t_indptr.begin(), t_indptr.end() - 1);
  for (int64_t i = 0; i < dim; ++i)
    for (int64_t j = indptr[i]; j < indptr[i + 1]; ++j) {
      int64_t q = fill[indices[j]]++;
      t_indices[q] = (int32_t)i;
      t_data[q] = data[j];
    }
  CsrTerm T;
  T.amp_wf = amp_wf;
  T.phase_wf = phase_wf;
  auto up = [&](const void* src, size_t bytes, const void** dst) -> int {
    void* d = nullptr;
    CK(cudaMalloc(&d, std::max<size_t>(bytes, 8)));
    if (bytes) CK(cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice));
    p->owned.push_back(d);
    *dst = d;
    return 0;
  };
  if (up(data, sizeof(double) * nnz, (const void**)&T.data)) return 1;
  if (up(indices, sizeof(int32_t) * nnz, (const void**)&T.indices)) return 1;
  if (up(indptr, sizeof(int64_t) * (dim + 1), (const void**)&T.indptr)) return 1;
  if (up(t_data.data(), sizeof(double) * nnz, (const void**)&T.t_data)) return 1;
  if (up(t_indices.data(), sizeof(int32_t) * nnz, (const void**)&T.t_indices)) return 1;
  if (up(t_indptr.data(), sizeof(int64_t) * (dim + 1), (const void**)&T.t_indptr)) return 1;
  p->csr.push_back(T);
  if (p->d_csr) cudaFree(p->d_csr);
  CK(cudaMalloc(&p->d_csr, sizeof(CsrTerm) * p->csr.size()));
  CK(cudaMemcpy(p->d_csr, p->csr.data(), sizeof(CsrTerm) * p->csr.size(), cudaMemcpyHostToDevice));
  p->P.csr = p->d_csr;
  p->P.n_csr = (int)p->csr.size();
  return 0;
}

int bqa_apply(bqa_plan* p, const double* record, int64_t record_len, double t, int32_t interaction, int32_t raw,
              const void* psi_in, void* psi_out, int32_t mem) {
  BQA_ENTER_PLAN(p);
  if (!p || !record || !psi_in || !psi_out) return fail("bqa_apply: null argument");
  const int64_t dim = p->P.dim;
  Staged rec, in, out;
  if (rec.in(record, sizeof(double) * record_len, BQA_MEM_HOST, p->stream)) return 1;
  if (in.in(psi_in, sizeof(cplx) * dim, mem, p->stream)) return 1;
  if (out.out(psi_out, sizeof(cplx) * dim, mem)) return 1;
  const int diag_id = (int)record[1];
  const double* V = (diag_id >= 0 && diag_id < (int)p->diags.size()) ? p->diags[diag_id] : nullptr;
  const cplx* src = (const cplx*)in.dev;
  Staged rot;
  if (interaction && V) {  // phi = u psi (SV:347-349)
    if (rot.out((void*)1, sizeof(cplx) * dim, BQA_MEM_HOST)) return 1;
    rot.host = nullptr;
    rotate_kernel<<<grid_for(p, dim, 256), 256, 0, p->stream>>>(dim, V, t, src, (cplx*)rot.dev);
    LAUNCHED();
    src = (const cplx*)rot.dev;
  }
  const int mode = (raw ? 1 : 0) | (interaction ? 2 : 0);
  std::vector<TileSpec> passes;
  if (!interaction && src != (const cplx*)out.dev && tiled_passes(p, raw ? 1 : 0, passes)) {
    if (apply_tiled_launch(p, passes, record, (const double*)rec.dev, V, t, src, (cplx*)out.dev)) return 1;
  } else {
    apply_kernel<<<grid_for(p, dim, kApplyThreads), kApplyThreads, 0, p->stream>>>(
        p->P, (const double*)rec.dev, V, nullptr, t, mode, src, (cplx*)out.dev, nullptr);
    LAUNCHED();
  }
  CK(cudaGetLastError());
  if (out.flush(p->stream)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

int bqa_norm(bqa_plan* p, const void* psi, int32_t mem, double* norm) {
  BQA_ENTER_PLAN(p);
  if (!p || !psi || !norm) return fail("bqa_norm: null argument");
  const int64_t dim = p->P.dim;
  Staged in;
  if (in.in(psi, sizeof(cplx) * dim, mem, p->stream)) return 1;
  const int blocks = grid_for(p, dim, 256, 4);
  void* sc = nullptr;
  if (plan_scratch(p, sizeof(double) * (blocks + 1), &sc)) return 1;
  double* partial = (double*)sc;
  if (norm2_segments(p, (const cplx*)in.dev, 1, partial, blocks, partial + blocks)) return 1;
  double n2 = 0.0;
  CK(cudaMemcpyAsync(&n2, partial + blocks, sizeof(double), cudaMemcpyDeviceToHost, p->stream));
  CK(cudaStreamSynchronize(p->stream));
  *norm = sqrt(n2);
  return 0;
}

int bqa_normalize(bqa_plan* p, void* psi, int32_t mem) {
  BQA_ENTER_PLAN(p);
  if (!p || !psi) return fail("bqa_normalize: null argument");
  const int64_t dim = p->P.dim;
  Staged buf;
  if (buf.in(psi, sizeof(cplx) * dim, mem, p->stream)) return 1;
  if (mem == BQA_MEM_HOST) buf.host = psi;
  const int blocks = grid_for(p, dim, 256, 4);
  void* sc = nullptr;
  if (plan_scratch(p, sizeof(double) * (blocks + 1), &sc)) return 1;
  double* partial = (double*)sc;
  if (norm2_segments(p, (const cplx*)buf.dev, 1, partial, blocks, partial + blocks)) return 1;
  scale_by_norm_kernel<<<dim3(blocks, 1), 256, 0, p->stream>>>((cplx*)buf.dev, dim, dim, partial + blocks);
  LAUNCHED();
  CK(cudaGetLastError());
  if (buf.flush(p->stream)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

int bqa_expectation(bqa_plan* p, const double* record, int64_t record_len, double t, const void* psi, int32_t mem,
                    double* average, double* variance) {
  BQA_ENTER_PLAN(p);
  if (!p || !record || !psi) return fail("bqa_expectation: null argument");
  const int64_t dim = p->P.dim;
  Staged rec, in;
  if (rec.in(record, sizeof(double) * record_len, BQA_MEM_HOST, p->stream)) return 1;
  if (in.in(psi, sizeof(cplx) * dim, mem, p->stream)) return 1;
  const int diag_id = (int)record[1];
  const double* V = (diag_id >= 0 && diag_id < (int)p->diags.size()) ? p->diags[diag_id] : nullptr;
  const int blocks = grid_for(p, dim, kApplyThreads, 4);
  void* sc = nullptr;
  size_t need = sizeof(cplx) * dim + sizeof(double) * (3 * blocks + 3);
  if (plan_scratch(p, need, &sc)) return 1;
  cplx* hpsi = (cplx*)sc;
  double* dots = (double*)(hpsi + dim);
  apply_kernel<<<blocks, kApplyThreads, 0, p->stream>>>(p->P, (const double*)rec.dev, V, nullptr, t, 1,
                                                        (const cplx*)in.dev, hpsi, dots);
  LAUNCHED();
  finish_sum_kernel<<<1, 256, 0, p->stream>>>(dots, blocks, 3, dots + 3 * blocks);
  LAUNCHED();
  CK(cudaGetLastError());
  double r[3];
  CK(cudaMemcpyAsync(r, dots + 3 * blocks, sizeof(r), cudaMemcpyDeviceToHost, p->stream));
  CK(cudaStreamSynchronize(p->stream));
  if (average) *average = r[0];
  if (variance) *variance = r[2] - r[0] * r[0];
  return 0;
}

int bqa_local_trace(bqa_plan* p, const void* psi, int32_t mem, int32_t n_sites, int32_t site_a, int32_t site_b,
                    const double* op, double* result) {
  BQA_ENTER_PLAN(p);
  if (!p || !psi || !op || !result) return fail("bqa_local_trace: null argument");
  if (n_sites != 1 && n_sites != 2) return fail("bqa_local_trace: n_sites must be 1 or 2");
  if (site_a < 0 || site_a >= p->P.n_atoms || (n_sites == 2 && (site_b < 0 || site_b >= p->P.n_atoms)))
    return fail("bqa_local_trace: site out of range");
  const int64_t dim = p->P.dim;
  const int m = n_sites == 1 ? p->P.n_level : p->P.n_level * p->P.n_level;
  Staged in, dop;
  if (in.in(psi, sizeof(cplx) * dim, mem, p->stream)) return 1;
  if (dop.in(op, sizeof(cplx) * m * m, BQA_MEM_HOST, p->stream)) return 1;
  const int blocks = grid_for(p, dim, 256, 4);
  void* sc = nullptr;
  if (plan_scratch(p, sizeof(double) * (2 * blocks + 2), &sc)) return 1;
  double* partial = (double*)sc;
  local_trace_kernel<<<blocks, 256, 0, p->stream>>>(p->P, (const cplx*)in.dev, n_sites, site_a, site_b,
                                                    (const cplx*)dop.dev, partial);
  LAUNCHED();
  finish_sum_kernel<<<1, 256, 0, p->stream>>>(partial, blocks, 2, partial + 2 * blocks);
  LAUNCHED();
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(result, partial + 2 * blocks, 2 * sizeof(double), cudaMemcpyDeviceToHost, p->stream));
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

int bqa_sample(bqa_plan* p, const void* psi, int32_t n_states, const double* uniforms, int32_t n_shots,
               int32_t project, uint8_t* shots, int32_t mem) {
  BQA_ENTER_PLAN(p);
  if (!p || !psi || !uniforms || !shots) return fail("bqa_sample: null argument");
  const int64_t dim = p->P.dim;
  Staged in, un, out;
  if (in.in(psi, sizeof(cplx) * dim * n_states, mem, p->stream)) return 1;
  if (un.in(uniforms, sizeof(double) * (size_t)n_states * n_shots, mem, p->stream)) return 1;
  if (out.out(shots, (size_t)n_states * n_shots * p->P.n_atoms, mem)) return 1;
  if (sample_device(p, (const cplx*)in.dev, n_states, (const double*)un.dev, n_shots, project, (uint8_t*)out.dev))
    return 1;
  if (out.flush(p->stream)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

}  // extern "C"

#include "bqa_api_evolve.cuh"
