// bqa_kernels.cuh - streaming (multi-CTA) kernels of the B200 emulator backend.
//
//   apply_kernel          north_star kernel (1): fused out = [-i] H(t) psi, all terms in one pass,
//                         coefficients evaluated on-device from the task record
//                         (state_vector.py:323-359,375-398; sparse_operator.py:15-34,99-128)
//   reduce kernels        kernel (3): norm, <psi|H psi>, <H psi|H psi>, local_trace
//                         (state_vector.py:210-217,401-463,30-80)
//   scan / search kernels kernel (4): Born sampling by prefix-sum CDF + binary search
//                         (space.py:211-229)
//   build_rydberg_kernel  on-device van-der-Waals diagonal, same pair order as hamiltonian.py:76-89
#pragma once
#include <cuda_runtime.h>

#include "bqa_core.cuh"

namespace bqa {

constexpr int kApplyThreads = 256;

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < nw; ++w) t += red[w];
  return t;
}

// ---- kernel (1), generic form ------------------------------------------------------------------
// mode bit 0: raw (no -i factor); bit 1: interaction picture.  When `dots` != nullptr the kernel also
// accumulates per-CTA partial sums of Re<psi|out>, Im<psi|out>, <out|out> (fused expectation
// epilogue for RydbergHamiltonian.average / variance, SV:401-463) into dots[blockIdx.x][3].
__global__ void __launch_bounds__(kApplyThreads)
apply_kernel(const PlanView P, const double* __restrict__ rec, const double* __restrict__ V, const double* t_ptr,
             double t_val, int mode, const cplx* __restrict__ in, cplx* __restrict__ out, double* dots) {
  __shared__ Coefs cf;
  __shared__ double red[32];
  const double t = t_ptr ? *t_ptr : t_val;
  const TaskView tv = task_view(rec);
  const int ne = tv.n_atoms * (P.n_level - 1);
  for (int e = threadIdx.x; e < ne + P.n_csr; e += blockDim.x) {
    if (e < ne)
      coef_entry(tv, t, e, cf);
    else
      coef_csr_entry(P, tv, t, e - ne, cf);
  }
  __syncthreads();
  const bool raw = mode & 1, inter = (mode & 2) != 0;
  double d_re = 0.0, d_im = 0.0, d_hh = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.dim; i += (int64_t)gridDim.x * blockDim.x) {
    cplx hz = h_apply_dispatch(P, cf, inter ? nullptr : V, in, i);
    if (inter && V) hz = cmul_conj_a(phase_u(V[i], t), hz);
    if (dots) {
      cplx z = in[i];
      d_re += z.re * hz.re + z.im * hz.im;   // conj(psi) * Hpsi
      d_im += z.re * hz.im - z.im * hz.re;
      d_hh += hz.re * hz.re + hz.im * hz.im;
    }
    out[i] = raw ? hz : times_minus_i(hz);
  }
  if (dots) {
    d_re = block_sum(d_re, red);
    d_im = block_sum(d_im, red);
    d_hh = block_sum(d_hh, red);
    if (threadIdx.x == 0) {
      dots[blockIdx.x * 3 + 0] = d_re;
      dots[blockIdx.x * 3 + 1] = d_im;
      dots[blockIdx.x * 3 + 2] = d_hh;
    }
  }
}

// rotate into the interaction frame: out = exp(-i V t) * in  (SV:347-349, 636-638)
__global__ void rotate_kernel(int64_t dim, const double* __restrict__ V, double t, const cplx* __restrict__ in,
                              cplx* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = cmul(phase_u(V[i], t), in[i]);
}

// ---- kernel (3): reductions ----------------------------------------------------------------------
// partial[blockIdx.y][blockIdx.x] = sum over this CTA's slice of segment blockIdx.y of |psi|^2
__global__ void __launch_bounds__(256)
norm2_partial_kernel(const cplx* __restrict__ psi, int64_t dim, int64_t seg_stride, double* partial) {
  __shared__ double red[32];
  const cplx* p = psi + (int64_t)blockIdx.y * seg_stride;
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    cplx z = p[i];
    s += z.re * z.re + z.im * z.im;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) partial[(int64_t)blockIdx.y * gridDim.x + blockIdx.x] = s;
}

// out[seg][c] = sum_b partial[seg][b][c]   (deterministic second pass, one CTA per segment)
__global__ void __launch_bounds__(256)
finish_sum_kernel(const double* __restrict__ partial, int n_blocks, int n_comp, double* out) {
  __shared__ double red[32];
  for (int c = 0; c < n_comp; ++c) {
    double s = 0.0;
    for (int b = threadIdx.x; b < n_blocks; b += blockDim.x) s += partial[((int64_t)blockIdx.x * n_blocks + b) * n_comp + c];
    s = block_sum(s, red);
    if (threadIdx.x == 0) out[(int64_t)blockIdx.x * n_comp + c] = s;
  }
}

// psi[seg] *= 1/sqrt(norm2[seg])   (StateVector.normalize, SV:210-213)
__global__ void scale_by_norm_kernel(cplx* psi, int64_t dim, int64_t seg_stride, const double* __restrict__ norm2) {
  cplx* p = psi + (int64_t)blockIdx.y * seg_stride;
  const double inv = 1.0 / sqrt(norm2[blockIdx.y]);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    cplx z = p[i];
    p[i] = cmake(z.re * inv, z.im * inv);
  }
}

// local_trace (SV:30-80): sum_i sum_row op[row, col_i] psi_i conj(psi_j), j = basis index of the
// configuration of i with the digits of the measured site(s) replaced by `row`.  `n_sites` = 1 or 2,
// `op` is the dense (n_level^n_sites)^2 complex matrix, row-major, in device memory.
// two sites: col = col_1 + n_level*col_2 with col_1 the digit of sites[1], col_2 of sites[0] (SV:53-59)
__global__ void __launch_bounds__(256)
local_trace_kernel(const PlanView P, const cplx* __restrict__ psi, int n_sites, int site_a, int site_b,
                   const cplx* __restrict__ op, double* partial) {
  __shared__ double red[32];
  const uint32_t nl = (uint32_t)P.n_level;
  const int m = (n_sites == 1) ? (int)nl : (int)(nl * nl);
  const uint32_t div_1 = P.pow_level[n_sites == 1 ? site_a : site_b];  // n_level ** sites[1]
  const uint32_t div_2 = (n_sites == 2) ? P.pow_level[site_a] : 0u;    // n_level ** sites[0]
  double sr = 0.0, si = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.dim; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t cfg = P.full_space ? (uint32_t)i : P.configs[i];
    const uint32_t c1 = (cfg / div_1) % nl;
    const uint32_t c2 = (n_sites == 2) ? (cfg / div_2) % nl : 0u;
    const int col = (int)(c1 + c2 * nl);
    const cplx zi = psi[i];
    for (int row = 0; row < m; ++row) {
      const cplx ele = op[row * m + col];
      if (ele.re == 0.0 && ele.im == 0.0) continue;
      const uint32_t r2 = (n_sites == 2) ? (uint32_t)row / nl : 0u;
      const uint32_t r1 = (n_sites == 2) ? (uint32_t)row % nl : (uint32_t)row;
      uint32_t ncfg = cfg - c1 * div_1 + r1 * div_1;
      if (n_sites == 2) ncfg = ncfg - c2 * div_2 + r2 * div_2;
      int64_t j;
      if (P.full_space) {
        j = ncfg;
      } else {  // binary search in the sorted configuration list (np.searchsorted, SV:41-43)
        int64_t lo = 0, hi = P.dim;
        while (lo < hi) {
          int64_t mid = (lo + hi) >> 1;
          if (P.configs[mid] < ncfg) lo = mid + 1; else hi = mid;
        }
        j = (lo < P.dim && P.configs[lo] == ncfg) ? lo : -1;
      }
      if (j < 0) continue;
      const cplx zj = psi[j];
      // ele * psi_i * conj(psi_j)
      const double pr = zi.re * zj.re + zi.im * zj.im;
      const double pi = zi.im * zj.re - zi.re * zj.im;
      sr += ele.re * pr - ele.im * pi;
      si += ele.re * pi + ele.im * pr;
    }
  }
  sr = block_sum(sr, red);
  si = block_sum(si, red);
  if (threadIdx.x == 0) {
    partial[blockIdx.x * 2 + 0] = sr;
    partial[blockIdx.x * 2 + 1] = si;
  }
}

// ---- kernel (4): Born sampling ---------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048 probabilities per CTA

// tile_sums[seg][tile] = sum of |psi|^2 over the tile (coalesced: a warp reads 512 contiguous bytes per load)
__global__ void __launch_bounds__(kScanThreads)
scan_tile_sums_kernel(const cplx* __restrict__ psi, int64_t dim, int64_t seg_stride, int n_tiles, double* tile_sums) {
  __shared__ double red[32];
  const cplx* p = psi + (int64_t)blockIdx.y * seg_stride;
  const int64_t base = (int64_t)blockIdx.x * kScanTile + threadIdx.x;
  cplx z[kScanItems];
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    const int64_t i = base + (int64_t)k * kScanThreads;
    z[k] = i < dim ? p[i] : cmake(0.0, 0.0);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) s += z[k].re * z[k].re + z[k].im * z[k].im;
  s = block_sum(s, red);
  if (threadIdx.x == 0) tile_sums[(int64_t)blockIdx.y * n_tiles + blockIdx.x] = s;
}

// exclusive scan of tile_sums per segment (one CTA per segment, chunks of 1024 tiles: warp-shuffle scans + a scan of
// the 32 warp totals); totals[seg]
__global__ void __launch_bounds__(1024)
scan_tile_offsets_kernel(double* tile_sums, int n_tiles, double* totals) {
  __shared__ double wtot[32];
  __shared__ double carry;
  double* ts = tile_sums + (int64_t)blockIdx.x * n_tiles;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0.0;
  __syncthreads();
  for (int base = 0; base < n_tiles; base += 1024) {
    const int idx = base + threadIdx.x;
    const double v = idx < n_tiles ? ts[idx] : 0.0;
    double incl = v;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const double n = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += n;
    }
    if (lane == 31) wtot[wid] = incl;
    __syncthreads();
    if (wid == 0) {  // inclusive scan of the warp totals
      double w = wtot[lane];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const double n = __shfl_up_sync(0xffffffffu, w, off);
        if (lane >= off) w += n;
      }
      wtot[lane] = w;
    }
    __syncthreads();
    const double c = carry;
    const double before = wid > 0 ? wtot[wid - 1] : 0.0;
    if (idx < n_tiles) ts[idx] = c + before + incl - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == 1023) carry = c + before + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

// In-tile part of the CDF, shared by the two sampling forms: value k of this thread is
// cdf[base + k] = (tile_offset + inclusive in-tile prefix of |psi|^2) / total, in one fixed summation order
// (np.random.choice: cdf = p.cumsum(); cdf /= cdf[-1]).  Needs kScanThreads threads and a shared double[8].
__device__ __forceinline__ void tile_cdf_values(const cplx* __restrict__ p, int64_t dim, int64_t tile, double tile_offset,
                                                double inv, double* warp_tot, double (&c)[kScanItems]) {
  const int64_t base = tile * kScanTile + (int64_t)threadIdx.x * kScanItems;
  double v[kScanItems];
  double run = 0.0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    int64_t i = base + k;
    double pr = 0.0;
    if (i < dim) {
      cplx z = p[i];
      pr = z.re * z.re + z.im * z.im;
    }
    run += pr;
    v[k] = run;
  }
  // exclusive scan of the per-thread totals across the CTA
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double incl = run;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    double n = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += n;
  }
  if (lane == 31) warp_tot[wid] = incl;
  __syncthreads();
  double wbase = 0.0;
  for (int w = 0; w < wid; ++w) wbase += warp_tot[w];
  const double excl = wbase + incl - run + tile_offset;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) c[k] = (excl + v[k]) * inv;
}

// cdf[seg][i] for every i: the form for batches of small states (the CDF of a task stays in L2 for its shots)
__global__ void __launch_bounds__(kScanThreads)
scan_write_cdf_kernel(const cplx* __restrict__ psi, int64_t dim, int64_t seg_stride, int n_tiles,
                      const double* __restrict__ tile_offsets, const double* __restrict__ totals, double* cdf) {
  __shared__ double warp_tot[kScanThreads / 32];
  const cplx* p = psi + (int64_t)blockIdx.y * seg_stride;
  double* c = cdf + (int64_t)blockIdx.y * dim;
  double val[kScanItems];
  tile_cdf_values(p, dim, blockIdx.x, tile_offsets[(int64_t)blockIdx.y * n_tiles + blockIdx.x], 1.0 / totals[blockIdx.y],
                  warp_tot, val);
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    int64_t i = base + k;
    if (i < dim) c[i] = val[k];
  }
}

// digits of configuration `lo` -> shots row (space.py:219-227; three-level projection h->0, r->1 when `project`)
__device__ __forceinline__ void write_shot(int64_t lo, const uint32_t* __restrict__ configs, int n_atoms, int n_level,
                                           int project, uint8_t* o) {
  uint32_t cfg = configs ? configs[lo] : (uint32_t)lo;
  for (int a = 0; a < n_atoms; ++a) {
    uint32_t d = cfg % (uint32_t)n_level;
    cfg /= (uint32_t)n_level;
    if (project) d = (d == 2u) ? 1u : 0u;
    o[a] = (uint8_t)d;
  }
}

// The form for LARGE states with few shots (no CDF array: 8 bytes per amplitude less traffic and scratch): one CTA
// per shot finds the tile by a binary search over the tile offsets and re-evaluates that tile's CDF values - the
// same arithmetic as scan_write_cdf_kernel - to find the first index with cdf > u (searchsorted side="right").
__global__ void __launch_bounds__(kScanThreads)
sample_search_tiles_kernel(const cplx* __restrict__ psi, int64_t dim, int64_t seg_stride, int n_tiles,
                           const double* __restrict__ tile_offsets, const double* __restrict__ totals,
                           const uint32_t* __restrict__ configs, int n_atoms, int n_level, int project,
                           const double* __restrict__ uniforms, int n_shots, uint8_t* shots) {
  __shared__ double warp_tot[kScanThreads / 32];
  __shared__ unsigned long long first;
  const int seg = blockIdx.y, s = blockIdx.x;
  const cplx* p = psi + (int64_t)seg * seg_stride;
  const double* off = tile_offsets + (int64_t)seg * n_tiles;
  const double total = totals[seg], inv = 1.0 / total;
  const double u = uniforms[(int64_t)seg * n_shots + s];
  // first tile whose end value (the offset of the next tile) exceeds u
  int lo = 0, hi = n_tiles - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (off[mid + 1] * inv <= u) lo = mid + 1; else hi = mid;
  }
  if (threadIdx.x == 0) first = ~0ull;
  double val[kScanItems];
  tile_cdf_values(p, dim, lo, off[lo], inv, warp_tot, val);  // has a barrier after `first` is set
  const int64_t base = (int64_t)lo * kScanTile + (int64_t)threadIdx.x * kScanItems;
  unsigned long long mine = ~0ull;
#pragma unroll
  for (int k = kScanItems - 1; k >= 0; --k)
    if (base + k < dim && val[k] > u) mine = (unsigned long long)(base + k);
  if (mine != ~0ull) atomicMin(&first, mine);
  __syncthreads();
  if (threadIdx.x == 0) {
    int64_t idx = first == ~0ull ? ((int64_t)(lo + 1) * kScanTile) : (int64_t)first;  // rounding at a tile edge
    if (idx >= dim) idx = dim - 1;
    write_shot(idx, configs, n_atoms, n_level, project, shots + ((int64_t)seg * n_shots + s) * n_atoms);
  }
}

// shots[seg][s][atom] = digit `atom` of the configuration found by searchsorted(cdf, u, side="right")
// (space.py:217-227; three-level projection h->0, r->1 when `project`)
__global__ void sample_search_kernel(const double* __restrict__ cdf, int64_t dim, const uint32_t* __restrict__ configs,
                                     int n_atoms, int n_level, int project, const double* __restrict__ uniforms,
                                     int n_shots, uint8_t* shots) {
  const int seg = blockIdx.y;
  const double* c = cdf + (int64_t)seg * dim;
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < n_shots; s += gridDim.x * blockDim.x) {
    const double u = uniforms[(int64_t)seg * n_shots + s];
    int64_t lo = 0, hi = dim;  // first index with cdf > u
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (c[mid] <= u) lo = mid + 1; else hi = mid;
    }
    if (lo >= dim) lo = dim - 1;
    write_shot(lo, configs, n_atoms, n_level, project, shots + ((int64_t)seg * n_shots + s) * n_atoms);
  }
}

// ---- on-device van-der-Waals diagonal (HC:76-89) ---------------------------------------------------
// V[i] = sum over pairs a<b (a ascending, then b ascending - the reference's loop order) of pair[a][b]
// when both atoms are Rydberg in basis state i.  Pure additions in the same order => bit-identical to
// the host diagonal.  pair entries cut by the eps test are 0 and skipped.
__global__ void build_rydberg_kernel(const PlanView P, const double* __restrict__ pair, double* V) {
  const int n = P.n_atoms;
  const uint32_t nl = (uint32_t)P.n_level, ryd = nl - 1u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.dim; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t cfg = P.full_space ? (uint32_t)i : P.configs[i];
    uint32_t mask = 0u, rem = cfg;
    for (int a = 0; a < n; ++a) {
      if (rem % nl == ryd) mask |= (1u << a);
      rem /= nl;
    }
    double v = 0.0;
    for (int a = 0; a < n; ++a) {
      if (!((mask >> a) & 1u)) continue;
      for (int b = a + 1; b < n; ++b) {
        const double c = pair[a * n + b];
        if (((mask >> b) & 1u) && c != 0.0) v = __dadd_rn(v, c);
      }
    }
    V[i] = v;
  }
}

}  // namespace bqa
