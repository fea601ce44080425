// bqa_apply_tiled.cuh - north_star kernel (1) for LARGE two-level full-space states: the fused H(t) psi as
// shared-memory-tiled passes instead of one gather pass.
//
// The one-pass kernel (apply_kernel) gathers 25 partners per element straight from global memory at 2^25:
// ncu shows 4.2x the algorithmic DRAM traffic (profiles/r01_summary.md).  Here a CTA stages a TILE of 2^13
// amplitudes (128 KB) in shared memory whose index bits are a contiguous low group [0, lo_bits) plus a high
// group [hb, hb + tb): every flip of a tile bit is a conflict-free LDS.128.  Pass A (lo_bits = 13) adds the
// diagonal and the flips of atoms 0..12 and writes H psi partially; pass B (lo_bits = 2 for 64-byte DRAM
// segments, high group = atoms 13..23) accumulates the remaining flips into it and applies the -i factor;
// one leftover atom per pass is gathered from global memory (atom 24 at n = 25).  DRAM traffic by design:
// 40 + 48 + 16 = 104 B per amplitude.
//
// One CTA per SM (the tile takes most of the shared memory), so the CTA overlaps its own DRAM reads with its
// shared-memory work: while tile i is computed, tile i+1 arrives by cp.async into the spare 88 KB of shared
// memory (11 of the 16 amplitudes each thread owns) and in 20 registers (the other 5); the third stream of
// the pass (V, or the partial result) is read one group of amplitudes ahead.  (L2 prefetch instructions were
// tried instead - bulk and per line - and doubled the DRAM reads without producing hits: profiles/r01_summary.md.)
//
// Reference code replaced: RydbergHamiltonian._ode_complex_kernel / _apply (emulate/ir/state_vector.py:323-334,
// 375-398) with IndexMapping / CSR matvecs of bit-flip form (emulate/sparse_operator.py:24-34,99-128).
#pragma once
#include <cuda_runtime.h>

#include "bqa_core.cuh"

namespace bqa {

constexpr int kTileBits = 13;
constexpr int kTileThreads = 512;
constexpr int kTileGroup = 2;
constexpr int kTilePer = (1 << kTileBits) / kTileThreads;  // 16 amplitudes per thread
constexpr int kTileSpare = 11;                             // of which this many are prefetched into shared memory
constexpr size_t kTileDynSmem = sizeof(double) * 2 * kTileThreads * (kTilePer + kTileSpare);

struct TileSpec {
  int lo_bits;   // tile bits [0, lo_bits); <= 9 unless first
  int hb, tb;    // tile bits [hb, hb + tb);  lo_bits + tb == kTileBits
  int p0;        // tile bits [p0, 13) are flipped in this pass (0 in the first pass)
  int n_x;       // 0 or 1 leftover atom whose partner is gathered from global memory in this pass
  int xbit;
  int first;     // first pass: out = (diag psi + flips); otherwise out += flips
  int last;      // last pass: multiply by -i unless raw
  int raw;
  // sharded state (bqa_api_shard.cuh): this rank holds the amplitudes whose index bits >= n_atoms(local) equal
  // idx_hi; a leftover atom above the local ones has its partner amplitudes at x_ptr (the peer rank's shard)
  int n_total;           // atoms of the whole register (== local atoms when not sharded)
  long long idx_hi;      // rank << n_local
  const cplx* x_ptr;     // null: the leftover atom is local (partner = in[i ^ bit])
};

struct TileShared {
  Coefs cf;
  double crp[16], cip[16];     // coefficient of tile bit p (atom p, or hb + p - lo_bits)
  double tabL[128], tabH[64];  // detuning sums of the tile-index bits (low 7 / high 6), first pass
};

__device__ __forceinline__ int64_t tile_base(const TileSpec& S, int64_t tile) {
  const int mid = S.hb - S.lo_bits;  // free bits between the two tile groups
  const int64_t lowpart = tile & (((int64_t)1 << mid) - 1);
  return (lowpart << S.lo_bits) | ((tile >> mid) << (S.hb + S.tb));
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <bool kReal, bool kFirst>
__device__ __forceinline__ void apply_tiled_body(const int64_t dim, const int n, const double* __restrict__ V,
                                                 const TileSpec& S, const cplx* __restrict__ in,
                                                 cplx* __restrict__ out, unsigned char* tile_bytes, TileShared& sh) {
  const uint32_t tid = threadIdx.x;
  const int p0 = kFirst ? 0 : S.p0;
  const bool has_x = S.n_x > 0;
  const int xb = S.xbit;
  const bool rot = S.last && !S.raw;
  const int64_t n_tiles = dim >> kTileBits;
  int64_t tile = blockIdx.x;
  if (tile >= n_tiles) return;
  cplx* T = reinterpret_cast<cplx*>(tile_bytes);
  cplx* Sp = T + (1 << kTileBits);  // spare: amplitudes k < kTileSpare of the next tile
  // amplitude k of this thread is tile element tid + 512 k, at global offset off0 + k * stride from the tile base
  const uint32_t off0 = kFirst ? tid : ((tid & ((1u << S.lo_bits) - 1u)) | ((tid >> S.lo_bits) << S.hb));
  const uint32_t stride = kFirst ? (uint32_t)kTileThreads : (((uint32_t)kTileThreads >> S.lo_bits) << S.hb);
  const uint32_t sp_addr = (uint32_t)__cvta_generic_to_shared(Sp + tid);
  {
    const cplx* src = in + tile_base(S, tile) + off0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      cplx tmp[kTilePer / 2];
#pragma unroll
      for (int k = 0; k < kTilePer / 2; ++k) tmp[k] = src[(size_t)(h * (kTilePer / 2) + k) * stride];
#pragma unroll
      for (int k = 0; k < kTilePer / 2; ++k) T[tid + (h * (kTilePer / 2) + k) * kTileThreads] = tmp[k];
    }
  }
  __syncthreads();
  for (; tile < n_tiles; tile += gridDim.x) {
    const int64_t base = tile_base(S, tile);
    const int64_t next = tile + gridDim.x;
    const bool has_next = next < n_tiles;
    cplx pre[kTilePer - kTileSpare];
    if (has_next) {
      const cplx* src = in + tile_base(S, next) + off0;
#pragma unroll
      for (int k = 0; k < kTileSpare; ++k)
        cp_async16(sp_addr + k * kTileThreads * (uint32_t)sizeof(cplx), src + (size_t)k * stride);
      cp_async_commit();
#pragma unroll
      for (int k = kTileSpare; k < kTilePer; ++k) pre[k - kTileSpare] = src[(size_t)k * stride];
    }
    // per-tile constants: detuning of the atoms outside the tile, leftover-atom partner and sign
    const int64_t gbase = base | (int64_t)S.idx_hi;  // index in the whole register
    double dbase = 0.0;
    if (kFirst) {
      for (int a = kTileBits; a < n; ++a)
        if ((gbase >> a) & 1) dbase += sh.cf.delta[0][a];
    }
    const cplx* xin = (S.x_ptr ? S.x_ptr + base : in + (has_x ? (base ^ ((int64_t)1 << xb)) : base)) + off0;
    double xr = 0.0, xi = 0.0;
    if (has_x) {
      xr = sh.cf.c_re[0][xb];
      xi = ((gbase >> xb) & 1) ? sh.cf.c_im[0][xb] : -sh.cf.c_im[0][xb];
    }
    const double* Vt = V ? V + base + off0 : nullptr;
    cplx* out_t = out + base + off0;
    // third stream of the pass (V in the first pass, the partial result afterwards) and the leftover-atom
    // partner are read one group of amplitudes ahead of the shared-memory work
    cplx oc[kTileGroup], xc[kTileGroup], on[kTileGroup], xn[kTileGroup];
    auto load_group = [&](int k0, cplx* o, cplx* x) {
#pragma unroll
      for (int j = 0; j < kTileGroup; ++j) {
        const size_t off = (size_t)(k0 + j) * stride;
        if (kFirst)
          o[j] = cmake(Vt ? Vt[off] : 0.0, 0.0);
        else
          o[j] = out_t[off];
        if (has_x) x[j] = xin[off];
      }
    };
    load_group(0, oc, xc);
#pragma unroll 1
    for (int k0 = 0; k0 < kTilePer; k0 += kTileGroup) {
      if (k0 + kTileGroup < kTilePer) load_group(k0 + kTileGroup, on, xn);
#pragma unroll
      for (int j = 0; j < kTileGroup; ++j) {
        const uint32_t e = tid + (k0 + j) * kTileThreads;
        const uint32_t eb = e * (uint32_t)sizeof(cplx);
        double ar, ai;
        if (kFirst) {
          const cplx own = *reinterpret_cast<const cplx*>(tile_bytes + eb);
          const double diag = oc[j].re - (dbase + sh.tabL[e & 127] + sh.tabH[e >> 7]);
          ar = diag * own.re;
          ai = diag * own.im;
        } else {
          ar = oc[j].re;
          ai = oc[j].im;
        }
#pragma unroll
        for (int p = 0; p < kTileBits; ++p) {
          if (!kFirst && p < p0) continue;
          const cplx z = *reinterpret_cast<const cplx*>(tile_bytes + (eb ^ ((uint32_t)sizeof(cplx) << p)));
          const double cr = sh.crp[p];
          if (kReal) {
            ar = fma(cr, z.re, ar);
            ai = fma(cr, z.im, ai);
          } else {
            const double ci = ((e >> p) & 1) ? sh.cip[p] : -sh.cip[p];
            ar = fma(cr, z.re, fma(-ci, z.im, ar));
            ai = fma(cr, z.im, fma(ci, z.re, ai));
          }
        }
        if (has_x) {
          ar = fma(xr, xc[j].re, fma(-xi, xc[j].im, ar));
          ai = fma(xr, xc[j].im, fma(xi, xc[j].re, ai));
        }
        out_t[(size_t)(k0 + j) * stride] = rot ? cmake(ai, -ar) : cmake(ar, ai);
      }
#pragma unroll
      for (int j = 0; j < kTileGroup; ++j) {
        oc[j] = on[j];
        xc[j] = xn[j];
      }
    }
    __syncthreads();  // everyone is done reading this tile
    if (has_next) {
      cp_async_wait_all();  // each thread moves only the slots it fetched itself: no barrier needed in between
#pragma unroll
      for (int k = 0; k < kTileSpare; ++k) T[tid + k * kTileThreads] = Sp[tid + k * kTileThreads];
#pragma unroll
      for (int k = kTileSpare; k < kTilePer; ++k) T[tid + k * kTileThreads] = pre[k - kTileSpare];
    }
    __syncthreads();
  }
}

// One CTA's share of a tiled pass: coefficients at time t, then the tile loop
__device__ __forceinline__ void apply_tiled_cta(const PlanView& P, const double* __restrict__ rec,
                                                const double* __restrict__ V, double t, const TileSpec& S,
                                                const cplx* __restrict__ in, cplx* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char tile_smem[];
  __shared__ TileShared sh;
  __shared__ int s_complex;
  const int tid = threadIdx.x;
  const TaskView tv = task_view(rec);
  const int n = tv.n_atoms;  // whole register (> P.n_atoms for a sharded state)
  if (tid == 0) s_complex = 0;
  __syncthreads();
  for (int e = tid; e < n; e += blockDim.x) {
    coef_entry(tv, t, e, sh.cf);
    if (sh.cf.c_im[0][e] != 0.0) s_complex = 1;
  }
  __syncthreads();
  auto atom_of = [&](int p) { return p < S.lo_bits ? p : S.hb + (p - S.lo_bits); };
  if (tid < kTileBits) {
    sh.crp[tid] = sh.cf.c_re[0][atom_of(tid)];
    sh.cip[tid] = sh.cf.c_im[0][atom_of(tid)];
  }
  if (S.first) {
    if (tid >= 32 && tid < 160) {
      const int q = tid - 32;
      double s = 0.0;
      for (int p = 0; p < 7; ++p)
        if ((q >> p) & 1) s += sh.cf.delta[0][atom_of(p)];
      sh.tabL[q] = s;
    } else if (tid >= 160 && tid < 224) {
      const int q = tid - 160;
      double s = 0.0;
      for (int p = 0; p < 6; ++p)
        if ((q >> p) & 1) s += sh.cf.delta[0][atom_of(7 + p)];
      sh.tabH[q] = s;
    }
  }
  __syncthreads();
  if (S.first) {
    if (s_complex)
      apply_tiled_body<false, true>(P.dim, n, V, S, in, out, tile_smem, sh);
    else
      apply_tiled_body<true, true>(P.dim, n, V, S, in, out, tile_smem, sh);
  } else {
    if (s_complex)
      apply_tiled_body<false, false>(P.dim, n, V, S, in, out, tile_smem, sh);
    else
      apply_tiled_body<true, false>(P.dim, n, V, S, in, out, tile_smem, sh);
  }
}

__global__ void __launch_bounds__(kTileThreads, 1)
apply_tiled_kernel(const PlanView P, const double* __restrict__ rec, const double* __restrict__ V, double t,
                   const TileSpec S, const cplx* __restrict__ in, cplx* __restrict__ out) {
  apply_tiled_cta(P, rec, V, t, S, in, out);
}

}  // namespace bqa
