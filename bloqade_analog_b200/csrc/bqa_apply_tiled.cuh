// bqa_apply_tiled.cuh - north_star kernel (1) for LARGE two-level full-space states: the fused H(t) psi as
// shared-memory-tiled passes instead of one gather pass.
//
// The one-pass kernel (apply_kernel) gathers 25 partners per element straight from global memory at 2^25:
// ncu shows 4.2x the algorithmic DRAM traffic (profiles/r01_summary.md).  Here a CTA stages a TILE of 2^13
// amplitudes (128 KB) in shared memory whose index bits are a contiguous low group [0, lo_bits) plus a high
// group [hb, hb + tb): every flip of a tile bit is a conflict-free shared-memory load (or a register move).
// Pass A (lo_bits = 13) adds the diagonal and the flips of atoms 0..12 and writes H psi partially; a later
// pass (lo_bits = 3: full 128-byte DRAM lines, or 2 when the 11th fresh atom saves a pass) accumulates the
// flips of its high group into it, the last one applies the -i factor.  Each pass may also take ONE leftover
// atom whose partner amplitudes are gathered from global memory: atom 13 in pass A (the neighbouring tile),
// a top atom (or, for a sharded state, a rank bit = the peer GPU's shard) in later passes.  DRAM traffic by
// design at n = 25: 40 + 48 B per amplitude + what the two leftover gathers miss in L2 (measured 98 B).
//
// One CTA per SM (the tile takes most of the shared memory), so the CTA overlaps its own DRAM reads with its
// shared-memory work: see apply_tiled_body for the ring of slots and the per-quad request order.  (L2
// prefetch instructions were tried - bulk and per line - and doubled the DRAM reads without producing hits;
// side streams requested a further quad ahead cost more in registers than they hide: profiles/r01_summary.md.)
//
// Reference code replaced: RydbergHamiltonian._ode_complex_kernel / _apply (emulate/ir/state_vector.py:323-334,
// 375-398) with IndexMapping / CSR matvecs of bit-flip form (emulate/sparse_operator.py:24-34,99-128).
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>

#include "bqa_core.cuh"

namespace bqa {

constexpr int kTileBits = 13;
constexpr int kTileThreads = 512;
constexpr int kTileGroup = 2;
constexpr int kTilePer = (1 << kTileBits) / kTileThreads;  // 16 amplitudes per thread
constexpr int kTileSpare = 12;                             // of which this many are prefetched into shared memory
constexpr int kTileSlots = kTilePer + kTileSpare;          // ring of 8 KB slots (512 amplitudes each)
constexpr size_t kTileSlotBytes = sizeof(double) * 2 * kTileThreads * kTileSlots;

struct TileSpec {
  int lo_bits;   // tile bits [0, lo_bits); <= 9 unless first
  int hb, tb;    // tile bits [hb, hb + tb);  lo_bits + tb == kTileBits
  int p0;        // tile bits [p0, 13) are flipped in this pass (0 in the first pass)
  int n_x;       // 0 or 1 leftover atom whose partner is gathered from global memory in this pass
  int xbit;
  int pair;      // the leftover atom is local: tiles are walked in partner pairs (t, t ^ (1 << pair_b))
  int pair_b;    // bit of the TILE index that the leftover atom's index bit maps to
  int hint;      // L2 eviction hints: bit 0 side streams and result stores evict_first, bit 1 psi tiles evict_last
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
  double tabK[16];             // first pass: detuning sums of tile bits 9..12 (the amplitude number k of a thread)
};
constexpr size_t kTileDynSmem = kTileSlotBytes + sizeof(TileShared);  // the ring, then the tables

__device__ __forceinline__ int64_t tile_base(const TileSpec& S, int64_t tile) {
  const int mid = S.hb - S.lo_bits;  // free bits between the two tile groups
  const int64_t lowpart = tile & (((int64_t)1 << mid) - 1);
  return (lowpart << S.lo_bits) | ((tile >> mid) << (S.hb + S.tb));
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
// Explicit state-space loads for the tile loop.  `volatile` (no memory clobber) keeps them in program order
// relative to each other and to the barriers: the side-stream requests of a quad are issued before its
// shared-memory loads, and no generic-to-shared address conversion is generated.
__device__ __forceinline__ cplx lds_cplx(uint32_t saddr) {
  cplx z;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(z.re), "=d"(z.im) : "r"(saddr));
  return z;
}
__device__ __forceinline__ double lds_f64(uint32_t saddr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(saddr));
  return v;
}
__device__ __forceinline__ cplx ldg_cplx(const cplx* g) {
  cplx z;
  asm volatile("ld.global.v2.f64 {%0, %1}, [%2];" : "=d"(z.re), "=d"(z.im) : "l"(g));
  return z;
}
__device__ __forceinline__ double ldg_f64(const double* g) {
  double v;
  asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(g));
  return v;
}
// The same accesses with an L2 eviction policy: the tile a CTA has just fetched is the leftover-atom partner of
// the next one, so the streams that are touched once (V, partial result, stores) should not push it out of L2.
__device__ __forceinline__ uint64_t l2_policy(int kind) {  // 0 normal, 1 evict_first, 2 evict_last
  uint64_t pol;
  if (kind == 1)
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  else if (kind == 2)
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  else
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* g, uint64_t pol) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(smem_addr), "l"(g), "l"(pol) : "memory");
}
__device__ __forceinline__ cplx ldg_cplx(const cplx* g, uint64_t pol) {
  cplx z;
  asm volatile("ld.global.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(z.re), "=d"(z.im) : "l"(g), "l"(pol));
  return z;
}
__device__ __forceinline__ double ldg_f64(const double* g, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(g), "l"(pol));
  return v;
}
__device__ __forceinline__ void stg_cplx(cplx* g, double re, double im, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(g), "d"(re), "d"(im), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Tile walked by a CTA at iteration `it` (-1: none left).  Plain passes stride the tiles by the grid; a pass
// with a LOCAL leftover atom (S.pair) walks the two tiles that are each other's partner under that atom back
// to back, so the leftover gather of either finds the tile this CTA is fetching / has just fetched in L2.
__device__ __forceinline__ int64_t tile_seq(const TileSpec& S, int64_t it, int64_t n_tiles) {
  if (S.pair) {
    const int64_t q = blockIdx.x + (it >> 1) * (int64_t)gridDim.x;
    if (q >= (n_tiles >> 1)) return -1;
    const int b = S.pair_b;
    const int64_t t = ((q >> b) << (b + 1)) | (q & (((int64_t)1 << b) - 1));
    return t | ((it & 1) << b);
  }
  const int64_t t = blockIdx.x + it * (int64_t)gridDim.x;
  return t < n_tiles ? t : -1;
}

// A thread owns tile elements tid + 512 k, k = 0..15, and works on them in QUADS k = 4 g + j: the four
// amplitudes of a quad differ in tile bits 9 and 10, so those two flips are register moves; the other flips
// are one LDS.128 per amplitude and the coefficient of a flip is fetched once per quad.  The side streams of
// a quad (V or the partial result, and the leftover-atom partner) are requested when the quad starts and
// consumed when it ends.
//
// Shared memory is a ring of 28 slots of 512 amplitudes: the current tile sits in 16 of them, the next one
// arrives by cp.async in the 12 free ones while quads 0..2 are computed and in the first 4 slots of the
// current tile while its last quad is computed, so nothing is copied between tiles or parked in registers.
//
// One kernel per variant (real / complex coefficients, first / later pass, leftover atom or not; the host picks):
// operands that are kernel parameters stay in the constant bank, and every variant gets its own register
// allocation - folded into one kernel, the variants spilled inside the tile loop.
template <bool kReal, bool kFirst, bool kHasX, int kP0, bool kRemX = false>
__device__ __forceinline__ void apply_tiled_body(const int64_t dim, const int n, const double* __restrict__ V,
                                                 const TileSpec& S, const cplx* __restrict__ in,
                                                 cplx* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char tile_bytes[];
  const TileShared& sh = *reinterpret_cast<const TileShared*>(tile_bytes + kTileSlotBytes);
  constexpr int kQuad = 4, kQuads = kTilePer / kQuad;
  constexpr uint32_t kSlotBytes = kTileThreads * (uint32_t)sizeof(cplx);
  const uint32_t tid = threadIdx.x;
  const uint32_t tb16 = tid * (uint32_t)sizeof(cplx);
  // kP0 = 2 / 3: a later pass with 11 / 10 fresh atoms (the shapes of the later passes of a big state) gets
  // compile-time loop bounds; kP0 = 0: read it from the pass description
  const int p0 = kFirst ? 0 : (kP0 > 0 ? kP0 : S.p0);
  const int xb = S.xbit;
  const bool rot = S.last && !S.raw;
  const int64_t n_tiles = dim >> kTileBits;
  int64_t it = 0;
  int64_t tile = tile_seq(S, 0, n_tiles);
  if (tile < 0) return;
  // amplitude k of this thread is tile element tid + 512 k, at global offset off0 + k * stride from the tile base
  const uint32_t off0 = kFirst ? tid : ((tid & ((1u << S.lo_bits) - 1u)) | ((tid >> S.lo_bits) << S.hb));
  const uint32_t stride = kFirst ? (uint32_t)kTileThreads : (((uint32_t)kTileThreads >> S.lo_bits) << S.hb);
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(tile_bytes);  // not 8 KB aligned: XOR offsets, then add
  const uint32_t smem0 = sbase + tb16;
  const uint32_t s_crp = sbase + (uint32_t)(kTileSlotBytes + offsetof(TileShared, crp));
  const uint32_t s_cip = sbase + (uint32_t)(kTileSlotBytes + offsetof(TileShared, cip));
  const uint32_t s_tabk = sbase + (uint32_t)(kTileSlotBytes + offsetof(TileShared, tabK));
  const uint64_t pol_once = l2_policy((S.hint & 1) ? 1 : 0);  // V, partial result, stores: touched once per pass
  const uint64_t pol_psi = l2_policy((S.hint & 2) ? 2 : 0);   // psi tiles: re-read as leftover-atom partners
  uint32_t s0 = 0;  // ring slot of amplitude k = 0 of the current tile (uniform over the CTA)
  // byte offset of the slot holding amplitude k of the current tile; k in [16, 32) addresses the next tile
  auto slot = [&](int k) -> uint32_t {
    uint32_t q = s0 + (uint32_t)k;
    q -= q >= 2u * kTileSlots ? 2u * kTileSlots : (q >= (uint32_t)kTileSlots ? (uint32_t)kTileSlots : 0u);
    return q * kSlotBytes;
  };
  {
    const cplx* src = in + tile_base(S, tile) + off0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      cplx tmp[kTilePer / 2];
#pragma unroll
      for (int k = 0; k < kTilePer / 2; ++k) tmp[k] = src[(size_t)(h * (kTilePer / 2) + k) * stride];
#pragma unroll
      for (int k = 0; k < kTilePer / 2; ++k)
        *reinterpret_cast<cplx*>(tile_bytes + (h * (kTilePer / 2) + k) * kSlotBytes + tb16) = tmp[k];
    }
  }
  double dl = 0.0;  // detuning of the tile bits held in the thread index (first pass: atoms 0..8)
  if (kFirst) {
#pragma unroll
    for (int p = 0; p < 9; ++p)
      if ((tid >> p) & 1) dl += sh.cf.delta[0][p];
  }
  // kRemX (sharded state, the leftover atom is a rank bit): the partner amplitudes come from the PEER GPU over NVLink
  // (~2 us away, L2 bypassed).  They are requested ONE QUAD AHEAD of the quad that uses them (xn), across tile
  // boundaries, so the link always has one or two quads (32 - 64 KB per SM) in flight instead of a burst per quad.
  cplx xn[kQuad];
  if (kRemX) {
    const cplx* x0 = S.x_ptr + tile_base(S, tile) + off0;
#pragma unroll
    for (int j = 0; j < kQuad; ++j) xn[j] = ldg_cplx(x0 + (size_t)j * stride);
  }
  __syncthreads();
  for (;;) {
    const int64_t base = tile_base(S, tile);
    const int64_t next = tile_seq(S, it + 1, n_tiles);
    const bool has_next = next >= 0;
    const cplx* nsrc = in + (has_next ? tile_base(S, next) : base) + off0;
    // per-tile constants: detuning of the atoms outside the tile, leftover-atom coefficient and sign
    const int64_t gbase = base | (int64_t)S.idx_hi;  // index in the whole register
    double dbase = dl;
    if (kFirst) {
      for (int a = kTileBits; a < n; ++a)
        if ((gbase >> a) & 1) dbase += sh.cf.delta[0][a];
    }
    double xr = 0.0, xi = 0.0;
    const cplx* xsrc = in;
    if (kHasX) {
      xr = sh.cf.c_re[0][xb];
      xi = ((gbase >> xb) & 1) ? sh.cf.c_im[0][xb] : -sh.cf.c_im[0][xb];
      xsrc = (S.x_ptr ? S.x_ptr + base : in + (base ^ ((int64_t)1 << xb))) + off0;
    }
    const double* Vt = V + base + off0;
    cplx* out_t = out + base + off0;
#pragma unroll
    for (int g = 0; g < kQuads; ++g) {
      // the last quad reads ring slots 4..15 of this tile only: once every thread is past quad 2, slots
      // 0..3 are dead and take the last 4 slots of the next tile
      if (g == kQuads - 1 && has_next) __syncthreads();
      // side streams of this quad: requested now, consumed after the flips
      double vq[kQuad];
      cplx oq[kQuad], xq[kQuad];
#pragma unroll
      for (int j = 0; j < kQuad; ++j) {
        const size_t off = (size_t)(g * kQuad + j) * stride;
        if (kFirst)
          vq[j] = V ? ldg_f64(Vt + off, pol_once) : 0.0;
        else
          oq[j] = ldg_cplx(out_t + off, pol_once);
        if (kHasX && !kRemX) xq[j] = ldg_cplx(xsrc + off, pol_psi);
      }
      if (kRemX) {
#pragma unroll
        for (int j = 0; j < kQuad; ++j) xq[j] = xn[j];
        if (g < kQuads - 1) {
#pragma unroll
          for (int j = 0; j < kQuad; ++j) xn[j] = ldg_cplx(xsrc + (size_t)((g + 1) * kQuad + j) * stride);
        } else if (has_next) {
          const cplx* xnext = S.x_ptr + tile_base(S, next) + off0;
#pragma unroll
          for (int j = 0; j < kQuad; ++j) xn[j] = ldg_cplx(xnext + (size_t)j * stride);
        }
      }
      // a quarter of the next tile per quad, behind the latency-critical side streams (the whole tile at once
      // is a 96 KB burst per SM that the side streams then queue behind)
      if (has_next) {
#pragma unroll
        for (int k = g * kQuad; k < (g + 1) * kQuad; ++k) cp_async16(smem0 + slot(kTilePer + k), nsrc + (size_t)k * stride, pol_psi);
        cp_async_commit();
      }
      cplx own[kQuad];
      uint32_t eb[kQuad];
      double ar[kQuad], ai[kQuad];
      const bool need_own = kFirst || p0 <= 10;
#pragma unroll
      for (int j = 0; j < kQuad; ++j) {
        eb[j] = slot(g * kQuad + j) + tb16;
        if (need_own) own[j] = lds_cplx(sbase + eb[j]);
        ar[j] = 0.0;
        ai[j] = 0.0;
      }
      // tile bits 9, 10: partners inside the quad
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int p = 9 + q;
        if (!kFirst && p < p0) continue;
        const double cr = lds_f64(s_crp + 8 * p), cim = kReal ? 0.0 : lds_f64(s_cip + 8 * p);
#pragma unroll
        for (int j = 0; j < kQuad; ++j) {
          const cplx z = own[j ^ (1 << q)];
          if (kReal) {
            ar[j] = fma(cr, z.re, ar[j]);
            ai[j] = fma(cr, z.im, ai[j]);
          } else {
            const double ci = ((j >> q) & 1) ? cim : -cim;
            ar[j] = fma(cr, z.re, fma(-ci, z.im, ar[j]));
            ai[j] = fma(cr, z.im, fma(ci, z.re, ai[j]));
          }
        }
      }
      // the other tile bits: conflict-free LDS.128 partners
#pragma unroll
      for (int p = 0; p < kTileBits; ++p) {
        if (p == 9 || p == 10) continue;
        if (!kFirst && p < p0) continue;
        const double cr = lds_f64(s_crp + 8 * p);
        double ci = 0.0;
        if (!kReal) {
          const bool up = p < 9 ? ((tid >> p) & 1) : ((g >> (p - 11)) & 1);
          const double cim = lds_f64(s_cip + 8 * p);
          ci = up ? cim : -cim;
        }
#pragma unroll
        for (int j = 0; j < kQuad; ++j) {
          const uint32_t pb = p < 9 ? (eb[j] ^ ((uint32_t)sizeof(cplx) << p))
                                    : (slot((g * kQuad + j) ^ (1 << (p - 9))) + tb16);
          const cplx z = lds_cplx(sbase + pb);
          if (kReal) {
            ar[j] = fma(cr, z.re, ar[j]);
            ai[j] = fma(cr, z.im, ai[j]);
          } else {
            ar[j] = fma(cr, z.re, fma(-ci, z.im, ar[j]));
            ai[j] = fma(cr, z.im, fma(ci, z.re, ai[j]));
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kQuad; ++j) {
        if (kFirst) {
          const double diag = vq[j] - (dbase + lds_f64(s_tabk + 8 * (g * kQuad + j)));
          ar[j] = fma(diag, own[j].re, ar[j]);
          ai[j] = fma(diag, own[j].im, ai[j]);
        } else {
          ar[j] += oq[j].re;
          ai[j] += oq[j].im;
        }
        if (kHasX) {
          ar[j] = fma(xr, xq[j].re, fma(-xi, xq[j].im, ar[j]));
          ai[j] = fma(xr, xq[j].im, fma(xi, xq[j].re, ai[j]));
        }
        stg_cplx(out_t + (size_t)(g * kQuad + j) * stride, rot ? ai[j] : ar[j], rot ? -ar[j] : ai[j], pol_once);
      }
    }
    if (!has_next) break;
    cp_async_wait_all();
    __syncthreads();  // everyone is done reading this tile and the next one has landed
    s0 += kTilePer;
    if (s0 >= (uint32_t)kTileSlots) s0 -= kTileSlots;
    tile = next;
    ++it;
  }
}

// One CTA's share of a tiled pass: coefficients at time t, then the tile loop.  kReal is the host's promise that
// no Rabi term of the record carries a phase waveform (record_has_phase), i.e. every c_im is exactly zero.
template <bool kReal, bool kFirst, bool kHasX, int kP0, bool kRemX = false>
__device__ __forceinline__ void apply_tiled_cta(const PlanView& P, const double* __restrict__ rec,
                                                const double* __restrict__ V, double t, const TileSpec& S,
                                                const cplx* __restrict__ in, cplx* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char tile_bytes[];
  TileShared& sh = *reinterpret_cast<TileShared*>(tile_bytes + kTileSlotBytes);
  const int tid = threadIdx.x;
  const TaskView tv = task_view(rec);
  const int n = tv.n_atoms;  // whole register (> P.n_atoms for a sharded state)
  for (int e = tid; e < n; e += blockDim.x) coef_entry(tv, t, e, sh.cf);
  __syncthreads();
  auto atom_of = [&](int p) { return p < S.lo_bits ? p : S.hb + (p - S.lo_bits); };
  if (tid < kTileBits) {
    sh.crp[tid] = sh.cf.c_re[0][atom_of(tid)];
    sh.cip[tid] = sh.cf.c_im[0][atom_of(tid)];
  }
  if (kFirst && tid >= 32 && tid < 48) {
    const int q = tid - 32;
    double d = 0.0;
    for (int p = 0; p < 4; ++p)
      if ((q >> p) & 1) d += sh.cf.delta[0][atom_of(9 + p)];
    sh.tabK[q] = d;
  }
  __syncthreads();
  apply_tiled_body<kReal, kFirst, kHasX, kP0, kRemX>(P.dim, n, V, S, in, out);
}

template <bool kReal, bool kFirst, bool kHasX, int kP0, bool kRemX = false>
__global__ void __launch_bounds__(kTileThreads, 1)
apply_tiled_kernel(const PlanView P, const double* __restrict__ rec, const double* __restrict__ V, double t,
                   const TileSpec S, const cplx* __restrict__ in, cplx* __restrict__ out) {
  apply_tiled_cta<kReal, kFirst, kHasX, kP0, kRemX>(P, rec, V, t, S, in, out);
}

// kernel variant index of a pass: 8 shapes {later p0 runtime, later p0 = 2, later p0 = 3, first} x {no x, x},
// then real coefficients; 16..23: the four shapes with a REMOTE leftover partner (complex, then real coefficients)
inline int tiled_variant(bool real, const TileSpec& s) {
  const int shape = s.first ? 6 : (s.p0 == 2 ? 2 : (s.p0 == 3 ? 4 : 0));
  // remote leftover partner (kRemX); BQA_NO_REMX: measurement hook, the partner is requested like a local one
  if (s.n_x > 0 && s.x_ptr != nullptr && getenv("BQA_NO_REMX") == nullptr) return 16 + (real ? 4 : 0) + shape / 2;
  return (real ? 8 : 0) + shape + (s.n_x > 0 ? 1 : 0);
}

// host: does any Rabi term of this task record carry a phase waveform?  (No: all coefficients are real.)
inline bool record_has_phase(const double* rec_host) { return task_has_phase(task_view(rec_host)); }

}  // namespace bqa
