// bqa_cluster_fast.cuh - fused DOP853 stage loop of the cluster engine for the common operator shape:
// two-level atoms, full 2^n space, bit-flip Rabi terms only (no generic CSR term), Schroedinger picture.
//
// Same algorithm as the generic TaskEvolver (bqa_evolver.cuh / bqa_evolver_steps.cuh) - the partner sums and
// stage combines add their terms in the same order - but organised for the machine:
//   * T compute threads (compile time) own K elements l = tid + T*k of the CTA's chunk each; flips of the
//     bits that index k are register moves, flips of the low log2(T) bits are conflict-free LDS.128 with
//     immediate offsets, flips of the bits above the chunk are DSMEM loads from a peer CTA (the peer's
//     window address is mapped once per task);
//   * the RHS of stage s and the stage combine of stage s+1 are ONE pass over the chunk (the fresh K value
//     never leaves the thread), writing the next stage input into the OTHER of two ys buffers: a stage
//     costs one pass + one cluster barrier;
//   * real drive coefficients (no phase waveform) take a 2-FMA-per-partner variant;
//   * one extra SERVICE WARP evaluates the waveform coefficients of the NEXT stage time into the other of
//     two Coefs buffers while the compute warps work, so that latency chain is off the critical path;
//   * the static diagonal V of the chunk is cached in shared memory, Butcher tables come from the
//     constant bank.
// Reference flow replaced: scipy dop853 driving RydbergHamiltonian._ode_complex_kernel
// (emulate/ir/state_vector.py:323-334,589-599).
#pragma once
// -DBQA_PHASE_CLOCKS: per-phase clock64 accounting printed by CTAs 0/1 (profiles/r01_summary.md)
#ifdef BQA_PHASE_CLOCKS
#include <cstdio>
#endif
#include "bqa_cluster.cuh"

namespace bqa {

constexpr int kVecYsB = 12;  // 13th vector of the fast path: second stage-input buffer
constexpr int kWfCache = 64;  // waveforms of a task the service warp evaluates one per lane

// Butcher rows in the constant bank (uniform LDC instead of L1-flushed global loads)
namespace ctab {
#undef BQA_TABLE_QUAL
#define BQA_TABLE_QUAL __constant__ const
#include "dop_coeffs.h"
}  // namespace ctab

// ---- explicit shared-space accesses (32-bit shared addresses; LDS/STS instead of generic LD/ST) ----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ cplx lds_c(uint32_t a) {
  cplx z;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(z.re), "=d"(z.im) : "r"(a) : "memory");
  return z;
}
__device__ __forceinline__ void sts_c(uint32_t a, cplx z) {
  asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(z.re), "d"(z.im) : "memory");
}
__device__ __forceinline__ double lds_d(uint32_t a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
// peer CTA's shared memory (DSMEM): map a shared address into the cluster window, load from it
__device__ __forceinline__ uint32_t mapa_u32(uint32_t a, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ cplx ldsc_c(uint32_t a) {
  cplx z;
  asm volatile("ld.shared::cluster.v2.f64 {%0, %1}, [%2];" : "=d"(z.re), "=d"(z.im) : "r"(a) : "memory");
  return z;
}

// this thread's view of one vector chunk: element l = tid + idx lives at a + 16*idx
struct SVec {
  uint32_t a;
  __device__ __forceinline__ cplx ld(int idx) const { return lds_c(a + 16u * (uint32_t)idx); }
  __device__ __forceinline__ void st(int idx, cplx z) const { sts_c(a + 16u * (uint32_t)idx, z); }
};

template <int K, int T>
struct Fast853 {
  static constexpr int LT = (T == 512) ? 9 : (T == 256) ? 8 : (T == 128) ? 7 : (T == 64) ? 6 : 5;
  static constexpr int LK = (K == 4) ? 2 : (K == 2) ? 1 : 0;
  static_assert((1 << LT) == T, "T must be 32..512, a power of two");
  static constexpr uint32_t kOffDelta = 0, kOffCre = sizeof(double) * kMaxSlots * kMaxAtoms,
                            kOffCim = 2 * sizeof(double) * kMaxSlots * kMaxAtoms;

  ClusterCtx& ctx;
  TaskEvolver<ClusterCtx>& ev;  // generic evolver of the same task (prologue of every integrate call)
  const PlanView& P;
  const TaskView& tv;
  cplx** vec;           // shared vector table (pre-offset generic pointers, used by the generic evolver)
  uint32_t* addr_gen;   // the u32 address table as a generic pointer (for the swaps)
  Coefs* coefs;         // shared [2]
  int* realflag;        // shared [2]: all drive coefficients of coefs[b] are real
  double* wfval;        // shared [kWfCache]: waveform values of the service warp
  const SolverOpts& o;
  Controller& c;
  int n, tid;
  bool compute;         // tid < T (the service warp has compute == false)
  uint32_t sm_base;     // shared address of the dynamic shared memory (vectors, then the V chunk)
  uint32_t addr_tab;    // shared address of the u32 table: shared address of every vector's chunk
  uint32_t coef_base;   // shared address of coefs[0]
  uint32_t vs_addr;     // shared address of the V chunk, 0 = no static diagonal
  uint32_t peer_base[3];  // cluster-window address of the peers' dynamic shared memory (rank bits 0..2)
  int cur;              // which ys buffer holds the current stage input (0: kVecYs, 1: kVecYsB)
  int ccur;             // which Coefs buffer holds the coefficients of the current stage time
#ifdef BQA_PHASE_CLOCKS
  long long pc_work = 0, pc_sync = 0, pc_final = 0, pc_post = 0, pc_passes = 0, pc_steps = 0;
#endif

  __device__ __forceinline__ uint32_t addr(int v) const { return lds_u(addr_tab + 4u * (uint32_t)v); }
  __device__ __forceinline__ SVec svec(int v) const {
    SVec s;
    s.a = addr(v) + 16u * (uint32_t)tid;
    return s;
  }
  __device__ __forceinline__ int ysv(int which) const { return which ? kVecYsB : kVecYs; }

  __device__ __forceinline__ void init_peers() {
    peer_base[0] = peer_base[1] = peer_base[2] = sm_base;
    if (ctx.csize > 1) peer_base[0] = mapa_u32(sm_base, (uint32_t)(ctx.rank ^ 1));
    if (ctx.csize > 2) peer_base[1] = mapa_u32(sm_base, (uint32_t)(ctx.rank ^ 2));
    if (ctx.csize > 4) peer_base[2] = mapa_u32(sm_base, (uint32_t)(ctx.rank ^ 4));
  }

  // kv[k] = (-i H(t) ys)[l_k] for this thread's K elements; ya = shared address of the stage input chunk,
  // cfa = shared address of the coefficient table
  template <bool kReal>
  __device__ __forceinline__ void rhs_elems(uint32_t ya, uint32_t cfa, cplx (&kv)[K]) {
    cplx own[K];
    double ar[K], ai[K];
    const uint32_t me = ya + 16u * (uint32_t)tid;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      own[k] = lds_c(me + 16u * T * k);
      ar[k] = 0.0;
      ai[k] = 0.0;
    }
    double dlow = 0.0;  // detuning of the set bits among the low (thread-index) bits
    // bits below log2(T): partner in this CTA's shared memory
#pragma unroll
    for (int b = 0; b < LT; ++b) {
      if (b < n) {
        const double cr = lds_d(cfa + kOffCre + 8u * b);
        const uint32_t pp = ya + 16u * (uint32_t)(tid ^ (1 << b));
        const bool up = (tid >> b) & 1;
        if (up) dlow += lds_d(cfa + kOffDelta + 8u * b);
        cplx z[K];
#pragma unroll
        for (int k = 0; k < K; ++k) z[k] = lds_c(pp + 16u * T * k);
        if (kReal) {
#pragma unroll
          for (int k = 0; k < K; ++k) {
            ar[k] += cr * z[k].re;
            ai[k] += cr * z[k].im;
          }
        } else {
          const double cim = lds_d(cfa + kOffCim + 8u * b);
          const double ci = up ? cim : -cim;
#pragma unroll
          for (int k = 0; k < K; ++k) {
            ar[k] += cr * z[k].re - ci * z[k].im;
            ai[k] += cr * z[k].im + ci * z[k].re;
          }
        }
      }
    }
    double dk[K];
#pragma unroll
    for (int k = 0; k < K; ++k) dk[k] = dlow;
    // bits that index k: partner is another register of this thread
#pragma unroll
    for (int jb = 0; jb < LK; ++jb) {
      const int b = LT + jb;
      if (b < n) {
        const double cr = lds_d(cfa + kOffCre + 8u * b), dl = lds_d(cfa + kOffDelta + 8u * b);
        const double cim = kReal ? 0.0 : lds_d(cfa + kOffCim + 8u * b);
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = own[k ^ (1 << jb)];
          if ((k >> jb) & 1) dk[k] += dl;
          if (kReal) {
            ar[k] += cr * z.re;
            ai[k] += cr * z.im;
          } else {
            const double ci = ((k >> jb) & 1) ? cim : -cim;
            ar[k] += cr * z.re - ci * z.im;
            ai[k] += cr * z.im + ci * z.re;
          }
        }
      }
    }
    // bits above the chunk: partner at the same offset in a peer CTA (DSMEM)
#pragma unroll
    for (int rb = 0; rb < 3; ++rb) {
      const int b = LT + LK + rb;
      if (b < n) {
        const double cr = lds_d(cfa + kOffCre + 8u * b);
        const bool up = (ctx.rank >> rb) & 1;
        const uint32_t pp = peer_base[rb] + (me - sm_base);
        const double dl = up ? lds_d(cfa + kOffDelta + 8u * b) : 0.0;
        cplx z[K];
#pragma unroll
        for (int k = 0; k < K; ++k) z[k] = ldsc_c(pp + 16u * T * k);
        if (kReal) {
#pragma unroll
          for (int k = 0; k < K; ++k) {
            dk[k] += dl;
            ar[k] += cr * z[k].re;
            ai[k] += cr * z[k].im;
          }
        } else {
          const double cim = lds_d(cfa + kOffCim + 8u * b);
          const double ci = up ? cim : -cim;
#pragma unroll
          for (int k = 0; k < K; ++k) {
            dk[k] += dl;
            ar[k] += cr * z[k].re - ci * z[k].im;
            ai[k] += cr * z[k].im + ci * z[k].re;
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const double v = vs_addr ? lds_d(vs_addr + 8u * (uint32_t)(tid + T * k)) : 0.0;
      const double diag = v - dk[k];
      kv[k] = cmake(diag * own[k].im + ai[k], -(diag * own[k].re + ar[k]));  // -i * (H ys)
    }
  }

  // next stage input  ysn = y + h * (row of the Butcher table . K)  for this thread's elements.
  // row >= 0: DOP853 a-row; row == -1: the hinit probe  y + h*K[0]
  __device__ __forceinline__ void combine_elems(double h, int row, int yn_vec) {
    double sr[K], si[K];
#pragma unroll
    for (int k = 0; k < K; ++k) sr[k] = si[k] = 0.0;
    const int nnz = row >= 0 ? ctab::DOP853_a_nnz[row] : 1;
    for (int j = 0; j < nnz; ++j) {
      const int slot = row >= 0 ? ctab::DOP853_a_slot[row][j] : 0;
      const double v = row >= 0 ? ctab::DOP853_a_val[row][j] : 1.0;
      const SVec kj = svec(kVecK + slot);
      cplx z[K];
#pragma unroll
      for (int k = 0; k < K; ++k) z[k] = kj.ld(T * k);
#pragma unroll
      for (int k = 0; k < K; ++k) {
        sr[k] += v * z[k].re;
        si[k] += v * z[k].im;
      }
    }
    const SVec yy = svec(kVecY), out = svec(yn_vec);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const cplx y0 = yy.ld(T * k);
      out.st(T * k, cmake(y0.re + h * sr[k], y0.im + h * si[k]));
    }
  }

  // service warp: coefficient table of time t into buffer b (+ its "all real" flag)
  // Every waveform of the task is evaluated once, one per lane (wf_eval is a dependent chain of ~700 cycles; a lane
  // per ATOM repeated it for every term, which made this warp the critical path of a stage at dim <= 1024); then a
  // lane per atom combines the values.  Same functions, same order: the same bits as coef_entry.
  __device__ __forceinline__ void service_coefs(int b, double t) {
    const int a = tid - T;
    bool im = false;
    if (tv.n_wf <= kWfCache) {
      for (int w = a; w < tv.n_wf; w += 32) wfval[w] = wf_eval(tv, w, t);
      __syncwarp();
      if (a < n) {
        coef_entry_from_values(tv, wfval, a, coefs[b]);
        im = coefs[b].c_im[0][a] != 0.0;
      }
      __syncwarp();
    } else if (a < n) {
      coef_entry(tv, t, a, coefs[b]);
      im = coefs[b].c_im[0][a] != 0.0;
    }
    const unsigned any = __ballot_sync(0xffffffffu, im);
    if (a == 0) realflag[b] = any ? 0 : 1;
  }

  // One fused pass: [K[dst] = f(ys_cur) with coefs[ccur]] ; [ys_other = y + h*row] by the compute warps,
  // [coefs[other] at t_next] by the service warp; then the cluster barrier.
  __device__ __forceinline__ void pass(bool do_rhs, int dst, int row, double h, bool do_coefs, double t_next) {
    const bool do_combine = row >= -1;
#ifdef BQA_PHASE_CLOCKS
    const long long pc0 = clock64();
#endif
    if (compute) {
      if (do_rhs) {
        const uint32_t cfa = coef_base + (uint32_t)ccur * (uint32_t)sizeof(Coefs);
        const bool real = realflag[ccur] != 0;
        const uint32_t ya = addr(ysv(cur));
        cplx kv[K];
        if (real)
          rhs_elems<true>(ya, cfa, kv);
        else
          rhs_elems<false>(ya, cfa, kv);
        const SVec kd = svec(kVecK + dst);
#pragma unroll
        for (int k = 0; k < K; ++k) kd.st(T * k, kv[k]);
      }
      if (do_combine) combine_elems(h, row, ysv(cur ^ 1));
    } else if (do_coefs) {
      service_coefs(ccur ^ 1, t_next);
    }
#ifdef BQA_PHASE_CLOCKS
    const long long pc1 = clock64();
    ctx.sync();
    pc_work += pc1 - pc0;
    pc_sync += clock64() - pc1;
    pc_passes++;
#else
    ctx.sync();
#endif
    if (do_combine) cur ^= 1;
    if (do_coefs) ccur ^= 1;
  }

  // new solution ynew -> K slot 4 and -> the other ys buffer; error sums; stiffness denominator
  __device__ __forceinline__ void final_elems(double h, double& err, double& err3, double& stden) {
      const int bn = ctab::DOP853_b_nnz[0], e5n = ctab::DOP853_e5_nnz[0], e3n = ctab::DOP853_e3_nnz[0];
      const double atol = o.atol, rtol = o.rtol;
      const SVec yy = svec(kVecY), old = svec(ysv(cur)), ynew = svec(kVecK + 4), ysn = svec(ysv(cur ^ 1));
      double sr[K], si[K], e5r[K], e5i[K], e3r[K], e3i[K];
#pragma unroll
      for (int k = 0; k < K; ++k) sr[k] = si[k] = e5r[k] = e5i[k] = e3r[k] = e3i[k] = 0.0;
      for (int j = 0; j < bn; ++j) {
        const SVec kj = svec(kVecK + ctab::DOP853_b_slot[0][j]);
        const double v = ctab::DOP853_b_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj.ld(T * k);
          sr[k] += v * z.re;
          si[k] += v * z.im;
        }
      }
      for (int j = 0; j < e5n; ++j) {
        const SVec kj = svec(kVecK + ctab::DOP853_e5_slot[0][j]);
        const double v = ctab::DOP853_e5_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj.ld(T * k);
          e5r[k] += v * z.re;
          e5i[k] += v * z.im;
        }
      }
      for (int j = 0; j < e3n; ++j) {
        const SVec kj = svec(kVecK + ctab::DOP853_e3_slot[0][j]);
        const double v = ctab::DOP853_e3_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj.ld(T * k);
          e3r[k] += v * z.re;
          e3i[k] += v * z.im;
        }
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const cplx yo = yy.ld(T * k);
        const cplx z = cmake(yo.re + h * sr[k], yo.im + h * si[k]);
        double sk = atol + rtol * fmax(fabs(yo.re), fabs(z.re));
        double q = e5r[k] / sk;
        err += q * q;
        q = e3r[k] / sk;
        err3 += q * q;
        sk = atol + rtol * fmax(fabs(yo.im), fabs(z.im));
        q = e5i[k] / sk;
        err += q * q;
        q = e3i[k] / sk;
        err3 += q * q;
        ynew.st(T * k, z);
        const cplx o12 = old.ld(T * k);
        const double dr = z.re - o12.re, di = z.im - o12.im;
        stden += dr * dr + di * di;
        ysn.st(T * k, z);
      }
  }

  // One attempted DOP853 step as 13 fused passes through ONE call site of pass() (code size):
  //   it = 0      stage-2 input
  //   it = 1..11  K of stage it+1 from the current stage input, then the next stage input (it <= 10)
  //   it = 12     new solution + error norm + decision; if accepted, K4 = f(x+h, ynew) (next step's k1)
  __device__ __forceinline__ void step(int& accepted) {
    const double x = c.x, h = c.h, tn = x + h;
    const int64_t n_real = 2 * P.dim;
    double stden = 0.0;
    accepted = 0;
    for (int it = 0; it <= 12; ++it) {
      int dst = 0, row = -2;
      bool do_coefs = false;
      double t_next = 0.0;
      if (it < 12) {
        const int s = it - 1;
        dst = (s == 9) ? 1 : (s == 10) ? 2 : s + 1;
        if (it <= 10) {
          row = it;
          do_coefs = true;
          t_next = (it == 10) ? tn : x + ctab::DOP853_c[it + 1] * h;
        }
      } else {
#ifdef BQA_PHASE_CLOCKS
        const long long pf0 = clock64();
#endif
        c.nfcn += 11;
        double err = 0.0, err3 = 0.0;
        if (compute) final_elems(h, err, err3, stden);
        ctx.sum3(err, err3, stden);  // contains the barrier that publishes ynew in the other ys buffer
        cur ^= 1;
        accepted = c.decide(o, c.err_dop853(err, err3, n_real));
#ifdef BQA_PHASE_CLOCKS
        pc_final += clock64() - pf0;
        pc_steps++;
#endif
        if (!accepted) break;
        dst = 3;  // the coefficients at tn are those of stage 12, still in coefs[ccur]
      }
      pass(it > 0, dst, row, h, do_coefs, t_next);
    }
#ifdef BQA_PHASE_CLOCKS
    const long long pp0 = clock64();
#endif
    if (accepted) {
      c.nfcn += 1;
      if (c.stiff_due(o)) {
        double stnum = 0.0;
        if (compute) {
          const SVec k4 = svec(kVecK + 3), k3 = svec(kVecK + 2);
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const cplx a = k4.ld(T * k), b = k3.ld(T * k);
            const double dr = a.re - b.re, di = a.im - b.im;
            stnum += dr * dr + di * di;
          }
        }
        ctx.sum1(stnum);
        c.stiff_update(o, stnum, stden);
      }
      cplx** vv = vec;
      uint32_t* aa = addr_gen;
      ctx.single([&]() {
        cplx* t0 = vv[kVecK + 0];
        vv[kVecK + 0] = vv[kVecK + 3];
        vv[kVecK + 3] = t0;
        t0 = vv[kVecY];
        vv[kVecY] = vv[kVecK + 4];
        vv[kVecK + 4] = t0;
        uint32_t a0 = aa[kVecK + 0];
        aa[kVecK + 0] = aa[kVecK + 3];
        aa[kVecK + 3] = a0;
        a0 = aa[kVecY];
        aa[kVecY] = aa[kVecK + 4];
        aa[kVecK + 4] = a0;
      });
    }
#ifdef BQA_PHASE_CLOCKS
    pc_post += clock64() - pp0;
#endif
  }

  // one integrate(x0 -> x1) call (fresh start like scipy's ode.integrate).  The prologue (k1, automatic
  // initial step: 2 RHS per call) runs through the generic evolver; the step loop is the fused one.
  __device__ __forceinline__ void integrate(double x0, double x1) {
#ifdef BQA_PHASE_CLOCKS
    const long long pi0 = clock64();
#endif
    c.begin(o, x0, x1);
    ev.stage_input(x0, 0.0, 0, nullptr, nullptr);
    ev.rhs_from_ys(x0, ev.k(0));
    ev.hinit();
#ifdef BQA_PHASE_CLOCKS
    const long long pi1 = clock64();
#endif
    while (!c.done) {
      if (!c.step_begin(o)) break;
      int accepted = 0;
      step(accepted);
      c.step_end(o, accepted);
    }
#ifdef BQA_PHASE_CLOCKS
    if (blockIdx.x < 2 && (tid == 0 || tid == T || tid == 33))
      printf("PHASE blk %d tid %d: total %lld prologue %lld | passes %lld work %lld sync %lld | steps %lld final %lld post %lld\n",
             (int)blockIdx.x, tid, clock64() - pi0, pi1 - pi0, pc_passes, pc_work, pc_sync, pc_steps, pc_final, pc_post);
#endif
  }
};

// ---- the kernel: persistent clusters pulling tasks from the work queue ---------------------------------
__global__ void __launch_bounds__(544, 1) evolve_cluster_kernel(const ClusterArgs CA) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ Coefs s_coefs[2];
  __shared__ int s_real[2];
  __shared__ double s_wfval[kWfCache];
  __shared__ uint32_t s_addr[kNumVec + 1];  // shared address of every vector's chunk (fast path)
  __shared__ double s_red[3 * 32];
  __shared__ double s_part[8];
  __shared__ cplx* s_vec[kNumVec + 1];
  __shared__ int s_task;
  __shared__ double s_rec[kRecSmemDoubles];

  const SmallArgs& A = CA.S;
  cg::cluster_group cl = cg::this_cluster();
  const int csize = CA.csize;
  const int rank = csize > 1 ? (int)cl.block_rank() : 0;
  const int64_t dim = A.P.dim;

  ClusterCtx ctx;
  ctx.red = s_red;
  ctx.part = s_part;
  ctx.parity = 0;
  ctx.rank = rank;
  ctx.csize = csize;
  ctx.chunk = CA.chunk;
  ctx.chunk_shift = CA.chunk_shift;
  ctx.lo = (int64_t)rank * CA.chunk;
  ctx.hi = ctx.lo + CA.chunk < dim ? ctx.lo + CA.chunk : dim;
  if (ctx.lo > dim) ctx.lo = dim;

  cplx* smem_vecs = reinterpret_cast<cplx*>(dyn_smem);

  for (;;) {
    // the cluster's rank-0 CTA pulls the next task; the others read it through DSMEM
    if (rank == 0 && threadIdx.x == 0) s_task = atomicAdd(A.counter, 1);
    ctx.sync();
    const int q = csize > 1 ? *cl.map_shared_rank(&s_task, 0) : s_task;
    ctx.sync();  // everyone has read s_task before rank 0 may overwrite it
    if (q >= A.n_tasks) break;
    const int task = A.order ? A.order[q] : q;
    const double* rec = A.records + A.offsets[task];
    const int64_t rec_len = A.offsets[task + 1] - A.offsets[task];
    if (rec_len <= kRecSmemDoubles) {
      for (int64_t k = threadIdx.x; k < rec_len; k += blockDim.x) s_rec[k] = rec[k];
      rec = s_rec;
    }
    // vector table: every vector's chunk lives in this CTA's shared memory, pointers pre-offset so
    // that vec[v][i] addresses global row i for lo <= i < hi
    if (threadIdx.x == 0) {
      for (int v = 0; v < CA.n_alloc; ++v) {
        s_vec[v] = smem_vecs + (int64_t)v * CA.chunk - ctx.lo;
        s_addr[v] = smem_u32(smem_vecs + (int64_t)v * CA.chunk);
      }
    }
    __syncthreads();
    {
      cplx* y = s_vec[kVecY];
      if (A.psi0_mode == 0) {
        for (int64_t i = ctx.lo + threadIdx.x; i < ctx.hi; i += blockDim.x) y[i] = cmake(i == 0 ? 1.0 : 0.0, 0.0);
      } else {
        const cplx* src = A.psi0 + (A.psi0_mode == 2 ? (int64_t)task * dim : 0);
        for (int64_t i = ctx.lo + threadIdx.x; i < ctx.hi; i += blockDim.x) y[i] = src[i];
      }
    }
    __syncthreads();

    const int diag_id = (int)rec[1];
    const double* V = (diag_id >= 0 && diag_id < A.n_diags) ? A.diags[diag_id] : nullptr;
    Workspace w;
    w.vec = s_vec;
    w.coefs = &s_coefs[0];
    TaskEvolver<ClusterCtx> ev(ctx, A.P, rec, V, w, A.o);
    if (CA.fastK == 0) {
      cluster_run_times(ev, ev, A, task, rank);
    } else {
      // chunk of the static diagonal -> shared memory (behind the vectors)
      double* Vs = reinterpret_cast<double*>(smem_vecs + (int64_t)CA.n_alloc * CA.chunk);
      if (V)
        for (int64_t l = threadIdx.x; l < (int64_t)CA.chunk; l += blockDim.x) Vs[l] = V[ctx.lo + l];
      __syncthreads();
      const int T = (int)blockDim.x - 32;
      const int tid = (int)threadIdx.x;
#define BQA_FAST_CASE(KK, TT)                                                                               \
  {                                                                                                         \
    Fast853<KK, TT> fe{ctx, ev, A.P, ev.tv, s_vec, s_addr, s_coefs, s_real, s_wfval, ev.o, ev.c, A.P.n_atoms, tid, \
                       tid < TT, smem_u32(smem_vecs), smem_u32(s_addr), smem_u32(s_coefs),                  \
                       V ? smem_u32(Vs) : 0u, {0u, 0u, 0u}, 0, 0};                                          \
    fe.init_peers();                                                                                        \
    cluster_run_times(fe, ev, A, task, rank);                                                               \
  }
      if (CA.fastK == 2 && T == 512) BQA_FAST_CASE(2, 512)
      else if (CA.fastK == 4 && T == 256) BQA_FAST_CASE(4, 256)
      else if (CA.fastK == 2 && T == 256) BQA_FAST_CASE(2, 256)
      else if (CA.fastK == 2 && T == 128) BQA_FAST_CASE(2, 128)
#undef BQA_FAST_CASE
    }
    ctx.sync();
  }
  // no CTA may exit while a peer could still read its shared memory
  ctx.sync();
}

}  // namespace bqa
