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
#include "bqa_cluster.cuh"

namespace bqa {

constexpr int kVecYsB = 12;  // 13th vector of the fast path: second stage-input buffer

// Butcher rows in the constant bank (uniform LDC instead of L1-flushed global loads)
namespace ctab {
#undef BQA_TABLE_QUAL
#define BQA_TABLE_QUAL __constant__ const
#include "dop_coeffs.h"
}  // namespace ctab

template <int K, int T>
struct Fast853 {
  static constexpr int LT = (T == 512) ? 9 : (T == 256) ? 8 : (T == 128) ? 7 : (T == 64) ? 6 : 5;
  static constexpr int LK = (K == 4) ? 2 : (K == 2) ? 1 : 0;
  static_assert((1 << LT) == T, "T must be 32..512, a power of two");

  ClusterCtx& ctx;
  TaskEvolver<ClusterCtx>& ev;  // generic evolver of the same task (prologue of every integrate call)
  const PlanView& P;
  const TaskView& tv;
  cplx* sm;             // dynamic shared memory base (vectors, then the V chunk)
  const double* Vs;     // chunk of the static diagonal in shared memory, or nullptr
  cplx** vec;           // shared vector table (pre-offset pointers, shared with the generic evolver)
  Coefs* coefs;         // shared [2]
  int* realflag;        // shared [2]: all drive coefficients of coefs[b] are real
  const SolverOpts& o;
  Controller& c;
  int n, tid;
  bool compute;         // tid < T (the service warp has compute == false)
  const cplx* peer[3];  // peers' dynamic shared memory base for rank bits 0..2 (DSMEM window)
  int cur;              // which ys buffer holds the current stage input (0: kVecYs, 1: kVecYsB)
  int ccur;             // which Coefs buffer holds the coefficients of the current stage time

  // element offset of vector v inside `sm`
  __device__ __forceinline__ int off(int v) const { return (int)((vec[v] + ctx.lo) - sm); }
  __device__ __forceinline__ int ys_off(int which) const { return off(which ? kVecYsB : kVecYs); }

  __device__ __forceinline__ void init_peers() {
    peer[0] = peer[1] = peer[2] = sm;
    if (ctx.csize > 1) {
      cg::cluster_group cl = cg::this_cluster();
      for (int rb = 0; (1 << rb) < ctx.csize; ++rb) peer[rb] = cl.map_shared_rank(sm, ctx.rank ^ (1 << rb));
    }
  }

  // kv[k] = (-i H(t) ys)[l_k] for this thread's K elements
  template <bool kReal>
  __device__ __forceinline__ void rhs_elems(int yo, const Coefs& cf, cplx (&kv)[K]) {
    cplx own[K];
    double ar[K], ai[K];
    const cplx* ysl = sm + yo + tid;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      own[k] = ysl[T * k];
      ar[k] = 0.0;
      ai[k] = 0.0;
    }
    double dlow = 0.0;  // detuning of the set bits among the low (thread-index) bits
    // bits below log2(T): partner in this CTA's shared memory
#pragma unroll
    for (int b = 0; b < LT; ++b) {
      if (b < n) {
        const double cr = cf.c_re[0][b];
        const cplx* pp = sm + yo + (tid ^ (1 << b));
        if ((tid >> b) & 1) dlow += cf.delta[0][b];
        if (kReal) {
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const cplx z = pp[T * k];
            ar[k] += cr * z.re;
            ai[k] += cr * z.im;
          }
        } else {
          const double ci = ((tid >> b) & 1) ? cf.c_im[0][b] : -cf.c_im[0][b];
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const cplx z = pp[T * k];
            ar[k] += cr * z.re - ci * z.im;
            ai[k] += cr * z.im + ci * z.re;
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
        const double cr = cf.c_re[0][b], cim = kReal ? 0.0 : cf.c_im[0][b], dl = cf.delta[0][b];
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
        const double cr = cf.c_re[0][b];
        const bool up = (ctx.rank >> rb) & 1;
        const cplx* pp = peer[rb] + yo + tid;
        const double dl = up ? cf.delta[0][b] : 0.0;
        cplx z[K];
#pragma unroll
        for (int k = 0; k < K; ++k) z[k] = pp[T * k];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          dk[k] += dl;
          if (kReal) {
            ar[k] += cr * z[k].re;
            ai[k] += cr * z[k].im;
          } else {
            const double ci = up ? cf.c_im[0][b] : -cf.c_im[0][b];
            ar[k] += cr * z[k].re - ci * z[k].im;
            ai[k] += cr * z[k].im + ci * z[k].re;
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const double diag = (Vs ? Vs[tid + T * k] : 0.0) - dk[k];
      kv[k] = cmake(diag * own[k].im + ai[k], -(diag * own[k].re + ar[k]));  // -i * (H ys)
    }
  }

  // next stage input  ysn = y + h * (row of the Butcher table . K)  for this thread's elements.
  // row >= 0: DOP853 a-row; row == -1: the hinit probe  y + h*K[0]
  __device__ __forceinline__ void combine_elems(double h, int row, int yn) {
    double sr[K], si[K];
#pragma unroll
    for (int k = 0; k < K; ++k) sr[k] = si[k] = 0.0;
    const int nnz = row >= 0 ? ctab::DOP853_a_nnz[row] : 1;
    for (int j = 0; j < nnz; ++j) {
      const int slot = row >= 0 ? ctab::DOP853_a_slot[row][j] : 0;
      const double v = row >= 0 ? ctab::DOP853_a_val[row][j] : 1.0;
      const cplx* kj = sm + off(kVecK + slot) + tid;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const cplx z = kj[T * k];
        sr[k] += v * z.re;
        si[k] += v * z.im;
      }
    }
    const cplx* yy = sm + off(kVecY) + tid;
    cplx* out = sm + yn + tid;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const cplx y0 = yy[T * k];
      out[T * k] = cmake(y0.re + h * sr[k], y0.im + h * si[k]);
    }
  }

  // service warp: coefficient table of time t into buffer b (+ its "all real" flag)
  __device__ __forceinline__ void service_coefs(int b, double t) {
    const int a = tid - T;
    bool im = false;
    if (a < n) {
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
    if (compute) {
      if (do_rhs) {
        const Coefs& cf = coefs[ccur];
        const bool real = realflag[ccur] != 0;
        cplx kv[K];
        if (real)
          rhs_elems<true>(ys_off(cur), cf, kv);
        else
          rhs_elems<false>(ys_off(cur), cf, kv);
        cplx* kd = sm + off(kVecK + dst) + tid;
#pragma unroll
        for (int k = 0; k < K; ++k) kd[T * k] = kv[k];
      }
      if (do_combine) combine_elems(h, row, ys_off(cur ^ 1));
    } else if (do_coefs) {
      service_coefs(ccur ^ 1, t_next);
    }
    ctx.sync();
    if (do_combine) cur ^= 1;
    if (do_coefs) ccur ^= 1;
  }

  // new solution ynew -> K slot 4 and -> the other ys buffer; error sums; stiffness denominator
  __device__ __forceinline__ void final_elems(double h, double& err, double& err3, double& stden) {
      const int bn = ctab::DOP853_b_nnz[0], e5n = ctab::DOP853_e5_nnz[0], e3n = ctab::DOP853_e3_nnz[0];
      const double atol = o.atol, rtol = o.rtol;
      const cplx* yy = sm + off(kVecY) + tid;
      const cplx* old = sm + ys_off(cur) + tid;
      cplx* ynew = sm + off(kVecK + 4) + tid;
      cplx* ysn = sm + ys_off(cur ^ 1) + tid;
      double sr[K], si[K], e5r[K], e5i[K], e3r[K], e3i[K];
#pragma unroll
      for (int k = 0; k < K; ++k) sr[k] = si[k] = e5r[k] = e5i[k] = e3r[k] = e3i[k] = 0.0;
      for (int j = 0; j < bn; ++j) {
        const cplx* kj = sm + off(kVecK + ctab::DOP853_b_slot[0][j]) + tid;
        const double v = ctab::DOP853_b_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj[T * k];
          sr[k] += v * z.re;
          si[k] += v * z.im;
        }
      }
      for (int j = 0; j < e5n; ++j) {
        const cplx* kj = sm + off(kVecK + ctab::DOP853_e5_slot[0][j]) + tid;
        const double v = ctab::DOP853_e5_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj[T * k];
          e5r[k] += v * z.re;
          e5i[k] += v * z.im;
        }
      }
      for (int j = 0; j < e3n; ++j) {
        const cplx* kj = sm + off(kVecK + ctab::DOP853_e3_slot[0][j]) + tid;
        const double v = ctab::DOP853_e3_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj[T * k];
          e3r[k] += v * z.re;
          e3i[k] += v * z.im;
        }
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const cplx yo = yy[T * k];
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
        ynew[T * k] = z;
        const cplx o12 = old[T * k];
        const double dr = z.re - o12.re, di = z.im - o12.im;
        stden += dr * dr + di * di;
        ysn[T * k] = z;
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
        c.nfcn += 11;
        double err = 0.0, err3 = 0.0;
        if (compute) final_elems(h, err, err3, stden);
        ctx.sum3(err, err3, stden);  // contains the barrier that publishes ynew in the other ys buffer
        cur ^= 1;
        accepted = c.decide(o, c.err_dop853(err, err3, n_real));
        if (!accepted) break;
        dst = 3;  // the coefficients at tn are those of stage 12, still in coefs[ccur]
      }
      pass(it > 0, dst, row, h, do_coefs, t_next);
    }
    if (accepted) {
      c.nfcn += 1;
      if (c.stiff_due(o)) {
        double stnum = 0.0;
        if (compute) {
          const cplx* k4 = sm + off(kVecK + 3) + tid;
          const cplx* k3 = sm + off(kVecK + 2) + tid;
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const cplx a = k4[T * k], b = k3[T * k];
            const double dr = a.re - b.re, di = a.im - b.im;
            stnum += dr * dr + di * di;
          }
        }
        ctx.sum1(stnum);
        c.stiff_update(o, stnum, stden);
      }
      cplx** vv = vec;
      ctx.single([&]() {
        cplx* t0 = vv[kVecK + 0];
        vv[kVecK + 0] = vv[kVecK + 3];
        vv[kVecK + 3] = t0;
        t0 = vv[kVecY];
        vv[kVecY] = vv[kVecK + 4];
        vv[kVecK + 4] = t0;
      });
    }
  }

  // one integrate(x0 -> x1) call (fresh start like scipy's ode.integrate).  The prologue (k1, automatic
  // initial step: 2 RHS per call) runs through the generic evolver; the step loop is the fused one.
  __device__ __forceinline__ void integrate(double x0, double x1) {
    c.begin(o, x0, x1);
    ev.stage_input(x0, 0.0, 0, nullptr, nullptr);
    ev.rhs_from_ys(x0, ev.k(0));
    ev.hinit();
    while (!c.done) {
      if (!c.step_begin(o)) break;
      int accepted = 0;
      step(accepted);
      c.step_end(o, accepted);
    }
  }
};

// ---- the kernel: persistent clusters pulling tasks from the work queue ---------------------------------
__global__ void __launch_bounds__(544, 1) evolve_cluster_kernel(const ClusterArgs CA) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ Coefs s_coefs[2];
  __shared__ int s_real[2];
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
      for (int v = 0; v < CA.n_alloc; ++v) s_vec[v] = smem_vecs + (int64_t)v * CA.chunk - ctx.lo;
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
    Fast853<KK, TT> fe{ctx, ev, A.P, ev.tv, smem_vecs, V ? Vs : nullptr, s_vec, s_coefs, s_real, ev.o, ev.c,    \
                       A.P.n_atoms, tid, tid < TT, {nullptr, nullptr, nullptr}, 0, 0};                      \
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
