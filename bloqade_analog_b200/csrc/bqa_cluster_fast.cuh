// bqa_cluster_fast.cuh - fused DOP853 stage loop of the cluster engine for the common operator shape:
// two-level atoms, full 2^n space, bit-flip Rabi terms only (no generic CSR term), Schroedinger picture.
//
// Same arithmetic, in the same order, as the generic TaskEvolver (bqa_evolver.cuh / bqa_evolver_steps.cuh)
// - the K vectors and stage inputs are bitwise identical - but organised for the machine:
//   * a thread owns K elements l = tid + T*k of its CTA's chunk; flips of the bits that index k are
//     register moves, flips of the low log2(T) bits are conflict-free LDS.128, flips of the bits above the
//     chunk are DSMEM loads from the peer CTA (map_shared_rank);
//   * the RHS of stage s and the stage combine of stage s+1 are ONE pass over the chunk (the fresh K value
//     never leaves the thread), writing the next stage input into the OTHER of two ys buffers, so a stage
//     costs one pass + one cluster barrier instead of two of each;
//   * the coefficients of the next stage time are evaluated by a few threads of the same pass into the
//     other of two Coefs buffers.
// Reference flow replaced: scipy dop853 driving RydbergHamiltonian._ode_complex_kernel
// (emulate/ir/state_vector.py:323-334,589-599).
#pragma once
#include "bqa_cluster.cuh"

namespace bqa {

constexpr int kVecYsB = 12;  // 13th vector of the fast path: second stage-input buffer

template <int K>
struct Fast853 {
  ClusterCtx& ctx;
  const PlanView& P;
  const TaskView& tv;
  const double* V;      // static diagonal (global memory) or nullptr
  cplx** vec;           // shared vector table (pre-offset pointers)
  Coefs* coefs;         // shared [2]
  const SolverOpts& o;
  Controller& c;
  int T, lt, m, n, tid;
  int cur;              // which ys buffer holds the current stage input (0: kVecYs, 1: kVecYsB)
  int ccur;             // which Coefs buffer holds the coefficients of the current stage time

  __device__ __forceinline__ cplx* raw(int v) const { return vec[v] + ctx.lo; }
  __device__ __forceinline__ cplx* ys_raw(int which) const { return raw(which ? kVecYsB : kVecYs); }

  // coefficients at time t into buffer `which` (threads 0..ne-1); visible after the next barrier
  __device__ __forceinline__ void coefs_into(int which, double t) {
    if (tid < n) coef_entry(tv, t, tid, coefs[which]);
  }

  // kout[l] = -i H(t) ys  for this thread's K elements; `own` = ys values of those elements
  __device__ __forceinline__ void rhs_elems(const cplx* ysr, const Coefs& cf, cplx (&kv)[K]) {
    cplx own[K];
    double ar[K], ai[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      own[k] = ysr[tid + T * k];
      ar[k] = 0.0;
      ai[k] = 0.0;
    }
    // bits below log2(T): partner in this CTA's shared memory
    const int nlow = lt < n ? lt : n;
    for (int b = 0; b < nlow; ++b) {
      const double cr = cf.c_re[0][b];
      const double ci = ((tid >> b) & 1) ? cf.c_im[0][b] : -cf.c_im[0][b];
      const cplx* pp = ysr + (tid ^ (1 << b));
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const cplx z = pp[T * k];
        ar[k] += cr * z.re - ci * z.im;
        ai[k] += cr * z.im + ci * z.re;
      }
    }
    // bits that index k: partner is another register of this thread
#pragma unroll
    for (int jb = 0; (1 << jb) < K; ++jb) {
      const int b = lt + jb;
      if (b < n) {
        const double cr = cf.c_re[0][b], cim = cf.c_im[0][b];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = own[k ^ (1 << jb)];
          const double ci = ((k >> jb) & 1) ? cim : -cim;
          ar[k] += cr * z.re - ci * z.im;
          ai[k] += cr * z.im + ci * z.re;
        }
      }
    }
    // bits above the chunk: partner at the same offset in a peer CTA (DSMEM)
    for (int b = m; b < n; ++b) {
      const int rb = b - m;
      const double cr = cf.c_re[0][b];
      const double ci = ((ctx.rank >> rb) & 1) ? cf.c_im[0][b] : -cf.c_im[0][b];
      const cplx* pp = cg::this_cluster().map_shared_rank(const_cast<cplx*>(ysr), ctx.rank ^ (1 << rb)) + tid;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const cplx z = pp[T * k];
        ar[k] += cr * z.re - ci * z.im;
        ai[k] += cr * z.im + ci * z.re;
      }
    }
    // diagonal: V[i] - sum over set bits of delta (same order as h_apply_element)
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const uint32_t i = (uint32_t)ctx.lo + tid + T * k;
      double diag = V ? V[i] : 0.0;
      for (int a = 0; a < n; ++a)
        if ((i >> a) & 1u) diag -= cf.delta[0][a];
      const cplx hz = cmake(diag * own[k].re + ar[k], diag * own[k].im + ai[k]);
      kv[k] = times_minus_i(hz);
    }
  }

  // next stage input  ysn = y + h * sum_j val[j] K[slot[j]]  for this thread's elements
  __device__ __forceinline__ void combine_elems(double h, int nnz, const int* slot, const double* val, cplx* ysn) {
    double sr[K], si[K];
#pragma unroll
    for (int k = 0; k < K; ++k) sr[k] = si[k] = 0.0;
    for (int j = 0; j < nnz; ++j) {
      const cplx* kj = raw(kVecK + slot[j]) + tid;
      const double v = val[j];
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const cplx z = kj[T * k];
        sr[k] += v * z.re;
        si[k] += v * z.im;
      }
    }
    const cplx* yy = raw(kVecY) + tid;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const cplx y0 = yy[T * k];
      ysn[tid + T * k] = cmake(y0.re + h * sr[k], y0.im + h * si[k]);
    }
  }

  // One fused pass: [K[dst] = f(ys_cur) with coefs[ccur]] ; [ys_other = y + h*row] ; [coefs[other] at t_next]
  // then the cluster barrier; flips cur / ccur when a new stage input was produced.
  __device__ __forceinline__ void pass(bool do_rhs, int dst, bool do_combine, double h, int nnz, const int* slot,
                                       const double* val, bool do_coefs, double t_next) {
    if (do_coefs) coefs_into(ccur ^ 1, t_next);
    if (do_rhs) {
      cplx kv[K];
      rhs_elems(ys_raw(cur), coefs[ccur], kv);
      cplx* kd = raw(kVecK + dst) + tid;
#pragma unroll
      for (int k = 0; k < K; ++k) kd[T * k] = kv[k];
    }
    if (do_combine) combine_elems(h, nnz, slot, val, ys_raw(cur ^ 1));
    ctx.sync();
    if (do_combine) cur ^= 1;
    if (do_coefs) ccur ^= 1;
  }

  __device__ void hinit() {
    const double atol = o.atol, rtol = o.rtol;
    const cplx* yy = raw(kVecY) + tid;
    const cplx* f0 = raw(kVecK + 0) + tid;
    double dnf = 0.0, dny = 0.0;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const cplx y0 = yy[T * k], f = f0[T * k];
      double sk = atol + rtol * fabs(y0.re);
      double a = f.re / sk, b = y0.re / sk;
      dnf += a * a;
      dny += b * b;
      sk = atol + rtol * fabs(y0.im);
      a = f.im / sk;
      b = y0.im / sk;
      dnf += a * a;
      dny += b * b;
    }
    ctx.sum2(dnf, dny);
    const double h0 = c.hinit_first(dnf, dny);
    const int one_slot[1] = {0};
    const double one_val[1] = {1.0};
    pass(false, 0, true, h0, 1, one_slot, one_val, true, c.x + h0);  // ys = y + h0*f0, coefs(x+h0)
    pass(true, 1, false, 0.0, 0, nullptr, nullptr, false, 0.0);      // K[1] = f(x+h0, ys)
    const cplx* f1 = raw(kVecK + 1) + tid;
    double der2 = 0.0;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const cplx y0 = yy[T * k], f = f0[T * k], g = f1[T * k];
      double sk = atol + rtol * fabs(y0.re);
      double a = (g.re - f.re) / sk;
      der2 += a * a;
      sk = atol + rtol * fabs(y0.im);
      a = (g.im - f.im) / sk;
      der2 += a * a;
    }
    ctx.sum1(der2);
    c.hinit_finish(o, h0, dnf, der2);
  }

  __device__ void step(int& accepted) {
    const double x = c.x, h = c.h, tn = x + h;
    const int64_t n_real = 2 * P.dim;
    // stage-2 input and its coefficients
    pass(false, 0, true, h, tables::DOP853_a_nnz[0], tables::DOP853_a_slot[0], tables::DOP853_a_val[0], true,
         x + tables::DOP853_c[1] * h);
    for (int s = 0; s < 11; ++s) {
      const int dst = (s == 9) ? 1 : (s == 10) ? 2 : s + 1;
      const bool more = s < 10;
      const double t_next = (s + 1 == 10) ? tn : x + tables::DOP853_c[more ? s + 2 : 1] * h;
      pass(true, dst, more, h, more ? tables::DOP853_a_nnz[s + 1] : 0, more ? tables::DOP853_a_slot[s + 1] : nullptr,
           more ? tables::DOP853_a_val[s + 1] : nullptr, more, t_next);
    }
    c.nfcn += 11;
    // new solution, error sums, stiffness denominator; ynew -> slot 4 and -> the other ys buffer
    double err = 0.0, err3 = 0.0, stden = 0.0;
    {
      const int bn = tables::DOP853_b_nnz[0], e5n = tables::DOP853_e5_nnz[0], e3n = tables::DOP853_e3_nnz[0];
      const double atol = o.atol, rtol = o.rtol;
      const cplx* yy = raw(kVecY) + tid;
      const cplx* old = ys_raw(cur) + tid;
      cplx* ynew = raw(kVecK + 4) + tid;
      cplx* ysn = ys_raw(cur ^ 1) + tid;
      double sr[K], si[K], e5r[K], e5i[K], e3r[K], e3i[K];
#pragma unroll
      for (int k = 0; k < K; ++k) sr[k] = si[k] = e5r[k] = e5i[k] = e3r[k] = e3i[k] = 0.0;
      for (int j = 0; j < bn; ++j) {
        const cplx* kj = raw(kVecK + tables::DOP853_b_slot[0][j]) + tid;
        const double v = tables::DOP853_b_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj[T * k];
          sr[k] += v * z.re;
          si[k] += v * z.im;
        }
      }
      for (int j = 0; j < e5n; ++j) {
        const cplx* kj = raw(kVecK + tables::DOP853_e5_slot[0][j]) + tid;
        const double v = tables::DOP853_e5_val[0][j];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx z = kj[T * k];
          e5r[k] += v * z.re;
          e5i[k] += v * z.im;
        }
      }
      for (int j = 0; j < e3n; ++j) {
        const cplx* kj = raw(kVecK + tables::DOP853_e3_slot[0][j]) + tid;
        const double v = tables::DOP853_e3_val[0][j];
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
    ctx.sum3(err, err3, stden);  // contains the cluster barrier that publishes ynew in the other ys buffer
    cur ^= 1;
    accepted = c.decide(o, c.err_dop853(err, err3, n_real));
    if (accepted) {
      // coefficients at tn are the ones of stage 12, still in coefs[ccur]
      pass(true, 3, false, 0.0, 0, nullptr, nullptr, false, 0.0);  // K4 = f(x+h, ynew): next step's k1
      c.nfcn += 1;
      if (c.stiff_due(o)) {
        const cplx* k4 = raw(kVecK + 3) + tid;
        const cplx* k3 = raw(kVecK + 2) + tid;
        double stnum = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const cplx a = k4[T * k], b = k3[T * k];
          const double dr = a.re - b.re, di = a.im - b.im;
          stnum += dr * dr + di * di;
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

  // one integrate(x0 -> x1) call (fresh start like scipy's ode.integrate)
  __device__ void integrate(double x0, double x1) {
    c.begin(o, x0, x1);
    // ys = y ; coefs(x0)
    coefs_into(ccur, x0);
    {
      const cplx* yy = raw(kVecY) + tid;
      cplx* yn = ys_raw(cur) + tid;
#pragma unroll
      for (int k = 0; k < K; ++k) yn[T * k] = yy[T * k];
    }
    ctx.sync();
    pass(true, 0, false, 0.0, 0, nullptr, nullptr, false, 0.0);  // K[0] = f(x0, y)
    hinit();
    while (!c.done) {
      if (!c.step_begin(o)) break;
      int accepted = 0;
      step(accepted);
      c.step_end(o, accepted);
    }
  }
};

// ---- the kernel: persistent clusters pulling tasks from the work queue ---------------------------------
__global__ void __launch_bounds__(512, 1) evolve_cluster_kernel(const ClusterArgs CA) {
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ Coefs s_coefs[2];
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
      const int T = blockDim.x;
      const int lt = 31 - __clz(T);
      if (CA.fastK == 4) {
        Fast853<4> fe{ctx, A.P, ev.tv, V, s_vec, s_coefs, ev.o, ev.c, T, lt, lt + 2, A.P.n_atoms, (int)threadIdx.x, 0, 0};
        cluster_run_times(fe, ev, A, task, rank);
      } else if (CA.fastK == 2) {
        Fast853<2> fe{ctx, A.P, ev.tv, V, s_vec, s_coefs, ev.o, ev.c, T, lt, lt + 1, A.P.n_atoms, (int)threadIdx.x, 0, 0};
        cluster_run_times(fe, ev, A, task, rank);
      } else {
        Fast853<1> fe{ctx, A.P, ev.tv, V, s_vec, s_coefs, ev.o, ev.c, T, lt, lt, A.P.n_atoms, (int)threadIdx.x, 0, 0};
        cluster_run_times(fe, ev, A, task, rank);
      }
    }
    ctx.sync();
  }
  // no CTA may exit while a peer could still read its shared memory
  ctx.sync();
}

}  // namespace bqa
