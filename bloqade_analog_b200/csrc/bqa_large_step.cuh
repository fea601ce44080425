// bqa_large_step.cuh - step-closing kernels and the host launch sequence of the streamed engine.
#pragma once

namespace bqa {

// DOP853: ynew = y + h sum(b K) -> slot 4 and (rotated) -> ys; partial sums (err5, err3, stden)
__global__ void __launch_bounds__(kLgThreads) lg_final853(LargeCtrl* ctl, int64_t dim, const double* __restrict__ V) {
  if (!ctl->active) return;
  __shared__ double red[32];
  const double h = ctl->c.h, tn = ctl->c.x + ctl->c.h;
  const double atol = ctl->o.atol, rtol = ctl->o.rtol;
  const bool rot = ctl->o.interaction != 0 && V != nullptr;
  const cplx* y = ctl->vec[kVecY];
  cplx* ys = ctl->vec[kVecYs];
  cplx* ynew = ctl->vec[kVecK + 4];
  cplx* const* kk = ctl->vec + kVecK;
  const int bn = dev_tables::DOP853_b_nnz[0], e5n = dev_tables::DOP853_e5_nnz[0], e3n = dev_tables::DOP853_e3_nnz[0];
  double err = 0.0, err3 = 0.0, stden = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    double sr = 0.0, si = 0.0, e5r = 0.0, e5i = 0.0, e3r = 0.0, e3i = 0.0;
    // b, e5 and e3 rows share the slot pattern {0,5,6,7,8,9,1,2}; e3 additionally uses slot 8 (k9)
    for (int j = 0; j < bn; ++j) {
      cplx kz = kk[dev_tables::DOP853_b_slot[0][j]][i];
      sr += dev_tables::DOP853_b_val[0][j] * kz.re;
      si += dev_tables::DOP853_b_val[0][j] * kz.im;
    }
    for (int j = 0; j < e5n; ++j) {
      cplx kz = kk[dev_tables::DOP853_e5_slot[0][j]][i];
      e5r += dev_tables::DOP853_e5_val[0][j] * kz.re;
      e5i += dev_tables::DOP853_e5_val[0][j] * kz.im;
    }
    for (int j = 0; j < e3n; ++j) {
      cplx kz = kk[dev_tables::DOP853_e3_slot[0][j]][i];
      e3r += dev_tables::DOP853_e3_val[0][j] * kz.re;
      e3i += dev_tables::DOP853_e3_val[0][j] * kz.im;
    }
    const cplx yo = y[i];
    const cplx z = cmake(yo.re + h * sr, yo.im + h * si);
    double sk = atol + rtol * fmax(fabs(yo.re), fabs(z.re));
    double q = e5r / sk; err += q * q;
    q = e3r / sk; err3 += q * q;
    sk = atol + rtol * fmax(fabs(yo.im), fabs(z.im));
    q = e5i / sk; err += q * q;
    q = e3i / sk; err3 += q * q;
    ynew[i] = z;
    const cplx old = ys[i];
    const cplx zr = rot ? cmul(phase_u(V[i], tn), z) : z;
    const double dr = zr.re - old.re, di = zr.im - old.im;
    stden += dr * dr + di * di;
    ys[i] = zr;
  }
  lg_store3(ctl, err, err3, stden, red);
}

// DOPRI5 part a: ynew -> slot 6 and (rotated) -> ys; partial stden
__global__ void __launch_bounds__(kLgThreads) lg_final5a(LargeCtrl* ctl, int64_t dim, const double* __restrict__ V) {
  if (!ctl->active) return;
  __shared__ double red[32];
  const double h = ctl->c.h, tn = ctl->c.x + ctl->c.h;
  const bool rot = ctl->o.interaction != 0 && V != nullptr;
  const cplx* y = ctl->vec[kVecY];
  cplx* ys = ctl->vec[kVecYs];
  cplx* ynew = ctl->vec[kVecK + 6];
  cplx* const* kk = ctl->vec + kVecK;
  const int nnz = dev_tables::DOPRI5_a_nnz[5];
  double stden = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    double sr = 0.0, si = 0.0;
    for (int j = 0; j < nnz; ++j) {
      cplx kz = kk[dev_tables::DOPRI5_a_slot[5][j]][i];
      sr += dev_tables::DOPRI5_a_val[5][j] * kz.re;
      si += dev_tables::DOPRI5_a_val[5][j] * kz.im;
    }
    const cplx yo = y[i];
    const cplx z = cmake(yo.re + h * sr, yo.im + h * si);
    ynew[i] = z;
    const cplx old = ys[i];
    const cplx zr = rot ? cmul(phase_u(V[i], tn), z) : z;
    const double dr = zr.re - old.re, di = zr.im - old.im;
    stden += dr * dr + di * di;
    ys[i] = zr;
  }
  stden = block_sum(stden, red);
  if (threadIdx.x == 0) ctl->partial[3 * ctl->n_blocks + blockIdx.x] = stden;
}

// DOPRI5 part b (after k7 = f(ynew) in slot 1): partial (err, stnum)
__global__ void __launch_bounds__(kLgThreads) lg_final5b(LargeCtrl* ctl, int64_t dim) {
  if (!ctl->active) return;
  __shared__ double red[32];
  const double h = ctl->c.h;
  const double atol = ctl->o.atol, rtol = ctl->o.rtol;
  const cplx* y = ctl->vec[kVecY];
  const cplx* ynew = ctl->vec[kVecK + 6];
  const cplx* k2 = ctl->vec[kVecK + 1];
  const cplx* k6 = ctl->vec[kVecK + 5];
  cplx* const* kk = ctl->vec + kVecK;
  const int en = dev_tables::DOPRI5_e_nnz[0];
  double err = 0.0, stnum = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    double er = 0.0, ei = 0.0;
    for (int j = 0; j < en; ++j) {
      cplx kz = kk[dev_tables::DOPRI5_e_slot[0][j]][i];
      er += dev_tables::DOPRI5_e_val[0][j] * kz.re;
      ei += dev_tables::DOPRI5_e_val[0][j] * kz.im;
    }
    er *= h;
    ei *= h;
    double sk = atol + rtol * fmax(fabs(y[i].re), fabs(ynew[i].re));
    double q = er / sk; err += q * q;
    sk = atol + rtol * fmax(fabs(y[i].im), fabs(ynew[i].im));
    q = ei / sk; err += q * q;
    const double dr = k2[i].re - k6[i].re, di = k2[i].im - k6[i].im;
    stnum += dr * dr + di * di;
  }
  lg_store3(ctl, err, stnum, 0.0, red);
}

// DOP853 stiffness numerator sum (K4new - k12)^2, only when due
__global__ void __launch_bounds__(kLgThreads) lg_stnum853(LargeCtrl* ctl, int64_t dim) {
  if (!ctl->active || !ctl->need_stnum) return;
  __shared__ double red[32];
  const cplx* k4 = ctl->vec[kVecK + 3];
  const cplx* k3 = ctl->vec[kVecK + 2];
  double stnum = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    const double dr = k4[i].re - k3[i].re, di = k4[i].im - k3[i].im;
    stnum += dr * dr + di * di;
  }
  lg_store3(ctl, stnum, 0.0, 0.0, red);
}

__device__ __forceinline__ void lg_swap(LargeCtrl* ctl, int a, int b) {
  cplx* t = ctl->vec[a];
  ctl->vec[a] = ctl->vec[b];
  ctl->vec[b] = t;
}

// accept / reject decision (DOP853: before the 13th RHS; DOPRI5: closes the step)
__global__ void lg_ctl_decide(LargeCtrl* ctl, int64_t dim) {
  if (!ctl->active) return;
  Controller& c = ctl->c;
  const SolverOpts& o = ctl->o;
  double a, b, s3;
  lg_total3(ctl, a, b, s3);
  if (o.solver == kSolverDopri5) {
    double stden = 0.0;
    for (int r = 0; r < (ctl->world > 1 ? ctl->world : 1); ++r) {
      const volatile double* part = lg_partial_of(ctl, r);
      for (int k = 0; k < ctl->n_blocks; ++k) stden += part[3 * ctl->n_blocks + k];
    }
    c.nfcn += 6;
    const int acc = c.decide(o, c.err_dopri5(a, 2 * dim));
    ctl->accepted = acc;
    if (acc) {
      if (c.stiff_due(o)) c.stiff_update(o, b, stden);
      lg_swap(ctl, kVecK + 0, kVecK + 1);
      lg_swap(ctl, kVecY, kVecK + 6);
    }
    c.step_end(o, acc);
  } else {
    c.nfcn += 11;
    ctl->stden = s3;
    const int acc = c.decide(o, c.err_dop853(a, b, 2 * dim));
    ctl->accepted = acc;
    if (acc) {
      c.nfcn += 1;
      ctl->need_stnum = c.stiff_due(o) ? 1 : 0;
    }
  }
}

// DOP853 only: closes the step after the 13th RHS (+ optional stiffness sum)
__global__ void lg_ctl_end853(LargeCtrl* ctl) {
  if (!ctl->active) return;
  Controller& c = ctl->c;
  if (ctl->accepted) {
    if (ctl->need_stnum) {
      double a, b, s3;
      lg_total3(ctl, a, b, s3);
      c.stiff_update(ctl->o, a, ctl->stden);
    }
    lg_swap(ctl, kVecK + 0, kVecK + 3);
    lg_swap(ctl, kVecY, kVecK + 4);
  }
  c.step_end(ctl->o, ctl->accepted);
}

// out = (u(t)*) y, stats.  ran = 0: no integration for this output time (t equals the previous one, or an earlier
// output time of the task failed - then the entry is marked kStatusInconsistent like evolve_small_kernel does)
__global__ void __launch_bounds__(kLgThreads)
lg_emit(LargeCtrl* ctl, int64_t dim, const double* __restrict__ V, double t, cplx* out, TaskStats* st, int ran, int failed) {
  const bool rot = ctl->o.interaction != 0 && V != nullptr;
  const cplx* y = ctl->vec[kVecY];
  if (out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
      const cplx z = y[i];
      out[i] = rot ? cmul(phase_u(V[i], t), z) : z;
    }
  }
  if (st && blockIdx.x == 0 && threadIdx.x == 0) {
    TaskStats s;
    s.pad = 0;
    if (ran) {
      s.status = ctl->c.status;
      s.nfcn = ctl->c.nfcn;
      s.nstep = ctl->c.nstep;
      s.naccpt = ctl->c.naccpt;
      s.nrejct = ctl->c.nrejct;
      s.h_last = ctl->c.h;
      s.t_last = ctl->c.x;
    } else {
      s.status = failed ? kStatusInconsistent : kStatusOk;
      s.nfcn = s.nstep = s.naccpt = s.nrejct = 0;
      s.h_last = 0.0;
      s.t_last = t;
    }
    *st = s;
  }
}

// |0...0> (of a sharded state: the unit amplitude is on rank 0)
__global__ void lg_init_state(cplx* y, int64_t dim, int rank) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = cmake((i == 0 && rank == 0) ? 1.0 : 0.0, 0.0);
}

}  // namespace bqa
