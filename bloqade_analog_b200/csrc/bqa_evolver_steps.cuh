// bqa_evolver_steps.cuh - one attempted step of DOPRI5 and DOP853 (Hairer's cores as driven by
// scipy.integrate.ode; reference call site emulate/ir/state_vector.py:589-599).
// Included from bqa_evolver.cuh.
#pragma once

namespace bqa {

// ---- DOPRI5: 6 stages + FSAL stage at the new solution; K slots 0..5, k7 -> slot 1, ynew -> slot 6
template <class Ctx>
BQA_HD void TaskEvolver<Ctx>::step_dopri5(int& accepted) {
  const double x = c.x, h = c.h, tn = x + h;
  const int64_t n_real = 2 * P.dim;
  auto stage_time = [=](int s) { return (s == 5) ? tn : x + tables::DOPRI5_c[s + 1] * h; };
  step_coefs(6, stage_time);  // entries 0..4: stages 2..6, entry 5: the FSAL evaluation at x + h
  for (int s = 0; s < 5; ++s) {
    double ts = stage_time(s);
    stage_input(ts, h, tables::DOPRI5_a_nnz[s], tables::DOPRI5_a_slot[s], tables::DOPRI5_a_val[s], s);
    rhs_from_ys(ts, k(s + 1), s);
  }
  double err = 0.0, stden = 0.0, stnum = 0.0;
  {
    // new solution y1 = y + h*sum(b k); ys still holds the stage-6 input YSTI (rotated with the
    // same u(x+h) because c6 = 1, and |u| = 1 keeps the norm of the difference)
    cplx* out = ys();
    const cplx* yy = y();
    cplx* const* kk = w.vec + kVecK;
    cplx* ynew = k(6);
    const double* Vd = V;
    const bool rot = rotating();
    const int nnz = tables::DOPRI5_a_nnz[5];
    const int* slot = tables::DOPRI5_a_slot[5];
    const double* val = tables::DOPRI5_a_val[5];
    ctx.for_each(P.dim, [&](int64_t i) {
      double sr = 0.0, si = 0.0;
      for (int j = 0; j < nnz; ++j) {
        cplx kz = kk[slot[j]][i];
        sr += val[j] * kz.re;
        si += val[j] * kz.im;
      }
      cplx z = cmake(yy[i].re + h * sr, yy[i].im + h * si);
      ynew[i] = z;
      cplx old = out[i];
      cplx zr = rot ? cmul(phase_u(Vd[i], tn), z) : z;
      double dr = zr.re - old.re, di = zr.im - old.im;
      stden += dr * dr + di * di;
      out[i] = zr;
    });
    ctx.sync();
  }
  rhs_from_ys(tn, k(1), 5);  // k7 (FSAL), evaluated before the error test like Hairer's DOPRI5
  {
    const cplx* yy = y();
    cplx* const* kk = w.vec + kVecK;
    const cplx* ynew = k(6);
    const cplx* k2 = k(1);
    const cplx* k6 = k(5);
    const int en = tables::DOPRI5_e_nnz[0];
    const int* es = tables::DOPRI5_e_slot[0];
    const double* ev = tables::DOPRI5_e_val[0];
    const double atol = o.atol, rtol = o.rtol;
    ctx.for_each(P.dim, [&](int64_t i) {
      double er = 0.0, ei = 0.0;
      for (int j = 0; j < en; ++j) {
        cplx kz = kk[es[j]][i];
        er += ev[j] * kz.re;
        ei += ev[j] * kz.im;
      }
      er *= h;
      ei *= h;
      double sk = atol + rtol * fmax(fabs(yy[i].re), fabs(ynew[i].re));
      double q = er / sk;
      err += q * q;
      sk = atol + rtol * fmax(fabs(yy[i].im), fabs(ynew[i].im));
      q = ei / sk;
      err += q * q;
      double dr = k2[i].re - k6[i].re, di = k2[i].im - k6[i].im;
      stnum += dr * dr + di * di;
    });
  }
  ctx.sum3(err, stden, stnum);
  c.nfcn += 6;
  accepted = c.decide(o, c.err_dopri5(err, n_real));
  if (accepted) {
    if (c.stiff_due(o)) c.stiff_update(o, stnum, stden);
    swap_vec(kVecK + 0, kVecK + 1);  // k1 = k7
    swap_vec(kVecY, kVecK + 6);      // y = y1
  }
}

// ---- DOP853: 12 stages; K slots 0..9, k11 -> slot 1, k12 -> slot 2, ynew -> slot 4 (k5 is dead)
template <class Ctx>
BQA_HD void TaskEvolver<Ctx>::step_dop853(int& accepted) {
  if constexpr (Ctx::kFastRhs) {
    // CTA engine, two-level full space with real coefficients: register-carry form of the same step
    if (fast_rhs && w.fused_layout && ctx.step853_fused(*this, accepted)) return;
  }
  const double x = c.x, h = c.h, tn = x + h;
  const int64_t n_real = 2 * P.dim;
  auto stage_time = [=](int s) { return (s == 10) ? tn : x + tables::DOP853_c[s + 1] * h; };
  step_coefs(kCoefTab, stage_time);  // entry s = coefficients of stage s + 2; entry 10 (time x + h) also serves K4
  for (int s = 0; s < 11; ++s) {
    double ts = stage_time(s);
    stage_input(ts, h, tables::DOP853_a_nnz[s], tables::DOP853_a_slot[s], tables::DOP853_a_val[s], s);
    int dst = (s == 9) ? 1 : (s == 10) ? 2 : s + 1;
    rhs_from_ys(ts, k(dst), s);
  }
  c.nfcn += 11;
  double err = 0.0, err3 = 0.0, stden = 0.0;
  {
    cplx* out = ys();
    const cplx* yy = y();
    cplx* const* kk = w.vec + kVecK;
    cplx* ynew = k(4);
    const double* Vd = V;
    const bool rot = rotating();
    const int bn = tables::DOP853_b_nnz[0];
    const int* bs = tables::DOP853_b_slot[0];
    const double* bv = tables::DOP853_b_val[0];
    const int e5n = tables::DOP853_e5_nnz[0];
    const int* e5s = tables::DOP853_e5_slot[0];
    const double* e5v = tables::DOP853_e5_val[0];
    const int e3n = tables::DOP853_e3_nnz[0];
    const int* e3s = tables::DOP853_e3_slot[0];
    const double* e3v = tables::DOP853_e3_val[0];
    const double atol = o.atol, rtol = o.rtol;
    bool done = false;
    if constexpr (Ctx::kFastRhs) {
      if (!rot && ctx.exact_dim(P.dim)) {  // same sums in the same order, four elements per thread
        ctx.final853_fast(P.dim, h, atol, rtol, kk, yy, ynew, out, err, err3, stden);
        done = true;
      }
    }
    if (!done) ctx.for_each(P.dim, [&](int64_t i) {
      double sr = 0.0, si = 0.0, e5r = 0.0, e5i = 0.0, e3r = 0.0, e3i = 0.0;
      for (int j = 0; j < bn; ++j) {
        cplx kz = kk[bs[j]][i];
        sr += bv[j] * kz.re;
        si += bv[j] * kz.im;
      }
      for (int j = 0; j < e5n; ++j) {
        cplx kz = kk[e5s[j]][i];
        e5r += e5v[j] * kz.re;
        e5i += e5v[j] * kz.im;
      }
      for (int j = 0; j < e3n; ++j) {
        cplx kz = kk[e3s[j]][i];
        e3r += e3v[j] * kz.re;
        e3i += e3v[j] * kz.im;
      }
      cplx yo = yy[i];
      cplx z = cmake(yo.re + h * sr, yo.im + h * si);
      double sk = atol + rtol * fmax(fabs(yo.re), fabs(z.re));
      double q = e5r / sk;
      err += q * q;
      q = e3r / sk;
      err3 += q * q;
      sk = atol + rtol * fmax(fabs(yo.im), fabs(z.im));
      q = e5i / sk;
      err += q * q;
      q = e3i / sk;
      err3 += q * q;
      ynew[i] = z;  // slot 4 is referenced by none of the b / e5 / e3 rows
      cplx old = out[i];
      cplx zr = rot ? cmul(phase_u(Vd[i], tn), z) : z;
      double dr = zr.re - old.re, di = zr.im - old.im;
      stden += dr * dr + di * di;  // sum (ynew - y1_stage12)^2 for the stiffness test
      out[i] = zr;
    });
  }
  ctx.sum3(err, err3, stden);
  accepted = c.decide(o, c.err_dop853(err, err3, n_real));
  if (accepted) {
    rhs_from_ys(tn, k(3), 10);  // K4 = f(x+h, ynew): next step's k1
    c.nfcn += 1;
    if (c.stiff_due(o)) {
      const cplx* k4 = k(3);
      const cplx* k3 = k(2);
      double stnum = 0.0;
      ctx.for_each(P.dim, [&](int64_t i) {
        double dr = k4[i].re - k3[i].re, di = k4[i].im - k3[i].im;
        stnum += dr * dr + di * di;
      });
      ctx.sum1(stnum);
      c.stiff_update(o, stnum, stden);
    }
    swap_vec(kVecK + 0, kVecK + 3);  // k1 = K4
    swap_vec(kVecY, kVecK + 4);      // y = ynew
  }
}

}  // namespace bqa
