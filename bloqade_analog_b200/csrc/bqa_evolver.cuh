// bqa_evolver.cuh - whole-task adaptive evolution on ONE execution context.
//
// This is north_star kernel (5) ("batched small-state kernels that evolve many parameter-sweep tasks
// per launch") together with kernel (2) (DOPRI5/DOP853 stage combine + error norm, step resident on
// the GPU): one CTA owns one task and runs the whole AnalogGate._apply /
// _apply_interaction_picture flow (emulate/ir/state_vector.py:574-641) - including SciPy's restart
// at every requested output time - without returning to the host.
//
// Ctx contract (device: one CTA, bqa_small.cuh; host: serial loop, tests/hostsim):
//   for_each(n, f)   run f(i) for all 0 <= i < n, partitioned over the context's threads
//   sync()           barrier + memory visibility inside the context
//   sum1/2/3(...)    replace each thread's partial by the total over the context (same value and
//                    same summation order in every thread, so control flow stays uniform)
//   single(f)        f() executed by exactly one thread (per CTA), followed by a barrier
//   for_small(n, f)  like for_each but NOT partitioned over the CTAs of a cluster: every CTA runs all n
//                    (per-call coefficient tables are replicated in each CTA's shared memory)
//   sync_local()     barrier that only has to order this CTA's own shared-memory traffic
//   gather(ptr)      accessor `in[p]` for the stage vector, valid for ANY index p of the state (a
//                    cluster context resolves indices owned by a peer CTA through DSMEM)
#pragma once
#include "bqa_core.cuh"

namespace bqa {

// vec[0] = y (current solution), vec[1] = ys (stage input / gather source), vec[2+s] = K slot s.
// The table itself lives in shared memory on the GPU so that slot indices read from the Butcher
// tables can index it dynamically without spilling pointers to local memory.
constexpr int kVecY = 0;
constexpr int kVecYs = 1;
constexpr int kVecK = 2;
constexpr int kNumVec = 12;

constexpr int kCoefTab = 11;  // stage times of one DOP853 step after the first stage (c2 h ... c11 h, h)

struct Workspace {
  cplx** vec;     // [kNumVec] pointer table (shared)
  Coefs* coefs;   // shared scratch for the per-call coefficients
  // optional (contexts with room for it): coefficient tables of ALL stage times of the current step,
  // evaluated in one pass at the start of the step (step_coefs) instead of one serial evaluation in front of
  // every stage - ncu showed the whole CTA waiting at the stage barrier for the one warp that evaluates the
  // waveforms (19 % of the stall samples of evolve_small_kernel).  nullptr: evaluate per stage.
  Coefs* coefs_tab = nullptr;
  // y, ys and K slot 0 sit in on-chip memory under pointers that must not be swapped (CtaCtxT::step853_fused)
  bool fused_layout = false;
};

template <class Ctx>
struct TaskEvolver {
  Ctx& ctx;
  const PlanView& P;
  TaskView tv;
  const double* V;   // static (Rydberg) diagonal or nullptr
  Workspace w;
  SolverOpts o;
  Controller c;
  // two-level full space, no CSR term, Schroedinger picture, no phase waveform: contexts with kFastRhs evaluate
  // the RHS with their own fused routine (same arithmetic in the same order, so the results are bit-identical)
  bool fast_rhs;
  bool sub_rhs;  // the same for a two-level blockade subspace (int32 gather partner table): Ctx::rhs_sub2_real
  bool fast_rhs_phase;  // the same operator shape with phase waveforms (complex coefficients), either picture
  bool fast_rhs_ip;  // fast_rhs's operator shape in the INTERACTION picture: the same routine on phi = u psi + rotation
  bool coefs_ready = false;  // the shared coefficient scratch holds the values of time coefs_t
  double coefs_t = 0.0;
  int tab_n = 0;             // entries of w.coefs_tab valid for the current step

  BQA_HD TaskEvolver(Ctx& ctx_, const PlanView& P_, const double* rec, const double* V_, Workspace w_,
                     SolverOpts o_)
      : ctx(ctx_), P(P_), tv(task_view(rec)), V(V_), w(w_), o(o_) {
    fast_rhs = P.full_space && P.n_level == 2 && P.n_csr == 0 && o.interaction == 0 && P.dim <= ((int64_t)1 << 30) &&
               !task_has_phase(tv);
    fast_rhs_phase = P.full_space && P.n_level == 2 && P.n_csr == 0 && P.dim <= ((int64_t)1 << 30) && task_has_phase(tv);
    fast_rhs_ip = P.full_space && P.n_level == 2 && P.n_csr == 0 && o.interaction != 0 && P.dim <= ((int64_t)1 << 30) &&
                  !task_has_phase(tv);
    sub_rhs = !P.full_space && P.n_level == 2 && P.n_csr == 0 && o.interaction == 0 && P.dim <= ((int64_t)1 << 28) &&
              P.configs != nullptr && P.partners != nullptr && !task_has_phase(tv);
  }

  BQA_HD cplx* y() const { return w.vec[kVecY]; }
  BQA_HD cplx* ys() const { return w.vec[kVecYs]; }
  BQA_HD cplx* k(int s) const { return w.vec[kVecK + s]; }
  BQA_HD bool rotating() const { return o.interaction != 0 && V != nullptr; }

  BQA_HD void swap_vec(int a, int b) {
    cplx** vec = w.vec;
    ctx.single([&]() {
      cplx* t = vec[a];
      vec[a] = vec[b];
      vec[b] = t;
    });
  }

  // coefficients at time t into the shared scratch; visible after the caller's next barrier
  BQA_HD void eval_coefs_nosync(double t) {
    int ne = tv.n_atoms * (P.n_level - 1);
    Coefs* cf = w.coefs;
    const TaskView& tvr = tv;
    const PlanView& Pr = P;
    ctx.for_small(ne + P.n_csr, [&](int64_t e) {
      if (e < ne)
        coef_entry(tvr, t, (int)e, *cf);
      else
        coef_csr_entry(Pr, tvr, t, (int)(e - ne), *cf);
    });
  }
  BQA_HD void eval_coefs(double t) {
    eval_coefs_nosync(t);
    ctx.sync_local();
  }

  // coefficient tables of the stage times  time_of(0) ... time_of(n-1)  of the step that starts now, all in one
  // pass (one table entry per thread).  Same function of the same time values as eval_coefs: identical numbers.
  template <class TimeOf>
  BQA_HD void step_coefs(int n, TimeOf time_of) {
    tab_n = 0;
    if (w.coefs_tab == nullptr) return;
    const int ne = tv.n_atoms * (P.n_level - 1), per = ne + P.n_csr;
    Coefs* tab = w.coefs_tab;
    const TaskView& tvr = tv;
    const PlanView& Pr = P;
    ctx.for_small((int64_t)n * per, [&](int64_t q) {
      const int j = (int)(q / per), e = (int)(q - (int64_t)j * per);
      const double t = time_of(j);
      if (e < ne)
        coef_entry(tvr, t, e, tab[j]);
      else
        coef_csr_entry(Pr, tvr, t, e - ne, tab[j]);
    });
    ctx.sync_local();
    tab_n = n;
  }
  BQA_HD const Coefs* tab_entry(int idx) const { return (idx >= 0 && idx < tab_n) ? w.coefs_tab + idx : nullptr; }

  // kout = f(t, .) evaluated on the stage input held in ys (already multiplied by u(t) in the
  // interaction picture)
  BQA_HD void rhs_from_ys(double t, cplx* kout, int tab = -1) {
    const Coefs* pre = tab_entry(tab);  // evaluated by step_coefs at the start of the step
    // otherwise stage_input(t, ...) has already evaluated the coefficients of this time (one barrier less per stage)
    if (pre == nullptr && !(coefs_ready && coefs_t == t)) eval_coefs(t);
    coefs_ready = false;
    const PlanView& Pr = P;
    const Coefs& cf = pre ? *pre : *w.coefs;
    const auto in = ctx.gather(ys());
    const double* Vd = V;
    const bool inter = o.interaction != 0;
    if constexpr (Ctx::kFastRhs) {
      if (fast_rhs) {
        ctx.rhs_full2_real(P.n_atoms, P.dim, cf, Vd, ys(), kout);
        ctx.sync();
        return;
      }
      if (fast_rhs_phase) {
        const bool rotv = inter && Vd != nullptr;
        ctx.rhs_full2_cplx(P.n_atoms, P.dim, cf, rotv ? nullptr : Vd, ys(), kout, rotv ? Vd : nullptr, t);
        ctx.sync();
        return;
      }
      if (fast_rhs_ip) {
        ctx.rhs_full2_real(P.n_atoms, P.dim, cf, nullptr, ys(), kout, Vd, t);
        ctx.sync();
        return;
      }
      if (sub_rhs) {
        ctx.rhs_sub2_real(P.n_atoms, P.dim, cf, Vd, P.configs, P.partners, ys(), kout);
        ctx.sync();
        return;
      }
    }
    ctx.for_each(P.dim, [&](int64_t i) {
      cplx hz = h_apply_dispatch(Pr, cf, inter ? nullptr : Vd, in, i);
      if (inter && Vd) hz = cmul_conj_a(phase_u(Vd[i], t), hz);
      kout[i] = times_minus_i(hz);
    });
    ctx.sync();
  }

  // ys = (u(t) *) [ y + h * sum_j val[j] * K[slot[j]] ]     (nnz = 0: ys = (u(t)*) y)
  BQA_HD void stage_input(double t, double h, int nnz, const int* slot, const double* val, int tab = -1) {
    // the RHS that follows is evaluated at the same time t: its coefficients are computed here (the previous RHS
    // is complete) and become visible with the barrier that ends this pass - unless step_coefs has them already
    if (tab_entry(tab) == nullptr) {
      eval_coefs_nosync(t);
      coefs_t = t;
      coefs_ready = true;
    }
    cplx* out = ys();
    const cplx* yy = y();
    cplx* const* kk = w.vec + kVecK;
    const double* Vd = V;
    const bool rot = rotating();
    if constexpr (Ctx::kFastRhs) {
      // same sums in the same order, four elements per thread and the term loop outermost (+ the u(t) rotation in
      // the interaction picture)
      ctx.stage_input_plain(P.dim, h, nnz, slot, val, kk, yy, out, rot ? Vd : nullptr, t);
      ctx.sync();
      return;
    }
    ctx.for_each(P.dim, [&](int64_t i) {
      double sr = 0.0, si = 0.0;
      for (int j = 0; j < nnz; ++j) {
        cplx kz = kk[slot[j]][i];
        sr += val[j] * kz.re;
        si += val[j] * kz.im;
      }
      cplx z = cmake(yy[i].re + h * sr, yy[i].im + h * si);
      out[i] = rot ? cmul(phase_u(Vd[i], t), z) : z;
    });
    ctx.sync();
  }

  // automatic initial step (Hairer HINIT); K[0] = f(x, y) must be valid.  Uses K[1] as f1.
  BQA_HD void hinit() {
    const cplx* yy = y();
    const cplx* f0 = k(0);
    double dnf = 0.0, dny = 0.0;
    const double atol = o.atol, rtol = o.rtol;
    ctx.for_each(P.dim, [&](int64_t i) {
      double sk = atol + rtol * fabs(yy[i].re);
      double a = f0[i].re / sk, b = yy[i].re / sk;
      dnf += a * a;
      dny += b * b;
      sk = atol + rtol * fabs(yy[i].im);
      a = f0[i].im / sk;
      b = yy[i].im / sk;
      dnf += a * a;
      dny += b * b;
    });
    ctx.sum2(dnf, dny);
    double h0 = c.hinit_first(dnf, dny);
    const int one_slot[1] = {0};
    const double one_val[1] = {1.0};
    stage_input(c.x + h0, h0, 1, one_slot, one_val);  // y1 = y + h0 * f0
    rhs_from_ys(c.x + h0, k(1));
    const cplx* f1 = k(1);
    double der2 = 0.0;
    ctx.for_each(P.dim, [&](int64_t i) {
      double sk = atol + rtol * fabs(yy[i].re);
      double a = (f1[i].re - f0[i].re) / sk;
      der2 += a * a;
      sk = atol + rtol * fabs(yy[i].im);
      a = (f1[i].im - f0[i].im) / sk;
      der2 += a * a;
    });
    ctx.sum1(der2);
    c.hinit_finish(o, h0, dnf, der2);
  }

  BQA_HD void step_dopri5(int& accepted);
  BQA_HD void step_dop853(int& accepted);

  // one integrate(x0 -> x1) call of scipy's ode (a fresh start: k1, hinit, counters)
  BQA_HD void integrate(double x0, double x1) {
    c.begin(o, x0, x1);
    stage_input(x0, 0.0, 0, nullptr, nullptr);
    rhs_from_ys(x0, k(0));
    hinit();
    while (!c.done) {
      if (!c.step_begin(o)) break;
      int accepted = 0;
      if (o.solver == kSolverDopri5)
        step_dopri5(accepted);
      else
        step_dop853(accepted);
      c.step_end(o, accepted);
    }
  }

  // Schroedinger-picture state at the current time into `out` (SV:636-638 for the rotation)
  BQA_HD void emit_state(double t, cplx* out) {
    const cplx* yy = y();
    const double* Vd = V;
    const bool rot = rotating();
    ctx.for_each(P.dim, [&](int64_t i) {
      cplx z = yy[i];
      out[i] = rot ? cmul(phase_u(Vd[i], t), z) : z;
    });
    ctx.sync();
  }
};

}  // namespace bqa

#include "bqa_evolver_steps.cuh"
