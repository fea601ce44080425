// hostsim.cpp - TEST-ONLY serial host build of the device algorithm headers.
//
// Compiles bloqade_analog_b200/csrc/bqa_core.cuh + bqa_evolver.cuh with g++ against a serial
// execution context so that the integrator / controller / waveform / RHS logic can be checked
// against the oracle on machines WITHOUT a GPU (`pytest -m "not gpu"`).  It is never loaded by the
// product package, bench.py or smoke(): the product path is CUDA only and fails loudly without it.
//
// Build: g++ -O2 -shared -fPIC -I bloqade_analog_b200/csrc tests/hostsim/hostsim.cpp -o tests/hostsim/libbqa_hostsim.so
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "bqa_evolver.cuh"

using namespace bqa;

struct HostCtx {
  static constexpr bool kFastRhs = false;
  template <class F>
  void for_each(int64_t n, F f) {
    for (int64_t i = 0; i < n; ++i) f(i);
  }
  void sync() {}
  void sync_local() {}
  template <class F>
  void for_small(int64_t n, F f) {
    for (int64_t i = 0; i < n; ++i) f(i);
  }
  PtrIn gather(const cplx* p) const {
    PtrIn g;
    g.p = p;
    return g;
  }
  void sum1(double&) {}
  void sum2(double&, double&) {}
  void sum3(double&, double&, double&) {}
  template <class F>
  void single(F f) {
    f();
  }
};

static PlanView make_plan(int n_atoms, int n_level, int64_t dim, const uint32_t* configs, const int32_t* partners) {
  PlanView P;
  memset(&P, 0, sizeof(P));
  P.n_atoms = n_atoms;
  P.n_level = n_level;
  P.dim = dim;
  P.full_space = (partners == nullptr);
  P.configs = configs;
  P.partners = partners;
  P.n_csr = 0;
  P.csr = nullptr;
  uint32_t p = 1;
  for (int a = 0; a < kMaxAtoms; ++a) {
    P.pow_level[a] = p;
    p *= (uint32_t)n_level;
  }
  return P;
}

extern "C" {

double hostsim_wf_eval(const double* rec, int w, double t) {
  TaskView tv = task_view(rec);
  return wf_eval(tv, w, t);
}

int64_t hostsim_record_size(const double* rec) { return task_record_size(rec); }

// out = H(t) psi (raw != 0) or -i H(t) psi (raw == 0); interaction picture when `interaction`
void hostsim_apply(int n_atoms, int n_level, int64_t dim, const uint32_t* configs, const int32_t* partners,
                   const double* V, const double* rec, double t, int interaction, int raw, const double* psi,
                   double* out) {
  PlanView P = make_plan(n_atoms, n_level, dim, configs, partners);
  TaskView tv = task_view(rec);
  Coefs cf;
  memset(&cf, 0, sizeof(cf));
  int ne = n_atoms * (n_level - 1);
  for (int e = 0; e < ne; ++e) coef_entry(tv, t, e, cf);
  const cplx* in = (const cplx*)psi;
  std::vector<cplx> rot;
  if (interaction && V) {
    rot.resize(dim);
    for (int64_t i = 0; i < dim; ++i) rot[i] = cmul(phase_u(V[i], t), in[i]);
    in = rot.data();
  }
  cplx* o = (cplx*)out;
  for (int64_t i = 0; i < dim; ++i) {
    cplx hz = h_apply_dispatch(P, cf, interaction ? nullptr : V, in, i);
    if (interaction && V) hz = cmul_conj_a(phase_u(V[i], t), hz);
    o[i] = raw ? hz : times_minus_i(hz);
  }
}

// Evolve psi0 through the output times; y_out[n_times][dim]; stats[n_times][6] =
// {status, nfcn, nstep, naccpt, nrejct, h_last}.  Returns the first non-OK status or 1.
int hostsim_evolve(int n_atoms, int n_level, int64_t dim, const uint32_t* configs, const int32_t* partners,
                   const double* V, const double* rec, int solver, double atol, double rtol, int64_t nsteps,
                   int interaction, int reject_mode, const double* psi0, const double* times, int n_times,
                   double* y_out, double* stats) {
  PlanView P = make_plan(n_atoms, n_level, dim, configs, partners);
  std::vector<cplx> store((size_t)kNumVec * dim);
  cplx* vec[kNumVec];
  for (int v = 0; v < kNumVec; ++v) vec[v] = store.data() + (size_t)v * dim;
  memcpy(vec[kVecY], psi0, sizeof(cplx) * dim);
  Coefs cf;
  memset(&cf, 0, sizeof(cf));
  Workspace w;
  w.vec = vec;
  w.coefs = &cf;
  HostCtx ctx;
  SolverOpts o = default_opts(solver, atol, rtol, nsteps, interaction);
  o.reject_mode = reject_mode;
  TaskEvolver<HostCtx> ev(ctx, P, rec, V, w, o);
  double t_cur = 0.0;
  int ret = 1;
  for (int k = 0; k < n_times; ++k) {
    double* st = stats + 6 * k;
    if (times[k] != t_cur) {
      ev.integrate(t_cur, times[k]);
      st[0] = ev.c.status;
      st[1] = (double)ev.c.nfcn;
      st[2] = (double)ev.c.nstep;
      st[3] = (double)ev.c.naccpt;
      st[4] = (double)ev.c.nrejct;
      st[5] = ev.c.h;
      if (ev.c.status != kStatusOk) return ev.c.status;
      t_cur = times[k];
    } else {
      st[0] = 1;
      st[1] = st[2] = st[3] = st[4] = st[5] = 0;
    }
    ev.emit_state(t_cur, (cplx*)y_out + (size_t)k * dim);
  }
  return ret;
}
}
