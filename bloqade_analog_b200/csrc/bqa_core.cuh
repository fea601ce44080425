// bqa_core.cuh - device-side building blocks shared by every kernel of the B200 emulator backend.
//
//   * task record decoding + on-device waveform evaluation (north_star kernel (1): "per-term waveform
//     coefficients evaluated on-device"); mirrors the interpreter semantics of
//     ir/control/waveform.py:282-296,330-335,369-382,668-708 through the flattened piece table built
//     by bloqade_analog_b200/ir.py:flatten_waveform
//   * the fused RHS element  out[i] = -i * (H(t) psi)[i]  of
//     RydbergHamiltonian._ode_complex_kernel / _ode_complex_kernel_int
//     (emulate/ir/state_vector.py:323-359), with DetuningOperator.get_diagonal (:266-270),
//     RabiOperator.dot (:279-288) and the IndexMapping / CSR / CSC matvecs
//     (emulate/sparse_operator.py:15-34,99-128) folded into ONE gather pass
//   * Hairer's DOPRI5 / DOP853 step controller exactly as SciPy's `ode` drives it
//     (state_vector.py:589-599; scipy.integrate._ode.dopri5/dop853: safety 0.9, beta 0,
//     ifactor/dfactor 10/0.2 resp. 6/0.3, automatic initial step, restart per integrate() call)
//
// Everything here is written against an execution context `Ctx` (for_each / sum / sync) so that the
// same source runs (a) as one CTA per task on the GPU (bqa_small.cuh) and (b) serially on the host
// inside the test-only simulator tests/hostsim (compiled with g++, never loaded by the product).
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define BQA_HD __host__ __device__ __forceinline__
#define BQA_D __device__ __forceinline__
#else
#define BQA_HD inline
#define BQA_D inline
#endif

namespace bqa {

struct alignas(16) cplx {
  double re, im;
};

BQA_HD cplx cmake(double re, double im) {
  cplx z;
  z.re = re;
  z.im = im;
  return z;
}

constexpr int kMaxAtoms = 32;   // uint32 configurations (space.py:59-60)
constexpr int kMaxSlots = 2;    // exchange slots: two-level {g<->r}; three-level {g<->h, h<->r}

// status codes of scipy's dopri5/dop853 (state_vector.py:494-508)
enum : int {
  kStatusOk = 1,
  kStatusInconsistent = -1,
  kStatusNeedSteps = -2,
  kStatusStepTooSmall = -3,
  kStatusStiff = -4,
  kStatusPeerAbort = -6,  // sharded state only (no scipy counterpart): a peer rank failed / a cross-rank barrier timed out
};

enum : int { kSolverDopri5 = 5, kSolverDop853 = 8 };

// ------------------------------------------------------------------------------------------------
// Plan: the structure of the Hamiltonian shared by all tasks of a sweep (device pointers)
// ------------------------------------------------------------------------------------------------
struct CsrTerm {           // generic SparseMatrixCSR fallback term: out += A*M*psi (+ conj(A)*M^T*psi)
  int amp_wf;              // waveform index of the amplitude (Omega), value is halved like SV:280
  int phase_wf;            // waveform index of the phase or -1 (then only M is applied, SV:281-282)
  const double* data;      // [nnz]      M in CSR
  const int32_t* indices;  // [nnz]
  const int64_t* indptr;   // [dim+1]
  const double* t_data;    // M^T in CSR (the reference's CSC scatter, SO:15-21, turned into a gather)
  const int32_t* t_indices;
  const int64_t* t_indptr;
};

struct PlanView {
  int n_atoms;
  int n_level;               // 2 or 3
  int full_space;            // 1: configuration == index, partner by digit arithmetic
  int n_csr;
  int64_t dim;
  const uint32_t* configs;   // [dim] (subspace) or nullptr
  const int32_t* partners;   // [n_atoms][n_level-1][dim] gather partners, -1 = none; nullptr if full
  const CsrTerm* csr;        // [n_csr]
  uint32_t pow_level[kMaxAtoms];  // n_level^a
};

// ------------------------------------------------------------------------------------------------
// Task record (see include/bqa_b200.h "task record layout"): a flat array of float64
// ------------------------------------------------------------------------------------------------
constexpr int kTaskHeader = 8;
constexpr int kPieceStride = 6;  // lo, hi, offset, scale, coef_start, n_coef

struct TaskView {
  double duration;
  int diag_id;
  int n_wf, n_pieces, n_det, n_rabi, n_coef, n_atoms;
  const double* wf_start;  // [n_wf+1] first piece of each waveform
  const double* pieces;    // [n_pieces][6]
  const double* coefs;     // [n_coef]
  const double* det;       // [n_det][2+n_atoms]   wf, slot, scale[a]
  const double* rabi;      // [n_rabi][3+n_atoms]  amp_wf, phase_wf, slot, scale[a]
};

BQA_HD TaskView task_view(const double* rec) {
  TaskView tv;
  tv.duration = rec[0];
  tv.diag_id = (int)rec[1];
  tv.n_wf = (int)rec[2];
  tv.n_pieces = (int)rec[3];
  tv.n_det = (int)rec[4];
  tv.n_rabi = (int)rec[5];
  tv.n_coef = (int)rec[6];
  tv.n_atoms = (int)rec[7];
  const double* p = rec + kTaskHeader;
  tv.wf_start = p;
  p += tv.n_wf + 1;
  tv.pieces = p;
  p += (int64_t)tv.n_pieces * kPieceStride;
  tv.coefs = p;
  p += tv.n_coef;
  tv.det = p;
  p += (int64_t)tv.n_det * (2 + tv.n_atoms);
  tv.rabi = p;
  return tv;
}

// does any Rabi term of the task carry a phase waveform?  (No: every coefficient c[s][a] is real, c_im == 0 exactly.)
BQA_HD bool task_has_phase(const TaskView& tv) {
  for (int k = 0; k < tv.n_rabi; ++k)
    if ((int)tv.rabi[(int64_t)k * (3 + tv.n_atoms) + 1] >= 0) return true;
  return false;
}

BQA_HD int64_t task_record_size(const double* rec) {
  TaskView tv = task_view(rec);
  return (tv.rabi - rec) + (int64_t)tv.n_rabi * (3 + tv.n_atoms);
}

// value of waveform `w` at time t: sum of the active polynomial pieces (Horner in local time)
BQA_HD double wf_eval(const TaskView& tv, int w, double t) {
  int p0 = (int)tv.wf_start[w], p1 = (int)tv.wf_start[w + 1];
  double total = 0.0;
  for (int p = p0; p < p1; ++p) {
    const double* pc = tv.pieces + (int64_t)p * kPieceStride;
    if (t > pc[0] && t <= pc[1]) {
      double u = t - pc[2];
      int c0 = (int)pc[4], nc = (int)pc[5];
      double acc = 0.0;
      for (int k = nc - 1; k >= 0; --k) acc = acc * u + tv.coefs[c0 + k];
      total += pc[3] * acc;
    }
  }
  return total;
}

// Per-RHS-call coefficients: every reference term collapses onto per-atom numbers.
//   delta[s][a] : sum over detuning terms acting on digit s+1 of  f(t) * scale_a      (HC:98-118, SV:266-270)
//   c[s][a]     : sum over Rabi terms exchanging digits s<->s+1 of (Omega(t)/2) e^{-i phi(t)} * scale_a
//                 (SV:279-288; a symmetric term is the phi = 0 case because swap = T + T^T)
struct Coefs {
  double delta[kMaxSlots][kMaxAtoms];
  double c_re[kMaxSlots][kMaxAtoms];
  double c_im[kMaxSlots][kMaxAtoms];
  double csr_re[8], csr_im[8];  // amplitudes of the generic CSR terms
};

// compute entry `e` (0 <= e < kMaxSlots*n_atoms) of the coefficient tables at time t
BQA_HD void coef_entry(const TaskView& tv, double t, int e, Coefs& out) {
  int n = tv.n_atoms;
  int s = e / n, a = e - s * n;
  double d = 0.0, cr = 0.0, ci = 0.0;
  int dstride = 2 + n, rstride = 3 + n;
  for (int k = 0; k < tv.n_det; ++k) {
    const double* r = tv.det + (int64_t)k * dstride;
    if ((int)r[1] != s) continue;
    double sc = r[2 + a];
    if (sc == 0.0) continue;
    d += wf_eval(tv, (int)r[0], t) * sc;
  }
  for (int k = 0; k < tv.n_rabi; ++k) {
    const double* r = tv.rabi + (int64_t)k * rstride;
    if ((int)r[2] != s) continue;
    double sc = r[3 + a];
    if (sc == 0.0) continue;
    double amp = wf_eval(tv, (int)r[0], t) * 0.5;
    int pw = (int)r[1];
    if (pw >= 0) {
      double ph = wf_eval(tv, pw, t);
      double sn, cs;
      sincos(ph, &sn, &cs);
      cr += amp * cs * sc;   // amp * exp(-i ph)
      ci -= amp * sn * sc;
    } else {
      cr += amp * sc;
    }
  }
  out.delta[s][a] = d;
  out.c_re[s][a] = cr;
  out.c_im[s][a] = ci;
}

// coef_entry with the waveform values at time t already evaluated (wfv[w] = wf_eval(tv, w, t) for every waveform of
// the task): the same sums in the same order, so the same bits.  Lets a warp evaluate every waveform ONCE, one per
// lane, instead of once per atom.
BQA_HD void coef_entry_from_values(const TaskView& tv, const double* wfv, int e, Coefs& out) {
  int n = tv.n_atoms;
  int s = e / n, a = e - s * n;
  double d = 0.0, cr = 0.0, ci = 0.0;
  int dstride = 2 + n, rstride = 3 + n;
  for (int k = 0; k < tv.n_det; ++k) {
    const double* r = tv.det + (int64_t)k * dstride;
    if ((int)r[1] != s) continue;
    double sc = r[2 + a];
    if (sc == 0.0) continue;
    d += wfv[(int)r[0]] * sc;
  }
  for (int k = 0; k < tv.n_rabi; ++k) {
    const double* r = tv.rabi + (int64_t)k * rstride;
    if ((int)r[2] != s) continue;
    double sc = r[3 + a];
    if (sc == 0.0) continue;
    double amp = wfv[(int)r[0]] * 0.5;
    int pw = (int)r[1];
    if (pw >= 0) {
      double ph = wfv[pw];
      double sn, cs;
      sincos(ph, &sn, &cs);
      cr += amp * cs * sc;   // amp * exp(-i ph)
      ci -= amp * sn * sc;
    } else {
      cr += amp * sc;
    }
  }
  out.delta[s][a] = d;
  out.c_re[s][a] = cr;
  out.c_im[s][a] = ci;
}

BQA_HD void coef_csr_entry(const PlanView& P, const TaskView& tv, double t, int k, Coefs& out) {
  const CsrTerm& T = P.csr[k];
  double amp = wf_eval(tv, T.amp_wf, t) * 0.5;
  double cr = amp, ci = 0.0;
  if (T.phase_wf >= 0) {
    double ph = wf_eval(tv, T.phase_wf, t), sn, cs;
    sincos(ph, &sn, &cs);
    cr = amp * cs;
    ci = -amp * sn;
  }
  out.csr_re[k] = cr;
  out.csr_im[k] = ci;
}

// ------------------------------------------------------------------------------------------------
// fused RHS element.  `in` is the gather source (psi, or u*psi in the interaction picture).
//   schroedinger : out = -i [ (V + sum delta) psi_i + sum_a coef * psi_partner ]              (SV:323-334)
//   interaction  : out = -i conj(u_i) [ (sum delta) phi_i + sum_a coef * phi_partner ],  phi = u psi,
//                  u = exp(-i V t)                                                              (SV:342-359)
// Returns H-applied value WITHOUT the -i factor when `raw` (used by RydbergHamiltonian._apply, SV:375-398)
// ------------------------------------------------------------------------------------------------
// Gather source: `in[p]` for plain (global / CTA-shared) vectors.  Execution contexts whose stage
// vector is partitioned over the CTAs of a thread-block cluster provide their own accessor with the
// same operator[] (bqa_cluster.cuh: distributed shared memory).
struct PtrIn {
  const cplx* p;
  BQA_HD cplx operator[](int64_t i) const { return p[i]; }
};

template <bool kFull, int kLevel, class In>
BQA_HD cplx h_apply_element(const PlanView& P, const Coefs& C, const double* V, const In& in, int64_t i) {
  const int n = P.n_atoms;
  uint32_t cfg = kFull ? (uint32_t)i : P.configs[i];
  double diag = V ? V[i] : 0.0;
  double ar = 0.0, ai = 0.0;
  if (kLevel == 2) {
    for (int a = 0; a < n; ++a) {
      uint32_t bit = (cfg >> a) & 1u;
      int64_t p;
      if (kFull) {
        p = i ^ ((int64_t)1 << a);
      } else {
        p = P.partners[(int64_t)a * P.dim + i];
      }
      if (bit) diag -= C.delta[0][a];
      if (kFull || p >= 0) {
        cplx z = in[p];
        double cr = C.c_re[0][a];
        double ci = bit ? C.c_im[0][a] : -C.c_im[0][a];
        ar += cr * z.re - ci * z.im;
        ai += cr * z.im + ci * z.re;
      }
    }
  } else {
    uint32_t rem = cfg;
    for (int a = 0; a < n; ++a) {
      uint32_t dg = rem % 3u;
      rem /= 3u;
      if (dg >= 1) {  // upper level of slot dg-1: detuning + coefficient A towards the lower partner
        int s = dg - 1;
        diag -= C.delta[s][a];
        int64_t p = kFull ? i - (int64_t)P.pow_level[a] : (int64_t)P.partners[((int64_t)a * 2 + s) * P.dim + i];
        if (kFull || p >= 0) {
          cplx z = in[p];
          double cr = C.c_re[s][a], ci = C.c_im[s][a];
          ar += cr * z.re - ci * z.im;
          ai += cr * z.im + ci * z.re;
        }
      }
      if (dg <= 1) {  // lower level of slot dg: coefficient conj(A) towards the upper partner
        int s = dg;
        int64_t p = kFull ? i + (int64_t)P.pow_level[a] : (int64_t)P.partners[((int64_t)a * 2 + s) * P.dim + i];
        if (kFull || p >= 0) {
          cplx z = in[p];
          double cr = C.c_re[s][a], ci = -C.c_im[s][a];
          ar += cr * z.re - ci * z.im;
          ai += cr * z.im + ci * z.re;
        }
      }
    }
  }
  // generic CSR fallback terms (gather form for both M and M^T)
  for (int k = 0; k < P.n_csr; ++k) {
    const CsrTerm& T = P.csr[k];
    double sr = 0.0, si = 0.0;
    for (int64_t j = T.indptr[i]; j < T.indptr[i + 1]; ++j) {
      cplx z = in[T.indices[j]];
      sr += T.data[j] * z.re;
      si += T.data[j] * z.im;
    }
    double cr = C.csr_re[k], ci = C.csr_im[k];
    ar += cr * sr - ci * si;
    ai += cr * si + ci * sr;
    if (T.phase_wf >= 0) {
      sr = 0.0;
      si = 0.0;
      for (int64_t j = T.t_indptr[i]; j < T.t_indptr[i + 1]; ++j) {
        cplx z = in[T.t_indices[j]];
        sr += T.t_data[j] * z.re;
        si += T.t_data[j] * z.im;
      }
      ar += cr * sr + ci * si;  // conj(A)
      ai += cr * si - ci * sr;
    }
  }
  cplx zi = in[i];
  return cmake(diag * zi.re + ar, diag * zi.im + ai);
}

template <class In>
BQA_HD cplx h_apply_dispatch(const PlanView& P, const Coefs& C, const double* V, const In& in, int64_t i) {
  if (P.n_level == 2) {
    return P.full_space ? h_apply_element<true, 2>(P, C, V, in, i) : h_apply_element<false, 2>(P, C, V, in, i);
  }
  return P.full_space ? h_apply_element<true, 3>(P, C, V, in, i) : h_apply_element<false, 3>(P, C, V, in, i);
}
BQA_HD cplx h_apply_dispatch(const PlanView& P, const Coefs& C, const double* V, const cplx* in, int64_t i) {
  PtrIn g;
  g.p = in;
  return h_apply_dispatch(P, C, V, g, i);
}

// u = exp(-i * v * t)
BQA_HD cplx phase_u(double v, double t) {
  double sn, cs;
  sincos(v * t, &sn, &cs);
  return cmake(cs, -sn);
}

BQA_HD cplx cmul(cplx a, cplx b) { return cmake(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
BQA_HD cplx cmul_conj_a(cplx a, cplx b) {  // conj(a) * b
  return cmake(a.re * b.re + a.im * b.im, a.re * b.im - a.im * b.re);
}
BQA_HD cplx times_minus_i(cplx z) { return cmake(z.im, -z.re); }

// ------------------------------------------------------------------------------------------------
// Runge-Kutta tables
// ------------------------------------------------------------------------------------------------
}  // namespace bqa

// Two copies of the tables: `dev_tables` (device memory) for kernels, `host_tables` for host code
// (bqa_large's launch logic and the g++-compiled test simulator).  dop_coeffs.h has no include guard.
namespace bqa {
#if defined(__CUDACC__)
namespace dev_tables {
#undef BQA_TABLE_QUAL
#define BQA_TABLE_QUAL __device__ const
#include "dop_coeffs.h"
}  // namespace dev_tables
#endif
namespace host_tables {
#undef BQA_TABLE_QUAL
#define BQA_TABLE_QUAL static const
#include "dop_coeffs.h"
}  // namespace host_tables
}  // namespace bqa

namespace bqa {

#if defined(__CUDA_ARCH__)
namespace tables = dev_tables;
#else
namespace tables = host_tables;
#endif

struct SolverOpts {
  int solver;        // kSolverDopri5 / kSolverDop853
  int interaction;   // interaction picture
  double atol, rtol;
  double safety, fac1, fac2, beta;   // scipy: 0.9, dfactor, ifactor, 0.0
  int64_t nmax;      // nsteps
  int nstiff;        // 1000
  // DOP853 step-size after a REJECTED step.
  //   0 (default): Hairer's published rule  hnew = h / min(1/fac1, err^(1/8)/safety)  - this is
  //      SciPy's Fortran dop853.f (<= 1.16; the reference's uv.lock pins 1.15.3).
  //   1: hnew = h * fac1, which reproduces SciPy 1.18.1's C translation step for step (measured in
  //      this image: identical nfcn/nstep/naccpt/nrejct, states to 1e-11); it shrinks by the full
  //      factor fac1 on every rejection and needs ~3.5x more steps on stability-limited problems.
  //      DOPRI5 is identical in both SciPy generations.
  int reject_mode;
};

BQA_HD SolverOpts default_opts(int solver, double atol, double rtol, int64_t nsteps, int interaction) {
  SolverOpts o;
  o.solver = solver;
  o.interaction = interaction;
  o.atol = atol;
  o.rtol = rtol;
  o.safety = 0.9;
  o.beta = 0.0;
  if (solver == kSolverDopri5) {
    // scipy passes beta = 0.0 and Hairer's DOPRI5 reads WORK(5) == 0 as "use the default 0.04"
    // (DOP853 reads it as 0.0)
    o.beta = 0.04;
    o.fac1 = 0.2;
    o.fac2 = 10.0;
  } else {
    o.fac1 = 0.3;
    o.fac2 = 6.0;
  }
  o.nmax = nsteps;
  o.nstiff = 1000;
  o.reject_mode = 0;
  return o;
}

// Per-task integrator statistics returned to the host
struct TaskStats {
  int32_t status;
  int32_t pad;
  int64_t nfcn, nstep, naccpt, nrejct;
  double h_last, t_last;
};

// ------------------------------------------------------------------------------------------------
// Scalar controller of Hairer's DOPRI5/DOP853 core loop (one integrate() call).  All threads of a
// CTA run it redundantly on identical inputs; bqa_large runs it in a single-thread kernel.
// ------------------------------------------------------------------------------------------------
struct Controller {
  double x, xend, h, hmax, posneg, facold, hlamb, expo1, facc1, facc2;
  int64_t nstep, naccpt, nrejct, nfcn;
  int iasti, nonsti, reject, last, status, done;

  BQA_HD void begin(const SolverOpts& o, double x0, double x1) {
    x = x0;
    xend = x1;
    h = 0.0;
    hmax = fabs(x1 - x0);
    posneg = (x1 - x0) >= 0 ? 1.0 : -1.0;
    facold = 1e-4;
    hlamb = 0.0;
    expo1 = (o.solver == kSolverDopri5) ? (0.2 - o.beta * 0.75) : (1.0 / 8.0 - o.beta * 0.2);
    facc1 = 1.0 / o.fac1;
    facc2 = 1.0 / o.fac2;
    nstep = naccpt = nrejct = nfcn = 0;
    iasti = nonsti = reject = last = 0;
    status = 0;
    done = 0;
  }

  // first guess of hinit: h = 0.01 * ||y|| / ||f0||
  BQA_HD double hinit_first(double dnf, double dny) const {
    double hh = (dnf <= 1e-10 || dny <= 1e-10) ? 1e-6 : sqrt(dny / dnf) * 0.01;
    hh = fmin(hh, hmax);
    return copysign(hh, posneg);
  }
  // second part of hinit given der2 = sum(((f1-f0)/sk)^2)
  BQA_HD void hinit_finish(const SolverOpts& o, double h0, double dnf, double der2sum) {
    double der2 = sqrt(der2sum) / h0;
    double der12 = fmax(fabs(der2), sqrt(dnf));
    int iord = (o.solver == kSolverDopri5) ? 5 : 8;
    double h1;
    if (der12 <= 1e-15)
      h1 = fmax(1e-6, fabs(h0) * 1e-3);
    else
      h1 = pow(0.01 / der12, 1.0 / iord);
    double hh = fmin(fmin(100.0 * fabs(h0), h1), hmax);
    h = copysign(hh, posneg);
    nfcn += 2;
  }

  // top of the basic integration step; returns false when the loop must stop (status set)
  BQA_HD bool step_begin(const SolverOpts& o) {
    const double uround = 2.3e-16;
    if (nstep > o.nmax) {
      status = kStatusNeedSteps;
      done = 1;
      return false;
    }
    if (0.1 * fabs(h) <= fabs(x) * uround) {
      status = kStatusStepTooSmall;
      done = 1;
      return false;
    }
    if ((x + 1.01 * h - xend) * posneg > 0.0) {
      h = xend - x;
      last = 1;
    }
    nstep += 1;
    return true;
  }

  // normalised error of the step from the raw sums
  BQA_HD double err_dop853(double err5sum, double err3sum, int64_t n_real) const {
    double deno = err5sum + 0.01 * err3sum;
    if (deno <= 0.0) deno = 1.0;
    return fabs(h) * err5sum * sqrt(1.0 / ((double)n_real * deno));
  }
  BQA_HD double err_dopri5(double errsum, int64_t n_real) const { return sqrt(errsum / (double)n_real); }

  // decide accept/reject.  Returns 1 if accepted.  `hnew` is kept internally until step_end.
  double hnew, fac11;
  BQA_HD int decide(const SolverOpts& o, double err) {
    fac11 = pow(err, expo1);
    // beta = 0 (SciPy's setting for dop853): pow(facold, 0) is exactly 1 and x / 1 is x - skip the call
    double fac = (o.beta == 0.0) ? fac11 : fac11 / pow(facold, o.beta);
    fac = fmax(facc2, fmin(facc1, fac / o.safety));
    hnew = h / fac;
    if (err <= 1.0) {
      facold = fmax(err, 1e-4);
      naccpt += 1;
      return 1;
    }
    return 0;
  }
  // 32-bit operands: naccpt <= nmax < 2^31, and a 64-bit remainder is a ~100-instruction software routine per thread
  BQA_HD bool stiff_due(const SolverOpts& o) const { return ((uint32_t)naccpt % (uint32_t)o.nstiff == 0u) || iasti > 0; }
  // stiffness test after an accepted step (sums are over all real components)
  BQA_HD void stiff_update(const SolverOpts& o, double stnum, double stden) {
    if (stden > 0.0) hlamb = ((o.solver == kSolverDopri5) ? h : fabs(h)) * sqrt(stnum / stden);
    double lim = (o.solver == kSolverDopri5) ? 3.25 : 6.1;
    if (hlamb > lim) {
      nonsti = 0;
      iasti += 1;
      if (iasti == 15) {
        status = kStatusStiff;
        done = 1;
      }
    } else {
      nonsti += 1;
      if (nonsti == 6) iasti = 0;
    }
  }
  // bookkeeping after accept (x advanced) or reject
  BQA_HD void step_end(const SolverOpts& o, int accepted) {
    if (accepted) {
      x = x + h;
      if (last) {
        x = xend;
        h = hnew;
        if (!done) {
          status = kStatusOk;
          done = 1;
        }
        return;
      }
      if (fabs(hnew) > hmax) hnew = posneg * hmax;
      if (reject) hnew = posneg * fmin(fabs(hnew), fabs(h));
      reject = 0;
    } else {
      if (o.reject_mode == 1 && o.solver != kSolverDopri5)
        hnew = h / facc1;
      else
        hnew = h / fmin(facc1, fac11 / o.safety);
      reject = 1;
      if (naccpt >= 1) nrejct += 1;
      last = 0;
    }
    h = hnew;
  }
};


}  // namespace bqa
