// bqa_large.cuh - streamed engine for ONE large state (north_star kernels (1)+(2)).
//
// The state is too big for a CTA, so every phase of Hairer's DOPRI5/DOP853 step is a grid-wide
// kernel on one stream: stage combine, fused H(t)*psi, error-norm partial sums, and a one-thread
// controller kernel that owns the scalar state (`Controller`, bqa_core.cuh) in DEVICE memory.  The
// host only enqueues a fixed launch sequence per step and polls a `done` flag every few steps, so
// the whole adaptive step - accept/reject included - is resident on the GPU
// (reference flow: AnalogGate._apply, state_vector.py:574-602, driving scipy's dop853/dopri5).
#pragma once
#include "bqa_apply_tiled.cuh"
#include "bqa_kernels.cuh"

namespace bqa {

struct LargeCtrl {
  Controller c;
  SolverOpts o;
  cplx* vec[kNumVec];
  double* partial;   // [3 * n_blocks] per-CTA partial sums
  int n_blocks;
  int active;        // this step is being attempted
  int accepted;      // the attempted step was accepted
  int need_stnum;
  double h0, dnf;    // hinit scratch
  double stden;
  // sharded state (bqa_api_shard.cuh): `world` ranks run this same controller on identical sums, so they
  // take identical decisions.  peer_partial[r] / peer_flag[r] are rank r's partial-sum array and barrier
  // counter, mapped into this process (CUDA IPC, NVLink peer access); world <= 1: not sharded.
  int world, rank;
  int aborted;                           // a barrier saw the abort word or timed out (lg_mark_abort)
  unsigned long long barrier_timeout_ns; // bound of one barrier's spin
  double* peer_partial[8];
  unsigned long long* peer_flag[8];
};

enum : int { kTimeX = 0, kTimeHinit = 1, kTimeStage = 2, kTimeNew = 3 };

__device__ __forceinline__ double lg_time(const LargeCtrl* ctl, int tmode, int s) {
  const Controller& c = ctl->c;
  if (tmode == kTimeX) return c.x;
  if (tmode == kTimeHinit) return c.x + ctl->h0;
  if (tmode == kTimeNew) return c.x + c.h;
  const bool d5 = ctl->o.solver == kSolverDopri5;
  if (d5) return c.x + dev_tables::DOPRI5_c[s + 1] * c.h;
  return (s == 10) ? c.x + c.h : c.x + dev_tables::DOP853_c[s + 1] * c.h;
}

constexpr int kLgThreads = 256;

__global__ void lg_ctl_begin(LargeCtrl* ctl, double x0, double x1) {
  ctl->c.begin(ctl->o, x0, x1);
  ctl->active = 1;
  ctl->accepted = 0;
  ctl->need_stnum = 0;
  ctl->stden = 0.0;
}

// ys = (u(t)*) [ y + h * sum a_j K_j ]     row: -1 = plain load, -2 = hinit (y + h0*K0), >= 0 = stage row
__global__ void __launch_bounds__(kLgThreads)
lg_combine(LargeCtrl* ctl, int64_t dim, const double* __restrict__ V, int tmode, int row) {
  if (!ctl->active) return;
  const bool d5 = ctl->o.solver == kSolverDopri5;
  int nnz = 0;
  const int* slot = nullptr;
  const double* val = nullptr;
  double h = ctl->c.h;
  __shared__ int s_slot[BQA_RK_MAXNZ];
  __shared__ double s_val[BQA_RK_MAXNZ];
  if (row == -2) {
    nnz = 1;
    h = ctl->h0;
    if (threadIdx.x == 0) {
      s_slot[0] = 0;
      s_val[0] = 1.0;
    }
  } else if (row >= 0) {
    nnz = d5 ? dev_tables::DOPRI5_a_nnz[row] : dev_tables::DOP853_a_nnz[row];
    if (threadIdx.x < nnz) {
      s_slot[threadIdx.x] = d5 ? dev_tables::DOPRI5_a_slot[row][threadIdx.x] : dev_tables::DOP853_a_slot[row][threadIdx.x];
      s_val[threadIdx.x] = d5 ? dev_tables::DOPRI5_a_val[row][threadIdx.x] : dev_tables::DOP853_a_val[row][threadIdx.x];
    }
  }
  __syncthreads();
  slot = s_slot;
  val = s_val;
  const double t = lg_time(ctl, tmode, row);
  const bool rot = ctl->o.interaction != 0 && V != nullptr;
  const cplx* y = ctl->vec[kVecY];
  cplx* ys = ctl->vec[kVecYs];
  const cplx* kk[BQA_RK_MAXNZ];
  for (int j = 0; j < nnz; ++j) kk[j] = ctl->vec[kVecK + slot[j]];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    double sr = 0.0, si = 0.0;
    for (int j = 0; j < nnz; ++j) {
      cplx kz = kk[j][i];
      sr += val[j] * kz.re;
      si += val[j] * kz.im;
    }
    cplx yo = y[i];
    cplx z = cmake(yo.re + h * sr, yo.im + h * si);
    ys[i] = rot ? cmul(phase_u(V[i], t), z) : z;
  }
}

// K[dst] = f(t, ys).  cond 0: when the step is active; cond 1: only when it was accepted
__global__ void __launch_bounds__(kApplyThreads)
lg_rhs(LargeCtrl* ctl, const PlanView P, const double* __restrict__ rec, const double* __restrict__ V, int tmode,
       int s, int dst, int cond) {
  if (!ctl->active || (cond == 1 && !ctl->accepted)) return;
  __shared__ Coefs cf;
  const double t = lg_time(ctl, tmode, s);
  const TaskView tv = task_view(rec);
  const int ne = tv.n_atoms * (P.n_level - 1);
  for (int e = threadIdx.x; e < ne + P.n_csr; e += blockDim.x) {
    if (e < ne)
      coef_entry(tv, t, e, cf);
    else
      coef_csr_entry(P, tv, t, e - ne, cf);
  }
  __syncthreads();
  const bool inter = ctl->o.interaction != 0;
  const cplx* in = ctl->vec[kVecYs];
  cplx* out = ctl->vec[kVecK + dst];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.dim; i += (int64_t)gridDim.x * blockDim.x) {
    cplx hz = h_apply_dispatch(P, cf, inter ? nullptr : V, in, i);
    if (inter && V) hz = cmul_conj_a(phase_u(V[i], t), hz);
    out[i] = times_minus_i(hz);
  }
}

// Interaction picture with the tiled RHS: the tiled passes apply the operator WITHOUT the static diagonal to
// phi = u ys (lg_combine has rotated the stage input) and write -i (H - V) phi; this pass finishes
// K[dst] = conj(u) * that, u = exp(-i V t)  (SV:342-359).  Multiplying by -i is exact, so conj(u) (-i z) has the bits of
// -i (conj(u) z), the order of lg_rhs.
__global__ void __launch_bounds__(kLgThreads)
lg_rotate_conj(LargeCtrl* ctl, int64_t dim, const double* __restrict__ V, int tmode, int s, int dst, int cond) {
  if (!ctl->active || (cond == 1 && !ctl->accepted)) return;
  const double t = lg_time(ctl, tmode, s);
  cplx* k = ctl->vec[kVecK + dst];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x)
    k[i] = cmul_conj_a(phase_u(V[i], t), k[i]);
}

__device__ __forceinline__ void lg_store3(LargeCtrl* ctl, double a, double b, double c, double* red) {
  a = block_sum(a, red);
  b = block_sum(b, red);
  c = block_sum(c, red);
  if (threadIdx.x == 0) {
    ctl->partial[blockIdx.x * 3 + 0] = a;
    ctl->partial[blockIdx.x * 3 + 1] = b;
    ctl->partial[blockIdx.x * 3 + 2] = c;
  }
}

// rank r's per-CTA partial sums (this rank's own array when not sharded); every rank adds them in the same order
__device__ __forceinline__ const double* lg_partial_of(const LargeCtrl* ctl, int r) {
  return ctl->world > 1 ? ctl->peer_partial[r] : ctl->partial;
}
__device__ __forceinline__ void lg_total3(const LargeCtrl* ctl, double& a, double& b, double& c) {
  a = b = c = 0.0;
  const int world = ctl->world > 1 ? ctl->world : 1;
  for (int r = 0; r < world; ++r) {
    const volatile double* part = lg_partial_of(ctl, r);
    for (int k = 0; k < ctl->n_blocks; ++k) {
      a += part[k * 3 + 0];
      b += part[k * 3 + 1];
      c += part[k * 3 + 2];
    }
  }
}

// Sync block of a rank (bqa_api_shard.cuh), first 64 bytes: u64 [0] barrier counter (every rank adds to it),
// [1] barrier count of this rank (only its owner touches it), [2] abort word (any rank may set it).
enum : int { kFlagCounter = 0, kFlagEpoch = 1, kFlagAbort = 2 };

__device__ __forceinline__ unsigned long long lg_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// a peer failed or a barrier timed out: this rank's evolution ends with kStatusPeerAbort, every kernel of the
// launch list that follows returns at once (active = 0) and the host's next poll sees `done`
__device__ __forceinline__ void lg_mark_abort(LargeCtrl* ctl) {
  ctl->aborted = 1;
  ctl->active = 0;
  ctl->accepted = 0;
  if (ctl->c.status >= 0) ctl->c.status = kStatusPeerAbort;
  ctl->c.done = 1;
}

// Cross-rank barrier of a sharded run, device initiated: every rank adds 1 to every rank's counter over
// NVLink, then spins on its own counter.  Writes of the kernels before it (stage input ys, partial sums) are
// published by the system-scope fence; launched <<<1, 32>>> on the engine's stream.  The spin is BOUNDED
// (LargeCtrl::barrier_timeout_ns) and polls the abort word of this rank's sync block: a rank that fails on the
// host (lg_abort_peers) or times out sets the abort word of EVERY rank, so one failed rank ends the others with
// kStatusPeerAbort instead of hanging them.  Counters are reset by bqa_shard_reset before each evolve call.
__global__ void lg_barrier(LargeCtrl* ctl) {
  const int world = ctl->world;
  if (world <= 1) return;
  __shared__ unsigned long long target;
  __shared__ int dead;
  unsigned long long* own = ctl->peer_flag[ctl->rank];
  if (threadIdx.x == 0) {
    dead = (ctl->aborted || *(volatile unsigned long long*)(own + kFlagAbort) != 0ull) ? 1 : 0;
    if (!dead) {
      unsigned long long* epoch = own + kFlagEpoch;
      *epoch += 1;
      target = *epoch * (unsigned long long)world;
    }
  }
  __syncthreads();
  if (dead) {
    if (threadIdx.x == 0) lg_mark_abort(ctl);
    return;
  }
  __threadfence_system();
  if ((int)threadIdx.x < world) atomicAdd_system(ctl->peer_flag[threadIdx.x] + kFlagCounter, 1ull);
  if (threadIdx.x == 0) {
    const volatile unsigned long long* mine = own + kFlagCounter;
    const volatile unsigned long long* abort_word = own + kFlagAbort;
    const unsigned long long t0 = lg_globaltimer();
    unsigned int polls = 0;
    while (*mine < target) {
      __nanosleep(200);
      if ((++polls & 63u) == 0u && (*abort_word != 0ull || lg_globaltimer() - t0 > ctl->barrier_timeout_ns)) {
        dead = 1;
        break;
      }
    }
  }
  __syncthreads();
  if (dead) {
    if ((int)threadIdx.x < world) atomicExch_system(ctl->peer_flag[threadIdx.x] + kFlagAbort, 1ull);
    if (threadIdx.x == 0) lg_mark_abort(ctl);
  }
  __threadfence_system();
}

// host-side failure of a rank inside a sharded evolve: tell every rank (this one included)
struct LgPeerFlags {
  unsigned long long* f[8];
};
__global__ void lg_abort_peers(const LgPeerFlags F, int world) {
  if ((int)threadIdx.x < world && F.f[threadIdx.x]) atomicExch_system(F.f[threadIdx.x] + kFlagAbort, 1ull);
  __threadfence_system();
}

// K[dst] = f(t, ys) by one shared-memory-tiled pass (two-level full space, Schroedinger picture); the
// first pass writes K[dst], later passes accumulate into it (bqa_apply_tiled.cuh)
template <bool kReal, bool kFirst, bool kHasX, int kP0, bool kRemX = false>
__global__ void __launch_bounds__(kTileThreads, 1)
lg_rhs_tiled(LargeCtrl* ctl, const PlanView P, const double* __restrict__ rec, const double* __restrict__ V, int tmode,
             int s, int dst, int cond, const TileSpec S) {
  if (!ctl->active || (cond == 1 && !ctl->accepted)) return;
  apply_tiled_cta<kReal, kFirst, kHasX, kP0, kRemX>(P, rec, V, lg_time(ctl, tmode, s), S, ctl->vec[kVecYs], ctl->vec[kVecK + dst]);
}

// hinit reductions: phase 0 -> (dnf, dny); phase 1 -> der2 with f1 = K[1]
__global__ void __launch_bounds__(kLgThreads) lg_hinit_reduce(LargeCtrl* ctl, int64_t dim, int phase) {
  __shared__ double red[32];
  const cplx* y = ctl->vec[kVecY];
  const cplx* f0 = ctl->vec[kVecK + 0];
  const cplx* f1 = ctl->vec[kVecK + 1];
  const double atol = ctl->o.atol, rtol = ctl->o.rtol;
  double a = 0.0, b = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += (int64_t)gridDim.x * blockDim.x) {
    const cplx yi = y[i], f = f0[i];
    const double skr = atol + rtol * fabs(yi.re), ski = atol + rtol * fabs(yi.im);
    if (phase == 0) {
      double q = f.re / skr; a += q * q;
      q = f.im / ski; a += q * q;
      q = yi.re / skr; b += q * q;
      q = yi.im / ski; b += q * q;
    } else {
      const cplx g = f1[i];
      double q = (g.re - f.re) / skr; a += q * q;
      q = (g.im - f.im) / ski; a += q * q;
    }
  }
  lg_store3(ctl, a, b, 0.0, red);
}

__global__ void lg_ctl_hinit(LargeCtrl* ctl, int phase) {
  double a, b, c;
  lg_total3(ctl, a, b, c);
  if (phase == 0) {
    ctl->dnf = a;
    ctl->h0 = ctl->c.hinit_first(a, b);
  } else {
    ctl->c.hinit_finish(ctl->o, ctl->h0, ctl->dnf, a);
  }
}

__global__ void lg_ctl_step_begin(LargeCtrl* ctl) {
  ctl->accepted = 0;
  ctl->need_stnum = 0;
  if (ctl->c.done) {
    ctl->active = 0;
    return;
  }
  ctl->active = ctl->c.step_begin(ctl->o) ? 1 : 0;
}

}  // namespace bqa

#include "bqa_large_step.cuh"
