// bqa_cluster.cuh - north_star kernel (5), B200 form: ONE THREAD-BLOCK CLUSTER PER TASK with every
// vector of the integrator resident in (distributed) shared memory.
//
// The CTA-per-task kernel of bqa_small.cuh keeps only ~3 of the 12 DOP853 vectors in shared memory
// when dim = 4096 (64 KB each); the stage combines then stream the other K vectors from L2 at the
// ~80 GB/s an SM gets from it, which is what bounded it (profiles/r01_*).  Here a cluster of
// C = 1, 2, 4 or 8 CTAs owns a task; CTA r holds rows [r*chunk, (r+1)*chunk) of ALL vectors in its
// own shared memory (12 x 32 KB at dim 4096, C = 2), so stage combines, error norms and the solution
// update are pure local shared-memory passes.  Only the bit-flip gathers of the fused H(t) psi that
// cross a chunk boundary (the top log2(C) atom bits) read a peer CTA's stage vector through DSMEM.
// The task record (waveform pieces, per-atom scales) is cached in shared memory as well, so the
// per-stage coefficient evaluation never touches global memory.
//
// Replaces the same reference code as bqa_small.cuh (task/batch.py:251-252 ->
// AnalogGate._apply, emulate/ir/state_vector.py:574-641).
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "bqa_small.cuh"

namespace bqa {

namespace cg = cooperative_groups;

constexpr int kMaxCluster = 8;
constexpr int kRecSmemDoubles = 384;  // task records up to 3 KB are cached in shared memory

// accessor for the stage vector partitioned over the cluster: row p lives in the shared memory of
// CTA r = p / chunk at the same offset as in this CTA (identical layout and swap history), reached
// with one `mapa` (map_shared_rank) when r is a peer
struct ClusterIn {
  const cplx* raw;   // this CTA's chunk of the vector (row `lo` at raw[0])
  int rank;
  int chunk_shift;   // >= 0: chunk is a power of two; < 0: divide
  uint32_t chunk;
  __device__ __forceinline__ cplx operator[](int64_t p) const {
    uint32_t r, off;
    if (chunk_shift >= 0) {
      r = (uint32_t)p >> chunk_shift;
      off = (uint32_t)p & (chunk - 1u);
    } else {
      r = (uint32_t)p / chunk;
      off = (uint32_t)p - r * chunk;
    }
    const cplx* q = raw + off;
    if ((int)r != rank) q = cg::this_cluster().map_shared_rank(const_cast<cplx*>(q), r);
    return *q;
  }
};

struct ClusterCtx {
  static constexpr bool kFastRhs = false;
  double* red;        // shared scratch [3 * 32]
  double* part;       // shared [2][4] cluster-exchange slots (double buffered)
  int parity;
  int rank, csize;
  int64_t lo, hi;     // rows owned by this CTA
  int chunk_shift;
  uint32_t chunk;

  template <class F>
  __device__ __forceinline__ void for_each(int64_t n, F f) {
    const int64_t end = hi < n ? hi : n;
    for (int64_t i = lo + threadIdx.x; i < end; i += blockDim.x) f(i);
  }
  template <class F>
  __device__ __forceinline__ void for_small(int64_t n, F f) {
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) f(i);
  }
  __device__ __forceinline__ void sync_local() { __syncthreads(); }
  __device__ __forceinline__ void sync() {
    if (csize > 1)
      cg::this_cluster().sync();
    else
      __syncthreads();
  }
  template <class F>
  __device__ __forceinline__ void single(F f) {
    if (threadIdx.x == 0) f();
    __syncthreads();
  }

  // `p` is this CTA's pre-offset pointer of a vector
  __device__ __forceinline__ ClusterIn gather(const cplx* p) const {
    ClusterIn g;
    g.raw = p + lo;
    g.rank = rank;
    g.chunk_shift = csize == 1 ? 31 : chunk_shift;  // one CTA: every row maps to rank 0 at offset p
    g.chunk = csize == 1 ? 0u : chunk;
    return g;
  }

  __device__ __forceinline__ void sum3(double& a, double& b, double& c) {
    a = CtaCtx::warp_sum(a);
    b = CtaCtx::warp_sum(b);
    c = CtaCtx::warp_sum(c);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    if (lane == 0) {
      red[wid * 3 + 0] = a;
      red[wid * 3 + 1] = b;
      red[wid * 3 + 2] = c;
    }
    __syncthreads();
    double ta = 0.0, tb = 0.0, tc = 0.0;
    for (int w = 0; w < nw; ++w) {
      ta += red[w * 3 + 0];
      tb += red[w * 3 + 1];
      tc += red[w * 3 + 2];
    }
    if (csize == 1) {
      __syncthreads();
      a = ta;
      b = tb;
      c = tc;
      return;
    }
    // cluster stage: every CTA adds the per-CTA totals in rank order => bitwise identical results in
    // all CTAs (the step controller must take the same decision everywhere)
    cg::cluster_group cl = cg::this_cluster();
    double* slot = part + parity * 4;
    if (threadIdx.x == 0) {
      slot[0] = ta;
      slot[1] = tb;
      slot[2] = tc;
    }
    cl.sync();
    ta = tb = tc = 0.0;
    for (int r = 0; r < csize; ++r) {
      const double* q = cl.map_shared_rank(slot, r);
      ta += q[0];
      tb += q[1];
      tc += q[2];
    }
    parity ^= 1;  // the other slot is used next; a cluster barrier always separates two uses of a slot
    a = ta;
    b = tb;
    c = tc;
  }
  __device__ __forceinline__ void sum2(double& a, double& b) {
    double c = 0.0;
    sum3(a, b, c);
  }
  __device__ __forceinline__ void sum1(double& a) {
    double b = 0.0, c = 0.0;
    sum3(a, b, c);
  }
};

struct ClusterArgs {
  SmallArgs S;       // same task description as the CTA-per-task kernel (gws / n_smem_vecs unused)
  int csize;         // CTAs per cluster
  uint32_t chunk;    // rows per CTA
  int chunk_shift;   // log2(chunk) or -1
  int n_alloc;       // vectors held per CTA (n_vec, +1 for the fast path's second stage-input buffer)
  int fastK;         // 0: generic TaskEvolver; 1/2/4: Fast853<K> (bqa_cluster_fast.cuh), blockDim.x*K == chunk
};

// the output-time loop of AnalogGate._apply (state_vector.py:594-602) around any integrator
template <class Integrator>
__device__ __forceinline__ void cluster_run_times(Integrator& I, TaskEvolver<ClusterCtx>& ev, const SmallArgs& A, int task,
                                                  int rank) {
  const int64_t dim = A.P.dim;
  double t_cur = A.t_start;
  bool failed = false;
  for (int k = 0; k < A.n_times; ++k) {
    const double t_next = A.times[(int64_t)task * A.n_times + k];
    TaskStats st;
    st.status = kStatusOk;
    st.pad = 0;
    st.nfcn = st.nstep = st.naccpt = st.nrejct = 0;
    st.h_last = 0.0;
    st.t_last = t_cur;
    if (!failed && t_next != t_cur) {
      I.integrate(t_cur, t_next);
      st.status = ev.c.status;
      st.nfcn = ev.c.nfcn;
      st.nstep = ev.c.nstep;
      st.naccpt = ev.c.naccpt;
      st.nrejct = ev.c.nrejct;
      st.h_last = ev.c.h;
      st.t_last = ev.c.x;
      if (ev.c.status != kStatusOk) failed = true;
      t_cur = t_next;
    } else if (failed) {
      st.status = kStatusInconsistent;
    }
    if (A.out) ev.emit_state(t_cur, A.out + ((int64_t)task * A.n_times + k) * dim);
    if (rank == 0 && threadIdx.x == 0 && A.stats) A.stats[(int64_t)task * A.n_times + k] = st;
  }
}

}  // namespace bqa
