// bqa_api_evolve.cuh - batched evolution entry points (included from bqa_api_ops.cuh)
#pragma once

#include <stdlib.h>

#include "bqa_cluster_launch.h"
#include "bqa_small_launch.h"
#include "bqa_large.cuh"
#include "bqa_large_host.cuh"

// smem priority of the 12 vectors: ys first (every gather reads it), then k1, then the K slots that
// are re-read by the most stage combines, then y.
static const int kOrder853[kNumVec] = {kVecYs, kVecK + 0, kVecK + 3, kVecK + 4, kVecY, kVecK + 5,
                                       kVecK + 6, kVecK + 7, kVecK + 8, kVecK + 9, kVecK + 1, kVecK + 2};
// fused DOP853 step of the CTA engine (CtaCtxT::step853_fused): y instead of k4 - with the just-computed K
// carried in registers, y (read by every stage) is the vector whose L2 reads are worth most
static const int kOrder853Fused[kNumVec] = {kVecYs, kVecK + 0, kVecY, kVecK + 3, kVecK + 4, kVecK + 5,
                                            kVecK + 6, kVecK + 7, kVecK + 8, kVecK + 9, kVecK + 1, kVecK + 2};
static const int kOrder5[kNumVec] = {kVecYs, kVecK + 0, kVecK + 1, kVecY, kVecK + 6, kVecK + 2,
                                     kVecK + 3, kVecK + 4, kVecK + 5, kVecK + 7, kVecK + 8, kVecK + 9};

template <int kThreads, int kMinCtas = 1>
static int launch_small(bqa_plan* p, SmallArgs& A, int n_tasks) {
  cudaFuncAttributes fa;
  CK((small_kernel_attributes<kThreads, kMinCtas>(&fa)));
  const int64_t vec_bytes = (int64_t)sizeof(cplx) * A.P.dim;
  const int64_t dyn_max = p->smem_optin - (int64_t)fa.sharedSizeBytes - 1024;
  int n_smem = (int)std::min<int64_t>(A.n_vec, dyn_max / vec_bytes);
  if (kMinCtas > 1)  // several CTAs (tasks) per SM: each one's share of the shared memory
    n_smem = (int)std::min<int64_t>(A.n_vec, (p->smem_optin / kMinCtas - (int64_t)fa.sharedSizeBytes - 1024) / vec_bytes);
  if (n_smem < 0) n_smem = 0;
  // tiny states: leave room for several CTAs per SM instead of hoarding shared memory
  size_t dyn = (size_t)n_smem * vec_bytes;
  int occ = 1;
  CK((small_kernel_occupancy<kThreads, kMinCtas>(dyn, &occ)));
  if (occ < 1) return fail("evolve kernel does not fit on an SM (threads=%d, smem=%zu)", kThreads, dyn);
  const int grid = (int)std::min<int64_t>(n_tasks, (int64_t)occ * p->sm_count);
  A.n_smem_vecs = n_smem;
  A.fused_layout = (A.fused_layout && n_smem >= 3 && kMinCtas == 1) ? 1 : 0;
  // tensor memory as storage of four K vectors: one CTA per SM (a second one could not allocate its columns),
  // 16 * dim / threads columns per thread <= 512 * 128 / threads, i.e. dim <= 4096
  A.use_tmem = (A.fused_layout && occ == 1 && A.P.dim <= 4096 && A.P.dim % kThreads == 0 &&
                getenv("BQA_NO_TMEM") == nullptr) ? 1 : 0;
  A.gws_stride = (int64_t)(A.n_vec - n_smem) * A.P.dim;
  cplx* gws = nullptr;
  if (A.gws_stride > 0) {
    CK(cudaMalloc(&gws, sizeof(cplx) * A.gws_stride * grid));
  }
  A.gws = gws;
  if (p->timing) cudaEventRecord(p->ev0, p->stream);
  cudaError_t e = small_kernel_launch<kThreads, kMinCtas>(A, grid, dyn, p->stream);
  LAUNCHED();
  if (p->timing) cudaEventRecord(p->ev1, p->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
  if (p->timing && e == cudaSuccess) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->ev0, p->ev1) == cudaSuccess) {
      p->evolve_ms += ms;
      p->evolve_launches += 1;
    }
  }
  if (gws) cudaFree(gws);
  if (e != cudaSuccess) return fail("evolve_small_kernel: %s", cudaGetErrorString(e));
  return 0;
}

// ---- cluster engine: smallest cluster whose distributed shared memory holds all vectors ---------------
static bool cluster_fit(const bqa_plan* p, int64_t dim, int n_vec, int* csize, uint32_t* chunk, int extra_per_row = 0) {
  const size_t stat = cluster_kernel_static_smem();
  if (stat == (size_t)-1) return false;
  const int64_t budget = p->smem_optin - (int64_t)stat - 512;
  for (int c = 1; c <= kMaxCluster; c *= 2) {
    int64_t ch = (dim + c - 1) / c;
    ch = (ch + 31) / 32 * 32;
    if ((int64_t)n_vec * ch * (int64_t)sizeof(cplx) + ch * extra_per_row <= budget) {
      *csize = c;
      *chunk = (uint32_t)ch;
      return true;
    }
  }
  return false;
}

static int launch_cluster(bqa_plan* p, SmallArgs& A, int n_tasks) {
  ClusterArgs CA;
  memset(&CA, 0, sizeof(CA));
  int csize = 1;
  uint32_t chunk = 0;
  // fused DOP853 path (bqa_cluster_fast.cuh): two-level full space, bit-flip terms only, Schroedinger picture
  bool fast = A.P.full_space && A.P.n_level == 2 && A.P.n_csr == 0 && !A.o.interaction && A.o.solver == kSolverDop853 &&
              A.P.dim == ((int64_t)1 << A.P.n_atoms) && getenv("BQA_NO_FAST") == nullptr;
  if (fast) {
    fast = cluster_fit(p, A.P.dim, A.n_vec + 1, &csize, &chunk, 8) && (int64_t)chunk * csize == A.P.dim &&
           chunk >= 256 && chunk <= 1024;
  }
  if (!fast && !cluster_fit(p, A.P.dim, A.n_vec, &csize, &chunk))
    return fail("cluster engine: state does not fit in 8 CTAs of shared memory");
  CA.S = A;
  CA.csize = csize;
  CA.chunk = chunk;
  CA.chunk_shift = -1;
  for (int b = 0; b < 31; ++b)
    if ((1u << b) == chunk) CA.chunk_shift = b;
  CA.n_alloc = A.n_vec + (fast ? 1 : 0);
  CA.fastK = 0;
  int threads = (int)std::min<uint32_t>(512u, std::max<uint32_t>(128u, (chunk / 2 + 31) / 32 * 32));
  if (const char* e = getenv("BQA_CLUSTER_THREADS")) {
    int v = atoi(e);
    if (v >= 32 && v <= 512 && v % 32 == 0) threads = v;
  }
  if (fast) {
    // (K elements per thread, T compute threads) instantiated in bqa_cluster_fast.cuh:
    //   chunk 1024: (2,512) or (4,256); 512: (2,256); 256: (2,128); 128: (1,128);  + one service warp
    CA.fastK = 2;
    if (chunk == 1024 && getenv("BQA_FAST_K4") != nullptr) CA.fastK = 4;
    threads = (int)chunk / CA.fastK + 32;
  }
  const size_t dyn = (size_t)CA.n_alloc * chunk * sizeof(cplx) + (fast ? (size_t)chunk * sizeof(double) : 0);
  int max_clusters = 0;
  CK(cluster_kernel_max_clusters(csize, threads, dyn, p->sm_count, &max_clusters));
  if (max_clusters < 1) return fail("cluster engine: no cluster of %d CTAs x %zu B fits", csize, dyn);
  const int n_clusters = std::min(n_tasks, max_clusters);
  if (p->timing) cudaEventRecord(p->ev0, p->stream);
  cudaError_t e = cluster_kernel_launch(CA, n_clusters, threads, dyn, p->stream);
  LAUNCHED();
  if (p->timing) cudaEventRecord(p->ev1, p->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
  if (p->timing && e == cudaSuccess) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->ev0, p->ev1) == cudaSuccess) {
      p->evolve_ms += ms;
      p->evolve_launches += 1;
    }
  }
  if (e != cudaSuccess) return fail("evolve_cluster_kernel (cluster %d, %d threads, %zu B): %s", csize, threads, dyn, cudaGetErrorString(e));
  return 0;
}

// all pointers are device pointers; h_times (host copy) is used for the longest-first ordering
static int evolve_small_device(bqa_plan* p, const double* d_records, const int64_t* d_offsets, int n_tasks,
                               const bqa_solver_opts* o, const cplx* d_psi0, int psi0_mode, const double* d_times,
                               const double* h_times, int n_times, cplx* d_out, TaskStats* d_stats, bool use_cluster) {
  SmallArgs A;
  memset(&A, 0, sizeof(A));
  A.P = p->P;
  A.diags = p->d_diag_table;
  A.n_diags = (int)p->diags.size();
  A.records = d_records;
  A.offsets = d_offsets;
  A.n_tasks = n_tasks;
  A.psi0 = d_psi0;
  A.psi0_mode = psi0_mode;
  A.times = d_times;
  A.n_times = n_times;
  A.out = d_out;
  A.stats = d_stats;
  A.o = default_opts(o->solver, o->atol, o->rtol, o->nsteps, o->interaction_picture);
  A.o.reject_mode = o->reject_mode;
  A.t_start = o->t_start;
  const bool d5 = o->solver == BQA_SOLVER_DOPRI5;
  A.n_vec = d5 ? 9 : 12;
  // the fused step needs: dop853, two-level full space, no CSR term, Schroedinger picture (the device also checks
  // that the task has no phase waveform and that dim = 4 or 8 elements per thread)
  const bool fused = !d5 && !use_cluster && p->P.full_space && p->P.n_level == 2 && p->P.n_csr == 0 &&
                     !o->interaction_picture && getenv("BQA_NO_FUSED_STEP") == nullptr;
  A.fused_layout = fused ? 1 : 0;
  memcpy(A.vec_order, d5 ? kOrder5 : (fused ? kOrder853Fused : kOrder853), sizeof(A.vec_order));
  // work queue: longest final time first
  std::vector<int32_t> order(n_tasks);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    return h_times[(size_t)a * n_times + n_times - 1] > h_times[(size_t)b * n_times + n_times - 1];
  });
  int32_t* d_order = nullptr;
  int* d_counter = nullptr;
  CK(cudaMalloc(&d_order, sizeof(int32_t) * n_tasks + sizeof(int)));
  d_counter = (int*)(d_order + n_tasks);
  CK(cudaMemcpyAsync(d_order, order.data(), sizeof(int32_t) * n_tasks, cudaMemcpyHostToDevice, p->stream));
  CK(cudaMemsetAsync(d_counter, 0, sizeof(int), p->stream));
  A.order = d_order;
  A.counter = d_counter;
  const int64_t dim = p->P.dim;
  int rc;
  // measured at dim 4096 (profiles/r02_summary.md): 512 threads x 8 elements with 128 registers per thread keep more
  // K-vector loads in flight than 1024 x 4 with 64 (75 vs 89 ms on the 148-task profile case); 256 x 16: 86 ms
  int threads = dim <= 256 ? 128 : (dim <= 1024 ? 256 : (dim <= 4096 ? 512 : 1024));
  // blockade subspaces run the gather RHS (rhs_sub2_real / h_apply_element): a latency chain per element, more warps
  // win (dim 2584: 12.7 us per RHS on 1024 threads, 18.6 on 512, 31 on 256; profiles/r02_summary.md)
  if (!p->P.full_space && dim > 1024) threads = 1024;
  // the plain step (interaction picture, dopri5, three-level atoms: no fused register-carry step) is a latency chain
  // too: at dim 4096 1024 threads win (interaction-picture config-3 subset 226 against 241 ms; profiles/r01_summary.md
  // had 127 against 143 ms for the plain Schroedinger step)
  if (!fused && dim > 2048) threads = 1024;
  if (const char* e = getenv("BQA_SMALL_THREADS")) threads = atoi(e);  // measurement hook (profiles/r01_summary.md)
  const bool pair = getenv("BQA_SMALL_PAIR") != nullptr;
  if (use_cluster)
    rc = launch_cluster(p, A, n_tasks);
  else if (pair && threads == 1024)
    rc = launch_small<512, 2>(p, A, n_tasks);
  else if (threads <= 128)
    rc = launch_small<128>(p, A, n_tasks);
  else if (threads <= 256)
    rc = launch_small<256>(p, A, n_tasks);
  else if (threads <= 512)
    rc = launch_small<512>(p, A, n_tasks);
  else
    rc = launch_small<1024>(p, A, n_tasks);
  cudaFree(d_order);
  return rc;
}

static int check_opts(const bqa_solver_opts* o) {
  if (!o) return fail("null solver options");
  if (o->solver != BQA_SOLVER_DOPRI5 && o->solver != BQA_SOLVER_DOP853)
    return fail("solver %d not supported (dop853 = 8, dopri5 = 5; lsoda has no GPU implementation)", o->solver);
  if (!(o->atol >= 0.0) || !(o->rtol >= 0.0) || (o->atol == 0.0 && o->rtol == 0.0)) return fail("bad tolerances");
  if (!(o->t_start == o->t_start) || o->t_start < 0.0) return fail("bad t_start");
  return 0;
}

// pick the engine: the CTA-per-task kernel for batches / small states, the streamed kernels for one
// large state
static int pick_engine(const bqa_plan* p, const bqa_solver_opts* o, int n_tasks) {
  if (o->engine == BQA_ENGINE_CTA_PER_TASK || o->engine == BQA_ENGINE_STREAMED || o->engine == BQA_ENGINE_CLUSTER)
    return o->engine;
  int cs = 0;
  uint32_t ch = 0;
  // measured on B200 (profiles/r01_summary.md): with all vectors on chip the cluster engine is 1.5-2.2x faster
  // than the CTA-per-task kernel while one or two CTAs hold a task (dim <= 2048 for dop853); at 4 CTAs per task
  // (dim 4096) the DSMEM gathers and cluster barriers cost more than the CTA kernel's L2 traffic (-10 %)
  if (cluster_fit(p, p->P.dim, o->solver == BQA_SOLVER_DOPRI5 ? 9 : 12, &cs, &ch) && cs <= 2) {
    // outside the cluster engine's fused dop853 path (interaction picture, dopri5, subspaces, three-level atoms) it
    // gathers through DSMEM with the generic loop, and from dim 1024 on the CTA engine's fused RHS routines win
    // (296 tasks: 10 atoms 12.8 against 16.1 ms in the interaction picture, 13.5 against 31.8 ms with dopri5; 11 atoms
    // 19.7 / 39.4 and 27.5 / 88.0 ms; 8 - 9 atoms the cluster engine stays ahead; profiles/r02_small_ip_engines.txt)
    const bool cluster_fast = p->P.full_space && p->P.n_level == 2 && p->P.n_csr == 0 && !o->interaction_picture &&
                              o->solver == BQA_SOLVER_DOP853;
    if (cluster_fast || p->P.dim < 1024) return BQA_ENGINE_CLUSTER;
    return BQA_ENGINE_CTA_PER_TASK;
  }
  if (p->P.dim <= 8192) return BQA_ENGINE_CTA_PER_TASK;
  if ((int64_t)n_tasks >= p->sm_count && p->P.dim <= 65536) return BQA_ENGINE_CTA_PER_TASK;
  return BQA_ENGINE_STREAMED;
}

static int evolve_any_device(bqa_plan* p, const double* d_records, const int64_t* d_offsets,
                             const std::vector<int64_t>& h_offsets, const double* h_records, int n_tasks,
                             const bqa_solver_opts* o, const cplx* d_psi0, int psi0_mode, const double* d_times,
                             const double* h_times, int n_times, cplx* d_out, TaskStats* d_stats) {
  const int engine = pick_engine(p, o, n_tasks);
  if (engine == BQA_ENGINE_CTA_PER_TASK || engine == BQA_ENGINE_CLUSTER)
    return evolve_small_device(p, d_records, d_offsets, n_tasks, o, d_psi0, psi0_mode, d_times, h_times, n_times,
                               d_out, d_stats, engine == BQA_ENGINE_CLUSTER);
  return evolve_large_device(p, d_records, h_offsets, h_records, n_tasks, o, d_psi0, psi0_mode, h_times, n_times,
                             d_out, d_stats);
}

extern "C" {

int bqa_evolve_batch(bqa_plan* p, const double* records, const int64_t* offsets, int32_t n_tasks,
                     const bqa_solver_opts* o, const void* psi0, int32_t psi0_shared, const double* times,
                     int32_t n_times, void* psi_out, bqa_task_stats* stats, int32_t mem) {
  BQA_ENTER_PLAN(p);
  if (!p || !records || !offsets || !times || n_tasks < 1 || n_times < 1) return fail("bqa_evolve_batch: bad argument");
  if (check_opts(o)) return 1;
  const int64_t dim = p->P.dim;
  // host copies of the small control arrays (needed for scheduling whatever `mem` is)
  std::vector<int64_t> h_off(n_tasks + 1);
  std::vector<double> h_times((size_t)n_tasks * n_times);
  if (mem == BQA_MEM_DEVICE) {
    CK(cudaMemcpy(h_off.data(), offsets, sizeof(int64_t) * (n_tasks + 1), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(h_times.data(), times, sizeof(double) * h_times.size(), cudaMemcpyDeviceToHost));
  } else {
    memcpy(h_off.data(), offsets, sizeof(int64_t) * (n_tasks + 1));
    memcpy(h_times.data(), times, sizeof(double) * h_times.size());
  }
  std::vector<double> h_rec_copy;
  const double* h_records = records;
  if (mem == BQA_MEM_DEVICE && pick_engine(p, o, n_tasks) == BQA_ENGINE_STREAMED) {
    h_rec_copy.resize(h_off[n_tasks]);
    CK(cudaMemcpy(h_rec_copy.data(), records, sizeof(double) * h_rec_copy.size(), cudaMemcpyDeviceToHost));
    h_records = h_rec_copy.data();
  }
  Staged rec, off, tm, p0, out, st;
  if (rec.in(records, sizeof(double) * h_off[n_tasks], mem, p->stream)) return 1;
  if (off.in(offsets, sizeof(int64_t) * (n_tasks + 1), mem, p->stream)) return 1;
  if (tm.in(times, sizeof(double) * h_times.size(), mem, p->stream)) return 1;
  const int psi0_mode = psi0 ? (psi0_shared ? 1 : 2) : 0;
  if (p0.in(psi0, sizeof(cplx) * dim * (psi0_mode == 2 ? n_tasks : 1), mem, p->stream)) return 1;
  if (out.out(psi_out, sizeof(cplx) * dim * (size_t)n_tasks * n_times, mem)) return 1;
  if (st.out(stats, sizeof(TaskStats) * (size_t)n_tasks * n_times, BQA_MEM_HOST)) return 1;
  if (evolve_any_device(p, (const double*)rec.dev, (const int64_t*)off.dev, h_off, h_records, n_tasks, o,
                        (const cplx*)p0.dev, psi0_mode, (const double*)tm.dev, h_times.data(), n_times,
                        (cplx*)out.dev, (TaskStats*)st.dev))
    return 1;
  if (out.flush(p->stream)) return 1;
  if (st.flush(p->stream)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

int bqa_run_batch(bqa_plan* p, const double* records, const int64_t* offsets, int32_t n_tasks,
                  const bqa_solver_opts* o, const double* times, const double* uniforms, int32_t n_shots,
                  int32_t project, uint8_t* shots, void* psi_out, bqa_task_stats* stats, int32_t mem) {
  BQA_ENTER_PLAN(p);
  if (!p || !records || !offsets || !times || !uniforms || !shots || n_tasks < 1 || n_shots < 1)
    return fail("bqa_run_batch: bad argument");
  if (check_opts(o)) return 1;
  const int64_t dim = p->P.dim;
  std::vector<int64_t> h_off(n_tasks + 1);
  std::vector<double> h_times((size_t)n_tasks);
  if (mem == BQA_MEM_DEVICE) {
    CK(cudaMemcpy(h_off.data(), offsets, sizeof(int64_t) * (n_tasks + 1), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(h_times.data(), times, sizeof(double) * n_tasks, cudaMemcpyDeviceToHost));
  } else {
    memcpy(h_off.data(), offsets, sizeof(int64_t) * (n_tasks + 1));
    memcpy(h_times.data(), times, sizeof(double) * n_tasks);
  }
  std::vector<double> h_rec_copy;
  const double* h_records = records;
  if (mem == BQA_MEM_DEVICE && pick_engine(p, o, n_tasks) == BQA_ENGINE_STREAMED) {
    h_rec_copy.resize(h_off[n_tasks]);
    CK(cudaMemcpy(h_rec_copy.data(), records, sizeof(double) * h_rec_copy.size(), cudaMemcpyDeviceToHost));
    h_records = h_rec_copy.data();
  }
  Staged rec, off, tm, un, sh, st, psi;
  if (rec.in(records, sizeof(double) * h_off[n_tasks], mem, p->stream)) return 1;
  if (off.in(offsets, sizeof(int64_t) * (n_tasks + 1), mem, p->stream)) return 1;
  if (tm.in(times, sizeof(double) * n_tasks, mem, p->stream)) return 1;
  if (un.in(uniforms, sizeof(double) * (size_t)n_tasks * n_shots, mem, p->stream)) return 1;
  if (sh.out(shots, (size_t)n_tasks * n_shots * p->P.n_atoms, mem)) return 1;
  if (st.out(stats, sizeof(TaskStats) * (size_t)n_tasks, BQA_MEM_HOST)) return 1;
  // final states: the caller's buffer, or an internal one
  if (psi_out) {
    if (psi.out(psi_out, sizeof(cplx) * dim * (size_t)n_tasks, mem)) return 1;
  } else {
    if (psi.out((void*)1, sizeof(cplx) * dim * (size_t)n_tasks, BQA_MEM_HOST)) return 1;
    psi.host = nullptr;
  }
  if (evolve_any_device(p, (const double*)rec.dev, (const int64_t*)off.dev, h_off, h_records, n_tasks, o, nullptr, 0,
                        (const double*)tm.dev, h_times.data(), 1, (cplx*)psi.dev, (TaskStats*)st.dev))
    return 1;
  // result.normalize() (SV:697) then result.sample(shots) (SV:699)
  {
    const int blocks = grid_for(p, dim, 256, 1);
    double* partial = nullptr;
    CK(cudaMalloc(&partial, sizeof(double) * ((size_t)blocks * n_tasks + n_tasks)));
    double* n2 = partial + (size_t)blocks * n_tasks;
    int rc = norm2_segments(p, (const cplx*)psi.dev, n_tasks, partial, blocks, n2);
    if (!rc) {
      for (int s0 = 0; s0 < n_tasks; s0 += kMaxSegPerLaunch) {
        const int ns = std::min(kMaxSegPerLaunch, n_tasks - s0);
        scale_by_norm_kernel<<<dim3(blocks, ns), 256, 0, p->stream>>>((cplx*)psi.dev + (size_t)s0 * dim, dim, dim, n2 + s0);
        LAUNCHED();
      }
      rc = sample_device(p, (const cplx*)psi.dev, n_tasks, (const double*)un.dev, n_shots, project, (uint8_t*)sh.dev);
    }
    cudaError_t e = cudaStreamSynchronize(p->stream);
    cudaFree(partial);
    if (rc) return 1;
    if (e != cudaSuccess) return fail("bqa_run_batch: %s", cudaGetErrorString(e));
  }
  if (sh.flush(p->stream)) return 1;
  if (psi.flush(p->stream)) return 1;
  if (st.flush(p->stream)) return 1;
  CK(cudaStreamSynchronize(p->stream));
  return 0;
}

}  // extern "C"
