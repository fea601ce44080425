// bqa_large_host.cuh - host launch sequence of the streamed engine (included from bqa_api_evolve.cuh)
#pragma once
#include <stddef.h>

// how the RHS of the streamed engine is launched: gather kernel or tiled passes; sharded or not
struct LgRhsPlan {
  bool tiled = false;
  std::vector<TileSpec> passes;
  int world = 1;
  int tile_grid = 1;
  bool real = false;  // no phase waveform in the current task record: real-coefficient kernels
  bool interaction = false;  // interaction picture: tiled passes without the static diagonal + lg_rotate_conj
  int rot_grid = 1;
};

typedef void (*LgTiledKernel)(LargeCtrl*, const PlanView, const double*, const double*, int, int, int, int, const TileSpec);

static LgTiledKernel lg_tiled_kernel_at(int variant) {
  static const LgTiledKernel tab[24] = {BQA_TILED_VARIANTS(lg_rhs_tiled, false), BQA_TILED_VARIANTS(lg_rhs_tiled, true),
                                        BQA_TILED_REMOTE(lg_rhs_tiled, false), BQA_TILED_REMOTE(lg_rhs_tiled, true)};
  return tab[variant];
}

static void lg_sync_ranks(bqa_plan* p, LargeCtrl* ctl, const LgRhsPlan& R) {
  if (R.world <= 1) return;
  lg_barrier<<<1, 32, 0, p->stream>>>(ctl);
  LAUNCHED();
}

// K[dst] = f(t, ys); sharded: a barrier before (ys complete on every rank) and after (peers done reading it)
static void lg_launch_rhs(bqa_plan* p, LargeCtrl* ctl, const double* rec, const double* V, const LgRhsPlan& R, int ga,
                          int tmode, int st, int dst, int cond) {
  cudaStream_t s = p->stream;
  lg_sync_ranks(p, ctl, R);
  if (R.tiled) {
    const bool rot = R.interaction && V != nullptr;
    for (const TileSpec& S : R.passes) {
      lg_tiled_kernel_at(tiled_variant(R.real, S))<<<R.tile_grid, kTileThreads, kTileDynSmem, s>>>(ctl, p->P, rec, rot ? nullptr : V, tmode, st, dst, cond, S);
      LAUNCHED();
    }
    if (rot) {
      lg_rotate_conj<<<R.rot_grid, kLgThreads, 0, s>>>(ctl, p->P.dim, V, tmode, st, dst, cond);
      LAUNCHED();
    }
  } else {
    lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, tmode, st, dst, cond);
    LAUNCHED();
  }
  lg_sync_ranks(p, ctl, R);
}

static void lg_enqueue_step(bqa_plan* p, LargeCtrl* ctl, const double* rec, const double* V, bool d5, int gs, int ga,
                            const LgRhsPlan& R, int64_t dim_total) {
  const int64_t dim = p->P.dim;
  cudaStream_t s = p->stream;
  lg_ctl_step_begin<<<1, 1, 0, s>>>(ctl);
  LAUNCHED();
  if (d5) {
    for (int st = 0; st < 5; ++st) {
      lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeStage, st);
      LAUNCHED();
      lg_launch_rhs(p, ctl, rec, V, R, ga, kTimeStage, st, st + 1, 0);
    }
    lg_final5a<<<gs, kLgThreads, 0, s>>>(ctl, dim, V);
    LAUNCHED();
    lg_launch_rhs(p, ctl, rec, V, R, ga, kTimeNew, 0, 1, 0);
    lg_final5b<<<gs, kLgThreads, 0, s>>>(ctl, dim);
    LAUNCHED();
    lg_sync_ranks(p, ctl, R);  // partial sums of every rank are complete
    lg_ctl_decide<<<1, 1, 0, s>>>(ctl, dim_total);
    LAUNCHED();
  } else {
    for (int st = 0; st < 11; ++st) {
      const int dst = (st == 9) ? 1 : (st == 10) ? 2 : st + 1;
      lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeStage, st);
      LAUNCHED();
      lg_launch_rhs(p, ctl, rec, V, R, ga, kTimeStage, st, dst, 0);
    }
    lg_final853<<<gs, kLgThreads, 0, s>>>(ctl, dim, V);
    LAUNCHED();
    lg_sync_ranks(p, ctl, R);
    lg_ctl_decide<<<1, 1, 0, s>>>(ctl, dim_total);
    LAUNCHED();
    lg_launch_rhs(p, ctl, rec, V, R, ga, kTimeNew, 0, 3, 1);
    lg_stnum853<<<gs, kLgThreads, 0, s>>>(ctl, dim);
    LAUNCHED();
    lg_sync_ranks(p, ctl, R);
    lg_ctl_end853<<<1, 1, 0, s>>>(ctl);
    LAUNCHED();
  }
}

// shard context of one evolve call (null = the state is not sharded); pointers are device pointers
struct LgShard {
  int world, rank, n_total;
  cplx* ys;                       // this rank's stage-input buffer (peer-visible)
  const cplx* partners[3];        // ys of rank ^ (1 << b)
  double* partial;                // this rank's partial-sum array (peer-visible)
  double* peer_partial[8];
  unsigned long long* peer_flag[8];
};

// scratch of one streamed-engine call, released on every exit path
struct LgScratch {
  cplx* work = nullptr;
  LargeCtrl* ctl = nullptr;
  double* partial = nullptr;
  bool own_partial = false;
  int* h_done = nullptr;
  ~LgScratch() {
    cudaFree(work);
    cudaFree(ctl);
    if (own_partial) cudaFree(partial);
    if (h_done) cudaFreeHost(h_done);
  }
};

// bound of one cross-rank barrier spin (BQA_BARRIER_TIMEOUT_MS, default 60 s)
static unsigned long long lg_barrier_timeout_ns() {
  const char* e = getenv("BQA_BARRIER_TIMEOUT_MS");
  const double ms = e ? atof(e) : 60000.0;
  return (unsigned long long)(std::max(1.0, ms) * 1e6);
}

static int evolve_large_run(bqa_plan* p, LgScratch& X, const double* d_records, const std::vector<int64_t>& h_off,
                            const double* h_records, int n_tasks, const bqa_solver_opts* o, const cplx* d_psi0,
                            int psi0_mode, const double* h_times, int n_times, cplx* d_out, TaskStats* d_stats,
                            const LgShard* sh) {
  const int64_t dim = p->P.dim;
  const int world = sh ? sh->world : 1;
  const int64_t dim_total = dim * world;
  const bool d5 = o->solver == BQA_SOLVER_DOPRI5;
  const int n_vec = d5 ? 9 : 12;
  cudaStream_t s = p->stream;
  // reduction-style kernels: a multiple of the SM count, capped by the amount of work
  const int gs = (int)std::max<int64_t>(1, std::min<int64_t>((dim + kLgThreads - 1) / kLgThreads, (int64_t)p->sm_count * 4));
  const int ga = grid_for(p, dim, kApplyThreads);
  // RHS form: shared-memory-tiled passes when the operator allows it (always for a sharded state)
  LgRhsPlan R;
  R.world = world;
  R.tile_grid = (int)std::min<int64_t>(std::max<int64_t>(dim >> kTileBits, 1), p->sm_count);
  R.interaction = o->interaction_picture != 0;
  R.rot_grid = gs;
  // (a sharded state stays in the Schroedinger picture: bqa_evolve_sharded refuses the interaction picture)
  if ((!o->interaction_picture || !sh) && p->P.full_space && p->P.n_level == 2 && p->P.n_csr == 0 &&
      (sh || !getenv("BQA_NO_TILED")))
    R.tiled = tiled_plan(p->P.n_atoms, sh ? sh->n_total : p->P.n_atoms, sh ? sh->rank : 0, sh ? sh->partners : nullptr,
                         0, R.passes);
  if (sh && !R.tiled)
    return fail("sharded evolve needs a two-level full space with >= 14 local atoms, no CSR term, Schroedinger picture");
  if (R.tiled) {
    static std::atomic<uint64_t> attr_set{0};
    if (!once_per_device(attr_set)) {
      for (int v = 0; v < 24; ++v)
        CK(cudaFuncSetAttribute(lg_tiled_kernel_at(v), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTileDynSmem));
      once_per_device(attr_set, true);
    }
  }
  CK(cudaMalloc(&X.work, sizeof(cplx) * dim * n_vec));
  CK(cudaMalloc(&X.ctl, sizeof(LargeCtrl)));
  if (sh) {
    X.partial = sh->partial;
  } else {
    CK(cudaMalloc(&X.partial, sizeof(double) * 4 * gs));
    X.own_partial = true;
  }
  CK(cudaMallocHost(&X.h_done, 2 * sizeof(int)));
  cplx* work = X.work;
  LargeCtrl* ctl = X.ctl;
  int* h_done = X.h_done;
  int rc = 0;
  const int steps_per_poll = dim >= (1 << 22) ? 4 : (dim >= (1 << 16) ? 16 : 64);
  for (int task = 0; task < n_tasks && !rc; ++task) {
    const double* rec = d_records + h_off[task];
    const int diag_id = (int)h_records[h_off[task] + 1];
    R.real = !record_has_phase(h_records + h_off[task]);
    const double* V = (diag_id >= 0 && diag_id < (int)p->diags.size()) ? p->diags[diag_id] : nullptr;
    LargeCtrl hc;
    memset(&hc, 0, sizeof(hc));
    hc.o = default_opts(o->solver, o->atol, o->rtol, o->nsteps, o->interaction_picture);
    hc.o.reject_mode = o->reject_mode;
    for (int v = 0; v < n_vec; ++v) hc.vec[v] = work + (int64_t)v * dim;
    if (sh) {
      hc.vec[kVecYs] = sh->ys;
      hc.world = sh->world;
      hc.rank = sh->rank;
      hc.barrier_timeout_ns = lg_barrier_timeout_ns();
      for (int r = 0; r < sh->world; ++r) {
        hc.peer_partial[r] = sh->peer_partial[r];
        hc.peer_flag[r] = sh->peer_flag[r];
      }
    }
    hc.partial = X.partial;
    hc.n_blocks = gs;
    hc.active = 1;
    if (cudaMemcpyAsync(ctl, &hc, sizeof(hc), cudaMemcpyHostToDevice, s) != cudaSuccess) { rc = fail("ctl upload failed"); break; }
    if (cudaStreamSynchronize(s) != cudaSuccess) { rc = fail("sync failed"); break; }  // hc is a stack object
    if (psi0_mode == 0) {
      lg_init_state<<<gs, kLgThreads, 0, s>>>(hc.vec[kVecY], dim, sh ? sh->rank : 0);
      LAUNCHED();
    } else {
      const cplx* src = d_psi0 + (psi0_mode == 2 ? (int64_t)task * dim : 0);
      cudaMemcpyAsync(hc.vec[kVecY], src, sizeof(cplx) * dim, cudaMemcpyDeviceToDevice, s);
    }
    double t_cur = o->t_start;
    bool failed = false;
    for (int k = 0; k < n_times; ++k) {
      const double t_next = h_times[(size_t)task * n_times + k];
      int ran = 0;
      if (!failed && t_next != t_cur) {
        ran = 1;
        lg_ctl_begin<<<1, 1, 0, s>>>(ctl, t_cur, t_next);
        lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeX, -1);
        lg_launch_rhs(p, ctl, rec, V, R, ga, kTimeX, 0, 0, 0);
        lg_hinit_reduce<<<gs, kLgThreads, 0, s>>>(ctl, dim, 0);
        lg_sync_ranks(p, ctl, R);
        lg_ctl_hinit<<<1, 1, 0, s>>>(ctl, 0);
        lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeHinit, -2);
        lg_launch_rhs(p, ctl, rec, V, R, ga, kTimeHinit, 0, 1, 0);
        lg_hinit_reduce<<<gs, kLgThreads, 0, s>>>(ctl, dim, 1);
        lg_sync_ranks(p, ctl, R);
        lg_ctl_hinit<<<1, 1, 0, s>>>(ctl, 1);
        for (int q = 0; q < 7; ++q) LAUNCHED();
        for (;;) {
          for (int q = 0; q < steps_per_poll; ++q) lg_enqueue_step(p, ctl, rec, V, d5, gs, ga, R, dim_total);
          cudaMemcpyAsync(h_done, (const char*)ctl + offsetof(LargeCtrl, c) + offsetof(Controller, status),
                          2 * sizeof(int), cudaMemcpyDeviceToHost, s);
          cudaError_t e = cudaStreamSynchronize(s);
          if (e != cudaSuccess) { rc = fail("streamed engine: %s", cudaGetErrorString(e)); break; }
          if (h_done[1]) break;  // Controller::done follows Controller::status
        }
        if (rc) break;
        if (h_done[0] != kStatusOk) failed = true;
        t_cur = t_next;
      }
      lg_emit<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, t_cur, d_out ? d_out + ((int64_t)task * n_times + k) * dim : nullptr,
                                        d_stats ? d_stats + (int64_t)task * n_times + k : nullptr, ran, failed ? 1 : 0);
      LAUNCHED();
    }
    cudaError_t e = cudaStreamSynchronize(s);
    if (!rc && e != cudaSuccess) rc = fail("streamed engine: %s", cudaGetErrorString(e));
  }
  return rc;
}

static int evolve_large_device(bqa_plan* p, const double* d_records, const std::vector<int64_t>& h_off,
                               const double* h_records, int n_tasks, const bqa_solver_opts* o, const cplx* d_psi0,
                               int psi0_mode, const double* h_times, int n_times, cplx* d_out, TaskStats* d_stats,
                               const LgShard* sh = nullptr) {
  LgScratch X;
  const int rc = evolve_large_run(p, X, d_records, h_off, h_records, n_tasks, o, d_psi0, psi0_mode, h_times, n_times, d_out,
                                  d_stats, sh);
  if (rc && sh) {
    // this rank failed on the host (allocation, launch, ...): release the peers from their barriers.  The error
    // text of the failure itself stays in bqa_last_error.
    LgPeerFlags F;
    for (int r = 0; r < 8; ++r) F.f[r] = r < sh->world ? sh->peer_flag[r] : nullptr;
    lg_abort_peers<<<1, 32, 0, p->stream>>>(F, sh->world);
    cudaStreamSynchronize(p->stream);
  }
  return rc;
}
