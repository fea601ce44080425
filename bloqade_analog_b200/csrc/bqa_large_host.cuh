// bqa_large_host.cuh - host launch sequence of the streamed engine (included from bqa_api_evolve.cuh)
#pragma once
#include <stddef.h>

static void lg_enqueue_step(bqa_plan* p, LargeCtrl* ctl, const double* rec, const double* V, bool d5, int gs, int ga) {
  const int64_t dim = p->P.dim;
  cudaStream_t s = p->stream;
  lg_ctl_step_begin<<<1, 1, 0, s>>>(ctl);
  LAUNCHED();
  if (d5) {
    for (int st = 0; st < 5; ++st) {
      lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeStage, st);
      lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, kTimeStage, st, st + 1, 0);
      LAUNCHED();
      LAUNCHED();
    }
    lg_final5a<<<gs, kLgThreads, 0, s>>>(ctl, dim, V);
    lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, kTimeNew, 0, 1, 0);
    lg_final5b<<<gs, kLgThreads, 0, s>>>(ctl, dim);
    lg_ctl_decide<<<1, 1, 0, s>>>(ctl, dim);
    for (int k = 0; k < 4; ++k) LAUNCHED();
  } else {
    for (int st = 0; st < 11; ++st) {
      const int dst = (st == 9) ? 1 : (st == 10) ? 2 : st + 1;
      lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeStage, st);
      lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, kTimeStage, st, dst, 0);
      LAUNCHED();
      LAUNCHED();
    }
    lg_final853<<<gs, kLgThreads, 0, s>>>(ctl, dim, V);
    lg_ctl_decide<<<1, 1, 0, s>>>(ctl, dim);
    lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, kTimeNew, 0, 3, 1);
    lg_stnum853<<<gs, kLgThreads, 0, s>>>(ctl, dim);
    lg_ctl_end853<<<1, 1, 0, s>>>(ctl);
    for (int k = 0; k < 5; ++k) LAUNCHED();
  }
}

static int evolve_large_device(bqa_plan* p, const double* d_records, const std::vector<int64_t>& h_off,
                               const double* h_records, int n_tasks, const bqa_solver_opts* o, const cplx* d_psi0,
                               int psi0_mode, const double* h_times, int n_times, cplx* d_out, TaskStats* d_stats) {
  const int64_t dim = p->P.dim;
  const bool d5 = o->solver == BQA_SOLVER_DOPRI5;
  const int n_vec = d5 ? 9 : 12;
  cudaStream_t s = p->stream;
  // reduction-style kernels: a multiple of the SM count, capped by the amount of work
  const int gs = (int)std::max<int64_t>(1, std::min<int64_t>((dim + kLgThreads - 1) / kLgThreads, (int64_t)p->sm_count * 4));
  const int ga = grid_for(p, dim, kApplyThreads);
  cplx* work = nullptr;
  LargeCtrl* ctl = nullptr;
  double* partial = nullptr;
  int* h_done = nullptr;
  CK(cudaMalloc(&work, sizeof(cplx) * dim * n_vec));
  CK(cudaMalloc(&ctl, sizeof(LargeCtrl)));
  CK(cudaMalloc(&partial, sizeof(double) * 4 * gs));
  CK(cudaMallocHost(&h_done, 2 * sizeof(int)));
  int rc = 0;
  const int steps_per_poll = dim >= (1 << 22) ? 4 : (dim >= (1 << 16) ? 16 : 64);
  for (int task = 0; task < n_tasks && !rc; ++task) {
    const double* rec = d_records + h_off[task];
    const int diag_id = (int)h_records[h_off[task] + 1];
    const double* V = (diag_id >= 0 && diag_id < (int)p->diags.size()) ? p->diags[diag_id] : nullptr;
    LargeCtrl hc;
    memset(&hc, 0, sizeof(hc));
    hc.o = default_opts(o->solver, o->atol, o->rtol, o->nsteps, o->interaction_picture);
    hc.o.reject_mode = o->reject_mode;
    for (int v = 0; v < n_vec; ++v) hc.vec[v] = work + (int64_t)v * dim;
    hc.partial = partial;
    hc.n_blocks = gs;
    hc.active = 1;
    if (cudaMemcpyAsync(ctl, &hc, sizeof(hc), cudaMemcpyHostToDevice, s) != cudaSuccess) { rc = fail("ctl upload failed"); break; }
    if (cudaStreamSynchronize(s) != cudaSuccess) { rc = fail("sync failed"); break; }  // hc is a stack object
    if (psi0_mode == 0) {
      lg_init_state<<<gs, kLgThreads, 0, s>>>(hc.vec[kVecY], dim);
      LAUNCHED();
    } else {
      const cplx* src = d_psi0 + (psi0_mode == 2 ? (int64_t)task * dim : 0);
      cudaMemcpyAsync(hc.vec[kVecY], src, sizeof(cplx) * dim, cudaMemcpyDeviceToDevice, s);
    }
    double t_cur = 0.0;
    bool failed = false;
    for (int k = 0; k < n_times; ++k) {
      const double t_next = h_times[(size_t)task * n_times + k];
      int ran = 0;
      if (!failed && t_next != t_cur) {
        ran = 1;
        lg_ctl_begin<<<1, 1, 0, s>>>(ctl, t_cur, t_next);
        lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeX, -1);
        lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, kTimeX, 0, 0, 0);
        lg_hinit_reduce<<<gs, kLgThreads, 0, s>>>(ctl, dim, 0);
        lg_ctl_hinit<<<1, 1, 0, s>>>(ctl, 0);
        lg_combine<<<gs, kLgThreads, 0, s>>>(ctl, dim, V, kTimeHinit, -2);
        lg_rhs<<<ga, kApplyThreads, 0, s>>>(ctl, p->P, rec, V, kTimeHinit, 0, 1, 0);
        lg_hinit_reduce<<<gs, kLgThreads, 0, s>>>(ctl, dim, 1);
        lg_ctl_hinit<<<1, 1, 0, s>>>(ctl, 1);
        for (int q = 0; q < 9; ++q) LAUNCHED();
        for (;;) {
          for (int q = 0; q < steps_per_poll; ++q) lg_enqueue_step(p, ctl, rec, V, d5, gs, ga);
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
                                        d_stats ? d_stats + (int64_t)task * n_times + k : nullptr, ran);
      LAUNCHED();
    }
    cudaError_t e = cudaStreamSynchronize(s);
    if (!rc && e != cudaSuccess) rc = fail("streamed engine: %s", cudaGetErrorString(e));
  }
  cudaFree(work);
  cudaFree(ctl);
  cudaFree(partial);
  cudaFreeHost(h_done);
  return rc;
}
