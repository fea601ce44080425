// bqa_api.cu - C ABI (include/bqa_b200.h) over the CUDA kernels.  Built for sm_100a only:
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC ...
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/bqa_b200.h"
#include "bqa_kernels.cuh"
#include "bqa_apply_tiled.cuh"
#include "bqa_small.cuh"

using namespace bqa;

static_assert(sizeof(bqa_task_stats) == sizeof(TaskStats), "stats layout");

// ---------------------------------------------------------------------------------------------------
// errors, launch counter
// ---------------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

static int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) return fail("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)

#define LAUNCHED() g_launches.fetch_add(1, std::memory_order_relaxed)

// Entry of a C-ABI function: drop (and, with BQA_DEBUG set, report) a CUDA error left behind by an earlier
// unchecked call, so that the cudaGetLastError() after this function's own launches speaks about them only.
static void api_enter(const char* name) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess && getenv("BQA_DEBUG"))
    fprintf(stderr, "[bqa] stale CUDA error at entry of %s: %s\n", name, cudaGetErrorString(e));
}
#define BQA_ENTER() api_enter(__func__)

struct bqa_plan;
static void api_enter_plan(const char* name, const bqa_plan* p);
// entry points that take a plan make the plan's device current in the calling thread (one host thread per
// GPU may drive plans of different devices concurrently)
#define BQA_ENTER_PLAN(p) api_enter_plan(__func__, (p))

// per-device "done once" flags (kernel attributes are per device)
static bool once_per_device(std::atomic<uint64_t>& mask, bool set = false) {
  int dev = 0;
  cudaGetDevice(&dev);
  const uint64_t bit = 1ull << (dev & 63);
  if (set) mask.fetch_or(bit);
  return (mask.load() & bit) != 0;
}

// ---------------------------------------------------------------------------------------------------
// plan
// ---------------------------------------------------------------------------------------------------
struct bqa_plan {
  PlanView P;
  cudaStream_t stream = 0;
  int device = 0;
  int sm_count = 148;
  int64_t smem_optin = 0;
  uint32_t* d_configs = nullptr;
  int32_t* d_partners = nullptr;
  std::vector<double*> diags;          // device diagonals
  double** d_diag_table = nullptr;     // device array of the pointers above
  std::vector<CsrTerm> csr;            // device pointers inside
  CsrTerm* d_csr = nullptr;
  std::vector<void*> owned;            // misc device allocations to free
  // scratch reused across calls
  void* scratch = nullptr;
  size_t scratch_bytes = 0;
  // instrumentation (bqa_plan_enable_timing)
  bool timing = false;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  double evolve_ms = 0.0;
  int64_t evolve_launches = 0;
};

static void api_enter_plan(const char* name, const bqa_plan* p) {
  if (p) {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || cur != p->device) cudaSetDevice(p->device);
  }
  api_enter(name);
}

static int plan_scratch(bqa_plan* p, size_t bytes, void** out) {
  if (bytes > p->scratch_bytes) {
    if (p->scratch) cudaFree(p->scratch);
    p->scratch = nullptr;
    p->scratch_bytes = 0;
    CK(cudaMalloc(&p->scratch, bytes));
    p->scratch_bytes = bytes;
  }
  *out = p->scratch;
  return 0;
}

static int refresh_diag_table(bqa_plan* p) {
  if (p->d_diag_table) cudaFree(p->d_diag_table);
  p->d_diag_table = nullptr;
  if (p->diags.empty()) return 0;
  CK(cudaMalloc(&p->d_diag_table, sizeof(double*) * p->diags.size()));
  CK(cudaMemcpy(p->d_diag_table, p->diags.data(), sizeof(double*) * p->diags.size(), cudaMemcpyHostToDevice));
  return 0;
}

// RAII staging of a data buffer that may live on the host or on the device
struct Staged {
  void* dev = nullptr;
  void* host = nullptr;
  size_t bytes = 0;
  bool owned = false;
  ~Staged() {
    if (owned && dev) cudaFree(dev);
  }
  // input buffer: copy H2D when it is host memory
  int in(const void* p, size_t n, int mem, cudaStream_t s) {
    bytes = n;
    if (!p || n == 0) return 0;
    if (mem == BQA_MEM_DEVICE) {
      dev = const_cast<void*>(p);
      return 0;
    }
    owned = true;
    CK(cudaMalloc(&dev, n));
    CK(cudaMemcpyAsync(dev, p, n, cudaMemcpyHostToDevice, s));
    return 0;
  }
  // output buffer: allocate a device mirror when it is host memory
  int out(void* p, size_t n, int mem) {
    bytes = n;
    if (!p || n == 0) return 0;
    if (mem == BQA_MEM_DEVICE) {
      dev = p;
      return 0;
    }
    owned = true;
    host = p;
    CK(cudaMalloc(&dev, n));
    return 0;
  }
  int flush(cudaStream_t s) {
    if (host && dev) CK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s));
    return 0;
  }
};

extern "C" {

int bqa_version(void) { return 100; }
const char* bqa_last_error(void) { return g_err.c_str(); }
int64_t bqa_launch_count(void) { return g_launches.load(); }

int bqa_device_count(int32_t* count) {
  BQA_ENTER();
  int c = 0;
  CK(cudaGetDeviceCount(&c));
  *count = c;
  return 0;
}

int bqa_set_device(int32_t device) {
  BQA_ENTER();
  CK(cudaSetDevice(device));
  return 0;
}

int bqa_current_device(int32_t* device) {
  BQA_ENTER();
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (device) *device = dev;
  return 0;
}

int bqa_device_info(int32_t* sm_count, int64_t* smem_per_block_optin, int64_t* l2_bytes) {
  BQA_ENTER();
  int dev = 0, v = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
  if (sm_count) *sm_count = v;
  CK(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  if (smem_per_block_optin) *smem_per_block_optin = v;
  CK(cudaDeviceGetAttribute(&v, cudaDevAttrL2CacheSize, dev));
  if (l2_bytes) *l2_bytes = v;
  return 0;
}

int bqa_plan_create(const bqa_space_desc* sp, bqa_plan** out) {
  BQA_ENTER();
  if (!sp || !out) return fail("bqa_plan_create: null argument");
  if (sp->n_level != 2 && sp->n_level != 3) return fail("bqa_plan_create: n_level must be 2 or 3");
  if (sp->n_atoms < 1 || sp->n_atoms > kMaxAtoms) return fail("bqa_plan_create: n_atoms out of range");
  if (sp->dim < 1) return fail("bqa_plan_create: dim < 1");
  if ((sp->configs == nullptr) != (sp->partners == nullptr))
    return fail("bqa_plan_create: configs and partners must both be given (subspace) or both NULL (full space)");
  bqa_plan* p = new bqa_plan();
  memset(&p->P, 0, sizeof(PlanView));
  p->P.n_atoms = sp->n_atoms;
  p->P.n_level = sp->n_level;
  p->P.dim = sp->dim;
  p->P.full_space = sp->configs == nullptr;
  uint64_t pw = 1;
  for (int a = 0; a < kMaxAtoms; ++a) {
    p->P.pow_level[a] = (uint32_t)std::min<uint64_t>(pw, 0xffffffffu);
    pw *= (uint64_t)sp->n_level;
    if (pw > 0xffffffffull) pw = 0x100000000ull;
  }
  if (p->P.full_space) {
    double full = 1.0;
    for (int a = 0; a < sp->n_atoms; ++a) full *= sp->n_level;
    if ((double)sp->dim != full) {
      delete p;
      return fail("bqa_plan_create: full space needs dim == n_level^n_atoms");
    }
  }
  cudaGetDevice(&p->device);
  int v = 0;
  cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, p->device);
  p->sm_count = v > 0 ? v : 148;
  cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device);
  p->smem_optin = v;
  if (!p->P.full_space) {
    size_t nc = sizeof(uint32_t) * sp->dim, np = sizeof(int32_t) * sp->dim * sp->n_atoms * (sp->n_level - 1);
    if (cudaMalloc(&p->d_configs, nc) != cudaSuccess || cudaMalloc(&p->d_partners, np) != cudaSuccess ||
        cudaMemcpy(p->d_configs, sp->configs, nc, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(p->d_partners, sp->partners, np, cudaMemcpyHostToDevice) != cudaSuccess) {
      bqa_plan_destroy(p);
      return fail("bqa_plan_create: device allocation/copy failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    p->P.configs = p->d_configs;
    p->P.partners = p->d_partners;
  }
  *out = p;
  return 0;
}

void bqa_plan_destroy(bqa_plan* p) {
  if (!p) return;
  if (p->d_configs) cudaFree(p->d_configs);
  if (p->d_partners) cudaFree(p->d_partners);
  for (double* d : p->diags) cudaFree(d);
  if (p->d_diag_table) cudaFree(p->d_diag_table);
  if (p->d_csr) cudaFree(p->d_csr);
  for (void* q : p->owned) cudaFree(q);
  if (p->scratch) cudaFree(p->scratch);
  if (p->ev0) cudaEventDestroy(p->ev0);
  if (p->ev1) cudaEventDestroy(p->ev1);
  delete p;
}

int bqa_plan_enable_timing(bqa_plan* p, int32_t on) {
  BQA_ENTER_PLAN(p);
  if (!p) return fail("null plan");
  if (on && !p->ev0) {
    CK(cudaEventCreate(&p->ev0));
    CK(cudaEventCreate(&p->ev1));
  }
  p->timing = on != 0;
  p->evolve_ms = 0.0;
  p->evolve_launches = 0;
  return 0;
}

int bqa_plan_get_timing(bqa_plan* p, double* evolve_kernel_ms, int64_t* evolve_kernel_launches) {
  BQA_ENTER_PLAN(p);
  if (!p) return fail("null plan");
  if (evolve_kernel_ms) *evolve_kernel_ms = p->evolve_ms;
  if (evolve_kernel_launches) *evolve_kernel_launches = p->evolve_launches;
  return 0;
}

int bqa_plan_set_stream(bqa_plan* p, void* s) {
  BQA_ENTER_PLAN(p);
  if (!p) return fail("null plan");
  p->stream = (cudaStream_t)s;
  return 0;
}

int bqa_plan_add_diagonal(bqa_plan* p, const double* v_host, int32_t* diag_id) {
  BQA_ENTER_PLAN(p);
  if (!p || !v_host) return fail("bqa_plan_add_diagonal: null argument");
  double* d = nullptr;
  CK(cudaMalloc(&d, sizeof(double) * p->P.dim));
  CK(cudaMemcpy(d, v_host, sizeof(double) * p->P.dim, cudaMemcpyHostToDevice));
  p->diags.push_back(d);
  if (diag_id) *diag_id = (int32_t)p->diags.size() - 1;
  return refresh_diag_table(p);
}

int bqa_plan_build_rydberg(bqa_plan* p, const double* pair_host, int32_t* diag_id) {
  BQA_ENTER_PLAN(p);
  if (!p || !pair_host) return fail("bqa_plan_build_rydberg: null argument");
  const int n = p->P.n_atoms;
  double *d = nullptr, *d_pair = nullptr;
  CK(cudaMalloc(&d, sizeof(double) * p->P.dim));
  CK(cudaMalloc(&d_pair, sizeof(double) * n * n));
  CK(cudaMemcpyAsync(d_pair, pair_host, sizeof(double) * n * n, cudaMemcpyHostToDevice, p->stream));
  int blocks = (int)std::min<int64_t>((p->P.dim + 255) / 256, (int64_t)p->sm_count * 16);
  build_rydberg_kernel<<<blocks, 256, 0, p->stream>>>(p->P, d_pair, d);
  LAUNCHED();
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(p->stream));
  cudaFree(d_pair);
  p->diags.push_back(d);
  if (diag_id) *diag_id = (int32_t)p->diags.size() - 1;
  return refresh_diag_table(p);
}

int bqa_plan_get_diagonal(bqa_plan* p, int32_t id, double* v_host) {
  BQA_ENTER_PLAN(p);
  if (!p || id < 0 || id >= (int)p->diags.size()) return fail("bqa_plan_get_diagonal: bad id");
  CK(cudaMemcpy(v_host, p->diags[id], sizeof(double) * p->P.dim, cudaMemcpyDeviceToHost));
  return 0;
}

}  // extern "C"

#include "bqa_api_ops.cuh"
#include "bqa_api_shard.cuh"
#include "bqa_api_noalloc.cuh"
