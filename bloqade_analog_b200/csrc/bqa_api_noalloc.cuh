// bqa_api_noalloc.cuh - allocation-free variant of the norm reduction (included at the end of bqa_api.cu, so
// the scratch symbol, cudaGetSymbolAddress and the launches all live in ONE translation unit).
//
// bqa_norm (bqa_api_ops.cuh) keeps its per-CTA partial sums in plan scratch obtained with cudaMalloc.  Callers
// that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the
// same two deterministic passes with the partial sums in a file-scope __device__ array instead.  Every GPU that
// loads the library holds the array (512 KiB) until the process exits.
#pragma once

#include <mutex>

// kNoAllocMaxPartials == BQA_NOALLOC_MAX_PARTIALS (include/bqa_b200.h)
constexpr int kNoAllocMaxPartials = BQA_NOALLOC_MAX_PARTIALS;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) double g_noalloc_partials[kNoAllocMaxPartials];

// norm[seg] = sqrt(sum_b partial[seg][b]): second pass of norm2_segments with the square root on the device
// (IEEE sqrt, the value bqa_norm computes on the host), so the result never has to visit the host.
__global__ void __launch_bounds__(256)
finish_norm_noalloc_kernel(const double* __restrict__ partial, int n_blocks, double* __restrict__ norm) {
  __shared__ double red[32];
  double s = 0.0;
  for (int b = threadIdx.x; b < n_blocks; b += blockDim.x) s += partial[(int64_t)blockIdx.x * n_blocks + b];
  s = block_sum(s, red);
  if (threadIdx.x == 0) norm[blockIdx.x] = sqrt(s);
}

// Device address of g_noalloc_partials on `device` (which must be current).  In host code the array's name is
// only a host-side stand-in for the symbol, never a device address: the address is looked up once per device.
static int noalloc_scratch(int device, double** out) {
  static std::mutex mu;
  static double* addr[kNoAllocMaxDevices] = {};
  if (device < 0 || device >= kNoAllocMaxDevices) return fail("bqa_norm_noalloc: device %d out of range", device);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    double* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_partials);
    if (e != cudaSuccess || !ptr)
      return fail("bqa_norm_noalloc: cudaGetSymbolAddress(g_noalloc_partials) on device %d: %s", device,
                  cudaGetErrorString(e));
    addr[device] = ptr;
  }
  *out = addr[device];
  return 0;
}

extern "C" {

int bqa_norm_noalloc(bqa_plan* p, const void* psi_dev, int32_t n_states, double* norm_dev) {
  BQA_ENTER();
  if (!p || !psi_dev || !norm_dev) return fail("bqa_norm_noalloc: null argument");
  if (n_states < 1) return fail("bqa_norm_noalloc: n_states must be >= 1");
  int cur = -1;
  CK(cudaGetDevice(&cur));
  if (cur != p->device)
    return fail("bqa_norm_noalloc: the plan lives on device %d but device %d is current", p->device, cur);
  const int64_t dim = p->P.dim;
  const int blocks = grid_for(p, dim, 256, 4);  // the grid of bqa_norm: same partial sums, same order
  const int64_t need = (int64_t)n_states * blocks;
  if (need > kNoAllocMaxPartials || n_states > 65535)
    return fail("bqa_norm_noalloc: %d states x %d partial sums = %lld doubles exceed the static scratch of %d doubles "
                "(at most %d states of dimension %lld per call)",
                n_states, blocks, (long long)need, kNoAllocMaxPartials, std::min(kNoAllocMaxPartials / blocks, 65535),
                (long long)dim);
  double* partial = nullptr;
  if (noalloc_scratch(cur, &partial)) return 1;
  // every CTA of the first pass overwrites its own slot and the second pass reads exactly those slots, so the
  // contents the previous call left in the array are never read: nothing has to be cleared
  norm2_partial_kernel<<<dim3(blocks, n_states), 256, 0, p->stream>>>((const cplx*)psi_dev, dim, dim, partial);
  LAUNCHED();
  finish_norm_noalloc_kernel<<<n_states, 256, 0, p->stream>>>(partial, blocks, norm_dev);
  LAUNCHED();
  CK(cudaGetLastError());
  return 0;
}

}  // extern "C"
