// bqa_small_launch.h - host interface of the CTA-per-task kernel's translation units (bqa_small_k*.cu): every
// CTA-size variant of evolve_small_kernel is compiled in its own TU so that the library builds in parallel
// and a change to the kernel does not recompile the rest of the C ABI.
#pragma once
#include <cuda_runtime.h>

#include "bqa_small.cuh"

namespace bqa {
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_attributes(cudaFuncAttributes* fa);
// raises the kernel's dynamic shared-memory limit to `dyn`, then asks how many CTAs fit on an SM
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_occupancy(size_t dyn, int* ctas_per_sm);
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_launch(const SmallArgs& A, int grid, size_t dyn, cudaStream_t stream);

#define BQA_SMALL_VARIANTS(X) X(128, 1) X(256, 1) X(512, 1) X(1024, 1) X(512, 2)
#define BQA_SMALL_EXTERN(T, M)                                                              \
  extern template cudaError_t small_kernel_attributes<T, M>(cudaFuncAttributes*);          \
  extern template cudaError_t small_kernel_occupancy<T, M>(size_t, int*);                  \
  extern template cudaError_t small_kernel_launch<T, M>(const SmallArgs&, int, size_t, cudaStream_t);
BQA_SMALL_VARIANTS(BQA_SMALL_EXTERN)
#undef BQA_SMALL_EXTERN
}  // namespace bqa
