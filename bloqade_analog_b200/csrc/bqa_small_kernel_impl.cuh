// bqa_small_kernel_impl.cuh - bodies of the host wrappers of bqa_small_launch.h; included by bqa_small_k*.cu,
// each of which instantiates ONE variant with BQA_SMALL_INSTANTIATE.
#pragma once
#include <algorithm>

#include "bqa_small.cuh"

namespace bqa {
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_attributes(cudaFuncAttributes* fa);
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_occupancy(size_t dyn, int* ctas_per_sm);
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_launch(const SmallArgs& A, int grid, size_t dyn, cudaStream_t stream);

template <int kThreads, int kMinCtas>
cudaError_t small_kernel_attributes(cudaFuncAttributes* fa) {
  return cudaFuncGetAttributes(fa, evolve_small_kernel<kThreads, kMinCtas>);
}
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_occupancy(size_t dyn, int* ctas_per_sm) {
  auto kern = evolve_small_kernel<kThreads, kMinCtas>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(dyn, 1024));
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, kern, kThreads, dyn);
}
template <int kThreads, int kMinCtas>
cudaError_t small_kernel_launch(const SmallArgs& A, int grid, size_t dyn, cudaStream_t stream) {
  evolve_small_kernel<kThreads, kMinCtas><<<grid, kThreads, dyn, stream>>>(A);
  return cudaGetLastError();
}
}  // namespace bqa

#define BQA_SMALL_INSTANTIATE(T, M)                                                                   \
  namespace bqa {                                                                                     \
  template cudaError_t small_kernel_attributes<T, M>(cudaFuncAttributes*);                           \
  template cudaError_t small_kernel_occupancy<T, M>(size_t, int*);                                   \
  template cudaError_t small_kernel_launch<T, M>(const SmallArgs&, int, size_t, cudaStream_t);       \
  }
