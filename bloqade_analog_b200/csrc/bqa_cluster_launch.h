// bqa_cluster_launch.h - host interface of bqa_cluster_kernel.cu (the cluster engine's kernel lives in its
// own translation unit so that the library builds in parallel)
#pragma once
#include <cuda_runtime.h>

#include "bqa_cluster.cuh"

namespace bqa {
size_t cluster_kernel_static_smem();
cudaError_t cluster_kernel_max_clusters(int csize, int threads, size_t dyn, int sm_count, int* max_clusters);
cudaError_t cluster_kernel_launch(const ClusterArgs& CA, int n_clusters, int threads, size_t dyn, cudaStream_t stream);
}  // namespace bqa
