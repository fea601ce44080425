// bqa_cluster_kernel.cu - translation unit of the cluster engine's kernel (compiled in parallel with
// bqa_api.cu; both are linked into libbqa_b200.so).  Host wrappers only: query + launch.
#include <string.h>

#include "bqa_cluster_fast.cuh"
#include "bqa_cluster_launch.h"

namespace bqa {

size_t cluster_kernel_static_smem() {
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, evolve_cluster_kernel) != cudaSuccess) return (size_t)-1;
  return fa.sharedSizeBytes;
}

static void fill_cfg(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, int csize, int threads, size_t dyn, int grid,
                     cudaStream_t stream) {
  memset(&cfg, 0, sizeof(cfg));
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = dyn;
  cfg.stream = stream;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cfg.gridDim = dim3(grid);
}

cudaError_t cluster_kernel_max_clusters(int csize, int threads, size_t dyn, int sm_count, int* max_clusters) {
  cudaError_t e = cudaFuncSetAttribute(evolve_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  fill_cfg(cfg, attr, csize, threads, dyn, csize * sm_count, 0);
  return cudaOccupancyMaxActiveClusters(max_clusters, evolve_cluster_kernel, &cfg);
}

cudaError_t cluster_kernel_launch(const ClusterArgs& CA, int n_clusters, int threads, size_t dyn, cudaStream_t stream) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  fill_cfg(cfg, attr, CA.csize, threads, dyn, n_clusters * CA.csize, stream);
  return cudaLaunchKernelEx(&cfg, evolve_cluster_kernel, CA);
}

}  // namespace bqa
