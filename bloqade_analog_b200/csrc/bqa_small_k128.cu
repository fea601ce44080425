// bqa_small_k128.cu - translation unit of evolve_small_kernel<128, 1> (see bqa_small_launch.h)
#include "bqa_small_kernel_impl.cuh"
BQA_SMALL_INSTANTIATE(128, 1)
