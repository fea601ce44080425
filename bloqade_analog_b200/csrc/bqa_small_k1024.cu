// bqa_small_k1024.cu - translation unit of evolve_small_kernel<1024, 1> (see bqa_small_launch.h)
#include "bqa_small_kernel_impl.cuh"
BQA_SMALL_INSTANTIATE(1024, 1)
