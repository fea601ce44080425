// bqa_small_k256.cu - translation unit of evolve_small_kernel<256, 1> (see bqa_small_launch.h)
#include "bqa_small_kernel_impl.cuh"
BQA_SMALL_INSTANTIATE(256, 1)
