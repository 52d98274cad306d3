#define RT_NS fast
#define RT_FAST 1
#include "rt_kernels.cuh"
