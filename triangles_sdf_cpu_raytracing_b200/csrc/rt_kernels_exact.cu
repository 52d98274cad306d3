#define RT_NS exact
#include "rt_kernels.cuh"
