// placeholder until the matcher kernels land
#include "common.cuh"
