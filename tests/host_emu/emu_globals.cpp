// TEST INFRASTRUCTURE ONLY: storage for the emulated CUDA built-in variables.
#include "cuda_runtime.h"
thread_local uint3 threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0};
thread_local dim3 blockDim, gridDim;
