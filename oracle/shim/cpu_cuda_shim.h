// Test infrastructure (oracle/): minimal CUDA-on-CPU shim.
//
// The reference's solver kernels are `__global__` templates that use only
// blockIdx.x/blockDim.x/threadIdx.x, plain pointer arithmetic, printf and min/max.
// Defining those names as thread-local host variables lets g++ compile the kernel
// bodies unchanged; oracle/ref_launch.cpp then calls each kernel once per "thread".
#pragma once
#include <cstdint>
#include <cstdio>
#include <cmath>
#include <algorithm>

#define __global__
#define __device__
#define __host__

struct jrl_shim_dim3 { unsigned int x, y, z; };
extern thread_local jrl_shim_dim3 blockIdx;
extern thread_local jrl_shim_dim3 blockDim;
extern thread_local jrl_shim_dim3 threadIdx;

using std::max;
using std::min;
