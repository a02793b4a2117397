// TEST INFRASTRUCTURE ONLY (tests/host_emu): stand-in for <cuda_runtime.h> that lets g++ compile
// the kernel sources under jrl-lbm_b200/csrc/ for the host.  The resulting tests/host_emu/
// libjrl_lbm_emu.so runs each "CUDA thread" as one iteration of a host loop, so the CPU test
// suite (-m "not gpu") can check kernel logic and the Python host layer against the oracle
// where no GPU exists.  The product never loads it: jrl_native.load() only ever opens the nvcc
// build, and raises when that is missing.
#pragma once
#include <cstring>
static inline int __float_as_int(float x) { int i; std::memcpy(&i, &x, 4); return i; }
static inline float __int_as_float(int i) { float x; std::memcpy(&x, &i, 4); return x; }
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __launch_bounds__(...)
#define __grid_constant__
#define __align__(n) __attribute__((aligned(n)))

struct uint3
{
    unsigned int x, y, z;
};
struct dim3
{
    unsigned int x, y, z;
    dim3(unsigned int x_ = 1, unsigned int y_ = 1, unsigned int z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct float2
{
    float x, y;
};
struct alignas(16) float4
{
    float x, y, z, w;
};
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
extern thread_local uint3 threadIdx, blockIdx;
extern thread_local dim3 blockDim, gridDim;

typedef void *cudaStream_t;
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaDevAttrMultiProcessorCount = 16 };
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline const char *cudaGetErrorString(cudaError_t) { return "host emulation"; }
inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int *v, int, int) { *v = 148; return cudaSuccess; }

using std::max;
using std::min;

namespace jrl_emu {
// JRL_EMU_REVERSE (tests/test_emu_order_independence.py): blocks and threads run serially in DESCENDING order instead
// of ascending / OpenMP-parallel.  A kernel whose result depends on the order in which its threads run -- a thread reading
// what another thread of the same launch writes, like the reference's OUTFLOW rows (propagation.cu:195-255) -- gives
// different bits in the two builds.
template <typename F>
void launch(dim3 grid, dim3 block, F body)
{
    const long long nblocks = (long long)grid.x * grid.y * grid.z;
#ifdef JRL_EMU_REVERSE
    for (long long bi = nblocks - 1; bi >= 0; --bi)
#else
#pragma omp parallel for schedule(static)
    for (long long bi = 0; bi < nblocks; ++bi)
#endif
    {
        gridDim = grid;
        blockDim = block;
        blockIdx.x = (unsigned)(bi % grid.x);
        blockIdx.y = (unsigned)((bi / grid.x) % grid.y);
        blockIdx.z = (unsigned)(bi / ((long long)grid.x * grid.y));
        const long long nthreads = (long long)block.x * block.y * block.z;
        for (long long k = 0; k < nthreads; ++k)
        {
#ifdef JRL_EMU_REVERSE
            const long long t = nthreads - 1 - k;
#else
            const long long t = k;
#endif
            threadIdx.x = (unsigned)(t % block.x);
            threadIdx.y = (unsigned)((t / block.x) % block.y);
            threadIdx.z = (unsigned)(t / ((long long)block.x * block.y));
            body();
        }
    }
}
}  // namespace jrl_emu
