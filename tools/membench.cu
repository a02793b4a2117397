// Experiment (not product): what bandwidth do D2Q9-like access patterns reach on this GPU?
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__global__ void k_copy(const float4 *__restrict__ a, float4 *__restrict__ b, long long n)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) b[i] = a[i];
}
// NP planes, plane stride S (float4 units); thread copies element i of every plane
template <int NP>
__global__ void k_planes(const float4 *__restrict__ a, float4 *__restrict__ b, long long n, long long S)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) v[q] = a[q * S + i];
#pragma unroll
    for (int q = 0; q < NP; ++q) b[q * S + i] = v[q];
}
// same, reading rows y-1,y,y+1 like the pull scheme (row length W4 float4), periodic
__global__ void k_stencil(const float4 *__restrict__ a, float4 *__restrict__ b, int H, int W4, long long S)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)H * W4) return;
    int y = (int)(i / W4), x = (int)(i - (long long)y * W4);
    int yn = y == 0 ? H - 1 : y - 1, yp = y + 1 == H ? 0 : y + 1;
    const int ey[9] = {0, 0, 1, 0, -1, 1, 1, -1, -1};
    float4 v[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) { int yy = ey[q] > 0 ? yn : (ey[q] < 0 ? yp : y); v[q] = a[q * S + (long long)yy * W4 + x]; }
#pragma unroll
    for (int q = 0; q < 9; ++q) b[q * S + i] = v[q];
}
template <int NP>
__global__ void k_read(const float4 *__restrict__ a, float *out, long long n, long long S)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0;
#pragma unroll
    for (int q = 0; q < NP; ++q) { float4 v = a[q * S + i]; s += v.x + v.y + v.z + v.w; }
    if (s == 123.456f) out[0] = s;
}
template <int NP>
__global__ void k_write(float4 *__restrict__ b, long long n, long long S)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int q = 0; q < NP; ++q) b[q * S + i] = make_float4(1, 2, 3, (float)q);
}
template <typename F>
float timeit(F f, int reps = 10)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) f();
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) f();
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); CK(cudaGetLastError());
    return ms / reps;
}
int main(int argc, char **argv)
{
    int G = argc > 1 ? atoi(argv[1]) : 8192;
    int H = G, W = G;
    long long n4 = (long long)H * W / 4;
    long long pad4 = argc > 2 ? atoll(argv[2]) : 0;  // plane padding in float4 units
    long long S = n4 + pad4;
    float4 *a, *b; float *out;
    CK(cudaMalloc(&a, 9 * S * 16)); CK(cudaMalloc(&b, 9 * S * 16)); CK(cudaMalloc(&out, 4));
    CK(cudaMemset(a, 0, 9 * S * 16)); CK(cudaMemset(b, 0, 9 * S * 16));
    const int T = 256;
    double bytes9 = 2.0 * 9 * n4 * 16;
    float ms;
    ms = timeit([&] { k_copy<<<(unsigned)((9 * n4 + T - 1) / T), T>>>(a, b, 9 * n4); });
    printf("G=%d pad4=%lld\n copy (1 stream, %0.1f GB)      : %7.1f GB/s\n", G, pad4, bytes9 / 1e9, bytes9 / ms / 1e6);
    ms = timeit([&] { k_planes<9><<<(unsigned)((n4 + T - 1) / T), T>>>(a, b, n4, S); });
    printf(" 9-plane copy                     : %7.1f GB/s\n", bytes9 / ms / 1e6);
    ms = timeit([&] { k_planes<3><<<(unsigned)((n4 + T - 1) / T), T>>>(a, b, n4, S); });
    printf(" 3-plane copy                     : %7.1f GB/s\n", bytes9 / 3 / ms / 1e6);
    ms = timeit([&] { k_stencil<<<(unsigned)((n4 + T - 1) / T), T>>>(a, b, H, W / 4, S); });
    printf(" 9-plane pull-stencil copy        : %7.1f GB/s\n", bytes9 / ms / 1e6);
    ms = timeit([&] { k_read<9><<<(unsigned)((n4 + T - 1) / T), T>>>(a, out, n4, S); });
    printf(" 9-plane read only                : %7.1f GB/s\n", bytes9 / 2 / ms / 1e6);
    ms = timeit([&] { k_write<9><<<(unsigned)((n4 + T - 1) / T), T>>>(b, n4, S); });
    printf(" 9-plane write only               : %7.1f GB/s\n", bytes9 / 2 / ms / 1e6);
    CK(cudaFuncSetAttribute(k_stencil, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int bps : {1, 2, 3, 4, 6, 8}) {
        size_t sm = (size_t)(200 * 1024 / bps) & ~(size_t)1023;
        ms = timeit([&] { k_stencil<<<(unsigned)((n4 + T - 1) / T), T, sm>>>(a, b, H, W / 4, S); });
        printf(" pull-stencil copy, %d blocks(256)/SM : %7.1f GB/s\n", bps, bytes9 / ms / 1e6);
    }
    for (int T2 : {128, 512, 1024}) {
        ms = timeit([&] { k_planes<9><<<(unsigned)((n4 + T2 - 1) / T2), T2>>>(a, b, n4, S); });
        printf(" 9-plane copy, %4d thr/block      : %7.1f GB/s\n", T2, bytes9 / ms / 1e6);
    }
    return 0;
}
