// Experiment (not product): what bandwidth does a TMA-bulk (cp.async.bulk) multi-plane streaming
// structure reach on this GPU, as a function of chunk size, ring depth and CTAs per SM?
//   read-only and copy (smem -> registers -> st.global.v4 / or TMA bulk store) over NP = 9 planes.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(void *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, uint32_t parity)
{
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(s32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, void *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(s32(src)), "r"(bytes) : "memory");
}

// MODE 0: read only; 1: copy via registers + st.global.v4; 2: copy via TMA bulk store straight from the stage
template <int ROWF, int NP, int STAGES, int MODE, int HALO>
__global__ void __launch_bounds__(ROWF / 4 + 32) k_tma(const float *__restrict__ a, float *__restrict__ b, long long S, unsigned total, float *sink)
{
    constexpr int NCONS = ROWF / 4;
    extern __shared__ __align__(128) unsigned char raw[];
    constexpr int PITCH = ROWF + 8;
    float(*buf)[NP][PITCH] = reinterpret_cast<float(*)[NP][PITCH]>(raw);
    unsigned long long *full = reinterpret_cast<unsigned long long *>(raw + sizeof(float) * STAGES * NP * PITCH);
    unsigned long long *empty = full + STAGES;
    const int tid = threadIdx.x;
    if (tid == 0)
    {
        for (int s = 0; s < STAGES; ++s)
        {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], MODE == 2 ? 1 : NCONS / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid >= NCONS)
    {
        // producer warp
        if (tid == NCONS)
        {
            unsigned it = 0;
            for (unsigned t = blockIdx.x; t < total; t += gridDim.x, ++it)
            {
                const int s = it % STAGES;
                if (it >= STAGES)
                    mbar_wait(&empty[s], ((it / STAGES) - 1) & 1u);
                if (HALO == 0)
                {
                    mbar_expect_tx(&full[s], NP * ROWF * 4u);
#pragma unroll
                    for (int q = 0; q < NP; ++q)
                        bulk_g2s(&buf[s][q][4], a + q * S + (long long)t * ROWF, ROWF * 4u, &full[s]);
                }
                else if (HALO == 1)
                {
                    mbar_expect_tx(&full[s], 3 * ROWF * 4u + 6 * (ROWF + 8) * 4u);
#pragma unroll
                    for (int q = 0; q < NP; ++q)
                        if (q < 3)
                            bulk_g2s(&buf[s][q][4], a + q * S + (long long)t * ROWF, ROWF * 4u, &full[s]);
                        else
                            bulk_g2s(&buf[s][q][0], a + q * S + (long long)t * ROWF - 4, (ROWF + 8) * 4u, &full[s]);
                }
                else
                {
                    mbar_expect_tx(&full[s], NP * ROWF * 4u + 12 * 16u);
#pragma unroll
                    for (int q = 0; q < NP; ++q)
                    {
                        bulk_g2s(&buf[s][q][4], a + q * S + (long long)t * ROWF, ROWF * 4u, &full[s]);
                        if (q >= 3)
                        {
                            bulk_g2s(&buf[s][q][0], a + q * S + (long long)t * ROWF - 4, 16u, &full[s]);
                            bulk_g2s(&buf[s][q][4 + ROWF], a + q * S + (long long)t * ROWF + ROWF, 16u, &full[s]);
                        }
                    }
                }
            }
        }
        return;
    }
    if (MODE == 2 && tid != 0)
        return;
    float acc = 0.f;
    unsigned it = 0;
    for (unsigned t = blockIdx.x; t < total; t += gridDim.x, ++it)
    {
        const int s = it % STAGES;
        mbar_wait(&full[s], (it / STAGES) & 1u);
        if (MODE == 2)
        {
            if (tid == 0)
            {
#pragma unroll
                for (int q = 0; q < NP; ++q)
                    bulk_s2g(b + q * S + (long long)t * ROWF, &buf[s][q][4], ROWF * 4u);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                mbar_arrive(&empty[s]);
            }
            continue;
        }
        float4 v[NP];
#pragma unroll
        for (int q = 0; q < NP; ++q)
            v[q] = *reinterpret_cast<const float4 *>(&buf[s][q][4 + 4 * tid]);
        __syncwarp();
        if ((tid & 31) == 0)
            mbar_arrive(&empty[s]);
        if (MODE == 1)
        {
#pragma unroll
            for (int q = 0; q < NP; ++q)
                *reinterpret_cast<float4 *>(b + q * S + (long long)t * ROWF + 4 * tid) = v[q];
        }
        else
        {
#pragma unroll
            for (int q = 0; q < NP; ++q)
                acc += v[q].x + v[q].y + v[q].z + v[q].w;
        }
    }
    if (MODE == 0 && acc == 123.456f)
        *sink = acc;
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

static float *A, *B, *sink;
static long long S;
static long long N;  // floats per plane
static int nsm;

template <int ROWF, int STAGES, int MODE, int HALO = 0>
void run(int ctas_per_sm)
{
    constexpr int NP = 9;
    auto kern = k_tma<ROWF, NP, STAGES, MODE, HALO>;
    const size_t smem = sizeof(float) * STAGES * NP * (ROWF + 8) + 16 * STAGES;
    if (smem * ctas_per_sm > 227 * 1024) return;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, ROWF / 4 + 32, smem));
    if (occ < ctas_per_sm) return;
    const unsigned total = (unsigned)(N / ROWF);
    float ms = timeit([&] { kern<<<ctas_per_sm * nsm, ROWF / 4 + 32, smem>>>(A, B, S, total, sink); });
    const double bytes = (MODE == 0 ? 1.0 : 2.0) * 9 * N * 4;
    printf(" halo %d chunk %5d B x9, %d stages, %d CTA/SM (%3zu KB in flight/SM), %s : %7.1f GB/s\n", HALO, ROWF * 4, STAGES, ctas_per_sm,
           smem * ctas_per_sm / 1024, MODE == 0 ? "read only " : (MODE == 1 ? "copy LDS+STG" : "copy TMA store"), bytes / ms / 1e6);
}

template <int MODE>
void sweep()
{
    run<512, 2, MODE>(4); run<512, 4, MODE>(2); run<512, 3, MODE>(4); run<512, 6, MODE>(2);
    run<1024, 2, MODE>(1); run<1024, 2, MODE>(2); run<1024, 3, MODE>(2); run<1024, 4, MODE>(1); run<1024, 6, MODE>(1);
    run<2048, 2, MODE>(1); run<2048, 3, MODE>(1);
    run<256, 4, MODE>(4); run<256, 8, MODE>(2); run<256, 4, MODE>(8);
}

int main(int argc, char **argv)
{
    int G = argc > 1 ? atoi(argv[1]) : 8192;
    N = (long long)G * G;
    S = N;
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); nsm = p.multiProcessorCount;
    CK(cudaMalloc(&A, 9 * S * 4 + 1024)); CK(cudaMalloc(&B, 9 * S * 4 + 1024)); A += 64; B += 64; CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(A, 0, 9 * S * 4)); CK(cudaMemset(B, 0, 9 * S * 4));
    printf("G=%d, %d SMs\n", G, nsm);
    if (argc > 2)
    {
        sweep<0>();
        sweep<1>();
        sweep<2>();
        return 0;
    }
    run<1024, 2, 0, 0>(1); run<1024, 2, 0, 1>(1); run<1024, 2, 0, 2>(1);
    run<1024, 2, 0, 0>(2); run<1024, 2, 0, 1>(2); run<1024, 2, 0, 2>(2);
    run<1024, 2, 1, 0>(1); run<1024, 2, 1, 1>(1); run<1024, 2, 1, 2>(1);
    run<1024, 2, 1, 0>(2); run<1024, 2, 1, 1>(2); run<1024, 2, 1, 2>(2);
    run<1024, 3, 1, 0>(1); run<1024, 3, 1, 1>(1); run<1024, 3, 1, 2>(1);
    run<2048, 2, 1, 0>(1); run<2048, 2, 1, 1>(1); run<2048, 2, 1, 2>(1);
    return 0;
}
