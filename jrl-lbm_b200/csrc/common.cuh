// Host-side plumbing shared by the translation units of libjrl_lbm.so: error reporting,
// argument checks, launch geometry.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cuda_runtime.h>

#include "../../include/jrl_lbm.h"
#include "lattice.cuh"

namespace jrl {

void set_error(const char *fmt, ...);  // abi.cu; thread-local message for jrl_last_error()

inline int fail_arg(const char *fn, const char *what)
{
    set_error("%s: %s", fn, what);
    return JRL_ERR_ARG;
}

inline int check_launch(const char *fn)
{
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
    {
        set_error("%s: CUDA error: %s", fn, cudaGetErrorString(e));
        return JRL_ERR_CUDA;
    }
    return JRL_OK;
}

#define JRL_REQUIRE(cond, fn, msg) \
    do                             \
    {                              \
        if (!(cond))               \
            return jrl::fail_arg(fn, msg); \
    } while (0)

inline bool dims_ok(int B, int H, int W) { return B > 0 && H > 1 && W > 1; }

inline Grid make_grid(int H, int W, int axisym)
{
    Grid g;
    g.H = H;
    g.W = W;
    g.axisym = axisym;
    g.HW = (long long)H * W;
    g.slab = 0;
    g.y_off = 0;
    return g;
}

constexpr int kThreads = 256;

// Kernel launch.  JRL_HOST_EMU is defined only by tests/host_emu/build.sh, which compiles these
// same kernel sources for the host so that the CPU test suite can exercise kernel logic and the
// Python host layer without a GPU; the product library is always the nvcc build.
#ifdef JRL_HOST_EMU
#define JRL_LAUNCH(kern, grid, block, stream, ...) \
    jrl_emu::launch(dim3(grid), dim3(block), [&]() { kern(__VA_ARGS__); })
#else
#define JRL_LAUNCH(kern, grid, block, stream, ...) kern<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__)
#endif

// Programmatic dependent launch for the small, latency-bound grids (the native 130^2 / 128 x 64 runs are chains of
// dependent 3-6 us kernels): a kernel launched with JRL_LAUNCH_PDL may be scheduled while its predecessor in the stream
// is still running, and blocks in pdl_enter() -- the FIRST statement of every kernel launched this way -- until the
// predecessor has completed and its writes are visible (griddepcontrol.wait).  pdl_enter() also releases the kernel's
// own dependents (griddepcontrol.launch_dependents), so successive launches overlap their launch latency, nothing else.
// Without the launch attribute both instructions are no-ops.
#ifdef JRL_HOST_EMU
#define JRL_LAUNCH_PDL(kern, grid, block, stream, ...) JRL_LAUNCH(kern, grid, block, stream, __VA_ARGS__)
inline void pdl_enter() {}
#else
__device__ __forceinline__ void pdl_enter()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
// Set while jrl_run_f32 captures its step sequence into a CUDA graph: the captured launches are plain (fully
// serialised) kernel nodes, without the programmatic-serialisation attribute.
inline bool &pdl_disabled()
{
    static bool off = false;
    return off;
}
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t stream, Args &&...args)
{
    if (pdl_disabled())
    {
        kern<<<grid, block, 0, stream>>>(KArgs(args)...);
        return;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);  // errors are picked up by check_launch() as for <<< >>>
}
#define JRL_LAUNCH_PDL(kern, grid, block, stream, ...) jrl::launch_pdl(kern, dim3(grid), dim3(block), stream, __VA_ARGS__)
#endif

inline unsigned int blocks_for(long long n, int threads = kThreads)
{
    return (unsigned int)((n + threads - 1) / threads);
}

// shared between ops.cu and fused.cu (defined in ops.cu)
struct EqParams;
struct CollideParams;
struct SourceParams;
struct WallParams;
void make_collide(CollideParams &p, SourceParams &sp, int op, int kbc_type, int is_convection, int axisym,
                  double dx, double dt, double tau, const float *mrt_host);
WallParams make_wall(double dt, double dx, double tau);
__global__ void k_contact_ring(Grid g, const float *__restrict__ rho_old, const uint8_t *__restrict__ flags,
                               float *__restrict__ rho_new, float tn);

// `nstrips` consecutive strips y0, y0 + 1, ... in ONE launch (one warp each), out2[2 * k] = strip k (defined in ops.cu)
int strip_flux_n(const float *C, const float *mesh_grid, int y0, int nstrips, int ncol, float *out2, int H, int W, void *stream);

// Number of SMs of the current device (148 on B200); cached.
int sm_count();

}  // namespace jrl
