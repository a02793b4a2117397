// Test infrastructure (oracle/): host-side launchers for the REFERENCE's own kernel bodies.
//
// oracle/build_ref.sh cuts the `__global__` kernel templates out of the four reference
// .cu files where they lie under /root/reference (everything above the first host
// launcher, i.e. above the first line starting with `void `), writes them to a scratch
// directory as ref_*.inc, and compiles this file against them with oracle/shim/ standing
// in for the CUDA/torch headers. The result, oracle/_ref/libjrl_ref.so, runs the
// reference's arithmetic on the CPU, one call per CUDA "thread". It is used only as
//   * the authority the numpy restatement (oracle/lbm_oracle.py) is pinned against,
//   * the backend of the stub extension modules that let the reference's Python
//     solver classes be imported here to generate tests/golden/*.npz,
//   * the streaming / wall / gradient legs of bench.py's CPU baseline.
// Nothing under jrl-lbm_b200/ links, loads or calls it.
//
// Kernels wrapped (reference file:line):
//   kernel_prop_2d_forward        src/LBM/LBM_propagation/cuda/propagation.cu:22-488
//   kernel_get_grad_2d_forward    src/LBM/LBM_collision/cuda/collision_2d.cu:22-148
//   kernel_get_div_2d_forward     src/LBM/LBM_collision/cuda/collision_2d.cu:150-220
//   kernel_macompute_C_2d_forward src/LBM/LBM_macro_compute/cuda/macro_compute.cu:22-78
//   kernel_macompute_2d_forward   src/LBM/LBM_macro_compute/cuda/macro_compute.cu:80-157
//   kernel_2d_forward             src/LBM/LBM_solid_fluid_coupling/cuda/solid_fluid_coupling.cu:25-252
#include "shim/cpu_cuda_shim.h"
#ifdef _OPENMP
#include <omp.h>
#endif

thread_local jrl_shim_dim3 blockIdx = {0, 0, 0};
thread_local jrl_shim_dim3 blockDim = {512, 1, 1};
thread_local jrl_shim_dim3 threadIdx = {0, 0, 0};

#include "ref_prop.inc"
#include "ref_collision.inc"
#include "ref_macro.inc"
#include "ref_wall.inc"

namespace {

const int kThreads = 512;  // the reference's block size (propagation.cu:506)

inline void set_thread(long idx)
{
    blockDim.x = kThreads;
    blockIdx.x = (unsigned int)(idx / kThreads);
    threadIdx.x = (unsigned int)(idx % kThreads);
}

template <typename F>
void for_each_thread(long n, F body)
{
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < n; ++idx)
    {
        set_thread(idx);
        body();
    }
}

template <typename T>
void run_prop(const uint8_t *flags, const T *f, const T *phi_obs, T *f_new, T *rho, T *vel,
              const T *inflow_vel, double inflow_density, int is_convection, int axisym,
              int B, int H, int W, int outflow_last)
{
    const long n = (long)B * H * W;
    auto body = [&]() {
        kernel_prop_2d_forward<T>(flags, f, phi_obs, f_new, rho, vel, inflow_vel, inflow_density,
                                  is_convection != 0, axisym, B, H, W);
    };
    if (!outflow_last)
    {
        // One possible GPU schedule: threads retire in index order.
        for (long idx = 0; idx < n; ++idx)
        {
            set_thread(idx);
            body();
        }
        return;
    }
    // Another possible GPU schedule, and the one the new kernels define as THE semantics:
    // every non-OUTFLOW thread retires before any OUTFLOW thread reads its neighbour's
    // f_new (propagation.cu:250-254 is a read of data another thread writes).
#pragma omp parallel for schedule(static)
    for (long idx = 0; idx < n; ++idx)
    {
        if (flags[idx] == CELLTYPE_OUTFLOW)
            continue;
        set_thread(idx);
        body();
    }
    // OUTFLOW threads retire afterwards, in index order. (An OUTFLOW cell whose inward
    // neighbour is itself OUTFLOW stays schedule-dependent; no reference demo builds one.)
    for (long idx = 0; idx < n; ++idx)
    {
        if (flags[idx] != CELLTYPE_OUTFLOW)
            continue;
        set_thread(idx);
        body();
    }
}

}  // namespace

extern "C" {

int jrl_ref_prop(int is_double, const uint8_t *flags, const void *f, const void *phi_obs, void *f_new,
                 void *rho, void *vel, const void *inflow_vel, double inflow_density,
                 int is_convection, int axisym, int B, int H, int W, int outflow_last)
{
    if (is_double)
        run_prop<double>(flags, (const double *)f, (const double *)phi_obs, (double *)f_new, (double *)rho,
                         (double *)vel, (const double *)inflow_vel, inflow_density, is_convection, axisym,
                         B, H, W, outflow_last);
    else
        run_prop<float>(flags, (const float *)f, (const float *)phi_obs, (float *)f_new, (float *)rho,
                        (float *)vel, (const float *)inflow_vel, inflow_density, is_convection, axisym,
                        B, H, W, outflow_last);
    return 0;
}

int jrl_ref_grad(int is_double, const void *rho, const uint8_t *flags, void *grad, double dx, int axisym,
                 int B, int H, int W)
{
    const long n = (long)B * H * W;
    if (is_double)
        for_each_thread(n, [&]() { kernel_get_grad_2d_forward<double>((const double *)rho, flags, (double *)grad, dx, axisym, B, H, W); });
    else
        for_each_thread(n, [&]() { kernel_get_grad_2d_forward<float>((const float *)rho, flags, (float *)grad, dx, axisym, B, H, W); });
    return 0;
}

int jrl_ref_div(int is_double, const void *vel, const uint8_t *flags, void *div, double dx, int axisym,
                int B, int H, int W)
{
    const long n = (long)B * H * W;
    if (is_double)
        for_each_thread(n, [&]() { kernel_get_div_2d_forward<double>((const double *)vel, flags, (double *)div, dx, axisym, B, H, W); });
    else
        for_each_thread(n, [&]() { kernel_get_div_2d_forward<float>((const float *)vel, flags, (float *)div, dx, axisym, B, H, W); });
    return 0;
}

int jrl_ref_macro_C(int is_double, double dx, double dt, const void *h, const void *C, const uint8_t *flags,
                    const void *vel, const void *mesh_grid, void *C_new, int axisym, int B, int H, int W)
{
    const long n = (long)B * H * W;
    if (is_double)
        for_each_thread(n, [&]() { kernel_macompute_C_2d_forward<double>(dx, dt, (const double *)h, (const double *)C, flags, (const double *)vel, (const double *)mesh_grid, (double *)C_new, axisym, B, H, W); });
    else
        for_each_thread(n, [&]() { kernel_macompute_C_2d_forward<float>(dx, dt, (const float *)h, (const float *)C, flags, (const float *)vel, (const float *)mesh_grid, (float *)C_new, axisym, B, H, W); });
    return 0;
}

int jrl_ref_macro(int is_double, double dx, double dt, const void *f, const void *rho, const void *vel,
                  const uint8_t *flags, void *rho_new, void *vel_new, void *density_new, int axisym,
                  int B, int H, int W)
{
    const long n = (long)B * H * W;
    if (is_double)
        for_each_thread(n, [&]() { kernel_macompute_2d_forward<double>(dx, dt, (const double *)f, (const double *)rho, (const double *)vel, flags, (double *)rho_new, (double *)vel_new, nullptr, nullptr, nullptr, (double *)density_new, axisym, B, H, W); });
    else
        for_each_thread(n, [&]() { kernel_macompute_2d_forward<float>(dx, dt, (const float *)f, (const float *)rho, (const float *)vel, flags, (float *)rho_new, (float *)vel_new, nullptr, nullptr, nullptr, (float *)density_new, axisym, B, H, W); });
    return 0;
}

int jrl_ref_wall(int is_double, double dt, double dx, double tau, const void *rho, const void *vel,
                 const uint8_t *flags, const void *f, void *f_new, const void *phi_obs, const void *obs_vel,
                 int is_convection, int axisym, int B, int H, int W)
{
    const long n = (long)B * H * W;
    if (is_double)
        for_each_thread(n, [&]() { kernel_2d_forward<double>(dt, dx, tau, (const double *)rho, (const double *)vel, flags, (const double *)f, (double *)f_new, (const double *)phi_obs, (const double *)obs_vel, is_convection != 0, axisym, B, H, W); });
    else
        for_each_thread(n, [&]() { kernel_2d_forward<float>(dt, dx, tau, (const float *)rho, (const float *)vel, flags, (const float *)f, (float *)f_new, (const float *)phi_obs, (const float *)obs_vel, is_convection != 0, axisym, B, H, W); });
    return 0;
}

void jrl_ref_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0)
        omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int jrl_ref_num_threads(void)
{
    int n = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp single
        n = omp_get_num_threads();
    }
#endif
    return n;
}

}  // extern "C"
