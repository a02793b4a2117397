// Per-operator kernels of libjrl_lbm.so: one launch per reference solver-object method
// (streaming, moments, equilibrium, stencils, collision, moving wall, strip reductions).
// They exist for API parity -- each replaces one pybind call or one chain of 10-100 torch
// ops -- and as the building blocks the fused step (fused.cu) is checked against bit for bit.
// One thread per cell; blockIdx.y = batch element.
#include "common.cuh"
#include "stencil.cuh"

namespace jrl {

#define JRL_CELL_PROLOGUE                                                       \
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;       \
    if (i >= g.HW)                                                              \
        return;                                                                 \
    const long long b = blockIdx.y;                                             \
    const int y = (int)(i / g.W);                                               \
    const int x = (int)(i - (long long)y * g.W);                                \
    (void)x;                                                                    \
    (void)y;

// ---- R1 ------------------------------------------------------------------------------------
template <bool HAS_PHI>
__global__ void __launch_bounds__(kThreads) k_stream(Grid g, const uint8_t *__restrict__ flags,
                                                     const float *__restrict__ f, const float *__restrict__ phi,
                                                     float *__restrict__ f_new, Inflow in)
{
    pdl_enter();
    JRL_CELL_PROLOGUE
    float out[9];
    stream_cell<HAS_PHI>(g, f + b * 9 * g.HW, flags + b * g.HW, HAS_PHI ? phi + b * g.HW : nullptr, in,
                         in.rho ? in.rho + b * g.HW : nullptr, in.vel ? in.vel + b * 2 * g.HW : nullptr, y, x, out);
    float *o = f_new + b * 9 * g.HW + i;
#pragma unroll
    for (int q = 0; q < 9; ++q)
        o[q * g.HW] = out[q];
}

__global__ void __launch_bounds__(kThreads) k_rebounce(Grid g, const float *__restrict__ f,
                                                       const uint8_t *__restrict__ flags, float *__restrict__ f_new)
{
    JRL_CELL_PROLOGUE
    const bool obs = flags[b * g.HW + i] == CT_OBSTACLE;
    const float *s = f + b * 9 * g.HW + i;
    float *o = f_new + b * 9 * g.HW + i;
#pragma unroll
    for (int q = 0; q < 9; ++q)
        o[q * g.HW] = s[(obs ? opp(q) : q) * g.HW];
}

// Streaming's in-place side effects alone (propagation.cu:300, 320-321, 362): rho = inflow_density and
// vel = inflow_vel at INFLOW cells, rho = 1 at INFLOW_2 cells.  The fused step never mutates the previous-step
// fields (lattice.cuh stream_cell<.., WRITE = false>); a caller that still holds them after a fused step gets the
// reference's mutation from this launch.
__global__ void __launch_bounds__(kThreads) k_inflow_apply(Grid g, const uint8_t *__restrict__ flags, float *rho, float *vel,
                                                           const float *__restrict__ inflow_vel, int has_density, float density)
{
    pdl_enter();
    JRL_CELL_PROLOGUE
    const uint8_t me = flags[b * g.HW + i];
    if (me == CT_INFLOW)
    {
        if (rho != nullptr && has_density)
            rho[b * g.HW + i] = density;
        if (vel != nullptr && inflow_vel != nullptr)
        {
            vel[b * 2 * g.HW + i] = inflow_vel[0];
            vel[b * 2 * g.HW + g.HW + i] = inflow_vel[1];
        }
    }
    else if (me == CT_INFLOW_2 && rho != nullptr)
        rho[b * g.HW + i] = 1.f;
}

// ---- R2 / R7 ---------------------------------------------------------------------------------
// mode 0: LBMMacroCompute2d.macro_compute (torch-op path); mode 1: kernel_macompute_2d_forward
template <int MODE>
__global__ void __launch_bounds__(kThreads) k_moments(Grid g, const float *__restrict__ f,
                                                      const float *__restrict__ rho_old,
                                                      const float *__restrict__ vel_old,
                                                      const uint8_t *__restrict__ flags, float *__restrict__ rho_new,
                                                      float *__restrict__ vel_new, int use_contact, float c)
{
    pdl_enter();
    JRL_CELL_PROLOGUE
    const float *s = f + b * 9 * g.HW + i;
    float fc[9];
#pragma unroll
    for (int q = 0; q < 9; ++q)
        fc[q] = s[q * g.HW];
    const bool obs = flags[b * g.HW + i] == CT_OBSTACLE;
    float r, vx, vy;
    if (obs)
    {
        r = rho_old[b * g.HW + i];
        vx = vel_old[b * 2 * g.HW + i];
        vy = vel_old[b * 2 * g.HW + g.HW + i];
        // contact_angle_correctness zeroes obstacle cells off the outer ring; ring cells are
        // finished by k_contact_ring
        if (MODE == 0 && use_contact && x != 0 && y != 0 && x != g.W - 1 && y != g.H - 1)
            r = 0.f;
    }
    else if (MODE == 0)
    {
        r = sum9(fc);
        float mx, my;
        momentum9(fc, mx, my);
        const float inv = c / r;
        vx = mx * inv;
        vy = my * inv;
    }
    else
    {
        r = 0.f + fc[0];
#pragma unroll
        for (int q = 1; q < 9; ++q)
            r = r + fc[q];
        float mx = 0.f, my = 0.f;
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            mx = mx + (fc[q] * (float)JRL_EX(q)) * c;
            my = my + (fc[q] * (float)JRL_EY(q)) * c;
        }
        vx = mx / r;
        vy = my / r;
    }
    rho_new[b * g.HW + i] = r;
    vel_new[b * 2 * g.HW + i] = vx;
    vel_new[b * 2 * g.HW + g.HW + i] = vy;
}

// contact_angle_correctness on the outer ring, LBM_macro_compute_2d.py:237-271.  One thread per
// ring cell.  `pre(cell)` is the density before the correction: rho_old at OBSTACLE cells (the
// caller guarantees rho_new does not alias it), the freshly summed rho_new elsewhere.
__global__ void k_contact_ring(Grid g, const float *__restrict__ rho_old, const uint8_t *__restrict__ flags,
                               float *__restrict__ rho_new, float tn)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int ring = 2 * (g.H + g.W) - 4;
    if (t >= ring)
        return;
    const long long b = blockIdx.y;
    const uint8_t *fl = flags + b * g.HW;
    const float *ro = rho_old + b * g.HW;
    float *rn = rho_new + b * g.HW;
    int y, x;
    if (t < g.W) { y = 0; x = t; }
    else if (t < 2 * g.W) { y = g.H - 1; x = t - g.W; }
    else if (t < 2 * g.W + g.H - 2) { y = t - 2 * g.W + 1; x = 0; }
    else { y = t - (2 * g.W + g.H - 2) + 1; x = g.W - 1; }
    if (fl[(long long)y * g.W + x] != CT_OBSTACLE)
        return;
    auto pre = [&](int yy, int xx) -> float {
        const long long k = (long long)yy * g.W + xx;
        return fl[k] == CT_OBSTACLE ? ro[k] : rn[k];
    };
    const int H = g.H, W = g.W;
    auto left = [&](int yy) { return pre(yy, 2) + tn * fabsf(pre(yy + 1, 1) - pre(yy - 1, 1)); };
    auto right = [&](int yy) { return pre(yy, W - 3) + tn * fabsf(pre(yy + 1, W - 2) - pre(yy - 1, W - 2)); };
    auto bottom = [&](int xx) { return pre(2, xx) + tn * fabsf(pre(1, xx + 1) - pre(1, xx - 1)); };
    auto top = [&](int xx) { return pre(H - 3, xx) + tn * fabsf(pre(H - 2, xx + 1) - pre(H - 2, xx - 1)); };
    float v;
    const bool cy0 = y == 0, cy1 = y == H - 1, cx0 = x == 0, cx1 = x == W - 1;
    if (cy0 && cx0)
        v = 0.5f * (left(1) + bottom(1));
    else if (cy1 && cx0)
        v = 0.5f * (left(H - 2) + top(1));
    else if (cy0 && cx1)
        v = 0.5f * (bottom(W - 2) + right(1));
    else if (cy1 && cx1)
        v = 0.5f * (0.f + 0.f);  // sic: the reference averages two interior (zero) cells (:269)
    else if (cx0)
        v = left(y);
    else if (cx1)
        v = right(y);
    else if (cy0)
        v = bottom(x);
    else
        v = top(x);
    rn[(long long)y * g.W + x] = v;
}

// mode 0: macro_compute_C (torch path, optional clamp); mode 1: kernel_macompute_C_2d_forward
template <int MODE>
__global__ void __launch_bounds__(kThreads) k_moments_C(Grid g, const float *__restrict__ h,
                                                        const float *__restrict__ C_old,
                                                        const uint8_t *__restrict__ flags, float *__restrict__ C_new,
                                                        int clamp01, const float *__restrict__ vel,
                                                        const float *__restrict__ mesh, double half_dt)
{
    JRL_CELL_PROLOGUE
    float v;
    if (flags[b * g.HW + i] == CT_OBSTACLE)
        v = C_old[b * g.HW + i];
    else
    {
        const float *s = h + b * 9 * g.HW + i;
        float fc[9];
#pragma unroll
        for (int q = 0; q < 9; ++q)
            fc[q] = s[q * g.HW];
        if (MODE == 0)
            v = sum9(fc);
        else
        {
            v = 0.f + fc[0];
#pragma unroll
            for (int q = 1; q < 9; ++q)
                v = v + fc[q];
            if (g.axisym == AXI_X)
            {
                const double den = 1.0 + (half_dt * (double)vel[b * 2 * g.HW + i]) / (double)mesh[b * 2 * g.HW + i];
                v = (float)((double)v / den);
            }
        }
    }
    if (MODE == 0 && clamp01)
        v = fminf(fmaxf(v, 0.f), 1.f);
    C_new[b * g.HW + i] = v;
}

// ---- R3 ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_feq(Grid g, const float *__restrict__ rho, const float *__restrict__ vel,
                                                  const float *__restrict__ force, float *__restrict__ feq, EqParams p)
{
    JRL_CELL_PROLOGUE
    const float r = rho[b * g.HW + i];
    float ux = vel[b * 2 * g.HW + i], uy = vel[b * 2 * g.HW + g.HW + i];
    if (force)
        half_force_shift(r, ux, uy, force[b * 2 * g.HW + i], force[b * 2 * g.HW + g.HW + i]);
    float e[9];
    equilibrium(p, r, ux, uy, e);
    float *o = feq + b * 9 * g.HW + i;
#pragma unroll
    for (int q = 0; q < 9; ++q)
        o[q * g.HW] = e[q];
}

// ---- R6 stencil ops ------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_grad(Grid g, const float *__restrict__ s, const uint8_t *__restrict__ flags,
                                                   float *__restrict__ grad, double twelve_dx)
{
    JRL_CELL_PROLOGUE
    const uint8_t *fl = flags + b * g.HW;
    const float *sb = s + b * g.HW;
    float gx = 0.f, gy = 0.f;
    if (fl[i] != CT_OBSTACLE)
    {
        const Stencil st = make_stencil(g, fl, y, x);
        float v[9];
#pragma unroll
        for (int q = 0; q < 9; ++q)
            v[q] = sb[st.at[q]];
        grad_sums(st, v, gx, gy);
        gx = div12(gx, twelve_dx);
        gy = div12(gy, twelve_dx);
    }
    grad[b * 2 * g.HW + i] = gx;
    grad[b * 2 * g.HW + g.HW + i] = gy;
}

__global__ void __launch_bounds__(kThreads) k_div(Grid g, const float *__restrict__ v2, const uint8_t *__restrict__ flags,
                                                  float *__restrict__ dv, double twelve_dx)
{
    JRL_CELL_PROLOGUE
    const uint8_t *fl = flags + b * g.HW;
    float d = 0.f;
    if (fl[i] != CT_OBSTACLE)
    {
        const Stencil st = make_stencil(g, fl, y, x);
        float a[9], c[9];
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            a[q] = v2[b * 2 * g.HW + st.at[q]];
            c[q] = v2[b * 2 * g.HW + g.HW + st.at[q]];
        }
        d = div12(div_sum(a, c), twelve_dx);
    }
    dv[b * g.HW + i] = d;
}

// ---- R4 / R4m / R4k / R6: collision ------------------------------------------------------------
// SRC: 0 = no stencil source; 1 = fluid axisymmetric source (3x3 stencil of rho / vel).
template <int CONV, int SRC>
__global__ void __launch_bounds__(kThreads) k_collide(Grid g, const __grid_constant__ CollideParams p,
                                                      const __grid_constant__ SourceParams sp,
                                                      const float *__restrict__ f, const float *__restrict__ rho,
                                                      const float *__restrict__ vel, const uint8_t *__restrict__ flags,
                                                      const float *__restrict__ force, const float *__restrict__ mesh,
                                                      float *__restrict__ f_out, int row0, int rows)
{
    pdl_enter();
    // global rows [row0, row0 + rows): the whole grid, or the owned rows of a row slab (cell() maps them to storage;
    // whole grid: i == idx)
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)rows * g.W)
        return;
    const long long b = blockIdx.y;
    const int y = row0 + (int)(idx / g.W);
    const int x = (int)(idx - (long long)(y - row0) * g.W);
    const long long i = cell(g, y, y, x);
    const float *s = f + b * 9 * g.HW + i;
    float *o = f_out + b * 9 * g.HW + i;
    float fc[9];
#pragma unroll
    for (int q = 0; q < 9; ++q)
        fc[q] = s[q * g.HW];
    if (flags[b * g.HW + i] != CT_OBSTACLE)
    {
        const float r0 = rho[b * g.HW + i];
        const float ux = vel[b * 2 * g.HW + i], uy = vel[b * 2 * g.HW + g.HW + i];
        float fx = 0.f, fy = 0.f;
        if (force)
        {
            fx = force[b * 2 * g.HW + i];
            fy = force[b * 2 * g.HW + g.HW + i];
        }
        float r = 1.f;
        if (CONV == 1)
            r = mesh[b * 2 * g.HW + i];
        else if (CONV == 2)
            r = y == 0 ? 1.f : mesh[b * 2 * g.HW + g.HW + i] - sp.half_dx;
        float G[9];
        if (SRC)
            axisym_fluid_source(sp, g, flags + b * g.HW, rho + b * g.HW, vel + b * 2 * g.HW, mesh + b * 2 * g.HW, y, x, G);
        collide_cell<-1, CONV>(p, r0, ux, uy, force != nullptr, fx, fy, r, SRC ? G : nullptr, fc);
    }
#pragma unroll
    for (int q = 0; q < 9; ++q)
        o[q * g.HW] = fc[q];
}

// ---- R5 --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_wall(Grid g, WallParams p, const float *__restrict__ vel,
                                                   const uint8_t *__restrict__ flags, const float *__restrict__ f,
                                                   float *__restrict__ f_new, const float *__restrict__ phi,
                                                   const float *__restrict__ obs_vel)
{
    pdl_enter();
    JRL_CELL_PROLOGUE
    const float *s = f + b * 9 * g.HW + i;
    float *o = f_new + b * 9 * g.HW + i;
    float fc[9];
#pragma unroll
    for (int q = 0; q < 9; ++q)
        fc[q] = s[q * g.HW];
    wall_cell(p, g, flags + b * g.HW, vel + b * 2 * g.HW, phi + b * g.HW, obs_vel ? obs_vel + b * 2 * g.HW : nullptr, y, x, fc);
#pragma unroll
    for (int q = 0; q < 9; ++q)
        o[q * g.HW] = fc[q];
}

// ---- R8: strip reductions, one warp, shuffle tree --------------------------------------------------
__device__ __forceinline__ void strip_terms(const float *__restrict__ C, const float *__restrict__ mesh, int y0, int ncol, int W,
                                            int k, float &flux, float &plain)
{
    flux = 0.f;
    plain = 0.f;
    for (int c = k; c < ncol; c += 32)
    {
        // area_k = pi (k+1)^2 - pi k^2 in fp32 (demo_2d_LBM_JRL_C_cylinder.py:176-180)
        const float pi = 3.14159274101257324f;
        const float area = pi * (float)((c + 1) * (c + 1)) - pi * (float)(c * c);
        const float c0 = C[(long long)y0 * W + c], c1 = C[(long long)y0 * W + c + 1];
        const float dz = c0 - C[(long long)(y0 - 1) * W + c];
        const float r0 = mesh[(long long)y0 * W + c], r1 = mesh[(long long)y0 * W + c + 1];
        const float dr = (r1 * c1 - r0 * c0) / r0;
        flux += (dz + dr) * area;
        plain += c0 * area;
    }
}

__global__ void k_strip_flux(const float *__restrict__ C, const float *__restrict__ mesh, int y0, int ncol, int W,
                             float *__restrict__ out2)
{
    pdl_enter();
    // block k = strip y0 + k (the walk loop records two adjacent strips per step in one launch)
    y0 += blockIdx.x;
    out2 += 2 * blockIdx.x;
#ifdef JRL_HOST_EMU
    // one host "thread" plays the 32 lanes and the same xor tree, so the bits equal the GPU's
    float fl[32], pl[32];
    for (int k = 0; k < 32; ++k)
        strip_terms(C, mesh, y0, ncol, W, k, fl[k], pl[k]);
    for (int off = 16; off > 0; off >>= 1)
        for (int k = 0; k < off; ++k)
        {
            fl[k] += fl[k + off];
            pl[k] += pl[k + off];
        }
    out2[0] = fl[0];
    out2[1] = pl[0];
#else
    const int k = threadIdx.x;
    float flux, plain;
    strip_terms(C, mesh, y0, ncol, W, k, flux, plain);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
    {
        flux += __shfl_xor_sync(0xffffffffu, flux, off);
        plain += __shfl_xor_sync(0xffffffffu, plain, off);
    }
    if (k == 0)
    {
        out2[0] = flux;
        out2[1] = plain;
    }
#endif
}

#ifndef JRL_HOST_EMU  // warp-shuffle kernels: GPU only
// ---- R9 (extension): momentum exchange over halfway bounce-back links -------------------------------
__global__ void __launch_bounds__(kThreads) k_momentum_exchange(Grid g, const float *__restrict__ f,
                                                                const uint8_t *__restrict__ flags, double *out2)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long b = blockIdx.y;
    double fx = 0.0, fy = 0.0;
    if (i < g.HW)
    {
        const uint8_t *fl = flags + b * g.HW;
        const int y = (int)(i / g.W), x = (int)(i - (long long)y * g.W);
        if (fl[i] != CT_OBSTACLE)
        {
#pragma unroll
            for (int q = 1; q < 9; ++q)
            {
                const int xx = x + JRL_EX(q), yy = y + JRL_EY(q);
                if (xx < 0 || yy < 0 || xx >= g.W || yy >= g.H)
                    continue;
                if (fl[(long long)yy * g.W + xx] == CT_OBSTACLE)
                {
                    const double v = 2.0 * (double)f[b * 9 * g.HW + q * g.HW + i];
                    fx += JRL_EX(q) * v;
                    fy += JRL_EY(q) * v;
                }
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
    {
        fx += __shfl_xor_sync(0xffffffffu, fx, off);
        fy += __shfl_xor_sync(0xffffffffu, fy, off);
    }
    __shared__ double sx[kThreads / 32], sy[kThreads / 32];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0)
    {
        sx[w] = fx;
        sy[w] = fy;
    }
    __syncthreads();
    if (w == 0)
    {
        fx = l < kThreads / 32 ? sx[l] : 0.0;
        fy = l < kThreads / 32 ? sy[l] : 0.0;
#pragma unroll
        for (int off = 4; off > 0; off >>= 1)
        {
            fx += __shfl_xor_sync(0xffffffffu, fx, off);
            fy += __shfl_xor_sync(0xffffffffu, fy, off);
        }
        if (l == 0 && (fx != 0.0 || fy != 0.0))
        {
            atomicAdd(out2, fx);
            atomicAdd(out2 + 1, fy);
        }
    }
}

#endif  // !JRL_HOST_EMU

// ---- device-side tip motion: the per-step input preparation of the JRL concentration
//      demos -- ~15 host-driven torch ops per step (demo_2d_LBM_JRL_C_cylinder.py:283-307) -- in one launch.
//      Arithmetic follows the torch expressions operation for operation (fp32 tensors, python scalars cast
//      to fp32), so the fields equal the demo's bit for bit.
// phi of the tip body before the union with the top wall: box (cylinder.py:34-36) or 45-degree cone (45degree.py:34-43)
__device__ __forceinline__ float tip_phi(int shape, float mx, float my, float box_cx, float box_cy, float box_radius)
{
    const float ty = (my - box_cy) - box_radius;
    if (shape == JRL_TIP_CONE45)
    {
        const float x1 = box_cx + box_radius, y1 = box_cy + box_radius;  // (box_center + box_radius)[0], [1]: fp32 tensor
        const float cc = -(x1 + y1);
        const float align = ((mx + my) + cc) * 0.70710678118654752440f;  // / math.sqrt(2): reciprocal multiply on CUDA
        return ty > align ? ty : align;
    }
    const float tx = (mx - box_cx) - box_radius;
    return tx > ty ? tx : ty;
}

__global__ void __launch_bounds__(kThreads) k_tip_conc(jrl_tip_desc d)
{
    pdl_enter();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long HW = (long long)d.H * d.W;
    if (i >= HW)
        return;
    const int y = (int)(i / d.W), x = (int)(i - (long long)y * d.W);
    const float mx = d.mesh_grid[i], my = d.mesh_grid[HW + i];
    // CalculateObsPhi_forC (:31-41): signed distance of the tip body, united (min) with the top wall
    float phi = tip_phi(d.shape, mx, my, d.box_cx, d.box_cy, d.box_radius);
    if (d.is_negative)
    {
        const float top = -(my - (float)d.H);
        phi = phi < top ? phi : top;
    }
    uint8_t fl = d.flags[i];
    if (!d.is_negative && y == d.H - 1)
        fl = CT_INFLOW_2;                                   // :287-289
    if (phi < 0.f)
        fl = CT_OBSTACLE;                                   // :290
    if (y == d.walks_int - 1 && x < d.radius_obs_int)
        fl = CT_INFLOW;                                     // :291
    d.flags[i] = fl;
    if (fl == CT_OBSTACLE)
        d.C[i] = 0.f;                                       // :292
    d.phi_obs[i] = phi;
    // :294-307  vel = v[b] * (1 - t) + v[b + 1] * t  (or v[b]);  C_vel = vel / 1 * vel_obs_conv; tip interior stamped
    float vx = d.vel0[i], vy = d.vel0[HW + i];
    if (d.vel1)
    {
        vx = vx * d.w0 + d.vel1[i] * d.w1;
        vy = vy * d.w0 + d.vel1[HW + i] * d.w1;
    }
    vx = vx * d.vel_obs_conv;
    vy = vy * d.vel_obs_conv;
    if (y < d.walks_int && x < d.Rg_radius)
    {
        vx = 0.f;
        vy = d.vel_obs_conv;
    }
    d.C_vel[i] = vx;
    d.C_vel[HW + i] = vy;
}

// Stage A (axisymmetric tip flow): per-step host work of demo_2d_LBM_JRL_fluid_cylinder.py:205-214 in one
// launch -- CalculateObsPhi (:39-49), the wall velocity field, and the OBSTACLE / velocity stamps inside the tip.
__global__ void __launch_bounds__(kThreads) k_tip_fluid(jrl_tip_fluid_desc d)
{
    pdl_enter();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long HW = (long long)d.H * d.W;
    if (i >= HW)
        return;
    const int y = (int)(i / d.W), x = (int)(i - (long long)y * d.W);
    const float mx = d.mesh_grid[i], my = d.mesh_grid[HW + i];
    float phi = tip_phi(d.shape, mx, my, d.box_cx, d.box_cy, d.box_radius);
    const float top = -(my - (float)d.H);
    phi = phi < top ? phi : top;
    d.phi_obs[i] = phi;
    if (d.shape == JRL_TIP_CONE45)
    {
        // fluid_45degree.py:203-220: is_obs = phi <= 0; obs_vel = (0, vel_obs) there (top row zeroed); flags and vel
        // stamped wherever is_obs (vel takes obs_vel, i.e. 0 on the top row)
        const bool is_obs = phi <= 0.f;
        const float wv = (is_obs && y != d.H - 1) ? d.vel_obs : 0.f;
        d.obs_vel[i] = 0.f;
        d.obs_vel[HW + i] = wv;
        if (is_obs)
        {
            d.flags[i] = CT_OBSTACLE;
            d.vel[i] = 0.f;
            d.vel[HW + i] = wv;
        }
        return;
    }
    const bool inside = y < d.walks_int && x < d.Rg_radius;
    d.obs_vel[i] = 0.f;
    d.obs_vel[HW + i] = inside ? d.vel_obs : 0.f;
    if (inside)
    {
        d.flags[i] = CT_OBSTACLE;
        d.vel[i] = 0.f;
        d.vel[HW + i] = d.vel_obs;
    }
}

// ---- halo staging ----------------------------------------------------------------------------------
__global__ void k_halo_copy(float *__restrict__ dst, const float *__restrict__ src, int C, long long plane_dst,
                            long long plane_src, long long off_dst, long long off_src, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const int c = blockIdx.y;
    (void)C;
    dst[c * plane_dst + off_dst + i] = src[c * plane_src + off_src + i];
}

}  // namespace jrl

// ====================================================================================================
// C ABI launchers
// ====================================================================================================
using namespace jrl;

namespace {

inline dim3 cell_grid(const Grid &g, int B) { return dim3(blocks_for(g.HW), (unsigned)B, 1); }

}  // namespace

namespace jrl {

EqParams make_eq(double dx, double dt)
{
    const double c = dx / dt;
    const double cs2 = c * c / 3.0;
    EqParams p;
    p.c = (float)c;
    p.inv_cs2 = (float)(1.0 / cs2);
    return p;
}

void make_collide(CollideParams &p, SourceParams &sp, int op, int kbc_type, int is_convection, int axisym,
                  double dx, double dt, double tau, const float *mrt_host)
{
    const double c = dx / dt;
    const double cs2 = c * c / 3.0;
    p.eq = make_eq(dx, dt);
    p.op = op;
    p.kbc_type = kbc_type;
    p.convection = is_convection;
    p.axisym = axisym;
    p.omega = (float)(1.0 / tau);
    p.dt = (float)dt;
    p.conv_k1 = (float)(dt * (1.0 - 0.5 / tau));
    p.conv_k2 = (float)(1.0 - 0.5 / tau);
    const double beta = 1.0 / (2.0 * tau);
    p.beta = (float)beta;
    p.inv_beta = (float)(1.0 / beta);
    p.two_m_inv_beta = (float)(2.0 - 1.0 / beta);
    for (int k = 0; k < 81; ++k)
        p.mrt[k] = mrt_host ? mrt_host[k] : 0.f;
    sp.axisym = axisym;
    sp.visc = (float)(cs2 * (tau - 0.5));
    sp.cs2 = (float)cs2;
    sp.inv_cs2 = (float)(1.0 / cs2);
    sp.dt = (float)dt;
    sp.half_dx = (float)(0.5 * dx);
    sp.k3 = (float)(dt * (tau - 1.0) * cs2);
    sp.twelve_dx = 12.0 * dx;
}

WallParams make_wall(double dt, double dx, double tau)
{
    WallParams p;
    p.c = (float)(dx / dt);
    p.cs2 = (float)((double)(p.c * p.c) / 3.0);
    p.tau = tau;
    return p;
}

}  // namespace jrl

extern "C" {

int jrl_stream_f32(const uint8_t *flags, const float *f, const float *phi_obs, float *f_new, float *rho, float *vel,
                   const float *inflow_vel, double inflow_density, int axisym, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_stream_f32";
    JRL_REQUIRE(flags && f && f_new, fn, "flags, f and f_new are required");
    JRL_REQUIRE(f != f_new, fn, "f_new must not alias f");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "B >= 1, H >= 2, W >= 2 required");
    JRL_REQUIRE(axisym >= 0 && axisym <= 2, fn, "axisym must be 0, 1 or 2");
    const Grid g = make_grid(H, W, axisym);
    Inflow in;
    in.rho = rho;
    in.vel = vel;
    in.inflow_vel = inflow_vel;
    in.has_density = inflow_density >= 0.0;
    in.density = (float)inflow_density;
    cudaStream_t s = (cudaStream_t)stream;
    if (phi_obs)
        JRL_LAUNCH_PDL(k_stream<true>, cell_grid(g, B), kThreads, s, g, flags, f, phi_obs, f_new, in);
    else
        JRL_LAUNCH_PDL(k_stream<false>, cell_grid(g, B), kThreads, s, g, flags, f, nullptr, f_new, in);
    return check_launch(fn);
}

int jrl_rebounce_f32(const float *f, const uint8_t *flags, float *f_new, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_rebounce_f32";
    JRL_REQUIRE(flags && f && f_new && f != f_new, fn, "f, flags, f_new required; f_new must not alias f");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, 0);
    JRL_LAUNCH(k_rebounce, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, f, flags, f_new);
    return check_launch(fn);
}

int jrl_inflow_apply_f32(const uint8_t *flags, float *rho, float *vel, const float *inflow_vel, double inflow_density, int B,
                         int H, int W, void *stream)
{
    const char *fn = "jrl_inflow_apply_f32";
    JRL_REQUIRE(flags && (rho || vel), fn, "flags and at least one of rho / vel required");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, 0);
    JRL_LAUNCH_PDL(k_inflow_apply, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, flags, rho, vel, inflow_vel,
                   inflow_density >= 0.0 ? 1 : 0, (float)inflow_density);
    return check_launch(fn);
}

int jrl_moments_f32(const float *f, const float *rho_old, const float *vel_old, const uint8_t *flags, float *rho_new,
                    float *vel_new, int use_contact, float contact_tan, double c, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_moments_f32";
    JRL_REQUIRE(f && rho_old && vel_old && flags && rho_new && vel_new, fn, "null pointer");
    JRL_REQUIRE(rho_new != rho_old && vel_new != vel_old, fn, "outputs must not alias the previous-step moments");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    JRL_REQUIRE(!use_contact || (H >= 4 && W >= 4), fn, "contact-angle wall density needs H, W >= 4");
    const Grid g = make_grid(H, W, 0);
    cudaStream_t s = (cudaStream_t)stream;
    JRL_LAUNCH_PDL(k_moments<0>, cell_grid(g, B), kThreads, s, g, f, rho_old, vel_old, flags, rho_new, vel_new, use_contact, (float)c);
    if (use_contact)
    {
        const int ring = 2 * (H + W) - 4;
        JRL_LAUNCH(k_contact_ring, dim3(blocks_for(ring, 128), B), 128, s, g, rho_old, flags, rho_new, contact_tan);
    }
    return check_launch(fn);
}

int jrl_moments_kernel_f32(const float *f, const float *rho_old, const float *vel_old, const uint8_t *flags,
                           float *rho_new, float *vel_new, double c, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_moments_kernel_f32";
    JRL_REQUIRE(f && rho_old && vel_old && flags && rho_new && vel_new, fn, "null pointer");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, 0);
    JRL_LAUNCH_PDL(k_moments<1>, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, f, rho_old, vel_old, flags, rho_new, vel_new, 0, (float)c);
    return check_launch(fn);
}

int jrl_moments_C_f32(const float *h, const float *C_old, const uint8_t *flags, float *C_new, int clamp01, int B, int H,
                      int W, void *stream)
{
    const char *fn = "jrl_moments_C_f32";
    JRL_REQUIRE(h && C_old && flags && C_new, fn, "null pointer");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, 0);
    JRL_LAUNCH(k_moments_C<0>, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, h, C_old, flags, C_new, clamp01, nullptr, nullptr, 0.0);
    return check_launch(fn);
}

int jrl_moments_C_kernel_f32(double dt, const float *h, const float *C_old, const uint8_t *flags, const float *vel,
                             const float *mesh_grid, float *C_new, int axisym, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_moments_C_kernel_f32";
    JRL_REQUIRE(h && C_old && flags && C_new, fn, "null pointer");
    JRL_REQUIRE(axisym != JRL_AXI_X || (vel && mesh_grid), fn, "x-axis symmetry needs vel and mesh_grid");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, axisym);
    JRL_LAUNCH(k_moments_C<1>, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, h, C_old, flags, C_new, 0, vel, mesh_grid, 0.5 * dt);
    return check_launch(fn);
}

int jrl_feq_f32(const float *rho, const float *vel, const float *force, float *feq, double dx, double dt, int B, int H,
                int W, void *stream)
{
    const char *fn = "jrl_feq_f32";
    JRL_REQUIRE(rho && vel && feq, fn, "null pointer");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, 0);
    JRL_LAUNCH(k_feq, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, rho, vel, force, feq, make_eq(dx, dt));
    return check_launch(fn);
}

int jrl_grad_f32(const float *s, const uint8_t *flags, float *grad, double dx, int axisym, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_grad_f32";
    JRL_REQUIRE(s && flags && grad, fn, "null pointer");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, axisym);
    JRL_LAUNCH(k_grad, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, s, flags, grad, 12.0 * dx);
    return check_launch(fn);
}

int jrl_div_f32(const float *v, const uint8_t *flags, float *div, double dx, int axisym, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_div_f32";
    JRL_REQUIRE(v && flags && div, fn, "null pointer");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, axisym);
    JRL_LAUNCH(k_div, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, v, flags, div, 12.0 * dx);
    return check_launch(fn);
}

int jrl_collide_f32(const jrl_collide_desc *d, void *stream)
{
    const char *fn = "jrl_collide_f32";
    JRL_REQUIRE(d, fn, "null descriptor");
    JRL_REQUIRE(d->f && d->rho && d->vel && d->flags && d->f_out, fn, "f, rho, vel, flags, f_out required");
    JRL_REQUIRE(dims_ok(d->B, d->H, d->W), fn, "bad dimensions");
    JRL_REQUIRE(d->op >= 0 && d->op <= 2, fn, "unknown operator");
    JRL_REQUIRE(d->op != JRL_OP_MRT || d->mrt, fn, "MRT needs the 9x9 matrix");
    JRL_REQUIRE(!(d->is_convection && d->op != JRL_OP_BGK), fn, "convection is BGK only (LBM_collision_MRT_2d.py:138)");
    JRL_REQUIRE(d->op != JRL_OP_KBC || d->kbc_type == JRL_KBC_LBGK || d->kbc_type == JRL_KBC_A || d->kbc_type == JRL_KBC_B ||
                    d->kbc_type == JRL_KBC_C || d->kbc_type == JRL_KBC_D,
                fn, "Please specific a right KBC Type value");
    JRL_REQUIRE(d->axisym == 0 || d->mesh_grid, fn, "axisymmetric collision needs mesh_grid");
    const bool stencil_src = d->axisym != 0 && !d->is_convection;
    Grid g = make_grid(d->H, d->W, d->axisym);
    int row0 = 0, rows = d->H;
    if (d->slab_rows > 0)  // row slab: every plane is [slab_rows + 2, W]; the owned rows are collided, ghost rows only read
    {
        JRL_REQUIRE(d->slab_rows >= 3 && d->slab_row0 >= 0 && d->slab_row0 + d->slab_rows <= d->H, fn, "bad slab");
        JRL_REQUIRE(d->f_out != d->f || !stencil_src, fn, "f_out must not alias f");
        g.slab = 1;
        g.y_off = d->slab_row0;
        g.HW = (long long)(d->slab_rows + 2) * d->W;
        row0 = d->slab_row0;
        rows = d->slab_rows;
    }
    CollideParams p;
    SourceParams sp;
    make_collide(p, sp, d->op, d->kbc_type, d->is_convection, d->axisym, d->dx, d->dt, d->tau, d->mrt);
    cudaStream_t s = (cudaStream_t)stream;
    const dim3 grid(blocks_for((long long)rows * d->W), (unsigned)d->B, 1);
#define JRL_LAUNCH_COLLIDE(CONV, SRC)                                                                         \
    do                                                                                                        \
    {                                                                                                         \
        auto kern = k_collide<CONV, SRC>;                                                                     \
        JRL_LAUNCH_PDL(kern, grid, kThreads, s, g, p, sp, d->f, d->rho, d->vel, d->flags, d->force, d->mesh_grid, \
                   d->f_out, row0, rows);                                                                     \
    } while (0)
    if (d->is_convection)
    {
        if (d->axisym == JRL_AXI_X)
            JRL_LAUNCH_COLLIDE(1, 0);
        else if (d->axisym == JRL_AXI_Y)
            JRL_LAUNCH_COLLIDE(2, 0);
        else
            JRL_LAUNCH_COLLIDE(3, 0);
    }
    else if (stencil_src)
        JRL_LAUNCH_COLLIDE(0, 1);
    else
        JRL_LAUNCH_COLLIDE(0, 0);
#undef JRL_LAUNCH_COLLIDE
    return check_launch(fn);
}

int jrl_wall_f32(double dt, double dx, double tau, const float *vel, const uint8_t *flags, const float *f, float *f_new,
                 const float *phi_obs, const float *obs_vel, int axisym, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_wall_f32";
    JRL_REQUIRE(vel && flags && f && f_new && phi_obs, fn, "vel, flags, f, f_new, phi_obs required");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
    const Grid g = make_grid(H, W, axisym);
    JRL_LAUNCH_PDL(k_wall, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, make_wall(dt, dx, tau), vel, flags, f, f_new, phi_obs, obs_vel);
    return check_launch(fn);
}

int jrl_strip_flux_f32(const float *C, const float *mesh_grid, int y0, int ncol, float *out2, int H, int W, void *stream)
{
    const char *fn = "jrl_strip_flux_f32";
    JRL_REQUIRE(C && mesh_grid && out2, fn, "null pointer");
    JRL_REQUIRE(y0 >= 1 && y0 < H && ncol >= 1 && ncol + 1 <= W, fn, "strip outside the grid");
    return jrl::strip_flux_n(C, mesh_grid, y0, 1, ncol, out2, H, W, stream);
}

extern "C++"
{
namespace jrl
{
int strip_flux_n(const float *C, const float *mesh_grid, int y0, int nstrips, int ncol, float *out2, int H, int W, void *stream)
{
    const char *fn = "jrl_strip_flux_f32";
    JRL_REQUIRE(C && mesh_grid && out2, fn, "null pointer");
    JRL_REQUIRE(nstrips >= 1 && y0 >= 1 && y0 + nstrips - 1 < H && ncol >= 1 && ncol + 1 <= W, fn, "strip outside the grid");
#ifdef JRL_HOST_EMU
    JRL_LAUNCH(k_strip_flux, nstrips, 1, (cudaStream_t)stream, C, mesh_grid, y0, ncol, W, out2);
#else
    JRL_LAUNCH_PDL(k_strip_flux, nstrips, 32, (cudaStream_t)stream, C, mesh_grid, y0, ncol, W, out2);
#endif
    return check_launch(fn);
}
}  // namespace jrl
}

int jrl_momentum_exchange_f32(const float *f_post, const uint8_t *flags, double *out2, int B, int H, int W, void *stream)
{
    const char *fn = "jrl_momentum_exchange_f32";
    JRL_REQUIRE(f_post && flags && out2, fn, "null pointer");
    JRL_REQUIRE(dims_ok(B, H, W), fn, "bad dimensions");
#ifdef JRL_HOST_EMU
    return fail_arg(fn, "warp-shuffle kernel: not part of the host emulation build");
#else
    const Grid g = make_grid(H, W, 0);
    JRL_LAUNCH(k_momentum_exchange, cell_grid(g, B), kThreads, (cudaStream_t)stream, g, f_post, flags, out2);
    return check_launch(fn);
#endif
}

int jrl_tip_prepare_conc_f32(const jrl_tip_desc *d, void *stream)
{
    const char *fn = "jrl_tip_prepare_conc_f32";
    JRL_REQUIRE(d, fn, "null descriptor");
    JRL_REQUIRE(d->H >= 3 && d->W >= 3, fn, "H, W >= 3 required");
    JRL_REQUIRE(d->mesh_grid && d->flags && d->C && d->vel0 && d->phi_obs && d->C_vel, fn,
                "mesh_grid, flags, C, vel0, phi_obs, C_vel required");
    JRL_REQUIRE(d->C_vel != d->vel0 && d->C_vel != d->vel1, fn, "C_vel must not alias the velocity snapshots");
    JRL_LAUNCH_PDL(k_tip_conc, blocks_for((long long)d->H * d->W), kThreads, (cudaStream_t)stream, *d);
    return check_launch(fn);
}

int jrl_tip_prepare_fluid_f32(const jrl_tip_fluid_desc *d, void *stream)
{
    const char *fn = "jrl_tip_prepare_fluid_f32";
    JRL_REQUIRE(d, fn, "null descriptor");
    JRL_REQUIRE(d->H >= 3 && d->W >= 3, fn, "H, W >= 3 required");
    JRL_REQUIRE(d->mesh_grid && d->flags && d->vel && d->phi_obs && d->obs_vel, fn, "mesh_grid, flags, vel, phi_obs, obs_vel required");
    JRL_LAUNCH_PDL(k_tip_fluid, blocks_for((long long)d->H * d->W), kThreads, (cudaStream_t)stream, *d);
    return check_launch(fn);
}

// Stage-A walk loop below the ABI (demo_2d_LBM_JRL_fluid_cylinder.py:203-264): per step tip motion -> streaming
// with interpolated bounce-back -> moving-wall reconstruction -> moments -> BGK collision with the axisymmetric
// source (the one operator chain that cannot be a single pass, see jrl_step_f32) -> walks += vel_obs dt.
int jrl_fluid_walk_f32(jrl_fluid_walk_desc *w, int nsteps, void *stream)
{
    const char *fn = "jrl_fluid_walk_f32";
    JRL_REQUIRE(w, fn, "null descriptor");
    JRL_REQUIRE(nsteps >= 0, fn, "nsteps must be >= 0");
    JRL_REQUIRE(w->f[0] && w->f[1] && w->f[2] && w->rho[0] && w->rho[1] && w->vel[0] && w->vel[1], fn,
                "three population buffers and two (rho, vel) pairs required");
    jrl_tip_fluid_desc t = w->tip;
    jrl_collide_desc c = w->collide;
    JRL_REQUIRE(c.B == 1 && c.H == t.H && c.W == t.W && c.mesh_grid && c.flags, fn, "collide descriptor: B = 1, same extents, mesh_grid and flags required");
    double walks = w->walks;
    int cur = 0;  // index of the current (rho, vel) pair
    float *fa = w->f[0], *fb = w->f[1], *fc = w->f[2];
    for (int i = 0; i < nsteps; ++i)
    {
        const int walks_int = (int)(walks + 0.5 * c.dx);
        JRL_REQUIRE(walks_int >= 1 && walks_int < t.H, fn, "the tip left the grid");
        t.walks_int = walks_int;
        t.box_cy = (float)(walks - w->box_radius);
        t.flags = const_cast<uint8_t *>(c.flags);
        t.vel = w->vel[cur];
        int rc = jrl_tip_prepare_fluid_f32(&t, stream);
        if (rc != 0)
            return rc;
        // first half: streaming (interpolated bounce-back) + moving wall + moments in ONE launch (JRL_OP_NONE), the
        // same device functions as jrl_stream_f32 -> jrl_wall_f32 -> jrl_moments_f32, so the same bits; the collision
        // needs the neighbours' new moments and is the second launch (28.0 -> 19.8 us with 5 launches; 3 now)
        jrl_step_desc sd = {};
        sd.B = 1;
        sd.H = c.H;
        sd.W = c.W;
        sd.axisym = c.axisym;
        sd.op = JRL_OP_NONE;
        sd.dx = c.dx;
        sd.dt = c.dt;
        sd.tau = c.tau;
        sd.flags = c.flags;
        sd.f_in = fa;
        sd.f_out = fc;
        sd.phi_obs = t.phi_obs;
        sd.rho = w->rho[cur];
        sd.vel = w->vel[cur];
        sd.rho_out = w->rho[1 - cur];
        sd.vel_out = w->vel[1 - cur];
        sd.inflow_density = -1.0;
        sd.use_wall = 1;
        sd.obs_vel = t.obs_vel;
        rc = jrl_step_f32(&sd, stream);
        if (rc != 0)
            return rc;
        cur = 1 - cur;
        c.f = fc;
        c.rho = w->rho[cur];
        c.vel = w->vel[cur];
        c.f_out = fa;
        rc = jrl_collide_f32(&c, stream);
        if (rc != 0)
            return rc;
        walks += w->vel_obs * c.dt;
    }
    (void)fb;
    w->walks = walks;
    w->current = cur;
    return JRL_OK;
}

int jrl_halo_pack_f32(const float *field, float *staging, int C, int Hloc, int W, int row, int nrows, void *stream)
{
    const char *fn = "jrl_halo_pack_f32";
    JRL_REQUIRE(field && staging && C > 0 && nrows > 0 && row >= 0 && row + nrows <= Hloc, fn, "bad arguments");
    const long long n = (long long)nrows * W;
    JRL_LAUNCH(k_halo_copy, dim3(blocks_for(n), C), kThreads, (cudaStream_t)stream, staging, field, C, n, (long long)Hloc * W, 0,
                                                                               (long long)row * W, n);
    return check_launch(fn);
}

int jrl_halo_unpack_f32(float *field, const float *staging, int C, int Hloc, int W, int row, int nrows, void *stream)
{
    const char *fn = "jrl_halo_unpack_f32";
    JRL_REQUIRE(field && staging && C > 0 && nrows > 0 && row >= 0 && row + nrows <= Hloc, fn, "bad arguments");
    const long long n = (long long)nrows * W;
    JRL_LAUNCH(k_halo_copy, dim3(blocks_for(n), C), kThreads, (cudaStream_t)stream, field, staging, C, (long long)Hloc * W, n,
                                                                               (long long)row * W, 0, n);
    return check_launch(fn);
}

}  // extern "C"
