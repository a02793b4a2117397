// 9-point isotropic stencils (R6) and the regularised moving-wall reconstruction (R5) as
// per-cell device functions.  Clamped neighbour rule (collision_2d.cu:59-72,
// solid_fluid_coupling.cu:98-112).
#pragma once
#include "lattice.cuh"

namespace jrl {

// Outcome-neighbour cell indices of one cell under the clamped rule, q = 0..8.
struct Stencil
{
    long long at[9];
    bool ob[9];  // neighbour is OBSTACLE
};

__device__ __forceinline__ Stencil make_stencil(const Grid &g, const uint8_t *__restrict__ fl, int y, int x)
{
    Stencil s;
    const uint8_t me = fl[cell(g, y, y, x)];
    const Nbr n = neighbours<false>(g, me, y, x);
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        s.at[q] = outcome(g, n, y, x, q);
        s.ob[q] = fl[s.at[q]] == CT_OBSTACLE;
    }
    return s;
}

// kernel_get_grad_2d_forward, collision_2d.cu:22-148: (4*d_axis + d_diag + d_diag) / (12 dx),
// a component is zero when any of its six stencil cells is an obstacle.  v[q] = field value at
// the q-th outcome neighbour.  Returns the raw fp32 sums; the caller divides (in fp64) by 12 dx.
__device__ __forceinline__ void grad_sums(const Stencil &s, const float v[9], float &gx, float &gy)
{
    gx = 0.f;
    gy = 0.f;
    if (s.ob[1] || s.ob[3] || s.ob[5] || s.ob[6] || s.ob[7] || s.ob[8])
        gx = 0.f;
    else
        gx = ((0.f + 4.f * (v[1] - v[3])) + (v[5] - v[6])) + (v[8] - v[7]);
    if (s.ob[2] || s.ob[4] || s.ob[5] || s.ob[8] || s.ob[6] || s.ob[7])
        gy = 0.f;
    else
        gy = ((0.f + 4.f * (v[2] - v[4])) + (v[5] - v[8])) + (v[6] - v[7]);
}

__device__ __forceinline__ float div12(float v, double twelve_dx) { return (float)((double)v / twelve_dx); }

// kernel_get_div_2d_forward, collision_2d.cu:150-220 (no obstacle test on neighbours).
__device__ __forceinline__ float div_sum(const float a[9], const float c[9])
{
    float d = 0.f;
    d = d + 4.f * (a[1] - a[3]);
    d = d + (a[5] - a[6]);
    d = d + (a[8] - a[7]);
    d = d + 4.f * (c[2] - c[4]);
    d = d + (c[5] - c[8]);
    d = d + (c[6] - c[7]);
    return d;
}

// ------------------------------------------------------------------------------------------
// Fluid axisymmetric source, get_G_source_from_axisymmetric, LBM_collision_2d.py:386-469 (the
// non-convection branches).  Evaluated for one non-obstacle... and obstacle cells alike (the
// caller discards obstacle cells).  rho/vel/mesh are batch-element bases.  G[q] includes the
// lattice weight.
// ------------------------------------------------------------------------------------------
struct SourceParams
{
    int axisym;
    float visc;      // (float)(cs2 (tau - 0.5))
    float cs2;       // (float)cs2
    float inv_cs2;   // (float)(1/cs2)
    float dt;        // (float)dt
    float half_dx;   // (float)(0.5 dx)
    float k3;        // (float)(dt (tau - 1) cs2)
    double twelve_dx;
};

__device__ __forceinline__ void axisym_fluid_source(const SourceParams &p, const Grid &g,
                                                    const uint8_t *__restrict__ fl, const float *__restrict__ rho,
                                                    const float *__restrict__ vel, const float *__restrict__ mesh,
                                                    int y, int x, float G[9])
{
    const long long self = cell(g, y, y, x);
    const Stencil s = make_stencil(g, fl, y, x);
    const bool me_obs = fl[self] == CT_OBSTACLE;
    float vx[9], vy[9], rh[9];
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        vx[q] = vel[s.at[q]];
        vy[q] = vel[g.HW + s.at[q]];
        rh[q] = rho[s.at[q]];
    }
    float gux, guy, gvx, gvy;
    grad_sums(s, vx, gux, guy);
    grad_sums(s, vy, gvx, gvy);
    gux = me_obs ? 0.f : div12(gux, p.twelve_dx);
    guy = me_obs ? 0.f : div12(guy, p.twelve_dx);
    gvx = me_obs ? 0.f : div12(gvx, p.twelve_dx);
    gvy = me_obs ? 0.f : div12(gvy, p.twelve_dx);
    const float rho0 = rh[0], ux = vx[0], uy = vy[0];
    if (p.axisym == AXI_X)
    {
        const float r = mesh[self];
        const float A1 = div_ieee((-rho0) * ux, r);
        const float mu = p.visc * rho0;
        // A2 = (d_r(rho cs2) + div(rho u_r u)) * dt / (2 r)
        float pr[9], ax[9], ay[9], a1n[9];
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            pr[q] = rh[q] * p.cs2;
            const float ru = rh[q] * vx[q];
            ax[q] = ru * vx[q];
            ay[q] = ru * vy[q];
            a1n[q] = div_ieee((-rh[q]) * vx[q], mesh[s.at[q]]);
        }
        float gpx, gpy, ga1x, ga1y;
        grad_sums(s, pr, gpx, gpy);
        grad_sums(s, a1n, ga1x, ga1y);
        gpx = me_obs ? 0.f : div12(gpx, p.twelve_dx);
        ga1x = me_obs ? 0.f : div12(ga1x, p.twelve_dx);
        ga1y = me_obs ? 0.f : div12(ga1y, p.twelve_dx);
        const float dv = me_obs ? 0.f : div12(div_sum(ax, ay), p.twelve_dx);
        float A2 = gpx + dv;
        A2 = A2 * (p.dt / (2.f * r));
        float F2x = (mu / r) * (gux - div_ieee(ux, r));
        float F2y = (mu / r) * gvx;
        F2x = F2x + ux * A1;
        F2y = F2y + uy * A1;
        F2x = F2x - p.k3 * ga1x;
        F2y = F2y - p.k3 * ga1y;
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            const float Fe = F2x * (float)JRL_EX(q) + F2y * (float)JRL_EY(q);
            G[q] = ((A1 + A2) + Fe * p.inv_cs2) * weight(q);
        }
    }
    else
    {
        // y-axis variant, LBM_collision_2d.py:428-469
        float r = mesh[g.HW + self] - p.half_dx;
        if (y == 0)
            r = 1.f;
        const float A1 = div_ieee((-rho0) * uy, r);
        const float mu = p.visc * rho0;
        const float H1 = div_ieee(mu * (guy + gvx), r) - div_ieee((rho0 * ux) * uy, r);
        const float H2 = (div_ieee(mu * (gvy * 2.f), r) - div_ieee((2.f * mu) * uy, r * r)) - div_ieee((rho0 * uy) * uy, r);
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            const float eh = (float)JRL_EX(q) * H1 + (float)JRL_EY(q) * H2;
            G[q] = (A1 + eh * p.inv_cs2) * weight(q);
        }
    }
}

// ------------------------------------------------------------------------------------------
// R5: regularised moving-wall reconstruction of one cell, solid_fluid_coupling.cu:85-251.
// `fc` holds the cell's post-stream populations and is updated in place.  vel = previous-step
// velocity field (batch base), obs_vel nullable.  Mixed precision follows the reference's
// double scalars (dt, dx, tau) and literals.
// ------------------------------------------------------------------------------------------
struct WallParams
{
    float c;     // (float)(dx/dt)
    float cs2;   // (float)((double)(c*c) / 3.0)
    double tau;
};

__device__ __forceinline__ void wall_cell(const WallParams &p, const Grid &g, const uint8_t *__restrict__ fl,
                                          const float *__restrict__ vel, const float *__restrict__ phi,
                                          const float *__restrict__ obs_vel, int y, int x, float fc[9],
                                          const float *__restrict__ inflow_vel = nullptr)
{
    const long long self = cell(g, y, y, x);
    const uint8_t me = fl[self];
    if (me == CT_OBSTACLE || x == g.W - 1 || y == g.H - 1)
        return;
    // The per-operator path runs after streaming has overwritten vel at INFLOW cells with
    // inflow_vel (propagation.cu:320-321).  The fused step never mutates the previous-step
    // field, so it passes inflow_vel and the substitution happens on read.
    auto vel_at = [&](long long k, int comp) -> float {
        if (inflow_vel != nullptr && fl[k] == CT_INFLOW)
            return inflow_vel[comp];
        return vel[comp * g.HW + k];
    };
    const Nbr n = neighbours<false>(g, me, y, x);
    const float eps = 1e-10f;
    const float t = phi[self];
    const bool t_ok = (t > -eps) && (t <= 1.f + eps);
    float vt0 = 0.f, vt1 = 0.f, rho_t = 0.f;
    unsigned hit_mask = 0;
    int count = 0;
#pragma unroll
    for (int i = 0; i < 9; ++i)
    {
        const int inv = opp(i);
        const long long o = outcome(g, n, y, x, i);
        if (t_ok && fl[o] == CT_OBSTACLE)
        {
            hit_mask |= 1u << i;
            ++count;
            const long long src = income(g, n, y, x, i);
            float ux = vel_at(src, 0), uy = vel_at(src, 1);
            if (n.on_x && (i == 1 || i == 5 || i == 8))
                ux = -ux;
            if (n.on_y && (i == 2 || i == 5 || i == 6))
                uy = -uy;
            const double tp1 = (double)t + 1.0;
            if (obs_vel != nullptr)
            {
                const float ox = obs_vel[o], oy = obs_vel[g.HW + o];
                vt0 = (float)((double)vt0 + (double)(t * ux + ox) / tp1);
                vt1 = (float)((double)vt1 + (double)(t * uy + oy) / tp1);
                const float dot = (float)JRL_EX(inv) * ox + (float)JRL_EY(inv) * oy;
                rho_t = (float)((double)rho_t + 6.0 * (double)weight(i) * (double)dot);
            }
            else
            {
                vt0 = (float)((double)vt0 + (double)(t * ux) / tp1);
                vt1 = (float)((double)vt1 + (double)(t * uy) / tp1);
            }
        }
        rho_t = rho_t + fc[inv];
    }
    if (count == 0)
        return;
    vt0 = vt0 / (float)count;
    vt1 = vt1 / (float)count;
    // velocity gradient, 9-point stencil / 12, evaluated in fp64
    double sx[9], sy[9];
#pragma unroll
    for (int q = 1; q < 9; ++q)
    {
        const long long o = outcome(g, n, y, x, q);
        sx[q] = (double)vel_at(o, 0);
        sy[q] = (double)vel_at(o, 1);
    }
    const float g00 = (float)((4.0 * sx[1] - 4.0 * sx[3] + sx[5] - sx[6] + sx[8] - sx[7]) / 12.0);
    const float g01 = (float)((4.0 * sx[2] - 4.0 * sx[4] + sx[5] - sx[8] + sx[6] - sx[7]) / 12.0);
    const float g10 = (float)((4.0 * sy[1] - 4.0 * sy[3] + sy[5] - sy[6] + sy[8] - sy[7]) / 12.0);
    const float g11 = (float)((4.0 * sy[2] - 4.0 * sy[4] + sy[5] - sy[8] + sy[6] - sy[7]) / 12.0);
    const double rc = (double)(rho_t * p.cs2) * p.tau;
    const double a00 = (double)((rho_t * vt0) * vt0) - rc * (double)(g00 + g00);
    const double a10 = (double)((rho_t * vt1) * vt0) - rc * (double)(g10 + g01);
    const double a01 = (double)((rho_t * vt0) * vt1) - rc * (double)(g01 + g10);
    const double a11 = (double)((rho_t * vt1) * vt1) - rc * (double)(g11 + g11);
    const double cs2d = (double)p.cs2;
#pragma unroll
    for (int i = 0; i < 9; ++i)
    {
        if (!(hit_mask & (1u << i)))
            continue;
        const int inv = opp(i);
        const int ex = JRL_EX(inv), ey = JRL_EY(inv);
        float pab = 0.f;
        pab = (float)((double)pab + a00 * (double)((float)(ex * ex) - p.cs2));
        pab = (float)((double)pab + a10 * (double)(ey * ex));
        pab = (float)((double)pab + a01 * (double)(ex * ey));
        pab = (float)((double)pab + a11 * (double)((float)(ey * ey) - p.cs2));
        const float lin = rho_t + ((rho_t * p.c) / p.cs2) * (vt0 * (float)ex + vt1 * (float)ey);
        fc[inv] = (float)((double)weight(inv) * ((double)lin + 0.5 * (double)pab / cs2d / cs2d));
    }
}

}  // namespace jrl
