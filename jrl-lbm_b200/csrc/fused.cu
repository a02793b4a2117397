// Fused lattice update: pull streaming + link boundaries + inflow/outflow + (moving wall) +
// moments + collision in ONE pass over the populations (jrl_step_f32).
//
// Data movement per lattice update: 9 fp32 reads + 9 fp32 writes of the populations (72 B, the
// algorithmic minimum for a double-buffered D2Q9 update) plus 1 B of flags; optional streams:
// rho/vel stores (+12 B), force (+8 B), concentration inputs (u: +8 B, mesh: +4 B).
//
// Work decomposition: a thread owns VEC (4, 2 or 1) x-adjacent cells of one row.  For a group
// whose 3 x (VEC+2) neighbourhood is all FLUID ("bulk") the nine upstream rows are read as
// aligned 128/64-bit vectors plus one scalar per x-shifted direction, and written back as
// aligned vectors; that is every access of the periodic shear-layer configs and >99% of any
// large grid.  Groups that touch an obstacle, an inlet/outlet, the symmetry axis or a domain
// edge of a non-FLUID cell take the generic per-cell path built from the same device functions
// as the per-operator kernels (lattice.cuh, stencil.cuh), so both paths are bit-identical to
// the reference's operator chain.  The previous-step moment fields are never mutated.
#include "common.cuh"
#include "stencil.cuh"

namespace jrl {

struct StepArgs
{
    Grid g;
    int row_begin, row_end;  // global rows updated by this launch
    const uint8_t *flags;
    const float *f_in;
    float *f_out;
    const float *phi;
    const float *wall_phi;
    const float *rho;  // previous-step density / concentration (kept at OBSTACLE cells)
    const float *vel;  // previous-step velocity, or the advecting velocity when convection
    float *rho_out;
    float *vel_out;
    const float *force;
    const float *mesh;
    const float *obs_vel;
    const float *inflow_vel;
    float inflow_density;
    int has_inflow_density;
    int use_contact, clamp01, use_wall;
    float half_dx;
    CollideParams cp;
    WallParams wp;
};

template <int VEC>
struct VecIO;
template <>
struct VecIO<4>
{
    static __device__ __forceinline__ void load(const float *p, float v[4])
    {
        const float4 t = *reinterpret_cast<const float4 *>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static __device__ __forceinline__ void store(float *p, const float v[4])
    {
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
    static __device__ __forceinline__ bool all_fluid(const uint8_t *p)
    {
        return *reinterpret_cast<const uint32_t *>(p) == 0x01010101u;
    }
};
template <>
struct VecIO<2>
{
    static __device__ __forceinline__ void load(const float *p, float v[2])
    {
        const float2 t = *reinterpret_cast<const float2 *>(p);
        v[0] = t.x; v[1] = t.y;
    }
    static __device__ __forceinline__ void store(float *p, const float v[2])
    {
        *reinterpret_cast<float2 *>(p) = make_float2(v[0], v[1]);
    }
    static __device__ __forceinline__ bool all_fluid(const uint8_t *p)
    {
        return *reinterpret_cast<const uint16_t *>(p) == 0x0101u;
    }
};
template <>
struct VecIO<1>
{
    static __device__ __forceinline__ void load(const float *p, float v[1]) { v[0] = *p; }
    static __device__ __forceinline__ void store(float *p, const float v[1]) { *p = v[0]; }
    static __device__ __forceinline__ bool all_fluid(const uint8_t *p) { return *p == CT_FLUID; }
};

template <int VEC, int OP, int CONV>
__global__ void __launch_bounds__(kThreads) k_step(const __grid_constant__ StepArgs a)
{
    const Grid &g = a.g;
    const int gpr = g.W / VEC;  // groups per row
    const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= (long long)(a.row_end - a.row_begin) * gpr)
        return;
    const int yl = (int)(gi / gpr);
    const int y = a.row_begin + yl;
    const int x0 = (int)(gi - (long long)yl * gpr) * VEC;
    const long long b = blockIdx.y;
    const uint8_t *fl = a.flags + b * g.HW;
    const float *f = a.f_in + b * 9 * g.HW;
    const float *phi = a.phi ? a.phi + b * g.HW : nullptr;
    const float *rho = a.rho ? a.rho + b * g.HW : nullptr;
    const float *vel = a.vel ? a.vel + b * 2 * g.HW : nullptr;

    // ---- classify the group -----------------------------------------------------------------
    const int yn = y == 0 ? g.H - 1 : y - 1, yp = y + 1 == g.H ? 0 : y + 1;
    const long long r0 = cell(g, y, y, 0), rn = cell(g, y, yn, 0), rp = cell(g, y, yp, 0);
    const int xl = x0 == 0 ? g.W - 1 : x0 - 1, xr = x0 + VEC == g.W ? 0 : x0 + VEC;
    bool bulk = !(g.axisym == AXI_X && x0 == 0) && !(g.axisym == AXI_Y && y == 0);
    bulk = bulk && VecIO<VEC>::all_fluid(fl + r0 + x0) && VecIO<VEC>::all_fluid(fl + rn + x0) &&
           VecIO<VEC>::all_fluid(fl + rp + x0);
    bulk = bulk && fl[r0 + xl] == CT_FLUID && fl[r0 + xr] == CT_FLUID && fl[rn + xl] == CT_FLUID &&
           fl[rn + xr] == CT_FLUID && fl[rp + xl] == CT_FLUID && fl[rp + xr] == CT_FLUID;

    float fc[VEC][9];
    uint8_t me[VEC];
    if (bulk)
    {
        // ---- streaming, bulk: out[q](x) = f[q](y - e_qy, x - e_qx), periodic ---------------------
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            const int ex = JRL_EX(q), ey = JRL_EY(q);
            const float *src = f + q * g.HW + (ey > 0 ? rn : (ey < 0 ? rp : r0));
            float v[VEC];
            if (VEC == 1)
                v[0] = src[ex > 0 ? xl : (ex < 0 ? xr : x0)];
            else
            {
                float t[VEC];
                VecIO<VEC>::load(src + x0, t);
                if (ex == 0)
                {
#pragma unroll
                    for (int k = 0; k < VEC; ++k)
                        v[k] = t[k];
                }
                else if (ex > 0)
                {
                    v[0] = src[xl];
#pragma unroll
                    for (int k = 1; k < VEC; ++k)
                        v[k] = t[k - 1];
                }
                else
                {
#pragma unroll
                    for (int k = 0; k < VEC - 1; ++k)
                        v[k] = t[k + 1];
                    v[VEC - 1] = src[xr];
                }
            }
#pragma unroll
            for (int k = 0; k < VEC; ++k)
                fc[k][q] = v[k];
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            me[k] = CT_FLUID;
    }
    else
    {
        // ---- streaming, generic: every link rule, inflow/outflow, moving wall --------------------
        Inflow in;
        in.rho = const_cast<float *>(a.rho);
        in.vel = const_cast<float *>(a.vel);
        in.inflow_vel = a.inflow_vel;
        in.density = a.inflow_density;
        in.has_density = a.has_inflow_density;
#pragma unroll
        for (int k = 0; k < VEC; ++k)
        {
            const int x = x0 + k;
            me[k] = fl[r0 + x];
            if (phi)
                stream_cell<true, false>(g, f, fl, phi, in, const_cast<float *>(rho), const_cast<float *>(vel), y, x, fc[k]);
            else
                stream_cell<false, false>(g, f, fl, nullptr, in, const_cast<float *>(rho), const_cast<float *>(vel), y, x, fc[k]);
            if (a.use_wall)
                wall_cell(a.wp, g, fl, vel, a.wall_phi + b * g.HW, a.obs_vel ? a.obs_vel + b * 2 * g.HW : nullptr, y, x, fc[k],
                          a.inflow_vel);
        }
    }

    // ---- moments + collision ------------------------------------------------------------------------
    float r_out[VEC], ux_out[VEC], uy_out[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k)
    {
        const int x = x0 + k;
        const long long self = r0 + x;
        const bool obs = me[k] == CT_OBSTACLE;
        float r, ux, uy;
        if (CONV != 0)
        {
            // concentration: C = sum h (obstacle keeps C), optional clamp, advecting u is an input
            r = obs ? (rho ? rho[self] : 0.f) : sum9(fc[k]);
            if (a.clamp01)
                r = fminf(fmaxf(r, 0.f), 1.f);
            ux = vel ? vel[self] : 0.f;
            uy = vel ? vel[g.HW + self] : 0.f;
        }
        else if (obs)
        {
            r = rho ? rho[self] : 0.f;
            ux = vel ? vel[self] : 0.f;
            uy = vel ? vel[g.HW + self] : 0.f;
            if (a.use_contact && x != 0 && y != 0 && x != g.W - 1 && y != g.H - 1)
                r = 0.f;
        }
        else
        {
            r = sum9(fc[k]);
            float mx, my;
            momentum9(fc[k], mx, my);
            const float inv = a.cp.eq.c / r;
            ux = mx * inv;
            uy = my * inv;
        }
        r_out[k] = r;
        ux_out[k] = ux;
        uy_out[k] = uy;
        if (!obs)
        {
            float fx = 0.f, fy = 0.f;
            if (a.force)
            {
                fx = a.force[b * 2 * g.HW + self];
                fy = a.force[b * 2 * g.HW + g.HW + self];
            }
            float rr = 1.f;
            if (CONV == 1)
                rr = a.mesh[b * 2 * g.HW + self];
            else if (CONV == 2)
                rr = y == 0 ? 1.f : a.mesh[b * 2 * g.HW + g.HW + self] - a.half_dx;
            collide_cell<OP, CONV>(a.cp, r, ux, uy, a.force != nullptr, fx, fy, rr, nullptr, fc[k]);
        }
    }

    // ---- stores ----------------------------------------------------------------------------------------
    float *fo = a.f_out + b * 9 * g.HW + r0 + x0;
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        float v[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            v[k] = fc[k][q];
        VecIO<VEC>::store(fo + q * g.HW, v);
    }
    if (a.rho_out)
        VecIO<VEC>::store(a.rho_out + b * g.HW + r0 + x0, r_out);
    if (CONV == 0 && a.vel_out)
    {
        VecIO<VEC>::store(a.vel_out + b * 2 * g.HW + r0 + x0, ux_out);
        VecIO<VEC>::store(a.vel_out + b * 2 * g.HW + g.HW + r0 + x0, uy_out);
    }
}


template <int VEC>
static int launch_step(const StepArgs &a, int B, cudaStream_t s)
{
    const Grid &g = a.g;
    const long long groups = (long long)(a.row_end - a.row_begin) * (g.W / VEC);
    const dim3 grid(blocks_for(groups), (unsigned)B, 1);
#define JRL_STEP(OPV, CONVV)                         \
    do                                               \
    {                                                \
        auto kern = k_step<VEC, OPV, CONVV>;         \
        JRL_LAUNCH(kern, grid, kThreads, s, a);      \
    } while (0)
    if (a.cp.convection)
    {
        if (g.axisym == AXI_X)
            JRL_STEP(OP_BGK, 1);
        else if (g.axisym == AXI_Y)
            JRL_STEP(OP_BGK, 2);
        else
            JRL_STEP(OP_BGK, 3);
    }
    else if (a.cp.op == OP_BGK)
        JRL_STEP(OP_BGK, 0);
    else if (a.cp.op == OP_MRT)
        JRL_STEP(OP_MRT, 0);
    else
        JRL_STEP(OP_KBC, 0);
#undef JRL_STEP
    return 0;
}

}  // namespace jrl

using namespace jrl;

static inline bool aligned(const void *p, size_t n) { return p == nullptr || ((uintptr_t)p % n) == 0; }

extern "C" int jrl_step_f32(const jrl_step_desc *d, void *stream)
{
    const char *fn = "jrl_step_f32";
    JRL_REQUIRE(d, fn, "null descriptor");
    JRL_REQUIRE(d->flags && d->f_in && d->f_out, fn, "flags, f_in, f_out required");
    JRL_REQUIRE(d->f_in != d->f_out, fn, "f_out must not alias f_in (double-buffered pull scheme)");
    JRL_REQUIRE(dims_ok(d->B, d->H, d->W) && d->H >= 3 && d->W >= 3, fn, "B >= 1, H >= 3, W >= 3 required");
    JRL_REQUIRE(d->op >= 0 && d->op <= 2, fn, "unknown operator");
    JRL_REQUIRE(d->op != JRL_OP_MRT || d->mrt, fn, "MRT needs the 9x9 matrix");
    JRL_REQUIRE(!(d->is_convection && d->op != JRL_OP_BGK), fn, "convection is BGK only");
    JRL_REQUIRE(d->op != JRL_OP_KBC || d->kbc_type == JRL_KBC_LBGK || d->kbc_type == JRL_KBC_A || d->kbc_type == JRL_KBC_B ||
                    d->kbc_type == JRL_KBC_C || d->kbc_type == JRL_KBC_D,
                fn, "Please specific a right KBC Type value");
    JRL_REQUIRE(d->axisym >= 0 && d->axisym <= 2, fn, "axisym must be 0, 1 or 2");
    JRL_REQUIRE(d->axisym == 0 || d->is_convection, fn,
                "the axisymmetric FLUID source needs neighbours' post-stream moments: use the two-launch path");
    JRL_REQUIRE(!(d->is_convection && d->axisym != 0) || d->mesh_grid, fn, "axisymmetric concentration step needs mesh_grid");
    JRL_REQUIRE(!d->is_convection || d->vel, fn, "concentration step needs the advecting velocity in `vel`");
    JRL_REQUIRE(!d->use_wall || ((d->phi_obs || d->wall_phi) && d->vel), fn, "moving wall needs phi_obs (or wall_phi) and the previous-step vel");
    JRL_REQUIRE(!d->use_wall || !d->vel_out || d->vel_out != d->vel, fn, "moving wall reads neighbours' previous vel: vel_out must not alias vel");
    JRL_REQUIRE(!d->use_contact || (d->rho && d->rho_out && d->rho_out != d->rho), fn,
                "contact-angle wall density needs rho and a distinct rho_out");
    JRL_REQUIRE(!d->use_contact || (d->H >= 4 && d->W >= 4), fn, "contact-angle wall density needs H, W >= 4");
    const bool slab = d->slab_rows > 0;
    JRL_REQUIRE(!slab || !d->use_contact, fn, "contact-angle wall density is not available on row slabs");
    StepArgs a;
    a.g = make_grid(d->H, d->W, d->axisym);
    if (slab)
    {
        JRL_REQUIRE(d->slab_rows >= 3 && d->slab_row0 >= 0 && d->slab_row0 + d->slab_rows <= d->H, fn, "bad slab");
        a.g.slab = 1;
        a.g.y_off = d->slab_row0;
        a.g.HW = (long long)(d->slab_rows + 2) * d->W;
    }
    const int own0 = slab ? d->slab_row0 : 0, own1 = slab ? d->slab_row0 + d->slab_rows : d->H;
    a.row_begin = d->row_end > d->row_begin ? d->row_begin : own0;
    a.row_end = d->row_end > d->row_begin ? d->row_end : own1;
    JRL_REQUIRE(a.row_begin >= own0 && a.row_end <= own1, fn, "row range outside the stored rows");
    a.flags = d->flags;
    a.f_in = d->f_in;
    a.f_out = d->f_out;
    a.phi = d->phi_obs;
    a.wall_phi = d->wall_phi ? d->wall_phi : d->phi_obs;
    a.rho = d->rho;
    a.vel = d->vel;
    a.rho_out = d->rho_out;
    a.vel_out = d->vel_out;
    a.force = d->force;
    a.mesh = d->mesh_grid;
    a.obs_vel = d->obs_vel;
    a.inflow_vel = d->inflow_vel;
    a.has_inflow_density = d->inflow_density >= 0.0;
    a.inflow_density = (float)d->inflow_density;
    a.use_contact = d->use_contact;
    a.clamp01 = d->clamp01;
    a.use_wall = d->use_wall;
    a.half_dx = (float)(0.5 * d->dx);
    SourceParams sp;
    make_collide(a.cp, sp, d->op, d->kbc_type, d->is_convection, d->axisym, d->dx, d->dt, d->tau, d->mrt);
    a.wp = make_wall(d->dt, d->dx, d->tau);
    cudaStream_t s = (cudaStream_t)stream;
    // widest vector the extents and pointer alignments allow
    int vec = 1;
    const void *f32s[] = {d->f_in, d->f_out, d->rho_out, d->vel_out};
    for (int v = 4; v >= 2; v /= 2)
    {
        bool ok = d->W % v == 0 && d->W >= 2 * v && aligned(d->flags, v);
        for (const void *p : f32s)
            ok = ok && aligned(p, 4 * v);
        if (ok)
        {
            vec = v;
            break;
        }
    }
    if (d->force_vec == 1 || d->force_vec == 2 || d->force_vec == 4)
    {
        JRL_REQUIRE(d->force_vec <= vec, fn, "force_vec wider than extents / alignment allow");
        vec = d->force_vec;
    }
    if (vec == 4)
        launch_step<4>(a, d->B, s);
    else if (vec == 2)
        launch_step<2>(a, d->B, s);
    else
        launch_step<1>(a, d->B, s);
    if (d->use_contact)
    {
        const int ring = 2 * (d->H + d->W) - 4;
        JRL_LAUNCH(k_contact_ring, dim3(blocks_for(ring, 128), d->B), 128, s, a.g, d->rho, d->flags, d->rho_out, d->contact_tan);
    }
    return check_launch(fn);
}
