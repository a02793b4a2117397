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

// Pipelined-kernel geometry (see `namespace pipe` below): cells per tile, ring slots, CTAs per SM.
// Measured on B200 at 8192^2 BGK (tools/exp_step.py, tools/tmabench.cu): 512 x 2 x 2 = 6.39 TB/s;
// 1024 x 2 x 1 = 6.26; 1024 x 3 x 1 = 6.00; 1024 x 4 x 1 = 5.67; 1024 x 2 x 2 = 5.70 -- about 74 KB of
// loads in flight per SM is the optimum, more starves the stores.
#ifndef JRL_PIPE_CW
#define JRL_PIPE_CW 512
#endif
#ifndef JRL_EXP_FAST_OP
#define JRL_EXP_FAST_OP OP_BGK
#endif
#ifndef JRL_PIPE_STAGES
#define JRL_PIPE_STAGES 2
#endif
#ifndef JRL_PIPE_MINB
#define JRL_PIPE_MINB 2
#endif

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
    const uint8_t *tile_mask;  // [B][stored rows][mask_tpr] from jrl_tile_mask_u8, or nullptr
    int mask_tpr;
    unsigned *worklist;        // [2 + cap]: [0] = entries appended by this launch, [2..] = (tile, thread) of deferred groups
    unsigned worklist_cap;
    double *sums;
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
#if defined(JRL_STORE_POLICY) && JRL_STORE_POLICY == 1 && !defined(JRL_HOST_EMU)
        __stcs(reinterpret_cast<float4 *>(p), make_float4(v[0], v[1], v[2], v[3]));
#elif defined(JRL_STORE_POLICY) && JRL_STORE_POLICY == 2 && !defined(JRL_HOST_EMU)
        __stwt(reinterpret_cast<float4 *>(p), make_float4(v[0], v[1], v[2], v[3]));
#elif defined(JRL_STORE_POLICY) && JRL_STORE_POLICY == 3 && !defined(JRL_HOST_EMU)
        __stcg(reinterpret_cast<float4 *>(p), make_float4(v[0], v[1], v[2], v[3]));
#else
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
#endif
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

// ---- streaming of one group of VEC cells by the generic path: every link rule, inflow / outflow,
//      moving wall; reads global memory directly (lattice.cuh, stencil.cuh) ------------------------------
template <int VEC>
__device__ __forceinline__ void step_generic(const StepArgs &a, const Grid &g, long long b, int y, int x0,
                                             float fc[VEC][9], uint8_t me[VEC])
{
    const uint8_t *fl = a.flags + b * g.HW;
    const float *f = a.f_in + b * 9 * g.HW;
    const float *phi = a.phi ? a.phi + b * g.HW : nullptr;
    const float *rho = a.rho ? a.rho + b * g.HW : nullptr;
    const float *vel = a.vel ? a.vel + b * 2 * g.HW : nullptr;
    const long long r0 = cell(g, y, y, 0);
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

// ---- moments + collision + stores of one group whose post-stream populations are in fc ----------------
template <int VEC, int OP, int CONV>
__device__ __forceinline__ void step_finish(const StepArgs &a, const Grid &g, long long b, int y, int x0, long long r0,
                                            float fc[VEC][9], const uint8_t me[VEC], bool active, float acc[3],
                                            const float (*pre)[VEC] = nullptr)
{
    // pre (concentration form, pipelined kernel): this group's advecting velocity [0], [1] and mesh_grid row
    // [2] (x for CONV 1, y for CONV 2), fetched one tile ahead so that their latency is off the critical path
    const float *rho = a.rho ? a.rho + b * g.HW : nullptr;
    const float *vel = a.vel ? a.vel + b * 2 * g.HW : nullptr;
    float r_out[VEC], ux_out[VEC], uy_out[VEC];
#ifdef JRL_EXP_NO_MATH  // experiment only (tools/): data movement without the arithmetic
#pragma unroll
    for (int k = 0; k < VEC; ++k)
        r_out[k] = ux_out[k] = uy_out[k] = fc[k][0];
#else
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
            ux = pre ? pre[0][k] : (vel ? vel[self] : 0.f);
            uy = pre ? pre[1][k] : (vel ? vel[g.HW + self] : 0.f);
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
            acc[0] += r;
            acc[1] += r * ux;
            acc[2] += r * uy;
            float fx = 0.f, fy = 0.f;
            if (a.force)
            {
                fx = a.force[b * 2 * g.HW + self];
                fy = a.force[b * 2 * g.HW + g.HW + self];
            }
            float rr = 1.f;
            if (CONV == 1)
                rr = pre ? pre[2][k] : a.mesh[b * 2 * g.HW + self];
            else if (CONV == 2)
                rr = y == 0 ? 1.f : (pre ? pre[2][k] : a.mesh[b * 2 * g.HW + g.HW + self]) - a.half_dx;
            collide_cell<OP, CONV>(a.cp, r, ux, uy, a.force != nullptr, fx, fy, rr, nullptr, fc[k]);
        }
    }
#endif

    // ---- stores ----------------------------------------------------------------------------------------
    if (!active)
        return;
#ifdef JRL_EXP_NO_STORE  // experiment only: keep the loads alive, never store
    if (fc[0][0] != 123.456f)
        return;
#endif
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

// ---- R8-style diagnostic: {sum rho, sum rho ux, sum rho uy, cells} of the cells a launch updated ------
// acc[] are this thread's fp32 partial sums (zero for threads without cells); warp-shuffle tree in
// fp64, one atomicAdd per warp and quantity into slot (warp id mod JRL_SUM_SLOTS).
__device__ __forceinline__ void add_sums(double *sums, const float acc[3], float cells)
{
#ifdef JRL_HOST_EMU
    double *s = sums;  // slot 0
#pragma omp atomic
    s[0] += (double)acc[0];
#pragma omp atomic
    s[1] += (double)acc[1];
#pragma omp atomic
    s[2] += (double)acc[2];
#pragma omp atomic
    s[3] += (double)cells;
#else
    double v[4] = {(double)acc[0], (double)acc[1], (double)acc[2], (double)cells};
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            v[j] += __shfl_xor_sync(0xffffffffu, v[j], off);
    if ((threadIdx.x & 31u) == 0)
    {
        const unsigned warp = ((blockIdx.x + 13u * blockIdx.y) * blockDim.y + threadIdx.y) * (blockDim.x >> 5) + (threadIdx.x >> 5);
        double *s = sums + 4 * (warp % JRL_SUM_SLOTS);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            atomicAdd(s + j, v[j]);
    }
#endif
}

// ---- prefetch primitives -------------------------------------------------------------------------
// The bulk path's loads are issued BEFORE the flag classification resolves (their addresses are
// valid for every group: rows and columns wrap periodically), as `asm volatile` so that the
// compiler neither sinks them into the bulk branch nor orders them behind the flag compare:
// one DRAM round trip per thread instead of two (profiles/README.md, r1 first capture).
// f_in is read-only for the whole launch (f_out never aliases it), hence ld.global.nc.
#ifndef JRL_HOST_EMU
template <int VEC>
__device__ __forceinline__ void ldg_vec(const float *p, float v[VEC])
{
    if constexpr (VEC == 4)
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "l"(p));
    else if constexpr (VEC == 2)
        asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(v[0]), "=f"(v[1]) : "l"(p));
    else
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v[0]) : "l"(p));
}
__device__ __forceinline__ float ldg_f32_if(const float *p, bool pred)
{
    float v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@p ld.global.nc.f32 %0, [%1];\n\t}"
                 : "=f"(v) : "l"(p), "r"((int)pred));
    return v;
}
template <int VEC>
__device__ __forceinline__ uint32_t ldg_flags(const uint8_t *p)
{
    uint32_t v;
    if constexpr (VEC == 4)
        asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if constexpr (VEC == 2)
    {
        unsigned short h;
        asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(h) : "l"(p));
        v = h;
    }
    else
        asm volatile("ld.global.nc.u8 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ldg_u8_if(const uint8_t *p, bool pred)
{
    uint32_t v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.u32 %0, 1;\n\t@p ld.global.nc.u8 %0, [%1];\n\t}"
                 : "=r"(v) : "l"(p), "r"((int)pred));
    return v;
}
#endif

#ifndef JRL_STEP_MIN_BLOCKS
#define JRL_STEP_MIN_BLOCKS 2
#endif

#ifndef JRL_STEP_MIN_BLOCKS_V1
#define JRL_STEP_MIN_BLOCKS_V1 3  // one cell per thread = the small, latency-bound grids: 3 CTAs per SM (80 registers)
                                  // instead of 2 is 15-20 % faster there (800 x 400 BGK: 12.4 -> 10.3 us per step; 4 = 3)
#endif

template <int VEC, int OP, int CONV>
__global__ void __launch_bounds__(kThreads, VEC == 1 ? JRL_STEP_MIN_BLOCKS_V1 : JRL_STEP_MIN_BLOCKS) k_step(const __grid_constant__ StepArgs a)
{
    // Thread -> cells: blockDim = (TX, TY), TX a multiple of 32 so that a warp never straddles rows;
    // x-group = blockIdx.x * TX + threadIdx.x, row = row_begin + blockIdx.y * TY + threadIdx.y,
    // batch element = blockIdx.z.  No integer division.
    pdl_enter();  // small grids are launched with JRL_LAUNCH_PDL (launch_step); a no-op otherwise
    const Grid &g = a.g;
    const int gpr = g.W / VEC;  // groups per row
    int gx = blockIdx.x * blockDim.x + threadIdx.x;
    int y = a.row_begin + blockIdx.y * blockDim.y + threadIdx.y;
#ifdef JRL_HOST_EMU
    if (gx >= gpr || y >= a.row_end)
        return;
    const bool active = true;
#else
    // Out-of-range threads stay in the warp (clamped to the last group / row, stores suppressed) so
    // that the full-mask shuffles below are well defined.
    const bool active = gx < gpr && y < a.row_end;
    gx = min(gx, gpr - 1);
    y = min(y, a.row_end - 1);
#endif
    const int x0 = gx * VEC;
    const long long b = blockIdx.z;
    const uint8_t *fl = a.flags + b * g.HW;
    const float *f = a.f_in + b * 9 * g.HW;
    const float *phi = a.phi ? a.phi + b * g.HW : nullptr;
    const float *rho = a.rho ? a.rho + b * g.HW : nullptr;
    const float *vel = a.vel ? a.vel + b * 2 * g.HW : nullptr;

    const int yn = y == 0 ? g.H - 1 : y - 1, yp = y + 1 == g.H ? 0 : y + 1;
    const long long r0 = cell(g, y, y, 0), rn = cell(g, y, yn, 0), rp = cell(g, y, yp, 0);
    const int xl = x0 == 0 ? g.W - 1 : x0 - 1, xr = x0 + VEC == g.W ? 0 : x0 + VEC;
    bool bulk = !(g.axisym == AXI_X && x0 == 0) && !(g.axisym == AXI_Y && y == 0);

    float fc[VEC][9];
    uint8_t me[VEC];
#ifndef JRL_HOST_EMU
    // ---- prefetch: 9 population vectors, the x-edge scalars only where no neighbour lane holds them,
    //      3 flag words + edge flag bytes -- all in flight together ------------------------------------
    const unsigned lane = threadIdx.x & 31u;
    // lane-1 / lane+1 hold the x-adjacent group of the SAME row unless this group starts / ends the row
    const bool own_l = lane == 0 || x0 == 0;
    const bool own_r = lane == 31 || x0 + VEC == g.W;
    float t[9][VEC];
    float edge[9];
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        const int ex = JRL_EX(q), ey = JRL_EY(q);
        const float *src = f + q * g.HW + (ey > 0 ? rn : (ey < 0 ? rp : r0));
        if (VEC == 1)
            ldg_vec<1>(src + (ex > 0 ? xl : (ex < 0 ? xr : x0)), t[q]);
        else
        {
            ldg_vec<VEC>(src + x0, t[q]);
            if (ex > 0)
                edge[q] = ldg_f32_if(src + xl, own_l);
            else if (ex < 0)
                edge[q] = ldg_f32_if(src + xr, own_r);
        }
    }
    const uint32_t fluid_word = VEC == 4 ? 0x01010101u : (VEC == 2 ? 0x0101u : 0x01u);
    const uint32_t w0 = ldg_flags<VEC>(fl + r0 + x0), wn = ldg_flags<VEC>(fl + rn + x0), wp = ldg_flags<VEC>(fl + rp + x0);
    const bool e_l = VEC == 1 || own_l, e_r = VEC == 1 || own_r;
    const uint32_t bl0 = ldg_u8_if(fl + r0 + xl, e_l), bln = ldg_u8_if(fl + rn + xl, e_l), blp = ldg_u8_if(fl + rp + xl, e_l);
    const uint32_t br0 = ldg_u8_if(fl + r0 + xr, e_r), brn = ldg_u8_if(fl + rn + xr, e_r), brp = ldg_u8_if(fl + rp + xr, e_r);
    // ---- classify: the whole 3 x (VEC + 2) neighbourhood must be FLUID.  Inside a warp the x-adjacent
    //      columns belong to the neighbour lanes' groups, so "neighbour lane's 3 x VEC block is all
    //      FLUID" (a superset of the two columns needed) is exchanged by shuffle; a group that fails
    //      this stricter test only takes the generic path, which computes identical values.
    const bool mine = w0 == fluid_word && wn == fluid_word && wp == fluid_word;
    bool left_ok = bl0 == CT_FLUID && bln == CT_FLUID && blp == CT_FLUID;
    bool right_ok = br0 == CT_FLUID && brn == CT_FLUID && brp == CT_FLUID;
    if (VEC > 1)
    {
        const bool up = __shfl_up_sync(0xffffffffu, mine, 1), dn = __shfl_down_sync(0xffffffffu, mine, 1);
        left_ok = own_l ? left_ok : up;
        right_ok = own_r ? right_ok : dn;
    }
    bulk = bulk && mine && left_ok && right_ok;
    // x-shifted directions: element x0-1 is the neighbour lane's last element, x0+VEC its first
    if (VEC > 1)
    {
#pragma unroll
        for (int q = 1; q < 9; ++q)
        {
            const int ex = JRL_EX(q);
            if (ex > 0)
            {
                const float s = __shfl_up_sync(0xffffffffu, t[q][VEC - 1], 1);
                edge[q] = own_l ? edge[q] : s;
            }
            else if (ex < 0)
            {
                const float s = __shfl_down_sync(0xffffffffu, t[q][0], 1);
                edge[q] = own_r ? edge[q] : s;
            }
        }
    }
    if (bulk)
    {
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            const int ex = JRL_EX(q);
#pragma unroll
            for (int k = 0; k < VEC; ++k)
            {
                float v;
                if (VEC == 1 || ex == 0)
                    v = t[q][k];
                else if (ex > 0)
                    v = k == 0 ? edge[q] : t[q][k > 0 ? k - 1 : 0];
                else
                    v = k == VEC - 1 ? edge[q] : t[q][k < VEC - 1 ? k + 1 : 0];
                fc[k][q] = v;
            }
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            me[k] = CT_FLUID;
    }
#else
    // ---- host emulation (tests only): classify, then plain loads ----------------------------------
    bulk = bulk && VecIO<VEC>::all_fluid(fl + r0 + x0) && VecIO<VEC>::all_fluid(fl + rn + x0) &&
           VecIO<VEC>::all_fluid(fl + rp + x0);
    bulk = bulk && fl[r0 + xl] == CT_FLUID && fl[r0 + xr] == CT_FLUID && fl[rn + xl] == CT_FLUID &&
           fl[rn + xr] == CT_FLUID && fl[rp + xl] == CT_FLUID && fl[rp + xr] == CT_FLUID;
    if (bulk)
    {
        // out[q](x) = f[q](y - e_qy, x - e_qx), periodic
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            const int ex = JRL_EX(q), ey = JRL_EY(q);
            const float *src = f + q * g.HW + (ey > 0 ? rn : (ey < 0 ? rp : r0));
#pragma unroll
            for (int k = 0; k < VEC; ++k)
            {
                const int x = x0 + k - ex;
                fc[k][q] = src[x < 0 ? g.W - 1 : (x >= g.W ? 0 : x)];
            }
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            me[k] = CT_FLUID;
    }
#endif
    else
        step_generic<VEC>(a, g, b, y, x0, fc, me);
    float acc[3] = {0.f, 0.f, 0.f};
    step_finish<VEC, OP, CONV>(a, g, b, y, x0, r0, fc, me, active, acc);
    if (a.sums)
    {
        float cells = 0.f;
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            cells += me[k] != CT_OBSTACLE ? 1.f : 0.f;
        if (!active)
            acc[0] = acc[1] = acc[2] = cells = 0.f;
        add_sums(a.sums, acc, cells);
    }
}

#ifndef JRL_HOST_EMU
// =====================================================================================================
// Pipelined variant for wide grids: persistent, warp-specialised CTAs; TMA bulk copies (cp.async.bulk,
// SASS UBLKCP) into a shared-memory ring behind full / empty mbarrier pairs.
//
// Why: the register-prefetch kernel above keeps only (resident threads x 144 B) of loads in flight and
// none while a thread is in its collision arithmetic (4.0-4.5 TB/s).  Here a producer warp issues the
// next tile's copies (9 direction rows, + 3 flag rows unless the tile mask says the neighbourhood is all
// FLUID, + velocity / mesh rows in the concentration form) as soon as the consumer warps have pulled the
// stage's previous contents into registers, so loads are in flight regardless of what the consumers do.
// About 74 KB in flight per SM is the optimum (2 CTAs x 2 stages x 18.6 KB): more loads in flight starve
// the stores (tools/tmabench.cu).
//
// Tile = one row segment of CW = 512 cells (128 consumer threads x 4 cells).  Direction q's buffer holds
// f[q][y - e_qy][x0 - 4 .. x0 + cw + 4) so that the pull f[q](x - e_qx) is a shared-memory read at
// offset -e_qx; the 4-float halos wrap periodically at the row ends (separate 16-byte copies there).
// Groups whose 3 x 6 neighbourhood is not all FLUID (or that sit on a symmetry axis) ignore the staged
// data and take the out-of-line generic path, built from the same device functions as k_step: both
// kernels produce identical bits.
// Requirements (checked by the launcher): W % 16 == 0, W >= JRL_PIPE_CW, 16-byte aligned planes.
// =====================================================================================================
namespace pipe {

constexpr int CW = JRL_PIPE_CW, HALO = 4, ROWF = CW + 2 * HALO, FHALO = 16, ROWB = CW + 2 * FHALO;

constexpr unsigned kPoison = 0xffffffffu;  // Stage::tile value that tells the consumers there is no more work

// NEX: extra staged rows (concentration form: advecting velocity x, y and the mesh row)
template <int NEX>
struct Stage
{
    alignas(16) float f[9][ROWF];
    alignas(16) float ex[NEX > 0 ? NEX : 1][NEX > 0 ? CW : 4];
    alignas(16) uint8_t fl[3][ROWB];
    alignas(8) unsigned long long full;   // mbarrier: tile landed (1 arrival + transaction bytes)
    unsigned long long empty;             // mbarrier: every consumer warp has pulled this fill into registers
    unsigned int uniform;                 // this fill: 3 x (cw + 2) neighbourhood all FLUID (tile mask) -> fl[] not loaded
    unsigned int tile;                    // this fill's tile index, or kPoison
    unsigned int pad[2];
};
static_assert(sizeof(Stage<0>) % 16 == 0 && sizeof(Stage<3>) % 16 == 0, "stage stride must keep 16-byte alignment");
__host__ __device__ constexpr int conv_rows(int conv) { return conv == 0 ? 0 : (conv == 3 ? 2 : 3); }


__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
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
#ifdef JRL_TMA_EVICT_FIRST  // experiment: the populations are read exactly once per step
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)), "l"(pol) : "memory");
    return;
#endif
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_arrive(void *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
}

struct Tile
{
    int b, y, x0, cw;
};

__device__ __forceinline__ Tile tile_of(const StepArgs &a, unsigned t, unsigned tpr, unsigned rows)
{
    const unsigned R = t / tpr, c = t - R * tpr, b = R / rows;
    Tile tl;
    tl.b = (int)b;
    tl.y = a.row_begin + (int)(R - b * rows);
    tl.x0 = (int)c * CW;
    tl.cw = min(CW, a.g.W - tl.x0);
    return tl;
}

// All-FLUID 3 x (cw + 2) neighbourhood of tile `tl`, read from the tile mask (three one-byte loads).
__device__ __forceinline__ bool tile_uniform(const StepArgs &a, const Tile &tl)
{
    const Grid &g = a.g;
    const int y = tl.y;
    if (a.tile_mask == nullptr || (g.axisym == AXI_X && tl.x0 == 0) || (g.axisym == AXI_Y && y == 0))
        return false;
    const int yn = y == 0 ? g.H - 1 : y - 1, yp = y + 1 == g.H ? 0 : y + 1;
    const uint8_t *m = a.tile_mask + (long long)tl.b * (g.HW / g.W) * a.mask_tpr + tl.x0 / CW;
    const uint8_t m0 = m[(cell(g, y, yn, 0) / g.W) * a.mask_tpr], m1 = m[(cell(g, y, y, 0) / g.W) * a.mask_tpr],
                  m2 = m[(cell(g, y, yp, 0) / g.W) * a.mask_tpr];
    return m0 && m1 && m2;
}

// one thread: arm the barrier and issue every bulk copy of tile `tl` into `st`; `uniform`: the tile mask
// says the whole neighbourhood is FLUID, so the flag rows are neither copied nor looked at
template <int CONV>
__device__ __forceinline__ void issue_tile(const StepArgs &a, const Tile &tl, Stage<conv_rows(CONV)> &st, bool uniform)
{
    const Grid &g = a.g;
    const int W = g.W, y = tl.y, x0 = tl.x0, cw = tl.cw;
    const int yn = y == 0 ? g.H - 1 : y - 1, yp = y + 1 == g.H ? 0 : y + 1;
    const long long row[3] = {cell(g, y, yn, 0), cell(g, y, y, 0), cell(g, y, yp, 0)};  // index = 1 - e_qy
    const bool f_in = x0 >= HALO && x0 + cw + HALO <= W;      // halos contiguous with the segment
    const bool b_in = x0 >= FHALO && x0 + cw + FHALO <= W;
    const uint32_t fbytes = (uint32_t)(cw + 2 * HALO) * 4u, bbytes = (uint32_t)(cw + 2 * FHALO);
    st.uniform = uniform ? 1u : 0u;  // ordered before the consumers' reads by the arrive (release) below
    mbar_expect_tx(&st.full, 9u * fbytes - 3u * 2u * HALO * 4u + (uniform ? 0u : 3u * bbytes) +
                                 (uint32_t)conv_rows(CONV) * (uint32_t)cw * 4u);  // the 3 e_qx == 0 directions carry no halo
    if (CONV != 0)
    {
        // concentration form: this row segment of the advecting velocity (and of the mesh coordinate the
        // symmetry axis needs) rides along with the populations
        const float *v = a.vel + (long long)tl.b * 2 * g.HW + row[1] + x0;
        bulk_g2s(&st.ex[0][0], v, (uint32_t)cw * 4u, &st.full);
        bulk_g2s(&st.ex[1][0], v + g.HW, (uint32_t)cw * 4u, &st.full);
        if (CONV == 1 || CONV == 2)
            bulk_g2s(&st.ex[2][0], a.mesh + (long long)tl.b * 2 * g.HW + (CONV == 2 ? g.HW : 0) + row[1] + x0, (uint32_t)cw * 4u,
                     &st.full);
    }
    const float *f = a.f_in + (long long)tl.b * 9 * g.HW;
    const uint8_t *fl = a.flags + (long long)tl.b * g.HW;
    const int xl = x0 == 0 ? W - HALO : x0 - HALO, xr = x0 + cw == W ? 0 : x0 + cw;
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        const int ex = JRL_EX(q), ey = JRL_EY(q);
        const float *src = f + q * g.HW + row[ey > 0 ? 0 : (ey < 0 ? 2 : 1)];
        if (ex == 0)
            bulk_g2s(&st.f[q][HALO], src + x0, (uint32_t)cw * 4u, &st.full);
        else if (f_in)
            bulk_g2s(&st.f[q][0], src + x0 - HALO, fbytes, &st.full);
        else
        {
            bulk_g2s(&st.f[q][HALO], src + x0, (uint32_t)cw * 4u, &st.full);
            bulk_g2s(&st.f[q][0], src + xl, HALO * 4u, &st.full);
            bulk_g2s(&st.f[q][HALO + cw], src + xr, HALO * 4u, &st.full);
        }
    }
    const int bl = x0 == 0 ? W - FHALO : x0 - FHALO;
    if (uniform)
        return;
#pragma unroll
    for (int r = 0; r < 3; ++r)
    {
        const uint8_t *src = fl + row[r];
        if (b_in)
            bulk_g2s(&st.fl[r][0], src + x0 - FHALO, bbytes, &st.full);
        else
        {
            bulk_g2s(&st.fl[r][FHALO], src + x0, (uint32_t)cw, &st.full);
            bulk_g2s(&st.fl[r][0], src + bl, FHALO, &st.full);
            bulk_g2s(&st.fl[r][FHALO + cw], src + xr, FHALO, &st.full);
        }
    }
}

// All-OBSTACLE group away from the domain edges (see the consumer loop): store the reflected populations and carry the
// moments over, exactly as step_finish() does for OBSTACLE cells (old rho / u / C kept; concentration clamp applied;
// interior solid cells get rho = 0 under the contact-angle option, LBM_macro_compute_2d.py:237-271).
template <int VEC>
struct SolidPops
{
    float v[VEC][9];
};
template <int VEC, int CONV>
__device__ __noinline__ void solid_group_store(const StepArgs &a, int b, int y, int x0, const SolidPops<VEC> sp)
{
    const Grid &g = a.g;
    const long long r0 = cell(g, y, y, 0);
    const float *rho = a.rho ? a.rho + b * g.HW : nullptr;
    const float *vel = a.vel ? a.vel + b * 2 * g.HW : nullptr;
    float *fo = a.f_out + (long long)b * 9 * g.HW + r0 + x0;
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        float v[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            v[k] = sp.v[k][q];
        VecIO<VEC>::store(fo + q * g.HW, v);
    }
    float r[VEC], ux[VEC], uy[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k)
    {
        const long long self = r0 + x0 + k;
        r[k] = rho ? rho[self] : 0.f;
        if (CONV != 0)
        {
            if (a.clamp01)
                r[k] = fminf(fmaxf(r[k], 0.f), 1.f);
        }
        else
        {
            ux[k] = vel ? vel[self] : 0.f;
            uy[k] = vel ? vel[g.HW + self] : 0.f;
            if (a.use_contact)
                r[k] = 0.f;  // interior by construction (1 <= y <= H - 2, 1 <= x <= W - 2)
        }
    }
    if (a.rho_out)
        VecIO<VEC>::store(a.rho_out + b * g.HW + r0 + x0, r);
    if (CONV == 0 && a.vel_out)
    {
        VecIO<VEC>::store(a.vel_out + b * 2 * g.HW + r0 + x0, ux);
        VecIO<VEC>::store(a.vel_out + b * 2 * g.HW + g.HW + r0 + x0, uy);
    }
}

// The rare path of the pipelined kernel, kept out of line so that its registers (neighbour indices, link
// rules, wall stencil) do not count against the bulk path: one group of VEC cells by the generic rules,
// moments, collision and stores included.  Returns {sum rho, sum rho ux, sum rho uy, cells}.
template <int VEC, int OP, int CONV>
__device__ __noinline__ float4 step_group_generic(const StepArgs &a, int b, int y, int x0)
{
    float fc[VEC][9];
    uint8_t me[VEC];
    float acc[3] = {0.f, 0.f, 0.f};
    step_generic<VEC>(a, a.g, b, y, x0, fc, me);
    step_finish<VEC, OP, CONV>(a, a.g, b, y, x0, cell(a.g, y, y, 0), fc, me, true, acc);
    float cells = 0.f;
#pragma unroll
    for (int k = 0; k < VEC; ++k)
        cells += me[k] != CT_OBSTACLE ? 1.f : 0.f;
    return make_float4(acc[0], acc[1], acc[2], cells);
}

// VEC cells per consumer thread (CW / VEC consumer threads + one producer warp per CTA), STAGES ring
// slots, MINB CTAs per SM.  Warp-specialised: lane 0 of the last warp only issues TMA copies (it waits
// on a stage's `empty` barrier, arms `full` and issues the tile); the consumer warps wait on `full`,
// pull their cells into registers, arrive on `empty` (one arrival per warp) and then do the arithmetic
// and the stores -- so a stage is refilled while its previous contents are still being computed on,
// and no warp ever waits for another consumer warp.
// Registers per thread that let MINB CTAs of `threads` threads be resident: each of the SM's four
// sub-partitions has 16 K registers and takes every fourth warp.  (ptxas' own choice under
// __launch_bounds__(threads, MINB) is more conservative and spills.)
constexpr int pipe_maxreg(int threads, int minb)
{
    const int warps = minb * ((threads + 31) / 32);
    const int r = 16384 / (32 * ((warps + 3) / 4)) / 8 * 8;
    return r > 255 ? 255 : r;
}

template <int OP, int CONV, int VEC, int STAGES, int MINB>
__global__ void __launch_bounds__(CW / VEC + 32) __maxnreg__(pipe_maxreg(CW / VEC + 32, MINB))
k_step_pipe(const __grid_constant__ StepArgs a, unsigned total, unsigned tpr, unsigned rows)
{
    constexpr int kCons = CW / VEC;
    using StageT = Stage<conv_rows(CONV)>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    StageT *stages = reinterpret_cast<StageT *>(smem_raw);
    const Grid &g = a.g;
    const int tid = threadIdx.x;
    if (tid == 0)
    {
#pragma unroll
        for (int s = 0; s < STAGES; ++s)
        {
            mbar_init(&stages[s].full, 1);
            mbar_init(&stages[s].empty, kCons / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid >= kCons)
    {
        // ---- producer warp ------------------------------------------------------------------------
        // Static round-robin (tile blockIdx.x + k gridDim.x): at any moment the CTAs work on neighbouring
        // 2 KB chunks of every plane, which is what the DRAM likes -- a dynamic scheduler handing out 1, 2, 4
        // or 8 consecutive tiles per atomicAdd measured 4.9 / 5.7 / 4.3 / 4.5 TB/s against 5.9 TB/s for this.
        // Every 32 tiles all lanes look up the tile mask of this CTA's next 32 tiles at once (one L2 / DRAM
        // round trip per batch, off lane 0's critical path); lane 0 alone waits for a free stage and issues.
        const unsigned lane = (unsigned)(tid - kCons);
        unsigned it = 0, ubits = 0u;
        for (unsigned t = blockIdx.x; t < total; t += gridDim.x, ++it)
        {
            if ((it & 31u) == 0u)
            {
                const unsigned long long tj = (unsigned long long)t + (unsigned long long)lane * gridDim.x;
                const bool u = tj < total && tile_uniform(a, tile_of(a, (unsigned)tj, tpr, rows));
                ubits = __ballot_sync(0xffffffffu, u);
            }
            if (lane == 0)
            {
                StageT &st = stages[it % STAGES];
                if (it >= (unsigned)STAGES)
                    mbar_wait(&st.empty, ((it / STAGES) - 1u) & 1u);
                st.tile = t;
                issue_tile<CONV>(a, tile_of(a, t, tpr, rows), st, ((ubits >> (it & 31u)) & 1u) != 0u);
            }
            __syncwarp();
        }
        if (lane == 0)
        {
            // no more work: complete the next stage's `full` phase with a poison tile
            StageT &st = stages[it % STAGES];
            if (it >= (unsigned)STAGES)
                mbar_wait(&st.empty, ((it / STAGES) - 1u) & 1u);
            st.tile = kPoison;
            mbar_arrive(&st.full);
        }
        return;
    }
    // ---- consumer warps -------------------------------------------------------------------------------
    for (unsigned it = 0;; ++it)
    {
        StageT &st = stages[it % STAGES];
        mbar_wait(&st.full, (it / STAGES) & 1u);
        const unsigned t = st.tile;
        if (t == kPoison)
            break;
        const Tile tl = tile_of(a, t, tpr, rows);
        const int x0 = tl.x0 + VEC * tid, y = tl.y;
        const bool active = VEC * tid < tl.cw;
        // ---- shared memory -> registers (everything this thread needs from the stage) ---------------
        float tv[9][VEC], edge[9];
        float cur[3][VEC];
        uint32_t w[3] = {0u, 0u, 0u}, el[3] = {0u, 0u, 0u}, er[3] = {0u, 0u, 0u};
        const bool uniform = st.uniform != 0u;  // CTA-uniform
        // Group on the x symmetry axis: its negative-x neighbour is the cell itself and the three links that
        // cross the axis pull the x-mirrored population (propagation.cu:74-87, 135-143) -- which sits in the
        // same stage, same row, column 0 of the mirrored direction's buffer.  Handled here so that one group
        // per ROW does not fall to the generic path (19 us each; 4x the whole step on the scaled C4 grid).
        const bool axis_x = g.axisym == AXI_X && x0 == 0;
        if (active)
        {
#pragma unroll
            for (int q = 0; q < 9; ++q)
            {
                const float *row = &st.f[q][HALO + VEC * tid];
                VecIO<VEC>::load(row, tv[q]);
                const int ex = JRL_EX(q);
                edge[q] = ex > 0 ? (axis_x ? st.f[mirror_x(q)][HALO] : row[-1]) : (ex < 0 ? row[VEC] : 0.f);
            }
#pragma unroll
            for (int j = 0; j < conv_rows(CONV); ++j)
                VecIO<VEC>::load(&st.ex[j][VEC * tid], cur[j]);
            if (!uniform)
#pragma unroll
            for (int r = 0; r < 3; ++r)
            {
                const uint8_t *row = &st.fl[r][FHALO + VEC * tid];
                w[r] = VEC == 4 ? *reinterpret_cast<const uint32_t *>(row) : (uint32_t) * reinterpret_cast<const uint16_t *>(row);
                el[r] = row[-1];
                er[r] = row[VEC];
            }
        }
        // this warp is done with the stage: __syncwarp() orders the other lanes' reads before lane 0's arrive
        __syncwarp();
        if ((tid & 31) == 0)
            mbar_arrive(&st.empty);
        float acc[3] = {0.f, 0.f, 0.f}, cells = 0.f;
        if (active)
        {
            // ---- classify (rows: 0 = y-1, 1 = y, 2 = y+1) ---------------------------------------------
            bool bulk = !(g.axisym == AXI_Y && y == 0);
            const uint32_t fluid_word = VEC == 4 ? 0x01010101u : 0x0101u;
#pragma unroll
            for (int r = 0; r < 3; ++r)
                bulk = bulk && w[r] == fluid_word && (axis_x || el[r] == CT_FLUID) && er[r] == CT_FLUID;
            if (bulk || uniform)
            {
                float fc[VEC][9];
                uint8_t me[VEC];
#pragma unroll
                for (int q = 0; q < 9; ++q)
                {
                    const int ex = JRL_EX(q);
#pragma unroll
                    for (int k = 0; k < VEC; ++k)
                        fc[k][q] = ex == 0 ? tv[q][k] : (ex > 0 ? (k == 0 ? edge[q] : tv[q][k > 0 ? k - 1 : 0])
                                                                : (k == VEC - 1 ? edge[q] : tv[q][k < VEC - 1 ? k + 1 : 0]));
                }
#pragma unroll
                for (int k = 0; k < VEC; ++k)
                    me[k] = CT_FLUID;
                step_finish<VEC, OP, CONV>(a, g, tl.b, y, x0, cell(g, y, y, 0), fc, me, true, acc, CONV != 0 ? cur : nullptr);
                cells = (float)VEC;
            }
            else if (w[1] == (VEC == 4 ? 0x02020202u : 0x0202u) && y >= 1 && y <= g.H - 2 && x0 >= 1 && x0 + VEC <= g.W - 1)
            {
                // A group whose OWN cells are all OBSTACLE, away from the domain edges: a solid cell stores the
                // reflected incoming set, out[opp(q)] = f[q] @ upstream(q), whatever its neighbours are
                // (propagation.cu:111-128; lattice.cuh stream_links), keeps its moments and is not collided -- pure data
                // movement of exactly the values the bulk path pulls.  Without this the inside of a large body (the probe
                // tip of the scaled C4 grid: 0.5 M groups) goes to the work list -- and, where the list is full (one slab
                // of an 8-GPU split holds the whole tip), is walked in place at ~20 us per group.  Kept out of line and
                // out of the bulk branch: a runtime select inside the bulk loops cost the BGK headline 17 %.
                SolidPops<VEC> sp;
#pragma unroll
                for (int q = 0; q < 9; ++q)
                {
                    const int ex = JRL_EX(q);
#pragma unroll
                    for (int k = 0; k < VEC; ++k)
                        sp.v[k][opp(q)] = ex == 0 ? tv[q][k] : (ex > 0 ? (k == 0 ? edge[q] : tv[q][k > 0 ? k - 1 : 0])
                                                                       : (k == VEC - 1 ? edge[q] : tv[q][k < VEC - 1 ? k + 1 : 0]));
                }
                solid_group_store<VEC, CONV>(a, tl.b, y, x0, sp);
            }
            else
            {
                // A boundary group.  The generic rules are ~45 k instructions that the persistent CTAs would fetch
                // cold and walk serially (measured: ~20 us per group, 8 consumer warps hiding nothing -- the 16 k
                // boundary groups of the scaled C4 grid cost more than its 16.7 M bulk groups).  Defer it: append
                // (tile, thread) to the work list and let k_step_worklist run all of them, one cell per thread, hot
                // in the instruction cache and across the whole GPU, right after this kernel.  f_in is read-only
                // and every group writes only its own cells of f_out, so the order does not matter.
                bool deferred = false;
                if (a.worklist != nullptr)
                {
                    const unsigned slot = atomicAdd(a.worklist, 1u);
                    if (slot < a.worklist_cap)
                    {
                        a.worklist[2 + slot] = t * (unsigned)(CW / VEC) + (unsigned)tid;
                        deferred = true;
                    }
                }
                if (!deferred)
                {
                    const float4 r = step_group_generic<VEC, OP, CONV>(a, tl.b, y, x0);
                    acc[0] = r.x;
                    acc[1] = r.y;
                    acc[2] = r.z;
                    cells = r.w;
                }
            }
        }
        if (a.sums)
            add_sums(a.sums, acc, cells);
    }
}

// Second pass of the pipelined step: the boundary groups k_step_pipe deferred, one CELL per thread by the generic
// rules (the same device functions as everywhere else: bit-identical results), grid-stride over the list.
template <int OP, int CONV, int VEC>
__global__ void __launch_bounds__(256) k_step_worklist(const __grid_constant__ StepArgs a, unsigned tpr, unsigned rows)
{
    const unsigned n = min(a.worklist[0], a.worklist_cap);
    const unsigned long long cells_total = (unsigned long long)n * VEC;
    float acc[3] = {0.f, 0.f, 0.f}, cells = 0.f;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < cells_total;
         i += (unsigned long long)gridDim.x * blockDim.x)
    {
        const unsigned e = a.worklist[2 + i / VEC], k = (unsigned)(i % VEC);
        const unsigned t = e / (unsigned)(CW / VEC), tid = e % (unsigned)(CW / VEC);
        const Tile tl = tile_of(a, t, tpr, rows);
        const float4 r = step_group_generic<1, OP, CONV>(a, tl.b, tl.y, tl.x0 + VEC * (int)tid + (int)k);
        acc[0] += r.x;
        acc[1] += r.y;
        acc[2] += r.z;
        cells += r.w;
    }
    if (a.sums)
        add_sums(a.sums, acc, cells);
}

template <int OP, int CONV, int VEC, int STAGES, int MINB>
static int launch_cfg(const StepArgs &a, unsigned total, unsigned tpr, unsigned rows, cudaStream_t s)
{
    auto kern = k_step_pipe<OP, CONV, VEC, STAGES, MINB>;
    const size_t smem = sizeof(Stage<conv_rows(CONV)>) * STAGES;
    static unsigned long long configured = 0ull;  // per instantiation, one bit per device (a multi-GPU process)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 64 || !((configured >> dev) & 1ull))
    {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return JRL_ERR_CUDA;
        if (dev < 64)
            configured |= 1ull << dev;
    }
    // CTA b works on tiles b, b + nb, b + 2 nb, ...: with nb coprime to the tiles per row every CTA visits every
    // tile column equally often, so a boundary that costs extra in one column (a side wall, an inlet, the
    // symmetry axis) is shared by all CTAs instead of landing on the nb / gcd(nb, tpr) ones that own that column.
    unsigned nb = min(total, (unsigned)(MINB * sm_count()));
    auto gcd = [](unsigned x, unsigned y) { while (y) { const unsigned r = x % y; x = y; y = r; } return x; };
    while (nb > 1 && gcd(nb, tpr) != 1)
        --nb;
    const bool defer = a.worklist != nullptr && total < (1u << 23);  // (tile, thread) packs into 32 bits
    StepArgs b = a;
    if (!defer)
        b.worklist = nullptr;
    else if (cudaMemsetAsync(a.worklist, 0, sizeof(unsigned), s) != cudaSuccess)
        return JRL_ERR_CUDA;
    kern<<<nb, CW / VEC + 32, smem, s>>>(b, total, tpr, rows);
    if (defer)
        k_step_worklist<OP, CONV, VEC><<<4 * sm_count(), 256, 0, s>>>(b, tpr, rows);
    return 0;
}

// Measured on B200 at 8192^2 BGK (tools/exp_step.py): 4 cells/thread, 2 stages, 2 CTAs/SM = 4.95 TB/s;
// 5 stages with 1 CTA/SM (256 or 512 threads) = 4.1-4.3 TB/s (too few warps for the collision
// arithmetic); 2 cells/thread at 2 CTAs/SM spills (64-register cap) = 3.6 TB/s.
template <int OP, int CONV>
static int launch_one(const StepArgs &a, unsigned total, unsigned tpr, unsigned rows, cudaStream_t s)
{
#ifndef JRL_PIPE_VEC
#define JRL_PIPE_VEC 4
#endif
    return launch_cfg<OP, CONV, JRL_PIPE_VEC, JRL_PIPE_STAGES, JRL_PIPE_MINB>(a, total, tpr, rows, s);
}

static bool eligible(const StepArgs &a, int B)
{
    const Grid &g = a.g;
    if (g.W % 16 != 0 || g.W < CW)
        return false;
    if (((uintptr_t)a.f_in | (uintptr_t)a.flags | (uintptr_t)a.f_out) % 16 != 0)
        return false;
    if ((a.rho_out && (uintptr_t)a.rho_out % 16 != 0) || (a.vel_out && (uintptr_t)a.vel_out % 16 != 0))
        return false;
    if (a.cp.convection && (((uintptr_t)a.vel | (uintptr_t)a.mesh) % 16 != 0))  // prefetched as 128-bit vectors
        return false;
    const long long tiles = (long long)B * (a.row_end - a.row_begin) * ((g.W + CW - 1) / CW);
    return tiles < (1ll << 31);
}

static int launch(const StepArgs &a, int B, cudaStream_t s)
{
    const unsigned tpr = (unsigned)((a.g.W + CW - 1) / CW), rows = (unsigned)(a.row_end - a.row_begin);
    const unsigned total = (unsigned)B * rows * tpr;
#ifdef JRL_EXP_FAST_BUILD  // experiment builds (tools/): one operator only (JRL_EXP_FAST_OP, default BGK), compiles in a fraction of the time
    return launch_one<JRL_EXP_FAST_OP, 0>(a, total, tpr, rows, s);
#else
    // x-axis concentration form: three IEEE divisions and the per-direction relaxation per cell make it issue-bound
    // with 8 consumer warps per SM; 2 cells per consumer thread and 3 stages, as for MRT (scaled C4 at 8192^2:
    // 1.54 -> 1.32 ms; all-FLUID 1.39 -> 1.22 ms)
    if (a.cp.convection)
        return a.g.axisym == AXI_X ? launch_cfg<OP_BGK, 1, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s)
                                   : (a.g.axisym == AXI_Y ? launch_one<OP_BGK, 2>(a, total, tpr, rows, s)
                                                          : launch_one<OP_BGK, 3>(a, total, tpr, rows, s));
    if (a.cp.op == OP_BGK)
        return launch_one<OP_BGK, 0>(a, total, tpr, rows, s);
    if (a.cp.op == OP_MRT)  // 2 cells per consumer thread, 3 stages (see the dispatch note in step_once)
        return launch_cfg<OP_MRT, 0, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s);
    // KBC (force_vec = 8 only; the register kernel is the default): variant as a template value, 2 cells per consumer
    // thread and 3 stages as for MRT -- the operator is issue-bound (~800 instructions per cell)
    switch (kbc_index_of(a.cp.kbc_type))
    {
    case 1: return launch_cfg<OP_KBC_T0 + 1, 0, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s);
    case 2: return launch_cfg<OP_KBC_T0 + 2, 0, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s);
    case 3: return launch_cfg<OP_KBC_T0 + 3, 0, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s);
    case 4: return launch_cfg<OP_KBC_T0 + 4, 0, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s);
    default: return launch_cfg<OP_KBC_T0 + 0, 0, 2, 3, JRL_PIPE_MINB>(a, total, tpr, rows, s);
    }
#endif
}

}  // namespace pipe
#endif  // !JRL_HOST_EMU


// ---- tile mask: which row segments of JRL_PIPE_CW cells (plus one cell either side, periodic in x) are
//      all FLUID.  The pipelined kernel skips the flag rows of tiles whose three row segments are set:
//      the flags are 1.4 % of the bytes but cost 20 % of the step when every tile copies and tests them.
__global__ void k_tile_mask(const uint8_t *__restrict__ flags, uint8_t *__restrict__ mask, int W, int tpr, long long total)
{
    constexpr int TW = JRL_PIPE_CW;
#ifdef JRL_HOST_EMU
    const long long gw = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gw >= total)
        return;
    const long long R = gw / tpr;
    const int x0 = (int)(gw - R * tpr) * TW, cw = min(TW, W - x0);
    const uint8_t *row = flags + R * W;
    bool ok = row[x0 == 0 ? W - 1 : x0 - 1] == CT_FLUID && row[x0 + cw == W ? 0 : x0 + cw] == CT_FLUID;
    for (int i = 0; i < cw; ++i)
        ok = ok && row[x0 + i] == CT_FLUID;
    mask[gw] = ok ? 1 : 0;
#else
    // one warp per segment, 16 bytes per lane and load
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= total)
        return;
    const unsigned lane = threadIdx.x & 31u;
    const long long R = gw / tpr;
    const int x0 = (int)(gw - R * tpr) * TW, cw = min(TW, W - x0);
    const uint8_t *row = flags + R * W;
    bool ok = true;
    for (int i = (int)lane * 16; i < cw; i += 512)
    {
        const uint4 v = *reinterpret_cast<const uint4 *>(row + x0 + i);
        ok = ok && v.x == 0x01010101u && v.y == 0x01010101u && v.z == 0x01010101u && v.w == 0x01010101u;
    }
    if (lane == 0)
        ok = ok && row[x0 == 0 ? W - 1 : x0 - 1] == CT_FLUID;
    if (lane == 1)
        ok = ok && row[x0 + cw == W ? 0 : x0 + cw] == CT_FLUID;
    ok = __all_sync(0xffffffffu, ok);
    if (lane == 0)
        mask[gw] = ok ? 1 : 0;
#endif
}

template <int VEC>
static int launch_step(const StepArgs &a, int B, cudaStream_t s)
{
    const Grid &g = a.g;
    const int gpr = g.W / VEC;
    int tx = 32;
    while (tx < gpr && tx < kThreads)
        tx <<= 1;
    const int ty = kThreads / tx;
    const dim3 block(tx, ty, 1);
    const int rows = a.row_end - a.row_begin;
    const dim3 grid((gpr + tx - 1) / tx, (rows + ty - 1) / ty, (unsigned)B);
    if (grid.y > 65535u || grid.z > 65535u)
        return fail_arg("jrl_step_f32", "more than 65535 row blocks (or batch elements) in one launch: split it with row_begin / row_end");
    // latency-bound sizes: overlap this launch with the tail of the previous kernel in the stream (common.cuh)
    const bool pdl = (long long)B * rows * g.W <= (1ll << 19);
#define JRL_STEP(OPV, CONVV)                         \
    do                                               \
    {                                                \
        auto kern = k_step<VEC, OPV, CONVV>;         \
        if (pdl)                                     \
            JRL_LAUNCH_PDL(kern, grid, block, s, a); \
        else                                         \
            JRL_LAUNCH(kern, grid, block, s, a);     \
    } while (0)
#ifdef JRL_EXP_FAST_BUILD
    JRL_STEP(JRL_EXP_FAST_OP, 0);
    return 0;
#endif
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
    else if (a.cp.op == OP_NONE)
        JRL_STEP(OP_NONE, 0);
    else
        switch (kbc_index_of(a.cp.kbc_type))  // one instantiation per KBC variant (lattice.cuh OP_KBC_T0)
        {
        case 1: JRL_STEP(OP_KBC_T0 + 1, 0); break;
        case 2: JRL_STEP(OP_KBC_T0 + 2, 0); break;
        case 3: JRL_STEP(OP_KBC_T0 + 3, 0); break;
        case 4: JRL_STEP(OP_KBC_T0 + 4, 0); break;
        default: JRL_STEP(OP_KBC_T0 + 0, 0); break;
        }
#undef JRL_STEP
    return 0;
}

}  // namespace jrl

using namespace jrl;

static inline bool aligned(const void *p, size_t n) { return p == nullptr || ((uintptr_t)p % n) == 0; }


extern "C" int jrl_tile_mask_bytes(int B, int Hs, int W)
{
    if (B < 1 || Hs < 1 || W < 1)
        return 0;
    const long long n = (long long)B * Hs * ((W + JRL_PIPE_CW - 1) / JRL_PIPE_CW);
    return n > 0x7fffffffll ? 0 : (int)n;
}

extern "C" long long jrl_worklist_bytes(int B, int Hs, int W)
{
    if (B < 1 || Hs < 1 || W < 1)
        return 0;
    const long long groups = (long long)B * Hs * ((W + 3) / 4);
    return (long long)sizeof(unsigned) * (2 + std::max<long long>(4096, groups / 8));
}

extern "C" int jrl_tile_mask_u8(const uint8_t *flags, uint8_t *mask, int B, int Hs, int W, void *stream)
{
    const char *fn = "jrl_tile_mask_u8";
    JRL_REQUIRE(flags && mask, fn, "flags and mask required");
    JRL_REQUIRE(B >= 1 && Hs >= 1 && W >= 16 && W % 16 == 0, fn, "B, Hs >= 1 and W a multiple of 16 required");
    JRL_REQUIRE((uintptr_t)flags % 16 == 0, fn, "flags must be 16-byte aligned");
    const int tpr = (W + JRL_PIPE_CW - 1) / JRL_PIPE_CW;
    const long long total = (long long)B * Hs * tpr;
#ifdef JRL_HOST_EMU
    JRL_LAUNCH(k_tile_mask, blocks_for(total), kThreads, (cudaStream_t)stream, flags, mask, W, tpr, total);
#else
    JRL_LAUNCH(k_tile_mask, blocks_for(total * 32), kThreads, (cudaStream_t)stream, flags, mask, W, tpr, total);
#endif
    return check_launch(fn);
}

static int step_once(const jrl_step_desc *d, void *stream)
{
    const char *fn = "jrl_step_f32";
    JRL_REQUIRE(d, fn, "null descriptor");
    JRL_REQUIRE(d->flags && d->f_in && d->f_out, fn, "flags, f_in, f_out required");
    JRL_REQUIRE(d->f_in != d->f_out, fn, "f_out must not alias f_in (double-buffered pull scheme)");
    JRL_REQUIRE(dims_ok(d->B, d->H, d->W) && d->H >= 3 && d->W >= 3, fn, "B >= 1, H >= 3, W >= 3 required");
    JRL_REQUIRE(d->op >= 0 && d->op <= 3, fn, "unknown operator");
    JRL_REQUIRE(d->op != JRL_OP_NONE || d->force_vec != 8, fn, "JRL_OP_NONE has no pipelined variant");
    JRL_REQUIRE(d->op != JRL_OP_NONE || (!d->is_convection && d->rho_out && d->vel_out), fn,
                "JRL_OP_NONE (no collision) is the fluid form's first half: rho_out and vel_out required");
    JRL_REQUIRE(d->op != JRL_OP_MRT || d->mrt, fn, "MRT needs the 9x9 matrix");
    JRL_REQUIRE(!(d->is_convection && d->op != JRL_OP_BGK), fn, "convection is BGK only");
    JRL_REQUIRE(d->op != JRL_OP_KBC || d->kbc_type == JRL_KBC_LBGK || d->kbc_type == JRL_KBC_A || d->kbc_type == JRL_KBC_B ||
                    d->kbc_type == JRL_KBC_C || d->kbc_type == JRL_KBC_D,
                fn, "Please specific a right KBC Type value");
    JRL_REQUIRE(d->axisym >= 0 && d->axisym <= 2, fn, "axisym must be 0, 1 or 2");
    JRL_REQUIRE(d->axisym == 0 || d->is_convection || d->op == JRL_OP_NONE, fn,
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
    a.sums = d->sums;
    a.tile_mask = d->tile_mask;
    a.mask_tpr = (d->W + JRL_PIPE_CW - 1) / JRL_PIPE_CW;
    a.worklist = d->worklist_bytes >= 3 * (long long)sizeof(unsigned) ? (unsigned *)d->worklist : nullptr;
    a.worklist_cap = a.worklist ? (unsigned)std::min<long long>(d->worklist_bytes / (long long)sizeof(unsigned) - 2, 1ll << 30) : 0u;
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
#ifndef JRL_HOST_EMU
    // automatic choice: the pipelined kernel wins for BGK (memory-bound, 4 cells per consumer thread) and for MRT
    // with 2 cells per consumer thread (16 consumer warps per SM for the 81-FMA matvec: 4.7 TB/s at 8192^2 against
    // 4.1 in the register kernel and 3.95 with 4 cells per thread); KBC is bound by its instruction count (~940 per
    // cell, 19 IEEE divisions) and runs faster in the register kernel (1.6 vs 0.9-1.4 TB/s)
    // The concentration form takes the pipeline too when the caller provides the work list through which boundary
    // groups (tip block, inlet column, edge rows) are deferred to the second launch: scaled C4 grid 1.89 ms in the
    // register kernel, 3.54 ms in the pipeline with the boundary groups computed in place, 1.32 ms with them deferred.
    if (d->force_vec == 8 || (d->force_vec == 0 && vec == 4 && (d->op == JRL_OP_BGK || d->op == JRL_OP_MRT) &&
                                (!d->is_convection || a.worklist != nullptr)))
    {
        const bool ok = vec == 4 && pipe::eligible(a, d->B);
        JRL_REQUIRE(ok || d->force_vec == 0, fn, "pipelined kernel needs W % 16 == 0, W >= 512 and 16-byte aligned planes");
        if (ok)
        {
            const int prc = pipe::launch(a, d->B, s);
            if (prc != 0)
                return prc;
            if (d->use_contact)
            {
                const int ring = 2 * (d->H + d->W) - 4;
                JRL_LAUNCH(k_contact_ring, dim3(blocks_for(ring, 128), d->B), 128, s, a.g, d->rho, d->flags, d->rho_out, d->contact_tan);
            }
            return check_launch(fn);
        }
    }
#endif
    // Small grids are latency-bound, not bandwidth-bound: a group that touches a boundary walks its VEC cells
    // through the generic rules one after the other (dependent L2 round trips), so one cell per thread is
    // faster there (800 x 400 cylinder, MRT + moving wall: 27 us per step instead of 40)
    if (d->force_vec == 0 && (long long)d->B * (a.row_end - a.row_begin) * d->W <= (1ll << 19))
        vec = 1;
    if (d->force_vec == 1 || d->force_vec == 2 || d->force_vec == 4)
    {
        JRL_REQUIRE(d->force_vec <= vec, fn, "force_vec wider than extents / alignment allow");
        vec = d->force_vec;
    }
    const int rc = vec == 4 ? launch_step<4>(a, d->B, s) : (vec == 2 ? launch_step<2>(a, d->B, s) : launch_step<1>(a, d->B, s));
    if (rc != 0)
        return rc;
    if (d->use_contact)
    {
        const int ring = 2 * (d->H + d->W) - 4;
        JRL_LAUNCH(k_contact_ring, dim3(blocks_for(ring, 128), d->B), 128, s, a.g, d->rho, d->flags, d->rho_out, d->contact_tan);
    }
    return check_launch(fn);
}

extern "C" int jrl_step_f32(const jrl_step_desc *d, void *stream) { return step_once(d, stream); }

// ---- allocation-free variant of the pipelined step (no counterpart in the reference)
//      -----------------------------------------------------------------------------------------
// The pipelined kernel needs two pieces of scratch: the packed tile mask and the work list through which it defers
// its boundary groups (jrl_step_desc.tile_mask / .worklist, normally device buffers the caller allocates).  Callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) use this entry point,
// whose scratch is the file-scope array below: JRL_NOALLOC_SCRATCH_BYTES per GPU that loads the library, held until
// exit, SHARED by every call on that device (calls that overlap on the GPU race; the array keeps the last call's
// contents: the mask is fully rewritten and the work-list counter cleared by every call).
constexpr long long kNoAllocScratchBytes = JRL_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// Device address of g_noalloc_scratch on the current device.  In host code the array's NAME is only a host-side
// stand-in for the symbol, not a device address: it must go through cudaGetSymbolAddress, once per device, with that
// device current.
static int noalloc_scratch(unsigned char **out)
{
#ifdef JRL_HOST_EMU
    *out = g_noalloc_scratch;
    return JRL_OK;
#else
    static unsigned char *ptrs[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess)
        return JRL_ERR_CUDA;
    if (dev < 0 || dev >= 64)
        return fail_arg("jrl_step_noalloc_f32", "device ordinal beyond 63");
    if (ptrs[dev] == nullptr)
    {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
        {
            cudaGetLastError();
            return fail_arg("jrl_step_noalloc_f32", "cudaGetSymbolAddress(g_noalloc_scratch) failed");
        }
        ptrs[dev] = static_cast<unsigned char *>(p);
    }
    *out = ptrs[dev];
    return JRL_OK;
#endif
}

extern "C" long long jrl_step_noalloc_max_bytes(void) { return kNoAllocScratchBytes; }

extern "C" int jrl_step_noalloc_f32(const jrl_step_desc *d0, void *stream)
{
    const char *fn = "jrl_step_noalloc_f32";
    JRL_REQUIRE(d0, fn, "null descriptor");
    JRL_REQUIRE(d0->tile_mask == nullptr && d0->worklist == nullptr, fn,
                "tile_mask and worklist must be NULL: this entry point supplies them from its static scratch");
    JRL_REQUIRE(dims_ok(d0->B, d0->H, d0->W), fn, "B >= 1, H, W >= 2 required");
    const int Hs = d0->slab_rows > 0 ? d0->slab_rows + 2 : d0->H;
    // scratch layout: [0, mask_bytes) tile mask, then (256-byte aligned) the work list of jrl_worklist_bytes()
    const long long mask_bytes = ((long long)jrl_tile_mask_bytes(d0->B, Hs, d0->W) + 255) / 256 * 256;
    const long long wl_bytes = jrl_worklist_bytes(d0->B, Hs, d0->W);
    JRL_REQUIRE(mask_bytes > 0 && wl_bytes > 0 && mask_bytes + wl_bytes <= kNoAllocScratchBytes, fn,
                "grid too large for the static scratch (tile mask + work list exceed JRL_NOALLOC_SCRATCH_BYTES = 4 MiB, "
                "about 4096 x 4096 cells): use jrl_step_f32 with caller-provided scratch");
    jrl_step_desc d = *d0;
    const bool masked = d.flags != nullptr && d.W % 16 == 0 && d.W >= JRL_PIPE_CW && (uintptr_t)d.flags % 16 == 0;
    if (masked)  // otherwise the step takes the register kernel, which uses no scratch
    {
        unsigned char *base = nullptr;
        const int rc = noalloc_scratch(&base);
        if (rc != JRL_OK)
            return rc;
        const int mrc = jrl_tile_mask_u8(d.flags, base, d.B, Hs, d.W, stream);
        if (mrc != JRL_OK)
            return mrc;
        d.tile_mask = base;
        d.worklist = base + mask_bytes;
        d.worklist_bytes = wl_bytes;
    }
    return step_once(&d, stream);
}

// Multi-step driver: the time loop of the reference's demos (demo_2d_LBM.py:101-119 and siblings) moved
// below the C ABI, so that small, launch-bound grids are not paced by the Python interpreter.
extern "C" int jrl_run_f32(const jrl_step_desc *d0, int nsteps, int store_every, long long sums_stride, void *stream)
{
    const char *fn = "jrl_run_f32";
    JRL_REQUIRE(d0, fn, "null descriptor");
    JRL_REQUIRE(nsteps >= 0, fn, "nsteps must be >= 0");
    JRL_REQUIRE(sums_stride >= 0, fn, "sums_stride must be >= 0");
    jrl_step_desc d = *d0;
    // previous-step moments are state (OBSTACLE cells keep them, INFLOW cells and the moving wall read
    // them) whenever both buffers of a pair are given: they ping-pong with the populations
    const bool pp_rho = d.rho != nullptr && d.rho_out != nullptr;
    const bool pp_vel = !d.is_convection && d.vel != nullptr && d.vel_out != nullptr;
    float *const rho_last = d.rho_out, *const vel_last = d.vel_out;
    int first = 0;
#ifndef JRL_HOST_EMU
    // Small grids are paced by the host's launch rate (2.6 us per cudaLaunchKernelEx at 130^2, where the kernel itself
    // takes ~1.5 us): replay the loop from a CUDA graph.  The step sequence has period 2 (ping-pong buffers), so
    // kGraphSteps (even) steps are captured once on a private stream and the executable graph is launched on the
    // caller's stream as many times as it fits; the remainder -- at least the last step, which stores the moments --
    // runs as plain launches.  Any failure of the capture machinery falls back to plain launches for everything.
    constexpr int kGraphSteps = 256;
    const char *genv = getenv("JRL_RUN_GRAPH");  // 0: never replay; 1: captured launches keep PDL; default 2: plain nodes
    const int gmode = genv ? atoi(genv) : 2;
    const bool periodic = gmode != 0 && (pp_rho || pp_vel || store_every == 0) && (d.sums == nullptr || sums_stride == 0);
    if (periodic && nsteps >= 4 * kGraphSteps && (long long)d.B * d.H * d.W <= (1ll << 19) && d.slab_rows == 0)
    {
        cudaStream_t cs = nullptr;
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        int launched = 0;
        bool ok = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
        if (ok)
        {
            jrl_step_desc c = d;
            pdl_disabled() = gmode != 1;
            for (int i = 0; i < kGraphSteps && ok; ++i)
            {
                if (!pp_rho && !pp_vel)
                    c.rho_out = nullptr, c.vel_out = nullptr;
                ok = step_once(&c, cs) == 0;
                const float *fi = c.f_in;
                c.f_in = c.f_out;
                c.f_out = const_cast<float *>(fi);
                if (pp_rho)
                    std::swap(c.rho, c.rho_out);
                if (pp_vel)
                    std::swap(c.vel, c.vel_out);
            }
            pdl_disabled() = false;
            ok = (cudaStreamEndCapture(cs, &graph) == cudaSuccess) && ok && graph != nullptr;
        }
        if (ok)
            ok = cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess;
        if (ok)
        {
            const int reps = (nsteps - 1) / kGraphSteps;  // leaves 1 .. kGraphSteps steps for the plain loop
            for (int r = 0; r < reps && ok; ++r)
            {
                ok = cudaGraphLaunch(exec, (cudaStream_t)stream) == cudaSuccess;
                launched += ok ? 1 : 0;
            }
            if (ok)
                first = reps * kGraphSteps;  // an even count: the buffers are back in their starting roles
        }
        if (exec)
            cudaGraphExecDestroy(exec);
        if (graph)
            cudaGraphDestroy(graph);
        if (cs)
            cudaStreamDestroy(cs);
        if (!ok)
        {
            cudaGetLastError();  // clear the error; fall back to plain launches unless part of the run is already enqueued
            JRL_REQUIRE(launched == 0, fn, "CUDA graph replay failed after some of its launches");
        }
    }
#endif
    for (int i = first; i < nsteps; ++i)
    {
        if (!pp_rho && !pp_vel)
        {
            // moments are pure outputs: store them every `store_every`-th step and at the last step
            const bool store = i == nsteps - 1 || (store_every > 0 && (i + 1) % store_every == 0);
            d.rho_out = store ? rho_last : nullptr;
            d.vel_out = store ? vel_last : nullptr;
        }
        const int rc = step_once(&d, stream);
        if (rc != 0)
            return rc;
        const float *fi = d.f_in;
        d.f_in = d.f_out;
        d.f_out = const_cast<float *>(fi);
        if (pp_rho)
        {
            float *t = d.rho;
            d.rho = d.rho_out;
            d.rho_out = t;
        }
        if (pp_vel)
        {
            float *t = d.vel;
            d.vel = d.vel_out;
            d.vel_out = t;
        }
        if (d.sums)
            d.sums += sums_stride;
    }
    return JRL_OK;
}

// Stage-B walk loop of the JRL concentration demos below the C ABI (SURVEY 8f ranks 1 + 2):
// per step  tip motion (1 launch) -> [tile mask] -> fused concentration step (1 launch) -> the two strip
// fluxes the demo records (2 launches) -> walks += vel_obs_conv dt, with every host scalar evaluated as the
// demo's python evaluates it (demo_2d_LBM_JRL_C_cylinder.py:281-362).
extern "C" int jrl_conc_walk_f32(jrl_walk_desc *w, int nsteps, void *stream)
{
    const char *fn = "jrl_conc_walk_f32";
    JRL_REQUIRE(w, fn, "null descriptor");
    JRL_REQUIRE(nsteps >= 0, fn, "nsteps must be >= 0");
    JRL_REQUIRE(w->vel_snaps && w->n_snaps >= 1 && w->x_interval > 0.0, fn, "velocity snapshots required");
    JRL_REQUIRE(w->records, fn, "records required");
    jrl_step_desc d = w->step;
    jrl_tip_desc t = w->tip;
    JRL_REQUIRE(d.is_convection && d.B == 1 && d.rho && d.rho_out && d.vel && d.phi_obs && d.mesh_grid, fn,
                "step descriptor: concentration form, B = 1, C ping-pong pair, C_vel, phi_obs and mesh_grid buffers required");
    JRL_REQUIRE(d.slab_rows == 0, fn, "single-domain only");
    const long long HW = (long long)d.H * d.W;
    double walks = w->walks, conv_timer = w->conv_timer;
    const bool demo_clock = w->dt_conv_real != 0.0;
    w->steps_done = 0;
    for (int i = 0; i < nsteps; ++i)
    {
        // -- host scalars, as the demo computes them (:283-284, :294-303)
        const int walks_int = (int)(walks + 0.5 * w->dx);
        JRL_REQUIRE(walks_int >= 1 && walks_int + 1 < d.H, fn, "the tip left the grid");
        t.walks_int = walks_int;
        t.box_cy = (float)(walks - w->box_radius);
        double x_comp;
        int b;
        if (demo_clock)
        {
            const double x_real = conv_timer * w->vel_obs_real;          // :295
            x_comp = x_real / w->radius_obs_real * w->radius_obs;        // :296
            b = (int)(x_comp / w->x_interval);                           // :297, not clamped: the demo raises
            if (b < 0 || b > w->n_snaps - 1)
            {
                w->walks = walks;
                w->conv_timer = conv_timer;
                JRL_REQUIRE(false, fn, "velocity snapshot index out of range (the demo raises IndexError here)");
            }
        }
        else
        {
            x_comp = walks - w->inflow_height;
            b = (int)(x_comp / w->x_interval);
            if (b > w->n_snaps - 1)
                b = w->n_snaps - 1;
            if (b < 0)
                b = 0;
        }
        t.vel0 = w->vel_snaps + (long long)b * 2 * HW;
        if (b + 1 < w->n_snaps)
        {
            const double tw = x_comp / w->x_interval - b;
            t.vel1 = t.vel0 + 2 * HW;
            t.w0 = (float)(1.0 - tw);
            t.w1 = (float)tw;
        }
        else
        {
            t.vel1 = nullptr;
            t.w0 = 1.f;
            t.w1 = 0.f;
        }
        t.flags = const_cast<uint8_t *>(d.flags);
        t.C = d.rho;
        t.phi_obs = const_cast<float *>(d.phi_obs);
        t.C_vel = d.vel;
        int rc = jrl_tip_prepare_conc_f32(&t, stream);
        if (rc != 0)
            return rc;
        if (d.tile_mask)
        {
            rc = jrl_tile_mask_u8(d.flags, const_cast<uint8_t *>(d.tile_mask), 1, d.H, d.W, stream);
            if (rc != 0)
                return rc;
        }
        rc = step_once(&d, stream);
        if (rc != 0)
            return rc;
        // -- records on the new C (:344-362): inflow strip at walks_int, current strip at walks_int + 1
        float *rec = w->records + 4ll * i;
        rc = strip_flux_n(d.rho_out, d.mesh_grid, walks_int, 2, t.radius_obs_int, rec, d.H, d.W, stream);  // one launch, two strips
        if (rc != 0)
            return rc;
        const float *fi = d.f_in;
        d.f_in = d.f_out;
        d.f_out = const_cast<float *>(fi);
        float *c = d.rho;
        d.rho = d.rho_out;
        d.rho_out = c;
        walks += w->vel_obs_conv * w->dt;
        conv_timer += w->dt_conv_real;                                   // :342
        w->steps_done = i + 1;
    }
    w->walks = walks;
    w->conv_timer = conv_timer;
    return JRL_OK;
}
