// D2Q9 lattice constants, neighbour rules and per-cell arithmetic shared by every kernel of
// libjrl_lbm.so (sm_100a only).
//
// Everything here is `__device__ __forceinline__` and works on registers; the kernels in
// ops.cu (one reference operator per launch) and fused.cu (stream + boundaries + moments +
// collide in one pass) are built from these pieces, so the fused path and the per-operator
// path compute bit-identical results.
//
// Arithmetic contract: fp32 storage; every expression keeps the operation ORDER of the
// reference's path AS IT RUNS ON CUDA (cited per function, paths relative to the reference tree):
// fp64 intermediates exactly where the reference's C++ literals force them; torch's CUDA
// semantics for the torch-op parts (division by a python scalar = multiplication by the fp32
// reciprocal; `sum(dim=1)` = four interleaved accumulators; the batched 9x9 mat-vec = two FMA
// chains) -- each measured against the reference-as-run on a B200 (tools/bench_ref_cuda.py).
// The library is built with -fmad=false, so the only fused multiply-adds are the ones written
// as fmaf() here.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace jrl {

// utils/types.py:5-15 -- compared with ==, never as bit sets.
enum : uint8_t { CT_NOTHING = 0, CT_FLUID = 1, CT_OBSTACLE = 2, CT_EMPTY = 4, CT_INFLOW = 8, CT_OUTFLOW = 16, CT_INFLOW_2 = 32 };
enum : int { AXI_NOT = 0, AXI_X = 1, AXI_Y = 2 };
// utils/types.py:18-23
enum : int { KBC_LBGK = 0, KBC_A = 0x85, KBC_B = 0x86, KBC_C = 0x89, KBC_D = 0x8A };
enum : int { OP_BGK = 0, OP_MRT = 1, OP_KBC = 2, OP_NONE = 3 };
// Template values OP_KBC_T0 + i select KBC with the variant known at compile time (i: 0 LBGK, 1 A, 2 B, 3 C, 4 D), which
// removes the run-time select chains and the dead central-moment code of the other variants from the hot kernels
// (~100 instructions per cell of ~940); OP_KBC itself keeps the run-time dispatch on p.kbc_type.
enum : int { OP_KBC_T0 = 16 };
__host__ __device__ constexpr int kbc_code_of(int i) { return i == 0 ? KBC_LBGK : (i == 1 ? KBC_A : (i == 2 ? KBC_B : (i == 3 ? KBC_C : KBC_D))); }
inline int kbc_index_of(int code) { return code == KBC_A ? 1 : (code == KBC_B ? 2 : (code == KBC_C ? 3 : (code == KBC_D ? 4 : 0))); }

// propagation.cu:9-20, 49-64: rest, +x, +y, -x, -y, ++, -+, --, +-
#define JRL_EX(q) ((q) == 1 || (q) == 5 || (q) == 8 ? 1 : ((q) == 3 || (q) == 6 || (q) == 7 ? -1 : 0))
#define JRL_EY(q) ((q) == 2 || (q) == 5 || (q) == 6 ? 1 : ((q) == 4 || (q) == 7 || (q) == 8 ? -1 : 0))

__device__ __forceinline__ constexpr int opp(int q)
{
    return q == 0 ? 0 : (q <= 4 ? ((q + 1) % 4) + 1 : ((q - 3) % 4) + 5);  // {0,3,4,1,2,7,8,5,6}
}
__device__ __forceinline__ constexpr int mirror_x(int q)
{
    // {0,3,2,1,4,6,5,8,7}
    return q == 1 ? 3 : q == 3 ? 1 : q == 5 ? 6 : q == 6 ? 5 : q == 7 ? 8 : q == 8 ? 7 : q;
}
__device__ __forceinline__ constexpr int mirror_y(int q)
{
    // {0,1,4,3,2,8,7,6,5}
    return q == 2 ? 4 : q == 4 ? 2 : q == 5 ? 8 : q == 8 ? 5 : q == 6 ? 7 : q == 7 ? 6 : q;
}
__device__ __forceinline__ constexpr float weight(int q)
{
    return q == 0 ? (float)(4.0 / 9.0) : (q <= 4 ? (float)(1.0 / 9.0) : (float)(1.0 / 36.0));
}

// One batch element's view of the lattice.  Plane stride is 64-bit: 9*H*W exceeds 2^31 at
// 32768^2 (the reference's int indexing overflows there, propagation.cu:38, 66-72).
//
// Row-slab storage (multi-GPU): a rank stores global rows [y_off, y_off + Hloc) of every plane as
// local rows 1..Hloc of a [Hloc + 2, W] plane; local rows 0 and Hloc + 1 are ghost rows filled by
// the halo exchange.  H, W stay the GLOBAL extents, so every neighbour rule (periodic wrap of
// FLUID cells, mirror / clamp at the domain edge) is evaluated in global coordinates, and
// cell() maps the result -- always within one row of the current cell, modulo H -- to storage.
struct Grid
{
    int H, W;
    int axisym;
    long long HW;  // plane stride in elements: H*W, or (Hloc+2)*W for a slab
    int slab;      // 0: whole grid stored
    int y_off;     // slab: global row of local row 1
};

// Storage index of cell (sy, sx), a neighbour (or self) of the cell in global row y.
__device__ __forceinline__ long long cell(const Grid &g, int y, int sy, int sx)
{
    int row = sy;
    if (g.slab)
    {
        int d = sy - y;
        if (d > 1)
            d -= g.H;
        else if (d < -1)
            d += g.H;
        row = y - g.y_off + 1 + d;
    }
    return (long long)row * g.W + sx;
}

// Neighbour coordinates of one cell.
//   mirror = true : streaming rule, propagation.cu:74-87 (FLUID wraps, others mirror to 1 / W-2)
//   mirror = false: stencil / wall rule, collision_2d.cu:59-72, solid_fluid_coupling.cu:98-112
struct Nbr
{
    int xn, xp, yn, yp;
    bool on_x, on_y;
};

template <bool MIRROR>
__device__ __forceinline__ Nbr neighbours(const Grid &g, uint8_t me, int y, int x)
{
    Nbr n;
    const bool fluid = me == CT_FLUID;
    n.on_x = (x == 0) && g.axisym == AXI_X;
    n.on_y = (y == 0) && g.axisym == AXI_Y;
    int xn_o, xp_o, yn_o, yp_o;
    if (MIRROR)
    {
        xn_o = x - 1 < 0 ? 1 : x - 1;
        xp_o = x + 1 >= g.W ? g.W - 2 : x + 1;
        yn_o = y - 1 < 0 ? 1 : y - 1;
        yp_o = y + 1 >= g.H ? g.H - 2 : y + 1;
    }
    else
    {
        xn_o = max(0, x - 1);
        xp_o = min(g.W - 1, x + 1);
        yn_o = max(0, y - 1);
        yp_o = min(g.H - 1, y + 1);
    }
    n.xn = n.on_x ? 0 : (fluid ? (x == 0 ? g.W - 1 : x - 1) : xn_o);
    n.xp = fluid ? (x + 1 == g.W ? 0 : x + 1) : xp_o;
    n.yn = n.on_y ? 0 : (fluid ? (y == 0 ? g.H - 1 : y - 1) : yn_o);
    n.yp = fluid ? (y + 1 == g.H ? 0 : y + 1) : yp_o;
    return n;
}

// Upstream ("income") and downstream ("outcome") cell of direction q, propagation.cu:88-108.
__device__ __forceinline__ long long income(const Grid &g, const Nbr &n, int y, int x, int q)
{
    const int ex = JRL_EX(q), ey = JRL_EY(q);
    const int sx = ex > 0 ? n.xn : (ex < 0 ? n.xp : x);
    const int sy = ey > 0 ? n.yn : (ey < 0 ? n.yp : y);
    return cell(g, y, sy, sx);
}
__device__ __forceinline__ long long outcome(const Grid &g, const Nbr &n, int y, int x, int q)
{
    const int ex = JRL_EX(q), ey = JRL_EY(q);
    const int sx = ex > 0 ? n.xp : (ex < 0 ? n.xn : x);
    const int sy = ey > 0 ? n.yp : (ey < 0 ? n.yn : y);
    return cell(g, y, sy, sx);
}

// ------------------------------------------------------------------------------------------
// R1: pull streaming of one cell with every link rule except the INFLOW / INFLOW_2 / OUTFLOW
// overwrites (propagation.cu:110-193).  f, fl, phi are batch-element bases.
// ------------------------------------------------------------------------------------------
template <bool HAS_PHI>
__device__ __forceinline__ void stream_links(const Grid &g, const float *__restrict__ f,
                                             const uint8_t *__restrict__ fl, const float *__restrict__ phi,
                                             int y, int x, uint8_t me, float out[9])
{
    const Nbr n = neighbours<true>(g, me, y, x);
    const long long self = cell(g, y, y, x);
    out[0] = f[self];
    const float t = HAS_PHI ? phi[self] : 0.f;
    // All eight upstream populations and flags are fetched before any link rule is evaluated: the loads are
    // independent (valid for every cell), so they overlap instead of forming a chain of sixteen dependent
    // round trips behind the rules' branches (19 us per boundary group measured on B200 before this).
    float pulled_q[9], self_q[9];
    uint8_t up_q[9];
#pragma unroll
    for (int q = 1; q < 9; ++q)
    {
        int qi = q;  // income_Q_idx: axis-crossing links read the mirrored population
        if (n.on_x && (q == 5 || q == 8 || q == 1))
            qi = mirror_x(q);
        else if (n.on_y && (q == 5 || q == 6 || q == 2))
            qi = mirror_y(q);
        const long long src = income(g, n, y, x, q);
        pulled_q[q] = f[qi * g.HW + src];
        up_q[q] = fl[src];
        self_q[q] = f[q * g.HW + self];  // f_this of the opposite link, should its upstream cell be an obstacle
    }
#pragma unroll
    for (int q = 1; q < 9; ++q)
    {
        const float pulled = pulled_q[q];
        if (me == CT_OBSTACLE)
        {
            out[opp(q)] = pulled;  // solid cells store the reflected incoming set (:111-128)
            continue;
        }
        if (up_q[q] != CT_OBSTACLE)
        {
            out[q] = pulled;
            continue;
        }
        const float f_this = self_q[opp(q)];
        float v = f_this;  // halfway bounce-back (:155-159)
        if (HAS_PHI)
        {
            if (t >= 0.5f && t <= 1.f)
            {
                const float t1 = 2.f * t - 1.f;
                v = t1 * pulled + (1.f - t1) * f_this;  // :163-169
            }
            else if (t <= 0.5f && t >= 0.f)
            {
                int qo = q;
                if (n.on_x && (q == 6 || q == 7 || q == 3))
                    qo = mirror_x(q);
                else if (n.on_y && (q == 7 || q == 8 || q == 4))
                    qo = mirror_y(q);
                const float t2 = 2.f * t;
                const float f_next = f[opp(qo) * g.HW + outcome(g, n, y, x, q)];
                v = t2 * f_this + (1.f - t2) * f_next;  // :170-184
            }
        }
        out[q] = v;
    }
}

// Equilibrium of the INFLOW / INFLOW_2 reset with the reference kernel's mixed precision
// (propagation.cu:340-350): cs2, uv, eu fp32; `1.0 + eu / cs2` and the second-order term fp64;
// each of the two statements rounds to fp32 on store.
__device__ __forceinline__ void inflow_equilibrium(float rho_t, float ux, float uy, float out[9])
{
    const float cs2 = (float)(1.0 / 3.0);
    const float uv = ux * ux + uy * uy;
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        const float eu = ux * (float)JRL_EX(q) + uy * (float)JRL_EY(q);
        const float rw = rho_t * weight(q);
        const float first = (float)((double)rw * (1.0 + (double)(eu / cs2)));
        const double second = (double)rw * (0.5 * (double)eu * (double)eu / (double)cs2 / (double)cs2 - 0.5 * (double)uv / (double)cs2);
        out[q] = (float)((double)first + second);
    }
}

// Inflow description of one streaming call (LBM_propagation_2d.py:52-62).
struct Inflow
{
    float *rho;               // [B,1,H,W] or null; written at INFLOW (if density >= 0) and INFLOW_2
    float *vel;               // [B,2,H,W] or null; written at INFLOW when inflow_vel is given
    const float *inflow_vel;  // device [2] or null
    float density;            // < 0: none
    int has_density;
};

// Streaming of one cell followed by its INFLOW / INFLOW_2 reset (propagation.cu:295-377).
// `rho_b`, `vel_b` are this batch element's bases inside in.rho / in.vel (or null).
// WRITE = false evaluates the same values without the rho/vel side effects (used when an
// OUTFLOW cell re-evaluates its inward neighbour).
template <bool HAS_PHI, bool WRITE>
__device__ __forceinline__ void stream_with_inflow(const Grid &g, const float *__restrict__ f,
                                                   const uint8_t *__restrict__ fl, const float *__restrict__ phi,
                                                   const Inflow &in, float *rho_b, float *vel_b, int y, int x,
                                                   float out[9])
{
    const long long self = cell(g, y, y, x);
    const uint8_t me = fl[self];
    if (me == CT_INFLOW)
    {
        float rho_t = 0.f;
        if (rho_b != nullptr && in.has_density)
        {
            rho_t = in.density;
            if (WRITE)
                rho_b[self] = rho_t;
        }
        else if (rho_b != nullptr)
            rho_t = rho_b[self];
        else if (in.has_density)
            rho_t = in.density;
        float ux = 0.f, uy = 0.f;
        if (vel_b != nullptr && in.inflow_vel != nullptr)
        {
            ux = in.inflow_vel[0];
            uy = in.inflow_vel[1];
            if (WRITE)
            {
                vel_b[self] = ux;
                vel_b[g.HW + self] = uy;
            }
        }
        else if (vel_b != nullptr)
        {
            ux = vel_b[self];
            uy = vel_b[g.HW + self];
        }
        else if (in.inflow_vel != nullptr)
        {
            ux = in.inflow_vel[0];
            uy = in.inflow_vel[1];
        }
        inflow_equilibrium(rho_t, ux, uy, out);
        return;
    }
    if (me == CT_INFLOW_2)
    {
        const float ux = vel_b == nullptr ? 0.f : vel_b[self];
        const float uy = vel_b == nullptr ? 0.f : vel_b[g.HW + self];
        if (WRITE && rho_b != nullptr)
            rho_b[self] = 1.f;
        inflow_equilibrium(1.f, ux, uy, out);
        return;
    }
    stream_links<HAS_PHI>(g, f, fl, phi, y, x, me, out);
}

// Complete streaming of one cell including the OUTFLOW copy (propagation.cu:195-255).  The
// reference reads the neighbour's f_new while another thread writes it; here, as in the
// oracle, the value is defined as the neighbour's populations after streaming and inflow
// reset.  (The neighbour's own OUTFLOW copy, if it is one, is not chained.)
template <bool HAS_PHI, bool WRITE = true>
__device__ __forceinline__ void stream_cell(const Grid &g, const float *__restrict__ f,
                                            const uint8_t *__restrict__ fl, const float *__restrict__ phi,
                                            const Inflow &in, float *rho_b, float *vel_b, int y, int x,
                                            float out[9])
{
    const uint8_t me = fl[cell(g, y, y, x)];
    if (me == CT_OUTFLOW)
    {
        const int sy = y == 0 ? y + 1 : (y == g.H - 1 ? y - 1 : y);
        const int sx = x == 0 ? x + 1 : (x == g.W - 1 ? x - 1 : x);
        if (sy != y || sx != x)
        {
            stream_with_inflow<HAS_PHI, false>(g, f, fl, phi, in, rho_b, vel_b, sy, sx, out);
            return;
        }
    }
    stream_with_inflow<HAS_PHI, WRITE>(g, f, fl, phi, in, rho_b, vel_b, y, x, out);
}

// ------------------------------------------------------------------------------------------
// R2 / R7: moments (LBM_macro_compute_2d.py:316-351, 273-295).
// The reference evaluates `f.sum(dim=1)` and `(f.unsqueeze(2) * e).sum(1)` with torch on CUDA (it hard-requires
// CUDA, readme.md:7).  ATen's reduction kernel gives every output its own thread there (9 inputs per output, far
// below the split threshold) and that thread keeps vt0 = 4 interleaved accumulators, element k into accumulator
// k % 4, combined in order at the end (ATen/native/cuda/Reduce.cuh, thread_reduce_impl):
//     ((((a0 + a4) + a8) + (a1 + a5)) + (a2 + a6)) + (a3 + a7)
// Established on a B200 against the reference-as-run (tools/bench_ref_cuda.py, tools/exp_torch_order.py: 0 differing
// bits in rho, u, C over every golden case) and pinned by tests/golden/refcuda_*.npz.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float torch_sum9(float a0, float a1, float a2, float a3, float a4, float a5, float a6,
                                            float a7, float a8)
{
    return ((((a0 + a4) + a8) + (a1 + a5)) + (a2 + a6)) + (a3 + a7);
}

__device__ __forceinline__ float sum9(const float f[9])
{
    return torch_sum9(f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8]);
}

__device__ __forceinline__ void momentum9(const float f[9], float &mx, float &my)
{
    // sum_q f_q e_q in the same accumulator order; e is 0 or +-1, so the products are exact and the zero terms
    // (which only add +-0 to an accumulator) are dropped:
    //   x: acc0 = f8, acc1 = f1 + f5, acc2 = -f6, acc3 = -(f3 + f7)
    //   y: acc0 = -(f4 + f8), acc1 = f5, acc2 = f2 + f6, acc3 = -f7
    mx = ((f[8] + (f[1] + f[5])) - f[6]) - (f[3] + f[7]);
    my = ((f[5] - (f[4] + f[8])) + (f[2] + f[6])) - f[7];
}

// ------------------------------------------------------------------------------------------
// R3: equilibrium (LBM_collision_2d.py:236-263).  Host-prepared fp32 scalars in EqParams.
// ------------------------------------------------------------------------------------------
struct EqParams
{
    float c;        // (float)(dx/dt)
    float inv_cs2;  // (float)(1/cs2): torch divides by a python scalar as a reciprocal multiply
};

// a / b with IEEE bits, keeping EXACT-ZERO numerators off the division's slow path.  nvcc's fp32 division is a
// reciprocal + Newton sequence guarded by FCHK, which sends zero / denormal / huge operands to a ~60-instruction
// subroutine (CALL); with a zero numerator in one lane the whole warp waits for it.  Lattice states are full of exact
// zeros (a force field of zeros, u = 0 during the concentration balance phase, f - feq = 0 in the saturated parts
// of a shear layer: ncu showed 7 CALLs per cell in the KBC kernel, profiles/r1_kstep_vec4_kbc_details.txt), and
// the quotient is then known: (+-0) / b = +-0 with the XOR of the signs for every b except 0 and NaN, which give NaN.
// The division itself is fed a harmless numerator in those lanes and its result discarded, so the bits never change.
__device__ __forceinline__ float div_ieee(float a, float b)
{
    const bool az = a == 0.f;
    const float q = (az ? 1.f : a) / b;
    const float z = __int_as_float(__float_as_int(a) ^ (__float_as_int(b) & (int)0x80000000));  // +-0, sign(a) ^ sign(b)
    const bool nan = (b != b) | (b == 0.f);
    return az ? (nan ? __int_as_float(0x7fffffff) : z) : q;
}

__device__ __forceinline__ void half_force_shift(float rho, float &ux, float &uy, float fx, float fy)
{
    // LBM_collision_2d.py:246-250: u += where(rho <= 1e-10, 0, 0.5 * force / rho)
    const float ax = rho <= 1e-10f ? 0.f : div_ieee(0.5f * fx, rho);
    const float ay = rho <= 1e-10f ? 0.f : div_ieee(0.5f * fy, rho);
    ux = ux + ax;
    uy = uy + ay;
}

__device__ __forceinline__ void equilibrium(const EqParams &p, float rho, float ux, float uy, float feq[9])
{
    // Reference order (LBM_collision_2d.py:252-261), per direction q:
    //   eu = (ux e_qx) c + (uy e_qy) c ;  feq = rho w (1 + eu/cs2) + rho w (0.5 eu eu /cs2 /cs2 - 0.5 uv /cs2)
    // evaluated here once per OPPOSITE PAIR.  IEEE arithmetic is sign-symmetric, so with e in {-1,0,1}:
    // (u * -1) * c == -(u * c), eu(opp q) == -eu(q), the quadratic term is identical for both members
    // of a pair, and a zero term only affects the sign of a zero eu, which neither 1 + eu/cs2 nor
    // eu * eu can see.  Finite inputs give bit-identical results to the per-direction form with
    // roughly half the operations.
    const float uv = ux * ux + uy * uy;
    const float huv = (0.5f * uv) * p.inv_cs2;
    const float uxc = ux * p.c, uyc = uy * p.c;
    const float rw0 = rho * weight(0), rw1 = rho * weight(1), rw5 = rho * weight(5);
    feq[0] = rw0 * 1.f + rw0 * (0.f - huv);
#define JRL_FEQ_PAIR(q, qo, eu_expr, rw)                                            \
    do                                                                              \
    {                                                                               \
        const float eu = (eu_expr);                                                 \
        const float t1 = eu * p.inv_cs2;                                            \
        const float bq = (rw) * ((((0.5f * eu) * eu) * p.inv_cs2) * p.inv_cs2 - huv); \
        feq[q] = (rw) * (1.f + t1) + bq;                                            \
        feq[qo] = (rw) * (1.f - t1) + bq;                                           \
    } while (0)
    JRL_FEQ_PAIR(1, 3, uxc, rw1);
    JRL_FEQ_PAIR(2, 4, uyc, rw1);
    JRL_FEQ_PAIR(5, 7, uxc + uyc, rw5);
    JRL_FEQ_PAIR(8, 6, uxc - uyc, rw5);
#undef JRL_FEQ_PAIR
}

// ------------------------------------------------------------------------------------------
// R4 / R4m / R4k: collision operators.
// ------------------------------------------------------------------------------------------
struct CollideParams
{
    EqParams eq;
    int op;          // OP_BGK / OP_MRT / OP_KBC
    int kbc_type;    // KBCType code (OP_KBC only)
    int convection;  // BGK only: advection-diffusion form, per-direction relaxation
    int axisym;
    float omega;     // (float)(1/tau)   (tau_D when convection)
    float dt;        // (float)dt
    float conv_k1;   // (float)(dt * (1 - 0.5/tau_D))      LBM_collision_2d.py:516-521
    float conv_k2;   // (float)(1 - 0.5/tau_D)             LBM_collision_2d.py:375-384
    float beta;      // (float)(1/(2 tau))                 LBM_collision_KBC_2d.py:237
    float inv_beta;  // (float)(1/beta)
    float two_m_inv_beta;  // (float)(2 - 1/beta)
    float mrt[81];   // M^-1 S M, row-major (LBM_collision_MRT_2d.py:59-69), built on the host
};

// One row of a torch-CUDA batched 9x9 mat-vec (`mat @ x.unsqueeze(-1)`, LBM_collision_MRT_2d.py:97-100,
// LBM_collision_KBC_2d.py:107-130): cuBLAS's kernel for this shape accumulates k = 0..4 and k = 5..8 as two FMA
// chains starting from 0 and adds the two halves.  Identified on a B200 by tools/exp_torch_order.py (0 differing bits
// over 10^5 products, shared and per-cell matrices, batch 1 and 2) and pinned by tests/golden/refcuda_*.npz.
__device__ __forceinline__ float dot9_cublas(const float *a, const float *x)
{
    float lo = 0.f, hi = 0.f;
#pragma unroll
    for (int k = 0; k < 5; ++k)
        lo = fmaf(a[k], x[k], lo);
#pragma unroll
    for (int k = 5; k < 9; ++k)
        hi = fmaf(a[k], x[k], hi);
    return lo + hi;
}

// Entropic KBC relaxation, LBM_collision_KBC_2d.py:70-249.  Returns Omega in om[].
template <int KT = -1>
__device__ __forceinline__ void kbc_omega(const CollideParams &p, float rho, float ux, float uy,
                                          const float f[9], const float feq[9], float om[9])
{
    const int type = KT < 0 ? p.kbc_type : KT;
    if (type == KBC_LBGK)
    {
        // ds = f - feq, dh = 0, gamma = 2: Omega = -beta * (2 ds + 2 * 0)
#pragma unroll
        for (int q = 0; q < 9; ++q)
            om[q] = -p.beta * (2.f * (f[q] - feq[q]) + 2.f * 0.f);
        return;
    }
    const bool central = (type == KBC_A || type == KBC_B);
    // second-order moments M20, M11, M02 of f and feq: rows 3, 4, 5 of `C_mat @ f` / `euev @ f` (:100-130), a batched
    // 9x9 mat-vec that torch-CUDA hands to cuBLAS, whose kernel sums k = 0..4 and k = 5..8 as two FMA chains and adds
    // the halves (matvec9 below; tools/exp_torch_order.py).  With raw moments the coefficients are 0 / +-1 and the
    // FMAs reduce to exact adds.
    float m20, m11, m02, q20, q11, q02;
    {
        float a20[9], a11[9], a02[9];
#pragma unroll
        for (int q = 0; q < 9; ++q)
        {
            if (central)
            {
                const float dxm = (float)JRL_EX(q) - ux;  // (c e - u), c = 1 (:96-105)
                const float dym = (float)JRL_EY(q) - uy;
                a20[q] = (dxm * dxm) * 1.f;
                a11[q] = dxm * dym;
                a02[q] = 1.f * (dym * dym);
            }
            else
            {
                a20[q] = (float)(JRL_EX(q) * JRL_EX(q));
                a11[q] = (float)(JRL_EX(q) * JRL_EY(q));
                a02[q] = (float)(JRL_EY(q) * JRL_EY(q));
            }
        }
        m20 = dot9_cublas(a20, f);
        m11 = dot9_cublas(a11, f);
        m02 = dot9_cublas(a02, f);
        q20 = dot9_cublas(a20, feq);
        q11 = dot9_cublas(a11, feq);
        q02 = dot9_cublas(a02, feq);
    }
    const float T = m20 + m02, N = m20 - m02, P = m11;
    const float Te = q20 + q02, Ne = q20 - q02, Pe = q11;
    float s[9], se[9];
    if (type == KBC_A || type == KBC_C)
    {
        s[0] = rho - T;
        s[1] = 0.5f * (0.5f * (T + N) + rho * ux);
        s[2] = 0.5f * (0.5f * (T - N) + rho * uy);
        s[3] = 0.5f * (0.5f * (T + N) - rho * ux);
        s[4] = 0.5f * (0.5f * (T - N) - rho * uy);
        se[0] = rho - Te;
        se[1] = 0.5f * (0.5f * (Te + Ne) + rho * ux);
        se[2] = 0.5f * (0.5f * (Te - Ne) + rho * uy);
        se[3] = 0.5f * (0.5f * (Te + Ne) - rho * ux);
        se[4] = 0.5f * (0.5f * (Te - Ne) - rho * uy);
    }
    else
    {
        s[0] = rho + 0.f * N;
        s[1] = 0.5f * (0.5f * N + rho * ux);
        s[2] = 0.5f * (0.5f * (-N) + rho * uy);
        s[3] = 0.5f * (0.5f * N - rho * ux);
        s[4] = 0.5f * (0.5f * (-N) - rho * uy);
        se[0] = rho + 0.f * Ne;
        se[1] = 0.5f * (0.5f * Ne + rho * ux);
        se[2] = 0.5f * (0.5f * (-Ne) + rho * uy);
        se[3] = 0.5f * (0.5f * Ne - rho * ux);
        se[4] = 0.5f * (0.5f * (-Ne) - rho * uy);
    }
    s[5] = 0.25f * P;
    s[6] = 0.25f * (-P);
    s[7] = 0.25f * P;
    s[8] = 0.25f * (-P);
    se[5] = 0.25f * Pe;
    se[6] = 0.25f * (-Pe);
    se[7] = 0.25f * Pe;
    se[8] = 0.25f * (-Pe);
    const float eps = 1e-10f;
    float ds[9], dh[9], a[9], b[9];
#pragma unroll
    for (int q = 0; q < 9; ++q)
    {
        ds[q] = s[q] - se[q];
        dh[q] = (f[q] - feq[q]) - ds[q];
        const float den = feq[q] + eps;
        a[q] = div_ieee(ds[q] * dh[q], den);
        b[q] = div_ieee(dh[q] * dh[q], den);
    }
    const float dsdh = sum9(a), dhdh = sum9(b);  // `.sum(dim=1, keepdim=True)` (:240-241): torch-CUDA order
    const float gamma = p.inv_beta - p.two_m_inv_beta * div_ieee(dsdh, dhdh + eps);
#pragma unroll
    for (int q = 0; q < 9; ++q)
        om[q] = -p.beta * (2.f * ds[q] + gamma * dh[q]);
}

// Per-cell collision of a non-obstacle cell given its moments.  `r` is mesh_grid x (radius) of
// the cell, used by the axisymmetric concentration form only.  `G` (nullable) is the fluid
// axisymmetric source already evaluated for this cell (LBM_collision_2d.py:386-427).
//   BGK        LBM_collision_2d.py:473-555
//   MRT        LBM_collision_MRT_2d.py:84-102, 121-184
//   KBC        LBM_collision_KBC_2d.py:251-317
//   OP   : OP_BGK / OP_MRT / OP_KBC, or -1 to dispatch on p.op at run time
//   CONV : 0 = momentum form; 1 = advection-diffusion with x-axis symmetry; 2 = advection-
//          diffusion with y-axis symmetry (r = r'); 3 = advection-diffusion, no symmetry
template <int OP, int CONV>
__device__ __forceinline__ void collide_cell(const CollideParams &p, float rho, float ux, float uy,
                                             bool has_force, float fx, float fy, float r,
                                             const float *G, float f[9])
{
    const int op = OP < 0 ? p.op : (OP >= OP_KBC_T0 ? (int)OP_KBC : OP);
    if (op == OP_NONE)
        return;  // streaming + boundaries + moments only (first half of the two-launch axisymmetric fluid step)
    const float ux_raw = ux, uy_raw = uy;  // the KBC moments use the unshifted velocity (LBM_collision_KBC_2d.py:295-305)
    if (has_force)
        half_force_shift(rho, ux, uy, fx, fy);
    float feq[9];
    equilibrium(p.eq, rho, ux, uy, feq);
    if (op == OP_BGK)
    {
        if (CONV == 1)
        {
            // omega_q = 1/tau_D + dt(1 - 0.5/tau_D) e_qx / r ; G_q = -(u_r/r) heq_q (1 - 0.5/tau_D - 0.5 s_q)
            // The reference divides per direction, (k e_qx) / r.  e_qx is +1, -1 or 0 and IEEE division is
            // sign-symmetric, so three quotients give the same bits: k/r, -(k/r), and (k 0)/r = 0 for every
            // finite non-zero r (NaN for r = 0 or NaN, kept).
            const float ur_r = div_ieee(-ux, r);
            const float w1 = p.conv_k1 / r, s1 = p.conv_k2 / r;
            const float z = (r == 0.f || r != r) ? 0.f / r : 0.f;
#pragma unroll
            for (int q = 0; q < 9; ++q)
            {
                const int ex = JRL_EX(q);
                const float w = p.omega + (ex > 0 ? w1 : (ex < 0 ? -w1 : z));
                const float s = ex > 0 ? s1 : (ex < 0 ? -s1 : z);
                const float g = (ur_r * feq[q]) * (p.conv_k2 - 0.5f * s);
                f[q] = (f[q] + w * (feq[q] - f[q])) + p.dt * g;
            }
        }
        else if (CONV == 2)
        {
            // LBM_collision_2d.py:428-436: G = w_q * A1, A1 = -rho u_z / r', uniform omega
            // (the per-direction branch for the y axis is unreachable, :522).  r here is r'.
            const float a1 = div_ieee((-rho) * uy, r);
#pragma unroll
            for (int q = 0; q < 9; ++q)
                f[q] = (f[q] + p.omega * (feq[q] - f[q])) + p.dt * (a1 * weight(q));
        }
        else
        {
#pragma unroll
            for (int q = 0; q < 9; ++q)
                f[q] = (f[q] + p.omega * (feq[q] - f[q])) + (G ? p.dt * G[q] : 0.f);
        }
        return;
    }
    float om[9];
    if (op == OP_MRT)
    {
        float d[9];
#pragma unroll
        for (int q = 0; q < 9; ++q)
            d[q] = feq[q] - f[q];
#pragma unroll
        for (int i = 0; i < 9; ++i)
            om[i] = dot9_cublas(p.mrt + i * 9, d);
    }
    else
        kbc_omega<(OP >= OP_KBC_T0 ? kbc_code_of(OP - OP_KBC_T0) : -1)>(p, rho, ux_raw, uy_raw, f, feq, om);
#pragma unroll
    for (int q = 0; q < 9; ++q)
        f[q] = (f[q] + om[q]) + (G ? p.dt * G[q] : 0.f);
}

}  // namespace jrl
