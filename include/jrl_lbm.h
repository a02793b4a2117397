/* libjrl_lbm.so -- C ABI of the B200-native D2Q9 lattice update (sm_100a).
 *
 * This is the drop-in boundary for JRL-LBM's per-step hot path.  Every entry point replaces
 * one of the reference's pybind/TORCH_LIBRARY functions or one chain of torch ops inside a
 * solver-object method (reference file:line cited per function, relative to the reference
 * tree).  Conventions (reference: propagation.cu:66-72):
 *
 *   - all pointers are DEVICE pointers owned by the caller; the library never allocates or
 *     frees user-visible memory and never synchronises the stream;
 *   - layout [B, C, H, W] contiguous, fp32; flags uint8 [B,1,H,W];
 *     f/h: C = 9, rho/C/phi_obs: C = 1, vel/force/mesh_grid/obs_vel: C = 2 (channel 0 = x);
 *   - optional tensors are NULL (the reference passes empty tensors, propagation.cu:513-517);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - return value 0 = success, negative = error; jrl_last_error() returns the message of
 *     the calling thread's last failure.
 *
 * There is no CPU fallback: without a CUDA device every compute entry point fails with
 * JRL_ERR_CUDA.
 */
#ifndef JRL_LBM_H
#define JRL_LBM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define JRL_LBM_VERSION 100

enum { JRL_OK = 0, JRL_ERR_ARG = -1, JRL_ERR_CUDA = -2 };

/* utils/types.py:5-23, 69-73 */
enum { JRL_CELL_NOTHING = 0, JRL_CELL_FLUID = 1, JRL_CELL_OBSTACLE = 2, JRL_CELL_EMPTY = 4,
       JRL_CELL_INFLOW = 8, JRL_CELL_OUTFLOW = 16, JRL_CELL_INFLOW_2 = 32 };
enum { JRL_AXI_NOT = 0, JRL_AXI_X = 1, JRL_AXI_Y = 2 };
enum { JRL_OP_BGK = 0, JRL_OP_MRT = 1, JRL_OP_KBC = 2,
       JRL_OP_NONE = 3 /* jrl_step_f32 only: stop before the collision -- f_out = the streamed (and wall-corrected)
                          populations, rho_out / vel_out = their moments.  First half of the two-launch step of the
                          axisymmetric FLUID form, whose source term needs the neighbours' post-stream moments
                          (jrl_collide_f32 is the second half). */ };
#define JRL_SUM_SLOTS 64
enum { JRL_KBC_LBGK = 0, JRL_KBC_A = 0x85, JRL_KBC_B = 0x86, JRL_KBC_C = 0x89, JRL_KBC_D = 0x8A };

int jrl_version(void);
const char *jrl_last_error(void);

/* ---- R1: streaming with all link-wise boundary rules --------------------------------------
 * Replaces propagation.propagation_2d_forward (propagation.cpp:15-33 -> propagation.cu:490-524,
 * kernel :22-488) AND the wrapper's `f_new = f + 0.0` copy (LBM_propagation_2d.py:49): f_new is
 * fully written, it need not be initialised.  rho / vel are mutated in place at INFLOW and
 * INFLOW_2 cells exactly as the reference does (propagation.cu:300, 320-321, 362).
 * inflow_density < 0 means "none" (LBM_propagation_2d.py:61-62); inflow_vel is a device [2].
 * OUTFLOW cells copy their inward neighbour's post-stream populations (the reference has a
 * data race there, propagation.cu:250-254; this is the deterministic reading). */
int jrl_stream_f32(const uint8_t *flags, const float *f, const float *phi_obs, float *f_new,
                   float *rho, float *vel, const float *inflow_vel, double inflow_density,
                   int axisym, int B, int H, int W, void *stream);

/* The in-place side effects of streaming alone (propagation.cu:300, 320-321, 362): rho = inflow_density (when >= 0)
 * and vel = inflow_vel (when given) at INFLOW cells, rho = 1 at INFLOW_2 cells.  jrl_step_f32 leaves the
 * previous-step rho / vel untouched; the host layer issues this launch when the caller still holds those tensors. */
int jrl_inflow_apply_f32(const uint8_t *flags, float *rho, float *vel, const float *inflow_vel, double inflow_density,
                         int B, int H, int W, void *stream);

/* prop.rebounce_obstacle, LBM_propagation_2d.py:181-186 */
int jrl_rebounce_f32(const float *f, const uint8_t *flags, float *f_new, int B, int H, int W, void *stream);

/* ---- R2 / R7: moments ---------------------------------------------------------------------
 * jrl_moments_f32 replaces the torch-op chain of LBMMacroCompute2d.macro_compute
 * (LBM_macro_compute_2d.py:316-351) including contact_angle_correctness (:237-271) when
 * use_contact != 0 (contact_tan = tan(pi/2 - contact_angle) as fp32).  `c` = dx/dt.
 * rho_new/vel_new must not alias rho_old/vel_old. */
int jrl_moments_f32(const float *f, const float *rho_old, const float *vel_old, const uint8_t *flags,
                    float *rho_new, float *vel_new, int use_contact, float contact_tan, double c,
                    int B, int H, int W, void *stream);
/* macro_compute.macro_compute_2d_forward (macro_compute.cpp:43-60, kernel macro_compute.cu:80-157) */
int jrl_moments_kernel_f32(const float *f, const float *rho_old, const float *vel_old, const uint8_t *flags,
                           float *rho_new, float *vel_new, double c, int B, int H, int W, void *stream);
/* LBMMacroCompute2d.macro_compute_C (LBM_macro_compute_2d.py:273-295); clamp01 != 0 fuses the
 * demo's torch.clamp(C, 0, 1) (demo_2d_LBM_JRL_C_cylinder.py:324) */
int jrl_moments_C_f32(const float *h, const float *C_old, const uint8_t *flags, float *C_new, int clamp01,
                      int B, int H, int W, void *stream);
/* macro_compute.macro_compute_C_2d_forward (macro_compute.cpp:29-41, kernel macro_compute.cu:22-78) */
int jrl_moments_C_kernel_f32(double dt, const float *h, const float *C_old, const uint8_t *flags,
                             const float *vel, const float *mesh_grid, float *C_new, int axisym,
                             int B, int H, int W, void *stream);

/* ---- R3: equilibrium, LBMCollision2d.get_feq_static_ (LBM_collision_2d.py:236-263) ---------- */
int jrl_feq_f32(const float *rho, const float *vel, const float *force, float *feq, double dx, double dt,
                int B, int H, int W, void *stream);

/* ---- R6: 9-point stencils, collision_2d.get_grad_2d_forward / get_div_2d_forward
 * (collision_2d.cpp:32-50, kernels collision_2d.cu:22-220) ------------------------------------ */
int jrl_grad_f32(const float *s, const uint8_t *flags, float *grad, double dx, int axisym,
                 int B, int H, int W, void *stream);
int jrl_div_f32(const float *v, const uint8_t *flags, float *div, double dx, int axisym,
                int B, int H, int W, void *stream);

/* ---- R4 / R4m / R4k / R6: collision ----------------------------------------------------------
 * One launch replaces LBMCollision2d.collision (LBM_collision_2d.py:473-555),
 * LBMCollisionMRT2d.collision (LBM_collision_MRT_2d.py:121-184) or LBMCollisionKBC2d.collision
 * (LBM_collision_KBC_2d.py:251-317), including get_feq_, the half-force velocity shift and the
 * axisymmetric source get_G_source_from_axisymmetric (LBM_collision_2d.py:343-471) with its
 * four gradient and one divergence stencils evaluated in registers. */
typedef struct jrl_collide_desc
{
    int B, H, W;
    int axisym;           /* JRL_AXI_* */
    int op;               /* JRL_OP_* */
    int kbc_type;         /* JRL_KBC_* (op == JRL_OP_KBC) */
    int is_convection;    /* BGK only */
    double dx, dt;
    double tau;           /* relaxation time used by the operator (tau_D when is_convection) */
    const float *mrt;     /* HOST pointer to the 81 fp32 entries of M^-1 S M (op == JRL_OP_MRT) */
    const float *f;       /* [B,9,H,W] post-stream populations */
    const float *rho;     /* [B,1,H,W] */
    const float *vel;     /* [B,2,H,W] */
    const uint8_t *flags; /* [B,1,H,W] */
    const float *force;   /* [B,2,H,W] or NULL */
    const float *mesh_grid; /* [B,2,H,W] (cell centres) or NULL; required when axisym != 0 */
    float *f_out;         /* [B,9,H,W]; may alias f when axisym == 0 or is_convection */
    /* Multi-GPU row slab, as in jrl_step_desc below: slab_rows == 0 = whole grid stored; slab_rows > 0 = every plane
       is a local [slab_rows + 2, W] array (H stays the GLOBAL height) whose owned rows are collided; the axisymmetric
       source reads rho / vel / flags / mesh_grid of the ghost rows, which the caller has exchanged. */
    int slab_row0, slab_rows;
} jrl_collide_desc;
int jrl_collide_f32(const jrl_collide_desc *d, void *stream);

/* ---- R5: regularised moving-wall reconstruction ----------------------------------------------
 * solid_fluid_coupling.solve_obstacle_2d_forward (solid_fluid_coupling.cpp:17-38 ->
 * solid_fluid_coupling.cu:254-290, kernel :25-252).  f_new is fully written. */
int jrl_wall_f32(double dt, double dx, double tau, const float *vel, const uint8_t *flags, const float *f,
                 float *f_new, const float *phi_obs, const float *obs_vel, int axisym,
                 int B, int H, int W, void *stream);

/* ---- R8: strip reductions feeding the validation plots ---------------------------------------
 * out[0] = sum_k ((C[y0,k]-C[y0-1,k]) + (r[k+1]C[y0,k+1]-r[k]C[y0,k])/r[k]) * area[k], k < ncol
 *          (demo_2d_LBM_JRL_C_cylinder.py:344-348, 355-359), r = mesh_grid[0,0,y0,:]
 * out[1] = sum_k C[y0,k] * area[k]   (:239-241; demo_2d_LBM_JRL_C_45degree.py:353-376)
 * area[k] = pi((k+1)^2 - k^2) (:176-180).  One warp, shuffle reduction; batch element 0. */
int jrl_strip_flux_f32(const float *C, const float *mesh_grid, int y0, int ncol, float *out2,
                       int H, int W, void *stream);

/* ---- R9 (extension; absent from the reference): momentum-exchange force on OBSTACLE cells.
 * out2[0..1] += sum over non-obstacle cells x and links q with x+e_q OBSTACLE of 2 e_q f_q(x)
 * (post-collision f).  out2 must be zeroed by the caller. */
int jrl_momentum_exchange_f32(const float *f_post, const uint8_t *flags, double *out2,
                              int B, int H, int W, void *stream);

/* ---- fused step: stream + boundaries + moments + collide in one pass ---------------------------
 * Semantically identical (bit for bit) to jrl_stream_f32 -> [jrl_wall_f32] -> jrl_moments_f32 /
 * jrl_moments_C_f32 -> jrl_collide_f32 on the same inputs, for the cases where the collision
 * is local: axisym == 0, or is_convection (the axisymmetric FLUID source needs post-stream
 * moments of neighbours and stays a two-launch path).  Pull scheme, double-buffered:
 * f_in -> f_out, 9 fp32 reads + 9 fp32 writes per cell (72 B per lattice update). */
typedef struct jrl_step_desc
{
    int B, H, W;
    int axisym;
    int op, kbc_type, is_convection;
    int use_contact;        /* fluid: contact-angle wall density on OBSTACLE cells of rho_out */
    float contact_tan;
    int clamp01;            /* concentration: clamp C to [0,1] before the collision */
    double dx, dt, tau;
    const float *mrt;       /* HOST pointer, 81 floats */
    const uint8_t *flags;
    const float *f_in;
    float *f_out;           /* must not alias f_in */
    const float *phi_obs;   /* NULL: halfway bounce-back */
    float *rho;             /* in/out: previous-step density (kept at OBSTACLE cells, read at INFLOW).
                               NULL allowed when no cell needs it (all FLUID) */
    float *vel;             /* in/out likewise; read-only input when is_convection */
    float *rho_out;         /* NULL: moments are not stored (pure 72 B/LUP path) */
    float *vel_out;
    const float *force;     /* [B,2,H,W] or NULL */
    const float *mesh_grid;
    const float *inflow_vel;
    double inflow_density;  /* < 0: none */
    /* moving wall (R5) between streaming and moments; needs phi_obs or wall_phi, and the previous-step vel */
    int use_wall;
    const float *obs_vel;
    /* Multi-GPU row slab.  slab_rows == 0: every plane holds the whole grid [H, W].
       slab_rows > 0: every plane (f, flags, phi, rho, vel, force, mesh, obs_vel, outputs) is a
       local [slab_rows + 2, W] array holding global rows slab_row0 .. slab_row0+slab_rows-1 in
       local rows 1 .. slab_rows; local rows 0 and slab_rows+1 are ghost rows the caller fills
       by halo exchange (the periodic wrap of FLUID cells makes the rank topology a ring). */
    int slab_row0, slab_rows;
    /* Global rows [row_begin, row_end) updated by this launch (lets the caller update the
       slab's edge rows first, start the halo exchange, then update the interior);
       row_begin == row_end == 0: all stored rows. */
    int row_begin, row_end;
    int force_vec;          /* 0 = automatic (pipelined TMA kernel on wide grids, else the widest cells-per-thread
                               the extents / alignment allow); 1, 2, 4 = register kernel with that many cells per
                               thread; 8 = pipelined kernel (error if the grid is not eligible) */
    double *sums;           /* NULL, or device [JRL_SUM_SLOTS][4] doubles that this launch ADDS to (caller zeroes):
                               per slot {sum rho, sum rho u_x, sum rho u_y, cells} over the non-OBSTACLE cells it
                               updated, with the post-stream moments (C, C u for the concentration form).  Warp-
                               shuffle reduction inside the step kernel, one atomicAdd per warp spread over the
                               slots; the total is the sum over slots (the mass / momentum diagnostic). */
    const float *wall_phi;  /* phi_obs seen by the moving-wall step; NULL: same as phi_obs.
                               (demo_2d_LBM_cylinder.py:139-157 streams WITHOUT phi_obs -- plain halfway
                               bounce-back -- but passes phi_obs to solve_boundary) */
    void *worklist;           /* NULL, or `worklist_bytes` of device scratch (4-byte aligned; jrl_worklist_bytes gives a
                                 sufficient size): the pipelined kernel defers its boundary groups to a second launch
                                 through it.  Contents are meaningless between calls. */
    long long worklist_bytes;
    const uint8_t *tile_mask; /* NULL, or the packed summary of `flags` written by jrl_tile_mask_u8 for exactly
                               these flags (same B, stored rows, W): lets the pipelined kernel skip the flag
                               rows of all-FLUID tiles.  Results are identical with and without it; it must be
                               recomputed whenever `flags` changes. */
} jrl_step_desc;
int jrl_step_f32(const jrl_step_desc *d, void *stream);

/* jrl_step_f32 without any allocation anywhere.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same descriptor, same results, but `tile_mask` and `worklist` must be NULL: the scratch the
 * pipelined kernel needs (packed tile mask + boundary-group work list) lives in a file-scope `static __device__`
 * array of JRL_NOALLOC_SCRATCH_BYTES (4 MiB) that every GPU loading the library holds until exit; the tile mask is
 * rebuilt from `flags` by every call.  LIMIT: jrl_tile_mask_bytes() rounded up to 256 + jrl_worklist_bytes() must fit
 * in the 4 MiB (grids up to about 4096 x 4096 cells per call); larger grids are rejected with JRL_ERR_ARG and a
 * message -- use jrl_step_f32 with caller-provided scratch there.  The call issues no cudaMalloc / cudaFree and no
 * allocating library routine (kernel launches and one 4-byte cudaMemsetAsync only).  All calls on a device SHARE the
 * array, which keeps the last call's contents (everything that must start at zero is cleared by the call itself):
 * two calls that overlap on the GPU (different streams) race; calls on one stream are safe.
 * jrl_step_noalloc_max_bytes() returns JRL_NOALLOC_SCRATCH_BYTES. */
#define JRL_NOALLOC_SCRATCH_BYTES (4ll << 20)
int jrl_step_noalloc_f32(const jrl_step_desc *d, void *stream);
long long jrl_step_noalloc_max_bytes(void);

/* Multi-step driver: `nsteps` lattice updates with the descriptor's buffers used as ping-pong pairs --
 * the time loop of the reference's demos (demo_2d_LBM.py:101-119, demo_2d_LBM_cylinder.py:138-172)
 * below the ABI, for grids small enough to be paced by the host.  Step i reads (f_in, rho, vel) and writes
 * (f_out, rho_out, vel_out), then the pairs swap: after the call the newest populations are in f_out when
 * nsteps is odd and in f_in when it is even; likewise rho / rho_out and vel / vel_out when BOTH members
 * of a pair are given (then they are state: OBSTACLE cells keep their previous moments, INFLOW cells and
 * the moving wall read them).  When rho or vel is NULL the moments are pure outputs, written to rho_out /
 * vel_out every `store_every`-th step (0: never) and at the last step.  In the concentration form vel is
 * the advecting velocity (an input, never swapped).  sums (optional): step i adds into sums + i *
 * sums_stride doubles (stride 0: one accumulator; JRL_SUM_SLOTS * 4: a per-step series).  Inputs that the
 * reference's demos rewrite every step (flags, phi_obs, ...) stay fixed for the whole call. */
int jrl_run_f32(const jrl_step_desc *d, int nsteps, int store_every, long long sums_stride, void *stream);

/* Packed flag summary for jrl_step_desc.tile_mask: one byte per row segment of the library's tile width,
 * 1 = every cell of the segment and its two x-neighbours (periodic) is FLUID.  Hs = stored rows (H, or
 * slab_rows + 2).  jrl_tile_mask_bytes returns the size of `mask` in bytes (0 on bad arguments).
 * Needs W % 16 == 0 and 16-byte aligned flags (the pipelined kernel's own requirements). */
int jrl_tile_mask_bytes(int B, int Hs, int W);
/* Size in bytes of the scratch jrl_step_desc.worklist may point to: room for every boundary group of a grid in which
 * up to 1/8 of the groups are boundary groups (groups beyond the capacity are simply computed in place). */
long long jrl_worklist_bytes(int B, int Hs, int W);
int jrl_tile_mask_u8(const uint8_t *flags, uint8_t *mask, int B, int Hs, int W, void *stream);

/* ---- device-side tip motion ----------------------------------------------------
 * One launch replaces the per-step host work of the JRL concentration demos' walk loop
 * (demo_2d_LBM_JRL_C_cylinder.py:283-307): signed distance of the moving tip (CalculateObsPhi_forC, :31-41),
 * flags restamp (top row when not is_negative, OBSTACLE where phi < 0, the INFLOW strip on the tip face),
 * C = 0 in OBSTACLE cells, and the advecting velocity C_vel = (vel0 * w0 + vel1 * w1) * vel_obs_conv with the
 * tip interior stamped to (0, vel_obs_conv).  vel1 == NULL: C_vel = vel0 * vel_obs_conv.  Batch size 1.
 * flags and C are updated in place; phi_obs [1,1,H,W] and C_vel [1,2,H,W] are fully written.  The scalars
 * are what the demo's python holds at that step: box_cx / box_cy = the fp32 box_center tensor, walks_int =
 * int(walks + 0.5 dx), w0 = (float)(1 - t), w1 = (float)t. */
/* Tip geometry (batch_validation.py:26-55 runs three of them): JRL_TIP_BOX = the box of
 * demo_2d_LBM_JRL_*_cylinder.py:31-49 (phi = max(tx, ty)); demo_2d_LBM_JRL_*_InfinitePlane.py is the same box with
 * Rg = 200, i.e. Rg_radius wider than the grid.  JRL_TIP_CONE45 = demo_2d_LBM_JRL_*_45degree.py:31-46:
 * phi = max(ty, ((x + y) - (x1 + y1)) / sqrt(2)), (x1, y1) = box_center + box_radius in fp32 (the division by the python
 * scalar sqrt(2) is torch-CUDA's multiplication by the fp32 reciprocal). */
enum { JRL_TIP_BOX = 0, JRL_TIP_CONE45 = 1 };

typedef struct jrl_tip_desc
{
    int H, W;
    int shape;              /* JRL_TIP_BOX / JRL_TIP_CONE45 */
    int is_negative;
    int walks_int, Rg_radius, radius_obs_int;
    float box_cx, box_cy, box_radius;
    float vel_obs_conv, w0, w1;
    const float *mesh_grid; /* [1,2,H,W] cell centres */
    uint8_t *flags;         /* [1,1,H,W] in/out */
    float *C;               /* [1,1,H,W] in/out */
    const float *vel0;      /* [2,H,W] velocity snapshot, normalised by the tip speed */
    const float *vel1;      /* [2,H,W] next snapshot, or NULL */
    float *phi_obs;         /* [1,1,H,W] out */
    float *C_vel;           /* [1,2,H,W] out */
} jrl_tip_desc;
int jrl_tip_prepare_conc_f32(const jrl_tip_desc *d, void *stream);

/* Stage-B walk loop below the ABI (SURVEY 8f ranks 1 + 2; demo_2d_LBM_JRL_C_cylinder.py:281-362): per step
 * jrl_tip_prepare_conc_f32 -> [jrl_tile_mask_u8 when step.tile_mask is set] -> jrl_step_f32 (concentration
 * form) -> two strip fluxes -> walks += vel_obs_conv * dt.  `step`: f_in / f_out = the h ping-pong pair,
 * rho / rho_out = the C pair, vel = the C_vel buffer, phi_obs = the phi buffer (both rewritten every step),
 * flags (rewritten in place), mesh_grid, inflow_density, clamp01 as for jrl_step_f32.  `tip`: the static
 * fields (H, W, is_negative, Rg_radius, radius_obs_int, box_cx, box_radius, vel_obs_conv, mesh_grid); the
 * per-step ones are filled here.  vel_snaps [n_snaps,2,H,W]: velocity snapshots at tip travel 0,
 * x_interval, ... interpolated as the demo does (:294-303).  records [nsteps][4] (device) = {inflow flux,
 * inflow strip sum, current flux, current strip sum} per step.  After the call the newest h / C are in
 * f_out / rho_out when nsteps is odd, in f_in / rho when even, and `walks` holds the new tip position. */
typedef struct jrl_walk_desc
{
    jrl_step_desc step;
    jrl_tip_desc tip;
    double walks, vel_obs_conv, dx, dt, box_radius, inflow_height, x_interval;
    /* The demo's own snapshot clock (:294-303, :342): x_comp = conv_timer * vel_obs_real / radius_obs_real * radius_obs,
     * conv_timer += dt_conv_real per step, and NO clamp of the snapshot index (the demo ends with an IndexError when it
     * runs past the stage-A cache; here: JRL_ERR_ARG, `walks` / `conv_timer` / `steps_done` hold the state at that
     * step).  dt_conv_real == 0 selects the algebraically equal x_comp = walks - inflow_height with a clamped index. */
    double conv_timer, dt_conv_real, vel_obs_real, radius_obs_real, radius_obs;
    int steps_done;
    const float *vel_snaps;
    int n_snaps;
    float *records;
} jrl_walk_desc;
int jrl_conc_walk_f32(jrl_walk_desc *w, int nsteps, void *stream);

/* Stage A (axisymmetric tip flow, demo_2d_LBM_JRL_fluid_cylinder.py).  jrl_tip_prepare_fluid_f32: the per-step host
 * work of :205-214 in one launch -- phi_obs (CalculateObsPhi :39-49, united with the top wall), the wall velocity field
 * obs_vel = (0, vel_obs) inside the tip and 0 outside, flags = OBSTACLE and vel = (0, vel_obs) inside the tip
 * (rows < walks_int, columns < Rg_radius).  flags / vel in place; phi_obs [1,1,H,W], obs_vel [1,2,H,W] fully written.
 * shape = JRL_TIP_CONE45 (demo_2d_LBM_JRL_fluid_45degree.py:31-46, 201-220): "inside" is phi <= 0 instead of the box,
 * obs_vel is zeroed on the top row, and vel = obs_vel inside. */
typedef struct jrl_tip_fluid_desc
{
    int H, W;
    int shape;              /* JRL_TIP_BOX / JRL_TIP_CONE45 */
    int walks_int, Rg_radius;
    float box_cx, box_cy, box_radius, vel_obs;
    const float *mesh_grid;
    uint8_t *flags;
    float *vel;
    float *phi_obs;
    float *obs_vel;
} jrl_tip_fluid_desc;
int jrl_tip_prepare_fluid_f32(const jrl_tip_fluid_desc *d, void *stream);

/* Stage-A walk loop below the ABI (:203-264): per step jrl_tip_prepare_fluid_f32 -> jrl_stream_f32 (phi_obs, rho, vel)
 * -> jrl_wall_f32 -> jrl_moments_f32 -> jrl_collide_f32 (BGK + axisymmetric source) -> walks += vel_obs * dt.
 * f[3]: population buffers, f[0] holds the state before and after the call; rho[2] / vel[2]: moment pairs, pair 0
 * current on entry, pair `current` on return.  `collide`: op, axisym, dx, dt, tau, flags, force, mesh_grid (f, rho, vel,
 * f_out are filled per step); `tip`: the static fields (H, W, Rg_radius, box_cx, box_radius, vel_obs, mesh_grid, phi_obs,
 * obs_vel buffers). */
typedef struct jrl_fluid_walk_desc
{
    jrl_collide_desc collide;
    jrl_tip_fluid_desc tip;
    double walks, vel_obs, box_radius;
    float *f[3];
    float *rho[2];
    float *vel[2];
    int current;
} jrl_fluid_walk_desc;
int jrl_fluid_walk_f32(jrl_fluid_walk_desc *w, int nsteps, void *stream);

/* ---- row-slab halo helpers (multi-GPU) ----------------------------------------------------------
 * Pack / unpack `nrows` full rows of all C planes of a [C,Hloc,W] slab starting at row `row`
 * into a contiguous [C,nrows,W] staging buffer (the message NCCL sends to the neighbour rank). */
int jrl_halo_pack_f32(const float *field, float *staging, int C, int Hloc, int W, int row, int nrows, void *stream);
int jrl_halo_unpack_f32(float *field, const float *staging, int C, int Hloc, int W, int row, int nrows, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* JRL_LBM_H */
