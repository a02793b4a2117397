"""Row-slab decomposition of one lattice across the GPUs of a node.

The reference is single-device (no nccl / mpi / torch.distributed anywhere); this is new work.
One process per GPU.  Rank r owns global rows [row0, row0 + nrows) of every field and stores
them as local rows 1..nrows of a [B, C, nrows + 2, W] tensor; local rows 0 and nrows + 1 are
ghost rows holding the neighbour ranks' edge rows.  Streaming is a radius-1 stencil, so ONE
exchange of the post-collision populations' edge rows per step is the only communication; the
periodic wrap of FLUID cells (propagation.cu:86-87) makes the rank topology a ring.

Per step (CUDA):   comm stream : wait for main, update the two edge rows (one launch each, one cell per thread), pack them,
                                 ncclSend / ncclRecv with both neighbours, unpack into the ghost rows
                   main stream : update the interior rows (side by side with all of the above), then wait for comm
The exchange moves 2 x 9 x W fp32 per rank per step (2.4 MB at W = 32768) against
72 B x nrows x W of HBM traffic for the interior, so it hides completely behind the interior.

A slab step is ~15 host calls (three step launches with their work-list passes, pack / NCCL send-recv / unpack, stream
switches) issued from Python: ~200 us of host time, which the 1.7 ms interior of the 32768^2 grid hides but the 185 us
of kernels of the 8192^2 concentration grid on 8 GPUs do not.  `step()` can therefore capture the whole step -- both
streams and the NCCL point-to-point calls -- into a CUDA graph per ping-pong parity once the same call has been seen
twice, and replay it: one graph launch per step (1024 x 8192 per rank: 0.229 -> 0.202 ms; profiles/r2_slab_trace.md).
OFF by default (`graph=True` or JRL_SLAB_GRAPH=1 enables it; bench.py's C4 slab does).  The replayed run is bit-exact
against the single domain (tools/check_slab_nccl.py at N = 2), but the captured graphs hold NCCL operations:
`release_all_graphs()` (or `SlabLattice.close()`) MUST run before `dist.destroy_process_group()`, which otherwise never
returns.  `AxisymFluidSlab` (below) runs JRL stage A's two-pass step on the same slabs.
"""
from __future__ import annotations

import ctypes
import os
import weakref

import torch
import torch.distributed as dist

import jrl_native as N


def slab_rows(H, world, rank):
    """Contiguous, near-equal split of H rows: (row0, nrows) of `rank`."""
    base, rem = divmod(H, world)
    n = base + (1 if rank < rem else 0)
    row0 = rank * base + min(rank, rem)
    return row0, n


_LATTICES = weakref.WeakSet()


def release_all_graphs():
    """Drop every captured slab-step graph of this process.  MUST run before `dist.destroy_process_group()`: a CUDA graph
    that holds captured NCCL operations keeps the communicator busy and the destroy call never returns (seen on 2 x B200)."""
    for lat in list(_LATTICES):
        lat.close()


class SlabLattice:
    def __init__(self, fused, H, W, rank=None, world=None, device=None, group=None, B=1, overlap=True, graph=None):
        self.fs = fused
        self.H, self.W, self.B = int(H), int(W), int(B)
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.device = torch.device(device)
        self.row0, self.nrows = slab_rows(self.H, self.world, self.rank)
        if self.nrows < 3:
            raise RuntimeError("a slab needs at least 3 rows per rank")
        self.prev = (self.rank - 1) % self.world
        self.next = (self.rank + 1) % self.world
        self.cuda = self.device.type == "cuda"
        self.overlap = overlap and self.cuda
        self.comm_stream = torch.cuda.Stream(self.device, priority=-1) if self.overlap else None
        self._staging = {}
        self.f = None       # current populations  [B,9,nrows+2,W]
        self.f_next = None  # double buffer
        self.fields = {}    # name -> slab tensor of the static / per-step inputs (flags, phi_obs, vel, ...)
        self.use_graph = self.cuda and (os.environ.get("JRL_SLAB_GRAPH", "0") == "1" if graph is None else bool(graph))
        self._descs = {}    # (parity, argument set) -> the three cached jrl_step_desc of a step
        self._graphs = {}   # call signature -> [times seen eagerly, CUDAGraph or None]
        self.graph_replays = 0
        # host order of the interior launch and the exchange's calls (both orders give the same device-side dependencies)
        self.interior_first = os.environ.get("JRL_SLAB_ORDER", "interior_first") == "interior_first"
        _LATTICES.add(self)

    def close(self):
        """Release the captured graphs (see release_all_graphs) and the cached descriptors."""
        if self._graphs:
            if self.cuda:
                torch.cuda.synchronize(self.device)
            self._graphs.clear()
        self._descs.clear()

    # -- storage ---------------------------------------------------------------------------------
    def empty(self, C, dtype=torch.float32):
        return torch.empty((self.B, C, self.nrows + 2, self.W), dtype=dtype, device=self.device)

    def scatter_global(self, full, exchange=True):
        """Slab (with ghost rows) of a global [B,C,H,W] tensor that every rank can index (tests and
        small grids; large grids build their slab directly from a formula)."""
        C = full.shape[1]
        out = self.empty(C, full.dtype)
        idx = torch.arange(self.row0 - 1, self.row0 + self.nrows + 1) % self.H
        out.copy_(full[:, :, idx.to(full.device), :].to(self.device))
        return out

    def _moment_buffer(self, name, C):
        if name not in self._staging:
            self._staging[name] = self.empty(C)
        return self._staging[name]

    def owned(self, slab):
        return slab[:, :, 1:self.nrows + 1, :]

    # -- halo exchange ---------------------------------------------------------------------------
    def _stage(self, C, dtype, key):
        k = (C, dtype, key)
        if k not in self._staging:
            self._staging[k] = torch.empty((self.B * C, 1, self.W), dtype=dtype, device=self.device)
        return self._staging[k]

    def _pack(self, field, row, staging):
        BC = field.shape[0] * field.shape[1]
        if field.dtype == torch.float32:
            N.call("jrl_halo_pack_f32", ctypes.c_void_p(field.data_ptr()), ctypes.c_void_p(staging.data_ptr()),
                   BC, self.nrows + 2, self.W, row, 1, like=field)
        else:
            staging.copy_(field.reshape(BC, self.nrows + 2, self.W)[:, row:row + 1, :])

    def _unpack(self, field, row, staging):
        BC = field.shape[0] * field.shape[1]
        if field.dtype == torch.float32:
            N.call("jrl_halo_unpack_f32", ctypes.c_void_p(field.data_ptr()), ctypes.c_void_p(staging.data_ptr()),
                   BC, self.nrows + 2, self.W, row, 1, like=field)
        else:
            field.reshape(BC, self.nrows + 2, self.W)[:, row:row + 1, :].copy_(staging)

    def exchange(self, field):
        """Fill `field`'s two ghost rows from the neighbour ranks' edge rows (ring)."""
        C = field.shape[1]
        n = self.nrows
        up, down = self._stage(C, field.dtype, "s_up"), self._stage(C, field.dtype, "s_down")
        self._pack(field, n, up)      # my last owned row  -> next rank's ghost row 0
        self._pack(field, 1, down)    # my first owned row -> prev rank's ghost row n+1
        if self.world == 1:
            self._unpack(field, 0, up)
            self._unpack(field, n + 1, down)
            return
        r_lo, r_hi = self._stage(C, field.dtype, "r_lo"), self._stage(C, field.dtype, "r_hi")
        ops = [dist.P2POp(dist.isend, up, self._peer(self.next), self.group, tag=1),
               dist.P2POp(dist.isend, down, self._peer(self.prev), self.group, tag=2),
               dist.P2POp(dist.irecv, r_lo, self._peer(self.prev), self.group, tag=1),
               dist.P2POp(dist.irecv, r_hi, self._peer(self.next), self.group, tag=2)]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        self._unpack(field, 0, r_lo)
        self._unpack(field, n + 1, r_hi)

    def _peer(self, group_rank):
        return group_rank if self.group is None else dist.get_global_rank(self.group, group_rank)

    # -- time stepping -----------------------------------------------------------------------------
    def set_state(self, f_slab, **fields):
        """f_slab: [B,9,nrows+2,W] with valid owned rows; ghost rows are filled here.  `fields`:
        slab tensors passed to FusedStep.step by keyword (flags required; rho, vel, phi_obs, force,
        mesh_grid, obs_vel ... as the configuration needs), ghost rows already valid."""
        self.f = f_slab
        self.f_next = torch.empty_like(f_slab)
        self.fields = dict(fields)
        self.exchange(self.f)

    def step(self, **kw):
        """One lattice update of the whole distributed grid; returns this rank's new slab.  Identical calls (same
        tensors, same options) are replayed from a CUDA graph from their third occurrence per ping-pong parity."""
        if not self.use_graph:
            return self._step(**kw)
        flags = kw.get("flags", self.fields.get("flags"))
        key = (self.f.data_ptr(), flags._version if flags is not None else 0,
               tuple(sorted((k, (v.data_ptr(), v._version) if isinstance(v, torch.Tensor) else v) for k, v in kw.items())))
        ent = self._graphs.get(key)
        if ent is None:
            if len(self._graphs) > 16:   # per-step-varying arguments (e.g. a new sums slice every step): stay eager
                self._graphs.clear()
            ent = self._graphs[key] = [0, None, 0]
        if ent[1] is not None:
            ent[1].replay()
            self.graph_replays += 1
            N.launch_count += ent[2]   # the library launches this graph holds (counted while it was captured)
            self.f, self.f_next = self.f_next, self.f
            return self.f
        ent[0] += 1
        if ent[0] < 2:
            return self._step(**kw)
        # second time this exact call is seen: capture it (NCCL communicators, tile masks, work lists and the
        # caching allocator have been warmed by the eager calls)
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize(self.device)
        n0 = N.launch_count
        try:
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                self._step(**kw)
        except Exception:  # noqa: BLE001 -- capture not possible in this configuration: stay eager for good
            self.use_graph = False
            self._graphs.clear()
            torch.cuda.synchronize(self.device)
            self.f, self.f_next = self._f_before_capture
            return self._step(**kw)
        ent[1], ent[2] = g, N.launch_count - n0
        # the capture only recorded; run it once for this step's effect (the eager swap already happened in _step)
        g.replay()
        self.graph_replays += 1
        return self.f

    def _step(self, **kw):
        self._f_before_capture = (self.f, self.f_next)
        fs, r0, n = self.fs, self.row0, self.nrows
        args = dict(self.fields)
        args.update(kw)
        flags = args.pop("flags")
        rho, vel = args.pop("rho", None), args.pop("vel", None)
        store = args.pop("store_moments", False)
        exchange = args.pop("exchange", True)   # False: the caller exchanges the ghost rows itself (AxisymFluidSlab)
        if store:  # the three launches of a step share one pair of moment buffers
            if "rho_out" not in args:
                args["rho_out"] = self._moment_buffer("rho_out", 1)
            if "vel_out" not in args and not fs.is_convection:
                args["vel_out"] = self._moment_buffer("vel_out", 2)
        # The three step descriptors (two edge rows, interior) are built once per ping-pong parity and argument set
        # and reused: building one costs ~40 us of Python (pointer validation), which at 8 GPUs is as long as the
        # interior kernel of the 8192^2 concentration grid.  The edge rows take the register kernel (force_vec = 1:
        # ONE launch each; the pipelined kernel would add its tile-mask, memset and work-list launches for a single row).
        key = (self.f.data_ptr(), flags.data_ptr(), flags._version, store,
               tuple(sorted((k, (v.data_ptr(), v._version) if isinstance(v, torch.Tensor) else v) for k, v in args.items())),
               None if rho is None else rho.data_ptr(), None if vel is None else vel.data_ptr())
        descs = self._descs.get(key)
        if descs is None:
            if len(self._descs) > 8:
                self._descs.clear()
            fv_edge = 1   # one cell per thread: a one-row launch is latency-bound (32 blocks instead of 8 at W = 8192)
            descs = []
            for rows, fv in (((r0, r0 + 1), fv_edge), ((r0 + n - 1, r0 + n), fv_edge), ((r0 + 1, r0 + n - 1), 0)):
                d, _, _, _, keep = fs._desc(self.f, flags, rho, vel, self.f_next, args.get("rho_out"), args.get("vel_out"), store,
                                            args.get("phi_obs"), args.get("force"), args.get("mesh_grid"), args.get("inflow_vel"),
                                            args.get("inflow_density"), args.get("use_wall", False), args.get("obs_vel"),
                                            args.get("wall_phi"), args.get("clamp01", False), (r0, n, self.H), rows, fv,
                                            args.get("sums"))
                descs.append((d, keep))
            self._descs[key] = descs
        like = self.f
        if not exchange:
            for d, _ in descs:
                N.call("jrl_step_f32", ctypes.byref(d), like=like)
        elif self.overlap:
            # comm stream (high priority): edge rows -> pack -> NCCL p2p -> unpack; main stream: the interior, which
            # reads only owned rows of the previous step and writes rows the edge launches do not touch, so the two
            # run side by side.  (Until round 2 the edge launches sat on the main stream IN FRONT of the interior:
            # tools/trace_slab.py showed 2 x 40-60 us of one-row kernels serialised with a 185 us interior.)
            main = torch.cuda.current_stream(self.device)
            self.comm_stream.wait_stream(main)
            with torch.cuda.stream(self.comm_stream):
                N.call("jrl_step_f32", ctypes.byref(descs[0][0]), like=like)
                N.call("jrl_step_f32", ctypes.byref(descs[1][0]), like=like)
                if not self.interior_first:
                    self.exchange(self.f_next)
            N.call("jrl_step_f32", ctypes.byref(descs[2][0]), like=like)
            if self.interior_first:
                with torch.cuda.stream(self.comm_stream):
                    self.exchange(self.f_next)
            main.wait_stream(self.comm_stream)
        else:
            # 1. edge rows first: they are what the neighbours wait for
            N.call("jrl_step_f32", ctypes.byref(descs[0][0]), like=like)
            N.call("jrl_step_f32", ctypes.byref(descs[1][0]), like=like)
            N.call("jrl_step_f32", ctypes.byref(descs[2][0]), like=like)
            self.exchange(self.f_next)
        self.f, self.f_next = self.f_next, self.f
        return self.f


class AxisymFluidSlab:
    """JRL stage A -- the axisymmetric FLUID step (demo_2d_LBM_JRL_fluid_cylinder.py:216-247) -- on row slabs.

    Its collision source takes gradients of the NEW moments (LBM_collision_2d.py:386-427), so the step is two passes with
    an exchange in between ("stage A"):

        1. streaming + moving wall + moments in one fused pass (JRL_OP_NONE) on the owned rows -> f', rho', vel'
        2. ghost rows of rho', vel' from the neighbours (the source's 3 x 3 stencils; next step's wall reads vel' too)
        3. collision with the axisymmetric source on the owned rows (jrl_collide_f32, slab form) -> f"
        4. ghost rows of f" for the next step's streaming

    `lat`: a SlabLattice over a FusedStep built with operator="none" and the stage's axisymmetric_type; `collision`: the
    LBMCollision2d of the same SimulationRunner.  Bit-identical to the single domain (tests/test_slab_gloo.py, world 2
    and 3 over gloo); not yet timed on GPUs -- all four stages run back to back on the current stream."""

    def __init__(self, lat, collision, dx=1.0, dt=1.0):
        if lat.fs.op != N.OP_NONE:
            raise RuntimeError("AxisymFluidSlab needs a SlabLattice over FusedStep(operator='none')")
        self.lat, self.col, self.dx, self.dt = lat, collision, float(dx), float(dt)
        self._f_post = None

    def step(self, rho, vel, mesh_grid, force=None, **kw):
        """rho, vel: slab tensors of the previous step's moments with valid ghost rows; mesh_grid (cell centres) and force
        (density * gravity, or None) slabs go to the collision only, as in the demo.  Per-step fields of the first pass
        (flags, phi_obs, obs_vel, ...) by keyword as in SlabLattice.step, or once through set_state.  Returns the new
        (rho, vel) slabs, ghost rows exchanged; the new populations are lat.f."""
        lat = self.lat
        lat.step(rho=rho, vel=vel, store_moments=True, exchange=False, **kw)
        rho_new, vel_new = lat._staging["rho_out"].clone(), lat._staging["vel_out"].clone()
        lat.exchange(rho_new)
        lat.exchange(vel_new)
        args = dict(lat.fields)
        args.update(kw)
        op, kbc, tau = self.col._operator(False, None)
        f_post = lat.f_next   # the buffer the previous populations lived in: free until the next streaming pass
        self.col._collide(self.dx, self.dt, lat.f, rho_new, vel_new, args["flags"], force, mesh_grid, False, op, kbc, tau,
                          slab=(lat.row0, lat.nrows, lat.H), f_out=f_post)
        lat.f, lat.f_next = f_post, lat.f
        lat.exchange(lat.f)
        return rho_new, vel_new
