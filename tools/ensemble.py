#!/usr/bin/env python
"""batch_validation.py as an N-GPU ensemble (BASELINE.json configs[4]):
independent stage-A (axisymmetric tip flow) -> stage-B (concentration + tip-current record) pairs,
one pair per GPU, no communication on the data path -- replicas, as the reference's sweep is eight
sequential `os.system` runs (demo/batch_validation.py:26-55).

    python tools/ensemble.py [--steps-a 2000] [--balance 200] [--walk 1000]                (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29540 tools/ensemble.py ...                                          (N GPUs)

Each member runs the native 128 x 64 grid for a FIXED step budget (the full runs are 11-14 million
steps) through the drop-in solver objects exactly as the demos drive them (host-side tip
motion included); stage B is advected by stage A's own velocity field, normalised by the tip speed
(the demos pass it through an HDF5 file, demo_2d_LBM_JRL_C_cylinder.py:276-303).  Members differ in
gravity and in the sign convention (--is_negative), the two knobs batch_validation.py turns for the
cylinder geometry; the 45-degree and infinite-plane geometries are separate demo scripts whose set-up
code is not reproduced here.  Prints one JSON line (rank 0): per-member wall time and the
aggregate member-steps per second."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies

import torch  # noqa: E402

MEMBERS = [  # (gravity_strength in lattice units, is_negative)
    (2e-5, True), (0.0, True), (2e-5, False), (0.0, False),
    (1e-5, True), (4e-5, True), (1e-5, False), (4e-5, False),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps-a", type=int, default=2000)
    ap.add_argument("--balance", type=int, default=200)
    ap.add_argument("--walk", type=int, default=1000)
    ap.add_argument("--members", type=int, default=0, help="members in total (default: one per rank)")
    ap.add_argument("--out", default="", help="directory: per member, hand stage A's velocity to stage B through the field cache "
                                              "(vel.h5 / vel.h5.npz, as the demos do), keep the final C / h cache and a Tecplot .dat "
                                              "of the final state streamed off the device by jrl_io.AsyncSnapshotter")
    ap.add_argument("--mode", default="fast", choices=["demo", "fast"],
                    help="demo: stage B as the demo drives it (host torch ops for the tip every step); fast: both stages' "
                         "loops in one library call each (jrl_fluid_walk_f32, jrl_conc_walk_f32)")
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    import jrl_native
    jrl_native.load()
    import jrl_scenarios as S
    import jrl_io
    api = S.api_from_src()
    from src.LBM.utils import io as lbm_io
    snap = jrl_io.AsyncSnapshotter(depth=2) if args.out else None
    total = args.members or world
    mine = [m for m in range(total) if m % world == rank]
    vel_obs = 0.02  # fast tip so that it crosses cells inside the step budget (tests use the same value)

    def member(m):
        g, neg = MEMBERS[m % len(MEMBERS)]
        a = S.jrl_fluid(api, dev, res=(128, 64), steps=args.steps_a, gravity_strength=g, vel_obs=vel_obs,
                        device_tip="run" if args.mode == "fast" else False)
        u = torch.from_numpy(a["vel"]).to(dev) / vel_obs
        if args.out:  # the stage A -> B coupling file (demo_2d_LBM_JRL_fluid_cylinder.py:266-268, ..._C_cylinder.py:276)
            d = os.path.join(args.out, "member_%02d" % m)
            jrl_io.save_fields(os.path.join(d, "vel.h5"), vel=u.reshape(1, 2, 128, 64))
            u = jrl_io.load_fields(os.path.join(d, "vel.h5"), names=["vel"], device=dev, dtype=torch.float32)["vel"][0]
        b = S.jrl_conc(api, dev, res=(128, 64), balance_steps=args.balance, walk_steps=args.walk, is_negative=neg,
                       clamp=neg, vel_snaps=u.reshape(1, 2, 128, 64).contiguous(), fused=True,
                       device_tip="run" if args.mode == "fast" else False)
        if args.out:
            C, h = torch.from_numpy(b["C"]).to(dev), torch.from_numpy(b["h"]).to(dev)
            jrl_io.save_fields(os.path.join(d, "C_final.h5"), C=C, h=h)
            zero = torch.zeros_like(C)
            snap.submit(lambda hst, p=os.path.join(d, "final.dat"): lbm_io.dump_2d_plt_file_C_rho(p, hst["C"], hst["C"], hst["u"], hst["v"], 0, 0),
                        C=C, u=zero, v=zero)
        return float(b["current"][-1]), float(b["C"].sum())

    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    results = [member(m) for m in mine]
    if snap is not None:
        snap.close()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    ts = [t.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(ts, t)
    if rank == 0:
        wall = max(float(x.item()) for x in ts)
        steps = total * (args.steps_a + args.balance + args.walk)
        print(json.dumps({"config": "batch_validation ensemble (cylinder geometry, native 128x64)", "n_gpus": world, "members": total,
                          "steps_per_member": {"stage_a": args.steps_a, "balance": args.balance, "walk": args.walk},
                          "wall_s": round(wall, 3), "per_rank_s": [round(float(x.item()), 3) for x in ts],
                          "member_steps_per_s": round(steps / wall, 1), "communication": "none (replicas)", "mode": args.mode,
                          "rank0_last_current": results[0][0] if results else None}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
