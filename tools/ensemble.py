#!/usr/bin/env python
"""batch_validation.py as an N-GPU ensemble (BASELINE.json configs[4]):
independent stage-A (axisymmetric tip flow) -> stage-B (concentration + tip-current record) pairs,
one pair per GPU, no communication on the data path -- replicas, as the reference's sweep is eight
sequential `os.system` runs (demo/batch_validation.py:26-55).

    python tools/ensemble.py [--steps-a 2000] [--balance 200] [--walk 1000]                (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29540 tools/ensemble.py ...                                          (N GPUs)

--mode scripts (no torchrun; `--gpus N`): runs EXACTLY the eight commands of batch_validation.py:26-55 (`--fluid water
--is_negative`: 45 degree with and without gravity, cylinder, infinite plane; stage A then stage B of each pair) as
subprocesses of the reference's own, unmodified demo scripts with this package first on PYTHONPATH, one pair per GPU
(CUDA_VISIBLE_DEVICES), stage B starting when its stage A has written the velocity cache (vel `Re_<Re>.h5`; the
cylinder pair needs the file copied from `_g9/` to `_cylinder_g9/`).  The full runs are 11-14 million
steps, so `--re` scales the tip speed (steps = travel / (Re visc / r)); `--re 3e-4` is batch_validation's own value.

The other modes: each member runs the native 128 x 64 grid for a FIXED step budget (the full runs are 11-14 million
steps) through the drop-in solver objects exactly as the demos drive them (host-side tip
motion included); stage B is advected by stage A's own velocity field, normalised by the tip speed
(the demos pass it through an HDF5 file, demo_2d_LBM_JRL_C_cylinder.py:276-303).  Members are batch_validation.py's
four pairs -- 45 degree with and without gravity, cylinder, infinite plane -- and, from the fifth member on, the same
matrix with --no-is_negative.  Prints one JSON line (rank 0): per-member wall time and the
aggregate member-steps per second."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies

import torch  # noqa: E402

# batch_validation.py:26-55: (geometry, gravity in lattice units); members 4-7 = the same matrix with --no-is_negative
PAIRS = [("45degree", 2e-5), ("45degree", 0.0), ("cylinder", 2e-5), ("plane", 2e-5)]
MEMBERS = [(g, grav, True) for g, grav in PAIRS] + [(g, grav, False) for g, grav in PAIRS]
GEOM_KW = {"cylinder": dict(Rg=4), "plane": dict(Rg=200), "45degree": dict(Rg=4, geom="45degree", inflow_height=40.50000000000001)}
SCRIPT_GEOMS = [("45degree", -9.8), ("45degree", 0), ("cylinder", -9.8), ("InfinitePlane", -9.8)]


def run_scripts(args):
    """batch_validation.py's eight commands on the drop-in, pairs spread over the GPUs."""
    import shutil
    import subprocess
    import threading
    src = next((d for d in (os.environ.get("JRL_REFERENCE_DEMO_DIR"), "/root/reference/demo",
                            os.path.join(ROOT, "baseline", "_ref", "demo")) if d and os.path.isdir(d)), None)
    if src is None:
        raise SystemExit("ensemble --mode scripts: the reference's demo scripts were not found")
    out = os.path.abspath(args.out or os.path.join(ROOT, "gpurun_out", "ensemble_scripts"))
    demo = os.path.join(out, "demo")
    shutil.rmtree(demo, ignore_errors=True)
    os.makedirs(demo)
    for f in os.listdir(src):
        if f.startswith("demo_2d_LBM_JRL_") and f.endswith(".py"):
            shutil.copyfile(os.path.join(src, f), os.path.join(demo, f))
    env0 = dict(os.environ, TQDM_DISABLE="1",
                PYTHONPATH=os.pathsep.join([os.path.join(ROOT, "jrl-lbm_b200"), os.path.join(ROOT, "tests", "standins"),
                                            os.environ.get("PYTHONPATH", "")]))
    # batch_validation.py:17-24 with `--fluid water`
    visc, D, r_real = 0.89e-6, 7.85e-10, 12.5e-6
    Re = args.re
    vel_obs_real = args.vel_obs_real if args.vel_obs_real else Re * visc / r_real
    Pe = args.pe if args.pe else vel_obs_real * r_real / D
    times, lock = {}, threading.Lock()

    def pair(i, geom, g):
        env = dict(env0, CUDA_VISIBLE_DEVICES=str(i % args.gpus))
        gdir = int(abs(g))
        for stage, cmd in (("A", f"demo_2d_LBM_JRL_fluid_{geom}.py --Re {Re} --vel_obs_real {vel_obs_real} --gravity_strength_real {g}"),
                           ("B", f"demo_2d_LBM_JRL_C_{geom}.py --Re {Re} --Pe {Pe} --vel_obs_real {vel_obs_real} --gravity_strength_real {g} "
                                 f"--is_negative" + (f" --balance_time {args.balance_time}" if args.balance_time else ""))):
            if stage == "B" and geom == "cylinder":
                os.makedirs(os.path.join(demo, f"_cylinder_g{gdir}"), exist_ok=True)
                shutil.copyfile(os.path.join(demo, f"_g{gdir}", f"Re_{Re}.h5"), os.path.join(demo, f"_cylinder_g{gdir}", f"Re_{Re}.h5"))
            t0 = time.perf_counter()
            with open(os.path.join(out, f"{geom}_g{gdir}_{stage}.log"), "w") as log:
                rc = subprocess.run([sys.executable] + cmd.split(), cwd=demo, env=env, stdout=log, stderr=subprocess.STDOUT).returncode
            with lock:
                times[f"{geom}_g{gdir}_{stage}"] = dict(seconds=round(time.perf_counter() - t0, 2), rc=rc, gpu=i % args.gpus)

    t0 = time.perf_counter()
    groups = [[] for _ in range(args.gpus)]
    for i, (geom, g) in enumerate(SCRIPT_GEOMS):
        groups[i % args.gpus].append((i, geom, g))
    threads = [threading.Thread(target=lambda grp=grp: [pair(*p) for p in grp]) for grp in groups if grp]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    print(json.dumps({"config": "batch_validation.py:26-55, the eight commands (--fluid water --is_negative), unmodified demo scripts on "
                                "the drop-in", "n_gpus": args.gpus, "Re": Re, "Pe": Pe, "vel_obs_real": vel_obs_real,
                      "wall_s": round(time.perf_counter() - t0, 2), "members": times, "communication": "none (replicas)"}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps-a", type=int, default=2000)
    ap.add_argument("--balance", type=int, default=200)
    ap.add_argument("--walk", type=int, default=1000)
    ap.add_argument("--members", type=int, default=0, help="members in total (default: one per rank)")
    ap.add_argument("--out", default="", help="directory: per member, hand stage A's velocity to stage B through the field cache "
                                              "(vel.h5 / vel.h5.npz, as the demos do), keep the final C / h cache and a Tecplot .dat "
                                              "of the final state streamed off the device by jrl_io.AsyncSnapshotter")
    ap.add_argument("--gpus", type=int, default=1, help="--mode scripts: GPUs to spread the four pairs over")
    ap.add_argument("--re", type=float, default=2.0, help="--mode scripts: Re of the runs (batch_validation: 3e-4 = 13.8 M steps)")
    ap.add_argument("--vel-obs-real", dest="vel_obs_real", type=float, default=0.05)
    ap.add_argument("--pe", type=float, default=1.95)
    ap.add_argument("--balance-time", dest="balance_time", type=float, default=0.000962)
    ap.add_argument("--mode", default="fast", choices=["demo", "fast", "scripts"],
                    help="demo: stage B as the demo drives it (host torch ops for the tip every step); fast: both stages' "
                         "loops in one library call each (jrl_fluid_walk_f32, jrl_conc_walk_f32)")
    args = ap.parse_args()
    if args.mode == "scripts":
        return run_scripts(args)
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
        geom, g, neg = MEMBERS[m % len(MEMBERS)]
        gk = GEOM_KW[geom]
        a = S.jrl_fluid(api, dev, res=(128, 64), steps=args.steps_a, gravity_strength=g, vel_obs=vel_obs,
                        device_tip="run" if args.mode == "fast" else False, **gk)
        u = torch.from_numpy(a["vel"]).to(dev) / vel_obs
        if args.out:  # the stage A -> B coupling file (demo_2d_LBM_JRL_fluid_cylinder.py:266-268, ..._C_cylinder.py:276)
            d = os.path.join(args.out, "member_%02d" % m)
            jrl_io.save_fields(os.path.join(d, "vel.h5"), vel=u.reshape(1, 2, 128, 64))
            u = jrl_io.load_fields(os.path.join(d, "vel.h5"), names=["vel"], device=dev, dtype=torch.float32)["vel"][0]
        b = S.jrl_conc(api, dev, res=(128, 64), balance_steps=args.balance, walk_steps=args.walk, is_negative=neg,
                       clamp=neg and geom != "45degree", vel_snaps=u.reshape(1, 2, 128, 64).contiguous(), fused=True,
                       device_tip="run" if args.mode == "fast" else False, **gk)
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
        print(json.dumps({"config": "batch_validation ensemble: 45 degree +-gravity, cylinder, infinite plane (members 4-7: --no-is_negative), native 128x64", "n_gpus": world, "members": total,
                          "steps_per_member": {"stage_a": args.steps_a, "balance": args.balance, "walk": args.walk},
                          "wall_s": round(wall, 3), "per_rank_s": [round(float(x.item()), 3) for x in ts],
                          "member_steps_per_s": round(steps / wall, 1), "communication": "none (replicas)", "mode": args.mode,
                          "rank0_last_current": results[0][0] if results else None}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
