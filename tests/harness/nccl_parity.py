"""Run under torch.distributed.run with N ranks (one GPU each): the point-sharded multi-GPU solve through the C ABI
(NCCL inside the library) against the single-GPU solve and the CPU oracle.  Rank 0 prints one JSON line.
Used by tests/test_gpu_multi.py."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from sfm_b200 import api, synth

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    linear = {"dense": api.LINEAR_DENSE_SCHUR, "pcg": api.LINEAR_PCG}[sys.argv[1]]
    which = sys.argv[2]
    if which == "trajectory":
        from tests.test_gpu_parity import _trajectory
        full = _trajectory(n_cams=640, n_pts=9000, closure=True)
    else:
        full = synth.make_config("c2", scale=0.1)
    shards = synth.shard_by_point(full, world)
    prob, gids = shards[rank]
    ctx = api.Context(local)
    uid = [api.Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(uid[0], rank, world)
    opts = api.default_options(linear_solver=linear, pcg_max_iterations=3000)
    # a refused upload is refused by EVERY rank (a bad index in one shard must not leave the others in a collective)
    if rank == world - 1:
        import dataclasses
        bad_idx = prob.cam_idx.copy()
        bad_idx[0] = prob.n_cams + 5
        bad = dataclasses.replace(prob, cam_idx=bad_idx)
    else:
        bad = prob
    try:
        ctx.upload(bad)
        refused = 0
    except api.BAError as e:
        refused = int(e.code == -1)  # HSFM_BA_ERR_ARG
    all_refused = [None] * world
    dist.all_gather_object(all_refused, refused)
    ctx.upload(prob)
    s = ctx.solve_resident(opts)
    cam, pts = ctx.download()
    # every rank holds the same cameras; points are gathered back to the global numbering
    all_pts = [None] * world
    dist.all_gather_object(all_pts, (gids, pts))
    cams = [None] * world
    dist.all_gather_object(cams, cam)
    out = None
    if rank == 0:
        gp = np.empty_like(full.pts)
        for g, p in all_pts:
            gp[g] = p
        cam_spread = max(float(np.abs(c - cam).max()) for c in cams)
        single = api.Context(local)
        c1, p1, s1 = single.solve_packed(full, opts=opts)
        single.close()
        from oracle import pyoracle as po
        c0, p0, so, _ = po.solve(full, opts=po.default_options(num_threads=4))
        out = dict(n_ranks=s["n_ranks"], iterations=s["num_iterations"], iterations_single=s1["num_iterations"], iterations_oracle=so["num_iterations"],
                   final_cost=s["final_cost"], final_cost_single=s1["final_cost"], final_cost_oracle=so["final_cost"], rms=s["rms_px"],
                   rms_oracle=so["rms_px"], cam_spread_across_ranks=cam_spread, cam_err_single=float(np.abs(cam - c1).max()),
                   pts_err_single=float(np.abs(gp - p1).max()), cam_err_oracle=float(np.abs(cam - c0).max()),
                   pts_err_oracle=float(np.abs(gp - p0).max()), linear_solver_used=s["linear_solver_used"], refused_by=sum(all_refused),
                   trace_cost=[t["cost"] for t in s["trace"]], trace_cost_single=[t["cost"] for t in s1["trace"]])
    ctx.close()
    dist.barrier()
    if rank == 0:
        print("NCCL_PARITY " + json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
