"""Per-keyframe global BA over a growing scan (CSfM::mapping, CSfM.cpp:252-259): the reference's way — re-send everything
every keyframe (hsfm_ba_solve_packed on the problem so far) — against the scan (hsfm_ba_scan_append_keyframe + resident
solve + download).  python scripts/bench_scan.py [c1|c2] [first keyframes]"""
import dataclasses, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sfm_b200 import api, synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "c1"
k0 = int(sys.argv[2]) if len(sys.argv) > 2 else 5
p = synth.make_config(cfg)
start, stages, final = synth.keyframe_stream(p, k0)
o = api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR)
ctx = api.Context(0)
nbytes = lambda q: q.obs.nbytes + q.cam_idx.nbytes + q.pt_idx.nbytes + q.K.nbytes + q.cam.nbytes + q.pts.nbytes  # noqa: E731

def grow(cur, st, cam, pts):
    return dataclasses.replace(cur, obs=np.concatenate([cur.obs, st["obs"]]), cam_idx=np.concatenate([cur.cam_idx, st["cam_idx"]]),
                               pt_idx=np.concatenate([cur.pt_idx, st["pt_idx"]]), K=np.concatenate([cur.K, st["K"]]),
                               cam=np.concatenate([cam, st["cam"]]), pts=np.concatenate([pts, st["pts"]]))

for rep in range(2):  # first pass warms everything up
    # (a) one-shot per keyframe
    cur = start
    cam, pts, s = ctx.solve_packed(cur, opts=o)
    t_a, sent_a, its_a = 0.0, 0, 0
    for st in stages:
        cur = grow(cur, st, cam, pts)
        t0 = time.perf_counter()
        cam, pts, s = ctx.solve_packed(cur, opts=o)
        t_a += time.perf_counter() - t0
        sent_a += nbytes(cur); its_a += s["num_iterations"]
    cost_a = s["final_cost"]
    # (b) scan
    ctx.scan_begin(start)
    ctx.solve_resident(o)
    t_b, sent_b, its_b = 0.0, 0, 0
    for st in stages:
        t0 = time.perf_counter()
        sent_b += ctx.scan_append_keyframe(st["K"], st["cam"], st["pts"], st["obs"], st["cam_idx"], st["pt_idx"])
        s = ctx.solve_resident(o)
        camb, ptsb = ctx.download()
        t_b += time.perf_counter() - t0
        its_b += s["num_iterations"]
    cost_b = s["final_cost"]
n = len(stages)
print(json.dumps({"config": cfg, "keyframes_appended": n, "final": {"n_cams": final.n_cams, "n_pts": final.n_pts, "n_obs": final.n_obs},
                  "one_shot": {"ms_per_keyframe": 1e3 * t_a / n, "h2d_bytes_per_keyframe": sent_a / n, "lm_iterations": its_a, "final_cost": cost_a},
                  "scan": {"ms_per_keyframe": 1e3 * t_b / n, "h2d_bytes_per_keyframe": sent_b / n, "lm_iterations": its_b, "final_cost": cost_b},
                  "rel_cost_diff": abs(cost_a - cost_b) / cost_a, "max_cam_diff": float(np.abs(cam - camb).max())}))
