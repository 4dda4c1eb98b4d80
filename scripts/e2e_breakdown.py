"""Where a host-buffer call (hsfm_ba_solve_packed on pinned memory) spends its time: python scripts/e2e_breakdown.py [config] [reps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sfm_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
prob = synth.make_config(cfg)
pin = lambda a: torch.from_numpy(a).pin_memory().numpy()  # noqa: E731
hp = synth.Problem(obs=pin(prob.obs), cam_idx=pin(prob.cam_idx), pt_idx=pin(prob.pt_idx), K=pin(prob.K), cam=pin(prob.cam), pts=pin(prob.pts),
                   cam_true=prob.cam_true, pts_true=prob.pts_true, dist=None, name=prob.name)
ctx = api.Context(0)
opts = api.default_options()
cam0, pts0 = prob.cam.copy(), prob.pts.copy()
ctx.solve_packed(hp, opts=opts, inplace=True)
for r in range(reps):
    np.copyto(hp.cam, cam0); np.copyto(hp.pts, pts0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, _, s = ctx.solve_packed(hp, opts=opts, inplace=True)
    torch.cuda.synchronize()
    wall = 1e3 * (time.perf_counter() - t0)
    other = wall - s["h2d_ms"] - s["setup_ms"] - s["solve_ms"] - s["d2h_ms"]
    print(f"{cfg} call {r}: wall {wall:.2f} ms = h2d {s['h2d_ms']:.2f} + index/order {s['setup_ms']:.2f} + solve {s['solve_ms']:.2f} ({s['num_iterations']} it) "
          f"+ d2h {s['d2h_ms']:.2f} + other {other:.2f}")
