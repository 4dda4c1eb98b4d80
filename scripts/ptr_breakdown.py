"""Where hsfm_ba_solve_ptr (the drop-in signature, pageable storage) spends its time: HSFM_PTR_TIMING=1 python scripts/ptr_breakdown.py [config] [reps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sfm_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
prob = synth.make_config(cfg)
ctx = api.Context(0)
opts = api.default_options()
frames = np.ascontiguousarray(prob.cam.copy()); store = np.ascontiguousarray(prob.pts.copy())
obs_v, cam_v, K_v = prob.obs.copy(), prob.cam_idx.copy(), prob.K.copy()
R_ptrs = frames.ctypes.data + 48 * np.arange(prob.n_cams, dtype=np.uint64)
t_ptrs = R_ptrs + np.uint64(24)
pt_ptrs = store.ctypes.data + np.uint64(24) * prob.pt_idx.astype(np.uint64)
ctx.solve_ptr(obs_v, cam_v, K_v, R_ptrs, t_ptrs, pt_ptrs, opts=opts)
for r in range(reps):
    np.copyto(frames, prob.cam); np.copyto(store, prob.pts)
    t0 = time.perf_counter()
    s = ctx.solve_ptr(obs_v, cam_v, K_v, R_ptrs, t_ptrs, pt_ptrs, opts=opts)
    print(f"{cfg} solve_ptr call {r}: wall {1e3*(time.perf_counter()-t0):.2f} ms, {s['num_iterations']} it", flush=True)
