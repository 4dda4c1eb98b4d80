"""Print the per-phase times of the one-shot host-buffer call (hsfm_ba_solve_packed) for a config."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sfm_b200 import api, synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
lin = {"auto": api.LINEAR_AUTO, "dense": api.LINEAR_DENSE_SCHUR, "pcg": api.LINEAR_PCG}[sys.argv[2] if len(sys.argv) > 2 else "auto"]
prob = synth.make_config(cfg)
pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
hp = synth.Problem(obs=pin(prob.obs), cam_idx=pin(prob.cam_idx), pt_idx=pin(prob.pt_idx), K=pin(prob.K), cam=pin(prob.cam), pts=pin(prob.pts),
                   cam_true=prob.cam_true, pts_true=prob.pts_true, name=prob.name)
ctx = api.Context(0)
opts = api.default_options(linear_solver=lin)
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    np.copyto(hp.cam, prob.cam); np.copyto(hp.pts, prob.pts)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, _, s = ctx.solve_packed(hp, opts=opts, inplace=True)
    dt = (time.perf_counter() - t0) * 1e3
    print(f"{cfg} rep {rep}: wall {dt:.1f} ms | h2d {s['h2d_ms']:.1f} setup {s['setup_ms']:.1f} solve {s['solve_ms']:.1f} d2h {s['d2h_ms']:.1f} | "
          f"its {s['num_iterations']} lin {s['total_linear_iterations']} launches {s['gpu_launches']} cost {s['final_cost']:.6f} | "
          f"phases lin/form/factor/solve/step/sync = " + "/".join(f"{v:.1f}" for v in s["phase_ms"][:6]))
ctx.close()
