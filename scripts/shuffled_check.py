"""A config with its cameras renumbered at random: does the upload find the band again?  python scripts/shuffled_check.py [config]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sfm_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
p = synth.make_config(cfg)
q, _ = synth.shuffle_cameras(p, seed=1)
ctx = api.Context(0)
for name, prob in (("ordered", p), ("shuffled", q)):
    t0 = time.perf_counter(); ctx.upload(prob); t_up = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.upload(prob); t_up2 = time.perf_counter() - t0
    s = ctx.solve_resident(api.default_options())
    s = ctx.solve_resident(api.default_options()) if False else s
    print(f"{cfg} {name}: upload {1e3*t_up:.1f} ms (again, order cached: {1e3*t_up2:.1f}), solver {s['linear_solver_used']}, {s['num_iterations']} it in {s['solve_ms']:.1f} ms "
          f"= {1e3*s['num_iterations']/s['solve_ms']:.1f} it/s, final cost {s['final_cost']:.6f}, panels {s['chol_panels']} critical {s['chol_critical_panels']}, "
          f"{s['chol_flops']/1e9:.1f} GFLOP per factorisation", flush=True)
