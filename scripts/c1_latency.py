"""Per-call latency of the small-object-scan problem (c1: 30 keyframes x 2k points x 40k observations — what hulop/SfM
solves at every keyframe): wall clock of upload / solve / download against the device time of the solve."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sfm_b200 import api, synth
p = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "c1")
ctx = api.Context(0)
o = api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR)
for _ in range(5):
    ctx.solve_packed(p, opts=o)
N = 50
t_up = t_solve = t_down = dev = 0.0
its = launches = 0
for _ in range(N):
    t0 = time.perf_counter(); ctx.upload(p); t1 = time.perf_counter()
    s = ctx.solve_resident(o); t2 = time.perf_counter()
    ctx.download(); t3 = time.perf_counter()
    t_up += t1 - t0; t_solve += t2 - t1; t_down += t3 - t2; dev += s["solve_ms"]; its += s["num_iterations"]; launches += s["gpu_launches"]
print(f"per call: upload+index {1e3*t_up/N:.3f} ms, solve wall {1e3*t_solve/N:.3f} ms (device {dev/N:.3f} ms), download {1e3*t_down/N:.3f} ms; "
      f"{its/N:.1f} LM iterations, {launches/N:.0f} launches per solve")
print("phases ms per solve:", [round(x, 3) for x in s["phase_ms"][:6]])
