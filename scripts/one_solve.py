"""One upload + resident solve with wall-clock stamps: python scripts/one_solve.py [config] [max_lm_iterations]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sfm_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
its = int(sys.argv[2]) if len(sys.argv) > 2 else 3
t0 = time.time(); p = synth.make_config(cfg); print(f"synth {time.time()-t0:.1f} s", flush=True)
ctx = api.Context(0)
t0 = time.time(); ctx.upload(p); print(f"upload {time.time()-t0:.2f} s", flush=True)
for rep in range(2):
    t0 = time.time()
    s = ctx.solve_resident(api.default_options(max_num_iterations=its))
    print(f"solve {time.time()-t0:.2f} s", {k: v for k, v in s.items() if k != "trace"}, flush=True)
