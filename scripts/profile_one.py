"""One resident solve for profiling: python scripts/profile_one.py [config] [dense|pcg] [max_lm_iterations]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sfm_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
lin = {"dense": api.LINEAR_DENSE_SCHUR, "pcg": api.LINEAR_PCG}[sys.argv[2] if len(sys.argv) > 2 else "dense"]
its = int(sys.argv[3]) if len(sys.argv) > 3 else 2
p = synth.make_config(cfg)
ctx = api.Context(0)
ctx.upload(p)
s = ctx.solve_resident(api.default_options(linear_solver=lin, max_num_iterations=its, pcg_max_iterations=30))
print({k: v for k, v in s.items() if k != "trace"})
