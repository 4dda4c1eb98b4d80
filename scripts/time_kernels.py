"""python scripts/time_kernels.py [config] [kernel ...] — live CUDA-event timings of single kernels on a resident problem."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sfm_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
names = sys.argv[2:] or ["rj_form", "rj", "schur_matvec", "schur_rhs", "backsub", "cam_form", "cost"]
p = synth.make_config(cfg)
ctx = api.Context(0)
ctx.upload(p)
ctx.solve_resident(api.default_options(max_num_iterations=1))
peak = 6552.3
for k in names:
    t = ctx.time_kernel(k, 20); b = ctx.algorithmic_bytes(k)
    print(f"{k:14s} {t*1e3:8.1f} us  {b/t/1e6:8.1f} GB/s  {b/t/1e6/peak:.3f}")
