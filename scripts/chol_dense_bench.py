"""Dense Cholesky tap (hsfm_ba_bench_cholesky): factor time and DMMA fraction at a few orders.  python scripts/chol_dense_bench.py [n ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sfm_b200 import api
ctx = api.Context(0)
for n in [int(a) for a in sys.argv[1:]] or [4096, 10668]:
    f, s, r = ctx.bench_cholesky(n, 3)
    print(f"n={n}: factor {f:.3f} ms = {n**3/3/f/1e9:.2f} TFLOP/s ({n**3/3/f/1e9/37.1:.2f} of DMMA peak), solve {s:.3f} ms, residual {r:.2e}")
