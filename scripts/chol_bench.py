"""Dense Cholesky timing tap: python scripts/chol_bench.py [n ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sfm_b200 import api
ns = [int(a) for a in sys.argv[1:]] or [1200, 4096, 10668]
with api.Context(0) as ctx:
    for n in ns:
        f, s, r = ctx.bench_cholesky(n, 3)
        print(f"n={n}: factor {f:.3f} ms ({n**3/3/f*1e-9:.2f} TFLOP/s), solve {s:.3f} ms, max|x-1| {r:.2e}")
