import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sfm_b200 import api
ctx = api.Context(0)
for n in (128, 130, 256, 384):
    rng = np.random.default_rng(n)
    A = rng.normal(size=(n, n)); A = A @ A.T + n * np.eye(n)
    b = rng.normal(size=n)
    x = ctx.dense_spd_solve(A, b)
    x0 = np.linalg.solve(A, b)
    bad = np.nonzero(~(np.abs(x - x0) < 1e-8))[0]
    print(n, "bad entries:", len(bad), bad[:8], x[bad[:4]])
