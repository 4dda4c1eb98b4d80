"""In-situ timing of the Cholesky chain: builds a copy of the library with -DHSFM_CHAIN_TIMING (globaltimer stamps at
the start and end of every diagonal-block kernel), runs one resident solve and prints, per panel, the diagonal-block
time and the gap to the next one (panel solve of row block K+1 + next diagonal tile + launch hand-overs + waits).
    python scripts/chain_timing.py [config]"""
import ctypes as C, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sfm_b200 import build as B
out = os.path.join(ROOT, "sfm_b200", "build", "libhsfm_ba_timing.so")
objs = []
os.makedirs(os.path.dirname(out), exist_ok=True)
for src in B.SOURCES:
    obj = os.path.join(ROOT, "sfm_b200", "build", "t_" + os.path.splitext(src)[0] + ".o")
    if not os.environ.get("CHAIN_TIMING_PREBUILT"):
        subprocess.run([B.NVCC, *B.FLAGS, "-DHSFM_CHAIN_TIMING", *(["-x", "cu"] if src.endswith(".cpp") else []), "-c", os.path.join(B.CSRC, src), "-o", obj], check=True, stderr=subprocess.DEVNULL)
    objs.append(obj)
if not os.environ.get("CHAIN_TIMING_PREBUILT"):
    subprocess.run([B.NVCC, "-shared", "-o", out, *objs, "-ccbin", "/usr/bin/g++", "-lcudart", "-ldl"], check=True, stderr=subprocess.DEVNULL)
if os.environ.get("CHAIN_TIMING_BUILD_ONLY"):
    sys.exit(0)
from sfm_b200 import api, synth
api.LIB_PATH = out
import numpy as np
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
p = synth.make_config(cfg)
ctx = api.Context(0)
ctx.upload(p)
s = ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR, max_num_iterations=3))
lib = api.load_library()
panels = (6 * p.n_cams + 127) // 128
ts = np.zeros(2 * panels, dtype=np.uint64)
lib.hsfm_ba_debug_chain_ts.argtypes = [C.c_void_p, C.c_int]
assert lib.hsfm_ba_debug_chain_ts(ts.ctypes.data, 2 * panels) == 0
t = ts.astype(np.int64).reshape(-1, 2)
diag = (t[:, 1] - t[:, 0]) * 1e-3
gap = (t[1:, 0] - t[:-1, 1]) * 1e-3
print(f"{cfg}: {panels} panels; diagonal-block kernel {np.median(diag):.1f} us median ({diag.min():.1f}..{diag.max():.1f}); "
      f"gap to the next {np.median(gap):.1f} us median ({gap.min():.1f}..{gap.max():.1f}); last factorisation {1e-3 * (t[-1, 1] - t[0, 0]):.0f} us")
print("diag:", " ".join(f"{x:.0f}" for x in diag))
print("gap: ", " ".join(f"{x:.0f}" for x in gap))
