"""The schedules behind the experiment knobs are fallbacks too (no cooperative launch, no graph capture, one chain): each
must give the solve of the default schedule.  The knobs are read once per process, so every variant runs in a process of
its own."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SNIPPET = r"""
import json, sys
sys.path.insert(0, %r)
import numpy as np
from sfm_b200 import api
from tests.test_gpu_parity import _trajectory
p = _trajectory(n_cams=640, n_pts=9000, closure=True)
ctx = api.Context(0)
cam, pts, s = ctx.solve_packed(p, opts=api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR, max_num_iterations=5))
rng = np.random.default_rng(1)
A = rng.normal(size=(300, 300)); A = A @ A.T + 300 * np.eye(300); b = rng.normal(size=300)
x = ctx.dense_spd_solve(A, b)
print("RESULT " + json.dumps(dict(cost=[t["cost"] for t in s["trace"]], chains=s["chol_chains"], critical=s["chol_critical_panels"],
                                  cam=float(np.abs(cam).sum()), spd_err=float(np.abs(x - np.linalg.solve(A, b)).max()))))
""" % ROOT


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, "-c", SNIPPET], capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


@pytest.mark.parametrize("knob", ["HSFM_TRSV_LAUNCHES", "HSFM_NO_GRAPH", "HSFM_NO_LOOKAHEAD", "HSFM_ONE_CHAIN", "HSFM_SYRK_BIG_TILES", "HSFM_SCHUR_UNSORTED"])
def test_fallback_schedule_gives_the_same_solve(knob):
    ref = _run({})
    alt = _run({knob: "1"})
    assert ref["chains"] >= 3 and ref["spd_err"] < 1e-9 and alt["spd_err"] < 1e-9
    if knob in ("HSFM_ONE_CHAIN", "HSFM_NO_LOOKAHEAD"):
        assert alt["chains"] >= 1
    assert len(ref["cost"]) == len(alt["cost"])
    for a, b in zip(ref["cost"], alt["cost"]):
        assert abs(a - b) <= 1e-9 * abs(a)
    assert abs(ref["cam"] - alt["cam"]) <= 1e-7 * ref["cam"]


DEPTH_SNIPPET = r"""
import json, sys
sys.path.insert(0, %r)
from sfm_b200 import api, synth
p = synth.make_config("c3")
ctx = api.Context(0)
ctx.upload(p)
s = ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR))
print("RESULT " + json.dumps(dict(cost=[t["cost"] for t in s["trace"]], chains=s["chol_chains"], critical=s["chol_critical_panels"],
                                  panels=s["chol_panels"], final=s["final_cost"], its=s["num_iterations"])))
""" % ROOT


@pytest.mark.parametrize("depth,min_chains", [(0, 3), (1, 6), (2, 12)])
def test_every_dissection_depth_follows_the_oracle_table(depth, min_chains):
    """Full-size c3 (1723 cameras) with the order depth forced: 2, 4 and 8 concurrent chains all reproduce the oracle's
    per-iteration table (tests/golden/lm_fullsize.npz) — the segment-tree schedule on the device, not only its emulation."""
    import numpy as np

    g = np.load(os.path.join(ROOT, "tests", "golden", "lm_fullsize.npz"))
    env = dict(os.environ, HSFM_ORDER_FORCE=str(depth))
    r = subprocess.run([sys.executable, "-c", DEPTH_SNIPPET], capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][len("RESULT "):])
    assert d["chains"] >= min_chains and d["critical"] < d["panels"]
    want = g["c3_cost"]
    assert len(d["cost"]) == len(want)
    for a, b in zip(d["cost"], want):
        assert abs(a - b) <= 1e-9 * abs(b)
