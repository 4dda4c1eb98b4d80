"""GPU suite: the C++ drop-in CTracker::bundleAdjustmentStructAndPose (sfm_b200/host) driven from a
C++ program that assembles vector<Point2d>/vector<double*> arguments like CSfM::bundleAdjustment
(reference CSfM.cpp:311-343), compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

from oracle import pyoracle as po
from sfm_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("shim") / "ctracker_host")
    lib = os.path.join(ROOT, "sfm_b200")
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", os.path.join(ROOT, "tests", "harness", "ctracker_host_main.cpp"),
                    os.path.join(lib, "host", "CTrackerBA.cpp"), "-o", out, f"-L{lib}", "-lhsfm_ba", f"-Wl,-rpath,{lib}",
                    "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64"], check=True)
    return out


@pytest.mark.parametrize("mode", [2, 1, 0])
def test_cpp_drop_in_matches_oracle(exe, tmp_path, mode):
    p = synth.make_config("c1")
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        np.array([p.n_obs, p.n_cams, p.n_pts], dtype=np.int64).tofile(f)
        for a in (p.obs, p.cam_idx, p.pt_idx, p.K, p.cam, p.pts):
            a.tofile(f)
    subprocess.run([exe, fin, fout, str(mode)], check=True)
    out = np.fromfile(fout, dtype=np.float64)
    cam = out[: 6 * p.n_cams].reshape(-1, 6)
    pts = out[6 * p.n_cams: 6 * p.n_cams + 3 * p.n_pts].reshape(-1, 3)
    final_cost, its, term = out[-3:]
    cam0, pts0, so, _ = po.solve(p, mode=mode)
    assert int(its) == so["num_iterations"] and int(term) == so["termination_type"]
    assert final_cost == pytest.approx(so["final_cost"], rel=1e-6)
    assert np.abs(cam - cam0).max() < 1e-6 and np.abs(pts - pts0).max() < 1e-6
