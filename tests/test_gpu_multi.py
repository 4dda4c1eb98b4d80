"""Multi-GPU parity through the CUDA/NCCL path: the point-sharded solve on 2 ranks against the
single-GPU solve and the CPU oracle.  Skips on a box with one GPU (the round's default GPU run); run with
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(n, linear, which):
    import torch

    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tests", "harness", "nccl_parity.py"), linear, which]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("NCCL_PARITY ")][-1]
    return json.loads(line[len("NCCL_PARITY "):])


@pytest.mark.parametrize("linear,which", [("dense", "c2"), ("pcg", "c2"), ("dense", "trajectory")])
def test_two_rank_solve_matches_single_gpu_and_oracle(linear, which):
    d = _run(2, linear, which)
    assert d["n_ranks"] == 2
    assert d["refused_by"] == 2  # the upload with a bad index in rank 1's shard was refused by both ranks
    assert d["iterations"] == d["iterations_single"] == d["iterations_oracle"]
    assert abs(d["final_cost"] - d["final_cost_single"]) <= 1e-10 * d["final_cost_single"]
    assert abs(d["final_cost"] - d["final_cost_oracle"]) <= 1e-6 * d["final_cost_oracle"]      # north star: cost within 1e-6 relative
    assert abs(d["rms"] - d["rms_oracle"]) <= 1e-6 * d["rms_oracle"]
    assert d["cam_spread_across_ranks"] == 0.0                                                  # replicated cameras stay bit-identical on all ranks
    assert d["cam_err_single"] < 1e-8 and d["pts_err_single"] < 1e-8
    assert d["cam_err_oracle"] < 1e-6 and d["pts_err_oracle"] < 1e-6
    for a, b in zip(d["trace_cost"], d["trace_cost_single"]):
        assert abs(a - b) <= 1e-10 * b
