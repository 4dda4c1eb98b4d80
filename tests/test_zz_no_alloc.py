"""hsfm_ba_cost_noalloc — the allocation-free entry point (static device scratch; include/hsfm_ba.h) — against the
existing hsfm_ba_eval tap on the same inputs, two calls in a row.  Collected last on purpose.
The test owns the device buffers (torch tensors); the entry point itself allocates nothing."""
import numpy as np
import pytest

from sfm_b200 import api, synth

pytestmark = pytest.mark.gpu


def _device_args(torch, p, cam, pts):
    dev = torch.device("cuda", 0)
    t = dict(obs=torch.from_numpy(p.obs).to(dev), cam_idx=torch.from_numpy(p.cam_idx).to(dev), pt_idx=torch.from_numpy(p.pt_idx).to(dev),
             K=torch.from_numpy(p.K).to(dev), cam=torch.from_numpy(np.ascontiguousarray(cam)).to(dev),
             pts=torch.from_numpy(np.ascontiguousarray(pts)).to(dev), cost=torch.full((1,), -1.0, dtype=torch.float64, device=dev))
    return t


def test_cost_noalloc_matches_eval_twice_in_a_row(ctx):
    import torch

    big = synth.make_config("c2", scale=0.2)      # 120k observations: 469 partial sums in the static scratch array
    small = synth.make_ring(7, 50, 4, seed=1)     # 200 observations: ONE partial; the array still holds the first call's 469
    ta = _device_args(torch, big, big.cam, big.pts)
    tb = _device_args(torch, small, small.cam_true, small.pts_true)
    torch.cuda.synchronize()
    for t, p in ((ta, big), (tb, small)):          # two calls in a row, no reset of anything in between
        ctx.cost_noalloc(p.n_obs, t["obs"].data_ptr(), t["cam_idx"].data_ptr(), t["pt_idx"].data_ptr(), p.n_cams, t["K"].data_ptr(),
                         t["cam"].data_ptr(), p.n_pts, t["pts"].data_ptr(), t["cost"].data_ptr())
    torch.cuda.synchronize()
    want_a = ctx.evaluate(big, want_jac=False)[2]
    want_b = ctx.evaluate(small, cam=small.cam_true, pts=small.pts_true, want_jac=False)[2]
    assert float(ta["cost"].item()) == pytest.approx(want_a, rel=1e-12)   # the tolerance the cost comparisons of the suite use
    assert float(tb["cost"].item()) == pytest.approx(want_b, rel=1e-12)
    # and once more on the first problem: same answer bit for bit (fixed summation order, nothing stale is read)
    first = float(ta["cost"].item())
    ctx.cost_noalloc(big.n_obs, ta["obs"].data_ptr(), ta["cam_idx"].data_ptr(), ta["pt_idx"].data_ptr(), big.n_cams, ta["K"].data_ptr(),
                     ta["cam"].data_ptr(), big.n_pts, ta["pts"].data_ptr(), ta["cost"].data_ptr())
    torch.cuda.synchronize()
    assert float(ta["cost"].item()) == first


def test_cost_noalloc_rejects_sizes_beyond_the_static_scratch(ctx):
    import torch

    one = torch.full((16,), 5.0, dtype=torch.float64, device="cuda:0")
    with pytest.raises(api.BAError) as e:
        ctx.cost_noalloc(api.NOALLOC_MAX_OBS + 1, one.data_ptr(), one.data_ptr(), one.data_ptr(), 1, one.data_ptr(), one.data_ptr(), 1,
                         one.data_ptr(), one.data_ptr())
    assert e.value.code == -4 and "HSFM_BA_NOALLOC_MAX_OBS" in str(e.value)
    # empty problem: cost 0, nothing read
    ctx.cost_noalloc(0, 0, 0, 0, 0, 0, 0, 0, 0, one.data_ptr())
    torch.cuda.synchronize()
    assert float(one[0].item()) == 0.0
