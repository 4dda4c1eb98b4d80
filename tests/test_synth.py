"""CPU suite: synthetic generator and point sharding (§8e) host logic."""
import numpy as np

from sfm_b200 import synth


def test_c1_shape_and_layout():
    p = synth.make_config("c1")
    assert (p.n_cams, p.n_pts, p.n_obs) == (30, 2000, 40000)
    assert np.all(p.cam[0, :3] == 0.0)  # keyframe 0 at identity rotation exactly (CFrame.cpp:229-235)
    assert np.all(np.diff(p.cam_idx) >= 0)  # camera-major delivery (CSfM.cpp:321-341)
    assert np.all(np.bincount(p.pt_idx) == 20)
    assert p.obs.dtype == np.float64 and p.cam_idx.dtype == np.int32 and p.obs.flags.c_contiguous
    q = synth.make_config("c1")
    assert np.array_equal(p.obs, q.obs) and np.array_equal(p.cam, q.cam)  # seeded


def test_path_config_exact_counts():
    p = synth.make_config("c3", scale=0.05)
    cnt = np.bincount(p.pt_idx, minlength=p.n_pts)
    assert cnt.min() >= 2 and abs(cnt.mean() - 678718 / 156502) < 0.01
    uv, z = synth.project(p.cam_true, p.K, p.pts_true, p.cam_idx, p.pt_idx)
    assert z.min() > 0.1 and uv.min() >= 0 and uv[:, 0].max() < synth.WIDTH and uv[:, 1].max() < synth.HEIGHT


def test_shard_by_point_is_a_partition():
    p = synth.make_config("c3", scale=0.03)
    for n in (2, 3, 8):
        shards = synth.shard_by_point(p, n)
        gids = np.concatenate([g for _, g in shards])
        assert np.array_equal(np.sort(gids), np.arange(p.n_pts))
        assert sum(s.n_obs for s, _ in shards) == p.n_obs
        sizes = [s.n_obs for s, _ in shards]
        assert max(sizes) - min(sizes) <= 0.02 * p.n_obs / n + 64  # balanced by observation count
        for s, g in shards:
            assert np.array_equal(s.pts, p.pts[g]) and s.n_cams == p.n_cams
            cnt_local = np.bincount(s.pt_idx, minlength=s.n_pts)
            assert np.array_equal(cnt_local, np.bincount(p.pt_idx, minlength=p.n_pts)[g])  # whole tracks
