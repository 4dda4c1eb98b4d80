"""CPU suite, world_size 2 over gloo: the N>1 host logic.  Each rank owns a point shard
(synth.shard_by_point), evaluates its observations, and the camera-keyed normal-equation blocks are
summed with an all-reduce — exactly the exchange the GPU path performs with NCCL.
The reduced blocks and the cost must equal the single-rank ones."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _camera_blocks(prob, r, J):
    H = np.zeros((prob.n_cams, 6, 6))
    g = np.zeros((prob.n_cams, 6))
    Jc = J[:, :, :6]
    np.add.at(H, prob.cam_idx, np.einsum("nij,nik->njk", Jc, Jc))
    np.add.at(g, prob.cam_idx, np.einsum("nij,ni->nj", Jc, r))
    return H, g


def _worker(rank, world, port, out):
    import sys

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch

    from oracle import pyoracle as po
    from sfm_b200 import synth

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = synth.make_ring(7, 150, 5, seed=21)
    shard, gids = synth.shard_by_point(full, world)[rank]
    r, J, cost = po.evaluate(shard)
    H, g = _camera_blocks(shard, r, J)
    buf = torch.from_numpy(np.concatenate([H.reshape(-1), g.reshape(-1), [cost]]))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    # replicated camera vectors must be bit-identical on every rank after the exchange
    gathered = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(gathered, buf)
    same = all(torch.equal(gathered[0], t) for t in gathered)
    if rank == 0:
        r0, J0, cost0 = po.evaluate(full)
        H0, g0 = _camera_blocks(full, r0, J0)
        ref = np.concatenate([H0.reshape(-1), g0.reshape(-1), [cost0]])
        err = float(np.abs(buf.numpy() - ref).max() / np.abs(ref).max())
        np.save(out, np.array([err, float(same)]))
    dist.destroy_process_group()


def test_two_rank_camera_block_allreduce(tmp_path):
    out = str(tmp_path / "res.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    err, same = np.load(out)
    assert err < 1e-12 and same == 1.0
