"""World-size-2 (and 3) CPU test of the frame-range sharding used by bench.py --gpus N: the ranges tile the
video, the boundary descriptor block reaches the lower rank, and the union of the ranks' match lists
(computed here with the oracle's matcher; on the GPU box the same exchange feeds the CUDA matcher)
equals the unsharded run.  gloo backend, rendezvous on 127.0.0.1."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _frame_desc(g, cap=96):
    """Deterministic descriptors of global frame g: frame g+1 is a noisy permutation of frame g."""
    rng = np.random.default_rng(1000)
    d = rng.integers(0, 256, (cap, 32), dtype=np.uint8)
    for k in range(1, g + 1):
        r = np.random.default_rng(2000 + k)
        flip = r.random((cap, 32, 8)) < 0.03
        d = (d ^ np.packbits(flip, axis=2, bitorder="little")[..., 0])[r.permutation(cap)]
    return d, cap - (g % 5)          # ragged live counts


def _worker(rank, world, port, total, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    from orbslam2 import sharding
    from oracle import oracle as O
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, e = sharding.frame_range(rank, world, total)
        descs = {g: _frame_desc(g) for g in range(b, e)}
        cap = 96
        first_desc = torch.from_numpy(descs[b][0].copy())
        first_count = torch.tensor([descs[b][1]], dtype=torch.int32)
        nb_desc = torch.zeros((cap, 32), dtype=torch.uint8)
        nb_count = torch.zeros(1, dtype=torch.int32)
        reqs = sharding.boundary_exchange(dist, rank, world, first_desc, first_count, nb_desc, nb_count)
        for r in reqs:
            r.wait()
        if rank < world - 1:
            ref_d, ref_n = _frame_desc(e)
            assert int(nb_count[0]) == ref_n and np.array_equal(nb_desc.numpy(), ref_d), "boundary block mismatch"
            descs[e] = (nb_desc.numpy(), int(nb_count[0]))
        else:
            assert int(nb_count[0]) == 0
        res = {}
        for i, j in sharding.owned_pairs(rank, world, total):
            (di, ni), (dj, nj) = descs[i], descs[j]
            res[i] = O.match(di[:ni], dj[:nj], 0.75)
        # the slowest rank's time is the step time
        t = sharding.max_over_ranks(dist, world, torch.tensor([float(rank + 1)], dtype=torch.float64))
        assert float(t[0]) == float(world)
        np.save(os.path.join(out_dir, "rank%d.npy" % rank), np.array([(i, len(m), sum(a * 31 + b for a, b, _ in m)) for i, m in sorted(res.items())], np.int64))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_frame_ranges_tile_the_video():
    sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
    from orbslam2 import sharding
    for total in (1, 7, 8, 64, 513):
        for world in (1, 2, 3, 4, 8):
            rs = [sharding.frame_range(r, world, total) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == total
            assert all(rs[r][1] == rs[r + 1][0] for r in range(world - 1))
            sizes = [e - b for b, e in rs]
            assert max(sizes) - min(sizes) <= 1
            pairs = sorted(p for r in range(world) for p in sharding.owned_pairs(r, world, total))
            assert pairs == [(i, i + 1) for i in range(total - 1)], (total, world)
    with pytest.raises(ValueError):
        sharding.frame_range(2, 2, 10)


@pytest.mark.parametrize("world", [2, 3])
def test_boundary_exchange_and_sharded_matching_equal_unsharded(world, tmp_path, oracle):
    import torch.multiprocessing as mp
    total = 11
    port = _free_port()
    mp.spawn(_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
    got = np.concatenate([np.load(str(tmp_path / ("rank%d.npy" % r))).reshape(-1, 3) for r in range(world)])
    got = got[np.argsort(got[:, 0])]
    ref = []
    for i in range(total - 1):
        (di, ni), (dj, nj) = _frame_desc(i), _frame_desc(i + 1)
        m = oracle.match(di[:ni], dj[:nj], 0.75)
        ref.append((i, len(m), sum(a * 31 + b for a, b, _ in m)))
    assert np.array_equal(got, np.array(ref, np.int64))
    assert int(got[:, 1].min()) > 20          # the ratio test keeps a real match list on every pair
