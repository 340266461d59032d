"""GPU parity of brute-force Hamming kNN-2 + ratio filter (C ABI) against the oracle.

Bar: bit-exact distances and indices given identical descriptors (BASELINE.json).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rand_desc(n, seed):
    return np.random.default_rng(seed).integers(0, 256, (n, 32), dtype=np.uint8)


@pytest.mark.parametrize("nq,nt", [(1, 1), (1, 2), (5, 1), (127, 129), (128, 128), (2000, 2000),
                                   (2049, 777), (300, 5000), (8192, 8192), (257, 63), (33, 65),
                                   (16000, 16000), (70, 65535), (70, 65536),   # 16-bit train index limit
                                   (40, 65537), (40, 70000)])   # nt > 65536 takes the scalar POPC kernel
def test_knn2_random_exact(oracle, nq, nt):
    from orbslam2 import _native
    q, t = _rand_desc(nq, nq), _rand_desc(nt, nt + 1)
    idx, dist = _native.bf_knn2_host(q, t)
    ridx, rdist = oracle.bf_knn2(q, t)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(dist, rdist)


def test_knn2_ties_lowest_train_index(oracle):
    from orbslam2 import _native
    t = _rand_desc(700, 5)
    t[[2, 5, 7, 300, 650]] = t[2]          # duplicates: first and second place tie
    q = np.concatenate([t[2:3], t[650:651], _rand_desc(10, 9)])
    q[5] = 0
    t[10] = 0
    t[400] = 0
    idx, dist = _native.bf_knn2_host(q, t)
    ridx, rdist = oracle.bf_knn2(q, t)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    assert idx[0].tolist() == [2, 5] and dist[0].tolist() == [0, 0]
    assert idx[5].tolist() == [10, 400]
    # low-entropy descriptors: many equal distances
    q = (_rand_desc(500, 1) & 1) * 255
    t = (_rand_desc(3000, 2) & 1) * 255
    idx, dist = _native.bf_knn2_host(q, t)
    ridx, rdist = oracle.bf_knn2(q, t)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)


def test_knn2_empty_train_and_masked(oracle):
    from orbslam2 import _native
    q = _rand_desc(10, 3)
    idx, dist = _native.bf_knn2_host(q, np.zeros((0, 32), np.uint8))
    assert (idx == -1).all() and (dist == -1).all()
    t = _rand_desc(257, 4)
    rng = np.random.default_rng(0)
    mask = (rng.random((10, 257)) < 0.3).astype(np.uint8)
    mask[3] = 0
    for m in (None, mask):
        i1, d1 = _native.bf_match_masked_host(q, t, m)
        r1, rd1 = oracle.bf_match_masked(q, t, m)
        assert np.array_equal(i1, r1) and np.array_equal(d1, rd1)


def test_ratio_filter_device_matches_oracle(oracle):
    """Batched device path: knn2 + ratio over several pairs with per-pair live counts."""
    import torch
    from orbslam2 import _native
    L = _native.lib()
    pairs, cap = 3, 600
    rng = np.random.default_rng(7)
    desc = rng.integers(0, 256, (pairs + 1, cap, 32), dtype=np.uint8)
    # make frame i+1 a noisy copy of frame i so that the ratio test keeps many rows
    for i in range(1, pairs + 1):
        flip = (rng.random((cap, 32, 8)) < 0.04)
        desc[i] = desc[i - 1] ^ np.packbits(flip, axis=2, bitorder="little")[..., 0]
        desc[i] = desc[i][rng.permutation(cap)]
    counts = np.array([600, 555, 1, 432], np.int32)
    d_desc = torch.from_numpy(desc).cuda()
    d_cnt = torch.from_numpy(counts).cuda()
    d_idx = torch.empty((pairs, cap, 2), dtype=torch.int32, device="cuda")
    d_dist = torch.empty_like(d_idx)
    d_match = torch.empty((pairs, cap, 3), dtype=torch.int32, device="cuda")
    d_mc = torch.empty(pairs, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for ratio in (0.75, 0.8, 0.85):
        _native.check(L.vslam_bf_knn2_hamming256(d_desc.data_ptr(), cap, cap, d_desc.data_ptr() + cap * 32, cap, cap,
                                                 d_cnt.data_ptr(), d_cnt.data_ptr() + 4, pairs,
                                                 d_idx.data_ptr(), d_dist.data_ptr(), C.c_void_p(s)))
        _native.check(L.vslam_ratio_filter(d_idx.data_ptr(), d_dist.data_ptr(), cap, d_cnt.data_ptr(), pairs,
                                           ratio, 1, d_match.data_ptr(), d_mc.data_ptr(), C.c_void_p(s)))
        torch.cuda.synchronize()
        match, mc = d_match.cpu().numpy(), d_mc.cpu().numpy()
        for p in range(pairs):
            ref = oracle.match(desc[p][:counts[p]], desc[p + 1][:counts[p + 1]], ratio)
            got = [(int(a), int(b), float(c)) for a, b, c in match[p][:mc[p]]]
            assert got == ref, (p, ratio, len(got), len(ref))


def test_frontend_host_chunked_pipeline_equals_unfused(oracle):
    """vslam_frontend_run_host splits the batch into chunks that overlap copies and compute; every
    row must equal what separate extract calls + the oracle's matcher give, chunk boundaries included."""
    from orbslam2._native import OrbHandle
    from orbslam2.synth import SyntheticVideo
    v = SyntheticVideo(320, 240, 3)
    frames = v.frames(0, 40)                                  # 40 frames -> 2 chunks of 20
    h = OrbHandle(320, 240, n_features=500, max_batch=40)
    kp, desc, cnt, match, mc = h.frontend_host(frames, 0.75)
    h1 = OrbHandle(320, 240, n_features=500, max_batch=1)
    for f in (0, 19, 20, 21, 39):
        k1, d1, c1 = h1.detect_and_compute_host(frames[f])
        n = int(c1[0])
        assert n == int(cnt[f]) and np.array_equal(d1[0][:n], desc[f][:n]) and k1[0][:n].tobytes() == kp[f][:n].tobytes()
    for p in (0, 18, 19, 20, 38):                             # 19 = the pair across the chunk boundary
        ref = oracle.match(desc[p][:cnt[p]], desc[p + 1][:cnt[p + 1]], 0.75)
        got = [(int(a), int(b), float(c)) for a, b, c in match[p][:mc[p]]]
        assert got == ref, (p, len(got), len(ref))
    # a batch too small to split takes the same path with one chunk
    kp2, desc2, cnt2, match2, mc2 = h.frontend_host(frames[:3], 0.75)
    assert np.array_equal(cnt2, cnt[:3]) and np.array_equal(mc2, mc[:2])
    assert np.array_equal(match2[0][:mc2[0]], match[0][:mc[0]])
    h.close(); h1.close()


@pytest.mark.parametrize("engine", ["popc"])
def test_knn2_popc_engine_still_exact(tmp_path, engine):
    """The kNN-2 engine is the tcgen05 kernel up to 65 536 train rows and the XOR + POPC kernel beyond; VSLAM_KNN=popc
    (read once per process) forces the POPC kernel -- the formulation north_star names -- at every size, and it must
    give the same bit-exact answers."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from orbslam2 import _native\n"
        "from oracle import oracle as O\n"
        "for nq, nt in ((1, 1), (257, 63), (2000, 2000), (300, 5000)):\n"
        "    q = np.random.default_rng(nq).integers(0, 256, (nq, 32), dtype=np.uint8)\n"
        "    t = np.random.default_rng(nt + 1).integers(0, 256, (nt, 32), dtype=np.uint8)\n"
        "    i, d = _native.bf_knn2_host(q, t); ri, rd = O.bf_knn2(q, t)\n"
        "    assert np.array_equal(i, ri) and np.array_equal(d, rd), (nq, nt)\n"
        "print('ok')\n" % (os.path.join(root, "visual-slam_b200"), root))
    env = dict(os.environ, VSLAM_KNN=engine)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]
