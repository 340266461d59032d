"""GPU parity of vslam_undistort (cv2.undistort, the reference's utils.undistort_image, utils.py:40-52) through
the C ABI: bit-exact against the golden cv2 outputs and against the oracle on larger seeded images."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "undistort.npz")


def test_undistort_equals_cv2_golden():
    from orbslam2 import utils
    g = np.load(GOLD)
    for n in range(int(g["n"])):
        out = utils.undistort_image(g["img%d" % n], g["K%d" % n], g["dist%d" % n])
        assert out.dtype == np.uint8 and out.shape == g["img%d" % n].shape
        assert np.array_equal(out, g["out%d" % n]), n
    img = g["img0"]
    assert np.array_equal(utils.undistort_image(img, g["K0"], np.zeros(5)), img)
    assert np.array_equal(utils.undistort_image(img, g["K0"], None), img)


@pytest.mark.parametrize("h,w,ch", [(480, 640, 1), (478, 850, 3), (1080, 1920, 1)])
def test_undistort_equals_oracle_frame_sizes(h, w, ch):
    from orbslam2 import utils
    from oracle import undistort_oracle as U
    rng = np.random.default_rng(h + w)
    img = rng.integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)
    K = np.array([[0.8 * w, 0, w / 2 + 3.3], [0, 0.82 * w, h / 2 - 2.1], [0, 0, 1]])
    dist = np.array([-0.28, 0.07, 0.0012, -0.0007, 0.01])
    got = utils.undistort_image(img, K, dist)
    assert np.array_equal(got, U.undistort(img, K, dist))
    try:
        import cv2
    except ImportError:
        return
    assert np.array_equal(got, cv2.undistort(img, K, dist))


def test_undistort_device_batch_then_extract(synth640):
    """The device entry point on a batch with a pitch, feeding the extractor without a host hop."""
    import ctypes as C
    import torch
    from orbslam2 import _native
    from oracle import undistort_oracle as U, oracle as O
    L = _native.lib()
    frames = synth640.frames(0, 2)
    K = np.array([[500.0, 0, 320], [0, 500.0, 240], [0, 0, 1]])
    dist = np.array([-0.2, 0.05, 0.001, 0.001, 0.0])
    d_src = torch.from_numpy(frames).cuda()
    d_dst = torch.zeros((2, 480, 672), dtype=torch.uint8, device="cuda")      # pitch 672
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _native.check(L.vslam_undistort(d_src.data_ptr(), 2, 480, 640, 1, 640, _native.ptr(K), _native.ptr(dist), 5,
                                    d_dst.data_ptr(), 672, s))
    hd = _native.OrbHandle(640, 480, n_features=1000, max_batch=2)
    kp = torch.empty((2, hd.kp_capacity, 6), dtype=torch.float32, device="cuda")
    desc = torch.empty((2, hd.kp_capacity, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
    _native.check(L.vslam_orb_detect_and_compute(hd.handle, d_dst.data_ptr(), 2, 672, kp.data_ptr(), desc.data_ptr(),
                                                 cnt.data_ptr(), s))
    torch.cuda.synchronize()
    und = d_dst.cpu().numpy()[:, :, :640]
    for f in range(2):
        ref_img = U.undistort(frames[f], K, dist)
        assert np.array_equal(und[f], ref_img)
        ref = O.orb_detect_and_compute(ref_img, 1000)
        n = int(cnt[f])
        assert n == len(ref["x"]) and np.array_equal(desc[f, :n].cpu().numpy(), ref["desc"])
    hd.close()
    assert L.vslam_undistort(d_src.data_ptr(), 1, 480, 640, 2, 1280, _native.ptr(K), None, 0, d_dst.data_ptr(), 1280, s) != 0
