"""Device-pointer forms of the single-shot calls (vslam_orb_compute, vslam_gftt_grid,
vslam_score_fundamental / _essential, vslam_ransac_essential, vslam_recover_pose, vslam_triangulate): points,
frames and masks stay on the device, small matrices are host pointers.  Each must give exactly what its *_host
form gives (the host forms are wrappers over them) -- and those are pinned against the reference's goldens in
tests/test_gpu_geometry.py / test_gpu_dropin.py."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu

K_GERARD = np.array([796.7, 0.0, 239.0, 0.0, 787.8, 425.0, 0.0, 0.0, 1.0]).reshape(3, 3)


def _lib():
    from orbslam2 import _native
    _native.require_gpu()
    L = _native.lib()
    vp, i32, f64 = C.c_void_p, C.c_int, C.c_double
    L.vslam_score_fundamental.argtypes = [vp, vp, vp, i32, f64, vp, vp, vp]
    L.vslam_score_essential.argtypes = [vp, vp, vp, vp, i32, f64, vp, vp, vp]
    L.vslam_ransac_essential.argtypes = [vp, vp, i32, vp, f64, f64, i32, C.c_uint64, vp, vp, vp, vp]
    L.vslam_recover_pose.argtypes = [vp, vp, vp, i32, vp, vp, vp, vp, vp, vp]
    L.vslam_triangulate.argtypes = [vp, vp, vp, vp, i32, vp, vp]
    L.vslam_gftt_grid.argtypes = [vp, i32, i32, i32, i32, i32, f64, f64, vp, vp, vp, vp]
    L.vslam_orb_compute.argtypes = [vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, vp]
    return L


def test_geometry_device_forms_equal_host_forms():
    import torch
    from orbslam2 import _native, utils
    L = _lib()
    g = load_golden("pair01.npz")
    p1, p2 = g["pair01_p1"], g["pair01_p2"]
    n = len(p1)
    d1, d2 = torch.from_numpy(p1).cuda(), torch.from_numpy(p2).cuda()
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    P = lambda t: C.c_void_p(t.data_ptr())
    K = np.ascontiguousarray(K_GERARD)
    # scoring
    F = np.ascontiguousarray(g["pair01_F"])
    mask = torch.zeros(n, dtype=torch.uint8, device="cuda")
    cnt = C.c_int()
    _native.check(L.vslam_score_fundamental(_native.ptr(F), P(d1), P(d2), n, 3.0, P(mask), C.byref(cnt), s))
    assert np.array_equal(mask.cpu().numpy(), g["pair01_Fmask"]) and cnt.value == int(g["pair01_Fmask"].sum())
    E = np.ascontiguousarray(g["pair01_E_1"])
    _native.check(L.vslam_score_essential(_native.ptr(E), _native.ptr(K), P(d1), P(d2), n, 1.0, P(mask), C.byref(cnt), s))
    assert np.array_equal(mask.cpu().numpy(), g["pair01_Emask_1"]) and cnt.value == int(g["pair01_Emask_1"].sum())
    # E-RANSAC: mask equals the reference's, E equals the host form's bit for bit
    Eo = np.zeros(9)
    _native.check(L.vslam_ransac_essential(P(d1), P(d2), n, _native.ptr(K), 0.999, 3.0, 1000, 0, _native.ptr(Eo), P(mask),
                                           C.byref(cnt), s))
    Eh, mh = utils.calculate_essential_matrix(p1, p2, K_GERARD, threshold=3.0)
    assert np.array_equal(Eo.reshape(3, 3), Eh) and np.array_equal(mask.cpu().numpy(), mh.ravel())
    assert np.array_equal(mask.cpu().numpy(), g["pair01_Emask_3"]) and cnt.value == int(mh.sum())
    # recoverPose on the device mask
    R, t, good = np.zeros(9), np.zeros(3), C.c_int()
    _native.check(L.vslam_recover_pose(_native.ptr(Eo), P(d1), P(d2), n, _native.ptr(K), P(mask), _native.ptr(R),
                                       _native.ptr(t), C.byref(good), s))
    torch.cuda.synchronize()
    nh, Rh, th, pm = utils.recover_pose(Eh, p1, p2, K_GERARD, mh)
    assert good.value == nh and np.array_equal(R.reshape(3, 3), Rh) and np.array_equal(t.reshape(3, 1), th)
    assert np.array_equal(mask.cpu().numpy(), pm.ravel())
    assert good.value == int(g["pair01_pose_n_3"]) and np.array_equal(pm.ravel() != 0, g["pair01_pose_mask_3"] != 0)
    # triangulation
    sel = pm.ravel() != 0
    q1, q2 = torch.from_numpy(p1[sel]).cuda(), torch.from_numpy(p2[sel]).cuda()
    m = int(sel.sum())
    X = torch.zeros((4, m), dtype=torch.float32, device="cuda")
    P1, P2 = np.ascontiguousarray(g["pair01_tri_P1"]), np.ascontiguousarray(g["pair01_tri_P2"])
    _native.check(L.vslam_triangulate(_native.ptr(P1), _native.ptr(P2), P(q1), P(q2), m, P(X), s))
    Xh = utils.triangulate_points(p1[sel], p2[sel], P1, P2)
    assert np.array_equal(X.cpu().numpy(), Xh)


def test_gftt_and_compute_device_forms_equal_host_forms(synth640):
    import torch
    from orbslam2 import _native
    L = _lib()
    img = synth640.frame(3)
    h, w = img.shape
    d_img = torch.from_numpy(img).cuda()
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    P = lambda t: C.c_void_p(t.data_ptr())
    per = 2000 // 64
    xy = torch.zeros((64, per, 2), dtype=torch.float32, device="cuda")
    cc = torch.zeros(64, dtype=torch.int32, device="cuda")
    st = torch.zeros(1, dtype=torch.int32, device="cuda")
    _native.check(L.vslam_gftt_grid(P(d_img), h, w, 8, 8, per, 0.01, 10.0, P(xy), P(cc), P(st), s))
    torch.cuda.synchronize()
    h_xy = np.zeros((64 * per, 2), np.float32)
    h_cc = np.zeros(64, np.int32)
    n_out = C.c_int()
    _native.check(L.vslam_gftt_grid_host(_native.ptr(img), h, w, 8, 8, per, 0.01, 10.0, _native.ptr(h_xy), _native.ptr(h_cc),
                                         C.byref(n_out)))
    cc_np, xy_np = cc.cpu().numpy(), xy.cpu().numpy()
    assert int(st.item()) == 0 and np.array_equal(cc_np, h_cc) and n_out.value == int(h_cc.sum()) > 500
    flat = np.concatenate([xy_np[c, :cc_np[c]] for c in range(64)])
    assert np.array_equal(flat, h_xy[:n_out.value])
    # descriptors at those corners (angle -1, octave 0, like distribute_keypoints) + a few out-of-border / bad-octave rows
    x = np.concatenate([flat[:, 0], [5.0, 320.0, 100.0]]).astype(np.float32)
    y = np.concatenate([flat[:, 1], [100.0, 470.0, 100.0]]).astype(np.float32)
    ang = np.full(len(x), -1.0, np.float32)
    octv = np.zeros(len(x), np.int32)
    octv[-1] = 9
    octv[5] = 2
    ang[7] = 123.5
    orb = _native.OrbHandle(w, h, n_features=2000, max_batch=1)
    dx, dy, da, do_ = (torch.from_numpy(a).cuda() for a in (x, y, ang, octv))
    desc = torch.zeros((len(x), 32), dtype=torch.uint8, device="cuda")
    keep = torch.zeros(len(x), dtype=torch.uint8, device="cuda")
    _native.check(L.vslam_orb_compute(orb.handle, P(d_img), w, P(dx), P(dy), P(da), P(do_), len(x), P(desc), P(keep), s))
    torch.cuda.synchronize()
    keep_np, desc_np = keep.cpu().numpy(), desc.cpu().numpy()
    hd = np.zeros((len(x), 32), np.uint8)
    hk = np.zeros(len(x), np.int32)
    m = C.c_int()
    octv_ok = octv.copy()
    octv_ok[-1] = 0                      # the host form rejects a bad octave outright; compare on the valid rows
    _native.check(L.vslam_orb_compute_host(orb.handle, _native.ptr(img), _native.ptr(x), _native.ptr(y), _native.ptr(ang),
                                           _native.ptr(octv_ok), len(x), _native.ptr(hd), _native.ptr(hk), C.byref(m)))
    assert keep_np[-1] == 0 and keep_np[-3] == 0 and keep_np[-2] == 0 and not desc_np[-1].any()
    kept_host = hk[:m.value]
    for row, i in enumerate(kept_host):
        if i == len(x) - 1:
            continue
        assert keep_np[i] == 1 and np.array_equal(desc_np[i], hd[row]), i
    assert int(keep_np.sum()) == m.value - (1 if (len(x) - 1) in kept_host else 0)
    orb.close()
