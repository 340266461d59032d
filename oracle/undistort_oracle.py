"""CPU oracle of cv2.undistort for uint8 images (the reference's utils.undistort_image, utils.py:40-52).

TEST INFRASTRUCTURE ONLY (imported by tests/ and nothing else).

The arithmetic lives in OpenCV (opencv-python-headless 4.13.0.92, not vendored by the reference); this is a
numpy restatement of cv::undistort -> initUndistortRectifyMap(CV_16SC2) -> remap(INTER_LINEAR, BORDER_CONSTANT):
stripes of min(max(1, 4096 // width), height) rows with their own inverse camera matrix (cv::invert's closed 3x3
form), the distortion model in double, a 5-bit fixed-point map and 15-bit bilinear weights (weight 32768
saturates to 32767 and the table fix-up adds 1 to the bottom-right tap).  Pinned: tests/golden/undistort.npz holds
cv2.undistort outputs produced by tests/golden/make_undistort_golden.py in the build container, and
tests/test_oracle_golden.py checks this file against them bit for bit.
"""
import numpy as np


def _inv3(S):
    d = (S[0, 0] * (S[1, 1] * S[2, 2] - S[1, 2] * S[2, 1]) - S[0, 1] * (S[1, 0] * S[2, 2] - S[1, 2] * S[2, 0])
         + S[0, 2] * (S[1, 0] * S[2, 1] - S[1, 1] * S[2, 0]))
    d = 1.0 / d
    t = np.zeros(9)
    t[0] = (S[1, 1] * S[2, 2] - S[1, 2] * S[2, 1]) * d
    t[1] = (S[0, 2] * S[2, 1] - S[0, 1] * S[2, 2]) * d
    t[2] = (S[0, 1] * S[1, 2] - S[0, 2] * S[1, 1]) * d
    t[3] = (S[1, 2] * S[2, 0] - S[1, 0] * S[2, 2]) * d
    t[4] = (S[0, 0] * S[2, 2] - S[0, 2] * S[2, 0]) * d
    t[5] = (S[0, 2] * S[1, 0] - S[0, 0] * S[1, 2]) * d
    t[6] = (S[1, 0] * S[2, 1] - S[1, 1] * S[2, 0]) * d
    t[7] = (S[0, 1] * S[2, 0] - S[0, 0] * S[2, 1]) * d
    t[8] = (S[0, 0] * S[1, 1] - S[0, 1] * S[1, 0]) * d
    return t


def undistort(src, camera_matrix, distortion):
    """uint8 [h, w] or [h, w, 3] -> the image cv2.undistort(src, camera_matrix, distortion) returns."""
    src = np.asarray(src, np.uint8)
    if src.ndim == 3:
        return np.stack([undistort(src[:, :, c], camera_matrix, distortion) for c in range(src.shape[2])], axis=2)
    h, w = src.shape
    K = np.asarray(camera_matrix, np.float64).reshape(3, 3)
    k = np.zeros(12)
    dist = np.zeros(0) if distortion is None else np.asarray(distortion, np.float64).ravel()
    k[:len(dist)] = dist
    k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4 = k
    fx, fy, u0, v0 = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    stripe = min(max(1, (1 << 12) // max(w, 1)), h)
    out = np.zeros_like(src)
    S = src.astype(np.int64)

    def fetch(yy, xx):
        ok = (xx >= 0) & (xx < w) & (yy >= 0) & (yy < h)
        return np.where(ok, S[np.clip(yy, 0, h - 1), np.clip(xx, 0, w - 1)], 0)

    for y0 in range(0, h, stripe):
        n = min(stripe, h - y0)
        Ar = np.array([[fx, 0.0, u0], [0.0, fy, v0 - y0], [0.0, 0.0, 1.0]])
        ir = _inv3(Ar)
        i = np.arange(n, dtype=np.float64)[:, None]
        j = np.arange(w, dtype=np.float64)[None, :]
        _x = i * ir[1] + ir[2] + j * ir[0]
        _y = i * ir[4] + ir[5] + j * ir[3]
        _w = i * ir[7] + ir[8] + j * ir[6]
        ww = 1.0 / _w
        x, y = _x * ww, _y * ww
        x2, y2 = x * x, y * y
        r2, _2xy = x2 + y2, 2 * x * y
        kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((k6 * r2 + k5) * r2 + k4) * r2)
        xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2) + s1 * r2 + s2 * r2 * r2
        yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy + s3 * r2 + s4 * r2 * r2
        u, v = fx * xd + u0, fy * yd + v0
        iu = np.rint(np.clip(u * 32, -2147483648.0, 2147483647.0)).astype(np.int64)
        iv = np.rint(np.clip(v * 32, -2147483648.0, 2147483647.0)).astype(np.int64)
        sx, sy = np.clip(iu >> 5, -32768, 32767), np.clip(iv >> 5, -32768, 32767)
        ax, ay = iu & 31, iv & 31
        w00, w01 = (32 - ax) * (32 - ay) * 32, ax * (32 - ay) * 32
        w10, w11 = (32 - ax) * ay * 32, ax * ay * 32
        whole = (ax == 0) & (ay == 0)
        w00, w11 = np.where(whole, 32767, w00), np.where(whole, 1, w11)
        val = (fetch(sy, sx) * w00 + fetch(sy, sx + 1) * w01 + fetch(sy + 1, sx) * w10
               + fetch(sy + 1, sx + 1) * w11 + (1 << 14)) >> 15
        out[y0:y0 + n] = np.clip(val, 0, 255).astype(np.uint8)
    return out
