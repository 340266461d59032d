"""Video ingest on the device (csrc/nvdec.cu + orbslam2/ingest.py).

Where the NVDEC engine is exposed: a short clip written with cv2 is decoded by cv2 on the CPU (the reference's
cv2.VideoCapture.read(), tester_map.py:58-66) and by NVDEC; luma planes, BGR frames and the reference's gray must be
equal bit for bit.  Where it is not (this pool's boxes run with NVIDIA_DRIVER_CAPABILITIES=compute,utility: libnvcuvid
loads, cuvidGetDecoderCaps answers CUDA_ERROR_NO_DEVICE -- profiles/r2_nvdec_probe.txt), the call has to fail loudly:
there is no CPU decode fallback behind this API."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _write_clip(path, n=8):
    import cv2
    from orbslam2.synth import SyntheticVideo
    v = SyntheticVideo(640, 480, 0)
    rng = np.random.default_rng(0)
    w = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"VP90"), 30.0, (640, 480))
    assert w.isOpened()
    for k in range(n):
        bgr = cv2.cvtColor(v.frame(k), cv2.COLOR_GRAY2BGR).astype(np.int16)
        tint = rng.integers(-60, 61, (15, 20, 3)).astype(np.int16)
        w.write(np.clip(bgr + np.kron(tint, np.ones((32, 32, 1), np.int16)), 0, 255).astype(np.uint8))
    w.release()


def test_nvdec_ingest_equals_cv2_or_fails_loudly(tmp_path):
    import cv2
    from orbslam2 import _native, ingest
    path = str(tmp_path / "clip.webm")
    _write_clip(path)
    if not ingest.available():
        with pytest.raises(_native.VslamError) as e:
            ingest.VideoReader(path, mode="gray")
        assert "NVDEC" in str(e.value) and "fallback" in str(e.value)
        return
    cap = cv2.VideoCapture(path, cv2.CAP_FFMPEG)
    ref = []
    while True:
        ok, f = cap.read()
        if not ok:
            break
        ref.append(f)
    cap.release()
    for mode in ("bgr", "gray"):
        rd = ingest.VideoReader(path, mode=mode)
        got = [f.cpu().numpy() for f in rd]
        rd.close()
        assert len(got) == len(ref)
        for a, b in zip(got, ref):
            want = b if mode == "bgr" else cv2.cvtColor(b, cv2.COLOR_BGR2GRAY)
            assert np.array_equal(a, want), mode
