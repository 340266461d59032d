"""pytest configuration: the `gpu` marker, import paths, shared fixtures.

`-m "not gpu"` runs here (no GPU): oracle vs golden vectors, host logic, ABI symbols.
`-m gpu` runs on a B200: the CUDA path through the C ABI against the oracle.
Nothing in the gpu tests reads /root/reference (it does not exist on the GPU box).
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "visual-slam_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu():
    try:
        from orbslam2 import _native
        return _native.lib().vslam_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # a gpu test that cannot run must FAIL loudly on a GPU run, but when the whole
    # suite is run on a CPU box without -m it is skipped
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    sel = config.getoption("-m") or ""
    for it in items:
        if "gpu" in it.keywords and "gpu" not in sel.replace("not gpu", ""):
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def synth640():
    from orbslam2.synth import SyntheticVideo
    return SyntheticVideo(640, 480, 0)


def golden_path(name):
    return os.path.join(GOLDEN, name)


def load_golden(name):
    return np.load(golden_path(name), allow_pickle=False)


def canonical_kp_hash(x, y, size, angle, response, octave, desc):
    """sha256 of a keypoint set in canonical order (octave, y, x) -- the construction of
    tests/golden/make_fixture_golden.py: every cv2.KeyPoint field as float32 / int32 plus the descriptor bytes."""
    import hashlib
    x, y = np.asarray(x, np.float32), np.asarray(y, np.float32)
    o = np.asarray(octave, np.int32)
    order = np.lexsort((x, y, o))
    h = hashlib.sha256()
    for c in (x, y, np.asarray(size, np.float32), np.asarray(angle, np.float32), np.asarray(response, np.float32), o):
        h.update(np.ascontiguousarray(c[order]).tobytes())
    h.update(np.ascontiguousarray(np.asarray(desc, np.uint8)[order]).tobytes())
    return np.frombuffer(h.digest(), np.uint8)


def fixture_image(pair, which):
    """gray fixture image (tests/golden/fixture_images: the reference's own BGR2GRAY of its data/ images)"""
    import cv2
    img = cv2.imread(os.path.join(GOLDEN, "fixture_images", "%s_img%d.png" % (pair, which)), cv2.IMREAD_UNCHANGED)
    assert img is not None and img.ndim == 2 and img.dtype == np.uint8
    return img


def match_row_set(rows):
    return {tuple(float(v) for v in r) for r in np.asarray(rows, np.float32).reshape(-1, 7)}
