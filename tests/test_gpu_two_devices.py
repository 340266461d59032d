"""One process, two GPUs: a handle belongs to the device that was current when it was created, kernel attributes,
workspaces and scratch are per device (per-process state must be per device).  Needs 2 visible GPUs."""
import numpy as np
import pytest


@pytest.mark.gpu
def test_handles_on_two_devices_in_one_process(synth640):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from orbslam2 import _native
    frames = synth640.frames(0, 3)
    h, w = frames.shape[1:]
    results = []
    handles = []
    for dev in (0, 1):
        torch.cuda.set_device(dev)
        orb = _native.OrbHandle(w, h, n_features=1500, max_batch=3)
        handles.append(orb)
        results.append(orb.frontend_host(frames, 0.75))
    # interleave: use the first handle again after the second device has been configured
    torch.cuda.set_device(0)
    again = handles[0].frontend_host(frames, 0.75)
    for a, b in ((results[0], results[1]), (results[0], again)):
        for x, y in zip(a, b):
            assert np.array_equal(np.asarray(x), np.asarray(y))
    # a handle used on the wrong device is refused, not run
    torch.cuda.set_device(1)
    with pytest.raises(_native.VslamError):
        handles[0].frontend_host(frames, 0.75)
    # the device-pointer matcher on the second device (its own attribute configuration and scratch)
    L = _native.lib()
    import ctypes as C
    g = torch.Generator(device="cuda:1").manual_seed(5)
    q = torch.randint(0, 256, (700, 32), dtype=torch.uint8, device="cuda:1", generator=g)
    t = torch.randint(0, 256, (900, 32), dtype=torch.uint8, device="cuda:1", generator=g)
    idx = torch.empty((700, 2), dtype=torch.int32, device="cuda:1")
    dist = torch.empty_like(idx)
    _native.check(L.vslam_bf_knn2_hamming256(q.data_ptr(), 700, 700, t.data_ptr(), 900, 900, None, None, 1,
                                             idx.data_ptr(), dist.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    qa, ta = q.cpu().numpy(), t.cpu().numpy()
    d = np.unpackbits(qa[:, None, :] ^ ta[None, :, :], axis=2).sum(axis=2).astype(np.int32)
    order = np.lexsort((np.arange(900)[None, :].repeat(700, 0), d), axis=1)[:, :2]
    assert np.array_equal(idx.cpu().numpy(), order.astype(np.int32))
    assert np.array_equal(dist.cpu().numpy(), np.take_along_axis(d, order, axis=1))
    for orb in handles:
        torch.cuda.set_device(0)
        orb.close()
