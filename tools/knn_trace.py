"""Timeline of one CTA of vs_knn2_tc5_kernel (build with VSLAM_NVCC_FLAGS=-DT5_TRACE into tools/ubench/build/).

usage (GPU box): python tools/knn_trace.py [n] [pairs]
Prints, per train tile of CTA (0,0,0): when the producer got the TMEM buffer, committed the MMAs, when consumer
warp 0 saw the accumulators and when it finished the tile (SM cycles relative to the first stamp)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
import numpy as np
import torch
from orbslam2 import _native

_native.LIB_PATH = os.path.join(ROOT, "tools", "ubench", "build", os.environ.get("T5_TRACE_LIB", "libvslam_trace.so"))
L = _native.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
pairs = int(sys.argv[2]) if len(sys.argv) > 2 else 512
g = torch.Generator(device="cuda").manual_seed(n)
desc = torch.randint(0, 256, (pairs + 1, n, 32), dtype=torch.uint8, device="cuda", generator=g)
idx = torch.empty((pairs, n, 2), dtype=torch.int32, device="cuda")
dist = torch.empty_like(idx)
s = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    _native.check(L.vslam_bf_knn2_hamming256(desc.data_ptr(), n, n, desc.data_ptr() + n * 32, n, n, None, None,
                                             pairs, idx.data_ptr(), dist.data_ptr(), C.c_void_p(s)))
torch.cuda.synchronize()
buf = np.zeros((8, 64), np.int64)
L.vslam_debug_t5_trace.argtypes = [C.c_void_p]
assert L.vslam_debug_t5_trace(buf.ctypes.data_as(C.c_void_p)) == 0
tiles = min((n + 127) // 128, 64)
t0 = buf[:, :tiles][buf[:, :tiles] > 0].min()
print("tile  free_seen  loaded  tile_full  mma_commit  acc_seen  w0_done  all_done   period")
for i in range(tiles):
    fs, f, fu, c, a, d, ad = (int(buf[r, i] - t0) if buf[r, i] else -1 for r in (4, 3, 5, 0, 1, 2, 6))
    per = int(buf[2, i] - buf[2, i - 1]) if i else 0
    print("%4d  %9d %7d  %9d  %10d  %8d  %7d  %8d   %6d" % (i, fs, f, fu, c, a, d, ad, per))
