"""BASELINE.json config 4: batched BF-Hamming kNN-2 sweep, 2k x 2k to 64k x 64k 256-bit descriptors.

usage (GPU box): python tools/knn_sweep.py [--engine tcgen05|popc] [--popc-peak FILE] > gpurun_out/knn_sweep.json

Prints one JSON line per size: descriptor pairs per second through vslam_bf_knn2_hamming256 (device-resident
descriptors, CUDA events, best of 5 after 2 warm-ups) and the ratio to the POPC-pipe roofline the reference
arithmetic (8 x XOR+POPC per pair) would be bound by: the MEASURED rate of tools/ubench/popc.cu (--popc-peak: the JSON
line it printed on the same box; `pairs256_per_s_at_8_popc`), else the paper figure 16 lanes/clk/SM x 148 SMs x SM clock / 8.
The batch is sized so that every launch covers ~4e9 pairs (one pair of frames at 64k, 1024 pairs at 2k).
"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--engine", default="tcgen05", choices=["tcgen05", "popc"])
    ap.add_argument("--popc-peak", default=None, help="JSON line printed by tools/ubench/popc.cu on this box")
    args = ap.parse_args()
    if args.engine != "tcgen05":
        os.environ["VSLAM_KNN"] = args.engine
    import torch
    from orbslam2 import _native
    L = _native.lib()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sm_hz = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
    popc_pairs_peak, peak_src = 16 * 148 * sm_hz / 8.0, "paper: 16 lanes/clk/SM"
    if args.popc_peak:
        pk = json.loads(open(args.popc_peak).read().strip().splitlines()[-1])
        popc_pairs_peak, peak_src = float(pk["pairs256_per_s_at_8_popc"]), "measured: tools/ubench/popc.cu"
    s = torch.cuda.current_stream().cuda_stream
    for n in (2000, 4000, 8000, 16000, 32000, 65536):
        pairs = max(1, min(1024, int(4.3e9 // (n * n))))
        g = torch.Generator(device="cuda").manual_seed(n)
        desc = torch.randint(0, 256, (pairs + 1, n, 32), dtype=torch.uint8, device="cuda", generator=g)
        idx = torch.empty((pairs, n, 2), dtype=torch.int32, device="cuda")
        dist = torch.empty_like(idx)

        def run():
            _native.check(L.vslam_bf_knn2_hamming256(desc.data_ptr(), n, n, desc.data_ptr() + n * 32, n, n, None, None,
                                                     pairs, idx.data_ptr(), dist.data_ptr(), C.c_void_p(s)))
        run(); run()
        best = 1e30
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(); run(); e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        pps = pairs * float(n) * n / (best * 1e-3)
        print(json.dumps({"n": n, "frame_pairs": pairs, "ms": round(best, 4), "gpairs_per_s": round(pps / 1e9, 1),
                          "engine": args.engine, "popc_pipe_peak_gpairs_per_s": round(popc_pairs_peak / 1e9, 1),
                          "popc_peak_source": peak_src,
                          "vs_popc_pipe_peak": round(pps / popc_pairs_peak, 3),
                          "u8_tmacs_per_s": round(pps * 256 / 1e12, 1)}), flush=True)


if __name__ == "__main__":
    main()
