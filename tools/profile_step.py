"""One device-resident step of the front end (extract B frames + match), for ncu / launch-list captures.

usage: python tools/profile_step.py [--workload c2] [--batch 512] [--steps 2] [--ransac] [--track]
Runs `steps` calls of vslam_frontend_run on synthetic frames and nothing else (no CPU leg, no e2e), so that
`ncu -k regex:... -s <skip> -c <n>` lands on the launches of a warm step.
"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    import numpy as np
    import torch
    import bench as BN
    from orbslam2 import _native
    from orbslam2.synth import SyntheticVideo
    w, h, nf, B0 = BN.WORKLOADS[args.workload]
    B = args.batch or B0
    L = _native.lib()
    v = SyntheticVideo(w, h, 0)
    base = v.frames(0, min(B, 64))
    frames = torch.from_numpy(np.concatenate([base] * ((B + len(base) - 1) // len(base)))[:B].copy()).cuda()
    orb = _native.OrbHandle(w, h, n_features=nf, max_batch=B)
    cap = orb.kp_capacity
    kp = torch.empty((B, cap, 6), dtype=torch.float32, device="cuda")
    desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    count = torch.zeros(B, dtype=torch.int32, device="cuda")
    match = torch.empty((B, cap, 3), dtype=torch.int32, device="cuda")
    mcount = torch.zeros(B, dtype=torch.int32, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for i in range(args.steps):
        _native.check(L.vslam_frontend_run(orb.handle, frames.data_ptr(), B, w, None, None, 0.75, kp.data_ptr(),
                                           desc.data_ptr(), count.data_ptr(), match.data_ptr(), mcount.data_ptr(), s))
        ev[i + 1].record()
    torch.cuda.synchronize()
    print("ms per step:", [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(args.steps)],
          "kp", int(count.sum()), "matches", int(mcount[:B - 1].sum()))


if __name__ == "__main__":
    main()
