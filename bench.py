#!/usr/bin/env python3
"""bench.py -- frames/sec of the ORB front end (extract + BF-Hamming kNN-2 + ratio).

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`,
launched under torch.distributed.run for N > 1; rank 0 prints ONE JSON line.

Workload (BASELINE.json config 2, "C2"): synthetic 640x480 gray video, 2000 ORB features,
8 levels, scale 1.2, FAST threshold 7, BF-Hamming kNN-2, Lowe ratio 0.75.  One *step* is one
pass of the hot path over one batch of B consecutive frames per GPU: extract the B frames,
match every frame against its predecessor (the predecessor of frame 0 is the previous
step's last frame, whose descriptors stay on the device) -> B extractions + B matches.
`value` = frames/sec over all ranks with the frames already resident in HBM;
`e2e` = the same call through the host-buffer C ABI (pinned host frames in, keypoints,
descriptors and matches out, copies inside the timed region).

Multi-GPU: frames shard by contiguous range (weak scaling, B*K frames per rank); the only
exchange is one boundary descriptor block per rank boundary (rank r+1 sends its first
frame's descriptors to rank r, which matches its last frame against them) over NCCL P2P.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "visual-slam_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (width, height, n_features, default batch)
    "c2": (640, 480, 2000, 512),
    "c3": (1920, 1080, 8000, 64),
    "c5": (3840, 2160, 16000, 16),
}
N_LEVELS, SCALE, FAST_THR, RATIO = 8, 1.2, 7, 0.75
METRIC = "frames/sec ORB extract+BF-Hamming match, 640x480 2k feats"
PROF_NAMES = {16: "harris", 17: "select", 18: "describe", 19: "knn2", 20: "ratio"}


def level_sizes(w, h):
    out = []
    for l in range(N_LEVELS):
        s = np.float32(np.float64(np.float32(SCALE)) ** l)
        out.append((int(np.rint(np.float32(w) / s)), int(np.rint(np.float32(h) / s))))
    return out


def level_kernel_bytes(w, h):
    """Algorithmic HBM bytes of the fused level kernel per frame:
    sum_{l>=1} (P_{l-1} + P_l) [resize: read parent, write child] + sum P [FAST reads the
    level] + 2 sum P [blur: read level, write blurred level] -- SURVEY.md 8(d)."""
    px = [a * b for a, b in level_sizes(w, h)]
    return sum(px[l - 1] + px[l] for l in range(1, N_LEVELS)) + 3 * sum(px)


def extraction_bytes(w, h, n):
    """B_ext of SURVEY.md 8(d): level kernel bytes + per-keypoint gathers and outputs."""
    return level_kernel_bytes(w, h) + 2 * n * 81 + n * 749 + n * 512 + n * 56


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------
# CPU arm: the reference's own path.  The reference is pure Python over OpenCV
# (extractor.py:38-65, matcher.py:29,70-81); its .py files are not on the GPU box, so the
# three cv2 calls it makes are issued here verbatim with cv2's own thread pool (kind
# "reference").  Without cv2 the C oracle port is timed on all cores (kind "port").
def _cv2_front_end(frames, n_features):
    import cv2
    orb = cv2.ORB_create(nfeatures=n_features, scaleFactor=SCALE, nlevels=N_LEVELS, edgeThreshold=31,
                         firstLevel=0, WTA_K=2, scoreType=cv2.ORB_HARRIS_SCORE, patchSize=31,
                         fastThreshold=FAST_THR)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
    prev, n_match = None, 0
    for f in frames:
        _, desc = orb.detectAndCompute(f, None)
        if prev is not None and desc is not None:
            for pair in bf.knnMatch(prev, desc, k=2):
                if len(pair) >= 2:
                    if pair[0].distance < RATIO * pair[1].distance:
                        n_match += 1
                elif len(pair) == 1:
                    n_match += 1
        prev = desc
    return n_match


def _oracle_chunk(args):
    w, h, seed, start, count, n_features = args
    from oracle import oracle as O
    from orbslam2.synth import SyntheticVideo
    v = SyntheticVideo(w, h, seed)
    prev, n_match = None, 0
    for k in range(start, start + count):
        r = O.orb_detect_and_compute(v.frame(k), n_features, SCALE, N_LEVELS, FAST_THR)
        if prev is not None:
            n_match += len(O.match(prev, r["desc"], RATIO))
        prev = r["desc"]
    return n_match


def cpu_front_end_fps(w, h, n_features, frames, budget_s=15.0):
    """Times the CPU path on a bounded sample of `frames`; returns (fps, cores, kind, sample)."""
    cores = os.cpu_count() or 1
    try:
        import cv2
        cv2.setNumThreads(cores)
        _cv2_front_end(frames[:3], n_features)                    # warm-up
        t0 = time.perf_counter()
        _cv2_front_end(frames[:4], n_features)
        per = (time.perf_counter() - t0) / 4
        n = int(max(8, min(len(frames), budget_s / max(per, 1e-4))))
        t0 = time.perf_counter()
        _cv2_front_end(frames[:n], n_features)
        dt = time.perf_counter() - t0
        return n / dt, cv2.getNumThreads(), "reference", ("%d frames of the workload through cv2 %s "
                                                           "detectAndCompute + knnMatch + ratio loop "
                                                           "(the reference's calls), %d threads"
                                                           % (n, cv2.__version__, cv2.getNumThreads()))
    except ImportError:
        import multiprocessing as mp
        per_proc = 4
        with mp.get_context("spawn").Pool(cores) as pool:
            t0 = time.perf_counter()
            pool.map(_oracle_chunk, [(w, h, 0, i * per_proc, per_proc, n_features) for i in range(cores)])
            dt = time.perf_counter() - t0
        n = cores * per_proc
        return n / dt, cores, "port", "%d frames through the C oracle port, %d processes" % (n, cores)


def run_reference_arm(args, rank):
    """--impl reference: the reference's CPU path on this box's host cores (rank 0 only)."""
    if rank != 0:
        return
    from orbslam2.synth import SyntheticVideo
    w, h, nf, _ = WORKLOADS[args.workload]
    v = SyntheticVideo(w, h, 0)
    frames = [v.frame(k) for k in range(64)]
    cores = os.cpu_count() or 1
    try:
        import cv2
        cv2.setNumThreads(cores)
        kind, threads = "reference", cv2.getNumThreads()
        run = lambda fr: _cv2_front_end(fr, nf)
        sample = "cv2 %s calls of extractor.py:38-65 + matcher.py:70-81" % cv2.__version__
    except ImportError:
        from oracle import oracle as O
        kind, threads = "port", 1

        def run(fr):
            prev = None
            for f in fr:
                d = O.orb_detect_and_compute(f, nf, SCALE, N_LEVELS, FAST_THR)["desc"]
                if prev is not None:
                    O.match(prev, d, RATIO)
                prev = d
        sample = "C oracle port, single thread"
    per_step = 32 if w <= 640 else 4      # frames per step: a bounded sample of the workload
    for _ in range(max(args.warmup, 1)):
        run(frames[:per_step])
    t0 = time.perf_counter()
    for s in range(args.steps):
        run(frames[(s % 2) * per_step:(s % 2) * per_step + per_step])
    dt = time.perf_counter() - t0
    fps = args.steps * per_step / dt
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_block(args.workload, args.batch or WORKLOADS[args.workload][3], max(args.gpus, 1), args.steps),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                             "sample": "bounded sample: %d frames of the workload per step (the GPU arm's step is the full "
                                       "batch); %s" % (per_step, sample)},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def config_block(workload, B, world, K):
    """The `config` object of both arms (the reference arm times a bounded sample of the same workload)."""
    w, h, _, _ = WORKLOADS[workload]
    return {"workload": workload_name(workload), "frames_per_step_per_gpu": B,
            "l2_policy": "inputs larger than L2 (%.0f MB of frames per step)" % (B * w * h / 1e6),
            "sharding": "one clip of world x B x steps frames, contiguous range per rank (sharding.frame_range), "
                        "1 boundary descriptor block per rank boundary"}


def matching_block(pairs_per_s, ms_per_step):
    """kNN-2 = binary GEMM on tcgen05 kind::i8 (256 u8 MACs per descriptor pair).  Two denominators: the nominal
    dense int8 tensor rate (B200_PROFILING.md, 4.5 Pop/s; MEASURED_PEAKS.json holds no integer-MMA peak) and the
    MEASURED POPC-pipe rate of tools/ubench/popc.cu (profiles/r2_popc_peak.json) -- the roofline of the XOR + POPC
    formulation north_star names, which the tensor-core kernel is measured against as a multiple."""
    blk = {"ms_per_step": ms_per_step, "gpairs_per_s": pairs_per_s / 1e9,
           "engine": os.environ.get("VSLAM_KNN", "tcgen05"),
           "tensor_pops": pairs_per_s * 512.0 / 1e15, "tensor_peak_pops": 4.5,
           "tensor_frac": pairs_per_s * 512.0 / 4.5e15}
    try:
        pk = json.load(open(os.path.join(ROOT, "profiles", "r2_popc_peak.json")))
        peak = float(pk["pairs256_per_s_at_8_popc"])
        blk["popc_pipe_peak_gpairs_per_s"] = peak / 1e9
        blk["vs_popc_pipe_peak"] = pairs_per_s / peak
        blk["popc_peak_source"] = "profiles/r2_popc_peak.json (tools/ubench/popc.cu, measured on this pool's B200)"
    except Exception:
        pass
    return blk


def _kp_table(kps):
    return np.array([(k.pt[0], k.pt[1], k.octave) for k in kps], np.float64).reshape(-1, 3)


def parity_gates(f0, f1, n_features, out, cap):
    """BASELINE.md section 4 gates on the first frame pair of the step, against cv2 itself (the reference's calls):
    kp_agree = fraction of cv2's keypoints with one of ours at the same octave within 0.5 px, desc_identical =
    fraction of those whose 32 descriptor bytes are equal, match_lists_equal = cv2's ratio-test match list on its own
    descriptors equals ours as a set of (pt1, pt2, distance).  `out` = the pinned results of the e2e call."""
    try:
        import cv2
    except ImportError:
        return {"unavailable": "cv2 is not importable on this box"}
    orb = cv2.ORB_create(nfeatures=n_features, scaleFactor=SCALE, nlevels=N_LEVELS, edgeThreshold=31, firstLevel=0,
                         WTA_K=2, scoreType=cv2.ORB_HARRIS_SCORE, patchSize=31, fastThreshold=FAST_THR)
    res = {"against": "cv2 %s on frames 0 and 1 of the shard" % cv2.__version__}
    kp_np, desc_np, cnt_np, match_np, mcnt_np = [x.numpy() for x in out]
    agree, ident, total = 0, 0, 0
    ref_desc, ref_kp = [], []
    for f, img in enumerate((f0, f1)):
        kps, desc = orb.detectAndCompute(img, None)
        ref_kp.append(kps)
        ref_desc.append(desc)
        n = int(cnt_np[f])
        mine = {}
        for i in range(min(n, cap)):
            x, y, octave = float(kp_np[f, i, 0]), float(kp_np[f, i, 1]), int(kp_np[f, i, 5:6].view(np.int32)[0])
            mine[(round(x * 4), round(y * 4), octave)] = i
        for j, k in enumerate(kps):
            total += 1
            i = mine.get((round(k.pt[0] * 4), round(k.pt[1] * 4), k.octave))
            if i is not None and abs(kp_np[f, i, 0] - k.pt[0]) <= 0.5 and abs(kp_np[f, i, 1] - k.pt[1]) <= 0.5:
                agree += 1
                ident += bool(np.array_equal(desc_np[f, i], desc[j]))
    res["kp_agree"] = agree / max(total, 1)
    res["desc_identical"] = ident / max(agree, 1)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
    ref = set()
    for pair in bf.knnMatch(ref_desc[0], ref_desc[1], k=2):
        if len(pair) >= 2 and pair[0].distance < RATIO * pair[1].distance or len(pair) == 1:
            m = pair[0]
            ref.add((ref_kp[0][m.queryIdx].pt, ref_kp[1][m.trainIdx].pt, ref_kp[0][m.queryIdx].octave, float(m.distance)))
    got = set()
    for q, t, d in match_np[0][:int(mcnt_np[0])]:
        got.add(((float(kp_np[0, q, 0]), float(kp_np[0, q, 1])), (float(kp_np[1, t, 0]), float(kp_np[1, t, 1])),
                 int(kp_np[0, q, 5:6].view(np.int32)[0]), float(d)))
    res["match_lists_equal"] = ref == got
    res["match_symmetric_difference"] = len(ref ^ got)
    res["matches"] = len(got)
    return res


def cpu_baseline_block(w, h, n_features, frames):
    """cpu_baseline of the bench line: the reference's calls on all host cores (the headline figure), plus a
    1-thread column and per-stage milliseconds on a short sample."""
    fps, cores, kind, sample = cpu_front_end_fps(w, h, n_features, frames)
    blk = {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample}
    try:
        import cv2
        orb = cv2.ORB_create(nfeatures=n_features, scaleFactor=SCALE, nlevels=N_LEVELS, edgeThreshold=31, firstLevel=0,
                             WTA_K=2, scoreType=cv2.ORB_HARRIS_SCORE, patchSize=31, fastThreshold=FAST_THR)
        bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
        for threads, key in ((cores, "all_threads"), (1, "one_thread")):
            cv2.setNumThreads(threads)
            n = 6 if w <= 640 else 3
            te = tm = 0.0
            prev = None
            for f in frames[:n + 1]:
                t0 = time.perf_counter()
                _, d = orb.detectAndCompute(f, None)
                t1 = time.perf_counter()
                if prev is not None:
                    pairs = bf.knnMatch(prev, d, k=2)
                    _ = [p[0] for p in pairs if len(p) >= 2 and p[0].distance < RATIO * p[1].distance]
                    tm += time.perf_counter() - t1
                    te += t1 - t0
                prev = d
            blk[key] = {"threads": threads, "extract_ms_per_frame": 1e3 * te / n, "match_ms_per_pair": 1e3 * tm / n,
                        "frames_per_s": n / (te + tm)}
        cv2.setNumThreads(cores)
    except ImportError:
        pass
    return blk


def workload_name(key):
    w, h, nf, _ = WORKLOADS[key]
    return ("%s: synthetic %dx%d gray video, %d ORB feats, %d levels, scale %.1f, FAST thr %d, "
            "BF-Hamming kNN-2 ratio %.2f" % (key.upper(), w, h, nf, N_LEVELS, SCALE, FAST_THR, RATIO))


_RESULT_OUT = None


def emit(line):
    """the one JSON line of a run, on the real stdout"""
    out = _RESULT_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="frames per step per GPU (0 = workload default)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg and the cv2 parity gates")
    ap.add_argument("--no-affinity", action="store_true", help="do not bind the rank to the CPUs near its GPU")
    ap.add_argument("--latency", action="store_true",
                    help="latency mode: p50 / p99 of the per-frame calls the reference's callers make (C2, 1 GPU)")
    ap.add_argument("--distribution", default="cv2", choices=["cv2", "orbslam2"],
                    help="keypoint distribution: cv2 = the reference's cv2.ORB retain-best (the metric's config), "
                         "orbslam2 = ini/min cell thresholds + quadtree (SURVEY 8(f) N4)")
    ap.add_argument("--track", action="store_true", help="add the batched tracking step (gates + E-RANSAC + recoverPose) to every step")
    ap.add_argument("--ransac", action="store_true", help="add the batched RANSAC-F stage to every step (default for c3)")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: file descriptor 1 is pointed at stderr for the whole run (NCCL prints
    # its version banner to fd 1, torchrun / libraries may chat), and the result goes to the saved descriptor
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        sys.stderr.write("bench.py: --gpus %d needs torch.distributed.run with %d ranks\n" % (args.gpus, args.gpus))
        sys.exit(2)
    B = args.batch or WORKLOADS[args.workload][3]
    sys.modules.setdefault("bench", sys.modules[__name__])
    if args.latency:
        from bench_latency import run_latency
        run_latency(args)
        return
    from bench_gpu_arm import run_gpu_arm
    run_gpu_arm(args, rank, world, B)


if __name__ == "__main__":
    main()
