"""GPU arm of bench.py (kept in its own file; bench.py is the entry point)."""
import ctypes as C
import json
import os
import time

import numpy as np


def run_gpu_arm(args, rank, world, B):
    import torch
    import torch.distributed as dist
    import bench as BN
    from orbslam2 import _native, sharding
    from orbslam2.synth import SyntheticVideo

    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _native.lib()
    w, h, nf, _ = BN.WORKLOADS[args.workload]
    K, W = args.steps, args.warmup

    # ---- synthetic frames: B distinct consecutive frames of this rank's range ----------
    video = SyntheticVideo(w, h, 0)
    host_frames = np.empty((B, h, w), np.uint8)
    canvas_period = 64                      # the synthetic pan repeats every 64 frames
    uniq = min(B, canvas_period)
    base = [video.frame(rank * 7 + k) for k in range(uniq)]
    rng = np.random.default_rng(1234 + rank)
    for k in range(B):
        if k < uniq:
            host_frames[k] = base[k]
        else:                               # re-noise so that no two frames of a batch are identical
            f = base[k % uniq].astype(np.int16) + rng.integers(-2, 3, (h, w), dtype=np.int16)
            host_frames[k] = np.clip(f, 0, 255).astype(np.uint8)
    frames = torch.from_numpy(host_frames).to(dev)

    dist_mode = getattr(args, "distribution", "cv2")
    orb = _native.OrbHandle(w, h, n_features=nf, scale_factor=BN.SCALE, n_levels=BN.N_LEVELS,
                            fast_threshold=BN.FAST_THR, max_batch=B, distribution=dist_mode, ini_threshold=20)
    cap = orb.kp_capacity
    kp = torch.empty((B, cap, 6), dtype=torch.float32, device=dev)
    desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    count = torch.zeros(B, dtype=torch.int32, device=dev)
    match = torch.empty((B, cap, 3), dtype=torch.int32, device=dev)
    mcount = torch.zeros(B, dtype=torch.int32, device=dev)
    prev_desc = torch.zeros((cap, 32), dtype=torch.uint8, device=dev)
    prev_count = torch.zeros(1, dtype=torch.int32, device=dev)
    nb_desc = torch.zeros((cap, 32), dtype=torch.uint8, device=dev)      # neighbour's first frame
    nb_count = torch.zeros(1, dtype=torch.int32, device=dev)
    first_desc = torch.zeros((cap, 32), dtype=torch.uint8, device=dev)
    first_count = torch.zeros(1, dtype=torch.int32, device=dev)
    b_idx = torch.empty((cap, 2), dtype=torch.int32, device=dev)
    b_dist = torch.empty((cap, 2), dtype=torch.int32, device=dev)
    b_match = torch.empty((cap, 3), dtype=torch.int32, device=dev)
    b_mcount = torch.zeros(1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)
    # "full front-end" workloads add the batched fundamental-matrix RANSAC over every match list of the step
    ransac = bool(args.ransac) or args.workload == "c3"
    if ransac:
        r_F = torch.zeros((B, 9), dtype=torch.float64, device=dev)
        r_mask = torch.zeros((B, cap), dtype=torch.uint8, device=dev)
        r_ninl = torch.zeros(B, dtype=torch.int32, device=dev)
        r_ws = torch.empty(max(int(L.vslam_ransac_fundamental_batched_workspace(B, cap)), 16), dtype=torch.uint8, device=dev)
        prev_kp = torch.zeros((cap, 6), dtype=torch.float32, device=dev)
    ransac_events = []
    # --track: the reference's tracking step (tracker.py:214-254: both gates, E-RANSAC, recoverPose) on every pair
    track = bool(getattr(args, "track", False))
    if track:
        t_match = torch.zeros((B, cap, 3), dtype=torch.int32, device=dev)
        t_count = torch.zeros(B, dtype=torch.int32, device=dev)
        t_E = torch.zeros((B, 9), dtype=torch.float64, device=dev)
        t_R = torch.zeros((B, 9), dtype=torch.float64, device=dev)
        t_t = torch.zeros((B, 3), dtype=torch.float64, device=dev)
        t_rmask = torch.zeros((B, cap), dtype=torch.uint8, device=dev)
        t_pmask = torch.zeros((B, cap), dtype=torch.uint8, device=dev)
        t_status = torch.zeros((B, 4), dtype=torch.int32, device=dev)
        t_ws = torch.empty(max(int(L.vslam_track_batched_workspace(B, cap)), 16), dtype=torch.uint8, device=dev)
        t_K = np.array([[0.5 * w, 0, 0.5 * w], [0, 0.5 * w, 0.5 * h], [0, 0, 1]], np.float64)
        if not ransac:
            prev_kp = torch.zeros((cap, 6), dtype=torch.float32, device=dev)
    track_events = []

    def track_call(kq, kt, m_off, pairs, o):
        """pairs frame pairs starting at match row m_off; outputs at row o"""
        _native.check(L.vslam_track_batched(
            kq, kt, cap, match.data_ptr() + m_off * cap * 12, mcount.data_ptr() + m_off * 4, cap, pairs,
            _native.ptr(t_K), w, h, 0.02, 0.999, 1.0, 1000, C.c_uint64(0x5EED), t_ws.data_ptr(),
            t_match.data_ptr() + o * cap * 12, t_count.data_ptr() + o * 4, t_E.data_ptr() + o * 72,
            t_rmask.data_ptr() + o * cap, t_pmask.data_ptr() + o * cap, t_R.data_ptr() + o * 72,
            t_t.data_ptr() + o * 24, t_status.data_ptr() + o * 16, sptr))

    def step(first):
        _native.check(L.vslam_frontend_run(orb.handle, frames.data_ptr(), B, w,
                                           None if first else prev_desc.data_ptr(),
                                           None if first else prev_count.data_ptr(), BN.RATIO,
                                           kp.data_ptr(), desc.data_ptr(), count.data_ptr(),
                                           match.data_ptr(), mcount.data_ptr(), sptr))
        if ransac:
            # pairs of this step: (prev, 0) when there is a previous step, then (i, i+1); match row p belongs to
            # pair p.  Without a previous frame pair 0 is (0, 1) and there are B - 1 pairs.
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            if first:
                _native.check(L.vslam_ransac_fundamental_batched(
                    kp.data_ptr(), kp.data_ptr() + cap * 24, cap, match.data_ptr(), mcount.data_ptr(), cap, B - 1,
                    3.0, 0.99, 1000, C.c_uint64(0x5EED), r_ws.data_ptr(), r_F.data_ptr(), r_mask.data_ptr(),
                    r_ninl.data_ptr(), sptr))
            else:
                _native.check(L.vslam_ransac_fundamental_batched(
                    prev_kp.data_ptr(), kp.data_ptr(), cap, match.data_ptr(), mcount.data_ptr(), cap, 1,
                    3.0, 0.99, 1000, C.c_uint64(0x5EED), r_ws.data_ptr(), r_F.data_ptr(), r_mask.data_ptr(),
                    r_ninl.data_ptr(), sptr))
                _native.check(L.vslam_ransac_fundamental_batched(
                    kp.data_ptr(), kp.data_ptr() + cap * 24, cap, match.data_ptr() + cap * 12, mcount.data_ptr() + 4,
                    cap, B - 1, 3.0, 0.99, 1000, C.c_uint64(0x5EED), r_ws.data_ptr() + 2 * cap * 8,
                    r_F.data_ptr() + 72, r_mask.data_ptr() + cap, r_ninl.data_ptr() + 4, sptr))
            ev1.record()
            ransac_events.append((ev0, ev1))
        if track:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            if first:
                track_call(kp.data_ptr(), kp.data_ptr() + cap * 24, 0, B - 1, 0)
            else:
                track_call(prev_kp.data_ptr(), kp.data_ptr(), 0, 1, 0)
                track_call(kp.data_ptr(), kp.data_ptr() + cap * 24, 1, B - 1, 1)
            ev1.record()
            track_events.append((ev0, ev1))
        if ransac or track:
            prev_kp.copy_(kp[B - 1], non_blocking=True)
        prev_desc.copy_(desc[B - 1], non_blocking=True)
        prev_count.copy_(count[B - 1:B], non_blocking=True)

    def boundary_exchange():
        """rank r+1 -> rank r: descriptors of its first frame (one block per rank boundary)."""
        if world == 1:
            return []
        first_desc.copy_(desc[0])
        first_count.copy_(count[0:1])
        return sharding.boundary_exchange(dist, rank, world, first_desc, first_count, nb_desc, nb_count)

    def boundary_match():
        if world == 1 or rank == world - 1:
            return
        _native.check(L.vslam_bf_knn2_hamming256(prev_desc.data_ptr(), cap, cap, nb_desc.data_ptr(), cap, cap,
                                                 prev_count.data_ptr(), nb_count.data_ptr(), 1,
                                                 b_idx.data_ptr(), b_dist.data_ptr(), sptr))
        _native.check(L.vslam_ratio_filter(b_idx.data_ptr(), b_dist.data_ptr(), cap, prev_count.data_ptr(), 1,
                                           BN.RATIO, 1, b_match.data_ptr(), b_mcount.data_ptr(), sptr))

    def job(steps):
        reqs = []
        for s in range(steps):
            step(first=(s == 0))
            if s == 0:
                reqs = boundary_exchange()
        for r in reqs:
            r.wait()
        boundary_match()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then the timed region (device events, max over ranks) -----------------
    job(max(W, 3))
    barrier()
    _native.check(orb_status(L, orb))
    launches0 = L.vslam_launch_count()
    sampler = BN.ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    prof = np.zeros(24, np.float64)
    buf = np.zeros(24, np.float32)
    L.vslam_orb_set_profiling(orb.handle, 1)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reqs = []
    for s in range(K):
        step(first=(s == 0))
        if s == 0:
            reqs = boundary_exchange()
        _native.check(L.vslam_orb_read_profile(orb.handle, _native.ptr(buf), 24))   # waits for this step
        prof += buf
    for r in reqs:
        r.wait()
    boundary_match()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    L.vslam_orb_set_profiling(orb.handle, 0)
    launches = L.vslam_launch_count() - launches0
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    ms = float(sharding.max_over_ranks(dist, world, t).item())
    value = world * B * K / (ms * 1e-3)
    n_kp = int(count.clamp(max=cap).sum().item())
    n_match = int(mcount.sum().item())
    ransac_ms = sum(a.elapsed_time(b_) for a, b_ in ransac_events[-K:]) if ransac else 0.0
    n_inl = int(r_ninl.clamp(min=0).sum().item()) if ransac else None
    track_ms = sum(a.elapsed_time(b_) for a, b_ in track_events[-K:]) if track else 0.0

    # ---- e2e through the host-buffer C ABI (pinned host memory, copies timed) -----------
    # Two host threads, each with its own handle (own streams and workspace) and its own pinned result
    # buffers, call vslam_frontend_run_host back to back (ctypes releases the GIL): the copies of one call
    # overlap the kernels of the other.  Every step's H2D, kernels and D2H are inside the timed region.
    import threading
    pin_frames = torch.from_numpy(host_frames).pin_memory()
    n_workers = 2
    handles = [orb, _native.OrbHandle(w, h, n_features=nf, scale_factor=BN.SCALE, n_levels=BN.N_LEVELS,
                                      fast_threshold=BN.FAST_THR, max_batch=B, distribution=dist_mode,
                                      ini_threshold=20)]
    outs = []
    for _ in range(n_workers):
        outs.append((torch.empty((B, cap, 6), dtype=torch.float32).pin_memory(),
                     torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory(),
                     torch.empty(B, dtype=torch.int32).pin_memory(),
                     torch.empty((B, cap, 3), dtype=torch.int32).pin_memory(),
                     torch.empty(B, dtype=torch.int32).pin_memory()))

    def e2e_step(k):
        o = outs[k]
        _native.check(L.vslam_frontend_run_host(handles[k].handle, pin_frames.data_ptr(), B, BN.RATIO,
                                                o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(),
                                                o[3].data_ptr(), o[4].data_ptr()))

    def e2e_run(steps_per_worker):
        errs = []

        def work(k):
            try:
                torch.cuda.set_device(local)
                for _ in range(steps_per_worker):
                    e2e_step(k)
            except Exception as e:      # surfaced after the join
                errs.append(e)
        th = [threading.Thread(target=work, args=(k,)) for k in range(n_workers)]
        for t_ in th:
            t_.start()
        for t_ in th:
            t_.join()
        if errs:
            raise errs[0]

    per_worker = max(2, min(K, 6) // 2)
    Ke = per_worker * n_workers
    e2e_step(0)
    e2e_step(1)
    e2e_run(1)
    barrier()
    t0 = time.perf_counter()
    e2e_run(per_worker)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if not (torch.equal(outs[0][2], outs[1][2]) and torch.equal(outs[0][4][:B - 1], outs[1][4][:B - 1])):
        raise RuntimeError("e2e workers disagree on keypoint / match counts")
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    e2e_fps = world * B * Ke / float(sharding.max_over_ranks(dist, world, t).item())
    # --track: the clip call (frames in, 4x4 poses out: vslam_track_clip_host) through the same two host threads
    e2e_track_fps = None
    if track:
        pose_out = [(torch.empty((B, 16), dtype=torch.float64).pin_memory(),
                     torch.empty((B, 4), dtype=torch.int32).pin_memory()) for _ in range(n_workers)]

        def clip_step(k):
            _native.check(L.vslam_track_clip_host(handles[k].handle, pin_frames.data_ptr(), B, BN.RATIO, _native.ptr(t_K),
                                                  0.02, 0.999, 1.0, 1000, C.c_uint64(0x5EED),
                                                  pose_out[k][0].data_ptr(), pose_out[k][1].data_ptr(), None))

        def clip_run(n):
            errs = []

            def work(k):
                try:
                    torch.cuda.set_device(local)
                    for _ in range(n):
                        clip_step(k)
                except Exception as e:
                    errs.append(e)
            th = [threading.Thread(target=work, args=(k,)) for k in range(n_workers)]
            for t_ in th:
                t_.start()
            for t_ in th:
                t_.join()
            if errs:
                raise errs[0]
        clip_run(1)
        barrier()
        t0 = time.perf_counter()
        clip_run(per_worker)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        e2e_track_fps = world * B * Ke / float(sharding.max_over_ranks(dist, world, t).item())
    handles[1].close()
    h2d = B * w * h
    d2h = B * cap * (24 + 32) + B * 4 + (B - 1) * (cap * 12 + 4)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(BN.ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        # measured DRAM traffic of the level launches of one step (ncu --set full capture committed under
        # profiles/); only quoted when it was taken on this workload and batch
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(BN.ROOT, "profiles", "r1_level_traffic.json")))
            if tj.get("workload") == args.workload and int(tj.get("frames_per_step", 0)) == B:
                traffic = int(tj["dram_read_bytes"]) + int(tj["dram_write_bytes"])
                traffic_src = tj.get("source")
        except Exception:
            pass
        level_ms = float(prof[:16].sum())
        lvl_bytes = BN.level_kernel_bytes(w, h) * B * K
        achieved = lvl_bytes / (level_ms * 1e-3) / 1e9 if level_ms > 0 else 0.0
        breakdown = {("level%d" % i): round(float(prof[i]) / K, 4) for i in range(BN.N_LEVELS)}
        breakdown.update({BN.PROF_NAMES[i]: round(float(prof[i]) / K, 4) for i in BN.PROF_NAMES})
        pairs_per_s = (float(n_kp) / B) ** 2 * B * K / (prof[19] * 1e-3) if prof[19] > 0 else 0.0
        line = {
            "metric": BN.METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": BN.workload_name(args.workload), "frames_per_step_per_gpu": B,
                       "l2_policy": "inputs larger than L2 (%.0f MB of frames per step)" % (B * w * h / 1e6),
                       "sharding": "contiguous frame ranges, 1 boundary descriptor block per rank boundary",
                       "keypoints_per_frame": n_kp / B, "matches_last_step": n_match},
            "clocks": clocks,
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": Ke, "host_threads": n_workers,
                    "call": "vslam_frontend_run_host, one handle per host thread"},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "vs_level_kernel (fused pyramid+FAST+NMS+blur, %d launches/step)" % BN.N_LEVELS,
                         "bound": "hbm", "achieved": achieved, "peak": hbm_peak,
                         "unit": "GB/s", "frac": achieved / hbm_peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                         "algorithmic_bytes_per_frame": BN.level_kernel_bytes(w, h),
                         "algorithmic_bytes_per_step": BN.level_kernel_bytes(w, h) * B, "traffic": traffic,
                         "traffic_note": traffic_src,
                         "ms_per_step": level_ms / K},
            "extraction": {"ms_per_step": float(prof[:19].sum()) / K,
                           "algorithmic_bytes_per_frame": BN.extraction_bytes(w, h, nf),
                           "achieved_gbs": BN.extraction_bytes(w, h, nf) * B * K / (prof[:19].sum() * 1e-3) / 1e9
                           if prof[:19].sum() > 0 else 0.0},
            "matching": {"ms_per_step": float(prof[19] + prof[20]) / K, "gpairs_per_s": pairs_per_s / 1e9,
                         # kNN-2 = binary GEMM on tcgen05 kind::i8 (256 u8 MACs per descriptor pair); the tensor
                         # reference is the nominal dense int8/fp8 rate of B200_PROFILING.md (4.5 Pop/s). The kernel
                         # is bound by the ALU-pipe top-2 reduction of the distances, not by the MMA.
                         "engine": os.environ.get("VSLAM_KNN", "tcgen05"),
                         "tensor_pops": pairs_per_s * 512.0 / 1e15, "tensor_peak_pops": 4.5,
                         "tensor_frac": pairs_per_s * 512.0 / 4.5e15},
            "kernel_ms_per_step": breakdown,
        }
        if ransac:
            breakdown["ransac_f"] = round(ransac_ms / K, 4)
            line["config"]["ransac"] = "fundamental matrix, thr 3.0, conf 0.99, <= 1000 iterations, every match list"
            line["config"]["inliers_last_step"] = n_inl
        if dist_mode != "cv2":
            line["config"]["distribution"] = ("orbslam2: FAST ini 20 / min 7 per 30-px cell, quadtree to the level quota "
                                              "(non-default option; the `select` slot times the quadtree kernel)")
        if track:
            breakdown["track"] = round(track_ms / K, 4)
            line["e2e_track"] = {"value": e2e_track_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d,
                                 "d2h_bytes_per_step": (B - 1) * (12 * 8 + 16),
                                 "call": "vslam_track_clip_host: frames in, 4x4 relative poses out"}
            st = t_status.cpu().numpy()
            line["config"]["track"] = ("tracker.py:214-254 per pair: displacement gate 0.02, 2 x median distance gate, "
                                       "findEssentialMat(RANSAC, 0.999, 1.0), recoverPose")
            line["config"]["track_last_step"] = {"pairs_ok": int((st[:, 0] == 1).sum()),
                                                 "gated_matches": int(t_count.sum().item()),
                                                 "ransac_inliers": int(st[:, 2].sum()), "pose_inliers": int(st[:, 1].sum())}
        if world == 1 and not args.no_cpu:
            fps, cores, kind, sample = BN.cpu_front_end_fps(w, h, nf, [host_frames[k] for k in range(min(B, 64))] * 8)
            line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample}
        BN.emit(line)
    orb.close()
    if world > 1:
        dist.destroy_process_group()


def orb_status(L, orb):
    return L.vslam_orb_last_status(orb.handle)
