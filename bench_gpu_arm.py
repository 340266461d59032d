"""GPU arm of bench.py (kept in its own file; bench.py is the entry point)."""
import ctypes as C
import json
import os
import threading
import time

import numpy as np


def pin_near_gpu(torch, local):
    """Bind this process (and the threads it starts later) to the CPUs NVML lists as local to the GPU, so that the
    pinned staging buffers are first touched on the GPU's NUMA node and the host threads of the e2e leg do not
    migrate across sockets.  Best effort: returns a description, never raises."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(local).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        before = len(os.sched_getaffinity(0))
        pynvml.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        return {"policy": "nvmlDeviceSetCpuAffinity", "cpus_before": before, "cpus": len(cpus),
                "first_cpu": cpus[0], "last_cpu": cpus[-1]}
    except Exception as e:            # no NVML, cpuset without the GPU's CPUs, ...
        return {"policy": "none", "why": "%s: %s" % (type(e).__name__, e), "cpus": len(os.sched_getaffinity(0))}


def match_checksums(torch, match, mcount, cap):
    """Per pair: (match count, order-sensitive 61-bit checksum of the (queryIdx, trainIdx, distance) rows)."""
    P = match.shape[0]
    n = mcount[:P].to(torch.int64).clamp(0, cap)
    rows = torch.arange(cap, device=match.device, dtype=torch.int64)[None, :]
    live = rows < n[:, None]
    m = match.to(torch.int64)
    h = (m[:, :, 0] * 1000003 + m[:, :, 1] * 10007 + m[:, :, 2] * 101 + 7) * (rows + 1)
    h = torch.where(live, h % 2305843009213693951, torch.zeros_like(h))
    return torch.stack([n, h.sum(dim=1) % 2305843009213693951], dim=1)


def run_gpu_arm(args, rank, world, B):
    import torch
    import torch.distributed as dist
    import bench as BN
    from orbslam2 import _native, sharding
    from orbslam2.synth import SyntheticVideo

    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    affinity = pin_near_gpu(torch, local) if not args.no_affinity else {"policy": "off"}
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _native.lib()
    w, h, nf, _ = BN.WORKLOADS[args.workload]
    K, W = args.steps, max(args.warmup, 3)

    # ---- ONE synthetic clip, sharded by contiguous frame range (sharding.frame_range) ------------------------
    # The job is T = world * B * K frames; rank r owns [r*B*K, (r+1)*B*K) and walks it in K steps of B frames.
    # Content: the clip holds world * B distinct frames and every shard loops over its own B of them (frame g of
    # shard r shows video frame r*B + (g - begin_r) mod B), so one device buffer per rank holds the shard's frames.
    T = world * B * K
    begin, end = sharding.frame_range(rank, world, T)
    assert (begin, end) == (rank * B * K, (rank + 1) * B * K)
    video = SyntheticVideo(w, h, 0)
    host_frames = video.frames(rank * B, B)
    frames = torch.from_numpy(host_frames).to(dev)

    dist_mode = getattr(args, "distribution", "cv2")

    def new_handle(max_batch):
        return _native.OrbHandle(w, h, n_features=nf, scale_factor=BN.SCALE, n_levels=BN.N_LEVELS,
                                 fast_threshold=BN.FAST_THR, max_batch=max_batch, distribution=dist_mode,
                                 ini_threshold=20)
    orb = new_handle(B)
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
    b_match = torch.zeros((1, cap, 3), dtype=torch.int32, device=dev)
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
    SEED = C.c_uint64(0)              # 0 = OpenCV's own RANSAC sample stream (what the reference runs)

    def track_call(kq, kt, m_off, pairs, o):
        """pairs frame pairs starting at match row m_off; outputs at row o"""
        _native.check(L.vslam_track_batched(
            kq, kt, cap, match.data_ptr() + m_off * cap * 12, mcount.data_ptr() + m_off * 4, cap, pairs,
            _native.ptr(t_K), w, h, 0.02, 0.999, 1.0, 1000, SEED, t_ws.data_ptr(),
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
                    3.0, 0.99, 1000, SEED, r_ws.data_ptr(), r_F.data_ptr(), r_mask.data_ptr(),
                    r_ninl.data_ptr(), sptr))
            else:
                _native.check(L.vslam_ransac_fundamental_batched(
                    prev_kp.data_ptr(), kp.data_ptr(), cap, match.data_ptr(), mcount.data_ptr(), cap, 1,
                    3.0, 0.99, 1000, SEED, r_ws.data_ptr(), r_F.data_ptr(), r_mask.data_ptr(),
                    r_ninl.data_ptr(), sptr))
                _native.check(L.vslam_ransac_fundamental_batched(
                    kp.data_ptr(), kp.data_ptr() + cap * 24, cap, match.data_ptr() + cap * 12, mcount.data_ptr() + 4,
                    cap, B - 1, 3.0, 0.99, 1000, SEED, r_ws.data_ptr() + 2 * cap * 8,
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

    def boundary_exchange(src_desc, src_count):
        """rank r+1 -> rank r: descriptors of its first frame (one block per rank boundary)."""
        if world == 1:
            return []
        first_desc.copy_(src_desc)
        first_count.copy_(src_count)
        return sharding.boundary_exchange(dist, rank, world, first_desc, first_count, nb_desc, nb_count)

    def boundary_match(last_desc, last_count):
        """the pair across this rank's upper boundary: (its last frame, the neighbour's first frame)"""
        if world == 1 or rank == world - 1:
            return
        _native.check(L.vslam_bf_knn2_hamming256(last_desc.data_ptr(), cap, cap, nb_desc.data_ptr(), cap, cap,
                                                 last_count.data_ptr(), nb_count.data_ptr(), 1,
                                                 b_idx.data_ptr(), b_dist.data_ptr(), sptr))
        _native.check(L.vslam_ratio_filter(b_idx.data_ptr(), b_dist.data_ptr(), cap, last_count.data_ptr(), 1,
                                           BN.RATIO, 1, b_match.data_ptr(), b_mcount.data_ptr(), sptr))

    def job(steps):
        reqs = []
        for s in range(steps):
            step(first=(s == 0))
            if s == 0:
                reqs = boundary_exchange(desc[0], count[0:1])
        for r in reqs:
            r.wait()
        boundary_match(prev_desc, prev_count)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then the timed region (device events, max over ranks) -----------------
    job(W)
    barrier()
    _native.check(orb_status(L, orb))
    launches0 = L.vslam_launch_count()
    sampler = BN.ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    prof = np.zeros(24, np.float64)
    buf = np.zeros(24, np.float32)
    # per-kernel events are recorded between the launches of the timed steps themselves; they accumulate on the
    # handle (mode 2) and are read once behind the timed region, so that the steps queue back to back like a
    # caller's would (a read per step is a host synchronisation per step: ~0.08 ms of idle GPU at C2)
    accumulate = K * 24 <= 8000
    L.vslam_orb_set_profiling(orb.handle, 2 if accumulate else 1)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reqs = []
    for s in range(K):
        step(first=(s == 0))
        if s == 0:
            reqs = boundary_exchange(desc[0], count[0:1])
        if not accumulate:
            _native.check(L.vslam_orb_read_profile(orb.handle, _native.ptr(buf), 24))   # waits for this step
            prof += buf
    for r in reqs:
        r.wait()
    boundary_match(prev_desc, prev_count)
    e1.record()
    barrier()
    if accumulate:
        _native.check(L.vslam_orb_read_profile(orb.handle, _native.ptr(buf), 24))
        prof += buf
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    L.vslam_orb_set_profiling(orb.handle, 0)
    launches = L.vslam_launch_count() - launches0
    _native.check(orb_status(L, orb))
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    ms = float(sharding.max_over_ranks(dist, world, t).item())
    value = world * B * K / (ms * 1e-3)
    n_kp = int(count.clamp(max=cap).sum().item())
    n_match = int(mcount.sum().item())
    ransac_ms = sum(a.elapsed_time(b_) for a, b_ in ransac_events[-K:]) if ransac else 0.0
    n_inl = int(r_ninl.clamp(min=0).sum().item()) if ransac else None
    track_ms = sum(a.elapsed_time(b_) for a, b_ in track_events[-K:]) if track else 0.0

    # ---- the sharded run equals the unsharded one (outside the timed region) ----------------------------------
    # A short clip of world * Bv frames (the first Bv frames of every shard): every rank runs the sharded path on
    # its Bv frames -- interior pairs, boundary exchange, boundary pair -- and rank 0 runs ONE pass over the whole
    # clip; per pair the match count and an order-sensitive checksum of the match rows must be equal.
    Bv = max(2, min(8, B // max(world, 1)))
    _native.check(L.vslam_frontend_run(orb.handle, frames.data_ptr(), Bv, w, None, None, BN.RATIO,
                                       kp.data_ptr(), desc.data_ptr(), count.data_ptr(),
                                       match.data_ptr(), mcount.data_ptr(), sptr))
    reqs = boundary_exchange(desc[0], count[0:1])
    for r in reqs:
        r.wait()
    boundary_match(desc[Bv - 1], count[Bv - 1:Bv])
    mine = match_checksums(torch, match[:Bv - 1], mcount, cap)
    if world > 1:
        bsum = match_checksums(torch, b_match, b_mcount, cap) if rank < world - 1 else torch.zeros((1, 2), dtype=torch.int64, device=dev)
        mine = torch.cat([mine, bsum])                                  # Bv rows per rank (last rank: a zero row)
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        clip = [torch.empty((Bv, h, w), dtype=torch.uint8, device=dev) for _ in range(world)]
        dist.all_gather(clip, frames[:Bv].contiguous())
        sharded = torch.cat(gathered)[:world * Bv - 1]
    else:
        clip = [frames[:Bv]]
        sharded = mine
    sharded_ok, n_checked = None, world * Bv - 1
    if rank == 0:
        whole = torch.cat(clip)
        TB = world * Bv
        vorb = orb if TB <= B else new_handle(TB)
        v_kp = torch.empty((TB, cap, 6), dtype=torch.float32, device=dev)
        v_desc = torch.empty((TB, cap, 32), dtype=torch.uint8, device=dev)
        v_count = torch.zeros(TB, dtype=torch.int32, device=dev)
        v_match = torch.zeros((TB, cap, 3), dtype=torch.int32, device=dev)
        v_mcount = torch.zeros(TB, dtype=torch.int32, device=dev)
        _native.check(L.vslam_frontend_run(vorb.handle, whole.data_ptr(), TB, w, None, None, BN.RATIO,
                                           v_kp.data_ptr(), v_desc.data_ptr(), v_count.data_ptr(),
                                           v_match.data_ptr(), v_mcount.data_ptr(), sptr))
        unsharded = match_checksums(torch, v_match[:TB - 1], v_mcount, cap)
        sharded_ok = bool(torch.equal(unsharded, sharded)) and int(unsharded[:, 0].min().item()) > 0
        if vorb is not orb:
            vorb.close()
        if not sharded_ok:
            raise RuntimeError("sharded run differs from the unsharded one: %s vs %s"
                               % (sharded.cpu().tolist(), unsharded.cpu().tolist()))
    barrier()

    # ---- e2e through the host-buffer C ABI (pinned host memory, copies timed) -----------
    # A few host threads (n_workers), each with its own handle (own streams and workspace) and its own pinned result
    # buffers, call vslam_frontend_run_host back to back (ctypes releases the GIL): the copies of one call
    # overlap the kernels of the other.  Every step's H2D, kernels and D2H are inside the timed region.
    pin_frames = torch.from_numpy(host_frames).pin_memory()
    # three calls in flight keep the GPU busier than two while one of them is in its copy-only head or tail
    # (measured at 1 GPU: 1 / 2 / 3 / 4 threads = 93 / 115 / 119 / 112 k frames/s); with more ranks than two the box's
    # copy path is the limit and more host threads only contend for its cores
    n_workers = int(os.environ.get("VSLAM_BENCH_E2E_WORKERS", "3" if world <= 2 else "2"))
    handles = [orb] + [new_handle(B) for _ in range(n_workers - 1)]
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

    def threaded(fn, n):
        errs = []

        def work(k):
            try:
                torch.cuda.set_device(local)
                for _ in range(n):
                    fn(k)
            except Exception as e:      # surfaced after the join
                errs.append(e)
        th = [threading.Thread(target=work, args=(k,)) for k in range(n_workers)]
        for t_ in th:
            t_.start()
        for t_ in th:
            t_.join()
        if errs:
            raise errs[0]

    def timed(fn, n):
        barrier()
        t0 = time.perf_counter()
        threaded(fn, n)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        return float(sharding.max_over_ranks(dist, world, tt).item())

    per_worker = max(2, (K + n_workers - 1) // n_workers)     # the driver's --steps, split over the host threads
    Ke = per_worker * n_workers
    threaded(e2e_step, 2)
    e2e_fps = world * B * Ke / timed(e2e_step, per_worker)
    if n_workers > 1 and not (torch.equal(outs[0][2], outs[1][2]) and torch.equal(outs[0][4][:B - 1], outs[1][4][:B - 1])):
        raise RuntimeError("e2e workers disagree on keypoint / match counts")
    h2d = B * w * h
    d2h = B * cap * (24 + 32) + B * 4 + (B - 1) * (cap * 12 + 4)
    # ---- parity gates on the first frame pair of the shard, against cv2 itself (rank 0, outside every timed region;
    # read from the pinned results of the e2e call before the copy-ceiling leg overwrites them)
    gates = None
    if rank == 0 and not args.no_cpu:
        gates = BN.parity_gates(host_frames[0], host_frames[1], nf, outs[0], cap)

    # copy-only ceiling: the same pinned buffers and the same bytes per step in the same chunking, no kernels --
    # what this box's PCIe / host memory system gives N ranks x 2 threads for this traffic
    d_in = [torch.empty((B, h, w), dtype=torch.uint8, device=dev) for _ in range(n_workers)]
    d_out = [(torch.empty((B, cap, 6), dtype=torch.float32, device=dev), torch.empty((B, cap, 32), dtype=torch.uint8, device=dev),
              torch.empty((B, cap, 3), dtype=torch.int32, device=dev)) for _ in range(n_workers)]
    cstreams = [(torch.cuda.Stream(dev), torch.cuda.Stream(dev)) for _ in range(n_workers)]
    nch = max(1, min(8, B // 16))
    csz = (B + nch - 1) // nch

    def copy_step(k):
        si, so = cstreams[k]
        for f0 in range(0, B, csz):
            f1 = min(B, f0 + csz)
            with torch.cuda.stream(si):
                d_in[k][f0:f1].copy_(pin_frames[f0:f1], non_blocking=True)
            with torch.cuda.stream(so):
                outs[k][0][f0:f1].copy_(d_out[k][0][f0:f1], non_blocking=True)
                outs[k][1][f0:f1].copy_(d_out[k][1][f0:f1], non_blocking=True)
                outs[k][3][f0:f1].copy_(d_out[k][2][f0:f1], non_blocking=True)
        si.synchronize()
        so.synchronize()

    threaded(copy_step, 2)
    copy_fps = world * B * Ke / timed(copy_step, per_worker)
    del d_in, d_out

    # matches only: the same call with NULL keypoint / descriptor outputs (25 KB per frame back instead of 140 KB)
    def e2e_matches_only(k):
        o = outs[k]
        _native.check(L.vslam_frontend_run_host(handles[k].handle, pin_frames.data_ptr(), B, BN.RATIO,
                                                None, None, o[2].data_ptr(), o[3].data_ptr(), o[4].data_ptr()))
    threaded(e2e_matches_only, 1)
    e2e_mo_fps = world * B * Ke / timed(e2e_matches_only, per_worker)

    # --track: the clip call (frames in, 4x4 poses out: vslam_track_clip_host) through the same host threads
    e2e_track_fps = None
    if track:
        pose_out = [(torch.empty((B, 16), dtype=torch.float64).pin_memory(),
                     torch.empty((B, 4), dtype=torch.int32).pin_memory()) for _ in range(n_workers)]

        def clip_step(k):
            _native.check(L.vslam_track_clip_host(handles[k].handle, pin_frames.data_ptr(), B, BN.RATIO, _native.ptr(t_K),
                                                  0.02, 0.999, 1.0, 1000, SEED,
                                                  pose_out[k][0].data_ptr(), pose_out[k][1].data_ptr(), None))
        threaded(clip_step, 1)
        e2e_track_fps = world * B * Ke / timed(clip_step, per_worker)

    for hh in handles[1:]:
        hh.close()

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(BN.ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        # measured DRAM traffic of the level launches of one step (ncu --set full capture committed under
        # profiles/); only quoted when it was taken on this workload and batch
        traffic, traffic_src, level_winst = None, None, None
        for name in ("r2_level_traffic.json", "r1_level_traffic.json"):
            try:
                tj = json.load(open(os.path.join(BN.ROOT, "profiles", name)))
                if tj.get("workload") == args.workload and int(tj.get("frames_per_step", 0)) == B:
                    traffic = int(tj["dram_read_bytes"]) + int(tj["dram_write_bytes"])
                    traffic_src = tj.get("source")
                    level_winst = tj.get("warp_instructions")
                    break
            except Exception:
                pass
        level_ms = float(prof[:16].sum())
        issue_slots = None
        if level_winst and level_ms > 0:
            sm_hz = float((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
            sms = torch.cuda.get_device_properties(0).multi_processor_count
            ach = level_winst / (level_ms / K * 1e-3)
            issue_slots = {"warp_instructions_per_step": int(level_winst), "achieved_ginst_s": ach / 1e9,
                           "peak_ginst_s": sms * 4 * sm_hz / 1e9, "frac": ach / (sms * 4 * sm_hz)}
        lvl_bytes = BN.level_kernel_bytes(w, h) * B * K
        achieved = lvl_bytes / (level_ms * 1e-3) / 1e9 if level_ms > 0 else 0.0
        breakdown = {("level%d" % i): round(float(prof[i]) / K, 4) for i in range(BN.N_LEVELS)}
        breakdown.update({BN.PROF_NAMES[i]: round(float(prof[i]) / K, 4) for i in BN.PROF_NAMES})
        pairs_per_s = (float(n_kp) / B) ** 2 * B * K / (prof[19] * 1e-3) if prof[19] > 0 else 0.0
        line = {
            "metric": BN.METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": BN.config_block(args.workload, B, world, K),
            "workload_stats": {"keypoints_per_frame": n_kp / B, "matches_last_step": n_match},
            "clocks": clocks,
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": Ke, "host_threads": n_workers,
                    "call": "vslam_frontend_run_host, one handle per host thread",
                    "copy_ceiling_fps": copy_fps, "frac_of_copy_ceiling": e2e_fps / copy_fps if copy_fps else None,
                    "copy_ceiling_note": "same pinned buffers, bytes and chunking, no kernels, same ranks and threads",
                    "matches_only_fps": e2e_mo_fps,
                    "matches_only_d2h_bytes_per_step": B * 4 + (B - 1) * (cap * 12 + 4),
                    "host_affinity": affinity},
            "gpu_launches": int(launches),
            "sharded_equals_unsharded": sharded_ok,
            "sharded_check": "%d pairs of a %d-frame clip (%d per rank, incl. %d boundary pairs): match count and "
                             "checksum equal to one 1-rank pass" % (n_checked, world * Bv, Bv, world - 1),
            "roofline": {"kernel": "vs_level_kernel (fused pyramid+FAST+NMS+blur, %d launches/step)" % BN.N_LEVELS,
                         "bound": "hbm", "achieved": achieved, "peak": hbm_peak,
                         "unit": "GB/s", "frac": achieved / hbm_peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                         "algorithmic_bytes_per_frame": BN.level_kernel_bytes(w, h),
                         "algorithmic_bytes_per_step": BN.level_kernel_bytes(w, h) * B, "traffic": traffic,
                         "traffic_note": traffic_src,
                         "ms_per_step": level_ms / K,
                         # what actually bounds it: the warp-instruction issue slots (4 schedulers per SM, one warp
                         # instruction per clock each).  Instruction count of the 8 launches from the same ncu capture as
                         # `traffic`; rate over the live-timed duration at the SM clock sampled during the timed region
                         "issue_slots": issue_slots},
            "extraction": {"ms_per_step": float(prof[:19].sum()) / K,
                           "algorithmic_bytes_per_frame": BN.extraction_bytes(w, h, nf),
                           "achieved_gbs": BN.extraction_bytes(w, h, nf) * B * K / (prof[:19].sum() * 1e-3) / 1e9
                           if prof[:19].sum() > 0 else 0.0},
            "matching": BN.matching_block(pairs_per_s, float(prof[19] + prof[20]) / K),
            "kernel_ms_per_step": breakdown,
        }
        if gates is not None:
            line["parity"] = gates
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
            line["cpu_baseline"] = BN.cpu_baseline_block(w, h, nf, [host_frames[k] for k in range(min(B, 64))] * 8)
        BN.emit(line)
    orb.close()
    if world > 1:
        dist.destroy_process_group()


def orb_status(L, orb):
    return L.vslam_orb_last_status(orb.handle)
