// abi_frontend.cu -- matching entry points, the fused front end and the debug taps.
#include <vector>

#include "common.cuh"

int vs_ensure_host_staging(vslam_orb* o);
int vs_frontend_match(vslam_orb* o, int batch, const uint8_t* d_prev_desc, const int32_t* d_prev_count, double ratio,
                      const uint8_t* d_desc, const int32_t* d_count, int32_t* d_match, int32_t* d_match_count,
                      cudaStream_t s);
int vs_launch_match_masked(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, const uint8_t* d_mask,
                           int32_t* d_idx, int32_t* d_dist, cudaStream_t s);

extern "C" int vslam_bf_knn2_hamming256(const uint8_t* d_query, int nq, long q_stride_rows,
                                        const uint8_t* d_train, int nt, long t_stride_rows,
                                        const int32_t* d_nq, const int32_t* d_nt, int pairs,
                                        int32_t* d_idx, int32_t* d_dist, void* stream)
{
    VS_REQUIRE(pairs >= 0 && nq >= 0 && nt >= 0, "negative size");
    if (pairs == 0 || nq == 0) return VSLAM_OK;
    VS_REQUIRE(d_query && d_idx && d_dist && (d_train || nt == 0), "null argument");
    VS_REQUIRE(((uintptr_t)d_query & 15) == 0 && ((uintptr_t)d_train & 15) == 0, "descriptors must be 16-byte aligned");
    return vs_launch_knn2(nullptr, d_query, nq, q_stride_rows, d_train, nt, t_stride_rows, d_nq, d_nt, pairs,
                          d_idx, d_dist, (cudaStream_t)stream);
}

extern "C" int vslam_ratio_filter(const int32_t* d_idx, const int32_t* d_dist, int nq, const int32_t* d_nq,
                                  int pairs, double ratio, int ratio_test, int32_t* d_match,
                                  int32_t* d_match_count, void* stream)
{
    VS_REQUIRE(pairs >= 0 && nq >= 0, "negative size");
    VS_REQUIRE(d_match_count && (nq == 0 || (d_idx && d_dist && d_match)), "null argument");
    return vs_launch_ratio(nullptr, d_idx, d_dist, nq, d_nq, pairs, ratio, ratio_test, d_match, d_match_count,
                           (cudaStream_t)stream);
}

// host-buffer forms: descriptors and results staged in the calling thread's scratch (no cudaMalloc per call)
extern "C" int vslam_bf_knn2_hamming256_host(const uint8_t* h_query, int nq, const uint8_t* h_train, int nt,
                                             int32_t* h_idx, int32_t* h_dist)
{
    VS_REQUIRE(nq >= 0 && nt >= 0, "negative size");
    if (nq == 0) return VSLAM_OK;
    VS_REQUIRE(h_query && h_idx && h_dist && (h_train || nt == 0), "null argument");
    void *q, *t, *i;
    int rc;
    if ((rc = vs_scratch(0, (size_t)nq * 32, &q)) || (rc = vs_scratch(1, (size_t)nt * 32, &t)) ||
        (rc = vs_scratch(2, (size_t)nq * 16, &i)))
        return rc;
    int32_t* d_idx = (int32_t*)i;
    int32_t* d_dist = d_idx + (size_t)nq * 2;
    VS_CUDA(cudaMemcpy(q, h_query, (size_t)nq * 32, cudaMemcpyHostToDevice));
    if (nt) VS_CUDA(cudaMemcpy(t, h_train, (size_t)nt * 32, cudaMemcpyHostToDevice));
    rc = vs_launch_knn2(nullptr, (const uint8_t*)q, nq, nq, (const uint8_t*)t, nt, nt, nullptr, nullptr, 1, d_idx, d_dist, 0);
    if (rc) return rc;
    VS_CUDA(cudaMemcpy(h_idx, d_idx, (size_t)nq * 8, cudaMemcpyDeviceToHost));
    VS_CUDA(cudaMemcpy(h_dist, d_dist, (size_t)nq * 8, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}

extern "C" int vslam_bf_match_masked_host(const uint8_t* h_query, int nq, const uint8_t* h_train, int nt,
                                          const uint8_t* h_mask, int32_t* h_idx, int32_t* h_dist)
{
    VS_REQUIRE(nq >= 0 && nt >= 0, "negative size");
    if (nq == 0) return VSLAM_OK;
    VS_REQUIRE(h_query && h_idx && h_dist && (h_train || nt == 0), "null argument");
    void *q, *t, *i, *m = nullptr;
    int rc;
    if ((rc = vs_scratch(0, (size_t)nq * 32, &q)) || (rc = vs_scratch(1, (size_t)nt * 32, &t)) ||
        (rc = vs_scratch(2, (size_t)nq * 8, &i)))
        return rc;
    int32_t* d_idx = (int32_t*)i;
    int32_t* d_dist = d_idx + nq;
    VS_CUDA(cudaMemcpy(q, h_query, (size_t)nq * 32, cudaMemcpyHostToDevice));
    if (nt) VS_CUDA(cudaMemcpy(t, h_train, (size_t)nt * 32, cudaMemcpyHostToDevice));
    if (h_mask) {
        if ((rc = vs_scratch(3, (size_t)nq * nt, &m))) return rc;
        if (nt) VS_CUDA(cudaMemcpy(m, h_mask, (size_t)nq * nt, cudaMemcpyHostToDevice));
    }
    rc = vs_launch_match_masked((const uint8_t*)q, nq, (const uint8_t*)t, nt, (const uint8_t*)m, d_idx, d_dist, 0);
    if (rc) return rc;
    VS_CUDA(cudaMemcpy(h_idx, d_idx, (size_t)nq * 4, cudaMemcpyDeviceToHost));
    VS_CUDA(cudaMemcpy(h_dist, d_dist, (size_t)nq * 4, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}

// ---------------------------------------------------------------------------
static int frontend_body(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch,
                         const uint8_t* d_prev_desc, const int32_t* d_prev_count, double ratio,
                         vslam_keypoint* d_kp, uint8_t* d_desc, int32_t* d_count,
                         int32_t* d_match, int32_t* d_match_count, void* stream)
{
    int rc = vslam_orb_detect_and_compute(o, d_frames, batch, pitch, d_kp, d_desc, d_count, stream);
    if (rc) return rc;
    return vs_frontend_match(o, batch, d_prev_desc, d_prev_count, ratio, d_desc, d_count, d_match, d_match_count,
                             (cudaStream_t)stream);
}

// batches up to this size go through a captured graph once their argument set has been seen (VSLAM_NO_GRAPH=1
// turns it off): the per-frame call is ~18 launches of a few microseconds each
#define VS_GRAPH_MAX_BATCH 8

extern "C" int vslam_frontend_run(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch,
                                  const uint8_t* d_prev_desc, const int32_t* d_prev_count, double ratio,
                                  vslam_keypoint* d_kp, uint8_t* d_desc, int32_t* d_count,
                                  int32_t* d_match, int32_t* d_match_count, void* stream)
{
    VS_REQUIRE(o && d_match && d_match_count, "null argument");
    VS_REQUIRE(!d_prev_desc || d_prev_count, "d_prev_count required with d_prev_desc");
    static const int no_graph = getenv("VSLAM_NO_GRAPH") && atoi(getenv("VSLAM_NO_GRAPH")) != 0;
    cudaStream_t s = (cudaStream_t)stream;
    // the legacy default stream cannot be captured; profiling records events between the launches
    const bool graphable = batch >= 1 && batch <= VS_GRAPH_MAX_BATCH && !no_graph && !o->graph_off && !o->prof_on &&
                           s != nullptr && s != cudaStreamLegacy && !o->knn_no_split;
    if (!graphable)
        return frontend_body(o, d_frames, batch, pitch, d_prev_desc, d_prev_count, ratio, d_kp, d_desc, d_count, d_match,
                             d_match_count, stream);
    uint64_t key[12] = {(uint64_t)(uintptr_t)d_frames, (uint64_t)batch, (uint64_t)pitch, (uint64_t)(uintptr_t)d_prev_desc,
                        (uint64_t)(uintptr_t)d_prev_count, 0, (uint64_t)(uintptr_t)d_kp, (uint64_t)(uintptr_t)d_desc,
                        (uint64_t)(uintptr_t)d_count, (uint64_t)(uintptr_t)d_match, (uint64_t)(uintptr_t)d_match_count,
                        ((uint64_t)o->dist_mode << 32) | (uint32_t)o->ini_threshold};
    memcpy(&key[5], &ratio, sizeof(double));
    vslam_orb::GraphSlot* slot = nullptr;
    for (auto& g : o->graph)
        if (g.seen && !memcmp(g.key, key, sizeof(key))) { slot = &g; break; }
    if (slot && slot->exec) {
        VS_CUDA(cudaGraphLaunch(slot->exec, s));
        o->last_stream = s;
        o->last_frames = d_frames; o->last_pitch = pitch; o->last_batch = batch;
        return VSLAM_OK;
    }
    if (!slot) {
        // first sight of this argument set: run it launch by launch (this also leaves every lazily configured kernel
        // attribute and scratch buffer in place) and remember it
        slot = &o->graph[o->graph_next];
        o->graph_next = (o->graph_next + 1) % (int)(sizeof(o->graph) / sizeof(o->graph[0]));
        if (slot->exec) { cudaGraphExecDestroy(slot->exec); slot->exec = nullptr; }
        memcpy(slot->key, key, sizeof(key));
        slot->seen = 1;
        return frontend_body(o, d_frames, batch, pitch, d_prev_desc, d_prev_count, ratio, d_kp, d_desc, d_count, d_match,
                             d_match_count, stream);
    }
    // second sight: capture the same launches, instantiate, replay from now on
    if (cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        o->graph_off = 1;
        return frontend_body(o, d_frames, batch, pitch, d_prev_desc, d_prev_count, ratio, d_kp, d_desc, d_count, d_match,
                             d_match_count, stream);
    }
    const int rc = frontend_body(o, d_frames, batch, pitch, d_prev_desc, d_prev_count, ratio, d_kp, d_desc, d_count, d_match,
                                 d_match_count, stream);
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamEndCapture(s, &graph);
    if (rc == VSLAM_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&slot->exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (rc != VSLAM_OK || e != cudaSuccess || !slot->exec) {
        // capture is not available here (or the body failed): nothing was executed, so run it plainly and stop trying
        cudaGetLastError();
        slot->exec = nullptr;
        o->graph_off = 1;
        return frontend_body(o, d_frames, batch, pitch, d_prev_desc, d_prev_count, ratio, d_kp, d_desc, d_count, d_match,
                             d_match_count, stream);
    }
    VS_CUDA(cudaGraphLaunch(slot->exec, s));
    return VSLAM_OK;
}

// matching half of the front end: pair 0 = (previous boundary frame, frame 0) when given, then
// pair i = (frame i, frame i+1)
int vs_frontend_match(vslam_orb* o, int batch, const uint8_t* d_prev_desc, const int32_t* d_prev_count, double ratio,
                      const uint8_t* d_desc, const int32_t* d_count, int32_t* d_match, int32_t* d_match_count,
                      cudaStream_t s)
{
    int rc;
    const int cap = o->plan.kp_cap;
    int32_t* m = d_match;
    int32_t* mc = d_match_count;
    if (d_prev_desc) {
        // pair 0 = (previous boundary frame, frame 0)
        rc = vs_launch_knn2(o, d_prev_desc, cap, cap, d_desc, cap, cap, d_prev_count, d_count, 1,
                            o->d_knn_idx, o->d_knn_dist, s);
        if (rc) return rc;
        rc = vs_launch_ratio(o, o->d_knn_idx, o->d_knn_dist, cap, d_prev_count, 1, ratio, 1, m, mc, s);
        if (rc) return rc;
        m += (size_t)cap * 3;
        mc += 1;
    }
    if (batch > 1) {
        // pair i = (frame i, frame i+1): the train block is the query block shifted by one frame
        rc = vs_launch_knn2(o, d_desc, cap, cap, d_desc + (size_t)cap * 32, cap, cap, d_count, d_count + 1,
                            batch - 1, o->d_knn_idx, o->d_knn_dist, s);
        if (rc) return rc;
        rc = vs_launch_ratio(o, o->d_knn_idx, o->d_knn_dist, cap, d_count, batch - 1, ratio, 1, m, mc, s);
        if (rc) return rc;
    }
    vs_prof_mark(o, -1, s);
    return VSLAM_OK;
}

extern "C" int vslam_frontend_run_host(vslam_orb* o, const uint8_t* h_frames, int batch, double ratio,
                                       vslam_keypoint* h_kp, uint8_t* h_desc, int32_t* h_count,
                                       int32_t* h_match, int32_t* h_match_count)
{
    VS_REQUIRE(o && h_frames, "null argument");
    VS_REQUIRE(batch >= 1 && batch <= o->prm.max_batch, "batch exceeds max_batch");
    int rc = vs_ensure_host_staging(o);
    if (rc) return rc;
    const VsPlan& p = o->plan;
    const size_t fbytes = (size_t)p.width * p.height;
    const size_t cap = p.kp_cap;
    // Chunked pipeline: the frames of chunk k+1 are copied in (stream_in) and the results of chunk
    // k-1 copied out (stream_out) while chunk k is computed.  Chunks alternate between two compute
    // streams on disjoint slices of the workspace, so that the tail of one chunk's kernels overlaps
    // the head of the next.  A chunk after the first is matched against the last frame of its
    // predecessor (its descriptors are ready at ev_ext[k-1]), so the match rows come out exactly as
    // for one undivided batch.  Host buffers should be pinned.
    static const int max_chunks = getenv("VSLAM_HOST_CHUNKS") ? atoi(getenv("VSLAM_HOST_CHUNKS")) : VS_HOST_CHUNKS;
    int nchunks = batch / 16;
    nchunks = nchunks < 1 ? 1 : nchunks > VS_HOST_CHUNKS ? VS_HOST_CHUNKS : nchunks;
    if (max_chunks >= 1 && nchunks > max_chunks) nchunks = max_chunks;
    const int chunk = (batch + nchunks - 1) / nchunks;
    const int saved_prof = o->prof_on;
    o->prof_on = 0;   // the per-kernel marks describe one undivided run
    int k = 0;
    for (int f0 = 0; f0 < batch; f0 += chunk, ++k) {
        const int n = batch - f0 < chunk ? batch - f0 : chunk;
        cudaStream_t cs = (k & 1) ? o->stream2 : o->stream;
        VS_CUDA(cudaMemcpyAsync(o->d_frames + f0 * fbytes, h_frames + f0 * fbytes, n * fbytes, cudaMemcpyHostToDevice, o->stream_in));
        VS_CUDA(cudaEventRecord(o->ev_in[k], o->stream_in));
        VS_CUDA(cudaStreamWaitEvent(cs, o->ev_in[k], 0));
        // view of the handle whose workspace starts at frame slot f0
        vslam_orb view = *o;
        view.d_pyr += f0 * p.pyr_frame_bytes;
        view.d_cand += f0 * p.cand_frame_entries;
        view.d_stage += f0 * p.stage_frame_entries;
        view.d_counters += (size_t)f0 * p.n_levels * VS_CTR_STRIDE;
        view.d_knn_idx += f0 * cap * 2;
        view.d_knn_dist += f0 * cap * 2;
        view.d_knn_ws += f0 * vs_knn2_tc5_workspace_per_pair((int)cap);
        view.knn_ws_bytes = (size_t)(n + (nchunks == 1 ? 1 : 0)) * vs_knn2_tc5_workspace_per_pair((int)cap);   // chunks own disjoint slices
        view.knn_no_split = 1;
        rc = vslam_orb_detect_and_compute(&view, o->d_frames + f0 * fbytes, n, p.width, o->d_kp + f0 * cap,
                                          o->d_desc + f0 * cap * 32, o->d_count + f0, cs);
        if (rc) { o->prof_on = saved_prof; return rc; }
        VS_CUDA(cudaEventRecord(o->ev_ext[k], cs));
        if (k > 0) VS_CUDA(cudaStreamWaitEvent(cs, o->ev_ext[k - 1], 0));
        const int row0 = f0 > 0 ? f0 - 1 : 0;                    // first match row this chunk produces
        const int rows = f0 > 0 ? n : n - 1;
        rc = vs_frontend_match(&view, n, f0 > 0 ? o->d_desc + (f0 - 1) * cap * 32 : nullptr,
                               f0 > 0 ? o->d_count + f0 - 1 : nullptr, ratio, o->d_desc + f0 * cap * 32, o->d_count + f0,
                               o->d_match + row0 * cap * 3, o->d_match_count + row0, cs);
        if (rc) { o->prof_on = saved_prof; return rc; }
        VS_CUDA(cudaEventRecord(o->ev_done[k], cs));
        VS_CUDA(cudaStreamWaitEvent(o->stream_out, o->ev_done[k], 0));
        cudaStream_t so = o->stream_out;
        // an output the caller does not want (NULL) is not copied: a caller that consumes the match lists only
        // moves 25 KB per frame back at C2 instead of 140 KB
        if (h_kp) VS_CUDA(cudaMemcpyAsync(h_kp + f0 * cap, o->d_kp + f0 * cap, n * cap * sizeof(vslam_keypoint), cudaMemcpyDeviceToHost, so));
        if (h_desc) VS_CUDA(cudaMemcpyAsync(h_desc + f0 * cap * 32, o->d_desc + f0 * cap * 32, n * cap * 32, cudaMemcpyDeviceToHost, so));
        if (h_count) VS_CUDA(cudaMemcpyAsync(h_count + f0, o->d_count + f0, n * sizeof(int32_t), cudaMemcpyDeviceToHost, so));
        if (rows > 0) {
            if (h_match) VS_CUDA(cudaMemcpyAsync(h_match + row0 * cap * 3, o->d_match + row0 * cap * 3, rows * cap * 3 * sizeof(int32_t), cudaMemcpyDeviceToHost, so));
            if (h_match_count) VS_CUDA(cudaMemcpyAsync(h_match_count + row0, o->d_match_count + row0, rows * sizeof(int32_t), cudaMemcpyDeviceToHost, so));
        }
        o->last_stream = cs;
    }
    o->prof_on = saved_prof;
    // stream_out is ordered behind every chunk (ev_done), so its tail is the end of the call.  Large batches wait on
    // a blocking event: a spinning cudaStreamSynchronize per host thread starves the other ranks' threads when eight
    // processes share the host's cores (8 ranks x 2 threads on 16 cores); small batches keep the low-latency spin.
    if (batch >= 64 && o->ev_host_done) {
        VS_CUDA(cudaEventRecord(o->ev_host_done, o->stream_out));
        VS_CUDA(cudaEventSynchronize(o->ev_host_done));
    } else {
        VS_CUDA(cudaStreamSynchronize(o->stream_out));
    }
    VS_CUDA(cudaStreamSynchronize(o->stream));
    VS_CUDA(cudaStreamSynchronize(o->stream2));
    return vslam_orb_last_status(o);
}

// ---------------------------------------------------------------------------
extern "C" int vslam_orb_debug_level(vslam_orb* o, int frame, int level, int blurred, uint8_t* h_out)
{
    VS_REQUIRE(o && h_out, "null argument");
    VS_REQUIRE(frame >= 0 && frame < o->last_batch && level >= 0 && level < o->plan.n_levels, "bad frame/level");
    const VsLevel& L = o->plan.lv[level];
    VS_CUDA(cudaStreamSynchronize(o->last_stream));
    const uint8_t* src;
    size_t pitch;
    if (level == 0 && !blurred) {
        src = o->last_frames + (size_t)frame * o->last_pitch * o->plan.height;
        pitch = o->last_pitch;
    } else {
        src = o->d_pyr + (size_t)frame * o->plan.pyr_frame_bytes + (blurred ? L.blur_off : L.raw_off);
        pitch = L.pitch;
    }
    if (L.w == 0) return VSLAM_OK;
    VS_CUDA(cudaMemcpy2D(h_out, L.w, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}

extern "C" int vslam_orb_debug_candidates(vslam_orb* o, int frame, int level, int cap, int32_t* h_x,
                                          int32_t* h_y, int32_t* h_score, int* n_out)
{
    VS_REQUIRE(o && n_out, "null argument");
    VS_REQUIRE(frame >= 0 && frame < o->last_batch && level >= 0 && level < o->plan.n_levels, "bad frame/level");
    const VsPlan& p = o->plan;
    const VsLevel& L = p.lv[level];
    VS_CUDA(cudaStreamSynchronize(o->last_stream));
    int32_t n = 0;
    VS_CUDA(cudaMemcpy(&n, o->d_counters + ((size_t)frame * p.n_levels + level) * VS_CTR_STRIDE + VS_CTR_NCAND,
                       sizeof(n), cudaMemcpyDeviceToHost));
    *n_out = n;
    int m = n < cap ? n : cap;
    if (m <= 0) return VSLAM_OK;
    std::vector<VsCand> tmp(m);
    VS_CUDA(cudaMemcpy(tmp.data(), o->d_cand + (size_t)frame * p.cand_frame_entries + L.cand_off,
                       (size_t)m * sizeof(VsCand), cudaMemcpyDeviceToHost));
    for (int i = 0; i < m; ++i) {
        if (h_x) h_x[i] = tmp[i].x;
        if (h_y) h_y[i] = tmp[i].y;
        if (h_score) h_score[i] = (int32_t)tmp[i].v;   // Harris key after a full extraction
    }
    return VSLAM_OK;
}
