// abi_track.cu -- a clip in, a trajectory out: the reference's per-frame loop (Tracker.process_frame,
// tracker.py:73-146: extract -> match with the previous frame -> gates -> findEssentialMat -> recoverPose ->
// 4x4 relative pose) for a batch of consecutive frames with host buffers.  Frames go up once; keypoints,
// descriptors, match lists, masks and essential matrices stay in HBM; 28 numbers per frame pair come back
// (the 4x4 pose and the status words of vslam_track_batched).  SURVEY.md 8(f) N1.
#include "common.cuh"

int vs_ensure_host_staging(vslam_orb* o);

namespace {
inline size_t up256(size_t n) { return (n + 255) & ~(size_t)255; }
}

extern "C" int vslam_track_clip_host(vslam_orb* o, const uint8_t* h_frames, int batch, double ratio, const double* h_K,
                                     double threshold_percent, double prob, double thr, int max_iters, uint64_t seed,
                                     double* h_T, int32_t* h_status, int32_t* h_count)
{
    VS_REQUIRE(o && h_frames && h_K && h_T && h_status, "null argument");
    VS_REQUIRE(batch >= 2 && batch <= o->prm.max_batch, "batch must lie in [2, max_batch]");
    int rc = vs_ensure_host_staging(o);
    if (rc) return rc;
    const VsPlan& p = o->plan;
    const size_t cap = p.kp_cap, B = o->prm.max_batch;
    const int pairs = batch - 1;
    // one block: gated matches, counts, E, R, t, both masks, status, workspace
    const size_t sz_match = up256(B * cap * 3 * sizeof(int32_t)), sz_cnt = up256(B * sizeof(int32_t)),
                 sz_m9 = up256(B * 9 * sizeof(double)), sz_t = up256(B * 3 * sizeof(double)), sz_mask = up256(B * cap),
                 sz_st = up256(B * 4 * sizeof(int32_t)), sz_ws = up256(vslam_track_batched_workspace((int)B, (int)cap));
    if (!o->d_track)
        VS_CUDA(cudaMalloc((void**)&o->d_track, sz_match + sz_cnt + 2 * sz_m9 + sz_t + 2 * sz_mask + sz_st + sz_ws));
    uint8_t* q = o->d_track;
    int32_t* d_tm = (int32_t*)q; q += sz_match;
    int32_t* d_tc = (int32_t*)q; q += sz_cnt;
    double* d_E = (double*)q; q += sz_m9;
    double* d_R = (double*)q; q += sz_m9;
    double* d_t = (double*)q; q += sz_t;
    uint8_t* d_rm = q; q += sz_mask;
    uint8_t* d_pm = q; q += sz_mask;
    int32_t* d_st = (int32_t*)q; q += sz_st;
    void* d_ws = q;
    cudaStream_t s = o->stream;
    VS_CUDA(cudaMemcpyAsync(o->d_frames, h_frames, (size_t)batch * p.width * p.height, cudaMemcpyHostToDevice, s));
    rc = vslam_frontend_run(o, o->d_frames, batch, p.width, nullptr, nullptr, ratio, o->d_kp, o->d_desc, o->d_count,
                            o->d_match, o->d_match_count, s);
    if (rc) return rc;
    rc = vslam_track_batched(o->d_kp, o->d_kp + cap, (int)cap, o->d_match, o->d_match_count, (int)cap, pairs, h_K,
                             p.width, p.height, threshold_percent, prob, thr, max_iters, seed, d_ws, d_tm, d_tc, d_E,
                             d_rm, d_pm, d_R, d_t, d_st, s);
    if (rc) return rc;
    // R, t and status travel back; the pose is assembled like tracker.py:250-253 (T = [R t; 0 0 0 1])
    double* h_R = h_T;                                   // staged in the caller's buffer: 16 doubles per pair >= 9 + 3
    VS_CUDA(cudaMemcpy2DAsync(h_R, 16 * sizeof(double), d_R, 9 * sizeof(double), 9 * sizeof(double), pairs,
                              cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaMemcpy2DAsync(h_T + 9, 16 * sizeof(double), d_t, 3 * sizeof(double), 3 * sizeof(double), pairs,
                              cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaMemcpyAsync(h_status, d_st, (size_t)pairs * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    if (h_count) VS_CUDA(cudaMemcpyAsync(h_count, o->d_count, batch * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaStreamSynchronize(s));
    for (int k = 0; k < pairs; ++k) {
        double* T = h_T + (size_t)k * 16;
        double R[9], t[3];
        for (int e = 0; e < 9; ++e) R[e] = T[e];
        for (int e = 0; e < 3; ++e) t[e] = T[9 + e];
        const bool ok = h_status[4 * k] == 1;
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) T[4 * r + c] = ok ? R[3 * r + c] : (r == c ? 1.0 : 0.0);
            T[4 * r + 3] = ok ? t[r] : 0.0;
        }
        T[12] = T[13] = T[14] = 0.0; T[15] = 1.0;
    }
    return vslam_orb_last_status(o);
}
