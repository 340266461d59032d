// geom_track.cu -- the reference's per-frame tracking step, batched and device-resident.
//
// tracker.py:214-254 runs, for every frame, on Python lists of cv2.DMatch / cv2.KeyPoint:
//   matches = matcher.match(last_desc, desc)                                  (kNN-2 + ratio: bf_match*.cu)
//   matches = matcher.filter_matches_by_geometric_distance(.., 0.02, shape)   matcher.py:109-142
//   matches = matcher.filter_matches_by_distance(matches)                     matcher.py:144-169
//   if len(matches) < 8: fail                                                 tracker.py:233
//   E, mask = cv2.findEssentialMat(p1, p2, K, RANSAC, 0.999, 1.0)             tracker.py:242
//   _, R, t, mask = cv2.recoverPose(E, p1, p2, K, mask=mask)                  tracker.py:249
// Here one CTA takes one frame pair from the match list to (gated matches, E, masks, R, t) without leaving the
// device:
//   A  displacement gate in double (hypot of the float32 keypoint differences, `<=`), order-preserving compaction;
//      decisions within 1e-9 of the limit are counted so that the host can re-decide them with math.hypot;
//   B  descriptor-distance gate: distances are integers 0..256, so the stable sort by distance is a counting
//      sort, 2 x np.median is the sum of the two middle order statistics, and `<` keeps a prefix of the sorted list;
//   C  K-normalised double points of the gated matches;
//   D  5-point RANSAC in rounds of TK_R iterations (one thread per minimal sample, one warp per model for the
//      Sampson inlier counts, OpenCV's adaptive-termination rule replayed by one thread) -- the same sampler,
//      solver and error function as vslam_ransac_essential_host, so a batch entry equals the host call;
//   E  recoverPose: decomposeEssentialMat by one thread, cheirality of the four candidates per point.
#include <float.h>
#include <limits.h>
#include <math.h>

#include "common.cuh"
#include "geom_5pt.cuh"
#include "geom_ransac.cuh"
#include "geom_pose.cuh"

#define TK_THREADS 128
#define TK_R 32                  // RANSAC iterations per round ...
#define TK_R0 8                  // ... after a first round of 8 (the 0.999 horizon is 5 iterations at 95 % inliers, 8 at 90 %)
#ifndef TK_CTAS
#define TK_CTAS 4                // resident CTAs per SM the register budget is capped for
#endif

namespace {

struct TrackParams {
    double fx, fy, cx, cy;
    double max_dist, border_eps;     // displacement gate: limit, and the band the host re-decides
    double prob;
    float thr2;                      // (thr / ((fx + fy) / 2))^2 as float, like cv::findEssentialMat
    int max_iters;
    uint64_t seed;
};

__global__ void __launch_bounds__(TK_THREADS, TK_CTAS)
vs_track_batched_kernel(const vslam_keypoint* __restrict__ kp_q, const vslam_keypoint* __restrict__ kp_t, int kp_cap,
                        const int32_t* __restrict__ match, const int32_t* __restrict__ mcount, int cap,
                        const TrackParams P, int32_t* __restrict__ ws_list, double2* __restrict__ ws_pts,
                        int32_t* __restrict__ out_match, int32_t* __restrict__ out_count, double* __restrict__ out_E,
                        uint8_t* __restrict__ out_rmask, uint8_t* __restrict__ out_pmask, double* __restrict__ out_R,
                        double* __restrict__ out_t, int32_t* __restrict__ out_status)
{
    __shared__ double s_models[TK_R * 10 * 9];
    __shared__ int s_nm[TK_R], s_counts[TK_R * 10], s_idx[TK_R * 5], s_ns[TK_R], s_trouble[TK_R];
    __shared__ signed char s_ok[TK_R * 10];
    __shared__ double s_basis[TK_R * 36];            // per sample: E's entries as linear forms in the pass's basis
    __shared__ unsigned long long s_rng_cursor;
    __shared__ int s_fail;
    __shared__ double s_best[9], s_R1[9], s_R2[9], s_t[3];
    __shared__ Proj s_cand[4];
    __shared__ int s_cursor[258], s_warp[TK_THREADS / 32];
    __shared__ int s_niters, s_maxgood, s_found, s_n2, s_thr, s_border, s_good[4], s_cnt;
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const vslam_keypoint* kq = kp_q + (size_t)pair * kp_cap;
    const vslam_keypoint* kt = kp_t + (size_t)pair * kp_cap;
    const int32_t* m = match + (size_t)pair * cap * 3;
    int32_t* l1 = ws_list + (size_t)pair * cap * 3;
    int32_t* om = out_match + (size_t)pair * cap * 3;
    double2* p1 = ws_pts + (size_t)pair * 2 * cap;
    double2* p2 = p1 + cap;
    uint8_t* rmask = out_rmask + (size_t)pair * cap;
    uint8_t* pmask = out_pmask + (size_t)pair * cap;
    int32_t* st = out_status + (size_t)pair * 4;
    const int n0 = min(mcount[pair], cap);

    for (int i = tid; i < 258; i += TK_THREADS) s_cursor[i] = 0;
    if (tid == 0) { s_border = 0; s_found = 0; s_cnt = 0; s_good[0] = s_good[1] = s_good[2] = s_good[3] = 0; }
    __syncthreads();

    // ---- A. displacement gate (matcher.py:131-140), list order kept
    int n1 = 0;
    for (int c0 = 0; c0 < n0; c0 += TK_THREADS) {
        const int i = c0 + tid;
        bool keep = false;
        int q = 0, t = 0, d = 0;
        if (i < n0) {
            q = m[3 * i]; t = m[3 * i + 1]; d = m[3 * i + 2];
            const vslam_keypoint a = kq[q], b = kt[t];
            const double dx = (double)b.x - (double)a.x, dy = (double)b.y - (double)a.y;
            const double dist = hypot(dx, dy);
            keep = dist <= P.max_dist;
            if (fabs(dist - P.max_dist) <= P.border_eps) atomicAdd(&s_border, 1);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[wid] = __popc(bal);
        __syncthreads();
        int off = 0, tot = 0;
        for (int w = 0; w < TK_THREADS / 32; ++w) { off += w < wid ? s_warp[w] : 0; tot += s_warp[w]; }
        if (keep) {
            const int o = n1 + off + __popc(bal & ((1u << lane) - 1u));
            l1[3 * o] = q; l1[3 * o + 1] = t; l1[3 * o + 2] = d;
            atomicAdd(&s_cursor[min(max(d, 0), 256) + 1], 1);       // histogram, shifted by one for the prefix sum
        }
        n1 += tot;
        __syncthreads();
    }

    // ---- B. descriptor-distance gate (matcher.py:156-169): threshold = 2 x median = sum of the two middle
    // order statistics; s_cursor[d] becomes the number of matches with a smaller distance (start of d's run)
    if (tid == 0) {
        int acc = 0, a = -1, b = -1;
        const int klo = (n1 - 1) / 2, khi = n1 / 2;
        for (int d = 0; d <= 256; ++d) {
            const int c = s_cursor[d + 1];
            if (n1 > 0 && a < 0 && klo < acc + c) a = d;
            if (n1 > 0 && b < 0 && khi < acc + c) b = d;
            s_cursor[d] = acc;
            acc += c;
        }
        s_cursor[257] = acc;
        const int thr = n1 > 0 ? a + b : 0;
        s_thr = thr;
        s_n2 = n1 > 0 ? s_cursor[min(thr, 257)] : 0;
    }
    __syncthreads();
    const int n2 = s_n2, dthr = s_thr;
    if (wid == 0) {           // stable ranks, 32 list entries at a time in list order
        for (int c0 = 0; c0 < n1; c0 += 32) {
            const int i = c0 + lane;
            const bool act = i < n1;
            int q = 0, t = 0, d = -1 - lane;          // idle lanes: distinct keys, groups of one
            if (act) { q = l1[3 * i]; t = l1[3 * i + 1]; d = l1[3 * i + 2]; }
            const unsigned same = __match_any_sync(0xffffffffu, d);
            const int dc = min(max(d, 0), 256);
            if (act) {
                const int rank = s_cursor[dc] + __popc(same & ((1u << lane) - 1u));
                if (d < dthr) { om[3 * rank] = q; om[3 * rank + 1] = t; om[3 * rank + 2] = d; }
            }
            __syncwarp();
            if (act && (same & ((1u << lane) - 1u)) == 0) s_cursor[dc] += __popc(same);
            __syncwarp();
        }
    }
    if (tid == 0) { out_count[pair] = n2; st[3] = s_border; }
    __syncthreads();
    if (n2 < 8) {             // tracker.py:233-234
        if (tid < 9) { out_E[(size_t)pair * 9 + tid] = 0.0; out_R[(size_t)pair * 9 + tid] = 0.0; }
        if (tid < 3) out_t[(size_t)pair * 3 + tid] = 0.0;
        if (tid == 0) { st[0] = 0; st[1] = 0; st[2] = 0; }
        return;
    }

    // ---- C. K-normalised points (cv::findEssentialMat / recoverPose: double, (x - cx) / fx)
    for (int i = tid; i < n2; i += TK_THREADS) {
        const vslam_keypoint a = kq[om[3 * i]], b = kt[om[3 * i + 1]];
        p1[i] = make_double2(((double)a.x - P.cx) / P.fx, ((double)a.y - P.cy) / P.fy);
        p2[i] = make_double2(((double)b.x - P.cx) / P.fx, ((double)b.y - P.cy) / P.fy);
    }
    if (tid == 0) { s_niters = P.max_iters; s_maxgood = 0; s_rng_cursor = 0ull; s_fail = INT_MAX; }
    const uint32_t salt = vs_seed_salt(P.seed);
    __syncthreads();

    // ---- D. RANSAC rounds
    // The minimal samples of a round are the lanes of warp 0 (the solver is one long double-precision chain per
    // sample: packed into one warp it costs one instruction stream per round instead of one per warp; the other
    // warps wait at the barrier and the SM's other CTAs fill the issue slots).
    for (int r0 = 0, R = TK_R0; r0 < s_niters; r0 += R, R = TK_R) {
        // the round's subsets from OpenCV's stream (EMEstimatorCallback has no checkSubset), drawn by warp 0
        if (wid == 0) cv_sample_round<5, false>(&s_rng_cursor, salt, n2, r0, R, s_idx, &s_fail, nullptr, nullptr, 10000);
        __syncthreads();
        // The five-point solver in three stages (geom_5pt.cuh).  Stage 1 -- elimination to the degree-10 polynomial,
        // its real roots, one start vector per root -- is one long serial chain per sample: lanes of warp 0.
        // Stage 2 -- Gauss-Newton polish of every root on the cubic constraints -- is independent per (sample, root)
        // and goes to the whole CTA (it was the other half of that chain while three warps waited at the barrier).
        // Stage 3 -- de-duplication, the "was a root missed" tests and, rarely, the second pass -- back on warp 0.
        if (tid < R) {
            const int j = tid, it = r0 + j;
            int ns = 0;
            s_trouble[j] = 0;
            if (it < P.max_iters && it < s_fail) {
                const int* idx = s_idx + j * 5;
                double x1[5][2], x2[5][2];
                for (int k = 0; k < 5; ++k) {
                    const double2 a = p1[idx[k]], b = p2[idx[k]];
                    x1[k][0] = a.x; x1[k][1] = a.y; x2[k][0] = b.x; x2[k][1] = b.y;
                }
                double nb[4][9], rb[4][9], L[9][4], starts[10][4];
                five_pt_null_space(x1, x2, nb);
                five_pt_rotate_basis(nb, rb);
                bool deflated = false;
                ns = five_pt_starts(rb, L, starts, &deflated);
                if (deflated) s_trouble[j] = 1;
                for (int e = 0; e < 36; ++e) s_basis[j * 36 + e] = (&L[0][0])[e];
                for (int k = 0; k < ns; ++k)
                    for (int e = 0; e < 4; ++e) s_models[(j * 10 + k) * 9 + e] = starts[k][e];
            }
            s_ns[j] = ns;
        }
        __syncthreads();
        for (int t = tid; t < R * 10; t += TK_THREADS) {
            const int j = t / 10, k = t - 10 * j;
            if (k >= s_ns[j]) continue;
            double* slot = s_models + t * 9;
            const double st[4] = {slot[0], slot[1], slot[2], slot[3]};
            double Ek[9];
            const int ok = five_pt_finish((const double(*)[4])(s_basis + j * 36), st, Ek);
            if (ok > 0)
                for (int e = 0; e < 9; ++e) slot[e] = Ek[e];
            s_ok[t] = ok;
        }
        __syncthreads();
        if (tid < R) {
            const int j = tid, ns = s_ns[j];
            double (*Ej)[9] = (double(*)[9])(s_models + j * 90);
            bool trouble = s_trouble[j] != 0;
            int n = 0;
            for (int k = 0; k < ns; ++k) {
                const int ok = s_ok[j * 10 + k];
                if (ok == 0) trouble = true;
                if (ok <= 0) continue;
                double Ek[9];
                for (int e = 0; e < 9; ++e) Ek[e] = Ej[k][e];
                bool dup;
                n = five_pt_merge(Ej, n, Ek, &dup);           // in place: slot n <= k has been read already
                if (dup) trouble = true;
            }
            if (n & 1) trouble = true;
            if (trouble && ns + (int)s_trouble[j] > 0) {      // second pass in the un-mixed basis (rare), serial
                const int* idx = s_idx + j * 5;
                double x1[5][2], x2[5][2];
                for (int k = 0; k < 5; ++k) {
                    const double2 a = p1[idx[k]], b = p2[idx[k]];
                    x1[k][0] = a.x; x1[k][1] = a.y; x2[k][0] = b.x; x2[k][1] = b.y;
                }
                double nb[4][9];
                five_pt_null_space(x1, x2, nb);
                bool t2 = false;
                n = solve_5pt_in_basis(nb, Ej, n, &t2);
            }
            s_nm[j] = n;
        }
        __syncthreads();
        for (int slot = wid; slot < R * 10; slot += TK_THREADS / 32) {
            if (slot % 10 >= s_nm[slot / 10]) continue;
            const double* M = s_models + slot * 9;
            int cnt = 0;
            for (int i = lane; i < n2; i += 32) {
                const double2 a = p1[i], b = p2[i];
                cnt += e_error(M, a.x, a.y, b.x, b.y) <= P.thr2;
            }
            for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
            if (lane == 0) s_counts[slot] = cnt;
        }
        __syncthreads();
        if (tid == 0) {       // cv::RANSACPointSetRegistrator::run over this round, model_points = 5
            int niters = min(s_niters, s_fail), max_good = s_maxgood;
            for (int j = 0; j < R && r0 + j < niters; ++j)
                for (int k = 0; k < s_nm[j]; ++k) {
                    const int good = s_counts[j * 10 + k];
                    if (good > (max_good > 4 ? max_good : 4)) {
                        for (int e = 0; e < 9; ++e) s_best[e] = s_models[(j * 10 + k) * 9 + e];
                        s_found = 1;
                        max_good = good;
                        niters = ransac_update_iters_dev(P.prob, (double)(n2 - good) / n2, 5, niters);
                    }
                }
            s_niters = niters; s_maxgood = max_good;
        }
        __syncthreads();
    }
    if (!s_found) {           // tracker.py:244-245: no essential matrix
        for (int i = tid; i < n2; i += TK_THREADS) { rmask[i] = 0; pmask[i] = 0; }
        if (tid < 9) { out_E[(size_t)pair * 9 + tid] = 0.0; out_R[(size_t)pair * 9 + tid] = 0.0; }
        if (tid < 3) out_t[(size_t)pair * 3 + tid] = 0.0;
        if (tid == 0) { st[0] = -1; st[1] = 0; st[2] = 0; }
        return;
    }

    // ---- E. winner's mask, decomposition, cheirality
    if (tid == 0) {
        decompose_essential(s_best, s_R1, s_R2, s_t);
        s_cand[0] = make_proj(s_R1, s_t, 1.0); s_cand[1] = make_proj(s_R2, s_t, 1.0);
        s_cand[2] = make_proj(s_R1, s_t, -1.0); s_cand[3] = make_proj(s_R2, s_t, -1.0);
    }
    if (tid < 9) out_E[(size_t)pair * 9 + tid] = s_best[tid];
    __syncthreads();
    int cnt = 0, g0 = 0, g1 = 0, g2 = 0, g3 = 0;
    for (int i = tid; i < n2; i += TK_THREADS) {
        const double2 a = p1[i], b = p2[i];
        const int in = e_error(s_best, a.x, a.y, b.x, b.y) <= P.thr2;
        rmask[i] = (uint8_t)in;
        cnt += in;
        const unsigned bits = in ? pose_bits(s_cand, a.x, a.y, b.x, b.y, 50.0) : 0u;
        pmask[i] = (uint8_t)bits;
        g0 += bits & 1; g1 += (bits >> 1) & 1; g2 += (bits >> 2) & 1; g3 += (bits >> 3) & 1;
    }
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_down_sync(0xffffffffu, cnt, o);
        g0 += __shfl_down_sync(0xffffffffu, g0, o); g1 += __shfl_down_sync(0xffffffffu, g1, o);
        g2 += __shfl_down_sync(0xffffffffu, g2, o); g3 += __shfl_down_sync(0xffffffffu, g3, o);
    }
    if (lane == 0) {
        atomicAdd(&s_cnt, cnt);
        atomicAdd(&s_good[0], g0); atomicAdd(&s_good[1], g1); atomicAdd(&s_good[2], g2); atomicAdd(&s_good[3], g3);
    }
    __syncthreads();
    const int w = pose_winner(s_good);
    for (int i = tid; i < n2; i += TK_THREADS) pmask[i] = (pmask[i] >> w) & 1;
    if (tid < 9) out_R[(size_t)pair * 9 + tid] = (w == 0 || w == 2) ? s_R1[tid] : s_R2[tid];
    if (tid < 3) out_t[(size_t)pair * 3 + tid] = (w >= 2 ? -1.0 : 1.0) * s_t[tid];
    if (tid == 0) { st[0] = 1; st[1] = s_good[w]; st[2] = s_cnt; }
}

}  // namespace

extern "C" size_t vslam_track_batched_workspace(int pairs, int match_capacity)
{
    const size_t p = pairs > 0 ? pairs : 0, c = match_capacity > 0 ? match_capacity : 0;
    return p * c * (2 * sizeof(double2) + 3 * sizeof(int32_t)) + 256;
}

extern "C" int vslam_track_batched(const vslam_keypoint* d_kp_query, const vslam_keypoint* d_kp_train, int kp_capacity,
                                   const int32_t* d_match, const int32_t* d_match_count, int match_capacity, int pairs,
                                   const double* h_K, int image_width, int image_height, double threshold_percent,
                                   double prob, double thr, int max_iters, uint64_t seed, void* d_workspace,
                                   int32_t* d_track_match, int32_t* d_track_count, double* d_E, uint8_t* d_ransac_mask,
                                   uint8_t* d_pose_mask, double* d_R, double* d_t, int32_t* d_status, void* stream)
{
    VS_REQUIRE(d_kp_query && d_kp_train && d_match && d_match_count && h_K && d_workspace && d_track_match &&
               d_track_count && d_E && d_ransac_mask && d_pose_mask && d_R && d_t && d_status, "null argument");
    VS_REQUIRE(pairs >= 0 && match_capacity > 0 && kp_capacity > 0, "bad sizes");
    VS_REQUIRE(max_iters >= 1 && max_iters <= 100000, "max_iters out of range");
    VS_REQUIRE(h_K[0] != 0.0 && h_K[4] != 0.0, "singular camera matrix");
    if (pairs == 0) return VSLAM_OK;
    TrackParams P;
    P.fx = h_K[0]; P.fy = h_K[4]; P.cx = h_K[2]; P.cy = h_K[5];
    P.max_dist = ((image_width + image_height) / 2.0) * threshold_percent;      // matcher.py:129
    P.border_eps = 1e-9 * (P.max_dist > 1.0 ? P.max_dist : 1.0);
    P.prob = prob;
    const double t = thr / ((h_K[0] + h_K[4]) / 2);
    P.thr2 = (float)(t * t);
    P.max_iters = max_iters;
    P.seed = seed;
    // workspace: K-normalised points (double2, 16-byte aligned) first, then the gate-1 lists
    uintptr_t base = ((uintptr_t)d_workspace + 15) & ~(uintptr_t)15;
    double2* pts = (double2*)base;
    int32_t* list = (int32_t*)(pts + (size_t)pairs * 2 * match_capacity);
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_track_batched_kernel<<<pairs, TK_THREADS, 0, (cudaStream_t)stream>>>(
        d_kp_query, d_kp_train, kp_capacity, d_match, d_match_count, match_capacity, P, list, pts, d_track_match,
        d_track_count, d_E, d_ransac_mask, d_pose_mask, d_R, d_t, d_status);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return vs_cuda_fail(e, "vs_track_batched_kernel", __FILE__, __LINE__);
    return VSLAM_OK;
}
