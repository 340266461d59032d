// geom.cu -- RANSAC fundamental / essential matrix, pose recovery and triangulation.
//
// Replaces, for the reference's call sites, the OpenCV calib3d entry points
//   cv2.findFundamentalMat(p1, p2, FM_RANSAC, thr, 0.99)      matcher.py:192, local_mapper.py:136
//   cv2.findEssentialMat(p1, p2, K, RANSAC, prob, thr)        utils.py:124-125, tracker.py:242
//   cv2.recoverPose(E, p1, p2, K, mask)                       utils.py:133, tracker.py:249
//   cv2.triangulatePoints(P1, P2, p1.T, p2.T)                 utils.py:59
// Scoring follows OpenCV's error functions exactly: F scores with
// max(d1^2, d2^2) of the two point-to-epipolar-line distances, E with the Sampson error in
// K-normalised coordinates; both in double, rounded to float and compared with (float)thr^2.
// Hypotheses: every RANSAC iteration is one thread (7-point / 5-point minimal solver) on the subset OpenCV itself
// would draw at that iteration (geom_ransac.cuh: cv::RNG(-1) stream, duplicate / collinearity rejection), every model
// is scored by one CTA / warp, and OpenCV's sequential adaptive-termination rule is replayed over the per-model inlier
// counts (by one thread of the batched F kernel, on the host for E): the inlier mask is the reference's own
// (tests/test_gpu_geometry.py: exact on 27 golden cases), not merely a mask of similar quality.
#include <float.h>
#include <limits.h>
#include <math.h>

#include <vector>

#include "common.cuh"
#include "geom_5pt.cuh"
#include "geom_ransac.cuh"

namespace {

// ---- scoring: one CTA per model -----------------------------------------------------------
// mode 0: fundamental (float2 points), mode 1: essential (double2 normalised points)
template <int MODE>
__global__ void __launch_bounds__(128)
vs_score_models_kernel(const double* __restrict__ models, const int* __restrict__ n_models, int per_iter,
                       const void* __restrict__ p1v, const void* __restrict__ p2v, int n, float thr2,
                       int* __restrict__ counts, uint8_t* __restrict__ mask)
{
    const int m = blockIdx.x;
    if (n_models && (m % per_iter) >= n_models[m / per_iter]) {
        if (threadIdx.x == 0 && counts) counts[m] = -1;
        return;
    }
    __shared__ double M[9];
    __shared__ int s_cnt;
    if (threadIdx.x < 9) M[threadIdx.x] = models[(size_t)m * 9 + threadIdx.x];
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float err;
        if (MODE == 0) {
            const float2 a = ((const float2*)p1v)[i], b = ((const float2*)p2v)[i];
            err = f_error(M, a.x, a.y, b.x, b.y);
        } else {
            const double2 a = ((const double2*)p1v)[i], b = ((const double2*)p2v)[i];
            err = e_error(M, a.x, a.y, b.x, b.y);
        }
        const int in = err <= thr2;
        cnt += in;
        if (mask) mask[i] = (uint8_t)in;
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (threadIdx.x == 0 && counts) counts[m] = s_cnt;
}

// scoring of ONE model (kernel argument) on float pixel coordinates; MODE 1 normalises by K in double first, as
// cv::findEssentialMat does.  count (may be null) += inliers, one atomic per warp.
struct Model9 { double m[9]; };
struct Intr4 { double fx, fy, cx, cy; };

template <int MODE>
__global__ void __launch_bounds__(256)
vs_score_points_kernel(Model9 M, Intr4 K, const float2* __restrict__ p1, const float2* __restrict__ p2, int n, float thr2,
                       uint8_t* __restrict__ mask, int32_t* __restrict__ count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int in = 0;
    if (i < n) {
        const float2 a = p1[i], b = p2[i];
        float err;
        if (MODE == 0) err = f_error(M.m, a.x, a.y, b.x, b.y);
        else err = e_error(M.m, ((double)a.x - K.cx) / K.fx, ((double)a.y - K.cy) / K.fy,
                           ((double)b.x - K.cx) / K.fx, ((double)b.y - K.cy) / K.fy);
        in = err <= thr2;
        if (mask) mask[i] = (uint8_t)in;
    }
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if (count && (threadIdx.x & 31) == 0 && m) atomicAdd(count, __popc(m));
}

__global__ void vs_normalise_points_kernel(Intr4 K, const float2* __restrict__ p, int n, double2* __restrict__ q)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) q[i] = make_double2(((double)p[i].x - K.cx) / K.fx, ((double)p[i].y - K.cy) / K.fy);
}

// ---- hypotheses -------------------------------------------------------------------------
// Hartley normalisation of one point set: x' = (x - cx) * s
struct Norm2 { double cx, cy, s; };

// one thread per RANSAC iteration: the 5-point solver on the subset `table[it]` (drawn on the host from OpenCV's
// stream: EMEstimatorCallback has no checkSubset, so the table depends on n alone)
__global__ void vs_ransac_e_hyp_kernel(const double2* __restrict__ p1, const double2* __restrict__ p2,
                                       const int* __restrict__ table, int iters, double* __restrict__ models,
                                       int* __restrict__ n_models)
{
    const int it = blockIdx.x * blockDim.x + threadIdx.x;
    if (it >= iters) return;
    double x1[5][2], x2[5][2];
    for (int j = 0; j < 5; ++j) {
        const int i = table[it * 5 + j];
        double2 a = p1[i], b = p2[i];
        x1[j][0] = a.x; x1[j][1] = a.y; x2[j][0] = b.x; x2[j][1] = b.y;
    }
    double E[10][9];
    const int ne = solve_5pt(x1, x2, E);
    for (int k = 0; k < ne; ++k)
        for (int e = 0; e < 9; ++e) models[((size_t)it * 10 + k) * 9 + e] = E[k][e];
    n_models[it] = ne;
}

// cv::RANSACUpdateNumIters
int ransac_update_iters(double p, double ep, int model_points, int max_iters)
{
    p = fmax(p, 0.); p = fmin(p, 1.);
    ep = fmax(ep, 0.); ep = fmin(ep, 1.);
    double num = fmax(1. - p, DBL_MIN);
    double denom = 1. - pow(1. - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)lrint(num / denom);
}

// Replays OpenCV's sequential RANSAC loop over per-model inlier counts; returns the winning
// model index (iteration * per_iter + k) or -1.
int ransac_replay(const std::vector<int>& counts, const std::vector<int>& n_models, int per_iter, int n,
                  int model_points, double conf, int max_iters)
{
    int niters = max_iters, max_good = 0, best = -1;
    for (int it = 0; it < niters && it < (int)n_models.size(); ++it)
        for (int k = 0; k < n_models[it]; ++k) {
            const int good = counts[(size_t)it * per_iter + k];
            if (good > (max_good > model_points - 1 ? max_good : model_points - 1)) {
                best = it * per_iter + k;
                max_good = good;
                niters = ransac_update_iters(conf, (double)(n - good) / n, model_points, niters);
            }
        }
    return best;
}

// ---- batched, device-resident fundamental-matrix RANSAC: one CTA per frame pair ---------------
// The CTA gathers the pair's matched points, normalises them (Hartley), then works in rounds of FB_R
// RANSAC iterations: 7-point hypotheses (one thread each, spread over the warps), inlier counts of the
// round's models (one warp per model), and the sequential replay of OpenCV's adaptive-termination rule
// by one thread.  Models past the horizon are never consulted by the replay, so stopping after the round
// that crosses it gives the same winner as evaluating all max_iters iterations -- at ~95 % inliers the
// horizon is a handful of iterations and one round is all a pair costs.
#define FB_R 32                  // iterations per round ...
#define FB_R0 8                  // ... after a first round of 8: at >= 90 % inliers the horizon is below 8 iterations
#define FB_THREADS 256

// deterministic block sum (fixed tree), result valid in every thread
__device__ inline double fb_block_sum(double v, double* s_red)
{
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0;
    for (int w = 0; w < FB_THREADS / 32; ++w) t += s_red[w];
    return t;
}

__global__ void __launch_bounds__(FB_THREADS)
vs_ransac_f_batched_kernel(const vslam_keypoint* __restrict__ kp_q, const vslam_keypoint* __restrict__ kp_t, int kp_cap,
                           const int32_t* __restrict__ match, const int32_t* __restrict__ mcount, int cap,
                           float thr2, double conf, int max_iters, uint64_t seed, int min_n, float2* __restrict__ ws_pts,
                           double* __restrict__ out_F, uint8_t* __restrict__ out_mask, int32_t* __restrict__ out_ninl)
{
    __shared__ double s_models[FB_R * 3 * 9];
    __shared__ int s_nm[FB_R], s_counts[FB_R * 3], s_idx[FB_R * 7];
    __shared__ float s_med[FB_R * 3];
    __shared__ double s_best[9], s_red[FB_THREADS / 32];
    __shared__ unsigned long long s_cursor;
    __shared__ int s_niters, s_maxgood, s_found, s_cnt, s_fail;
    __shared__ double s_minmed;
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n = min(mcount ? mcount[pair] : cap, cap);
    float2* p1 = ws_pts + (size_t)pair * 2 * cap;
    float2* p2 = p1 + cap;
    uint8_t* mask = out_mask + (size_t)pair * cap;
    double* Fo = out_F + (size_t)pair * 9;
    if (n < min_n) {             // matcher.py:184-185 (min_n = 8): fewer than 8 matches -> (matches, ones), no estimate
        for (int i = tid; i < n; i += FB_THREADS) mask[i] = 1;
        if (tid < 9) Fo[tid] = 0.0;
        if (tid == 0) out_ninl[pair] = -1;
        return;
    }
    // cv::findFundamentalMat: FM_RANSAC with fewer than 15 points silently runs the LMedS registrator instead
    const bool lmeds = n < 15;
    const uint32_t salt = vs_seed_salt(seed);
    // ---- matched points (kp == NULL: the caller placed them in ws_pts)
    if (match) {
        const vslam_keypoint* kq = kp_q + (size_t)pair * kp_cap;
        const vslam_keypoint* kt = kp_t + (size_t)pair * kp_cap;
        for (int i = tid; i < n; i += FB_THREADS) {
            const int32_t* m = match + ((size_t)pair * cap + i) * 3;
            const vslam_keypoint a = kq[m[0]], b = kt[m[1]];
            p1[i] = make_float2(a.x, a.y);
            p2[i] = make_float2(b.x, b.y);
        }
        __syncthreads();
    }
    // ---- Hartley normalisation over all points: centroid, then mean distance to it.  (OpenCV's run7Point normalises
    // the 7 sample points; the solutions of the 7-point problem do not depend on the coordinate frame.)
    double sx1 = 0, sy1 = 0, sx2 = 0, sy2 = 0;
    for (int i = tid; i < n; i += FB_THREADS) {
        const float2 a = p1[i], b = p2[i];
        sx1 += a.x; sy1 += a.y; sx2 += b.x; sy2 += b.y;
    }
    Norm2 n1, n2;
    n1.cx = fb_block_sum(sx1, s_red) / n; n1.cy = fb_block_sum(sy1, s_red) / n;
    n2.cx = fb_block_sum(sx2, s_red) / n; n2.cy = fb_block_sum(sy2, s_red) / n;
    double d1 = 0, d2 = 0;
    for (int i = tid; i < n; i += FB_THREADS) {
        const float2 a = p1[i], b = p2[i];
        d1 += hypot(a.x - n1.cx, a.y - n1.cy);
        d2 += hypot(b.x - n2.cx, b.y - n2.cy);
    }
    d1 = fb_block_sum(d1, s_red) / n; d2 = fb_block_sum(d2, s_red) / n;
    n1.s = d1 > 0 ? sqrt(2.0) / d1 : 1.0;
    n2.s = d2 > 0 ? sqrt(2.0) / d2 : 1.0;
    if (tid == 0) {
        // LMedS: a fixed horizon for 45 % outliers (LMeDSPointSetRegistrator::run), at least 3 iterations
        s_niters = lmeds ? max(ransac_update_iters_dev(conf, 0.45, 7, max_iters), 3) : max_iters;
        s_maxgood = 0; s_found = 0; s_cursor = 0ull; s_fail = INT_MAX; s_minmed = DBL_MAX;
    }
    __syncthreads();

    for (int r0 = 0, R = FB_R0; r0 < s_niters; r0 += R, R = FB_R) {
        // ---- the round's subsets from OpenCV's stream (warp 0); getSubset tries 10000 times for RANSAC, 1000 for LMedS
        if (n > 7) {
            if (wid == 0) cv_sample_round<7, true>(&s_cursor, salt, n, r0, R, s_idx, &s_fail, p1, p2, lmeds ? 1000 : 10000);
        } else if (tid < R * 7) {
            s_idx[tid] = tid % 7;         // exactly 7 points: the only subset
        }
        __syncthreads();
        // ---- hypotheses of iterations r0 .. r0 + R - 1, one thread each, spread evenly over the warps
        const int stride = FB_THREADS / R;
        if (tid % stride == 0) {
            const int j = tid / stride, it = r0 + j;
            int kept = 0;
            if (it < max_iters && it < s_fail) {
                double x1[7][2], x2[7][2];
                for (int q = 0; q < 7; ++q) {
                    const float2 a = p1[s_idx[j * 7 + q]], b = p2[s_idx[j * 7 + q]];
                    x1[q][0] = (a.x - n1.cx) * n1.s; x1[q][1] = (a.y - n1.cy) * n1.s;
                    x2[q][0] = (b.x - n2.cx) * n2.s; x2[q][1] = (b.y - n2.cy) * n2.s;
                }
                double F[3][9];
                const int nf = solve_7pt(x1, x2, F);
                for (int k = 0; k < nf; ++k) {
                    const double* f = F[k];
                    double A[9], G[9];   // Fn * T1, then T2^T * A  (T = [s 0 -s cx; 0 s -s cy; 0 0 1])
                    for (int r = 0; r < 3; ++r) {
                        A[3 * r] = f[3 * r] * n1.s;
                        A[3 * r + 1] = f[3 * r + 1] * n1.s;
                        A[3 * r + 2] = f[3 * r + 2] - n1.s * (f[3 * r] * n1.cx + f[3 * r + 1] * n1.cy);
                    }
                    for (int c = 0; c < 3; ++c) {
                        G[c] = n2.s * A[c];
                        G[3 + c] = n2.s * A[3 + c];
                        G[6 + c] = A[6 + c] - n2.s * (n2.cx * A[c] + n2.cy * A[3 + c]);
                    }
                    bool ok = true;
                    for (int e = 0; e < 9; ++e) ok &= isfinite(G[e]);
                    if (!ok) continue;
                    const double mu = fabs(G[8]) > DBL_EPSILON ? 1.0 / G[8] : 1.0;    // F[2][2] = 1 as OpenCV's run7Point
                    double* o = s_models + (j * 3 + kept) * 9;
                    for (int e = 0; e < 9; ++e) o[e] = G[e] * mu;
                    ++kept;
                }
            }
            s_nm[j] = kept;
        }
        __syncthreads();
        // ---- one warp per model slot: inlier count (RANSAC) or median error (LMedS, n <= 14: one point per lane)
        for (int slot = wid; slot < R * 3; slot += FB_THREADS / 32) {
            if (slot % 3 >= s_nm[slot / 3]) continue;
            const double* M = s_models + slot * 9;
            if (!lmeds) {
                int cnt = 0;
                for (int i = lane; i < n; i += 32) {
                    const float2 a = p1[i], b = p2[i];
                    cnt += f_error(M, a.x, a.y, b.x, b.y) <= thr2;
                }
                for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
                if (lane == 0) s_counts[slot] = cnt;
            } else {
                // std::nth_element(err, err + n / 2, err + n): the (n / 2)-th smallest
                float e = 0.f;
                if (lane < n) { const float2 a = p1[lane], b = p2[lane]; e = f_error(M, a.x, a.y, b.x, b.y); }
                int rank = 0;
                for (int j = 0; j < n; ++j) {
                    const float ej = __shfl_sync(0xffffffffu, e, j);
                    rank += (ej < e) || (ej == e && j < lane);
                }
                if (lane < n && rank == n / 2) s_med[slot] = e;
                if (lane < n && e != e) s_med[slot] = e;            // a NaN error poisons the comparison as in OpenCV
            }
        }
        __syncthreads();
        // ---- OpenCV's sequential loop over this round (cv::RANSACPointSetRegistrator::run / LMeDS...::run)
        if (tid == 0) {
            int niters = min(s_niters, s_fail), max_good = s_maxgood;
            for (int j = 0; j < R && r0 + j < niters; ++j)
                for (int k = 0; k < s_nm[j]; ++k) {
                    if (lmeds) {
                        const double med = (double)s_med[j * 3 + k];
                        if (med < s_minmed) {
                            s_minmed = med;
                            for (int e = 0; e < 9; ++e) s_best[e] = s_models[(j * 3 + k) * 9 + e];
                            s_found = 1;
                        }
                        continue;
                    }
                    const int good = s_counts[j * 3 + k];
                    if (good > (max_good > 6 ? max_good : 6)) {
                        for (int e = 0; e < 9; ++e) s_best[e] = s_models[(j * 3 + k) * 9 + e];
                        s_found = 1;
                        max_good = good;
                        niters = ransac_update_iters_dev(conf, (double)(n - good) / n, 7, niters);
                    }
                }
            s_niters = niters; s_maxgood = max_good;
        }
        __syncthreads();
    }
    // ---- mask of the winner
    if (!s_found) {
        for (int i = tid; i < n; i += FB_THREADS) mask[i] = 0;
        if (tid < 9) Fo[tid] = 0.0;
        if (tid == 0) out_ninl[pair] = 0;
        return;
    }
    float t2 = thr2;
    if (lmeds) {                 // sigma = 2.5 * 1.4826 * (1 + 5 / (n - 7)) * sqrt(minMedian), at least 0.001
        double sigma = 2.5 * 1.4826 * (1 + 5. / (n - 7)) * sqrt(s_minmed);
        sigma = fmax(sigma, 0.001);
        t2 = (float)(sigma * sigma);
    }
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < n; i += FB_THREADS) {
        const float2 a = p1[i], b = p2[i];
        const int in = f_error(s_best, a.x, a.y, b.x, b.y) <= t2;
        mask[i] = (uint8_t)in;
        cnt += in;
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
    // LMedS reports failure (cv2 then returns F = None) when fewer than 7 points pass its own gate; the mask stands
    __syncthreads();
    const bool have_F = !lmeds || s_cnt >= 7;
    if (tid < 9) Fo[tid] = have_F ? s_best[tid] : 0.0;
    if (tid == 0) out_ninl[pair] = s_cnt;
}

}  // namespace

#define GEOM_CUDA(call)                                                                  \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) return vs_cuda_fail(e__, #call, __FILE__, __LINE__);     \
    } while (0)

// scoring of one model on device points: mask on the device, inlier count to the host when asked for (then the call
// synchronises the stream)
static int score_points(int mode, const double* h_M, const double* h_K, const float* d_p1, const float* d_p2, int n,
                        float thr2, uint8_t* d_mask, int* n_inliers, cudaStream_t s)
{
    if (n_inliers) *n_inliers = 0;
    if (n == 0) return VSLAM_OK;
    Model9 M;
    memcpy(M.m, h_M, sizeof(M.m));
    Intr4 K = {1, 1, 0, 0};
    if (h_K) K = {h_K[0], h_K[4], h_K[2], h_K[5]};
    void* cnt = nullptr;
    if (n_inliers) {
        int rc = vs_scratch(5, sizeof(int32_t), &cnt);
        if (rc) return rc;
        GEOM_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int32_t), s));
    }
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    if (mode == 0)
        vs_score_points_kernel<0><<<(n + 255) / 256, 256, 0, s>>>(M, K, (const float2*)d_p1, (const float2*)d_p2, n, thr2,
                                                                 d_mask, (int32_t*)cnt);
    else
        vs_score_points_kernel<1><<<(n + 255) / 256, 256, 0, s>>>(M, K, (const float2*)d_p1, (const float2*)d_p2, n, thr2,
                                                                 d_mask, (int32_t*)cnt);
    GEOM_CUDA(cudaGetLastError());
    if (n_inliers) {
        int32_t c = 0;
        GEOM_CUDA(cudaMemcpyAsync(&c, cnt, sizeof(c), cudaMemcpyDeviceToHost, s));
        GEOM_CUDA(cudaStreamSynchronize(s));
        *n_inliers = c;
    }
    return VSLAM_OK;
}

extern "C" int vslam_score_fundamental(const double* h_F, const float* d_p1, const float* d_p2, int n, double thr,
                                       uint8_t* d_mask, int* n_inliers, void* stream)
{
    VS_REQUIRE(h_F && n >= 0 && (n == 0 || (d_p1 && d_p2)), "bad argument");
    return score_points(0, h_F, nullptr, d_p1, d_p2, n, (float)(thr * thr), d_mask, n_inliers, (cudaStream_t)stream);
}

extern "C" int vslam_score_essential(const double* h_E, const double* h_K, const float* d_p1, const float* d_p2, int n,
                                     double thr, uint8_t* d_mask, int* n_inliers, void* stream)
{
    VS_REQUIRE(h_E && h_K && n >= 0 && (n == 0 || (d_p1 && d_p2)), "bad argument");
    VS_REQUIRE(h_K[0] != 0.0 && h_K[4] != 0.0, "singular camera matrix");
    const double t = thr / ((h_K[0] + h_K[4]) / 2);
    return score_points(1, h_E, h_K, d_p1, d_p2, n, (float)(t * t), d_mask, n_inliers, (cudaStream_t)stream);
}

// host points -> per-thread device scratch (slots 0, 1: points; 2: mask)
static int stage_points(const float* h_p1, const float* h_p2, int n, void** a, void** b, void** m)
{
    int rc;
    if ((rc = vs_scratch(0, (size_t)n * 8, a)) || (rc = vs_scratch(1, (size_t)n * 8, b)) || (rc = vs_scratch(2, (size_t)n, m)))
        return rc;
    GEOM_CUDA(cudaMemcpy(*a, h_p1, (size_t)n * 8, cudaMemcpyHostToDevice));
    GEOM_CUDA(cudaMemcpy(*b, h_p2, (size_t)n * 8, cudaMemcpyHostToDevice));
    return VSLAM_OK;
}

extern "C" int vslam_score_fundamental_host(const double* h_F, const float* h_p1, const float* h_p2, int n,
                                            double thr, uint8_t* h_mask, int* n_inliers)
{
    VS_REQUIRE(h_F && h_p1 && h_p2 && n >= 0, "bad argument");
    if (n_inliers) *n_inliers = 0;
    if (n == 0) return VSLAM_OK;
    void *a, *b, *m;
    int rc = stage_points(h_p1, h_p2, n, &a, &b, &m);
    if (rc) return rc;
    int c = 0;
    if ((rc = vslam_score_fundamental(h_F, (const float*)a, (const float*)b, n, thr, (uint8_t*)m, &c, nullptr))) return rc;
    if (h_mask) GEOM_CUDA(cudaMemcpy(h_mask, m, n, cudaMemcpyDeviceToHost));
    if (n_inliers) *n_inliers = c;
    return VSLAM_OK;
}

extern "C" int vslam_score_essential_host(const double* h_E, const double* h_K, const float* h_p1,
                                          const float* h_p2, int n, double thr, uint8_t* h_mask, int* n_inliers)
{
    VS_REQUIRE(h_E && h_K && h_p1 && h_p2 && n >= 0, "bad argument");
    if (n_inliers) *n_inliers = 0;
    if (n == 0) return VSLAM_OK;
    void *a, *b, *m;
    int rc = stage_points(h_p1, h_p2, n, &a, &b, &m);
    if (rc) return rc;
    int c = 0;
    if ((rc = vslam_score_essential(h_E, h_K, (const float*)a, (const float*)b, n, thr, (uint8_t*)m, &c, nullptr))) return rc;
    if (h_mask) GEOM_CUDA(cudaMemcpy(h_mask, m, n, cudaMemcpyDeviceToHost));
    if (n_inliers) *n_inliers = c;
    return VSLAM_OK;
}

// essential-matrix RANSAC on device points: subsets from OpenCV's stream (host) -> hypotheses -> per-model counts ->
// sequential replay of cv::RANSACPointSetRegistrator::run -> final mask.  E and the inlier count are host results, so
// the call synchronises the stream (twice: per-model counts, final count).
extern "C" int vslam_ransac_essential(const float* d_p1, const float* d_p2, int n, const double* h_K, double prob,
                                      double thr, int max_iters, uint64_t seed, double* h_E, uint8_t* d_mask,
                                      int* n_inliers, void* stream)
{
    VS_REQUIRE(d_p1 && d_p2 && h_K && h_E, "null argument");
    VS_REQUIRE(n >= 5, "need at least 5 correspondences");
    VS_REQUIRE(max_iters >= 1 && max_iters <= 100000, "max_iters out of range");
    VS_REQUIRE(h_K[0] != 0.0 && h_K[4] != 0.0, "singular camera matrix");
    cudaStream_t s = (cudaStream_t)stream;
    const int per = 10, mp = 5;
    const double t = thr / ((h_K[0] + h_K[4]) / 2);
    const float thr2 = (float)(t * t);
    // EMEstimatorCallback has no checkSubset: the subsets are a function of n alone
    std::vector<int> table((size_t)max_iters * mp);
    int n_rows = max_iters;
    {
        CvReader rd;
        rd.init(0, vs_seed_salt(seed));
        for (int it = 0; it < max_iters; ++it) {
            if (n > mp) {
                if (!cv_draw_subset<5, false>(rd, n, &table[(size_t)it * mp], nullptr, nullptr, 10000)) { n_rows = it; break; }
            } else {
                for (int j = 0; j < mp; ++j) table[(size_t)it * mp + j] = j;
            }
        }
    }
    if (n_inliers) *n_inliers = 0;
    // one scratch block: normalised points (2 x n double2), models, model counts, per-iteration counts, subset table
    const size_t o_q2 = (size_t)n * 16, o_models = 2 * o_q2, o_counts = o_models + (size_t)max_iters * per * 72,
                 o_nm = o_counts + (size_t)max_iters * per * 4, o_tab = o_nm + (size_t)max_iters * 4,
                 total = o_tab + table.size() * 4;
    void* blk = nullptr;
    int rc = vs_scratch(6, total, &blk);
    if (rc) return rc;
    uint8_t* base = (uint8_t*)blk;
    double2* q1 = (double2*)base;
    double2* q2 = (double2*)(base + o_q2);
    double* models = (double*)(base + o_models);
    int* counts = (int*)(base + o_counts);
    int* nm = (int*)(base + o_nm);
    int* tab = (int*)(base + o_tab);
    const Intr4 K = {h_K[0], h_K[4], h_K[2], h_K[5]};
    GEOM_CUDA(cudaMemcpyAsync(tab, table.data(), table.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    GEOM_CUDA(cudaMemsetAsync(nm, 0, (size_t)max_iters * sizeof(int), s));
    __atomic_fetch_add(&g_vs_launches, 4ull, __ATOMIC_RELAXED);
    vs_normalise_points_kernel<<<(n + 255) / 256, 256, 0, s>>>(K, (const float2*)d_p1, n, q1);
    vs_normalise_points_kernel<<<(n + 255) / 256, 256, 0, s>>>(K, (const float2*)d_p2, n, q2);
    if (n_rows > 0)
        vs_ransac_e_hyp_kernel<<<(n_rows + 31) / 32, 32, 0, s>>>(q1, q2, tab, n_rows, models, nm);
    GEOM_CUDA(cudaGetLastError());
    vs_score_models_kernel<1><<<max_iters * per, 128, 0, s>>>(models, nm, per, q1, q2, n, thr2, counts, nullptr);
    GEOM_CUDA(cudaGetLastError());
    std::vector<int> h_counts((size_t)max_iters * per), h_nm(max_iters);
    GEOM_CUDA(cudaMemcpyAsync(h_counts.data(), counts, h_counts.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    GEOM_CUDA(cudaMemcpyAsync(h_nm.data(), nm, h_nm.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    GEOM_CUDA(cudaStreamSynchronize(s));
    h_nm.resize(n_rows);
    const int best = ransac_replay(h_counts, h_nm, per, n, mp, prob, max_iters);
    if (best < 0) {
        for (int e = 0; e < 9; ++e) h_E[e] = 0.0;
        if (d_mask) GEOM_CUDA(cudaMemsetAsync(d_mask, 0, n, s));
        return VSLAM_OK;
    }
    GEOM_CUDA(cudaMemcpyAsync(h_E, models + (size_t)best * 9, 9 * sizeof(double), cudaMemcpyDeviceToHost, s));
    GEOM_CUDA(cudaStreamSynchronize(s));
    return score_points(1, h_E, h_K, d_p1, d_p2, n, thr2, d_mask, n_inliers, s);
}

extern "C" size_t vslam_ransac_fundamental_batched_workspace(int pairs, int cap)
{
    return (size_t)(pairs > 0 ? pairs : 0) * 2 * (size_t)(cap > 0 ? cap : 0) * sizeof(float2);
}

extern "C" int vslam_ransac_fundamental_batched(const vslam_keypoint* d_kp_query, const vslam_keypoint* d_kp_train,
                                                int kp_capacity, const int32_t* d_match,
                                                const int32_t* d_match_count, int match_capacity, int pairs,
                                                double thr, double conf, int max_iters, uint64_t seed,
                                                void* d_workspace, double* d_F, uint8_t* d_mask,
                                                int32_t* d_n_inliers, void* stream)
{
    VS_REQUIRE(d_kp_query && d_kp_train && d_match && d_match_count && d_workspace && d_F && d_mask && d_n_inliers,
               "null argument");
    VS_REQUIRE(pairs >= 0 && match_capacity > 0 && kp_capacity > 0, "bad sizes");
    VS_REQUIRE(max_iters >= 1 && max_iters <= 100000, "max_iters out of range");
    if (pairs == 0) return VSLAM_OK;
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_ransac_f_batched_kernel<<<pairs, FB_THREADS, 0, (cudaStream_t)stream>>>(
        d_kp_query, d_kp_train, kp_capacity, d_match, d_match_count, match_capacity, (float)(thr * thr), conf, max_iters,
        seed, 8, (float2*)d_workspace, d_F, d_mask, d_n_inliers);
    GEOM_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

extern "C" int vslam_ransac_fundamental_host(const float* h_p1, const float* h_p2, int n, double thr, double conf,
                                             int max_iters, uint64_t seed, double* h_F, uint8_t* h_mask,
                                             int* n_inliers)
{
    VS_REQUIRE(h_p1 && h_p2 && h_F, "null argument");
    VS_REQUIRE(n >= 7, "need at least 7 correspondences");
    VS_REQUIRE(max_iters >= 1 && max_iters <= 100000, "max_iters out of range");
    // one pair through the batched kernel (same arithmetic, so a batch entry equals a single call); per-thread
    // scratch: points (both sets), then F, count + n, mask
    const int cap = n;
    const size_t o_F = (size_t)2 * cap * sizeof(float2), o_cnt = o_F + 9 * sizeof(double), o_mask = o_cnt + 2 * sizeof(int);
    void* blk = nullptr;
    int rc = vs_scratch(6, o_mask + cap, &blk);
    if (rc) return rc;
    uint8_t* base = (uint8_t*)blk;
    float2* pts = (float2*)base;
    double* F = (double*)(base + o_F);
    int* cnt = (int*)(base + o_cnt);
    uint8_t* mask = base + o_mask;
    GEOM_CUDA(cudaMemcpy(pts, h_p1, (size_t)n * 8, cudaMemcpyHostToDevice));
    GEOM_CUDA(cudaMemcpy(pts + cap, h_p2, (size_t)n * 8, cudaMemcpyHostToDevice));
    const int n_eff = n;
    GEOM_CUDA(cudaMemcpy(cnt + 1, &n_eff, sizeof(int), cudaMemcpyHostToDevice));
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_ransac_f_batched_kernel<<<1, FB_THREADS>>>(nullptr, nullptr, 0, nullptr, cnt + 1, cap, (float)(thr * thr),
                                                  conf, max_iters, seed, 7, pts, F, mask, cnt);
    GEOM_CUDA(cudaGetLastError());
    int c = 0;
    GEOM_CUDA(cudaMemcpy(&c, cnt, sizeof(int), cudaMemcpyDeviceToHost));
    GEOM_CUDA(cudaMemcpy(h_F, F, 9 * sizeof(double), cudaMemcpyDeviceToHost));
    if (h_mask) GEOM_CUDA(cudaMemcpy(h_mask, mask, n, cudaMemcpyDeviceToHost));
    if (n_inliers) *n_inliers = c < 0 ? 0 : c;
    return VSLAM_OK;
}

extern "C" int vslam_ransac_essential_host(const float* h_p1, const float* h_p2, int n, const double* h_K,
                                           double prob, double thr, int max_iters, uint64_t seed, double* h_E,
                                           uint8_t* h_mask, int* n_inliers)
{
    VS_REQUIRE(h_p1 && h_p2 && h_K && h_E, "null argument");
    VS_REQUIRE(n >= 5, "need at least 5 correspondences");
    void *a, *b, *m;
    int rc = stage_points(h_p1, h_p2, n, &a, &b, &m);
    if (rc) return rc;
    rc = vslam_ransac_essential((const float*)a, (const float*)b, n, h_K, prob, thr, max_iters, seed, h_E, (uint8_t*)m,
                                n_inliers, nullptr);
    if (rc) return rc;
    if (h_mask) GEOM_CUDA(cudaMemcpy(h_mask, m, n, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}
