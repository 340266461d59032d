// geom_ransac.cuh -- pieces shared by the RANSAC kernels (geom.cu, geom_track.cu): the counter-based sampler,
// OpenCV's error functions for F and E and cv::RANSACUpdateNumIters.
#pragma once
#include <float.h>
#include <math.h>
#include <stdint.h>

namespace {

__host__ __device__ inline uint64_t mix64(uint64_t x)
{
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

// ---- OpenCV's sample stream -----------------------------------------------------------------------------------
// cv2.findFundamentalMat(FM_RANSAC) / findEssentialMat(RANSAC) are deterministic: RANSACPointSetRegistrator::run
// seeds `RNG rng((uint64)-1)` on every call and draws each subset with
//     idx_i = rng.uniform(0, count)   (multiply-with-carry: state = (uint32)state * 4164903690 + (state >> 32),
//                                      output (uint32)state, uniform = output % count), duplicates redrawn,
// redrawing the whole subset when FMEstimatorCallback::checkSubset finds its last point collinear with two earlier
// ones (F only).  The raw outputs do not depend on the data, so the first VS_CVRNG_N of them are a compile-time table
// (no per-device set-up) and a subset is "a cursor into the stream": iteration j of a round speculates that no draw
// before it was rejected and the warp repairs the cursors by a prefix sum when one was (cv_sample_round).
// tools/ransac_agreement.py / oracle/tools/probe_cv2_ransac_stream.py: the 7 points cv2's returned F interpolates are
// a row of this table on every golden pair.
#define VS_CVRNG_N 16384
#define VS_CVRNG_COEFF 4164903690u
struct VsCvStream {
    uint32_t v[VS_CVRNG_N];     // v[k] = output of draw k (0-based)
    uint64_t end_state;         // generator state after VS_CVRNG_N draws
};
constexpr VsCvStream vs_make_cv_stream()
{
    VsCvStream s{};
    uint64_t st = ~0ull;        // RNG((uint64)-1)
    for (int i = 0; i < VS_CVRNG_N; ++i) {
        st = (uint64_t)(uint32_t)st * VS_CVRNG_COEFF + (st >> 32);
        s.v[i] = (uint32_t)st;
    }
    s.end_state = st;
    return s;
}
static __device__ const VsCvStream g_vs_cv_stream_dev = vs_make_cv_stream();
static const VsCvStream g_vs_cv_stream_host = vs_make_cv_stream();

// `seed` of the C ABI: 0 = OpenCV's stream itself (the reference ignores cv2.setRNGSeed, so this is what it runs);
// any other value xors a 32-bit hash of the seed into every raw draw -- a different, equally deterministic stream.
__host__ __device__ inline uint32_t vs_seed_salt(uint64_t seed) { return seed ? (uint32_t)mix64(seed) : 0u; }

// Reader of the raw stream at an arbitrary position; past the table it steps the generator from the table's end state.
struct CvReader {
    uint64_t pos;               // index of the next draw
    uint64_t st, st_n;          // generator state after st_n draws (used only past the table)
    uint32_t salt;
    __host__ __device__ void seek(uint64_t p) { pos = p; }
    __host__ __device__ void init(uint64_t p, uint32_t salt_)
    {
        pos = p; salt = salt_; st = 0; st_n = 0;
    }
    __host__ __device__ uint32_t next()
    {
#ifdef __CUDA_ARCH__
        const VsCvStream& S = g_vs_cv_stream_dev;
#else
        const VsCvStream& S = g_vs_cv_stream_host;
#endif
        uint32_t r;
        if (pos < VS_CVRNG_N) {
            r = S.v[pos];
        } else {
            if (st_n < VS_CVRNG_N || st_n > pos + 1) { st = S.end_state; st_n = VS_CVRNG_N; }
            while (st_n < pos + 1) { st = (uint64_t)(uint32_t)st * VS_CVRNG_COEFF + (st >> 32); ++st_n; }
            r = (uint32_t)st;
        }
        ++pos;
        return r ^ salt;
    }
};

// haveCollinearPoints(m, count) of OpenCV's fundam.cpp for the LAST point of the subset: float differences widened to
// double, |cross| <= FLT_EPSILON * (|dx1| + |dy1| + |dx2| + |dy2|).
__host__ __device__ inline bool cv_last_point_collinear(const float2* p, const int* idx, int count)
{
    const int i = count - 1;
    const float2 pi = p[idx[i]];
    for (int j = 0; j < i; ++j) {
        const float2 pj = p[idx[j]];
        const double dx1 = (double)(pj.x - pi.x), dy1 = (double)(pj.y - pi.y);
        for (int k = 0; k < j; ++k) {
            const float2 pk = p[idx[k]];
            const double dx2 = (double)(pk.x - pi.x), dy2 = (double)(pk.y - pi.y);
            if (fabs(dx2 * dy1 - dy2 * dx1) <= (double)FLT_EPSILON * (fabs(dx1) + fabs(dy1) + fabs(dx2) + fabs(dy2)))
                return true;
        }
    }
    return false;
}

// RANSACPointSetRegistrator::getSubset: K distinct indices in [0, n) from the reader's position; with CHECK the subset
// is redrawn while its last point is collinear in either image (p1, p2: the float points).  false = no subset in
// max_attempts tries (OpenCV then stops iterating).
template <int K, bool CHECK>
__host__ __device__ inline bool cv_draw_subset(CvReader& rd, int n, int* idx, const float2* p1, const float2* p2,
                                               int max_attempts)
{
    for (int attempt = 0; attempt < max_attempts; ++attempt) {
        for (int i = 0; i < K; ++i) {
            int v;
            for (;;) {
                v = (int)(rd.next() % (uint32_t)n);
                bool dup = false;
                for (int t = 0; t < i; ++t) dup |= idx[t] == v;
                if (!dup) break;
            }
            idx[i] = v;
        }
        if (!CHECK || !(cv_last_point_collinear(p1, idx, K) || cv_last_point_collinear(p2, idx, K))) return true;
    }
    return false;
}

#ifdef __CUDACC__
// The subsets of RANSAC iterations r0 .. r0 + R - 1 (R <= 32), drawn by ONE warp: lane j draws iteration r0 + j from
// the cursor it would have if no earlier lane had a rejected draw, the warp compares with the prefix sum of the draws
// really used and re-draws the lanes whose cursor moved (lane 0 is always right, so this ends; with ~1500 points a
// round of 32 sees a duplicate in about a third of the rounds and is repaired in one more pass).
// s_idx [32][K] receives the subsets, *s_cursor (shared) is advanced, *s_fail gets the first iteration without a
// subset (or INT_MAX).  All 32 lanes must call.
template <int K, bool CHECK>
__device__ inline void cv_sample_round(unsigned long long* s_cursor, uint32_t salt, int n, int r0, int R, int* s_idx,
                                       int* s_fail, const float2* p1, const float2* p2, int max_attempts)
{
    const int lane = threadIdx.x & 31;
    const unsigned long long base = *s_cursor;
    __syncwarp();
    unsigned long long start = base + (unsigned long long)K * lane;
    int idx[K];
    for (int i = 0; i < K; ++i) idx[i] = 0;
    unsigned used = 0;
    bool found = true, need = lane < R;
    CvReader rd;
    rd.init(start, salt);
    for (int pass = 0; pass < 33; ++pass) {
        if (need) {
            rd.seek(start);
            found = cv_draw_subset<K, CHECK>(rd, n, idx, p1, p2, max_attempts);
            used = (unsigned)(rd.pos - start);
        }
        unsigned incl = lane < R ? used : 0u;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const unsigned long long true_start = base + incl - (lane < R ? used : 0u);
        need = lane < R && true_start != start;
        start = true_start;
        if (!__any_sync(0xffffffffu, need)) break;
    }
    if (lane < R)
        for (int i = 0; i < K; ++i) s_idx[lane * K + i] = idx[i];
    const unsigned fail_mask = __ballot_sync(0xffffffffu, lane < R && !found);
    if (lane == R - 1) *s_cursor = start + used;
    if (lane == 0 && fail_mask) *s_fail = min(*s_fail, r0 + __ffs(fail_mask) - 1);
    __syncwarp();
}
#endif

// OpenCV's FMEstimatorCallback::computeError for one correspondence (double, then float)
__device__ __forceinline__ float f_error(const double* F, float x1, float y1, float x2, float y2)
{
    double a = F[0] * x1 + F[1] * y1 + F[2];
    double b = F[3] * x1 + F[4] * y1 + F[5];
    double c = F[6] * x1 + F[7] * y1 + F[8];
    double s2 = 1. / (a * a + b * b);
    double d2 = x2 * a + y2 * b + c;
    a = F[0] * x2 + F[3] * y2 + F[6];
    b = F[1] * x2 + F[4] * y2 + F[7];
    c = F[2] * x2 + F[5] * y2 + F[8];
    double s1 = 1. / (a * a + b * b);
    double d1 = x1 * a + y1 * b + c;
    return (float)fmax(d1 * d1 * s1, d2 * d2 * s2);
}

// OpenCV's EMEstimatorCallback::computeError (Sampson) on normalised double coordinates
__device__ __forceinline__ float e_error(const double* E, double x1, double y1, double x2, double y2)
{
    double Ex0 = E[0] * x1 + E[1] * y1 + E[2];
    double Ex1 = E[3] * x1 + E[4] * y1 + E[5];
    double Ex2 = E[6] * x1 + E[7] * y1 + E[8];
    double Et0 = E[0] * x2 + E[3] * y2 + E[6];
    double Et1 = E[1] * x2 + E[4] * y2 + E[7];
    double x2tEx1 = x2 * Ex0 + y2 * Ex1 + Ex2;
    double a = Ex0 * Ex0, b = Ex1 * Ex1, c = Et0 * Et0, d = Et1 * Et1;
    return (float)(x2tEx1 * x2tEx1 / (a + b + c + d));
}

__device__ inline int ransac_update_iters_dev(double p, double ep, int model_points, int max_iters)
{
    p = fmax(p, 0.); p = fmin(p, 1.);
    ep = fmax(ep, 0.); ep = fmin(ep, 1.);
    double num = fmax(1. - p, DBL_MIN);
    double denom = 1. - pow(1. - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)llrint(num / denom);
}

}  // namespace
