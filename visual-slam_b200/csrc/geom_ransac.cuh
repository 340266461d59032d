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

// k distinct indices in [0, n) for RANSAC iteration `it` (duplicate rejection, like cv::RNG sampling)
template <int K>
__device__ inline void draw_sample(uint64_t seed, int it, int n, int idx[K])
{
    uint64_t ctr = 0;
    for (int j = 0; j < K;) {
        uint64_t r = mix64(seed ^ mix64(((uint64_t)it << 20) + ctr));
        ++ctr;
        int v = (int)(r % (uint64_t)n);
        bool dup = false;
        for (int t = 0; t < j; ++t) dup |= idx[t] == v;
        if (!dup) idx[j++] = v;
    }
}

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
