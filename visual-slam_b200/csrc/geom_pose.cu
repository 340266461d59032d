// geom_pose.cu -- cv2.recoverPose (utils.py:133, tracker.py:249) and cv2.triangulatePoints
// (utils.py:59) on the GPU: one thread per correspondence solves the 4x4 DLT system
//   [x1 P1_2 - P1_0; y1 P1_2 - P1_1; x2 P2_2 - P2_0; y2 P2_2 - P2_1] X = 0
// by a one-sided Jacobi SVD (null vector = right singular vector of the smallest singular
// value), exactly the linear system OpenCV's triangulatePoints solves per point.  recoverPose
// evaluates the four (R, t) candidates of decomposeEssentialMat with OpenCV's cheirality
// rules (depth > 0 in both views and < 50) and the host picks the winner with OpenCV's
// tie order.  The 3x3 decomposition of E is scalar work and runs on the host.
#include <math.h>

#include <vector>

#include "common.cuh"
#include "geom_5pt.cuh"
#include "geom_pose.cuh"

namespace {

__global__ void vs_triangulate_kernel(Proj P1, Proj P2, const float2* __restrict__ p1, const float2* __restrict__ p2,
                                      int n, float* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double X[4];
    dlt_point(P1.m, P2.m, p1[i].x, p1[i].y, p2[i].x, p2[i].y, X);
    for (int r = 0; r < 4; ++r) out[(size_t)r * n + i] = (float)X[r];
}

// cheirality bits of the 4 candidates for each correspondence: float pixel coordinates, normalised by K in double as
// cv::recoverPose does; rows with mask == 0 get no bits.  counts[k] += points with bit k (one atomic per warp).
struct Intr { double fx, fy, cx, cy; };

__global__ void vs_pose_points_kernel(Proj c0, Proj c1, Proj c2, Proj c3, Intr K, const float2* __restrict__ p1,
                                      const float2* __restrict__ p2, int n, double dist,
                                      const uint8_t* __restrict__ mask, uint8_t* __restrict__ bits,
                                      int32_t* __restrict__ counts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned b = 0;
    if (i < n && (!mask || mask[i])) {
        const Proj cand[4] = {c0, c1, c2, c3};
        const float2 a = p1[i], q = p2[i];
        b = pose_bits(cand, ((double)a.x - K.cx) / K.fx, ((double)a.y - K.cy) / K.fy,
                      ((double)q.x - K.cx) / K.fx, ((double)q.y - K.cy) / K.fy, dist);
    }
    if (i < n) bits[i] = (uint8_t)b;
    for (int k = 0; k < 4; ++k) {
        const unsigned m = __ballot_sync(0xffffffffu, (b >> k) & 1u);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counts[k], __popc(m));
    }
}

__global__ void vs_pose_pick_kernel(const uint8_t* __restrict__ bits, int n, int w, uint8_t* __restrict__ mask)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) mask[i] = (bits[i] >> w) & 1;
}

}  // namespace

#define GEOM_CUDA(call)                                                                  \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) return vs_cuda_fail(e__, #call, __FILE__, __LINE__);     \
    } while (0)

extern "C" int vslam_triangulate(const double* h_P1, const double* h_P2, const float* d_p1, const float* d_p2, int n,
                                 float* d_X, void* stream)
{
    VS_REQUIRE(h_P1 && h_P2 && n >= 0, "bad argument");
    if (n == 0) return VSLAM_OK;
    VS_REQUIRE(d_p1 && d_p2 && d_X, "null argument");
    Proj P1, P2;
    memcpy(P1.m, h_P1, sizeof(P1.m));
    memcpy(P2.m, h_P2, sizeof(P2.m));
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_triangulate_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P1, P2, (const float2*)d_p1,
                                                                            (const float2*)d_p2, n, d_X);
    GEOM_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

extern "C" int vslam_triangulate_host(const double* h_P1, const double* h_P2, const float* h_p1, const float* h_p2,
                                      int n, float* h_X)
{
    VS_REQUIRE(h_P1 && h_P2 && n >= 0, "bad argument");
    if (n == 0) return VSLAM_OK;
    VS_REQUIRE(h_p1 && h_p2 && h_X, "null argument");
    void *a, *b, *x;
    int rc;
    if ((rc = vs_scratch(0, (size_t)n * 8, &a)) || (rc = vs_scratch(1, (size_t)n * 8, &b)) ||
        (rc = vs_scratch(2, (size_t)n * 16, &x)))
        return rc;
    GEOM_CUDA(cudaMemcpy(a, h_p1, (size_t)n * 8, cudaMemcpyHostToDevice));
    GEOM_CUDA(cudaMemcpy(b, h_p2, (size_t)n * 8, cudaMemcpyHostToDevice));
    if ((rc = vslam_triangulate(h_P1, h_P2, (const float*)a, (const float*)b, n, (float*)x, nullptr))) return rc;
    GEOM_CUDA(cudaMemcpy(h_X, x, (size_t)n * 16, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}

extern "C" int vslam_recover_pose(const double* h_E, const float* d_p1, const float* d_p2, int n, const double* h_K,
                                  uint8_t* d_mask_inout, double* h_R, double* h_t, int* n_good, void* stream)
{
    VS_REQUIRE(h_E && h_K && h_R && h_t && n >= 0, "bad argument");
    VS_REQUIRE(n == 0 || (d_p1 && d_p2), "null points");
    VS_REQUIRE(h_K[0] != 0.0 && h_K[4] != 0.0, "singular camera matrix");
    cudaStream_t s = (cudaStream_t)stream;
    double R1[9], R2[9], t[3];
    decompose_essential(h_E, R1, R2, t);
    const Proj c[4] = {make_proj(R1, t, 1.0), make_proj(R2, t, 1.0), make_proj(R1, t, -1.0), make_proj(R2, t, -1.0)};
    int good[4] = {0, 0, 0, 0};
    void *bits = nullptr, *cnt = nullptr;
    if (n > 0) {
        int rc;
        if ((rc = vs_scratch(3, (size_t)n, &bits)) || (rc = vs_scratch(4, 4 * sizeof(int32_t), &cnt))) return rc;
        const Intr K = {h_K[0], h_K[4], h_K[2], h_K[5]};
        GEOM_CUDA(cudaMemsetAsync(cnt, 0, 4 * sizeof(int32_t), s));
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_pose_points_kernel<<<(n + 127) / 128, 128, 0, s>>>(c[0], c[1], c[2], c[3], K, (const float2*)d_p1,
                                                             (const float2*)d_p2, n, 50.0, d_mask_inout,
                                                             (uint8_t*)bits, (int32_t*)cnt);
        GEOM_CUDA(cudaGetLastError());
        int32_t h_cnt[4];
        GEOM_CUDA(cudaMemcpyAsync(h_cnt, cnt, sizeof(h_cnt), cudaMemcpyDeviceToHost, s));
        GEOM_CUDA(cudaStreamSynchronize(s));                       // the winner is decided on the host (4 numbers)
        for (int k = 0; k < 4; ++k) good[k] = h_cnt[k];
    }
    const int w = pose_winner(good);
    memcpy(h_R, (w == 0 || w == 2) ? R1 : R2, sizeof(R1));
    const double sgn = w >= 2 ? -1.0 : 1.0;
    for (int k = 0; k < 3; ++k) h_t[k] = sgn * t[k];
    if (d_mask_inout && n > 0) {
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_pose_pick_kernel<<<(n + 255) / 256, 256, 0, s>>>((const uint8_t*)bits, n, w, d_mask_inout);
        GEOM_CUDA(cudaGetLastError());
    }
    if (n_good) *n_good = good[w];
    return VSLAM_OK;
}

extern "C" int vslam_recover_pose_host(const double* h_E, const float* h_p1, const float* h_p2, int n,
                                       const double* h_K, uint8_t* h_mask_inout, double* h_R, double* h_t,
                                       int* n_good)
{
    VS_REQUIRE(h_E && h_K && h_R && h_t && n >= 0, "bad argument");
    VS_REQUIRE(n == 0 || (h_p1 && h_p2), "null points");
    void *a = nullptr, *b = nullptr, *m = nullptr;
    if (n > 0) {
        int rc;
        if ((rc = vs_scratch(0, (size_t)n * 8, &a)) || (rc = vs_scratch(1, (size_t)n * 8, &b)) ||
            (rc = vs_scratch(2, (size_t)n, &m)))
            return rc;
        GEOM_CUDA(cudaMemcpy(a, h_p1, (size_t)n * 8, cudaMemcpyHostToDevice));
        GEOM_CUDA(cudaMemcpy(b, h_p2, (size_t)n * 8, cudaMemcpyHostToDevice));
        if (h_mask_inout) GEOM_CUDA(cudaMemcpy(m, h_mask_inout, (size_t)n, cudaMemcpyHostToDevice));
    }
    int rc = vslam_recover_pose(h_E, (const float*)a, (const float*)b, n, h_K, h_mask_inout && n > 0 ? (uint8_t*)m : nullptr,
                                h_R, h_t, n_good, nullptr);
    if (rc) return rc;
    if (h_mask_inout && n > 0) GEOM_CUDA(cudaMemcpy(h_mask_inout, m, (size_t)n, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}
