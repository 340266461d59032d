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

struct DevMem {
    void* p = nullptr;
    ~DevMem() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 16); }
    template <class T> T* as() { return (T*)p; }
};

__global__ void vs_triangulate_kernel(Proj P1, Proj P2, const float2* __restrict__ p1, const float2* __restrict__ p2,
                                      int n, float* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double X[4];
    dlt_point(P1.m, P2.m, p1[i].x, p1[i].y, p2[i].x, p2[i].y, X);
    for (int r = 0; r < 4; ++r) out[(size_t)r * n + i] = (float)X[r];
}

// cheirality bits of the 4 candidates for each correspondence (normalised double points)
__global__ void vs_pose_kernel(Proj c0, Proj c1, Proj c2, Proj c3, const double2* __restrict__ p1,
                               const double2* __restrict__ p2, int n, double dist, uint8_t* __restrict__ bits)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Proj cand[4] = {c0, c1, c2, c3};
    const unsigned b = pose_bits(cand, p1[i].x, p1[i].y, p2[i].x, p2[i].y, dist);
    bits[i] = (uint8_t)b;
}

}  // namespace

#define GEOM_CUDA(call)                                                                  \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) return vs_cuda_fail(e__, #call, __FILE__, __LINE__);     \
    } while (0)

extern "C" int vslam_triangulate_host(const double* h_P1, const double* h_P2, const float* h_p1, const float* h_p2,
                                      int n, float* h_X)
{
    VS_REQUIRE(h_P1 && h_P2 && n >= 0, "bad argument");
    if (n == 0) return VSLAM_OK;
    VS_REQUIRE(h_p1 && h_p2 && h_X, "null argument");
    Proj P1, P2;
    memcpy(P1.m, h_P1, sizeof(P1.m));
    memcpy(P2.m, h_P2, sizeof(P2.m));
    DevMem a, b, x;
    GEOM_CUDA(a.alloc((size_t)n * 8));
    GEOM_CUDA(b.alloc((size_t)n * 8));
    GEOM_CUDA(x.alloc((size_t)n * 16));
    GEOM_CUDA(cudaMemcpy(a.p, h_p1, (size_t)n * 8, cudaMemcpyHostToDevice));
    GEOM_CUDA(cudaMemcpy(b.p, h_p2, (size_t)n * 8, cudaMemcpyHostToDevice));
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_triangulate_kernel<<<(n + 127) / 128, 128>>>(P1, P2, a.as<float2>(), b.as<float2>(), n, x.as<float>());
    GEOM_CUDA(cudaGetLastError());
    GEOM_CUDA(cudaMemcpy(h_X, x.p, (size_t)n * 16, cudaMemcpyDeviceToHost));
    return VSLAM_OK;
}

extern "C" int vslam_recover_pose_host(const double* h_E, const float* h_p1, const float* h_p2, int n,
                                       const double* h_K, uint8_t* h_mask_inout, double* h_R, double* h_t,
                                       int* n_good)
{
    VS_REQUIRE(h_E && h_K && h_R && h_t && n >= 0, "bad argument");
    VS_REQUIRE(n == 0 || (h_p1 && h_p2), "null points");
    double R1[9], R2[9], t[3];
    decompose_essential(h_E, R1, R2, t);
    const Proj c[4] = {make_proj(R1, t, 1.0), make_proj(R2, t, 1.0), make_proj(R1, t, -1.0), make_proj(R2, t, -1.0)};
    std::vector<uint8_t> bits(n > 0 ? n : 1, 0);
    if (n > 0) {
        const double fx = h_K[0], fy = h_K[4], cx = h_K[2], cy = h_K[5];
        std::vector<double> q1((size_t)2 * n), q2((size_t)2 * n);
        for (int i = 0; i < n; ++i) {
            q1[2 * i] = ((double)h_p1[2 * i] - cx) / fx; q1[2 * i + 1] = ((double)h_p1[2 * i + 1] - cy) / fy;
            q2[2 * i] = ((double)h_p2[2 * i] - cx) / fx; q2[2 * i + 1] = ((double)h_p2[2 * i + 1] - cy) / fy;
        }
        DevMem a, b, d;
        GEOM_CUDA(a.alloc((size_t)n * 16));
        GEOM_CUDA(b.alloc((size_t)n * 16));
        GEOM_CUDA(d.alloc(n));
        GEOM_CUDA(cudaMemcpy(a.p, q1.data(), (size_t)n * 16, cudaMemcpyHostToDevice));
        GEOM_CUDA(cudaMemcpy(b.p, q2.data(), (size_t)n * 16, cudaMemcpyHostToDevice));
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_pose_kernel<<<(n + 63) / 64, 64>>>(c[0], c[1], c[2], c[3], a.as<double2>(), b.as<double2>(), n, 50.0,
                                              d.as<uint8_t>());
        GEOM_CUDA(cudaGetLastError());
        GEOM_CUDA(cudaMemcpy(bits.data(), d.p, n, cudaMemcpyDeviceToHost));
    }
    int good[4] = {0, 0, 0, 0};
    for (int i = 0; i < n; ++i) {
        if (h_mask_inout && !h_mask_inout[i]) bits[i] = 0;
        for (int k = 0; k < 4; ++k) good[k] += (bits[i] >> k) & 1;
    }
    const int w = pose_winner(good);
    memcpy(h_R, (w == 0 || w == 2) ? R1 : R2, sizeof(R1));
    const double sgn = w >= 2 ? -1.0 : 1.0;
    for (int k = 0; k < 3; ++k) h_t[k] = sgn * t[k];
    if (h_mask_inout)
        for (int i = 0; i < n; ++i) h_mask_inout[i] = (bits[i] >> w) & 1;
    if (n_good) *n_good = good[w];
    return VSLAM_OK;
}
