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

namespace {

struct DevMem {
    void* p = nullptr;
    ~DevMem() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 16); }
    template <class T> T* as() { return (T*)p; }
};

struct Proj { double m[12]; };

__device__ inline void dlt_point(const double* P1, const double* P2, double x1, double y1, double x2, double y2,
                                 double X[4])
{
    double A[4][4], V[4][4];
    for (int c = 0; c < 4; ++c) {
        A[0][c] = x1 * P1[8 + c] - P1[c];
        A[1][c] = y1 * P1[8 + c] - P1[4 + c];
        A[2][c] = x2 * P2[8 + c] - P2[c];
        A[3][c] = y2 * P2[8 + c] - P2[4 + c];
    }
    jacobi_svd<4>(A, V);
    int best = 0;
    double bn = 1e300;
    for (int c = 0; c < 4; ++c) {
        double s = 0.0;
        for (int r = 0; r < 4; ++r) s += A[r][c] * A[r][c];
        if (s < bn) { bn = s; best = c; }
    }
    for (int r = 0; r < 4; ++r) X[r] = V[r][best];
}

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
    const double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    const Proj* cand[4] = {&c0, &c1, &c2, &c3};
    unsigned b = 0;
    for (int c = 0; c < 4; ++c) {
        double Q[4];
        dlt_point(P0, cand[c]->m, p1[i].x, p1[i].y, p2[i].x, p2[i].y, Q);
        bool ok = Q[2] * Q[3] > 0;
        const double x = Q[0] / Q[3], y = Q[1] / Q[3], z = Q[2] / Q[3];
        ok = ok && (z < dist);
        const double* P = cand[c]->m;
        const double z2 = P[8] * x + P[9] * y + P[10] * z + P[11];
        ok = ok && (z2 > 0) && (z2 < dist);
        b |= (unsigned)ok << c;
    }
    bits[i] = (uint8_t)b;
}

void mat3_mul(const double* A, const double* B, double* C)
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += A[3 * i + k] * B[3 * k + j];
            C[3 * i + j] = s;
        }
}

// cv::decomposeEssentialMat: E = U diag(s) Vt, det(U), det(Vt) forced positive,
// R1 = U W Vt, R2 = U W^T Vt, t = U[:, 2]
void decompose_essential(const double* E, double* R1, double* R2, double* t)
{
    double A[3][3], V[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) A[i][j] = E[3 * i + j];
    jacobi_svd<3>(A, V);
    double s[3];
    int order[3] = {0, 1, 2};
    for (int c = 0; c < 3; ++c) s[c] = sqrt(A[0][c] * A[0][c] + A[1][c] * A[1][c] + A[2][c] * A[2][c]);
    for (int a = 0; a < 3; ++a)
        for (int b = a + 1; b < 3; ++b)
            if (s[order[b]] > s[order[a]]) { int tt = order[a]; order[a] = order[b]; order[b] = tt; }
    double U[9], Vm[9];
    for (int c = 0; c < 3; ++c) {
        const int k = order[c];
        for (int r = 0; r < 3; ++r) {
            U[3 * r + c] = s[k] > 1e-300 ? A[r][k] / s[k] : 0.0;
            Vm[3 * r + c] = V[r][k];
        }
    }
    // third left singular vector: orthogonal complement (singular value ~ 0)
    U[2] = U[3] * U[7] - U[6] * U[4];
    U[5] = U[6] * U[1] - U[0] * U[7];
    U[8] = U[0] * U[4] - U[3] * U[1];
    double nn = sqrt(U[2] * U[2] + U[5] * U[5] + U[8] * U[8]);
    if (nn > 0) { U[2] /= nn; U[5] /= nn; U[8] /= nn; }
    if (det3(U) < 0) for (double& v : U) v = -v;
    double Vt[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) Vt[3 * i + j] = Vm[3 * j + i];
    if (det3(Vt) < 0) for (double& v : Vt) v = -v;
    const double W[9] = {0, 1, 0, -1, 0, 0, 0, 0, 1}, Wt[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1};
    double T[9];
    mat3_mul(U, W, T); mat3_mul(T, Vt, R1);
    mat3_mul(U, Wt, T); mat3_mul(T, Vt, R2);
    t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

Proj make_proj(const double* R, const double* t, double sign)
{
    Proj P;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) P.m[4 * r + c] = R[3 * r + c];
        P.m[4 * r + 3] = sign * t[r];
    }
    return P;
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
    int w;
    if (good[0] >= good[1] && good[0] >= good[2] && good[0] >= good[3]) w = 0;
    else if (good[1] >= good[0] && good[1] >= good[2] && good[1] >= good[3]) w = 1;
    else if (good[2] >= good[0] && good[2] >= good[1] && good[2] >= good[3]) w = 2;
    else w = 3;
    memcpy(h_R, (w == 0 || w == 2) ? R1 : R2, sizeof(R1));
    const double sgn = w >= 2 ? -1.0 : 1.0;
    for (int k = 0; k < 3; ++k) h_t[k] = sgn * t[k];
    if (h_mask_inout)
        for (int i = 0; i < n; ++i) h_mask_inout[i] = (bits[i] >> w) & 1;
    if (n_good) *n_good = good[w];
    return VSLAM_OK;
}
