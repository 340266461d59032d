// map_cull.cu -- the per-point test of LocalMapper._cull_map_points (local_mapper.py:205-237) on the device
// ("the per-point reprojection culling loop is a trivially parallel kernel").
//
// The reference walks a Python list of map-point dicts; for each point with at least min_observations
// observations it projects the position into every observing keyframe, P = K [R | t] (utils.py:137-159),
// proj = P [X 1], err = |proj_xy / proj_z - kp.pt| in float64, and drops the point at the first err > 5.0.
// Here the map is flat: points [n][3], CSR observation lists (offset [n+1], keyframe [m], keypoint xy [m][2] in
// float32 like cv2.KeyPoint.pt) and one 3x4 projection matrix per keyframe.  One thread per map point; the lists
// are 2..tens of entries long.  Compiled with -fmad=false: products and sums round like numpy's float64 loop
// (BLAS may order the four-term dot product differently -- the tolerance is written in tests/test_gpu_cull.py).
// A NaN error (point on a keyframe's principal plane) does not reject, exactly like `err > 5.0` in the reference.
#include "common.cuh"


namespace {

__global__ void __launch_bounds__(256)
vs_cull_map_points_kernel(const double* __restrict__ pts, const int32_t* __restrict__ off, const int32_t* __restrict__ okf,
                          const float2* __restrict__ oxy, const double* __restrict__ P, int n, int n_kf, int min_obs,
                          double max_err, uint8_t* __restrict__ valid, double* __restrict__ err)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int lo = off[i], hi = off[i + 1];
    uint8_t ok = 0;
    double worst = 0.0;
    if (hi - lo >= min_obs) {
        const double X = pts[3 * (size_t)i], Y = pts[3 * (size_t)i + 1], Z = pts[3 * (size_t)i + 2];
        ok = 1;
        for (int j = lo; j < hi; ++j) {
            const int k = okf[j];
            if ((unsigned)k >= (unsigned)n_kf) { ok = VSLAM_CULL_BAD_KEYFRAME; break; }
            const double* p = P + 12 * (size_t)k;
            const double px = __ldg(p + 0) * X + __ldg(p + 1) * Y + __ldg(p + 2) * Z + __ldg(p + 3);
            const double py = __ldg(p + 4) * X + __ldg(p + 5) * Y + __ldg(p + 6) * Z + __ldg(p + 7);
            const double pw = __ldg(p + 8) * X + __ldg(p + 9) * Y + __ldg(p + 10) * Z + __ldg(p + 11);
            const float2 kp = oxy[j];
            const double dx = px / pw - (double)kp.x, dy = py / pw - (double)kp.y;
            const double e = sqrt(dx * dx + dy * dy);
            if (e > worst) worst = e;
            if (e > max_err) { ok = 0; break; }
        }
    }
    valid[i] = ok;
    if (err) err[i] = worst;
}

}  // namespace

extern "C" int vslam_cull_map_points(const double* d_points, const int32_t* d_obs_offset, const int32_t* d_obs_keyframe,
                                     const float* d_obs_xy, const double* d_P, int n_points, int n_keyframes,
                                     int min_observations, double max_error, uint8_t* d_valid, double* d_err, void* stream)
{
    VS_REQUIRE(n_points >= 0 && n_keyframes >= 0, "negative count");
    if (n_points == 0) return VSLAM_OK;
    VS_REQUIRE(d_points && d_obs_offset && d_valid, "null pointer");
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_cull_map_points_kernel<<<(n_points + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        d_points, d_obs_offset, d_obs_keyframe, (const float2*)d_obs_xy, d_P, n_points, n_keyframes, min_observations,
        max_error, d_valid, d_err);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

extern "C" int vslam_cull_map_points_host(const double* h_points, const int32_t* h_obs_offset, const int32_t* h_obs_keyframe,
                                          const float* h_obs_xy, const double* h_P, int n_points, int n_keyframes,
                                          int min_observations, double max_error, uint8_t* h_valid, double* h_err)
{
    VS_REQUIRE(n_points >= 0 && n_keyframes >= 0, "negative count");
    if (n_points == 0) return VSLAM_OK;
    VS_REQUIRE(h_points && h_obs_offset && h_valid, "null pointer");
    const int m = h_obs_offset[n_points];
    VS_REQUIRE(h_obs_offset[0] == 0 && m >= 0, "observation offsets must start at 0 and not decrease");
    for (int i = 0; i < n_points; ++i) VS_REQUIRE(h_obs_offset[i + 1] >= h_obs_offset[i], "observation offsets must not decrease");
    VS_REQUIRE(m == 0 || (h_obs_keyframe && h_obs_xy && h_P), "null pointer");
    // one allocation: points | P | err | xy | offsets | keyframes | valid (8-byte types first)
    const size_t b_pts = sizeof(double) * 3 * (size_t)n_points, b_P = sizeof(double) * 12 * (size_t)n_keyframes,
                 b_err = sizeof(double) * (size_t)n_points, b_xy = sizeof(float) * 2 * (size_t)m,
                 b_off = sizeof(int32_t) * ((size_t)n_points + 1), b_kf = sizeof(int32_t) * (size_t)m;
    uint8_t* d = nullptr;
    VS_CUDA(cudaMalloc(&d, b_pts + b_P + b_err + b_xy + b_off + b_kf + (size_t)n_points));
    double* d_pts = (double*)d;
    double* d_P = (double*)(d + b_pts);
    double* d_err = (double*)(d + b_pts + b_P);
    float* d_xy = (float*)(d + b_pts + b_P + b_err);
    int32_t* d_off = (int32_t*)(d + b_pts + b_P + b_err + b_xy);
    int32_t* d_kf = d_off + n_points + 1;
    uint8_t* d_valid = (uint8_t*)(d_kf + m);
    cudaError_t e = cudaMemcpy(d_pts, h_points, b_pts, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && b_P) e = cudaMemcpy(d_P, h_P, b_P, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && m) e = cudaMemcpy(d_xy, h_obs_xy, b_xy, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_off, h_obs_offset, b_off, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && m) e = cudaMemcpy(d_kf, h_obs_keyframe, b_kf, cudaMemcpyHostToDevice);
    int rc = VSLAM_OK;
    if (e == cudaSuccess)
        rc = vslam_cull_map_points(d_pts, d_off, d_kf, d_xy, d_P, n_points, n_keyframes, min_observations, max_error,
                                   d_valid, h_err ? d_err : nullptr, nullptr);
    if (e == cudaSuccess && rc == VSLAM_OK) e = cudaMemcpy(h_valid, d_valid, (size_t)n_points, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && rc == VSLAM_OK && h_err) e = cudaMemcpy(h_err, d_err, b_err, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return vs_cuda_fail(e, "cull_map_points", __FILE__, __LINE__);
    if (rc != VSLAM_OK) return rc;
    for (int i = 0; i < n_points; ++i)
        if (h_valid[i] == VSLAM_CULL_BAD_KEYFRAME) {      // the reference raises IndexError on self.keyframes[kf_id]
            vs_set_error("vslam_cull_map_points_host: map point %d observes a keyframe outside [0, %d)", i, n_keyframes);
            return VSLAM_E_INVALID;
        }
    return VSLAM_OK;
}
