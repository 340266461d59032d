// orb_compute.cu -- BGR->gray (K0) and descriptors at caller-supplied keypoints (a3).
//
// vslam_bgr2gray_host   replaces cv2.cvtColor(image, COLOR_BGR2GRAY) (extractor.py:62,81,101):
//                       gray = (B*3735 + G*19235 + R*9798 + 16384) >> 15  
// vslam_orb_compute_host replaces cv2.ORB.compute(image, keypoints) (extractor.py:83,140):
//                       border filter on the rounded level-0 point, level-major stable order,
//                       rBRIEF on the blurred level with the keypoint's OWN angle and octave.
#include <math.h>

#include <vector>

#include "brief.cuh"

// four pixels per thread: three aligned 32-bit loads (12 BGR bytes), one 32-bit store
__global__ void __launch_bounds__(256)
vs_bgr2gray_kernel(const uint8_t* __restrict__ bgr, uint8_t* __restrict__ gray, size_t n)
{
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // pixel quad
    const size_t i = 4 * q;
    if (i >= n) return;
    if (i + 4 <= n && ((((size_t)bgr) | ((size_t)gray)) & 3) == 0) {
        const uint32_t* p = (const uint32_t*)bgr + 3 * q;
        const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
        // bytes: w0 = B0 G0 R0 B1, w1 = G1 R1 B2 G2, w2 = R2 B3 G3 R3
        const uint32_t g0 = ((w0 & 255) * 3735 + ((w0 >> 8) & 255) * 19235 + ((w0 >> 16) & 255) * 9798 + 16384) >> 15;
        const uint32_t g1 = ((w0 >> 24) * 3735 + (w1 & 255) * 19235 + ((w1 >> 8) & 255) * 9798 + 16384) >> 15;
        const uint32_t g2 = (((w1 >> 16) & 255) * 3735 + (w1 >> 24) * 19235 + (w2 & 255) * 9798 + 16384) >> 15;
        const uint32_t g3 = (((w2 >> 8) & 255) * 3735 + ((w2 >> 16) & 255) * 19235 + (w2 >> 24) * 9798 + 16384) >> 15;
        ((uint32_t*)gray)[q] = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
        return;
    }
    for (size_t k = i; k < n && k < i + 4; ++k) {
        const uint8_t* p = bgr + 3 * k;
        gray[k] = (uint8_t)((p[0] * 3735 + p[1] * 19235 + p[2] * 9798 + 16384) >> 15);
    }
}

int vs_launch_bgr2gray(const uint8_t* d_bgr, uint8_t* d_gray, size_t n, cudaStream_t s)
{
    if (n == 0) return VSLAM_OK;
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_bgr2gray_kernel<<<(unsigned)((n / 4 + 256) / 256), 256, 0, s>>>(d_bgr, d_gray, n);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

// device-pointer form for callers that already hold BGR frames in HBM (e.g. a decoder's output)
extern "C" int vslam_bgr2gray(const uint8_t* d_bgr, size_t n_pixels, uint8_t* d_gray, void* stream)
{
    VS_REQUIRE(d_bgr && d_gray, "null argument");
    return vs_launch_bgr2gray(d_bgr, d_gray, n_pixels, (cudaStream_t)stream);
}

extern "C" int vslam_bgr2gray_host(const uint8_t* h_bgr, int height, int width, uint8_t* h_gray)
{
    VS_REQUIRE(h_bgr && h_gray && height > 0 && width > 0, "bad argument");
    const size_t n = (size_t)height * width;
    uint8_t *d_in = nullptr, *d_out = nullptr;
    VS_CUDA(cudaMalloc(&d_in, 3 * n));
    cudaError_t e = cudaMalloc(&d_out, n);
    if (e == cudaSuccess) e = cudaMemcpy(d_in, h_bgr, 3 * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && vs_launch_bgr2gray(d_in, d_out, n, 0) != VSLAM_OK) e = cudaErrorLaunchFailure;
    if (e == cudaSuccess) e = cudaMemcpy(h_gray, d_out, n, cudaMemcpyDeviceToHost);
    cudaFree(d_in);
    if (d_out) cudaFree(d_out);
    if (e != cudaSuccess) return vs_cuda_fail(e, "bgr2gray", __FILE__, __LINE__);
    return VSLAM_OK;
}

struct VsGivenKp { float x, y, angle; int32_t octave; };

// one warp per keypoint; pyramid (raw + blurred levels) of frame 0 of the last level pass
__global__ void __launch_bounds__(256)
vs_describe_given_kernel(const __grid_constant__ VsPlan plan, const uint8_t* __restrict__ frame, int fpitch,
                         const uint8_t* __restrict__ pyr, const VsGivenKp* __restrict__ kps, int n,
                         uint8_t* __restrict__ desc_out)
{
    __shared__ uint32_t s_pat[256];
    load_pattern(s_pat);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= n) return;
    const VsGivenKp k = kps[i];
    const VsLevel& L = plan.lv[k.octave];
    const uint8_t* raw;
    int rpitch;
    if (k.octave == 0) { raw = frame; rpitch = fpitch; }
    else { raw = pyr + L.raw_off; rpitch = L.pitch; }
    const float inv = __fdiv_rn(1.f, L.scale);
    const int cx = __float2int_rn(__fmul_rn(k.x, inv)), cy = __float2int_rn(__fmul_rn(k.y, inv));
    const float rad = __fmul_rn(k.angle, (float)(3.14159265358979323846 / 180.f));
    const float ca = (float)cos((double)rad), sa = (float)sin((double)rad);
    uint32_t byte = brief_byte(pyr + L.blur_off, raw, L.w, L.h, L.pitch, rpitch, cx, cy, ca, sa, s_pat, lane);
    desc_out[(size_t)i * 32 + lane] = (uint8_t)byte;
}

// Device form of the same: keypoints as device arrays in the caller's order.  Row i of desc_out belongs to keypoint
// i; keep[i] = 1 when the keypoint passes cv2's border test (rounded level-0 point inside [31, W-31) x [31, H-31))
// and its octave is a planned level, else 0 and the row is zero.
__global__ void __launch_bounds__(256)
vs_describe_points_kernel(const __grid_constant__ VsPlan plan, const uint8_t* __restrict__ frame, int fpitch,
                          const uint8_t* __restrict__ pyr, const float* __restrict__ kx, const float* __restrict__ ky,
                          const float* __restrict__ kangle, const int32_t* __restrict__ koct, int n,
                          uint8_t* __restrict__ desc_out, uint8_t* __restrict__ keep)
{
    __shared__ uint32_t s_pat[256];
    load_pattern(s_pat);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= n) return;
    const float x = kx[i], y = ky[i];
    const int oct = koct[i];
    const int rx = __float2int_rn(x), ry = __float2int_rn(y);
    const bool ok = oct >= 0 && oct < plan.n_levels && rx >= VS_EDGE && rx < plan.width - VS_EDGE && ry >= VS_EDGE &&
                    ry < plan.height - VS_EDGE;
    uint32_t byte = 0;
    if (ok) {
        const VsLevel& L = plan.lv[oct];
        const uint8_t* raw;
        int rpitch;
        if (oct == 0) { raw = frame; rpitch = fpitch; }
        else { raw = pyr + L.raw_off; rpitch = L.pitch; }
        const float inv = __fdiv_rn(1.f, L.scale);
        const int cx = __float2int_rn(__fmul_rn(x, inv)), cy = __float2int_rn(__fmul_rn(y, inv));
        const float rad = __fmul_rn(kangle[i], (float)(3.14159265358979323846 / 180.f));
        const float ca = (float)cos((double)rad), sa = (float)sin((double)rad);
        byte = brief_byte(pyr + L.blur_off, raw, L.w, L.h, L.pitch, rpitch, cx, cy, ca, sa, s_pat, lane);
    }
    desc_out[(size_t)i * 32 + lane] = (uint8_t)byte;
    if (lane == 0 && keep) keep[i] = ok ? 1 : 0;
}

extern "C" int vslam_orb_compute(vslam_orb* o, const uint8_t* d_frame, int pitch, const float* d_x, const float* d_y,
                                 const float* d_angle, const int32_t* d_octave, int n, uint8_t* d_desc,
                                 uint8_t* d_keep_mask, void* stream)
{
    VS_REQUIRE(o && d_frame && n >= 0, "bad argument");
    if (n == 0) return VSLAM_OK;
    VS_REQUIRE(d_x && d_y && d_angle && d_octave && d_desc, "null argument");
    VS_REQUIRE(pitch >= o->prm.width, "pitch smaller than width");
    const VsPlan& p = o->plan;
    cudaStream_t s = (cudaStream_t)stream;
    o->last_stream = s; o->last_frames = d_frame; o->last_pitch = pitch; o->last_batch = 1;
    if (o->prof_on != 2) o->prof_count = 0;
    int rc = vs_launch_levels(o, d_frame, 1, pitch, s);           // pyramid + blurred levels of this frame
    if (rc) return rc;
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_describe_points_kernel<<<(n + 7) / 8, 256, 0, s>>>(p, d_frame, pitch, o->d_pyr, d_x, d_y, d_angle, d_octave, n,
                                                        d_desc, d_keep_mask);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

int vs_ensure_host_staging(vslam_orb* o);

extern "C" int vslam_orb_compute_host(vslam_orb* o, const uint8_t* h_frame, const float* h_x, const float* h_y,
                                      const float* h_angle, const int32_t* h_octave, int n, uint8_t* h_desc,
                                      int32_t* h_keep_index, int* n_out)
{
    VS_REQUIRE(o && h_frame && n_out && n >= 0, "bad argument");
    *n_out = 0;
    if (n == 0) return VSLAM_OK;
    VS_REQUIRE(h_x && h_y && h_angle && h_octave && h_desc && h_keep_index, "null argument");
    const VsPlan& p = o->plan;
    // runByImageBorder on the rounded level-0 point, then level-major stable order
    std::vector<VsGivenKp> kept;
    std::vector<int32_t> keep_idx;
    kept.reserve(n);
    for (int l = 0; l < p.n_levels; ++l)
        for (int i = 0; i < n; ++i) {
            if (h_octave[i] != l) continue;
            const int rx = (int)lrintf(h_x[i]), ry = (int)lrintf(h_y[i]);
            if (!(rx >= VS_EDGE && rx < p.width - VS_EDGE && ry >= VS_EDGE && ry < p.height - VS_EDGE)) continue;
            kept.push_back({h_x[i], h_y[i], h_angle[i], l});
            keep_idx.push_back(i);
        }
    for (int i = 0; i < n; ++i)
        if (h_octave[i] < 0 || h_octave[i] >= p.n_levels) {
            vs_set_error("vslam_orb_compute_host: keypoint octave %d outside the %d planned levels", h_octave[i], p.n_levels);
            return VSLAM_E_INVALID;
        }
    const int m = (int)kept.size();
    if (m == 0) return VSLAM_OK;
    int rc = vs_ensure_host_staging(o);
    if (rc) return rc;
    cudaStream_t s = o->stream;
    VS_CUDA(cudaMemcpyAsync(o->d_frames, h_frame, (size_t)p.width * p.height, cudaMemcpyHostToDevice, s));
    o->last_stream = s; o->last_frames = o->d_frames; o->last_pitch = p.width; o->last_batch = 1;
    if (o->prof_on != 2) o->prof_count = 0;
    rc = vs_launch_levels(o, o->d_frames, 1, p.width, s);
    if (rc) return rc;
    VsGivenKp* d_k = nullptr;
    uint8_t* d_d = nullptr;
    if ((rc = vs_scratch(0, (size_t)m * sizeof(VsGivenKp), (void**)&d_k)) || (rc = vs_scratch(1, (size_t)m * 32, (void**)&d_d)))
        return rc;
    cudaError_t e = cudaMemcpyAsync(d_k, kept.data(), (size_t)m * sizeof(VsGivenKp), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) {
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_describe_given_kernel<<<(m + 7) / 8, 256, 0, s>>>(p, o->d_frames, p.width, o->d_pyr, d_k, m, d_d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_desc, d_d, (size_t)m * 32, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return vs_cuda_fail(e, "orb_compute", __FILE__, __LINE__);
    memcpy(h_keep_index, keep_idx.data(), (size_t)m * sizeof(int32_t));
    *n_out = m;
    return VSLAM_OK;
}
