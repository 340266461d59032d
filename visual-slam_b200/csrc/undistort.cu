// undistort.cu -- cv2.undistort(image, camera_matrix, distortion) (utils.py:40-52), the step the reference's
// scripts run between gray conversion and the extractor (tests/test_matcher.py:67-72, test_orb_extractor.py:77,
// run_video.py:149), on the device (the ingest in front of the path).
//
// OpenCV's arithmetic, restated (oracle/undistort_oracle.py holds the numpy form, pinned against cv2 outputs):
//   * the image is processed in stripes of min(max(1, 4096 / width), height) rows; per stripe the new camera
//     matrix (= the camera matrix) gets cy - stripe_start and is inverted with cv::invert's closed 3x3 form;
//   * per pixel, in double: (x, y) = iR (j, i, 1) normalised, radial / tangential / thin-prism model
//     (k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4), u = fx xd + cx, v = fy yd + cy;
//   * the map is fixed point: iu = cvRound(32 u), iv = cvRound(32 v) (half to even), source pixel (iu >> 5, iv >> 5)
//     saturated to int16, 5 fractional bits per axis;
//   * bilinear weights are int16 with 15 fractional bits, w = (32 - a)(32 - b) * 32 ...; at a = b = 0 the weight
//     32768 saturates to 32767 and OpenCV's table fix-up adds the missing 1 to the bottom-right tap;
//   * out = (sum w * src + 2^14) >> 15 with BORDER_CONSTANT 0 outside the image.
// The tilt coefficients (tauX, tauY) are not supported (the reference's configs carry five coefficients).
#include "common.cuh"

namespace {

struct UndistortParams {
    double fx, fy, u0, v0;
    double k[12];        // k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4
    int w, h, pitch_src, pitch_dst, channels, stripe;
};

__device__ __forceinline__ int fetch0(const uint8_t* img, int pitch, int ch, int c, int x, int y, int w, int h)
{
    return ((unsigned)x < (unsigned)w && (unsigned)y < (unsigned)h) ? (int)__ldg(img + (size_t)y * pitch + x * ch + c) : 0;
}

__global__ void __launch_bounds__(256)
vs_undistort_kernel(const UndistortParams P, const uint8_t* __restrict__ src, size_t src_stride,
                    uint8_t* __restrict__ dst, size_t dst_stride)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= P.w) return;
    // inverse of [[fx 0 u0] [0 fy v0 - y0] [0 0 1]] exactly as cv::invert writes it out for 3x3 (row-major S)
    const int y0 = (y / P.stripe) * P.stripe;
    const double S00 = P.fx, S01 = 0.0, S02 = P.u0, S10 = 0.0, S11 = P.fy, S12 = P.v0 - (double)y0, S20 = 0.0, S21 = 0.0, S22 = 1.0;
    double d = S00 * (S11 * S22 - S12 * S21) - S01 * (S10 * S22 - S12 * S20) + S02 * (S10 * S21 - S11 * S20);
    d = 1.0 / d;
    const double ir0 = (S11 * S22 - S12 * S21) * d, ir1 = (S02 * S21 - S01 * S22) * d, ir2 = (S01 * S12 - S02 * S11) * d;
    const double ir3 = (S12 * S20 - S10 * S22) * d, ir4 = (S00 * S22 - S02 * S20) * d, ir5 = (S02 * S10 - S00 * S12) * d;
    const double ir6 = (S10 * S21 - S11 * S20) * d, ir7 = (S01 * S20 - S00 * S21) * d, ir8 = (S00 * S11 - S01 * S10) * d;
    const double i = (double)(y - y0), j = (double)x;
    const double _x = i * ir1 + ir2 + j * ir0, _y = i * ir4 + ir5 + j * ir3, _w = i * ir7 + ir8 + j * ir6;
    const double ww = 1.0 / _w, xn = _x * ww, yn = _y * ww;
    const double x2 = xn * xn, y2 = yn * yn, r2 = x2 + y2, _2xy = 2.0 * xn * yn;
    const double* k = P.k;
    const double kr = (1.0 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2) / (1.0 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2);
    const double xd = xn * kr + k[2] * _2xy + k[3] * (r2 + 2.0 * x2) + k[8] * r2 + k[9] * r2 * r2;
    const double yd = yn * kr + k[2] * (r2 + 2.0 * y2) + k[3] * _2xy + k[10] * r2 + k[11] * r2 * r2;
    const double u = P.fx * xd + P.u0, v = P.fy * yd + P.v0;
    // cvRound(32 u): saturating to int like saturate_cast<int>(double)
    const double us = fmin(fmax(u * 32.0, -2147483648.0), 2147483647.0), vs = fmin(fmax(v * 32.0, -2147483648.0), 2147483647.0);
    const long long iu = isfinite(us) ? __double2ll_rn(us) : 0, iv = isfinite(vs) ? __double2ll_rn(vs) : 0;
    const int sx = (int)max(min(iu >> 5, 32767LL), -32768LL), sy = (int)max(min(iv >> 5, 32767LL), -32768LL);
    const int ax = (int)(iu & 31), ay = (int)(iv & 31);
    int w00 = (32 - ax) * (32 - ay) * 32, w01 = ax * (32 - ay) * 32, w10 = (32 - ax) * ay * 32, w11 = ax * ay * 32;
    if ((ax | ay) == 0) { w00 = 32767; w11 = 1; }
    const uint8_t* s = src + (size_t)b * src_stride;
    uint8_t* o = dst + (size_t)b * dst_stride + (size_t)y * P.pitch_dst + x * P.channels;
    for (int c = 0; c < P.channels; ++c) {
        const int val = (fetch0(s, P.pitch_src, P.channels, c, sx, sy, P.w, P.h) * w00 +
                         fetch0(s, P.pitch_src, P.channels, c, sx + 1, sy, P.w, P.h) * w01 +
                         fetch0(s, P.pitch_src, P.channels, c, sx, sy + 1, P.w, P.h) * w10 +
                         fetch0(s, P.pitch_src, P.channels, c, sx + 1, sy + 1, P.w, P.h) * w11 + (1 << 14)) >> 15;
        o[c] = (uint8_t)min(max(val, 0), 255);
    }
}

int fill_params(UndistortParams& P, int h, int w, int channels, int pitch_src, int pitch_dst, const double* K,
                const double* dist, int n_dist)
{
    VS_REQUIRE(K && h > 0 && w > 0, "bad argument");
    VS_REQUIRE(channels == 1 || channels == 3, "1 or 3 channels");
    VS_REQUIRE(n_dist == 0 || dist, "null distortion coefficients");
    VS_REQUIRE(n_dist == 0 || n_dist == 4 || n_dist == 5 || n_dist == 8 || n_dist == 12,
               "4, 5, 8 or 12 distortion coefficients (the tilt model is not supported)");
    VS_REQUIRE(K[0] != 0.0 && K[4] != 0.0, "singular camera matrix");
    P.fx = K[0]; P.fy = K[4]; P.u0 = K[2]; P.v0 = K[5];
    for (int i = 0; i < 12; ++i) P.k[i] = i < n_dist ? dist[i] : 0.0;
    P.w = w; P.h = h; P.channels = channels; P.pitch_src = pitch_src; P.pitch_dst = pitch_dst;
    int stripe = 4096 / w;
    stripe = stripe < 1 ? 1 : stripe;
    P.stripe = stripe < h ? stripe : h;
    return VSLAM_OK;
}

}  // namespace

extern "C" int vslam_undistort(const uint8_t* d_src, int batch, int height, int width, int channels, int pitch_src,
                               const double* h_K, const double* h_dist, int n_dist, uint8_t* d_dst, int pitch_dst,
                               void* stream)
{
    VS_REQUIRE(d_src && d_dst && batch >= 0, "bad argument");
    VS_REQUIRE(pitch_src >= width * channels && pitch_dst >= width * channels, "pitch smaller than a row");
    UndistortParams P;
    int rc = fill_params(P, height, width, channels, pitch_src, pitch_dst, h_K, h_dist, n_dist);
    if (rc) return rc;
    if (batch == 0) return VSLAM_OK;
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_undistort_kernel<<<dim3((width + 255) / 256, height, batch), 256, 0, (cudaStream_t)stream>>>(
        P, d_src, (size_t)pitch_src * height, d_dst, (size_t)pitch_dst * height);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

extern "C" int vslam_undistort_host(const uint8_t* h_src, int height, int width, int channels, const double* h_K,
                                    const double* h_dist, int n_dist, uint8_t* h_dst)
{
    VS_REQUIRE(h_src && h_dst && height > 0 && width > 0, "bad argument");
    VS_REQUIRE(channels == 1 || channels == 3, "1 or 3 channels");
    const size_t n = (size_t)height * width * channels;
    uint8_t *d_in = nullptr, *d_out = nullptr;
    VS_CUDA(cudaMalloc(&d_in, n));
    cudaError_t e = cudaMalloc(&d_out, n);
    int rc = VSLAM_OK;
    if (e == cudaSuccess) e = cudaMemcpy(d_in, h_src, n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        rc = vslam_undistort(d_in, 1, height, width, channels, width * channels, h_K, h_dist, n_dist, d_out,
                             width * channels, nullptr);
    if (e == cudaSuccess && rc == VSLAM_OK) e = cudaMemcpy(h_dst, d_out, n, cudaMemcpyDeviceToHost);
    cudaFree(d_in);
    if (d_out) cudaFree(d_out);
    if (e != cudaSuccess) return vs_cuda_fail(e, "undistort", __FILE__, __LINE__);
    return rc;
}
