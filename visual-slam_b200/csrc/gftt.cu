// gftt.cu -- the corner stage of ORBExtractor.distribute_keypoints (extractor.py:85-136):
// rows x cols cv2.goodFeaturesToTrack(image, n//64, 0.01, 10, mask=cell) calls.
//
// The reference recomputes the same full-image min-eigenvalue map 64 times (once per cell
// mask).  Here it is computed ONCE and every cell is then one CTA:
//   vs_gftt_sobel_kernel   Dx, Dy = 3x3 Sobel * 1/(4*3*255) in cv2's float operation order
//   vs_gftt_eig_kernel     cov = (DxDx, DxDy, DyDy); 3x3 box sum accumulated in double with
//                          OpenCV's running column sum -- a sequential recurrence down each
//                          column (one thread per column) that has to be replayed for
//                          bit-exactness -- then eig = (a + c) - sqrt((a - c)^2 + b^2)
//   vs_gftt_cell_kernel    per cell: max -> quality threshold, 3x3 local maxima, sort by
//                          (value desc, address desc), greedy min-distance selection
#include <math.h>

#include <vector>

#include "common.cuh"

#define GFTT_MAX_PER_CELL 1024

__global__ void vs_gftt_sobel_kernel(const uint8_t* __restrict__ img, int w, int h, float2* __restrict__ dxy)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const float sc = (float)(1.0 / (4 * 3 * 255.0));
    const float s1 = sc, s2 = __fmul_rn(2.0f, sc);
    const int xm = vs_reflect101(x - 1, w), xp = vs_reflect101(x + 1, w);
    const bool fma_zone = x < w - (w & 31);     // cv2's vector body; the scalar tail has no FMA
    float rowd[3], rows[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const uint8_t* p = img + (size_t)vs_reflect101(y - 1 + k, h) * w;
        const float a = (float)__ldg(p + xm), b = (float)__ldg(p + x), c = (float)__ldg(p + xp);
        rowd[k] = __fsub_rn(c, a);
        float s = __fmul_rn(s1, a);
        if (fma_zone) { s = __fmaf_rn(s2, b, s); s = __fmaf_rn(s1, c, s); }
        else { s = __fadd_rn(s, __fmul_rn(s2, b)); s = __fadd_rn(s, __fmul_rn(s1, c)); }
        rows[k] = s;
    }
    float2 o;
    o.x = __fmaf_rn(s1, __fadd_rn(rowd[0], rowd[2]), __fmul_rn(s2, rowd[1]));
    o.y = __fsub_rn(rows[2], rows[0]);
    dxy[(size_t)y * w + x] = o;
}

__device__ __forceinline__ void gftt_rowsum(const float2* __restrict__ dxy, int w, int h, int yy, int xm, int x,
                                            int xp, double out[3])
{
    const float2* r = dxy + (size_t)vs_reflect101(yy, h) * w;
    const float2 a = __ldg(r + xm), b = __ldg(r + x), c = __ldg(r + xp);
    out[0] = __dadd_rn(__dadd_rn((double)__fmul_rn(a.x, a.x), (double)__fmul_rn(b.x, b.x)), (double)__fmul_rn(c.x, c.x));
    out[1] = __dadd_rn(__dadd_rn((double)__fmul_rn(a.x, a.y), (double)__fmul_rn(b.x, b.y)), (double)__fmul_rn(c.x, c.y));
    out[2] = __dadd_rn(__dadd_rn((double)__fmul_rn(a.y, a.y), (double)__fmul_rn(b.y, b.y)), (double)__fmul_rn(c.y, c.y));
}

__global__ void vs_gftt_eig_kernel(const float2* __restrict__ dxy, int w, int h, float* __restrict__ eig)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= w) return;
    const int xm = vs_reflect101(x - 1, w), xp = vs_reflect101(x + 1, w);
    double r0[3], r1[3], r2[3], sum[3];
    gftt_rowsum(dxy, w, h, -1, xm, x, xp, r0);
    gftt_rowsum(dxy, w, h, 0, xm, x, xp, r1);
    for (int k = 0; k < 3; ++k) sum[k] = __dadd_rn(r0[k], r1[k]);
    for (int y = 0; y < h; ++y) {
        gftt_rowsum(dxy, w, h, y + 1, xm, x, xp, r2);
        float cv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double s0 = __dadd_rn(sum[k], r2[k]);
            cv[k] = (float)s0;
            sum[k] = __dsub_rn(s0, r0[k]);
            r0[k] = r1[k];
            r1[k] = r2[k];
        }
        const float a = __fmul_rn(cv[0], 0.5f), b = cv[1], c = __fmul_rn(cv[2], 0.5f);
        const float t = __fsub_rn(a, c);
        const float u = __fadd_rn(__fmul_rn(t, t), __fmul_rn(b, b));
        eig[(size_t)y * w + x] = __fsub_rn(__fadd_rn(a, c), __fsqrt_rn(u));
    }
}

// One CTA per grid cell.  Dynamic shared memory: cand_cap 64-bit keys (value bits << 32 | address).
__global__ void __launch_bounds__(256)
vs_gftt_cell_kernel(const float* __restrict__ eig, int w, int h, int rows, int cols, int max_per_cell, double quality,
                    double md2, int cand_cap, float* __restrict__ xy, int32_t* __restrict__ cell_count,
                    int32_t* __restrict__ status)
{
    extern __shared__ unsigned long long s_key[];
    __shared__ float s_red[8];
    __shared__ float s_thr;
    __shared__ int s_n;
    __shared__ int s_ax[GFTT_MAX_PER_CELL], s_ay[GFTT_MAX_PER_CELL];
    const int cell = blockIdx.x, ci = cell / cols, cj = cell % cols, tid = threadIdx.x;
    const int ch = h / rows, cw = w / cols;
    const int y0 = ci * ch, x0 = cj * cw;
    // 1. cell maximum -> threshold (float)((double)max * quality)
    float mx = -INFINITY;
    for (int i = tid; i < ch * cw; i += 256) mx = fmaxf(mx, eig[(size_t)(y0 + i / cw) * w + x0 + i % cw]);
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((tid & 31) == 0) s_red[tid >> 5] = mx;
    if (tid == 0) s_n = 0;
    __syncthreads();
    if (tid == 0) {
        float m = s_red[0];
        for (int k = 1; k < 8; ++k) m = fmaxf(m, s_red[k]);
        s_thr = (float)((double)m * quality);
    }
    __syncthreads();
    const float thr = s_thr;
    // 2. candidates: above the threshold, >= all 8 neighbours, not on the image border
    for (int i = tid; i < ch * cw; i += 256) {
        const int y = y0 + i / cw, x = x0 + i % cw;
        if (y < 1 || y >= h - 1 || x < 1 || x >= w - 1) continue;
        const float* p = eig + (size_t)y * w + x;
        const float v = p[0];
        if (!(v > thr)) continue;
        const float nb = fmaxf(fmaxf(fmaxf(p[-w - 1], p[-w]), fmaxf(p[-w + 1], p[-1])),
                               fmaxf(fmaxf(p[1], p[w - 1]), fmaxf(p[w], p[w + 1])));
        if (v >= nb) {
            const int slot = atomicAdd(&s_n, 1);
            if (slot < cand_cap)
                s_key[slot] = ((unsigned long long)__float_as_uint(v) << 32) | (unsigned)(y * w + x);
        }
    }
    __syncthreads();
    int n = s_n;
    if (n > cand_cap) {
        if (tid == 0) atomicExch(status, VSLAM_E_CAPACITY);
        n = cand_cap;
    }
    // 3. sort descending by (value, address): bitonic on the padded list
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    for (int i = n + tid; i < np2; i += 256) s_key[i] = 0ull;
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += 256) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long a = s_key[i], b = s_key[ixj];
                    const bool desc = (i & k) == 0;
                    if ((a < b) == desc) { s_key[i] = b; s_key[ixj] = a; }
                }
            }
            __syncthreads();
        }
    // 4. greedy min-distance selection (acceptance order is the output order)
    if (tid < 32) {
        int acc = 0;
        for (int i = 0; i < n && acc < max_per_cell; ++i) {
            const unsigned addr = (unsigned)s_key[i];
            const int y = addr / w, x = addr % w;
            bool bad = false;
            for (int j = tid; j < acc; j += 32) {
                const double dx = x - s_ax[j], dy = y - s_ay[j];
                bad |= dx * dx + dy * dy < md2;
            }
            if (__any_sync(0xffffffffu, bad)) continue;
            if (tid == 0) {
                s_ax[acc] = x; s_ay[acc] = y;
                xy[((size_t)cell * max_per_cell + acc) * 2] = (float)x;
                xy[((size_t)cell * max_per_cell + acc) * 2 + 1] = (float)y;
            }
            ++acc;
            __syncwarp();
        }
        if (tid == 0) cell_count[cell] = acc;
    }
}

#define GFTT_CUDA(call)                                                                  \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) return vs_cuda_fail(e__, #call, __FILE__, __LINE__);     \
    } while (0)

// Device form: frame on the device (tightly packed, pitch = width), corners per cell in d_xy
// [rows*cols][max_per_cell][2] (cell c's corners in acceptance order at row c), d_cell_count [rows*cols];
// d_status (1 int, may be NULL) is set non-zero when a cell held more local maxima than the candidate staging.
// The min-eigenvalue maps live in the calling thread's scratch (12 bytes per pixel).
extern "C" int vslam_gftt_grid(const uint8_t* d_frame, int height, int width, int rows, int cols, int max_per_cell,
                               double quality, double min_distance, float* d_xy, int32_t* d_cell_count,
                               int32_t* d_status, void* stream)
{
    VS_REQUIRE(d_frame && d_cell_count, "null argument");
    VS_REQUIRE(height >= 3 && width >= 3 && rows >= 1 && cols >= 1 && rows <= height && cols <= width, "bad geometry");
    VS_REQUIRE(max_per_cell >= 0 && max_per_cell <= GFTT_MAX_PER_CELL, "max_per_cell out of range (0..1024)");
    cudaStream_t s = (cudaStream_t)stream;
    const int cells = rows * cols;
    GFTT_CUDA(cudaMemsetAsync(d_cell_count, 0, cells * sizeof(int32_t), s));
    if (d_status) GFTT_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), s));
    if (max_per_cell == 0) return VSLAM_OK;
    VS_REQUIRE(d_xy, "null d_xy");
    const size_t n = (size_t)width * height;
    void* blk = nullptr;
    int rc = vs_scratch(7, n * (sizeof(float2) + sizeof(float)) + 256, &blk);
    if (rc) return rc;
    float2* dxy = (float2*)blk;
    float* eig = (float*)(dxy + n);
    int32_t* st = d_status ? d_status : (int32_t*)(eig + n);
    if (!d_status) GFTT_CUDA(cudaMemsetAsync(st, 0, sizeof(int32_t), s));
    __atomic_fetch_add(&g_vs_launches, 3ull, __ATOMIC_RELAXED);
    vs_gftt_sobel_kernel<<<dim3((width + 127) / 128, height), 128, 0, s>>>(d_frame, width, height, dxy);
    vs_gftt_eig_kernel<<<(width + 63) / 64, 64, 0, s>>>(dxy, width, height, eig);
    const int cell_area = (height / rows) * (width / cols);
    int cand_cap = cell_area < 20480 ? cell_area : 20480;
    if (cand_cap < 32) cand_cap = 32;
    int p2 = 1;
    while (p2 < cand_cap) p2 <<= 1;
    const size_t smem = (size_t)p2 * sizeof(unsigned long long);
    if (smem > 48 * 1024)
        GFTT_CUDA(cudaFuncSetAttribute(vs_gftt_cell_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    vs_gftt_cell_kernel<<<cells, 256, smem, s>>>(eig, width, height, rows, cols, max_per_cell, quality,
                                                 min_distance * min_distance, cand_cap, d_xy, d_cell_count, st);
    GFTT_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

extern "C" int vslam_gftt_grid_host(const uint8_t* h_frame, int height, int width, int rows, int cols,
                                    int max_per_cell, double quality, double min_distance, float* h_xy,
                                    int32_t* h_cell_count, int* n_out)
{
    VS_REQUIRE(h_frame && h_cell_count && n_out, "null argument");
    VS_REQUIRE(height >= 3 && width >= 3 && rows >= 1 && cols >= 1 && rows <= height && cols <= width, "bad geometry");
    VS_REQUIRE(max_per_cell >= 0 && max_per_cell <= GFTT_MAX_PER_CELL, "max_per_cell out of range (0..1024)");
    *n_out = 0;
    const int cells = rows * cols;
    for (int c = 0; c < cells; ++c) h_cell_count[c] = 0;
    if (max_per_cell == 0) return VSLAM_OK;
    VS_REQUIRE(h_xy, "null h_xy");
    const size_t n = (size_t)width * height, xy_bytes = (size_t)cells * max_per_cell * 2 * sizeof(float);
    void *img = nullptr, *out = nullptr;
    int rc;
    if ((rc = vs_scratch(0, n, &img)) || (rc = vs_scratch(1, xy_bytes + (cells + 1) * sizeof(int32_t), &out))) return rc;
    float* d_xy = (float*)out;
    int32_t* d_cnt = (int32_t*)((uint8_t*)out + xy_bytes);
    int32_t* d_st = d_cnt + cells;
    GFTT_CUDA(cudaMemcpy(img, h_frame, n, cudaMemcpyHostToDevice));
    rc = vslam_gftt_grid((const uint8_t*)img, height, width, rows, cols, max_per_cell, quality, min_distance, d_xy, d_cnt,
                         d_st, nullptr);
    if (rc) return rc;
    std::vector<float> tmp((size_t)cells * max_per_cell * 2);
    int32_t status = 0;
    GFTT_CUDA(cudaMemcpy(h_cell_count, d_cnt, cells * sizeof(int32_t), cudaMemcpyDeviceToHost));
    GFTT_CUDA(cudaMemcpy(tmp.data(), d_xy, tmp.size() * sizeof(float), cudaMemcpyDeviceToHost));
    GFTT_CUDA(cudaMemcpy(&status, d_st, sizeof(status), cudaMemcpyDeviceToHost));
    if (status) {
        vs_set_error("vslam_gftt_grid_host: more local maxima in a cell than the candidate staging holds");
        return status;
    }
    int total = 0;
    for (int c = 0; c < cells; ++c) {
        memcpy(h_xy + 2 * (size_t)total, tmp.data() + (size_t)c * max_per_cell * 2,
               (size_t)h_cell_count[c] * 2 * sizeof(float));
        total += h_cell_count[c];
    }
    *n_out = total;
    return VSLAM_OK;
}
