// orb_select.cu -- keypoint selection, orientation and description.
//
//   vs_harris_kernel   first retain-best cut (2*quota by FAST score, ties kept; threshold
//                      from the 256-bin histogram) + Harris response of the survivors (K3/K4)
//   vs_select_kernel   second retain-best cut (quota by Harris response, ties kept; exact
//                      radix select on the order-preserving key), (y, x) sort, staging (K5)
//   vs_describe_kernel one warp per keypoint: intensity-centroid angle with cv::fastAtan2's
//                      polynomial (K6), keypoint record, steered rBRIEF on the blurred level (K7b)
// All of this restates what cv2.ORB.detectAndCompute does behind extractor.py:65.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "brief.cuh"

__device__ __constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__device__ __forceinline__ uint32_t f2key(float f)
{
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// Inclusive suffix sum over 256 values, one per thread (blockDim.x == 256):
// returns sum_{u >= tid} val[u].  s_w: 8 ints of scratch.
__device__ __forceinline__ int suffix_sum_256(int val, int* s_w)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int s = val;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_down_sync(0xffffffffu, s, o);
        if (lane + o < 32) s += t;
    }
    __syncthreads();
    if (lane == 0) s_w[w] = s;
    __syncthreads();
    for (int j = w + 1; j < 8; ++j) s += s_w[j];
    return s;
}

// Harris response of cv2.ORB (7x7 block of 3x3 Sobel, k = 0.04), SURVEY.md 8(a) K4.
// ALIGNED: rows are 4-byte aligned, so the 9 bytes of a row come from three aligned 32-bit loads
// (funnel-shifted by the column's byte offset) instead of nine byte loads -- the kernel is bound by
// load latency, not arithmetic.
template <bool ALIGNED>
__device__ __forceinline__ float harris_response(const uint8_t* img, int pitch, int x0, int y0)
{
    int a = 0, b = 0, c = 0;
    int dxm[7], sxm[7], dxc[7], sxc[7];
    const int xa = (x0 - 4) & ~3, sh = ((x0 - 4) & 3) * 8;
    // rows y0-4 .. y0+4; Sobel at row r uses rows r-1, r, r+1
#pragma unroll
    for (int r = 0; r < 9; ++r) {
        int v[9];
        if (ALIGNED) {
            const uint32_t* p = (const uint32_t*)(img + (size_t)(y0 - 4 + r) * pitch + xa);
            const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
            const uint32_t q0 = __funnelshift_r(w0, w1, sh), q1 = __funnelshift_r(w1, w2, sh), q2 = w2 >> sh;
            v[0] = q0 & 0xff; v[1] = (q0 >> 8) & 0xff; v[2] = (q0 >> 16) & 0xff; v[3] = q0 >> 24;
            v[4] = q1 & 0xff; v[5] = (q1 >> 8) & 0xff; v[6] = (q1 >> 16) & 0xff; v[7] = q1 >> 24;
            v[8] = q2 & 0xff;
        } else {
            const uint8_t* p = img + (size_t)(y0 - 4 + r) * pitch + (x0 - 4);
#pragma unroll
            for (int j = 0; j < 9; ++j) v[j] = __ldg(p + j);
        }
        int dxn[7], sxn[7];
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            dxn[j] = v[j + 2] - v[j];
            sxn[j] = v[j] + 2 * v[j + 1] + v[j + 2];
        }
        if (r >= 2) {
#pragma unroll
            for (int j = 0; j < 7; ++j) {
                int Ix = dxm[j] + 2 * dxc[j] + dxn[j];
                int Iy = sxn[j] - sxm[j];
                a += Ix * Ix;
                b += Iy * Iy;
                c += Ix * Iy;
            }
        }
#pragma unroll
        for (int j = 0; j < 7; ++j) { dxm[j] = dxc[j]; sxm[j] = sxc[j]; dxc[j] = dxn[j]; sxc[j] = sxn[j]; }
    }
    const float scale = 1.f / (4 * 7 * 255.f);
    const float s2 = __fmul_rn(scale, scale);
    const float s4 = __fmul_rn(__fmul_rn(s2, scale), scale);
    float fa = (float)a, fb = (float)b, fc = (float)c;
    float sum = __fadd_rn(fa, fb);
    float r = __fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc));
    r = __fsub_rn(r, __fmul_rn(__fmul_rn(0.04f, sum), sum));
    return __fmul_rn(r, s4);
}

#define VS_HARRIS_CTAS 4
#define VS_HARRIS_LIST 1024

__global__ void __launch_bounds__(256)
vs_harris_kernel(const __grid_constant__ VsPlan plan, const uint8_t* __restrict__ frames, int fpitch,
                 size_t fstride, const uint8_t* __restrict__ pyr, VsCand* __restrict__ cand,
                 int32_t* __restrict__ counters)
{
    __shared__ int s_w[8];
    __shared__ int s_ge[256];
    __shared__ int s_thr;
    const int level = blockIdx.y, b = blockIdx.z, tid = threadIdx.x;
    const VsLevel& L = plan.lv[level];
    int32_t* ctr = counters + ((size_t)b * plan.n_levels + level) * VS_CTR_STRIDE;
    const int ncand = ctr[VS_CTR_NCAND];
    if (ncand == 0) return;
    const int keep = 2 * L.quota;
    // first cut: threshold = keep-th largest FAST score, everything >= it survives
    if (tid == 0) s_thr = keep == 0 ? 256 : 0;
    __syncthreads();
    if (ncand > keep && keep > 0) {
        int ge = suffix_sum_256(ctr[VS_CTR_HIST + tid], s_w);        // #scores >= tid
        s_ge[tid] = ge;
        __syncthreads();
        int ge1 = tid < 255 ? s_ge[tid + 1] : 0;
        if (ge >= keep && ge1 < keep) s_thr = tid;
        __syncthreads();
    }
    const int thr = s_thr;
    const uint8_t* img;
    int pitch;
    if (level == 0) { img = frames + (size_t)b * fstride; pitch = fpitch; }
    else { img = pyr + (size_t)b * plan.pyr_frame_bytes + L.raw_off; pitch = L.pitch; }
    VsCand* cl = cand + (size_t)b * plan.cand_frame_entries + L.cand_off;
    const bool aligned = (((size_t)img | (size_t)pitch) & 3) == 0;
    // Only the candidates at or above the cut get a response (a quarter of them on textured frames), so they are
    // compacted into a shared list first and the 400-instruction response runs on full warps.
    __shared__ uint2 s_todo[VS_HARRIS_LIST];      // (index in the candidate list, y << 16 | x)
    __shared__ int s_ntodo;
    if (tid == 0) s_ntodo = 0;
    __syncthreads();
    int nsurv = 0;
    const int lane = tid & 31;
    for (int base = blockIdx.x * 256; base < ncand; base += VS_HARRIS_CTAS * 256) {
        const int i = base + tid;
        bool sv = false;
        VsCand c;
        if (i < ncand) {
            c = cl[i];
            sv = (int)c.v >= thr;
            if (!sv) cl[i].v = 0;
        }
        const uint32_t m = __ballot_sync(0xffffffffu, sv);
        int pos = 0;
        if (lane == 0 && m) pos = atomicAdd(&s_ntodo, __popc(m));
        pos = __shfl_sync(0xffffffffu, pos, 0);
        if (sv) s_todo[pos + __popc(m & ((1u << lane) - 1u))] = make_uint2((uint32_t)i, ((uint32_t)c.y << 16) | c.x);
        __syncthreads();
        const int n = s_ntodo;
        const bool last = base + VS_HARRIS_CTAS * 256 >= ncand;
        if (n > VS_HARRIS_LIST - 256 || last) {       // the next chunk might not fit: drain
            for (int j = tid; j < n; j += 256) {
                const uint2 e = s_todo[j];
                const int x = (int)(e.y & 0xffff), y = (int)(e.y >> 16);
                uint32_t key = f2key(aligned ? harris_response<true>(img, pitch, x, y) : harris_response<false>(img, pitch, x, y));
                if (key == 0) key = 1;
                cl[e.x].v = key;
                ++nsurv;
            }
            __syncthreads();
            if (tid == 0) s_ntodo = 0;
            __syncthreads();
        }
    }
    // survivor count (warp-aggregated)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nsurv += __shfl_down_sync(0xffffffffu, nsurv, o);
    if ((tid & 31) == 0 && nsurv) atomicAdd(&ctr[VS_CTR_NSURV], nsurv);
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
vs_select_kernel(const __grid_constant__ VsPlan plan, const VsCand* __restrict__ cand,
                 VsCand* __restrict__ stage, int32_t* __restrict__ counters,
                 int32_t* __restrict__ status)
{
    extern __shared__ unsigned long long s_sort[];
    __shared__ int s_hist[256];
    __shared__ int s_w[8];
    __shared__ int s_digit, s_k, s_cnt;
    const int level = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const VsLevel& L = plan.lv[level];
    int32_t* ctr = counters + ((size_t)b * plan.n_levels + level) * VS_CTR_STRIDE;
    const int ncand = ctr[VS_CTR_NCAND], nsurv = ctr[VS_CTR_NSURV];
    if (nsurv == 0 || L.quota == 0) return;   // NFINAL stays 0 (counters are zeroed per batch)
    const VsCand* cl = cand + (size_t)b * plan.cand_frame_entries + L.cand_off;

    uint32_t thr_key = 1;
    if (nsurv > L.quota) {
        // exact quota-th largest key by 4 passes of 8-bit radix select
        uint32_t prefix = 0, mask = 0;
        if (tid == 0) s_k = L.quota;
        for (int shift = 24; shift >= 0; shift -= 8) {
            s_hist[tid] = 0;
            __syncthreads();
            for (int i = tid; i < ncand; i += 256) {
                uint32_t k = cl[i].v;
                if (k != 0 && (k & mask) == prefix) atomicAdd(&s_hist[(k >> shift) & 255], 1);
            }
            __syncthreads();
            const int k_need = s_k;
            int ge = suffix_sum_256(s_hist[tid], s_w);
            int gt = ge - s_hist[tid];
            if (ge >= k_need && gt < k_need) { s_digit = tid; s_k = k_need - gt; }
            __syncthreads();
            prefix |= (uint32_t)s_digit << shift;
            mask |= 255u << shift;
        }
        thr_key = prefix;
    }
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    const int cap = L.stage_cap;
    for (int i = tid; i < ncand; i += 256) {
        VsCand c = cl[i];
        if (c.v >= thr_key && c.v != 0) {
            int slot = atomicAdd(&s_cnt, 1);
            if (slot < cap)
                s_sort[slot] = ((unsigned long long)(((uint32_t)c.y << 16) | c.x) << 32) | c.v;
        }
    }
    __syncthreads();
    int n = s_cnt;
    if (n > cap) {
        if (tid == 0) atomicExch(status, VSLAM_E_CAPACITY);
        n = cap;
    }
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    for (int i = n + tid; i < np2; i += 256) s_sort[i] = ~0ull;
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += 256) {
                int ixj = i ^ j;
                if (ixj > i) {
                    unsigned long long a = s_sort[i], c = s_sort[ixj];
                    bool up = (i & k) == 0;
                    if ((a > c) == up) { s_sort[i] = c; s_sort[ixj] = a; }
                }
            }
            __syncthreads();
        }
    VsCand* st = stage + (size_t)b * plan.stage_frame_entries + L.stage_off;
    for (int i = tid; i < n; i += 256) {
        unsigned long long e = s_sort[i];
        VsCand c;
        c.x = (uint16_t)((e >> 32) & 0xffff);
        c.y = (uint16_t)(e >> 48);
        c.v = __float_as_uint(key2f((uint32_t)e));
        st[i] = c;
    }
    if (tid == 0) ctr[VS_CTR_NFINAL] = n;
}

// ---------------------------------------------------------------------------
// cv::fastAtan2 in degrees, float32 without contraction.
__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float kdeg = (float)(180.0 / 3.14159265358979323846);
    const float p1 = __fmul_rn(0.9997878412794807f, kdeg), p3 = __fmul_rn(-0.3258083974640975f, kdeg),
                p5 = __fmul_rn(0.1555786518463281f, kdeg), p7 = __fmul_rn(-0.04432655554792128f, kdeg);
    const float eps = (float)2.2204460492503131e-16;
    float ax = fabsf(x), ay = fabsf(y), a, c, c2, t;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        t = __fmul_rn(p7, c2); t = __fadd_rn(t, p5); t = __fmul_rn(t, c2); t = __fadd_rn(t, p3);
        t = __fmul_rn(t, c2); t = __fadd_rn(t, p1);
        a = __fmul_rn(t, c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        t = __fmul_rn(p7, c2); t = __fadd_rn(t, p5); t = __fmul_rn(t, c2); t = __fadd_rn(t, p3);
        t = __fmul_rn(t, c2); t = __fadd_rn(t, p1);
        a = __fsub_rn(90.f, __fmul_rn(t, c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// rBRIEF pattern as float4 {x0,y0,x1,y1}, transposed: lane j (descriptor byte j) reads test 8j+t
// at [t*32 + j] with one coalesced 128-bit load (4 KB, L1 resident)
__device__ const float4 g_patf_t[256] = {
#include "rbrief_pattern_t.inc"
};

#define DESC_R 18                     // blurred window radius: a rotated pattern point stays within |(13,13)| = 18.4
#define DESC_WIN (2 * DESC_R + 1)     // 37 rows
#define DESC_WIN_S 40                 // bytes per row (37 + up to 3 of alignment slack)
#define IC_WIN (2 * VS_HALF_PATCH + 1)   // 31 rows of the unblurred level for the intensity centroid
#define IC_WIN_S 36                   // bytes per row (31 + up to 3 of alignment slack)

// Copy a rows x (4*words)-byte window starting at the 4-byte aligned column xs of row ys into shared
// memory.  Lane = (row group, word): the row pointer advances by a fixed stride, no per-load index math.
template <int ROWS, int WORDS, bool ALIGNED>
__device__ __forceinline__ void stage_window(uint8_t* win, const uint8_t* img, int pitch, int xs, int ys, int lane)
{
    constexpr int RG = 32 / WORDS;                  // rows per pass
    const int r0 = lane / WORDS, cw = lane - r0 * WORDS;
    if (r0 >= RG) return;
    const uint8_t* src = img + (size_t)(ys + r0) * pitch + xs + 4 * cw;
    uint32_t* dst = (uint32_t*)win + r0 * WORDS + cw;
    const size_t step = (size_t)pitch * RG;
    if (ALIGNED) {
        // aligned rows: asynchronous 4-byte copies global -> shared (LDGSTS), one instruction per word instead of a
        // load and a store; the caller's __syncwarp follows the wait
        uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
#pragma unroll
        for (int k = 0; k < (ROWS + RG - 1) / RG; ++k) {
            if (k * RG + r0 < ROWS)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src) : "memory");
            src += step;
            d += RG * WORDS * 4;
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
        return;
    }
    uint32_t v[(ROWS + RG - 1) / RG];
#pragma unroll
    for (int k = 0; k < (ROWS + RG - 1) / RG; ++k) {            // all loads first, then the stores
        v[k] = 0;
        if (k * RG + r0 < ROWS)
            v[k] = (uint32_t)__ldg(src) | ((uint32_t)__ldg(src + 1) << 8) | ((uint32_t)__ldg(src + 2) << 16) | ((uint32_t)__ldg(src + 3) << 24);
        src += step;
    }
#pragma unroll
    for (int k = 0; k < (ROWS + RG - 1) / RG; ++k)
        if (k * RG + r0 < ROWS) dst[k * RG * WORDS] = v[k];
}

// One warp per keypoint, VS_DESC_PER_WARP keypoints in turn.  Keypoints are at least 31 px inside
// their level, so both windows lie inside the image.
#define VS_DESC_PER_WARP 8

__global__ void __launch_bounds__(256)
vs_describe_kernel(const __grid_constant__ VsPlan plan, const uint8_t* __restrict__ frames, int fpitch,
                   size_t fstride, const uint8_t* __restrict__ pyr, const VsCand* __restrict__ stage,
                   const int32_t* __restrict__ counters, vslam_keypoint* __restrict__ kp_out,
                   uint8_t* __restrict__ desc_out, int32_t* __restrict__ count_out)
{
    __shared__ __align__(16) uint8_t s_blur[8][DESC_WIN * DESC_WIN_S];
    __shared__ __align__(16) uint8_t s_raw[8][IC_WIN * IC_WIN_S];
    const int b = blockIdx.y, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int32_t* ctr = counters + (size_t)b * plan.n_levels * VS_CTR_STRIDE;
    // start of each level's keypoints in the frame's output rows (prefix sums of the final counts)
    __shared__ int s_start[VS_MAX_LEVELS + 1];
    if (threadIdx.x < 32) {
        const int l = threadIdx.x;
        int n = l < plan.n_levels ? ctr[l * VS_CTR_STRIDE + VS_CTR_NFINAL] : 0, incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (l >= o) incl += t;
        }
        if (l <= VS_MAX_LEVELS) s_start[l] = incl - n;          // s_start[n_levels..] = total
    }
    __syncthreads();
    const int total = s_start[plan.n_levels];
    if (blockIdx.x == 0 && threadIdx.x == 0) count_out[b] = total;
    const bool frames_aligned = ((((size_t)frames) | (size_t)fpitch | fstride) & 3) == 0;
    const uint8_t* fp = pyr + (size_t)b * plan.pyr_frame_bytes;
    uint8_t* wb = s_blur[wid];
    uint8_t* wr = s_raw[wid];

    // disc mask of this lane's patch row (pass A): byte k of word j <-> u = 4j + k - 15
    uint32_t ic_mask[8];
    {
        const int v = lane - VS_HALF_PATCH;
        const int hw = lane < IC_WIN ? c_umax[v < 0 ? -v : v] : -1;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint32_t mk = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int uu = 4 * j + k - VS_HALF_PATCH;
                if ((uu < 0 ? -uu : uu) <= hw) mk |= 255u << (8 * k);
            }
            ic_mask[j] = mk;
        }
    }
    // Pass A: intensity-centroid angle of this warp's keypoints, one after the other (lane k keeps keypoint
    // k's angle, centre and level).  Then sin/cos in double for all of them at once -- lane k evaluates
    // keypoint k, so the warp pays for one evaluation instead of one per keypoint.  Pass B: steered rBRIEF.
    const int i0 = (blockIdx.x * 8 + wid) * VS_DESC_PER_WARP;
    const int nk = min(VS_DESC_PER_WARP, min(total, plan.kp_cap) - i0);   // warp-uniform; <= 0: nothing to do
    int level = 0;
    float m_ang = 0.f;
    int m_cx = 0, m_cy = 0, m_level = 0;
    for (int k = 0; k < nk; ++k) {
        const int i = i0 + k;
        while (level + 1 < plan.n_levels && i >= s_start[level + 1]) ++level;   // warp-uniform, monotone in k
        const int j = i - s_start[level];
        const bool raw_aligned = level > 0 || frames_aligned;
        const VsLevel& L = plan.lv[level];
        const VsCand c = stage[(size_t)b * plan.stage_frame_entries + L.stage_off + j];
        const uint8_t* raw;
        int rpitch;
        if (level == 0) { raw = frames + (size_t)b * fstride; rpitch = fpitch; }
        else { raw = fp + L.raw_off; rpitch = L.pitch; }
        // ---- intensity-centroid angle (K6) on the unblurred level
        const int x0 = c.x, y0 = c.y;
        const int rxs = (x0 - VS_HALF_PATCH) & ~3;
        int m10 = 0, m01 = 0;
        __syncwarp();
        if (raw_aligned) stage_window<IC_WIN, IC_WIN_S / 4, true>(wr, raw, rpitch, rxs, y0 - VS_HALF_PATCH, lane);
        else stage_window<IC_WIN, IC_WIN_S / 4, false>(wr, raw, rpitch, rxs, y0 - VS_HALF_PATCH, lane);
        __syncwarp();
        // lane = patch row: the row's nine staged words (row stride 9 words: conflict-free), funnel-shifted to the
        // patch's first column, masked to the disc and reduced with two dot products per word (sum of u*I, sum of I)
        if (lane < IC_WIN) {
            const uint32_t* p = (const uint32_t*)(wr + lane * IC_WIN_S);
            uint32_t w[9];
#pragma unroll
            for (int j = 0; j < 9; ++j) w[j] = p[j];
            const int sh = ((x0 - VS_HALF_PATCH) & 3) * 8;
            int rs = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t q = __funnelshift_r(w[j], w[j + 1], sh) & ic_mask[j];
                // bytes of word j <-> u = 4j-15 .. 4j-12
                const uint32_t wu = (uint32_t)((4 * j - 15) & 255) | ((uint32_t)((4 * j - 14) & 255) << 8) |
                                    ((uint32_t)((4 * j - 13) & 255) << 16) | ((uint32_t)((4 * j - 12) & 255) << 24);
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(q), "r"(wu));
                rs = (int)__dp4a(q, 0x01010101u, (uint32_t)rs);
            }
            m01 = (lane - VS_HALF_PATCH) * rs;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, o);
            m01 += __shfl_xor_sync(0xffffffffu, m01, o);
        }
        const float ang = fast_atan2_deg((float)m01, (float)m10);
        const float kx = __fmul_rn((float)x0, L.scale), ky = __fmul_rn((float)y0, L.scale);
        if (lane == 0) {
            vslam_keypoint* ko = kp_out + (size_t)b * plan.kp_cap + i;
            ko->x = kx; ko->y = ky; ko->size = __fmul_rn(31.f, L.scale);
            ko->angle = ang; ko->response = __uint_as_float(c.v); ko->octave = level;
        }
        // centre of the rBRIEF pattern: cvRound(pt * (1/scale))
        const float inv = __fdiv_rn(1.f, L.scale);
        if (lane == k) {
            m_ang = ang; m_level = level;
            m_cx = __float2int_rn(__fmul_rn(kx, inv)); m_cy = __float2int_rn(__fmul_rn(ky, inv));
        }
    }
    if (nk <= 0) return;
    // rotation from the angle in radians (K7b): (float)cos / (float)sin of the float32 angle, evaluated in double
    float m_ca, m_sa;
    {
        const float rad = __fmul_rn(m_ang, (float)(3.14159265358979323846 / 180.f));
        double sd, cd;
        sincos((double)rad, &sd, &cd);
        m_ca = (float)cd; m_sa = (float)sd;
    }
    for (int k = 0; k < nk; ++k) {
        const int i = i0 + k;
        const float ca = __shfl_sync(0xffffffffu, m_ca, k), sa = __shfl_sync(0xffffffffu, m_sa, k);
        const int cx = __shfl_sync(0xffffffffu, m_cx, k), cy = __shfl_sync(0xffffffffu, m_cy, k);
        const VsLevel& L = plan.lv[__shfl_sync(0xffffffffu, m_level, k)];
        const int xs = (cx - DESC_R) & ~3;
        __syncwarp();
        stage_window<DESC_WIN, DESC_WIN_S / 4, true>(wb, fp + L.blur_off, L.pitch, xs, cy - DESC_R, lane);
        __syncwarp();
        // cvRound(f) as the low mantissa bits of f + 1.5 * 2^23 (round half to even, |f| < 2^22)
        const float magic = 12582912.0f;
        const uint32_t off0 = (uint32_t)(DESC_R * DESC_WIN_S + (cx - xs)) - 0x4B400000u * (DESC_WIN_S + 1);
        uint32_t byte = 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const float4 p = __ldg(&g_patf_t[t * 32 + lane]);
            const float fx0 = __fsub_rn(__fmul_rn(p.x, ca), __fmul_rn(p.y, sa)), fy0 = __fadd_rn(__fmul_rn(p.x, sa), __fmul_rn(p.y, ca));
            const float fx1 = __fsub_rn(__fmul_rn(p.z, ca), __fmul_rn(p.w, sa)), fy1 = __fadd_rn(__fmul_rn(p.z, sa), __fmul_rn(p.w, ca));
            const uint32_t i0w = __float_as_uint(__fadd_rn(fy0, magic)) * DESC_WIN_S + __float_as_uint(__fadd_rn(fx0, magic)) + off0;
            const uint32_t i1w = __float_as_uint(__fadd_rn(fy1, magic)) * DESC_WIN_S + __float_as_uint(__fadd_rn(fx1, magic)) + off0;
            byte |= (uint32_t)(wb[i0w] < wb[i1w]) << t;
        }
        desc_out[((size_t)b * plan.kp_cap + i) * 32 + lane] = (uint8_t)byte;
    }
}

// ---------------------------------------------------------------------------
// The same kernel with both windows staged by TMA (cp.async.bulk.tensor.3d over (x, y, frame), one instruction per
// window issued by one lane).  The hand-staged kernel above is bound by the LSU data pipe (85 % busy: ~135 of its
// ~200 wavefronts per keypoint are the LDGSTS copies); TMA writes shared memory through the async proxy and takes
// that traffic off the pipe.  A TMA box must start on a 16-byte boundary of x, so a window carries up to 15 bytes of
// slack per row: the unblurred patch is a 31 x 48-byte box (lane = row reads it as three 128-bit words, conflict-free
// at that stride, and picks its nine words by the warp-uniform word offset), the blurred one 37 x 64 bytes.  The
// unblurred windows are double-buffered: lane k looks keypoint k up at the start and issues its loads when a buffer
// frees, so the load of keypoint k + 1 flies while k's centroid is reduced.
#define TD_RAW_W 48
#define TD_RAW_BYTES 1536                 // 31 rows x 48 B = 1488, padded to the 128-byte TMA destination alignment
#define TD_BLUR_W 64
// The blurred window is 37 rows of 64 B, of which a keypoint's patch touches ~9 words per row: without a swizzle the
// 512 byte gathers of a keypoint land on 18 of the 32 banks (bank = 16 (y & 1) + (x >> 2)) and cost 4.8 wavefronts
// per warp load.  With the 64-byte TMA swizzle (address bits 4-5 ^= bits 7-8) the four 16-byte chunks of a row are
// permuted differently in successive row pairs, the gathers spread over every bank, and the reader applies the same
// XOR to its address.
#ifndef VS_DESC_SWIZZLE
#define VS_DESC_SWIZZLE 1
#endif
__device__ __forceinline__ uint32_t td_lds_u8_swz(uint32_t saddr)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr ^ ((saddr >> 3) & 0x30u)));
    return v;
}
#define TD_BLUR_BYTES 2432                // 37 rows x 64 B = 2368, padded likewise
struct alignas(128) VsDescMaps {
    CUtensorMap raw[8], blur[8];          // per level; raw[0] is the caller's frame stack
};

__device__ __forceinline__ void td_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n .reg .pred p;\n"
        "W%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra D%=;\n bra W%=;\n"
        "D%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}

__device__ __forceinline__ void td_load(uint32_t dst, const CUtensorMap* tm, int x, int y, int b, uint32_t bar, uint32_t bytes)
{
    // the window was last read through the generic proxy: order those reads before the async-proxy write
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"((uint64_t)tm), "r"(x), "r"(y), "r"(b), "r"(bar) : "memory");
}

__global__ void __launch_bounds__(256, 5)
vs_describe_tma_kernel(const __grid_constant__ VsPlan plan, const __grid_constant__ VsDescMaps maps,
                       const VsCand* __restrict__ stage, const int32_t* __restrict__ counters,
                       vslam_keypoint* __restrict__ kp_out, uint8_t* __restrict__ desc_out, int32_t* __restrict__ count_out)
{
    __shared__ __align__(128) uint8_t s_blur[8][TD_BLUR_BYTES];
    __shared__ __align__(128) uint8_t s_raw[8][2][TD_RAW_BYTES];
    __shared__ __align__(8) unsigned long long s_bar[8][4];      // per warp: raw buffer 0 / 1, blurred window A, blurred window B
    const int b = blockIdx.y, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int32_t* ctr = counters + (size_t)b * plan.n_levels * VS_CTR_STRIDE;
    __shared__ int s_start[VS_MAX_LEVELS + 1];
    if (threadIdx.x < 32) {
        const int l = threadIdx.x;
        int n = l < plan.n_levels ? ctr[l * VS_CTR_STRIDE + VS_CTR_NFINAL] : 0, incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (l >= o) incl += t;
        }
        if (l <= VS_MAX_LEVELS) s_start[l] = incl - n;          // s_start[n_levels..] = total
    }
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[wid][0]);
    if (lane == 0) {
        for (int k = 0; k < 4; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * k));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int total = s_start[plan.n_levels];
    if (blockIdx.x == 0 && threadIdx.x == 0) count_out[b] = total;
    uint8_t* wb = s_blur[wid];
    const uint32_t wb_u = (uint32_t)__cvta_generic_to_shared(wb);
    const uint32_t wr_u = (uint32_t)__cvta_generic_to_shared(&s_raw[wid][0][0]);

    // disc mask of this lane's patch row (pass A): byte k of word j <-> u = 4j + k - 15
    uint32_t ic_mask[8];
    {
        const int v = lane - VS_HALF_PATCH;
        const int hw = lane < IC_WIN ? c_umax[v < 0 ? -v : v] : -1;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint32_t mk = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int uu = 4 * j + k - VS_HALF_PATCH;
                if ((uu < 0 ? -uu : uu) <= hw) mk |= 255u << (8 * k);
            }
            ic_mask[j] = mk;
        }
    }
    const int i0 = (blockIdx.x * 8 + wid) * VS_DESC_PER_WARP;
    const int nk = min(VS_DESC_PER_WARP, min(total, plan.kp_cap) - i0);   // warp-uniform; <= 0: nothing to do
    if (nk <= 0) return;
    // lane k < nk looks up keypoint k: level, level coordinates, response, pattern centre
    int m_level = 0, m_x = 0, m_y = 0, m_cx = 0, m_cy = 0;
    uint32_t m_resp = 0;
    float m_scale = 1.f;
    if (lane < nk) {
        const int i = i0 + lane;
        while (m_level + 1 < plan.n_levels && i >= s_start[m_level + 1]) ++m_level;
        const VsLevel& L = plan.lv[m_level];
        const VsCand c = stage[(size_t)b * plan.stage_frame_entries + L.stage_off + (i - s_start[m_level])];
        m_x = c.x; m_y = c.y; m_resp = c.v; m_scale = L.scale;
        // centre of the rBRIEF pattern: cvRound(pt * (1/scale))
        const float inv = __fdiv_rn(1.f, L.scale);
        m_cx = __float2int_rn(__fmul_rn(__fmul_rn((float)m_x, L.scale), inv));
        m_cy = __float2int_rn(__fmul_rn(__fmul_rn((float)m_y, L.scale), inv));
        // the first loads: unblurred windows of keypoints 0 and 1, blurred window of keypoint 0
        if (lane < 2) td_load(wr_u + lane * TD_RAW_BYTES, &maps.raw[m_level], (m_x - VS_HALF_PATCH) & ~15, m_y - VS_HALF_PATCH, b,
                              bar0 + 8 * lane, TD_RAW_W * IC_WIN);
        if (lane == 0) td_load(wb_u, &maps.blur[m_level], (m_cx - DESC_R) & ~15, m_cy - DESC_R, b, bar0 + 16, TD_BLUR_W * DESC_WIN);
    }
    // Pass A: intensity-centroid angles, one keypoint after the other (lane k keeps keypoint k's angle)
    float m_ang = 0.f;
    for (int k = 0; k < nk; ++k) {
        const int px0 = __shfl_sync(0xffffffffu, m_x, k) - VS_HALF_PATCH;     // first column of the patch
        td_wait(bar0 + 8 * (k & 1), (uint32_t)(k >> 1) & 1u);
        int m10 = 0, m01 = 0;
        if (lane < IC_WIN) {
            const uint4* p4 = (const uint4*)(&s_raw[wid][k & 1][0] + lane * TD_RAW_W);
            const uint4 a = p4[0], c4 = p4[1], e = p4[2];
            const uint32_t r[12] = {a.x, a.y, a.z, a.w, c4.x, c4.y, c4.z, c4.w, e.x, e.y, e.z, e.w};
            uint32_t w[9];
            switch ((px0 & 15) >> 2) {
            case 0:
#pragma unroll
                for (int j = 0; j < 9; ++j) w[j] = r[j];
                break;
            case 1:
#pragma unroll
                for (int j = 0; j < 9; ++j) w[j] = r[j + 1];
                break;
            case 2:
#pragma unroll
                for (int j = 0; j < 9; ++j) w[j] = r[j + 2];
                break;
            default:
#pragma unroll
                for (int j = 0; j < 9; ++j) w[j] = r[j + 3];
                break;
            }
            const int sh = (px0 & 3) * 8;
            int rs = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t q = __funnelshift_r(w[j], w[j + 1], sh) & ic_mask[j];
                // bytes of word j <-> u = 4j-15 .. 4j-12
                const uint32_t wu = (uint32_t)((4 * j - 15) & 255) | ((uint32_t)((4 * j - 14) & 255) << 8) |
                                    ((uint32_t)((4 * j - 13) & 255) << 16) | ((uint32_t)((4 * j - 12) & 255) << 24);
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(q), "r"(wu));
                rs = (int)__dp4a(q, 0x01010101u, (uint32_t)rs);
            }
            m01 = (lane - VS_HALF_PATCH) * rs;
        }
        __syncwarp();
        if (lane == k + 2 && lane < nk)       // the buffer is free again: keypoint k + 2's window
            td_load(wr_u + (k & 1) * TD_RAW_BYTES, &maps.raw[m_level], (m_x - VS_HALF_PATCH) & ~15, m_y - VS_HALF_PATCH, b,
                    bar0 + 8 * (k & 1), TD_RAW_W * IC_WIN);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, o);
            m01 += __shfl_xor_sync(0xffffffffu, m01, o);
        }
        const float ang = fast_atan2_deg((float)m01, (float)m10);
        if (lane == k) {
            m_ang = ang;
            vslam_keypoint* ko = kp_out + (size_t)b * plan.kp_cap + i0 + k;
            ko->x = __fmul_rn((float)m_x, m_scale); ko->y = __fmul_rn((float)m_y, m_scale); ko->size = __fmul_rn(31.f, m_scale);
            ko->angle = ang; ko->response = __uint_as_float(m_resp); ko->octave = m_level;
        }
    }
    // rotation from the angle in radians (K7b): (float)cos / (float)sin of the float32 angle, evaluated in double
    float m_ca, m_sa;
    {
        const float rad = __fmul_rn(m_ang, (float)(3.14159265358979323846 / 180.f));
        double sd, cd;
        sincos((double)rad, &sd, &cd);
        m_ca = (float)cd; m_sa = (float)sd;
    }
    // Pass B: steered rBRIEF on the blurred window.  Two window buffers per warp: A = s_blur[wid] (keypoint 0's window
    // has been in flight since the start), B = the two unblurred-window buffers, which pass A no longer needs
    // (2 x 1536 B >= 2432 B).  Keypoint k uses buffer k & 1 and, once its tests are done, hands the buffer to keypoint
    // k + 2, so a window's TMA load flies while the previous keypoint's 256 tests run.
    __syncwarp();
    if (lane == 1 && lane < nk)
        td_load(wr_u, &maps.blur[m_level], (m_cx - DESC_R) & ~15, m_cy - DESC_R, b, bar0 + 24, TD_BLUR_W * DESC_WIN);
    for (int k = 0; k < nk; ++k) {
        const float ca = __shfl_sync(0xffffffffu, m_ca, k), sa = __shfl_sync(0xffffffffu, m_sa, k);
        const int bx0 = __shfl_sync(0xffffffffu, m_cx, k) - DESC_R;           // first column of the window
        const uint8_t* win = (k & 1) ? &s_raw[wid][0][0] : wb;
        const uint32_t win_u = (k & 1) ? wr_u : wb_u;
        (void)win; (void)win_u;
        td_wait(bar0 + 16 + 8 * (k & 1), (uint32_t)(k >> 1) & 1u);
        // cvRound(f) as the low mantissa bits of f + 1.5 * 2^23 (round half to even, |f| < 2^22)
        const float magic = 12582912.0f;
        const uint32_t off0 = (uint32_t)(DESC_R * TD_BLUR_W + DESC_R + (bx0 & 15)) - 0x4B400000u * (TD_BLUR_W + 1);
        uint32_t byte = 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const float4 p = __ldg(&g_patf_t[t * 32 + lane]);
            const float fx0 = __fsub_rn(__fmul_rn(p.x, ca), __fmul_rn(p.y, sa)), fy0 = __fadd_rn(__fmul_rn(p.x, sa), __fmul_rn(p.y, ca));
            const float fx1 = __fsub_rn(__fmul_rn(p.z, ca), __fmul_rn(p.w, sa)), fy1 = __fadd_rn(__fmul_rn(p.z, sa), __fmul_rn(p.w, ca));
            const uint32_t i0w = __float_as_uint(__fadd_rn(fy0, magic)) * TD_BLUR_W + __float_as_uint(__fadd_rn(fx0, magic)) + off0;
            const uint32_t i1w = __float_as_uint(__fadd_rn(fy1, magic)) * TD_BLUR_W + __float_as_uint(__fadd_rn(fx1, magic)) + off0;
#if VS_DESC_SWIZZLE
            byte |= (uint32_t)(td_lds_u8_swz(win_u + i0w) < td_lds_u8_swz(win_u + i1w)) << t;
#else
            byte |= (uint32_t)(win[i0w] < win[i1w]) << t;
#endif
        }
        __syncwarp();
        if (lane == k + 2 && lane < nk)
            td_load((k & 1) ? wr_u : wb_u, &maps.blur[m_level], (m_cx - DESC_R) & ~15, m_cy - DESC_R, b, bar0 + 16 + 8 * (k & 1),
                    TD_BLUR_W * DESC_WIN);
        desc_out[((size_t)b * plan.kp_cap + i0 + k) * 32 + lane] = (uint8_t)byte;
    }
}

// The attribute belongs to (function, device): it is only ever raised, per device, so that a second handle with a
// smaller plan cannot pull it below what an older handle's launches need.
void vs_configure_select_smem(int bytes)
{
    static int cur[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (bytes > 48 * 1024 && bytes > __atomic_load_n(&cur[dev], __ATOMIC_ACQUIRE)) {
        if (cudaFuncSetAttribute(vs_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess)
            __atomic_store_n(&cur[dev], bytes, __ATOMIC_RELEASE);
    }
}

int vs_launch_select(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch,
                     vslam_keypoint* d_kp, uint8_t* d_desc, int32_t* d_count, cudaStream_t s)
{
    const VsPlan& p = o->plan;
    const size_t fstride = (size_t)pitch * p.height;
    if (o->dist_mode == 1) {
        vs_prof_mark(o, VS_PROF_SELECT, s);
        int rc = vs_launch_octree(o, batch, s);
        if (rc) return rc;
    } else {
        vs_prof_mark(o, VS_PROF_HARRIS, s);
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_harris_kernel<<<dim3(VS_HARRIS_CTAS, p.n_levels, batch), 256, 0, s>>>(
            p, d_frames, pitch, fstride, o->d_pyr, o->d_cand, o->d_counters);
        int max_stage = 0;
        for (int l = 0; l < p.n_levels; ++l) max_stage = max(max_stage, p.lv[l].stage_cap);
        size_t smem = (size_t)max_stage * sizeof(unsigned long long);
        vs_prof_mark(o, VS_PROF_SELECT, s);
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_select_kernel<<<dim3(p.n_levels, batch), 256, smem, s>>>(p, o->d_cand, o->d_stage,
                                                                   o->d_counters, o->d_status);
    }
    vs_prof_mark(o, VS_PROF_DESCRIBE, s);
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    // both windows by TMA when every level's images can be described by a tensor map (16-byte aligned base and
    // pitches: always for the pyramid, for the caller's frames when they are aligned), else the hand-staged kernel
    VsDescMaps maps;
    bool tma = p.n_levels <= 8 && getenv("VSLAM_DESCRIBE_NO_TMA") == nullptr;
    for (int l = 0; tma && l < p.n_levels; ++l) {
        const VsLevel& L = p.lv[l];
        if (L.w <= 0 || L.h <= 0) { memset(&maps.raw[l], 0, sizeof(CUtensorMap)); memset(&maps.blur[l], 0, sizeof(CUtensorMap)); continue; }
        tma = l == 0 ? vs_image_tensor_map(&maps.raw[0], d_frames, p.width, p.height, (size_t)pitch, fstride, batch, TD_RAW_W, IC_WIN)
                     : vs_image_tensor_map(&maps.raw[l], o->d_pyr + L.raw_off, L.w, L.h, (size_t)L.pitch, p.pyr_frame_bytes, batch,
                                           TD_RAW_W, IC_WIN);
        tma = tma && vs_image_tensor_map(&maps.blur[l], o->d_pyr + L.blur_off, L.w, L.h, (size_t)L.pitch, p.pyr_frame_bytes, batch,
                                         TD_BLUR_W, DESC_WIN, VS_DESC_SWIZZLE != 0);
    }
    const dim3 dgrid((p.kp_cap + 8 * VS_DESC_PER_WARP - 1) / (8 * VS_DESC_PER_WARP), batch);
    if (tma)
        vs_describe_tma_kernel<<<dgrid, 256, 0, s>>>(p, maps, o->d_stage, o->d_counters, d_kp, d_desc, d_count);
    else
        vs_describe_kernel<<<dgrid, 256, 0, s>>>(p, d_frames, pitch, fstride, o->d_pyr, o->d_stage, o->d_counters, d_kp,
                                                  d_desc, d_count);
    vs_prof_mark(o, -1, s);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
