// orb_level.cu -- fused per-level kernel of the ORB extraction path.
//
// One launch per pyramid level over a (tile_x, tile_y, frame) grid.  A CTA
//   1. produces the (TW+8) x (TH+8) tile of level l in shared memory -- copied from
//      the input frame for l = 0, else resized from level l-1 with OpenCV's fixed-point
//      INTER_LINEAR_EXACT (cv2.ORB builds its pyramid that way
//      behind extractor.py:65) -- halo pixels outside the level are REFLECT_101 copies,
//   2. stores the raw level (needed by the next level, Harris and the IC angle),
//   3. scores FAST-9-16 on the tile (+1 ring), applies 3x3 non-max suppression and the
//      31-px border and appends the survivors to the (frame, level) candidate list while
//      accumulating the 256-bin score histogram used by the first retain-best cut (K2/K3),
//   4. applies the 7x7 sigma=2 float Gaussian in cv2's exact operation order (K7a) and
//      stores the blurred level used by rBRIEF.
// Level l therefore costs one read of level l-1 and two writes of level l in HBM.
#include "common.cuh"

#define RAW_W (VS_TW + 2 * VS_HALO)   // 72
#define RAW_H (VS_TH + 2 * VS_HALO)   // 40
#define RAW_S 80                      // smem row stride of the raw tile
#define SC_W (VS_TW + 2)
#define SC_H (VS_TH + 2)
#define SC_S 68
#define BL_ROWS (VS_TH + 6)
#define LIST_CAP ((VS_TW / 2) * (VS_TH / 2))   // NMS leaves at most one corner per 2x2 block

// getGaussianKernel(7, 2, CV_32F) of cv2 4.13, taps outer..centre (bit patterns)
__device__ __constant__ uint32_t c_gauss_bits[4] = {0x3d8fafb1u, 0x3e06387eu, 0x3e434a39u, 0x3e5d4ae0u};

__device__ __forceinline__ bool run9(uint32_t m)
{
    m |= m << 16;
    m &= m >> 1;
    m &= m >> 2;
    m &= m >> 4;
    m &= m >> 1;
    return m != 0;
}

// FAST-9-16 score S-1 of the pixel at p (0 when not a corner), SURVEY.md 8(a) K2.
__device__ __forceinline__ int fast_score(const uint8_t* p, int thr)
{
    // circle offsets (dx, dy) in the order of SURVEY.md K2
    const int off[16] = {3 * RAW_S, 3 * RAW_S + 1, 2 * RAW_S + 2, RAW_S + 3, 3, -RAW_S + 3,
                         -2 * RAW_S + 2, -3 * RAW_S + 1, -3 * RAW_S, -3 * RAW_S - 1,
                         -2 * RAW_S - 2, -RAW_S - 3, -3, RAW_S - 3, 2 * RAW_S - 2, 3 * RAW_S - 1};
    const int c = p[0];
    int v[16];
    uint32_t bm = 0, dm = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        v[k] = p[off[k]];
        bm |= (uint32_t)(v[k] > c + thr) << k;
        dm |= (uint32_t)(v[k] < c - thr) << k;
    }
    if (!run9(bm) && !run9(dm)) return 0;
    // exact score: max over the 16 arcs of max(min d, min -d), d = c - v
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = c - v[k];
    int lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = min(d[k], min(d[(k + 1) & 15], d[(k + 2) & 15]));
        hi3[k] = max(d[k], max(d[(k + 1) & 15], d[(k + 2) & 15]));
    }
    // NOTE: max(best, max(mn, -mx)) per arc is miscompiled by ptxas 12.9 (and the 580 driver
    // JIT) for sm_100a: the negation is dropped when the chain is fused into VIMNMX3
    // (tools/repro/ptxas_vimnmx3_neg.cu reproduces it).  Keep the two reductions apart and negate once,
    // behind an opaque barrier.
    int best_lo = -256, best_hi = 256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        int mn = min(lo3[k], min(lo3[(k + 3) & 15], lo3[(k + 6) & 15]));
        int mx = max(hi3[k], max(hi3[(k + 3) & 15], hi3[(k + 6) & 15]));
        best_lo = max(best_lo, mn);
        best_hi = min(best_hi, mx);
    }
    int neg_hi = -best_hi;
    asm volatile("" : "+r"(neg_hi));
    const int best = max(best_lo, neg_hi);
    return best > thr ? best - 1 : 0;
}

__global__ void __launch_bounds__(256)
vs_level_kernel(const __grid_constant__ VsPlan plan, int level, const uint8_t* __restrict__ frames,
                int fpitch, size_t fstride, uint8_t* __restrict__ pyr,
                const uint32_t* __restrict__ tab, VsCand* __restrict__ cand,
                int32_t* __restrict__ counters)
{
    __shared__ __align__(16) uint8_t s_raw[RAW_H * RAW_S];
    __shared__ __align__(16) uint8_t s_sc[SC_H * SC_S];
    __shared__ float s_row[BL_ROWS * VS_TW];
    __shared__ uint32_t s_hist[256];
    __shared__ VsCand s_list[LIST_CAP];
    __shared__ int s_n, s_base;

    const VsLevel& L = plan.lv[level];
    const int tid = threadIdx.x;
    const int b = blockIdx.z;
    const int x0 = blockIdx.x * VS_TW, y0 = blockIdx.y * VS_TH;
    const int W = L.w, H = L.h;
    uint8_t* fpyr = pyr + (size_t)b * plan.pyr_frame_bytes;

    s_hist[tid] = 0;
    if (tid == 0) s_n = 0;

    // ---- 1. raw tile -------------------------------------------------------
    if (level == 0) {
        const uint8_t* src = frames + (size_t)b * fstride;
        for (int i = tid; i < RAW_H * RAW_W; i += 256) {
            int sy = i / RAW_W, sx = i - sy * RAW_W;
            int gx = vs_reflect101(x0 - VS_HALO + sx, W);
            int gy = vs_reflect101(y0 - VS_HALO + sy, H);
            s_raw[sy * RAW_S + sx] = __ldg(src + (size_t)gy * fpitch + gx);
        }
    } else {
        const VsLevel& P = plan.lv[level - 1];
        const uint8_t* src;
        int sp;
        if (level == 1) { src = frames + (size_t)b * fstride; sp = fpitch; }
        else { src = fpyr + P.raw_off; sp = P.pitch; }
        const uint32_t* tx = tab + L.tabx_off;
        const uint32_t* ty = tab + L.taby_off;
        const int sw = P.w, sh = P.h;
        for (int i = tid; i < RAW_H * RAW_W; i += 256) {
            int sy = i / RAW_W, sx = i - sy * RAW_W;
            int gx = vs_reflect101(x0 - VS_HALO + sx, W);
            int gy = vs_reflect101(y0 - VS_HALO + sy, H);
            uint32_t ex = __ldg(tx + gx), ey = __ldg(ty + gy);
            int i0 = ex & 0xffff, a = ex >> 16, j0 = ey & 0xffff, bq = ey >> 16;
            int i1 = min(i0 + 1, sw - 1), j1 = min(j0 + 1, sh - 1);
            const uint8_t* r0 = src + (size_t)j0 * sp;
            const uint8_t* r1 = src + (size_t)j1 * sp;
            uint32_t h0 = (uint32_t)__ldg(r0 + i0) * (256 - a) + (uint32_t)__ldg(r0 + i1) * a;
            uint32_t h1 = (uint32_t)__ldg(r1 + i0) * (256 - a) + (uint32_t)__ldg(r1 + i1) * a;
            uint32_t v = h0 * (256 - bq) + h1 * bq;
            s_raw[sy * RAW_S + sx] = (uint8_t)((v + 32768u) >> 16);
        }
    }
    __syncthreads();

    // ---- 2. store the raw level -------------------------------------------
    if (level > 0 && tid < 128) {
        int row = tid >> 2, seg = tid & 3;
        int gx = x0 + seg * 16, gy = y0 + row;
        if (gx < L.pitch && gy < H) {
            const uint32_t* s = (const uint32_t*)(s_raw + (row + VS_HALO) * RAW_S + VS_HALO + seg * 16);
            uint4 v = make_uint4(s[0], s[1], s[2], s[3]);
            *(uint4*)(fpyr + L.raw_off + (size_t)gy * L.pitch + gx) = v;
        }
    }

    // ---- 3. FAST scores on the tile + 1 ring --------------------------------
    const int thr = plan.fast_thr;
    for (int i = tid; i < SC_H * SC_W; i += 256) {
        int sy = i / SC_W, sx = i - sy * SC_W;
        int gx = x0 - 1 + sx, gy = y0 - 1 + sy;
        int sc = 0;
        if (gx >= 3 && gx < W - 3 && gy >= 3 && gy < H - 3)
            sc = fast_score(s_raw + (sy + 3) * RAW_S + (sx + 3), thr);
        s_sc[sy * SC_S + sx] = (uint8_t)sc;
    }
    __syncthreads();
    // non-max suppression (strict >, non-corners count as 0) + 31-px border
    for (int i = tid; i < VS_TH * VS_TW; i += 256) {
        int ty = i / VS_TW, tx = i - ty * VS_TW;
        int gx = x0 + tx, gy = y0 + ty;
        const uint8_t* q = s_sc + (ty + 1) * SC_S + (tx + 1);
        int v = q[0];
        bool keep = v > 0 && gx >= VS_EDGE && gx < W - VS_EDGE && gy >= VS_EDGE && gy < H - VS_EDGE;
        if (keep)
            keep = v > q[-SC_S - 1] && v > q[-SC_S] && v > q[-SC_S + 1] && v > q[-1] && v > q[1] &&
                   v > q[SC_S - 1] && v > q[SC_S] && v > q[SC_S + 1];
        if (keep) {
            int slot = atomicAdd(&s_n, 1);
            VsCand c;
            c.x = (uint16_t)gx; c.y = (uint16_t)gy; c.v = (uint32_t)v;
            s_list[slot] = c;
            atomicAdd(&s_hist[v], 1u);
        }
    }
    __syncthreads();
    int32_t* ctr = counters + ((size_t)b * plan.n_levels + level) * VS_CTR_STRIDE;
    const int n = s_n;
    if (n > 0) {
        if (tid == 0) s_base = atomicAdd(&ctr[VS_CTR_NCAND], n);
        uint32_t hv = s_hist[tid];
        if (hv) atomicAdd(&ctr[VS_CTR_HIST + tid], (int)hv);
    }
    __syncthreads();
    if (n > 0) {
        VsCand* out = cand + (size_t)b * plan.cand_frame_entries + L.cand_off + s_base;
        for (int i = tid; i < n; i += 256) out[i] = s_list[i];
    }

    // ---- 4. Gaussian 7x7 (row pass, then column pass) -----------------------
    float k0 = __uint_as_float(c_gauss_bits[0]), k1 = __uint_as_float(c_gauss_bits[1]),
          k2 = __uint_as_float(c_gauss_bits[2]), k3 = __uint_as_float(c_gauss_bits[3]);
    const int wrow = W & ~31, wcol = W & ~7;   // cv2's vector blocks use FMA, its scalar tails do not
    for (int i = tid; i < BL_ROWS * VS_TW; i += 256) {
        int r = i / VS_TW, tx = i - r * VS_TW;
        const uint8_t* p = s_raw + (r + 1) * RAW_S + (tx + VS_HALO);   // row r <-> tile row r-3
        float x0f = (float)p[-3], x1f = (float)p[-2], x2f = (float)p[-1], x3f = (float)p[0],
              x4f = (float)p[1], x5f = (float)p[2], x6f = (float)p[3];
        float s = __fmul_rn(k0, x0f);
        if (x0 + tx < wrow) {
            s = __fmaf_rn(k1, x1f, s); s = __fmaf_rn(k2, x2f, s); s = __fmaf_rn(k3, x3f, s);
            s = __fmaf_rn(k2, x4f, s); s = __fmaf_rn(k1, x5f, s); s = __fmaf_rn(k0, x6f, s);
        } else {
            s = __fadd_rn(s, __fmul_rn(k1, x1f)); s = __fadd_rn(s, __fmul_rn(k2, x2f));
            s = __fadd_rn(s, __fmul_rn(k3, x3f)); s = __fadd_rn(s, __fmul_rn(k2, x4f));
            s = __fadd_rn(s, __fmul_rn(k1, x5f)); s = __fadd_rn(s, __fmul_rn(k0, x6f));
        }
        s_row[i] = s;
    }
    __syncthreads();
    uint8_t* blur = fpyr + L.blur_off;
    for (int i = tid; i < VS_TH * VS_TW / 4; i += 256) {
        int ty = i / (VS_TW / 4), tx = (i - ty * (VS_TW / 4)) * 4;
        uint32_t packed = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float* c = s_row + (ty + 3) * VS_TW + tx + j;
            float s = __fmul_rn(k3, c[0]);
            float p1 = __fadd_rn(c[-VS_TW], c[VS_TW]);
            float p2 = __fadd_rn(c[-2 * VS_TW], c[2 * VS_TW]);
            float p3 = __fadd_rn(c[-3 * VS_TW], c[3 * VS_TW]);
            if (x0 + tx + j < wcol) {
                s = __fmaf_rn(k2, p1, s); s = __fmaf_rn(k1, p2, s); s = __fmaf_rn(k0, p3, s);
            } else {
                s = __fadd_rn(s, __fmul_rn(k2, p1)); s = __fadd_rn(s, __fmul_rn(k1, p2));
                s = __fadd_rn(s, __fmul_rn(k0, p3));
            }
            int v = __float2int_rn(s);
            v = min(max(v, 0), 255);
            packed |= (uint32_t)v << (8 * j);
        }
        int gx = x0 + tx, gy = y0 + ty;
        if (gx < L.pitch && gy < H) *(uint32_t*)(blur + (size_t)gy * L.pitch + gx) = packed;
    }
}

int vs_launch_levels(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch, cudaStream_t s)
{
    const VsPlan& p = o->plan;
    VS_CUDA(cudaMemsetAsync(o->d_counters, 0,
                            (size_t)batch * p.n_levels * VS_CTR_STRIDE * sizeof(int32_t), s));
    for (int l = 0; l < p.n_levels; ++l) {
        const VsLevel& L = p.lv[l];
        if (L.w <= 0 || L.h <= 0) continue;
        dim3 grid(L.tiles_x, L.tiles_y, batch);
        vs_prof_mark(o, l, s);
        ++g_vs_launches;
        vs_level_kernel<<<grid, 256, 0, s>>>(p, l, d_frames, pitch, (size_t)pitch * p.height,
                                             o->d_pyr, o->d_tab, o->d_cand, o->d_counters);
    }
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
