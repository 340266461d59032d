// orb_level.cu -- fused per-level kernel of the ORB extraction path.
//
// One launch per pyramid level over a (tile_x, tile_y, frame) grid.  A CTA
//   1. produces the (TW+8) x (TH+8) tile of level l in shared memory as 16-bit lanes --
//      widened from the input frame for l = 0, else resized from level l-1 with OpenCV's
//      fixed-point INTER_LINEAR_EXACT (cv2.ORB builds its pyramid that
//      way behind extractor.py:65): the parent region is staged in shared memory with
//      aligned 32-bit loads and the resize runs as a horizontal then a vertical pass --
//      halo pixels outside the level are REFLECT_101 copies,
//   2. stores the raw level (needed by the next level, Harris and the IC angle),
//   3. scores FAST-9-16 on the tile (+1 ring) four pixels per thread with packed 16x2
//      min/max networks (VIMNMX3.U16x2), applies 3x3 non-max suppression and the 31-px
//      border and appends the survivors to the (frame, level) candidate list while
//      accumulating the 256-bin score histogram used by the first retain-best cut (K2/K3),
//   4. applies the 7x7 sigma=2 float Gaussian in cv2's exact operation order (K7a) and
//      stores the blurred level used by rBRIEF.
// Level l therefore costs one read of level l-1 and two writes of level l in HBM.
#include "common.cuh"

#define RAW_W (VS_TW + 2 * VS_HALO)   // 72 columns: x0-4 .. x0+67
#define RAW_H (VS_TH + 2 * VS_HALO)   // 38 rows:    y0-4 .. y0+33
#define RAW_S 80                      // row stride of the 16-bit tile (elements)
#define SC_H (VS_TH + 2)              // score rows  y0-1 .. y0+TH
#define SC_S 72                       // score row stride; column index = tx + 4
#define BL_ROWS (VS_TH + 6)
#define LIST_CAP ((VS_TW / 2) * (VS_TH / 2))   // NMS leaves at most one corner per 2x2 block

// getGaussianKernel(7, 2, CV_32F) of cv2 4.13, taps outer..centre (bit patterns)
__device__ __constant__ uint32_t c_gauss_bits[4] = {0x3d8fafb1u, 0x3e06387eu, 0x3e434a39u, 0x3e5d4ae0u};

// ---------------------------------------------------------------------------------------
// FAST-9-16, SURVEY.md 8(a) K2.  With v[k] the 16 circle pixels and c the centre,
//   S = max( c - min_arcs max_arc v ,  max_arcs min_arc v - c )   over the 16 arcs of 9,
// the pixel is a corner iff S > thr and cv2's score is S - 1.  Both routines return
// max(S - thr, 0) (1 byte; order-preserving, 0 = not a corner) so that the non-max
// suppression can compare bytes; the caller adds thr - 1 back.
// ---------------------------------------------------------------------------------------

// two pixels at once: 16x2 lanes hold raw pixel values (0..255), no negation anywhere
// (see tools/repro/ptxas_vimnmx3_neg.cu for why the signed form is avoided).
__device__ __forceinline__ uint32_t fast_pair(const uint32_t (&v)[16], uint32_t c, uint32_t kbias)
{
    uint32_t lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = __vimin3_u16x2(v[k], v[(k + 1) & 15], v[(k + 2) & 15]);
        hi3[k] = __vimax3_u16x2(v[k], v[(k + 1) & 15], v[(k + 2) & 15]);
    }
    uint32_t lo9[16], hi9[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo9[k] = __vimin3_u16x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
        hi9[k] = __vimax3_u16x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    }
    uint32_t a0 = __vimax3_u16x2(lo9[0], lo9[1], lo9[2]), a1 = __vimax3_u16x2(lo9[3], lo9[4], lo9[5]),
             a2 = __vimax3_u16x2(lo9[6], lo9[7], lo9[8]), a3 = __vimax3_u16x2(lo9[9], lo9[10], lo9[11]),
             a4 = __vimax3_u16x2(lo9[12], lo9[13], lo9[14]);
    uint32_t Lo = __vimax3_u16x2(__vimax3_u16x2(a0, a1, a2), __vimax3_u16x2(a3, a4, lo9[15]), 0u);
    uint32_t b0 = __vimin3_u16x2(hi9[0], hi9[1], hi9[2]), b1 = __vimin3_u16x2(hi9[3], hi9[4], hi9[5]),
             b2 = __vimin3_u16x2(hi9[6], hi9[7], hi9[8]), b3 = __vimin3_u16x2(hi9[9], hi9[10], hi9[11]),
             b4 = __vimin3_u16x2(hi9[12], hi9[13], hi9[14]);
    uint32_t Hi = __vminu2(__vimin3_u16x2(b0, b1, b2), __vimin3_u16x2(b3, b4, hi9[15]));
    // lanes biased by 512 - thr so that neither subtraction can borrow across the lanes
    uint32_t t1 = (c + kbias) - Hi;          // (c - Hi - thr) + 512
    uint32_t t2 = (Lo + kbias) - c;          // (Lo - c - thr) + 512
    uint32_t m = __vmaxu2(__vmaxu2(t1, t2), 0x02000200u);
    return m - 0x02000200u;                  // max(S - thr, 0) per lane
}

// one pixel (ring columns of the tile); p points into the 16-bit tile
__device__ __forceinline__ int fast_single(const uint16_t* p, int thr)
{
    const int off[16] = {3 * RAW_S, 3 * RAW_S + 1, 2 * RAW_S + 2, RAW_S + 3, 3, -RAW_S + 3,
                         -2 * RAW_S + 2, -3 * RAW_S + 1, -3 * RAW_S, -3 * RAW_S - 1,
                         -2 * RAW_S - 2, -RAW_S - 3, -3, RAW_S - 3, 2 * RAW_S - 2, 3 * RAW_S - 1};
    const int c = p[0];
    int v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = p[off[k]];
    int lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = min(v[k], min(v[(k + 1) & 15], v[(k + 2) & 15]));
        hi3[k] = max(v[k], max(v[(k + 1) & 15], v[(k + 2) & 15]));
    }
    int Lo = 0, Hi = 255;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        Lo = max(Lo, min(lo3[k], min(lo3[(k + 3) & 15], lo3[(k + 6) & 15])));
        Hi = min(Hi, max(hi3[k], max(hi3[(k + 3) & 15], hi3[(k + 6) & 15])));
    }
    const int s = max(c - Hi, Lo - c) - thr;
    return s > 0 ? s : 0;
}

__device__ __forceinline__ uint32_t fsr16(uint32_t lo, uint32_t hi) { return __funnelshift_r(lo, hi, 16); }

// exact u16 -> float without the conversion pipe: 2^23 + v has v in its low mantissa bits
__device__ __forceinline__ float lane_lo_f(uint32_t w)
{
    return __fadd_rn(__uint_as_float((w & 0xffffu) | 0x4B000000u), -8388608.0f);
}
__device__ __forceinline__ float lane_hi_f(uint32_t w)
{
    return __fadd_rn(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7632)), -8388608.0f);
}

// four source bytes at column gx of a row (reflecting / byte-wise when the word is not
// wholly inside the row or the row is not 4-byte aligned)
__device__ __forceinline__ uint32_t load4_reflect(const uint8_t* row, int gx, int W, bool aligned)
{
    if (aligned && gx >= 0 && gx + 3 < W) return __ldg((const uint32_t*)(row + gx));
    uint32_t u = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) u |= (uint32_t)__ldg(row + vs_reflect101(gx + j, W)) << (8 * j);
    return u;
}

__global__ void __launch_bounds__(256)
vs_level_kernel(const __grid_constant__ VsPlan plan, int level, const uint8_t* __restrict__ frames,
                int fpitch, size_t fstride, uint8_t* __restrict__ pyr,
                const uint32_t* __restrict__ tab, VsCand* __restrict__ cand,
                int32_t* __restrict__ counters)
{
    __shared__ __align__(16) uint16_t s_raw[RAW_H * RAW_S];
    __shared__ __align__(16) uint8_t s_sc[SC_H * SC_S];
    __shared__ __align__(16) float s_row[BL_ROWS * VS_TW];
    __shared__ uint32_t s_hist[256];
    __shared__ VsCand s_list[LIST_CAP];
    __shared__ uint32_t s_tx[RAW_W], s_ty[RAW_H];
    __shared__ int s_bnd[4];
    __shared__ int s_n, s_base;
    extern __shared__ __align__(16) uint8_t s_dyn[];   // parent region + horizontal pass (levels >= 1)

    const VsLevel& L = plan.lv[level];
    const int tid = threadIdx.x;
    const int b = blockIdx.z;
    const int x0 = blockIdx.x * VS_TW, y0 = blockIdx.y * VS_TH;
    const int W = L.w, H = L.h;
    uint8_t* fpyr = pyr + (size_t)b * plan.pyr_frame_bytes;

    s_hist[tid] = 0;
    if (tid == 0) s_n = 0;

    // ---- 1. raw tile (16-bit lanes) ------------------------------------------
    if (level == 0) {
        const uint8_t* src = frames + (size_t)b * fstride;
        const bool aligned = (((size_t)src | (size_t)fpitch) & 3) == 0;
        for (int i = tid; i < RAW_H * (RAW_W / 4); i += 256) {
            int r = i / (RAW_W / 4), wd = i - r * (RAW_W / 4);
            int gy = vs_reflect101(y0 - VS_HALO + r, H);
            uint32_t u = load4_reflect(src + (size_t)gy * fpitch, x0 - VS_HALO + 4 * wd, W, aligned);
            uint2 o = make_uint2(__byte_perm(u, 0, 0x4140), __byte_perm(u, 0, 0x4342));
            *(uint2*)(s_raw + r * RAW_S + 4 * wd) = o;
        }
    } else {
        const VsLevel& P = plan.lv[level - 1];
        const uint8_t* src;
        int sp;
        if (level == 1) { src = frames + (size_t)b * fstride; sp = fpitch; }
        else { src = fpyr + P.raw_off; sp = P.pitch; }
        const bool aligned = (((size_t)src | (size_t)sp) & 3) == 0;
        const int sw = P.w, sh = P.h;
        const int PW = plan.par_w;
        uint8_t* s_par = s_dyn;                                         // [par_h][par_w] parent pixels
        uint16_t* s_h = (uint16_t*)(s_dyn + (size_t)plan.par_h * PW);   // [par_h][RAW_W] horizontal pass
        // table entries of this tile's (reflected) columns / rows and the parent bounding box
        const int wid = tid >> 5, lane = tid & 31;
        if (wid < 2) {
            const uint32_t* t = tab + (wid == 0 ? L.tabx_off : L.taby_off);
            const int n = wid == 0 ? RAW_W : RAW_H, o0 = (wid == 0 ? x0 : y0) - VS_HALO, lim = wid == 0 ? W : H;
            uint32_t* dst = wid == 0 ? s_tx : s_ty;
            int lo = 0x7fffffff, hi = 0;
            for (int c = lane; c < n; c += 32) {
                uint32_t e = __ldg(t + vs_reflect101(o0 + c, lim));
                dst[c] = e;
                lo = min(lo, (int)(e & 0xffff));
                hi = max(hi, (int)(e & 0xffff));
            }
            lo = __reduce_min_sync(0xffffffffu, lo);
            hi = __reduce_max_sync(0xffffffffu, hi);
            if (lane == 0) { s_bnd[2 * wid] = lo; s_bnd[2 * wid + 1] = min(hi + 1, (wid == 0 ? sw : sh) - 1); }
        }
        __syncthreads();
        const int pxa = s_bnd[0] & ~3, py0 = s_bnd[2];
        const int nwords = (s_bnd[1] - pxa) / 4 + 1, nrows = s_bnd[3] - py0 + 1;
        for (int i = tid; i < nrows * nwords; i += 256) {
            int r = i / nwords, wd = i - r * nwords;
            const uint8_t* row = src + (size_t)(py0 + r) * sp;
            int gx = pxa + 4 * wd;
            uint32_t u;
            if (aligned && gx + 3 < sw) u = __ldg((const uint32_t*)(row + gx));
            else {
                u = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) u |= (uint32_t)__ldg(row + min(gx + j, sw - 1)) << (8 * j);
            }
            *(uint32_t*)(s_par + r * PW + 4 * wd) = u;
        }
        __syncthreads();
        for (int i = tid; i < nrows * RAW_W; i += 256) {
            int r = i / RAW_W, c = i - r * RAW_W;
            uint32_t e = s_tx[c];
            int i0 = (int)(e & 0xffff), a = (int)(e >> 16);
            int i1 = min(i0 + 1, sw - 1);
            const uint8_t* p = s_par + r * PW - pxa;
            s_h[i] = (uint16_t)((uint32_t)p[i0] * (256 - a) + (uint32_t)p[i1] * a);
        }
        __syncthreads();
        for (int i = tid; i < RAW_H * RAW_W; i += 256) {
            int r = i / RAW_W, c = i - r * RAW_W;
            uint32_t e = s_ty[r];
            int j0 = (int)(e & 0xffff), bq = (int)(e >> 16);
            int j1 = min(j0 + 1, sh - 1);
            uint32_t v = (uint32_t)s_h[(j0 - py0) * RAW_W + c] * (256 - bq) + (uint32_t)s_h[(j1 - py0) * RAW_W + c] * bq;
            s_raw[r * RAW_S + c] = (uint16_t)((v + 32768u) >> 16);
        }
    }
    __syncthreads();

    // ---- 2. store the raw level -------------------------------------------
    if (level > 0 && tid < VS_TH * 4) {
        int row = tid >> 2, seg = tid & 3;
        int gx = x0 + seg * 16, gy = y0 + row;
        if (gx < L.pitch && gy < H) {
            const uint2* s = (const uint2*)(s_raw + (row + VS_HALO) * RAW_S + VS_HALO + seg * 16);
            uint2 q0 = s[0], q1 = s[1], q2 = s[2], q3 = s[3];
            uint4 v = make_uint4(__byte_perm(q0.x, q0.y, 0x6420), __byte_perm(q1.x, q1.y, 0x6420),
                                 __byte_perm(q2.x, q2.y, 0x6420), __byte_perm(q3.x, q3.y, 0x6420));
            *(uint4*)(fpyr + L.raw_off + (size_t)gy * L.pitch + gx) = v;
        }
    }

    // ---- 3. FAST scores on the tile + 1 ring --------------------------------
    // Only pixels that can be a kept corner or the 3x3 neighbour of one are scored
    // (columns 30 .. W-31, rows 30 .. H-31); everything else stays 0.
    const int thr = plan.fast_thr;
    const uint32_t kbias = (uint32_t)(512 - thr) * 0x00010001u;
    const int cx_lo = VS_EDGE - 1, cx_hi = W - VS_EDGE, cy_lo = VS_EDGE - 1, cy_hi = H - VS_EDGE;
    for (int i = tid; i < SC_H * (VS_TW / 4); i += 256) {
        const int w = i & 15, r = i >> 4;
        const int gy = y0 - 1 + r, gx = x0 + 4 * w;
        uint32_t packed = 0;
        if (gy >= cy_lo && gy <= cy_hi && gx + 3 >= cx_lo && gx <= cx_hi) {
            const uint16_t* base = s_raw + (r + 3) * RAW_S + 4 * w;   // word 2 of this row = pixels 0,1
            uint32_t vA[16], vB[16], cA, cB;
            {
                const uint2* q = (const uint2*)(base + 3 * RAW_S);
                uint32_t w1 = q[0].y, w2 = q[1].x, w3 = q[1].y, w4 = q[2].x;
                vA[0] = w2; vB[0] = w3;
                vA[1] = fsr16(w2, w3); vB[1] = fsr16(w3, w4);
                vA[15] = fsr16(w1, w2); vB[15] = vA[1];
            }
            {
                const uint2* q = (const uint2*)(base - 3 * RAW_S);
                uint32_t w1 = q[0].y, w2 = q[1].x, w3 = q[1].y, w4 = q[2].x;
                vA[8] = w2; vB[8] = w3;
                vA[7] = fsr16(w2, w3); vB[7] = fsr16(w3, w4);
                vA[9] = fsr16(w1, w2); vB[9] = vA[7];
            }
            {
                const uint2* q = (const uint2*)(base + 2 * RAW_S);
                vA[14] = q[0].y; vB[14] = q[1].x; vA[2] = q[1].y; vB[2] = q[2].x;
            }
            {
                const uint2* q = (const uint2*)(base - 2 * RAW_S);
                vA[10] = q[0].y; vB[10] = q[1].x; vA[6] = q[1].y; vB[6] = q[2].x;
            }
            {
                const uint2* q = (const uint2*)(base + RAW_S);
                uint2 a = q[0], m = q[1], e = q[2];
                vA[3] = fsr16(m.y, e.x); vB[3] = fsr16(e.x, e.y);
                vA[13] = fsr16(a.x, a.y); vB[13] = fsr16(a.y, m.x);
            }
            {
                const uint2* q = (const uint2*)(base - RAW_S);
                uint2 a = q[0], m = q[1], e = q[2];
                vA[5] = fsr16(m.y, e.x); vB[5] = fsr16(e.x, e.y);
                vA[11] = fsr16(a.x, a.y); vB[11] = fsr16(a.y, m.x);
            }
            {
                const uint2* q = (const uint2*)base;
                uint2 a = q[0], m = q[1], e = q[2];
                vA[4] = fsr16(m.y, e.x); vB[4] = fsr16(e.x, e.y);
                vA[12] = fsr16(a.x, a.y); vB[12] = fsr16(a.y, m.x);
                cA = m.x; cB = m.y;
            }
            const uint32_t oA = fast_pair(vA, cA, kbias), oB = fast_pair(vB, cB, kbias);
            packed = __byte_perm(oA, oB, 0x6420);
            if (gx < cx_lo || gx + 3 > cx_hi) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (gx + j < cx_lo || gx + j > cx_hi) packed &= ~(0xffu << (8 * j));
            }
        }
        *(uint32_t*)(s_sc + r * SC_S + 4 + 4 * w) = packed;
    }
    if (tid < 2 * SC_H) {   // ring columns tx = -1 and tx = TW
        const int r = tid % SC_H, side = tid / SC_H;
        const int tx = side ? VS_TW : -1;
        const int gx = x0 + tx, gy = y0 - 1 + r;
        int sc = 0;
        if (gy >= cy_lo && gy <= cy_hi && gx >= cx_lo && gx <= cx_hi)
            sc = fast_single(s_raw + (r + 3) * RAW_S + tx + VS_HALO, thr);
        s_sc[r * SC_S + 4 + tx] = (uint8_t)sc;
    }
    __syncthreads();

    // non-max suppression (strict >, non-corners count as 0) + 31-px border
    for (int i = tid; i < VS_TH * (VS_TW / 4); i += 256) {
        const int w = i & 15, ty = i >> 4;
        const uint8_t* q = s_sc + (ty + 1) * SC_S + 4 + 4 * w;
        const uint32_t C = *(const uint32_t*)q;
        if (C == 0) continue;
        const int gy = y0 + ty;
        if (gy < VS_EDGE || gy >= H - VS_EDGE) continue;
        const unsigned long long up = (unsigned long long)q[-SC_S - 1] | ((unsigned long long)*(const uint32_t*)(q - SC_S) << 8) |
                                      ((unsigned long long)q[-SC_S + 4] << 40);
        const unsigned long long mid = (unsigned long long)q[-1] | ((unsigned long long)C << 8) | ((unsigned long long)q[4] << 40);
        const unsigned long long dn = (unsigned long long)q[SC_S - 1] | ((unsigned long long)*(const uint32_t*)(q + SC_S) << 8) |
                                      ((unsigned long long)q[SC_S + 4] << 40);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int v = (int)((C >> (8 * j)) & 0xff);
            const int gx = x0 + 4 * w + j;
            if (v == 0 || gx < VS_EDGE || gx >= W - VS_EDGE) continue;
            const uint32_t u = (uint32_t)(up >> (8 * j)), m = (uint32_t)(mid >> (8 * j)), d = (uint32_t)(dn >> (8 * j));
            const int nb = max(max(max((int)(u & 0xff), (int)((u >> 8) & 0xff)), max((int)((u >> 16) & 0xff), (int)(m & 0xff))),
                               max(max((int)((m >> 16) & 0xff), (int)(d & 0xff)), max((int)((d >> 8) & 0xff), (int)((d >> 16) & 0xff))));
            if (v > nb) {
                const int score = v + thr - 1;
                int slot = atomicAdd(&s_n, 1);
                VsCand c;
                c.x = (uint16_t)gx; c.y = (uint16_t)gy; c.v = (uint32_t)score;
                s_list[slot] = c;
                atomicAdd(&s_hist[score], 1u);
            }
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
    // cv2's vector blocks use FMA, its scalar tails do not: rows switch at W - W%32,
    // columns at W - W%8; a 4-pixel group never straddles either boundary.
    const float k0 = __uint_as_float(c_gauss_bits[0]), k1 = __uint_as_float(c_gauss_bits[1]),
                k2 = __uint_as_float(c_gauss_bits[2]), k3 = __uint_as_float(c_gauss_bits[3]);
    const int wrow = W & ~31, wcol = W & ~7;
    for (int i = tid; i < BL_ROWS * (VS_TW / 4); i += 256) {
        const int w = i & 15, rr = i >> 4;                            // s_row row rr <-> gy = y0 - 3 + rr
        const uint2* q = (const uint2*)(s_raw + (rr + 1) * RAW_S + 4 * w);
        const uint2 a = q[0], m = q[1], e = q[2];
        float x[10];                                                  // columns 4w+1 .. 4w+10 of the tile
        x[0] = lane_hi_f(a.x); x[1] = lane_lo_f(a.y); x[2] = lane_hi_f(a.y);
        x[3] = lane_lo_f(m.x); x[4] = lane_hi_f(m.x); x[5] = lane_lo_f(m.y); x[6] = lane_hi_f(m.y);
        x[7] = lane_lo_f(e.x); x[8] = lane_hi_f(e.x); x[9] = lane_lo_f(e.y);
        float o[4];
        if (x0 + 4 * w < wrow) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float s = __fmul_rn(k0, x[j]);
                s = __fmaf_rn(k1, x[j + 1], s); s = __fmaf_rn(k2, x[j + 2], s); s = __fmaf_rn(k3, x[j + 3], s);
                s = __fmaf_rn(k2, x[j + 4], s); s = __fmaf_rn(k1, x[j + 5], s); s = __fmaf_rn(k0, x[j + 6], s);
                o[j] = s;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float s = __fmul_rn(k0, x[j]);
                s = __fadd_rn(s, __fmul_rn(k1, x[j + 1])); s = __fadd_rn(s, __fmul_rn(k2, x[j + 2]));
                s = __fadd_rn(s, __fmul_rn(k3, x[j + 3])); s = __fadd_rn(s, __fmul_rn(k2, x[j + 4]));
                s = __fadd_rn(s, __fmul_rn(k1, x[j + 5])); s = __fadd_rn(s, __fmul_rn(k0, x[j + 6]));
                o[j] = s;
            }
        }
        *(float4*)(s_row + rr * VS_TW + 4 * w) = make_float4(o[0], o[1], o[2], o[3]);
    }
    __syncthreads();
    uint8_t* blur = fpyr + L.blur_off;
    for (int i = tid; i < (VS_TH / 2) * (VS_TW / 4); i += 256) {
        const int w = i & 15, ty = (i >> 4) * 2;                      // two output rows per item
        const int gx = x0 + 4 * w;
        if (gx >= L.pitch || y0 + ty >= H) continue;
        float4 r[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = *(const float4*)(s_row + (ty + k) * VS_TW + 4 * w);
        const bool fma = gx < wcol;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            uint32_t u[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#define RJ(k) (j == 0 ? r[k].x : j == 1 ? r[k].y : j == 2 ? r[k].z : r[k].w)
                float s = __fmul_rn(k3, RJ(h + 3));
                const float p1 = __fadd_rn(RJ(h + 2), RJ(h + 4));
                const float p2 = __fadd_rn(RJ(h + 1), RJ(h + 5));
                const float p3 = __fadd_rn(RJ(h + 0), RJ(h + 6));
#undef RJ
                if (fma) {
                    s = __fmaf_rn(k2, p1, s); s = __fmaf_rn(k1, p2, s); s = __fmaf_rn(k0, p3, s);
                } else {
                    s = __fadd_rn(s, __fmul_rn(k2, p1)); s = __fadd_rn(s, __fmul_rn(k1, p2));
                    s = __fadd_rn(s, __fmul_rn(k0, p3));
                }
                // round half to even in the mantissa of 1.5 * 2^23 + s  (0 <= s < 255.5)
                u[j] = __float_as_uint(__fadd_rn(s, 12582912.0f));
            }
            const uint32_t packed = __byte_perm(__byte_perm(u[0], u[1], 0x0040), __byte_perm(u[2], u[3], 0x0040), 0x5410);
            if (y0 + ty + h < H) *(uint32_t*)(blur + (size_t)(y0 + ty + h) * L.pitch + gx) = packed;
        }
    }
}

size_t vs_level_dyn_smem(const VsPlan& p)
{
    return (size_t)p.par_h * p.par_w + (size_t)p.par_h * RAW_W * sizeof(uint16_t);
}

int vs_configure_level_smem(const VsPlan& p)
{
    const size_t dyn = vs_level_dyn_smem(p);
    if (dyn > 160 * 1024) {
        vs_set_error("scale_factor too large for the level kernel's parent staging (%zu bytes)", dyn);
        return VSLAM_E_INVALID;
    }
    VS_CUDA(cudaFuncSetAttribute(vs_level_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    return VSLAM_OK;
}

int vs_launch_levels(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch, cudaStream_t s)
{
    const VsPlan& p = o->plan;
    VS_CUDA(cudaMemsetAsync(o->d_counters, 0,
                            (size_t)batch * p.n_levels * VS_CTR_STRIDE * sizeof(int32_t), s));
    const size_t dyn = vs_level_dyn_smem(p);
    for (int l = 0; l < p.n_levels; ++l) {
        const VsLevel& L = p.lv[l];
        if (L.w <= 0 || L.h <= 0) continue;
        dim3 grid(L.tiles_x, L.tiles_y, batch);
        vs_prof_mark(o, l, s);
        ++g_vs_launches;
        vs_level_kernel<<<grid, 256, l > 0 ? dyn : 0, s>>>(p, l, d_frames, pitch, (size_t)pitch * p.height,
                                                         o->d_pyr, o->d_tab, o->d_cand, o->d_counters);
    }
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
