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
// Phase order between barriers: tile | compass pre-test + Gaussian row pass | queued FAST networks |
// NMS into the tile list | list-slot atomic + Gaussian column pass | list write-out -- the row pass shares
// the pre-test's interval (both only read the tile) and the round trip of the one global atomic per tile
// hides behind the column pass (measured: work stealing from a shared chunk counter was slower than
// static striding, and 6 CTAs/SM at 40 registers bought 2 % on level 0 only).
#include <cuda.h>

#include "common.cuh"

#define RAW_W (VS_TW + 2 * VS_HALO)   // 72 columns: x0-4 .. x0+67
#define RAW_H (VS_TH + 2 * VS_HALO)   // 38 rows:    y0-4 .. y0+33
#define RAW_S 84                      // row stride of the 16-bit tile (elements): 42 words, so a column of words spreads over 16 banks
#define SC_H (VS_TH + 2)              // score rows  y0-1 .. y0+TH
#define SC_S 72                       // score row stride; column index = tx + 4
#define BL_ROWS (VS_TH + 6)
#define LIST_CAP ((VS_TW / 2) * (VS_TH / 2))   // NMS leaves at most one corner per 2x2 block
#define L0_BOX_W 96                   // level-0 TMA box: columns x0-16 .. x0+79 (16-byte aligned start, tile = x0-4 .. x0+67)
#ifndef VS_LEVEL_CTAS
#define VS_LEVEL_CTAS 6                // resident CTAs per SM the register budget is capped for (40 registers)
#endif

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

// plain shared-memory atomic add (inline PTX keeps the compiler from wrapping it in its own warp aggregation)
__device__ __forceinline__ int smem_atomic_add(int* p, int v)
{
    int old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    return old;
}

// Compass pre-test for a pixel pair: a, b, c, d = the circle points (0,+3), (0,-3), (+3,0), (-3,0), cen = centre.
// An arc of 9 of the 16 circle positions contains two or three of the compass positions 0, 4, 8, 12, and they are
// NEIGHBOURS on the compass ring: a corner needs an adjacent compass pair (a,c), (c,b), (b,d) or (d,a) with both
// values above cen + thr, or both below cen - thr.  Returns bit 15 / 31 set for the lanes that pass.  (The weaker
// "second-largest above / second-smallest below" test of round 1 queued 28 % of the pairs of the bench clip; the
// networks and the NMS walk run on the queue only.)
__device__ __forceinline__ uint32_t compass_need(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t cen, uint32_t kbias15)
{
    // the ring is a - c - b - d - a: every adjacent pair takes one point of {a, b} and one of {c, d}, so "some
    // adjacent pair is above T" <=> max(a, b) > T and max(c, d) > T, and likewise below
    const uint32_t hi = __vminu2(__vmaxu2(a, b), __vmaxu2(c, d));     // brightest adjacent pair (its darker point)
    const uint32_t lo = __vmaxu2(__vminu2(a, b), __vminu2(c, d));     // darkest adjacent pair (its brighter point)
    // lanes biased by kbias15 = 0x7fff - thr: bit 15 of a lane <=> difference > thr.  Each lane sum stays inside
    // [0x7f00 - thr, 0x80fe - thr], so the 32-bit adds are lane-wise adds
    const uint32_t m = __vmaxu2(hi + kbias15 - cen, cen + kbias15 - lo);
    return m & 0x80008000u;
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

__device__ __forceinline__ void sts_u16(uint32_t saddr, uint32_t v)
{
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(saddr), "h"((uint16_t)v) : "memory");
}

__device__ __forceinline__ uint32_t fsr16(uint32_t lo, uint32_t hi) { return __funnelshift_r(lo, hi, 16); }

// u16 lane -> float: I2F.U16 takes either half of a register directly (XU pipe, which is otherwise
// idle here; the ALU pipe is the one this kernel saturates)
__device__ __forceinline__ float lane_lo_f(uint32_t w) { return (float)(uint16_t)w; }
__device__ __forceinline__ float lane_hi_f(uint32_t w) { return (float)(uint16_t)(w >> 16); }

// four source bytes at column gx of a row, byte-wise with REFLECT_101 (words that are not
// wholly inside the row, or rows that are not 4-byte aligned)
__device__ __noinline__ uint32_t load4_reflect(const uint8_t* row, int gx, int W)
{
    uint32_t u = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) u |= (uint32_t)__ldg(row + vs_reflect101(gx + j, W)) << (8 * j);
    return u;
}

// same for rows of at least 8 pixels, where one reflection per side covers the 4-pixel halo (branch-free)
__device__ __forceinline__ uint32_t load4_edge(const uint8_t* row, int gx, int W)
{
    uint32_t u = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int x = (W - 1) - abs((W - 1) - abs(gx + j));
        u |= (uint32_t)__ldg(row + x) << (8 * j);
    }
    return u;
}

// Row pass of the 7-tap Gaussian for 8 adjacent pixels: q = 8 words of the 16-bit tile starting 4
// columns left of the first pixel, dst = 8 floats.  cv2 order: k0*x0, then one tap at a time.
template <bool FMA>
__device__ __forceinline__ void blur_row_item(const uint2* q, float* dst0, float* dst1)
{
    const float k0 = __uint_as_float(c_gauss_bits[0]), k1 = __uint_as_float(c_gauss_bits[1]),
                k2 = __uint_as_float(c_gauss_bits[2]), k3 = __uint_as_float(c_gauss_bits[3]);
    const uint2 a = q[0], b = q[1], c = q[2], d = q[3];
    float x[14];                                                      // tile columns +1 .. +14
    x[0] = lane_hi_f(a.x); x[1] = lane_lo_f(a.y); x[2] = lane_hi_f(a.y);
    x[3] = lane_lo_f(b.x); x[4] = lane_hi_f(b.x); x[5] = lane_lo_f(b.y); x[6] = lane_hi_f(b.y);
    x[7] = lane_lo_f(c.x); x[8] = lane_hi_f(c.x); x[9] = lane_lo_f(c.y); x[10] = lane_hi_f(c.y);
    x[11] = lane_lo_f(d.x); x[12] = lane_hi_f(d.x); x[13] = lane_lo_f(d.y);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float s = __fmul_rn(k0, x[j]);
        if (FMA) {
            s = __fmaf_rn(k1, x[j + 1], s); s = __fmaf_rn(k2, x[j + 2], s); s = __fmaf_rn(k3, x[j + 3], s);
            s = __fmaf_rn(k2, x[j + 4], s); s = __fmaf_rn(k1, x[j + 5], s); s = __fmaf_rn(k0, x[j + 6], s);
        } else {
            s = __fadd_rn(s, __fmul_rn(k1, x[j + 1])); s = __fadd_rn(s, __fmul_rn(k2, x[j + 2]));
            s = __fadd_rn(s, __fmul_rn(k3, x[j + 3])); s = __fadd_rn(s, __fmul_rn(k2, x[j + 4]));
            s = __fadd_rn(s, __fmul_rn(k1, x[j + 5])); s = __fadd_rn(s, __fmul_rn(k0, x[j + 6]));
        }
        o[j] = s;
    }
    *(float4*)dst0 = make_float4(o[0], o[1], o[2], o[3]);
    *(float4*)dst1 = make_float4(o[4], o[5], o[6], o[7]);
}

// Column pass for 4 columns x 2 output rows: p = row-pass results of the first output row's
// top tap (8 rows of VS_TW floats are read).  cv2 order: k3*c, then symmetric pairs inner..outer;
// the result is rounded half-to-even to a byte (0 <= s < 255.5) in the mantissa of 1.5*2^23 + s.
template <bool FMA>
__device__ __forceinline__ void blur_col_item(const float* p, uint32_t& out0, uint32_t& out1)
{
    const float k0 = __uint_as_float(c_gauss_bits[0]), k1 = __uint_as_float(c_gauss_bits[1]),
                k2 = __uint_as_float(c_gauss_bits[2]), k3 = __uint_as_float(c_gauss_bits[3]);
    float r[8][4];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float4 v = *(const float4*)(p + k * VS_TW);
        r[k][0] = v.x; r[k][1] = v.y; r[k][2] = v.z; r[k][3] = v.w;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t u[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float s = __fmul_rn(k3, r[h + 3][j]);
            const float p1 = __fadd_rn(r[h + 2][j], r[h + 4][j]);
            const float p2 = __fadd_rn(r[h + 1][j], r[h + 5][j]);
            const float p3 = __fadd_rn(r[h + 0][j], r[h + 6][j]);
            if (FMA) {
                s = __fmaf_rn(k2, p1, s); s = __fmaf_rn(k1, p2, s); s = __fmaf_rn(k0, p3, s);
            } else {
                s = __fadd_rn(s, __fmul_rn(k2, p1)); s = __fadd_rn(s, __fmul_rn(k1, p2));
                s = __fadd_rn(s, __fmul_rn(k0, p3));
            }
            u[j] = __float_as_uint(__fadd_rn(s, 12582912.0f));
        }
        const uint32_t packed = __byte_perm(__byte_perm(u[0], u[1], 0x0040), __byte_perm(u[2], u[3], 0x0040), 0x5410);
        if (h == 0) out0 = packed; else out1 = packed;
    }
}

// L0 = true: level 0 (tile widened from the frame); false: levels >= 1 (tile resized from level - 1).
// PWC: row pitch of the staged parent region when it is known at compile time (112 bytes at the reference's scale
// factor 1.2, whatever the frame size), 0 = read it from the plan.  With PWC fixed the two resize passes are
// straight-line code: every address is base + immediate and the per-thread trip counts are constants (the generic
// loops spent as many instructions on their bookkeeping as on the blend: 0.5 of 1.9 warp-instructions per pixel).
// Separate instantiations so that each gets its own register allocation and schedule.
#define VS_PAR_W_STD 112
#ifndef VS_RESIZE_UNROLL
#define VS_RESIZE_UNROLL 6             // full unrolling (5 / 6 predicated rows per thread): 4.31 vs 4.38 ms per step with 2
#endif
constexpr int kResizeUnroll = VS_RESIZE_UNROLL;
template <bool L0, int PWC>
__global__ void __launch_bounds__(256, VS_LEVEL_CTAS)
vs_level_kernel(const __grid_constant__ VsPlan plan, int level, const uint8_t* __restrict__ frames,
                int fpitch, size_t fstride, uint8_t* __restrict__ pyr,
                const uint32_t* __restrict__ tab, VsCand* __restrict__ cand,
                int32_t* __restrict__ counters, const __grid_constant__ CUtensorMap tmap, int use_tma)
{
    // 8 elements of front padding: the networks of the ring column read one word left of row 0
    __shared__ __align__(16) uint16_t s_raw_store[RAW_H * RAW_S + 8];
    uint16_t* const s_raw = s_raw_store + 8;
    __shared__ uint16_t s_queue[SC_H * (VS_TW / 2 + 2)];   // (row << 6) | word of the pixel pairs that pass the pre-test
    __shared__ uint16_t s_queue2[SC_H * (VS_TW / 2)];      // the queued pairs that hold a score and can hold a kept corner
    __shared__ int s_nq, s_nq2;
    __shared__ __align__(16) uint16_t s_sc[SC_H * SC_S];   // max(S - thr, 0), 16-bit lanes
    __shared__ __align__(16) float s_row[BL_ROWS * VS_TW];
    __shared__ uint32_t s_hist[256];
    __shared__ VsCand s_list[LIST_CAP];
    __shared__ int s_n, s_base;
    __shared__ __align__(8) uint32_t s_tx[RAW_W];
    __shared__ uint2 s_ty[RAW_H];       // per tile row: byte offsets of its two parent rows in the staged region, weight of the second
    __shared__ __align__(8) unsigned long long s_mbar;   // TMA completion barrier (levels >= 1)
    extern __shared__ __align__(16) uint8_t s_dyn[];   // parent region + vertical pass (levels >= 1)

    const VsLevel& L = plan.lv[level];
    const int tid = threadIdx.x;
    const int b = blockIdx.z;
    const int x0 = blockIdx.x * VS_TW, y0 = blockIdx.y * VS_TH;
    const int W = L.w, H = L.h;
    uint8_t* fpyr = pyr + (size_t)b * plan.pyr_frame_bytes;

    s_hist[tid] = 0;
    if (tid == 0) { s_nq = 0; s_nq2 = 0; s_n = 0; }

    // ---- 1. raw tile (16-bit lanes) ------------------------------------------
    if (L0 && use_tma) {
        // Level 0 by TMA, like the parent regions of the other levels: one thread issues a (L0_BOX_W x RAW_H) u8 box
        // at (x0 - 16, y0 - 4, frame) -- a box must start on a 16-byte boundary, the tile starts at x0 - 4 -- and the
        // CTA widens it to 16-bit lanes from shared memory.  Pixels outside the frame arrive as zeros; border tiles
        // read their REFLECT_101 halo from the mirrored position inside the same box (the halo is 4 wide, the box has
        // 12 spare columns on either side).
        __shared__ __align__(128) uint8_t s_box[L0_BOX_W * RAW_H];
        if (tid == 0) {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_box);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(L0_BOX_W * RAW_H) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"((uint64_t)&tmap), "r"(x0 - 16), "r"(y0 - VS_HALO), "r"(b), "r"(bar) : "memory");
        }
        __syncthreads();
        {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
            asm volatile(
                "{\n .reg .pred p;\n"
                "W%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                " @p bra D%=;\n bra W%=;\n"
                "D%=:\n}\n" ::"r"(bar) : "memory");
        }
        const int wd = tid % (RAW_W / 4), rg = tid / (RAW_W / 4);        // 18 word columns x 14 row groups
        const int gx = x0 - VS_HALO + 4 * wd;
        const bool yin = y0 >= VS_HALO && y0 + VS_TH + VS_HALO <= H;    // no mirrored rows in this tile
        const bool xin = gx >= 0 && gx + 3 < W;                          // this word lies inside the frame's columns
        const bool xleft = gx == -VS_HALO && W >= 8;                     // columns -4 .. -1 = columns 4, 3, 2, 1
        if (rg < 14) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int r = rg + 14 * k;
                if (r < RAW_H) {
                    // mirrored row, clamped into the box where the pixel cannot matter (beyond the halo)
                    int br = r;
                    if (!yin) br = min(max((H - 1) - abs((H - 1) - abs(y0 - VS_HALO + r)) - (y0 - VS_HALO), 0), RAW_H - 1);
                    const uint8_t* row = s_box + br * L0_BOX_W;
                    uint32_t u;
                    if (xin) {
                        u = *(const uint32_t*)(row + 12 + 4 * wd);
                    } else if (xleft) {
                        u = __byte_perm(*(const uint32_t*)(row + 16), *(const uint32_t*)(row + 20), 0x1234);
                    } else {
                        u = 0;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int bc = min(max((W - 1) - abs((W - 1) - abs(gx + j)) - (x0 - 16), 0), L0_BOX_W - 1);
                            u |= (uint32_t)row[bc] << (8 * j);
                        }
                    }
                    *(uint2*)(s_raw + r * RAW_S + 4 * wd) = make_uint2(__byte_perm(u, 0, 0x4140), __byte_perm(u, 0, 0x4342));
                }
            }
        }
    } else if (L0) {
        const uint8_t* src = frames + (size_t)b * fstride;
        const bool aligned = (((size_t)src | (size_t)fpitch) & 3) == 0;
        // thread = one 4-pixel word column x up to 3 rows; the loads are issued before any store
        const int wd = tid % (RAW_W / 4), rg = tid / (RAW_W / 4);        // 18 word columns x 14 row groups
        const bool interior = x0 >= VS_HALO && x0 + VS_TW + VS_HALO <= W && y0 >= VS_HALO && y0 + VS_TH + VS_HALO <= H;
        if (rg < 14) {
            uint32_t u[3];
            if (aligned && interior) {      // no reflection anywhere in the tile: three strided row loads
                const uint8_t* p0 = src + (size_t)(y0 - VS_HALO + rg) * fpitch + (x0 - VS_HALO) + 4 * wd;
                u[0] = __ldg((const uint32_t*)p0);
                u[1] = __ldg((const uint32_t*)(p0 + (size_t)14 * fpitch));
                u[2] = rg < RAW_H - 28 ? __ldg((const uint32_t*)(p0 + (size_t)28 * fpitch)) : 0u;
            } else {
                const int gx = x0 - VS_HALO + 4 * wd;
                const bool fast = aligned && gx >= 0 && gx + 3 < W;
                const bool small = W < 8 || H < 8;                      // several reflections per side possible
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int r = rg + 14 * k;
                    u[k] = 0;
                    if (r < RAW_H) {
                        const int gy = y0 - VS_HALO + r;
                        const int ry = small ? vs_reflect101(gy, H) : (H - 1) - abs((H - 1) - abs(gy));
                        const uint8_t* row = src + (size_t)ry * fpitch;
                        u[k] = fast ? __ldg((const uint32_t*)(row + gx)) : small ? load4_reflect(row, gx, W) : load4_edge(row, gx, W);
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int r = rg + 14 * k;
                if (r < RAW_H)
                    *(uint2*)(s_raw + r * RAW_S + 4 * wd) = make_uint2(__byte_perm(u[k], 0, 0x4140), __byte_perm(u[k], 0, 0x4342));
            }
        }
    } else {
        const VsLevel& P = plan.lv[level - 1];
        const uint8_t* src;
        int sp;
        if (level == 1) { src = frames + (size_t)b * fstride; sp = fpitch; }
        else { src = fpyr + P.raw_off; sp = P.pitch; }
        const bool aligned = (((size_t)src | (size_t)sp) & 3) == 0;
        const int sw = P.w, sh = P.h;
        const int PW = PWC ? PWC : plan.par_w;
        // dynamic smem: [par_h][par_w] parent pixels (128-byte aligned for TMA), then the horizontal pass
        // (offset arithmetic on s_dyn, not on a generic address, so that the accesses stay LDS/STS)
        uint8_t* s_par = s_dyn + ((128u - ((uint32_t)__cvta_generic_to_shared(s_dyn) & 127u)) & 127u);
        // [RAW_H][par_w] vertical pass (8.8 fixed point): dead before the Gaussian row pass writes s_row, so it
        // lives there when it fits (always at the reference's scale 1.2) and the CTA stays under 1/6 of the SM
        uint16_t* s_v = (size_t)RAW_H * PW * sizeof(uint16_t) <= sizeof(s_row) ? (uint16_t*)s_row
                                                                               : (uint16_t*)(s_par + (size_t)plan.par_h * PW);
        // bounding box of the parent pixels this tile reads (host-computed per tile column / row)
        const uint32_t bx = __ldg(tab + L.bndx_off + blockIdx.x), by = __ldg(tab + L.bndy_off + blockIdx.y);
        const int px_lo = (int)(bx & 0xffff), px_hi = (int)(bx >> 16), py0 = (int)(by & 0xffff);
        const int nrows = (int)(by >> 16) - py0 + 1;
        const int pxa = px_lo & ~15;                                    // parent column of s_par[.][0] (TMA needs 16-byte aligned starts)
        if (use_tma && tid == 0) {
            // one thread: arm the mbarrier with the box size and issue the 3-D tile load (x, y, frame);
            // rows / columns of the box beyond the parent image are zero-filled and never read
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_par);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(PW * plan.par_h) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"((uint64_t)&tmap), "r"(pxa), "r"(py0), "r"(b), "r"(bar) : "memory");
        }
        // table entries of this tile's (reflected) columns / rows
        const int wid = tid >> 5, lane = tid & 31;
        if (wid >= 6) {
            const bool isx = wid == 6;
            const uint32_t* t = tab + (isx ? L.tabx_off : L.taby_off);
            const int n = isx ? RAW_W : RAW_H, o0 = (isx ? x0 : y0) - VS_HALO, lim = isx ? W : H;
            if (isx) {
                for (int c = lane; c < n; c += 32) s_tx[c] = __ldg(t + vs_reflect101(o0 + c, lim));
            } else {
                // rows: what the vertical pass needs, ready-made -- the byte offsets of parent rows j0 and
                // j1 = min(j0 + 1, sh - 1) inside the staged region (low / high half) and the weight of j1
                const int sh = plan.lv[level - 1].h;
                for (int c = lane; c < n; c += 32) {
                    const uint32_t e = __ldg(t + vs_reflect101(o0 + c, lim));
                    const int j0 = (int)(e & 0xffff), j1 = min(j0 + 1, sh - 1);
                    s_ty[c] = make_uint2((uint32_t)((j0 - py0) * PW) | ((uint32_t)((j1 - py0) * PW) << 16), e >> 16);
                }
            }
        }
        if (PWC == 0 && !use_tma) {   // fallback (unaligned frames, very large scale factors; only in the generic instantiation): thread = one word column x every G-th row
            const int nwords = (px_hi - pxa) / 4 + 1;
            const int PWW = PW >> 2, G = 256 / PWW;
            const int wd = tid % PWW, rg = tid / PWW;
            if (rg < G && wd < nwords) {
                const int gx = pxa + 4 * wd;
                const bool fast = aligned && gx + 3 < sw;
                for (int r0 = rg; r0 < nrows; r0 += 4 * G) {
                    uint32_t u[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int r = r0 + k * G;
                        if (r < nrows) {
                            const uint8_t* row = src + (size_t)(py0 + r) * sp;
                            if (fast) u[k] = __ldg((const uint32_t*)(row + gx));
                            else {
                                u[k] = 0;
#pragma unroll
                                for (int j = 0; j < 4; ++j) u[k] |= (uint32_t)__ldg(row + min(gx + j, sw - 1)) << (8 * j);
                            }
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int r = r0 + k * G;
                        if (r < nrows) *(uint32_t*)(s_par + r * PW + 4 * wd) = u[k];
                    }
                }
            }
        }
        __syncthreads();
        if (use_tma) {    // wait for the tile (phase 0 of the single-use barrier)
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
            asm volatile(
                "{\n .reg .pred p;\n"
                "W%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                " @p bra D%=;\n bra W%=;\n"
                "D%=:\n}\n" ::"r"(bar) : "memory");
        }
        // separable INTER_LINEAR_EXACT, vertical pass first (the weights are exact integers, so the order of
        // the two passes does not change the result).  Vertical: thread = one parent word column x every G-th
        // output row; four bytes of parent rows j0 / j1 are widened to 16-bit lanes and blended two at a time --
        // a lane holds at most 255 * 256, so the packed multiply-adds cannot carry between lanes.
        {
            const int PWW = PW >> 2, G = 256 / PWW;                        // 28 word columns x 9 row groups at PW = 112
            const int wc = tid % PWW, rg = tid / PWW;
            const int nwords = min((px_hi + 1 - pxa) / 4 + 1, PWW);
            if (rg < G && wc < nwords) {
                const uint32_t* par32 = (const uint32_t*)s_par + wc;
                uint2* dst = (uint2*)s_v + wc;
                const int VIT = PWC ? (RAW_H + (256 / (PWC / 4)) - 1) / (256 / (PWC / 4)) : RAW_H;   // 5 at PW = 112
#pragma unroll kResizeUnroll
                for (int k = 0; k < VIT; ++k) {
                    const int row = rg + k * G;
                    if (row < RAW_H) {
                        const uint2 e = s_ty[row];
                        const uint32_t bq = e.y, wq = 256u - bq;
                        const uint32_t ua = *(const uint32_t*)((const uint8_t*)par32 + (e.x & 0xffffu));
                        const uint32_t ub = *(const uint32_t*)((const uint8_t*)par32 + (e.x >> 16));
                        const uint32_t lo = __byte_perm(ua, 0, 0x4140) * wq + __byte_perm(ub, 0, 0x4140) * bq;
                        const uint32_t hi = __byte_perm(ua, 0, 0x4342) * wq + __byte_perm(ub, 0, 0x4342) * bq;
                        dst[row * PWW] = make_uint2(lo, hi);
                    }
                }
            }
        }
        __syncthreads();
        // Horizontal: a column pair per thread, fixed over the rows it handles, so that the table entries are loop
        // invariant.  Lane = column pair 0 .. 31 of one row, warp w = rows w, w + 8, ...: the 32 lanes of a load read
        // neighbouring 16-bit entries of ONE row of the vertical pass (27 words: no bank conflict; with lanes spread
        // over two rows the second row sat 24 banks further and every load cost two wavefronts).  The four pairs
        // 32 .. 35 of all rows follow as one pass of 152 threads.
        // A column with i0 = sw-1 has weight 0 on its right neighbour, whatever that value is.
        auto hpass = [&](int cp, int row0, int rstep, int nit) {
            const uint2 et = *(const uint2*)(s_tx + 2 * cp);
            const uint32_t ea = et.x, eb = et.y;
            const int ia = (int)(ea & 0xffff) - pxa, ib = (int)(eb & 0xffff) - pxa;
            const uint32_t aa = ea >> 16, ab = eb >> 16, wa = 256u - aa, wb = 256u - ab;
            uint32_t* out = (uint32_t*)s_raw + row0 * (RAW_S / 2) + cp;
            const uint16_t* va = s_v + row0 * PW + ia;
            const uint16_t* vb = s_v + row0 * PW + ib;
#pragma unroll
            for (int k = 0; k < nit; ++k) {
                if (row0 + rstep * k < RAW_H) {
                    const uint32_t lo = (uint32_t)va[rstep * k * PW] * wa + (uint32_t)va[rstep * k * PW + 1] * aa + 32768u;
                    const uint32_t hi = (uint32_t)vb[rstep * k * PW] * wb + (uint32_t)vb[rstep * k * PW + 1] * ab + 32768u;
                    out[rstep * k * (RAW_S / 2)] = (lo >> 16) | (hi & 0xffff0000u);
                }
            }
        };
        hpass(tid & 31, tid >> 5, 8, (RAW_H + 7) / 8);
        if (tid < 4 * RAW_H) hpass(32 + (tid & 3), tid >> 2, 0, 1);
    }
    __syncthreads();

    // ---- 2. store the raw level -------------------------------------------
    if (!L0 && tid < VS_TH * 4) {
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
    const uint32_t kbias15 = (uint32_t)(0x7fff - thr) * 0x00010001u;
    const int cx_lo = VS_EDGE - 1, cx_hi = W - VS_EDGE, cy_lo = VS_EDGE - 1, cy_hi = H - VS_EDGE;
    const int lane = tid & 31;
    // 3a. compass pre-test.  Any arc of 9 contains at least two of the four compass points, so a corner
    // needs its second-largest compass value above c + thr or its second-smallest below c - thr.  Pixel
    // pairs that pass are queued; the full min/max networks (3b) then run on whole warps of queued pairs
    // only (a third to a half of the pairs on textured images) instead of on every pixel.
    {
        const int w = tid & 15;
        uint32_t need[4];                 // pass k = score row (tid >> 4) + 16k; [2k] = pair A, [2k+1] = pair B
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int r = (tid >> 4) + 16 * k;
            const int gy = y0 - 1 + r, gx = x0 + 4 * w;
            need[2 * k] = need[2 * k + 1] = 0;
            if (gy >= cy_lo && gy <= cy_hi && gx + 3 >= cx_lo && gx <= cx_hi) {
                const uint16_t* base = s_raw + (r + 3) * RAW_S + 4 * w;   // word 2 of this row = pixels 0,1
                const uint2 up = *(const uint2*)(base + 3 * RAW_S + 4), dn = *(const uint2*)(base - 3 * RAW_S + 4);
                const uint2* q = (const uint2*)base;
                const uint2 a = q[0], m = q[1], e = q[2];
                need[2 * k] = compass_need(up.x, dn.x, fsr16(m.y, e.x), fsr16(a.x, a.y), m.x, kbias15);
                need[2 * k + 1] = compass_need(up.y, dn.y, fsr16(e.x, e.y), fsr16(a.y, m.x), m.y, kbias15);
            }
            *(uint2*)(s_sc + r * SC_S + 4 + 4 * w) = make_uint2(0u, 0u);
        }
        // one warp-aggregated push of (row, word) entries for the four masks, in row-major order (lane = word pair
        // of row r0 or r0 + 1: pair A then pair B per lane, rows r0 + 16 after): a warp of the network pass then
        // reads runs of neighbouring words of one row instead of every second word of two rows -- half the bank
        // conflicts of the 20 loads per queued pair
        uint32_t mk[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) mk[j] = __ballot_sync(0xffffffffu, need[j] != 0);
        const int c01 = __popc(mk[0]) + __popc(mk[1]), c23 = __popc(mk[2]) + __popc(mk[3]);
        if (c01 + c23) {
            int basepos = 0;
            if (lane == 0) basepos = smem_atomic_add(&s_nq, c01 + c23);
            basepos = __shfl_sync(0xffffffffu, basepos, 0);
            const uint32_t lt = (1u << lane) - 1u;
            const int r0 = tid >> 4;
            // shared-window addresses and entry values once, then four predicated 16-bit stores (written as indexed
            // array stores the compiler rebuilt base, index and value for each of them: ~40 instructions)
            const uint32_t q_u = (uint32_t)__cvta_generic_to_shared(s_queue);
            const uint32_t a01 = q_u + 2u * (uint32_t)(basepos + __popc(mk[0] & lt) + __popc(mk[1] & lt));
            const uint32_t a23 = q_u + 2u * (uint32_t)(basepos + c01 + __popc(mk[2] & lt) + __popc(mk[3] & lt));
            const uint32_t e0 = (uint32_t)((r0 << 6) | (2 * w + 2));
            if (need[0]) sts_u16(a01, e0);
            if (need[1]) sts_u16(a01 + (need[0] ? 2u : 0u), e0 + 1u);
            if (need[2]) sts_u16(a23, e0 + (16u << 6));
            if (need[3]) sts_u16(a23 + (need[2] ? 2u : 0u), e0 + (16u << 6) + 1u);
        }
    }
    if (tid >= 256 - 2 * SC_H) {   // ring columns tx = -1 (word 1, high lane) and tx = TW (word TW/2 + 2, low lane)
        const int t2 = tid - (256 - 2 * SC_H);
        const int r = t2 % SC_H, side = t2 / SC_H;
        const int wi = side ? VS_TW / 2 + 2 : 1;
        const int gx = x0 + (side ? VS_TW : -1), gy = y0 - 1 + r;
        if (gy >= cy_lo && gy <= cy_hi && gx >= cx_lo && gx <= cx_hi) {
            const uint32_t* rw = (const uint32_t*)(s_raw + (r + 3) * RAW_S) + wi;
            const uint32_t need = compass_need(rw[3 * (RAW_S / 2)], rw[-3 * (RAW_S / 2)], fsr16(rw[1], rw[2]), fsr16(rw[-2], rw[-1]),
                                               rw[0], kbias15) & (side ? 0x00008000u : 0x80000000u);
            if (need) s_queue[smem_atomic_add(&s_nq, 1)] = (uint16_t)((r << 6) | wi);
        }
        ((uint32_t*)(s_sc + r * SC_S))[wi] = 0u;
    }

    // ---- 4a. Gaussian 7x7, row pass (reads the raw tile only, so it shares the barrier interval of 3a).
    // cv2's vector blocks use FMA, its scalar tails do not: rows switch at W - W%32,
    // columns at W - W%8; an aligned 8-pixel group never straddles either boundary.
    const int wrow = W & ~31, wcol = W & ~7;
    auto row_item = [&](int i) {
        const int g = i & 7, rr = i >> 3;                             // s_row row rr <-> gy = y0 - 3 + rr
        if (y0 - 3 + rr >= H + 3) return;                             // below every output row of the level
        const uint2* q = (const uint2*)(s_raw + (rr + 1) * RAW_S + 8 * g);
        // 4-float chunk c of a row sits at chunk position c ^ (c >> 3): the eight lanes of a store phase (g = 0..7,
        // chunks 2g) then cover all 32 banks instead of colliding pairwise (g, g + 4); the column pass reads chunk w
        // of 16 neighbouring lanes at w ^ (w >> 3), still conflict-free
        float* dst = s_row + rr * VS_TW;
        const int sw = g >> 2;
        if (x0 + 8 * g < wrow) blur_row_item<true>(q, dst + 4 * ((2 * g) ^ sw), dst + 4 * ((2 * g + 1) ^ sw));
        else blur_row_item<false>(q, dst + 4 * ((2 * g) ^ sw), dst + 4 * ((2 * g + 1) ^ sw));
    };
    // 288 items: one per thread here; the last 32 (rows 32 .. 35) run on the last warp in the NMS interval, where
    // only the one or two warps that hold scored pairs have work -- a second pass here would keep seven warps
    // waiting at the barrier for the one that has it
    row_item(tid);
    __syncthreads();

    // ---- 3b. exact scores of the queued pairs: word wi of score row r <-> tile columns 2wi, 2wi+1.  Pairs that
    // hold a score and can hold a kept corner (not the ring, rows inside the 31-px border) go to a second queue:
    // the NMS pass then runs on whole warps of scored pairs instead of walking the first queue with most lanes idle.
    {
        const int nqueue = s_nq;
        const int cyk_lo = max(VS_EDGE - (y0 - 1), 1), cyk_hi = min(H - VS_EDGE - (y0 - 1), SC_H - 1);   // kept score rows [lo, hi)
        for (int q0 = tid & ~31; q0 < nqueue; q0 += 256) {              // warp-uniform trip count (ballot below)
            const int qi = q0 + lane;
            bool keep = false;
            int e = 0;
            if (qi < nqueue) {
                e = s_queue[qi];
                const int r = e >> 6, wi = e & 63;
                const uint32_t* rw = (const uint32_t*)(s_raw + (r + 3) * RAW_S) + wi;
                const int RW = RAW_S / 2;
                uint32_t v[16];
                {
                    const uint32_t m1 = rw[3 * RW - 1], z = rw[3 * RW], p1 = rw[3 * RW + 1];
                    v[0] = z; v[1] = fsr16(z, p1); v[15] = fsr16(m1, z);
                }
                {
                    const uint32_t m1 = rw[-3 * RW - 1], z = rw[-3 * RW], p1 = rw[-3 * RW + 1];
                    v[8] = z; v[7] = fsr16(z, p1); v[9] = fsr16(m1, z);
                }
                v[2] = rw[2 * RW + 1]; v[14] = rw[2 * RW - 1];
                v[6] = rw[-2 * RW + 1]; v[10] = rw[-2 * RW - 1];
                v[3] = fsr16(rw[RW + 1], rw[RW + 2]); v[13] = fsr16(rw[RW - 2], rw[RW - 1]);
                v[5] = fsr16(rw[-RW + 1], rw[-RW + 2]); v[11] = fsr16(rw[-RW - 2], rw[-RW - 1]);
                v[4] = fsr16(rw[1], rw[2]); v[12] = fsr16(rw[-2], rw[-1]);
                uint32_t o = fast_pair(v, rw[0], kbias);
                const int gx = x0 - VS_HALO + 2 * wi;
                if (gx < cx_lo || gx > cx_hi) o &= 0xffff0000u;
                if (gx + 1 < cx_lo || gx + 1 > cx_hi) o &= 0x0000ffffu;
                ((uint32_t*)(s_sc + r * SC_S))[wi] = o;
                keep = o != 0 && r >= cyk_lo && r < cyk_hi && wi >= 2 && wi <= VS_TW / 2 + 1;
            }
            const uint32_t km = __ballot_sync(0xffffffffu, keep);
            if (km) {
                int base = 0;
                if (lane == 0) base = smem_atomic_add(&s_nq2, __popc(km));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (keep) s_queue2[base + __popc(km & ((1u << lane) - 1u))] = (uint16_t)e;
            }
        }
    }
    __syncthreads();

    if (tid >= 256 - (BL_ROWS * (VS_TW / 8) - 256)) row_item(tid + (BL_ROWS * (VS_TW / 8) - 256));   // the rest of 4a

    // ---- 5. non-max suppression (strict >, non-corners count as 0) + 31-px border -> tile list + histogram.
    // Word wi of score row r and its eight neighbouring words give the 3x3 neighbourhoods of both pixels at once.
    {
        const int nq2 = s_nq2;
        for (int qi = tid; qi < nq2; qi += 256) {
            const int e = s_queue2[qi], r = e >> 6, wi = e & 63;
            const uint32_t* q = (const uint32_t*)(s_sc + r * SC_S) + wi;
            const uint32_t cm = q[0];
            const int gy = y0 - 1 + r;
            const int SW = SC_S / 2;
            const uint32_t ml = q[-1], mr = q[1], ul = q[-SW - 1], uc = q[-SW], ur = q[-SW + 1],
                           dl = q[SW - 1], dc = q[SW], dr = q[SW + 1];
            // lanes = (left pixel, right pixel) of the pair; fsr16(l, c) = their left neighbours, fsr16(c, r) = right
            const uint32_t nb = __vimax3_u16x2(__vimax3_u16x2(fsr16(ul, uc), uc, fsr16(uc, ur)),
                                               __vimax3_u16x2(fsr16(dl, dc), dc, fsr16(dc, dr)),
                                               __vmaxu2(fsr16(ml, cm), fsr16(cm, mr)));
            // bit 15 of a lane of (c + 0x7fff - nb) is set iff c > nb (no carry between lanes: both < 256)
            const uint32_t kk = ((cm + 0x7fff7fffu) - nb) & 0x80008000u;
            if (kk == 0) continue;
            const int gx0 = x0 - VS_HALO + 2 * wi;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int gx = gx0 + j;
                if (!(kk & (j ? 0x80000000u : 0x8000u)) || gx < VS_EDGE || gx >= W - VS_EDGE) continue;
                const int score = (int)(j ? (cm >> 16) : (cm & 0xffff)) + thr - 1;
                const int slot = atomicAdd(&s_n, 1);
                VsCand c;
                c.x = (uint16_t)gx; c.y = (uint16_t)gy; c.v = (uint32_t)score;
                s_list[slot] = c;
                atomicAdd(&s_hist[score], 1u);
            }
        }
    }
    __syncthreads();
    // reserve the tile's slots in the (frame, level) list; the round trip of that atomic hides behind 4b
    int32_t* ctr = counters + ((size_t)b * plan.n_levels + level) * VS_CTR_STRIDE;
    const int n = s_n;
    if (n > 0) {
        if (tid == 0) s_base = atomicAdd(&ctr[VS_CTR_NCAND], n);
        const uint32_t hv = s_hist[tid];
        if (hv) atomicAdd(&ctr[VS_CTR_HIST + tid], (int)hv);
    }

    // ---- 4b. Gaussian 7x7, column pass
    uint8_t* blur = fpyr + L.blur_off;
    for (int i = tid; i < (VS_TH / 2) * (VS_TW / 4); i += 256) {
        const int w = i & 15, ty = (i >> 4) * 2;                      // two output rows per item
        const int gx = x0 + 4 * w, gy = y0 + ty;
        if (gx >= L.pitch || gy >= H) continue;
        uint32_t o0, o1;
        const float* src = s_row + ty * VS_TW + 4 * (w ^ (w >> 3));
        if (gx < wcol) blur_col_item<true>(src, o0, o1);
        else blur_col_item<false>(src, o0, o1);
        uint8_t* dst = blur + (size_t)gy * L.pitch + gx;
        *(uint32_t*)dst = o0;
        if (gy + 1 < H) *(uint32_t*)(dst + L.pitch) = o1;
    }
    if (n > 0) {
        __syncthreads();
        VsCand* out = cand + (size_t)b * plan.cand_frame_entries + L.cand_off + s_base;
        for (int i = tid; i < n; i += 256) out[i] = s_list[i];
    }
}

size_t vs_level_dyn_smem(const VsPlan& p)
{
    const size_t v = (size_t)RAW_H * p.par_w * sizeof(uint16_t);
    return 128 + (size_t)p.par_h * p.par_w + (v <= sizeof(float) * BL_ROWS * VS_TW ? 0 : v);
}

int vs_configure_level_smem(const VsPlan& p)
{
    const size_t dyn = vs_level_dyn_smem(p);
    if (dyn > 160 * 1024 || p.par_w > 1024 || (size_t)p.par_h * p.par_w > 65535) {   // the row table holds 16-bit byte offsets into the region
        vs_set_error("scale_factor too large for the level kernel's parent staging (%zu bytes)", dyn);
        return VSLAM_E_INVALID;
    }
    // only ever raised, per device: handles with different plans (scale factors) share the function attribute
    static int cur[64] = {};
    int dev = 0;
    VS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) dev = 0;
    if ((int)dyn > __atomic_load_n(&cur[dev], __ATOMIC_ACQUIRE)) {
        VS_CUDA(cudaFuncSetAttribute(vs_level_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        VS_CUDA(cudaFuncSetAttribute(vs_level_kernel<false, VS_PAR_W_STD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        __atomic_store_n(&cur[dev], (int)dyn, __ATOMIC_RELEASE);
    }
    return VSLAM_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point query (no link against libcuda)
typedef CUresult (*vs_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                       const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                       CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static vs_encode_tiled_fn vs_encode_tiled()
{
    static vs_encode_tiled_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (vs_encode_tiled_fn)f;
        cudaGetLastError();
    }
    return fn;
}

// Tensor map of an image stack uint8 [frames][h][pitch] with a (box_w, box_h, 1) box.  Returns false when TMA
// cannot describe it (unaligned base / strides, box too large); the kernels then stage by hand.  A box must start
// on a 16-byte boundary of the innermost dimension (an unaligned x faults with "illegal instruction").  swizzle64:
// the box lands in shared memory with bits 4-5 of every byte address XORed with bits 7-8 (CU_TENSOR_MAP_SWIZZLE_64B).
bool vs_image_tensor_map(struct CUtensorMap_st* tm, const uint8_t* base, int w, int h, size_t pitch, size_t frame_stride,
                         int frames, int box_w, int box_h, bool swizzle64)
{
    vs_encode_tiled_fn enc = vs_encode_tiled();
    if (!enc || box_w > 256 || box_h > 256 || (box_w & 15) || w <= 0 || h <= 0 || frames <= 0) return false;
    if (swizzle64 && box_w != 64) return false;   // the box row is exactly one swizzle span
    if (((uintptr_t)base & 15) || (pitch & 15) || (frame_stride & 15)) return false;
    const cuuint64_t gdim[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)frames};
    const cuuint64_t gstr[2] = {(cuuint64_t)pitch, (cuuint64_t)frame_stride};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               swizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
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
        CUtensorMap tm;
        memset(&tm, 0, sizeof(tm));
        int use_tma = 0;
        if (l == 0)       // frames of at least one tile's halo in each direction (smaller ones reflect more than once)
            use_tma = p.width >= 16 && p.height >= 16 &&
                      vs_image_tensor_map(&tm, d_frames, p.width, p.height, (size_t)pitch, (size_t)pitch * p.height, batch,
                                          L0_BOX_W, RAW_H);
        else if (l == 1)
            use_tma = vs_image_tensor_map(&tm, d_frames, p.width, p.height, (size_t)pitch, (size_t)pitch * p.height, batch,
                                          p.par_w, p.par_h);
        else if (l > 1)
            use_tma = vs_image_tensor_map(&tm, o->d_pyr + p.lv[l - 1].raw_off, p.lv[l - 1].w, p.lv[l - 1].h,
                                          (size_t)p.lv[l - 1].pitch, p.pyr_frame_bytes, batch, p.par_w, p.par_h);
        dim3 grid(L.tiles_x, L.tiles_y, batch);
        vs_prof_mark(o, l, s);
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        if (l == 0)
            vs_level_kernel<true, 0><<<grid, 256, 0, s>>>(p, l, d_frames, pitch, (size_t)pitch * p.height, o->d_pyr, o->d_tab,
                                                          o->d_cand, o->d_counters, tm, use_tma);
        else if (p.par_w == VS_PAR_W_STD && use_tma)
            vs_level_kernel<false, VS_PAR_W_STD><<<grid, 256, dyn, s>>>(p, l, d_frames, pitch, (size_t)pitch * p.height, o->d_pyr,
                                                                        o->d_tab, o->d_cand, o->d_counters, tm, use_tma);
        else
            vs_level_kernel<false, 0><<<grid, 256, dyn, s>>>(p, l, d_frames, pitch, (size_t)pitch * p.height, o->d_pyr, o->d_tab,
                                                             o->d_cand, o->d_counters, tm, use_tma);
    }
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
