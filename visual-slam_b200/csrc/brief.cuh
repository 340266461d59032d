// brief.cuh -- steered rBRIEF sampling shared by the extraction and compute() kernels.
#pragma once
#include "common.cuh"

struct VsPat { int8_t x0, y0, x1, y1; };
__device__ const VsPat g_pattern[256] = {
#include "rbrief_pattern.inc"
};

// One 256-bit steered rBRIEF descriptor by one warp: lane j produces byte j.
// s_pat: pattern transposed so that lane j reads test 8j+t at s_pat[t*32 + j].
__device__ __forceinline__ uint32_t brief_byte(const uint8_t* blur, const uint8_t* raw, int w, int h,
                                               int pitch, int raw_pitch, int cx, int cy, float a,
                                               float b, const uint32_t* s_pat, int lane)
{
    uint32_t byte = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        uint32_t pw = s_pat[t * 32 + lane];
        int val[2];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            float px = (float)(int8_t)(pw >> (16 * s)), py = (float)(int8_t)(pw >> (16 * s + 8));
            float fx = __fsub_rn(__fmul_rn(px, a), __fmul_rn(py, b));
            float fy = __fadd_rn(__fmul_rn(px, b), __fmul_rn(py, a));
            int ix = cx + __float2int_rn(fx), iy = cy + __float2int_rn(fy);
            if (ix >= 0 && ix < w && iy >= 0 && iy < h)
                val[s] = __ldg(blur + (size_t)iy * pitch + ix);
            else   // OpenCV's bordered pyramid holds REFLECT_101 copies of the unblurred level there
                val[s] = __ldg(raw + (size_t)vs_reflect101(iy, h) * raw_pitch + vs_reflect101(ix, w));
        }
        byte |= (uint32_t)(val[0] < val[1]) << t;
    }
    return byte;
}

__device__ __forceinline__ void load_pattern(uint32_t* s_pat)
{
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        VsPat p = g_pattern[i];
        uint32_t w = (uint32_t)(uint8_t)p.x0 | ((uint32_t)(uint8_t)p.y0 << 8) |
                     ((uint32_t)(uint8_t)p.x1 << 16) | ((uint32_t)(uint8_t)p.y1 << 24);
        s_pat[(i & 7) * 32 + (i >> 3)] = w;
    }
}

