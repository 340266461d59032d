// minimal check of the 3-D uint8 TMA tile load used by vs_level_kernel
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                           const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
#define BW 96
#define BH 49
__global__ void k(const __grid_constant__ CUtensorMap tmap, int x, int y, int z, uint8_t* out, int mode, const CUtensorMap* gmap)
{
    extern __shared__ __align__(16) uint8_t s_dyn[];
    __shared__ __align__(8) unsigned long long s_mbar;
    uint8_t* s_par = (uint8_t*)(((uintptr_t)s_dyn + 127) & ~(uintptr_t)127);
    if (threadIdx.x == 0) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_par);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (mode >= 1) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(BW * BH) : "memory");
        if (mode >= 2)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"(gmap ? (uint64_t)gmap : (uint64_t)&tmap), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
    }
    __syncthreads();
    if (mode >= 2) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
        asm volatile(
            "{\n .reg .pred p;\n"
            "W%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
            " @p bra D%=;\n bra W%=;\n"
            "D%=:\n}\n" ::"r"(bar) : "memory");
    }
    for (int i = threadIdx.x; i < BW * BH; i += blockDim.x) out[i] = s_par[i];
}
int main()
{
    const int W = 640, H = 480, B = 3, pitch = 640;
    std::vector<uint8_t> h((size_t)B * H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, BW * BH);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* f = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
    printf("entry point: %s q=%d f=%p\n", cudaGetErrorString(e), (int)q, f);
    CUtensorMap tm; memset(&tm, 0, sizeof(tm));
    const cuuint64_t gdim[3] = {W, H, B}; const cuuint64_t gstr[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * H};
    const cuuint32_t box[3] = {BW, BH, 1}; const cuuint32_t es[3] = {1, 1, 1};
    CUresult r = ((enc_fn)f)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    CUtensorMap* dmap; cudaMalloc(&dmap, sizeof(tm)); cudaMemcpy(dmap, &tm, sizeof(tm), cudaMemcpyHostToDevice);
    for (int mode = 0; mode <= 3; ++mode) {
        const int x = 571, y = 440, z = 2;
        k<<<1, 256, BW * BH + 128>>>(tm, x, y, z, o, mode >= 2 ? 2 : mode, mode == 2 ? dmap : nullptr);
        e = cudaDeviceSynchronize();
        printf("mode %d: %s\n", mode, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        if (mode >= 2) {
            std::vector<uint8_t> g(BW * BH); cudaMemcpy(g.data(), o, g.size(), cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int yy = 0; yy < BH; ++yy) for (int xx = 0; xx < BW; ++xx) {
                int gx = x + xx, gy = y + yy;
                uint8_t ref = (gx < W && gy < H) ? h[((size_t)z * H + gy) * pitch + gx] : 0;
                if (g[yy * BW + xx] != ref) ++bad;
            }
            printf("mismatches: %d\n", bad);
        }
    }
    return 0;
}
