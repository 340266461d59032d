// bisect the TMA tile load: ./tma_test2 rank(2|3) boxw boxh x y z fence(0|1)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                           const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void k(const __grid_constant__ CUtensorMap tmap, int rank, int x, int y, int z, int bytes, uint8_t* out, int fence)
{
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(8) unsigned long long s_mbar;
    uint8_t* s_par = (uint8_t*)(((uintptr_t)s_dyn + 127) & ~(uintptr_t)127);
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_par);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        if (fence) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        else asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        if (rank == 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst), "l"((uint64_t)&tmap), "r"(x), "r"(y), "r"(bar) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"((uint64_t)&tmap), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
    }
    asm volatile(
        "{\n .reg .pred p;\n"
        "W%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
        " @p bra D%=;\n bra W%=;\n"
        "D%=:\n}\n" ::"r"(bar) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = s_par[i];
}
int main(int argc, char** argv)
{
    const int rank = atoi(argv[1]), BW = atoi(argv[2]), BH = atoi(argv[3]), x = atoi(argv[4]), y = atoi(argv[5]), z = atoi(argv[6]),
              fence = atoi(argv[7]);
    const int W = 640, H = 480, B = 3, pitch = 640;
    std::vector<uint8_t> h((size_t)B * H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, BW * BH);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* f = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
    CUtensorMap tm; memset(&tm, 0, sizeof(tm));
    const cuuint64_t gdim[3] = {W, H, B}; const cuuint64_t gstr[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * H};
    const cuuint32_t box[3] = {(cuuint32_t)BW, (cuuint32_t)BH, 1}; const cuuint32_t es[3] = {1, 1, 1};
    CUresult r = ((enc_fn)f)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    k<<<1, 256, BW * BH + 256>>>(tm, rank, x, y, z, BW * BH, o, fence);
    cudaError_t e = cudaDeviceSynchronize();
    int bad = -1;
    if (e == cudaSuccess) {
        std::vector<uint8_t> g(BW * BH); cudaMemcpy(g.data(), o, g.size(), cudaMemcpyDeviceToHost);
        bad = 0;
        for (int yy = 0; yy < BH; ++yy) for (int xx = 0; xx < BW; ++xx) {
            int gx = x + xx, gy = y + yy;
            uint8_t ref = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? h[((size_t)(rank == 3 ? z : 0) * H + gy) * pitch + gx] : 0;
            if (g[yy * BW + xx] != ref) ++bad;
        }
    }
    printf("rank %d box %dx%d at (%d,%d,%d) fence %d: encode %d, %s, mismatches %d\n", rank, BW, BH, x, y, z, fence, (int)r,
           cudaGetErrorString(e), bad);
    return 0;
}
