// bf_match_noalloc.cu -- vslam_bf_knn2_hamming256_noalloc: brute-force Hamming kNN-2 for callers that
// must not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  It is the split-train
// XOR + POPC formulation of bf_match.cu (vs_knn2_kernel + vs_knn2_merge_kernel: K9 of SURVEY.md 8(a),
// ascending strict-'<' scan, ties -> lowest train index), whose per-split partial top-2 lists live in a
// file-scope __device__ array instead of cudaMalloc'ed scratch.  Every GPU that loads the library holds
// the array (VS_NA_SPLITS x VSLAM_NOALLOC_MAX_QUERY x 16 B = 2 MiB) until the process exits.
//
// The scratch is shared by every call: two calls that overlap on the GPU (different streams, or
// different host threads on different streams) race on it; callers serialise them.  Nothing in it has
// to start at zero -- a call writes every partial slot (all splits of rows [0, nq)) before its merge
// kernel reads them, so the previous call's contents are never observed.
//
// The entry point calls no cudaMalloc / cudaFree / allocating library routine.  The one piece of set-up,
// cudaGetSymbolAddress (the array's name is only a host-side stand-in in host code, not a device
// address), runs once per device on the first call made with that device current.
#include <mutex>

#include "common.cuh"

#define VS_NA_THREADS 128
#define VS_NA_TILE 128      // train descriptors per shared-memory tile (4 KiB)
#define VS_NA_SPLITS 16     // train-range splits at most (grid y)
#define VS_NA_MAX_DEVICES 64

static __device__ __align__(256) int4 g_na_partial[VS_NA_SPLITS * VSLAM_NOALLOC_MAX_QUERY];

// (d1, i1) <= (d2, i2): the two best of an ascending scan with strict '<'
struct NaBest2 { int d1, i1, d2, i2; };

__device__ __forceinline__ void na_insert(NaBest2& b, int d, int i)
{
    if (d < b.d2) {
        if (d < b.d1) { b.d2 = b.d1; b.i2 = b.i1; b.d1 = d; b.i1 = i; }
        else { b.d2 = d; b.i2 = i; }
    }
}

// grid (query blocks, nsplit): one query row per thread in 8 registers, the split's train tiles through
// shared memory (broadcast 128-bit reads); partial[split][q] = (d1, i1, d2, i2), i = -1 where missing
__global__ void __launch_bounds__(VS_NA_THREADS)
vs_knn2_noalloc_kernel(const uint8_t* __restrict__ query, int nq, const uint8_t* __restrict__ train, int nt,
                       int nsplit, int4* __restrict__ partial)
{
    __shared__ __align__(16) uint4 s_t[2][VS_NA_TILE * 2];
    const int split = blockIdx.y, tid = threadIdx.x;
    const int q = blockIdx.x * VS_NA_THREADS + tid;
    const uint4* qp = (const uint4*)(query + (size_t)min(q, nq - 1) * 32);
    const uint4 qa = __ldg(qp), qb = __ldg(qp + 1);
    const uint4* tp = (const uint4*)train;
    const int tiles = (nt + VS_NA_TILE - 1) / VS_NA_TILE;
    const int t_begin = (int)((long)tiles * split / nsplit), t_end = (int)((long)tiles * (split + 1) / nsplit);
    NaBest2 best = {1 << 30, -1, 1 << 30, -1};

    auto load_tile = [&](int tile, int buf) {
        const int base = tile * VS_NA_TILE;
        for (int i = tid; i < VS_NA_TILE * 2; i += VS_NA_THREADS) {
            const int row = base + (i >> 1);
            uint4 v = make_uint4(0, 0, 0, 0);
            if (row < nt) v = __ldg(tp + (size_t)row * 2 + (i & 1));
            s_t[buf][i] = v;
        }
    };
    if (t_begin < t_end) load_tile(t_begin, 0);
    __syncthreads();
    for (int tile = t_begin; tile < t_end; ++tile) {
        const int buf = (tile - t_begin) & 1;
        if (tile + 1 < t_end) load_tile(tile + 1, buf ^ 1);
        const int base = tile * VS_NA_TILE;
        const int lim = min(VS_NA_TILE, nt - base);
        const uint4* st = s_t[buf];
#pragma unroll 4
        for (int j = 0; j < lim; ++j) {
            const uint4 a = st[2 * j], b = st[2 * j + 1];
            const int d = __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) +
                          __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) + __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
            na_insert(best, d, base + j);
        }
        __syncthreads();
    }
    if (q < nq) partial[(size_t)split * VSLAM_NOALLOC_MAX_QUERY + q] = make_int4(best.d1, best.i1, best.d2, best.i2);
}

// splits are merged in ascending train order, which keeps the tie rule of the sequential scan
__global__ void vs_knn2_noalloc_merge_kernel(const int4* __restrict__ partial, int nq, int nsplit,
                                             int32_t* __restrict__ out_idx, int32_t* __restrict__ out_dist)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    NaBest2 best = {1 << 30, -1, 1 << 30, -1};
    for (int s = 0; s < nsplit; ++s) {
        const int4 p = partial[(size_t)s * VSLAM_NOALLOC_MAX_QUERY + q];
        if (p.y >= 0) na_insert(best, p.x, p.y);
        if (p.w >= 0) na_insert(best, p.z, p.w);
    }
    out_idx[2 * q] = best.i1; out_idx[2 * q + 1] = best.i2;
    out_dist[2 * q] = best.i1 >= 0 ? best.d1 : -1;
    out_dist[2 * q + 1] = best.i2 >= 0 ? best.d2 : -1;
}

// device address of the scratch on the current device (symbol lookup once per device, no allocation)
static int na_scratch(int4** out)
{
    static int4* ptrs[VS_NA_MAX_DEVICES] = {};
    static std::mutex mu;
    int dev = -1;
    VS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= VS_NA_MAX_DEVICES) {
        vs_set_error("vslam_bf_knn2_hamming256_noalloc: device ordinal %d is outside [0, %d)", dev, VS_NA_MAX_DEVICES);
        return VSLAM_E_INVALID;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        VS_CUDA(cudaGetSymbolAddress(&p, g_na_partial));
        if (!p) {
            vs_set_error("vslam_bf_knn2_hamming256_noalloc: cudaGetSymbolAddress returned a null scratch address");
            return VSLAM_E_CUDA;
        }
        ptrs[dev] = (int4*)p;
    }
    *out = ptrs[dev];
    return VSLAM_OK;
}

extern "C" int vslam_bf_knn2_hamming256_noalloc(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                                                int32_t* d_idx, int32_t* d_dist, void* stream)
{
    VS_REQUIRE(nq >= 0 && nt >= 0, "negative row count");
    if (nq > VSLAM_NOALLOC_MAX_QUERY) {
        vs_set_error("vslam_bf_knn2_hamming256_noalloc: nq = %d exceeds VSLAM_NOALLOC_MAX_QUERY = %d (the static "
                     "device scratch is sized for that many query rows; use vslam_bf_knn2_hamming256 for more)",
                     nq, VSLAM_NOALLOC_MAX_QUERY);
        return VSLAM_E_CAPACITY;
    }
    if (nq == 0) return VSLAM_OK;
    VS_REQUIRE(d_query && d_idx && d_dist, "null query / output pointer");
    VS_REQUIRE(d_train || nt == 0, "null train pointer");
    int4* scratch = nullptr;
    if (int rc = na_scratch(&scratch)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int qblocks = (nq + VS_NA_THREADS - 1) / VS_NA_THREADS;
    const int ttiles = (nt + VS_NA_TILE - 1) / VS_NA_TILE;
    // split the train range until the grid covers the 148 SMs a few times over
    int nsplit = (148 * 4 + qblocks - 1) / qblocks;
    if (nsplit > ttiles) nsplit = ttiles;
    if (nsplit > VS_NA_SPLITS) nsplit = VS_NA_SPLITS;
    if (nsplit < 1) nsplit = 1;
    __atomic_fetch_add(&g_vs_launches, 2ull, __ATOMIC_RELAXED);
    vs_knn2_noalloc_kernel<<<dim3(qblocks, nsplit), VS_NA_THREADS, 0, s>>>(d_query, nq, d_train, nt, nsplit, scratch);
    vs_knn2_noalloc_merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(scratch, nq, nsplit, d_idx, d_dist);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
