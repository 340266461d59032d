// bf_match.cu -- brute-force Hamming kNN-2 over 256-bit descriptors + Lowe ratio filter.
//
// Replaces cv2.BFMatcher(NORM_HAMMING).knnMatch(d1, d2, k=2) (matcher.py:70) and the
// ratio loop of DescriptorMatcher.match (matcher.py:73-81).  Semantics (SURVEY.md 8(a)
// K9): int32 distances, train scanned in ascending index with strict '<' insertion,
// so ties go to the lowest train index for first and for second place.
//
// Layout: one query descriptor per thread in 8 registers; the train set streams through
// shared memory in tiles that every lane reads with broadcast 128-bit loads; per pair
// 8 XOR + 8 POPC + adds and a predicated two-deep insert.  The train range can be split
// over blockIdx.y (partials merged in ascending split order, which preserves the tie
// rule); blockIdx.z runs over descriptor-set pairs (consecutive frames of a batch).
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

#define BF_THREADS 128
#define BF_TILE 128   // train descriptors per shared-memory tile (4 KiB)

struct Best2 { int d1, i1, d2, i2; };

__device__ __forceinline__ void best2_insert(Best2& b, int d, int i)
{
    if (d < b.d2) {
        if (d < b.d1) { b.d2 = b.d1; b.i2 = b.i1; b.d1 = d; b.i1 = i; }
        else { b.d2 = d; b.i2 = i; }
    }
}

__global__ void __launch_bounds__(BF_THREADS)
vs_knn2_kernel(const uint8_t* __restrict__ query, int nq_max, long q_stride,
               const uint8_t* __restrict__ train, int nt_max, long t_stride,
               const int32_t* __restrict__ d_nq, const int32_t* __restrict__ d_nt,
               int nsplit, int4* __restrict__ partial, int32_t* __restrict__ out_idx,
               int32_t* __restrict__ out_dist)
{
    __shared__ __align__(16) uint4 s_t[2][BF_TILE * 2];
    const int pair = blockIdx.z, split = blockIdx.y, tid = threadIdx.x;
    const int nq = d_nq ? min(d_nq[pair], nq_max) : nq_max;
    const int nt = d_nt ? min(d_nt[pair], nt_max) : nt_max;
    const int q = blockIdx.x * BF_THREADS + tid;
    if (blockIdx.x * BF_THREADS >= nq) {
        // rows beyond the live count of this pair still get a defined value
        if (q < nq_max && split == 0) {
            int4 e = make_int4(-1, -1, -1, -1);
            size_t o = ((size_t)pair * nq_max + q) * 2;
            out_idx[o] = e.x; out_idx[o + 1] = e.x; out_dist[o] = e.x; out_dist[o + 1] = e.x;
        }
        return;
    }
    const uint4* qp = (const uint4*)(query + ((size_t)pair * q_stride + min(q, nq - 1)) * 32);
    const uint4 qa = __ldg(qp), qb = __ldg(qp + 1);
    const uint4* tp = (const uint4*)(train + (size_t)pair * t_stride * 32);

    // train range of this split, in whole tiles
    const int tiles = (nt + BF_TILE - 1) / BF_TILE;
    const int t_begin = (int)((long)tiles * split / nsplit), t_end = (int)((long)tiles * (split + 1) / nsplit);
    Best2 best = {1 << 30, -1, 1 << 30, -1};

    auto load_tile = [&](int tile, int buf) {
        int base = tile * BF_TILE;
        for (int i = tid; i < BF_TILE * 2; i += BF_THREADS) {
            int row = base + (i >> 1);
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
        const int base = tile * BF_TILE;
        const int lim = min(BF_TILE, nt - base);
        const uint4* st = s_t[buf];
        if (lim == BF_TILE) {
#pragma unroll 4
            for (int j = 0; j < BF_TILE; ++j) {
                uint4 a = st[2 * j], b = st[2 * j + 1];
                int d = __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) +
                        __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) + __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
                best2_insert(best, d, base + j);
            }
        } else {
            for (int j = 0; j < lim; ++j) {
                uint4 a = st[2 * j], b = st[2 * j + 1];
                int d = __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) +
                        __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) + __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
                best2_insert(best, d, base + j);
            }
        }
        __syncthreads();
    }
    if (q >= nq) {
        if (q < nq_max && split == 0) {
            size_t o = ((size_t)pair * nq_max + q) * 2;
            out_idx[o] = -1; out_idx[o + 1] = -1; out_dist[o] = -1; out_dist[o + 1] = -1;
        }
        return;
    }
    if (nsplit == 1) {
        size_t o = ((size_t)pair * nq_max + q) * 2;
        out_idx[o] = best.i1; out_idx[o + 1] = best.i2;
        out_dist[o] = best.i1 >= 0 ? best.d1 : -1;
        out_dist[o + 1] = best.i2 >= 0 ? best.d2 : -1;
    } else {
        partial[((size_t)pair * nsplit + split) * nq_max + q] = make_int4(best.d1, best.i1, best.d2, best.i2);
    }
}

__global__ void vs_knn2_merge_kernel(const int4* __restrict__ partial, int nq_max, const int32_t* __restrict__ d_nq,
                                     int nsplit, int32_t* __restrict__ out_idx, int32_t* __restrict__ out_dist)
{
    const int pair = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
    const int nq = d_nq ? min(d_nq[pair], nq_max) : nq_max;
    if (q >= nq) return;
    Best2 best = {1 << 30, -1, 1 << 30, -1};
    for (int s = 0; s < nsplit; ++s) {
        int4 p = partial[((size_t)pair * nsplit + s) * nq_max + q];
        if (p.y >= 0) best2_insert(best, p.x, p.y);
        if (p.w >= 0) best2_insert(best, p.z, p.w);
    }
    size_t o = ((size_t)pair * nq_max + q) * 2;
    out_idx[o] = best.i1; out_idx[o + 1] = best.i2;
    out_dist[o] = best.i1 >= 0 ? best.d1 : -1;
    out_dist[o + 1] = best.i2 >= 0 ? best.d2 : -1;
}

int vs_knn2_tc5_launch(const uint8_t* d_q, int nq, long q_stride, const uint8_t* d_t, int nt, long t_stride,
                       const int32_t* d_nq, const int32_t* d_nt, int pairs, int nsplit, int4* partial,
                       int32_t* d_idx, int32_t* d_dist, uint8_t* ws, size_t ws_bytes, cudaStream_t s);

// kNN-2 engines: tcgen05 kind::i8 (bf_match_tc5.cu) up to 65 536 train rows -- its keys pack the train index into
// 16 bits -- and the XOR + POPC kernel above beyond.  VSLAM_KNN=popc forces the POPC kernel at every size: that is the
// formulation BASELINE.json's north_star names, and tools/knn_sweep.py measures its popc-pipe fraction with it.
static int knn_force_popc()
{
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("VSLAM_KNN");
        mode = (e && !strcmp(e, "popc")) ? 1 : 0;
    }
    return mode;
}

// scratch of the split-train path, one per device, shared by every caller of the process on that device: launches
// that use it are serialised
#define VS_MAX_DEVICES 64
struct KnnScratch {
    int4* partial = nullptr;
    size_t bytes = 0;
    cudaEvent_t ev = nullptr;          // end of the last launch pair that used the scratch
    cudaStream_t last = nullptr;
};
static KnnScratch g_knn_scratch[VS_MAX_DEVICES];
static std::mutex g_partial_mu;

int vs_launch_knn2(vslam_orb* prof, const uint8_t* d_q, int nq, long q_stride, const uint8_t* d_t, int nt,
                   long t_stride, const int32_t* d_nq, const int32_t* d_nt, int pairs,
                   int32_t* d_idx, int32_t* d_dist, cudaStream_t s)
{
    if (pairs <= 0 || nq <= 0) return VSLAM_OK;
    const bool tc5 = nt <= 65536 && !knn_force_popc();
    const int qtiles = tc5 ? (nq + 255) / 256 : (nq + BF_THREADS - 1) / BF_THREADS;
    const int ttiles = tc5 ? (nt + 127) / 128 : (nt + BF_TILE - 1) / BF_TILE;
    // split the train range only when the grid would not fill the 148 SMs a few times over
    int nsplit = 1;
    const long want = 148L * 8;
    if ((long)qtiles * pairs < want && ttiles > 1 && !(prof && prof->knn_no_split)) {
        nsplit = (int)((want + (long)qtiles * pairs - 1) / ((long)qtiles * pairs));
        if (nsplit > ttiles) nsplit = ttiles;
        if (nsplit > 64) nsplit = 64;
    }
    std::unique_lock<std::mutex> lock(g_partial_mu, std::defer_lock);
    KnnScratch* sc = nullptr;
    // a handle brings its own partial lists (no process-wide scratch, no cross-stream events: capturable)
    const bool own_partial = nsplit > 1 && prof && prof->d_knn_partial &&
                             (size_t)pairs * nsplit * nq * sizeof(int4) <= prof->knn_partial_bytes;
    if (nsplit > 1 && !own_partial) {
        int dev = 0;
        VS_CUDA(cudaGetDevice(&dev));
        VS_REQUIRE(dev >= 0 && dev < VS_MAX_DEVICES, "device ordinal out of range");
        lock.lock();
        sc = &g_knn_scratch[dev];
        size_t need = (size_t)pairs * nsplit * nq * sizeof(int4);
        if (need > sc->bytes) {
            if (sc->partial) { VS_CUDA(cudaStreamSynchronize(s)); VS_CUDA(cudaFree(sc->partial)); sc->partial = nullptr; sc->bytes = 0; }
            VS_CUDA(cudaMalloc(&sc->partial, need));
            sc->bytes = need;
        }
        // another stream may still be reading / writing the scratch: order this launch behind it
        if (sc->ev && sc->last != s) VS_CUDA(cudaStreamWaitEvent(s, sc->ev, 0));
    }
    int4* partial = own_partial ? prof->d_knn_partial : sc ? sc->partial : nullptr;
    vs_prof_mark(prof, VS_PROF_KNN2, s);
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    if (tc5) {
        // a handle owns the expanded-tile workspace of its own batches (no allocation call on the hot path)
        int rc = vs_knn2_tc5_launch(d_q, nq, q_stride, d_t, nt, t_stride, d_nq, d_nt, pairs, nsplit, partial, d_idx, d_dist,
                                    prof ? prof->d_knn_ws : nullptr, prof ? prof->knn_ws_bytes : 0, s);
        if (rc) return rc;
    } else
        vs_knn2_kernel<<<dim3(qtiles, nsplit, pairs), BF_THREADS, 0, s>>>(
            d_q, nq, q_stride, d_t, nt, t_stride, d_nq, d_nt, nsplit, partial, d_idx, d_dist);
    if (nsplit > 1) {
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_knn2_merge_kernel<<<dim3((nq + 255) / 256, pairs), 256, 0, s>>>(partial, nq, d_nq, nsplit, d_idx, d_dist);
        if (sc) {
            if (!sc->ev) VS_CUDA(cudaEventCreateWithFlags(&sc->ev, cudaEventDisableTiming));
            VS_CUDA(cudaEventRecord(sc->ev, s));
            sc->last = s;
        }
    }
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

// ---------------------------------------------------------------------------
// Ratio filter + ordered compaction: one CTA per pair, block-wide scan over query rows.
__global__ void __launch_bounds__(256)
vs_ratio_kernel(const int32_t* __restrict__ idx, const int32_t* __restrict__ dist, int nq_max,
                const int32_t* __restrict__ d_nq, double ratio, int ratio_test,
                int32_t* __restrict__ match, int32_t* __restrict__ match_count)
{
    __shared__ int s_w[8];
    __shared__ int s_base;
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int nq = d_nq ? min(d_nq[pair], nq_max) : nq_max;
    const int2* ip = (const int2*)(idx + (size_t)pair * nq_max * 2);
    const int2* dp = (const int2*)(dist + (size_t)pair * nq_max * 2);
    int32_t* mo = match + (size_t)pair * nq_max * 3;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int q0 = 0; q0 < nq; q0 += 256) {
        const int q = q0 + tid;
        int keep = 0;
        int2 i2 = make_int2(-1, -1), d2 = make_int2(-1, -1);
        if (q < nq) {
            i2 = ip[q]; d2 = dp[q];
            if (i2.x >= 0 && i2.y >= 0)
                keep = !ratio_test || (double)(float)d2.x < ratio * (double)(float)d2.y;
            else if (i2.x >= 0)
                keep = 1;
        }
        unsigned bal = __ballot_sync(0xffffffffu, keep);
        int wpre = __popc(bal & ((1u << lane) - 1));
        if (lane == 0) s_w[w] = __popc(bal);
        __syncthreads();
        int base = s_base;
        for (int j = 0; j < w; ++j) base += s_w[j];
        if (keep) {
            int o = (base + wpre) * 3;
            mo[o] = q; mo[o + 1] = i2.x; mo[o + 2] = d2.x;
        }
        __syncthreads();
        if (tid == 0) { int t = 0; for (int j = 0; j < 8; ++j) t += s_w[j]; s_base += t; }
        __syncthreads();
    }
    if (tid == 0) match_count[pair] = s_base;
}

int vs_launch_ratio(vslam_orb* prof, const int32_t* d_idx, const int32_t* d_dist, int nq, const int32_t* d_nq,
                    int pairs, double ratio, int ratio_test, int32_t* d_match,
                    int32_t* d_match_count, cudaStream_t s)
{
    if (pairs <= 0) return VSLAM_OK;
    vs_prof_mark(prof, VS_PROF_RATIO, s);
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_ratio_kernel<<<pairs, 256, 0, s>>>(d_idx, d_dist, nq, d_nq, ratio, ratio_test, d_match, d_match_count);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}

// ---------------------------------------------------------------------------
// k = 1 with an optional uint8 [nq][nt] mask (matcher.py:102).  One warp per query row.
__global__ void vs_match_masked_kernel(const uint8_t* __restrict__ query, int nq,
                                       const uint8_t* __restrict__ train, int nt,
                                       const uint8_t* __restrict__ mask, int32_t* __restrict__ out_idx,
                                       int32_t* __restrict__ out_dist)
{
    const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= nq) return;
    const uint4* qp = (const uint4*)(query + (size_t)q * 32);
    const uint4 qa = __ldg(qp), qb = __ldg(qp + 1);
    int bd = 1 << 30, bi = -1;
    for (int t = lane; t < nt; t += 32) {
        if (mask && !mask[(size_t)q * nt + t]) continue;
        const uint4* tp = (const uint4*)(train + (size_t)t * 32);
        uint4 a = __ldg(tp), b = __ldg(tp + 1);
        int d = __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) +
                __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) + __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
        if (d < bd) { bd = d; bi = t; }
    }
    // lowest (distance, index) wins, as a sequential ascending scan with strict '<' would
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        int od = __shfl_xor_sync(0xffffffffu, bd, o), oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi >= 0 && (bi < 0 || od < bd || (od == bd && oi < bi))) { bd = od; bi = oi; }
    }
    if (lane == 0) { out_idx[q] = bi; out_dist[q] = bi >= 0 ? bd : -1; }
}

int vs_launch_match_masked(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, const uint8_t* d_mask,
                           int32_t* d_idx, int32_t* d_dist, cudaStream_t s)
{
    if (nq <= 0) return VSLAM_OK;
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    vs_match_masked_kernel<<<(nq + 7) / 8, 256, 0, s>>>(d_q, nq, d_t, nt, d_mask, d_idx, d_dist);
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
