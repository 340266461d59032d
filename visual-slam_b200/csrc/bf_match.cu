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

#ifndef VS_KNN_DEFAULT_TC5
#define VS_KNN_DEFAULT_TC5 1
#endif

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

// ---------------------------------------------------------------------------
// Tensor-core variant (nt <= 65536).  Brute-force Hamming matching is a binary GEMM; the POPC
// pipe issues 16 lanes/clk/SM, which the scalar kernel above already saturates (93 % measured),
// while the int8 MMA pipe does the same pair 3.8x faster (tools/ubench/imma.cu: 0.12
// IMMA.16832 per clk per SMSP).  Bits are expanded to 0/1 bytes, so that
//     dot(a, b) = popc(a & b),   hamming = popc(a) + popc(b) - 2 dot   (exact int32).
// The k order of a dot product is free: k-step s of a descriptor is its 32-bit word s, thread
// t of a quad owns byte t of that word (low nibble -> a0/a1/b0, high nibble -> a2/a3/b1).
// One CTA = 256 query rows (8 warps x two m16 tiles, A fragments resident in 64 registers);
// train descriptors stream through shared memory 64 at a time, expanded once per CTA into
// fragment order (conflict-free 128-bit loads).  Each thread keeps the top-2 of its 4 rows as
// signed keys ((popc(b) - 2 dot) << 16 | train index): lexicographic (distance, index) order
// is exactly the ascending strict-'<' scan of K9, and popc(a) is a per-row constant added at
// the end.  The four threads of a quad merge by shuffle.
#define KT_Q 256
#define KT_T 64
#define KT_INVALID 0x7fff0000

__device__ __forceinline__ uint32_t expand_nibble(uint32_t n) { return (n * 0x00204081u) & 0x01010101u; }

__device__ __forceinline__ void imma_u8(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// (k1 <= k2) <- two smallest of {k1, k2, x, y}; all keys distinct
__device__ __forceinline__ void top2_merge(int& k1, int& k2, int x, int y)
{
    const int lo = min(x, y), hi = max(x, y);
    const int n2 = __vimin3_s32(max(k1, lo), k2, hi);
    k1 = min(k1, lo);
    k2 = n2;
}

__global__ void __launch_bounds__(256, 2)
vs_knn2_mma_kernel(const uint8_t* __restrict__ query, int nq_max, long q_stride,
                   const uint8_t* __restrict__ train, int nt_max, long t_stride,
                   const int32_t* __restrict__ d_nq, const int32_t* __restrict__ d_nt,
                   int nsplit, int4* __restrict__ partial, int32_t* __restrict__ out_idx,
                   int32_t* __restrict__ out_dist)
{
    __shared__ __align__(16) uint4 s_b[2][(KT_T / 8) * 4 * 32];   // [n-tile][k-step pair][lane]
    __shared__ __align__(8) int32_t s_c[2][KT_T];                 // (popc(b) << 16) | index, or KT_INVALID
    __shared__ int32_t s_pa[KT_Q];                                // popc(a) of the CTA's query rows
    const int pair = blockIdx.z, split = blockIdx.y, tid = threadIdx.x;
    const int lane = tid & 31, wid = tid >> 5, g = lane >> 2, t = lane & 3;
    const int nq = d_nq ? min(d_nq[pair], nq_max) : nq_max;
    const int nt = d_nt ? min(d_nt[pair], nt_max) : nt_max;
    const int q0 = blockIdx.x * KT_Q;
    if (q0 >= nq) {
        const int q = q0 + tid;
        if (q < nq_max && split == 0) {
            size_t o = ((size_t)pair * nq_max + q) * 2;
            out_idx[o] = -1; out_idx[o + 1] = -1; out_dist[o] = -1; out_dist[o + 1] = -1;
        }
        return;
    }
    // A fragments of this thread's 4 rows (m-tile m, half h: row q0 + wid*32 + m*16 + g + 8h)
    uint32_t a[2][8][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int row = min(q0 + wid * 32 + m * 16 + g + 8 * h, nq - 1);
            const uint4* qp = (const uint4*)(query + ((size_t)pair * q_stride + row) * 32);
            const uint4 x = __ldg(qp), y = __ldg(qp + 1);
            const uint32_t w[8] = {x.x, x.y, x.z, x.w, y.x, y.y, y.z, y.w};
            int pc = 0;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                pc += __popc(w[s]);
                const uint32_t byte = (w[s] >> (8 * t)) & 0xffu;
                a[m][s][h] = expand_nibble(byte & 15u);
                a[m][s][2 + h] = expand_nibble(byte >> 4);
            }
            if (t == 0) s_pa[wid * 32 + m * 16 + g + 8 * h] = pc;
        }
    int k1[4], k2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { k1[i] = 0x7fffffff; k2[i] = 0x7fffffff; }

    const uint8_t* tp = train + (size_t)pair * t_stride * 32;
    const int tiles = (nt + KT_T - 1) / KT_T;
    const int t_begin = (int)((long)tiles * split / nsplit), t_end = (int)((long)tiles * (split + 1) / nsplit);

    uint2 raw[4];
    auto fetch = [&](int tile) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int combo = wid + 8 * r;                       // (n-tile, k-step pair)
            const int d = tile * KT_T + (combo >> 2) * 8 + g;
            raw[r] = d < nt ? __ldg((const uint2*)(tp + (size_t)d * 32) + (combo & 3)) : make_uint2(0u, 0u);
        }
    };
    auto stash = [&](int tile, int buf) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int combo = wid + 8 * r;
            const uint32_t b0 = (raw[r].x >> (8 * t)) & 0xffu, b1 = (raw[r].y >> (8 * t)) & 0xffu;
            s_b[buf][combo * 32 + lane] = make_uint4(expand_nibble(b0 & 15u), expand_nibble(b0 >> 4),
                                                      expand_nibble(b1 & 15u), expand_nibble(b1 >> 4));
        }
        if (tid < KT_T) {   // two warps: popc(b) and the key constant of each train descriptor of the tile
            const int d = tile * KT_T + tid;
            uint4 cx = make_uint4(0u, 0u, 0u, 0u), cy = cx;
            if (d < nt) { cx = __ldg((const uint4*)(tp + (size_t)d * 32)); cy = __ldg((const uint4*)(tp + (size_t)d * 32) + 1); }
            const int pb = __popc(cx.x) + __popc(cx.y) + __popc(cx.z) + __popc(cx.w) +
                           __popc(cy.x) + __popc(cy.y) + __popc(cy.z) + __popc(cy.w);
            s_c[buf][tid] = d < nt ? ((pb << 16) | d) : (KT_INVALID | (d & 0xffff));
        }
    };

    if (t_begin < t_end) { fetch(t_begin); stash(t_begin, 0); }
    __syncthreads();
    for (int tile = t_begin; tile < t_end; ++tile) {
        const int buf = (tile - t_begin) & 1;
        const bool more = tile + 1 < t_end;
        if (more) fetch(tile + 1);
#pragma unroll 1
        for (int j = 0; j < KT_T / 8; ++j) {
            const uint4* bp = s_b[buf] + j * 128 + lane;
            const uint4 B0 = bp[0], B1 = bp[32], B2 = bp[64], B3 = bp[96];
            const int2 cc = *(const int2*)&s_c[buf][j * 8 + 2 * t];
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                int acc[4] = {0, 0, 0, 0};
                imma_u8(acc, a[m][0], B0.x, B0.y); imma_u8(acc, a[m][1], B0.z, B0.w);
                imma_u8(acc, a[m][2], B1.x, B1.y); imma_u8(acc, a[m][3], B1.z, B1.w);
                imma_u8(acc, a[m][4], B2.x, B2.y); imma_u8(acc, a[m][5], B2.z, B2.w);
                imma_u8(acc, a[m][6], B3.x, B3.y); imma_u8(acc, a[m][7], B3.z, B3.w);
                top2_merge(k1[2 * m], k2[2 * m], acc[0] * -131072 + cc.x, acc[1] * -131072 + cc.y);
                top2_merge(k1[2 * m + 1], k2[2 * m + 1], acc[2] * -131072 + cc.x, acc[3] * -131072 + cc.y);
            }
        }
        if (more) stash(tile + 1, buf ^ 1);
        __syncthreads();
    }
    // merge the quad's four partial top-2 lists, then thread t == 0 writes its 4 rows
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            const int o1 = __shfl_xor_sync(0xffffffffu, k1[i], o), o2 = __shfl_xor_sync(0xffffffffu, k2[i], o);
            const int n2 = __vimin3_s32(max(k1[i], o1), k2[i], o2);
            k1[i] = min(k1[i], o1);
            k2[i] = n2;
        }
    }
    if (t == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int lrow = wid * 32 + (i >> 1) * 16 + g + 8 * (i & 1), row = q0 + lrow;
            if (row >= nq_max) continue;
            const int pa = s_pa[lrow];
            const bool v1 = k1[i] < KT_INVALID, v2 = k2[i] < KT_INVALID;
            int d1 = v1 ? (k1[i] >> 16) + pa : -1, i1 = v1 ? (k1[i] & 0xffff) : -1;
            int d2 = v2 ? (k2[i] >> 16) + pa : -1, i2 = v2 ? (k2[i] & 0xffff) : -1;
            if (row >= nq) {
                if (split != 0) continue;
                d1 = d2 = i1 = i2 = -1;
            }
            if (nsplit == 1 || row >= nq) {
                size_t o = ((size_t)pair * nq_max + row) * 2;
                out_idx[o] = i1; out_idx[o + 1] = i2; out_dist[o] = d1; out_dist[o + 1] = d2;
            } else {
                partial[((size_t)pair * nsplit + split) * nq_max + row] = make_int4(d1, i1, d2, i2);
            }
        }
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
                       int32_t* d_idx, int32_t* d_dist, cudaStream_t s);

// kNN-2 engine for nt <= 65536: 1 = tcgen05 kind::i8 (bf_match_tc5.cu), 0 = mma.sync int8 (above).
// VSLAM_KNN=imma|tc5 overrides the default; VSLAM_KNN=popc forces the XOR + POPC kernel (the formulation
// BASELINE.json's north_star names) at every size, for the popc-pipe roofline of tools/knn_sweep.py.
static int knn_force_popc()
{
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("VSLAM_KNN");
        mode = (e && !strcmp(e, "popc")) ? 1 : 0;
    }
    return mode;
}

static int knn_use_tc5()
{
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("VSLAM_KNN");
        mode = VS_KNN_DEFAULT_TC5;
        if (e && !strcmp(e, "imma")) mode = 0;
        if (e && !strcmp(e, "tc5")) mode = 1;
    }
    return mode;
}

// scratch of the split-train path, shared by every caller of the process: launches that use it are serialised
static int4* g_partial = nullptr;
static size_t g_partial_bytes = 0;
static std::mutex g_partial_mu;
static cudaEvent_t g_partial_ev = nullptr;     // end of the last launch pair that used the scratch
static cudaStream_t g_partial_last = nullptr;

int vs_launch_knn2(vslam_orb* prof, const uint8_t* d_q, int nq, long q_stride, const uint8_t* d_t, int nt,
                   long t_stride, const int32_t* d_nq, const int32_t* d_nt, int pairs,
                   int32_t* d_idx, int32_t* d_dist, cudaStream_t s)
{
    if (pairs <= 0 || nq <= 0) return VSLAM_OK;
    const bool mma = nt <= 65536 && !knn_force_popc();  // the tensor-core kernels pack the train index into 16 bits
    const int qtiles = mma ? (nq + KT_Q - 1) / KT_Q : (nq + BF_THREADS - 1) / BF_THREADS;
    const bool tc5 = mma && knn_use_tc5();
    const int ttiles = tc5 ? (nt + 127) / 128 : mma ? (nt + KT_T - 1) / KT_T : (nt + BF_TILE - 1) / BF_TILE;
    // split the train range only when the grid would not fill the 148 SMs a few times over
    int nsplit = 1;
    const long want = 148L * 8;
    if ((long)qtiles * pairs < want && ttiles > 1 && !(prof && prof->knn_no_split)) {
        nsplit = (int)((want + (long)qtiles * pairs - 1) / ((long)qtiles * pairs));
        if (nsplit > ttiles) nsplit = ttiles;
        if (nsplit > 64) nsplit = 64;
    }
    std::unique_lock<std::mutex> lock(g_partial_mu, std::defer_lock);
    if (nsplit > 1) {
        lock.lock();
        size_t need = (size_t)pairs * nsplit * nq * sizeof(int4);
        if (need > g_partial_bytes) {
            if (g_partial) { VS_CUDA(cudaStreamSynchronize(s)); VS_CUDA(cudaFree(g_partial)); g_partial = nullptr; }
            VS_CUDA(cudaMalloc(&g_partial, need));
            g_partial_bytes = need;
        }
        // another stream may still be reading / writing the scratch: order this launch behind it
        if (g_partial_ev && g_partial_last != s) VS_CUDA(cudaStreamWaitEvent(s, g_partial_ev, 0));
    }
    vs_prof_mark(prof, VS_PROF_KNN2, s);
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
    if (tc5) {
        int rc = vs_knn2_tc5_launch(d_q, nq, q_stride, d_t, nt, t_stride, d_nq, d_nt, pairs, nsplit, g_partial, d_idx, d_dist, s);
        if (rc) return rc;
    } else if (mma)
        vs_knn2_mma_kernel<<<dim3(qtiles, nsplit, pairs), 256, 0, s>>>(
            d_q, nq, q_stride, d_t, nt, t_stride, d_nq, d_nt, nsplit, g_partial, d_idx, d_dist);
    else
        vs_knn2_kernel<<<dim3(qtiles, nsplit, pairs), BF_THREADS, 0, s>>>(
            d_q, nq, q_stride, d_t, nt, t_stride, d_nq, d_nt, nsplit, g_partial, d_idx, d_dist);
    if (nsplit > 1) {
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_knn2_merge_kernel<<<dim3((nq + 255) / 256, pairs), 256, 0, s>>>(g_partial, nq, d_nq, nsplit,
                                                                          d_idx, d_dist);
        if (!g_partial_ev) VS_CUDA(cudaEventCreateWithFlags(&g_partial_ev, cudaEventDisableTiming));
        VS_CUDA(cudaEventRecord(g_partial_ev, s));
        g_partial_last = s;
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
