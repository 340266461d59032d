// bf_match_tc5.cu -- brute-force Hamming kNN-2 on the 5th-generation tensor cores (tcgen05, kind::i8).
//
// Same contract as vs_knn2_kernel / vs_knn2_mma_kernel in bf_match.cu (cv2.BFMatcher(NORM_HAMMING).knnMatch
// k=2, matcher.py:70), same arithmetic as the mma.sync variant: descriptor bits are
// expanded to 0/1 bytes, dot(a, b) = popc(a & b), hamming = popc(a) + popc(b) - 2 dot.  Top-2 per query: inside
// a tile the 64 columns a warp reads back become 16-bit keys ((popc(b) - 2 dot + 256) << 6 | column), two per
// register, reduced with packed 16x2 min/max (1.25 ALU operations per distance instead of 2.5); the four
// survivors of a tile are widened to ((popc(b) - 2 dot + 256) << 16 | train index) and merged into the
// running pair.  Lexicographic (distance, index) order = cv2's ascending strict-< scan; popc(a) is added at the end.
//
// One CTA = 256 query rows = two M=128 UMMA tiles; train descriptors stream in tiles of N=128.
//   * operands live in shared memory in the canonical K-major no-swizzle UMMA layout (8-row x 16-byte
//     core matrices; LBO = 128 B between the K halves, SBO = 2048 B between 8-row groups); the bit
//     expansion writes that layout directly (the k order of a dot product is free, so k = bit index),
//   * a train set is expanded ONCE (vs_knn2_expand_kernel: operand tile + 16-bit key constants, 33 024 B
//     per 128 descriptors, L2 resident for the CTAs of the same pair) instead of by every CTA that reads it,
//   * a loader warp streams those tiles into a 3-stage shared-memory ring with bulk async copies (TMA,
//     mbarrier complete_tx),
//   * two MMA warps take turns (even / odd tiles = TMEM buffer 0 / 1): 2 x 8 tcgen05.mma (M128 N128 K32,
//     u8 x u8 -> s32) per tile into a 256-column TMEM buffer, one tcgen05.commit per M-tile.  The warp walks
//     its loop with warp-uniform values so that the descriptors are built on the uniform datapath (1-3
//     instructions between MMAs; a lane-0 branch costs ~16 with a waterfall loop per MMA and made the issue
//     thread, not the tensor pipe, the bottleneck -- tools/knn_trace.py shows the timeline),
//   * 16 consumer warps read finished tiles back with tcgen05.ld 32x32b.pack::16b (thread = TMEM lane = query
//     row, so the top-2 state is private to a thread; two dot products per register) and release the TMEM
//     buffer / ring stage through mbarriers; there is no CTA-wide barrier in the tile loop.
#include "common.cuh"

#define T5_ROWS 256          // query rows per CTA
#define T5_CONSUMERS 512     // 16 warps: warps w and w + 8 read the same TMEM lanes, one column half each
#define T5_THREADS 608       // + MMA warp 0 (16), the loader warp (17), MMA warp 1 (18)
#define T5_STAGES 3
#define T5_N 128             // train descriptors per tile
#define T5_OPER_BYTES 32768  // one 128-row operand tile: 16 groups x 2048 B
#define T5_TILE_BYTES (T5_OPER_BYTES + 256)   // + the tile's 64 words of 16-bit key constants
#define T5_INVALID (1023 << 16)   // widened key of a padding column (16-bit key 0xffff)

__device__ __forceinline__ uint32_t t5_expand(uint32_t n) { return (n * 0x00204081u) & 0x01010101u; }

// 32 descriptor bits -> 32 bytes of 0/1 = the two 16-byte core-matrix rows of one K=32 step
__device__ __forceinline__ void t5_expand_word(uint32_t w, uint4& lo, uint4& hi)
{
    lo = make_uint4(t5_expand(w & 15u), t5_expand((w >> 4) & 15u), t5_expand((w >> 8) & 15u), t5_expand((w >> 12) & 15u));
    hi = make_uint4(t5_expand((w >> 16) & 15u), t5_expand((w >> 20) & 15u), t5_expand((w >> 24) & 15u), t5_expand(w >> 28));
}

// (row, half) of a 128-row operand tile: expand words 4*half .. 4*half+3 of the descriptor
__device__ __forceinline__ void t5_store_half(uint8_t* tile, int row, int half, const uint4& d)
{
    uint8_t* base = tile + (row >> 3) * 2048 + (row & 7) * 16 + half * 1024;   // k-step s = 4*half + j at + j*256
    const uint32_t w[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint4 lo, hi;
        t5_expand_word(w[j], lo, hi);
        *(uint4*)(base + j * 256) = lo;
        *(uint4*)(base + j * 256 + 128) = hi;
    }
}

// (row, quarter) of a 128-row operand tile: expand words 2*quarter, 2*quarter+1 of the descriptor
__device__ __forceinline__ void t5_store_quarter(uint8_t* tile, int row, int quarter, const uint2& d)
{
    uint8_t* base = tile + (row >> 3) * 2048 + (row & 7) * 16 + quarter * 512;
    uint4 lo, hi;
    t5_expand_word(d.x, lo, hi);
    *(uint4*)base = lo; *(uint4*)(base + 128) = hi;
    t5_expand_word(d.y, lo, hi);
    *(uint4*)(base + 256) = lo; *(uint4*)(base + 384) = hi;
}

__device__ __forceinline__ uint64_t t5_smem_desc(uint32_t saddr)
{
    // start address, LBO = 128 B, SBO = 2048 B (all >> 4), descriptor version 1 (sm_100), no swizzle
    return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(128u >> 4) << 16) | ((uint64_t)(2048u >> 4) << 32) | (1ull << 46);
}

__device__ __forceinline__ void t5_top2(int& k1, int& k2, int x, int y)
{
    const int lo = min(x, y), hi = max(x, y);
    const int n2 = __vimin3_s32(max(k1, lo), k2, hi);
    k1 = min(k1, lo);
    k2 = n2;
}

// the same on two independent 16-bit lanes
__device__ __forceinline__ void t5_top2_u16x2(uint32_t& k1, uint32_t& k2, uint32_t x, uint32_t y)
{
    const uint32_t lo = __vminu2(x, y), hi = __vmaxu2(x, y);
    const uint32_t n2 = __vimin3_u16x2(__vmaxu2(k1, lo), k2, hi);
    k1 = __vminu2(k1, lo);
    k2 = n2;
}

// 32 consecutive accumulator columns of this thread's TMEM lane as 16 registers: .pack::16b keeps the low 16 bits
// of each column (a dot product is <= 256) and packs columns 2j / 2j+1 into the low / high half of register j --
// already the 16x2 lane form of the keys, at half the TMEM read time of a plain load (tools/ubench/ldtm.cu; TMEM
// reads and the MMAs' accumulator traffic share the TMEM port, so read time is MMA time).  Asynchronous until
// tcgen05.wait::ld.
__device__ __forceinline__ void t5_tmem_ld32p(uint32_t (&v)[16], uint32_t taddr)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}

// fold 32 dot products (16 packed registers) into the packed top-2 pair: c2[j] holds the 16-bit key constants of
// columns 2j (low lane) and 2j+1 (high lane); key = constant - 128 dot never borrows (it is >= 0 per lane), so one
// 32-bit multiply-add forms both lanes
__device__ __forceinline__ void t5_reduce32(const uint32_t (&v)[16], const uint32_t* c2, uint32_t& p1, uint32_t& p2)
{
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
        const uint4 cc = *(const uint4*)(c2 + j);
        const uint32_t a = v[j] * 0xffffff80u + cc.x;
        const uint32_t b = v[j + 1] * 0xffffff80u + cc.y;
        const uint32_t c = v[j + 2] * 0xffffff80u + cc.z;
        const uint32_t d = v[j + 3] * 0xffffff80u + cc.w;
        t5_top2_u16x2(p1, p2, a, b);
        t5_top2_u16x2(p1, p2, c, d);
    }
}

// 16-bit tile key -> ((distance part) << 16 | train index)
__device__ __forceinline__ int t5_widen(uint32_t x, int base) { return (int)((x & 0xffc0u) * 1024u + (x & 63u)) + base; }

__device__ __forceinline__ void t5_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n .reg .pred p;\n"
        "W%=: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra D%=;\n bra W%=;\n"
        "D%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}

#ifdef T5_TRACE
// timeline of CTA (0, 0, 0), clock64 per tile i: [0] MMAs issued + commit, [1] consumer warp 0 sees the
// accumulators, [2] consumer warp 0 done, [3] loader issued the tile, [4] MMA warp saw the free TMEM buffer,
// [5] MMA warp about to issue, [6] last consumer warp done (tools/knn_trace.py; build with -DT5_TRACE)
__device__ long long t5_trace_buf[8 * 64];
extern "C" int vslam_debug_t5_trace(long long* out) { return (int)cudaMemcpyFromSymbol(out, t5_trace_buf, sizeof(t5_trace_buf)); }
#define T5_STAMP(row, i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (i) < 64) t5_trace_buf[(row) * 64 + (i)] = clock64(); } while (0)
#else
#define T5_STAMP(row, i) do { } while (0)
#endif

// ---- train descriptors -> UMMA operand tiles + key constants (once per train set) ------------------
// thread = (train row, quarter) with consecutive lanes on consecutive rows; quarter 0 also counts the row's bits
__global__ void __launch_bounds__(512)
vs_knn2_expand_kernel(const uint8_t* __restrict__ train, int nt_max, long t_stride, const int32_t* __restrict__ d_nt,
                      int tiles_max, uint8_t* __restrict__ expanded)
{
    const int tile = blockIdx.x, pair = blockIdx.y, tid = threadIdx.x;
    const int nt = d_nt ? min(d_nt[pair], nt_max) : nt_max;
    if (tile * T5_N >= nt) return;                       // tiles past the end are never read
    const uint8_t* tp = train + (size_t)pair * t_stride * 32;
    uint8_t* out = expanded + ((size_t)pair * tiles_max + tile) * T5_TILE_BYTES;
    const int row = tid & 127, quarter = tid >> 7, d_idx = tile * T5_N + row;
    uint2 d = make_uint2(0u, 0u);
    if (d_idx < nt) d = __ldg((const uint2*)(tp + (size_t)d_idx * 32) + quarter);
    t5_store_quarter(out, row, quarter, d);
    if (quarter == 0) {
        int pb = 0;
        if (d_idx < nt) {
            const uint4 x = __ldg((const uint4*)(tp + (size_t)d_idx * 32)), y = __ldg((const uint4*)(tp + (size_t)d_idx * 32) + 1);
            pb = __popc(x.x) + __popc(x.y) + __popc(x.z) + __popc(x.w) + __popc(y.x) + __popc(y.y) + __popc(y.z) + __popc(y.w);
        }
        ((uint16_t*)(out + T5_OPER_BYTES))[row] = d_idx < nt ? (uint16_t)(((pb + 256) << 6) | (row & 63)) : (uint16_t)0xffffu;
    }
}

// Work item w of the search kernel = (pair, 256-row query tile, train split), pair-major.  The kernel is
// PERSISTENT: one CTA per SM walks the items w = blockIdx.x, + gridDim.x, ... with its TMEM allocation, barriers and
// role warps alive across items, and the A operand (two pre-expanded 128-row tiles of the query set, the same
// format as the train tiles) arrives by bulk copy like the train tiles do.  Between two items the tensor pipe only
// waits for that copy (the loader issues it the moment the last MMA of the previous item has completed, while the
// consumer warps are still reducing its last tiles); the non-persistent form paid CTA launch, TMEM allocation,
// barrier set-up, the query expansion by the consumer warps and the pipeline fill per item, with nothing of the
// same CTA to overlap them with (one CTA per SM): 39 % of its time.
struct T5Item {
    int pair, split, q0, nq, ntile, t_begin;
    bool live, has_a;
};

__device__ __forceinline__ T5Item t5_item(int w, int qtiles, int nsplit, int nq_max, int nt_max,
                                          const int32_t* __restrict__ d_nq, const int32_t* __restrict__ d_nt)
{
    T5Item it;
    it.pair = w / (qtiles * nsplit);
    const int rem = w - it.pair * (qtiles * nsplit);
    const int qt = rem / nsplit;
    it.split = rem - qt * nsplit;
    it.q0 = qt * T5_ROWS;
    it.nq = d_nq ? min(__ldg(d_nq + it.pair), nq_max) : nq_max;
    const int nt = d_nt ? min(__ldg(d_nt + it.pair), nt_max) : nt_max;
    it.live = it.q0 < it.nq;
    const int tiles = (nt + T5_N - 1) / T5_N;
    it.t_begin = (int)((long)tiles * it.split / nsplit);
    it.ntile = it.live ? (int)((long)tiles * (it.split + 1) / nsplit) - it.t_begin : 0;
    it.has_a = it.ntile > 0;
    return it;
}

__global__ void __launch_bounds__(T5_THREADS, 1)
vs_knn2_tc5_kernel(const uint8_t* __restrict__ query, int nq_max, long q_stride,
                   const uint8_t* __restrict__ exp_q, int qtiles_max,
                   const uint8_t* __restrict__ exp_t, int ttiles_max, int nt_max,
                   const int32_t* __restrict__ d_nq, const int32_t* __restrict__ d_nt,
                   int nsplit, int items, int4* __restrict__ partial, int32_t* __restrict__ out_idx,
                   int32_t* __restrict__ out_dist)
{
    extern __shared__ __align__(128) uint8_t t5_smem[];
    // full[s]: train tile landed in ring stage s; empty[s]: all 16 consumer warps are through with the tile in stage
    // s; done[2b + m]: the MMAs of M-tile m into TMEM buffer b finished; free[2b + m]: the 8 consumer warps of that
    // M-tile are through with it (hand-off per M-tile: four units in flight instead of two); turn[w]: MMA warp w
    // may issue (the other warp's tile is in the pipe); afull: the item's A operand landed; afree: both MMA warps
    // have their last MMAs of the item behind them (the A operand may be overwritten)
    __shared__ __align__(8) unsigned long long s_full[T5_STAGES], s_empty[T5_STAGES], s_done[4], s_free[4], s_turn[2],
        s_afull, s_afree;
    __shared__ uint32_t s_tmem;
    __shared__ int2 s_merge[2][T5_ROWS];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const bool consumer = wid < T5_CONSUMERS / 32;
    const int mtile = (wid >> 2) & 1, chalf = (wid >> 3) & 1;           // M-tile and column half this warp reads back
    const int lrow = mtile * 128 + (wid & 3) * 32 + lane;                // TMEM lane (wid & 3) * 32 + lane of that M-tile
    const int qtiles = (nq_max + T5_ROWS - 1) / T5_ROWS;
    uint8_t* sm = t5_smem + ((128u - ((uint32_t)__cvta_generic_to_shared(t5_smem) & 127u)) & 127u);
    uint8_t* sA = sm;                         // 2 tiles: query rows 0..127, 128..255 of the item (operand + key constants)
    uint8_t* sT = sm + 2 * T5_TILE_BYTES;     // T5_STAGES x (operand tile + key constants)
    const uint32_t sA_u = (uint32_t)__cvta_generic_to_shared(sA), sT_u = (uint32_t)__cvta_generic_to_shared(sT);
    const uint32_t full_u = (uint32_t)__cvta_generic_to_shared(&s_full[0]);
    const uint32_t done_u = (uint32_t)__cvta_generic_to_shared(&s_done[0]);
    const uint32_t free_u = (uint32_t)__cvta_generic_to_shared(&s_free[0]);
    const uint32_t empty_u = (uint32_t)__cvta_generic_to_shared(&s_empty[0]);
    const uint32_t turn_u = (uint32_t)__cvta_generic_to_shared(&s_turn[0]);
    const uint32_t afull_u = (uint32_t)__cvta_generic_to_shared(&s_afull);
    const uint32_t afree_u = (uint32_t)__cvta_generic_to_shared(&s_afree);

    // ---- once per CTA: TMEM (all 512 columns: two buffers of 2 x 128), barriers
    if (wid == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (wid == T5_CONSUMERS / 32 + 1 && lane == 0) {
        for (int k = 0; k < T5_STAGES; ++k) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(full_u + 8 * k));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty_u + 8 * k), "r"(T5_CONSUMERS / 32));
        }
        for (int k = 0; k < 4; ++k) {      // per (TMEM buffer, M-tile): 8 consumer warps read one M-tile
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(done_u + 8 * k));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(free_u + 8 * k), "r"(T5_CONSUMERS / 64));
        }
        for (int k = 0; k < 2; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(turn_u + 8 * k));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(afull_u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 2;" ::"r"(afree_u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    // every role walks the same item sequence and keeps the same running counters:
    //   T = train tiles of this CTA so far (ring stage T % STAGES, TMEM buffer / MMA warp T & 1), na = A loads so far
    int T = 0, na = 0;
    if (wid == T5_CONSUMERS / 32 + 1) {
        // ---- loader (one thread): the item's A tiles, then its train tiles into the ring; stage s is reusable once
        // all 16 consumer warps are through with the tile that was in it (which also means its MMAs are complete)
        if (lane == 0) {
            for (int w = blockIdx.x; w < items; w += gridDim.x) {
                const T5Item it = t5_item(w, qtiles, nsplit, nq_max, nt_max, d_nq, d_nt);
                if (it.has_a) {
                    if (na > 0) t5_wait(afree_u, (uint32_t)(na - 1) & 1u);      // the previous item's MMAs are done with A
                    const int qt128 = it.q0 / T5_N;
                    const int atiles = min(2, (it.nq + T5_N - 1) / T5_N - qt128);   // tiles past the end were never written
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(afull_u), "r"(atiles * T5_TILE_BYTES) : "memory");
                    for (int m = 0; m < atiles; ++m)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(sA_u + m * T5_TILE_BYTES),
                                       "l"(exp_q + ((size_t)it.pair * qtiles_max + qt128 + m) * T5_TILE_BYTES),
                                       "r"(T5_TILE_BYTES), "r"(afull_u) : "memory");
                    ++na;
                }
                const uint8_t* tile_src = exp_t + ((size_t)it.pair * ttiles_max + it.t_begin) * T5_TILE_BYTES;
                for (int i = 0; i < it.ntile; ++i) {
                    const int g = T + i, stage = g % T5_STAGES;
                    if (g >= T5_STAGES) t5_wait(empty_u + 8 * stage, (uint32_t)(g / T5_STAGES - 1) & 1u);
                    const uint32_t bar = full_u + 8 * stage, dst = sT_u + stage * T5_TILE_BYTES;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(T5_TILE_BYTES) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(dst), "l"(tile_src + (size_t)i * T5_TILE_BYTES), "r"(T5_TILE_BYTES), "r"(bar) : "memory");
                    T5_STAMP(3, g);
                }
                T += it.ntile;
            }
        }
    } else if (!consumer) {
        // ---- MMA issue: one thread in each of two warps, even / odd tiles (= TMEM buffer 0 / 1).  The tensor pipe
        // takes a few MMAs ahead and then blocks the issuing thread, so a single issuer reaches the barrier waits of
        // the next tile only when the pipe is already draining; with two, the other thread has those waits behind
        // it and only needs its turn (tiles must enter the pipe in order: the consumers want tile g first).
        // The whole warp walks the loop with warp-uniform values (REDUX results, constants), so the descriptors are
        // formed on the uniform datapath; one elected lane issues.
        const int me = (int)__reduce_min_sync(0xffffffffu, (unsigned)(wid == T5_CONSUMERS / 32 ? 0 : 1));
        const uint32_t tmem_u = __reduce_min_sync(0xffffffffu, tmem);
        const uint32_t idesc = (2u << 4) | ((uint32_t)(T5_N >> 3) << 17) | ((128u >> 4) << 24);   // s32 accum, u8 x u8, K-major both
        uint32_t elected;
        asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(elected));
        for (int w = blockIdx.x; w < items; w += gridDim.x) {
            const T5Item it = t5_item(w, qtiles, nsplit, nq_max, nt_max, d_nq, d_nt);
            const int ntile_u = (int)__reduce_min_sync(0xffffffffu, (unsigned)it.ntile);
            const int T_u = (int)__reduce_min_sync(0xffffffffu, (unsigned)T);
            if (ntile_u > 0) {
                bool issued = false;
                for (int g = T_u + ((T_u ^ me) & 1); g < T_u + ntile_u; g += 2) {
                    const int buf = me, stage = g % T5_STAGES;
                    t5_wait(full_u + 8 * stage, (uint32_t)(g / T5_STAGES) & 1u);
                    if (g >= 1) t5_wait(turn_u + 8 * me, (uint32_t)(((g + 1) >> 1) - 1) & 1u);   // tile g - 1 is in the pipe
                    if (!issued) t5_wait(afull_u, (uint32_t)na & 1u);                             // the item's A operand
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        // the 8 warps that read M-tile m of tile g - 2 are through with this half of the TMEM buffer
                        if (g >= 2) t5_wait(free_u + 8 * (2 * buf + m), (uint32_t)((g - 2) >> 1) & 1u);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        if (m == 0) T5_STAMP(5, g);
                        if (elected) {
                            const uint32_t tacc = tmem_u + buf * 256 + m * 128;
#pragma unroll
                            for (int s = 0; s < 8; ++s) {
                                const uint64_t da = t5_smem_desc(sA_u + m * T5_TILE_BYTES + s * 256);
                                const uint64_t db = t5_smem_desc(sT_u + stage * T5_TILE_BYTES + s * 256);
                                const uint32_t accumulate = s > 0;
                                asm volatile(
                                    "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                                    " tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n}\n"
                                    ::"r"(tacc), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
                            }
                            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                                         ::"r"(done_u + 8 * (2 * buf + m)) : "memory");
                            if (m == 1) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(turn_u + 8 * (me ^ 1)) : "memory");
                        }
                        __syncwarp();
                    }
                    issued = true;
                    T5_STAMP(0, g);
                }
                // this warp's share of "the item's MMAs are complete": a commit behind its last MMAs, or a plain
                // arrival when none of the item's tiles was its
                if (elected) {
                    if (issued)
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(afree_u) : "memory");
                    else
                        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(afree_u) : "memory");
                }
                __syncwarp();
                ++na;
            }
            T += it.ntile;
        }
    } else {
        // ---- consumers: finished tiles -> packed top-2 -> running pair -> the item's output rows
        int n_items = 0;
        for (int w = blockIdx.x; w < items; w += gridDim.x, ++n_items) {
            const T5Item it = t5_item(w, qtiles, nsplit, nq_max, nt_max, d_nq, d_nt);
            const int myrow = it.q0 + lrow;
            if (!it.live) {       // a query tile past the pair's rows: its output rows read "no neighbour"
                const int q = it.q0 + tid;
                if (tid < T5_ROWS && q < nq_max && it.split == 0) {
                    const size_t o = ((size_t)it.pair * nq_max + q) * 2;
                    out_idx[o] = -1; out_idx[o + 1] = -1; out_dist[o] = -1; out_dist[o + 1] = -1;
                }
                continue;
            }
            // popc(a) of this thread's row (used at the end of the item; the loads fly behind the tile loop)
            const uint4* qp = (const uint4*)(query + ((size_t)it.pair * q_stride + min(myrow, it.nq - 1)) * 32);
            const uint4 qx = __ldg(qp), qy = __ldg(qp + 1);
            int k1 = 0x7fffffff, k2 = 0x7fffffff;       // running pair of widened keys
            for (int i = 0; i < it.ntile; ++i) {
                const int g = T + i, buf = g & 1, stage = g % T5_STAGES;
                t5_wait(full_u + 8 * stage, (uint32_t)(g / T5_STAGES) & 1u);     // the tile's key constants (acquire)
                t5_wait(done_u + 8 * (2 * buf + mtile), (uint32_t)(g >> 1) & 1u);   // its accumulators (this warp's M-tile)
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (tid == 0) T5_STAMP(1, g);
                const uint32_t taddr = tmem + (((uint32_t)(wid & 3) * 32u) << 16) + buf * 256 + mtile * 128 + chalf * 64;
                const uint32_t* cbase = (const uint32_t*)(sT + stage * T5_TILE_BYTES + T5_OPER_BYTES) + chalf * 32;
                // this warp's 64 columns as two 32-column chunks, the second load in flight while the first is reduced
                uint32_t va[16], vb[16];
                uint32_t p1 = 0xffffffffu, p2 = 0xffffffffu;
                t5_tmem_ld32p(va, taddr);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                t5_tmem_ld32p(vb, taddr + 32);
                t5_reduce32(va, cbase, p1, p2);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                t5_reduce32(vb, cbase + 16, p1, p2);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(free_u + 8 * (2 * buf + mtile)) : "memory");
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty_u + 8 * stage) : "memory");
                }
                // the tile's four survivors (two per lane) join the running pair
                const int base = (it.t_begin + i) * T5_N + chalf * 64;
                t5_top2(k1, k2, t5_widen(p1 & 0xffffu, base), t5_widen(p1 >> 16, base));
                t5_top2(k1, k2, t5_widen(p2 & 0xffffu, base), t5_widen(p2 >> 16, base));
                if (tid == 0) T5_STAMP(2, g);
            }
            T += it.ntile;
            // merge the two column halves (two sorted pairs -> the two smallest of the four keys); the buffer of item
            // n is rewritten by item n + 2, behind the barrier of item n + 1
            int2* mg = s_merge[n_items & 1];
            if (chalf == 1) mg[lrow] = make_int2(k1, k2);
            asm volatile("bar.sync 1, %0;" ::"r"(T5_CONSUMERS) : "memory");
            if (chalf == 1) continue;
            {
                const int2 o = mg[lrow];
                const int n2 = __vimin3_s32(max(k1, o.x), k2, o.y);
                k1 = min(k1, o.x);
                k2 = n2;
            }
            if (myrow < nq_max) {
                const int pa = __popc(qx.x) + __popc(qx.y) + __popc(qx.z) + __popc(qx.w) + __popc(qy.x) + __popc(qy.y) + __popc(qy.z) + __popc(qy.w);
                const bool v1 = k1 < T5_INVALID, v2 = k2 < T5_INVALID;
                int d1 = v1 ? (k1 >> 16) - 256 + pa : -1, i1 = v1 ? (k1 & 0xffff) : -1;
                int d2 = v2 ? (k2 >> 16) - 256 + pa : -1, i2 = v2 ? (k2 & 0xffff) : -1;
                if (myrow >= it.nq) {
                    if (it.split != 0) continue;
                    d1 = d2 = i1 = i2 = -1;
                }
                if (nsplit == 1 || myrow >= it.nq) {
                    const size_t o = ((size_t)it.pair * nq_max + myrow) * 2;
                    out_idx[o] = i1; out_idx[o + 1] = i2; out_dist[o] = d1; out_dist[o + 1] = d2;
                } else {
                    partial[((size_t)it.pair * nsplit + it.split) * nq_max + myrow] = make_int4(d1, i1, d2, i2);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (wid == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

#define T5_DYN_SMEM ((2 + T5_STAGES) * T5_TILE_BYTES + 128)

// function attributes and memory pools belong to a device (context), not to the process: configure each device once
static int g_t5_sms[64] = {};
int vs_knn2_tc5_configure()
{
    static unsigned long long done_mask = 0;       // bit per device ordinal
    int dev = 0;
    VS_CUDA(cudaGetDevice(&dev));
    const unsigned long long bit = dev >= 0 && dev < 64 ? 1ull << dev : 0ull;
    if (!bit || !(__atomic_load_n(&done_mask, __ATOMIC_ACQUIRE) & bit)) {
        VS_CUDA(cudaFuncSetAttribute(vs_knn2_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, T5_DYN_SMEM));
        // the expanded tiles of callers without a handle live in stream-ordered allocations: keep freed blocks in the pool
        cudaMemPool_t pool;
        VS_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
        unsigned long long keep = ~0ull;
        VS_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        int sms = 0;
        VS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        if (bit) g_t5_sms[dev] = sms;
        __atomic_fetch_or(&done_mask, bit, __ATOMIC_RELEASE);
    }
    return VSLAM_OK;
}

// bytes of expanded-tile workspace one descriptor set of n rows needs
size_t vs_knn2_tc5_workspace_per_pair(int nt) { return (size_t)((nt + T5_N - 1) / T5_N) * T5_TILE_BYTES; }

// ws / ws_bytes: caller-owned workspace for the expanded tiles (a handle's, sized at creation); when it is missing
// or too small the tiles live in a stream-ordered allocation of this call.  Both operands are expanded: the query
// sets feed the A operand by bulk copy exactly as the train sets feed B.  When the train block is the query block
// shifted by one set (the front end: pair i = frame i against frame i + 1) every set is expanded once and serves
// as the train set of pair i - 1 and the query set of pair i.
int vs_knn2_tc5_launch(const uint8_t* d_q, int nq, long q_stride, const uint8_t* d_t, int nt, long t_stride,
                       const int32_t* d_nq, const int32_t* d_nt, int pairs, int nsplit, int4* partial,
                       int32_t* d_idx, int32_t* d_dist, uint8_t* ws, size_t ws_bytes, cudaStream_t s)
{
    int rc = vs_knn2_tc5_configure();
    if (rc) return rc;
    const int ttiles = (nt + T5_N - 1) / T5_N, qtiles128 = (nq + T5_N - 1) / T5_N;
    const bool shifted = nt == nq && q_stride == t_stride && d_t == d_q + (size_t)q_stride * 32 &&
                         ((!d_nq && !d_nt) || (d_nq && d_nt == d_nq + 1));
    const size_t q_sets = shifted ? (size_t)pairs + 1 : (size_t)pairs, t_sets = shifted ? 0 : (size_t)pairs;
    const size_t need = (q_sets * qtiles128 + t_sets * ttiles) * T5_TILE_BYTES;
    uint8_t* expanded = nullptr;
    bool own = false;
    if (ws && ws_bytes >= need) expanded = ws;
    else { VS_CUDA(cudaMallocAsync((void**)&expanded, need, s)); own = true; }
    uint8_t* exp_q = expanded;
    uint8_t* exp_t = shifted ? expanded + (size_t)qtiles128 * T5_TILE_BYTES : expanded + q_sets * qtiles128 * T5_TILE_BYTES;
    __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);    // the caller counts the search kernel
    vs_knn2_expand_kernel<<<dim3(qtiles128, (unsigned)q_sets), 512, 0, s>>>(d_q, nq, q_stride, d_nq, qtiles128, exp_q);
    if (!shifted && ttiles > 0) {
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_knn2_expand_kernel<<<dim3(ttiles, pairs), 512, 0, s>>>(d_t, nt, t_stride, d_nt, ttiles, exp_t);
    }
    const int items = ((nq + T5_ROWS - 1) / T5_ROWS) * nsplit * pairs;
    int dev = 0;
    VS_CUDA(cudaGetDevice(&dev));
    const int sms = dev >= 0 && dev < 64 && g_t5_sms[dev] > 0 ? g_t5_sms[dev] : 148;
    vs_knn2_tc5_kernel<<<items < sms ? items : sms, T5_THREADS, T5_DYN_SMEM, s>>>(
        d_q, nq, q_stride, exp_q, qtiles128, exp_t, shifted ? qtiles128 : ttiles, nt, d_nq, d_nt, nsplit, items, partial, d_idx, d_dist);
    if (own) VS_CUDA(cudaFreeAsync(expanded, s));
    return VSLAM_OK;
}
