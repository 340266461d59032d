// orb_octree.cu -- ORB-SLAM2-style keypoint distribution as an option of the extractor.
//
// The reference's ORBExtractor stores `ini_threshold` / `min_threshold` (extractor.py:19-35) and the project it
// is modelled on (ORB-SLAM2, Mur-Artal & Tardos 2017, ORBextractor::ComputeKeyPointsOctTree / DistributeOctTree)
// uses them like this, per pyramid level:
//   * FAST per 30-px cell of the region [EDGE-3, W-EDGE+3) with the initial threshold (20); a cell without a
//     single corner is searched again with the minimum threshold (7);
//   * the corners are spread by a quadtree: every node holding more than one corner is split into four until the
//     level's feature quota of nodes is reached (the last round splits the fullest nodes first and stops at the
//     quota), then each node keeps its corner with the highest FAST score;
//   * response = FAST score (no Harris ranking).
// The reference itself never implements this (cv2.ORB does its own Harris retain-best), so there is nothing in
// /root/reference to pin it against: PARITY UNPINNED.  The oracle is oracle/orbslam2_oracle.py, a sequential
// node-list restatement of the published algorithm; this kernel must equal it exactly.  Deviations from
// ORB-SLAM2 kept on purpose so that the rest of the path (pyramid, blur, orientation, rBRIEF) is shared with the
// cv2.ORB mode: the 31-px edge of cv2.ORB instead of 19, FAST non-max suppression over the whole level instead
// of per cell, and a deterministic tie order (fullest node first, then quadtree order) where ORB-SLAM2 sorts by
// pointer value.
//
// One CTA per (level, frame).  A corner's node at any depth is a pure function of its coordinates (the child
// boundaries are ceil-halvings of the root rectangle), so the whole tree is implied by the corners' path keys:
// sort the keys once, take the depth at which neighbours in the sorted order separate, and the number of nodes
// after a full round of splits at depth d is a prefix sum over the histogram of those depths.  Only the last,
// partial round needs per-node work (count, gain, order).
#include "common.cuh"

#define OT_THREADS 256
#define OT_MAXM 16384            // corners per (level, frame) the sort holds; above that the weakest are dropped first
#define OT_DEPTH 12              // quadtree depth: regions up to 4096 px
#define OT_CELL 30               // ORB-SLAM2's cell size W
#define OT_MAXREC 8192

namespace {

__device__ __forceinline__ void bitonic_sort_u64(unsigned long long* a, int np2, int tid)
{
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += OT_THREADS) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long x = a[i], y = a[ixj];
                    const bool up = (i & k) == 0;
                    if ((x > y) == up) { a[i] = y; a[ixj] = x; }
                }
            }
            __syncthreads();
        }
}

__device__ __forceinline__ void bitonic_sort_u32_desc(uint32_t* a, int np2, int tid)
{
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < np2; i += OT_THREADS) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const uint32_t x = a[i], y = a[ixj];
                    const bool up = (i & k) == 0;
                    if ((x < y) == up) { a[i] = y; a[ixj] = x; }
                }
            }
            __syncthreads();
        }
}

__global__ void __launch_bounds__(OT_THREADS)
vs_octree_kernel(const __grid_constant__ VsPlan plan, int ini_thr, int m_lo, int m_hi, const VsCand* __restrict__ cand,
                 VsCand* __restrict__ stage, int32_t* __restrict__ counters, int32_t* __restrict__ status)
{
    extern __shared__ __align__(16) unsigned long long s_key[];      // [np2] (key << 32) | candidate index
    __shared__ uint32_t s_cell[1024];                                // cells that hold a corner above ini_thr
    __shared__ uint32_t s_rec[OT_MAXREC];
    __shared__ int s_hist[OT_DEPTH + 3], s_w[OT_THREADS / 32];
    __shared__ int s_m, s_nrec, s_df, s_partial, s_size, s_tcap;
    const int level = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const VsLevel& L = plan.lv[level];
    int32_t* ctr = counters + ((size_t)b * plan.n_levels + level) * VS_CTR_STRIDE;
    const int ncand = ctr[VS_CTR_NCAND], N = L.quota;
    if (ncand == 0 || N == 0) return;            // NFINAL stays 0 (counters are zeroed per batch)
    // one launch per shared-memory size class: this CTA's class is the one whose sort buffer fits its corners
    if (min(ncand, OT_MAXM) <= m_lo || min(ncand, OT_MAXM) > m_hi) return;
    const VsCand* cl = cand + (size_t)b * plan.cand_frame_entries + L.cand_off;

    // ---- geometry of ComputeKeyPointsOctTree with EDGE = 31
    const int minb = VS_EDGE - 3;
    const int width = L.w - 2 * minb, height = L.h - 2 * minb;
    const int ncols = max(width / OT_CELL, 1), nrows = max(height / OT_CELL, 1);
    const int wcell = (width + ncols - 1) / ncols, hcell = (height + nrows - 1) / nrows;
    const int ncx = (width + wcell - 1) / wcell;                     // cells per row of the bitmap
    int nini = (int)roundf((float)width / (float)height);
    nini = min(max(nini, 1), 255);            // the root index has 8 bits of the key
    const float hx = (float)width / (float)nini;

    for (int i = tid; i < 1024; i += OT_THREADS) s_cell[i] = 0;
    if (tid < OT_DEPTH + 3) s_hist[tid] = 0;
    if (tid == 0) {
        s_m = 0; s_nrec = 0;
        // more corners than the sort holds: drop the weakest first (threshold from the level's score histogram)
        int t = 0;
        if (ncand > OT_MAXM) {
            int acc = 0;
            for (t = 255; t > 0; --t) {
                acc += ctr[VS_CTR_HIST + t];
                if (acc > OT_MAXM) { ++t; break; }
            }
        }
        s_tcap = t;
    }
    __syncthreads();
    const int tcap = s_tcap;
    // ---- 1. cells with a corner above the initial threshold (cv2.FAST keeps S > thr, the stored score is S - 1)
    for (int i = tid; i < ncand; i += OT_THREADS) {
        const VsCand c = cl[i];
        if ((int)c.v >= ini_thr) {
            const int cj = (c.x - minb - 3) / wcell, ci = (c.y - minb - 3) / hcell;
            const int cell = min(ci * ncx + cj, 32767);
            atomicOr(&s_cell[cell >> 5], 1u << (cell & 31));
        }
    }
    __syncthreads();
    // ---- 2. keep: initial threshold where the cell has such a corner, minimum threshold elsewhere; path keys
    for (int i = tid; i < ncand; i += OT_THREADS) {
        const VsCand c = cl[i];
        const int rx = c.x - minb, ry = c.y - minb;
        const int cj = (rx - 3) / wcell, ci = (ry - 3) / hcell;
        if (cj * wcell >= width - 6 || ci * hcell >= height - 3) continue;      // ORB-SLAM2 skips these border cells
        const int cell = min(ci * ncx + cj, 32767);
        const bool has_ini = (s_cell[cell >> 5] >> (cell & 31)) & 1u;
        if (((int)c.v < ini_thr && has_ini) || (int)c.v < tcap) continue;
        int r = (int)((float)rx / hx);
        r = min(r, nini - 1);
        int lox = (int)(hx * (float)r), hix = (int)(hx * (float)(r + 1)), loy = 0, hiy = height;
        uint32_t path = 0;
#pragma unroll
        for (int d = 0; d < OT_DEPTH; ++d) {
            const int mx = lox + ((hix - lox + 1) >> 1), my = loy + ((hiy - loy + 1) >> 1);
            const int cx = rx >= mx, cy = ry >= my;
            if (cx) lox = mx; else hix = mx;
            if (cy) loy = my; else hiy = my;
            path = (path << 2) | (uint32_t)(cx + 2 * cy);
        }
        const int slot = atomicAdd(&s_m, 1);
        if (slot < OT_MAXM) s_key[slot] = ((unsigned long long)(((uint32_t)r << 24) | path) << 32) | (uint32_t)i;
    }
    __syncthreads();
    const int M = min(s_m, OT_MAXM);
    if (M == 0) return;
    int np2 = 1;
    while (np2 < M) np2 <<= 1;
    for (int i = M + tid; i < np2; i += OT_THREADS) s_key[i] = ~0ull;
    __syncthreads();
    bitonic_sort_u64(s_key, np2, tid);
    uint8_t* s_sep = (uint8_t*)(s_key + np2);                         // depth at which i separates from i-1
    uint8_t* s_hf = s_sep + np2;                                      // final head flags
    // ---- 3. separation depths and their histogram
    for (int i = tid; i < M; i += OT_THREADS) {
        int sep = 0;
        if (i > 0) {
            const uint32_t x = (uint32_t)(s_key[i] >> 32) ^ (uint32_t)(s_key[i - 1] >> 32);
            if (x == 0) sep = OT_DEPTH + 1;
            else if (x >> 24) sep = 0;
            else sep = (23 - (31 - __clz(x))) / 2 + 1;
        }
        s_sep[i] = (uint8_t)sep;
        atomicAdd(&s_hist[sep], 1);
    }
    __syncthreads();
    // ---- 4. full rounds: split every node with more than one corner while the quota stays out of reach
    if (tid == 0) {
        int size = s_hist[0], d = 0, partial = 0;
        while (d < OT_DEPTH) {
            if (size >= M) break;                          // every node holds one corner
            if (d > 0 && size >= N) break;                 // the first round always splits every root node, like ORB-SLAM2
            const int G = s_hist[d + 1];
            if (d > 0 && size + G >= N) { partial = 1; break; }   // this round reaches the quota: fullest nodes first
            if (G == 0) break;                             // a round that adds no node ends the loop (lNodes.size() == prevSize)
            size += G;
            ++d;
        }
        s_df = d; s_partial = partial; s_size = size;
    }
    __syncthreads();
    const int df = s_df;
    for (int i = tid; i < M; i += OT_THREADS) s_hf[i] = s_sep[i] <= df;
    __syncthreads();
    if (s_partial) {
        // ---- 5. last round: nodes with more than one corner, ordered by (count desc, quadtree order), are split
        // until the quota is reached.  record = count << 17 | (16383 - head) << 3 | gain
        for (int i = tid; i < M; i += OT_THREADS) {
            if (s_sep[i] > df) continue;
            int cnt = 1, gain = 0;
            for (int j = i + 1; j < M && s_sep[j] > df; ++j) { ++cnt; gain += s_sep[j] == df + 1; }
            if (cnt > 1) {
                const int slot = atomicAdd(&s_nrec, 1);
                if (slot < OT_MAXREC) s_rec[slot] = ((uint32_t)cnt << 17) | ((uint32_t)(16383 - i) << 3) | (uint32_t)gain;
            }
        }
        __syncthreads();
        const int nrec = min(s_nrec, OT_MAXREC);
        int rp2 = 1;
        while (rp2 < nrec) rp2 <<= 1;
        for (int i = nrec + tid; i < rp2; i += OT_THREADS) s_rec[i] = 0;
        __syncthreads();
        bitonic_sort_u32_desc(s_rec, rp2, tid);
        if (tid == 0) {
            int size = s_size, k = 0;
            for (; k < nrec && size < N; ++k) size += s_rec[k] & 7;
            s_nrec = k;                                    // the first k records are split
        }
        __syncthreads();
        const int nsplit = s_nrec;
        for (int k = tid; k < nsplit; k += OT_THREADS) {
            const int head = 16383 - (int)((s_rec[k] >> 3) & 16383);
            for (int j = head + 1; j < M && s_sep[j] > df; ++j)
                if (s_sep[j] == df + 1) s_hf[j] = 1;
        }
        __syncthreads();
    }
    // ---- 6. one corner per node: the highest FAST score (first in quadtree order on ties); rank = node order
    const int chunk = (M + OT_THREADS - 1) / OT_THREADS;
    const int i0 = tid * chunk, i1 = min(i0 + chunk, M);
    int heads = 0;
    for (int i = i0; i < i1; ++i) heads += s_hf[i];
    int incl = heads;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[wid] = incl;
    __syncthreads();
    int base = incl - heads;
    for (int w = 0; w < wid; ++w) base += s_w[w];
    int total = 0;
    for (int w = 0; w < OT_THREADS / 32; ++w) total += s_w[w];
    VsCand* st = stage + (size_t)b * plan.stage_frame_entries + L.stage_off;
    const int cap = L.stage_cap;
    for (int i = i0; i < i1; ++i) {
        if (!s_hf[i]) continue;
        VsCand best = cl[(uint32_t)s_key[i]];
        for (int j = i + 1; j < M && !s_hf[j]; ++j) {
            const VsCand c = cl[(uint32_t)s_key[j]];
            if (c.v > best.v) best = c;
        }
        if (base < cap) {
            best.v = __float_as_uint((float)best.v);
            st[base] = best;
        }
        ++base;
    }
    if (tid == 0) {
        if (total > cap) { atomicExch(status, VSLAM_E_CAPACITY); total = cap; }
        ctr[VS_CTR_NFINAL] = total;
    }
}

}  // namespace

// shared-memory size classes (corners per level and frame): small levels run several CTAs per SM
static const int kOctreeClass[] = {0, 2048, 8192, OT_MAXM};

static size_t vs_octree_smem(int m) { return (size_t)m * (sizeof(unsigned long long) + 2); }

int vs_configure_octree_smem(void)
{
    VS_CUDA(cudaFuncSetAttribute(vs_octree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vs_octree_smem(OT_MAXM)));
    return VSLAM_OK;
}

// Replaces the Harris + retain-best pair of vs_launch_select when the handle is in ORB-SLAM2 distribution mode.
int vs_launch_octree(vslam_orb* o, int batch, cudaStream_t s)
{
    const VsPlan& p = o->plan;
    int max_cand = 0;
    for (int l = 0; l < p.n_levels; ++l) max_cand = max(max_cand, p.lv[l].cand_cap);
    for (int c = 1; c < 4; ++c) {
        if (kOctreeClass[c - 1] >= max_cand) break;       // no level can hold that many corners
        __atomic_fetch_add(&g_vs_launches, 1ull, __ATOMIC_RELAXED);
        vs_octree_kernel<<<dim3(p.n_levels, batch), OT_THREADS, vs_octree_smem(kOctreeClass[c]), s>>>(
            p, o->ini_threshold, kOctreeClass[c - 1], kOctreeClass[c], o->d_cand, o->d_stage, o->d_counters, o->d_status);
    }
    VS_CUDA(cudaGetLastError());
    return VSLAM_OK;
}
