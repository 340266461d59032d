// common.cuh -- shared declarations of libvslam_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/vslam_b200.h"

#define VS_MAX_LEVELS 16
#define VS_EDGE 31        // edgeThreshold of cv2.ORB_create, extractor.py:42
#define VS_HALF_PATCH 15  // patchSize 31, extractor.py:46

// Fused level kernel tile: TW x TH output pixels, halo of 4 (3 for the FAST circle
// and the 7-tap blur, +1 for the NMS neighbours' scores).  TH = 30 makes the score rows
// (TH + 2) x 16 four-pixel words exactly 2 items per thread of a 256-thread CTA.
#define VS_TW 64
#define VS_TH 30
#define VS_HALO 4

extern unsigned long long g_vs_launches;   // kernels launched by this library (bench.py reports it)
void vs_set_error(const char* fmt, ...);
int vs_cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define VS_CUDA(call)                                                      \
    do {                                                                   \
        cudaError_t e__ = (call);                                          \
        if (e__ != cudaSuccess) return vs_cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)

#define VS_REQUIRE(cond, msg)                 \
    do {                                      \
        if (!(cond)) {                        \
            vs_set_error("%s: %s", __func__, msg); \
            return VSLAM_E_INVALID;           \
        }                                     \
    } while (0)

// FAST candidate of one (frame, level): level coordinates + value.  `v` is the FAST
// score (as float) after the level kernel, then the order-preserving key of the
// Harris response (0 = dropped at the first retain-best cut) after the Harris kernel.
struct __align__(8) VsCand {
    uint16_t x, y;
    uint32_t v;
};

struct VsLevel {
    int w, h, pitch;        // level size and row pitch (bytes) of raw/blur storage
    float scale;            // (float)pow((double)(float)scale_factor, l)
    int quota;              // per-level feature quota
    int cand_cap;           // candidate capacity per frame
    int stage_cap;          // staging rows per frame for the final per-level selection
    size_t raw_off, blur_off;   // byte offsets inside one frame's pyramid block
    size_t cand_off;        // entry offset inside one frame's candidate block
    size_t stage_off;       // entry offset inside one frame's staging block
    int tiles_x, tiles_y;
    int tabx_off, taby_off; // offsets of this level's resize tables (from level l-1) in the table block
    int bndx_off, bndy_off; // offsets of the per-tile parent-region bounds (lo | hi << 16) in the table block
};

// Device-side plan (kernel argument, passed by value through constant bank).
struct VsPlan {
    int n_levels, width, height, n_features, fast_thr, kp_cap;
    VsLevel lv[VS_MAX_LEVELS];
    size_t pyr_frame_bytes;     // bytes of one frame's pyramid block (raw 1.. + blur 0..)
    size_t cand_frame_entries;
    size_t stage_frame_entries;
    int par_w, par_h;           // level kernel: shared-memory extent of the staged parent region
};

// profiling slots: 0..15 level kernels, then the per-keypoint kernels and matching
#define VS_PROF_HARRIS 16
#define VS_PROF_SELECT 17
#define VS_PROF_DESCRIBE 18
#define VS_PROF_KNN2 19
#define VS_PROF_RATIO 20
#define VS_PROF_SLOTS 24
#define VS_PROF_MAX_MARKS 8192     // marks of one call (enable = 1) or of the calls since the last read (enable = 2)

#define VS_HOST_CHUNKS 8

struct vslam_orb {
    vslam_orb_params prm;
    VsPlan plan;
    int device;
    // device workspace
    uint8_t* d_pyr;        // [max_batch][pyr_frame_bytes]
    VsCand* d_cand;        // [max_batch][cand_frame_entries]
    VsCand* d_stage;       // [max_batch][stage_frame_entries]  (x, y, response bits)
    int32_t* d_counters;   // [max_batch][n_levels][VS_CTR_STRIDE]
    uint32_t* d_tab;       // resize tables: (alpha_q << 16) | i0 per destination column / row
    int32_t* d_status;     // 1 int: sticky overflow flag
    // frontend scratch
    int32_t* d_knn_idx;    // [max_batch][kp_cap][2]
    int32_t* d_knn_dist;
    uint8_t* d_knn_ws;     // expanded descriptor tiles of the tcgen05 kNN-2: [max_batch + 1] x vs_knn2_tc5_workspace_per_pair(kp_cap)
    size_t knn_ws_bytes;   // bytes available from d_knn_ws (a chunk view of the handle sees its own slice)
    int4* d_knn_partial;   // partial top-2 lists of the split-train kNN-2 (small batches), owned by the handle so that
    size_t knn_partial_bytes;   // such a call needs neither the process-wide scratch nor its events (graph capture)
    // host staging for *_host entry points
    uint8_t* d_frames;     // [max_batch][h][pitch0]
    uint8_t* d_track;      // outputs + workspace of vslam_track_clip_host (lazy, one block)
    uint8_t* d_bgr;        // [max_batch][h][w][3] staging of vslam_orb_detect_and_compute_host_bgr (lazy)
    vslam_keypoint* d_kp;
    uint8_t* d_desc;
    int32_t* d_count;
    int32_t* d_match;
    int32_t* d_match_count;
    int frame_pitch;
    int dist_mode;             // 0: cv2.ORB (Harris retain-best), 1: ORB-SLAM2-style cells + quadtree (orb_octree.cu)
    int ini_threshold;         // initial FAST threshold of mode 1
    cudaStream_t stream;       // own stream for *_host calls
    // chunked host pipeline (vslam_frontend_run_host): copy-in / copy-out streams and per-chunk events
    cudaStream_t stream_in, stream_out, stream2;   // stream / stream2 alternate between compute chunks
    cudaEvent_t ev_in[VS_HOST_CHUNKS], ev_ext[VS_HOST_CHUNKS], ev_done[VS_HOST_CHUNKS];
    cudaEvent_t ev_host_done;  // blocking-sync event: end of a *_host call (the waiting host thread sleeps instead of spinning)
    int knn_no_split;          // set on the per-chunk views: the split-train scratch of kNN-2 is not per stream
    // CUDA graphs of vslam_frontend_run for small batches (the per-frame call of the reference's Tracker): a call is
    // a chain of ~18 short launches, and at batch 1 the launch gaps are as long as the kernels.  A call whose
    // arguments were seen before is captured once and replayed (VS_GRAPH_SLOTS argument sets, round robin).
    struct GraphSlot {
        uint64_t key[12];
        int seen;
        cudaGraphExec_t exec;
    } graph[4];
    int graph_next, graph_off;
    cudaStream_t last_stream;
    const uint8_t* last_frames;  // level-0 pixels of the last batch (for debug taps)
    int last_pitch, last_batch;
    // optional per-kernel timing (vslam_orb_set_profiling): events recorded between launches
    int prof_on, prof_count;
    cudaEvent_t prof_ev[VS_PROF_MAX_MARKS];
    int prof_slot[VS_PROF_MAX_MARKS];
};

// counters per (frame, level): [0] candidates, [1] survivors of cut 1, [2] final count,
// [3] unused, [4..4+256) FAST score histogram
#define VS_CTR_NCAND 0
#define VS_CTR_NSURV 1
#define VS_CTR_NFINAL 2
#define VS_CTR_HIST 4
#define VS_CTR_STRIDE (4 + 256)

// per-thread, per-device scratch buffers (abi.cu): *out = device pointer of at least `bytes` bytes in `slot`
#define VS_SCRATCH_SLOTS 8
int vs_scratch(int slot, size_t bytes, void** out);

// kernels / launchers implemented in the other translation units
int vs_configure_level_smem(const VsPlan& p);
int vs_launch_levels(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch, cudaStream_t s);
int vs_launch_bgr2gray(const uint8_t* d_bgr, uint8_t* d_gray, size_t n_pixels, cudaStream_t s);
int vs_configure_octree_smem(void);
int vs_launch_octree(vslam_orb* o, int batch, cudaStream_t s);
struct CUtensorMap_st;
bool vs_image_tensor_map(struct CUtensorMap_st* tm, const uint8_t* base, int w, int h, size_t pitch, size_t frame_stride,
                         int frames, int box_w, int box_h, bool swizzle64 = false);
int vs_launch_select(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch,
                     vslam_keypoint* d_kp, uint8_t* d_desc, int32_t* d_count, cudaStream_t s);
size_t vs_knn2_tc5_workspace_per_pair(int nt);
int vs_launch_knn2(vslam_orb* prof, const uint8_t* d_q, int nq, long q_stride, const uint8_t* d_t, int nt,
                   long t_stride, const int32_t* d_nq, const int32_t* d_nt, int pairs,
                   int32_t* d_idx, int32_t* d_dist, cudaStream_t s);
int vs_launch_ratio(vslam_orb* prof, const int32_t* d_idx, const int32_t* d_dist, int nq, const int32_t* d_nq,
                    int pairs, double ratio, int ratio_test, int32_t* d_match,
                    int32_t* d_match_count, cudaStream_t s);

// record a timing mark: the time from this mark to the next one is attributed to `slot`
// (slot < 0 = end of the run)
static inline void vs_prof_mark(vslam_orb* o, int slot, cudaStream_t s)
{
    if (!o || !o->prof_on || o->prof_count >= VS_PROF_MAX_MARKS) return;
    if (!o->prof_ev[o->prof_count]) cudaEventCreate(&o->prof_ev[o->prof_count]);
    cudaEventRecord(o->prof_ev[o->prof_count], s);
    o->prof_slot[o->prof_count++] = slot;
}

__host__ __device__ inline int vs_reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}
