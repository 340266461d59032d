// abi.cu -- extern "C" entry points of libvslam_b200.so (include/vslam_b200.h):
// handle/plan management, the extraction + matching front end, host-buffer wrappers.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include <new>
#include <vector>

#include "common.cuh"

static thread_local char g_err[512] = "";
unsigned long long g_vs_launches = 0;
extern "C" unsigned long long vslam_launch_count(void) { return g_vs_launches; }

void vs_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int vs_cuda_fail(cudaError_t e, const char* what, const char* file, int line)
{
    vs_set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return e == cudaErrorMemoryAllocation ? VSLAM_E_NOMEM : VSLAM_E_CUDA;
}

extern "C" const char* vslam_last_error(void) { return g_err; }
extern "C" int vslam_abi_version(void) { return 1000; }
extern "C" int vslam_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// Per-thread, per-device scratch for the entry points that need a little device memory of their own (the *_host
// wrappers, the device forms of the geometry calls): a few growing buffers per host thread instead of a
// cudaMalloc / cudaFree pair per call.  Slots are independent buffers; a pointer stays valid until the same thread
// asks the same slot for more.  Freed when the thread exits.
namespace {
struct TlsScratch {
    void* p[VS_SCRATCH_SLOTS] = {};
    size_t n[VS_SCRATCH_SLOTS] = {};
    int dev = -1;
    void release()
    {
        for (int i = 0; i < VS_SCRATCH_SLOTS; ++i) {
            if (p[i]) cudaFree(p[i]);
            p[i] = nullptr;
            n[i] = 0;
        }
    }
    ~TlsScratch() { release(); cudaGetLastError(); }
};
thread_local TlsScratch g_scratch;
}

int vs_scratch(int slot, size_t bytes, void** out)
{
    *out = nullptr;
    VS_REQUIRE(slot >= 0 && slot < VS_SCRATCH_SLOTS, "bad scratch slot");
    int dev = 0;
    VS_CUDA(cudaGetDevice(&dev));
    if (dev != g_scratch.dev) { g_scratch.release(); g_scratch.dev = dev; }
    if (bytes < 256) bytes = 256;
    if (bytes > g_scratch.n[slot]) {
        if (g_scratch.p[slot]) { VS_CUDA(cudaDeviceSynchronize()); VS_CUDA(cudaFree(g_scratch.p[slot])); }
        g_scratch.p[slot] = nullptr;
        g_scratch.n[slot] = 0;
        const size_t grow = bytes + bytes / 2;
        VS_CUDA(cudaMalloc(&g_scratch.p[slot], grow));
        g_scratch.n[slot] = grow;
    }
    *out = g_scratch.p[slot];
    return VSLAM_OK;
}

static int check_device()
{
    int dev = 0;
    VS_CUDA(cudaGetDevice(&dev));
    int major = 0;
    VS_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) {
        vs_set_error("libvslam_b200 is built for sm_100a only; device %d has compute capability %d.x", dev, major);
        return VSLAM_E_CUDA;
    }
    return VSLAM_OK;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// INTER_LINEAR_EXACT per-axis table: i0 and the 8-bit fraction.
static void linear_exact_table(int src_n, int dst_n, uint32_t* out)
{
    const double s = (double)src_n / (double)dst_n;
    for (int i = 0; i < dst_n; ++i) {
        double f = ((double)i + 0.5) * s - 0.5;
        int k = (int)floor(f);
        int q = (int)lrint((f - (double)k) * 256.0);
        if (k < 0) { k = 0; q = 0; }
        if (k >= src_n - 1) { k = src_n - 1; q = 0; }
        out[i] = ((uint32_t)q << 16) | (uint32_t)k;
    }
}

static int build_plan(const vslam_orb_params& prm, VsPlan& p, std::vector<uint32_t>& tab)
{
    memset(&p, 0, sizeof(p));
    p.n_levels = prm.n_levels;
    p.width = prm.width;
    p.height = prm.height;
    p.n_features = prm.n_features;
    p.fast_thr = prm.fast_threshold;
    p.kp_cap = prm.kp_capacity > 0 ? prm.kp_capacity : prm.n_features + 64;
    // per-level quotas
    const float factor = (float)(1.0 / (double)prm.scale_factor);
    float nd = (float)prm.n_features * (1.0f - factor) /
               (1.0f - (float)pow((double)factor, (double)prm.n_levels));
    int sum = 0;
    size_t pyr = 0, cand = 0, stage = 0;
    for (int l = 0; l < prm.n_levels; ++l) {
        VsLevel& L = p.lv[l];
        L.scale = (float)pow((double)prm.scale_factor, (double)l);
        L.w = (int)lrintf((float)prm.width / L.scale);
        L.h = (int)lrintf((float)prm.height / L.scale);
        if (l < prm.n_levels - 1) {
            L.quota = (int)lrintf(nd);
            sum += L.quota;
            nd *= factor;
        } else {
            L.quota = prm.n_features - sum > 0 ? prm.n_features - sum : 0;
        }
        if (L.w <= 0 || L.h <= 0) { L.w = L.h = 0; continue; }
        if (L.w > 65535 || L.h > 65535) { vs_set_error("frame too large"); return VSLAM_E_INVALID; }
        L.pitch = (int)align_up((size_t)L.w, 32);
        L.tiles_x = (L.w + VS_TW - 1) / VS_TW;
        L.tiles_y = (L.h + VS_TH - 1) / VS_TH;
        if (l > 0) { L.raw_off = pyr; pyr += align_up((size_t)L.pitch * L.h, 256); }
        L.blur_off = pyr;
        pyr += align_up((size_t)L.pitch * L.h, 256);
        const int rw = L.w - 2 * VS_EDGE, rh = L.h - 2 * VS_EDGE;
        L.cand_cap = (rw > 0 && rh > 0) ? (rw / 2 + 1) * (rh / 2 + 1) : 0;
        L.cand_off = cand;
        cand += align_up((size_t)L.cand_cap, 32);
        int sc = 64;
        while (sc < L.quota + 64) sc <<= 1;
        L.stage_cap = sc;
        L.stage_off = stage;
        stage += sc;
        if (l > 0) {
            L.tabx_off = (int)tab.size();
            tab.resize(tab.size() + L.w);
            linear_exact_table(p.lv[l - 1].w, L.w, tab.data() + L.tabx_off);
            L.taby_off = (int)tab.size();
            tab.resize(tab.size() + L.h);
            linear_exact_table(p.lv[l - 1].h, L.h, tab.data() + L.taby_off);
        }
    }
    // extent of the parent region one tile of the level kernel stages in shared memory
    // (same arithmetic as the kernel: reflected tile coordinates -> table entries -> bounding box)
    p.par_w = 4; p.par_h = 1;
    for (int l = 1; l < prm.n_levels; ++l) {
        const VsLevel& L = p.lv[l];
        if (L.w <= 0 || L.h <= 0) continue;
        for (int axis = 0; axis < 2; ++axis) {
            const int n = axis == 0 ? L.w : L.h, src_n = axis == 0 ? p.lv[l - 1].w : p.lv[l - 1].h;
            const int tile = axis == 0 ? VS_TW : VS_TH, tiles = axis == 0 ? L.tiles_x : L.tiles_y;
            // per-tile bounds (lo | hi << 16) of the staged parent region, read by the kernel
            const int bnd_off = (int)tab.size();
            tab.resize(tab.size() + tiles);
            if (axis == 0) p.lv[l].bndx_off = bnd_off; else p.lv[l].bndy_off = bnd_off;
            const uint32_t* t = tab.data() + (axis == 0 ? L.tabx_off : L.taby_off);
            for (int k = 0; k < tiles; ++k) {
                int lo = 0x7fffffff, hi = 0;
                for (int c = 0; c < tile + 2 * VS_HALO; ++c) {
                    int i0 = (int)(t[vs_reflect101(k * tile - VS_HALO + c, n)] & 0xffff);
                    lo = i0 < lo ? i0 : lo;
                    hi = i0 > hi ? i0 : hi;
                }
                hi = hi + 1 < src_n - 1 ? hi + 1 : src_n - 1;
                tab[bnd_off + k] = (uint32_t)lo | ((uint32_t)hi << 16);
                if (axis == 0) {
                    int wbytes = (int)align_up((size_t)(hi - (lo & ~15)) + 1, 16);   // TMA: box start and row length are multiples of 16 bytes
                    p.par_w = wbytes > p.par_w ? wbytes : p.par_w;
                } else {
                    p.par_h = hi - lo + 1 > p.par_h ? hi - lo + 1 : p.par_h;
                }
            }
        }
    }
    p.pyr_frame_bytes = align_up(pyr, 256);
    p.cand_frame_entries = align_up(cand, 32);
    p.stage_frame_entries = stage;
    return VSLAM_OK;
}

extern void vs_configure_select_smem(int bytes);

extern "C" int vslam_orb_create(const vslam_orb_params* prm, vslam_orb** out)
{
    VS_REQUIRE(prm && out, "null argument");
    VS_REQUIRE(prm->width > 0 && prm->height > 0, "bad frame size");
    VS_REQUIRE(prm->n_levels >= 1 && prm->n_levels <= VS_MAX_LEVELS, "n_levels out of range");
    VS_REQUIRE(prm->n_features >= 0 && prm->n_features <= 200000, "n_features out of range");
    VS_REQUIRE(prm->scale_factor > 1.0f, "scale_factor must be > 1");
    VS_REQUIRE(prm->max_batch >= 1, "max_batch must be >= 1");
    VS_REQUIRE(prm->fast_threshold >= 0 && prm->fast_threshold < 255, "fast_threshold out of range");
    int rc = check_device();
    if (rc) return rc;
    vslam_orb* o = new (std::nothrow) vslam_orb();
    if (!o) return VSLAM_E_NOMEM;
    memset(o, 0, sizeof(*o));
    o->prm = *prm;
    std::vector<uint32_t> tab;
    tab.reserve(8192);
    rc = build_plan(*prm, o->plan, tab);
    if (rc) { delete o; return rc; }
    if (tab.empty()) tab.push_back(0);
    const VsPlan& p = o->plan;
    const size_t B = prm->max_batch;
    int max_stage = 0;
    for (int l = 0; l < p.n_levels; ++l) max_stage = max_stage > p.lv[l].stage_cap ? max_stage : p.lv[l].stage_cap;
    if ((size_t)max_stage * 8 > 200 * 1024) {
        vs_set_error("n_features too large for the per-level selection staging");
        delete o;
        return VSLAM_E_INVALID;
    }
    vs_configure_select_smem(max_stage * 8);
    rc = vs_configure_level_smem(o->plan);
    if (rc) { delete o; return rc; }
    cudaGetDevice(&o->device);
#define ALLOC(ptr, bytes)                                                   \
    do {                                                                    \
        cudaError_t e = cudaMalloc((void**)&(ptr), (size_t)(bytes) > 0 ? (size_t)(bytes) : (size_t)256); \
        if (e != cudaSuccess) { vslam_orb_destroy(o); return vs_cuda_fail(e, "cudaMalloc " #ptr, __FILE__, __LINE__); } \
    } while (0)
    ALLOC(o->d_pyr, B * p.pyr_frame_bytes);
    ALLOC(o->d_cand, B * p.cand_frame_entries * sizeof(VsCand));
    ALLOC(o->d_stage, B * p.stage_frame_entries * sizeof(VsCand));
    ALLOC(o->d_counters, B * p.n_levels * VS_CTR_STRIDE * sizeof(int32_t));
    ALLOC(o->d_tab, tab.size() * sizeof(uint32_t));
    ALLOC(o->d_status, sizeof(int32_t));
    ALLOC(o->d_knn_idx, B * p.kp_cap * 2 * sizeof(int32_t));
    ALLOC(o->d_knn_dist, B * p.kp_cap * 2 * sizeof(int32_t));
    o->knn_ws_bytes = (B + 1) * vs_knn2_tc5_workspace_per_pair(p.kp_cap);   // + 1: a batch of one matched against the previous frame expands two sets
    ALLOC(o->d_knn_ws, o->knn_ws_bytes);
    {   // split-train kNN-2 of small batches: pairs * nsplit <= pairs + 148 * 8 / qtiles + 1 partial lists of kp_cap rows
        const size_t qtiles = ((size_t)p.kp_cap + 255) / 256;
        o->knn_partial_bytes = (B + 148 * 8 / qtiles + 2) * (size_t)p.kp_cap * sizeof(int4);
        ALLOC(o->d_knn_partial, o->knn_partial_bytes);
    }
#undef ALLOC
    cudaError_t e = cudaMemcpy(o->d_tab, tab.data(), tab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(o->d_status, 0, sizeof(int32_t));
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&o->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&o->stream_in, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&o->stream_out, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&o->stream2, cudaStreamNonBlocking);
    for (int i = 0; i < VS_HOST_CHUNKS && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&o->ev_in[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&o->ev_ext[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&o->ev_done[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&o->ev_host_done, cudaEventDisableTiming | cudaEventBlockingSync);
    if (e != cudaSuccess) { vslam_orb_destroy(o); return vs_cuda_fail(e, "orb_create init", __FILE__, __LINE__); }
    *out = o;
    return VSLAM_OK;
}

extern "C" int vslam_orb_destroy(vslam_orb* o)
{
    if (!o) return VSLAM_OK;
    cudaDeviceSynchronize();
    void* ptrs[] = {o->d_pyr, o->d_cand, o->d_stage, o->d_counters, o->d_tab, o->d_status, o->d_knn_idx,
                    o->d_knn_dist, o->d_knn_ws, o->d_knn_partial, o->d_frames, o->d_bgr, o->d_track, o->d_kp, o->d_desc, o->d_count, o->d_match, o->d_match_count};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (o->stream) cudaStreamDestroy(o->stream);
    if (o->stream_in) cudaStreamDestroy(o->stream_in);
    if (o->stream_out) cudaStreamDestroy(o->stream_out);
    if (o->stream2) cudaStreamDestroy(o->stream2);
    for (int i = 0; i < VS_HOST_CHUNKS; ++i) {
        if (o->ev_in[i]) cudaEventDestroy(o->ev_in[i]);
        if (o->ev_ext[i]) cudaEventDestroy(o->ev_ext[i]);
        if (o->ev_done[i]) cudaEventDestroy(o->ev_done[i]);
    }
    if (o->ev_host_done) cudaEventDestroy(o->ev_host_done);
    for (auto& g : o->graph) if (g.exec) cudaGraphExecDestroy(g.exec);
    for (cudaEvent_t e : o->prof_ev) if (e) cudaEventDestroy(e);
    delete o;
    return VSLAM_OK;
}

extern "C" int vslam_orb_set_distribution(vslam_orb* o, int mode, int ini_threshold)
{
    VS_REQUIRE(o, "null handle");
    VS_REQUIRE(mode == VSLAM_DIST_CV2_ORB || mode == VSLAM_DIST_ORBSLAM2, "unknown distribution mode");
    if (mode == VSLAM_DIST_ORBSLAM2) {
        VS_REQUIRE(ini_threshold >= o->prm.fast_threshold && ini_threshold < 255,
                   "ini_threshold must lie in [min_threshold, 255)");
        VS_REQUIRE(o->plan.width - 56 <= 4096 && o->plan.height - 56 <= 4096,
                   "quadtree distribution handles frames up to 4152 px per side");
        int rc = vs_configure_octree_smem();
        if (rc) return rc;
    }
    o->dist_mode = mode;
    o->ini_threshold = ini_threshold;
    return VSLAM_OK;
}

extern "C" int vslam_orb_kp_capacity(const vslam_orb* o) { return o ? o->plan.kp_cap : VSLAM_E_INVALID; }

extern "C" int vslam_orb_level_info(const vslam_orb* o, int level, int* w, int* h, int* pitch, float* scale,
                                    int* quota)
{
    VS_REQUIRE(o && level >= 0 && level < o->plan.n_levels, "bad level");
    const VsLevel& L = o->plan.lv[level];
    if (w) *w = L.w;
    if (h) *h = L.h;
    if (pitch) *pitch = L.pitch;
    if (scale) *scale = L.scale;
    if (quota) *quota = L.quota;
    return VSLAM_OK;
}

extern "C" int vslam_orb_detect_and_compute(vslam_orb* o, const uint8_t* d_frames, int batch, int pitch,
                                            vslam_keypoint* d_kp, uint8_t* d_desc, int32_t* d_count,
                                            void* stream)
{
    VS_REQUIRE(o && d_frames && d_kp && d_desc && d_count, "null argument");
    VS_REQUIRE(batch >= 1 && batch <= o->prm.max_batch, "batch exceeds max_batch");
    VS_REQUIRE(pitch >= o->prm.width, "pitch smaller than width");
    {   // the handle's workspace, streams and kernel attributes belong to the device it was created on
        int dev = -1;
        VS_CUDA(cudaGetDevice(&dev));
        if (dev != o->device) {
            vs_set_error("vslam_orb_detect_and_compute: handle was created on device %d, current device is %d "
                         "(cudaSetDevice first; one handle per device)", o->device, dev);
            return VSLAM_E_INVALID;
        }
    }
    cudaStream_t s = (cudaStream_t)stream;
    o->last_stream = s;
    o->last_frames = d_frames;
    o->last_pitch = pitch;
    o->last_batch = batch;
    if (o->prof_on != 2) o->prof_count = 0;      // 2: marks accumulate over calls until they are read
    int rc = vs_launch_levels(o, d_frames, batch, pitch, s);
    if (rc) return rc;
    return vs_launch_select(o, d_frames, batch, pitch, d_kp, d_desc, d_count, s);
}

extern "C" int vslam_orb_last_status(vslam_orb* o)
{
    VS_REQUIRE(o, "null handle");
    int32_t st = 0;
    VS_CUDA(cudaStreamSynchronize(o->last_stream));
    VS_CUDA(cudaMemcpy(&st, o->d_status, sizeof(st), cudaMemcpyDeviceToHost));
    if (st) {
        VS_CUDA(cudaMemset(o->d_status, 0, sizeof(st)));
        vs_set_error("retain-best ties exceeded the per-level staging capacity");
    }
    return st;
}

int vs_ensure_host_staging(vslam_orb* o)
{
    if (o->d_frames) return VSLAM_OK;
    const VsPlan& p = o->plan;
    const size_t B = o->prm.max_batch;
    o->frame_pitch = p.width;
    VS_CUDA(cudaMalloc((void**)&o->d_frames, B * (size_t)p.width * p.height + 256));
    VS_CUDA(cudaMalloc((void**)&o->d_kp, B * p.kp_cap * sizeof(vslam_keypoint)));
    VS_CUDA(cudaMalloc((void**)&o->d_desc, B * (size_t)p.kp_cap * 32));
    VS_CUDA(cudaMalloc((void**)&o->d_count, B * sizeof(int32_t)));
    VS_CUDA(cudaMalloc((void**)&o->d_match, B * (size_t)p.kp_cap * 3 * sizeof(int32_t)));
    VS_CUDA(cudaMalloc((void**)&o->d_match_count, B * sizeof(int32_t)));
    // rows past a frame's count are never written by the kernels: make them read as zeros
    VS_CUDA(cudaMemset(o->d_kp, 0, B * p.kp_cap * sizeof(vslam_keypoint)));
    VS_CUDA(cudaMemset(o->d_desc, 0, B * (size_t)p.kp_cap * 32));
    VS_CUDA(cudaMemset(o->d_match, 0, B * (size_t)p.kp_cap * 3 * sizeof(int32_t)));
    return VSLAM_OK;
}

extern "C" int vslam_orb_detect_and_compute_host(vslam_orb* o, const uint8_t* h_frames, int batch,
                                                 vslam_keypoint* h_kp, uint8_t* h_desc, int32_t* h_count)
{
    VS_REQUIRE(o && h_frames && h_kp && h_desc && h_count, "null argument");
    VS_REQUIRE(batch >= 1 && batch <= o->prm.max_batch, "batch exceeds max_batch");
    int rc = vs_ensure_host_staging(o);
    if (rc) return rc;
    const VsPlan& p = o->plan;
    cudaStream_t s = o->stream;
    const size_t fbytes = (size_t)p.width * p.height;
    VS_CUDA(cudaMemcpyAsync(o->d_frames, h_frames, batch * fbytes, cudaMemcpyHostToDevice, s));
    rc = vslam_orb_detect_and_compute(o, o->d_frames, batch, p.width, o->d_kp, o->d_desc, o->d_count, s);
    if (rc) return rc;
    VS_CUDA(cudaMemcpyAsync(h_kp, o->d_kp, (size_t)batch * p.kp_cap * sizeof(vslam_keypoint), cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaMemcpyAsync(h_desc, o->d_desc, (size_t)batch * p.kp_cap * 32, cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaMemcpyAsync(h_count, o->d_count, batch * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaStreamSynchronize(s));
    return vslam_orb_last_status(o);
}

// Ingest on the device: BGR frames are uploaded once, converted to gray (K0) into the frame staging and
// extracted from there -- no gray image travels back to the host as with cvtColor + detectAndCompute
// (extractor.py:61-65).
extern "C" int vslam_orb_detect_and_compute_host_bgr(vslam_orb* o, const uint8_t* h_bgr, int batch,
                                                     vslam_keypoint* h_kp, uint8_t* h_desc, int32_t* h_count)
{
    VS_REQUIRE(o && h_bgr && h_kp && h_desc && h_count, "null argument");
    VS_REQUIRE(batch >= 1 && batch <= o->prm.max_batch, "batch exceeds max_batch");
    int rc = vs_ensure_host_staging(o);
    if (rc) return rc;
    const VsPlan& p = o->plan;
    cudaStream_t s = o->stream;
    const size_t px = (size_t)p.width * p.height;
    if (!o->d_bgr) VS_CUDA(cudaMalloc((void**)&o->d_bgr, (size_t)o->prm.max_batch * px * 3 + 256));
    VS_CUDA(cudaMemcpyAsync(o->d_bgr, h_bgr, batch * px * 3, cudaMemcpyHostToDevice, s));
    rc = vs_launch_bgr2gray(o->d_bgr, o->d_frames, batch * px, s);
    if (rc) return rc;
    rc = vslam_orb_detect_and_compute(o, o->d_frames, batch, p.width, o->d_kp, o->d_desc, o->d_count, s);
    if (rc) return rc;
    VS_CUDA(cudaMemcpyAsync(h_kp, o->d_kp, (size_t)batch * p.kp_cap * sizeof(vslam_keypoint), cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaMemcpyAsync(h_desc, o->d_desc, (size_t)batch * p.kp_cap * 32, cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaMemcpyAsync(h_count, o->d_count, batch * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    VS_CUDA(cudaStreamSynchronize(s));
    return vslam_orb_last_status(o);
}

extern "C" int vslam_orb_debug_run_levels_host(vslam_orb* o, const uint8_t* h_frames, int batch)
{
    VS_REQUIRE(o && h_frames, "null argument");
    VS_REQUIRE(batch >= 1 && batch <= o->prm.max_batch, "batch exceeds max_batch");
    int rc = vs_ensure_host_staging(o);
    if (rc) return rc;
    const VsPlan& p = o->plan;
    cudaStream_t s = o->stream;
    VS_CUDA(cudaMemcpyAsync(o->d_frames, h_frames, (size_t)batch * p.width * p.height, cudaMemcpyHostToDevice, s));
    o->last_stream = s;
    o->last_frames = o->d_frames;
    o->last_pitch = p.width;
    o->last_batch = batch;
    rc = vs_launch_levels(o, o->d_frames, batch, p.width, s);
    if (rc) return rc;
    VS_CUDA(cudaStreamSynchronize(s));
    return VSLAM_OK;
}

extern "C" int vslam_orb_set_profiling(vslam_orb* o, int enable)
{
    VS_REQUIRE(o, "null handle");
    o->prof_on = enable == 2 ? 2 : enable ? 1 : 0;
    o->prof_count = 0;
    return VSLAM_OK;
}

extern "C" int vslam_orb_read_profile(vslam_orb* o, float* h_ms, int n_slots)
{
    VS_REQUIRE(o && h_ms && n_slots >= VS_PROF_SLOTS, "need room for 24 slots");
    for (int i = 0; i < n_slots; ++i) h_ms[i] = 0.f;
    if (o->prof_count < 2) return VSLAM_OK;
    VS_CUDA(cudaEventSynchronize(o->prof_ev[o->prof_count - 1]));
    for (int i = 0; i + 1 < o->prof_count; ++i) {
        if (o->prof_slot[i] < 0) continue;
        float ms = 0.f;
        VS_CUDA(cudaEventElapsedTime(&ms, o->prof_ev[i], o->prof_ev[i + 1]));
        h_ms[o->prof_slot[i]] += ms;
    }
    if (o->prof_on == 2) o->prof_count = 0;
    return VSLAM_OK;
}
