/*
 * vslam_b200.h -- C ABI of the B200-native ORB front end (libvslam_b200.so).
 *
 * This is the drop-in boundary for ONE hot path of p2004dr/visual-slam: ORB
 * extraction -> brute-force Hamming kNN-2 + Lowe ratio -> RANSAC F/E scoring, pose
 * recovery and triangulation.  The reference has no FFI of its own: the path is
 * reached from Python through cv2 calls.  Every entry point below therefore names the
 * reference call site (file:line under /root/reference) whose cv2 call it replaces;
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *   - plain C, no C++/torch types; all functions return 0 on success or a negative
 *     vslam_status; vslam_last_error() returns a thread-local message for the last
 *     failure.  No exceptions, no aborts.
 *   - pointers named d_* are DEVICE pointers (any allocator: cudaMalloc, torch, ...),
 *     h_* are HOST pointers, small fixed-size matrices are host pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *     Device-pointer entry points only enqueue work; *_host entry points copy in,
 *     run, copy out and synchronise the stream before returning.
 *   - there is NO CPU fallback: every entry point needs a CUDA device of compute
 *     capability 10.x and fails with VSLAM_E_CUDA otherwise.
 */
#ifndef VSLAM_B200_H
#define VSLAM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum vslam_status {
    VSLAM_OK = 0,
    VSLAM_E_INVALID = -1,   /* bad argument (null pointer, size out of range, ...)        */
    VSLAM_E_CUDA = -2,      /* CUDA runtime error; text in vslam_last_error()             */
    VSLAM_E_CAPACITY = -3,  /* an output or tie-staging capacity was exceeded             */
    VSLAM_E_NOMEM = -4
} vslam_status;

const char* vslam_last_error(void);
/* ABI version: major*1000 + minor. */
int vslam_abi_version(void);
/* Number of visible CUDA devices (0 when there is no usable driver). */
int vslam_device_count(void);
/* Number of CUDA kernels this library has launched in this process so far. */
unsigned long long vslam_launch_count(void);

/* ------------------------------------------------------------------ extraction */

/* One keypoint as cv2.KeyPoint carries it (extractor.py:65 returns a tuple of these):
 * pt = (x, y) in level-0 pixels, size = 31*scale, angle in degrees [0,360),
 * response = Harris score, octave = pyramid level.  class_id is always -1. */
typedef struct vslam_keypoint {
    float x, y, size, angle, response;
    int32_t octave;
} vslam_keypoint;

/* Parameters of ORBExtractor.__init__ (extractor.py:19-48).  ini_threshold is
 * stored by the reference and never used, so it has no field here.  Fixed by the
 * reference's cv2.ORB_create call: edgeThreshold 31, firstLevel 0, WTA_K 2,
 * HARRIS_SCORE, patchSize 31. */
typedef struct vslam_orb_params {
    int32_t width, height;   /* frame size the extractor is planned for                  */
    int32_t n_features;      /* nfeatures                                                */
    int32_t n_levels;        /* nlevels (1..16)                                          */
    int32_t fast_threshold;  /* fastThreshold = min_threshold                            */
    float scale_factor;      /* scaleFactor                                              */
    int32_t max_batch;       /* frames per launch the workspace is sized for (>= 1)      */
    int32_t kp_capacity;     /* keypoint rows per frame in the outputs; 0 = n_features + 64
                              * (retain-best keeps ties, so a level may exceed its quota) */
} vslam_orb_params;

typedef struct vslam_orb vslam_orb;

int vslam_orb_create(const vslam_orb_params* params, vslam_orb** out);
int vslam_orb_destroy(vslam_orb* orb);
/* Keypoint distribution of the detector.  VSLAM_DIST_CV2_ORB (default) is what the
 * reference runs: cv2.ORB's per-level retain-best by Harris response (extractor.py:38-48).
 * VSLAM_DIST_ORBSLAM2 gives `ini_threshold` (extractor.py:34, stored and unused by the
 * reference) the meaning it has in ORB-SLAM2: per 30-px cell FAST at ini_threshold with
 * fallback to min_threshold, quadtree distribution to the level quota, response = FAST
 * score.  The reference holds no implementation of it (parity unpinned, see
 * oracle/orbslam2_oracle.py); pyramid, blur, orientation and rBRIEF are shared. */
#define VSLAM_DIST_CV2_ORB 0
#define VSLAM_DIST_ORBSLAM2 1
int vslam_orb_set_distribution(vslam_orb* orb, int mode, int ini_threshold);
/* Effective keypoint capacity per frame and pyramid geometry of a handle. */
int vslam_orb_kp_capacity(const vslam_orb* orb);
int vslam_orb_level_info(const vslam_orb* orb, int level, int* w, int* h, int* pitch,
                         float* scale, int* quota);

/* Replaces cv2.ORB.detectAndCompute(image, None) at extractor.py:65 for a batch of gray
 * frames.  d_frames: uint8 [batch][height][pitch].  Outputs (capacity = kp_capacity
 * rows per frame): d_kp [batch][capacity], d_desc uint8 [batch][capacity][32],
 * d_count int32 [batch] (true count; rows beyond capacity are dropped but counted).
 * Order: level-major, ascending (y, x) of the level coordinates inside a level.
 * cv2 leaves the in-level order unspecified (std::nth_element). */
int vslam_orb_detect_and_compute(vslam_orb* orb, const uint8_t* d_frames, int batch,
                                 int pitch, vslam_keypoint* d_kp, uint8_t* d_desc,
                                 int32_t* d_count, void* stream);
/* Same call with host buffers (h_frames tightly packed: pitch = width). */
int vslam_orb_detect_and_compute_host(vslam_orb* orb, const uint8_t* h_frames, int batch,
                                      vslam_keypoint* h_kp, uint8_t* h_desc,
                                      int32_t* h_count);
/* The same with BGR host frames (uint8 [batch][height][width][3], tightly packed), for callers
 * that hold colour frames like Tracker.process_frame (tracker.py:84-87): one upload, the
 * cvtColor(BGR2GRAY) of extractor.py:61-62 runs on the device in front of the pyramid. */
int vslam_orb_detect_and_compute_host_bgr(vslam_orb* orb, const uint8_t* h_bgr, int batch,
                                          vslam_keypoint* h_kp, uint8_t* h_desc,
                                          int32_t* h_count);
/* Replaces cv2.undistort(image, camera_matrix, distortion) (utils.undistort_image,
 * utils.py:40-52; called between gray conversion and the extractor by tests/test_matcher.py:71,
 * test_orb_extractor.py:77, run_video.py:149).  uint8 images with 1 or 3 interleaved channels,
 * [batch][height][pitch]; h_K double [9] row-major, h_dist double [n_dist] with n_dist in
 * {0, 4, 5, 8, 12} (k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4), both on the host.  Bit-exact with
 * OpenCV's fixed-point remap (5 fractional map bits, 15-bit bilinear weights, constant border). */
int vslam_undistort(const uint8_t* d_src, int batch, int height, int width, int channels,
                    int pitch_src, const double* h_K, const double* h_dist, int n_dist,
                    uint8_t* d_dst, int pitch_dst, void* stream);
int vslam_undistort_host(const uint8_t* h_src, int height, int width, int channels,
                         const double* h_K, const double* h_dist, int n_dist, uint8_t* h_dst);
/* Video ingest on the device.  Replaces cv2.VideoCapture(path).read() + the extractor's
 * cvtColor(BGR2GRAY) (src/tests/tester_map.py:58-66, run_video.py:138-149, extractor.py:61-62) for the DECODE and
 * COLOUR part: compressed packets go to the GPU's NVDEC engine through libnvcuvid (dlopen'ed; the driver ships it)
 * and come out as frames in HBM in the form the extractor takes.  The container is still demuxed on the host
 * (cv2.VideoCapture in raw mode hands out the packets: orbslam2/ingest.py).  8-bit 4:2:0 streams.
 *   vslam_nvdec_available   1 when libnvcuvid is loadable, else 0 (then vslam_nvdec_create fails with VSLAM_E_CUDA).
 *   vslam_nvdec_decode      feeds one packet (h_packet may be NULL with bytes 0 and end_of_stream 1 to flush) and
 *                           writes the frames that become displayable, in display order, to d_out + i * frame_stride
 *                           (i < max_frames; *n_frames = how many; more than max_frames -> VSLAM_E_CAPACITY).
 *                           mode VSLAM_NVDEC_LUMA: the Y plane; VSLAM_NVDEC_GRAY: the reference's gray bit for bit
 *                           (swscale yuv420p -> bgr24, then OpenCV's BGR2GRAY); VSLAM_NVDEC_BGR: that BGR frame
 *                           (uint8 [h][w][3], what cv2.VideoCapture.read() returns).  The frames are complete when
 *                           the call returns (the decoder's surface is released at unmap).
 *   vslam_nvdec_set_colour  the 16-bit fixed-point constants of swscale's conversion (defaults: ITU-R BT.601
 *                           limited range, what cv2's FFmpeg backend selects). */
#define VSLAM_CODEC_MPEG4 1
#define VSLAM_CODEC_H264 2
#define VSLAM_CODEC_MJPEG 3
#define VSLAM_CODEC_HEVC 4
#define VSLAM_CODEC_VP8 5
#define VSLAM_CODEC_VP9 6
#define VSLAM_CODEC_AV1 7
#define VSLAM_NVDEC_LUMA 0
#define VSLAM_NVDEC_GRAY 1
#define VSLAM_NVDEC_BGR 2
#define VSLAM_NVDEC_NV12 3   /* the decoder's surface as is: uint8 [h + h/2][w], luma rows then interleaved U, V rows */
typedef struct vslam_nvdec vslam_nvdec;
int vslam_nvdec_available(void);
int vslam_nvdec_create(int codec, vslam_nvdec** out);
int vslam_nvdec_destroy(vslam_nvdec* dec);
int vslam_nvdec_set_colour(vslam_nvdec* dec, int cy, int cbu, int cgu, int cgv, int crv, int y_offset);
int vslam_nvdec_frame_size(const vslam_nvdec* dec, int* width, int* height);
int vslam_nvdec_decode(vslam_nvdec* dec, const uint8_t* h_packet, size_t bytes, int end_of_stream, int mode,
                       uint8_t* d_out, int out_pitch, size_t frame_stride, int max_frames, int* n_frames,
                       void* stream);

/* cv2.cvtColor(COLOR_BGR2GRAY) on device buffers (d_bgr uint8 [n_pixels][3] -> d_gray uint8
 * [n_pixels]); bit-exact with OpenCV's fixed-point formula. */
int vslam_bgr2gray(const uint8_t* d_bgr, size_t n_pixels, uint8_t* d_gray, void* stream);
/* Status of the last extraction on this handle: VSLAM_OK, or VSLAM_E_CAPACITY when
 * ties at a retain-best cut exceeded the staging capacity (results then differ from
 * the reference).  Synchronises the stream of that call. */
int vslam_orb_last_status(vslam_orb* orb);

/* Replaces cv2.ORB.compute(image, keypoints) at extractor.py:83,140: descriptors at
 * caller-supplied keypoints (level-0 coords, own angle in degrees used as given,
 * own octave).  Keypoints outside [31, W-31) x [31, H-31) are dropped; survivors are
 * ordered level-major (stable).  h_desc uint8 [n][32]; h_keep_index int32 [n] receives
 * the input index of each output row.  *n_out = rows written. */
int vslam_orb_compute_host(vslam_orb* orb, const uint8_t* h_frame, const float* h_x,
                           const float* h_y, const float* h_angle, const int32_t* h_octave,
                           int n, uint8_t* h_desc, int32_t* h_keep_index, int* n_out);

/* The same on device buffers: the pyramid of d_frame (uint8 [height][pitch]) is built in the
 * handle's workspace, then row i of d_desc (uint8 [n][32]) is the descriptor of keypoint i of the device arrays
 * d_x / d_y / d_angle (float [n]) / d_octave (int32 [n]), in the caller's order.  d_keep_mask uint8 [n] (may be
 * NULL): 1 where the keypoint passes cv2's border test and its octave is a planned level, else 0 and the row is
 * zero -- the host form above does the compaction and level-major ordering on top of this.  Enqueues only. */
int vslam_orb_compute(vslam_orb* orb, const uint8_t* d_frame, int pitch, const float* d_x, const float* d_y,
                      const float* d_angle, const int32_t* d_octave, int n, uint8_t* d_desc,
                      uint8_t* d_keep_mask, void* stream);

/* Replaces cv2.cvtColor(image, COLOR_BGR2GRAY) at extractor.py:62,81,101. */
int vslam_bgr2gray_host(const uint8_t* h_bgr, int height, int width, uint8_t* h_gray);

/* Replaces the 64 cv2.goodFeaturesToTrack(image, n//64, 0.01, 10, mask=cell) calls of
 * ORBExtractor.distribute_keypoints (extractor.py:115-129): one min-eigenvalue map,
 * per-cell quality threshold, 3x3 local maxima, greedy min-distance selection.
 * h_xy float [rows*cols*max_per_cell][2] in the reference's visiting order (cells
 * row-major, corners in acceptance order); h_cell_count int32 [rows*cols]. */
int vslam_gftt_grid_host(const uint8_t* h_frame, int height, int width, int rows, int cols,
                         int max_per_cell, double quality, double min_distance,
                         float* h_xy, int32_t* h_cell_count, int* n_out);

/* The same on a device frame (uint8 [height][width], tightly packed): d_xy float [rows*cols][max_per_cell][2]
 * holds cell c's corners in acceptance order at row c (not compacted), d_cell_count int32 [rows*cols]; d_status
 * (1 int32, may be NULL) becomes non-zero when a cell held more local maxima than the candidate staging.  The
 * min-eigenvalue maps live in per-thread scratch of the library.  Enqueues only. */
int vslam_gftt_grid(const uint8_t* d_frame, int height, int width, int rows, int cols, int max_per_cell,
                    double quality, double min_distance, float* d_xy, int32_t* d_cell_count, int32_t* d_status,
                    void* stream);

/* Per-kernel timing for bench.py's roofline figures: when enabled, CUDA events are
 * recorded on the launching stream between the kernels of every call on this handle.
 * vslam_orb_read_profile waits for the last call and returns its milliseconds per slot:
 * [0..15] fused level kernel per pyramid level, [16] Harris, [17] select, [18] describe,
 * [19] kNN-2 (+ merge), [20] ratio filter; n_slots must be >= 24.
 * enable = 2: the marks of successive calls accumulate (up to 8192 marks, ~340 front-end
 * calls) and vslam_orb_read_profile returns the sums since the last read -- a timed loop
 * then needs no host synchronisation per call. */
int vslam_orb_set_profiling(vslam_orb* orb, int enable);
int vslam_orb_read_profile(vslam_orb* orb, float* h_ms, int n_slots);

/* Debug/test taps used by the stage-wise parity tests: copy pyramid level `level`
 * of frame `frame` of the last batch (raw or blurred) to the host, tightly packed. */
int vslam_orb_debug_level(vslam_orb* orb, int frame, int level, int blurred, uint8_t* h_out);
/* Runs only the per-level kernels (pyramid, FAST + NMS, blur) on host frames so that the
 * candidate lists still hold FAST scores when they are read back. */
int vslam_orb_debug_run_levels_host(vslam_orb* orb, const uint8_t* h_frames, int batch);
/* FAST candidates (after NMS and the 31-px border) of one frame/level of the last
 * batch, unordered: x, y, score.  Returns the count through *n_out. */
int vslam_orb_debug_candidates(vslam_orb* orb, int frame, int level, int cap, int32_t* h_x,
                               int32_t* h_y, int32_t* h_score, int* n_out);

/* -------------------------------------------------------------------- matching */

/* Replaces cv2.BFMatcher(NORM_HAMMING).knnMatch(d1, d2, k=2) at matcher.py:70 (and
 * local_mapper.py:116-117) for a batch of descriptor-set pairs.
 * d_query  uint8 [pairs][q_stride_rows][32], d_train uint8 [pairs][t_stride_rows][32].
 * Row counts: nq / nt for every pair, or per pair from the device arrays d_nq / d_nt
 * when non-NULL (values are clamped to nq / nt, which then act as maxima).
 * Outputs per query row: d_idx int32 [pairs][nq][2], d_dist int32 [pairs][nq][2];
 * missing neighbours are (-1, -1).  Ties -> lowest train index (1st and 2nd place). */
int vslam_bf_knn2_hamming256(const uint8_t* d_query, int nq, long q_stride_rows,
                             const uint8_t* d_train, int nt, long t_stride_rows,
                             const int32_t* d_nq, const int32_t* d_nt, int pairs,
                             int32_t* d_idx, int32_t* d_dist, void* stream);
int vslam_bf_knn2_hamming256_host(const uint8_t* h_query, int nq, const uint8_t* h_train,
                                  int nt, int32_t* h_idx, int32_t* h_dist);

/* No counterpart in the reference: added for callers that
 * cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * kNN-2 of ONE descriptor-set pair with the outputs of vslam_bf_knn2_hamming256 (d_idx / d_dist
 * int32 [nq][2], same tie rule), computed by the split-train XOR + POPC kernel with its partial
 * results in a static __device__ array of the library (2 MiB, held by every GPU that loads the
 * library until exit) -- the call makes no cudaMalloc / cudaFree / allocating library call.
 * Limit: nq <= VSLAM_NOALLOC_MAX_QUERY (VSLAM_E_CAPACITY beyond, nothing launched); nt is unlimited.
 * All calls share the scratch: calls that overlap on the GPU (two streams) race -- serialise them.
 * The first call on a device looks the array's address up (cudaGetSymbolAddress); make one call
 * before starting a graph capture. */
#define VSLAM_NOALLOC_MAX_QUERY 8192
int vslam_bf_knn2_hamming256_noalloc(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                                     int32_t* d_idx, int32_t* d_dist, void* stream);

/* Replaces the Lowe ratio loop of DescriptorMatcher.match (matcher.py:73-81):
 * keep row q iff two neighbours exist and (!ratio_test or
 * (double)(float)d1 < ratio*(double)(float)d2), or exactly one neighbour exists.
 * Survivors are compacted in ascending query order into d_match int32 [pairs][nq][3]
 * = (queryIdx, trainIdx, distance) and counted in d_match_count int32 [pairs]. */
int vslam_ratio_filter(const int32_t* d_idx, const int32_t* d_dist, int nq,
                       const int32_t* d_nq, int pairs, double ratio, int ratio_test,
                       int32_t* d_match, int32_t* d_match_count, void* stream);

/* Replaces cv2.BFMatcher.match(d1, d2, mask=mask) at matcher.py:102 (k = 1;
 * h_mask uint8 [nq][nt] non-zero = allowed, may be NULL). Rows with no allowed
 * train descriptor get (-1, -1). */
int vslam_bf_match_masked_host(const uint8_t* h_query, int nq, const uint8_t* h_train, int nt,
                               const uint8_t* h_mask, int32_t* h_idx, int32_t* h_dist);

/* ------------------------------------------------------------------- front end */

/* The BASELINE.json hot path as one call: extract every frame of the batch, then
 * match frame i (query) against frame i+1 (train) for i = 0..batch-2 with the ratio
 * test.  Device buffers; outputs as in the calls above plus d_match
 * int32 [batch-1][kp_capacity][3] and d_match_count int32 [batch-1].
 * If d_prev_desc is non-NULL it is the descriptor block (kp_capacity rows, count
 * *d_prev_count) of the frame preceding frame 0 (the previous batch's or the
 * neighbouring rank's boundary frame); the outputs then hold `batch` pairs and
 * pair 0 is (prev, frame 0).
 * Small batches (<= 8 frames: the per-frame call of Tracker.process_frame,
 * tracker.py:87,214) on a stream other than the legacy default stream are captured
 * into a CUDA graph the second time an argument set is seen and replayed afterwards
 * (four argument sets per handle, e.g. ping-pong descriptor buffers); results are
 * those of the plain launches.  VSLAM_NO_GRAPH=1 turns this off. */
int vslam_frontend_run(vslam_orb* orb, const uint8_t* d_frames, int batch, int pitch,
                       const uint8_t* d_prev_desc, const int32_t* d_prev_count, double ratio,
                       vslam_keypoint* d_kp, uint8_t* d_desc, int32_t* d_count,
                       int32_t* d_match, int32_t* d_match_count, void* stream);
int vslam_frontend_run_host(vslam_orb* orb, const uint8_t* h_frames, int batch, double ratio,
                            vslam_keypoint* h_kp, uint8_t* h_desc, int32_t* h_count,
                            int32_t* h_match, int32_t* h_match_count);

/* -------------------------------------------------------------------- geometry */

/* Scoring kernel of cv2.findFundamentalMat(FM_RANSAC) (matcher.py:192,
 * local_mapper.py:136): inlier iff max(d1^2, d2^2) <= thr^2 with the two squared
 * point-to-epipolar-line distances.  h_F row-major double[9]; points float [n][2]. */
int vslam_score_fundamental_host(const double* h_F, const float* h_p1, const float* h_p2, int n,
                                 double thr, uint8_t* h_mask, int* n_inliers);
/* Scoring kernel of cv2.findEssentialMat(RANSAC) (utils.py:124-125, tracker.py:242):
 * Sampson error in K-normalised coordinates against (thr / ((fx+fy)/2))^2. */
int vslam_score_essential_host(const double* h_E, const double* h_K, const float* h_p1,
                               const float* h_p2, int n, double thr, uint8_t* h_mask,
                               int* n_inliers);
/* Device-pointer forms of the two scoring calls (points float [n][2] on the device, mask uint8 [n] on the device or
 * NULL; the models are host pointers).  n_inliers (host, may be NULL) makes the call wait for the stream. */
int vslam_score_fundamental(const double* h_F, const float* d_p1, const float* d_p2, int n, double thr,
                            uint8_t* d_mask, int* n_inliers, void* stream);
int vslam_score_essential(const double* h_E, const double* h_K, const float* d_p1, const float* d_p2, int n,
                          double thr, uint8_t* d_mask, int* n_inliers, void* stream);
/* Replaces cv2.findFundamentalMat(p1, p2, FM_RANSAC, thr, conf) (matcher.py:192):
 * 7-point hypotheses evaluated in parallel on the subsets OpenCV itself draws (its RNG(-1)
 * stream with duplicate and collinearity rejection), winner and stopping point by a replay of
 * its sequential loop, so the mask is the one cv2 returns; below 15 points cv2 runs LMedS
 * and so does this call.  h_F receives the winning minimal solution scaled to F[8] = 1
 * (all zeros when there is no model).  `seed` (here and below): 0 = OpenCV's own stream --
 * what the reference runs, cv2's estimators ignore cv2.setRNGSeed; any other value selects
 * another deterministic stream. */
int vslam_ransac_fundamental_host(const float* h_p1, const float* h_p2, int n, double thr,
                                  double conf, int max_iters, uint64_t seed, double* h_F,
                                  uint8_t* h_mask, int* n_inliers);
/* Batched, device-resident form of DescriptorMatcher.filter_matches_by_fundamental
 * (matcher.py:171-200) for the output of vslam_frontend_run: pair p takes its points from
 * d_match[p] (int32 [pairs][match_capacity][3], d_match_count[p] rows) as
 * d_kp_query[p * kp_capacity + queryIdx] and d_kp_train[p * kp_capacity + trainIdx]
 * (for consecutive frames of one keypoint array: d_kp_train = d_kp_query + kp_capacity).
 * One CTA per pair; pair p gives exactly what vslam_ransac_fundamental_host gives on
 * its point lists.  Outputs: d_F double [pairs][9], d_mask uint8 [pairs][match_capacity]
 * (row i <-> match i), d_n_inliers int32 [pairs].  Pairs with fewer than 8 matches
 * follow matcher.py:184-185: mask rows 1, d_n_inliers = -1, F = 0.  d_workspace:
 * vslam_ransac_fundamental_batched_workspace(pairs, match_capacity) bytes. */
size_t vslam_ransac_fundamental_batched_workspace(int pairs, int match_capacity);
int vslam_ransac_fundamental_batched(const vslam_keypoint* d_kp_query,
                                     const vslam_keypoint* d_kp_train, int kp_capacity,
                                     const int32_t* d_match, const int32_t* d_match_count,
                                     int match_capacity, int pairs, double thr, double conf,
                                     int max_iters, uint64_t seed, void* d_workspace,
                                     double* d_F, uint8_t* d_mask, int32_t* d_n_inliers,
                                     void* stream);
/* Replaces cv2.findEssentialMat(p1, p2, K, RANSAC, prob, thr) (utils.py:124-125). */
int vslam_ransac_essential_host(const float* h_p1, const float* h_p2, int n, const double* h_K,
                                double prob, double thr, int max_iters, uint64_t seed,
                                double* h_E, uint8_t* h_mask, int* n_inliers);
/* The same with the points and the mask on the device; E and the inlier count are host results, so
 * the call waits for the stream.  (vslam_track_batched is the fully device-resident, batched form.) */
int vslam_ransac_essential(const float* d_p1, const float* d_p2, int n, const double* h_K, double prob, double thr,
                           int max_iters, uint64_t seed, double* h_E, uint8_t* d_mask, int* n_inliers,
                           void* stream);
/* Replaces cv2.recoverPose(E, p1, p2, K, mask) (utils.py:133, tracker.py:249).
 * h_mask_inout uint8 [n] (NULL = all ones on input, no output). */
int vslam_recover_pose_host(const double* h_E, const float* h_p1, const float* h_p2, int n,
                            const double* h_K, uint8_t* h_mask_inout, double* h_R, double* h_t,
                            int* n_good);
/* Device-pointer form: points float [n][2] and d_mask_inout uint8 [n] (may be NULL) on the device; R, t and the
 * count are host results (the winner among the four candidates is picked on the host: the call waits for the
 * stream once), the mask is updated on the stream. */
int vslam_recover_pose(const double* h_E, const float* d_p1, const float* d_p2, int n, const double* h_K,
                       uint8_t* d_mask_inout, double* h_R, double* h_t, int* n_good, void* stream);
/* Batched, device-resident form of the reference's per-frame tracking step
 * (Tracker._track_frame, tracker.py:214-254) for the output of vslam_frontend_run: pair p
 * takes its match list d_match[p] (layout as above) through
 *   filter_matches_by_geometric_distance(threshold_percent, (image_height, image_width))   matcher.py:109-142
 *   filter_matches_by_distance()   (stable sort by distance, keep < 2 x median)          matcher.py:144-169
 *   fewer than 8 left -> failure                                                          tracker.py:233
 *   cv2.findEssentialMat(p1, p2, K, RANSAC, prob, thr)                                    tracker.py:242
 *   cv2.recoverPose(E, p1, p2, K, mask)                                                   tracker.py:249
 * One CTA per pair; the RANSAC and pose stages give what vslam_ransac_essential_host /
 * vslam_recover_pose_host give on the gated point lists.  Outputs per pair: d_track_match
 * int32 [match_capacity][3] + d_track_count (the gated matches in the reference's order,
 * ascending distance, stable), d_E double [9], d_ransac_mask / d_pose_mask uint8
 * [match_capacity] (row i <-> gated match i), d_R double [9], d_t double [3], d_status
 * int32 [4] = {1 ok | 0 fewer than 8 gated matches | -1 no essential matrix, points passing
 * the cheirality check, RANSAC inliers, displacement decisions within 1e-9 of the limit
 * (non-zero: the caller re-decides that pair on the host with math.hypot)}.  h_K double [9]
 * row-major (host).  d_workspace: vslam_track_batched_workspace(pairs, match_capacity) bytes. */
size_t vslam_track_batched_workspace(int pairs, int match_capacity);
int vslam_track_batched(const vslam_keypoint* d_kp_query, const vslam_keypoint* d_kp_train,
                        int kp_capacity, const int32_t* d_match, const int32_t* d_match_count,
                        int match_capacity, int pairs, const double* h_K, int image_width,
                        int image_height, double threshold_percent, double prob, double thr,
                        int max_iters, uint64_t seed, void* d_workspace, int32_t* d_track_match,
                        int32_t* d_track_count, double* d_E, uint8_t* d_ransac_mask,
                        uint8_t* d_pose_mask, double* d_R, double* d_t, int32_t* d_status,
                        void* stream);
/* A clip in, a trajectory out: the reference's per-frame loop (Tracker.process_frame,
 * tracker.py:73-146) for `batch` consecutive gray host frames (2 <= batch <= max_batch):
 * one upload, extraction, matching and vslam_track_batched on the device, and per frame
 * pair only the relative pose h_T double [batch-1][16] (row-major 4x4, T = [R t; 0 0 0 1]
 * as tracker.py:250-253; identity where the pair failed) and h_status int32 [batch-1][4]
 * (as vslam_track_batched) come back.  h_count int32 [batch] (keypoints per frame) may be
 * NULL.  The device-side results of the call (keypoints, descriptors, match lists) stay in
 * the handle's staging buffers. */
int vslam_track_clip_host(vslam_orb* orb, const uint8_t* h_frames, int batch, double ratio,
                          const double* h_K, double threshold_percent, double prob, double thr,
                          int max_iters, uint64_t seed, double* h_T, int32_t* h_status,
                          int32_t* h_count);
/* Replaces cv2.triangulatePoints(P1, P2, p1.T, p2.T) (utils.py:59): per-point DLT.
 * h_X float [4][n] (row-major, un-normalised homogeneous, as cv2 returns it). */
int vslam_triangulate_host(const double* h_P1, const double* h_P2, const float* h_p1,
                           const float* h_p2, int n, float* h_X);
/* Device-pointer form: d_X float [4][n] on the device.  Enqueues only. */
int vslam_triangulate(const double* h_P1, const double* h_P2, const float* d_p1, const float* d_p2, int n,
                      float* d_X, void* stream);
/* Replaces the per-point loop of LocalMapper._cull_map_points (local_mapper.py:205-237): a map
 * point is kept when it has at least min_observations observations and, walking them in stored
 * order, none has a reprojection error above max_error (5.0 in the reference):
 *   P = K [R | t] of the observing keyframe (utils.compute_projection_matrix, utils.py:137-159),
 *   proj = P [X 1];  err = |proj_xy / proj_z - kp.pt| in float64;  `err > 5.0` rejects (NaN does not).
 * The map as flat arrays: points double [n_points][3]; observation lists in CSR form --
 * obs_offset int32 [n_points + 1], obs_keyframe int32 [m] (index into P), obs_xy float [m][2]
 * (cv2.KeyPoint.pt of the observed keypoint); P double [n_keyframes][12] row-major 3x4.
 * valid uint8 [n_points]: 1 kept, 0 dropped, VSLAM_CULL_BAD_KEYFRAME for an observation whose
 * keyframe index is outside [0, n_keyframes) (the host call returns VSLAM_E_INVALID then, where
 * the reference raises IndexError).  err double [n_points] (may be NULL): the largest error
 * among the observations the reference looks at (it stops at the first one above the gate). */
#define VSLAM_CULL_BAD_KEYFRAME 255
int vslam_cull_map_points(const double* d_points, const int32_t* d_obs_offset,
                          const int32_t* d_obs_keyframe, const float* d_obs_xy, const double* d_P,
                          int n_points, int n_keyframes, int min_observations, double max_error,
                          uint8_t* d_valid, double* d_err, void* stream);
int vslam_cull_map_points_host(const double* h_points, const int32_t* h_obs_offset,
                               const int32_t* h_obs_keyframe, const float* h_obs_xy,
                               const double* h_P, int n_points, int n_keyframes,
                               int min_observations, double max_error, uint8_t* h_valid,
                               double* h_err);

#ifdef __cplusplus
}
#endif
#endif /* VSLAM_B200_H */
