/*
 * vxorb.h — C ABI of the B200-native ORB extract + Hamming match hot path.
 *
 * This is the drop-in boundary for VisionX-SLAM's front-end feature path.  Every entry
 * point is `extern "C"`, takes plain pointers and sizes, returns an int status
 * (0 = VXORB_OK, negative = error; text via vxorb_last_error) and never throws or aborts.
 * Host buffers are caller-owned; the library owns every byte of device memory.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   core/feature/feature_extractor.h:10-16   FeatureExtractor::Extract(Frame&)
 *   core/feature/orb_extractor.cpp:5-7       cv::ORB::create(n_features, scale, levels)
 *   core/feature/orb_extractor.cpp:13        cv::ORB::detectAndCompute(image, noArray, kps, desc)
 *   core/feature/feature_matcher.h:7-13      FeatureMatcher::Match(last, curr, matches)
 *   core/feature/orb_matcher.cpp:22-25       cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, knn, 2)
 *   core/feature/orb_matcher.cpp:27-36       ratio test  m1.distance < nn_ratio * m2.distance
 *   core/frontend/tracking.cpp:211-222       caller's post-filter  d <= max(2*min_dist, 30)
 *   core/frontend/tracking.cpp:93-118        CheckFeatureDistribution (5x5 grid occupancy)
 *   core/frontend/tracking.cpp:120-139       CheckImageQuality (cv::meanStdDev of the gray frame)
 *   core/frontend/tracking.cpp:505-514       matched point pairs pts_last / pts_curr
 *   core/frontend/tracking.cpp:548-560       ComputeParallax (mean distance of matched points)
 *
 * Output order is canonical: keypoints (and descriptor rows) are sorted by
 * (octave asc, y asc, x asc) in level coordinates; matches by ascending query index.
 * A handle is not thread-safe; distinct handles (one per GPU / host thread) are independent.
 */
#ifndef VXORB_H_
#define VXORB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VXORB_OK 0
#define VXORB_E_INVALID (-1)  /* bad argument (null pointer, size out of the handle's limits)   */
#define VXORB_E_CUDA (-2)     /* CUDA runtime / launch failure; see vxorb_last_error            */
#define VXORB_E_CAPACITY (-3) /* an output or internal list was too small; counts are still set */
#define VXORB_E_NOMEM (-4)    /* host or device allocation failed                               */

#define VXORB_DESC_BYTES 32   /* 256-bit rBRIEF descriptor                                      */
#define VXORB_MAX_LEVELS 16

/* Parameters of one extractor+matcher instance.  Defaults (vxorb_default_params) are the
 * reference's: ORBExtractor(1000, 1.2f, 8) (orb_extractor.h:11-13), the remaining cv::ORB
 * defaults (edge 31, patch 31, FAST threshold 20, Harris score, WTA_K 2) and
 * ORBMatcher::Options{0.8f, 50} (orb_matcher.h:13-16). */
typedef struct vxorb_params {
    int n_features;     /* 1000 */
    float scale_factor; /* 1.2f */
    int n_levels;       /* 8    */
    int edge_threshold; /* 31; only 31 is supported (tied to patch_size)                        */
    int patch_size;     /* 31; only 31 is supported (pattern + umax table are for 31)           */
    int fast_threshold; /* 20   */
    float nn_ratio;     /* 0.8f */
    int min_matches;    /* 50; advisory (the adapter logs a warning below it)                   */
    int device_id;      /* CUDA device ordinal                                                  */
    int max_width;      /* largest frame the handle accepts (device buffers are sized for it);
                           at most 4095 x 4095: corner positions travel as 12-bit fields          */
    int max_height;
    int max_batch;      /* largest number of frames per batched call                            */
    int max_keypoints;  /* per-frame output capacity; 0 = n_features + n_features/4 + 64        */
    int max_train;      /* sizing hint for the match scratch (it grows on demand; < 2^20 rows per side) */
} vxorb_params;

/* One keypoint as cv::ORB reports it (orb_extractor.cpp:19-24 keeps pt and response). */
typedef struct vxorb_keypoint {
    float x, y;     /* level-0 pixel coordinates: level position * scale(octave)              */
    float response; /* Harris response                                                         */
    float angle;    /* degrees in [0,360)                                                      */
    int octave;     /* pyramid level                                                           */
} vxorb_keypoint;

/* One accepted match; mirrors cv::DMatch{queryIdx, trainIdx, imgIdx=0, distance}. */
typedef struct vxorb_match_t {
    int query_idx;
    int train_idx;
    int distance; /* Hamming distance, 0..256 */
} vxorb_match_t;

/* Raw 2-NN row; idx = -1 where fewer than two neighbours exist (cv knn row shorter than 2). */
typedef struct vxorb_knn2_t {
    int idx[2];
    int dist[2];
} vxorb_knn2_t;

/* One FAST corner at level resolution (stage-isolation output). */
typedef struct vxorb_corner {
    int x, y, score;
} vxorb_corner;

/* Per-frame quantities the tracking front end derives from the extractor / matcher outputs
 * (tracking.cpp:93-139, 548-560), computed on the device next to the data. */
typedef struct vxorb_frame_info {
    double mean;       /* cv::meanStdDev(gray)[0]: S * (1/N)                                    */
    double stddev;     /* cv::meanStdDev(gray)[1]: sqrt(max(Q * (1/N) - mean^2, 0)), population */
    double parallax;   /* mean |p_last - p_curr| over the accepted matches (previous frame ->
                          this frame) in level-0 pixels; 0 when there are none                  */
    int grid_occupied; /* cells of the 5x5 grid holding at least one keypoint                   */
    unsigned grid_mask; /* bit (col * 5 + row), col = (int)((x / width) * 5) clamped, row alike  */
    int n_keypoints;   /* keypoints written for the frame (<= capacity)                          */
    int n_matches;     /* accepted matches of the pair (previous frame -> this frame)            */
} vxorb_frame_info;

typedef struct vxorb_handle vxorb_handle;

void vxorb_default_params(vxorb_params* p);
int vxorb_create(const vxorb_params* p, vxorb_handle** out);
void vxorb_destroy(vxorb_handle* h);
/* Message of the last failing call on this handle (or of the last failed vxorb_create when
 * h == NULL).  Never NULL. */
const char* vxorb_last_error(const vxorb_handle* h);
const char* vxorb_version(void);

/* Geometry helpers (pure host arithmetic, no device work): level sizes, scales and per-level
 * keypoint quotas exactly as cv::ORB derives them. */
int vxorb_level_geometry(const vxorb_handle* h, int width, int height, int* level_w, int* level_h,
                         float* level_scale, int* level_quota);

/* ---- extraction: replaces cv::ORB::detectAndCompute (orb_extractor.cpp:13) ----------------- */

/* One 8-bit gray frame, `stride` bytes per row.  Writes up to kp_cap keypoints and kp_cap*32
 * descriptor bytes in canonical order; *n_out = number of keypoints found.  If that exceeds
 * kp_cap or the handle's max_keypoints the call returns VXORB_E_CAPACITY after writing the
 * first min(...) entries.  Blank / too-small images give *n_out = 0 and VXORB_OK. */
int vxorb_extract(vxorb_handle* h, const uint8_t* gray, int width, int height, size_t stride,
                  int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out);

/* Same for a 3-channel BGR frame (system.cpp:46 reads IMREAD_COLOR); the gray conversion
 * (R*9798 + G*19235 + B*3735 + 2^14) >> 15 is fused into the level-0 load. */
int vxorb_extract_bgr(vxorb_handle* h, const uint8_t* bgr, int width, int height, size_t stride,
                      int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out);

/* vxorb_extract plus the frame's image statistics and grid coverage (CheckImageQuality and
 * CheckFeatureDistribution of tracking.cpp:93-139 evaluated on the device); parallax and n_matches
 * are 0.  channels = 1 (gray) or 3 (BGR, converted as vxorb_extract_bgr does). */
int vxorb_extract_info(vxorb_handle* h, const uint8_t* img, int channels, int width, int height,
                       size_t stride, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out,
                       vxorb_frame_info* info_out);

/* vxorb_extract_info (info_out may be NULL here) that also keeps the frame's descriptors and positions ON THE
 * DEVICE under *id_out, so that ORBMatcher::Match(last, curr) (tracking.cpp:289, 340, 863) can run on two ids
 * without re-uploading what Extract just produced.  Every single-frame extract call takes the next entry of a
 * ring of 8 resident frames; an id stays valid until 8 more single-frame extractions have run on the handle. */
int vxorb_extract_id(vxorb_handle* h, const uint8_t* img, int channels, int width, int height, size_t stride,
                     int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out,
                     vxorb_frame_info* info_out, long long* id_out);

/* `batch` independent gray frames of identical size.  frames[i] points at frame i.  Outputs are
 * frame-major: kp_out[i*kp_cap + k], desc_out[(i*kp_cap + k)*32], n_out[i]. */
int vxorb_extract_batch(vxorb_handle* h, const uint8_t* const* frames, int batch, int width,
                        int height, size_t stride, int kp_cap, vxorb_keypoint* kp_out,
                        uint8_t* desc_out, int* n_out);

/* ---- matching: replaces cv::BFMatcher(NORM_HAMMING).knnMatch(...,2) + ratio --------------- */

/* Raw 2-NN table, one row per query, ties broken by lowest train index. */
int vxorb_knn2(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt,
               vxorb_knn2_t* knn_out);

/* ORBMatcher::Match semantics (orb_matcher.cpp:11-43): 2-NN, keep the best iff a second
 * neighbour exists and d1 < nn_ratio*d2 (float compare, as the reference), ascending query
 * order.  nn_ratio <= 0 uses the handle's.  Either side empty gives *n_out = 0, VXORB_OK. */
int vxorb_match(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt,
                float nn_ratio, int out_cap, vxorb_match_t* out, int* n_out);

/* Windowed variants = cv knnMatch with mask[i][j] = |xq_i - xt_j| <= radius && |yq_i - yt_j|
 * <= radius (float compares on level-0 coordinates).  xy arrays are interleaved (x,y) floats. */
int vxorb_knn2_windowed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq,
                        const uint8_t* train, const float* train_xy, int nt, float radius,
                        vxorb_knn2_t* knn_out);
int vxorb_match_windowed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq,
                         const uint8_t* train, const float* train_xy, int nt, float radius,
                         float nn_ratio, int out_cap, vxorb_match_t* out, int* n_out);

/* Box-masked variants (BASELINE config 3, stereo left-right matching along the epipolar band) =
 * cv knnMatch with mask[i][j] = box[0] <= xq_i - xt_j <= box[1] && box[2] <= yq_i - yt_j <= box[3]
 * (float differences and compares).  box = {dx_min, dx_max, dy_min, dy_max}; the square window of
 * radius r is the box {-r, r, -r, r}.  For a rectified pair with the left image as query:
 * {0, max_disparity, -dy_tol, dy_tol}. */
int vxorb_knn2_boxed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq,
                     const uint8_t* train, const float* train_xy, int nt, const float* box,
                     vxorb_knn2_t* knn_out);
int vxorb_match_boxed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq,
                      const uint8_t* train, const float* train_xy, int nt, const float* box,
                      float nn_ratio, int out_cap, vxorb_match_t* out, int* n_out);

/* Arbitrary mask = cv::BFMatcher::knnMatch(query, train, 2, mask): mask is an nq x nt byte matrix (mask_stride bytes
 * per row), pair (i, j) is allowed iff mask[i][j] != 0.  The general (and slowest: the mask crosses the link) form of
 * the window and box variants above. */
int vxorb_knn2_masked(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt,
                      const uint8_t* mask, size_t mask_stride, vxorb_knn2_t* knn_out);
int vxorb_match_masked(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt,
                       const uint8_t* mask, size_t mask_stride, float nn_ratio, int out_cap, vxorb_match_t* out,
                       int* n_out);

/* vxorb_knn2 for callers that may not allocate.  No counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same rows
 * as vxorb_knn2 (brute force, ties to the lowest train index), host buffers in and out, no handle.  The device staging
 * and the kernel's partial-key scratch live in one static __device__ array (~1.7 MiB, held by every GPU that loads
 * the library until the process exits); the call runs on the current device's default stream and calls no cudaMalloc,
 * cudaFree or allocating library routine.  Limits: nq <= VXORB_STATIC_MAX_QUERY and nt <= VXORB_STATIC_MAX_TRAIN,
 * beyond which it returns VXORB_E_INVALID with the limits spelled out in `err` (NUL-terminated, at most err_cap
 * bytes; err may be NULL).  All calls share the array: calls that overlap on one GPU race, so serialise them. */
#define VXORB_STATIC_MAX_QUERY 4096
#define VXORB_STATIC_MAX_TRAIN 16384
int vxorb_knn2_static(const uint8_t* query, int nq, const uint8_t* train, int nt, vxorb_knn2_t* knn_out,
                      char* err, size_t err_cap);

/* cv::BFMatcher(NORM_HAMMING, crossCheck=true).match: (i, j) kept iff j is i's nearest train
 * row and i is j's nearest query row, both with lowest-index tie-break. */
int vxorb_match_crosscheck(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train,
                           int nt, int out_cap, vxorb_match_t* out, int* n_out);

/* vxorb_match followed by the caller's post-filter (tracking.cpp:211-222, 292-304, 343-355):
 * min_dist = min(100, min distance); keep distance <= max(2*min_dist, 30). */
int vxorb_match_filtered(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train,
                         int nt, float nn_ratio, int out_cap, vxorb_match_t* out, int* n_out);

/* ---- device-resident descriptor sets ------------------------------------------------------------
 * vxorb_set_upload keeps n descriptor rows (and, optionally, their level-0 positions: xy may be NULL) on the
 * device under *id_out until vxorb_set_free — a local map (BASELINE config 4) or a keyframe.  vxorb_match_ids
 * matches two resident sets (frame ids from vxorb_extract_id and / or uploaded sets) with the semantics of
 * vxorb_match (mode 0), vxorb_match_filtered (mode 1) or vxorb_match_crosscheck (mode 2); radius > 0 applies
 * the square window of vxorb_match_windowed.  Only the matches cross the link.  An id that is unknown, freed
 * or overwritten gives VXORB_E_INVALID: fall back to the host-buffer calls. */
int vxorb_set_upload(vxorb_handle* h, const uint8_t* desc, const float* xy, int n, long long* id_out);
int vxorb_set_free(vxorb_handle* h, long long id);
/* Rows of a resident set, or VXORB_E_INVALID when the id is not (or no longer) resident. */
int vxorb_set_size(const vxorb_handle* h, long long id);
int vxorb_match_ids(vxorb_handle* h, long long query_id, long long train_id, float radius, float nn_ratio,
                    int mode, int out_cap, vxorb_match_t* out, int* n_out);

/* Depth of every feature of a resident frame as Tracking::CreateLandmarksFromDepth reads it
 * (tracking.cpp:606-637): u = (int)(x + 0.5), v = (int)(y + 0.5); depth_m_out[i] = depth in metres (16-bit / 5000,
 * float or double image: depth_type 0 / 1 / 2), 0.0 where the reference skips the feature (outside the image,
 * 16-bit depth 0, metres outside [0.1, 10]).  *n_out = number of features of the frame. */
int vxorb_depth_lookup(vxorb_handle* h, long long frame_id, const void* depth, int depth_type, int width, int height,
                       size_t stride, int cap, double* depth_m_out, int* n_out);

/* ---- streaming sequence path (BASELINE config 2): batched extract + frame-to-frame match ---
 *
 * A sequence is fed in chunks of `batch` consecutive frames.  Every frame is extracted; frame t
 * is the train set and frame t-1 the query set of a windowed (radius > 0) or brute-force
 * (radius <= 0) ratio-tested match, which is ORBMatcher::Match(last, curr) for every
 * consecutive pair.  The last frame of a chunk stays on the device as the `last` of the next
 * chunk; vxorb_track_reset forgets it (the first frame after a reset has no matches).
 *
 * vxorb_track_submit copies the frames from host memory and queues the work on one of the
 * handle's pipeline slots (one CUDA stream each); vxorb_track_wait blocks until that slot is
 * done.  Up to vxorb_track_slots(h) submits may be in flight, so copies of one chunk overlap
 * kernels of another.  Outputs are frame-major with strides kp_cap / match_cap. */
int vxorb_track_slots(const vxorb_handle* h);
int vxorb_track_reset(vxorb_handle* h);
/* Output pointers are given at submit time (asynchronous I/O style) and must stay valid until the
 * matching vxorb_track_wait returns; any of them may be NULL to skip that copy.  kp_cap and
 * match_cap must equal the handle's max_keypoints (vxorb_keypoint_capacity).  n_kp_out[i] is the
 * true keypoint count of frame i (rows beyond the capacity are dropped and the wait returns
 * VXORB_E_CAPACITY).  For full copy/compute overlap pass pinned memory (vxorb_host_alloc). */
int vxorb_track_submit(vxorb_handle* h, const uint8_t* frames, int batch, int width, int height,
                       size_t stride, size_t frame_stride, float radius, int kp_cap,
                       vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_kp_out, int match_cap,
                       vxorb_match_t* match_out, int* n_match_out, int* slot_out);
/* vxorb_track_submit plus, per frame, vxorb_frame_info (info_out[batch]) and, per accepted match,
 * the matched positions pts_out[(i*match_cap + k)*4 .. +3] = (x_last, y_last, x_curr, y_curr) in
 * level-0 pixels — the pts_last / pts_curr arrays of tracking.cpp:505-514.  Either may be NULL. */
int vxorb_track_submit_info(vxorb_handle* h, const uint8_t* frames, int batch, int width, int height,
                            size_t stride, size_t frame_stride, float radius, int kp_cap,
                            vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_kp_out, int match_cap,
                            vxorb_match_t* match_out, int* n_match_out, vxorb_frame_info* info_out,
                            float* pts_out, int* slot_out);
/* vxorb_track_submit_info for 3-channel frames in BGR byte order, what System::Run hands to Tracking
 * (system.cpp:46) and cv::ORB converts internally (orb_extractor.cpp:13): stride >= 3 * width bytes per
 * row; the chunk is converted on the device with the conversion of vxorb_extract_bgr.  info_out and
 * pts_out may be NULL. */
int vxorb_track_submit_bgr(vxorb_handle* h, const uint8_t* bgr_frames, int batch, int width, int height,
                           size_t stride, size_t frame_stride, float radius, int kp_cap,
                           vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_kp_out, int match_cap,
                           vxorb_match_t* match_out, int* n_match_out, vxorb_frame_info* info_out,
                           float* pts_out, int* slot_out);
/* Rectified stereo sequences (BASELINE config 3).  `frames` holds `pairs` (left, right) pairs: frame 2k = left_k,
 * frame 2k + 1 = right_k (2 * pairs <= max_batch).  Every frame is extracted.  Match rows [0, pairs): left_k ->
 * right_k through the box mask of vxorb_match_boxed ({0, max_disparity, -dy_tol, dy_tol} for a rectified pair);
 * rows [pairs, 2 * pairs): the temporal leg left_(k-1) -> left_k through the square window `radius` (brute force
 * when radius <= 0), left_(-1) being the last left frame of the previous chunk (none after vxorb_track_reset).
 * kp_out / desc_out / n_kp_out have 2 * pairs frame rows, match_out / n_match_out 2 * pairs pair rows. */
int vxorb_track_submit_stereo(vxorb_handle* h, const uint8_t* frames, int pairs, int width, int height,
                              size_t stride, size_t frame_stride, const float* box, float radius, int kp_cap,
                              vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_kp_out, int match_cap,
                              vxorb_match_t* match_out, int* n_match_out, int* slot_out);
int vxorb_track_wait(vxorb_handle* h, int slot);
int vxorb_keypoint_capacity(const vxorb_handle* h);

/* Page-locked host memory for the streaming path (cudaHostAlloc / cudaFreeHost). */
void* vxorb_host_alloc(size_t bytes);
void vxorb_host_free(void* p);

/* Device-resident form used to time the kernels alone: the chunk uploaded with
 * vxorb_resident_upload is kept in every pipeline slot and re-processed by vxorb_resident_run (no host
 * copies).  Consecutive runs take the slots in turn, as consecutive vxorb_track_submit calls do, so
 * up to vxorb_track_slots(h) passes are in flight; every pass is the next chunk of the same sequence
 * (its first frame is matched against the previous pass's last frame).  Results stay on the device;
 * vxorb_resident_fetch returns those of the most recent pass. */
int vxorb_resident_upload(vxorb_handle* h, const uint8_t* frames, int batch, int width, int height,
                          size_t stride, size_t frame_stride);
int vxorb_resident_run(vxorb_handle* h, float radius);
/* The resident frames as (left, right) pairs with the matching of vxorb_track_submit_stereo. */
int vxorb_resident_run_stereo(vxorb_handle* h, const float* box, float radius);
int vxorb_resident_fetch(vxorb_handle* h, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out,
                         int* n_kp_out, int match_cap, vxorb_match_t* match_out, int* n_match_out);

/* ---- measurement hooks ---------------------------------------------------------------------
 * CUDA-event timing on the handle's own streams (torch.cuda.Event only sees torch's stream).
 * vxorb_timer_begin/end bracket a region; stage times are accumulated per kernel family
 * between vxorb_stage_times_reset and vxorb_stage_times. */
#define VXORB_STAGE_PYRAMID 0
#define VXORB_STAGE_FAST 1
#define VXORB_STAGE_SELECT 2
#define VXORB_STAGE_BLUR 3
#define VXORB_STAGE_DESCRIBE 4
#define VXORB_STAGE_MATCH 5
#define VXORB_STAGE_COUNT 6
int vxorb_sync(vxorb_handle* h);
int vxorb_timer_begin(vxorb_handle* h);
int vxorb_timer_end(vxorb_handle* h, float* ms_out);
int vxorb_stage_times_reset(vxorb_handle* h, int enable);
int vxorb_stage_times(vxorb_handle* h, float* ms_out /* [VXORB_STAGE_COUNT] */,
                      int* launches_out /* [VXORB_STAGE_COUNT] */);
/* Number of kernel launches issued by this handle since creation. */
long long vxorb_launch_count(const vxorb_handle* h);
/* Writes > L2-size bytes on the handle's stream so the next launch starts from a cold L2. */
int vxorb_flush_l2(vxorb_handle* h);

/* Integer-pipe calibration for the match roofline: runs a register-only XOR+POPC+IADD kernel on
 * every SM and reports the sustained 32-bit popcount rate (lane operations per second). */
int vxorb_measure_popc_rate(vxorb_handle* h, double* popc_per_s_out);

/* ---- stage isolation (parity tests only; each runs one kernel family on host buffers) ------ */
int vxorb_debug_resize(vxorb_handle* h, const uint8_t* src, int sw, int sh, size_t sstride,
                       uint8_t* dst, int dw, int dh, size_t dstride);
/* The production pyramid kernel (one level, one image): band_rows = 16 runs the throughput shape of the batched
 * path, 8 or 4 the latency shape (small batches, wide levels).  dw, dh >= 8 and scale <= ~2. */
int vxorb_debug_resize_band(vxorb_handle* h, const uint8_t* src, int sw, int sh, size_t sstride,
                            uint8_t* dst, int dw, int dh, size_t dstride, int band_rows);
/* FAST-9/16 + score + strict 3x3 NMS over one image; corners with border <= x < w-border (same
 * for y), border >= 3, are written in row-major order.  cv::FastFeatureDetector == border 3. */
int vxorb_debug_fast(vxorb_handle* h, const uint8_t* img, int width, int height, size_t stride,
                     int threshold, int border, int cap, vxorb_corner* out, int* n_out);
int vxorb_debug_blur(vxorb_handle* h, const uint8_t* src, int width, int height, size_t stride,
                     uint8_t* dst, size_t dstride);
/* Pyramid / blurred-pyramid level of frame `frame` of the most recent extract call.  A blurred level is defined only
 * where a descriptor can read it: pixels at least 13 px (edge threshold 31 - window reach 18) inside the level, and only
 * on levels that can hold keypoints; the rest of the plane is not computed (set VXORB_FULL_BLUR=1 before vxorb_create
 * for whole-level blurring, as cv::GaussianBlur(BORDER_REFLECT_101) would give it). */
int vxorb_debug_level(vxorb_handle* h, int frame, int level, int blurred, uint8_t* dst,
                      size_t dstride);

#ifdef __cplusplus
}
#endif
#endif /* VXORB_H_ */
