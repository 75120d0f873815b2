// vxorb_internal.h — shared declarations of the sm_100a ORB extract + match library.
//
// Data layout in HBM.  Everything is level-major so that one launch covers all
// frames of a batch with blockIdx.{y|z} = frame:
//   pyr / blur : for level l, frame f, pixel (x, y) : base + off[l] + f*fstride[l] + y*pitch[l] + x
//                (u8; off[l] is pixel (0,0) of frame 0, 16 B aligned; each plane has a 3-px
//                reflect-101 apron around it, see kApron*; planes 256 B aligned)
//   cand       : FAST corners after NMS, packed u32 = score<<24 | y<<12 | x, per (level, frame)
//   resp       : Harris response per candidate slot (f32), written only for retained candidates
//   fin        : final per-level keypoints (packed key, response), sorted by (y, x)
//   kp / desc  : frame-major outputs in canonical (octave, y, x) order
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../../include/vxorb.h"

// Checked build (make EXTRA=-DVXORB_CHECKED=1, scratch/ab.sh): every shared-memory / list / plane index the
// kernels compute is range-checked on the device and a violation traps, which fails the launch loudly.  This
// is the repo's own substitute for compute-sanitizer memcheck (closed on this GPU pool); the production build
// compiles the checks away.
#if defined(VXORB_CHECKED) && VXORB_CHECKED
#define VX_CHECK(cond)                                                                                     \
    do {                                                                                                   \
        if (!(cond)) {                                                                                     \
            printf("VX_CHECK failed: %s  (%s:%d, block %d,%d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, \
                   (int)blockIdx.y, (int)threadIdx.x);                                                     \
            __trap();                                                                                      \
        }                                                                                                  \
    } while (0)
#else
#define VX_CHECK(cond) ((void)0)
#endif

namespace vxorb {

constexpr int kMaxLevels = VXORB_MAX_LEVELS;
constexpr int kEdge = 31;        // cv::ORB edgeThreshold (only value supported)
constexpr int kBlurMargin = kEdge - 18;  // no descriptor window reads a blurred pixel closer than this to the border
constexpr int kHalfPatch = 15;   // patchSize 31
constexpr int kPitchAlign = 32;
// Every pyramid / blurred plane carries a reflect-101 apron so stencil kernels need no border
// logic: kApronY rows above and below, kApronX bytes left of pixel (0,0) (keeps it 16 B aligned)
// and at least kApronRight bytes after pixel (w-1).  The apron holds 3 valid reflected pixels.
constexpr int kApronY = 3;
constexpr int kApronX = 16;
constexpr int kApronRight = 12;
#ifndef VXORB_BLUR_ROWS
#define VXORB_BLUR_ROWS 64
#endif
constexpr int kBlurRows = VXORB_BLUR_ROWS;    // output rows per blur strip
#ifndef VXORB_BLUR_COLS
#define VXORB_BLUR_COLS 4
#endif
constexpr int kBlurCols = VXORB_BLUR_COLS;    // columns per blur strip (one thread): 4 or 8
static_assert(kBlurCols == 4 || kBlurCols == 8, "blur strips are 4 or 8 columns wide");

// FAST tiling (fast.cu): a CTA computes scores for kFastScoreW x kFastScoreH pixels and emits corners
// from the inner kFastEmitW x kFastEmitH; tile origins are multiples of 4 px in x.
#ifndef VXORB_FAST_ROWS
#define VXORB_FAST_ROWS 48  // score rows per tile (multiple of 8, <= 64: one row in eight per warp); swept 24..64 on B200
#endif
constexpr int kFastScoreW = 128, kFastScoreH = VXORB_FAST_ROWS;
constexpr int kFastEmitW = 124, kFastEmitH = kFastScoreH - 2;
inline int fast_tiles_x(int w, int border) {
    const int first = ((border - 1) & ~3) + 1;  // first emission column a tile row can cover
    const int span = (w - border) - first;
    return span > 0 ? (span + kFastEmitW - 1) / kFastEmitW : 0;
}
inline int fast_tiles_y(int h, int border) {
    const int span = h - 2 * border;
    return span > 0 ? (span + kFastEmitH - 1) / kFastEmitH : 0;
}

inline int plane_pitch(int w) { return (kApronX + w + kApronRight + kPitchAlign - 1) / kPitchAlign * kPitchAlign; }
inline size_t plane_bytes(int w, int h) { return (size_t)plane_pitch(w) * (size_t)(h + 2 * kApronY); }
// Fills the blur extent of level l (see LevelTab::blur_margin) and returns its strip count.
template <class LT>
inline int set_blur_extent(LT& lt, int l, int margin, bool active) {
    if (margin > 0 && !active) {  // no keypoints on this level: nothing reads its blurred plane
        lt.blur_x0[l] = lt.blur_y0[l] = lt.blur_y1[l] = lt.blur_tiles_x[l] = 0;
        return 0;
    }
    lt.blur_x0[l] = margin / kBlurCols * kBlurCols;
    lt.blur_y0[l] = margin;
    lt.blur_y1[l] = lt.h[l] - margin;
    lt.blur_tiles_x[l] = (lt.w[l] - margin - lt.blur_x0[l] + kBlurCols - 1) / kBlurCols;
    return lt.blur_tiles_x[l] * ((lt.blur_y1[l] - lt.blur_y0[l] + kBlurRows - 1) / kBlurRows);
}
inline size_t plane_origin(int w) { return (size_t)kApronY * plane_pitch(w) + kApronX; }

// Per-geometry tables passed to kernels by value (kernel parameters live in the constant bank).
struct LevelTab {
    int n_levels;
    int batch;                       // frames in this launch
    int w[kMaxLevels], h[kMaxLevels], pitch[kMaxLevels];
    unsigned long long off[kMaxLevels];      // byte offset of the level plane (frame 0)
    unsigned long long fstride[kMaxLevels];  // bytes between frames inside the level plane
    float scale[kMaxLevels];
    int quota[kMaxLevels];
    int active[kMaxLevels];          // level takes part in detection (w > 62 && h > 62 && quota > 0)
    int cand_cap[kMaxLevels];        // capacity of the candidate list of one (level, frame)
    unsigned long long cand_off[kMaxLevels];  // element offset of (level, frame 0) in cand/resp
    int fin_cap[kMaxLevels];
    unsigned long long fin_off[kMaxLevels];
    // tile bookkeeping: tiles of level l are [tile_begin[l], tile_begin[l+1]) in blockIdx.x
    int fast_tiles_x[kMaxLevels], fast_tile_begin[kMaxLevels + 1];
    // blur strips: blur_tiles_x = strips of kBlurCols px per row; CTAs of level l are [blur_tile_begin[l], [l+1])
    int blur_tiles_x[kMaxLevels], blur_tile_begin[kMaxLevels + 1];
    // The blurred planes are read by the descriptor windows alone: +-18 px (the rBRIEF pattern reaches 13 px, rotated
    // sqrt(2) * 13 -> 18) around keypoints that sit kEdge = 31 px inside the level.  With blur_margin = kBlurMargin only
    // the pixels [margin, w - margin) x [margin, h - margin) of the active levels are filtered (x rounded down to a
    // strip), nothing of the reflect-101 apron is ever read and nobody builds it; blur_margin = 0 filters whole levels
    // from aproned planes (stage-isolation entry points and VXORB_FULL_BLUR=1).  blur_x0[l] is the first strip's
    // column, blur_tiles_x[l] the strips per row, rows [blur_y0[l], blur_y1[l]).
    int blur_margin;
    int blur_x0[kMaxLevels], blur_y0[kMaxLevels], blur_y1[kMaxLevels];
    int coef_off[kMaxLevels];        // offset (in u32) of level l's resize tables: x[w_l] then y[h_l]
    int resize_fast[kMaxLevels];     // level l's x taps fit the register-resident resize kernel
    // band kernel (pyramid.cu): per-word tap table at coef + xw_off[l] (uint4[words] then uint2[words]),
    // rows per band, band count and the shared-memory bytes of the largest band's source chunk
    int xw_off[kMaxLevels], band_rows[kMaxLevels], n_bands[kMaxLevels], band_smem[kMaxLevels];
    int resize_band;                 // every level above 0 can use the band kernel
};

struct DeviceBuffers {
    uint8_t* pyr = nullptr;
    uint8_t* blur = nullptr;
    uint32_t* coef = nullptr;     // resize taps: ofs | c1 << 16
    uint32_t* cand = nullptr;
    int* cand_n = nullptr;        // [level][frame]
    float* resp = nullptr;
    uint32_t* fin_key = nullptr;
    float* fin_resp = nullptr;
    int* fin_n = nullptr;         // [level][frame]
    vxorb_keypoint* kp = nullptr; // [frame][kp_cap]
    uint8_t* desc = nullptr;      // [frame][kp_cap][32]
    float* xy = nullptr;          // [frame][kp_cap][2]  level-0 positions for windowed matching
    int* kp_n = nullptr;          // [frame] written count (<= kp_cap)
    int* kp_total = nullptr;      // [frame] true count (overflow detection)
    vxorb_knn2_t* knn = nullptr;  // [pair][kp_cap]
    vxorb_match_t* match = nullptr; // [pair][kp_cap]
    int* match_n = nullptr;       // [pair]
    int* status = nullptr;        // device-side overflow flags
};

// ---- kernel launchers (each in its own .cu) ----------------------------------------------------
// All take the stream explicitly; none synchronises.  Return the number of kernel launches issued.
int launch_gray_from_bgr(const uint8_t* bgr, size_t bgr_pitch, size_t bgr_fstride, uint8_t* dst,
                         int w, int h, int pitch, size_t fstride, int batch, cudaStream_t s);
// w must be a multiple of 16; src holds `batch` frames of w*h bytes back to back (16 B aligned)
int launch_unpack_frames(const uint8_t* src, uint8_t* dst, int w, int h, int pitch, size_t fstride, int batch,
                         cudaStream_t s);
int launch_pyramid(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s, bool aprons_inline = true);
int launch_resize_single(const uint8_t* src, int sw, int sh, int spitch, uint8_t* dst, int dw, int dh,
                         int dpitch, const uint32_t* xtab, const uint32_t* ytab, bool fast, cudaStream_t s);
bool resize_taps_fit_fast_path(const uint32_t* xtaps, int dw);
int launch_apron(const LevelTab& lt, const DeviceBuffers& db, int levels, cudaStream_t s);
void build_resize_words(const uint32_t* xtaps, int dw, uint32_t* out);  // 6 words per output word
bool plan_resize_bands(const uint32_t* ytaps, int dh, int sh, int spitch, int* band_rows, int* n_bands, int* smem);
void resize_band_extent(const uint32_t* ytaps, int dh, int sh, int spitch, int rows, int* n_bands, long long* smem);
// src_plane = start of the source plane allocation (apron included); rows 16 = throughput shape, 8 / 4 = latency shape
void launch_resize_band(const uint8_t* src_plane, int sh, int spitch, size_t sfstride, uint8_t* dst, int dw, int dh,
                        int dpitch, size_t dfstride, const uint32_t* ytab, const uint4* xw4, const uint2* xw2, int rows,
                        int n_bands, int smem, int batch, cudaStream_t s, bool write_apron = true);
cudaError_t init_pyramid_kernels();
// L1 / shared-memory split of the pipeline's kernels.  Left to the driver, every kernel gets the split that suits it
// alone (FAST all shared memory, the blur none), and CTAs of kernels with different splits cannot share an SM: it has
// to drain and be reconfigured.  In the production schedule of a chunk, where the pipeline slots overlap their
// kernels, every kernel therefore asks for the SAME split, kSharedCarveout percent shared memory (measured: 62 %
// 218.5 k, 68 % 219.7 k, 75 % 223.7 k, 82 % 222.1 k, 90 % 216.8 k, 100 % 215.7 k against 216.6-218.2 k frames/s
// with the driver's choice; windowed matcher 278.3 k against 267.5 k).  Alone — serial instrumented passes, single
// frames, per-call matching — the driver's choice stays, which is 1-6 % faster per kernel.
// carveout_*(pct) set the preference of one file's kernels (pct < 0: the driver's default); api.cu switches on change.
#ifndef VXORB_SHARED_CARVEOUT
#define VXORB_SHARED_CARVEOUT 75
#endif
constexpr int kSharedCarveout = VXORB_SHARED_CARVEOUT;
cudaError_t carveout_pyramid(int pct);
cudaError_t carveout_fast(int pct);
cudaError_t carveout_select(int pct);
cudaError_t carveout_blur(int pct);
cudaError_t carveout_describe(int pct);
cudaError_t carveout_match(int pct);
template <class K>
inline cudaError_t set_carveout(K kernel, int pct) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct < 0 ? -1 : pct);
}
int launch_fast(const LevelTab& lt, const DeviceBuffers& db, int threshold, int border, cudaStream_t s, bool shared = false);
int launch_select(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s);
int launch_blur(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s);
int launch_describe(const LevelTab& lt, const DeviceBuffers& db, int kp_cap, cudaStream_t s);

// knn over `pairs` independent (query, train) descriptor sets laid out with a fixed stride (rows).
struct MatchArgs {
    const uint8_t* q_desc; const uint8_t* t_desc;   // [pair][stride][32]
    const float* q_xy; const float* t_xy;           // [pair][stride][2] or nullptr
    const int* q_n; const int* t_n;                 // row counts (device): pair p reads q_n[p * n_stride], t_n[p * n_stride]
    int n_stride = 1;
    long long q_pair_stride, t_pair_stride;         // rows between consecutive pairs
    int max_q;                                      // upper bound of q_n (grid sizing)
    int max_t = 0;                                  // upper bound of t_n (shared-memory sizing of the windowed path)
    int pairs;
    float radius;                                   // <= 0: brute force
    // box mask (stereo / epipolar band): pair allowed iff box[0] <= xq - xt <= box[1] and
    // box[2] <= yq - yt <= box[3]; used instead of the square window when boxed != 0
    int boxed = 0;
    float box[4] = {0.f, 0.f, 0.f, 0.f};
    // arbitrary mask (cv knnMatch(mask)): row-major [max_q][mask_stride] bytes of ONE pair, pair (i, j) allowed iff != 0
    const uint8_t* mask = nullptr;
    long long mask_stride = 0;
    vxorb_knn2_t* knn;                              // [pair][knn_stride]
    int knn_stride;
    // brute-force path with few pairs: the train set is split over gridDim.z CTAs per query block, each
    // writing its packed top-2 keys to part[(pair * splits + split) * max_q + query]; a merge pass folds them
    uint2* part = nullptr;                          // scratch for >= pairs * kMaxTrainSplits * max_q entries, or nullptr
    int splits = 1;                                 // set by launch_knn2
};
constexpr int kMaxTrainSplits = 32;
int launch_knn2(const MatchArgs& a, cudaStream_t s);
// mode 0: ratio test; mode 1: ratio test + caller's post-filter; mode 2: cross-check with `back`
int launch_match_compact(const vxorb_knn2_t* knn, const vxorb_knn2_t* back, int knn_stride, int back_stride,
                         const int* q_n, int pairs, float nn_ratio, int mode, vxorb_match_t* out,
                         int out_stride, int* out_n, cudaStream_t s, int n_stride = 1);
// frame_info.cu: image sums over level 0, grid coverage, matched positions and parallax (2 launches)
int launch_frame_info(const LevelTab& lt, const DeviceBuffers& db, int kp_cap, unsigned long long* sums,
                      vxorb_frame_info* info, float4* pts, int grid_cols, int grid_rows, cudaStream_t s);
// Tracking::CreateLandmarksFromDepth's per-feature depth (tracking.cpp:606-637): depth_type 0 = u16 (/5000), 1 = f32, 2 = f64
int launch_depth_lookup(const float* xy, const int* n_dev, int max_n, const void* depth, int depth_type, int cols, int rows,
                        size_t stride, double* out, cudaStream_t s);
int launch_flush(uint8_t* buf, size_t bytes, cudaStream_t s);
int launch_popc_peak(uint32_t* out, int blocks, int iters, cudaStream_t s);

// CUtensorMap[kMaxLevels] over pyramid-shaped planes, box {box_w (multiple of 16), box_h, 1 frame}; fast.cu
bool build_plane_maps(const LevelTab& lt, const uint8_t* planes, int box_w, int box_h, bool active_only, void* out_maps);
cudaError_t init_device_tables();
cudaError_t init_match_kernels();
void build_resize_taps(int ssize, int dsize, uint32_t* taps);

}  // namespace vxorb
