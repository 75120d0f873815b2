// api.cu — the C ABI of include/vxorb.h: handle, geometry tables, device workspaces, pipeline
// slots (one CUDA stream each) and the host-buffer entry points.
//
// Boundary replaced (reference tree): FeatureExtractor::Extract (core/feature/feature_extractor.h:15,
// orb_extractor.cpp:9-27) and FeatureMatcher::Match (core/feature/feature_matcher.h:11-12,
// orb_matcher.cpp:11-43).  There is no CPU fallback anywhere in this file: every result comes from
// the kernels in pyramid.cu / fast.cu / select.cu / blur.cu / describe.cu / match.cu, and a failed
// CUDA call is reported as VXORB_E_CUDA.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <stdlib.h>
#include <string>
#include <vector>

#include "vxorb_internal.h"

using namespace vxorb;

namespace {

#ifndef VXORB_SLOTS
#define VXORB_SLOTS 4  // brute-force matcher: 3 slots 197.3 k, 4 slots 201.7 k, 5 slots 202.5 k resident frames/s (scratch/ab_run.py)
#endif
constexpr int kSlots = VXORB_SLOTS;
std::string g_create_error = "";
std::mutex g_create_mutex;

struct StageSpan {
    int stage;
    cudaEvent_t a, b;
};

struct Workspace {
    DeviceBuffers db;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    // side stream of the slot: the blur of a chunk runs here beside FAST + select (both only read the
    // pyramid); forked after the pyramid and joined before the descriptors
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int batch = 0;          // frames of the chunk in flight / last processed
    bool in_flight = false;
    int* h_status = nullptr;  // pinned: status word copied back with every chunk
    int* h_counts = nullptr;  // pinned: [max_batch] keypoint rows written, [max_batch] match rows of the chunk in flight
    struct Tail {             // host destinations of the chunk in flight: rows beyond `head` are fetched on demand
        vxorb_keypoint* kp;
        uint8_t* desc;
        vxorb_match_t* match;
        float* pts;
        size_t head;
    } tail = {nullptr, nullptr, nullptr, nullptr, 0};
    // caller-side per-frame quantities (frame_info.cu)
    unsigned long long* sums = nullptr;   // [max_batch][2]
    vxorb_frame_info* info = nullptr;     // [max_batch]
    float4* pts = nullptr;                // [max_batch][kp_cap]
    uint8_t* stage_in = nullptr;          // [max_batch][H][W] densely packed upload staging
    uint8_t* stage_bgr = nullptr;         // [max_batch][H][3 W] colour upload staging, allocated on the first BGR chunk
};

// A descriptor set that stays on the device under an id: the result of a single-frame extract call (ring of
// kRing entries, overwritten in turn) or rows uploaded with vxorb_set_upload (a local map; freed by the caller).
constexpr int kRing = 8;
constexpr long long kFirstSetId = 1LL << 40;  // frame ids count up from 1, uploaded sets from here
struct ResidentSet {
    long long id = 0;             // 0 = empty
    uint8_t* block = nullptr;     // ring entries: header | kp rows | desc rows in one allocation (n, kp, desc point into it)
    vxorb_keypoint* kp = nullptr; // ring entries only
    uint8_t* desc = nullptr;
    float* xy = nullptr;          // level-0 positions (nullptr for an uploaded set without positions)
    int* n = nullptr;             // device: [0] rows written, [1] true count
    int n_host = 0;
    // single-frame extraction is launch bound: its launches, memsets and the side-stream fork/join are captured
    // once per geometry and ring entry (the describe kernel writes straight into the entry) and replayed
    cudaGraphExec_t graph = nullptr;
    int graph_launches = 0;
};

}  // namespace

struct vxorb_handle {
    vxorb_params p;
    int kp_cap = 0;
    std::string err;
    // geometry
    int geo_w = 0, geo_h = 0;
    LevelTab lt;  // batch field is set per launch
    uint32_t* d_coef = nullptr;
    size_t pyr_bytes = 0;
    Workspace ws[kSlots];
    int next_slot = 0;
    int last_slot = -1;  // slot that processed the previous chunk; -1 after reset
    int res_batch = 0, res_next = 0, res_last = -1;  // resident mode: frames per slot, next / last slot used
    // last frame of the previous chunk (query side of the next chunk's first pair)
    uint8_t* carry_desc = nullptr;
    float* carry_xy = nullptr;
    int* carry_n = nullptr;
    // standalone match scratch (host-buffer API)
    uint8_t *m_q = nullptr, *m_t = nullptr;
    float *m_qxy = nullptr, *m_txy = nullptr;
    int* m_n = nullptr;  // [0] = nq, [1] = nt
    vxorb_knn2_t *m_knn = nullptr, *m_back = nullptr;
    uint8_t* m_out_blk = nullptr;  // m_out_n (header) and m_out (rows) live in this allocation
    vxorb_match_t* m_out = nullptr;
    uint2* m_part = nullptr;  // per-split partial top-2 keys of the single-pair brute-force launch
    uint8_t* m_mask = nullptr;  // nq x nt byte mask of vxorb_knn2_masked / vxorb_match_masked
    size_t m_mask_bytes = 0;
    const uint8_t* pending_mask = nullptr;
    int* m_out_n = nullptr;
    int m_cap = 0;
    // single-image stage scratch
    uint8_t *s_a = nullptr, *s_b = nullptr;
    size_t s_bytes = 0;
    // measurement
    cudaEvent_t t_begin = nullptr, t_end = nullptr;
    bool stage_enabled = false;
    std::vector<StageSpan> spans;
    std::vector<cudaEvent_t> event_pool;
    float stage_ms[VXORB_STAGE_COUNT] = {0};
    int stage_launches[VXORB_STAGE_COUNT] = {0};
    long long launches = 0;
    // device-resident descriptor sets (vxorb_extract_id / vxorb_set_upload / vxorb_match_ids)
    ResidentSet ring[kRing];
    bool ring_ready = false;
    long long next_frame_id = 1, next_set_id = kFirstSetId;
    std::map<long long, ResidentSet> sets;
    bool no_graph = false, no_overlap = false, no_flat_upload = false;  // VXORB_NO_* switches, read once in vxorb_create
    bool full_blur = false;  // VXORB_FULL_BLUR: blur whole levels (with aprons) instead of what descriptors can read
    uint8_t* x_stage = nullptr;  // pinned staging block of the synchronous extract calls
    size_t x_bytes = 0;
    uint8_t* flush_buf = nullptr;
    size_t flush_bytes = 0;
};

namespace {

int fail(vxorb_handle* h, int code, const std::string& msg) {
    if (h) h->err = msg;
    return code;
}

#define CK(expr)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(h, e__ == cudaErrorMemoryAllocation ? VXORB_E_NOMEM : VXORB_E_CUDA,        \
                        std::string(#expr) + ": " + cudaGetErrorString(e__));                      \
    } while (0)

inline int round_half_even(float v) { return (int)lrintf(v); }
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// cv::ORB level sizes, scales and quotas.
void level_geometry(const vxorb_params& p, int W, int H, int* lw, int* lh, float* ls, int* quota) {
    const double sf = (double)p.scale_factor;
    for (int l = 0; l < p.n_levels; ++l) {
        const float sc = (float)pow(sf, (double)l);
        const float inv = 1.0f / sc;
        ls[l] = sc;
        lw[l] = round_half_even((float)W * inv);
        lh[l] = round_half_even((float)H * inv);
    }
    const float factor = (float)(1.0 / sf);
    float nd = (float)p.n_features * (1.0f - factor) / (1.0f - (float)pow((double)factor, (double)p.n_levels));
    int sum = 0;
    for (int l = 0; l < p.n_levels - 1; ++l) {
        quota[l] = round_half_even(nd);
        sum += quota[l];
        nd *= factor;
    }
    quota[p.n_levels - 1] = std::max(p.n_features - sum, 0);
}

void free_workspace(Workspace& w) {
    DeviceBuffers& d = w.db;
    cudaFree(d.pyr); cudaFree(d.blur); cudaFree(d.cand); cudaFree(d.cand_n); cudaFree(d.resp);
    cudaFree(d.fin_key); cudaFree(d.fin_resp); cudaFree(d.fin_n); cudaFree(d.kp); cudaFree(d.desc);
    cudaFree(d.xy); cudaFree(d.kp_n); cudaFree(d.kp_total); cudaFree(d.knn); cudaFree(d.match);
    cudaFree(d.match_n); cudaFree(d.status);
    cudaFree(w.sums); cudaFree(w.info); cudaFree(w.pts); cudaFree(w.stage_in); cudaFree(w.stage_bgr);
    w.sums = nullptr; w.info = nullptr; w.pts = nullptr; w.stage_in = nullptr; w.stage_bgr = nullptr;
    d = DeviceBuffers();
}

void free_geometry(vxorb_handle* h) {
    for (auto& r : h->ring) {  // graph nodes point into the buffers freed below
        if (r.graph) cudaGraphExecDestroy(r.graph);
        r.graph = nullptr;
    }
    for (auto& w : h->ws) free_workspace(w);
    cudaFree(h->d_coef);
    cudaFree(h->carry_desc);
    cudaFree(h->carry_xy);
    cudaFree(h->carry_n);
    h->d_coef = nullptr;
    h->carry_desc = nullptr;
    h->carry_xy = nullptr;
    h->carry_n = nullptr;
    h->geo_w = h->geo_h = 0;
}

// (Re)builds the level table and device buffers for w x h frames.  All slots must be idle.
int ensure_geometry(vxorb_handle* h, int W, int H) {
    if (W == h->geo_w && H == h->geo_h) return VXORB_OK;
    if (W < 1 || H < 1 || W > h->p.max_width || H > h->p.max_height || W >= 4096 || H >= 4096)
        return fail(h, VXORB_E_INVALID, "frame size outside the handle's max_width/max_height (or >= 4096)");
    for (auto& w : h->ws)
        if (w.in_flight) return fail(h, VXORB_E_INVALID, "geometry change while a track slot is in flight");
    CK(cudaDeviceSynchronize());
    free_geometry(h);

    const vxorb_params& p = h->p;
    const int B = p.max_batch;
    LevelTab& lt = h->lt;
    memset(&lt, 0, sizeof(lt));
    lt.n_levels = p.n_levels;
    lt.batch = B;
    level_geometry(p, W, H, lt.w, lt.h, lt.scale, lt.quota);
    size_t off = 0, cand_off = 0, fin_off = 0;
    int coef_off = 0;
    std::vector<uint32_t> coef;
    lt.fast_tile_begin[0] = lt.blur_tile_begin[0] = 0;
    lt.resize_band = 1;
    lt.blur_margin = h->full_blur ? 0 : kBlurMargin;
    for (int l = 0; l < p.n_levels; ++l) {
        if (lt.w[l] < 1 || lt.h[l] < 1) {  // pyramid bottoms out: drop the remaining levels
            lt.n_levels = l;
            break;
        }
        lt.pitch[l] = plane_pitch(lt.w[l]);
        lt.fstride[l] = plane_bytes(lt.w[l], lt.h[l]);
        lt.off[l] = off + plane_origin(lt.w[l]);
        off += align_up((size_t)lt.fstride[l] * B, 256);
        lt.active[l] = (lt.w[l] > 2 * kEdge && lt.h[l] > 2 * kEdge && lt.quota[l] > 0) ? 1 : 0;
        const int ew = lt.w[l] - 2 * kEdge, eh = lt.h[l] - 2 * kEdge;  // emission domain
        lt.cand_cap[l] = lt.active[l] ? ((ew + 1) / 2 + 1) * ((eh + 1) / 2 + 1) : 0;
        lt.cand_off[l] = cand_off;
        cand_off += (size_t)lt.cand_cap[l] * B;
        lt.fin_cap[l] = std::min(lt.cand_cap[l], h->kp_cap);
        lt.fin_off[l] = fin_off;
        fin_off += (size_t)lt.fin_cap[l] * B;
        lt.fast_tiles_x[l] = lt.active[l] ? fast_tiles_x(lt.w[l], kEdge) : 0;
        const int fty = lt.active[l] ? fast_tiles_y(lt.h[l], kEdge) : 0;
        lt.fast_tile_begin[l + 1] = lt.fast_tile_begin[l] + lt.fast_tiles_x[l] * fty;
        lt.blur_tile_begin[l + 1] = lt.blur_tile_begin[l] + (set_blur_extent(lt, l, lt.blur_margin, lt.active[l] != 0) + 255) / 256;
        lt.coef_off[l] = coef_off;
        if (l > 0) {
            coef.resize((size_t)coef_off + lt.w[l] + lt.h[l]);
            build_resize_taps(lt.w[l - 1], lt.w[l], coef.data() + coef_off);
            build_resize_taps(lt.h[l - 1], lt.h[l], coef.data() + coef_off + lt.w[l]);
            lt.resize_fast[l] = resize_taps_fit_fast_path(coef.data() + coef_off, lt.w[l]) ? 1 : 0;
            const bool banded = lt.resize_fast[l] && lt.w[l] >= 8 &&
                                plan_resize_bands(coef.data() + coef_off + lt.w[l], lt.h[l], lt.h[l - 1], lt.pitch[l - 1],
                                                  &lt.band_rows[l], &lt.n_bands[l], &lt.band_smem[l]);
            if (!banded) lt.resize_band = 0;
            coef_off += lt.w[l] + lt.h[l];
            coef_off = (coef_off + 3) & ~3;  // uint4 alignment of the per-word table
            const int words = (lt.w[l] + 3) / 4;
            lt.xw_off[l] = coef_off;
            coef.resize((size_t)coef_off + (size_t)words * 6);
            build_resize_words(coef.data() + lt.coef_off[l], lt.w[l], coef.data() + coef_off);
            coef_off += words * 6;
        }
    }
    for (int l = lt.n_levels; l < kMaxLevels; ++l) {
        lt.fast_tile_begin[l + 1] = lt.fast_tile_begin[lt.n_levels];
        lt.blur_tile_begin[l + 1] = lt.blur_tile_begin[lt.n_levels];
    }
    h->pyr_bytes = off;
    if (!coef.empty()) {
        CK(cudaMalloc(&h->d_coef, coef.size() * 4));
        CK(cudaMemcpy(h->d_coef, coef.data(), coef.size() * 4, cudaMemcpyHostToDevice));
    }
    const size_t cap = (size_t)h->kp_cap;
    CK(cudaMalloc(&h->carry_desc, cap * 32));
    CK(cudaMalloc(&h->carry_xy, cap * 8));
    CK(cudaMalloc(&h->carry_n, sizeof(int)));
    CK(cudaMemset(h->carry_n, 0, sizeof(int)));
    for (auto& w : h->ws) {
        DeviceBuffers& d = w.db;
        d.coef = h->d_coef;
        CK(cudaMalloc(&d.pyr, std::max<size_t>(off, 256)));
        CK(cudaMalloc(&d.blur, std::max<size_t>(off, 256)));
        CK(cudaMemset(d.pyr, 0, std::max<size_t>(off, 256)));
        CK(cudaMemset(d.blur, 0, std::max<size_t>(off, 256)));
        CK(cudaMalloc(&d.cand, std::max<size_t>(cand_off, 1) * 4));
        CK(cudaMalloc(&d.resp, std::max<size_t>(cand_off, 1) * 4));
        CK(cudaMalloc(&d.fin_key, std::max<size_t>(fin_off, 1) * 4));
        CK(cudaMalloc(&d.fin_resp, std::max<size_t>(fin_off, 1) * 4));
        CK(cudaMalloc(&d.cand_n, sizeof(int) * kMaxLevels * B));
        CK(cudaMalloc(&d.fin_n, sizeof(int) * kMaxLevels * B));
        CK(cudaMalloc(&d.kp, sizeof(vxorb_keypoint) * cap * (B + 1)));
        CK(cudaMalloc(&d.desc, 32 * cap * (B + 1)));
        CK(cudaMalloc(&d.xy, 8 * cap * (B + 1)));
        CK(cudaMalloc(&d.kp_n, sizeof(int) * (B + 1)));
        CK(cudaMalloc(&d.kp_total, sizeof(int) * (B + 1)));
        CK(cudaMemset(d.kp_n, 0, sizeof(int) * (B + 1)));
        CK(cudaMemset(d.kp_total, 0, sizeof(int) * (B + 1)));
        CK(cudaMalloc(&d.knn, sizeof(vxorb_knn2_t) * cap * B));
        CK(cudaMalloc(&d.match, sizeof(vxorb_match_t) * cap * B));
        CK(cudaMalloc(&d.match_n, sizeof(int) * B));
        CK(cudaMalloc(&d.status, sizeof(int) * 4));
        CK(cudaMemset(d.status, 0, sizeof(int) * 4));
        CK(cudaMalloc(&w.sums, sizeof(unsigned long long) * 2 * B));
        CK(cudaMalloc(&w.info, sizeof(vxorb_frame_info) * B));
        CK(cudaMalloc(&w.pts, sizeof(float4) * cap * B));
        CK(cudaMalloc(&w.stage_in, (size_t)W * H * B));
    }
    h->geo_w = W;
    h->geo_h = H;
    h->last_slot = -1;
    h->res_batch = 0;  // resident frames belong to the previous geometry
    h->res_last = -1;
    return VXORB_OK;
}

cudaEvent_t take_event(vxorb_handle* h) {
    if (!h->event_pool.empty()) {
        cudaEvent_t e = h->event_pool.back();
        h->event_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

struct StageScope {
    vxorb_handle* h;
    cudaStream_t s;
    int stage;
    cudaEvent_t a = nullptr;
    StageScope(vxorb_handle* h_, cudaStream_t s_, int stage_) : h(h_), s(s_), stage(stage_) {
        if (h->stage_enabled) {
            a = take_event(h);
            cudaEventRecord(a, s);
        }
    }
    void end(int launches) {
        h->launches += launches;
        h->stage_launches[stage] += launches;
        if (h->stage_enabled) {
            cudaEvent_t b = take_event(h);
            cudaEventRecord(b, s);
            h->spans.push_back({stage, a, b});
        }
    }
};

// Switches the L1 / shared-memory preference of every pipeline kernel (vxorb_internal.h: kSharedCarveout) when the
// mode changes: pct = kSharedCarveout while chunks overlap over the pipeline slots, -1 (the driver's per-kernel
// choice) for everything that runs alone.  The preference is read when a launch is queued, so launches already in a
// stream keep theirs.  Function attributes are per device and per process: one entry per device here.
int apply_carveout(vxorb_handle* h, int pct) {
    static int current[64] = {0};  // 0 = never set (the library never asks for 0 %)
    static std::mutex guard;       // handles of different host threads may share a device
    const int dev = h->p.device_id & 63;
    std::lock_guard<std::mutex> lock(guard);
    if (current[dev] == pct) return VXORB_OK;
    cudaError_t e;
    if ((e = carveout_pyramid(pct)) != cudaSuccess || (e = carveout_fast(pct)) != cudaSuccess ||
        (e = carveout_select(pct)) != cudaSuccess || (e = carveout_blur(pct)) != cudaSuccess ||
        (e = carveout_describe(pct)) != cudaSuccess || (e = carveout_match(pct)) != cudaSuccess)
        return fail(h, VXORB_E_CUDA, cudaGetErrorString(e));
    current[dev] = pct;
    return VXORB_OK;
}

// Queues the extraction kernels for `batch` frames whose level 0 already sits in ws.db.pyr.
// Outputs go to frame slots 1..batch of kp/desc/xy/kp_n (slot 0 is the streaming carry).
int enqueue_extract(vxorb_handle* h, Workspace& ws, int batch, const ResidentSet* out = nullptr) {
    LevelTab lt = h->lt;
    // Frame strides inside a level plane do not depend on the batch, only the launch extent does.
    lt.batch = batch;
    cudaStream_t s = ws.stream;
    DeviceBuffers d = ws.db;
    CK(cudaMemsetAsync(d.cand_n, 0, sizeof(int) * kMaxLevels * h->p.max_batch, s));
    CK(cudaMemsetAsync(d.status, 0, sizeof(int) * 4, s));
    // cand_n / fin_n are indexed [level * batch + frame] with the launch's batch
    // The blur needs only the pyramid, FAST + select never touch the blurred planes: outside the
    // instrumented (per-stage timed) mode the blur runs on the slot's side stream beside them and fills
    // the issue slots FAST's barriers and select's tails leave idle.  Instrumented runs stay serial so a
    // stage's events bracket that stage alone.
    const bool overlap = !h->stage_enabled && ws.side != nullptr && !h->no_overlap;
    {
        const int rc = apply_carveout(h, (overlap && batch >= 8) ? kSharedCarveout : -1);
        if (rc) return rc;
    }
    // The reflect apron of the planes is read by the blur alone (FAST's tiles take it in but never score a ring that
    // touches it; the resampling, Harris and the descriptor windows stay inside the image).  For a few frames, where
    // the level chain is the critical path of the call, the level kernels therefore skip their apron phase and the
    // aprons of all levels are built on the side stream just before the blur, beside FAST + select.  Large batches
    // keep them in the band kernels (a separate pass over the planes costs more than it hides there).
    const bool need_apron = lt.blur_margin < kApronY;  // the interior-only blur never reaches the border
    const bool apron_aside = need_apron && overlap && lt.resize_band && batch < 8;
    {
        StageScope sc(h, s, VXORB_STAGE_PYRAMID);
        sc.end(launch_pyramid(lt, d, s, !apron_aside));
    }
    if (overlap) {
        CK(cudaEventRecord(ws.ev_fork, s));
        CK(cudaStreamWaitEvent(ws.side, ws.ev_fork, 0));
        if (apron_aside) {
            const int na = launch_apron(lt, d, lt.n_levels, ws.side);
            h->launches += na;
            h->stage_launches[VXORB_STAGE_PYRAMID] += na;
        }
        const int n = launch_blur(lt, d, ws.side);
        h->launches += n;
        h->stage_launches[VXORB_STAGE_BLUR] += n;
        CK(cudaEventRecord(ws.ev_join, ws.side));
    }
    {
        StageScope sc(h, s, VXORB_STAGE_FAST);
        const int n = launch_fast(lt, d, h->p.fast_threshold, kEdge, s, overlap && batch >= 8);
        if (n < 0) return fail(h, VXORB_E_CUDA, "cuTensorMapEncodeTiled failed for the FAST tile maps");
        sc.end(n);
    }
    {
        StageScope sc(h, s, VXORB_STAGE_SELECT);
        sc.end(launch_select(lt, d, s));
    }
    if (overlap) {
        CK(cudaStreamWaitEvent(s, ws.ev_join, 0));
    } else {
        StageScope sc(h, s, VXORB_STAGE_BLUR);
        sc.end(launch_blur(lt, d, s));
    }
    {
        StageScope sc(h, s, VXORB_STAGE_DESCRIBE);
        DeviceBuffers o = d;  // shift outputs by one frame slot
        o.kp = d.kp + h->kp_cap;
        o.desc = d.desc + (size_t)h->kp_cap * 32;
        o.xy = d.xy + (size_t)h->kp_cap * 2;
        o.kp_n = d.kp_n + 1;
        o.kp_total = d.kp_total + 1;
        if (out) {  // single frame: straight into the resident set
            o.kp = out->kp;
            o.desc = out->desc;
            o.xy = out->xy;
            o.kp_n = out->n;
            o.kp_total = out->n + 1;
        }
        const int n = launch_describe(lt, o, h->kp_cap, s);
        if (n < 0) return fail(h, VXORB_E_CUDA, "cuTensorMapEncodeTiled failed for the descriptor window maps");
        sc.end(n);
    }
    CK(cudaGetLastError());
    return VXORB_OK;
}

// Single-frame extraction on slot 0 through a CUDA graph (captured on first use for the current geometry and
// ring entry).
int launch_extract_graph(vxorb_handle* h, Workspace& ws, ResidentSet& rs) {
    if (!rs.graph) {
        const long long before = h->launches;
        int stage_before[VXORB_STAGE_COUNT];
        for (int i = 0; i < VXORB_STAGE_COUNT; ++i) stage_before[i] = h->stage_launches[i];
        {
            const int rc = apply_carveout(h, -1);  // a single frame runs alone; not inside the capture
            if (rc) return rc;
        }
        CK(cudaStreamBeginCapture(ws.stream, cudaStreamCaptureModeThreadLocal));
        const int rc = enqueue_extract(h, ws, 1, &rs);
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(ws.stream, &graph);
        if (rc) {
            if (graph) cudaGraphDestroy(graph);
            return rc;
        }
        if (e != cudaSuccess || !graph) return fail(h, VXORB_E_CUDA, std::string("graph capture: ") + cudaGetErrorString(e));
        const cudaError_t ei = cudaGraphInstantiate(&rs.graph, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) {
            rs.graph = nullptr;
            return fail(h, VXORB_E_CUDA, std::string("graph instantiate: ") + cudaGetErrorString(ei));
        }
        rs.graph_launches = (int)(h->launches - before);
        h->launches = before;  // nothing ran during the capture; every replay below counts its kernels
        for (int i = 0; i < VXORB_STAGE_COUNT; ++i) h->stage_launches[i] = stage_before[i];
    }
    CK(cudaGraphLaunch(rs.graph, ws.stream));
    h->launches += rs.graph_launches;
    return VXORB_OK;
}

// Allocates the ring of resident frame sets (kp_cap rows each) on first use.
int ensure_ring(vxorb_handle* h) {
    if (h->ring_ready) return VXORB_OK;
    const size_t cap = (size_t)h->kp_cap;
    // one block per entry, laid out like the pinned staging block of the synchronous extract call — 64-byte header
    // (rows written, true count), keypoint rows, descriptor rows — so the whole result is ONE device-to-host copy
    const size_t kp_bytes = align_up(sizeof(vxorb_keypoint) * cap, 64);
    for (auto& r : h->ring) {
        CK(cudaMalloc(&r.block, 64 + kp_bytes + 32 * cap));
        CK(cudaMemset(r.block, 0, 64));
        r.n = reinterpret_cast<int*>(r.block);
        r.kp = reinterpret_cast<vxorb_keypoint*>(r.block + 64);
        r.desc = r.block + 64 + kp_bytes;
        CK(cudaMalloc(&r.xy, 8 * cap));
    }
    h->ring_ready = true;
    return VXORB_OK;
}

// Looks an id up in the ring / the uploaded sets; nullptr when it is unknown or has been overwritten.
const ResidentSet* find_set(const vxorb_handle* h, long long id) {
    if (id <= 0) return nullptr;
    if (id >= kFirstSetId) {
        auto it = h->sets.find(id);
        return it == h->sets.end() ? nullptr : &it->second;
    }
    const ResidentSet& r = h->ring[id % kRing];
    return (h->ring_ready && r.id == id) ? &r : nullptr;
}

// Queues the frame-to-frame match of slots (p, p+1), p = 0..batch-1.
int enqueue_track_match(vxorb_handle* h, Workspace& ws, int batch, float radius) {
    cudaStream_t s = ws.stream;
    DeviceBuffers& d = ws.db;
    StageScope sc(h, s, VXORB_STAGE_MATCH);
    MatchArgs a;
    a.q_desc = d.desc;
    a.t_desc = d.desc + (size_t)h->kp_cap * 32;
    a.q_xy = d.xy;
    a.t_xy = d.xy + (size_t)h->kp_cap * 2;
    a.q_n = d.kp_n;
    a.t_n = d.kp_n + 1;
    a.q_pair_stride = a.t_pair_stride = h->kp_cap;
    a.max_q = h->kp_cap;
    a.max_t = h->kp_cap;
    a.pairs = batch;
    a.radius = radius;
    a.knn = d.knn;
    a.knn_stride = h->kp_cap;
    int n = launch_knn2(a, s);
    n += launch_match_compact(d.knn, nullptr, h->kp_cap, 0, d.kp_n, batch, h->p.nn_ratio, 0, d.match, h->kp_cap,
                              d.match_n, s);
    sc.end(n);
    CK(cudaGetLastError());
    return VXORB_OK;
}

// Stereo chunk (BASELINE config 3): frame slots 1..2P hold (left_0, right_0, left_1, right_1, ...).  Queues, into
// knn / match / match_n rows [0, P): left_k -> right_k through the box mask, and into rows [P, 2P): the temporal
// leg left_{k-1} -> left_k through the square window (brute force when radius <= 0); left_{-1} is the carry in
// slot 0.
int enqueue_stereo_match(vxorb_handle* h, Workspace& ws, int pairs, const float* box, float radius) {
    cudaStream_t s = ws.stream;
    DeviceBuffers& d = ws.db;
    const size_t cap = (size_t)h->kp_cap;
    StageScope sc(h, s, VXORB_STAGE_MATCH);
    int n = 0;
    auto leg = [&](int q_slot, int t_slot, int slot_stride, int n_pairs, int out_row, bool boxed) {
        if (n_pairs <= 0) return;
        MatchArgs a;
        a.q_desc = d.desc + cap * 32 * q_slot;
        a.t_desc = d.desc + cap * 32 * t_slot;
        a.q_xy = d.xy + cap * 2 * q_slot;
        a.t_xy = d.xy + cap * 2 * t_slot;
        a.q_n = d.kp_n + q_slot;
        a.t_n = d.kp_n + t_slot;
        a.n_stride = slot_stride;
        a.q_pair_stride = a.t_pair_stride = (long long)cap * slot_stride;
        a.max_q = a.max_t = h->kp_cap;
        a.pairs = n_pairs;
        a.radius = boxed ? 0.f : radius;
        if (boxed) {
            a.boxed = 1;
            for (int i = 0; i < 4; ++i) a.box[i] = box[i];
        }
        a.knn = d.knn + cap * out_row;
        a.knn_stride = h->kp_cap;
        n += launch_knn2(a, s);
        n += launch_match_compact(a.knn, nullptr, h->kp_cap, 0, a.q_n, n_pairs, h->p.nn_ratio, 0, d.match + cap * out_row,
                                  h->kp_cap, d.match_n + out_row, s, slot_stride);
    };
    // the two legs are independent (different output rows) and each is latency bound at one CTA per pair and query
    // split: outside the instrumented mode the temporal leg runs on the slot's side stream beside the stereo leg
    const bool aside = !h->stage_enabled && ws.side != nullptr && !h->no_overlap;
    if (aside) {
        CK(cudaEventRecord(ws.ev_fork, s));
        CK(cudaStreamWaitEvent(ws.side, ws.ev_fork, 0));
    }
    leg(1, 2, 2, pairs, 0, true);              // left_k -> right_k
    if (aside) s = ws.side;
    leg(0, 1, 1, 1, pairs, false);             // carry -> left_0
    leg(1, 3, 2, pairs - 1, pairs + 1, false); // left_{k-1} -> left_k, k >= 1
    if (aside) {
        CK(cudaEventRecord(ws.ev_join, ws.side));
        s = ws.stream;
        CK(cudaStreamWaitEvent(s, ws.ev_join, 0));
    }
    sc.end(n);
    CK(cudaGetLastError());
    return VXORB_OK;
}

// Queues the caller-side per-frame quantities (image statistics, grid coverage, matched positions,
// parallax) for frames 1..batch of the workspace; needs match / match_n of those pairs to be final.
int enqueue_frame_info(vxorb_handle* h, Workspace& ws, int batch, bool want_pts) {
    LevelTab lt = h->lt;
    lt.batch = batch;
    StageScope sc(h, ws.stream, VXORB_STAGE_MATCH);
    sc.end(launch_frame_info(lt, ws.db, h->kp_cap, ws.sums, ws.info, want_pts ? ws.pts : nullptr, 5, 5, ws.stream));
    CK(cudaGetLastError());
    return VXORB_OK;
}

// Copies `batch` host frames into level 0 of the workspace.
int upload_frames(vxorb_handle* h, Workspace& ws, const uint8_t* frames, int batch, int W, int H, size_t stride,
                  size_t frame_stride) {
    const LevelTab& lt = h->lt;
    uint8_t* dst = ws.db.pyr + lt.off[0];
    if (batch > 1 && stride == (size_t)W && frame_stride == (size_t)W * H && W % 16 == 0 && ws.stage_in &&
        !h->no_flat_upload) {
        // densely packed chunk: one flat copy over the link, rows scattered into the planes on the device
        CK(cudaMemcpyAsync(ws.stage_in, frames, (size_t)W * H * batch, cudaMemcpyHostToDevice, ws.stream));
        h->launches += launch_unpack_frames(ws.stage_in, dst, W, H, lt.pitch[0], (size_t)lt.fstride[0], batch, ws.stream);
        h->stage_launches[VXORB_STAGE_PYRAMID] += 1;
        return VXORB_OK;
    }
    if (stride > 0 && frame_stride % stride == 0 && frame_stride / stride >= (size_t)H) {
        // one strided 3-D copy: frame f, row y of the source -> plane f, row y (planes carry an apron)
        cudaMemcpy3DParms p;
        memset(&p, 0, sizeof(p));
        p.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t*>(frames), stride, (size_t)W, frame_stride / stride);
        p.dstPtr = make_cudaPitchedPtr(dst, (size_t)lt.pitch[0], (size_t)W, (size_t)(lt.fstride[0] / lt.pitch[0]));
        p.extent = make_cudaExtent((size_t)W, (size_t)H, (size_t)batch);
        p.kind = cudaMemcpyHostToDevice;
        CK(cudaMemcpy3DAsync(&p, ws.stream));
    } else {
        for (int f = 0; f < batch; ++f)
            CK(cudaMemcpy2DAsync(dst + (size_t)f * lt.fstride[0], lt.pitch[0], frames + (size_t)f * frame_stride,
                                 stride, W, H, cudaMemcpyHostToDevice, ws.stream));
    }
    return VXORB_OK;
}

// Copies `batch` host BGR frames (3 bytes per pixel) into the workspace's colour staging buffer and converts
// them into level 0 on the device (the conversion cv::ORB applies to a 3-channel input).
int upload_bgr_frames(vxorb_handle* h, Workspace& ws, const uint8_t* frames, int batch, int W, int H, size_t stride,
                      size_t frame_stride) {
    const LevelTab& lt = h->lt;
    const size_t row = (size_t)W * 3, frame = row * H;
    if (!ws.stage_bgr) {
        if (cudaMalloc(&ws.stage_bgr, frame * (size_t)h->p.max_batch) != cudaSuccess) {
            ws.stage_bgr = nullptr;
            cudaGetLastError();
            return fail(h, VXORB_E_NOMEM, "device allocation of the BGR staging buffer failed");
        }
    }
    if (stride == row && frame_stride == frame) {
        CK(cudaMemcpyAsync(ws.stage_bgr, frames, frame * batch, cudaMemcpyHostToDevice, ws.stream));
    } else {
        for (int f = 0; f < batch; ++f)
            CK(cudaMemcpy2DAsync(ws.stage_bgr + (size_t)f * frame, row, frames + (size_t)f * frame_stride, stride, row, H,
                                 cudaMemcpyHostToDevice, ws.stream));
    }
    h->launches += launch_gray_from_bgr(ws.stage_bgr, row, frame, ws.db.pyr + lt.off[0], W, H, lt.pitch[0],
                                        (size_t)lt.fstride[0], batch, ws.stream);
    h->stage_launches[VXORB_STAGE_PYRAMID] += 1;
    return VXORB_OK;
}

int check_status(vxorb_handle* h, int status) {
    if (status == 0) return VXORB_OK;
    char buf[160];
    snprintf(buf, sizeof(buf),
             "capacity exceeded (flags 0x%x: 1=candidate list, 2=per-level list, 4=max_keypoints per frame)", status);
    return fail(h, VXORB_E_CAPACITY, buf);
}

int ensure_match_scratch(vxorb_handle* h, int rows) {
    if (rows <= h->m_cap) return VXORB_OK;
    cudaFree(h->m_q); cudaFree(h->m_t); cudaFree(h->m_qxy); cudaFree(h->m_txy); cudaFree(h->m_n);
    cudaFree(h->m_knn); cudaFree(h->m_back); cudaFree(h->m_out_blk); cudaFree(h->m_part);
    h->m_q = h->m_t = nullptr;
    h->m_qxy = h->m_txy = nullptr;
    h->m_n = nullptr;
    h->m_knn = h->m_back = nullptr;
    h->m_out_blk = nullptr;
    h->m_out = nullptr;
    h->m_out_n = nullptr;
    h->m_part = nullptr;
    h->m_cap = 0;
    const size_t r = (size_t)rows;
    CK(cudaMalloc(&h->m_q, r * 32));
    CK(cudaMalloc(&h->m_t, r * 32));
    CK(cudaMalloc(&h->m_qxy, r * 8));
    CK(cudaMalloc(&h->m_txy, r * 8));
    CK(cudaMalloc(&h->m_n, sizeof(int) * 4));
    CK(cudaMalloc(&h->m_knn, sizeof(vxorb_knn2_t) * r));
    CK(cudaMalloc(&h->m_back, sizeof(vxorb_knn2_t) * r));
    // count (64-byte header) and rows in one block: the read-back of a match call is one copy
    CK(cudaMalloc(&h->m_out_blk, 64 + sizeof(vxorb_match_t) * r));
    h->m_out_n = reinterpret_cast<int*>(h->m_out_blk);
    h->m_out = reinterpret_cast<vxorb_match_t*>(h->m_out_blk + 64);
    CK(cudaMalloc(&h->m_part, sizeof(uint2) * r * kMaxTrainSplits));
    h->m_cap = rows;
    return VXORB_OK;
}

// Grows the handle's pinned host staging block (shared by the synchronous extract and match calls).
int ensure_host_stage(vxorb_handle* h, size_t need) {
    if (need <= h->x_bytes) return VXORB_OK;
    if (h->x_stage) cudaFreeHost(h->x_stage);
    h->x_stage = nullptr;
    h->x_bytes = 0;
    if (cudaHostAlloc(reinterpret_cast<void**>(&h->x_stage), need, cudaHostAllocDefault) != cudaSuccess)
        return fail(h, VXORB_E_NOMEM, "pinned staging allocation failed");
    h->x_bytes = need;
    return VXORB_OK;
}

// A descriptor set on the device as the match kernels see it.
struct DevSet {
    const uint8_t* desc;
    const float* xy;   // or nullptr
    const int* n_dev;  // device row count
    int n;             // the same count on the host (grid sizing)
};

// 2-NN of q against t on slot 0's stream into m_knn (and, for the cross-check, the reverse direction into m_back).
int run_knn_device(vxorb_handle* h, const DevSet& q, const DevSet& t, float radius, bool reverse_too, const float* box) {
    if (q.n >= (1 << 20) || t.n >= (1 << 20))
        return fail(h, VXORB_E_INVALID, "descriptor count must stay below 2^20 (index field of the packed 2-NN key)");
    int rc = ensure_match_scratch(h, std::max(std::max(q.n, t.n), 1));  // grows on demand: no fixed limit per side
    if (rc) return rc;
    if ((rc = apply_carveout(h, -1)) != VXORB_OK) return rc;  // a single pair runs alone
    cudaStream_t s = h->ws[0].stream;
    const bool win = (radius > 0.f || box) && q.xy && t.xy;
    if (box && !win) return fail(h, VXORB_E_INVALID, "box mask needs query and train positions");
    StageScope sc(h, s, VXORB_STAGE_MATCH);
    MatchArgs a;
    a.q_desc = q.desc; a.t_desc = t.desc;
    a.q_xy = win ? q.xy : nullptr; a.t_xy = win ? t.xy : nullptr;
    a.q_n = q.n_dev; a.t_n = t.n_dev;
    a.q_pair_stride = a.t_pair_stride = 0;
    a.max_q = q.n; a.max_t = t.n; a.pairs = 1; a.radius = win ? radius : 0.f;
    a.knn = h->m_knn; a.knn_stride = 0;
    a.part = h->m_part;
    if (h->pending_mask) {  // set by the masked entry points around their run_knn_host call
        a.mask = h->pending_mask;
        a.mask_stride = t.n;
    }
    if (box) {
        a.boxed = 1;
        for (int i = 0; i < 4; ++i) a.box[i] = box[i];
    }
    int launches = launch_knn2(a, s);
    if (reverse_too) {
        MatchArgs b = a;
        b.q_desc = t.desc; b.t_desc = q.desc;
        b.q_xy = a.t_xy; b.t_xy = a.q_xy;
        b.q_n = t.n_dev; b.t_n = q.n_dev;
        b.max_q = t.n; b.max_t = q.n; b.knn = h->m_back;
        launches += launch_knn2(b, s);
    }
    sc.end(launches);
    CK(cudaGetLastError());
    return VXORB_OK;
}

// Compaction of m_knn into accepted matches (mode 0: ratio test, 1: + the caller's post-filter, 2: cross-check)
// and the read-back: the count and the first min(out_cap, nq) rows reach the pinned staging block in one round trip.
int finish_match(vxorb_handle* h, const DevSet& q, float nn_ratio, int mode, int out_cap, vxorb_match_t* out, int* n_out) {
    cudaStream_t s = h->ws[0].stream;
    if (nn_ratio <= 0.f) nn_ratio = h->p.nn_ratio;
    h->launches += launch_match_compact(h->m_knn, mode == 2 ? h->m_back : nullptr, 0, 0, q.n_dev, 1, nn_ratio, mode,
                                        h->m_out, h->m_cap, h->m_out_n, s);
    h->stage_launches[VXORB_STAGE_MATCH] += 1;
    CK(cudaGetLastError());
    const int rows = std::min(out_cap, q.n);
    const size_t need = 64 + sizeof(vxorb_match_t) * (size_t)rows;
    int rc = ensure_host_stage(h, need);
    if (rc) return rc;
    int* n_st = reinterpret_cast<int*>(h->x_stage);
    vxorb_match_t* out_st = reinterpret_cast<vxorb_match_t*>(h->x_stage + 64);
    CK(cudaMemcpyAsync(n_st, h->m_out_blk, need, cudaMemcpyDeviceToHost, s));  // header + rows, one copy
    CK(cudaStreamSynchronize(s));
    const int n = *n_st;
    *n_out = n;
    const int m = std::min(n, rows);
    if (m > 0) memcpy(out, out_st, sizeof(vxorb_match_t) * (size_t)m);
    if (n > out_cap) return fail(h, VXORB_E_CAPACITY, "match output buffer too small");
    return VXORB_OK;
}

// Host-buffer knn: uploads both descriptor sets through the pinned staging block (pageable sources would make
// every copy a blocking one) and runs the knn kernel.
int run_knn_host(vxorb_handle* h, const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy,
                 int nt, float radius, bool reverse_too, const float* box = nullptr) {
    if (nq < 0 || nt < 0 || (nq > 0 && !q) || (nt > 0 && !t)) return fail(h, VXORB_E_INVALID, "bad descriptor arguments");
    if (nt >= (1 << 20) || nq >= (1 << 20))
        return fail(h, VXORB_E_INVALID, "descriptor count must stay below 2^20 (index field of the packed 2-NN key)");
    int rc = ensure_match_scratch(h, std::max(std::max(nq, nt), 1));
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    const bool win = (radius > 0.f || box) && qxy && txy;
    if (box && !win) return fail(h, VXORB_E_INVALID, "box mask needs query and train positions");
    // the block is also large enough for the outputs the callers stage afterwards
    const size_t q_off = 64, t_off = q_off + (size_t)nq * 32, qxy_off = t_off + (size_t)nt * 32;
    const size_t txy_off = qxy_off + (size_t)nq * 8, in_end = txy_off + (size_t)nt * 8;
    rc = ensure_host_stage(h, std::max(in_end, (size_t)64 + (size_t)nq * 16));
    if (rc) return rc;
    CK(cudaStreamSynchronize(s));  // the block may still feed an earlier call's copies
    int* n_st = reinterpret_cast<int*>(h->x_stage);
    n_st[0] = nq;
    n_st[1] = nt;
    CK(cudaMemcpyAsync(h->m_n, n_st, 2 * sizeof(int), cudaMemcpyHostToDevice, s));
    if (nq) {
        memcpy(h->x_stage + q_off, q, (size_t)nq * 32);
        CK(cudaMemcpyAsync(h->m_q, h->x_stage + q_off, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
    }
    if (nt) {
        memcpy(h->x_stage + t_off, t, (size_t)nt * 32);
        CK(cudaMemcpyAsync(h->m_t, h->x_stage + t_off, (size_t)nt * 32, cudaMemcpyHostToDevice, s));
    }
    if (win) {
        if (nq) {
            memcpy(h->x_stage + qxy_off, qxy, (size_t)nq * 8);
            CK(cudaMemcpyAsync(h->m_qxy, h->x_stage + qxy_off, (size_t)nq * 8, cudaMemcpyHostToDevice, s));
        }
        if (nt) {
            memcpy(h->x_stage + txy_off, txy, (size_t)nt * 8);
            CK(cudaMemcpyAsync(h->m_txy, h->x_stage + txy_off, (size_t)nt * 8, cudaMemcpyHostToDevice, s));
        }
    }
    const DevSet dq{h->m_q, win ? h->m_qxy : nullptr, h->m_n, nq}, dt{h->m_t, win ? h->m_txy : nullptr, h->m_n + 1, nt};
    return run_knn_device(h, dq, dt, radius, reverse_too, box);
}

int run_match_host(vxorb_handle* h, const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy,
                   int nt, float radius, float nn_ratio, int mode, int out_cap, vxorb_match_t* out, int* n_out,
                   const float* box = nullptr) {
    if (!n_out || out_cap < 0 || (out_cap > 0 && !out)) return fail(h, VXORB_E_INVALID, "bad output arguments");
    *n_out = 0;
    if (nq <= 0 || nt <= 0) {  // orb_matcher.cpp:18-20: either side empty -> no matches
        if (nq < 0 || nt < 0) return fail(h, VXORB_E_INVALID, "negative descriptor count");
        return VXORB_OK;
    }
    int rc = run_knn_host(h, q, qxy, nq, t, txy, nt, radius, mode == 2, box);
    if (rc) return rc;
    const DevSet dq{h->m_q, nullptr, h->m_n, nq};
    return finish_match(h, dq, nn_ratio, mode, out_cap, out, n_out);
}

int ensure_stage_scratch(vxorb_handle* h, size_t bytes) {
    if (bytes <= h->s_bytes) return VXORB_OK;
    cudaFree(h->s_a); cudaFree(h->s_b);
    h->s_a = h->s_b = nullptr;
    h->s_bytes = 0;
    CK(cudaMalloc(&h->s_a, bytes));
    CK(cudaMalloc(&h->s_b, bytes));
    h->s_bytes = bytes;
    return VXORB_OK;
}

}  // namespace

// ================================================================================================
extern "C" {

const char* vxorb_version(void) { return "vxorb 0.1 (sm_100a)"; }

void vxorb_default_params(vxorb_params* p) {
    if (!p) return;
    p->n_features = 1000;  // orb_extractor.h:11
    p->scale_factor = 1.2f;
    p->n_levels = 8;
    p->edge_threshold = 31;
    p->patch_size = 31;
    p->fast_threshold = 20;
    p->nn_ratio = 0.8f;    // orb_matcher.h:14
    p->min_matches = 50;   // orb_matcher.h:15
    p->device_id = 0;
    p->max_width = 640;
    p->max_height = 480;
    p->max_batch = 1;
    p->max_keypoints = 0;
    p->max_train = 16384;
}

const char* vxorb_last_error(const vxorb_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int vxorb_create(const vxorb_params* p, vxorb_handle** out) {
    std::lock_guard<std::mutex> lock(g_create_mutex);
    if (!p || !out) {
        g_create_error = "null argument";
        return VXORB_E_INVALID;
    }
    *out = nullptr;
    if (p->n_features < 1 || p->n_levels < 1 || p->n_levels > kMaxLevels || !(p->scale_factor > 1.0f) ||
        p->edge_threshold != 31 || p->patch_size != 31 || p->fast_threshold < 1 || p->fast_threshold > 254 ||
        p->max_batch < 1 || p->max_width < 1 || p->max_height < 1 || p->max_train < 1) {
        g_create_error = "unsupported parameters (edge_threshold and patch_size must be 31; scale_factor > 1)";
        return VXORB_E_INVALID;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                         " — this library has no CPU fallback";
        return VXORB_E_CUDA;
    }
    if (p->device_id < 0 || p->device_id >= ndev) {
        g_create_error = "device_id out of range";
        return VXORB_E_INVALID;
    }
    vxorb_handle* h = new (std::nothrow) vxorb_handle();
    if (!h) {
        g_create_error = "out of host memory";
        return VXORB_E_NOMEM;
    }
    h->p = *p;
    h->no_graph = getenv("VXORB_NO_GRAPH") != nullptr;
    h->no_overlap = getenv("VXORB_NO_OVERLAP") != nullptr;
    h->no_flat_upload = getenv("VXORB_NO_FLAT_UPLOAD") != nullptr;
    h->full_blur = getenv("VXORB_FULL_BLUR") != nullptr;
    h->kp_cap = p->max_keypoints > 0 ? p->max_keypoints : p->n_features + p->n_features / 4 + 64;
    auto bail = [&](cudaError_t err, const char* what) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
        vxorb_destroy(h);
        return VXORB_E_CUDA;
    };
    if ((e = cudaSetDevice(p->device_id)) != cudaSuccess) return bail(e, "cudaSetDevice");
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, p->device_id)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
    if (prop.major < 10) {
        g_create_error = "device is not sm_100 class; this library ships sm_100a code only";
        vxorb_destroy(h);
        return VXORB_E_CUDA;
    }
    for (auto& w : h->ws) {
        if ((e = cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "stream");
        if ((e = cudaEventCreateWithFlags(&w.done, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "event");
        if ((e = cudaStreamCreateWithFlags(&w.side, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "side stream");
        if ((e = cudaEventCreateWithFlags(&w.ev_fork, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "event");
        if ((e = cudaEventCreateWithFlags(&w.ev_join, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "event");
        if ((e = cudaHostAlloc(&w.h_status, sizeof(int) * 4, cudaHostAllocDefault)) != cudaSuccess) return bail(e, "pinned");
        w.h_status[0] = 0;
        if ((e = cudaHostAlloc(&w.h_counts, sizeof(int) * 2 * (size_t)p->max_batch, cudaHostAllocDefault)) != cudaSuccess)
            return bail(e, "pinned");
    }
    if ((e = cudaEventCreate(&h->t_begin)) != cudaSuccess) return bail(e, "event");
    if ((e = cudaEventCreate(&h->t_end)) != cudaSuccess) return bail(e, "event");
    if ((e = init_device_tables()) != cudaSuccess) return bail(e, "pattern upload");
    if ((e = init_match_kernels()) != cudaSuccess) return bail(e, "match kernel attributes");
    if ((e = init_pyramid_kernels()) != cudaSuccess) return bail(e, "pyramid kernel attributes");
    *out = h;
    return VXORB_OK;
}

void vxorb_destroy(vxorb_handle* h) {
    if (!h) return;
    cudaSetDevice(h->p.device_id);
    cudaDeviceSynchronize();
    free_geometry(h);
    for (auto& w : h->ws) {
        if (w.stream) cudaStreamDestroy(w.stream);
        if (w.done) cudaEventDestroy(w.done);
        if (w.side) cudaStreamDestroy(w.side);
        if (w.ev_fork) cudaEventDestroy(w.ev_fork);
        if (w.ev_join) cudaEventDestroy(w.ev_join);
        if (w.h_status) cudaFreeHost(w.h_status);
        if (w.h_counts) cudaFreeHost(w.h_counts);
    }
    for (auto& r : h->ring) {
        cudaFree(r.block); cudaFree(r.xy);
    }
    for (auto& kv : h->sets) {
        cudaFree(kv.second.desc); cudaFree(kv.second.xy); cudaFree(kv.second.n);
    }
    cudaFree(h->m_q); cudaFree(h->m_t); cudaFree(h->m_qxy); cudaFree(h->m_txy); cudaFree(h->m_n);
    cudaFree(h->m_knn); cudaFree(h->m_back); cudaFree(h->m_out_blk); cudaFree(h->m_part);
    cudaFree(h->s_a); cudaFree(h->s_b); cudaFree(h->flush_buf); cudaFree(h->m_mask);
    if (h->x_stage) cudaFreeHost(h->x_stage);
    for (auto& sp : h->spans) {
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    for (auto e : h->event_pool) cudaEventDestroy(e);
    if (h->t_begin) cudaEventDestroy(h->t_begin);
    if (h->t_end) cudaEventDestroy(h->t_end);
    delete h;
}

int vxorb_level_geometry(const vxorb_handle* h, int width, int height, int* level_w, int* level_h,
                         float* level_scale, int* level_quota) {
    if (!h || width < 1 || height < 1) return VXORB_E_INVALID;
    int lw[kMaxLevels], lh[kMaxLevels], q[kMaxLevels];
    float ls[kMaxLevels];
    level_geometry(h->p, width, height, lw, lh, ls, q);
    for (int l = 0; l < h->p.n_levels; ++l) {
        if (level_w) level_w[l] = lw[l];
        if (level_h) level_h[l] = lh[l];
        if (level_scale) level_scale[l] = ls[l];
        if (level_quota) level_quota[l] = q[l];
    }
    return VXORB_OK;
}

// ---- extraction ----------------------------------------------------------------------------------

static int extract_common(vxorb_handle* h, const uint8_t* const* frames, const uint8_t* bgr, int batch, int W, int H,
                          size_t stride, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out,
                          vxorb_frame_info* info_out = nullptr, long long* id_out = nullptr) {
    if (!h) return VXORB_E_INVALID;
    if (batch < 1 || batch > h->p.max_batch) return fail(h, VXORB_E_INVALID, "batch outside 1..max_batch");
    if (!n_out || kp_cap < 0 || (kp_cap > 0 && (!kp_out || !desc_out)))
        return fail(h, VXORB_E_INVALID, "null output pointer");
    if (id_out) *id_out = 0;
    CK(cudaSetDevice(h->p.device_id));
    int rc = ensure_geometry(h, W, H);
    if (rc) return rc;
    Workspace& ws = h->ws[0];
    if (ws.in_flight) return fail(h, VXORB_E_INVALID, "slot 0 busy with a track submit; wait for it first");
    const LevelTab& lt = h->lt;
    uint8_t* l0 = ws.db.pyr + lt.off[0];
    // a single frame goes straight into the next entry of the resident ring (its id stays valid until kRing
    // more single-frame extractions have run); batches use the slot's frame-major buffers
    ResidentSet* rs = nullptr;
    if (batch == 1) {
        rc = ensure_ring(h);
        if (rc) return rc;
        rs = &h->ring[h->next_frame_id % kRing];
        rs->id = 0;  // invalid while it is being overwritten
        rs->n_host = 0;
    }
    if (bgr) {
        if (stride < (size_t)W * 3) return fail(h, VXORB_E_INVALID, "stride smaller than a BGR row");
        rc = ensure_stage_scratch(h, stride * (size_t)H);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->s_a, bgr, stride * (size_t)H, cudaMemcpyHostToDevice, ws.stream));
        h->launches += launch_gray_from_bgr(h->s_a, stride, 0, l0, W, H, lt.pitch[0], (size_t)lt.fstride[0], 1, ws.stream);
        h->stage_launches[VXORB_STAGE_PYRAMID] += 1;
    } else {
        if (stride < (size_t)W) return fail(h, VXORB_E_INVALID, "stride smaller than a row");
        for (int f = 0; f < batch; ++f) {
            if (!frames[f]) return fail(h, VXORB_E_INVALID, "null frame pointer");
            CK(cudaMemcpy2DAsync(l0 + (size_t)f * lt.fstride[0], lt.pitch[0], frames[f], stride, W, H,
                                 cudaMemcpyHostToDevice, ws.stream));
        }
    }
    if (batch == 1 && !h->stage_enabled && !h->no_graph) {
        rc = launch_extract_graph(h, ws, *rs);
    } else {
        rc = enqueue_extract(h, ws, batch, rs);
    }
    if (rc) return rc;
    // where the rows of frame 0 sit, and the stride to the next frame
    const vxorb_keypoint* d_kp = rs ? rs->kp : ws.db.kp + h->kp_cap;
    const uint8_t* d_desc = rs ? rs->desc : ws.db.desc + (size_t)h->kp_cap * 32;
    const int* d_n = rs ? rs->n : ws.db.kp_n + 1;
    const int* d_total = rs ? rs->n + 1 : ws.db.kp_total + 1;
    if (info_out) {  // no pairs here: zero match counts, then statistics + grid coverage of every frame
        CK(cudaMemsetAsync(ws.db.match_n, 0, sizeof(int) * batch, ws.stream));
        LevelTab ltb = h->lt;
        ltb.batch = batch;
        DeviceBuffers dbi = ws.db;
        if (rs) {  // the kernel reads frame f's positions / count at slot f + 1
            dbi.xy = rs->xy - (size_t)h->kp_cap * 2;
            dbi.kp_n = rs->n - 1;
        }
        const int nl = launch_frame_info(ltb, dbi, h->kp_cap, ws.sums, ws.info, nullptr, 5, 5, ws.stream);
        h->launches += nl;
        h->stage_launches[VXORB_STAGE_MATCH] += nl;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(info_out, ws.info, sizeof(vxorb_frame_info) * batch, cudaMemcpyDeviceToHost, ws.stream));
    }
    // One round trip: counts, status and the first min(kp_cap, capacity) rows of every frame land in the
    // handle's pinned staging block, one synchronisation, then the exact row counts are copied out on the host.
    const int rows = std::min(kp_cap, h->kp_cap);
    const size_t n_ints = (size_t)batch * 2 + 4;
    const size_t kp_off = align_up(n_ints * sizeof(int), 64);
    const size_t desc_off = kp_off + align_up(sizeof(vxorb_keypoint) * (size_t)rows * batch, 64);
    const size_t need = desc_off + (size_t)32 * rows * batch;
    rc = ensure_host_stage(h, need);
    if (rc) return rc;
    int* counts = reinterpret_cast<int*>(h->x_stage);
    vxorb_keypoint* st_kp = reinterpret_cast<vxorb_keypoint*>(h->x_stage + kp_off);
    uint8_t* st_desc = h->x_stage + desc_off;
    const bool one_copy = rs != nullptr && rows == h->kp_cap;  // the ring entry has the staging block's layout
    if (one_copy) {
        CK(cudaMemcpyAsync(h->x_stage, rs->block, need, cudaMemcpyDeviceToHost, ws.stream));
    } else {
        CK(cudaMemcpyAsync(counts, d_n, sizeof(int) * batch, cudaMemcpyDeviceToHost, ws.stream));
        CK(cudaMemcpyAsync(counts + batch, d_total, sizeof(int) * batch, cudaMemcpyDeviceToHost, ws.stream));
    }
    CK(cudaMemcpyAsync(counts + 2 * batch, ws.db.status, sizeof(int) * 4, cudaMemcpyDeviceToHost, ws.stream));
    if (rows > 0 && !one_copy) {
        // frame f's rows sit h->kp_cap rows apart on the device, `rows` apart in the staging block
        CK(cudaMemcpy2DAsync(st_kp, sizeof(vxorb_keypoint) * (size_t)rows, d_kp,
                             sizeof(vxorb_keypoint) * (size_t)h->kp_cap, sizeof(vxorb_keypoint) * (size_t)rows, batch,
                             cudaMemcpyDeviceToHost, ws.stream));
        CK(cudaMemcpy2DAsync(st_desc, (size_t)32 * rows, d_desc, (size_t)32 * h->kp_cap,
                             (size_t)32 * rows, batch, cudaMemcpyDeviceToHost, ws.stream));
    }
    CK(cudaStreamSynchronize(ws.stream));
    bool truncated = false;
    for (int f = 0; f < batch; ++f) {
        const int n = counts[f], total = counts[batch + f];
        n_out[f] = total;
        const int m = std::min(n, rows);
        if (total > m) truncated = true;
        if (m > 0) {
            memcpy(kp_out + (size_t)f * kp_cap, st_kp + (size_t)f * rows, sizeof(vxorb_keypoint) * (size_t)m);
            memcpy(desc_out + (size_t)f * kp_cap * 32, st_desc + (size_t)f * rows * 32, (size_t)m * 32);
        }
    }
    h->last_slot = -1;  // the synchronous extract calls do not take part in a tracked sequence
    const int status = counts[2 * batch];
    if (status & 3) return check_status(h, status);
    if (rs) {  // the set is complete: publish it
        rs->n_host = counts[0];
        rs->id = h->next_frame_id++;
        if (id_out) *id_out = rs->id;
    }
    if (truncated) return fail(h, VXORB_E_CAPACITY, "more keypoints than kp_cap / max_keypoints; output truncated");
    return VXORB_OK;
}

int vxorb_extract(vxorb_handle* h, const uint8_t* gray, int width, int height, size_t stride, int kp_cap,
                  vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!gray) return fail(h, VXORB_E_INVALID, "null image");
    return extract_common(h, &gray, nullptr, 1, width, height, stride, kp_cap, kp_out, desc_out, n_out);
}

int vxorb_extract_info(vxorb_handle* h, const uint8_t* img, int channels, int width, int height, size_t stride,
                       int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out, vxorb_frame_info* info_out) {
    if (!h) return VXORB_E_INVALID;
    if (!img || !info_out) return fail(h, VXORB_E_INVALID, "null image or info pointer");
    if (channels != 1 && channels != 3) return fail(h, VXORB_E_INVALID, "channels must be 1 (gray) or 3 (BGR)");
    if (channels == 3)
        return extract_common(h, nullptr, img, 1, width, height, stride, kp_cap, kp_out, desc_out, n_out, info_out);
    return extract_common(h, &img, nullptr, 1, width, height, stride, kp_cap, kp_out, desc_out, n_out, info_out);
}

int vxorb_extract_id(vxorb_handle* h, const uint8_t* img, int channels, int width, int height, size_t stride, int kp_cap,
                     vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out, vxorb_frame_info* info_out, long long* id_out) {
    if (!h) return VXORB_E_INVALID;
    if (!img || !id_out) return fail(h, VXORB_E_INVALID, "null image or id pointer");
    if (channels != 1 && channels != 3) return fail(h, VXORB_E_INVALID, "channels must be 1 (gray) or 3 (BGR)");
    if (channels == 3)
        return extract_common(h, nullptr, img, 1, width, height, stride, kp_cap, kp_out, desc_out, n_out, info_out, id_out);
    return extract_common(h, &img, nullptr, 1, width, height, stride, kp_cap, kp_out, desc_out, n_out, info_out, id_out);
}

int vxorb_extract_bgr(vxorb_handle* h, const uint8_t* bgr, int width, int height, size_t stride, int kp_cap,
                      vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!bgr) return fail(h, VXORB_E_INVALID, "null image");
    return extract_common(h, nullptr, bgr, 1, width, height, stride, kp_cap, kp_out, desc_out, n_out);
}

int vxorb_extract_batch(vxorb_handle* h, const uint8_t* const* frames, int batch, int width, int height,
                        size_t stride, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!frames) return fail(h, VXORB_E_INVALID, "null frame array");
    return extract_common(h, frames, nullptr, batch, width, height, stride, kp_cap, kp_out, desc_out, n_out);
}

// ---- matching ------------------------------------------------------------------------------------

static int knn_common(vxorb_handle* h, const uint8_t* q, const float* qxy, int nq, const uint8_t* t,
                      const float* txy, int nt, float radius, vxorb_knn2_t* knn_out, const float* box = nullptr) {
    if (!h) return VXORB_E_INVALID;
    if (nq > 0 && !knn_out) return fail(h, VXORB_E_INVALID, "null output");
    CK(cudaSetDevice(h->p.device_id));
    if (nq <= 0) return nq < 0 ? fail(h, VXORB_E_INVALID, "negative count") : VXORB_OK;
    int rc = run_knn_host(h, q, qxy, nq, t, txy, nt, radius, false, box);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    CK(cudaMemcpyAsync(knn_out, h->m_knn, sizeof(vxorb_knn2_t) * (size_t)nq, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return VXORB_OK;
}

int vxorb_knn2(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, vxorb_knn2_t* knn_out) {
    return knn_common(h, query, nullptr, nq, train, nullptr, nt, 0.f, knn_out);
}

int vxorb_knn2_windowed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq, const uint8_t* train,
                        const float* train_xy, int nt, float radius, vxorb_knn2_t* knn_out) {
    if (h && (!query_xy || !train_xy || !(radius > 0.f)))
        return fail(h, VXORB_E_INVALID, "windowed match needs positions and radius > 0");
    return knn_common(h, query, query_xy, nq, train, train_xy, nt, radius, knn_out);
}

static bool box_ok(const float* box) {
    return box && box[0] <= box[1] && box[2] <= box[3];  // also rejects NaN bounds
}

int vxorb_knn2_boxed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq, const uint8_t* train,
                     const float* train_xy, int nt, const float* box, vxorb_knn2_t* knn_out) {
    if (h && (!query_xy || !train_xy || !box_ok(box)))
        return fail(h, VXORB_E_INVALID, "boxed match needs positions and box = {dx_min <= dx_max, dy_min <= dy_max}");
    return knn_common(h, query, query_xy, nq, train, train_xy, nt, 0.f, knn_out, box);
}

int vxorb_match_boxed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq, const uint8_t* train,
                      const float* train_xy, int nt, const float* box, float nn_ratio, int out_cap,
                      vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!query_xy || !train_xy || !box_ok(box))
        return fail(h, VXORB_E_INVALID, "boxed match needs positions and box = {dx_min <= dx_max, dy_min <= dy_max}");
    CK(cudaSetDevice(h->p.device_id));
    return run_match_host(h, query, query_xy, nq, train, train_xy, nt, 0.f, nn_ratio, 0, out_cap, out, n_out, box);
}

// Arbitrary mask = cv::BFMatcher::knnMatch(q, t, 2, mask): uploads the nq x nt byte mask next to the descriptors.
static int masked_common(vxorb_handle* h, const uint8_t* q, int nq, const uint8_t* t, int nt, const uint8_t* mask,
                         size_t mask_stride) {
    if (!mask || mask_stride < (size_t)nt) return fail(h, VXORB_E_INVALID, "masked match needs an nq x nt mask (stride >= nt)");
    CK(cudaSetDevice(h->p.device_id));
    const size_t bytes = (size_t)nq * nt;
    if (bytes > h->m_mask_bytes) {
        cudaFree(h->m_mask);
        h->m_mask = nullptr;
        h->m_mask_bytes = 0;
        CK(cudaMalloc(&h->m_mask, bytes));
        h->m_mask_bytes = bytes;
    }
    cudaStream_t s = h->ws[0].stream;
    CK(cudaMemcpy2DAsync(h->m_mask, (size_t)nt, mask, mask_stride, (size_t)nt, (size_t)nq, cudaMemcpyHostToDevice, s));
    h->pending_mask = h->m_mask;
    const int rc = run_knn_host(h, q, nullptr, nq, t, nullptr, nt, 0.f, false);
    h->pending_mask = nullptr;
    return rc;
}

int vxorb_knn2_masked(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, const uint8_t* mask,
                      size_t mask_stride, vxorb_knn2_t* knn_out) {
    if (!h) return VXORB_E_INVALID;
    if (nq > 0 && !knn_out) return fail(h, VXORB_E_INVALID, "null output");
    if (nq <= 0) return nq < 0 ? fail(h, VXORB_E_INVALID, "negative count") : VXORB_OK;
    if (nt <= 0) {  // no train rows: every knn row is empty
        if (nt < 0) return fail(h, VXORB_E_INVALID, "negative count");
        for (int i = 0; i < nq; ++i) knn_out[i] = vxorb_knn2_t{{-1, -1}, {0, 0}};
        return VXORB_OK;
    }
    int rc = masked_common(h, query, nq, train, nt, mask, mask_stride);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    CK(cudaMemcpyAsync(knn_out, h->m_knn, sizeof(vxorb_knn2_t) * (size_t)nq, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return VXORB_OK;
}

int vxorb_match_masked(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, const uint8_t* mask,
                       size_t mask_stride, float nn_ratio, int out_cap, vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!n_out || out_cap < 0 || (out_cap > 0 && !out)) return fail(h, VXORB_E_INVALID, "bad output arguments");
    *n_out = 0;
    if (nq <= 0 || nt <= 0) return (nq < 0 || nt < 0) ? fail(h, VXORB_E_INVALID, "negative descriptor count") : VXORB_OK;
    int rc = masked_common(h, query, nq, train, nt, mask, mask_stride);
    if (rc) return rc;
    const DevSet dq{h->m_q, nullptr, h->m_n, nq};
    return finish_match(h, dq, nn_ratio, 0, out_cap, out, n_out);
}

// Tracking::CreateLandmarksFromDepth's look-up (tracking.cpp:606-637) for a frame whose positions are still on the
// device: only the depth image goes up and one double per feature comes back.
int vxorb_depth_lookup(vxorb_handle* h, long long frame_id, const void* depth, int depth_type, int width, int height,
                       size_t stride, int cap, double* depth_m_out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    const size_t elem = depth_type == 0 ? 2 : depth_type == 1 ? 4 : 8;
    if (!depth || !n_out || depth_type < 0 || depth_type > 2 || width < 1 || height < 1 || stride < elem * (size_t)width ||
        cap < 0 || (cap > 0 && !depth_m_out))
        return fail(h, VXORB_E_INVALID, "bad depth arguments (depth_type 0 = u16, 1 = f32, 2 = f64)");
    *n_out = 0;
    const ResidentSet* fs = find_set(h, frame_id);
    if (!fs || !fs->xy) return fail(h, VXORB_E_INVALID, "unknown or overwritten frame id (or a set without positions)");
    CK(cudaSetDevice(h->p.device_id));
    const int n = fs->n_host;
    *n_out = n;
    if (n == 0) return VXORB_OK;
    const size_t img = stride * (size_t)height, out_off = align_up(img, 256);
    int rc = ensure_stage_scratch(h, out_off + sizeof(double) * (size_t)n);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    CK(cudaMemcpyAsync(h->s_a, depth, img, cudaMemcpyHostToDevice, s));
    double* d_out = reinterpret_cast<double*>(h->s_a + out_off);
    h->launches += launch_depth_lookup(fs->xy, fs->n, n, h->s_a, depth_type, width, height, stride, d_out, s);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(depth_m_out, d_out, sizeof(double) * (size_t)std::min(n, cap), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return n > cap ? fail(h, VXORB_E_CAPACITY, "depth output buffer too small") : VXORB_OK;
}

int vxorb_match(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, float nn_ratio,
                int out_cap, vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    return run_match_host(h, query, nullptr, nq, train, nullptr, nt, 0.f, nn_ratio, 0, out_cap, out, n_out);
}

int vxorb_match_windowed(vxorb_handle* h, const uint8_t* query, const float* query_xy, int nq, const uint8_t* train,
                         const float* train_xy, int nt, float radius, float nn_ratio, int out_cap,
                         vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!query_xy || !train_xy || !(radius > 0.f))
        return fail(h, VXORB_E_INVALID, "windowed match needs positions and radius > 0");
    CK(cudaSetDevice(h->p.device_id));
    return run_match_host(h, query, query_xy, nq, train, train_xy, nt, radius, nn_ratio, 0, out_cap, out, n_out);
}

int vxorb_match_crosscheck(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, int out_cap,
                           vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    return run_match_host(h, query, nullptr, nq, train, nullptr, nt, 0.f, 1.f, 2, out_cap, out, n_out);
}

int vxorb_match_filtered(vxorb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, float nn_ratio,
                         int out_cap, vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    return run_match_host(h, query, nullptr, nq, train, nullptr, nt, 0.f, nn_ratio, 1, out_cap, out, n_out);
}

// ---- device-resident descriptor sets ---------------------------------------------------------------

int vxorb_set_upload(vxorb_handle* h, const uint8_t* desc, const float* xy, int n, long long* id_out) {
    if (!h) return VXORB_E_INVALID;
    if (!id_out || n < 0 || (n > 0 && !desc) || n >= (1 << 20)) return fail(h, VXORB_E_INVALID, "bad set arguments");
    *id_out = 0;
    CK(cudaSetDevice(h->p.device_id));
    ResidentSet rs;
    const size_t rows = (size_t)std::max(n, 1);
    auto bail = [&](const char* what) {
        cudaFree(rs.desc); cudaFree(rs.xy); cudaFree(rs.n);
        cudaGetLastError();
        return fail(h, VXORB_E_NOMEM, what);
    };
    if (cudaMalloc(&rs.desc, rows * 32) != cudaSuccess) return bail("device allocation of the descriptor set failed");
    if (xy && cudaMalloc(&rs.xy, rows * 8) != cudaSuccess) return bail("device allocation of the set positions failed");
    if (cudaMalloc(&rs.n, sizeof(int) * 2) != cudaSuccess) return bail("device allocation of the set count failed");
    const int counts[2] = {n, n};
    CK(cudaMemcpy(rs.n, counts, sizeof(counts), cudaMemcpyHostToDevice));
    if (n) {
        CK(cudaMemcpy(rs.desc, desc, (size_t)n * 32, cudaMemcpyHostToDevice));
        if (xy) CK(cudaMemcpy(rs.xy, xy, (size_t)n * 8, cudaMemcpyHostToDevice));
    }
    rs.n_host = n;
    rs.id = h->next_set_id++;
    h->sets[rs.id] = rs;
    *id_out = rs.id;
    return VXORB_OK;
}

int vxorb_set_free(vxorb_handle* h, long long id) {
    if (!h) return VXORB_E_INVALID;
    auto it = h->sets.find(id);
    if (it == h->sets.end()) return fail(h, VXORB_E_INVALID, "unknown set id");
    CK(cudaSetDevice(h->p.device_id));
    CK(cudaStreamSynchronize(h->ws[0].stream));
    cudaFree(it->second.desc); cudaFree(it->second.xy); cudaFree(it->second.n);
    h->sets.erase(it);
    return VXORB_OK;
}

int vxorb_set_size(const vxorb_handle* h, long long id) {
    if (!h) return VXORB_E_INVALID;
    const ResidentSet* s = find_set(h, id);
    return s ? s->n_host : VXORB_E_INVALID;
}

int vxorb_match_ids(vxorb_handle* h, long long query_id, long long train_id, float radius, float nn_ratio, int mode,
                    int out_cap, vxorb_match_t* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!n_out || out_cap < 0 || (out_cap > 0 && !out) || mode < 0 || mode > 2)
        return fail(h, VXORB_E_INVALID, "bad output arguments or mode");
    *n_out = 0;
    const ResidentSet* q = find_set(h, query_id);
    const ResidentSet* t = find_set(h, train_id);
    if (!q || !t) return fail(h, VXORB_E_INVALID, "unknown or overwritten set id (frame ids live for 8 single-frame extractions)");
    if (radius > 0.f && (!q->xy || !t->xy)) return fail(h, VXORB_E_INVALID, "windowed match needs positions on both sets");
    if (q->n_host <= 0 || t->n_host <= 0) return VXORB_OK;  // orb_matcher.cpp:18-20
    CK(cudaSetDevice(h->p.device_id));
    const DevSet dq{q->desc, q->xy, q->n, q->n_host}, dt{t->desc, t->xy, t->n, t->n_host};
    int rc = run_knn_device(h, dq, dt, radius, mode == 2, nullptr);
    if (rc) return rc;
    return finish_match(h, dq, mode == 2 ? 1.f : nn_ratio, mode, out_cap, out, n_out);
}

// ---- streaming sequence path ---------------------------------------------------------------------

int vxorb_track_slots(const vxorb_handle* h) { return h ? kSlots : 0; }
int vxorb_keypoint_capacity(const vxorb_handle* h) { return h ? h->kp_cap : 0; }

int vxorb_track_reset(vxorb_handle* h) {
    if (!h) return VXORB_E_INVALID;
    for (auto& w : h->ws)
        if (w.in_flight) return fail(h, VXORB_E_INVALID, "reset while a slot is in flight");
    h->last_slot = -1;
    return VXORB_OK;
}

// Carry of the sequence across chunks.  The last frame of every chunk is copied to the handle's
// carry buffers on the producing stream (before its `done` event); the next chunk, on whatever
// stream it runs, waits for that event and copies the carry into its own frame slot 0.  Reads and
// writes of the carry are therefore ordered by stream order + the done events alone, independent
// of when the caller waits on a slot.
static int enqueue_carry_in(vxorb_handle* h, Workspace& ws) {
    cudaStream_t s = ws.stream;
    const size_t cap = (size_t)h->kp_cap;
    if (h->last_slot < 0) {
        CK(cudaMemsetAsync(ws.db.kp_n, 0, sizeof(int), s));
        return VXORB_OK;
    }
    Workspace& prev = h->ws[h->last_slot];
    if (&prev != &ws) CK(cudaStreamWaitEvent(s, prev.done, 0));
    CK(cudaMemcpyAsync(ws.db.desc, h->carry_desc, cap * 32, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(ws.db.xy, h->carry_xy, cap * 8, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(ws.db.kp_n, h->carry_n, sizeof(int), cudaMemcpyDeviceToDevice, s));
    return VXORB_OK;
}

static int enqueue_carry_out(vxorb_handle* h, Workspace& ws, int batch) {
    cudaStream_t s = ws.stream;
    const size_t cap = (size_t)h->kp_cap;
    CK(cudaMemcpyAsync(h->carry_desc, ws.db.desc + cap * 32 * batch, cap * 32, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(h->carry_xy, ws.db.xy + cap * 2 * batch, cap * 8, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(h->carry_n, ws.db.kp_n + batch, sizeof(int), cudaMemcpyDeviceToDevice, s));
    return VXORB_OK;
}

// Extraction, carry and pair matching of one chunk whose level 0 is already in ws.  The carry leaves
// as soon as the descriptors exist (the next chunk, on another slot, only needs this chunk's last
// frame), so `done` is recorded before this chunk's own matching starts.
static int enqueue_chunk(vxorb_handle* h, Workspace& ws, int batch, float radius, const float* stereo_box = nullptr) {
    int rc = enqueue_extract(h, ws, batch);
    if (rc) return rc;
    rc = enqueue_carry_in(h, ws);
    if (rc) return rc;
    rc = enqueue_carry_out(h, ws, stereo_box ? batch - 1 : batch);  // stereo: the last LEFT frame carries
    if (rc) return rc;
    CK(cudaEventRecord(ws.done, ws.stream));  // descriptors and the carry of this chunk are final
    if (stereo_box) return enqueue_stereo_match(h, ws, batch / 2, stereo_box, radius);
    return enqueue_track_match(h, ws, batch, radius);
}

int vxorb_track_submit(vxorb_handle* h, const uint8_t* frames, int batch, int width, int height, size_t stride,
                       size_t frame_stride, float radius, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out,
                       int* n_kp_out, int match_cap, vxorb_match_t* match_out, int* n_match_out, int* slot_out) {
    return vxorb_track_submit_info(h, frames, batch, width, height, stride, frame_stride, radius, kp_cap, kp_out, desc_out,
                                   n_kp_out, match_cap, match_out, n_match_out, nullptr, nullptr, slot_out);
}

static int track_submit_common(vxorb_handle* h, const uint8_t* frames, int channels, int batch, int width, int height,
                               size_t stride, size_t frame_stride, float radius, int kp_cap, vxorb_keypoint* kp_out,
                               uint8_t* desc_out, int* n_kp_out, int match_cap, vxorb_match_t* match_out,
                               int* n_match_out, vxorb_frame_info* info_out, float* pts_out, int* slot_out,
                               const float* stereo_box = nullptr) {
    if (!h) return VXORB_E_INVALID;
    if (!frames || !slot_out) return fail(h, VXORB_E_INVALID, "null argument");
    if (channels != 1 && channels != 3) return fail(h, VXORB_E_INVALID, "channels must be 1 (gray) or 3 (BGR)");
    if (stereo_box && (batch % 2 != 0 || !(stereo_box[0] <= stereo_box[1]) || !(stereo_box[2] <= stereo_box[3])))
        return fail(h, VXORB_E_INVALID, "stereo chunks hold (left, right) pairs and need box = {dx_min <= dx_max, dy_min <= dy_max}");
    if (batch < 1 || batch > h->p.max_batch) return fail(h, VXORB_E_INVALID, "batch outside 1..max_batch");
    if (width < 1 || height < 1 || stride < (size_t)width * channels) return fail(h, VXORB_E_INVALID, "stride smaller than a row");
    if ((kp_out || desc_out) && kp_cap != h->kp_cap)
        return fail(h, VXORB_E_INVALID, "track outputs must use kp_cap == the handle's max_keypoints");
    if ((match_out || pts_out) && match_cap != h->kp_cap)
        return fail(h, VXORB_E_INVALID, "track outputs must use match_cap == the handle's max_keypoints");
    CK(cudaSetDevice(h->p.device_id));
    int rc = ensure_geometry(h, width, height);
    if (rc) return rc;
    const int slot = h->next_slot;
    Workspace& ws = h->ws[slot];
    if (ws.in_flight) return fail(h, VXORB_E_INVALID, "all pipeline slots in flight; call vxorb_track_wait first");
    cudaStream_t s = ws.stream;
    // Read-back: the first `head` rows of every frame (a strided copy into the caller's [batch][cap] arrays) —
    // cv::ORB returns n_features keypoints unless response ties push a level over its quota, so the rows
    // between n_features and the capacity are almost always padding and stay on the device; vxorb_track_wait
    // fetches the tail of a frame that did overflow.
    const size_t cap = (size_t)h->kp_cap;
    const size_t head = std::min<size_t>(cap, (size_t)h->p.n_features);
    auto rows_d2h = [&](void* dst, const void* src, size_t row_bytes) {
        return cudaMemcpy2DAsync(dst, cap * row_bytes, src, cap * row_bytes, head * row_bytes, (size_t)batch,
                                 cudaMemcpyDeviceToHost, s);
    };
    auto queue = [&]() -> int {
        int rc2 = channels == 3 ? upload_bgr_frames(h, ws, frames, batch, width, height, stride, frame_stride)
                                : upload_frames(h, ws, frames, batch, width, height, stride, frame_stride);
        if (rc2) return rc2;
        rc2 = enqueue_chunk(h, ws, batch, radius, stereo_box);
        if (rc2) return rc2;
        if (info_out || pts_out) {
            rc2 = enqueue_frame_info(h, ws, batch, pts_out != nullptr);
            if (rc2) return rc2;
        }
        if (kp_out) CK(rows_d2h(kp_out, ws.db.kp + cap, sizeof(vxorb_keypoint)));
        if (desc_out) CK(rows_d2h(desc_out, ws.db.desc + cap * 32, 32));
        if (n_kp_out) CK(cudaMemcpyAsync(n_kp_out, ws.db.kp_total + 1, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
        if (match_out) CK(rows_d2h(match_out, ws.db.match, sizeof(vxorb_match_t)));
        if (n_match_out) CK(cudaMemcpyAsync(n_match_out, ws.db.match_n, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
        if (info_out) CK(cudaMemcpyAsync(info_out, ws.info, sizeof(vxorb_frame_info) * batch, cudaMemcpyDeviceToHost, s));
        if (pts_out) CK(rows_d2h(pts_out, ws.pts, sizeof(float4)));
        // counts of this chunk for the tail check in vxorb_track_wait (pinned, next to the status word)
        CK(cudaMemcpyAsync(ws.h_counts, ws.db.kp_n + 1, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(ws.h_counts + h->p.max_batch, ws.db.match_n, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(ws.h_status, ws.db.status, sizeof(int) * 4, cudaMemcpyDeviceToHost, s));
        return VXORB_OK;
    };
    rc = queue();
    if (rc) {
        // part of the chunk may already be queued (possibly a carry-out that overwrote the handle's carry): drain the
        // slot and restart the sequence, so the next submit neither races these buffers nor matches against a
        // carry that belongs to no completed chunk
        const std::string msg = h->err;
        cudaStreamSynchronize(s);
        cudaGetLastError();
        h->last_slot = -1;
        h->err = msg + " (chunk dropped; the tracked sequence was reset)";
        return rc;
    }
    ws.batch = batch;
    ws.in_flight = true;
    ws.tail = {kp_out, desc_out, match_out, pts_out, head};
    h->last_slot = slot;
    h->next_slot = (slot + 1) % kSlots;
    *slot_out = slot;
    return VXORB_OK;
}

int vxorb_track_submit_info(vxorb_handle* h, const uint8_t* frames, int batch, int width, int height, size_t stride,
                            size_t frame_stride, float radius, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out,
                            int* n_kp_out, int match_cap, vxorb_match_t* match_out, int* n_match_out,
                            vxorb_frame_info* info_out, float* pts_out, int* slot_out) {
    return track_submit_common(h, frames, 1, batch, width, height, stride, frame_stride, radius, kp_cap, kp_out, desc_out,
                               n_kp_out, match_cap, match_out, n_match_out, info_out, pts_out, slot_out);
}

int vxorb_track_submit_bgr(vxorb_handle* h, const uint8_t* bgr_frames, int batch, int width, int height, size_t stride,
                           size_t frame_stride, float radius, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out,
                           int* n_kp_out, int match_cap, vxorb_match_t* match_out, int* n_match_out,
                           vxorb_frame_info* info_out, float* pts_out, int* slot_out) {
    return track_submit_common(h, bgr_frames, 3, batch, width, height, stride, frame_stride, radius, kp_cap, kp_out,
                               desc_out, n_kp_out, match_cap, match_out, n_match_out, info_out, pts_out, slot_out);
}

int vxorb_track_submit_stereo(vxorb_handle* h, const uint8_t* frames, int pairs, int width, int height, size_t stride,
                              size_t frame_stride, const float* box, float radius, int kp_cap, vxorb_keypoint* kp_out,
                              uint8_t* desc_out, int* n_kp_out, int match_cap, vxorb_match_t* match_out, int* n_match_out,
                              int* slot_out) {
    if (!h) return VXORB_E_INVALID;
    if (!box || pairs < 1) return fail(h, VXORB_E_INVALID, "stereo submit needs pairs >= 1 and a box");
    return track_submit_common(h, frames, 1, 2 * pairs, width, height, stride, frame_stride, radius, kp_cap, kp_out, desc_out,
                               n_kp_out, match_cap, match_out, n_match_out, nullptr, nullptr, slot_out, box);
}

int vxorb_track_wait(vxorb_handle* h, int slot) {
    if (!h) return VXORB_E_INVALID;
    if (slot < 0 || slot >= kSlots) return fail(h, VXORB_E_INVALID, "bad slot");
    Workspace& ws = h->ws[slot];
    if (!ws.in_flight) return fail(h, VXORB_E_INVALID, "slot not in flight");
    CK(cudaSetDevice(h->p.device_id));
    CK(cudaStreamSynchronize(ws.stream));
    ws.in_flight = false;
    // frames whose keypoint or match rows ran past the `head` rows copied with the chunk (response ties): fetch
    // their tails now; the common case touches nothing
    const size_t cap = (size_t)h->kp_cap, head = ws.tail.head;
    bool extra = false;
    for (int f = 0; f < ws.batch; ++f) {
        const size_t nk = (size_t)std::min(ws.h_counts[f], h->kp_cap);
        const size_t nm = (size_t)std::min(ws.h_counts[h->p.max_batch + f], h->kp_cap);
        const size_t row = (size_t)f * cap + head;
        if (nk > head) {
            if (ws.tail.kp) CK(cudaMemcpyAsync(ws.tail.kp + row, ws.db.kp + cap + row, sizeof(vxorb_keypoint) * (nk - head), cudaMemcpyDeviceToHost, ws.stream));
            if (ws.tail.desc) CK(cudaMemcpyAsync(ws.tail.desc + row * 32, ws.db.desc + (cap + row) * 32, 32 * (nk - head), cudaMemcpyDeviceToHost, ws.stream));
            extra = true;
        }
        if (nm > head) {
            if (ws.tail.match) CK(cudaMemcpyAsync(ws.tail.match + row, ws.db.match + row, sizeof(vxorb_match_t) * (nm - head), cudaMemcpyDeviceToHost, ws.stream));
            if (ws.tail.pts) CK(cudaMemcpyAsync(ws.tail.pts + row * 4, reinterpret_cast<const float*>(ws.pts) + row * 4, sizeof(float4) * (nm - head), cudaMemcpyDeviceToHost, ws.stream));
            extra = true;
        }
    }
    if (extra) CK(cudaStreamSynchronize(ws.stream));
    return check_status(h, ws.h_status[0]);
}

// Resident mode keeps the same chunk in every pipeline slot; consecutive vxorb_resident_run calls take
// the slots in turn, exactly like consecutive vxorb_track_submit calls, so the kernels of up to kSlots
// passes are in flight together (each pass still waits for its predecessor's last-frame descriptors
// before its first pair is matched).  Nothing is copied from or to the host.
int vxorb_resident_upload(vxorb_handle* h, const uint8_t* frames, int batch, int width, int height, size_t stride,
                          size_t frame_stride) {
    if (!h) return VXORB_E_INVALID;
    if (!frames) return fail(h, VXORB_E_INVALID, "null frames");
    if (batch < 1 || batch > h->p.max_batch) return fail(h, VXORB_E_INVALID, "batch outside 1..max_batch");
    CK(cudaSetDevice(h->p.device_id));
    int rc = ensure_geometry(h, width, height);
    if (rc) return rc;
    for (auto& ws : h->ws)
        if (ws.in_flight) return fail(h, VXORB_E_INVALID, "a pipeline slot is in flight");
    for (auto& ws : h->ws) {
        rc = upload_frames(h, ws, frames, batch, width, height, stride, frame_stride);
        if (rc) return rc;
        CK(cudaMemsetAsync(ws.db.kp_n, 0, sizeof(int), ws.stream));
        CK(cudaStreamSynchronize(ws.stream));
        ws.batch = batch;
    }
    h->res_batch = batch;
    h->res_next = 0;
    h->res_last = -1;
    h->last_slot = -1;
    return VXORB_OK;
}

static int resident_run_common(vxorb_handle* h, float radius, const float* stereo_box);

int vxorb_resident_run(vxorb_handle* h, float radius) { return resident_run_common(h, radius, nullptr); }

int vxorb_resident_run_stereo(vxorb_handle* h, const float* box, float radius) {
    if (!h) return VXORB_E_INVALID;
    if (!box || !(box[0] <= box[1]) || !(box[2] <= box[3]) || h->res_batch % 2 != 0)
        return fail(h, VXORB_E_INVALID, "stereo pass needs an even number of resident frames and box = {dx_min <= dx_max, dy_min <= dy_max}");
    return resident_run_common(h, radius, box);
}

static int resident_run_common(vxorb_handle* h, float radius, const float* stereo_box) {
    if (!h) return VXORB_E_INVALID;
    if (h->res_batch < 1 || h->geo_w == 0) return fail(h, VXORB_E_INVALID, "no resident frames; call vxorb_resident_upload");
    CK(cudaSetDevice(h->p.device_id));
    // instrumented runs stay on slot 0 so that every stage's events sit on one stream, one pass at a time
    const int slot = h->stage_enabled ? 0 : h->res_next;
    Workspace& ws = h->ws[slot];
    if (ws.in_flight) return fail(h, VXORB_E_INVALID, "slot busy with a track submit");
    const int rc = enqueue_chunk(h, ws, h->res_batch, radius, stereo_box);
    if (rc) return rc;
    h->last_slot = slot;
    h->res_last = slot;
    h->res_next = (slot + 1) % kSlots;
    return VXORB_OK;
}

int vxorb_resident_fetch(vxorb_handle* h, int kp_cap, vxorb_keypoint* kp_out, uint8_t* desc_out, int* n_kp_out,
                         int match_cap, vxorb_match_t* match_out, int* n_match_out) {
    if (!h) return VXORB_E_INVALID;
    if (h->res_batch < 1 || h->res_last < 0) return fail(h, VXORB_E_INVALID, "no resident pass to fetch");
    Workspace& ws = h->ws[h->res_last];
    if ((kp_out || desc_out) && kp_cap != h->kp_cap) return fail(h, VXORB_E_INVALID, "kp_cap must equal max_keypoints");
    if (match_out && match_cap != h->kp_cap) return fail(h, VXORB_E_INVALID, "match_cap must equal max_keypoints");
    CK(cudaSetDevice(h->p.device_id));
    cudaStream_t s = ws.stream;
    const size_t cap = (size_t)h->kp_cap;
    const int batch = h->res_batch;
    int status[4] = {0, 0, 0, 0};
    if (kp_out) CK(cudaMemcpyAsync(kp_out, ws.db.kp + cap, sizeof(vxorb_keypoint) * cap * batch, cudaMemcpyDeviceToHost, s));
    if (desc_out) CK(cudaMemcpyAsync(desc_out, ws.db.desc + cap * 32, 32 * cap * batch, cudaMemcpyDeviceToHost, s));
    if (n_kp_out) CK(cudaMemcpyAsync(n_kp_out, ws.db.kp_total + 1, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
    if (match_out) CK(cudaMemcpyAsync(match_out, ws.db.match, sizeof(vxorb_match_t) * cap * batch, cudaMemcpyDeviceToHost, s));
    if (n_match_out) CK(cudaMemcpyAsync(n_match_out, ws.db.match_n, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(status, ws.db.status, sizeof(status), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return check_status(h, status[0]);
}

// ---- pinned host memory --------------------------------------------------------------------------

void* vxorb_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}

void vxorb_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

// ---- measurement ---------------------------------------------------------------------------------

int vxorb_sync(vxorb_handle* h) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    for (auto& w : h->ws) CK(cudaStreamSynchronize(w.stream));
    return VXORB_OK;
}

// Joins every slot stream into slot 0's stream and records there.
static int record_joined(vxorb_handle* h, cudaEvent_t ev) {
    cudaStream_t s0 = h->ws[0].stream;
    for (int i = 1; i < kSlots; ++i) {
        cudaEvent_t e = take_event(h);
        CK(cudaEventRecord(e, h->ws[i].stream));
        CK(cudaStreamWaitEvent(s0, e, 0));
        h->event_pool.push_back(e);  // safe: the wait has been enqueued
    }
    CK(cudaEventRecord(ev, s0));
    return VXORB_OK;
}

int vxorb_timer_begin(vxorb_handle* h) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    int rc = record_joined(h, h->t_begin);
    if (rc) return rc;
    // later work on the other slots must not start before the begin mark
    for (int i = 1; i < kSlots; ++i) CK(cudaStreamWaitEvent(h->ws[i].stream, h->t_begin, 0));
    return VXORB_OK;
}

int vxorb_timer_end(vxorb_handle* h, float* ms_out) {
    if (!h || !ms_out) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    int rc = record_joined(h, h->t_end);
    if (rc) return rc;
    CK(cudaEventSynchronize(h->t_end));
    CK(cudaEventElapsedTime(ms_out, h->t_begin, h->t_end));
    return VXORB_OK;
}

int vxorb_stage_times_reset(vxorb_handle* h, int enable) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    int rc = vxorb_sync(h);
    if (rc) return rc;
    for (auto& sp : h->spans) {
        h->event_pool.push_back(sp.a);
        h->event_pool.push_back(sp.b);
    }
    h->spans.clear();
    for (int i = 0; i < VXORB_STAGE_COUNT; ++i) {
        h->stage_ms[i] = 0.f;
        h->stage_launches[i] = 0;
    }
    h->stage_enabled = enable != 0;
    return VXORB_OK;
}

int vxorb_stage_times(vxorb_handle* h, float* ms_out, int* launches_out) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    int rc = vxorb_sync(h);
    if (rc) return rc;
    for (auto& sp : h->spans) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, sp.a, sp.b));
        h->stage_ms[sp.stage] += ms;
        h->event_pool.push_back(sp.a);
        h->event_pool.push_back(sp.b);
    }
    h->spans.clear();
    for (int i = 0; i < VXORB_STAGE_COUNT; ++i) {
        if (ms_out) ms_out[i] = h->stage_ms[i];
        if (launches_out) launches_out[i] = h->stage_launches[i];
    }
    return VXORB_OK;
}

long long vxorb_launch_count(const vxorb_handle* h) { return h ? h->launches : 0; }

int vxorb_flush_l2(vxorb_handle* h) {
    if (!h) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    if (!h->flush_buf) {
        h->flush_bytes = (size_t)256 << 20;  // 2x the 126 MB L2
        CK(cudaMalloc(&h->flush_buf, h->flush_bytes));
    }
    h->launches += launch_flush(h->flush_buf, h->flush_bytes, h->ws[0].stream);
    CK(cudaGetLastError());
    return VXORB_OK;
}

int vxorb_measure_popc_rate(vxorb_handle* h, double* popc_per_s_out) {
    if (!h || !popc_per_s_out) return VXORB_E_INVALID;
    CK(cudaSetDevice(h->p.device_id));
    int rc = ensure_match_scratch(h, 64);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    const int blocks = 148 * 8, iters = 4096;
    cudaEvent_t a = take_event(h), b = take_event(h);
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {  // first rep is the warm-up
        CK(cudaEventRecord(a, s));
        h->launches += launch_popc_peak(reinterpret_cast<uint32_t*>(h->m_out), blocks, iters, s);
        CK(cudaEventRecord(b, s));
        CK(cudaEventSynchronize(b));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, a, b));
        const double rate = (double)blocks * 256.0 * iters * 8.0 / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    h->event_pool.push_back(a);
    h->event_pool.push_back(b);
    CK(cudaGetLastError());
    *popc_per_s_out = best;
    return VXORB_OK;
}

// ---- stage isolation -----------------------------------------------------------------------------

int vxorb_debug_resize(vxorb_handle* h, const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw,
                       int dh, size_t dstride) {
    if (!h) return VXORB_E_INVALID;
    if (!src || !dst || sw < 1 || sh < 1 || dw < 1 || dh < 1 || sw >= 65536 || sh >= 65536)
        return fail(h, VXORB_E_INVALID, "bad resize arguments");
    CK(cudaSetDevice(h->p.device_id));
    // source and destination live in apron planes like the pyramid levels (the kernels read whole
    // aligned words around the taps)
    const int sp = plane_pitch(sw), dp = plane_pitch(dw);
    const size_t sbytes = align_up(plane_bytes(sw, sh), 256), dbytes = align_up(plane_bytes(dw, dh), 256);
    int rc = ensure_stage_scratch(h, std::max(sbytes, dbytes) + (size_t)(dw + dh) * 4 + 256);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    std::vector<uint32_t> taps((size_t)dw + dh);
    build_resize_taps(sw, dw, taps.data());
    build_resize_taps(sh, dh, taps.data() + dw);
    uint32_t* d_taps = reinterpret_cast<uint32_t*>(h->s_b + dbytes);
    uint8_t* s0 = h->s_a + plane_origin(sw);
    uint8_t* d0 = h->s_b + plane_origin(dw);
    CK(cudaMemsetAsync(h->s_a, 0, sbytes, s));
    CK(cudaMemcpy2DAsync(s0, sp, src, sstride, sw, sh, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_taps, taps.data(), taps.size() * 4, cudaMemcpyHostToDevice, s));
    h->launches += launch_resize_single(s0, sw, sh, sp, d0, dw, dh, dp, d_taps, d_taps + dw,
                                        resize_taps_fit_fast_path(taps.data(), dw), s);
    CK(cudaGetLastError());
    CK(cudaMemcpy2DAsync(dst, dstride, d0, dp, dw, dh, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return VXORB_OK;
}

// The production pyramid kernel (resize_band_kernel) on one image: band_rows 16 runs the throughput shape the
// batched path uses, 8 / 4 the latency shape of small batches and of wide levels.
int vxorb_debug_resize_band(vxorb_handle* h, const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw,
                            int dh, size_t dstride, int band_rows) {
    if (!h) return VXORB_E_INVALID;
    if (!src || !dst || sw < 1 || sh < 1 || dw < 8 || dh < 8 || sw >= 65536 || sh >= 65536 ||
        (band_rows != 4 && band_rows != 8 && band_rows != 16))
        return fail(h, VXORB_E_INVALID, "bad band-resize arguments (dw, dh >= 8; band_rows 4, 8 or 16)");
    CK(cudaSetDevice(h->p.device_id));
    const int sp = plane_pitch(sw), dp = plane_pitch(dw);
    const int words = (dw + 3) / 4;
    std::vector<uint32_t> taps((size_t)dw + dh);
    build_resize_taps(sw, dw, taps.data());
    build_resize_taps(sh, dh, taps.data() + dw);
    if (!resize_taps_fit_fast_path(taps.data(), dw))
        return fail(h, VXORB_E_INVALID, "scale too large for the band kernel (four output columns span more than 8 source bytes)");
    int nb = 0;
    long long smem = 0;
    resize_band_extent(taps.data() + dw, dh, sh, sp, band_rows, &nb, &smem);
    if (smem > 200 * 1024) return fail(h, VXORB_E_INVALID, "band does not fit shared memory");
    // device tables: ytab[dh] | pad to 16 B | uint4 xw4[words] | uint2 xw2[words]
    const size_t y_words = ((size_t)dh + 3) & ~(size_t)3;
    std::vector<uint32_t> tab(y_words + (size_t)words * 6);
    memcpy(tab.data(), taps.data() + dw, (size_t)dh * 4);
    build_resize_words(taps.data(), dw, tab.data() + y_words);
    const size_t sbytes = align_up(plane_bytes(sw, sh), 256), dbytes = align_up(plane_bytes(dw, dh), 256);
    int rc = ensure_stage_scratch(h, std::max(sbytes, dbytes) + tab.size() * 4 + 256);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    uint32_t* d_tab = reinterpret_cast<uint32_t*>(h->s_b + dbytes);
    uint8_t* d0 = h->s_b + plane_origin(dw);
    CK(cudaMemsetAsync(h->s_a, 0, sbytes, s));
    CK(cudaMemsetAsync(h->s_b, 0, dbytes, s));
    CK(cudaMemcpy2DAsync(h->s_a + plane_origin(sw), sp, src, sstride, sw, sh, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_tab, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, s));
    launch_resize_band(h->s_a, sh, sp, 0, d0, dw, dh, dp, 0, d_tab, reinterpret_cast<const uint4*>(d_tab + y_words),
                       reinterpret_cast<const uint2*>(d_tab + y_words + (size_t)words * 4), band_rows, nb, (int)smem, 1, s);
    h->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpy2DAsync(dst, dstride, d0, dp, dw, dh, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return VXORB_OK;
}

// FAST / blur stage tests run a one-level, one-frame LevelTab through the production kernels.
static int single_level_tab(vxorb_handle* h, int W, int H, int border, LevelTab* out, DeviceBuffers* db, size_t* cand_cap) {
    LevelTab lt;
    memset(&lt, 0, sizeof(lt));
    lt.n_levels = 1;
    lt.batch = 1;
    lt.w[0] = W;
    lt.h[0] = H;
    lt.pitch[0] = plane_pitch(W);
    lt.fstride[0] = plane_bytes(W, H);
    lt.off[0] = plane_origin(W);
    lt.scale[0] = 1.f;
    const int ew = W - 2 * border, eh = H - 2 * border;
    lt.active[0] = (ew > 0 && eh > 0) ? 1 : 0;
    lt.cand_cap[0] = lt.active[0] ? ((ew + 1) / 2 + 1) * ((eh + 1) / 2 + 1) : 0;
    lt.fast_tiles_x[0] = lt.active[0] ? fast_tiles_x(W, border) : 0;
    lt.fast_tile_begin[1] = lt.active[0] ? lt.fast_tiles_x[0] * fast_tiles_y(H, border) : 0;
    lt.blur_margin = 0;  // stage isolation: the whole image, from an aproned plane
    lt.blur_tile_begin[1] = (set_blur_extent(lt, 0, 0, true) + 255) / 256;
    for (int l = 1; l < kMaxLevels; ++l) {
        lt.fast_tile_begin[l + 1] = lt.fast_tile_begin[1];
        lt.blur_tile_begin[l + 1] = lt.blur_tile_begin[1];
    }
    const size_t img = align_up((size_t)lt.fstride[0], 256);
    const size_t need = img + (size_t)lt.cand_cap[0] * 4 + 256;
    int rc = ensure_stage_scratch(h, need);
    if (rc) return rc;
    *db = DeviceBuffers();
    db->pyr = h->s_a;
    db->blur = h->s_b;
    db->cand = reinterpret_cast<uint32_t*>(h->s_b + img);  // FAST test: candidates behind the (unused) blur plane
    db->cand_n = reinterpret_cast<int*>(h->s_a + img);
    *cand_cap = (size_t)lt.cand_cap[0];
    *out = lt;
    return VXORB_OK;
}

int vxorb_debug_fast(vxorb_handle* h, const uint8_t* img, int width, int height, size_t stride, int threshold,
                     int border, int cap, vxorb_corner* out, int* n_out) {
    if (!h) return VXORB_E_INVALID;
    if (!img || !n_out || width < 1 || height < 1 || width >= 4096 || height >= 4096 || border < 3 || threshold < 1 ||
        threshold > 254 || cap < 0 || (cap > 0 && !out))
        return fail(h, VXORB_E_INVALID, "bad FAST arguments");
    CK(cudaSetDevice(h->p.device_id));
    *n_out = 0;
    LevelTab lt;
    DeviceBuffers db;
    size_t ccap = 0;
    int rc = single_level_tab(h, width, height, border, &lt, &db, &ccap);
    if (rc) return rc;
    if (!lt.active[0]) return VXORB_OK;
    cudaStream_t s = h->ws[0].stream;
    CK(cudaMemsetAsync(h->s_a, 0, h->s_bytes, s));
    CK(cudaMemcpy2DAsync(db.pyr + lt.off[0], lt.pitch[0], img, stride, width, height, cudaMemcpyHostToDevice, s));
    {
        const int nl = launch_fast(lt, db, threshold, border, s);
        if (nl < 0) return fail(h, VXORB_E_CUDA, "cuTensorMapEncodeTiled failed for the FAST tile maps");
        h->launches += nl;
    }
    CK(cudaGetLastError());
    int n = 0;
    CK(cudaMemcpyAsync(&n, db.cand_n, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    if ((size_t)n > ccap) return fail(h, VXORB_E_CAPACITY, "candidate list overflow");
    std::vector<uint32_t> packed((size_t)n);
    if (n) {
        CK(cudaMemcpyAsync(packed.data(), db.cand, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
    }
    // the kernel emits per tile; present in row-major order like cv::FastFeatureDetector
    std::sort(packed.begin(), packed.end(), [](uint32_t a, uint32_t b) { return (a & 0xffffffu) < (b & 0xffffffu); });
    *n_out = n;
    for (int i = 0; i < n && i < cap; ++i) {
        out[i].x = (int)(packed[i] & 0xfffu);
        out[i].y = (int)((packed[i] >> 12) & 0xfffu);
        out[i].score = (int)(packed[i] >> 24);
    }
    return n > cap ? fail(h, VXORB_E_CAPACITY, "corner output buffer too small") : VXORB_OK;
}

int vxorb_debug_blur(vxorb_handle* h, const uint8_t* src, int width, int height, size_t stride, uint8_t* dst,
                     size_t dstride) {
    if (!h) return VXORB_E_INVALID;
    if (!src || !dst || width < 1 || height < 1 || width >= 65536 || height >= 65536)
        return fail(h, VXORB_E_INVALID, "bad blur arguments");
    CK(cudaSetDevice(h->p.device_id));
    LevelTab lt;
    DeviceBuffers db;
    size_t ccap = 0;
    int rc = single_level_tab(h, width, height, 3, &lt, &db, &ccap);
    if (rc) return rc;
    cudaStream_t s = h->ws[0].stream;
    CK(cudaMemcpy2DAsync(db.pyr + lt.off[0], lt.pitch[0], src, stride, width, height, cudaMemcpyHostToDevice, s));
    h->launches += launch_apron(lt, db, lt.n_levels, s);
    h->launches += launch_blur(lt, db, s);
    CK(cudaGetLastError());
    CK(cudaMemcpy2DAsync(dst, dstride, db.blur + lt.off[0], lt.pitch[0], width, height, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return VXORB_OK;
}

int vxorb_debug_level(vxorb_handle* h, int frame, int level, int blurred, uint8_t* dst, size_t dstride) {
    if (!h) return VXORB_E_INVALID;
    if (h->geo_w == 0 || !dst || level < 0 || level >= h->lt.n_levels || frame < 0 || frame >= h->p.max_batch)
        return fail(h, VXORB_E_INVALID, "bad level query");
    CK(cudaSetDevice(h->p.device_id));
    const LevelTab& lt = h->lt;
    const uint8_t* base = (blurred ? h->ws[0].db.blur : h->ws[0].db.pyr) + lt.off[level] + (size_t)frame * lt.fstride[level];
    CK(cudaMemcpy2D(dst, dstride, base, lt.pitch[level], lt.w[level], lt.h[level], cudaMemcpyDeviceToHost));
    return VXORB_OK;
}

}  // extern "C"
