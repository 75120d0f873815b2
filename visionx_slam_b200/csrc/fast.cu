// fast.cu — FAST-9/16 corner detection with cornerScore and strict 3x3 non-max suppression.
//
// Replaces cv::FastFeatureDetector(20, true)::detect on every pyramid level plus
// KeyPointsFilter::runByImageBorder(31), both inside cv::ORB::detectAndCompute
// (orb_extractor.cpp:13).  Semantics: SURVEY.md A.2/A.3 — a pixel is a corner iff >= 9 contiguous
// ring pixels are all > v+t or all < v-t; score = max over the 16 arcs of 9 of min|v - p| minus 1
// (so corner <=> score >= t); kept iff its score is strictly greater than the 8 neighbours'.
// Integer work, bit-exact.
//
// Persistent CTAs (8 warps) walk over tiles whose score region is 128 x 48 px (emission region: the
// inner 124 x 46); the level pixels a tile needs sit in shared memory as 32-bit words and the score
// map never leaves shared memory.  Warp w owns one row of every group of eight score rows (row w ^ i of group i,
// so its six rows sit in six different bank groups of the 176-byte-pitch tile buffer) from the first
// test to the emission, so a tile costs ONE CTA barrier (between scoring and NMS, which reads the
// neighbouring rows other warps scored); score maps and tile buffers are double buffered.
//   load    one TMA box per tile (cp.async.bulk.tensor.3d over a per-level {pitch, rows, frame}
//           tensor map: 54 rows x 176 B = 3-row ring halo, halo words left and right, 16 B aligned start),
//           issued two tiles ahead right after the barrier that frees the buffer and completed
//           through an mbarrier; pixels outside the plane are zero-filled by the TMA unit and are
//           never part of a valid ring
//   stage 1 every lane tests FOUR horizontally adjacent pixels per step with byte SIMD: a corner
//           needs, for each opposite ring pair {k, k+8}, one pixel with |p - v| > t, so
//           (F0|F8) & (F4|F12) with F = per-byte |p - v| > t (VABSDIFF4 + carry trick) is a
//           necessary condition; 16-byte pixel rows -> 5 LDS, 2 PRMT, 4 VABSDIFF4, ~16 logic ops
//           per 4 px.  Survivor bits of the warp's 6 rows collect in one register per lane and are
//           compacted into the warp's own list with one warp prefix sum (no atomics).
//   stage 2 one lane scores one listed pixel, both polarities at once in s16x2 halves:
//           D_k = (p_k - v + 256) | (v - p_k + 256) << 16 is ONE IMAD per ring pixel, min over 9
//           contiguous ring pixels = two rounds of 3-input VIMNMX3 (32 ops), max over the 16 arcs
//           (8 ops); score = max(lo, hi) - 257, written to the score map iff >= t.  Corners are
//           re-compacted to the front of the same list by ballot.
//   stage 3 strict 3x3 NMS of the warp's corners on the score map, survivors compacted by ballot,
//   emit    one global atomicAdd per warp and tile, packed corners (score<<24|y<<12|x) written directly
// Roofline: HBM, algorithmic bytes = one read of every level pixel (950 532 B per 640x480 frame);
// the kernel is integer-issue bound.
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)
#include <string.h>

#include <mutex>

#include "vxorb_internal.h"

namespace vxorb {

namespace {

constexpr int FW = kFastScoreW, FSR = kFastScoreH;   // score region of a tile
constexpr int TROWS = FSR + 6, TWORDS = FW / 4 + 12; // tile rows, 32-bit words per tile row: 136 B of pixels + up to 12 B of slack (a
                                                     // TMA box must START on a 16 B boundary of the plane row) = 148 -> 160 B would do,
                                                     // but 176 B = 44 words puts consecutive rows 12 banks apart, so the 8 rows of a
                                                     // row group land in 8 different bank groups (40 words: only 4)
constexpr int SCP = FW + 4;                          // score-map row pitch: 33 words, rows never share a bank at equal columns
constexpr uint32_t kTileBytes = TROWS * TWORDS * 4;
constexpr int TPITCH = TWORDS * 4;
constexpr int RPW = FSR / 8;                          // score rows per warp
static_assert(FSR % 8 == 0 && RPW >= 1 && RPW <= 8, "a lane keeps 4 survivor bits per row in one 32-bit register");
#ifndef VXORB_FAST_CTAS
#define VXORB_FAST_CTAS 5
#endif
constexpr int kFastCtasPerSm = VXORB_FAST_CTAS;                      // 41.4 KB of shared memory and 32 registers per thread; swept 3..8
constexpr int WQCAP = RPW * FW;                       // pixels one warp tests per tile

// ring offsets (dx, dy), k = 0..15, as byte offsets into the tile
__device__ __forceinline__ constexpr int ring_off(int k) {
    constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    return dy[k] * TPITCH + dx[k];
}

// per-byte flag (bit 7) of d > t
template <bool HI>
__device__ __forceinline__ uint32_t gt_flags(uint32_t d, uint32_t c) {
    const uint32_t lo = (d & 0x7f7f7f7fu) + c;
    return HI ? (lo & d) : (lo | d);
}

struct FastMaps {
    CUtensorMap m[kMaxLevels];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int c0, int c1, int c2,
                                            unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}

}  // namespace

template <bool HI>
__global__ void __launch_bounds__(256, kFastCtasPerSm)
fast_kernel(const LevelTab lt, const __grid_constant__ FastMaps maps, uint32_t* __restrict__ cand,
            int* __restrict__ cand_n, int threshold, int border, int tiles_per_frame, int total_tiles) {
    struct __align__(128) TileBuf {
        uint32_t w[TROWS][TWORDS];  // TMA destinations must be 128 B aligned: the struct alignment pads buffer 1
    };
    __shared__ TileBuf tiles[2];
    __shared__ __align__(16) uint8_t score[2][FSR][SCP];  // tile t uses map t & 1; a warp writes only its own rows
    __shared__ uint16_t wq[8][WQCAP];       // per-warp list: stage-1 survivors, then corners, then NMS survivors
    __shared__ __align__(8) unsigned long long mbar[2];
    __shared__ int4 info[4];                // tile descriptors (level, X, Y, frame), ring over 4 tiles

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int grid = (int)gridDim.x;
    const uint32_t lt_mask = (1u << lane) - 1u;

    // one thread: decode tile t, publish its descriptor and start its TMA load into buffer b
    auto issue = [&](int t, int b, int slot) {
        const int f = t / tiles_per_frame;
        int tl = t - f * tiles_per_frame;
        int l = 0;
        while (l + 1 < lt.n_levels && tl >= lt.fast_tile_begin[l + 1]) ++l;
        tl -= lt.fast_tile_begin[l];
        const int ty = tl / lt.fast_tiles_x[l], tx = tl - ty * lt.fast_tiles_x[l];
        // score region origin: X is a multiple of 4 so tile words are aligned words of the plane
        const int X = ((border - 1) & ~3) + tx * kFastEmitW;
        const int Y = (border - 1) + ty * kFastEmitH;
        info[slot] = make_int4(l, X, Y, f);
        mbar_expect_tx(&mbar[b], kTileBytes);
        // tensor map coordinates count from the plane allocation: kApronX bytes / kApronY rows before
        // pixel (0, 0).  The box starts at the 16 B boundary at or below the tile's first byte (X - 4).
        tma_load_3d(&tiles[b].w[0][0], &maps.m[l], (X - 4 + kApronX) & ~15, Y - 3 + kApronY, f, &mbar[b]);
    };

    if (tid == 0) {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if ((int)blockIdx.x < total_tiles) issue(blockIdx.x, 0, 0);
        if ((int)blockIdx.x + grid < total_tiles) issue(blockIdx.x + grid, 1, 1);
    }
    __syncthreads();

    uint16_t* q16 = wq[warp];
    const uint32_t relcol = (uint32_t)lane * 0x04040404u + 0x03020100u;  // this lane's four tile-relative columns
    int it = 0;
    for (int t = blockIdx.x; t < total_tiles; t += grid, ++it) {
        const int buf = it & 1;
        uint32_t (*tile)[TWORDS] = tiles[buf].w;
        uint8_t* sc = &score[buf][0][0];
        const int4 ti = info[it & 3];
        const int l = ti.x, X = ti.y, Y = ti.z, f = ti.w;
        const int w = lt.w[l], h = lt.h[l];
        const int wo = ((X - 4 + kApronX) & 15) >> 2;  // words of slack before the tile's first word in a buffer row

        // this warp's rows of the score map (last read by the NMS of tile t - 2, two barriers ago)
#pragma unroll
        for (int i = 0; i < RPW; ++i) reinterpret_cast<uint32_t*>(sc + (i * 8 + (warp ^ i)) * SCP)[lane] = 0u;

        // score domain (global coordinates, inclusive)
        const int sx_min = max(3, border - 1), sx_max = min(w - 4, w - border);
        const int sy_min = max(3, border - 1), sy_max = min(h - 4, h - border);
        // bit 7 of byte j <=> column X + 4*lane + j lies in [sx_min, sx_max].  Tile-relative columns are bytes
        // 0..127, so both comparisons are one packed add / subtract against a tile-uniform constant: byte + (128 - lo)
        // and (128 + hi) - byte carry into bit 7 exactly when byte >= lo resp. byte <= hi, never across bytes.
        const int clo = min(max(sx_min - X, 0), 128), chi = min(max(sx_max - X, -1), 127);
        const uint32_t colmask = (relcol + (uint32_t)(128 - clo) * 0x01010101u) &
                                 ((uint32_t)(128 + chi) * 0x01010101u - relcol) & 0x80808080u;
        mbar_wait(&mbar[buf], (uint32_t)(it >> 1) & 1u);

        // ---- stage 1: 4-pixel-wide necessary-condition test -----------------------------------
        int nq;
        {
            const uint32_t cadd = (uint32_t)(HI ? 255 - threshold : 127 - threshold) * 0x01010101u;
            uint32_t acc = 0;
            // rows of the score domain, tile-relative: [ylo, ylo + yspan]; an empty domain gets an unreachable ylo
            const int ylo = sy_max >= sy_min ? sy_min - Y : (1 << 20);
            const uint32_t yspan = sy_max >= sy_min ? (uint32_t)(sy_max - sy_min) : 0u;
#pragma unroll
            for (int i = 0; i < RPW; ++i) {
                // rows interleaved over warps (a corner cluster spreads over several warps); within row group i the warp
                // takes row warp ^ i, so its six rows sit in six different bank groups of the tile buffer: vertical
                // structures, whose survivors share columns, then do not collide in the ring gathers of stage 2
                const int r = i * 8 + (warp ^ i);
                uint32_t q = 0;
                if ((uint32_t)(r - ylo) <= yspan) {  // sy_min <= Y + r <= sy_max in one unsigned compare
                    const uint32_t* row = &tile[r + 3][lane + wo];
                    const uint32_t wl = row[0], v = row[1], wr = row[2];
                    const uint32_t up = tile[r][lane + 1 + wo], dn = tile[r + 6][lane + 1 + wo];
                    const uint32_t lf = __byte_perm(wl, v, 0x4321u), rt = __byte_perm(v, wr, 0x6543u);
                    const uint32_t f0 = gt_flags<HI>(__vabsdiffu4(dn, v), cadd);
                    const uint32_t f8 = gt_flags<HI>(__vabsdiffu4(up, v), cadd);
                    const uint32_t f4 = gt_flags<HI>(__vabsdiffu4(rt, v), cadd);
                    const uint32_t f12 = gt_flags<HI>(__vabsdiffu4(lf, v), cadd);
                    q = (f0 | f8) & (f4 | f12) & colmask;
                }
                // bits 7/15/23/31 -> bits 28..31 (distinct partial products, no carries)
                acc = (acc >> 4) | ((q * 0x00204081u) & 0xF0000000u);
            }
            acc >>= 32 - 4 * RPW;  // row i of this warp now sits in bits 4i .. 4i+3
            // compaction into the warp's own list: one inclusive prefix sum, no atomics
            const int cnt = __popc(acc);
            int incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int n = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += n;
            }
            nq = __shfl_sync(0xffffffffu, incl, 31);
            int pos = incl - cnt;
            const int colbase = 4 * lane;
            while (acc) {
                const int b = __ffs(acc) - 1;
                acc &= acc - 1;
                const int i = b >> 2;
                VX_CHECK(pos >= 0 && pos < WQCAP && i < RPW);
                q16[pos++] = (uint16_t)(((i * 8 + (warp ^ i)) << 7) + colbase + (b & 3));
            }
        }
        __syncwarp();

        // ---- stage 2: cornerScore, one listed pixel per lane ----------------------------------
        // Both polarities ride in one s16x2 register: lo = p_k - v + 256, hi = v - p_k + 256 (one IMAD
        // per ring pixel: p_k * (1 - 2^16) + C; both fields stay in 1..511, so no carry crosses).
        // min over 9 contiguous ring pixels by two rounds of 3-input minima, max over the 16 arcs:
        // lo -> 256 + max_arcs min(p - v) (bright), hi -> 256 + max_arcs min(v - p) (dark).
        int nc = 0;
        {
            const uint8_t* t8 = reinterpret_cast<const uint8_t*>(&tile[0][0]) + 3 * TPITCH + 4 + 4 * wo;
            for (int e0 = 0; e0 < nq; e0 += 32) {
                const int e = e0 + lane;
                bool corner = false;
                int idx = 0;
                if (e < nq) {
                    idx = q16[e];
                    VX_CHECK((idx >> 7) < FSR && 4 * wo + 4 + (idx & 127) + 3 < TPITCH);  // ring stays inside the tile buffer
                    const uint8_t* tp = t8 + (idx >> 7) * TPITCH + (idx & 127);
                    const uint32_t v = tp[0];
                    const uint32_t c0 = v * 0xFFFFu + 0x01000100u;  // (256 - v) | (256 + v) << 16
                    uint32_t d[16];
#pragma unroll
                    for (int k = 0; k < 16; ++k) d[k] = (uint32_t)tp[ring_off(k)] * 0xFFFF0001u + c0;
                    uint32_t m3[16];
#pragma unroll
                    for (int k = 0; k < 16; ++k) m3[k] = __vimin3_s16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
                    uint32_t m9[16];
#pragma unroll
                    for (int k = 0; k < 16; ++k) m9[k] = __vimin3_s16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
                    uint32_t best = __vimax3_s16x2(m9[0], m9[1], m9[2]);  // 16 values: seven 3-input maxima + one 2-input
#pragma unroll
                    for (int k = 3; k < 15; k += 2) best = __vimax3_s16x2(best, m9[k], m9[k + 1]);
                    best = __vmaxs2(best, m9[15]);
                    const int s = max((int)(best & 0xffffu), (int)(best >> 16)) - 257;
                    if (s >= threshold) {
                        sc[idx + (idx >> 7) * (SCP - FW)] = (uint8_t)s;
                        corner = true;
                    }
                }
                __syncwarp();  // every lane has read its entry before the front of the list is overwritten
                const uint32_t m = __ballot_sync(0xffffffffu, corner);
                VX_CHECK(nc + __popc(m) <= e0 + 32 && nc + __popc(m) <= WQCAP);  // compaction never overtakes the read cursor
                if (corner) q16[nc + __popc(m & lt_mask)] = (uint16_t)idx;
                nc += __popc(m);
            }
        }
        __syncthreads();  // the one CTA barrier per tile: all rows of the score map are final, tile buffer is free
        if (tid == ((it & 7) << 5)) {
            const int tn = t + 2 * grid;
            if (tn < total_tiles) issue(tn, buf, (it + 2) & 3);
        }

        // ---- stage 3: strict 3x3 non-max suppression of this warp's corners, emission ----------
        {
            const int ex_max = w - border, ey_max = h - border;  // exclusive
            int nk = 0;
            for (int e0 = 0; e0 < nc; e0 += 32) {
                const int e = e0 + lane;
                bool keep = false;
                int idx = 0;
                if (e < nc) {
                    idx = q16[e];
                    const int r = idx >> 7, c = idx & 127;
                    const int gx = X + c, gy = Y + r;
                    const bool inside = r >= 1 && r <= kFastEmitH && c >= 1 && c <= kFastEmitW && gx >= border &&
                                        gy >= border && gx < ex_max && gy < ey_max;
                    // branch-free: strictly greater than all eight neighbours <=> greater than their maximum
                    // (lanes outside the emission domain read around a safe interior cell instead)
                    VX_CHECK(r < FSR && (!inside || (r >= 1 && r + 1 < FSR && c >= 1 && c + 1 < FW)));
                    const uint8_t* p = sc + (inside ? idx + r * (SCP - FW) : SCP + 1);
                    const int s = p[0];
                    const int m0 = max(max((int)p[-SCP - 1], (int)p[-SCP]), (int)p[-SCP + 1]);
                    const int m1 = max(max((int)p[-1], (int)p[1]), (int)p[SCP - 1]);
                    const int m2 = max((int)p[SCP], (int)p[SCP + 1]);
                    keep = inside && s > max(max(m0, m1), m2);
                }
                __syncwarp();
                const uint32_t m = __ballot_sync(0xffffffffu, keep);
                if (keep) q16[nk + __popc(m & lt_mask)] = (uint16_t)idx;
                nk += __popc(m);
            }
            if (nk > 0) {  // warp-uniform
                __syncwarp();
                const int cap = lt.cand_cap[l];
                int obase = 0;
                if (lane == 0) obase = atomicAdd(cand_n + l * lt.batch + f, nk);
                obase = __shfl_sync(0xffffffffu, obase, 0);
                uint32_t* out = cand + lt.cand_off[l] + (size_t)f * cap;
                for (int e = lane; e < nk; e += 32) {
                    const int idx = q16[e];
                    const uint32_t s = sc[idx + (idx >> 7) * (SCP - FW)];
                    VX_CHECK(Y + (idx >> 7) < h && X + (idx & 127) < w);
                    if (obase + e < cap)
                        out[obase + e] = (s << 24) | ((uint32_t)(Y + (idx >> 7)) << 12) | (uint32_t)(X + (idx & 127));
                }
                __syncwarp();  // the list is reused by the next tile
            }
        }
    }
}

// Per-level tensor maps over pyramid-shaped planes (pyr or blur): dims {pitch bytes, rows incl. apron,
// frames}, u8, box {box_w, box_h, 1}, no swizzle, zero fill outside.  `out` is CUtensorMap[kMaxLevels].
// The encoder is a driver entry point fetched through the runtime, so the library still links
// against cudart only.
bool build_plane_maps(const LevelTab& lt, const uint8_t* planes, int box_w, int box_h, bool active_only, void* out_maps) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn)
            return false;
        encode = reinterpret_cast<EncodeFn>(fn);
    }
    CUtensorMap* out = static_cast<CUtensorMap*>(out_maps);
    memset(out, 0, sizeof(CUtensorMap) * kMaxLevels);
    for (int l = 0; l < lt.n_levels; ++l) {
        if (active_only && !lt.active[l]) continue;
        const size_t origin = plane_origin(lt.w[l]);
        void* base = const_cast<uint8_t*>(planes) + lt.off[l] - origin;
        const cuuint64_t dims[3] = {(cuuint64_t)lt.pitch[l], (cuuint64_t)(lt.h[l] + 2 * kApronY),
                                    (cuuint64_t)(lt.batch > 0 ? lt.batch : 1)};
        const cuuint64_t strides[2] = {(cuuint64_t)lt.pitch[l], (cuuint64_t)lt.fstride[l]};
        const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
        const cuuint32_t estr[3] = {1u, 1u, 1u};
        if (encode(&out[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return false;
    }
    return true;
}

namespace {

// The tensor maps of a launch depend only on the planes' base address, the level geometry and the batch: the
// streaming path launches the same few combinations (one per pipeline slot) over and over, so the last encodings
// are kept instead of calling the driver's encoder up to kMaxLevels times per launch.
struct MapKey {
    const uint8_t* planes;
    int n_levels, batch;
    int w[kMaxLevels], h[kMaxLevels], pitch[kMaxLevels], active[kMaxLevels];
    unsigned long long off[kMaxLevels], fstride[kMaxLevels];
};
struct MapEntry {
    MapKey key;
    FastMaps maps;
    unsigned long long stamp = 0;
    bool used = false;
};
constexpr int kMapCache = 16;
MapEntry g_map_cache[kMapCache];
unsigned long long g_map_stamp = 0;
std::mutex g_map_mutex;

bool cached_fast_maps(const LevelTab& lt, const uint8_t* planes, FastMaps* out) {
    MapKey key;
    memset(&key, 0, sizeof(key));
    key.planes = planes;
    key.n_levels = lt.n_levels;
    key.batch = lt.batch;
    for (int l = 0; l < lt.n_levels; ++l) {
        key.w[l] = lt.w[l]; key.h[l] = lt.h[l]; key.pitch[l] = lt.pitch[l]; key.active[l] = lt.active[l];
        key.off[l] = lt.off[l]; key.fstride[l] = lt.fstride[l];
    }
    std::lock_guard<std::mutex> lock(g_map_mutex);
    MapEntry* victim = nullptr;  // first unused entry, else the least recently used one
    for (auto& e : g_map_cache) {
        if (e.used && memcmp(&e.key, &key, sizeof(key)) == 0) {
            e.stamp = ++g_map_stamp;
            *out = e.maps;
            return true;
        }
        if (!victim || (victim->used && (!e.used || e.stamp < victim->stamp))) victim = &e;
    }
    if (!build_plane_maps(lt, planes, TWORDS * 4, TROWS, true, victim->maps.m)) {
        victim->used = false;
        return false;
    }
    victim->key = key;
    victim->used = true;
    victim->stamp = ++g_map_stamp;
    *out = victim->maps;
    return true;
}

}  // namespace

cudaError_t carveout_fast(int pct) {
    const cudaError_t e = set_carveout(fast_kernel<true>, pct);
    return e != cudaSuccess ? e : set_carveout(fast_kernel<false>, pct);
}

// shared: the launch runs beside the kernels of other pipeline slots (the production schedule of a chunk).  The
// persistent grid then takes kFastSharedCtas CTAs per SM instead of all kFastCtasPerSm: at five, FAST holds every
// byte of an SM's shared memory for its whole run and nothing but the register-only blur can join it; at three it is
// 8 % slower alone (0.205 against 0.190 ms per 128 frames) and the pipelined step 2 % faster (216 k against 211 k
// frames/s; two CTAs: the same, four: +1 %).  Alone — instrumented passes, single frames, stage isolation — it
// takes the whole SM.
int launch_fast(const LevelTab& lt, const DeviceBuffers& db, int threshold, int border, cudaStream_t s, bool shared) {
    const int tiles = lt.fast_tile_begin[lt.n_levels];
    if (tiles == 0 || lt.batch <= 0) return 0;
    FastMaps maps;
    if (!cached_fast_maps(lt, db.pyr, &maps)) return -1;
    const int total = tiles * lt.batch;
#ifndef VXORB_FAST_SHARED_CTAS
#define VXORB_FAST_SHARED_CTAS 3
#endif
    const int per_sm = shared ? VXORB_FAST_SHARED_CTAS : kFastCtasPerSm;
    const int grid = total < 148 * per_sm ? total : 148 * per_sm;  // persistent CTAs of 8 warps
    if (threshold >= 128)
        fast_kernel<true><<<grid, 256, 0, s>>>(lt, maps, db.cand, db.cand_n, threshold, border, tiles, total);
    else
        fast_kernel<false><<<grid, 256, 0, s>>>(lt, maps, db.cand, db.cand_n, threshold, border, tiles, total);
    return 1;
}

}  // namespace vxorb
