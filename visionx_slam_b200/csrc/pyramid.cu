// pyramid.cu — gray conversion and the chained fixed-point pyramid.
//
// Replaces, for orb_extractor.cpp:13 (cv::ORB::detectAndCompute):
//   * cvtColor(BGR2GRAY) applied when the frame has 3 channels (system.cpp:46 feeds IMREAD_COLOR)
//   * the 7 chained cv::resize(prev, INTER_LINEAR_EXACT) calls that build levels 1..7
// Arithmetic: 8.8 fixed-point taps per axis computed on the host in double,
// horizontal pass in 16 bits, vertical pass in 32 bits, (v + 2^15) >> 16.  Bit-exact integer work.
//
// Roofline: HBM.  Algorithmic bytes per frame = read(level l-1) + write(level l) summed over
// l = 1..7 (1 569 878 B at 640x480).  One launch per level covers the whole batch (grid.z = frame).
#include <math.h>

#include <algorithm>
#include <type_traits>

#include "vxorb_internal.h"

namespace vxorb {

// Taps for one axis: ofs | c1 << 16, with c0 = 256 - c1 (INTER_LINEAR_EXACT coefficient rule).
void build_resize_taps(int ssize, int dsize, uint32_t* taps) {
    const double scale = 1.0 / ((double)dsize / (double)ssize);
    for (int i = 0; i < dsize; ++i) {
        double f = scale * ((double)i + 0.5) - 0.5;
        int i0 = (int)floor(f);
        uint32_t ofs, c1;
        if (i0 < 0) {
            ofs = 0;
            c1 = 0;
        } else if (i0 >= ssize - 1) {
            ofs = (uint32_t)(ssize - 1);
            c1 = 0;
        } else {
            ofs = (uint32_t)i0;
            c1 = (uint32_t)lrint((f - (double)i0) * 256.0);
        }
        taps[i] = ofs | (c1 << 16);
    }
}

__global__ void __launch_bounds__(256) bgr2gray_kernel(const uint8_t* __restrict__ bgr, size_t bgr_pitch,
                                                       size_t bgr_fstride, uint8_t* __restrict__ dst,
                                                       int w, int h, int pitch, size_t fstride) {
    const int x0 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (x0 >= w || y >= h) return;
    const uint8_t* s = bgr + (size_t)blockIdx.z * bgr_fstride + (size_t)y * bgr_pitch + (size_t)x0 * 3;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (x0 + k < w) {
            uint32_t b = s[3 * k], g = s[3 * k + 1], r = s[3 * k + 2];
            out |= ((r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15) << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)blockIdx.z * fstride + (size_t)y * pitch + x0) = out;
}

// Generic path (any scale): one thread produces four horizontally adjacent output pixels with
// scalar taps.  Used only when four consecutive output columns span more than 8 source bytes
// (scale factor above ~2); cv::ORB's 1.2 never gets here.
__global__ void __launch_bounds__(256)
resize_generic_kernel(const uint8_t* __restrict__ src, int sw, int sh, int spitch, size_t sfstride,
                      uint8_t* __restrict__ dst, int dw, int dh, int dpitch, size_t dfstride,
                      const uint32_t* __restrict__ xtab, const uint32_t* __restrict__ ytab) {
    const int x0 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (x0 >= dw || y >= dh) return;
    src += (size_t)blockIdx.z * sfstride;
    dst += (size_t)blockIdx.z * dfstride;
    const uint32_t ty = __ldg(ytab + y);
    const int j0 = (int)(ty & 0xffffu);
    const uint32_t cy1 = ty >> 16, cy0 = 256u - cy1;
    const int j1 = min(j0 + 1, sh - 1);
    const uint8_t* r0 = src + (size_t)j0 * spitch;
    const uint8_t* r1 = src + (size_t)j1 * spitch;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        if (x < dw) {
            const uint32_t tx = __ldg(xtab + x);
            const int i0 = (int)(tx & 0xffffu);
            const uint32_t cx1 = tx >> 16, cx0 = 256u - cx1;
            const int i1 = min(i0 + 1, sw - 1);
            const uint32_t h0 = __ldg(r0 + i0) * cx0 + __ldg(r0 + i1) * cx1;  // 8.8
            const uint32_t h1 = __ldg(r1 + i0) * cx0 + __ldg(r1 + i1) * cx1;
            const uint32_t v = (h0 * cy0 + h1 * cy1 + (1u << 15)) >> 16;      // <= 255
            out |= v << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * dpitch + x0) = out;
}

// Fast path: one thread owns 4 output columns and walks down `strip_rows` output rows; a warp owns
// 128 adjacent output columns of one row band, so the row schedule is warp-uniform.
//   horizontal pass of one source row: three aligned 32-bit loads cover the <= 8 source bytes the
//   4 columns need; a funnel shift aligns them to the first tap, two PRMTs (loop-invariant
//   selectors) gather the (p[i0], p[i0+1]) byte pairs, and IDP.2A multiplies each pair with its
//   16-bit weights (256-c1, c1): H = p0*c0 + p1*c1 in ONE instruction per column;
//   the two most recent source rows' H values stay in registers (rows advance ~1.2 per output row);
//   vertical pass: 2 IMAD per pixel, +2^15, byte 2 of each result packed into one 32-bit store.
//   The source words of the next two rows are always in flight (2-deep register queue), and the
//   next row's y tap is fetched one row early, so the dependent load chain never stalls the strip.
// Rows per strip shrink with the level size (resize_rows_for) so that small levels still spread over
// every SM: they are latency bound, not bandwidth bound.
__global__ void __launch_bounds__(256)
resize_kernel(const uint8_t* __restrict__ src, int sh, int spitch, size_t sfstride,
              uint8_t* __restrict__ dst, int dw, int dh, int dpitch, size_t dfstride,
              const uint32_t* __restrict__ xtab, const uint32_t* __restrict__ ytab, int word_groups,
              int strip_rows) {
    const int item = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int band = item / word_groups;
    const int y0 = band * strip_rows;
    if (y0 >= dh) return;
    const int x0 = ((item - band * word_groups) * 32 + (threadIdx.x & 31)) * 4;
    if (x0 >= dw) return;
    src += (size_t)blockIdx.y * sfstride;
    dst += (size_t)blockIdx.y * dfstride + (size_t)y0 * dpitch + x0;

    // loop-invariant column taps
    uint32_t wgt[4];
    int ofs[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t t = __ldg(xtab + min(x0 + k, dw - 1));
        ofs[k] = (int)(t & 0xffffu);
        const uint32_t c1 = t >> 16;
        wgt[k] = (256u - c1) | (c1 << 16);
    }
    const uint32_t shift = (uint32_t)(ofs[0] & 3) * 8u;
    const int e0 = 0, e1 = ofs[1] - ofs[0], e2 = ofs[2] - ofs[0], e3 = ofs[3] - ofs[0];
    const uint32_t sel01 = (uint32_t)(e0 | ((e0 + 1) << 4) | (e1 << 8) | ((e1 + 1) << 12));
    const uint32_t sel23 = (uint32_t)(e2 | ((e2 + 1) << 4) | (e3 << 8) | ((e3 + 1) << 12));
    const uint32_t* sp = reinterpret_cast<const uint32_t*>(src) + (ofs[0] >> 2);
    const int spitch_w = spitch >> 2;
    const int rows = min(strip_rows, dh - y0);

    uint32_t hp[4] = {0, 0, 0, 0}, hc[4] = {0, 0, 0, 0};  // H of source rows cur-1 and cur
    uint32_t ty = __ldg(ytab + y0);
    int cur = (int)(ty & 0xffffu) - 1;
    // queue: q?[0] = words of source row cur+1, q?[1] = row cur+2
    uint32_t q0[2], q1[2], q2[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const uint32_t* p = sp + (size_t)min(cur + 1 + i, sh - 1) * spitch_w;
        q0[i] = __ldg(p);
        q1[i] = __ldg(p + 1);
        q2[i] = __ldg(p + 2);
    }
    for (int i = 0; i < rows; ++i) {
        const int j0 = (int)(ty & 0xffffu);
        const uint32_t cy1 = ty >> 16, cy0 = 256u - cy1;
        ty = __ldg(ytab + min(y0 + i + 1, dh - 1));  // next row's tap, one row early
        if (cur < j0 - 1) {  // a jump of more than one source row (scale > 2): restart the queue
            cur = j0 - 1;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint32_t* p = sp + (size_t)min(cur + 1 + k, sh - 1) * spitch_w;
                q0[k] = __ldg(p);
                q1[k] = __ldg(p + 1);
                q2[k] = __ldg(p + 2);
            }
        }
        while (cur < j0 + 1) {
            ++cur;
            const uint32_t w0 = q0[0], w1 = q1[0], w2 = q2[0];
            q0[0] = q0[1];
            q1[0] = q1[1];
            q2[0] = q2[1];
            const uint32_t* p = sp + (size_t)min(cur + 2, sh - 1) * spitch_w;
            q0[1] = __ldg(p);
            q1[1] = __ldg(p + 1);
            q2[1] = __ldg(p + 2);
            const uint32_t lo = __funnelshift_r(w0, w1, shift), hi = __funnelshift_r(w1, w2, shift);
            const uint32_t b01 = __byte_perm(lo, hi, sel01), b23 = __byte_perm(lo, hi, sel23);
#pragma unroll
            for (int k = 0; k < 4; ++k) hp[k] = hc[k];
            hc[0] = __dp2a_lo(wgt[0], b01, 0u);
            hc[1] = __dp2a_hi(wgt[1], b01, 0u);
            hc[2] = __dp2a_lo(wgt[2], b23, 0u);
            hc[3] = __dp2a_hi(wgt[3], b23, 0u);
        }
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = hc[k] * cy1 + (hp[k] * cy0 + (1u << 15));  // 16.16, byte 2 = result
        *reinterpret_cast<uint32_t*>(dst) =
            __byte_perm(__byte_perm(v[0], v[1], 0x0062u), __byte_perm(v[2], v[3], 0x0062u), 0x5410u);
        dst += dpitch;
    }
}

// Strip height for one level: 16 rows amortise the two start-up source rows best; halve it while the
// launch would leave SMs without a full set of resident warps (148 SMs x 5 CTAs of 8 warps).
static int resize_rows_for(int dw, int dh, int batch) {
    const int word_groups = ((dw + 3) / 4 + 31) / 32;
    int rows = 16;
    while (rows > 4) {
        const long long warps = (long long)word_groups * ((dh + rows - 1) / rows) * batch;
        if (warps >= 148LL * 40) break;
        rows >>= 1;
    }
    return rows;
}

// True when every group of 4 output columns reads at most 8 consecutive source bytes.
bool resize_taps_fit_fast_path(const uint32_t* xtaps, int dw) {
    for (int x0 = 0; x0 < dw; x0 += 4) {
        const int last = x0 + 3 < dw ? x0 + 3 : dw - 1;
        if ((int)(xtaps[last] & 0xffffu) - (int)(xtaps[x0] & 0xffffu) > 6) return false;
    }
    return true;
}

// ---- band kernel: the production path of the pyramid -----------------------------------------------
// One CTA produces a band of output rows of one frame.  The source rows the band needs are ONE
// contiguous chunk of the previous level's plane (rows are pitch-strided and a band takes whole
// rows), so thread 0 fetches them with a single bulk async copy (cp.async.bulk, completion on an
// mbarrier) into shared memory; 6-8 CTAs are resident per SM, so other bands compute while this
// one's copy is in flight.  Every (output row, output word) item is then independent: two source
// rows x (3 LDS, funnel shift to the first tap, 2 PRMT, 4 IDP.2A) for the horizontal pass and
// 2 IMAD per pixel for the vertical pass.  The per-word tap state (weights, PRMT selectors, word
// index, shift) comes precomputed from the host (6 words per output word, build_resize_words).
// Warp w of the CTA takes rows w, w+8, ... of the band; lanes take consecutive output words.
// The band's CTA also writes the reflect-101 apron of its rows (and of the plane's top / bottom for
// the first / last band), which the blur reads; the last band absorbs the remainder rows so that it
// always holds the rows the bottom apron mirrors.
void build_resize_words(const uint32_t* xtaps, int dw, uint32_t* out) {
    const int words = (dw + 3) / 4;
    for (int k = 0; k < words; ++k) {
        int ofs[4];
        uint32_t wgt[4];
        for (int j = 0; j < 4; ++j) {
            const uint32_t t = xtaps[std::min(4 * k + j, dw - 1)];
            ofs[j] = (int)(t & 0xffffu);
            const uint32_t c1 = t >> 16;
            wgt[j] = (256u - c1) | (c1 << 16);
        }
        const int e1 = ofs[1] - ofs[0], e2 = ofs[2] - ofs[0], e3 = ofs[3] - ofs[0];
        const uint32_t sel01 = (uint32_t)(0 | (1 << 4) | (e1 << 8) | ((e1 + 1) << 12));
        const uint32_t sel23 = (uint32_t)(e2 | ((e2 + 1) << 4) | (e3 << 8) | ((e3 + 1) << 12));
        uint32_t* o = out + (size_t)k * 4;
        o[0] = wgt[0]; o[1] = wgt[1]; o[2] = wgt[2]; o[3] = wgt[3];
        uint32_t* o2 = out + (size_t)words * 4 + (size_t)k * 2;
        o2[0] = sel01 | (sel23 << 16);
        o2[1] = (uint32_t)(kApronX / 4 + (ofs[0] >> 2)) | ((uint32_t)(ofs[0] & 3) * 8u) << 16;
    }
}

namespace {
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
}  // namespace

// The throughput shape's work on one band of one frame: output rows [y_begin, y_end) of `dst` from the source rows
// [jlo, jhi] that sit in shared memory at sm_w (spitch_w words per row).  Called by all 256 threads of the CTA.
__device__ __forceinline__ void band_rows_full(const uint32_t* sm_w, int jlo, int jhi, int sh, int spitch_w,
                                               uint8_t* __restrict__ dst, int dpitch, int y_begin, int y_end,
                                               const uint32_t* __restrict__ ytab, const uint4* __restrict__ xw4,
                                               const uint2* __restrict__ xw2, int words, int lane, int warp) {
    // when the last group of 32 words of a row is at most half full, it is not given a warp-iteration per row pair;
    // the tails of 2 (<= 16 words) or 4 (<= 8 words) row pairs share one iteration below
    const int rem = words & 31;
    const int seg = (rem != 0 && rem <= 16) ? (rem <= 8 ? 8 : 16) : 0;
    const int words_main = seg ? words - rem : words;
    (void)jhi;
    {
        // Throughput shape: a warp owns one row pair of the band (two in the taller last band), so everything that
        // depends on the rows alone — tap rows, their shared-memory bases, the vertical weights, which horizontal
        // passes the second row can reuse, the output row pointers — is worked out once per pair, and the loop over
        // the word groups inside carries only the per-word tables, the passes and the stores (the other loop order
        // redid the row bookkeeping for every group of 32 words: about 155 instructions per 8 pixels against 95).
        for (int y = y_begin + 2 * warp; y < y_end; y += 16) {
            const uint32_t ty = __ldg(ytab + y);
            const bool two = y + 1 < y_end;
            const uint32_t tz = two ? __ldg(ytab + y + 1) : ty;
            const int j0 = (int)(ty & 0xffffu), j1 = min(j0 + 1, sh - 1);
            const int i0 = (int)(tz & 0xffffu), i1 = min(i0 + 1, sh - 1);
            VX_CHECK(j0 >= jlo && i1 <= jhi);
            const uint32_t* const rj0 = sm_w + (j0 - jlo) * spitch_w;
            const uint32_t* const rj1 = sm_w + (j1 - jlo) * spitch_w;
            const uint32_t* const ri0 = sm_w + (i0 - jlo) * spitch_w;
            const uint32_t* const ri1 = sm_w + (i1 - jlo) * spitch_w;
            const bool reuse = i0 == j1, clamped = i1 == i0;  // warp-uniform: rows only
            const uint32_t cya1 = ty >> 16, cya0 = 256u - cya1, cyb1 = tz >> 16, cyb0 = 256u - cyb1;
            uint8_t* const drow = dst + (size_t)y * dpitch;
            // MODE 0: one output row (the odd last row of a band); 1: two rows, the second one's upper source row is
            // the first one's lower one (five pairs out of six at scale 1.2: three horizontal passes); 2: two rows
            // with four passes.  The mode depends on the rows only, so each is a straight-line loop over the word groups.
            auto groups = [&](auto mode_tag) {
                constexpr int MODE = decltype(mode_tag)::value;
                for (int k0 = 0; k0 < words_main; k0 += 32) {
                    const int k = k0 + lane;
                    const bool active = k < words;
                    const int kk = active ? k : words - 1;
                    const uint4 wg = __ldg(xw4 + kk);
                    const uint2 sx = __ldg(xw2 + kk);
                    const uint32_t sel01 = sx.x & 0xffffu, sel23 = sx.x >> 16;
                    const int wofs = (int)(sx.y & 0xffffu);
                    const uint32_t shift = sx.y >> 16;
                    VX_CHECK(wofs + 2 < spitch_w);
                    auto hpass = [&](const uint32_t* row, uint32_t (&hv)[4]) {
                        const uint32_t* p = row + wofs;
                        const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
                        const uint32_t lo = __funnelshift_r(w0, w1, shift), hi = __funnelshift_r(w1, w2, shift);
                        const uint32_t b01 = __byte_perm(lo, hi, sel01), b23 = __byte_perm(lo, hi, sel23);
                        hv[0] = __dp2a_lo(wg.x, b01, 0u);
                        hv[1] = __dp2a_hi(wg.y, b01, 0u);
                        hv[2] = __dp2a_lo(wg.z, b23, 0u);
                        hv[3] = __dp2a_hi(wg.w, b23, 0u);
                    };
                    auto vpass_store = [&](uint8_t* out, const uint32_t (&h0)[4], const uint32_t (&h1)[4], uint32_t c0,
                                           uint32_t c1) {
                        uint32_t v[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) v[j] = h1[j] * c1 + (h0[j] * c0 + (1u << 15));  // 16.16, byte 2 = result
                        if (active)
                            *reinterpret_cast<uint32_t*>(out + 4 * k) =
                                __byte_perm(__byte_perm(v[0], v[1], 0x0062u), __byte_perm(v[2], v[3], 0x0062u), 0x5410u);
                    };
                    uint32_t ha[4], hb[4];
                    hpass(rj0, ha);
                    hpass(rj1, hb);
                    vpass_store(drow, ha, hb, cya0, cya1);
                    if (MODE == 1) {
                        uint32_t hc[4];
                        hpass(ri1, hc);
                        vpass_store(drow + dpitch, hb, hc, cyb0, cyb1);
                    } else if (MODE == 2) {
                        uint32_t hc[4], hd[4];
                        hpass(ri0, hc);
                        hpass(ri1, hd);
                        vpass_store(drow + dpitch, hc, hd, cyb0, cyb1);
                    }
                }
            };
            if (!two) groups(std::integral_constant<int, 0>());
            else if (reuse && !clamped) groups(std::integral_constant<int, 1>());
            else groups(std::integral_constant<int, 2>());
        }
    }
    if (seg) {
        const int sub = lane / seg, kq = lane & (seg - 1), ppi = 32 / seg;
        const bool has_word = kq < rem;
        const int k = words_main + (has_word ? kq : 0);
        const uint4 wg = __ldg(xw4 + k);
        const uint2 sx = __ldg(xw2 + k);
        const uint32_t sel01 = sx.x & 0xffffu, sel23 = sx.x >> 16;
        const int wofs = (int)(sx.y & 0xffffu);
        const uint32_t shift = sx.y >> 16;
        uint8_t* const dcol = dst + 4 * k;
        auto hpass = [&](int j, uint32_t (&hv)[4]) {
            const uint32_t* p = sm_w + (j - jlo) * spitch_w + wofs;
            const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
            const uint32_t lo = __funnelshift_r(w0, w1, shift), hi = __funnelshift_r(w1, w2, shift);
            const uint32_t b01 = __byte_perm(lo, hi, sel01), b23 = __byte_perm(lo, hi, sel23);
            hv[0] = __dp2a_lo(wg.x, b01, 0u);
            hv[1] = __dp2a_hi(wg.y, b01, 0u);
            hv[2] = __dp2a_lo(wg.z, b23, 0u);
            hv[3] = __dp2a_hi(wg.w, b23, 0u);
        };
        auto row_out = [&](int y, bool on) {  // rows differ between the lanes of the warp here: no reuse of passes
            const uint32_t ty = __ldg(ytab + y);
            const int j0 = (int)(ty & 0xffffu);
            const int j1 = min(j0 + 1, sh - 1);
            uint32_t h0[4], h1[4], v[4];
            hpass(j0, h0);
            hpass(j1, h1);
            const uint32_t cy1 = ty >> 16, cy0 = 256u - cy1;
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = h1[j] * cy1 + (h0[j] * cy0 + (1u << 15));
            if (on)
                *reinterpret_cast<uint32_t*>(dcol + (uint32_t)(y * dpitch)) =
                    __byte_perm(__byte_perm(v[0], v[1], 0x0062u), __byte_perm(v[2], v[3], 0x0062u), 0x5410u);
        };
        const int n_pairs = (y_end - y_begin + 1) >> 1;
        for (int p0 = warp * ppi; p0 < n_pairs; p0 += 8 * ppi) {
            const int y = y_begin + 2 * (p0 + sub);
            row_out(min(y, y_end - 1), has_word && y < y_end);
            row_out(min(y + 1, y_end - 1), has_word && y + 1 < y_end);
        }
    }
}

// FULL: 16-row bands of a large batch — eight row pairs side by side, one warp each, rows outside and word groups
// inside (the throughput shape, 40 registers).  !FULL: any band height, pairs_par pairs side by side and the per-word tables prefetched while the
// band is in flight (the latency shape for a few frames: 39 registers, which would cost the batched case a CTA
// per SM).
#ifndef VXORB_PYR_FULL_CTAS
#define VXORB_PYR_FULL_CTAS 6  // resident CTAs per SM asked of the throughput shape: 40 registers (1: 83, 4: 64, 8: 32 + spills; measured)
#endif
template <bool FULL>
__global__ void __launch_bounds__(256, FULL ? VXORB_PYR_FULL_CTAS : 1)
resize_band_kernel(const uint8_t* __restrict__ src_plane, int sh, int spitch, size_t sfstride,
                   uint8_t* __restrict__ dst, int dw, int dh, int dpitch, size_t dfstride,
                   const uint32_t* __restrict__ ytab, const uint4* __restrict__ xw4, const uint2* __restrict__ xw2,
                   int words, int band_rows, int n_bands, int pairs_par, int write_apron) {
    extern __shared__ __align__(128) uint8_t rows_sm[];
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    const int y_begin = b * band_rows, y_end = (b == n_bands - 1) ? dh : y_begin + band_rows;
    const int jlo = (int)(__ldg(ytab + y_begin) & 0xffffu);
    const int jhi = min((int)(__ldg(ytab + y_end - 1) & 0xffffu) + 1, sh - 1);
    dst += (size_t)blockIdx.y * dfstride;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const uint32_t bytes = (uint32_t)(jhi - jlo + 1) * (uint32_t)spitch;
        const uint8_t* g = src_plane + (size_t)blockIdx.y * sfstride + (size_t)(jlo + kApronY) * spitch;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_addr(rows_sm)),
                     "l"(g), "r"(bytes), "r"(smem_addr(&bar))
                     : "memory");
    }
    // pairs_par (8, 4, 2 or 1) row pairs are worked on side by side; the 8 / pairs_par warps of a pair split
    // its words 32 at a time, so a short band (small batches: more, shorter bands) still occupies all 8 warps
    if (FULL) pairs_par = 8;
    const int pw = warp & (pairs_par - 1), gw = warp / pairs_par, gstep = (8 / pairs_par) * 32;
    // the per-word tables of the first group are fetched while the band is still in flight
    uint4 wg_next = make_uint4(0, 0, 0, 0);
    uint2 sx_next = make_uint2(0, 0);
    if (!FULL && gw * 32 < words) {
        const int kk = min(gw * 32 + lane, words - 1);
        wg_next = __ldg(xw4 + kk);
        sx_next = __ldg(xw2 + kk);
    }
    __syncthreads();  // barrier initialised before anyone polls it
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "RW_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra RD_%=;\n"
        "bra RW_%=;\n"
        "RD_%=:\n"
        "}\n" ::"r"(smem_addr(&bar)),
        "r"(0)
        : "memory");

    const uint32_t* sm_w = reinterpret_cast<const uint32_t*>(rows_sm);
    const int spitch_w = spitch >> 2;
    if (FULL) band_rows_full(sm_w, jlo, jhi, sh, spitch_w, dst, dpitch, y_begin, y_end, ytab, xw4, xw2, words, lane, warp);
    for (int k0 = FULL ? words : gw * 32; k0 < words; k0 += gstep) {
        const int k = k0 + lane;
        const bool active = k < words;
        uint4 wg;
        uint2 sx;
        if (FULL) {
            const int kk = active ? k : words - 1;
            wg = __ldg(xw4 + kk);
            sx = __ldg(xw2 + kk);
        } else {
            wg = wg_next;
            sx = sx_next;
            if (k0 + gstep < words) {
                const int kn = min(k + gstep, words - 1);
                wg_next = __ldg(xw4 + kn);
                sx_next = __ldg(xw2 + kn);
            }
        }
        const uint32_t sel01 = sx.x & 0xffffu, sel23 = sx.x >> 16;
        const int wofs = (int)(sx.y & 0xffffu);
        const uint32_t shift = sx.y >> 16;
        uint8_t* const dcol = dst + 4 * k;  // a level plane is far below 4 GB: 32-bit row offsets from here
        // horizontal pass of one source row for this lane's output word: 4 x (p0 * c0 + p1 * c1), 8.8 fixed point
        auto hpass = [&](int j, uint32_t (&hv)[4]) {
            VX_CHECK(j >= jlo && j <= jhi && wofs + 2 < spitch_w);  // inside the band's source chunk
            const uint32_t* p = sm_w + (j - jlo) * spitch_w + wofs;
            const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
            const uint32_t lo = __funnelshift_r(w0, w1, shift), hi = __funnelshift_r(w1, w2, shift);
            const uint32_t b01 = __byte_perm(lo, hi, sel01), b23 = __byte_perm(lo, hi, sel23);
            hv[0] = __dp2a_lo(wg.x, b01, 0u);
            hv[1] = __dp2a_hi(wg.y, b01, 0u);
            hv[2] = __dp2a_lo(wg.z, b23, 0u);
            hv[3] = __dp2a_hi(wg.w, b23, 0u);
        };
        auto vpass_store = [&](int y, const uint32_t (&h0)[4], const uint32_t (&h1)[4], uint32_t cy1) {
            const uint32_t cy0 = 256u - cy1;
            uint32_t v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = h1[j] * cy1 + (h0[j] * cy0 + (1u << 15));  // 16.16, byte 2 = result
            if (active)
                *reinterpret_cast<uint32_t*>(dcol + (uint32_t)(y * dpitch)) =
                    __byte_perm(__byte_perm(v[0], v[1], 0x0062u), __byte_perm(v[2], v[3], 0x0062u), 0x5410u);
        };
        // a warp takes PAIRS of consecutive output rows: at scale 1.2 the second row's upper source row is
        // the first row's lower one five times out of six, so its horizontal pass is reused
        for (int y = y_begin + 2 * pw; y < y_end; y += 2 * pairs_par) {
            const uint32_t ty = __ldg(ytab + y);
            const int j0 = (int)(ty & 0xffffu);
            const int j1 = min(j0 + 1, sh - 1);
            uint32_t ha[4], hb[4];
            hpass(j0, ha);
            hpass(j1, hb);
            vpass_store(y, ha, hb, ty >> 16);
            if (y + 1 < y_end) {
                const uint32_t tz = __ldg(ytab + y + 1);
                const int i0 = (int)(tz & 0xffffu);
                const int i1 = min(i0 + 1, sh - 1);
                uint32_t hc[4];
                if (i0 == j1) {          // warp-uniform: depends on the row only
                    if (i1 == i0) {      // clamped at the last source row
                        vpass_store(y + 1, hb, hb, tz >> 16);
                    } else {
                        hpass(i1, hc);
                        vpass_store(y + 1, hb, hc, tz >> 16);
                    }
                } else {
                    hpass(i0, ha);
                    hpass(i1, hc);
                    vpass_store(y + 1, ha, hc, tz >> 16);
                }
            }
        }
    }
    // A few frames: the apron is only read by the blur, which runs on the side stream — the caller builds it there
    // (apron_kernel) and the level chain, the critical path of a single-frame call, does not pay for reading its own
    // pixels back twice per level.
    if (!write_apron) return;
    __syncthreads();  // the band's pixels are visible to this CTA

    // reflect-101 apron, left / right: x = -1,-2,-3 <- 1,2,3 ; x = w,w+1,w+2 <- w-2,w-3,w-4
    const int nrows = y_end - y_begin;
    for (int i = tid; i < nrows * 6; i += 256) {
        const int r = i / 6, c = i - r * 6;
        uint8_t* row = dst + (size_t)(y_begin + r) * dpitch;
        if (c < 3) row[-1 - c] = row[1 + c];
        else row[dw + (c - 3)] = row[dw - 2 - (c - 3)];
    }
    if (b != 0 && b != n_bands - 1) return;
    __syncthreads();
    // top: rows -1,-2,-3 <- 1,2,3 ; bottom: rows h,h+1,h+2 <- h-2,h-3,h-4 (x from -3 to w+2)
    const int span = dw + 2 * kApronY;
    for (int i = tid; i < 3 * span; i += 256) {
        const int r = i / span, x = i - r * span - kApronY;
        if (b == 0) dst[(ptrdiff_t)(-1 - r) * dpitch + x] = dst[(ptrdiff_t)(1 + r) * dpitch + x];
        if (b == n_bands - 1) dst[(ptrdiff_t)(dh + r) * dpitch + x] = dst[(ptrdiff_t)(dh - 2 - r) * dpitch + x];
    }
}

static void launch_resize(const uint8_t* src, int sw, int sh, int spitch, size_t sfstride, uint8_t* dst, int dw,
                          int dh, int dpitch, size_t dfstride, const uint32_t* xtab, const uint32_t* ytab, int batch,
                          bool fast, cudaStream_t s) {
    if (fast) {
        const int word_groups = ((dw + 3) / 4 + 31) / 32;
        const int rows = resize_rows_for(dw, dh, batch);
        const int items = word_groups * ((dh + rows - 1) / rows);
        dim3 grid((items + 7) / 8, batch);
        resize_kernel<<<grid, 256, 0, s>>>(src, sh, spitch, sfstride, dst, dw, dh, dpitch, dfstride, xtab, ytab,
                                           word_groups, rows);
    } else {
        dim3 block(32, 8), grid((dw + 127) / 128, (dh + 7) / 8, batch);
        resize_generic_kernel<<<grid, block, 0, s>>>(src, sw, sh, spitch, sfstride, dst, dw, dh, dpitch, dfstride, xtab,
                                                     ytab);
    }
}

// Densely packed frames (row stride = width) -> level-0 planes.  The streaming path moves a chunk over
// PCIe as ONE flat copy into a staging buffer (a strided 3-D copy of 640-byte rows reaches ~2 % less
// of the link) and this kernel scatters the rows into the pitched, aproned planes: 16 B per thread.
__global__ void __launch_bounds__(256)
unpack_frames_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int w16, int h, int pitch,
                     size_t fstride) {
    const int i = blockIdx.x * 256 + threadIdx.x;  // 16-byte chunk within the frame
    if (i >= w16 * h) return;
    const int y = i / w16, c = i - y * w16;
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + (size_t)blockIdx.y * ((size_t)w16 * 16 * h)) + i);
    *reinterpret_cast<uint4*>(dst + (size_t)blockIdx.y * fstride + (size_t)y * pitch + (size_t)c * 16) = v;
}

int launch_unpack_frames(const uint8_t* src, uint8_t* dst, int w, int h, int pitch, size_t fstride, int batch,
                         cudaStream_t s) {
    const int w16 = w / 16;
    dim3 grid((w16 * h + 255) / 256, batch);
    unpack_frames_kernel<<<grid, 256, 0, s>>>(src, dst, w16, h, pitch, fstride);
    return 1;
}

int launch_gray_from_bgr(const uint8_t* bgr, size_t bgr_pitch, size_t bgr_fstride, uint8_t* dst, int w,
                         int h, int pitch, size_t fstride, int batch, cudaStream_t s) {
    dim3 block(32, 8), grid((w + 127) / 128, (h + 7) / 8, batch);
    bgr2gray_kernel<<<grid, block, 0, s>>>(bgr, bgr_pitch, bgr_fstride, dst, w, h, pitch, fstride);
    return 1;
}

int launch_resize_single(const uint8_t* src, int sw, int sh, int spitch, uint8_t* dst, int dw, int dh,
                         int dpitch, const uint32_t* xtab, const uint32_t* ytab, bool fast, cudaStream_t s) {
    launch_resize(src, sw, sh, spitch, 0, dst, dw, dh, dpitch, 0, xtab, ytab, 1, fast, s);
    return 1;
}

// Picks the band height of one level: the largest of 16 / 8 / 4 rows whose biggest source chunk
// (the last band takes the remainder rows, so it is the largest) stays under 40 KB of shared memory.
#ifndef VXORB_BAND_ROWS
#define VXORB_BAND_ROWS 16
#endif
// Band count and the shared-memory bytes of the largest source chunk for bands of `rows` output rows (the last
// band takes the remainder rows, so it is the largest).
void resize_band_extent(const uint32_t* ytaps, int dh, int sh, int spitch, int rows, int* n_bands, long long* smem) {
    const int nb = std::max(1, dh / rows);
    int worst = 0;
    for (int b = 0; b < nb; ++b) {
        const int y0 = b * rows, y1 = (b == nb - 1) ? dh : y0 + rows;
        const int jlo = (int)(ytaps[y0] & 0xffffu), jhi = std::min((int)(ytaps[y1 - 1] & 0xffffu) + 1, sh - 1);
        worst = std::max(worst, jhi - jlo + 1);
    }
    *n_bands = nb;
    *smem = (long long)worst * spitch;
}

bool plan_resize_bands(const uint32_t* ytaps, int dh, int sh, int spitch, int* band_rows, int* n_bands, int* smem) {
    if (dh < 8) return false;
    for (int rows = VXORB_BAND_ROWS; rows >= 4; rows >>= 1) {
        int nb;
        long long bytes;
        resize_band_extent(ytaps, dh, sh, spitch, rows, &nb, &bytes);
        if (bytes <= 40 * 1024 || (rows == 4 && bytes <= 200 * 1024)) {
            *band_rows = rows;
            *n_bands = nb;
            *smem = (int)bytes;
            return true;
        }
    }
    return false;
}

// One level through the production band kernel (launch_pyramid and the stage-isolation entry share this):
// rows == 16 selects the throughput shape, 8 / 4 the latency shape.
void launch_resize_band(const uint8_t* src_plane, int sh, int spitch, size_t sfstride, uint8_t* dst, int dw, int dh,
                        int dpitch, size_t dfstride, const uint32_t* ytab, const uint4* xw4, const uint2* xw2, int rows,
                        int n_bands, int smem, int batch, cudaStream_t s, bool write_apron) {
    const int words = (dw + 3) / 4;
    dim3 grid(n_bands, batch);
    auto kern = rows == 16 ? resize_band_kernel<true> : resize_band_kernel<false>;
    kern<<<grid, 256, smem, s>>>(src_plane, sh, spitch, sfstride, dst, dw, dh, dpitch, dfstride, ytab, xw4, xw2, words, rows,
                                 n_bands, rows / 2, write_apron ? 1 : 0);
}

cudaError_t init_pyramid_kernels() {
    cudaError_t e = cudaFuncSetAttribute(resize_band_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(resize_band_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

cudaError_t carveout_pyramid(int pct) {
    const cudaError_t e = set_carveout(resize_band_kernel<true>, pct);
    return e != cudaSuccess ? e : set_carveout(resize_band_kernel<false>, pct);
}

// Builds levels 1.. of every frame and the reflect-101 apron of every level.  Returns launches.
int launch_pyramid(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s, bool aprons_inline) {
    int launches = 0;
    // aprons_inline == false (only for fewer than 8 frames on the band path): the caller builds the apron of EVERY level
    // itself (launch_apron over all levels, beside the chain's consumers that do not read it)
    const bool need = lt.blur_margin < kApronY;  // the apron is read by a whole-level blur only
    const bool level0_apron = aprons_inline && need;
    if (lt.resize_band) {
        // level 0 comes from the upload; the bands write their own apron.  The resampling reads no apron byte of
        // its source, so a caller may build level 0's apron elsewhere (beside the chain) and pass false.
        if (level0_apron) launches += launch_apron(lt, db, 1, s);
        for (int l = 1; l < lt.n_levels; ++l) {
            const int words = (lt.w[l] + 3) / 4;
            const uint32_t* tab = db.coef + lt.coef_off[l];
            const uint4* xw4 = reinterpret_cast<const uint4*>(db.coef + lt.xw_off[l]);
            const uint2* xw2 = reinterpret_cast<const uint2*>(db.coef + lt.xw_off[l] + 4 * words);
            const uint8_t* src_plane = db.pyr + lt.off[l - 1] - plane_origin(lt.w[l - 1]);
            // A few frames alone leave most SMs without a band: 4-row bands then (4x the CTAs, the level chain is
            // the critical path of a single-frame call).  The planned shared memory covers the tallest planned
            // band, which is taller than any 4..7-row band.
            const bool small = lt.batch < 8 && lt.band_rows[l] > 4;
            const int rows = small ? 4 : lt.band_rows[l];
            const int nb = small ? std::max(1, lt.h[l] / 4) : lt.n_bands[l];
            launch_resize_band(src_plane, lt.h[l - 1], lt.pitch[l - 1], (size_t)lt.fstride[l - 1], db.pyr + lt.off[l], lt.w[l],
                               lt.h[l], lt.pitch[l], (size_t)lt.fstride[l], tab + lt.w[l], xw4, xw2, rows, nb, lt.band_smem[l],
                               lt.batch, s, need && (aprons_inline || lt.batch >= 8));
            ++launches;
        }
        return launches;
    }
    for (int l = 1; l < lt.n_levels; ++l) {
        launch_resize(db.pyr + lt.off[l - 1], lt.w[l - 1], lt.h[l - 1], lt.pitch[l - 1], (size_t)lt.fstride[l - 1],
                      db.pyr + lt.off[l], lt.w[l], lt.h[l], lt.pitch[l], (size_t)lt.fstride[l],
                      db.coef + lt.coef_off[l], db.coef + lt.coef_off[l] + lt.w[l], lt.batch, lt.resize_fast[l] != 0, s);
        ++launches;
    }
    if (need) launches += launch_apron(lt, db, lt.n_levels, s);
    return launches;
}

// Reflect-101 apron (3 px) around every level of every frame, so the blur reads straight through
// the image border exactly as cv::GaussianBlur(BORDER_REFLECT_101) sees it.  One thread per apron
// pixel; items of a level: rows -3..-1 and h..h+2 over x in [-3, w+2], then columns -3..-1 and
// w..w+2 over y in [0, h).
__device__ __forceinline__ int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}

__global__ void __launch_bounds__(256) apron_kernel(const LevelTab lt, uint8_t* __restrict__ pyr, int levels) {
    int idx = blockIdx.x * 256 + threadIdx.x;
    const int f = blockIdx.y;
    for (int l = 0; l < levels; ++l) {
        const int w = lt.w[l], h = lt.h[l];
        const int horiz = 2 * kApronY * (w + 2 * kApronY), vert = 2 * kApronY * h;
        if (idx < horiz + vert) {
            int x, y;
            if (idx < horiz) {
                const int r = idx / (w + 2 * kApronY);
                x = idx - r * (w + 2 * kApronY) - kApronY;
                y = r < kApronY ? r - kApronY : h + (r - kApronY);
            } else {
                const int j = idx - horiz;
                y = j / (2 * kApronY);
                const int c = j - y * (2 * kApronY);
                x = c < kApronY ? c - kApronY : w + (c - kApronY);
            }
            uint8_t* img = pyr + lt.off[l] + (size_t)f * lt.fstride[l];
            const int pitch = lt.pitch[l];
            img[(ptrdiff_t)y * pitch + x] = img[(ptrdiff_t)reflect101(y, h) * pitch + reflect101(x, w)];
            return;
        }
        idx -= horiz + vert;
    }
}

int launch_apron(const LevelTab& lt, const DeviceBuffers& db, int levels, cudaStream_t s) {
    int items = 0;
    levels = std::min(levels, lt.n_levels);
    for (int l = 0; l < levels; ++l) items += 2 * kApronY * (lt.w[l] + 2 * kApronY) + 2 * kApronY * lt.h[l];
    if (items == 0) return 0;
    dim3 grid((items + 255) / 256, lt.batch);
    apron_kernel<<<grid, 256, 0, s>>>(lt, db.pyr, levels);
    return 1;
}

// Streams `bytes` of writes so the next kernel starts from a cold L2 (bench hygiene).
__global__ void flush_kernel(uint4* p, size_t n16, uint32_t tag) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n16; i += stride) p[i] = make_uint4(tag, tag, tag, tag);
}

int launch_flush(uint8_t* buf, size_t bytes, cudaStream_t s) {
    flush_kernel<<<148 * 8, 256, 0, s>>>(reinterpret_cast<uint4*>(buf), bytes / 16, 0u);
    return 1;
}

}  // namespace vxorb
