// fast.cu — FAST-9/16 corner detection with cornerScore and strict 3x3 non-max suppression.
//
// Replaces cv::FastFeatureDetector(20, true)::detect on every pyramid level plus
// KeyPointsFilter::runByImageBorder(31), both inside cv::ORB::detectAndCompute
// (orb_extractor.cpp:13).  Semantics: SURVEY.md A.2/A.3 — a pixel is a corner iff >= 9 contiguous
// ring pixels are all > v+t or all < v-t; score = max over the 16 arcs of 9 of min|v - p| minus 1;
// kept iff its score is strictly greater than the 8 neighbours' scores.  Integer work, bit-exact.
//
// One CTA handles a 64x32 tile of one (level, frame); the score map never leaves shared memory.
//   load   : (32+8) rows x 96 B of 16-byte vectors (3-px ring + 1-px NMS halo, 16 B aligned start)
//   pass A : every pixel of the (64+2)x(32+2) score domain runs the 4 opposite-pair quick test;
//            survivors are pushed to a shared queue (ballot + popc), so later passes stay converged
//   pass B : full 16-pixel segment test + score for queued pixels (arc minima in packed s16x2)
//   pass C : NMS on the shared score map, survivors staged in shared memory
//   emit   : one global atomicAdd per CTA, coalesced write of packed corners
// Roofline: HBM, algorithmic bytes = one read of every level pixel (950 532 B per 640x480 frame).
#include "vxorb_internal.h"

namespace vxorb {

namespace {

constexpr int TW = 64, TH = 32;
constexpr int SMW = 96;           // shared row bytes (6 x uint4)
constexpr int SMH = TH + 8;
constexpr int SCW = TW + 2, SCH = TH + 2;
constexpr int SCP = SCW + 2;      // padded score row
constexpr int QCAP = SCW * SCH;
constexpr int OCAP = (TW / 2 + 1) * (TH / 2 + 1);

__device__ __forceinline__ unsigned has_run9(unsigned m) {
    m |= m << 16;
    unsigned x = m & (m >> 1);
    x &= x >> 2;
    x &= x >> 4;
    x &= m >> 8;
    return x & 0xffffu;
}

// class bits of ring pixel p against centre band [lo, hi]: 1 = darker, 2 = brighter
__device__ __forceinline__ unsigned cls(int p, int lo, int hi) {
    return (p < lo ? 1u : 0u) | (p > hi ? 2u : 0u);
}

}  // namespace

__global__ void __launch_bounds__(256)
fast_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, uint32_t* __restrict__ cand,
            int* __restrict__ cand_n, int threshold, int border) {
    __shared__ __align__(16) uint8_t tile[SMH][SMW];
    __shared__ uint8_t score[SCH][SCP];
    __shared__ uint16_t queue[QCAP];
    __shared__ uint32_t outq[OCAP];
    __shared__ int qn, on, obase;

    const int tid = threadIdx.x;
    int l = 0;
    {
        const int t = blockIdx.x;
        while (l + 1 < lt.n_levels && t >= lt.fast_tile_begin[l + 1]) ++l;
    }
    const int tl = blockIdx.x - lt.fast_tile_begin[l];
    const int tx = tl % lt.fast_tiles_x[l], ty = tl / lt.fast_tiles_x[l];
    const int f = blockIdx.y;
    const int w = lt.w[l], h = lt.h[l], pitch = lt.pitch[l];
    const uint8_t* img = pyr + lt.off[l] + (size_t)f * lt.fstride[l];

    const int ox = border + tx * TW, oy = border + ty * TH;  // tile origin in the emission domain
    const int x_al = (ox - 4) & ~15;                          // 16 B aligned left edge of the load
    const int ry0 = oy - 4;
    const int cx0 = ox - x_al;                                // shared column of pixel ox

    for (int i = tid; i < SMH * (SMW / 16); i += 256) {
        const int r = i / (SMW / 16), c = i % (SMW / 16);
        const int gy = ry0 + r, gx = x_al + c * 16;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (gy >= 0 && gy < h && gx >= 0 && gx < pitch)
            v = __ldg(reinterpret_cast<const uint4*>(img + (size_t)gy * pitch + gx));
        *reinterpret_cast<uint4*>(&tile[r][c * 16]) = v;
    }
    for (int i = tid; i < SCH * SCP; i += 256) (&score[0][0])[i] = 0;
    if (tid == 0) {
        qn = 0;
        on = 0;
    }
    __syncthreads();

    // score domain (global coordinates, inclusive)
    const int sx_min = max(3, border - 1), sx_max = min(w - 4, w - border);
    const int sy_min = max(3, border - 1), sy_max = min(h - 4, h - border);
    const int lane = tid & 31;

    // ---- pass A: quick reject on the 4 opposite pairs of compass / diagonal points ------------
    for (int i0 = 0; i0 < QCAP; i0 += 256) {
        const int i = i0 + tid;
        bool pass = false;
        int lx = 0, ly = 0;
        if (i < QCAP) {
            lx = i % SCW - 1;
            ly = i / SCW - 1;
            const int gx = ox + lx, gy = oy + ly;
            if (gx >= sx_min && gx <= sx_max && gy >= sy_min && gy <= sy_max) {
                const uint8_t* c = &tile[ly + 4][lx + cx0];
                const int v = c[0], lo = v - threshold, hi = v + threshold;
                unsigned d = cls(c[3 * SMW], lo, hi) | cls(c[-3 * SMW], lo, hi);          // 0, 8
                if (d) {
                    d &= cls(c[3], lo, hi) | cls(c[-3], lo, hi);                            // 4, 12
                    if (d) {
                        d &= cls(c[2 * SMW + 2], lo, hi) | cls(c[-2 * SMW - 2], lo, hi);   // 2, 10
                        d &= cls(c[-2 * SMW + 2], lo, hi) | cls(c[2 * SMW - 2], lo, hi);   // 6, 14
                        pass = d != 0;
                    }
                }
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (m) {
            int base = 0;
            if (lane == 0) base = atomicAdd(&qn, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (pass) queue[base + __popc(m & ((1u << lane) - 1u))] = (uint16_t)((ly + 1) * SCW + (lx + 1));
        }
    }
    __syncthreads();
    const int nq = qn;

    // ---- pass B: full segment test and score --------------------------------------------------
    for (int q = tid; q < nq; q += 256) {
        const int e = queue[q];
        const int lx = e % SCW - 1, ly = e / SCW - 1;
        const uint8_t* c = &tile[ly + 4][lx + cx0];
        const int v = c[0];
        int d[16];
        d[0] = v - c[3 * SMW];
        d[1] = v - c[3 * SMW + 1];
        d[2] = v - c[2 * SMW + 2];
        d[3] = v - c[SMW + 3];
        d[4] = v - c[3];
        d[5] = v - c[-SMW + 3];
        d[6] = v - c[-2 * SMW + 2];
        d[7] = v - c[-3 * SMW + 1];
        d[8] = v - c[-3 * SMW];
        d[9] = v - c[-3 * SMW - 1];
        d[10] = v - c[-2 * SMW - 2];
        d[11] = v - c[-SMW - 3];
        d[12] = v - c[-3];
        d[13] = v - c[SMW - 3];
        d[14] = v - c[2 * SMW - 2];
        d[15] = v - c[3 * SMW - 1];
        unsigned bm = 0, dm = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            bm |= (d[k] < -threshold ? 1u : 0u) << k;  // ring pixel brighter than v + t
            dm |= (d[k] > threshold ? 1u : 0u) << k;   // ring pixel darker than v - t
        }
        if (has_run9(bm) | has_run9(dm)) {
            // packed s16x2: low half d, high half -d; sliding minima over windows of 9
            unsigned p[16], m2[16], m4[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) p[k] = ((unsigned)d[k] & 0xffffu) | ((unsigned)(-d[k]) << 16);
#pragma unroll
            for (int k = 0; k < 16; ++k) m2[k] = __vmins2(p[k], p[(k + 1) & 15]);
#pragma unroll
            for (int k = 0; k < 16; ++k) m4[k] = __vmins2(m2[k], m2[(k + 2) & 15]);
            unsigned best = 0x80008000u;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                unsigned m8 = __vmins2(m4[k], m4[(k + 4) & 15]);
                best = __vmaxs2(best, __vmins2(m8, p[(k + 8) & 15]));
            }
            const int a = (int)(short)(best & 0xffffu), b = (int)(short)(best >> 16);
            score[ly + 1][lx + 1] = (uint8_t)(max(a, b) - 1);
        }
    }
    __syncthreads();

    // ---- pass C: strict 3x3 non-max suppression over the emission domain ----------------------
    const int ex_max = w - border, ey_max = h - border;  // exclusive
    for (int q = tid; q < nq; q += 256) {
        const int e = queue[q];
        const int sx = e % SCW, sy = e / SCW;  // score-map coordinates (= local + 1)
        const int lx = sx - 1, ly = sy - 1;
        const int s = score[sy][sx];
        if (s == 0 || lx < 0 || lx >= TW || ly < 0 || ly >= TH) continue;
        const int gx = ox + lx, gy = oy + ly;
        if (gx >= ex_max || gy >= ey_max) continue;
        const uint8_t* r = &score[sy][sx];
        if (s > r[-SCP - 1] && s > r[-SCP] && s > r[-SCP + 1] && s > r[-1] && s > r[1] && s > r[SCP - 1] &&
            s > r[SCP] && s > r[SCP + 1]) {
            const int o = atomicAdd(&on, 1);
            outq[o] = ((uint32_t)s << 24) | ((uint32_t)gy << 12) | (uint32_t)gx;
        }
    }
    __syncthreads();
    const int no = on;
    if (no == 0) return;
    const int cap = lt.cand_cap[l];
    int* cn = cand_n + l * lt.batch + f;
    if (tid == 0) obase = atomicAdd(cn, no);
    __syncthreads();
    uint32_t* out = cand + lt.cand_off[l] + (size_t)f * cap;
    for (int i = tid; i < no; i += 256)
        if (obase + i < cap) out[obase + i] = outq[i];
}

int launch_fast(const LevelTab& lt, const DeviceBuffers& db, int threshold, int border, cudaStream_t s) {
    const int tiles = lt.fast_tile_begin[lt.n_levels];
    if (tiles == 0) return 0;
    dim3 grid(tiles, lt.batch);
    fast_kernel<<<grid, 256, 0, s>>>(lt, db.pyr, db.cand, db.cand_n, threshold, border);
    return 1;
}

}  // namespace vxorb
