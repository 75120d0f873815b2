// match.cu — Hamming 2-NN (brute force, square-window, box or byte masked), ratio test, cross-check and the
// caller's post-filter, for a batch of independent (query, train) descriptor sets.
//
// Replaces cv::BFMatcher(cv::NORM_HAMMING).knnMatch(desc_last, desc_curr, knn, 2)
// (orb_matcher.cpp:22-25) and the ratio loop (orb_matcher.cpp:27-36); SURVEY.md A.9.  For each query
// row the two smallest (popcount(q ^ t), train index) keys are kept lexicographically, so ties go
// to the lowest train index exactly like OpenCV's batchDistance.  Integer work, bit-exact.
//
// knn kernel: one thread owns one query descriptor (8 x u32 in registers); the CTA streams the
// train set through shared memory in 256-row tiles which every thread reads as broadcast 16-byte
// vectors; per pair: 8 XOR + 4 carry-save adders (8 LOP3) + 4 POPC + 4 IMAD that assemble the packed key
// (dist * kKeyMul + idx), then a top-2 update of 2.5 min/max operations (two rows at a time).  Bound: the integer
// (alu) pipe, 80 % busy — not HBM (64 KB of descriptors per 1000x1000 pair).  Windowed / box masks over more than a
// few pairs go through the cell-binned kernel below.  compact kernel: one CTA per pair, ordered block scan.
#include <algorithm>

#include "vxorb_internal.h"

namespace vxorb {

namespace {

#ifndef VXORB_KNN_QPB
#define VXORB_KNN_QPB 128
#endif
#ifndef VXORB_KNN_TILE
#define VXORB_KNN_TILE 256
#endif
constexpr int kQPerBlock = VXORB_KNN_QPB;   // queries (threads) per CTA
constexpr int kTrainTile = VXORB_KNN_TILE;  // train rows staged in shared memory at a time
constexpr uint32_t kNoKey = 0x7fffffffu;
// Packed 2-NN key of the brute-force kernels: distance * kKeyMul + train index, ordered like the pair (distance,
// index), so ties go to the lowest index.  kKeyMul is 2^20 + 1 rather than 2^20: with a power of two ptxas turns
// the multiply-adds that assemble the key into shift-adds (LEA) on the alu pipe, which is the pipe that bounds the
// kernel; a non-power-of-two keeps them IMADs on the fma pipe.  Indices stay below 2^20.
#ifndef VXORB_KNN_PAIRUPD
#define VXORB_KNN_PAIRUPD 1
#endif
#ifndef VXORB_KEY_POW2
#define VXORB_KEY_POW2 0
#endif
constexpr uint32_t kKeyMul = VXORB_KEY_POW2 ? (1u << 20) : (1u << 20) + 1u;
__device__ __forceinline__ int key_dist(uint32_t k) { return (int)(k / kKeyMul); }
__device__ __forceinline__ int key_idx(uint32_t k) { return (int)(k - (k / kKeyMul) * kKeyMul); }
__device__ __forceinline__ int4 key_row(uint32_t k0, uint32_t k1) {
    return make_int4(k0 == kNoKey ? -1 : key_idx(k0), k1 == kNoKey ? -1 : key_idx(k1), k0 == kNoKey ? 0 : key_dist(k0),
                     k1 == kNoKey ? 0 : key_dist(k1));
}

// 256-bit Hamming distance of a query held in registers against a train row read as two 16-byte vectors, delivered
// as the packed 2-NN key (distance << 20 | train index).  The integer (alu) pipe is what bounds the kernel (85 % busy
// in the ncu capture, profiles/r2a_*), so the count is arranged to cost it as little as possible: the eight XOR words
// go through four carry-save adders (2 LOP3 each) down to two words of weight 1, one of weight 2 and one of weight 4
//     x0+x1+x2 = sa + 2 ca,  x3+x4+x5 = sb + 2 cb,  x6+x7+sa = sc + 2 cc,  ca+cb+cc = se + 2 ce
//     popc(x0..x7) = popc(sb) + popc(sc) + 2 popc(se) + 4 popc(ce)
// i.e. 16 LOP3 + 4 POPC per pair (8 LOP3 + 8 POPC + 7 IADD without the tree; the full Harley-Seal tree down to one word
// per weight costs 22 LOP3 for the same 4 POPC), and the weighted sum and the index go into the key with four IMADs,
// which run on the otherwise idle fma pipe.
__device__ __forceinline__ void csa(uint32_t a, uint32_t b, uint32_t c, uint32_t& sum, uint32_t& carry) {
    sum = a ^ b ^ c;
    carry = (a & b) | (c & (a ^ b));
}
__device__ __forceinline__ uint32_t hamming_key(const uint32_t (&q)[8], const uint4 lo, const uint4 hi, uint32_t idx) {
    const uint32_t x0 = q[0] ^ lo.x, x1 = q[1] ^ lo.y, x2 = q[2] ^ lo.z, x3 = q[3] ^ lo.w;
    const uint32_t x4 = q[4] ^ hi.x, x5 = q[5] ^ hi.y, x6 = q[6] ^ hi.z, x7 = q[7] ^ hi.w;
    uint32_t sa, ca, sb, cb, sc, cc, se, ce;
    csa(x0, x1, x2, sa, ca);
    csa(x3, x4, x5, sb, cb);
    csa(x6, x7, sa, sc, cc);
    csa(ca, cb, cc, se, ce);
    uint32_t key = (uint32_t)__popc(sb) * kKeyMul + idx;
    key = (uint32_t)__popc(sc) * kKeyMul + key;
    key = (uint32_t)__popc(se) * (2u * kKeyMul) + key;
    key = (uint32_t)__popc(ce) * (4u * kKeyMul) + key;
    return key;
}

}  // namespace

// QPT queries per thread (1 in the shipped build; 2 keeps two query descriptors in registers so that one pair of
// broadcast LDS.128 and one loop step serve two distances — see launch_knn2 for why it is not used).
#ifndef VXORB_KNN_MINCTAS
#define VXORB_KNN_MINCTAS 8  // asks ptxas for <= 64 registers; measured per 128 pairs of 1000 x 1000 alone / pipelined step: no second argument (64 registers, another schedule) 0.208 ms / 215.9 k frames/s, 1 (83) 0.185 / 215.2 k, 8 (64) 0.2015 / 217.8 k, 9 (56) 0.220 / 216.9 k, 10 (48) 0.222 / 217.0 k, 11 (40, spills) 0.241 / 214.0 k
#endif
template <int QPT>
__global__ void __launch_bounds__(kQPerBlock, VXORB_KNN_MINCTAS)
knn2_kernel(const MatchArgs a) {
    __shared__ __align__(16) uint32_t tdesc[kTrainTile][8];
    __shared__ float2 txy[kTrainTile];

    const int pair = blockIdx.y;
    const int nq = a.q_n[pair * a.n_stride], nt = a.t_n[pair * a.n_stride];
    const int q_first = blockIdx.x * (kQPerBlock * QPT);
    if (q_first >= nq) return;  // whole CTA idle
    const bool bytemask = a.mask != nullptr;
    const bool windowed = bytemask || ((a.radius > 0.0f || a.boxed) && a.q_xy != nullptr && a.t_xy != nullptr);
    // square window = the box [-r, r] x [-r, r] on (xq - xt, yq - yt): |d| <= r <=> -r <= d <= r in float
    const float bx0 = a.boxed ? a.box[0] : -a.radius, bx1 = a.boxed ? a.box[1] : a.radius;
    const float by0 = a.boxed ? a.box[2] : -a.radius, by1 = a.boxed ? a.box[3] : a.radius;

    const uint8_t* qbase = a.q_desc + (size_t)pair * a.q_pair_stride * 32;
    const uint8_t* tbase = a.t_desc + (size_t)pair * a.t_pair_stride * 32;
    uint32_t q[QPT][8];
    float qx[QPT], qy[QPT];
    int qi[QPT];
    bool live[QPT];
    uint32_t k0[QPT], k1[QPT];
#pragma unroll
    for (int u = 0; u < QPT; ++u) {
        qi[u] = q_first + u * kQPerBlock + (int)threadIdx.x;
        live[u] = qi[u] < nq;
        qx[u] = qy[u] = 0.f;
        k0[u] = k1[u] = kNoKey;
        if (live[u]) {
            const uint4 lo = __ldg(reinterpret_cast<const uint4*>(qbase + (size_t)qi[u] * 32));
            const uint4 hi = __ldg(reinterpret_cast<const uint4*>(qbase + (size_t)qi[u] * 32 + 16));
            q[u][0] = lo.x; q[u][1] = lo.y; q[u][2] = lo.z; q[u][3] = lo.w;
            q[u][4] = hi.x; q[u][5] = hi.y; q[u][6] = hi.z; q[u][7] = hi.w;
            if (windowed && !bytemask) {
                const float2 p = __ldg(reinterpret_cast<const float2*>(a.q_xy + ((size_t)pair * a.q_pair_stride + qi[u]) * 2));
                qx[u] = p.x;
                qy[u] = p.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) q[u][k] = 0;
        }
    }

    // this CTA's share of the train rows (whole tiles; the last split takes the remainder)
    const int tiles_all = (nt + kTrainTile - 1) / kTrainTile;
    const int tiles_per = (tiles_all + a.splits - 1) / a.splits;
    const int t_begin = min(nt, (int)blockIdx.z * tiles_per * kTrainTile);
    const int t_end = min(nt, t_begin + tiles_per * kTrainTile);
    for (int t0 = t_begin; t0 < t_end; t0 += kTrainTile) {
        const int tn = min(kTrainTile, t_end - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < tn * 2; i += kQPerBlock)
            reinterpret_cast<uint4*>(&tdesc[0][0])[i] =
                __ldg(reinterpret_cast<const uint4*>(tbase + (size_t)t0 * 32) + i);
        if (windowed && !bytemask)
            for (int i = threadIdx.x; i < tn; i += kQPerBlock)
                txy[i] = __ldg(reinterpret_cast<const float2*>(a.t_xy + ((size_t)pair * a.t_pair_stride + t0) * 2) + i);
        __syncthreads();
        if (!windowed) {
#if VXORB_KNN_PAIRUPD
            // two train rows per step: the two smallest of {k0, k1, a, b} in five min/max operations instead of six
            int j = 0;
#pragma unroll 2
            for (; j + 1 < tn; j += 2) {
                const uint4 lo = *reinterpret_cast<const uint4*>(&tdesc[j][0]);
                const uint4 hi = *reinterpret_cast<const uint4*>(&tdesc[j][4]);
                const uint4 lo2 = *reinterpret_cast<const uint4*>(&tdesc[j + 1][0]);
                const uint4 hi2 = *reinterpret_cast<const uint4*>(&tdesc[j + 1][4]);
#pragma unroll
                for (int u = 0; u < QPT; ++u) {
                    const uint32_t ka = hamming_key(q[u], lo, hi, (uint32_t)(t0 + j));
                    const uint32_t kb = hamming_key(q[u], lo2, hi2, (uint32_t)(t0 + j + 1));
                    const uint32_t m = min(ka, kb), M = max(ka, kb);
                    k1[u] = __vimin3_u32(max(k0[u], m), k1[u], M);
                    k0[u] = min(k0[u], m);
                }
            }
            if (j < tn) {
                const uint4 lo = *reinterpret_cast<const uint4*>(&tdesc[j][0]);
                const uint4 hi = *reinterpret_cast<const uint4*>(&tdesc[j][4]);
#pragma unroll
                for (int u = 0; u < QPT; ++u) {
                    const uint32_t key = hamming_key(q[u], lo, hi, (uint32_t)(t0 + j));
                    k1[u] = min(k1[u], max(k0[u], key));
                    k0[u] = min(k0[u], key);
                }
            }
#else
#pragma unroll 4
            for (int j = 0; j < tn; ++j) {
                const uint4 lo = *reinterpret_cast<const uint4*>(&tdesc[j][0]);
                const uint4 hi = *reinterpret_cast<const uint4*>(&tdesc[j][4]);
#pragma unroll
                for (int u = 0; u < QPT; ++u) {
                    const uint32_t key = hamming_key(q[u], lo, hi, (uint32_t)(t0 + j));
                    k1[u] = min(k1[u], max(k0[u], key));
                    k0[u] = min(k0[u], key);
                }
            }
#endif
        } else {
            for (int j = 0; j < tn; ++j) {
                const float2 p = bytemask ? make_float2(0.f, 0.f) : txy[j];
                bool in[QPT];
                bool any = false;
#pragma unroll
                for (int u = 0; u < QPT; ++u) {
                    if (bytemask) {  // a thread walks along its own mask row: consecutive bytes, cached lines
                        in[u] = live[u] && __ldg(a.mask + (size_t)qi[u] * a.mask_stride + (t0 + j)) != 0;
                    } else {
                        const float dx = __fsub_rn(qx[u], p.x), dy = __fsub_rn(qy[u], p.y);
                        in[u] = dx >= bx0 && dx <= bx1 && dy >= by0 && dy <= by1;
                    }
                    any = any || in[u];
                }
                if (__any_sync(0xffffffffu, any)) {
                    const uint4 lo = *reinterpret_cast<const uint4*>(&tdesc[j][0]);
                    const uint4 hi = *reinterpret_cast<const uint4*>(&tdesc[j][4]);
#pragma unroll
                    for (int u = 0; u < QPT; ++u) {
                        const uint32_t key = in[u] ? hamming_key(q[u], lo, hi, (uint32_t)(t0 + j)) : kNoKey;
                        k1[u] = min(k1[u], max(k0[u], key));
                        k0[u] = min(k0[u], key);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int u = 0; u < QPT; ++u) {
        if (!live[u]) continue;
        if (a.splits > 1) {
            a.part[((size_t)pair * a.splits + blockIdx.z) * a.max_q + qi[u]] = make_uint2(k0[u], k1[u]);
        } else {
            reinterpret_cast<int4*>(a.knn + (size_t)pair * a.knn_stride)[qi[u]] = key_row(k0[u], k1[u]);
        }
    }
}

// Folds the per-split packed top-2 keys of the split brute-force launch into the final 2-NN rows.  The
// keys carry the global train index, so the two smallest over all splits are the two nearest
// neighbours with OpenCV's lowest-index tie-break.
__global__ void __launch_bounds__(256)
knn2_merge_kernel(const MatchArgs a) {
    const int pair = blockIdx.y;
    const int qi = blockIdx.x * 256 + threadIdx.x;
    if (qi >= a.q_n[pair * a.n_stride]) return;
    uint32_t k0 = kNoKey, k1 = kNoKey;
    for (int sp = 0; sp < a.splits; ++sp) {
        const uint2 p = a.part[((size_t)pair * a.splits + sp) * a.max_q + qi];
        k1 = min(k1, max(k0, p.x));
        k0 = min(k0, p.x);
        k1 = min(k1, p.y);  // p.y >= p.x: it can only displace the second entry
    }
    reinterpret_cast<int4*>(a.knn + (size_t)pair * a.knn_stride)[qi] = key_row(k0, k1);
}

// ---- windowed 2-NN over a uniform cell grid ------------------------------------------------------
// The brute-force kernel above tests every (query, train) pair against the window.  With a square
// window of radius r only ~(2r)^2 / (W*H) of the train rows can pass (1.3 % at r = 32 on 640x480),
// so this kernel first bins the train keypoints of the pair into a kGridX x kGridY grid of
// cells at least r / 2 wide (counting sort in shared memory: histogram, block scan, scatter of
// descriptor + position + original index in cell order), and every query thread then visits only
// the cells its window touches: <= 5 x 5 cells (80 x 80 px at r = 32 against the 65 x 65 window), each cell row one contiguous range of the sorted
// list.  The exact window test (same float expression as the brute-force kernel) is still applied
// per candidate, and the packed key (distance << 20 | original train index) keeps OpenCV's
// tie-break (lowest train index) whatever the visiting order.  Cell lookup is conservative: the
// cell range is computed from a slightly widened radius with the same monotonic float expression
// that bins the train points, so no in-window row can be missed.
#ifndef VXORB_WIN_GX
#define VXORB_WIN_GX 64
#endif
#ifndef VXORB_WIN_GY
#define VXORB_WIN_GY 32
#endif
#ifndef VXORB_WIN_DIV
#define VXORB_WIN_DIV 2
#endif
constexpr int kGridX = VXORB_WIN_GX, kGridY = VXORB_WIN_GY;
constexpr int kCells = kGridX * kGridY;
constexpr int kWinThreads = 1024;
constexpr int kCellsPerThread = kCells / kWinThreads;
constexpr float kCellPerRadius = 1.0f / VXORB_WIN_DIV;  // cells are at least radius / 2 wide: a window touches <= 5 x 5
static_assert(kCells % kWinThreads == 0 && kCellsPerThread >= 1, "whole cells per thread in the block scan");

__device__ __forceinline__ int cell_of(float v, float org, float inv, int n) {
    const float c = floorf(__fmul_rn(__fsub_rn(v, org), inv));
    return (int)fminf(fmaxf(c, 0.0f), (float)(n - 1));
}

static size_t window_smem_bytes(int max_t, int max_q) {
    return (size_t)max_t * (32 + 8 + 4) + (size_t)(2 * kCells + 1) * 4 + 64 + (size_t)((max_q + 1) & ~1) * 2;
}

__global__ void __launch_bounds__(kWinThreads)
knn2_window_kernel(const MatchArgs a, int max_t, int max_q) {
    extern __shared__ __align__(16) uint8_t wsm[];
    uint4* tdesc = reinterpret_cast<uint4*>(wsm);                                  // [max_t][2]
    float2* txy = reinterpret_cast<float2*>(wsm + (size_t)max_t * 32);             // [max_t]
    int* tidx = reinterpret_cast<int*>(wsm + (size_t)max_t * 40);                  // [max_t]
    int* cell_start = tidx + max_t;                                                // [kCells + 1]
    int* cursor = cell_start + kCells + 1;                                         // [kCells]
    float* bounds = reinterpret_cast<float*>(cursor + kCells);                     // minx, miny, maxx, maxy
    uint16_t* qorder = reinterpret_cast<uint16_t*>(bounds + 16);                   // [max_q] query rows in cell order
    __shared__ int warp_tot[kWinThreads / 32];
    __shared__ int s_rank_lo, s_rank_hi;

    const int pair = blockIdx.x;
    const int nq = min(a.q_n[pair * a.n_stride], max_q), nt = min(a.t_n[pair * a.n_stride], max_t);
    if (nq <= 0) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t* qbase = a.q_desc + (size_t)pair * a.q_pair_stride * 32;
    const uint8_t* tbase = a.t_desc + (size_t)pair * a.t_pair_stride * 32;
    const float2* txy_g = reinterpret_cast<const float2*>(a.t_xy + (size_t)pair * a.t_pair_stride * 2);
    // the mask as a box on (query - train): the square window of radius r is {-r, r, -r, r} (|d| <= r <=> -r <= d <= r
    // in float), a stereo / epipolar band is e.g. {0, max_disparity, -dy_tol, dy_tol}
    const float bx0 = a.boxed ? a.box[0] : -a.radius, bx1 = a.boxed ? a.box[1] : a.radius;
    const float by0 = a.boxed ? a.box[2] : -a.radius, by1 = a.boxed ? a.box[3] : a.radius;

    // ---- 1. bounding box of the train positions -> grid geometry ------------------------------
    {
        float mnx = 3.0e38f, mny = 3.0e38f, mxx = -3.0e38f, mxy = -3.0e38f;
        for (int i = tid; i < nt; i += kWinThreads) {
            const float2 p = __ldg(txy_g + i);
            mnx = fminf(mnx, p.x); mxx = fmaxf(mxx, p.x);
            mny = fminf(mny, p.y); mxy = fmaxf(mxy, p.y);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o));
            mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
            mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
            mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
        }
        float* wb = reinterpret_cast<float*>(cursor);  // scratch: 8 warps x 4 floats
        if (lane == 0) {
            wb[warp * 4 + 0] = mnx; wb[warp * 4 + 1] = mny; wb[warp * 4 + 2] = mxx; wb[warp * 4 + 3] = mxy;
        }
        for (int i = tid; i < kCells + 1; i += kWinThreads) cell_start[i] = 0;
        __syncthreads();
        if (tid == 0) {
            for (int wv = 1; wv < kWinThreads / 32; ++wv) {
                mnx = fminf(mnx, wb[wv * 4 + 0]); mny = fminf(mny, wb[wv * 4 + 1]);
                mxx = fmaxf(mxx, wb[wv * 4 + 2]); mxy = fmaxf(mxy, wb[wv * 4 + 3]);
            }
            bounds[0] = mnx; bounds[1] = mny; bounds[2] = mxx; bounds[3] = mxy;
        }
        __syncthreads();
    }
    const float orgx = bounds[0], orgy = bounds[1];
    // cells at least a quarter of the box wide (= r / 2 for the square window), and wide enough that kGridX x kGridY
    // of them cover the bounding box; never zero (a degenerate box over coincident points)
    const float csx = fmaxf(fmaxf((bx1 - bx0) * (0.5f * kCellPerRadius), (bounds[2] - orgx) * (1.0f / kGridX) * 1.0001f), 1.0e-3f);
    const float csy = fmaxf(fmaxf((by1 - by0) * (0.5f * kCellPerRadius), (bounds[3] - orgy) * (1.0f / kGridY) * 1.0001f), 1.0e-3f);
    const float invx = 1.0f / csx, invy = 1.0f / csy;

    // ---- 2. counting sort of the train rows by cell --------------------------------------------
    for (int i = tid; i < nt; i += kWinThreads) {
        const float2 p = __ldg(txy_g + i);
        atomicAdd(&cell_start[cell_of(p.y, orgy, invy, kGridY) * kGridX + cell_of(p.x, orgx, invx, kGridX)], 1);
    }
    __syncthreads();
    // exclusive scan of the kCells counts in cell_start (one per thread: warp scan, then warp totals);
    // leaves the running offsets in cursor as well
    auto scan_cells = [&]() {
        int c[kCellsPerThread], sum = 0;
#pragma unroll
        for (int i = 0; i < kCellsPerThread; ++i) {
            c[i] = cell_start[tid * kCellsPerThread + i];
            sum += c[i];
        }
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int base = incl - sum;
        for (int wv = 0; wv < warp; ++wv) base += warp_tot[wv];
#pragma unroll
        for (int i = 0; i < kCellsPerThread; ++i) {
            cell_start[tid * kCellsPerThread + i] = base;
            cursor[tid * kCellsPerThread + i] = base;
            base += c[i];
        }
        if (tid == kWinThreads - 1) cell_start[kCells] = base;
        __syncthreads();
    };
    scan_cells();
    for (int i = tid; i < nt; i += kWinThreads) {
        const float2 p = __ldg(txy_g + i);
        const int cell = cell_of(p.y, orgy, invy, kGridY) * kGridX + cell_of(p.x, orgx, invx, kGridX);
        const int pos = atomicAdd(&cursor[cell], 1);
        VX_CHECK(cell >= 0 && cell < kCells && pos >= 0 && pos < nt);
        txy[pos] = p;
        tidx[pos] = i;
        tdesc[2 * pos] = __ldg(reinterpret_cast<const uint4*>(tbase + (size_t)i * 32));
        tdesc[2 * pos + 1] = __ldg(reinterpret_cast<const uint4*>(tbase + (size_t)i * 32 + 16));
    }
    __syncthreads();

    // ---- 3. queries in cell order ----------------------------------------------------------------
    // The lanes of a warp then look at the same or adjacent cells: equal trip counts (the warp runs
    // max-over-lanes iterations) and broadcast reads of the sorted train rows.  Counting sort of the
    // query rows by the cell of their position in the train grid; cell_start is saved first.
    const float2* qxy_g = reinterpret_cast<const float2*>(a.q_xy + (size_t)pair * a.q_pair_stride * 2);
    int my_start[kCellsPerThread + 1];
#pragma unroll
    for (int i = 0; i < kCellsPerThread; ++i) my_start[i] = cell_start[tid * kCellsPerThread + i];
    my_start[kCellsPerThread] = tid == 0 ? cell_start[kCells] : 0;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kCellsPerThread; ++i) cell_start[tid * kCellsPerThread + i] = 0;
    __syncthreads();
    for (int i = tid; i < nq; i += kWinThreads) {
        const float2 p = __ldg(qxy_g + i);
        atomicAdd(&cell_start[cell_of(p.y, orgy, invy, kGridY) * kGridX + cell_of(p.x, orgx, invx, kGridX)], 1);
    }
    __syncthreads();
    scan_cells();
    // gridDim.y CTAs share the queries of a pair (each bins the train rows itself: that part is a fraction of a
    // CTA's time, and a pair with many more than kWinThreads queries would otherwise need several sequential rounds).
    // The order of the queries INSIDE a cell depends on the order of the atomics and differs from CTA to CTA, so the
    // split is made on whole cells: CTA s owns the cells whose first rank lies in [s, s + 1) * per_split — a range
    // of ranks [rank_lo, rank_hi) that every CTA computes identically.
    const int per_split = (nq + (int)gridDim.y - 1) / (int)gridDim.y;
    const int own_lo = (int)blockIdx.y * per_split, own_hi = own_lo + per_split;
    const bool split = gridDim.y > 1;
    if (tid == 0) {
        s_rank_lo = split ? nq : 0;
        s_rank_hi = nq;
    }
    __syncthreads();
    for (int i = tid; i < nq; i += kWinThreads) {
        const float2 p = __ldg(qxy_g + i);
        const int cell = cell_of(p.y, orgy, invy, kGridY) * kGridX + cell_of(p.x, orgx, invx, kGridX);
        const int first = cell_start[cell];  // first rank of the cell (the scan of the query histogram)
        if (split && first >= own_lo && first < s_rank_lo) atomicMin(&s_rank_lo, first);
        if (split && first >= own_hi && first < s_rank_hi) atomicMin(&s_rank_hi, first);
        qorder[atomicAdd(&cursor[cell], 1)] = (uint16_t)i;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kCellsPerThread; ++i) cell_start[tid * kCellsPerThread + i] = my_start[i];  // back to the train offsets
    if (tid == 0) cell_start[kCells] = my_start[kCellsPerThread];
    __syncthreads();

    const int rank_end = s_rank_hi;
    for (int rank = s_rank_lo + tid; rank < rank_end; rank += kWinThreads) {
        const int qi = qorder[rank];
        uint32_t q[8];
        {
            const uint4 lo = __ldg(reinterpret_cast<const uint4*>(qbase + (size_t)qi * 32));
            const uint4 hi = __ldg(reinterpret_cast<const uint4*>(qbase + (size_t)qi * 32 + 16));
            q[0] = lo.x; q[1] = lo.y; q[2] = lo.z; q[3] = lo.w;
            q[4] = hi.x; q[5] = hi.y; q[6] = hi.z; q[7] = hi.w;
        }
        const float2 qp = __ldg(qxy_g + qi);
        const float qx = qp.x, qy = qp.y;
        // train rows that can pass have tx in [qx - bx1, qx - bx0] (ty alike); the cell range is taken from a slightly
        // widened interval (covers the rounding of qx - tx and of qx - bound)
        const float ew = (fmaxf(fabsf(bx0), fabsf(bx1)) + fmaxf(fabsf(by0), fabsf(by1))) * 1.0e-4f +
                         fmaxf(fabsf(qx), fabsf(qy)) * 1.0e-6f + 1.0e-6f;
        const int cx0 = cell_of(qx - bx1 - ew, orgx, invx, kGridX), cx1 = cell_of(qx - bx0 + ew, orgx, invx, kGridX);
        const int cy0 = cell_of(qy - by1 - ew, orgy, invy, kGridY), cy1 = cell_of(qy - by0 + ew, orgy, invy, kGridY);
        uint32_t k0 = kNoKey, k1 = kNoKey;
        for (int cy = cy0; cy <= cy1; ++cy) {
            const int j0 = cell_start[cy * kGridX + cx0], j1 = cell_start[cy * kGridX + cx1 + 1];
            VX_CHECK(j0 >= 0 && j1 <= nt && cy >= 0 && cy < kGridY && cx0 >= 0 && cx1 < kGridX);
            for (int j = j0; j < j1; ++j) {
                const float2 p = txy[j];
                const float dx = __fsub_rn(qx, p.x), dy = __fsub_rn(qy, p.y);
                if (dx >= bx0 && dx <= bx1 && dy >= by0 && dy <= by1) {
                    const uint4 lo = tdesc[2 * j], hi = tdesc[2 * j + 1];
                    const int d = __popc(q[0] ^ lo.x) + __popc(q[1] ^ lo.y) + __popc(q[2] ^ lo.z) + __popc(q[3] ^ lo.w) +
                                  __popc(q[4] ^ hi.x) + __popc(q[5] ^ hi.y) + __popc(q[6] ^ hi.z) + __popc(q[7] ^ hi.w);
                    const uint32_t key = ((uint32_t)d << 20) | (uint32_t)tidx[j];
                    k1 = min(k1, max(k0, key));
                    k0 = min(k0, key);
                }
            }
        }
        reinterpret_cast<int4*>(a.knn + (size_t)pair * a.knn_stride)[qi] =
            make_int4(k0 == kNoKey ? -1 : (int)(k0 & 0xfffffu), k1 == kNoKey ? -1 : (int)(k1 & 0xfffffu),
                      k0 == kNoKey ? 0 : (int)(k0 >> 20), k1 == kNoKey ? 0 : (int)(k1 >> 20));
    }
}

// One CTA per pair.  Walks the queries in ascending order in chunks of blockDim and appends the
// accepted ones with an ordered block scan, so matches come out sorted by query index exactly as
// the reference's loop over knn_matches (orb_matcher.cpp:27-36) produces them.
__global__ void __launch_bounds__(256)
match_compact_kernel(const vxorb_knn2_t* __restrict__ knn, const vxorb_knn2_t* __restrict__ back, int knn_stride,
                     int back_stride, const int* __restrict__ q_n, int n_stride, float nn_ratio, int mode,
                     vxorb_match_t* __restrict__ out, int out_stride, int* __restrict__ out_n) {
    __shared__ int warp_sum[8];
    __shared__ int s_base, s_min;
    const int pair = blockIdx.x;
    const int nq = q_n[pair * n_stride];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const vxorb_knn2_t* kp = knn + (size_t)pair * knn_stride;
    const vxorb_knn2_t* bp = back ? back + (size_t)pair * back_stride : nullptr;
    vxorb_match_t* op = out + (size_t)pair * out_stride;

    auto accept = [&](int i, int& tidx, int& dist) -> bool {
        const int4 r = reinterpret_cast<const int4*>(kp)[i];  // idx0, idx1, dist0, dist1
        tidx = r.x;
        dist = r.z;
        if (mode == 2) return r.x >= 0 && bp[r.x].idx[0] == i;
        if (r.y < 0) return false;  // fewer than two neighbours (orb_matcher.cpp:28)
        return (float)r.z < __fmul_rn(nn_ratio, (float)r.w);  // m1.distance < nn_ratio * m2.distance
    };

    float limit = 3.0e38f;
    if (mode == 1) {
        // caller's post-filter (tracking.cpp:211-222): min_dist starts at 100
        if (tid == 0) s_min = 100;
        __syncthreads();
        int m = 100;
        for (int i = tid; i < nq; i += 256) {
            int t, d;
            if (accept(i, t, d)) m = min(m, d);
        }
        m = __reduce_min_sync(0xffffffffu, m);
        if (lane == 0) atomicMin(&s_min, m);
        __syncthreads();
        limit = fmaxf(__fmul_rn(2.0f, (float)s_min), 30.0f);
    }

    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int i0 = 0; i0 < nq; i0 += 256) {
        const int i = i0 + tid;
        int tidx = -1, dist = 0;
        bool ok = i < nq && accept(i, tidx, dist);
        if (ok && mode == 1) ok = (float)dist <= limit;
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) warp_sum[warp] = __popc(m);
        __syncthreads();
        int before = s_base;
        for (int wv = 0; wv < warp; ++wv) before += warp_sum[wv];
        if (ok) {
            const int o = before + __popc(m & ((1u << lane) - 1u));
            if (o < out_stride) {
                op[o].query_idx = i;
                op[o].train_idx = tidx;
                op[o].distance = dist;
            }
        }
        __syncthreads();
        if (tid == 0) {
            int tot = 0;
            for (int wv = 0; wv < 8; ++wv) tot += warp_sum[wv];
            s_base += tot;
        }
        __syncthreads();
    }
    if (tid == 0) out_n[pair] = s_base;
}

// Integer-pipe calibration: every thread runs `iters` rounds of 8 independent XOR+POPC+IADD chains
// (the inner loop of the knn kernel without its memory traffic).  Used by bench.py to measure the
// popcount peak the match stage is normalised against.
__global__ void __launch_bounds__(256) popc_peak_kernel(uint32_t* out, int iters, uint32_t seed) {
    uint32_t a[8], acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        a[k] = seed * (threadIdx.x + 1u) + 0x9e3779b9u * (k + 1u) + blockIdx.x;
        acc[k] = 0;
    }
    uint32_t x = seed ^ threadIdx.x;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += __popc(a[k] ^ x);
        x = x * 1664525u + 1013904223u;
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) r += acc[k];
    if (r == 0xffffffffu) out[0] = r;  // keeps the chain alive; practically never taken
}

int launch_popc_peak(uint32_t* out, int blocks, int iters, cudaStream_t s) {
    popc_peak_kernel<<<blocks, 256, 0, s>>>(out, iters, 12345u);
    return 1;
}

// Opts the windowed kernel into large dynamic shared memory on the current device (vxorb_create).
cudaError_t init_match_kernels() {
    return cudaFuncSetAttribute(knn2_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
}

cudaError_t carveout_match(int pct) {
    cudaError_t e;
    if ((e = set_carveout(knn2_kernel<1>, pct)) != cudaSuccess) return e;
    if ((e = set_carveout(knn2_merge_kernel, pct)) != cudaSuccess) return e;
    if ((e = set_carveout(knn2_window_kernel, pct)) != cudaSuccess) return e;
    return set_carveout(match_compact_kernel, pct);
}

int launch_knn2(const MatchArgs& a, cudaStream_t s) {
    if (a.pairs <= 0 || a.max_q <= 0) return 0;
    const bool windowed = a.mask == nullptr && (a.radius > 0.0f || a.boxed) && a.q_xy != nullptr && a.t_xy != nullptr;
    // a lone box-masked pair is served faster by the split brute-force launch below (32 CTAs instead of one)
    const bool cells = windowed && !(a.boxed && a.pairs < 4);
    if (cells && a.max_t > 0 && a.max_q <= 65536 && window_smem_bytes(a.max_t, a.max_q) <= 224 * 1024) {
        // cell-binned path: the whole train set of a pair sits in shared memory in cell order; one CTA per pair and
        // kWinThreads queries
#ifndef VXORB_WIN_QSPLIT_MAX
#define VXORB_WIN_QSPLIT_MAX 8
#endif
        const int qsplit = std::min(VXORB_WIN_QSPLIT_MAX, (a.max_q + 1535) / 1536);  // 1000 queries: one CTA is faster (0.049 vs 0.066 ms per 128 pairs); 2000: three CTAs 0.20 vs 0.35 ms per 96 pairs
        dim3 grid(a.pairs, qsplit);
        knn2_window_kernel<<<grid, kWinThreads, window_smem_bytes(a.max_t, a.max_q), s>>>(a, a.max_t, a.max_q);
        return 1;
    }
    // few pairs: split the train rows over gridDim.z so that the launch fills the GPU (towards sixteen CTAs of
    // 4 warps per SM), then merge the partial top-2 keys
    MatchArgs b = a;
    const bool masked = a.mask != nullptr || ((a.radius > 0.0f || a.boxed) && a.q_xy != nullptr && a.t_xy != nullptr);
    // One query per thread everywhere.  Two per thread (one pair of broadcast loads and one loop step for two distances,
    // 72 registers) were in use for sets of 2048 queries and more; measured again with the split target below they lose
    // on every shape but one: 4000 x 4000 0.0437 -> 0.0385 ms, 2500 x 16384 0.106 -> 0.077, 4000 x 16384 0.135 -> 0.101,
    // 8000 x 8000 0.135 -> 0.112 (0.108 with two), 32 pairs of 4000 x 4000 0.0341 -> 0.0249 ms per pair.
#ifndef VXORB_KNN_QPT
#define VXORB_KNN_QPT 1
#endif
    const int qpt = masked ? 1 : VXORB_KNN_QPT;
    const int qblocks = (a.max_q + kQPerBlock * qpt - 1) / (kQPerBlock * qpt);
    const int tiles = (a.max_t + kTrainTile - 1) / kTrainTile;
    int splits = 1;
#ifndef VXORB_KNN_CTA_TARGET
#define VXORB_KNN_CTA_TARGET (16 * 148)  // a single pair is split towards 16 CTAs of 4 warps per SM (4000 x 16384: 2 per SM 0.154 ms, 4: 0.118, 8 and 16: 0.101; 8000 x 8000: 0.124 / 0.118 / 0.112 / 0.100)
#endif
    if (a.part && a.max_t > 0) {
        splits = (VXORB_KNN_CTA_TARGET + qblocks * a.pairs - 1) / (qblocks * a.pairs);
        splits = splits < 1 ? 1 : splits;
        splits = splits > tiles ? tiles : splits;
        splits = splits > kMaxTrainSplits ? kMaxTrainSplits : splits;
        splits = splits < 1 ? 1 : splits;
    }
    b.splits = splits;
    dim3 grid(qblocks, a.pairs, splits);
#if VXORB_KNN_QPT == 2
    if (qpt == 2)
        knn2_kernel<2><<<grid, kQPerBlock, 0, s>>>(b);
    else
#endif
        knn2_kernel<1><<<grid, kQPerBlock, 0, s>>>(b);
    if (splits == 1) return 1;
    dim3 mgrid((a.max_q + 255) / 256, a.pairs);
    knn2_merge_kernel<<<mgrid, 256, 0, s>>>(b);
    return 2;
}

int launch_match_compact(const vxorb_knn2_t* knn, const vxorb_knn2_t* back, int knn_stride, int back_stride,
                         const int* q_n, int pairs, float nn_ratio, int mode, vxorb_match_t* out, int out_stride,
                         int* out_n, cudaStream_t s, int n_stride) {
    if (pairs <= 0) return 0;
    match_compact_kernel<<<pairs, 256, 0, s>>>(knn, back, knn_stride, back_stride, q_n, n_stride, nn_ratio, mode, out,
                                               out_stride, out_n);
    return 1;
}

}  // namespace vxorb
