// select.cu — per-level keypoint selection: retainBest(2q) on the FAST score, Harris response,
// retainBest(q) on Harris, then canonical (y, x) ordering.
//
// Replaces, inside cv::ORB::detectAndCompute (orb_extractor.cpp:13), KeyPointsFilter::retainBest
// (twice) and HarrisResponses.  There is no grid/quadtree distribution in the
// reference path: cv::ORB keeps everything >= the n-th largest response per level, ties included,
// so the kept SET is order independent and is reproduced here with histogram / radix selection.
//
//   step 1  256-bin shared histogram of FAST scores -> threshold of the (2q)-th largest
//   step 2  Harris response, one warp per retained candidate (9x9 window staged in shared memory,
//           integer sums reduced with REDUX, float combine un-fused exactly as OpenCV's)
//   step 3  4-pass 8-bit radix select on the order-preserving float key -> q-th largest response
//   step 4  survivors staged in shared memory, rank-sorted by (y, x), written to the final list
// One CTA per (level, frame).  Working set is a few KB per CTA and L2 resident; the stage is
// latency/issue bound, not HBM bound.
#include "vxorb_internal.h"

namespace vxorb {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kSortCap = 2048;

__device__ __forceinline__ uint32_t float_key(float v) {
    if (v == 0.0f) v = 0.0f;  // -0 -> +0
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Harris response of the 7x7 block around (x, y); all 32 lanes of the warp take part.
// `patch` is this warp's 9 x 12 byte staging area.
__device__ __forceinline__ float harris_warp(const uint8_t* __restrict__ img, int pitch, int x, int y,
                                             uint8_t (*patch)[12], int lane) {
    __syncwarp();
    for (int i = lane; i < 81; i += 32) {
        const int r = i / 9, c = i - r * 9;
        patch[r][c] = __ldg(img + (size_t)(y - 4 + r) * pitch + (x - 4 + c));
    }
    __syncwarp();
    int a = 0, b = 0, c = 0;
#pragma unroll
    for (int rnd = 0; rnd < 2; ++rnd) {
        const int k = lane + rnd * 32;
        if (k < 49) {
            const int by = k / 7, bx = k - by * 7;  // block pixel, patch coords (bx+1, by+1)
            const uint8_t* p = &patch[by + 1][bx + 1];
            const int p00 = p[-12 - 1], p01 = p[-12], p02 = p[-12 + 1];
            const int p10 = p[-1], p12 = p[1];
            const int p20 = p[12 - 1], p21 = p[12], p22 = p[12 + 1];
            const int ix = (p12 - p10) * 2 + (p02 - p00) + (p22 - p20);
            const int iy = (p21 - p01) * 2 + (p20 - p00) + (p22 - p02);
            a += ix * ix;
            b += iy * iy;
            c += ix * iy;
        }
    }
    a = __reduce_add_sync(0xffffffffu, a);
    b = __reduce_add_sync(0xffffffffu, b);
    c = __reduce_add_sync(0xffffffffu, c);
    const float scale = 1.0f / (4 * 7 * 255.0f);
    const float s4 = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
    const float fa = (float)a, fb = (float)b, fc = (float)c;
    const float det = __fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc));
    const float tr = __fadd_rn(fa, fb);
    const float r = __fsub_rn(det, __fmul_rn(__fmul_rn(0.04f, tr), tr));
    return __fmul_rn(r, s4);
}

}  // namespace

__global__ void __launch_bounds__(kThreads)
select_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, const uint32_t* __restrict__ cand,
              const int* __restrict__ cand_n, float* __restrict__ resp, uint32_t* __restrict__ fin_key,
              float* __restrict__ fin_resp, int* __restrict__ fin_n, int* __restrict__ status) {
    __shared__ int hist[256];
    __shared__ uint8_t patch[kWarps][9][12];
    __shared__ uint32_t skey[kSortCap];
    __shared__ float sresp[kSortCap];
    __shared__ int s_thr1, s_ns, s_nf;
    __shared__ uint32_t s_prefix;
    __shared__ int s_k;

    const int l = blockIdx.x, f = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int* out_n = fin_n + l * lt.batch + f;
    if (!lt.active[l]) {
        if (tid == 0) *out_n = 0;
        return;
    }
    const int cap = lt.cand_cap[l];
    int n = cand_n[l * lt.batch + f];
    if (n > cap) {
        if (tid == 0) atomicOr(status, 1);
        n = cap;
    }
    const uint32_t* c = cand + lt.cand_off[l] + (size_t)f * cap;
    float* rs = resp + lt.cand_off[l] + (size_t)f * cap;
    const int q = lt.quota[l];
    const int pitch = lt.pitch[l];
    const uint8_t* img = pyr + lt.off[l] + (size_t)f * lt.fstride[l];

    // ---- step 1: retainBest(2q) on the FAST score ---------------------------------------------
    hist[tid] = 0;
    if (tid == 0) {
        s_thr1 = 0;
        s_ns = n;
        s_nf = 0;
    }
    __syncthreads();
    if (n > 2 * q) {
        for (int i = tid; i < n; i += kThreads) atomicAdd(&hist[c[i] >> 24], 1);
        __syncthreads();
        if (tid == 0) {
            int cum = 0, s = 255;
            for (; s > 0; --s) {
                cum += hist[s];
                if (cum >= 2 * q) break;
            }
            s_thr1 = s;
            s_ns = cum;
        }
        __syncthreads();
    }
    const uint32_t thr1 = (uint32_t)s_thr1;
    const int ns = s_ns;

    // ---- step 2: Harris response of every retained candidate ----------------------------------
    for (int i = warp; i < n; i += kWarps) {
        const uint32_t k = c[i];
        if ((k >> 24) >= thr1) {
            const float r = harris_warp(img, pitch, (int)(k & 0xfffu), (int)((k >> 12) & 0xfffu), patch[warp], lane);
            if (lane == 0) rs[i] = r;
        }
    }
    __syncthreads();

    // ---- step 3: retainBest(q) on the Harris response (radix select of the q-th largest) ------
    uint32_t thr2 = 0;  // key threshold; 0 keeps everything
    if (ns > q) {
        if (tid == 0) {
            s_prefix = 0;
            s_k = q;
        }
        uint32_t mask = 0;
        for (int pass = 3; pass >= 0; --pass) {
            hist[tid] = 0;
            __syncthreads();
            const uint32_t prefix = s_prefix;
            for (int i = tid; i < n; i += kThreads) {
                if ((c[i] >> 24) >= thr1) {
                    const uint32_t u = float_key(rs[i]);
                    if ((u & mask) == prefix) atomicAdd(&hist[(u >> (8 * pass)) & 255u], 1);
                }
            }
            __syncthreads();
            if (tid == 0) {
                int k = s_k, cum = 0, d = 255;
                for (; d > 0; --d) {
                    if (cum + hist[d] >= k) break;
                    cum += hist[d];
                }
                s_k = k - cum;
                s_prefix = prefix | ((uint32_t)d << (8 * pass));
            }
            mask |= 255u << (8 * pass);
            __syncthreads();
        }
        thr2 = s_prefix;
    }

    // ---- step 4: gather survivors, order by (y, x), write the final list ----------------------
    for (int i = tid; i < n; i += kThreads) {
        const uint32_t k = c[i];
        if ((k >> 24) >= thr1) {
            const float r = rs[i];
            if (float_key(r) >= thr2) {
                const int o = atomicAdd(&s_nf, 1);
                if (o < kSortCap) {
                    skey[o] = k & 0xffffffu;
                    sresp[o] = r;
                }
            }
        }
    }
    __syncthreads();
    const int nf = s_nf;
    const int fcap = lt.fin_cap[l];
    uint32_t* ok = fin_key + lt.fin_off[l] + (size_t)f * fcap;
    float* orr = fin_resp + lt.fin_off[l] + (size_t)f * fcap;
    if (nf <= kSortCap) {
        for (int i = tid; i < nf; i += kThreads) {
            const uint32_t k = skey[i];
            int rank = 0;
            for (int j = 0; j < nf; ++j) rank += skey[j] < k ? 1 : 0;
            if (rank < fcap) {
                ok[rank] = k;
                orr[rank] = sresp[i];
            }
        }
    } else {
        // Degenerate images (massive ties): rank against the candidate list in global memory.
        for (int i = tid; i < n; i += kThreads) {
            const uint32_t k = c[i];
            if ((k >> 24) < thr1) continue;
            const float r = rs[i];
            if (float_key(r) < thr2) continue;
            const uint32_t key = k & 0xffffffu;
            int rank = 0;
            for (int j = 0; j < n; ++j) {
                const uint32_t kj = c[j];
                if ((kj >> 24) >= thr1 && float_key(rs[j]) >= thr2 && (kj & 0xffffffu) < key) ++rank;
            }
            if (rank < fcap) {
                ok[rank] = key;
                orr[rank] = r;
            }
        }
    }
    if (tid == 0) {
        *out_n = nf;  // true count; consumers clamp to fin_cap
        if (nf > fcap) atomicOr(status, 2);
    }
}

int launch_select(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s) {
    dim3 grid(lt.n_levels, lt.batch);
    select_kernel<<<grid, kThreads, 0, s>>>(lt, db.pyr, db.cand, db.cand_n, db.resp, db.fin_key, db.fin_resp,
                                           db.fin_n, db.status);
    return 1;
}

}  // namespace vxorb
