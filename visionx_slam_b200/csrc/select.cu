// select.cu — per-level keypoint selection: retainBest(2q) on the FAST score, Harris response,
// retainBest(q) on Harris, then canonical (y, x) ordering.
//
// Replaces, inside cv::ORB::detectAndCompute (orb_extractor.cpp:13), KeyPointsFilter::retainBest
// (twice) and HarrisResponses.  There is no grid/quadtree distribution in the
// reference path: cv::ORB keeps everything >= the n-th largest response per level, ties included,
// so the kept SET is order independent and is reproduced here with histogram / radix selection.
//
//   step 1  256-bin shared histogram of FAST scores -> threshold of the (2q)-th largest
//           (bin walk done by one warp: 8 bins per lane + a shuffle suffix sum)
//   step 2  retained candidates compacted into a shared index list, then ONE THREAD per candidate
//           computes the Harris response: 27 independent word loads, separable gradients in
//           registers, float combine un-fused exactly as OpenCV's
//   step 3  4-pass 8-bit radix select on the order-preserving float key -> q-th largest response
//   step 4  survivors rank-sorted by (y, x) inside shared memory, written to the final list
// One CTA per (level, frame), level-major so the long level-0 CTAs are scheduled first.  Working set is a few KB per CTA and L2 resident; the stage is
// latency/issue bound, not HBM bound.
#include "vxorb_internal.h"

namespace vxorb {

namespace {

constexpr int kSortCap = 2048;

__device__ __forceinline__ uint32_t float_key(float v) {
    if (v == 0.0f) v = 0.0f;  // -0 -> +0
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Harris response of the 7x7 block around (x, y), computed by ONE thread: 9 rows x 3 aligned words
// are loaded up front (27 independent loads in flight), then the separable form is evaluated in
// registers: d[r][x] = p[x+1] - p[x-1], s[r][x] = p[x-1] + 2 p[x] + p[x+1] per row,
// Ix = d[y-1] + 2 d[y] + d[y+1], Iy = s[y+1] - s[y-1].  Integer sums are exact, so the grouping does
// not matter; the float combine is un-fused in OpenCV's association.
__device__ __forceinline__ float harris_thread(const uint8_t* __restrict__ img, int pitch, int x, int y) {
    VX_CHECK(x >= 4 && y >= 4);  // the 9 x 9 window of a candidate (border 31) never leaves the plane
    const uint8_t* base = img + (ptrdiff_t)(y - 4) * pitch + (x - 4);
    const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(base) & 3u) * 8u;
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(base) & ~(uintptr_t)3);
    const int pw = pitch >> 2;
    uint32_t w[9][3];
#pragma unroll
    for (int r = 0; r < 9; ++r) {
        w[r][0] = __ldg(wp + r * pw);
        w[r][1] = __ldg(wp + r * pw + 1);
        w[r][2] = __ldg(wp + r * pw + 2);
    }
    int d[3][7], s[3][7];  // rolling window of rows r-2, r-1, r (index r % 3)
    int a = 0, b = 0, c = 0;
#pragma unroll
    for (int r = 0; r < 9; ++r) {
        const uint32_t lo = __funnelshift_r(w[r][0], w[r][1], sh), mid = __funnelshift_r(w[r][1], w[r][2], sh);
        const uint32_t hi = w[r][2] >> sh;
        int p[9];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            p[k] = (int)((lo >> (8 * k)) & 0xffu);
            p[4 + k] = (int)((mid >> (8 * k)) & 0xffu);
        }
        p[8] = (int)(hi & 0xffu);
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            d[r % 3][k] = p[k + 2] - p[k];
            s[r % 3][k] = p[k] + 2 * p[k + 1] + p[k + 2];
        }
        if (r >= 2) {
            // block row centred on patch row r-1
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                const int ix = d[(r - 2) % 3][k] + 2 * d[(r - 1) % 3][k] + d[r % 3][k];
                const int iy = s[r % 3][k] - s[(r - 2) % 3][k];
                a += ix * ix;
                b += iy * iy;
                c += ix * iy;
            }
        }
    }
    const float scale = 1.0f / (4 * 7 * 255.0f);
    const float s4 = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
    const float fa = (float)a, fb = (float)b, fc = (float)c;
    const float det = __fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc));
    const float tr = __fadd_rn(fa, fb);
    const float rr = __fsub_rn(det, __fmul_rn(__fmul_rn(0.04f, tr), tr));
    return __fmul_rn(rr, s4);
}

// Warp-parallel replacement of the serial "walk the 256-bin histogram from the top until k entries
// are covered" loop.  Called by all 32 lanes of ONE warp.  Returns (via out params, valid in every
// lane) the largest bin d >= 1 with sum_{b >= d} hist[b] >= k, and above = sum_{b > d} hist[b]; if
// no such bin exists d = 0 and above = sum_{b >= 1} hist[b] (what the serial loop leaves behind).
__device__ __forceinline__ void top_bin_scan(const int* hist, int k, int lane, int& d_out, int& above_out) {
    int hb[8], tot = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        hb[j] = hist[8 * lane + j];
        tot += hb[j];
    }
    int suf = tot;  // becomes sum over lanes >= lane
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_down_sync(0xffffffffu, suf, o);
        if (lane + o < 32) suf += n;
    }
    int cum = suf - tot;  // bins of higher lanes
    const bool has = cum < k && suf >= k;
    int d = 0, above = 0;
    if (has) {
        d = -1;
#pragma unroll
        for (int j = 7; j >= 0; --j) {
            if (d < 0) {
                if (cum + hb[j] >= k) {
                    d = 8 * lane + j;
                    above = cum;
                } else {
                    cum += hb[j];
                }
            }
        }
        if (d == 0) {  // the serial loop never stops on bin 0
            above = cum;
        }
    }
    const unsigned who = __ballot_sync(0xffffffffu, has);
    if (who == 0) {
        const int all = __shfl_sync(0xffffffffu, suf, 0);
        d_out = 0;
        above_out = all - hist[0];
        return;
    }
    const int src = __ffs(who) - 1;
    d_out = __shfl_sync(0xffffffffu, d, src);
    above_out = __shfl_sync(0xffffffffu, above, src);
}

}  // namespace

#ifndef VXORB_SELECT_CTAS
#define VXORB_SELECT_CTAS 4
#endif
// kThreads = 256 (4 CTAs per SM) for batches; 1024 for a few frames, where the eight CTAs of a frame are alone on the GPU
// and the level-0 CTA's loops over ~2400 candidates are the critical path of the call.
template <int kThreads>
__global__ void __launch_bounds__(kThreads, kThreads == 256 ? VXORB_SELECT_CTAS : 1)
select_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, const uint32_t* __restrict__ cand,
              const int* __restrict__ cand_n, float* __restrict__ resp, uint32_t* __restrict__ fin_key,
              float* __restrict__ fin_resp, int* __restrict__ fin_n, int* __restrict__ status) {
    __shared__ int hist[256];
    __shared__ uint32_t skey[kSortCap];   // phase A: index of a retained candidate; phase B: (y, x) key
    __shared__ float sresp[kSortCap];
    __shared__ int s_thr1, s_ns, s_nf, s_nr;
    __shared__ uint32_t s_prefix;
    __shared__ int s_k;

    // heavy levels first: blockIdx.y = level, so all level-0 CTAs are scheduled before level 1 ...
    const int l = blockIdx.y, f = blockIdx.x;
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
    if (tid < 256) hist[tid] = 0;
    if (tid == 0) {
        s_thr1 = 0;
        s_ns = n;
        s_nf = 0;
        s_nr = 0;
    }
    __syncthreads();
    if (n > 2 * q) {
        for (int i = tid; i < n; i += kThreads) atomicAdd(&hist[c[i] >> 24], 1);
        __syncthreads();
        if (warp == 0) {
            int d, above;
            top_bin_scan(hist, 2 * q, lane, d, above);
            if (lane == 0) {
                s_thr1 = d;
                s_ns = above + hist[d];  // d >= 20 here: n > 2q and every score is >= the threshold
            }
        }
        __syncthreads();
    }
    const uint32_t thr1 = (uint32_t)s_thr1;
    const int ns = s_ns;
    const bool listed = ns <= kSortCap;  // retained candidates fit the shared index list

    // ---- step 2: Harris response of every retained candidate, one thread per candidate --------
    if (listed) {
        for (int i0 = 0; i0 < n; i0 += kThreads) {
            const int i = i0 + tid;
            const bool keep = i < n && (c[i] >> 24) >= thr1;
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            int base = 0;
            if (lane == 0 && m) base = atomicAdd(&s_nr, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            VX_CHECK(!keep || base + __popc(m & ((1u << lane) - 1u)) < kSortCap);
            if (keep) skey[base + __popc(m & ((1u << lane) - 1u))] = (uint32_t)i;
        }
        __syncthreads();
        for (int j = tid; j < ns; j += kThreads) {
            const uint32_t i = skey[j];
            const uint32_t k = c[i];
            sresp[j] = harris_thread(img, pitch, (int)(k & 0xfffu), (int)((k >> 12) & 0xfffu));
        }
    } else {
        for (int i = tid; i < n; i += kThreads) {
            const uint32_t k = c[i];
            if ((k >> 24) >= thr1) rs[i] = harris_thread(img, pitch, (int)(k & 0xfffu), (int)((k >> 12) & 0xfffu));
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
            if (tid < 256) hist[tid] = 0;
            __syncthreads();
            const uint32_t prefix = s_prefix;
            if (listed) {
                for (int j = tid; j < ns; j += kThreads) {
                    const uint32_t u = float_key(sresp[j]);
                    if ((u & mask) == prefix) atomicAdd(&hist[(u >> (8 * pass)) & 255u], 1);
                }
            } else {
                for (int i = tid; i < n; i += kThreads) {
                    if ((c[i] >> 24) >= thr1) {
                        const uint32_t u = float_key(rs[i]);
                        if ((u & mask) == prefix) atomicAdd(&hist[(u >> (8 * pass)) & 255u], 1);
                    }
                }
            }
            __syncthreads();
            if (warp == 0) {
                int d, above;
                top_bin_scan(hist, s_k, lane, d, above);
                __syncwarp();
                if (lane == 0) {
                    s_k -= above;
                    s_prefix = prefix | ((uint32_t)d << (8 * pass));
                }
            }
            mask |= 255u << (8 * pass);
            __syncthreads();
        }
        thr2 = s_prefix;
    }

    // ---- step 4: survivors in (y, x) order -> final list ----------------------------------------
    const int fcap = lt.fin_cap[l];
    uint32_t* ok = fin_key + lt.fin_off[l] + (size_t)f * fcap;
    float* orr = fin_resp + lt.fin_off[l] + (size_t)f * fcap;
    if (listed) {
        // survivors are first compacted to the front of the shared list (order does not matter), then
        // every survivor ranks its (y, x) key against the survivors only: dense lanes, nf-long loops
        uint32_t mykey[kSortCap / kThreads];
        float myresp[kSortCap / kThreads];
        int mypos[kSortCap / kThreads];
#pragma unroll
        for (int t = 0; t < kSortCap / kThreads; ++t) {
            const int j = tid + t * kThreads;
            bool keep = false;
            mykey[t] = 0;
            myresp[t] = 0.f;
            if (j < ns) {
                myresp[t] = sresp[j];
                keep = float_key(myresp[t]) >= thr2;
                if (keep) mykey[t] = c[skey[j]] & 0xffffffu;
            }
            // warp-aggregated append (uniform trip count: every lane runs all kSortCap / kThreads rounds)
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            int base = 0;
            if (lane == 0 && m) base = atomicAdd(&s_nf, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            mypos[t] = keep ? base + __popc(m & ((1u << lane) - 1u)) : -1;
        }
        __syncthreads();  // every slot has been read: the list can be overwritten
#pragma unroll
        for (int t = 0; t < kSortCap / kThreads; ++t)
            if (mypos[t] >= 0) {
                VX_CHECK(mypos[t] < kSortCap);
                skey[mypos[t]] = mykey[t];
                sresp[mypos[t]] = myresp[t];
            }
        __syncthreads();
        const int nf = s_nf;
        for (int i = tid; i < nf; i += kThreads) {
            const uint32_t k = skey[i];
            int rank = 0;
#pragma unroll 8
            for (int jj = 0; jj < nf; ++jj) rank += skey[jj] < k ? 1 : 0;
            if (rank < fcap) {
                ok[rank] = k;
                orr[rank] = sresp[i];
            }
        }
    } else {
        // Degenerate images (massive ties): rank against the candidate list in global memory.
        int cnt = 0;
        for (int i = tid; i < n; i += kThreads) {
            const uint32_t k = c[i];
            if ((k >> 24) < thr1) continue;
            const float r = rs[i];
            if (float_key(r) < thr2) continue;
            ++cnt;
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
        if (cnt) atomicAdd(&s_nf, cnt);
    }
    __syncthreads();
    if (tid == 0) {
        const int nf = s_nf;
        *out_n = nf;  // true count; consumers clamp to fin_cap
        if (nf > fcap) atomicOr(status, 2);
    }
}

cudaError_t carveout_select(int pct) {
    const cudaError_t e = set_carveout(select_kernel<256>, pct);
    return e != cudaSuccess ? e : set_carveout(select_kernel<1024>, pct);
}

int launch_select(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s) {
    dim3 grid(lt.batch, lt.n_levels);
    if (lt.batch < 8) {
        select_kernel<1024><<<grid, 1024, 0, s>>>(lt, db.pyr, db.cand, db.cand_n, db.resp, db.fin_key, db.fin_resp, db.fin_n,
                                                   db.status);
        return 1;
    }
    select_kernel<256><<<grid, 256, 0, s>>>(lt, db.pyr, db.cand, db.cand_n, db.resp, db.fin_key, db.fin_resp,
                                           db.fin_n, db.status);
    return 1;
}

}  // namespace vxorb
