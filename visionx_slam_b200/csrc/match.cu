// match.cu — Hamming 2-NN (brute force or square-window masked), ratio test, cross-check and the
// caller's post-filter, for a batch of independent (query, train) descriptor sets.
//
// Replaces cv::BFMatcher(cv::NORM_HAMMING).knnMatch(desc_last, desc_curr, knn, 2)
// (orb_matcher.cpp:22-25) and the ratio loop (orb_matcher.cpp:27-36); SURVEY.md A.9.  For each query
// row the two smallest (popcount(q ^ t), train index) keys are kept lexicographically, so ties go
// to the lowest train index exactly like OpenCV's batchDistance.  Integer work, bit-exact.
//
// knn kernel: one thread owns one query descriptor (8 x u32 in registers); the CTA streams the
// train set through shared memory in 256-row tiles which every thread reads as broadcast 16-byte
// vectors; per pair: 8 XOR + 8 POPC + 7 IADD, then the packed key (dist << 20 | idx) goes through a
// 3-instruction min/max top-2 update.  Bound: integer/popcount issue rate, not HBM (64 KB of
// descriptors per 1000x1000 pair).  compact kernel: one CTA per pair, ordered block scan.
#include "vxorb_internal.h"

namespace vxorb {

namespace {

constexpr int kQPerBlock = 128;
constexpr int kTrainTile = 256;
constexpr uint32_t kNoKey = 0x7fffffffu;

}  // namespace

__global__ void __launch_bounds__(kQPerBlock)
knn2_kernel(const MatchArgs a) {
    __shared__ __align__(16) uint32_t tdesc[kTrainTile][8];
    __shared__ float2 txy[kTrainTile];

    const int pair = blockIdx.y;
    const int nq = a.q_n[pair], nt = a.t_n[pair];
    const int qi = blockIdx.x * kQPerBlock + threadIdx.x;
    if (blockIdx.x * kQPerBlock >= nq) return;  // whole CTA idle
    const bool live = qi < nq;
    const bool windowed = a.radius > 0.0f && a.q_xy != nullptr && a.t_xy != nullptr;

    const uint8_t* qbase = a.q_desc + (size_t)pair * a.q_pair_stride * 32;
    const uint8_t* tbase = a.t_desc + (size_t)pair * a.t_pair_stride * 32;
    uint32_t q[8];
    float qx = 0.f, qy = 0.f;
    if (live) {
        const uint4 lo = __ldg(reinterpret_cast<const uint4*>(qbase + (size_t)qi * 32));
        const uint4 hi = __ldg(reinterpret_cast<const uint4*>(qbase + (size_t)qi * 32 + 16));
        q[0] = lo.x; q[1] = lo.y; q[2] = lo.z; q[3] = lo.w;
        q[4] = hi.x; q[5] = hi.y; q[6] = hi.z; q[7] = hi.w;
        if (windowed) {
            const float2 p = __ldg(reinterpret_cast<const float2*>(a.q_xy + ((size_t)pair * a.q_pair_stride + qi) * 2));
            qx = p.x;
            qy = p.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) q[k] = 0;
    }

    uint32_t k0 = kNoKey, k1 = kNoKey;
    for (int t0 = 0; t0 < nt; t0 += kTrainTile) {
        const int tn = min(kTrainTile, nt - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < tn * 2; i += kQPerBlock)
            reinterpret_cast<uint4*>(&tdesc[0][0])[i] =
                __ldg(reinterpret_cast<const uint4*>(tbase + (size_t)t0 * 32) + i);
        if (windowed)
            for (int i = threadIdx.x; i < tn; i += kQPerBlock)
                txy[i] = __ldg(reinterpret_cast<const float2*>(a.t_xy + ((size_t)pair * a.t_pair_stride + t0) * 2) + i);
        __syncthreads();
        if (!windowed) {
#pragma unroll 4
            for (int j = 0; j < tn; ++j) {
                const uint4 lo = *reinterpret_cast<const uint4*>(&tdesc[j][0]);
                const uint4 hi = *reinterpret_cast<const uint4*>(&tdesc[j][4]);
                const int d = __popc(q[0] ^ lo.x) + __popc(q[1] ^ lo.y) + __popc(q[2] ^ lo.z) + __popc(q[3] ^ lo.w) +
                              __popc(q[4] ^ hi.x) + __popc(q[5] ^ hi.y) + __popc(q[6] ^ hi.z) + __popc(q[7] ^ hi.w);
                const uint32_t key = ((uint32_t)d << 20) | (uint32_t)(t0 + j);
                k1 = min(k1, max(k0, key));
                k0 = min(k0, key);
            }
        } else {
            for (int j = 0; j < tn; ++j) {
                const float2 p = txy[j];
                const bool in = fabsf(__fsub_rn(qx, p.x)) <= a.radius && fabsf(__fsub_rn(qy, p.y)) <= a.radius;
                if (__any_sync(0xffffffffu, in)) {
                    const uint4 lo = *reinterpret_cast<const uint4*>(&tdesc[j][0]);
                    const uint4 hi = *reinterpret_cast<const uint4*>(&tdesc[j][4]);
                    const int d = __popc(q[0] ^ lo.x) + __popc(q[1] ^ lo.y) + __popc(q[2] ^ lo.z) + __popc(q[3] ^ lo.w) +
                                  __popc(q[4] ^ hi.x) + __popc(q[5] ^ hi.y) + __popc(q[6] ^ hi.z) + __popc(q[7] ^ hi.w);
                    const uint32_t key = in ? (((uint32_t)d << 20) | (uint32_t)(t0 + j)) : kNoKey;
                    k1 = min(k1, max(k0, key));
                    k0 = min(k0, key);
                }
            }
        }
    }
    if (live) {
        vxorb_knn2_t r;
        r.idx[0] = k0 == kNoKey ? -1 : (int)(k0 & 0xfffffu);
        r.idx[1] = k1 == kNoKey ? -1 : (int)(k1 & 0xfffffu);
        r.dist[0] = k0 == kNoKey ? 0 : (int)(k0 >> 20);
        r.dist[1] = k1 == kNoKey ? 0 : (int)(k1 >> 20);
        reinterpret_cast<int4*>(a.knn + (size_t)pair * a.knn_stride)[qi] = make_int4(r.idx[0], r.idx[1], r.dist[0], r.dist[1]);
    }
}

// One CTA per pair.  Walks the queries in ascending order in chunks of blockDim and appends the
// accepted ones with an ordered block scan, so matches come out sorted by query index exactly as
// the reference's loop over knn_matches (orb_matcher.cpp:27-36) produces them.
__global__ void __launch_bounds__(256)
match_compact_kernel(const vxorb_knn2_t* __restrict__ knn, const vxorb_knn2_t* __restrict__ back, int knn_stride,
                     int back_stride, const int* __restrict__ q_n, float nn_ratio, int mode,
                     vxorb_match_t* __restrict__ out, int out_stride, int* __restrict__ out_n) {
    __shared__ int warp_sum[8];
    __shared__ int s_base, s_min;
    const int pair = blockIdx.x;
    const int nq = q_n[pair];
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

int launch_knn2(const MatchArgs& a, cudaStream_t s) {
    if (a.pairs <= 0 || a.max_q <= 0) return 0;
    dim3 grid((a.max_q + kQPerBlock - 1) / kQPerBlock, a.pairs);
    knn2_kernel<<<grid, kQPerBlock, 0, s>>>(a);
    return 1;
}

int launch_match_compact(const vxorb_knn2_t* knn, const vxorb_knn2_t* back, int knn_stride, int back_stride,
                         const int* q_n, int pairs, float nn_ratio, int mode, vxorb_match_t* out, int out_stride,
                         int* out_n, cudaStream_t s) {
    if (pairs <= 0) return 0;
    match_compact_kernel<<<pairs, 256, 0, s>>>(knn, back, knn_stride, back_stride, q_n, nn_ratio, mode, out,
                                               out_stride, out_n);
    return 1;
}

}  // namespace vxorb
