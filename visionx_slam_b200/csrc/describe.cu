// describe.cu — intensity-centroid orientation, keypoint records and rotated-BRIEF descriptors.
//
// Replaces ICAngles + cv::fastAtan2, the pt *= scale step and computeOrbDescriptors (WTA_K = 2)
// inside cv::ORB::detectAndCompute (orb_extractor.cpp:13); SURVEY.md A.6, A.8.  Output rows are
// what ORBExtractor::Extract copies into Frame::Features() / Frame::Descriptors()
// (orb_extractor.cpp:19-26), in canonical (octave, y, x) order.
//
// One warp owns a group of G consecutive keypoints of one frame (G = 16, or 4 for tiny batches).
// Both per-keypoint windows are staged in shared memory: every lane copies 16-byte chunks of the
// window rows (row starts rounded down to 16 B) with cp.async, one commit group per keypoint,
// 4 (orientation) / 3 (descriptor) windows deep per warp, so the windows of the next keypoints are
// in flight while keypoint i is processed.  (TMA boxes were measured first: at 31-37 rows of 48-64 B
// per box the TMA unit's per-row rate, ~5 cycles/row/SM, became the limiter.)
//   orientation  31 rows x 48 B of the un-blurred level.  lane = patch row v; three LDS.128, a
//                funnel shift to the patch origin, then IDP.4A does the work:
//                m10 += dp4a(pixels, u-weights), rowsum += dp4a(pixels, mask), m01 += v*rowsum;
//                two REDUX adds per keypoint
//   angle        all lanes at once, lane i for keypoint i: fastAtan2 polynomial with un-fused
//                float ops (degrees), then cos/sin in double — once per group
//   descriptor   37 rows x 64 B of the blurred level (pattern radius 18.4 -> |offset| <= 18).
//                lane i produces byte i: 8 comparisons of blurred pixels at pattern points rotated
//                by the angle (float un-fused rotate, round-half-even via the 1.5*2^23 trick so the
//                rounded coordinates feed the shared-memory address IMAD directly); the 16 byte
//                gathers per lane are LDS.U8 instead of scattered global loads
// Float rules are load-bearing: every op is an explicit *_rn intrinsic so nvcc cannot contract.
#include <float.h>

#include <type_traits>

#include "vxorb_internal.h"

namespace vxorb {

namespace {

// pattern pairs transposed for conflict-free access: entry [k][i] = comparison 8*i + k of byte i,
// packed (x1, y1, x2, y2) as int8
struct PatternT {
    char4 p[8][32];
};

const int8_t kPatternHost[1024] = {
#include "../../include/vxorb_pattern.inc"
};

__device__ PatternT g_pattern;
// umax[0..15] = 15,15,15,15,14,14,14,13,13,12,11,10,9,8,6,3 packed one nibble per entry.  The
// disc is symmetric (umax[umax[v]] >= v), so column u spans rows |v| <= umax[|u|].
__device__ __forceinline__ int umax_of(int a) { return (int)((0x3689ABCDDEEEFFFFull >> (4 * a)) & 15ull); }

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float rad2deg = (float)(180.0 / 3.141592653589793238462643383279502884);
    const float p1 = __fmul_rn(0.9997878412794807f, rad2deg);
    const float p3 = __fmul_rn(-0.3258083974640975f, rad2deg);
    const float p5 = __fmul_rn(0.1555786518463281f, rad2deg);
    const float p7 = __fmul_rn(-0.04432655554792128f, rad2deg);
    const float eps = (float)DBL_EPSILON;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.0f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0.0f) a = __fsub_rn(180.0f, a);
    if (y < 0.0f) a = __fsub_rn(360.0f, a);
    return a;
}

}  // namespace

__device__ __forceinline__ int dp4a_us(uint32_t pixels, uint32_t weights_s8, int acc) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pixels), "r"(weights_s8), "r"(acc));
    return d;
}

namespace {

constexpr int kPaW = 48, kPaH = 31, kPaBytes = kPaW * kPaH, kPaStride = 1536;  // orientation window (un-blurred)
constexpr int kPbW = 64, kPbH = 37, kPbBytes = kPbW * kPbH, kPbStride = 2432;  // descriptor window (blurred)
#ifndef VXORB_DESC_WARPS
#define VXORB_DESC_WARPS 4
#endif
#ifndef VXORB_DESC_GROUP
#define VXORB_DESC_GROUP 16
#endif
constexpr int kDescWarps = VXORB_DESC_WARPS;
constexpr int kDepthA = 4, kDepthB = 3;            // windows in flight per warp
constexpr int kWinBytes = kDepthB * kPbStride;     // per-warp window memory, shared by both phases
static_assert(kDepthA * kPaStride <= kWinBytes, "orientation windows must fit the per-warp region");

__device__ __forceinline__ uint32_t s_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void lds_v4(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}

}  // namespace

__global__ void __launch_bounds__(32 * kDescWarps)
describe_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blur,
                const uint32_t* __restrict__ fin_key, const float* __restrict__ fin_resp,
                const int* __restrict__ fin_n, int kp_cap, int group, vxorb_keypoint* __restrict__ kp_out,
                uint8_t* __restrict__ desc_out, float* __restrict__ xy_out, int* __restrict__ kp_n,
                int* __restrict__ kp_total, int* __restrict__ status) {
    __shared__ __align__(128) uint8_t win[kDescWarps][kWinBytes];
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g0 = (blockIdx.x * kDescWarps + (threadIdx.x >> 5)) * group;  // first keypoint of this warp's group

    // ---- setup: locate keypoint g0 + lane from the per-level counts -----------------------------
    const int g = g0 + lane;
    int level = -1, idx = 0, total = 0;
    for (int l = 0; l < lt.n_levels; ++l) {
        const int cnt = min(fin_n[l * lt.batch + f], lt.fin_cap[l]);
        if (level < 0 && g < total + cnt) {
            level = l;
            idx = g - total;
        }
        total += cnt;
    }
    if (g0 == 0 && lane == 0) {
        kp_n[f] = min(total, kp_cap);
        kp_total[f] = total;
        if (total > kp_cap) atomicOr(status, 4);
    }
    const int n_here = min(group, min(total, kp_cap) - g0);  // keypoints this warp processes
    if (n_here <= 0) return;
    const bool mine = lane < n_here;
    int x = 0, y = 0;
    float response = 0.f;
    if (mine) {
        const int fcap = lt.fin_cap[level];
        const uint32_t key = fin_key[lt.fin_off[level] + (size_t)f * fcap + idx];
        response = fin_resp[lt.fin_off[level] + (size_t)f * fcap + idx];
        x = (int)(key & 0xfffu);
        y = (int)((key >> 12) & 0xfffu);
    } else {
        level = 0;
    }

    // this lane's pattern points (byte `lane` of the descriptor) as floats
    float px[16], py[16];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const char4 pp = g_pattern.p[k][lane];
        px[2 * k] = (float)pp.x;
        py[2 * k] = (float)pp.y;
        px[2 * k + 1] = (float)pp.z;
        py[2 * k + 1] = (float)pp.w;
    }
    // this lane's patch row v = lane - 15: weights for u = -15 .. 16 (8 words)
    const int v = lane - 15;
    const int um = lane < 31 ? umax_of(abs(v)) : -1;
    uint32_t wu[8], wm[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t a = 0, m = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int u = 4 * j + b - 15;
            if (abs(u) <= um) {
                a |= ((uint32_t)u & 0xffu) << (8 * b);
                m |= 1u << (8 * b);
            }
        }
        wu[j] = a;
        wm[j] = m;
    }

    const uint32_t win0 = s_addr(&win[wid][0]);
    // Window loads of keypoint i: every lane copies its 16-byte chunks (3 per orientation row, 4 per
    // descriptor row; rows start at the 16 B boundary at or below the window's first byte) with
    // cp.async, one commit group per keypoint.  Orientation windows cycle through kDepthA slots,
    // descriptor windows through kDepthB slots of the same per-warp region (the descriptor phase
    // starts after the last orientation window has been read).
    auto load_a = [&](int i) {
        const int kx = __shfl_sync(0xffffffffu, x, i), ky = __shfl_sync(0xffffffffu, y, i);
        const int kl = __shfl_sync(0xffffffffu, level, i);
        const int pitch = lt.pitch[kl];
        VX_CHECK(kl >= 0 && kl < lt.n_levels && kx >= kEdge && ky >= kEdge && kx < lt.w[kl] - kEdge && ky < lt.h[kl] - kEdge);
        const uint8_t* g = pyr + lt.off[kl] + (size_t)f * lt.fstride[kl] + (ptrdiff_t)(ky - 15) * pitch + ((kx - 15) & ~15);
        const uint32_t d = win0 + (i % kDepthA) * kPaStride;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int idx = lane + 32 * k;  // chunk index: row = idx / 3, 16-byte column = idx % 3
            const int row = idx / 3, ch = idx - row * 3;
            if (idx < kPaH * 3) cp_async16(d + idx * 16, g + (ptrdiff_t)row * pitch + ch * 16);
        }
    };
    auto load_b = [&](int i) {
        const int kx = __shfl_sync(0xffffffffu, x, i), ky = __shfl_sync(0xffffffffu, y, i);
        const int kl = __shfl_sync(0xffffffffu, level, i);
        const int pitch = lt.pitch[kl];
        VX_CHECK(kl >= 0 && kl < lt.n_levels && kx >= kEdge && ky >= kEdge && kx < lt.w[kl] - kEdge && ky < lt.h[kl] - kEdge);
        const uint8_t* g = blur + lt.off[kl] + (size_t)f * lt.fstride[kl] + (ptrdiff_t)(ky - 18) * pitch + ((kx - 18) & ~15);
        const uint32_t d = win0 + (i % kDepthB) * kPbStride;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const int idx = lane + 32 * k;  // row = idx / 4, 16-byte column = idx % 4
            if (idx < kPbH * 4) cp_async16(d + idx * 16, g + (ptrdiff_t)(idx >> 2) * pitch + (idx & 3) * 16);
        }
    };

    // ---- orientation moments, one keypoint at a time, lane = patch row --------------------------
#pragma unroll
    for (int i = 0; i < kDepthA; ++i) {
        if (i < n_here) load_a(i);
        cp_async_commit();
    }
    int my_m10 = 0, my_m01 = 0;
    const uint32_t my_row = (uint32_t)min(lane, kPaH - 1) * kPaW;
    for (int i = 0; i < n_here; ++i) {
        const int slot = i % kDepthA;
        const int kx = __shfl_sync(0xffffffffu, x, i);
        const uint32_t off = (uint32_t)(kx - 15) & 15u;
        const uint32_t sh = (off & 3u) * 8u;
        const uint32_t rw = win0 + slot * kPaStride + my_row;
        cp_async_wait<kDepthA - 1>();  // this lane's chunks of keypoint i have landed ...
        __syncwarp();                  // ... and so have everyone else's
        // three 16-byte loads per lane (rows are 48 B apart, so a quarter warp covers all 32 banks: no
        // conflicts, a third of the wavefronts of nine word loads); the window's word offset inside the
        // row is the same for every lane, so the word selection is a warp-uniform switch
        uint32_t W[12];
#pragma unroll
        for (int j = 0; j < 3; ++j) lds_v4(rw + 16 * j, W[4 * j], W[4 * j + 1], W[4 * j + 2], W[4 * j + 3]);
        int m10 = 0;
        uint32_t rs = 0;
        auto moments = [&](auto O) {
            constexpr int o = decltype(O)::value;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t a = __funnelshift_r(W[o + j], W[o + j + 1], sh);
                m10 = dp4a_us(a, wu[j], m10);
                rs = __dp4a(a, wm[j], rs);
            }
        };
        switch (off >> 2) {
            case 0: moments(std::integral_constant<int, 0>{}); break;
            case 1: moments(std::integral_constant<int, 1>{}); break;
            case 2: moments(std::integral_constant<int, 2>{}); break;
            default: moments(std::integral_constant<int, 3>{}); break;
        }
        int m01 = v * (int)rs;
        m10 = __reduce_add_sync(0xffffffffu, m10);
        m01 = __reduce_add_sync(0xffffffffu, m01);
        if (lane == i) {
            my_m10 = m10;
            my_m01 = m01;
        }
        __syncwarp();  // every lane is done with this slot before it is refilled
        if (i + kDepthA < n_here) load_a(i + kDepthA);
        cp_async_commit();  // one group per iteration (possibly empty) keeps the wait distance fixed
    }
    // the region is free: start the first descriptor windows, the angle math below hides their latency
#pragma unroll
    for (int i = 0; i < kDepthB; ++i) {
        if (i < n_here) load_b(i);
        cp_async_commit();
    }

    // ---- angle, cos/sin and the keypoint record: lane i for keypoint i --------------------------
    const float angle = fast_atan2_deg((float)my_m01, (float)my_m10);
    const float ang = __fmul_rn(angle, (float)(3.141592653589793238462643383279502884 / 180.f));
    float ca = (float)cos((double)ang), sa = (float)sin((double)ang);
    if (mine) {
        const float sc = lt.scale[level];
        const size_t row = (size_t)f * kp_cap + g;
        vxorb_keypoint kp;
        kp.x = __fmul_rn((float)x, sc);
        kp.y = __fmul_rn((float)y, sc);
        kp.response = response;
        kp.angle = angle;
        kp.octave = level;
        kp_out[row] = kp;
        xy_out[row * 2] = kp.x;
        xy_out[row * 2 + 1] = kp.y;
    }

    // ---- rotated BRIEF, one keypoint at a time, lane = descriptor byte ---------------------------
    const float magic = 12582912.0f;  // 1.5 * 2^23: low mantissa bits of (t + magic) = round-half-even(t)
    for (int i = 0; i < n_here; ++i) {
        const int slot = i % kDepthB;
        const int kx = __shfl_sync(0xffffffffu, x, i);
        const float c = __shfl_sync(0xffffffffu, ca, i), s = __shfl_sync(0xffffffffu, sa, i);
        const uint32_t off = (uint32_t)(kx - 18) & 15u;
        // address = window centre + iy*64 + ix with iy = bits(ty) - bits(magic), ix likewise (mod 2^32)
        const uint32_t bc = win0 + slot * kPbStride + 18u * kPbW + 18u + off - 0x4B400000u * (uint32_t)kPbW - 0x4B400000u;
        cp_async_wait<kDepthB - 1>();
        __syncwarp();
        unsigned byte = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            uint32_t t[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const float xx = px[2 * k + e], yy = py[2 * k + e];
                const float rx = __fsub_rn(__fmul_rn(xx, c), __fmul_rn(yy, s));
                const float ry = __fadd_rn(__fmul_rn(xx, s), __fmul_rn(yy, c));
                const uint32_t bx = __float_as_uint(__fadd_rn(rx, magic)), by = __float_as_uint(__fadd_rn(ry, magic));
                VX_CHECK(bc + by * (uint32_t)kPbW + bx >= win0 + slot * kPbStride &&
                         bc + by * (uint32_t)kPbW + bx < win0 + slot * kPbStride + kPbBytes);  // gather inside this keypoint's window
                t[e] = lds_u8(bc + by * (uint32_t)kPbW + bx);
            }
            byte |= (t[0] < t[1] ? 1u : 0u) << k;
        }
        VX_CHECK(g0 + i < kp_cap);
        desc_out[((size_t)f * kp_cap + g0 + i) * 32 + lane] = (uint8_t)byte;
        __syncwarp();  // every lane is done with this slot before it is refilled
        if (i + kDepthB < n_here) load_b(i + kDepthB);
        cp_async_commit();
    }
}

// Uploads the transposed pattern to the current device (called once from vxorb_create).
cudaError_t init_device_tables() {
    PatternT host;
    for (int i = 0; i < 32; ++i)
        for (int k = 0; k < 8; ++k) {
            const int8_t* pp = kPatternHost + (i * 8 + k) * 4;
            host.p[k][i] = make_char4(pp[0], pp[1], pp[2], pp[3]);
        }
    return cudaMemcpyToSymbol(g_pattern, &host, sizeof(host));
}

cudaError_t carveout_describe(int pct) { return set_carveout(describe_kernel, pct); }

int launch_describe(const LevelTab& lt, const DeviceBuffers& db, int kp_cap, cudaStream_t s) {
    // 16 keypoints per warp amortise the per-group trigonometry; tiny batches use 4 so that one
    // frame still spreads over ~80 CTAs
    const int group = lt.batch >= 8 ? VXORB_DESC_GROUP : 4;
    const int warps = (kp_cap + group - 1) / group;
    dim3 grid((warps + kDescWarps - 1) / kDescWarps, lt.batch);
    describe_kernel<<<grid, 32 * kDescWarps, 0, s>>>(lt, db.pyr, db.blur, db.fin_key, db.fin_resp, db.fin_n, kp_cap, group, db.kp,
                                                     db.desc, db.xy, db.kp_n, db.kp_total, db.status);
    return 1;
}

}  // namespace vxorb
