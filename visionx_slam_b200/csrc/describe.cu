// describe.cu — intensity-centroid orientation, keypoint records and rotated-BRIEF descriptors.
//
// Replaces ICAngles + cv::fastAtan2, the pt *= scale step and computeOrbDescriptors (WTA_K = 2)
// inside cv::ORB::detectAndCompute (orb_extractor.cpp:13); SURVEY.md A.6, A.8.  Output rows are
// what ORBExtractor::Extract copies into Frame::Features() / Frame::Descriptors()
// (orb_extractor.cpp:19-26), in canonical (octave, y, x) order.
//
// One warp per keypoint:
//   orientation  lane j owns column u = j-15 of the radius-15 disc, integer moments, REDUX sum,
//                fastAtan2 polynomial evaluated with un-fused float ops (degrees)
//   descriptor   lane i produces byte i: 8 comparisons of blurred pixels at pattern points rotated
//                by the angle (double cos/sin, float un-fused rotate, round-half-even)
// Float rules are load-bearing: every op is an explicit *_rn intrinsic so nvcc cannot contract.
#include <float.h>

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

__global__ void __launch_bounds__(256)
describe_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blur,
                const uint32_t* __restrict__ fin_key, const float* __restrict__ fin_resp,
                const int* __restrict__ fin_n, int kp_cap, vxorb_keypoint* __restrict__ kp_out,
                uint8_t* __restrict__ desc_out, float* __restrict__ xy_out, int* __restrict__ kp_n,
                int* __restrict__ kp_total, int* __restrict__ status) {
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5);  // keypoint index inside the frame

    // locate (level, index in level) from the per-level counts
    int level = -1, idx = 0, total = 0;
    for (int l = 0; l < lt.n_levels; ++l) {
        const int cnt = min(fin_n[l * lt.batch + f], lt.fin_cap[l]);
        if (level < 0 && g < total + cnt) {
            level = l;
            idx = g - total;
        }
        total += cnt;
    }
    if (g == 0 && lane == 0) {
        kp_n[f] = min(total, kp_cap);
        kp_total[f] = total;
        if (total > kp_cap) atomicOr(status, 4);
    }
    if (level < 0 || g >= kp_cap) return;

    const int fcap = lt.fin_cap[level];
    const uint32_t key = fin_key[lt.fin_off[level] + (size_t)f * fcap + idx];
    const float response = fin_resp[lt.fin_off[level] + (size_t)f * fcap + idx];
    const int x = (int)(key & 0xfffu), y = (int)((key >> 12) & 0xfffu);
    const int pitch = lt.pitch[level];
    const size_t plane = lt.off[level] + (size_t)f * lt.fstride[level];

    // ---- orientation: m10 = sum u*I, m01 = sum v*I over the radius-15 disc --------------------
    const uint8_t* c = pyr + plane + (size_t)y * pitch + x;
    const int u = lane - 15;
    const int au = abs(u);
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int vm = umax_of(au);
        for (int v = -vm; v <= vm; ++v) {
            const int p = __ldg(c + (ptrdiff_t)v * pitch + u);
            m10 += u * p;
            m01 += v * p;
        }
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    m01 = __reduce_add_sync(0xffffffffu, m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- rotated BRIEF ------------------------------------------------------------------------
    const float ang = __fmul_rn(angle, (float)(3.141592653589793238462643383279502884 / 180.f));
    const float ca = (float)cos((double)ang), sa = (float)sin((double)ang);
    const uint8_t* bc = blur + plane + (size_t)y * pitch + x;
    unsigned byte = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const char4 pp = g_pattern.p[k][lane];
        const float x1 = (float)pp.x, y1 = (float)pp.y, x2 = (float)pp.z, y2 = (float)pp.w;
        const int ix1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sa)));
        const int iy1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sa), __fmul_rn(y1, ca)));
        const int ix2 = __float2int_rn(__fsub_rn(__fmul_rn(x2, ca), __fmul_rn(y2, sa)));
        const int iy2 = __float2int_rn(__fadd_rn(__fmul_rn(x2, sa), __fmul_rn(y2, ca)));
        const int t0 = __ldg(bc + (ptrdiff_t)iy1 * pitch + ix1);
        const int t1 = __ldg(bc + (ptrdiff_t)iy2 * pitch + ix2);
        byte |= (t0 < t1 ? 1u : 0u) << k;
    }
    const size_t row = (size_t)f * kp_cap + g;
    desc_out[row * 32 + lane] = (uint8_t)byte;
    if (lane == 0) {
        const float sc = lt.scale[level];
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

int launch_describe(const LevelTab& lt, const DeviceBuffers& db, int kp_cap, cudaStream_t s) {
    dim3 grid((kp_cap + 7) / 8, lt.batch);
    describe_kernel<<<grid, 256, 0, s>>>(lt, db.pyr, db.blur, db.fin_key, db.fin_resp, db.fin_n, kp_cap, db.kp,
                                         db.desc, db.xy, db.kp_n, db.kp_total, db.status);
    return 1;
}

}  // namespace vxorb
