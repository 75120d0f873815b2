// frame_info.cu — the per-frame quantities the tracking front end derives from the extractor and
// matcher outputs ("next" rows), computed where the data already lives:
//
//   image statistics   Tracking::CheckImageQuality (tracking.cpp:120-139): cv::meanStdDev of the gray
//                      frame.  Integer sums S = sum p and Q = sum p^2 over level 0 (IDP.4A, exact),
//                      then OpenCV's double expression  mean = S * (1/N),
//                      sd = sqrt(max(Q * (1/N) - mean^2, 0))  un-fused — bit-exact against cv2.
//   grid coverage      Tracking::CheckFeatureDistribution (tracking.cpp:93-118): 5 x 5 grid over the
//                      frame, cell (col, row) = (int)((x / width) * 5), (int)((y / height) * 5) in
//                      double from the float keypoint position, clamped; a 25-bit occupancy mask and
//                      its popcount per frame.
//   correspondences    EstimatePoseByEssential's pts_last / pts_curr (tracking.cpp:505-514) and
//                      ComputeParallax (tracking.cpp:548-560): for every accepted match the two
//                      keypoint positions, and the mean of their Euclidean distances in double
//                      (distances un-fused; summed by a fixed-shape tree, so the value is
//                      deterministic and within 1e-12 relative of the reference's serial sum).
//
// Bound: the statistics kernel streams level 0 once (HBM, 307 200 B per 640x480 frame); the other two
// touch a few KB per frame and are latency bound.
#include "vxorb_internal.h"

namespace vxorb {

namespace {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = __dadd_rn(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace

// grid (slices, frames): each CTA sums kStatRows rows of level 0 of one frame
constexpr int kStatRows = 32;

__global__ void __launch_bounds__(256)
image_sums_kernel(const uint8_t* __restrict__ pyr, size_t off0, size_t fstride, int pitch, int w, int h,
                  unsigned long long* __restrict__ sums /* [frame][2] */) {
    const int f = blockIdx.y;
    const int y0 = blockIdx.x * kStatRows, y1 = min(y0 + kStatRows, h);
    const uint8_t* img = pyr + off0 + (size_t)f * fstride;
    const int words = (w + 3) >> 2;
    const int tail = w & 3;  // valid bytes of the last word when the width is not a multiple of 4
    unsigned long long s = 0, q = 0;
    for (int y = y0 + (int)(threadIdx.x >> 5); y < y1; y += 8) {
        const uint32_t* row = reinterpret_cast<const uint32_t*>(img + (size_t)y * pitch);
        uint32_t rs = 0, rq = 0;  // one row: <= 4096 * 255^2 < 2^32
        for (int k = threadIdx.x & 31; k < words; k += 32) {
            uint32_t v = __ldg(row + k);
            if (tail && k == words - 1) v &= (1u << (8 * tail)) - 1u;
            rs = __dp4a(v, 0x01010101u, rs);
            rq = __dp4a(v, v, rq);
        }
        s += rs;
        q += rq;
    }
    __shared__ unsigned long long sh[2][8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if ((threadIdx.x & 31) == 0) {
        sh[0][threadIdx.x >> 5] = s;
        sh[1][threadIdx.x >> 5] = q;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        unsigned long long t = 0;
        for (int i = 0; i < 8; ++i) t += sh[threadIdx.x][i];
        atomicAdd(sums + (size_t)f * 2 + threadIdx.x, t);
    }
}

// One CTA per frame: finishes the statistics, builds the grid mask from the frame's keypoints, gathers
// the matched positions of pair (frame - 1 -> frame) and reduces the parallax.
__global__ void __launch_bounds__(256)
frame_info_kernel(const unsigned long long* __restrict__ sums, int w, int h, const float* __restrict__ xy /* [frame+1][cap][2], slot 0 = carry */,
                  const int* __restrict__ kp_n /* [frame+1] */, int kp_cap, const vxorb_match_t* __restrict__ match,
                  const int* __restrict__ match_n, int grid_cols, int grid_rows, vxorb_frame_info* __restrict__ info,
                  float4* __restrict__ pts /* [frame][cap] or nullptr */) {
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ uint32_t s_mask[8];
    __shared__ double s_sum[8];

    // ---- grid coverage of this frame's keypoints (slot f + 1) ----
    const float2* my = reinterpret_cast<const float2*>(xy) + (size_t)(f + 1) * kp_cap;
    const int n = min(kp_n[f + 1], kp_cap);
    uint32_t mask = 0;
    for (int i = tid; i < n; i += 256) {
        const float2 p = __ldg(my + i);
        int col = (int)__dmul_rn(__ddiv_rn((double)p.x, (double)w), (double)grid_cols);
        int row = (int)__dmul_rn(__ddiv_rn((double)p.y, (double)h), (double)grid_rows);
        col = min(max(col, 0), grid_cols - 1);
        row = min(max(row, 0), grid_rows - 1);
        mask |= 1u << (col * grid_rows + row);
    }
    mask = __reduce_or_sync(0xffffffffu, mask);

    // ---- correspondences of pair f: query = slot f, train = slot f + 1 ----
    const float2* qxy = reinterpret_cast<const float2*>(xy) + (size_t)f * kp_cap;
    const vxorb_match_t* m = match + (size_t)f * kp_cap;
    const int nm = min(match_n[f], kp_cap);
    double acc = 0.0;
    for (int i = tid; i < nm; i += 256) {
        const vxorb_match_t e = m[i];
        const float2 a = __ldg(qxy + e.query_idx), b = __ldg(my + e.train_idx);
        if (pts) pts[(size_t)f * kp_cap + i] = make_float4(a.x, a.y, b.x, b.y);
        const double dx = __dsub_rn((double)a.x, (double)b.x), dy = __dsub_rn((double)a.y, (double)b.y);
        acc = __dadd_rn(acc, __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy))));
    }
    acc = warp_sum(acc);
    if (lane == 0) {
        s_mask[warp] = mask;
        s_sum[warp] = acc;
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t mk = 0;
        double tot = 0.0;
        for (int i = 0; i < 8; ++i) {
            mk |= s_mask[i];
            tot = __dadd_rn(tot, s_sum[i]);
        }
        vxorb_frame_info o;
        const double scale = __ddiv_rn(1.0, (double)((long long)w * h));
        const double mean = __dmul_rn((double)sums[(size_t)f * 2], scale);
        const double var = __dsub_rn(__dmul_rn((double)sums[(size_t)f * 2 + 1], scale), __dmul_rn(mean, mean));
        o.mean = mean;
        o.stddev = __dsqrt_rn(fmax(var, 0.0));
        o.parallax = nm > 0 ? __ddiv_rn(tot, (double)nm) : 0.0;
        o.grid_occupied = __popc(mk);
        o.grid_mask = mk;
        o.n_keypoints = n;
        o.n_matches = nm;
        info[f] = o;
    }
}

// Depth of every feature as Tracking::CreateLandmarksFromDepth reads it (tracking.cpp:606-637): u = (int)(x + 0.5),
// v = (int)(y + 0.5) on the position widened to double; 0 where the reference skips the feature (outside the image,
// 16-bit depth 0, depth in metres outside [0.1, 10]).
__global__ void __launch_bounds__(256)
depth_lookup_kernel(const float2* __restrict__ xy, const int* __restrict__ n_dev, int max_n, const uint8_t* __restrict__ depth,
                    int depth_type, int cols, int rows, size_t stride, double* __restrict__ out) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= max_n) return;
    double m = 0.0;
    if (i < *n_dev) {
        const float2 p = __ldg(xy + i);
        const int u = (int)__dadd_rn((double)p.x, 0.5), v = (int)__dadd_rn((double)p.y, 0.5);
        if (u >= 0 && u < cols && v >= 0 && v < rows) {
            const uint8_t* row = depth + (size_t)v * stride;
            bool ok = true;
            if (depth_type == 0) {
                const unsigned short d = __ldg(reinterpret_cast<const unsigned short*>(row) + u);
                ok = d != 0;
                m = __ddiv_rn((double)d, 5000.0);
            } else if (depth_type == 1) {
                m = (double)__ldg(reinterpret_cast<const float*>(row) + u);
            } else {
                m = __ldg(reinterpret_cast<const double*>(row) + u);
            }
            if (!ok || m < 0.1 || m > 10.0) m = 0.0;  // the reference's two comparisons (a NaN depth passes both, as there)
        }
    }
    out[i] = m;
}

int launch_depth_lookup(const float* xy, const int* n_dev, int max_n, const void* depth, int depth_type, int cols, int rows,
                        size_t stride, double* out, cudaStream_t s) {
    if (max_n <= 0) return 0;
    depth_lookup_kernel<<<(max_n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float2*>(xy), n_dev, max_n,
                                                            static_cast<const uint8_t*>(depth), depth_type, cols, rows, stride, out);
    return 1;
}

int launch_frame_info(const LevelTab& lt, const DeviceBuffers& db, int kp_cap, unsigned long long* sums,
                      vxorb_frame_info* info, float4* pts, int grid_cols, int grid_rows, cudaStream_t s) {
    if (lt.batch <= 0) return 0;
    cudaMemsetAsync(sums, 0, sizeof(unsigned long long) * 2 * lt.batch, s);
    dim3 grid((lt.h[0] + kStatRows - 1) / kStatRows, lt.batch);
    image_sums_kernel<<<grid, 256, 0, s>>>(db.pyr, (size_t)lt.off[0], (size_t)lt.fstride[0], lt.pitch[0], lt.w[0], lt.h[0], sums);
    frame_info_kernel<<<lt.batch, 256, 0, s>>>(sums, lt.w[0], lt.h[0], db.xy, db.kp_n, kp_cap, db.match, db.match_n,
                                               grid_cols, grid_rows, info, pts);
    return 2;
}

}  // namespace vxorb
