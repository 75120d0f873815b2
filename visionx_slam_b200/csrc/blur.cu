// blur.cu — 7x7 sigma=2 Gaussian of every pyramid level (the image rBRIEF samples).
//
// Replaces cv::GaussianBlur(level, 7x7, 2, 2, BORDER_REFLECT_101) inside
// cv::ORB::detectAndCompute (orb_extractor.cpp:13).  OpenCV resolves this to the *float*
// sepFilter2D path; the bytes depend on the exact operation order:
//   row pass     acc = k0*x0;  acc = fma(x_i, k_i, acc)            i = 1..6
//   column pass  acc = k3*r3;  acc = fma(r[3-j] + r[3+j], k[3+j], acc)   j = 1..3
//   output       saturate(round-half-even(acc))
// Tile: 64x32 outputs per CTA; (32+6) x (64+6) source bytes with the reflect-101 halo resolved at
// load time; row-pass results kept as float in shared memory; four outputs per thread per store.
// Roofline: HBM, algorithmic bytes = read + write of every level pixel (1 901 064 B per 640x480).
#include <math.h>

#include "vxorb_internal.h"

namespace vxorb {

namespace {

constexpr int BW = 64, BH = 32;
constexpr int SRC_W = BW + 6, SRC_H = BH + 6;
constexpr int SRC_P = 72;  // padded source row

__device__ __forceinline__ int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}

struct GaussK {
    float k[7];
};

}  // namespace

__global__ void __launch_bounds__(256)
blur_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur, const GaussK gk) {
    __shared__ uint8_t src[SRC_H][SRC_P];
    __shared__ float rows[SRC_H][BW];

    const int tid = threadIdx.x;
    int l = 0;
    while (l + 1 < lt.n_levels && (int)blockIdx.x >= lt.blur_tile_begin[l + 1]) ++l;
    const int tl = blockIdx.x - lt.blur_tile_begin[l];
    const int tx = tl % lt.blur_tiles_x[l], ty = tl / lt.blur_tiles_x[l];
    const int f = blockIdx.y;
    const int w = lt.w[l], h = lt.h[l], pitch = lt.pitch[l];
    const uint8_t* img = pyr + lt.off[l] + (size_t)f * lt.fstride[l];
    uint8_t* out = blur + lt.off[l] + (size_t)f * lt.fstride[l];
    const int ox = tx * BW, oy = ty * BH;

    for (int i = tid; i < SRC_H * SRC_W; i += 256) {
        const int r = i / SRC_W, c = i - r * SRC_W;
        const int gy = reflect101(oy - 3 + r, h), gx = reflect101(ox - 3 + c, w);
        src[r][c] = __ldg(img + (size_t)gy * pitch + gx);
    }
    __syncthreads();

    // row pass: rows[r][x] for the 38 source rows of the tile
    for (int i = tid; i < SRC_H * BW; i += 256) {
        const int r = i / BW, x = i - r * BW;
        const uint8_t* p = &src[r][x];
        float acc = __fmul_rn(gk.k[0], (float)p[0]);
#pragma unroll
        for (int t = 1; t < 7; ++t) acc = __fmaf_rn((float)p[t], gk.k[t], acc);
        rows[r][x] = acc;
    }
    __syncthreads();

    // column pass: each thread produces 4 adjacent pixels of one row, two rows per thread
    for (int i = tid; i < BH * (BW / 4); i += 256) {
        const int y = i / (BW / 4), x0 = (i - y * (BW / 4)) * 4;
        const int gy = oy + y, gx0 = ox + x0;
        if (gy >= h || gx0 >= w) continue;
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = x0 + k;
            float acc = __fmul_rn(gk.k[3], rows[y + 3][x]);
#pragma unroll
            for (int j = 1; j <= 3; ++j)
                acc = __fmaf_rn(__fadd_rn(rows[y + 3 - j][x], rows[y + 3 + j][x]), gk.k[3 + j], acc);
            int v = __float2int_rn(acc);
            v = min(max(v, 0), 255);
            if (gx0 + k < w) packed |= (uint32_t)v << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(out + (size_t)gy * pitch + gx0) = packed;
    }
}

int launch_blur(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s) {
    const int tiles = lt.blur_tile_begin[lt.n_levels];
    if (tiles == 0) return 0;
    GaussK gk;
    {
        // cv::getGaussianKernel(7, 2, CV_32F): normalised in double, then rounded to float
        double t[7], sum = 0;
        for (int i = 0; i < 7; ++i) {
            const double x = (double)(i - 3);
            t[i] = exp(-(x * x) / 8.0);
            sum += t[i];
        }
        for (int i = 0; i < 7; ++i) gk.k[i] = (float)(t[i] / sum);
    }
    dim3 grid(tiles, lt.batch);
    blur_kernel<<<grid, 256, 0, s>>>(lt, db.pyr, db.blur, gk);
    return 1;
}

}  // namespace vxorb
