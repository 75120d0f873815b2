// blur.cu — 7x7 sigma=2 Gaussian of every pyramid level (the image rBRIEF samples).
//
// Replaces cv::GaussianBlur(level, 7x7, 2, 2, BORDER_REFLECT_101) inside
// cv::ORB::detectAndCompute (orb_extractor.cpp:13).  OpenCV resolves this to the *float*
// sepFilter2D path; the bytes depend on the exact operation order:
//   row pass     acc = k0*x0;  acc = fma(x_i, k_i, acc)            i = 1..6
//   column pass  acc = k3*r3;  acc = fma(r[3-j] + r[3+j], k[3+j], acc)   j = 1..3
//   output       saturate(round-half-even(acc))
//
// Register-resident separable filter, no shared memory.  One thread owns a strip of 4 adjacent
// columns x kBlurRows output rows and walks down the source rows:
//   * per source row: three aligned 32-bit loads (columns x0-4 .. x0+7; the production extent stays kBlurMargin px
//     inside the level — only what a descriptor window can read — so no border is ever touched; whole-level runs
//     read the reflect-101 apron written by apron_kernel as plain pixels), ten byte->float conversions
//     (I2F.U8 with a byte selector, on the conversion unit), 4 x 7 row-pass FMAs with the taps as immediates;
//   * the last seven row-pass rows live in registers (window rotated by unrolling x7);
//   * per output row: 4 x 7 column-pass ops, round-half-even by adding 1.5*2^23, pack, one
//     32-bit store.  Consecutive lanes own consecutive words, so loads and stores coalesce.
// Strips are enumerated per level (band-major, then word) and dealt to threads linearly, so lane
// utilisation does not depend on the level width.
// Roofline: HBM, algorithmic bytes = read + write of every level pixel (1 901 064 B per 640x480);
// the kernel is bound by instruction issue (79 % of the issue slots, fma pipe 47 %): 15 FP operations, 2.5
// conversions and ~5.5 load / store / address / predicate instructions per pixel; 0.131 ms per 128 frames.
// Packed fma.rn.f32x2 / add.f32x2 (FFMA2 / FADD2) were tried: 25 % fewer instructions, bit-identical
// output, same time — the fma pipe spends two passes on a packed instruction.
#include <math.h>
#include <string.h>

#include "vxorb_internal.h"

namespace vxorb {

namespace {

struct GaussK {
    float k[7];
};

#ifndef VXORB_BLUR_I2F
#define VXORB_BLUR_I2F 1
#endif
__device__ __forceinline__ float byte_to_float(uint32_t word, uint32_t sel) {
#if VXORB_BLUR_I2F
    // one I2F.U8 with a byte selector on the conversion unit instead of PRMT (alu) + FADD (fma): the same time for
    // the kernel alone (0.131 ms per 128 frames), but 3 % more resident throughput once it runs beside the
    // alu-bound FAST / matching kernels of the other pipeline slots (scratch/ab_run.py on a B200)
    return (float)((word >> ((sel & 3u) * 8u)) & 0xffu);
#else
    // bytes [b, 0, 0, 0x4B] = 2^23 + b as a float; the subtraction is exact
    return __fsub_rn(__uint_as_float(__byte_perm(word, 0x4B000000u, sel)), 8388608.0f);
#endif
}

}  // namespace

// cv::getGaussianKernel(7, 2, CV_32F) as literals: with the taps as immediates the row and column FMAs take the
// FFMA-with-immediate form (one operand less to fetch from the register file per instruction); launch_blur
// checks the literals against the kernel computed the way OpenCV computes it and falls back to the
// parameter form if they ever differ.
constexpr float kG0 = 0.0701593235f, kG1 = 0.131074876f, kG2 = 0.190712824f, kG3 = 0.216105938f;

constexpr int kBC = kBlurCols;            // columns per thread (4 or 8)
constexpr int kBW = (kBC + 8) / 4;        // aligned source words per row: columns x0-4 .. x0+kBC+3
constexpr int kBF = kBC + 6;              // source pixels per row: x0-3 .. x0+kBC+2

template <bool IMM>
__global__ void __launch_bounds__(256, kBC == 8 ? 2 : 3)
blur_kernel(const LevelTab lt, const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur, const GaussK gkp) {
    GaussK gk;
    if (IMM) {
        gk.k[0] = kG0; gk.k[1] = kG1; gk.k[2] = kG2; gk.k[3] = kG3; gk.k[4] = kG2; gk.k[5] = kG1; gk.k[6] = kG0;
    } else {
        gk = gkp;
    }
    int l = 0;
    while (l + 1 < lt.n_levels && (int)blockIdx.x >= lt.blur_tile_begin[l + 1]) ++l;
    const int y_end = lt.blur_y1[l], pitch = lt.pitch[l];
    const int words = lt.blur_tiles_x[l];
    const int item = ((int)blockIdx.x - lt.blur_tile_begin[l]) * 256 + (int)threadIdx.x;
    const int band = item / words;
    const int y0 = lt.blur_y0[l] + band * kBlurRows;
    if (y0 >= y_end) return;
    const int x0 = lt.blur_x0[l] + (item - band * words) * kBC;
    const int rows_out = min(kBlurRows, y_end - y0);
    const size_t plane = lt.off[l] + (size_t)blockIdx.y * lt.fstride[l];
    // source row r (0-based in the strip) is image row y0 - 3 + r; the first word starts at x0 - 4
    const uint8_t* src = pyr + plane + (ptrdiff_t)(y0 - kApronY) * pitch + (x0 - 4);
    uint8_t* dst = blur + plane + (size_t)y0 * pitch + x0;
    const int last_src = rows_out + 2 * kApronY - 1;
    const int groups = (rows_out + 2 * kApronY + 6) / 7;

    float win[7][kBC];
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
        for (int j = 0; j < kBC; ++j) win[i][j] = 0.f;

    const uint32_t* p = reinterpret_cast<const uint32_t*>(src);
    const int pitch_w = pitch >> 2;  // pitch is a multiple of 32
    // software pipeline: the words of source rows r+1 and r+2 are in flight while row r is filtered
    // (the pointer stops at the last apron row, so the tail re-reads that row instead of running on)
    uint32_t q[2][kBW];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
        for (int k = 0; k < kBW; ++k) q[i][k] = __ldg(p + k);
        if (i < last_src) p += pitch_w;
    }
    int r = 0;
    dst -= (size_t)6 * pitch;        // output row of source row r is r - 6
    for (int g = 0; g < groups; ++g) {
#pragma unroll
        for (int s = 0; s < 7; ++s, ++r) {
            uint32_t w[kBW];
#pragma unroll
            for (int k = 0; k < kBW; ++k) {
                w[k] = q[0][k];
                q[0][k] = q[1][k];
                q[1][k] = __ldg(p + k);
            }
            if (r + 2 < last_src) p += pitch_w;  // never read below the apron
            float f[kBF];
#pragma unroll
            for (int i = 0; i < kBF; ++i) f[i] = byte_to_float(w[(i + 1) >> 2], 0x7540u + (uint32_t)((i + 1) & 3));
#pragma unroll
            for (int j = 0; j < kBC; ++j) {
                float acc = __fmul_rn(gk.k[0], f[j]);
#pragma unroll
                for (int t = 1; t < 7; ++t) acc = __fmaf_rn(f[j + t], gk.k[t], acc);
                win[s][j] = acc;
            }
            // rows r-6 .. r are in win[(s+1)%7] .. win[s]; the output row is r - 6
            const int o = r - 6;
            if (o >= 0 && o < rows_out) {
                uint32_t t4[kBC];
#pragma unroll
                for (int j = 0; j < kBC; ++j) {
                    float acc = __fmul_rn(gk.k[3], win[(s + 4) % 7][j]);
                    acc = __fmaf_rn(__fadd_rn(win[(s + 3) % 7][j], win[(s + 5) % 7][j]), gk.k[4], acc);
                    acc = __fmaf_rn(__fadd_rn(win[(s + 2) % 7][j], win[(s + 6) % 7][j]), gk.k[5], acc);
                    acc = __fmaf_rn(__fadd_rn(win[(s + 1) % 7][j], win[s][j]), gk.k[6], acc);
                    // 0 <= acc <= 255.001: adding 1.5 * 2^23 leaves round-half-even(acc) in the low byte
                    t4[j] = __float_as_uint(__fadd_rn(acc, 12582912.0f));
                }
                uint32_t outw[kBC / 4];
#pragma unroll
                for (int k = 0; k < kBC / 4; ++k) {
                    const uint32_t lo = __byte_perm(t4[4 * k], t4[4 * k + 1], 0x0040u);
                    const uint32_t hi = __byte_perm(t4[4 * k + 2], t4[4 * k + 3], 0x0040u);
                    outw[k] = __byte_perm(lo, hi, 0x5410u);
                }
                if (kBC == 8)
                    *reinterpret_cast<uint2*>(dst) = make_uint2(outw[0], outw[kBC / 4 - 1]);
                else
                    *reinterpret_cast<uint32_t*>(dst) = outw[0];
            }
            dst += pitch;
        }
    }
}

cudaError_t carveout_blur(int pct) {
    const cudaError_t e = set_carveout(blur_kernel<true>, pct);
    return e != cudaSuccess ? e : set_carveout(blur_kernel<false>, pct);
}

int launch_blur(const LevelTab& lt, const DeviceBuffers& db, cudaStream_t s) {
    const int ctas = lt.blur_tile_begin[lt.n_levels];
    if (ctas == 0) return 0;
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
    dim3 grid(ctas, lt.batch);
    const float lit[7] = {kG0, kG1, kG2, kG3, kG2, kG1, kG0};
    if (memcmp(lit, gk.k, sizeof(lit)) == 0)
        blur_kernel<true><<<grid, 256, 0, s>>>(lt, db.pyr, db.blur, gk);
    else
        blur_kernel<false><<<grid, 256, 0, s>>>(lt, db.pyr, db.blur, gk);
    return 1;
}

}  // namespace vxorb
