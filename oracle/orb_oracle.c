/*
 * orb_oracle.c — CPU restatement of the reference's ORB extract + Hamming match path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's CPU-baseline leg may load this, and only as the
 * checker / the timed CPU baseline.  The shipped path (visionx_slam_b200/csrc) never calls it.
 *
 * What it restates.  The reference's wrappers (core/feature/orb_extractor.cpp:5-27,
 * core/feature/orb_matcher.cpp:11-43) delegate all arithmetic to OpenCV 4 (vcpkg `opencv4`,
 * builtin-baseline 25b458671af03578e6a34edd8f0d1ac85e084df4, vcpkg.json:4,8-11), which is not in
 * the reference tree.  This file restates OpenCV's published algorithm for
 * cv::ORB(1000,1.2f,8)::detectAndCompute and cv::BFMatcher(NORM_HAMMING)::knnMatch as specified in
 * SURVEY.md Appendix A, stage by stage.
 *
 * Parity pin.  The reference holds no tests or golden vectors for this path ("parity unpinned"
 * by the reference itself).  The pin used instead is OpenCV itself: tests/golden/make_golden.py
 * (committed) runs cv2 4.13.0 — the same OpenCV entry points the reference calls, plus the per-stage
 * primitives — and freezes the outputs as the .npz files under tests/golden/; tests/test_oracle_golden.py replays
 * every one of them against the functions below, bit for bit.
 *
 * Build: gcc -O2 -ffp-contract=off (the float expressions below rely on un-fused rounding
 * except where fmaf is written out).
 */
#include "orb_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static const int8_t kPattern[1024] = {
#include "../include/vxorb_pattern.inc"
};

/* cvRound: round-half-to-even (SSE cvtss2si / lrint under the default rounding mode). */
static inline int round_f(float v) { return (int)lrintf(v); }
static inline int round_d(double v) { return (int)lrint(v); }

/* ---------------------------------------------------------------------------------------------
 * A.0  BGR -> gray, OpenCV's 15-bit fixed point (cvtColor inside detectAndCompute, reached from
 * orb_extractor.cpp:13 when system.cpp:46 supplies IMREAD_COLOR frames). */
void vxo_bgr2gray(const uint8_t* bgr, int w, int h, size_t stride, uint8_t* gray, size_t gstride) {
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = bgr + (size_t)y * stride;
        uint8_t* d = gray + (size_t)y * gstride;
        for (int x = 0; x < w; ++x) {
            int b = s[3 * x], g = s[3 * x + 1], r = s[3 * x + 2];
            d[x] = (uint8_t)((r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15);
        }
    }
}

/* ---------------------------------------------------------------------------------------------
 * A.1 / A.4  Level geometry and per-level quotas (cv::ORB, orb_extractor.cpp:6). */
int vxo_level_geometry(int W, int H, int n_features, float scale_factor, int n_levels, int* lw,
                       int* lh, float* lscale, int* quota) {
    if (n_levels < 1 || n_levels > 16) return -1;
    const double sf = (double)scale_factor;
    for (int l = 0; l < n_levels; ++l) {
        float sc = (float)pow(sf, (double)l);
        float inv = 1.0f / sc;
        if (lscale) lscale[l] = sc;
        if (lw) lw[l] = round_f((float)W * inv);
        if (lh) lh[l] = round_f((float)H * inv);
    }
    if (quota) {
        float factor = (float)(1.0 / sf);
        float nd = (float)n_features * (1.0f - factor) /
                   (1.0f - (float)pow((double)factor, (double)n_levels));
        int sum = 0;
        for (int l = 0; l < n_levels - 1; ++l) {
            quota[l] = round_f(nd);
            sum += quota[l];
            nd *= factor;
        }
        int last = n_features - sum;
        quota[n_levels - 1] = last > 0 ? last : 0;
    }
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * A.1  resize(INTER_LINEAR_EXACT): 8.8 fixed-point taps, 16.16 accumulate. */
static void linear_coeffs(int ssize, int dsize, int* ofs, int* c1) {
    double scale = 1.0 / ((double)dsize / (double)ssize);
    for (int i = 0; i < dsize; ++i) {
        double f = scale * ((double)i + 0.5) - 0.5;
        int i0 = (int)floor(f);
        if (i0 < 0) { /* left edge: single tap src[0] */
            ofs[i] = 0;
            c1[i] = 0;
        } else if (i0 >= ssize - 1) { /* right edge: single tap src[ssize-1] */
            ofs[i] = ssize - 1;
            c1[i] = 0;
        } else {
            ofs[i] = i0;
            c1[i] = round_d((f - (double)i0) * 256.0);
        }
    }
}

void vxo_resize_linear_exact(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst,
                             int dw, int dh, size_t dstride) {
    int* xo = (int*)malloc(sizeof(int) * (size_t)dw * 2);
    int* xc = xo + dw;
    int* yo = (int*)malloc(sizeof(int) * (size_t)dh * 2);
    int* yc = yo + dh;
    linear_coeffs(sw, dw, xo, xc);
    linear_coeffs(sh, dh, yo, yc);
    for (int y = 0; y < dh; ++y) {
        int j0 = yo[y], j1 = j0 + 1 < sh ? j0 + 1 : j0;
        uint32_t cy1 = (uint32_t)yc[y], cy0 = 256u - cy1;
        const uint8_t* r0 = src + (size_t)j0 * sstride;
        const uint8_t* r1 = src + (size_t)j1 * sstride;
        uint8_t* d = dst + (size_t)y * dstride;
        for (int x = 0; x < dw; ++x) {
            int i0 = xo[x], i1 = i0 + 1 < sw ? i0 + 1 : i0;
            uint32_t cx1 = (uint32_t)xc[x], cx0 = 256u - cx1;
            uint32_t h0 = r0[i0] * cx0 + r0[i1] * cx1; /* 8.8, <= 65280 */
            uint32_t h1 = r1[i0] * cx0 + r1[i1] * cx1;
            uint32_t v = h0 * cy0 + h1 * cy1; /* 16.16 */
            v = (v + (1u << 15)) >> 16;
            d[x] = (uint8_t)(v > 255u ? 255u : v);
        }
    }
    free(xo);
    free(yo);
}

/* ---------------------------------------------------------------------------------------------
 * A.2 / A.3  FAST-9/16 with cornerScore and strict 3x3 non-max suppression. */
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* Score of pixel (x,y) or 0 when it is not a FAST-9 corner at threshold t. */
static int fast_score(const uint8_t* img, size_t stride, int x, int y, int t) {
    const uint8_t* p = img + (size_t)y * stride + x;
    int v = p[0];
    int d[16];
    unsigned bright = 0, dark = 0;
    for (int k = 0; k < 16; ++k) {
        int q = p[(ptrdiff_t)kRingDy[k] * (ptrdiff_t)stride + kRingDx[k]];
        d[k] = v - q;
        if (q > v + t) bright |= 1u << k;
        if (q < v - t) dark |= 1u << k;
    }
    int corner = 0;
    for (int s = 0; s < 16 && !corner; ++s) {
        unsigned arc = 0;
        for (int j = 0; j < 9; ++j) arc |= 1u << ((s + j) & 15);
        if ((bright & arc) == arc || (dark & arc) == arc) corner = 1;
    }
    if (!corner) return 0;
    int best = -256;
    for (int s = 0; s < 16; ++s) {
        int mn = 256, mx = -256;
        for (int j = 0; j < 9; ++j) {
            int dv = d[(s + j) & 15];
            if (dv < mn) mn = dv;
            if (dv > mx) mx = dv;
        }
        if (mn > best) best = mn;   /* centre brighter than the whole arc by mn */
        if (-mx > best) best = -mx; /* centre darker than the whole arc by -mx  */
    }
    return best - 1;
}

int vxo_fast9_nms(const uint8_t* img, int w, int h, size_t stride, int threshold, int border,
                  vxorb_corner* out, int cap) {
    if (border < 3) border = 3;
    if (w < 7 || h < 7) return 0;
    /* score map over the valid FAST domain 3 <= x <= w-4 */
    uint8_t* sc = (uint8_t*)calloc((size_t)w * (size_t)h, 1);
    int x0 = border - 1 < 3 ? 3 : border - 1, y0 = x0;
    int x1 = w - border + 1 > w - 3 ? w - 3 : w - border + 1; /* exclusive */
    int y1 = h - border + 1 > h - 3 ? h - 3 : h - border + 1;
    for (int y = y0; y < y1; ++y)
        for (int x = x0; x < x1; ++x) sc[(size_t)y * w + x] = (uint8_t)fast_score(img, stride, x, y, threshold);
    int n = 0;
    for (int y = border; y < h - border; ++y) {
        for (int x = border; x < w - border; ++x) {
            int s = sc[(size_t)y * w + x];
            if (!s) continue;
            const uint8_t* a = sc + (size_t)(y - 1) * w + x;
            const uint8_t* b = sc + (size_t)(y + 1) * w + x;
            const uint8_t* c = sc + (size_t)y * w + x;
            if (s > a[-1] && s > a[0] && s > a[1] && s > c[-1] && s > c[1] && s > b[-1] &&
                s > b[0] && s > b[1]) {
                if (n < cap) {
                    out[n].x = x;
                    out[n].y = y;
                    out[n].score = s;
                }
                ++n;
            }
        }
    }
    free(sc);
    return n;
}

/* ---------------------------------------------------------------------------------------------
 * A.5  Harris response, 7x7 block, k = 0.04, un-fused float combine. */
float vxo_harris(const uint8_t* img, size_t stride, int x, int y) {
    const ptrdiff_t st = (ptrdiff_t)stride;
    int a = 0, b = 0, c = 0;
    for (int dy = -3; dy <= 3; ++dy) {
        for (int dx = -3; dx <= 3; ++dx) {
            const uint8_t* p = img + (ptrdiff_t)(y + dy) * st + (x + dx);
            int ix = (p[1] - p[-1]) * 2 + (p[-st + 1] - p[-st - 1]) + (p[st + 1] - p[st - 1]);
            int iy = (p[st] - p[-st]) * 2 + (p[st - 1] - p[-st - 1]) + (p[st + 1] - p[-st + 1]);
            a += ix * ix;
            b += iy * iy;
            c += ix * iy;
        }
    }
    const float scale = 1.0f / (4 * 7 * 255.0f);
    const float s4 = scale * scale * scale * scale;
    float fa = (float)a, fb = (float)b, fc = (float)c;
    float det = fa * fb - fc * fc;
    float tr = fa + fb;
    float r = det - (0.04f * tr) * tr;
    return r * s4;
}

/* ---------------------------------------------------------------------------------------------
 * A.6  Intensity-centroid orientation + cv::fastAtan2 (degrees). */
static const int kUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

float vxo_fast_atan2(float y, float x) {
    const float rad2deg = (float)(180.0 / 3.141592653589793238462643383279502884);
    const float p1 = 0.9997878412794807f * rad2deg;
    const float p3 = -0.3258083974640975f * rad2deg;
    const float p5 = 0.1555786518463281f * rad2deg;
    const float p7 = -0.04432655554792128f * rad2deg;
    float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.0f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.0f - a;
    if (y < 0) a = 360.0f - a;
    return a;
}

float vxo_ic_angle(const uint8_t* img, size_t stride, int x, int y) {
    const ptrdiff_t st = (ptrdiff_t)stride;
    const uint8_t* c = img + (ptrdiff_t)y * st + x;
    int m01 = 0, m10 = 0;
    for (int v = -15; v <= 15; ++v) {
        int um = kUmax[v < 0 ? -v : v];
        for (int u = -um; u <= um; ++u) {
            int p = c[(ptrdiff_t)v * st + u];
            m10 += u * p;
            m01 += v * p;
        }
    }
    return vxo_fast_atan2((float)m01, (float)m10);
}

/* ---------------------------------------------------------------------------------------------
 * A.7  7x7 sigma=2 Gaussian through the float sepFilter2D path, BORDER_REFLECT_101. */
static inline int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        if (i >= n) i = 2 * (n - 1) - i;
    }
    return i;
}

void vxo_gauss_kernel7(float* k) {
    /* cv::getGaussianKernel(7, 2, CV_32F): double weights normalised, then cast. */
    double t[7], s = 0;
    for (int i = 0; i < 7; ++i) {
        double x = (double)(i - 3);
        t[i] = exp(-(x * x) / 8.0);
        s += t[i];
    }
    for (int i = 0; i < 7; ++i) k[i] = (float)(t[i] / s);
}

void vxo_blur7(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride) {
    float k[7];
    vxo_gauss_kernel7(k);
    float* rows = (float*)malloc(sizeof(float) * (size_t)w * (size_t)h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * sstride;
        float* r = rows + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            float acc = k[0] * (float)s[reflect101(x - 3, w)];
            for (int i = 1; i < 7; ++i) acc = fmaf((float)s[reflect101(x - 3 + i, w)], k[i], acc);
            r[x] = acc;
        }
    }
    for (int y = 0; y < h; ++y) {
        const float* r[7];
        for (int i = 0; i < 7; ++i) r[i] = rows + (size_t)reflect101(y - 3 + i, h) * w;
        uint8_t* d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            float acc = k[3] * r[3][x];
            for (int j = 1; j <= 3; ++j) acc = fmaf(r[3 - j][x] + r[3 + j][x], k[3 + j], acc);
            int v = round_f(acc);
            d[x] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
        }
    }
    free(rows);
}

/* ---------------------------------------------------------------------------------------------
 * A.8  rotated BRIEF, WTA_K = 2, 256 bits. */
void vxo_brief(const uint8_t* blurred, size_t stride, int cx, int cy, float angle_deg,
               uint8_t* desc) {
    float ang = angle_deg * (float)(3.141592653589793238462643383279502884 / 180.f);
    float a = (float)cos((double)ang), b = (float)sin((double)ang);
    const uint8_t* center = blurred + (ptrdiff_t)cy * (ptrdiff_t)stride + cx;
    for (int i = 0; i < 32; ++i) {
        unsigned byte = 0;
        for (int k = 0; k < 8; ++k) {
            const int8_t* pp = kPattern + (i * 8 + k) * 4;
            float x0 = (float)pp[0] * a - (float)pp[1] * b;
            float y0 = (float)pp[0] * b + (float)pp[1] * a;
            float x1 = (float)pp[2] * a - (float)pp[3] * b;
            float y1 = (float)pp[2] * b + (float)pp[3] * a;
            int t0 = center[(ptrdiff_t)round_f(y0) * (ptrdiff_t)stride + round_f(x0)];
            int t1 = center[(ptrdiff_t)round_f(y1) * (ptrdiff_t)stride + round_f(x1)];
            byte |= (unsigned)(t0 < t1) << k;
        }
        desc[i] = (uint8_t)byte;
    }
}

/* ---------------------------------------------------------------------------------------------
 * A.4  KeyPointsFilter::retainBest on a float key: keep everything >= the n-th largest. */
static int cmp_float_desc(const void* a, const void* b) {
    float fa = *(const float*)a, fb = *(const float*)b;
    return fa > fb ? -1 : (fa < fb ? 1 : 0);
}

/* Returns the threshold; *keep_all is set when count <= n. */
static float retain_threshold(const float* keys, int count, int n, int* keep_all) {
    *keep_all = 0;
    if (count <= n) {
        *keep_all = 1;
        return 0.f;
    }
    float* tmp = (float*)malloc(sizeof(float) * (size_t)count);
    memcpy(tmp, keys, sizeof(float) * (size_t)count);
    qsort(tmp, (size_t)count, sizeof(float), cmp_float_desc);
    float thr = tmp[n - 1];
    free(tmp);
    return thr;
}

/* ---------------------------------------------------------------------------------------------
 * Whole extraction = cv::ORB::detectAndCompute (orb_extractor.cpp:13), canonical output order
 * (octave, y, x).  Returns the keypoint count (may exceed cap; only cap entries are written). */
int vxo_orb_extract(const uint8_t* gray, int W, int H, size_t stride, int n_features,
                    float scale_factor, int n_levels, int fast_threshold, vxorb_keypoint* kp_out,
                    uint8_t* desc_out, int cap) {
    int lw[16], lh[16], quota[16];
    float lscale[16];
    if (vxo_level_geometry(W, H, n_features, scale_factor, n_levels, lw, lh, lscale, quota)) return -1;
    uint8_t* lvl[16] = {0};
    int total = 0;
    for (int l = 0; l < n_levels; ++l) {
        if (lw[l] < 1 || lh[l] < 1) {
            n_levels = l;
            break;
        }
        lvl[l] = (uint8_t*)malloc((size_t)lw[l] * (size_t)lh[l]);
        if (l == 0)
            for (int y = 0; y < H; ++y) memcpy(lvl[0] + (size_t)y * W, gray + (size_t)y * stride, (size_t)W);
        else
            vxo_resize_linear_exact(lvl[l - 1], lw[l - 1], lh[l - 1], (size_t)lw[l - 1], lvl[l], lw[l],
                                    lh[l], (size_t)lw[l]);
    }
    for (int l = 0; l < n_levels; ++l) {
        const int w = lw[l], h = lh[l];
        int n_final = 0;
        vxorb_corner* fin = NULL;
        float* fresp = NULL;
        if (w > 62 && h > 62 && quota[l] > 0) {
            int capc = (w / 2 + 1) * (h / 2 + 1);
            vxorb_corner* cand = (vxorb_corner*)malloc(sizeof(vxorb_corner) * (size_t)capc);
            int nc = vxo_fast9_nms(lvl[l], w, h, (size_t)w, fast_threshold, 31, cand, capc);
            /* retainBest(2*quota) on the FAST score */
            float* keys = (float*)malloc(sizeof(float) * (size_t)(nc > 0 ? nc : 1));
            for (int i = 0; i < nc; ++i) keys[i] = (float)cand[i].score;
            int all;
            float thr = retain_threshold(keys, nc, 2 * quota[l], &all);
            int ns = 0;
            for (int i = 0; i < nc; ++i)
                if (all || keys[i] >= thr) cand[ns++] = cand[i];
            /* Harris, then retainBest(quota) */
            for (int i = 0; i < ns; ++i) keys[i] = vxo_harris(lvl[l], (size_t)w, cand[i].x, cand[i].y);
            thr = retain_threshold(keys, ns, quota[l], &all);
            fin = cand;
            fresp = keys;
            for (int i = 0; i < ns; ++i)
                if (all || keys[i] >= thr) {
                    fin[n_final] = cand[i];
                    fresp[n_final] = keys[i];
                    ++n_final;
                }
        }
        /* orientation on the un-blurred level, then blur, then descriptors */
        if (n_final > 0) {
            uint8_t* bl = (uint8_t*)malloc((size_t)w * (size_t)h);
            vxo_blur7(lvl[l], w, h, (size_t)w, bl, (size_t)w);
            for (int i = 0; i < n_final; ++i) { /* FAST output is row-major => canonical order */
                if (total < cap) {
                    float ang = vxo_ic_angle(lvl[l], (size_t)w, fin[i].x, fin[i].y);
                    vxorb_keypoint* kp = kp_out + total;
                    kp->x = (float)fin[i].x * lscale[l];
                    kp->y = (float)fin[i].y * lscale[l];
                    kp->response = fresp[i];
                    kp->angle = ang;
                    kp->octave = l;
                    vxo_brief(bl, (size_t)w, fin[i].x, fin[i].y, ang, desc_out + (size_t)total * 32);
                }
                ++total;
            }
            free(bl);
        }
        free(fin);
        free(fresp);
    }
    for (int l = 0; l < 16; ++l) free(lvl[l]);
    return total;
}

/* Pyramid level l of a frame (chained resizes), for stage tests.  dst is w_l*h_l tight. */
int vxo_pyramid_level(const uint8_t* gray, int W, int H, size_t stride, float scale_factor,
                      int n_levels, int level, uint8_t* dst) {
    int lw[16], lh[16];
    if (vxo_level_geometry(W, H, 0, scale_factor, n_levels, lw, lh, NULL, NULL)) return -1;
    if (level < 0 || level >= n_levels) return -1;
    uint8_t* prev = (uint8_t*)malloc((size_t)W * (size_t)H);
    for (int y = 0; y < H; ++y) memcpy(prev + (size_t)y * W, gray + (size_t)y * stride, (size_t)W);
    for (int l = 1; l <= level; ++l) {
        uint8_t* cur = (uint8_t*)malloc((size_t)lw[l] * (size_t)lh[l]);
        vxo_resize_linear_exact(prev, lw[l - 1], lh[l - 1], (size_t)lw[l - 1], cur, lw[l], lh[l],
                                (size_t)lw[l]);
        free(prev);
        prev = cur;
    }
    memcpy(dst, prev, (size_t)lw[level] * (size_t)lh[level]);
    free(prev);
    return 0;
}

/* ---------------------------------------------------------------------------------------------
 * A.9  Brute-force Hamming 2-NN (cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, knn, 2),
 * orb_matcher.cpp:22-25), optional square window mask, ties to the lowest train index. */
static inline int hamming256(const uint8_t* a, const uint8_t* b) {
    uint64_t x[4], y[4];
    memcpy(x, a, 32);
    memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
           __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}

void vxo_knn2(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
              float radius, vxorb_knn2_t* out) {
    const int windowed = (qxy && txy && radius > 0.f);
    for (int i = 0; i < nq; ++i) {
        int d0 = 1 << 30, d1 = 1 << 30, i0 = -1, i1 = -1;
        for (int j = 0; j < nt; ++j) {
            if (windowed) {
                if (!(fabsf(qxy[2 * i] - txy[2 * j]) <= radius && fabsf(qxy[2 * i + 1] - txy[2 * j + 1]) <= radius))
                    continue;
            }
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < d0) {
                d1 = d0;
                i1 = i0;
                d0 = d;
                i0 = j;
            } else if (d < d1) {
                d1 = d;
                i1 = j;
            }
        }
        out[i].idx[0] = i0;
        out[i].idx[1] = i1;
        out[i].dist[0] = i0 >= 0 ? d0 : 0;
        out[i].dist[1] = i1 >= 0 ? d1 : 0;
    }
}

/* ORBMatcher::Match (orb_matcher.cpp:11-43): rows with < 2 neighbours skipped (:28), keep iff
 * m1.distance < nn_ratio * m2.distance in float (:33). */
int vxo_match(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
              float radius, float nn_ratio, vxorb_match_t* out, int cap) {
    if (nq <= 0 || nt <= 0) return 0;
    vxorb_knn2_t* knn = (vxorb_knn2_t*)malloc(sizeof(vxorb_knn2_t) * (size_t)nq);
    vxo_knn2(q, qxy, nq, t, txy, nt, radius, knn);
    int n = 0;
    for (int i = 0; i < nq; ++i) {
        if (knn[i].idx[1] < 0) continue;
        float d1 = (float)knn[i].dist[0], d2 = (float)knn[i].dist[1];
        if (d1 < nn_ratio * d2) {
            if (n < cap) {
                out[n].query_idx = i;
                out[n].train_idx = knn[i].idx[0];
                out[n].distance = knn[i].dist[0];
            }
            ++n;
        }
    }
    free(knn);
    return n;
}

/* cv::BFMatcher(NORM_HAMMING, crossCheck=true).match. */
int vxo_match_crosscheck(const uint8_t* q, int nq, const uint8_t* t, int nt, vxorb_match_t* out,
                         int cap) {
    if (nq <= 0 || nt <= 0) return 0;
    int* back = (int*)malloc(sizeof(int) * (size_t)nt);
    for (int j = 0; j < nt; ++j) {
        int best = 1 << 30, bi = -1;
        for (int i = 0; i < nq; ++i) {
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < best) {
                best = d;
                bi = i;
            }
        }
        back[j] = bi;
    }
    int n = 0;
    for (int i = 0; i < nq; ++i) {
        int best = 1 << 30, bj = -1;
        for (int j = 0; j < nt; ++j) {
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < best) {
                best = d;
                bj = j;
            }
        }
        if (bj >= 0 && back[bj] == i) {
            if (n < cap) {
                out[n].query_idx = i;
                out[n].train_idx = bj;
                out[n].distance = best;
            }
            ++n;
        }
    }
    free(back);
    return n;
}

/* Caller's post-filter (tracking.cpp:211-222): min_dist starts at 100.0f. */
int vxo_postfilter(const vxorb_match_t* in, int n, vxorb_match_t* out) {
    float min_dist = 100.0f;
    for (int i = 0; i < n; ++i)
        if ((float)in[i].distance < min_dist) min_dist = (float)in[i].distance;
    float lim = 2.f * min_dist > 30.0f ? 2.f * min_dist : 30.0f;
    int m = 0;
    for (int i = 0; i < n; ++i)
        if ((float)in[i].distance <= lim) out[m++] = in[i];
    return m;
}

/* Box-masked 2-NN + ratio (cv knnMatch with mask[i][j] = box[0] <= xq-xt <= box[1] && box[2] <= yq-yt
 * <= box[3]; BASELINE config 3 left-right matching).  Same tie-break and ratio rule as vxo_match. */
void vxo_knn2_boxed(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
                    const float* box, vxorb_knn2_t* out) {
    for (int i = 0; i < nq; ++i) {
        int d0 = 1 << 30, d1 = 1 << 30, i0 = -1, i1 = -1;
        for (int j = 0; j < nt; ++j) {
            const float dx = qxy[2 * i] - txy[2 * j], dy = qxy[2 * i + 1] - txy[2 * j + 1];
            if (!(dx >= box[0] && dx <= box[1] && dy >= box[2] && dy <= box[3])) continue;
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < d0) {
                d1 = d0; i1 = i0; d0 = d; i0 = j;
            } else if (d < d1) {
                d1 = d; i1 = j;
            }
        }
        out[i].idx[0] = i0;
        out[i].idx[1] = i1;
        out[i].dist[0] = i0 >= 0 ? d0 : 0;
        out[i].dist[1] = i1 >= 0 ? d1 : 0;
    }
}

int vxo_match_boxed(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
                    const float* box, float nn_ratio, vxorb_match_t* out, int cap) {
    if (nq <= 0 || nt <= 0) return 0;
    vxorb_knn2_t* knn = (vxorb_knn2_t*)malloc(sizeof(vxorb_knn2_t) * (size_t)nq);
    vxo_knn2_boxed(q, qxy, nq, t, txy, nt, box, knn);
    int n = 0;
    for (int i = 0; i < nq; ++i) {
        if (knn[i].idx[1] < 0) continue;
        if ((float)knn[i].dist[0] < nn_ratio * (float)knn[i].dist[1]) {
            if (n < cap) {
                out[n].query_idx = i;
                out[n].train_idx = knn[i].idx[0];
                out[n].distance = knn[i].dist[0];
            }
            ++n;
        }
    }
    free(knn);
    return n;
}

/* Arbitrary mask: cv::BFMatcher::knnMatch(q, t, 2, mask) with mask an nq x nt CV_8U matrix, pair (i, j) allowed
 * iff mask[i][j] != 0 (the general form of the window / box masks above; SURVEY.md section 8b vxorb_match_masked). */
void vxo_knn2_masked(const uint8_t* q, int nq, const uint8_t* t, int nt, const uint8_t* mask, size_t mask_stride,
                     vxorb_knn2_t* out) {
    for (int i = 0; i < nq; ++i) {
        int d0 = 1 << 30, d1 = 1 << 30, i0 = -1, i1 = -1;
        for (int j = 0; j < nt; ++j) {
            if (!mask[(size_t)i * mask_stride + j]) continue;
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < d0) {
                d1 = d0; i1 = i0; d0 = d; i0 = j;
            } else if (d < d1) {
                d1 = d; i1 = j;
            }
        }
        out[i].idx[0] = i0;
        out[i].idx[1] = i1;
        out[i].dist[0] = i0 >= 0 ? d0 : 0;
        out[i].dist[1] = i1 >= 0 ? d1 : 0;
    }
}

/* ---- caller-side per-frame quantities --------------------------------- */

/* Depth of every feature as Tracking::CreateLandmarksFromDepth reads it (core/frontend/tracking.cpp:606-637):
 * u = (int)(x + 0.5), v = (int)(y + 0.5) on the double position; skipped when outside the image, when a 16-bit
 * depth is 0, or when the depth in metres (u16 / 5000, float, double) is outside [0.1, 10].  depth_type: 0 = CV_16U,
 * 1 = CV_32F, 2 = CV_64F.  out[i] = depth in metres, 0 where the reference skips the feature. */
void vxo_depth_lookup(const float* xy, int n, const void* depth, int depth_type, int cols, int rows, size_t stride,
                      double* out) {
    for (int i = 0; i < n; ++i) {
        out[i] = 0.0;
        const int u = (int)((double)xy[2 * i] + 0.5), v = (int)((double)xy[2 * i + 1] + 0.5);
        if (u < 0 || u >= cols || v < 0 || v >= rows) continue;
        const uint8_t* row = (const uint8_t*)depth + (size_t)v * stride;
        double m;
        if (depth_type == 0) {
            const uint16_t d = ((const uint16_t*)row)[u];
            if (d == 0) continue;
            m = (double)d / 5000.0;
        } else if (depth_type == 1) {
            m = (double)((const float*)row)[u];
        } else {
            m = ((const double*)row)[u];
        }
        if (m < 0.1 || m > 10.0) continue;
        out[i] = m;
    }
}

/* cv::meanStdDev of an 8-bit single-channel image, as Tracking::CheckImageQuality calls it
 * (core/frontend/tracking.cpp:126-127).  OpenCV accumulates integer sums, then evaluates
 * mean = S * (1/N), sd = sqrt(max(Q * (1/N) - mean^2, 0)) in double (checked against cv2 4.13 by
 * tests/golden/make_golden.py -> caller_kat.npz). */
void vxo_image_stats(const uint8_t* img, int w, int h, size_t stride, double* mean, double* sd) {
    uint64_t s = 0, q = 0;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const uint64_t p = img[(size_t)y * stride + x];
            s += p;
            q += p * p;
        }
    const double scale = 1.0 / (double)((int64_t)w * h);
    const double m = (double)s * scale;
    double var = (double)q * scale - m * m;
    if (var < 0.0) var = 0.0;
    *mean = m;
    *sd = sqrt(var);
}

/* Tracking::CheckFeatureDistribution (core/frontend/tracking.cpp:93-118): occupied cells of a
 * cols x rows grid; positions are the keypoints' float level-0 coordinates widened to double
 * (Feature::position, orb_extractor.cpp:21).  Returns the count; *mask bit = col * rows + row. */
int vxo_grid_coverage(const float* xy, int n, int width, int height, int cols, int rows, uint32_t* mask) {
    uint32_t m = 0;
    for (int i = 0; i < n; ++i) {
        int col = (int)(((double)xy[2 * i] / width) * cols);
        int row = (int)(((double)xy[2 * i + 1] / height) * rows);
        if (col < 0) col = 0;
        if (col > cols - 1) col = cols - 1;
        if (row < 0) row = 0;
        if (row > rows - 1) row = rows - 1;
        m |= 1u << (col * rows + row);
    }
    if (mask) *mask = m;
    int c = 0;
    for (uint32_t t = m; t; t &= t - 1) ++c;
    return c;
}

/* Matched point pairs (core/frontend/tracking.cpp:505-514) and Tracking::ComputeParallax
 * (tracking.cpp:548-560): pts[4i..4i+3] = (x_last, y_last, x_curr, y_curr); returns the mean
 * Euclidean distance in double, summed serially in match order as the reference does. */
double vxo_correspondences(const float* q_xy, const float* t_xy, const vxorb_match_t* m, int n, float* pts) {
    double sum = 0.0;
    for (int i = 0; i < n; ++i) {
        const float ax = q_xy[2 * m[i].query_idx], ay = q_xy[2 * m[i].query_idx + 1];
        const float bx = t_xy[2 * m[i].train_idx], by = t_xy[2 * m[i].train_idx + 1];
        if (pts) {
            pts[4 * i] = ax; pts[4 * i + 1] = ay; pts[4 * i + 2] = bx; pts[4 * i + 3] = by;
        }
        const double dx = (double)ax - (double)bx, dy = (double)ay - (double)by;
        sum += sqrt(dx * dx + dy * dy);
    }
    return n > 0 ? sum / n : 0.0;
}
