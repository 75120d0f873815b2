/* orb_oracle.h — declarations of the CPU oracle (TEST INFRASTRUCTURE ONLY; see orb_oracle.c). */
#ifndef VXO_ORB_ORACLE_H_
#define VXO_ORB_ORACLE_H_
#include <stddef.h>
#include <stdint.h>

#include "../include/vxorb.h" /* plain-data structs shared with the boundary */

#ifdef __cplusplus
extern "C" {
#endif
void vxo_bgr2gray(const uint8_t* bgr, int w, int h, size_t stride, uint8_t* gray, size_t gstride);
int vxo_level_geometry(int W, int H, int n_features, float scale_factor, int n_levels, int* lw,
                       int* lh, float* lscale, int* quota);
void vxo_resize_linear_exact(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst,
                             int dw, int dh, size_t dstride);
int vxo_fast9_nms(const uint8_t* img, int w, int h, size_t stride, int threshold, int border,
                  vxorb_corner* out, int cap);
float vxo_harris(const uint8_t* img, size_t stride, int x, int y);
float vxo_fast_atan2(float y, float x);
float vxo_ic_angle(const uint8_t* img, size_t stride, int x, int y);
void vxo_gauss_kernel7(float* k);
void vxo_blur7(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride);
void vxo_brief(const uint8_t* blurred, size_t stride, int cx, int cy, float angle_deg,
               uint8_t* desc);
int vxo_orb_extract(const uint8_t* gray, int W, int H, size_t stride, int n_features,
                    float scale_factor, int n_levels, int fast_threshold, vxorb_keypoint* kp_out,
                    uint8_t* desc_out, int cap);
int vxo_pyramid_level(const uint8_t* gray, int W, int H, size_t stride, float scale_factor,
                      int n_levels, int level, uint8_t* dst);
void vxo_knn2(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
              float radius, vxorb_knn2_t* out);
int vxo_match(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
              float radius, float nn_ratio, vxorb_match_t* out, int cap);
int vxo_match_crosscheck(const uint8_t* q, int nq, const uint8_t* t, int nt, vxorb_match_t* out,
                         int cap);
int vxo_postfilter(const vxorb_match_t* in, int n, vxorb_match_t* out);
void vxo_knn2_boxed(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
                    const float* box, vxorb_knn2_t* out);
int vxo_match_boxed(const uint8_t* q, const float* qxy, int nq, const uint8_t* t, const float* txy, int nt,
                    const float* box, float nn_ratio, vxorb_match_t* out, int cap);
void vxo_knn2_masked(const uint8_t* q, int nq, const uint8_t* t, int nt, const uint8_t* mask, size_t mask_stride,
                     vxorb_knn2_t* out);
void vxo_depth_lookup(const float* xy, int n, const void* depth, int depth_type, int cols, int rows, size_t stride,
                      double* out);
void vxo_image_stats(const uint8_t* img, int w, int h, size_t stride, double* mean, double* sd);
int vxo_grid_coverage(const float* xy, int n, int width, int height, int cols, int rows, uint32_t* mask);
double vxo_correspondences(const float* q_xy, const float* t_xy, const vxorb_match_t* m, int n, float* pts);
#ifdef __cplusplus
}
#endif
#endif
