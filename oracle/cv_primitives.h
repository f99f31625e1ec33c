/* TEST INFRASTRUCTURE — CPU oracle, never shipped, never on the product path.
 *
 * CPU restatement of the OpenCV arithmetic that the reference's ORB front-end calls.
 * OpenCV is NOT vendored under /root/reference (opencv-2.4.13.zip is a missing blob; any
 * system OpenCV is accepted by cslam/CMakeLists.txt:36-38,62-64), so these functions restate
 * the published OpenCV 4.13.0 algorithms and are PINNED bit-for-bit against the cv2 4.13.0
 * wheel by tests/test_oracle_cv.py and the golden vectors in tests/golden/.
 *
 * Call sites in the reference that these replace:
 *   cv::resize(INTER_LINEAR)            cslam/src/ORBextractor.cpp:1293
 *   cv::copyMakeBorder(REFLECT_101)     cslam/src/ORBextractor.cpp:1295,1300
 *   cv::FAST(img, kps, th, true)        cslam/src/ORBextractor.cpp:978,983
 *   cv::GaussianBlur(7x7, s=2, R101)    cslam/src/ORBextractor.cpp:1259
 *   cv::fastAtan2                       cslam/src/ORBextractor.cpp:94
 *   cv::undistortPoints                 cslam/src/Frame.cpp:295,321
 *   cvRound                             cslam/src/ORBextractor.cpp:72,107,111-124,611,629,1285
 */
#ifndef CCM_ORACLE_CV_PRIMITIVES_H_
#define CCM_ORACLE_CV_PRIMITIVES_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ocv_corner {
    int x, y;   /* pixel position inside the image handed to ocv_fast9_16 */
    int score;  /* cornerScore<16>: largest threshold at which the pixel is still a corner */
} ocv_corner;

/* round-half-to-even, as cvRound (lrint under the default rounding mode) */
int ocv_round_f(float v);
int ocv_round_d(double v);

/* cv::fastAtan2: degrees in [0,360), degree-7 odd polynomial, separately rounded float ops */
float ocv_fast_atan2(float y, float x);

/* cv::resize(..., INTER_LINEAR) for CV_8UC1: 11-bit fixed-point coefficients */
void ocv_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstep,
                          uint8_t* dst, int dw, int dh, size_t dstep);

/* cv::copyMakeBorder(..., BORDER_REFLECT_101) for CV_8UC1 (dst is (w+l+r) x (h+t+b)) */
void ocv_border_reflect101_u8(const uint8_t* src, int w, int h, size_t sstep,
                              uint8_t* dst, size_t dstep, int top, int bottom, int left, int right);

/* cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 (4.x fixed-point path) */
void ocv_gaussian7_sigma2_u8(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep);
/* the same call with OpenCV 2.4.13's taps {18,34,49,55,49,34,18} (restated from knowledge of the 2.4 sources, unverifiable here) */
void ocv24_gaussian7_sigma2_u8(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep);

/* FAST-9/16 corner score of one interior pixel (3 px away from every edge) */
int ocv_fast_score_9_16(const uint8_t* p, size_t step);

/* cv::FAST(img, kps, threshold, nonmax) TYPE_9_16.  Returns the number of corners found (may exceed
 * cap; only the first cap are written), in raster order, exactly as cv::FAST lists them. */
int ocv_fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold, int nonmax,
                 ocv_corner* out, int cap);

/* cv::undistortPoints(src, dst, K, dist, noArray(), P) for CV_32FC2 points, float K/dist/P, default 5 iterations;
 * dist may be NULL (no distortion model). Frame.cpp:295, :321 */
void ocv_undistort_points_f32(const float* src_xy, float* dst_xy, int n, const float K[9], const float* dist, int n_dist,
                              const float P[9]);

#ifdef __cplusplus
}
#endif
#endif
