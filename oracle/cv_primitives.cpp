/* TEST INFRASTRUCTURE — CPU oracle (see cv_primitives.h). Restates OpenCV 4.13.0 arithmetic;
 * pinned bit-for-bit against the cv2 4.13.0 wheel by tests/test_oracle_cv.py.
 * Build with -ffp-contract=off: every float op below must round separately (no FMA). */
#include "cv_primitives.h"

#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" {

int ocv_round_f(float v) { return (int)lrintf(v); }
int ocv_round_d(double v) { return (int)lrint(v); }

/* OpenCV modules/core/src/mathfuncs_core.simd.hpp atan_f32(): p1..p7 scaled to degrees. */
float ocv_fast_atan2(float y, float x) {
    static const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    static const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    static const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    static const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/* OpenCV modules/imgproc/src/resize.cpp, INTER_LINEAR 8U path: per-axis tables of
 * (source index, two short coefficients scaled by 2048), horizontal pass to int32,
 * vertical pass (((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2. */
static void linear_table(int ssize, int dsize, std::vector<int>& ofs, std::vector<short>& coef) {
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1. / inv_scale;
    ofs.resize(dsize);
    coef.resize(2 * (size_t)dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (s < 0) { f = 0; s = 0; }
        if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        ofs[d] = s;
        coef[2 * d] = (short)lrintf((1.f - f) * 2048.f);
        coef[2 * d + 1] = (short)lrintf(f * 2048.f);
    }
}

void ocv_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstep,
                          uint8_t* dst, int dw, int dh, size_t dstep) {
    std::vector<int> xo, yo;
    std::vector<short> xa, ya;
    linear_table(sw, dw, xo, xa);
    linear_table(sh, dh, yo, ya);
    std::vector<int> r0(dw), r1(dw);
    for (int y = 0; y < dh; y++) {
        const int sy0 = yo[y], sy1 = sy0 + 1 < sh ? sy0 + 1 : sh - 1;
        const uint8_t* s0 = src + (size_t)sy0 * sstep;
        const uint8_t* s1 = src + (size_t)sy1 * sstep;
        for (int x = 0; x < dw; x++) {
            const int sx0 = xo[x], sx1 = sx0 + 1 < sw ? sx0 + 1 : sw - 1;
            const int a0 = xa[2 * x], a1 = xa[2 * x + 1];
            r0[x] = s0[sx0] * a0 + s0[sx1] * a1;
            r1[x] = s1[sx0] * a0 + s1[sx1] * a1;
        }
        const int b0 = ya[2 * y], b1 = ya[2 * y + 1];
        uint8_t* d = dst + (size_t)y * dstep;
        for (int x = 0; x < dw; x++)
            d[x] = (uint8_t)((((b0 * (r0[x] >> 4)) >> 16) + ((b1 * (r1[x] >> 4)) >> 16) + 2) >> 2);
    }
}

static inline int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

void ocv_border_reflect101_u8(const uint8_t* src, int w, int h, size_t sstep,
                              uint8_t* dst, size_t dstep, int top, int bottom, int left, int right) {
    const int W = w + left + right, H = h + top + bottom;
    /* interior rows first (src may alias dst's interior, as in ORBextractor.cpp:1295), then borders */
    std::vector<uint8_t> row(w);
    for (int y = 0; y < h; y++) {
        std::memcpy(row.data(), src + (size_t)y * sstep, w);
        uint8_t* d = dst + (size_t)(y + top) * dstep;
        for (int x = 0; x < W; x++) d[x] = row[reflect101(x - left, w)];
    }
    for (int y = 0; y < H; y++) {
        if (y >= top && y < top + h) continue;
        const int sy = reflect101(y - top, h);
        std::memcpy(dst + (size_t)y * dstep, dst + (size_t)(sy + top) * dstep, W);
    }
}

/* OpenCV 4.x GaussianBlur CV_8U fixed-point path (smooth.simd.hpp): 8.8 ufixedpoint16 kernel
 * {18,34,48,56,48,34,18}/256 for ksize 7, sigma 2 (recovered from cv2 and pinned by test),
 * exact integer row+column accumulation, one final round-half-up: (sum + 2^15) >> 16. */
static void gaussian7_fixed(const int* k, const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep);
static int g_blur_variant = 0;                         /* 0: OpenCV 4.13 taps (pinned), 1: OpenCV 2.4.13 taps (see below) */
extern "C" void ocv_set_blur_variant(int v) { g_blur_variant = v; }
void ocv_gaussian7_sigma2_u8(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep) {
    static const int k[7] = {18, 34, 48, 56, 48, 34, 18};
    if (g_blur_variant == 1) { ocv24_gaussian7_sigma2_u8(src, w, h, sstep, dst, dstep); return; }
    gaussian7_fixed(k, src, w, h, sstep, dst, dstep);
}

/* OpenCV 2.4.13 (the version cslam/CMakeLists.txt:36-38,86-90 pins; its source zip is a missing blob, so this is a
 * restatement FROM KNOWLEDGE of the 2.4 sources that no build here can confirm): GaussianBlur on CV_8U goes through
 * createSeparableLinearFilter's integer branch for symmetric smoothing kernels (filter.cpp): getGaussianKernel(7, 2, CV_32F)
 * = {0.07015933, 0.13107488, 0.19071282, 0.21610594, ...}, each tap times 2^8 and rounded on its own (convertTo CV_32S) =
 * {18, 34, 49, 55, 49, 34, 18} - they sum to 257, not 256 - row pass 8U -> 32S exact, column pass 32S -> 8U through
 * FixedPtCastEx<int, uchar>(16): saturate_cast<uchar>((sum + 2^15) >> 16). (Its SSE2 column filter evaluates the same sum
 * in float and rounds to nearest-even: a difference only at exact halves.) Used by tests/test_oracle_cv.py to MEASURE how
 * far a reference built against 2.4.13 is from the 4.13 semantics this repo pins; selectable on the GPU with
 * CCM_ORB_BLUR=opencv2. */
void ocv24_gaussian7_sigma2_u8(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep) {
    static const int k[7] = {18, 34, 49, 55, 49, 34, 18};
    gaussian7_fixed(k, src, w, h, sstep, dst, dstep);
}

static void gaussian7_fixed(const int* k, const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep) {
    std::vector<uint16_t> hbuf((size_t)w * h);          /* row sums <= 255*257 fit u16 */
    std::vector<uint8_t> pad((size_t)w + 6);
    for (int y = 0; y < h; y++) {
        const uint8_t* s = src + (size_t)y * sstep;
        for (int x = -3; x < w + 3; x++) pad[x + 3] = s[reflect101(x, w)];
        uint16_t* hb = &hbuf[(size_t)y * w];
        const uint8_t* p = pad.data();
        for (int x = 0; x < w; x++)
            hb[x] = (uint16_t)(k[0] * (p[x] + p[x + 6]) + k[1] * (p[x + 1] + p[x + 5]) + k[2] * (p[x + 2] + p[x + 4]) + k[3] * p[x + 3]);
    }
    for (int y = 0; y < h; y++) {
        const uint16_t* r[7];
        for (int i = 0; i < 7; i++) r[i] = &hbuf[(size_t)reflect101(y + i - 3, h) * w];
        uint8_t* d = dst + (size_t)y * dstep;
        for (int x = 0; x < w; x++) {
            const int acc = k[0] * (r[0][x] + r[6][x]) + k[1] * (r[1][x] + r[5][x]) + k[2] * (r[2][x] + r[4][x]) + k[3] * r[3][x];
            const int v = (acc + 32768) >> 16;
            d[x] = (uint8_t)(v > 255 ? 255 : v);
        }
    }
}

/* OpenCV modules/features2d/src/fast_score.cpp cornerScore<16>: radius-3 Bresenham ring,
 * score = max over the 16 arcs of 9 contiguous ring pixels of min(|signed difference|) - 1,
 * taken separately for "all darker" and "all brighter" arcs. Independent of the threshold
 * for every pixel that passes the threshold test. */
int ocv_fast_score_9_16(const uint8_t* p, size_t step) {
    const ptrdiff_t s = (ptrdiff_t)step;
    const ptrdiff_t ring[16] = {3 * s,      3 * s + 1,  2 * s + 2,  s + 3,  3,       -s + 3,  -2 * s + 2, -3 * s + 1,
                                -3 * s,     -3 * s - 1, -2 * s - 2, -s - 3, -3,      s - 3,   2 * s - 2,  3 * s - 1};
    int d[16];
    const int v = p[0];
    for (int k = 0; k < 16; k++) d[k] = v - p[ring[k]];
    int best = -1000;
    for (int k = 0; k < 16; k++) {
        int mn = 1000, mx = -1000;
        for (int i = 0; i < 9; i++) {
            const int x = d[(k + i) & 15];
            if (x < mn) mn = x;
            if (x > mx) mx = x;
        }
        if (mn > best) best = mn;
        if (-mx > best) best = -mx;
    }
    return best - 1;
}

/* OpenCV modules/features2d/src/fast.cpp FAST_t<16>: scores exist only for pixels at least 3 px
 * from each edge of THIS image and only if they are corners at THIS threshold (else 0); with
 * nonmax, a corner survives iff its score is strictly greater than all 8 neighbours' entries. */
int ocv_fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold, int nonmax,
                 ocv_corner* out, int cap) {
    if (w < 7 || h < 7) return 0;
    if (threshold < 0) threshold = 0;
    if (threshold > 255) threshold = 255;
    std::vector<int> sc((size_t)w * h, 0);
    std::vector<uint8_t> is((size_t)w * h, 0);
    const ptrdiff_t s_ = (ptrdiff_t)step;
    const ptrdiff_t ring[16] = {3 * s_,  3 * s_ + 1,  2 * s_ + 2,  s_ + 3,  3,  -s_ + 3,  -2 * s_ + 2, -3 * s_ + 1,
                                -3 * s_, -3 * s_ - 1, -2 * s_ - 2, -s_ - 3, -3, s_ - 3,   2 * s_ - 2,  3 * s_ - 1};
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            /* early rejection, as OpenCV's scalar path does (same result: a pixel is a corner at t iff score >= t):
             * any arc of 9 ring pixels holds at least 2 of the 4 compass pixels {0,4,8,12} */
            const uint8_t* p = img + (size_t)y * step + x;
            const int v = p[0], hi = v + threshold, lo = v - threshold;
            const int c0 = p[ring[0]], c4 = p[ring[4]], c8 = p[ring[8]], c12 = p[ring[12]];
            const int nb = (c0 > hi) + (c4 > hi) + (c8 > hi) + (c12 > hi);
            const int nd = (c0 < lo) + (c4 < lo) + (c8 < lo) + (c12 < lo);
            if (nb < 2 && nd < 2) continue;
            unsigned mb = 0, md = 0;
            for (int k = 0; k < 16; k++) { const int q = p[ring[k]]; mb |= (unsigned)(q > hi) << k; md |= (unsigned)(q < lo) << k; }
            bool corner = false;
            for (unsigned m : {mb, md}) {
                unsigned z = m | (m << 16);
                z &= z >> 1; z &= z >> 2; z &= z >> 4; z &= z >> 1;
                if (z) corner = true;
            }
            if (!corner) continue;
            const int s = ocv_fast_score_9_16(p, step);
            if (s >= threshold) { sc[(size_t)y * w + x] = s; is[(size_t)y * w + x] = 1; }
        }
    int n = 0;
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            const size_t i = (size_t)y * w + x;
            if (!is[i]) continue;
            const int s = sc[i];
            if (nonmax) {
                if (!(s > sc[i - 1] && s > sc[i + 1] && s > sc[i - w - 1] && s > sc[i - w] && s > sc[i - w + 1] &&
                      s > sc[i + w - 1] && s > sc[i + w] && s > sc[i + w + 1]))
                    continue;
            }
            if (n < cap) { out[n].x = x; out[n].y = y; out[n].score = nonmax ? s : 0; }
            n++;
        }
    return n;
}

/* cv::undistortPoints(src, dst, K, distCoeffs, noArray(), P) on CV_32FC2 points, OpenCV 4.13.0
 * calib3d/src/undistort.dispatch.cpp cvUndistortPointsInternal with the 4-argument overload's default
 * TermCriteria(MAX_ITER, 5, 0.01): five fixed-point iterations in double, no early exit (the EPS bit is not set).
 * Call sites: Frame::UndistortKeyPoints cslam/src/Frame.cpp:295, Frame::ComputeImageBounds :321 (K = P = mK, float;
 * distCoeffs = (k1,k2,p1,p2[,k3]) float, Tracking.cpp:47-58). k[5..13] (rational, thin-prism, tilt) are 0 for the
 * 4/5-coefficient models the reference loads, so the tilt matrices are the identity; the zero terms are kept where
 * they take part in the arithmetic so the rounding sequence is OpenCV's. Every operation is a separately rounded
 * double operation (this file is built with -ffp-contract=off). */
void ocv_undistort_points_f32(const float* src_xy, float* dst_xy, int n, const float K[9], const float* dist, int n_dist,
                              const float P[9]) {
    double k[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < n_dist && i < 14; i++) k[i] = (double)dist[i];
    double RR[3][3];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) RR[r][c] = (double)P[3 * r + c];          /* PP * I */
    const double fx = (double)K[0], fy = (double)K[4], ifx = 1. / fx, ify = 1. / fy, cx = (double)K[2], cy = (double)K[5];
    for (int i = 0; i < n; i++) {
        double x = (double)src_xy[2 * i], y = (double)src_xy[2 * i + 1];
        const double u = x, v = y;
        x = (x - cx) * ifx;
        y = (y - cy) * ify;
        if (dist) {
            /* invMatTilt = I: vecUntilt = (x*1 + y*0 + 1*0, ..., 1); invProj = 1 */
            const double x0 = x, y0 = y;
            for (int j = 0; j < 5; j++) {
                const double r2 = x * x + y * y;
                const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
                if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
                const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
                const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
                x = (x0 - deltaX) * icdist;
                y = (y0 - deltaY) * icdist;
            }
        }
        const double xx = RR[0][0] * x + RR[0][1] * y + RR[0][2];
        const double yy = RR[1][0] * x + RR[1][1] * y + RR[1][2];
        const double ww = 1. / (RR[2][0] * x + RR[2][1] * y + RR[2][2]);
        dst_xy[2 * i] = (float)(xx * ww);
        dst_xy[2 * i + 1] = (float)(yy * ww);
    }
}

}  // extern "C"
