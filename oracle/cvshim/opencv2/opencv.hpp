/* TEST INFRASTRUCTURE — minimal stand-in for <opencv2/opencv.hpp>, used ONLY to compile the
 * reference's own cslam/src/ORBextractor.cpp unmodified into oracle/_ref (OpenCV C++ headers are
 * not installed here and the pinned opencv-2.4.13.zip is a missing blob). It provides just the
 * cv:: surface that file touches; the arithmetic (FAST, resize, copyMakeBorder, GaussianBlur,
 * fastAtan2, cvRound) forwards to oracle/cv_primitives.cpp, which is pinned against cv2 4.13.0.
 * Only CV_8UC1 matrices are supported. */
#ifndef CCM_ORACLE_CVSHIM_OPENCV_HPP_
#define CCM_ORACLE_CVSHIM_OPENCV_HPP_

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <vector>

#include "cv_primitives.h"

typedef unsigned char uchar;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

static inline int cvRound(double v) { return ocv_round_d(v); }
static inline int cvRound(float v) { return ocv_round_f(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum { INTER_LINEAR = 1 };
enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

/* u8 single-channel matrix with shared ownership and ROI views */
class Mat {
public:
    int rows, cols;
    size_t step;
    uchar* data;

    Mat() : rows(0), cols(0), step(0), data(nullptr) {}
    Mat(Size sz, int /*type*/) { alloc(sz.height, sz.width); }
    Mat(int r, int c, int /*type*/) { alloc(r, c); }
    Mat(int r, int c, int /*type*/, void* ext, size_t _step) : rows(r), cols(c), step(_step), data((uchar*)ext) {}

    /* cv::Mat::zeros returns a MatExpr; assigning it to a Mat of the same shape fills IN PLACE (no
     * reallocation) — computeDescriptors (ORBextractor.cpp:1208) relies on that to write into a row range. */
    struct ZerosExpr { int r, c; };
    static ZerosExpr zeros(int r, int c, int /*type*/) { return ZerosExpr{r, c}; }
    Mat(const ZerosExpr& e) { alloc(e.r, e.c); fill0(); }
    Mat& operator=(const ZerosExpr& e) { create(e.r, e.c, CV_8UC1); fill0(); return *this; }
    void create(int r, int c, int /*type*/) {
        if (data && r == rows && c == cols) return;
        alloc(r, c);
    }
    void release() { buf.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    size_t step1() const { return step; }

    Mat operator()(const Rect& r) const { return view(r.y, r.y + r.height, r.x, r.x + r.width); }
    Mat rowRange(int a, int b) const { return view(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return view(0, rows, a, b); }
    Mat clone() const {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
        return m;
    }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }

private:
    std::shared_ptr<uchar> buf;
    void fill0() { for (int y = 0; y < rows; y++) std::memset(data + (size_t)y * step, 0, cols); }
    void alloc(int r, int c) {
        rows = r; cols = c; step = (size_t)c;
        if (r > 0 && c > 0) {
            buf.reset(new uchar[(size_t)r * c], std::default_delete<uchar[]>());
            data = buf.get();
        } else {
            buf.reset(); data = nullptr;
        }
    }
    Mat view(int y0, int y1, int x0, int x1) const {
        Mat m;
        m.buf = buf; m.rows = y1 - y0; m.cols = x1 - x0; m.step = step;
        m.data = data + (size_t)y0 * step + x0;
        return m;
    }
};

/* proxies: only the uses in ORBextractor::operator() are supported */
class _InputArray {
public:
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& _m) : m(const_cast<Mat*>(&_m)) {}
    bool empty() const { return !m || m->empty(); }
    Mat getMat() const { return m ? *m : Mat(); }
protected:
    Mat* m;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray(Mat& _m) : _InputArray(_m) {}
    void release() const { m->release(); }
    void create(int r, int c, int type) const { m->create(r, c, type); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

/* ---- arithmetic, forwarded to oracle/cv_primitives.cpp ---- */
static inline float fastAtan2(float y, float x) { return ocv_fast_atan2(y, x); }

static inline void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int interpolation) {
    assert(interpolation == INTER_LINEAR);
    (void)interpolation;
    Mat src = _src.getMat();
    Mat dst = _dst.getMat();
    assert(dst.rows == dsize.height && dst.cols == dsize.width);
    ocv_resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dsize.width, dsize.height, dst.step);
}

static inline void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int) {
    Mat src = _src.getMat();
    Mat dst = _dst.getMat();
    assert(dst.rows == src.rows + top + bottom && dst.cols == src.cols + left + right);
    ocv_border_reflect101_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top, bottom, left, right);
}

static inline void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sx, double sy, int border) {
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 && border == BORDER_REFLECT_101);
    (void)ksize; (void)sx; (void)sy; (void)border;
    Mat src = _src.getMat();
    Mat dst = _dst.getMat();
    Mat tmp = src.clone(); /* in-place call at ORBextractor.cpp:1259 */
    ocv_gaussian7_sigma2_u8(tmp.data, tmp.cols, tmp.rows, tmp.step, dst.data, dst.step);
}

static inline void FAST(InputArray _img, std::vector<KeyPoint>& kps, int threshold, bool nonmax) {
    Mat img = _img.getMat();
    kps.clear();
    int cap = img.rows * img.cols;
    if (cap <= 0) return;
    std::vector<ocv_corner> c((size_t)cap);
    int n = ocv_fast9_16(img.data, img.cols, img.rows, img.step, threshold, nonmax ? 1 : 0, c.data(), cap);
    kps.reserve(n);
    for (int i = 0; i < n; i++) kps.push_back(KeyPoint((float)c[i].x, (float)c[i].y, 7.f, -1, (float)c[i].score));
}

/* only referenced from the dead ComputeKeyPointsOld (call commented out at ORBextractor.cpp:1230) */
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& kps, int n) {
        if (n >= 0 && (int)kps.size() > n) {
            std::stable_sort(kps.begin(), kps.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            kps.resize(n);
        }
    }
};

}  // namespace cv
#endif
