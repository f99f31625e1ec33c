/* TEST INFRASTRUCTURE — the few cv:: types cslam/src/ORBmatcher.cpp touches, so that the reference file compiles
 * UNMODIFIED (oracle/Makefile target _ref/libref_match.so). A dense row-major matrix of float (CV_32F) or bytes (CV_8U)
 * with value-semantics sub-views that alias their parent (rowRange/colRange/row/col keep a shared buffer, like cv::Mat),
 * float arithmetic evaluated entry by entry in float with the accumulation order of OpenCV's generic gemm for these tiny
 * sizes (sum over k ascending). Nothing here is shipped. */
#ifndef MATCHSHIM_OPENCV_HPP_
#define MATCHSHIM_OPENCV_HPP_
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

namespace cv {

template <typename T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T a, T b) : x(a), y(b) {} };
typedef Point_<float> Point2f;
typedef Point_<int> Point2i;
typedef Point2i Point;

struct KeyPoint {
    Point2f pt; float size, angle, response; int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1) : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

class Mat {
public:
    int rows, cols, tp;
    size_t step;                 /* bytes between rows */
    unsigned char* data;
    std::shared_ptr<std::vector<unsigned char>> buf;
    Mat() : rows(0), cols(0), tp(CV_8U), step(0), data(nullptr) {}
    Mat(int r, int c, int type) { create(r, c, type); }
    static size_t esz(int type) { return type == CV_32F ? 4 : 1; }
    void create(int r, int c, int type) {
        rows = r; cols = c; tp = type; step = (size_t)c * esz(type);
        buf = std::make_shared<std::vector<unsigned char>>((size_t)r * step, 0);
        data = buf->data();
    }
    void release() { rows = cols = 0; step = 0; data = nullptr; buf.reset(); }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
    static Mat eye(int r, int c, int type) { Mat m(r, c, type); for (int i = 0; i < r && i < c; i++) m.at<float>(i, i) = 1.f; return m; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return tp; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    unsigned char* ptr(int r = 0) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T& at(int r, int c) { return ((T*)(data + (size_t)r * step))[c]; }
    template <typename T> const T& at(int r, int c) const { return ((const T*)(data + (size_t)r * step))[c]; }
    /* single index: element i of a vector (row or column), as cv::Mat::at(int) does */
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    Mat view(int r0, int r1, int c0, int c1) const {
        Mat m; m.rows = r1 - r0; m.cols = c1 - c0; m.tp = tp; m.step = step; m.buf = buf;
        m.data = data + (size_t)r0 * step + (size_t)c0 * esz(tp);
        return m;
    }
    Mat rowRange(int a, int b) const { return view(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return view(0, rows, a, b); }
    Mat row(int r) const { return view(r, r + 1, 0, cols); }
    Mat col(int c) const { return view(0, rows, c, c + 1); }
    Mat clone() const {
        Mat m(rows, cols, tp);
        for (int r = 0; r < rows; r++) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * esz(tp));
        return m;
    }
    void copyTo(Mat& o) const { o = clone(); }
    Mat t() const { Mat m(cols, rows, CV_32F); for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) m.at<float>(c, r) = at<float>(r, c); return m; }
    double dot(const Mat& o) const {      /* cv::Mat::dot accumulates in double */
        double s = 0;
        for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) s += (double)at<float>(r, c) * (double)o.at<float>(r, c);
        return s;
    }
};

inline Mat operator*(const Mat& a, const Mat& b) {
    assert(a.cols == b.rows);
    Mat m(a.rows, b.cols, CV_32F);
    for (int r = 0; r < a.rows; r++)
        for (int c = 0; c < b.cols; c++) {
            float s = 0;
            for (int k = 0; k < a.cols; k++) s += a.at<float>(r, k) * b.at<float>(k, c);
            m.at<float>(r, c) = s;
        }
    return m;
}
inline Mat ew(const Mat& a, const Mat& b, float sb) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; r++) for (int c = 0; c < a.cols; c++) m.at<float>(r, c) = a.at<float>(r, c) + sb * b.at<float>(r, c);
    return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return ew(a, b, 1.f); }
inline Mat operator-(const Mat& a, const Mat& b) { return ew(a, b, -1.f); }
inline Mat scale(const Mat& a, double s) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; r++) for (int c = 0; c < a.cols; c++) m.at<float>(r, c) = (float)(a.at<float>(r, c) * s);
    return m;
}
inline Mat operator-(const Mat& a) { return scale(a, -1.0); }
inline Mat operator*(double s, const Mat& a) { return scale(a, s); }
inline Mat operator*(const Mat& a, double s) { return scale(a, s); }
inline Mat operator/(const Mat& a, double s) { return scale(a, 1.0 / s); }
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }

}  // namespace cv
#endif
