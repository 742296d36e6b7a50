/* cvshim2.h — a typed `cv` namespace large enough for the UNMODIFIED reference translation units
 * src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc, src/MapPoint.cc, src/ORBextractor.cc and the vendored
 * Thirdparty/DBoW2 (FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp, TemplatedVocabulary.h)
 * to compile without OpenCV (oracle/Makefile target `ref2` -> oracle/_ref/libref_matchers.so).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing here is reference code: it is a from-scratch stand-in for the OpenCV
 * types those files touch.  8-bit pixel primitives forward to the oracle's cv2-4.13-pinned models (as cvshim.h
 * does); cv::undistortPoints forwards to the pinned model in frame_oracle.cpp.  Float matrix arithmetic
 * (operator*, +, -, dot, norm) is plain IEEE float / double evaluation in index order: ref_driver2.cpp only feeds
 * the geometric paths operands for which every product and sum is exact (dyadic poses and points), so the
 * result does not depend on OpenCV's accumulation width or FMA use; everything else on these paths (Hamming
 * distances, grid queries, thresholds, ratio tests, histograms, SAD over integer-valued floats) is integer or
 * exactly-representable float work.
 */
#ifndef ORBX_CVSHIM2_H
#define ORBX_CVSHIM2_H

#include <algorithm>
#include <cassert>
#include <cmath>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <memory>
#include <string>
#include <vector>

#include "../orb_oracle.h"
extern "C" void oracle_undistort_points(const float* xy_in, int n, const float* cam, const float* dist, float* xy_out);

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_PI 3.1415926535897932384626433832795
#define CV_CN_SHIFT 3
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_16UC1 CV_MAKETYPE(CV_16U, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

static inline int cvRound(double v) { return oracle_cv_round(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return (int)std::floor(v); }
static inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

[[noreturn]] static inline void shim_abort(const char* what)
{
    std::fprintf(stderr, "cvshim2: %s is not part of the pinned paths\n", what);
    std::abort();
}

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };
enum { NORM_L1 = 2, NORM_L2 = 4 };

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
    template <typename S> Point_& operator*=(S s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};
struct Range { int start, end; Range(int s, int e) : start(s), end(e) {} };

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

static inline int depth_bytes(int type)
{
    switch (type & 7) { case CV_8U: case CV_8S: return 1; case CV_16U: case CV_16S: return 2; case CV_64F: return 8; default: return 4; }
}

/* dense 2-D matrix header over a shared, reference-counted buffer (views share it) */
class Mat {
public:
    int rows, cols;
    int flags;          /* the type code */
    size_t step;        /* bytes per row */
    uchar* data;
    uchar* block;       /* malloc block, first 64 bytes = reference count; nullptr = external / empty */

    Mat() : rows(0), cols(0), flags(0), step(0), data(nullptr), block(nullptr) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(Size s, int type) : Mat() { create(s.height, s.width, type); }
    Mat(int r, int c, int type, void* ext, size_t st = 0) : rows(r), cols(c), flags(type), step(st), data((uchar*)ext), block(nullptr)
    {
        if (!step) step = (size_t)c * elemSize();
    }
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), flags(o.flags), step(o.step), data(o.data), block(o.block) { retain(); }
    Mat& operator=(const Mat& o)
    {
        if (this != &o) {
            if (o.block) __atomic_add_fetch((int*)o.block, 1, __ATOMIC_RELAXED);
            drop();
            rows = o.rows; cols = o.cols; flags = o.flags; step = o.step; data = o.data; block = o.block;
        }
        return *this;
    }
    ~Mat() { drop(); }
    void retain() { if (block) __atomic_add_fetch((int*)block, 1, __ATOMIC_RELAXED); }
    void drop()
    {
        if (block && __atomic_sub_fetch((int*)block, 1, __ATOMIC_ACQ_REL) == 0) std::free(block);
        block = nullptr;
    }

    int type() const { return flags; }
    int depth() const { return flags & 7; }
    int channels() const { return (flags >> CV_CN_SHIFT) + 1; }
    size_t elemSize() const { return (size_t)depth_bytes(flags) * channels(); }
    size_t elemSize1() const { return (size_t)depth_bytes(flags); }
    size_t total() const { return (size_t)rows * cols; }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }
    size_t step1() const { return step / elemSize1(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }

    void create(int r, int c, int type)
    {
        if (data && rows == r && cols == c && flags == type) return;
        drop();
        rows = r; cols = c; flags = type; step = (size_t)c * elemSize();
        block = (uchar*)std::malloc((size_t)r * step + 128);
        *(int*)block = 1;
        data = block + 64;
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    void release() { drop(); rows = cols = 0; step = 0; data = nullptr; }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    /* single index: element i of a vector (row or column), or of a continuous matrix in row-major order */
    template <typename T> T& at(int i) { return const_cast<T&>(static_cast<const Mat*>(this)->at<T>(i)); }
    template <typename T> const T& at(int i) const
    {
        if (rows == 1) return *(const T*)(data + (size_t)i * sizeof(T));
        if (cols == 1) return *(const T*)(data + (size_t)i * step);
        return *(const T*)(data + (size_t)(i / cols) * step + (size_t)(i % cols) * sizeof(T));
    }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }

    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * elemSize(); m.cols = b - a; return m; }
    Mat rowRange(const Range& r) const { return rowRange(r.start, r.end); }
    Mat colRange(const Range& r) const { return colRange(r.start, r.end); }
    Mat row(int r) const { return rowRange(r, r + 1); }
    Mat col(int c) const { return colRange(c, c + 1); }
    Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }

    Mat clone() const { Mat m; copyTo(m); return m; }
    /* like OpenCV: dst is (re)allocated unless it already has this size and type — then its memory is written */
    void copyTo(Mat& dst) const
    {
        if (empty()) { dst.release(); return; }
        if (dst.data == data && dst.step == step && dst.rows == rows && dst.cols == cols) return;
        dst.create(rows, cols, flags);
        const size_t rb = (size_t)cols * elemSize();
        for (int y = 0; y < rows; ++y) std::memcpy(dst.data + (size_t)y * dst.step, data + (size_t)y * step, rb);
    }
    void copyTo(Mat&& view) const
    {
        if (view.rows != rows || view.cols != cols || view.flags != flags) shim_abort("copyTo into a view of another size");
        copyTo(static_cast<Mat&>(view));
    }
    /* saturating / rounding rules only matter for float -> integer, which the pinned paths never do */
    void convertTo(Mat& dst, int rtype, double alpha = 1.0, double beta = 0.0) const
    {
        const int ddepth = rtype & 7;
        Mat out(rows, cols, CV_MAKETYPE(ddepth, channels()));
        const int n = cols * channels();
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < n; ++x) {
                double v;
                switch (depth()) {
                    case CV_8U: v = ptr<uchar>(y)[x]; break;
                    case CV_16U: v = ptr<ushort>(y)[x]; break;
                    case CV_16S: v = ptr<short>(y)[x]; break;
                    case CV_32S: v = ptr<int>(y)[x]; break;
                    case CV_32F: v = ptr<float>(y)[x]; break;
                    case CV_64F: v = ptr<double>(y)[x]; break;
                    default: shim_abort("convertTo from this depth");
                }
                if (ddepth == CV_32F) out.ptr<float>(y)[x] = (alpha == 1.0 && beta == 0.0) ? (float)v : (float)(v * alpha + beta);
                else if (ddepth == CV_64F) out.ptr<double>(y)[x] = v * alpha + beta;
                else shim_abort("convertTo to an integer depth");
            }
        dst = out;
    }
    /* change the channel count of a continuous matrix, rows unchanged (Frame::UndistortKeyPoints) */
    Mat reshape(int cn, int newRows = 0) const
    {
        if (newRows != 0 || !isContinuous()) shim_abort("reshape with a row change / on a view");
        Mat m(*this);
        const int width = cols * channels();
        if (width % cn) shim_abort("reshape: width not divisible");
        m.cols = width / cn;
        m.flags = CV_MAKETYPE(depth(), cn);
        return m;
    }
    Mat t() const
    {
        if (flags != CV_32FC1) shim_abort("t() of a non-float matrix");
        Mat m(cols, rows, flags);
        for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) m.at<float>(x, y) = at<float>(y, x);
        return m;
    }
    Mat inv() const { shim_abort("Mat::inv"); }
    double dot(const Mat& o) const
    {
        if (flags != CV_32FC1 || o.flags != CV_32FC1 || total() != o.total()) shim_abort("dot of non-float / mismatched matrices");
        double s = 0;                                   /* OpenCV accumulates float dot products in double */
        const int n = (int)total();
        for (int i = 0; i < n; ++i) s += (double)at<float>(i) * (double)o.at<float>(i);
        return s;
    }
    Mat& setTo(double v)
    {
        for (int y = 0; y < rows; ++y) for (int x = 0; x < cols * channels(); ++x) {
            if (depth() == CV_32F) ptr<float>(y)[x] = (float)v; else if (depth() == CV_8U) ptr<uchar>(y)[x] = (uchar)v; else shim_abort("setTo depth");
        }
        return *this;
    }
    static Mat zeros(int r, int c, int t) { Mat m(r, c, t); for (int y = 0; y < r; ++y) std::memset(m.ptr(y), 0, (size_t)c * m.elemSize()); return m; }
    static Mat ones(int r, int c, int t) { Mat m = zeros(r, c, t); m.setTo(1.0); return m; }
    static Mat eye(int r, int c, int t)
    {
        Mat m = zeros(r, c, t);
        if (t != CV_32FC1) shim_abort("eye of a non-float type");
        for (int i = 0; i < std::min(r, c); ++i) m.at<float>(i, i) = 1.f;
        return m;
    }
};

/* `(cv::Mat_<float>(3,1) << x, y, z)` */
template <typename T> class Mat_;
template <typename T> struct MatCommaInitializer_ {
    Mat_<T>* m; int i;
    MatCommaInitializer_(Mat_<T>* mm) : m(mm), i(0) {}
    template <typename U> MatCommaInitializer_& operator,(U v);
    operator Mat() const;
};
template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c);
    Mat_(const Mat& o) : Mat(o) {}
    template <typename U> MatCommaInitializer_<T> operator<<(U v) { MatCommaInitializer_<T> ci(this); return (ci, v); }
    T& operator()(int r, int c) { return this->template at<T>(r, c); }
};
template <> inline Mat_<float>::Mat_(int r, int c) : Mat(r, c, CV_32FC1) {}
template <> inline Mat_<double>::Mat_(int r, int c) : Mat(r, c, CV_64FC1) {}
template <> inline Mat_<uchar>::Mat_(int r, int c) : Mat(r, c, CV_8UC1) {}
template <typename T> template <typename U> MatCommaInitializer_<T>& MatCommaInitializer_<T>::operator,(U v)
{
    m->template at<T>(i / m->cols, i % m->cols) = (T)v; ++i; return *this;
}
template <typename T> MatCommaInitializer_<T>::operator Mat() const { return Mat(*m); }

/* ---- float matrix arithmetic, evaluated eagerly in index order (see the header comment) ------------------ */
static inline void need_f32(const Mat& a, const char* what) { if (a.flags != CV_32FC1) shim_abort(what); }
inline Mat operator*(const Mat& a, const Mat& b)
{
    need_f32(a, "operator* on non-float"); need_f32(b, "operator* on non-float");
    if (a.cols != b.rows) shim_abort("operator*: inner dimensions differ");
    Mat d(a.rows, b.cols, CV_32FC1);
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < b.cols; ++j) {
            float s = 0.f;
            for (int k = 0; k < a.cols; ++k) s += a.at<float>(i, k) * b.at<float>(k, j);
            d.at<float>(i, j) = s;
        }
    return d;
}
template <typename F> inline Mat map1(const Mat& a, F f)
{
    need_f32(a, "element-wise operator on non-float");
    Mat d(a.rows, a.cols, CV_32FC1);
    for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) d.at<float>(i, j) = f(a.at<float>(i, j));
    return d;
}
template <typename F> inline Mat map2(const Mat& a, const Mat& b, F f)
{
    need_f32(a, "element-wise operator on non-float"); need_f32(b, "element-wise operator on non-float");
    if (a.rows != b.rows || a.cols != b.cols) shim_abort("element-wise operator: sizes differ");
    Mat d(a.rows, a.cols, CV_32FC1);
    for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) d.at<float>(i, j) = f(a.at<float>(i, j), b.at<float>(i, j));
    return d;
}
/* OpenCV scales float matrices by a DOUBLE factor: (float)(x * (double)s) */
inline Mat operator*(const Mat& a, double s) { return map1(a, [s](float x) { return (float)((double)x * s); }); }
inline Mat operator*(double s, const Mat& a) { return a * s; }
inline Mat operator/(const Mat& a, double s) { return map1(a, [s](float x) { return (float)((double)x * (1.0 / s)); }); }
inline Mat operator-(const Mat& a) { return map1(a, [](float x) { return -x; }); }
inline Mat operator+(const Mat& a, const Mat& b) { return map2(a, b, [](float x, float y) { return x + y; }); }
inline Mat operator-(const Mat& a, const Mat& b) { return map2(a, b, [](float x, float y) { return x - y; }); }

inline double norm(const Mat& a, int type = NORM_L2)
{
    need_f32(a, "norm of non-float");
    double s = 0;
    for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) {
        const double v = a.at<float>(i, j);
        s += type == NORM_L1 ? std::fabs(v) : v * v;
    }
    return type == NORM_L1 ? s : std::sqrt(s);
}
inline double norm(const Mat& a, const Mat& b, int type = NORM_L2)
{
    need_f32(a, "norm of non-float"); need_f32(b, "norm of non-float");
    double s = 0;
    for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) {
        const double v = (double)a.at<float>(i, j) - (double)b.at<float>(i, j);
        s += type == NORM_L1 ? std::fabs(v) : v * v;
    }
    return type == NORM_L1 ? s : std::sqrt(s);
}

class _InputArray {
public:
    Mat* m;
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& mm) : m(const_cast<Mat*>(&mm)) {}
    bool empty() const { return !m || m->empty(); }
    Mat getMat() const { return m ? *m : Mat(); }
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& mm) { m = &mm; }
    void create(int r, int c, int t) const { m->create(r, c, t); }
    void release() const { if (m) m->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline float fastAtan2(float y, float x) { return oracle_fast_atan2(y, x); }

/* ---- 8-bit pixel primitives (ORBextractor.cc): the cv2-pinned models ------------------------------------ */
inline void resize(const Mat& src, Mat& dst, Size sz, double, double, int)
{
    dst.create(sz.height, sz.width, CV_8UC1);
    oracle_resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, sz.width, sz.height, dst.step);
}
inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int)
{
    assert(top == bottom && top == left && top == right);
    dst.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
    if (src.data == dst.data + (size_t)top * dst.step + left) {
        Mat tmp = src.clone();
        oracle_copy_make_border101(tmp.data, tmp.cols, tmp.rows, tmp.step, dst.data, dst.step, top);
    } else
        oracle_copy_make_border101(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top);
}
inline void GaussianBlur(const Mat& src, Mat& dst, Size k, double sx, double sy, int)
{
    assert(k.width == 7 && k.height == 7 && sx == 2 && sy == 2);
    Mat in = (src.data == dst.data) ? src.clone() : src;
    dst.create(src.rows, src.cols, CV_8UC1);
    oracle_gaussian_blur7_u8(in.data, in.cols, in.rows, in.step, dst.data, dst.step);
}
inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int th, bool nms)
{
    std::vector<float> xyr(3 * 4096);
    kps.clear();
    int n = oracle_fast9_16(img.data, img.cols, img.rows, img.step, th, nms ? 1 : 0, xyr.data(), (int)xyr.size() / 3);
    if (n > (int)xyr.size() / 3) { xyr.resize((size_t)3 * n); n = oracle_fast9_16(img.data, img.cols, img.rows, img.step, th, nms ? 1 : 0, xyr.data(), n); }
    kps.reserve(n);
    for (int i = 0; i < n; ++i) kps.push_back(KeyPoint(xyr[3 * i], xyr[3 * i + 1], 7.f, -1.f, xyr[3 * i + 2]));
}
struct KeyPointsFilter { static void retainBest(std::vector<KeyPoint>&, int) { shim_abort("KeyPointsFilter::retainBest"); } };

/* cv::undistortPoints(mat, mat, K, D, Mat(), K) on an N x 1 two-channel float matrix (Frame.cc:422, :448) */
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P)
{
    if (src.flags != CV_32FC2 || !src.isContinuous() || !R.empty() || P.data != K.data) shim_abort("undistortPoints in another form than Frame.cc uses");
    const float cam[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    float dist[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < (int)D.total() && i < 5; ++i) dist[i] = D.at<float>(i);
    const int n = (int)src.total();
    std::vector<float> out((size_t)2 * n);
    oracle_undistort_points(src.ptr<float>(), n, cam, dist, out.data());
    dst.create(src.rows, src.cols, CV_32FC2);
    std::memcpy(dst.data, out.data(), out.size() * sizeof(float));
}

/* YAML storage: only named by the virtual save/load of DBoW2::TemplatedVocabulary; never called */
class FileNode {
public:
    FileNode operator[](const char*) const { shim_abort("FileNode"); }
    FileNode operator[](const std::string&) const { shim_abort("FileNode"); }
    FileNode operator[](int) const { shim_abort("FileNode"); }
    size_t size() const { shim_abort("FileNode"); }
    operator int() const { shim_abort("FileNode"); }
    operator float() const { shim_abort("FileNode"); }
    operator double() const { shim_abort("FileNode"); }
    operator std::string() const { shim_abort("FileNode"); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const char*, int) {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { shim_abort("FileStorage"); }
    FileNode operator[](const std::string&) const { shim_abort("FileStorage"); }
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { shim_abort("FileStorage <<"); return fs; }

} // namespace cv
#endif
