/* cvshim.h — the smallest `cv` namespace that lets the UNMODIFIED reference translation unit
 * /root/reference/src/ORBextractor.cc (and include/ORBextractor.h) compile without OpenCV.
 *
 * TEST INFRASTRUCTURE ONLY (oracle/_ref cross-check and the "reference" CPU baseline).  Nothing
 * here is reference code: it is a from-scratch stand-in for the OpenCV types the reference
 * touches (Mat views, KeyPoint, Point_, Size, Rect, Input/OutputArray) whose pixel primitives
 * (resize, GaussianBlur, FAST, fastAtan2, copyMakeBorder) forward to the oracle's models of
 * OpenCV 4.13.0 — the ones tests/golden/cv2_primitives.npz pins against the real cv2 wheel.
 */
#ifndef ORBX_CVSHIM_H
#define ORBX_CVSHIM_H

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "../orb_oracle.h"

typedef unsigned char uchar;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

static inline int cvRound(double v) { return oracle_cv_round(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return (int)std::floor(v); }
static inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

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

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

/* 8-bit single-channel matrix header over a shared buffer; views share the buffer. */
class Mat {
public:
    int rows, cols;
    size_t step;
    uchar* data;
    /* owned buffer = malloc block whose first 64 bytes hold the reference count (plain malloc, so
     * pixel storage never depends on operator new — see ref_driver.cpp's arena) */
    uchar* block;

    Mat() : rows(0), cols(0), step(0), data(nullptr), block(nullptr) {}
    Mat(int r, int c, int type) : rows(0), cols(0), step(0), data(nullptr), block(nullptr) { create(r, c, type); }
    Mat(Size s, int type) : rows(0), cols(0), step(0), data(nullptr), block(nullptr) { create(s.height, s.width, type); }
    /* external memory, not owned */
    Mat(int r, int c, int, void* ext, size_t st) : rows(r), cols(c), step(st), data((uchar*)ext), block(nullptr) {}
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), step(o.step), data(o.data), block(o.block) { retain(); }
    Mat& operator=(const Mat& o)
    {
        if (this != &o) {
            if (o.block) __atomic_add_fetch((int*)o.block, 1, __ATOMIC_RELAXED);
            drop();
            rows = o.rows; cols = o.cols; step = o.step; data = o.data; block = o.block;
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

    void create(int r, int c, int)
    {
        if (data && rows == r && cols == c) return;
        drop();
        rows = r; cols = c; step = (size_t)c;
        block = (uchar*)std::malloc((size_t)r * c + 128);
        *(int*)block = 1;
        data = block + 64;
    }
    void release() { drop(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    size_t step1() const { return step; }
    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step + c); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step + c); }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + a; m.cols = b - a; return m; }
    Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
    Mat clone() const
    {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
        return m;
    }
    /* Mat::zeros yields an expression in OpenCV: assigning it to a matrix that already has the
     * right size fills that matrix IN PLACE (the reference relies on this when it zeroes a row
     * range of the output descriptor matrix, src/ORBextractor.cc:1037). */
    struct ZerosExpr { int rows, cols, type; };
    static ZerosExpr zeros(int r, int c, int t) { return ZerosExpr{r, c, t}; }
    Mat(const ZerosExpr& e) : rows(0), cols(0), step(0), data(nullptr), block(nullptr) { *this = e; }
    Mat& operator=(const ZerosExpr& e)
    {
        create(e.rows, e.cols, e.type);
        for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step, 0, cols);
        return *this;
    }
};

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

inline void resize(const Mat& src, Mat& dst, Size sz, double, double, int)
{
    dst.create(sz.height, sz.width, CV_8UC1);
    oracle_resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, sz.width, sz.height, dst.step);
}

/* dst keeps its buffer when it already has the padded size (true at both reference call sites,
 * where the source is either a separate image or the interior view of dst itself). */
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
    /* scratch comes from malloc, not operator new, so it is reused instead of growing the arena */
    static thread_local float* xyr = nullptr;
    static thread_local int cap = 0;
    kps.clear();
    if (!xyr) { cap = 4096; xyr = (float*)std::malloc(sizeof(float) * 3 * cap); }
    int n = oracle_fast9_16(img.data, img.cols, img.rows, img.step, th, nms ? 1 : 0, xyr, cap);
    if (n > cap) {
        std::free(xyr);
        cap = n; xyr = (float*)std::malloc(sizeof(float) * 3 * cap);
        n = oracle_fast9_16(img.data, img.cols, img.rows, img.step, th, nms ? 1 : 0, xyr, cap);
    }
    kps.reserve(n);
    for (int i = 0; i < n; ++i) kps.push_back(KeyPoint(xyr[3 * i], xyr[3 * i + 1], 7.f, -1.f, xyr[3 * i + 2]));
}

struct KeyPointsFilter {
    /* only referenced by ORBextractor::ComputeKeyPointsOld, which operator() never calls */
    static void retainBest(std::vector<KeyPoint>&, int) { std::abort(); }
};

} // namespace cv
#endif
