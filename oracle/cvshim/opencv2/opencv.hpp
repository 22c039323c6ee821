/*
 * oracle/cvshim/opencv2/opencv.hpp — TEST INFRASTRUCTURE ONLY.
 *
 * A minimal stand-in for the slice of the OpenCV C++ API that the reference's hot-path sources use, so
 * that the UNMODIFIED files under /root/reference/src (ORBextractor.cc) compile with g++ in an image
 * that has no OpenCV C++ headers or libraries (only the cv2 Python wheel).  The container types (Mat
 * with ROI views, KeyPoint, Point_, Size, Rect, Range, InputArray/OutputArray) are written here; the
 * pixel primitives the reference calls (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2, cvRound)
 * forward to oracle/orc_prims.c, which is pinned bit-exact against the real OpenCV 4.13
 * (tests/test_oracle_vs_cv2.py, tests/golden/).
 *
 * Built only by oracle/build_ref.sh into oracle/_ref/; nothing in the product links or includes it.
 * Semantics follow the OpenCV 4.x documentation of each call; only what the reference reaches is
 * implemented, anything else aborts loudly (cvshim_fail).
 */
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../orc.h"

#define CV_8U 0
#define CV_8UC1 0
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_32FC1 5
#define CV_PI 3.1415926535897932384626433832795

typedef unsigned char uchar;

#include <stdexcept>
#include <string>
#include <climits>
#include <sstream>      /* the real core.hpp pulls it in; the vendored DBoW2 relies on that */
namespace cv {

typedef ::uchar uchar;

[[noreturn]] inline void cvshim_fail(const char* what)
{
    std::fprintf(stderr, "cvshim: unsupported use: %s\n", what);
    std::abort();
}

/* cvRound: round half to even (SSE cvtss2si / cvtsd2si in OpenCV builds for x86). */
inline int cvRound(float v) { return (int)lrintf(v); }
inline int cvRound(double v) { return (int)lrint(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }
inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

template <typename T> inline T saturate_cast(float v) { return (T)v; }
template <> inline int saturate_cast<int>(float v) { return cvRound(v); }
template <typename T> inline T saturate_cast(double v) { return (T)v; }
template <> inline int saturate_cast<int>(double v) { return cvRound(v); }
template <typename T> inline T saturate_cast(int v) { return (T)v; }

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U>& p) : x(saturate_cast<T>(p.x)), y(saturate_cast<T>(p.y)) {}
};
/* core/types.hpp: a *= b  ->  saturate_cast<T>(a.x * b) per component */
template <typename T> inline Point_<T>& operator*=(Point_<T>& a, float b)
{
    a.x = saturate_cast<T>(a.x * b);
    a.y = saturate_cast<T>(a.y * b);
    return a;
}
template <typename T> inline Point_<T>& operator*=(Point_<T>& a, double b)
{
    a.x = saturate_cast<T>(a.x * b);
    a.y = saturate_cast<T>(a.y * b);
    return a;
}
template <typename T> inline Point_<T>& operator*=(Point_<T>& a, int b)
{
    a.x = saturate_cast<T>(a.x * b);
    a.y = saturate_cast<T>(a.y * b);
    return a;
}
template <typename T> inline Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x + b.x, a.y + b.y); }
template <typename T> inline Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x - b.x, a.y - b.y); }
template <typename T> inline bool operator==(const Point_<T>& a, const Point_<T>& b) { return a.x == b.x && a.y == b.y; }
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;

template <typename T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
};
template <typename T> inline bool operator==(const Size_<T>& a, const Size_<T>& b) { return a.width == b.width && a.height == b.height; }
typedef Size_<int> Size;

template <typename T> struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T _x, T _y, T w, T h) : x(_x), y(_y), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Range {
    int start, end;
    Range() : start(0), end(0) {}
    Range(int s, int e) : start(s), end(e) {}
};

/* cv::KeyPoint: 28 bytes, the layout the product's ABI mirrors (include/orbx.h). */
struct KeyPoint {
    Point2f pt;
    float size;
    float angle;
    float response;
    int octave;
    int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f _pt, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(_pt), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

enum BorderTypes { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
                   BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };
enum NormTypes { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4, NORM_HAMMING = 6 };

inline int cvshim_elem_size(int type)
{
    switch (type) {
    case CV_8U: return 1;
    case CV_32S: case CV_32F: return 4;
    case CV_64F: return 8;
    default: cvshim_fail("Mat type");
    }
}

struct MatStep {
    size_t v;
    MatStep() : v(0) {}
    operator size_t() const { return v; }
    MatStep& operator=(size_t s) { v = s; return *this; }
};

class Mat;
class _InputArray;
class _OutputArray;
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

/* Reference-counted 2-D buffer with ROI views (the cv::Mat subset the reference touches). */
class Mat {
public:
    int rows, cols;
    uchar* data;
    MatStep step;

    Mat() : rows(0), cols(0), data(nullptr), type_(CV_8U), ref_(nullptr), datastart_(nullptr), wrows_(0), wcols_(0) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
    /* user-owned data (no reference counting), as cv::Mat(rows, cols, type, void* data, step) */
    Mat(int r, int c, int type, void* ext, size_t st = 0) : Mat()
    {
        rows = r; cols = c; type_ = type; data = (uchar*)ext;
        step = st ? st : (size_t)c * cvshim_elem_size(type);
        datastart_ = data; wrows_ = r; wcols_ = c;
    }
    Mat(const Mat& m) : rows(m.rows), cols(m.cols), data(m.data), step(m.step), type_(m.type_), ref_(m.ref_),
                        datastart_(m.datastart_), wrows_(m.wrows_), wcols_(m.wcols_)
    {
        if (ref_) ++*ref_;
    }
    Mat& operator=(const Mat& m)
    {
        if (this != &m) {
            if (m.ref_) ++*m.ref_;
            release();
            rows = m.rows; cols = m.cols; data = m.data; step = m.step; type_ = m.type_; ref_ = m.ref_;
            datastart_ = m.datastart_; wrows_ = m.wrows_; wcols_ = m.wcols_;
        }
        return *this;
    }
    ~Mat() { release(); }

    /* pixel buffers come from malloc, never operator new: the creation-order build of oracle/_ref replaces
     * operator new with a per-call arena (oracle/ref_wrap.cc) and mvImagePyramid outlives the call */
    void release()
    {
        if (ref_ && --*ref_ == 0) {
            std::free(datastart_);
            std::free(ref_);
        }
        rows = cols = 0; data = nullptr; ref_ = nullptr; datastart_ = nullptr; wrows_ = wcols_ = 0; step = 0;
    }
    /* cv::Mat::create: keeps the buffer when shape and type already match (also for ROI views) */
    void create(int r, int c, int type)
    {
        if (data && rows == r && cols == c && type_ == type) return;
        release();
        if (r <= 0 || c <= 0) { type_ = type; return; }
        rows = r; cols = c; type_ = type;
        step = (size_t)c * cvshim_elem_size(type);
        datastart_ = (uchar*)std::malloc((size_t)r * step + 64);
        ref_ = (int*)std::malloc(sizeof(int));
        *ref_ = 1;
        data = datastart_; wrows_ = r; wcols_ = c;
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }

    int type() const { return type_; }
    int depth() const { return type_; }
    int channels() const { return 1; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t elemSize() const { return (size_t)cvshim_elem_size(type_); }
    size_t step1() const { return step.v / elemSize(); }
    Size size() const { return Size(cols, rows); }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows <= 1 || step.v == (size_t)cols * elemSize(); }
    bool isSubmatrix() const { return rows != wrows_ || cols != wcols_; }
    /* position of this view inside the allocation it was cut from */
    void locateROI(Size& whole, Point& ofs) const
    {
        whole = Size(wcols_, wrows_);
        size_t d = (size_t)(data - datastart_);
        ofs.y = step.v ? (int)(d / step.v) : 0;
        ofs.x = step.v ? (int)((d - (size_t)ofs.y * step.v) / elemSize()) : 0;
    }

    Mat operator()(const Rect& r) const
    {
        if (r.x < 0 || r.y < 0 || r.x + r.width > cols || r.y + r.height > rows) cvshim_fail("Mat ROI out of range");
        Mat m(*this);
        m.data = data + (size_t)r.y * step.v + (size_t)r.x * elemSize();
        m.rows = r.height; m.cols = r.width;
        return m;
    }
    Mat operator()(const Range& rr, const Range& cr) const { return (*this)(Rect(cr.start, rr.start, cr.end - cr.start, rr.end - rr.start)); }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat rowRange(const Range& r) const { return rowRange(r.start, r.end); }
    Mat colRange(const Range& r) const { return colRange(r.start, r.end); }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }

    Mat clone() const
    {
        Mat m;
        copyToMat(m);
        return m;
    }
    void copyToMat(Mat& dst) const
    {
        dst.create(rows, cols, type_);
        const size_t w = (size_t)cols * elemSize();
        for (int y = 0; y < rows; y++) std::memmove(dst.data + (size_t)y * dst.step.v, data + (size_t)y * step.v, w);
    }
    inline void copyTo(OutputArray dst) const;

    uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.v); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.v); }
    template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step.v))[x]; }
    template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step.v))[x]; }

    /* single-index access on a row or column vector */
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }

    /* ---- CV_32F matrix algebra (only what ORBmatcher.cc's pose arithmetic needs) ---- */
    Mat t() const
    {
        if (type_ != CV_32F) cvshim_fail("Mat::t type");
        Mat m(cols, rows, CV_32F);
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) m.at<float>(x, y) = at<float>(y, x);
        return m;
    }
    /* cv::Mat::dot: double accumulation over all elements */
    double dot(const Mat& b) const
    {
        if (type_ != CV_32F || b.type_ != CV_32F || rows != b.rows || cols != b.cols) cvshim_fail("Mat::dot");
        double s = 0;
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) s += (double)at<float>(y, x) * b.at<float>(y, x);
        return s;
    }
    static Mat eye(int r, int c, int type)
    {
        Mat m = zeros(r, c, type);
        if (type != CV_32F) cvshim_fail("Mat::eye type");
        for (int i = 0; i < (r < c ? r : c); i++) m.at<float>(i, i) = 1.f;
        return m;
    }

    static Mat zeros(int r, int c, int type)
    {
        Mat m(r, c, type);
        for (int y = 0; y < r; y++) std::memset(m.ptr(y), 0, (size_t)c * m.elemSize());
        return m;
    }

    /* Frame::UndistortKeyPoints reshapes its N x 2 point list to N x 1 two-channel and back (src/Frame.cc:950-953): the
     * stand-in keeps point lists as N x 2 one-channel throughout (undistortPoints below takes them that way) */
    Mat reshape(int) const { return *this; }

private:
    int type_;
    int* ref_;
    uchar* datastart_;
    int wrows_, wcols_;
};

/* CV_32F matrix product: cv::gemm accumulates every output element in DOUBLE for float inputs of these small sizes
 * (core/src/matmul.simd.hpp GEMMSingleMul<float,double>) and rounds once. */
inline Mat operator*(const Mat& a, const Mat& b)
{
    if (a.type() != CV_32F || b.type() != CV_32F || a.cols != b.rows) cvshim_fail("Mat * Mat");
    Mat m(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < b.cols; x++) {
            double s = 0;
            for (int k = 0; k < a.cols; k++) s += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
            m.at<float>(y, x) = (float)s;
        }
    return m;
}
inline Mat cvshim_ew(const Mat& a, const Mat& b, int sign)
{
    if (a.type() != CV_32F || b.type() != CV_32F || a.rows != b.rows || a.cols != b.cols) cvshim_fail("Mat +- Mat");
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = sign > 0 ? a.at<float>(y, x) + b.at<float>(y, x) : a.at<float>(y, x) - b.at<float>(y, x);
    return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return cvshim_ew(a, b, +1); }
inline Mat operator-(const Mat& a, const Mat& b) { return cvshim_ew(a, b, -1); }
inline Mat operator-(const Mat& a)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = -a.at<float>(y, x);
    return m;
}
inline Mat operator*(const Mat& a, double s)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = (float)(a.at<float>(y, x) * s);
    return m;
}
inline Mat operator*(double s, const Mat& a) { return a * s; }
inline Mat operator/(const Mat& a, double s)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = (float)(a.at<float>(y, x) / s);
    return m;
}
/* cv::norm(Mat) on CV_32F: L2 with double accumulation */
inline double norm(const Mat& a)
{
    if (a.type() != CV_32F) cvshim_fail("norm(a): only CV_32F");
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) s += (double)a.at<float>(y, x) * (double)a.at<float>(y, x);
    return std::sqrt(s);
}

/* Fixed-size matrices: cv::Matx arithmetic is plain float (core/matx.hpp accumulates in _Tp). */
template <typename T, int M, int N> struct Matx {
    T val[M * N];
    Matx() { for (int i = 0; i < M * N; i++) val[i] = T(0); }
    Matx(T a, T b, T c) { static_assert(M * N == 3, "Matx3"); val[0] = a; val[1] = b; val[2] = c; }
    Matx(T a, T b, T c, T d, T e, T f, T g, T h, T i)
    {
        static_assert(M * N == 9, "Matx9");
        val[0] = a; val[1] = b; val[2] = c; val[3] = d; val[4] = e; val[5] = f; val[6] = g; val[7] = h; val[8] = i;
    }
    Matx(T a, T b, T c, T d, T e, T f, T g, T h, T i, T j, T k, T l)
    {
        static_assert(M * N == 12, "Matx12");
        val[0] = a; val[1] = b; val[2] = c; val[3] = d; val[4] = e; val[5] = f; val[6] = g; val[7] = h; val[8] = i; val[9] = j; val[10] = k; val[11] = l;
    }
    static Matx zeros() { return Matx(); }
    template <int M1, int N1> Matx<T, M1, N1> get_minor(int i0, int j0) const
    {
        Matx<T, M1, N1> r;
        for (int i = 0; i < M1; i++)
            for (int j = 0; j < N1; j++) r(i, j) = (*this)(i0 + i, j0 + j);
        return r;
    }
    T& operator()(int i, int j) { return val[i * N + j]; }
    const T& operator()(int i, int j) const { return val[i * N + j]; }
    T& operator()(int i) { return val[i]; }
    const T& operator()(int i) const { return val[i]; }
    Matx<T, N, M> t() const
    {
        Matx<T, N, M> r;
        for (int i = 0; i < M; i++)
            for (int j = 0; j < N; j++) r(j, i) = (*this)(i, j);
        return r;
    }
    T dot(const Matx<T, M, N>& b) const
    {
        T s = 0;
        for (int i = 0; i < M * N; i++) s += val[i] * b.val[i];
        return s;
    }
};
template <typename T, int M, int N, int L> inline Matx<T, M, L> operator*(const Matx<T, M, N>& a, const Matx<T, N, L>& b)
{
    Matx<T, M, L> r;
    for (int i = 0; i < M; i++)
        for (int j = 0; j < L; j++) {
            T s = 0;
            for (int k = 0; k < N; k++) s += a(i, k) * b(k, j);
            r(i, j) = s;
        }
    return r;
}
template <typename T, int M, int N> inline Matx<T, M, N> operator+(const Matx<T, M, N>& a, const Matx<T, M, N>& b)
{
    Matx<T, M, N> r;
    for (int i = 0; i < M * N; i++) r.val[i] = a.val[i] + b.val[i];
    return r;
}
template <typename T, int M, int N> inline Matx<T, M, N> operator-(const Matx<T, M, N>& a, const Matx<T, M, N>& b)
{
    Matx<T, M, N> r;
    for (int i = 0; i < M * N; i++) r.val[i] = a.val[i] - b.val[i];
    return r;
}
template <typename T, int M, int N> inline Matx<T, M, N> operator-(const Matx<T, M, N>& a)
{
    Matx<T, M, N> r;
    for (int i = 0; i < M * N; i++) r.val[i] = -a.val[i];
    return r;
}
template <typename T, int M, int N> inline Matx<T, M, N> operator*(const Matx<T, M, N>& a, float s)
{
    Matx<T, M, N> r;
    for (int i = 0; i < M * N; i++) r.val[i] = a.val[i] * s;
    return r;
}
template <typename T, int M, int N> inline Matx<T, M, N> operator*(float s, const Matx<T, M, N>& a) { return a * s; }
typedef Matx<float, 3, 3> Matx33f;
typedef Matx<float, 3, 1> Matx31f;
typedef Matx<float, 4, 4> Matx44f;
typedef Matx<float, 3, 4> Matx34f;
/* cv::norm(Matx): sqrt of the sum of squares accumulated in double (normL2Sqr<_Tp, double>) */
template <typename T, int M, int N> inline double norm(const Matx<T, M, N>& a)
{
    double s = 0;
    for (int i = 0; i < M * N; i++) s += (double)a.val[i] * (double)a.val[i];
    return std::sqrt(s);
}

class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
protected:
    Mat* m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& m) { m_ = &m; }
    _OutputArray(const Mat& m) { m_ = const_cast<Mat*>(&m); }
    void release() const { if (m_) m_->release(); }
    void create(int r, int c, int type) const { if (m_) m_->create(r, c, type); }
    void create(Size sz, int type) const { if (m_) m_->create(sz, type); }
    Mat& getMatRef() const { return *m_; }
};
inline void Mat::copyTo(OutputArray dst) const { copyToMat(dst.getMatRef()); }
inline const _InputArray& noArray() { static _InputArray n; return n; }

/* ---- pixel primitives: forward to the cv2-pinned restatements in oracle/orc_prims.c ---- */

inline void resize(InputArray _src, OutputArray _dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR)
{
    Mat src = _src.getMat();
    if (src.type() != CV_8UC1 || interpolation != INTER_LINEAR || fx != 0 || fy != 0) cvshim_fail("resize: only u8 INTER_LINEAR with dsize");
    _dst.create(dsize, src.type());
    Mat dst = _dst.getMat();
    orc_resize_linear_u8(src.data, src.rows, src.cols, src.step, dst.data, dst.rows, dst.cols, dst.step);
}

/* copyMakeBorder, BORDER_REFLECT_101 with or without BORDER_ISOLATED.  Without ISOLATED a source that is a
 * view takes real pixels from its parent where they exist (core/src/copy.cpp: locateROI + border shrink). */
inline void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int borderType)
{
    Mat src = _src.getMat();
    if (src.type() != CV_8UC1 || (borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101) cvshim_fail("copyMakeBorder: only u8 REFLECT_101");
    int dtop = 0, dbottom = 0, dleft = 0, dright = 0;
    if (src.isSubmatrix() && !(borderType & BORDER_ISOLATED)) {
        Size whole; Point ofs;
        src.locateROI(whole, ofs);
        dtop = std::min(ofs.y, top);
        dbottom = std::min(whole.height - src.rows - ofs.y, bottom);
        dleft = std::min(ofs.x, left);
        dright = std::min(whole.width - src.cols - ofs.x, right);
    }
    _dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
    Mat dst = _dst.getMat();
    /* effective source = the view grown by the real pixels around it */
    const uchar* es = src.data - (size_t)dtop * src.step - dleft;
    const int er = src.rows + dtop + dbottom, ec = src.cols + dleft + dright;
    const int etop = top - dtop, eleft = left - dleft;
    if (er < 2 || ec < 2) cvshim_fail("copyMakeBorder: image too small");
    std::vector<int> xmap(dst.cols);
    for (int x = 0; x < dst.cols; x++) {
        int sx = x - eleft;
        while (sx < 0 || sx >= ec) sx = sx < 0 ? -sx : 2 * (ec - 1) - sx;
        xmap[x] = sx;
    }
    /* interior rows first (memmove: src may live inside dst), then the mirrored columns and rows from dst itself */
    for (int y = 0; y < er; y++) {
        const uchar* s = es + (size_t)y * src.step;
        uchar* d = dst.ptr(y + etop);
        if (d + eleft != s) std::memmove(d + eleft, s, (size_t)ec);
    }
    for (int y = 0; y < er; y++) {
        uchar* d = dst.ptr(y + etop);
        for (int x = 0; x < eleft; x++) d[x] = d[eleft + xmap[x]];
        for (int x = eleft + ec; x < dst.cols; x++) d[x] = d[eleft + xmap[x]];
    }
    for (int y = 0; y < dst.rows; y++) {
        int sy = y - etop;
        if (sy >= 0 && sy < er) continue;
        while (sy < 0 || sy >= er) sy = sy < 0 ? -sy : 2 * (er - 1) - sy;
        std::memcpy(dst.ptr(y), dst.ptr(sy + etop), (size_t)dst.cols);
    }
}

/* cv::FAST(image, keypoints, threshold, nonmaxSuppression) — TYPE_9_16; KeyPoint(x, y, 7.f, -1, score) */
inline void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true)
{
    Mat img = _img.getMat();
    if (img.type() != CV_8UC1) cvshim_fail("FAST: only u8");
    keypoints.clear();
    if (img.rows < 7 || img.cols < 7) return;
    /* scratch from malloc, not operator new (see Mat::release) */
    static thread_local orc_cand* buf = nullptr;
    static thread_local int bufcap = 0;
    const int cap = img.rows * img.cols;
    if (bufcap < cap) { std::free(buf); buf = (orc_cand*)std::malloc(sizeof(orc_cand) * (size_t)cap); bufcap = cap; }
    const int n = orc_fast9_16(img.data, img.rows, img.cols, img.step, threshold, nonmaxSuppression ? 1 : 0, buf, cap);
    keypoints.reserve(n);
    for (int i = 0; i < n; i++) keypoints.push_back(KeyPoint((float)buf[i].x, (float)buf[i].y, 7.f, -1, (float)buf[i].score));
}

/* cv::GaussianBlur — only the call the reference makes: 7x7, sigma 2, u8, REFLECT_101, on a standalone Mat */
inline void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT)
{
    Mat src = _src.getMat();
    if (src.type() != CV_8UC1 || ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || sigmaY != 2 ||
        (borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101)
        cvshim_fail("GaussianBlur: only u8 7x7 sigma 2 REFLECT_101");
    if (src.isSubmatrix() && !(borderType & BORDER_ISOLATED)) cvshim_fail("GaussianBlur on a view without BORDER_ISOLATED");
    _dst.create(src.rows, src.cols, src.type());
    Mat dst = _dst.getMat();
    if (dst.data == src.data) {
        Mat tmp = src.clone();
        orc_gaussian7x7_s2(tmp.data, tmp.rows, tmp.cols, tmp.step, dst.data, dst.step);
    } else {
        orc_gaussian7x7_s2(src.data, src.rows, src.cols, src.step, dst.data, dst.step);
    }
}

/* features2d KeyPointsFilter::retainBest (keypoint.cpp): keep the n strongest plus everything tied with
 * the n-th.  Reached only from the reference's dead ComputeKeyPointsOld; present so that the file links. */
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& keypoints, int n_points)
    {
        if (n_points >= 0 && keypoints.size() > (size_t)n_points) {
            if (n_points == 0) { keypoints.clear(); return; }
            std::nth_element(keypoints.begin(), keypoints.begin() + n_points - 1, keypoints.end(),
                             [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            const float ambiguous = keypoints[n_points - 1].response;
            auto new_end = std::partition(keypoints.begin() + n_points, keypoints.end(),
                                          [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
            keypoints.resize(new_end - keypoints.begin());
        }
    }
};

/* ---- what src/Frame.cc needs on top of the extractor / matcher subset ---- */

/* cv::norm(a, b, NORM_L1) on CV_8U windows (the SAD of Frame::ComputeStereoMatches, src/Frame.cc:1197): integer sum, returned as double */
inline double norm(const Mat& a, const Mat& b, int normType)
{
    if (normType != NORM_L1 || a.type() != CV_8U || b.type() != CV_8U || a.rows != b.rows || a.cols != b.cols) cvshim_fail("norm(a, b): only NORM_L1 on CV_8U");
    long s = 0;
    for (int y = 0; y < a.rows; y++) {
        const uchar* pa = a.ptr(y); const uchar* pb = b.ptr(y);
        for (int x = 0; x < a.cols; x++) s += pa[x] > pb[x] ? pa[x] - pb[x] : pb[x] - pa[x];
    }
    return (double)s;
}

/* cv::Mat_<float>(r, c) << a, b, c  (src/Frame.cc:1321) */
template <typename T> struct cvshim_type;
template <> struct cvshim_type<float> { enum { value = CV_32F }; };
template <> struct cvshim_type<double> { enum { value = CV_64F }; };
template <> struct cvshim_type<uchar> { enum { value = CV_8U }; };
template <> struct cvshim_type<int> { enum { value = CV_32S }; };
template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, cvshim_type<T>::value) {}
};
template <typename T> class MatCommaInitializer_ {
public:
    MatCommaInitializer_(const Mat_<T>& m) : m_(m), i_(0) {}
    template <typename V> MatCommaInitializer_<T>& operator,(V v)
    {
        m_.template at<T>(i_ / m_.cols, i_ % m_.cols) = (T)v; i_++;
        return *this;
    }
    operator Mat() const { return m_; }
    operator Mat_<T>() const { return m_; }
private:
    Mat_<T> m_; int i_;
};
template <typename T, typename V> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, V v)
{
    MatCommaInitializer_<T> ci(m);
    return (ci, v);
}

inline void hconcat(const Mat& a, const Mat& b, Mat& dst)
{
    if (a.rows != b.rows || a.type() != b.type()) cvshim_fail("hconcat");
    Mat m(a.rows, a.cols + b.cols, a.type());
    for (int y = 0; y < a.rows; y++) {
        std::memcpy(m.ptr(y), a.ptr(y), (size_t)a.cols * a.elemSize());
        std::memcpy(m.ptr(y) + (size_t)a.cols * a.elemSize(), b.ptr(y), (size_t)b.cols * b.elemSize());
    }
    dst = m;
}
inline void vconcat(const Mat& a, const Mat& b, Mat& dst)
{
    if ((a.cols != b.cols && !a.empty() && !b.empty()) || a.type() != b.type()) cvshim_fail("vconcat");
    const int cols = a.empty() ? b.cols : a.cols;
    Mat m(a.rows + b.rows, cols, a.type());
    for (int y = 0; y < a.rows; y++) std::memcpy(m.ptr(y), a.ptr(y), (size_t)cols * a.elemSize());
    for (int y = 0; y < b.rows; y++) std::memcpy(m.ptr(a.rows + y), b.ptr(y), (size_t)cols * b.elemSize());
    dst = m;
}

/* cv::undistortPoints(src, dst, K, distCoeffs, R = empty, P) for N x 2 CV_32F point lists with P == K, 4 or 5 coefficients:
 * orc_undistort_points (oracle/orc_prepare.c), pinned against cv2.undistortPoints in tests/test_oracle_vs_cv2.py */
extern "C" void orc_undistort_points(const float* xy_in, int n, float fx, float fy, float cx, float cy, const float* dist, float* xy_out);
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& distCoeffs, const Mat& R, const Mat& P)
{
    if (src.type() != CV_32F || src.cols != 2 || !src.isContinuous() || !R.empty()) cvshim_fail("undistortPoints: N x 2 CV_32F, R empty");
    if (K.type() != CV_32F || P.type() != CV_32F) cvshim_fail("undistortPoints: K, P CV_32F");
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            if (K.at<float>(i, j) != P.at<float>(i, j)) cvshim_fail("undistortPoints: P != K");
    float d[5] = {0, 0, 0, 0, 0};
    const int nd = (int)distCoeffs.total();
    if (nd != 4 && nd != 5) cvshim_fail("undistortPoints: 4 or 5 distortion coefficients");
    for (int i = 0; i < nd; i++) d[i] = distCoeffs.at<float>(i);
    Mat out(src.rows, 2, CV_32F);
    orc_undistort_points(src.ptr<float>(), src.rows, K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2), d, out.ptr<float>());
    dst = out;
}

/* features2d: cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, matches, 2) = orc_bf_knn2 (oracle/orc_prims.c), pinned against
 * cv2.BFMatcher.knnMatch incl. its tie order in tests/test_oracle_vs_cv2.py */
struct DMatch {
    int queryIdx, trainIdx, imgIdx; float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(3.4e38f) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(0), distance(d) {}
};
extern "C" void orc_bf_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* out_idx, int32_t* out_dist);
class BFMatcher {
public:
    explicit BFMatcher(int normType = NORM_L2, bool crossCheck = false) : norm_(normType) { (void)crossCheck; }
    void knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch> >& matches, int k) const
    {
        if (norm_ != NORM_HAMMING || k != 2) cvshim_fail("BFMatcher: only NORM_HAMMING, k = 2");
        matches.clear();
        const int nq = q.rows, nt = t.rows;
        if (nq == 0) return;
        if (q.cols != 32 || (nt > 0 && t.cols != 32) || !q.isContinuous() || (nt > 0 && !t.isContinuous())) cvshim_fail("BFMatcher: 32-byte continuous rows");
        matches.resize((size_t)nq);
        if (nt == 0) return;
        std::vector<int32_t> idx((size_t)nq * 2), dist((size_t)nq * 2);
        orc_bf_knn2(q.ptr(), nq, t.ptr(), nt, idx.data(), dist.data());
        for (int i = 0; i < nq; i++)
            for (int j = 0; j < 2; j++)
                if (idx[2 * i + j] >= 0) matches[(size_t)i].push_back(DMatch(i, idx[2 * i + j], (float)dist[2 * i + j]));
    }
private:
    int norm_;
};

/* persistence (core/persistence.hpp): the vendored DBoW2's YAML save / load (TemplatedVocabulary.h:1456-1620) are virtual
 * members and must compile; ORB-SLAM3 only ever calls loadFromTextFile (plain fstream), so these throw if reached. */
class FileNode {
public:
    FileNode operator[](const char*) const { unreachable(); return FileNode(); }
    FileNode operator[](const std::string&) const { unreachable(); return FileNode(); }
    FileNode operator[](int) const { unreachable(); return FileNode(); }
    size_t size() const { unreachable(); return 0; }
    operator int() const { unreachable(); return 0; }
    operator float() const { unreachable(); return 0.f; }
    operator double() const { unreachable(); return 0.0; }
    operator std::string() const { unreachable(); return std::string(); }
private:
    static void unreachable() { throw std::runtime_error("cvshim: cv::FileStorage is not implemented (YAML vocabulary path is unused)"); }
};
class FileStorage {
public:
    enum Mode { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { return FileNode()["x"]; }
    FileNode operator[](const std::string&) const { return FileNode()["x"]; }
};
template <class T> inline FileStorage& operator<<(FileStorage& fs, const T&)
{
    throw std::runtime_error("cvshim: cv::FileStorage is not implemented (YAML vocabulary path is unused)");
    return fs;
}

} // namespace cv
