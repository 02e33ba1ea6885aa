/* cvstub.hpp — the small part of the OpenCV C++ API that /root/reference/src/ORBextractor.cc and the body of
 * Frame::ComputeStereoMatches (/root/reference/src/Frame.cc:797-967) touch.  TEST INFRASTRUCTURE ONLY.
 *
 * OpenCV's C++ headers are not installed in this image, so the reference cannot be built against the real
 * library.  This header lets the reference's own source files compile UNMODIFIED (oracle/Makefile, target
 * _ref/liborb_ref.so): containers and geometry types are restated here, and the five pixel primitives
 * (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2) forward to the restatements in oracle/orb_oracle.cpp
 * that tests/test_oracle_golden.py pins bit-exactly to the real cv2 4.13.  What the resulting library therefore
 * pins is the reference's own glue — cell loop, DistributeOctTree, ordering, lapping partition, IC_Angle sums,
 * rotated-BRIEF tap arithmetic as the compiler evaluates it, the stereo row search — which the oracle restates.
 *
 * Only the members the reference uses exist; anything else fails to compile, on purpose.
 */
#ifndef ORBFE_CVSTUB_HPP
#define ORBFE_CVSTUB_HPP

#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "../orb_oracle.h"

typedef unsigned char uchar;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5   /* the small float matrices of the matcher and camera-model bodies (cvstub_float.hpp); the pixel routines are 8-bit only */
#define CV_32FC1 5

/* cvRound = round half to even in the current (default) rounding mode, like OpenCV's SSE2 / lrint paths. */
static inline int cvRound(double v) { return (int)std::lrint(v); }
static inline int cvRound(float v) { return (int)std::lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

using ::cvRound;
using ::cvFloor;
using ::cvCeil;

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
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
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};
static_assert(sizeof(KeyPoint) == sizeof(orc_keypoint), "cv::KeyPoint is 7 x 4 bytes");

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };
enum { NORM_L1 = 2 };

class _OutputArray;
struct MatExpr;

/* 8-bit single-channel matrix header over a shared buffer.  `origin`/`wholeRows`/`wholeCols` describe the
 * allocation a view was cut from (what Mat::locateROI reports), which copyMakeBorder without BORDER_ISOLATED needs. */
class Mat {
public:
    struct Step {
        size_t v;
        operator size_t() const { return v; }
    };
    int rows, cols;
    uchar* data;
    Step step;

    Mat() : rows(0), cols(0), data(nullptr), step{0}, origin(nullptr), wholeRows(0), wholeCols(0), type_(CV_8UC1) {}
    Mat(Size sz, int type) { alloc(sz.height, sz.width, type); }
    Mat(int r, int c, int type) { alloc(r, c, type); }
    /* header over caller memory (used by the C entry points of ref_shim.cpp, not by the reference) */
    Mat(int r, int c, int type, void* ext, size_t extStep)
        : rows(r), cols(c), data((uchar*)ext), step{extStep}, origin((uchar*)ext), wholeRows(r), wholeCols(c), type_(type) { assert(type == CV_8UC1 || type == CV_32F); }

    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        if (m.data) std::memset(m.data, 0, (size_t)r * c * m.elemSize());
        return m;
    }

    int type() const { return type_; }
    size_t elemSize() const { return type_ == CV_32F ? 4 : 1; }
    size_t total() const { return (size_t)rows * cols; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step.v; }

    Mat operator()(const Rect& r) const { return view(r.y, r.y + r.height, r.x, r.x + r.width); }
    Mat rowRange(int a, int b) const { return view(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return view(0, rows, a, b); }
    Mat row(int y) const { return view(y, y + 1, 0, cols); }

    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step.v, data + (size_t)y * step.v, cols * elemSize());
        return m;
    }
    inline void copyTo(const _OutputArray& dst) const;

    template <typename T> T& at(int y, int x) { assert(sizeof(T) == elemSize()); return *(T*)(data + (size_t)y * step.v + (size_t)x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { assert(sizeof(T) == elemSize()); return *(const T*)(data + (size_t)y * step.v + (size_t)x * sizeof(T)); }
    /* element i of a row or column vector */
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    Mat col(int x) const { return view(0, rows, x, x + 1); }
    inline Mat t() const;
    inline Mat inv() const;
    inline double dot(const Mat& b) const;
    inline Mat& operator=(const MatExpr& e);
    Mat(const Mat&) = default;
    Mat& operator=(const Mat&) = default;
    inline Mat(const MatExpr& e);
    uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.v); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.v); }

    void create(int r, int c, int type) {
        if (data && r == rows && c == cols && type == type_) return;
        alloc(r, c, type);
    }
    void release() { *this = Mat(); }

    bool isSubmatrix() const { return rows != wholeRows || cols != wholeCols; }
    /* grow (or shrink) the view inside the allocation it was cut from, clamped to it */
    Mat& adjustROI(int dtop, int dbottom, int dleft, int dright) {
        Size whole; Point ofs;
        locateROI(whole, ofs);
        int y0 = std::max(ofs.y - dtop, 0), y1 = std::min(ofs.y + rows + dbottom, whole.height);
        int x0 = std::max(ofs.x - dleft, 0), x1 = std::min(ofs.x + cols + dright, whole.width);
        data = origin + (size_t)y0 * step.v + x0;
        rows = y1 - y0; cols = x1 - x0;
        return *this;
    }
    void locateROI(Size& whole, Point& ofs) const {
        whole = Size(wholeCols, wholeRows);
        size_t d = (size_t)(data - origin);
        ofs = Point((int)(d % step.v / elemSize()), (int)(d / step.v));
    }

private:
    std::shared_ptr<uchar> hold;
    uchar* origin;
    int wholeRows, wholeCols;
    int type_;

    void alloc(int r, int c, int type) {
        assert(type == CV_8UC1 || type == CV_32F);
        const size_t esz = type == CV_32F ? 4 : 1;   /* every view / pixel routine of this header is for the 8-bit case */
        type_ = type;
        rows = r; cols = c; step.v = (size_t)c * esz; wholeRows = r; wholeCols = c;
        size_t n = (size_t)r * c * esz;
        hold = std::shared_ptr<uchar>(n ? new uchar[n] : nullptr, std::default_delete<uchar[]>());
        data = origin = hold.get();
    }
    Mat view(int y0, int y1, int x0, int x1) const {
        assert(0 <= y0 && y0 <= y1 && y1 <= rows && 0 <= x0 && x0 <= x1 && x1 <= cols);
        Mat m(*this);
        m.rows = y1 - y0; m.cols = x1 - x0;
        m.data = data + (size_t)y0 * step.v + (size_t)x0 * elemSize();
        return m;
    }
};

class _InputArray {
public:
    _InputArray(const Mat& m) : mat(const_cast<Mat*>(&m)) {}
    bool empty() const { return mat->empty(); }
    Mat getMat() const { return *mat; }
protected:
    Mat* mat;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray(Mat& m) : _InputArray(m), fixed(false) {}
    _OutputArray(const Mat& m) : _InputArray(m), fixed(true) {}   /* a temporary header: the size must already fit */
    void create(int r, int c, int type) const {
        if (fixed) { assert(mat->rows == r && mat->cols == c); return; }
        mat->create(r, c, type);
    }
    void release() const { if (!fixed) mat->release(); }
private:
    bool fixed;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline void Mat::copyTo(const _OutputArray& dst) const {
    dst.create(rows, cols, CV_8UC1);
    Mat d = dst.getMat();
    for (int y = 0; y < rows; ++y) std::memmove(d.data + (size_t)y * d.step.v, data + (size_t)y * step.v, cols);
}

/* --- pixel primitives: forwards to the cv2-pinned restatements of oracle/orb_oracle.cpp --------------------------- */

inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

inline void resize(InputArray _src, OutputArray _dst, Size dsize, double fx, double fy, int interpolation) {
    assert(fx == 0 && fy == 0 && interpolation == INTER_LINEAR);
    (void)fx; (void)fy; (void)interpolation;
    Mat src = _src.getMat();
    _dst.create(dsize.height, dsize.width, CV_8UC1);
    Mat dst = _dst.getMat();
    orc_resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

inline int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

/* Only BORDER_REFLECT_101 [+ BORDER_ISOLATED].  Without BORDER_ISOLATED a view first widens into the pixels of the
 * matrix it was cut from, as OpenCV does (src/ORBextractor.cc:1172-1173 is such a call; for a whole image it is moot). */
inline void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int borderType) {
    Mat src = _src.getMat();
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    int rows = src.rows, cols = src.cols;
    const uchar* base = src.data;
    size_t sstep = src.step;
    if (src.isSubmatrix() && !(borderType & BORDER_ISOLATED)) {
        Size whole; Point ofs;
        src.locateROI(whole, ofs);
        int dt = std::min(ofs.y, top), db = std::min(whole.height - rows - ofs.y, bottom);
        int dl = std::min(ofs.x, left), dr = std::min(whole.width - cols - ofs.x, right);
        base -= (size_t)dt * sstep + dl;
        rows += dt + db; cols += dl + dr;
        top -= dt; bottom -= db; left -= dl; right -= dr;
    }
    _dst.create(rows + top + bottom, cols + left + right, CV_8UC1);
    Mat dst = _dst.getMat();
    /* interior first (it may already be in place: level > 0 passes the interior view of dst itself) */
    for (int y = 0; y < rows; ++y) {
        uchar* d = dst.data + (size_t)(y + top) * dst.step + left;
        const uchar* s = base + (size_t)y * sstep;
        if (d != s) std::memmove(d, s, cols);
    }
    for (int y = 0; y < rows; ++y) {
        uchar* d = dst.data + (size_t)(y + top) * dst.step;
        for (int x = 0; x < left; ++x) d[x] = d[left + reflect101(x - left, cols)];
        for (int x = 0; x < right; ++x) d[left + cols + x] = d[left + reflect101(cols + x, cols)];
    }
    for (int y = 0; y < top; ++y)
        std::memcpy(dst.data + (size_t)y * dst.step, dst.data + (size_t)(top + reflect101(y - top, rows)) * dst.step, dst.cols);
    for (int y = 0; y < bottom; ++y)
        std::memcpy(dst.data + (size_t)(top + rows + y) * dst.step, dst.data + (size_t)(top + reflect101(rows + y, rows)) * dst.step, dst.cols);
}

inline void FAST(InputArray _image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    assert(nonmaxSuppression);
    (void)nonmaxSuppression;
    Mat im = _image.getMat();
    /* scratch for the detector's output: one malloc'd block per thread, grown on demand and reused by every cell (plain malloc, so it
     * is neither zero-filled per call nor part of ref_shim.cpp's per-call arena) */
    static thread_local orc_keypoint* out = nullptr;
    static thread_local size_t outCap = 0;
    const size_t need = (size_t)im.rows * im.cols + 1;
    if (need > outCap) {
        std::free(out);
        out = (orc_keypoint*)std::malloc(need * sizeof(orc_keypoint));
        outCap = out ? need : 0;
        if (!out) std::abort();
    }
    int n = orc_fast9_16(im.data, im.cols, im.rows, (int)im.step, threshold, out, (int)outCap);
    keypoints.clear();
    keypoints.reserve(n);
    for (int i = 0; i < n; ++i) keypoints.push_back(KeyPoint(out[i].x, out[i].y, out[i].size, out[i].angle, out[i].response, out[i].octave, out[i].class_id));
}

inline void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sx, double sy, int borderType) {
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sx; (void)sy; (void)borderType;
    Mat src = _src.getMat().clone();
    _dst.create(src.rows, src.cols, CV_8UC1);
    Mat dst = _dst.getMat();
    orc_gaussian7_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, (int)dst.step);
}

/* cv::norm(a, b, NORM_L1) on 8-bit views: the exact integer sum of |a - b| as a double (src/Frame.cc:911). */
inline double norm(InputArray _a, InputArray _b, int normType) {
    assert(normType == NORM_L1);
    (void)normType;
    Mat a = _a.getMat(), b = _b.getMat();
    long s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += std::abs((int)a.at<uchar>(y, x) - (int)b.at<uchar>(y, x));
    return (double)s;
}

/* Only ORBextractor::ComputeKeyPointsOld (src/ORBextractor.cc:880-1057) names this, and its one call is commented out
 * (:1083); it has to link, it never runs. */
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>&, int) { std::abort(); }
};

/* cv::FileStorage / cv::FileNode: Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h names them in its YAML save / load members
 * (:1452-1660).  The vocabulary under test is read with loadFromTextFile (:1338-1426), which uses no OpenCV; these only have to
 * compile, and abort if anything ever calls them. */
class FileNode {
public:
    FileNode operator[](const char*) const { std::abort(); }
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](int) const { std::abort(); }
    size_t size() const { std::abort(); }
    operator int() const { std::abort(); }
    operator float() const { std::abort(); }
    operator double() const { std::abort(); }
    operator std::string() const { std::abort(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](const char*) const { std::abort(); }
    void release() {}
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { std::abort(); return fs; }

}  // namespace cv

#include "cvstub_float.hpp"

#endif
