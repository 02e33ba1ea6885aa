/* cvstub_float.hpp — the CV_32F matrix algebra that the matcher and camera-model bodies taken from the reference use
 * (src/ORBmatcher.cc SearchForTriangulation, src/CameraModels/Pinhole.cpp, KannalaBrandt8.cpp).  TEST INFRASTRUCTURE ONLY.
 *
 * Every operation evaluates with the arithmetic that was checked against OpenCV 4.13 through cv2 (oracle/triangulate_oracle.cpp lists
 * the checks): Mat * Mat = float products accumulated in float, in order, no FMA; `s * A - B` = one fused multiply-add per element
 * (MatExpr -> addWeighted); dot / norm in double; A / s = multiplication by the float reciprocal; SVD::compute = OpenCV's Jacobi
 * (orc_svd4x4_vt, bit-exact against cv2.SVDecomp).  inv() of a 3x3 is by cofactors: the bodies use it only to form the fundamental
 * matrix that the library takes as an input, so its rounding is outside what is compared. */
#ifndef ORB_REFSTUB_CVSTUB_FLOAT_HPP
#define ORB_REFSTUB_CVSTUB_FLOAT_HPP

namespace cv {

template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;

/* an evaluated matrix, possibly still to be multiplied by alpha (what `-A` and `s * A` are before they meet their consumer) */
struct MatExpr {
    Mat m;
    float alpha;
    MatExpr(const Mat& a, float s = 1.f) : m(a), alpha(s) {}
    Mat eval() const {
        if (alpha == 1.f) return m;
        Mat r(m.rows, m.cols, CV_32F);
        for (int i = 0; i < m.rows; i++)
            for (int j = 0; j < m.cols; j++) r.at<float>(i, j) = m.at<float>(i, j) * alpha;
        return r;
    }
    Mat t() const { return eval().t(); }
    Mat inv() const { return eval().inv(); }
};

inline Mat::Mat(const MatExpr& e) : Mat(e.eval()) {}
/* OpenCV evaluates an expression INTO a destination that already has the right shape (A.row(0) = ... writes the row) */
inline Mat& Mat::operator=(const MatExpr& e) {
    const Mat v = e.eval();
    if (data && rows == v.rows && cols == v.cols && type_ == v.type()) {
        for (int i = 0; i < rows; i++) std::memcpy(data + (size_t)i * step.v, v.data + (size_t)i * v.step.v, (size_t)cols * elemSize());
        return *this;
    }
    return *this = v;
}

inline Mat Mat::t() const {
    assert(type_ == CV_32F);
    Mat r(cols, rows, CV_32F);
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++) r.at<float>(j, i) = at<float>(i, j);
    return r;
}
inline Mat Mat::inv() const {
    assert(type_ == CV_32F && rows == 3 && cols == 3);
    const Mat& a = *this;
    auto A = [&](int i, int j) { return (double)a.at<float>(i, j); };
    const double det = A(0, 0) * (A(1, 1) * A(2, 2) - A(1, 2) * A(2, 1)) - A(0, 1) * (A(1, 0) * A(2, 2) - A(1, 2) * A(2, 0)) + A(0, 2) * (A(1, 0) * A(2, 1) - A(1, 1) * A(2, 0));
    Mat r(3, 3, CV_32F);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            const int i1 = (j + 1) % 3, i2 = (j + 2) % 3, j1 = (i + 1) % 3, j2 = (i + 2) % 3;
            r.at<float>(i, j) = (float)((A(i1, j1) * A(i2, j2) - A(i1, j2) * A(i2, j1)) / det);
        }
    return r;
}
inline double Mat::dot(const Mat& b) const {
    assert(type_ == CV_32F && total() == b.total());
    double s = 0;
    const int n = (int)total();
    for (int k = 0; k < n; k++) {
        const float x = rows == 1 ? at<float>(0, k) : at<float>(k / cols, k % cols);
        const float y = b.rows == 1 ? b.at<float>(0, k) : b.at<float>(k / b.cols, k % b.cols);
        s += (double)x * y;
    }
    return s;
}
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }

/* cv::gemm on these sizes: float products, float accumulation, in order */
inline Mat gemm_f32(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
    Mat r(a.rows, b.cols, CV_32F);
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < b.cols; j++) {
            volatile float s = 0.f;
            for (int k = 0; k < a.cols; k++) { const volatile float p = a.at<float>(i, k) * b.at<float>(k, j); s = s + p; }
            r.at<float>(i, j) = s;
        }
    return r;
}
inline MatExpr operator*(const Mat& a, const Mat& b) { return MatExpr(gemm_f32(a, b)); }
inline MatExpr operator*(const MatExpr& a, const Mat& b) { return MatExpr(gemm_f32(a.m, b), a.alpha); }      /* alpha * (A B) */
inline MatExpr operator*(const Mat& a, const MatExpr& b) { return MatExpr(gemm_f32(a, b.m), b.alpha); }
inline MatExpr operator*(const MatExpr& a, const MatExpr& b) { return MatExpr(gemm_f32(a.m, b.m), a.alpha * b.alpha); }
inline MatExpr operator-(const Mat& a) { return MatExpr(a, -1.f); }
inline MatExpr operator-(const MatExpr& a) { return MatExpr(a.m, -a.alpha); }
inline MatExpr operator*(float s, const Mat& a) { return MatExpr(a, s); }
inline MatExpr operator*(const Mat& a, float s) { return MatExpr(a, s); }
inline MatExpr operator*(double s, const Mat& a) { return MatExpr(a, (float)s); }

inline Mat add_f32(const Mat& a, float alpha, const Mat& b, float beta) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat r(a.rows, a.cols, CV_32F);
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) {
            const float x = a.at<float>(i, j), y = b.at<float>(i, j);
            /* alpha == 1: plain add / subtract; else addWeighted: one fused multiply-add (beta is +-1 here) */
            r.at<float>(i, j) = alpha == 1.f ? (beta == 1.f ? x + y : x - y) : std::fmaf(alpha, x, beta == 1.f ? y : -y);
        }
    return r;
}
inline MatExpr operator+(const Mat& a, const Mat& b) { return MatExpr(add_f32(a, 1.f, b, 1.f)); }
inline MatExpr operator-(const Mat& a, const Mat& b) { return MatExpr(add_f32(a, 1.f, b, -1.f)); }
inline MatExpr operator+(const MatExpr& a, const Mat& b) { return MatExpr(add_f32(a.m, a.alpha, b, 1.f)); }
inline MatExpr operator-(const MatExpr& a, const Mat& b) { return MatExpr(add_f32(a.m, a.alpha, b, -1.f)); }
inline MatExpr operator+(const MatExpr& a, const MatExpr& b) { return MatExpr(add_f32(a.m, a.alpha, b.eval(), 1.f)); }
inline MatExpr operator/(const Mat& a, float s) {   /* A / s: multiplication by the float reciprocal */
    const float inv = (float)(1.0 / (double)s);
    Mat r(a.rows, a.cols, CV_32F);
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) r.at<float>(i, j) = a.at<float>(i, j) * inv;
    return MatExpr(r);
}

inline void hconcat(const Mat& a, const Mat& b, Mat& dst) {
    assert(a.rows == b.rows);
    Mat r(a.rows, a.cols + b.cols, CV_32F);
    for (int i = 0; i < a.rows; i++) {
        for (int j = 0; j < a.cols; j++) r.at<float>(i, j) = a.at<float>(i, j);
        for (int j = 0; j < b.cols; j++) r.at<float>(i, a.cols + j) = b.at<float>(i, j);
    }
    dst = r;
}

/* cv::Mat_<float>(r, c) << a, b, c, ... */
template <typename T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) { static_assert(sizeof(T) == 4, "float only"); }
};
template <typename T> struct MatCommaInitializer_ {
    Mat m;
    int k;
    MatCommaInitializer_(const Mat& mm, T v) : m(mm), k(0) { put(v); }
    void put(T v) { assert(k < (int)m.total()); m.at<T>(k / m.cols, k % m.cols) = v; k++; }
    template <typename U> MatCommaInitializer_& operator,(U v) { put((T)v); return *this; }
    operator Mat() const { assert(k == (int)m.total()); return m; }
};
template <typename T, typename U> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v) { return MatCommaInitializer_<T>(m, (T)v); }

struct SVD {
    enum { MODIFY_A = 1, NO_UV = 2, FULL_UV = 4 };
    static void compute(const Mat& A, Mat& w, Mat& u, Mat& vt, int = 0) {
        assert(A.type() == CV_32F && A.rows == 4 && A.cols == 4);
        float a[16], ww[4], v[16];
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) a[4 * i + j] = A.at<float>(i, j);
        orc_svd4x4_vt(a, ww, v);
        w = Mat(4, 1, CV_32F); vt = Mat(4, 4, CV_32F); u = Mat();
        for (int i = 0; i < 4; i++) {
            w.at<float>(i, 0) = ww[i];
            for (int j = 0; j < 4; j++) vt.at<float>(i, j) = v[4 * i + j];
        }
    }
};

}  // namespace cv
#endif
