/* triangulate_oracle.cpp — CPU ORACLE of KannalaBrandt8::TriangulateMatches.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates /root/reference/src/CameraModels/KannalaBrandt8.cpp:
 *     unproject            :96-125   (Newton iteration on the distortion polynomial, float)
 *     project              :25-41
 *     TriangulateMatches   :338-407  (parallax test, Triangulate, depth signs, two reprojection tests)
 *     Triangulate          :499-513  (4x4 homogeneous system, cv::SVD::compute, last row of vt)
 * and the call site Frame::ComputeStereoFishEyeMatches, src/Frame.cc:1145-1157 (depth > 0.0001f keeps the match).
 *
 * The cv::Mat arithmetic of those functions lives in OpenCV (not vendored).  What is restated here, and how it is pinned:
 *   - cv::SVD::compute on a 4x4 CV_32F matrix = OpenCV's one-sided Jacobi (JacobiSVDImpl_<float>: column norms and dot products
 *     accumulated in double, rotations applied in float, sweeps until no pair rotates, singular values sorted descending);
 *     orc_svd4x4_vt is checked BIT-EXACTLY against cv2.SVDecomp 4.13 on random and on triangulation matrices
 *     (tools/make_goldens.py -> tests/golden/triangulate.npz -> tests/test_triangulate.py);
 *   - Mat * Mat on these sizes (cv::gemm): float products accumulated in float, in order, no FMA — bit-exact against cv2.gemm;
 *   - `p * row - row` (MatExpr -> addWeighted): one fused multiply-add per element — bit-exact against cv2.addWeighted;
 *   - Mat::dot and cv::norm on 3-vectors: summed in double (cv2.norm agrees on 99.4 % of random vectors to the last bit of the
 *     double; the rest differ by one unit in the last place of that double, far below the float the result is rounded to);
 *   - `vec / scalar`: multiplication by the float reciprocal.
 * libm calls (tanf, atan2f, cosf, sinf, sqrtf) are the host's.  Tolerance of everything that compares against this file is stated
 * in tests/test_triangulate.py (1e-4 relative on depth and point; identical accept / reject away from the thresholds).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>

#include "orb_oracle.h"

namespace {

/* JacobiSVDImpl_<float> for m = n = 4 on At (rows of At = columns of A); Vt accumulates the right singular vectors as rows. */
void jacobi_svd4(float At[4][4], float W_[4], float Vt[4][4]) {
    const int m = 4, n = 4;
    const float eps = FLT_EPSILON * 2;
    double W[4];
    for (int i = 0; i < n; i++) {
        double sd = 0;
        for (int k = 0; k < m; k++) { const float t = At[i][k]; sd += (double)t * t; }
        W[i] = sd;
        for (int k = 0; k < n; k++) Vt[i][k] = 0;
        Vt[i][i] = 1;
    }
    const int max_iter = std::max(m, 30);
    for (int iter = 0; iter < max_iter; iter++) {
        bool changed = false;
        for (int i = 0; i < n - 1; i++)
            for (int j = i + 1; j < n; j++) {
                float* Ai = At[i];
                float* Aj = At[j];
                double a = W[i], p = 0, b = W[j];
                for (int k = 0; k < m; k++) p += (double)Ai[k] * Aj[k];
                if (fabs(p) <= eps * sqrt((double)a * b)) continue;
                p *= 2;
                const double beta = a - b, gamma = hypot((double)p, beta);
                float c, s;
                if (beta < 0) {
                    const double delta = (gamma - beta) * 0.5;
                    s = (float)sqrt(delta / gamma);
                    c = (float)(p / (gamma * s * 2));
                } else {
                    c = (float)sqrt((gamma + beta) / (gamma * 2));
                    s = (float)(p / (gamma * c * 2));
                }
                a = b = 0;
                for (int k = 0; k < m; k++) {
                    const float t0 = c * Ai[k] + s * Aj[k];
                    const float t1 = -s * Ai[k] + c * Aj[k];
                    Ai[k] = t0; Aj[k] = t1;
                    a += (double)t0 * t0; b += (double)t1 * t1;
                }
                W[i] = a; W[j] = b;
                changed = true;
                float* Vi = Vt[i];
                float* Vj = Vt[j];
                for (int k = 0; k < n; k++) {
                    const float t0 = c * Vi[k] + s * Vj[k];
                    const float t1 = -s * Vi[k] + c * Vj[k];
                    Vi[k] = t0; Vj[k] = t1;
                }
            }
        if (!changed) break;
    }
    for (int i = 0; i < n; i++) {
        double sd = 0;
        for (int k = 0; k < m; k++) { const float t = At[i][k]; sd += (double)t * t; }
        W[i] = sqrt(sd);
    }
    for (int i = 0; i < n - 1; i++) {
        int j = i;
        for (int k = i + 1; k < n; k++)
            if (W[j] < W[k]) j = k;
        if (i != j) {
            std::swap(W[i], W[j]);
            for (int k = 0; k < m; k++) std::swap(At[i][k], At[j][k]);
            for (int k = 0; k < n; k++) std::swap(Vt[i][k], Vt[j][k]);
        }
    }
    for (int i = 0; i < n; i++) W_[i] = (float)W[i];
}

struct Kb8 {
    float fx, fy, cx, cy, k0, k1, k2, k3;
};

/* KannalaBrandt8::unproject, :96-125 (precision = 1e-6f, KannalaBrandt8.h:39-44) */
void kb8_unproject(const Kb8& c, float u, float v, float ray[3]) {
    const float pwx = (u - c.cx) / c.fx, pwy = (v - c.cy) / c.fy;
    float scale = 1.f;
    float theta_d = sqrtf(pwx * pwx + pwy * pwy);
    theta_d = fminf(fmaxf((float)(-M_PI / 2.f), theta_d), (float)(M_PI / 2.f));
    if ((double)theta_d > 1e-8) {
        float theta = theta_d;
        for (int j = 0; j < 10; j++) {
            const float theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
            const float k0t2 = c.k0 * theta2, k1t4 = c.k1 * theta4, k2t6 = c.k2 * theta6, k3t8 = c.k3 * theta8;
            const float fix = (theta * (1 + k0t2 + k1t4 + k2t6 + k3t8) - theta_d) / (1 + 3 * k0t2 + 5 * k1t4 + 7 * k2t6 + 9 * k3t8);
            theta = theta - fix;
            if (fabsf(fix) < 1e-6f) break;
        }
        scale = tanf(theta) / theta_d;
    }
    ray[0] = pwx * scale; ray[1] = pwy * scale; ray[2] = 1.f;
}

/* KannalaBrandt8::project, :25-41 */
void kb8_project(const Kb8& c, const float p[3], float uv[2]) {
    const float x2y2 = p[0] * p[0] + p[1] * p[1];
    const float theta = atan2f(sqrtf(x2y2), p[2]);
    const float psi = atan2f(p[1], p[0]);
    const float t2 = theta * theta, t3 = theta * t2, t5 = t3 * t2, t7 = t5 * t2, t9 = t7 * t2;
    const float r = theta + c.k0 * t3 + c.k1 * t5 + c.k2 * t7 + c.k3 * t9;
    /* `cos(psi)` / `sin(psi)` on a float, with no std overloads in scope in the reference's translation unit, are ::cos(double) / ::sin(double):
     * product and sum in double, one rounding in cv::Point2f's constructor (oracle/ref_camera_shim.cpp asserts the overload that is taken) */
    uv[0] = (float)((double)(c.fx * r) * cos((double)psi) + (double)c.cx);
    uv[1] = (float)((double)(c.fy * r) * sin((double)psi) + (double)c.cy);
}

/* Mat::dot / cv::norm on 3-vectors: products summed in double */
inline double dot3(const float* a, const float* b) { return (double)a[0] * b[0] + (double)a[1] * b[1] + (double)a[2] * b[2]; }
/* one output element of Mat * Mat for these sizes (cv::gemm, CV_32F): float products accumulated in float, in order, no FMA
 * (checked against cv2.gemm 4.13 on 20000 random 3x3 * 3x1 products: identical bits) */
inline float gemm3(const float* a, const float* b) {
    float s = 0.f;
    s = s + a[0] * b[0];
    s = s + a[1] * b[1];
    s = s + a[2] * b[2];
    return s;
}

}  // namespace

extern "C" {

/* cv::SVD::compute(A, w, u, vt, MODIFY_A | FULL_UV) on a row-major 4x4 float matrix: singular values and vt. */
void orc_svd4x4_vt(const float* A, float* w, float* vt) {
    float At[4][4], Vt[4][4];
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 4; k++) At[i][k] = A[k * 4 + i];   /* the routine works on the rows of A transposed */
    jacobi_svd4(At, w, Vt);
    memcpy(vt, Vt, sizeof Vt);
}

void orc_kb8_unproject(const float* params8, float u, float v, float* ray3) {
    Kb8 c; memcpy(&c, params8, sizeof c);
    kb8_unproject(c, u, v, ray3);
}

void orc_kb8_project(const float* params8, const float* p3, float* uv2) {
    Kb8 c; memcpy(&c, params8, sizeof c);
    kb8_project(c, p3, uv2);
}

/* KannalaBrandt8::TriangulateMatches for n keypoint pairs.  cam1 / cam2: the 8 calibration parameters of the two cameras;
 * R12 row-major 3x3, t12: mRlr, mtlr of the frame; sigma1[i] / sigma2[i]: mvLevelSigma2 of the two keypoints' octaves.
 * depth[i] = the function's return value (-1 when rejected); p3d[3*i..] = the triangulated point when depth[i] > 0. */
void orc_kb8_triangulate_matches(const float* cam1, const float* cam2, const float* R12, const float* t12, const float* uv1, const float* uv2,
                                 const float* sigma1, const float* sigma2, int n, float* depth, float* p3d) {
    Kb8 c1, c2;
    memcpy(&c1, cam1, sizeof c1);
    memcpy(&c2, cam2, sizeof c2);
    float R21[3][3], t21[3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) R21[i][j] = R12[j * 3 + i];
    for (int i = 0; i < 3; i++) t21[i] = -gemm3(R21[i], t12);
    for (int q = 0; q < n; q++) {
        depth[q] = -1.f;
        p3d[3 * q] = p3d[3 * q + 1] = p3d[3 * q + 2] = 0.f;
        float r1[3], r2[3], r21[3];
        kb8_unproject(c1, uv1[2 * q], uv1[2 * q + 1], r1);
        kb8_unproject(c2, uv2[2 * q], uv2[2 * q + 1], r2);
        for (int i = 0; i < 3; i++) r21[i] = gemm3(R12 + 3 * i, r2);
        const float cosParallax = (float)(dot3(r1, r21) / (sqrt(dot3(r1, r1)) * sqrt(dot3(r21, r21))));
        if ((double)cosParallax > 0.9998) continue;
        /* Triangulate: rows p1.x*T1[2]-T1[0], p1.y*T1[2]-T1[1], p2.x*T2[2]-T2[0], p2.y*T2[2]-T2[1]; Tcw1 = [I|0], Tcw2 = [R21|t21] */
        const float p1x = r1[0], p1y = r1[1], p2x = r2[0], p2y = r2[1];
        float A[16];
        const float T1[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}};
        float T2[3][4];
        for (int i = 0; i < 3; i++) { T2[i][0] = R21[i][0]; T2[i][1] = R21[i][1]; T2[i][2] = R21[i][2]; T2[i][3] = t21[i]; }
        for (int k = 0; k < 4; k++) {
            /* `p * row2 - row0` is a MatExpr that OpenCV evaluates with addWeighted: one fused multiply-add per element
             * (checked against cv2.addWeighted 4.13: identical bits on 20000 random rows) */
            A[k] = fmaf(p1x, T1[2][k], -T1[0][k]);
            A[4 + k] = fmaf(p1y, T1[2][k], -T1[1][k]);
            A[8 + k] = fmaf(p2x, T2[2][k], -T2[0][k]);
            A[12 + k] = fmaf(p2y, T2[2][k], -T2[1][k]);
        }
        float w[4], vt[16];
        orc_svd4x4_vt(A, w, vt);
        const float inv = (float)(1.0 / (double)vt[15]);
        const float x3D[3] = {vt[12] * inv, vt[13] * inv, vt[14] * inv};
        const float z1 = x3D[2];
        if (z1 <= 0) continue;
        const float z2 = (float)(dot3(R21[2], x3D) + (double)t21[2]);
        if (z2 <= 0) continue;
        float uvp[2];
        kb8_project(c1, x3D, uvp);
        const float ex1 = uvp[0] - uv1[2 * q], ey1 = uvp[1] - uv1[2 * q + 1];
        if ((double)(ex1 * ex1 + ey1 * ey1) > 5.991 * (double)sigma1[q]) continue;
        float x3D2[3];
        for (int i = 0; i < 3; i++) x3D2[i] = gemm3(R21[i], x3D) + t21[i];
        kb8_project(c2, x3D2, uvp);
        const float ex2 = uvp[0] - uv2[2 * q], ey2 = uvp[1] - uv2[2 * q + 1];
        if ((double)(ex2 * ex2 + ey2 * ey2) > 5.991 * (double)sigma2[q]) continue;
        depth[q] = z1;
        p3d[3 * q] = x3D[0]; p3d[3 * q + 1] = x3D[1]; p3d[3 * q + 2] = x3D[2];
    }
}

}  // extern "C"
