/* orbfe_kb8.cuh — KannalaBrandt8::unproject / project / TriangulateMatches and OpenCV's 4x4 Jacobi SVD as device functions, shared by the
 * fisheye stereo kernels (orbfe_match.cu) and the rig branch of SearchForTriangulation (orbfe_bow.cu). */
#ifndef ORBFE_KB8_CUH
#define ORBFE_KB8_CUH
#include <cuda_runtime.h>
#include <stdint.h>

/* KannalaBrandt8::TriangulateMatches (src/CameraModels/KannalaBrandt8.cpp:338-407).  The float / double split follows OpenCV's arithmetic as the oracle pins it
 * (oracle/triangulate_oracle.cpp): gemm = float products accumulated in float, Mat::dot / norm = double sums, the rows of the
 * 4x4 system by one fused multiply-add, cv::SVD::compute = one-sided Jacobi with double norms and float rotations. */
struct Kb8Cam { float fx, fy, cx, cy, k0, k1, k2, k3; };

static __device__ void kb8_unproject(const Kb8Cam& c, float u, float v, float (&ray)[3]) {   /* KannalaBrandt8.cpp:96-125 */
    const float pwx = (u - c.cx) / c.fx, pwy = (v - c.cy) / c.fy;
    float scale = 1.f;
    float theta_d = sqrtf(pwx * pwx + pwy * pwy);
    theta_d = fminf(fmaxf(-1.57079637f, theta_d), 1.57079637f);
    if ((double)theta_d > 1e-8) {
        float theta = theta_d;
        for (int j = 0; j < 10; j++) {
            const float t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
            const float a = c.k0 * t2, b = c.k1 * t4, cc = c.k2 * t6, d = c.k3 * t8;
            const float fix = (theta * (1 + a + b + cc + d) - theta_d) / (1 + 3 * a + 5 * b + 7 * cc + 9 * d);
            theta = theta - fix;
            if (fabsf(fix) < 1e-6f) break;
        }
        scale = tanf(theta) / theta_d;
    }
    ray[0] = pwx * scale; ray[1] = pwy * scale; ray[2] = 1.f;
}

static __device__ void kb8_project(const Kb8Cam& c, const float (&p)[3], float (&uv)[2]) {   /* KannalaBrandt8.cpp:25-41 */
    const float x2y2 = p[0] * p[0] + p[1] * p[1];
    const float theta = atan2f(sqrtf(x2y2), p[2]);
    const float psi = atan2f(p[1], p[0]);
    const float t2 = theta * theta, t3 = theta * t2, t5 = t3 * t2, t7 = t5 * t2, t9 = t7 * t2;
    const float r = theta + c.k0 * t3 + c.k1 * t5 + c.k2 * t7 + c.k3 * t9;
    /* the reference writes `cos(psi)` / `sin(psi)` on a float with no std overloads in scope (:39-40): the C library's double functions,
     * so the product and the sum are evaluated in double and rounded once by cv::Point2f's constructor */
    double sn, cs;
    sincos((double)psi, &sn, &cs);
    uv[0] = (float)((double)(c.fx * r) * cs + (double)c.cx);
    uv[1] = (float)((double)(c.fy * r) * sn + (double)c.cy);
}

static __device__ __forceinline__ double dot3d(const float* a, const float* b) { return (double)a[0] * b[0] + (double)a[1] * b[1] + (double)a[2] * b[2]; }
static __device__ __forceinline__ float gemm3f(const float* a, const float* b) {
    float s = 0.f;
    s = s + a[0] * b[0];
    s = s + a[1] * b[1];
    s = s + a[2] * b[2];
    return s;
}

/* cv::SVD::compute on the 4x4 system, last row of vt (JacobiSVDImpl_<float>, m = n = 4; At = A transposed) */
static __device__ void svd4_last_row(float (&At)[4][4], float (&last)[4]) {
    float Vt[4][4];
    double W[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) sd += (double)At[i][k] * At[i][k];
        W[i] = sd;
#pragma unroll
        for (int k = 0; k < 4; k++) Vt[i][k] = i == k ? 1.f : 0.f;
    }
    const float eps = 1.1920929e-07f * 2;
    for (int iter = 0; iter < 30; iter++) {
        bool changed = false;
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = i + 1; j < 4; j++) {
                double a = W[i], p = 0, b = W[j];
#pragma unroll
                for (int k = 0; k < 4; k++) p += (double)At[i][k] * At[j][k];
                if (fabs(p) <= eps * sqrt(a * b)) continue;
                p *= 2;
                const double beta = a - b, gamma = hypot(p, beta);
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
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const float t0 = c * At[i][k] + s * At[j][k];
                    const float t1 = -s * At[i][k] + c * At[j][k];
                    At[i][k] = t0; At[j][k] = t1;
                    a += (double)t0 * t0; b += (double)t1 * t1;
                }
                W[i] = a; W[j] = b;
                changed = true;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const float t0 = c * Vt[i][k] + s * Vt[j][k];
                    const float t1 = -s * Vt[i][k] + c * Vt[j][k];
                    Vt[i][k] = t0; Vt[j][k] = t1;
                }
            }
        if (!changed) break;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) sd += (double)At[i][k] * At[i][k];
        W[i] = sqrt(sd);
    }
    /* the selection sort of the singular values (descending, first maximum wins): only the row that ends up last matters */
    int idx[4] = {0, 1, 2, 3};
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int j = i;
#pragma unroll
        for (int k = i + 1; k < 4; k++)
            if (W[j] < W[k]) j = k;
        if (i != j) {
            const double tw = W[i]; W[i] = W[j]; W[j] = tw;
            const int ti = idx[i]; idx[i] = idx[j]; idx[j] = ti;
        }
    }
#pragma unroll
    for (int r = 0; r < 4; r++)
        if (idx[3] == r) {
#pragma unroll
            for (int k = 0; k < 4; k++) last[k] = Vt[r][k];
        }
}

struct FisheyeRig { Kb8Cam c1, c2; float R12[9], t12[3]; };

/* TriangulateMatches for one keypoint pair; returns the function's value (z1, or -1) and the point. */
static __device__ float kb8_triangulate_pair(const FisheyeRig& rig, float u1, float v1, float u2, float v2, float s1, float s2, float (&x3D)[3]) {
    float R21[3][3], t21[3];
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 3; b++) R21[a][b] = rig.R12[b * 3 + a];
#pragma unroll
    for (int a = 0; a < 3; a++) t21[a] = -gemm3f(R21[a], rig.t12);
    float r1[3], r2[3], r21[3];
    kb8_unproject(rig.c1, u1, v1, r1);
    kb8_unproject(rig.c2, u2, v2, r2);
#pragma unroll
    for (int a = 0; a < 3; a++) r21[a] = gemm3f(rig.R12 + 3 * a, r2);
    const float cosParallax = (float)(dot3d(r1, r21) / (sqrt(dot3d(r1, r1)) * sqrt(dot3d(r21, r21))));
    if ((double)cosParallax > 0.9998) return -1.f;
    /* At[col][row] of A = [p1.x*T1[2]-T1[0]; p1.y*T1[2]-T1[1]; p2.x*T2[2]-T2[0]; p2.y*T2[2]-T2[1]], Tcw1 = [I|0], Tcw2 = [R21|t21] */
    float At[4][4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float T1r2 = k == 2 ? 1.f : 0.f, T1r0 = k == 0 ? 1.f : 0.f, T1r1 = k == 1 ? 1.f : 0.f;
        const float T2r2 = k < 3 ? R21[2][k] : t21[2], T2r0 = k < 3 ? R21[0][k] : t21[0], T2r1 = k < 3 ? R21[1][k] : t21[1];
        At[k][0] = __fmaf_rn(r1[0], T1r2, -T1r0);
        At[k][1] = __fmaf_rn(r1[1], T1r2, -T1r1);
        At[k][2] = __fmaf_rn(r2[0], T2r2, -T2r0);
        At[k][3] = __fmaf_rn(r2[1], T2r2, -T2r1);
    }
    float v[4];
    svd4_last_row(At, v);
    const float inv = (float)(1.0 / (double)v[3]);
    x3D[0] = v[0] * inv; x3D[1] = v[1] * inv; x3D[2] = v[2] * inv;
    const float z1 = x3D[2];
    if (!(z1 > 0)) return -1.f;
    if (!((float)(dot3d(R21[2], x3D) + (double)t21[2]) > 0)) return -1.f;
    float uv[2];
    kb8_project(rig.c1, x3D, uv);
    float ex = uv[0] - u1, ey = uv[1] - v1;
    if ((double)(ex * ex + ey * ey) > 5.991 * (double)s1) return -1.f;
    float x2[3];
#pragma unroll
    for (int a = 0; a < 3; a++) x2[a] = gemm3f(R21[a], x3D) + t21[a];
    kb8_project(rig.c2, x2, uv);
    ex = uv[0] - u2; ey = uv[1] - v2;
    if ((double)(ex * ex + ey * ey) > 5.991 * (double)s2) return -1.f;
    return z1;
}

#endif
