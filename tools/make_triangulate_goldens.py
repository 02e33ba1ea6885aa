#!/usr/bin/env python3
"""Golden vectors for KannalaBrandt8::TriangulateMatches (/root/reference/src/CameraModels/KannalaBrandt8.cpp:338-407, 499-513), made
with the real OpenCV of this image (cv2 4.13): cv2.SVDecomp runs the 4x4 SVDs, cv2.gemm the matrix products, cv2.addWeighted the
`p * row - row` rows, cv2.norm the norms; the rest (unproject's Newton iteration, project, the tests) is restated in numpy float32.
Written to tests/golden/triangulate.npz; tests/test_triangulate.py checks the C++ oracle against it (the SVD bit-exactly).

    python tools/make_triangulate_goldens.py
"""
import math
import os

import cv2
import numpy as np

f32 = np.float32
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# Examples/Stereo-Inertial/TUM_512.yaml:9-27 (Camera / Camera2: fx fy cx cy k1..k4) and :29-35 (Tlr)
CAM1 = np.array([190.97847715128717, 190.9733070521226, 254.93170605935475, 256.8974428996504,
                 0.0034823894022493434, 0.0007150348452162257, -0.0020532361418706202, 0.00020293673591811182], f32)
CAM2 = np.array([190.44236969414825, 190.4344384721956, 252.59949716835982, 254.91723064636983,
                 0.0034003170790442797, 0.001766278153469831, -0.00266312569781606, 0.0003299517423931039], f32)
TLR = np.array([[0.999999445773493, 0.000791687752817, 0.000694034010224, 0.101063427414194],
                [-0.000823363992158, 0.998899461915674, 0.046895490788700, 0.001946204678584],
                [-0.000656143613644, -0.046896036240590, 0.998899560146304, 0.001015350132563]], f32)


def unproject(c, u, v):
    pwx, pwy = f32(f32(u - c[2]) / c[0]), f32(f32(v - c[3]) / c[1])
    scale = f32(1)
    theta_d = f32(math.sqrt(f32(f32(pwx * pwx) + f32(pwy * pwy))))
    theta_d = f32(min(max(f32(-math.pi / 2), theta_d), f32(math.pi / 2)))
    if float(theta_d) > 1e-8:
        theta = theta_d
        for _ in range(10):
            t2 = f32(theta * theta); t4 = f32(t2 * t2); t6 = f32(t4 * t2); t8 = f32(t4 * t4)
            a, b, cc, d = f32(c[4] * t2), f32(c[5] * t4), f32(c[6] * t6), f32(c[7] * t8)
            num = f32(f32(theta * f32(f32(f32(f32(f32(1) + a) + b) + cc) + d)) - theta_d)
            den = f32(f32(f32(f32(f32(1) + f32(f32(3) * a)) + f32(f32(5) * b)) + f32(f32(7) * cc)) + f32(f32(9) * d))
            fix = f32(num / den)
            theta = f32(theta - fix)
            if abs(fix) < f32(1e-6):
                break
        scale = f32(f32(math.tan(float(theta))) / theta_d)      # tanf: the double tan rounded to float (<= 1 ulp of libm's tanf)
    return np.array([f32(pwx * scale), f32(pwy * scale), f32(1)], f32)


def project(c, p):
    x2y2 = f32(f32(p[0] * p[0]) + f32(p[1] * p[1]))
    theta = f32(math.atan2(float(f32(math.sqrt(x2y2))), float(p[2])))
    psi = f32(math.atan2(float(p[1]), float(p[0])))
    t2 = f32(theta * theta); t3 = f32(theta * t2); t5 = f32(t3 * t2); t7 = f32(t5 * t2); t9 = f32(t7 * t2)
    r = f32(f32(f32(f32(theta + f32(c[4] * t3)) + f32(c[5] * t5)) + f32(c[6] * t7)) + f32(c[7] * t9))
    return np.array([f32(f32(f32(c[0] * r) * f32(math.cos(float(psi)))) + c[2]), f32(f32(f32(c[1] * r) * f32(math.sin(float(psi)))) + c[3])], f32)


def triangulate_matches(c1, c2, R12, t12, kp1, kp2, s1, s2):
    """One keypoint pair; returns (depth, p3D, A) with depth = -1 when the reference returns -1."""
    r1, r2 = unproject(c1, *kp1), unproject(c2, *kp2)
    r21 = cv2.gemm(R12, r2.reshape(3, 1), 1.0, None, 0.0).reshape(3)
    dot = float(r1[0]) * float(r21[0]) + float(r1[1]) * float(r21[1]) + float(r1[2]) * float(r21[2])
    cosp = f32(dot / (cv2.norm(r1) * cv2.norm(r21)))
    A = np.zeros((4, 4), f32)
    if float(cosp) > 0.9998:
        return f32(-1), np.zeros(3, f32), A
    T1 = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0]], f32)
    R21 = np.ascontiguousarray(R12.T)
    t21 = -cv2.gemm(R21, t12.reshape(3, 1), 1.0, None, 0.0)
    T2 = np.hstack([R21, t21]).astype(f32)
    A[0] = cv2.addWeighted(T1[2:3], float(r1[0]), T1[0:1], -1.0, 0.0)
    A[1] = cv2.addWeighted(T1[2:3], float(r1[1]), T1[1:2], -1.0, 0.0)
    A[2] = cv2.addWeighted(T2[2:3], float(r2[0]), T2[0:1], -1.0, 0.0)
    A[3] = cv2.addWeighted(T2[2:3], float(r2[1]), T2[1:2], -1.0, 0.0)
    w, u, vt = cv2.SVDecomp(A.copy(), flags=cv2.SVD_FULL_UV)
    inv = f32(1.0 / float(vt[3, 3]))
    x = np.array([f32(vt[3, 0] * inv), f32(vt[3, 1] * inv), f32(vt[3, 2] * inv)], f32)
    if x[2] <= 0:
        return f32(-1), np.zeros(3, f32), A
    z2 = f32(float(R21[2, 0]) * float(x[0]) + float(R21[2, 1]) * float(x[1]) + float(R21[2, 2]) * float(x[2]) + float(t21[2, 0]))
    if z2 <= 0:
        return f32(-1), np.zeros(3, f32), A
    uv = project(c1, x)
    ex, ey = f32(uv[0] - kp1[0]), f32(uv[1] - kp1[1])
    if float(f32(f32(ex * ex) + f32(ey * ey))) > 5.991 * float(s1):
        return f32(-1), np.zeros(3, f32), A
    x2 = (cv2.gemm(R21, x.reshape(3, 1), 1.0, None, 0.0) + t21).astype(f32).reshape(3)
    uv = project(c2, x2)
    ex, ey = f32(uv[0] - kp2[0]), f32(uv[1] - kp2[1])
    if float(f32(f32(ex * ex) + f32(ey * ey))) > 5.991 * float(s2):
        return f32(-1), np.zeros(3, f32), A
    return x[2], x, A


def make_pairs(rng, n):
    """Keypoint pairs as a fisheye stereo frame produces them: random scene points seen by both cameras, pixel noise of a fraction
    of a pixel up to several pixels (so that both reprojection tests accept and reject), points at infinity (parallax test),
    and mismatched pairs."""
    R12, t12 = np.ascontiguousarray(TLR[:, :3]), np.ascontiguousarray(TLR[:, 3])
    R21 = R12.T.astype(np.float64); t21 = -R21 @ t12.astype(np.float64)
    kp1, kp2, s1, s2 = [], [], [], []
    sig = [f32(1.2) ** (2 * l) for l in range(8)]
    while len(kp1) < n:
        kind = rng.randint(10)
        depth = 10 ** rng.uniform(-0.5, 1.7) if kind else 1e4          # 0.3 m .. 50 m, or practically at infinity
        d = rng.randn(3); d[2] = abs(d[2]) + 0.35; d /= np.linalg.norm(d)
        X1 = d * depth
        X2 = R21 @ X1 + t21
        if X2[2] <= 0.05:
            continue
        a = project(CAM1, X1.astype(f32)).astype(np.float64)
        b = project(CAM2, X2.astype(f32)).astype(np.float64)
        noise = [0.05, 0.3, 1.0, 3.0][rng.randint(4)]
        a += rng.randn(2) * noise; b += rng.randn(2) * noise
        if kind == 9:
            b = rng.uniform(20, 490, 2)                                   # a wrong match
        if not (16 <= a[0] < 496 and 16 <= a[1] < 496 and 16 <= b[0] < 496 and 16 <= b[1] < 496):
            continue
        o1, o2 = rng.randint(8), rng.randint(8)
        # keypoint coordinates are level pixels times the level's scale factor: floats, as the extractor emits them
        kp1.append(np.array(a, f32)); kp2.append(np.array(b, f32)); s1.append(sig[o1]); s2.append(sig[o2])
    return R12, t12, np.array(kp1, f32), np.array(kp2, f32), np.array(s1, f32), np.array(s2, f32)


def main():
    rng = np.random.RandomState(20261107)
    rec = {"cam1": CAM1, "cam2": CAM2}
    # 1. the SVD itself: random matrices, rank-deficient ones, and the triangulation matrices of the cases below
    mats = [rng.randn(4, 4).astype(f32) for _ in range(300)]
    for i in range(100):
        m = rng.randn(4, 4).astype(f32)
        m[rng.randint(4)] *= f32(10.0 ** rng.uniform(-6, -1))
        mats.append(m)
    R12, t12, kp1, kp2, s1, s2 = make_pairs(rng, 600)
    depth, p3d = np.zeros(len(kp1), f32), np.zeros((len(kp1), 3), f32)
    for i in range(len(kp1)):
        depth[i], p3d[i], A = triangulate_matches(CAM1, CAM2, R12, t12, kp1[i], kp2[i], s1[i], s2[i])
        if A.any():
            mats.append(A)
    mats = np.array(mats, f32)
    ws, vts = np.zeros((len(mats), 4), f32), np.zeros((len(mats), 4, 4), f32)
    for i, m in enumerate(mats):
        w, u, vt = cv2.SVDecomp(m.copy(), flags=cv2.SVD_FULL_UV)
        ws[i], vts[i] = w.reshape(4), vt
    rec.update(svd_A=mats, svd_w=ws, svd_vt=vts, R12=R12, t12=t12, kp1=kp1, kp2=kp2, sigma1=s1, sigma2=s2, depth=depth, p3d=p3d)
    # 2. unproject / project samples
    uv = rng.uniform(0, 512, (200, 2)).astype(f32)
    rec["unproject_uv"] = uv
    rec["unproject_ray"] = np.array([unproject(CAM1, *p) for p in uv], f32)
    pts = (rng.randn(200, 3) * [2, 2, 1] + [0, 0, 3]).astype(f32)
    rec["project_p"] = pts
    rec["project_uv"] = np.array([project(CAM2, p) for p in pts], f32)
    out = os.path.join(ROOT, "tests", "golden", "triangulate.npz")
    np.savez_compressed(out, **rec)
    print("wrote", out, "accepted", int((depth > 0).sum()), "of", len(depth), "; svd cases", len(mats))


if __name__ == "__main__":
    main()
