#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the ORB front-end with the REAL OpenCV primitives (cv2 4.13.0).

The reference (/root/reference/src/ORBextractor.cc, Frame.cc) cannot be compiled in this container (no OpenCV
C++ headers), so the pin for the CPU oracle is this black-box pipeline: every pixel primitive the reference
calls is executed by cv2 itself (cv2.resize, copyMakeBorder, FastFeatureDetector, GaussianBlur, fastAtan2,
norm, BFMatcher.knnMatch); only the glue of ORBextractor.cc / Frame.cc (cell loop, quadtree, moments, tap
rotation, output order, stereo search) is restated here in Python, independently of oracle/*.cpp.

Run in the build container:  python tools/make_goldens.py     (needs cv2; does NOT need /root/reference)
Outputs are committed; tests/test_oracle_golden.py checks oracle/liborb_oracle.so against them.
"""
import ctypes
import hashlib
import math
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam3_detailed_comments_kor_b200 import synth  # noqa: E402

f32 = np.float32
libm = ctypes.CDLL("libm.so.6")
libm.cosf.restype = ctypes.c_float
libm.cosf.argtypes = [ctypes.c_float]
libm.sinf.restype = ctypes.c_float
libm.sinf.argtypes = [ctypes.c_float]

EDGE = 19
_txt = open(os.path.join(ROOT, "include/orbfe_pattern31.inc")).read().split("*/")[1]
PATTERN = np.array([int(v) for v in _txt.replace("\n", "").split(",") if v.strip()], np.int32).reshape(512, 2)


def cv_round(v):
    return int(np.rint(f32(v)))


class CvExtractor:
    """ORBextractor (src/ORBextractor.cc:408-1177) with cv2 doing the pixel work."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        sf = float(f32(scale_factor))  # member is double, initialised from the float argument
        self.scale = [f32(1.0)]
        for i in range(1, nlevels):
            self.scale.append(f32(float(self.scale[-1]) * sf))
        self.inv_scale = [f32(1.0) / s for s in self.scale]
        factor = f32(1.0 / sf)
        nd = f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        self.quota, tot = [], 0
        for _ in range(nlevels - 1):
            self.quota.append(cv_round(nd))
            tot += self.quota[-1]
            nd = f32(nd * factor)
        self.quota.append(max(nfeatures - tot, 0))
        um = [0] * 16
        vmax = int(math.floor(f32(15) * f32(math.sqrt(f32(2.0))) / f32(2) + f32(1)))
        vmin = int(math.ceil(f32(15) * f32(math.sqrt(f32(2.0))) / f32(2)))
        for v in range(vmax + 1):
            um[v] = int(np.rint(math.sqrt(225.0 - v * v)))
        v0 = 0
        for v in range(15, vmin - 1, -1):
            while um[v0] == um[v0 + 1]:
                v0 += 1
            um[v] = v0
            v0 += 1
        self.umax = um
        self.fast = {t: cv2.FastFeatureDetector_create(threshold=t, nonmaxSuppression=True,
                                                       type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
                     for t in (ini_th, min_th)}

    # ORBextractor.cc:1152-1177
    def compute_pyramid(self, image):
        self.padded, self.pyr = [], []
        for l in range(self.nlevels):
            s = self.inv_scale[l]
            w, h = cv_round(f32(image.shape[1]) * s), cv_round(f32(image.shape[0]) * s)
            if l:
                lvl = cv2.resize(self.pyr[l - 1], (w, h), interpolation=cv2.INTER_LINEAR)
                temp = cv2.copyMakeBorder(lvl, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101 | cv2.BORDER_ISOLATED)
            else:
                temp = cv2.copyMakeBorder(image, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
            self.padded.append(temp)
            self.pyr.append(temp[EDGE:EDGE + h, EDGE:EDGE + w])

    # ORBextractor.cc:479-535 + 537-761, lists as Python lists (front = index 0), tie-break = creation order
    def distribute(self, keys, minX, maxX, minY, maxY, N):
        # std::round(float): half away from zero
        nIni = int(math.floor(float(f32(maxX - minX) / f32(maxY - minY)) + 0.5))
        hX = f32(maxX - minX) / f32(nIni)
        counter = [0]

        def mk(ulx, uly, brx, bry):
            counter[0] += 1
            return {"ul": (ulx, uly), "br": (brx, bry), "keys": [], "nomore": False, "id": counter[0]}

        nodes = [mk(int(hX * f32(i)), 0, int(hX * f32(i + 1)), maxY - minY) for i in range(nIni)]
        for k in keys:
            nodes[int(f32(k[0]) / hX)]["keys"].append(k)
        nodes = [n for n in nodes if n["keys"]]
        for n in nodes:
            if len(n["keys"]) == 1:
                n["nomore"] = True

        def divide(n):
            ulx, uly = n["ul"]
            brx, bry = n["br"]
            hx = int(math.ceil(float(f32(brx - ulx) / f32(2))))
            hy = int(math.ceil(float(f32(bry - uly) / f32(2))))
            ch = [mk(ulx, uly, ulx + hx, uly + hy), mk(ulx + hx, uly, brx, uly + hy),
                  mk(ulx, uly + hy, ulx + hx, bry), mk(ulx + hx, uly + hy, brx, bry)]
            mx, my = ulx + hx, uly + hy
            for k in n["keys"]:
                if k[0] < mx:
                    (ch[0] if k[1] < my else ch[2])["keys"].append(k)
                else:
                    (ch[1] if k[1] < my else ch[3])["keys"].append(k)
            for c in ch:
                if len(c["keys"]) == 1:
                    c["nomore"] = True
            return ch

        finish = False
        while not finish:
            prev = len(nodes)
            nexp = 0
            size_node = []
            i = 0
            while i < len(nodes):
                n = nodes[i]
                if n["nomore"]:
                    i += 1
                    continue
                for c in divide(n):
                    if c["keys"]:
                        nodes.insert(0, c)
                        i += 1
                        if len(c["keys"]) > 1:
                            nexp += 1
                            size_node.append((len(c["keys"]), c["id"], c))
                del nodes[i]
            if len(nodes) >= N or len(nodes) == prev:
                finish = True
            elif len(nodes) + nexp * 3 > N:
                while not finish:
                    prev = len(nodes)
                    pv = sorted(size_node, key=lambda t: (t[0], t[1]))
                    size_node = []
                    for j in range(len(pv) - 1, -1, -1):
                        n = pv[j][2]
                        for c in divide(n):
                            if c["keys"]:
                                nodes.insert(0, c)
                                if len(c["keys"]) > 1:
                                    size_node.append((len(c["keys"]), c["id"], c))
                        nodes.remove(n)
                        if len(nodes) >= N:
                            break
                    if len(nodes) >= N or len(nodes) == prev:
                        finish = True
        out = []
        for n in nodes:
            best = n["keys"][0]
            for k in n["keys"][1:]:
                if k[2] > best[2]:
                    best = k
            out.append(best)
        return out

    # ORBextractor.cc:763-878
    def compute_keypoints(self):
        self.cand, self.level_kp = [], []
        W = f32(35)
        for l in range(self.nlevels):
            im = self.pyr[l]
            minBX = minBY = EDGE - 3
            maxBX, maxBY = im.shape[1] - EDGE + 3, im.shape[0] - EDGE + 3
            width, height = f32(maxBX - minBX), f32(maxBY - minBY)
            nCols, nRows = int(width / W), int(height / W)
            wCell, hCell = int(math.ceil(float(width / f32(nCols)))), int(math.ceil(float(height / f32(nRows))))
            todist = []
            for i in range(nRows):
                iniY = minBY + i * hCell
                maxY = iniY + hCell + 6
                if iniY >= maxBY - 3:
                    continue
                maxY = min(maxY, maxBY)
                for j in range(nCols):
                    iniX = minBX + j * wCell
                    maxX = iniX + wCell + 6
                    if iniX >= maxBX - 6:
                        continue
                    maxX = min(maxX, maxBX)
                    roi = np.ascontiguousarray(im[iniY:maxY, iniX:maxX])
                    kps = self.fast[self.ini_th].detect(roi)
                    if not kps:
                        kps = self.fast[self.min_th].detect(roi)
                    for kp in kps:
                        todist.append((kp.pt[0] + j * wCell, kp.pt[1] + i * hCell, kp.response))
            self.cand.append(todist)
            sel = self.distribute(todist, minBX, maxBX, minBY, maxBY, self.quota[l])
            size = float(int(f32(31) * self.scale[l]))
            kl = []
            for (x, y, r) in sel:
                x, y = x + minBX, y + minBY
                kl.append([x, y, size, self.ic_angle(self.pyr[l], x, y), r, l])
            self.level_kp.append(kl)

    # ORBextractor.cc:75-102
    def ic_angle(self, im, x, y):
        cx, cy = cv_round(x), cv_round(y)
        m01 = m10 = 0
        for u in range(-15, 16):
            m10 += u * int(im[cy, cx + u])
        for v in range(1, 16):
            d = self.umax[v]
            vs = 0
            for u in range(-d, d + 1):
                vp, vm = int(im[cy + v, cx + u]), int(im[cy - v, cx + u])
                vs += vp - vm
                m10 += u * (vp + vm)
            m01 += v * vs
        return cv2.fastAtan2(float(m01), float(m10))

    # ORBextractor.cc:106-145
    def descriptor(self, blurred, x, y, angle):
        ang = f32(angle) * f32(math.pi / 180.0)
        a, b = f32(libm.cosf(ang)), f32(libm.sinf(ang))
        px, py = PATTERN[:, 0].astype(f32), PATTERN[:, 1].astype(f32)
        ry = np.rint(px * b + py * a).astype(np.int64)
        rx = np.rint(px * a - py * b).astype(np.int64)
        vals = blurred[cv_round(y) + ry, cv_round(x) + rx].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
        return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").reshape(32)

    # ORBextractor.cc:1068-1150
    def __call__(self, image, lapping):
        self.compute_pyramid(image)
        self.compute_keypoints()
        n = sum(len(k) for k in self.level_kp)
        kps = np.zeros((n, 6), np.float64)
        desc = np.zeros((n, 32), np.uint8)
        mono, stereo = 0, n - 1
        self.blurred = []
        for l in range(self.nlevels):
            if not self.level_kp[l]:
                self.blurred.append(None)
                continue
            work = cv2.GaussianBlur(self.pyr[l].copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            self.blurred.append(work)
            for (x, y, size, ang, r, o) in self.level_kp[l]:
                d = self.descriptor(work, x, y, ang)
                if l:
                    x, y = f32(x) * self.scale[l], f32(y) * self.scale[l]
                if f32(x) >= lapping[0] and f32(x) <= lapping[1]:
                    slot = stereo
                    stereo -= 1
                else:
                    slot = mono
                    mono += 1
                kps[slot] = (x, y, size, ang, r, o)
                desc[slot] = d
        return mono, kps, desc


def hamming(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def stereo_matches(exL, exR, kL, dL, kR, dR, mbf, maxD):
    """Frame::ComputeStereoMatches (src/Frame.cc:797-967) with cv2.norm for the SAD; minD = 0, maxD explicit."""
    N = len(kL)
    uR, depth = np.full(N, -1, f32), np.full(N, -1, f32)
    nRows = exL.pyr[0].shape[0]
    rows = [[] for _ in range(nRows)]
    for iR in range(len(kR)):
        y = f32(kR[iR][1])
        r = f32(2.0) * exR.scale[int(kR[iR][5])]
        for yi in range(int(math.floor(float(y - r))), int(math.ceil(float(y + r))) + 1):
            rows[yi].append(iR)
    minD, maxD = f32(0), f32(maxD)
    dist_idx = []
    for iL in range(N):
        uL, vL, lvl = f32(kL[iL][0]), f32(kL[iL][1]), int(kL[iL][5])
        cands = rows[int(vL)]
        if not cands:
            continue
        minU, maxU = uL - maxD, uL - minD
        if maxU < 0:
            continue
        best, bestR = 100, 0
        for iR in cands:
            o = int(kR[iR][5])
            if o < lvl - 1 or o > lvl + 1:
                continue
            u = f32(kR[iR][0])
            if minU <= u <= maxU:
                d = hamming(dL[iL], dR[iR])
                if d < best:
                    best, bestR = d, iR
        if best < 75:
            uR0 = f32(kR[bestR][0])
            sf = exL.inv_scale[lvl]
            rnd = lambda v: f32(math.floor(float(v) + 0.5)) if v >= 0 else f32(-math.floor(-float(v) + 0.5))
            su, sv, sr = rnd(uL * sf), rnd(vL * sf), rnd(uR0 * sf)
            w = Lh = 5
            IL = exL.pyr[lvl][int(sv - w):int(sv + w + 1), int(su - w):int(su + w + 1)]
            iniu, endu = sr + Lh - w, sr + Lh + w + 1
            if iniu < 0 or endu >= exR.pyr[lvl].shape[1]:
                continue
            bestD, bestInc, dists = 2 ** 31 - 1, 0, [f32(0)] * 11
            for inc in range(-Lh, Lh + 1):
                IR = exR.pyr[lvl][int(sv - w):int(sv + w + 1), int(sr + inc - w):int(sr + inc + w + 1)]
                d = f32(cv2.norm(IL, IR, cv2.NORM_L1))
                if d < bestD:
                    bestD, bestInc = int(d), inc
                dists[Lh + inc] = d
            if bestInc in (-Lh, Lh):
                continue
            d1, d2, d3 = dists[Lh + bestInc - 1], dists[Lh + bestInc], dists[Lh + bestInc + 1]
            with np.errstate(divide="ignore", invalid="ignore"):
                delta = (d1 - d3) / (f32(2.0) * (d1 + d3 - f32(2.0) * d2))
            if np.isnan(delta) or delta < -1 or delta > 1:
                continue  # (a NaN passes both C++ comparisons but then fails the disparity range test: same outcome)
            bestu = exL.scale[lvl] * (f32(sr) + f32(bestInc) + delta)
            disp = uL - bestu
            if disp >= minD and disp < maxD:
                if disp <= 0:
                    disp = f32(0.01)
                    bestu = f32(float(uL) - 0.01)
                depth[iL] = f32(mbf) / disp
                uR[iL] = bestu
                dist_idx.append((bestD, iL))
    if dist_idx:
        dist_idx.sort()
        median = f32(dist_idx[len(dist_idx) // 2][0])
        th = f32(f32(1.5) * f32(1.4)) * median
        for i in range(len(dist_idx) - 1, -1, -1):
            if dist_idx[i][0] < th:
                break
            uR[dist_idx[i][1]] = -1
            depth[dist_idx[i][1]] = -1
    return uR, depth


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def special_image(kind, w, h, seed):
    if kind == "synth":
        return synth.stereo_pair(seed, w, h)
    rng = np.random.RandomState(seed)  # stored in the npz, so the RNG need not be reproducible elsewhere
    if kind == "noise":
        a = rng.randint(0, 256, (h, w)).astype(np.uint8)
        return a, np.roll(a, -7, axis=1)
    if kind == "lowcontrast":
        L, R = synth.stereo_pair(seed, w, h)
        return (L // 6 + 100).astype(np.uint8), (R // 6 + 100).astype(np.uint8)
    raise ValueError(kind)


CASES = [
    # name, kind, W, H, seed, (nfeat, sf, nlevels, ini, min), lappings, stereo(mbf, maxD) or None, knn
    ("euroc", "synth", 752, 480, 0, (1200, 1.2, 8, 20, 7), [(0, 0), (200, 400)], (47.90639, 435.2047), False),
    ("tumvi", "synth", 512, 512, 1, (1500, 1.2, 8, 20, 7), [(0, 511)], None, True),
    ("kitti", "synth", 1241, 376, 2, (2000, 1.2, 8, 20, 7), [(0, 0), (0, 1000)], (386.1448, 718.856), False),
    ("small", "synth", 160, 120, 3, (300, 1.2, 4, 20, 7), [(0, 0)], None, False),
    ("noise", "noise", 320, 240, 4, (500, 1.2, 8, 20, 7), [(0, 0), (0, 1000)], None, True),
    ("lowc", "lowcontrast", 400, 300, 5, (600, 1.2, 6, 20, 7), [(0, 1000)], (40.0, 300.0), False),
]


def py_search_by_projection(c):
    """Independent restatement (plain Python, written from /root/reference/src/ORBmatcher.cc:44-214 and src/Frame.cc:380-411,
    643-721, not from oracle/project_oracle.cpp) of the projection-gated search on a synth/projection_cases.py case."""
    kps, desc, q, qd = c["kps"], c["desc"], c["queries"], c["qdesc"]
    n, nleft = len(kps), int(c["n_left"][0])
    minx, miny, invw, invh = [f32(v) for v in c["grid"]]
    ur, partner = c.get("u_right"), c.get("partner")
    taken = c["taken"].astype(bool).copy()
    th_high, ratio = int(c["th_high"][0]), f32(c["nnratio"][0])
    check_ori = int(c["check_orientation"][0]) != 0
    sig = c.get("inv_level_sigma2")
    rot_hist = [[] for _ in range(30)]
    popc = np.array([bin(i).count("1") for i in range(256)], np.int32)
    cells = {}
    for i in range(n):
        px = (f32(kps["x"][i]) - minx) * invw
        py = (f32(kps["y"][i]) - miny) * invh
        cx, cy = int(math.floor(abs(px) + 0.5) * (1 if px >= 0 else -1)), int(math.floor(abs(py) + 0.5) * (1 if py >= 0 else -1))   # C round()
        if 0 <= cx < 64 and 0 <= cy < 48:
            cells.setdefault((1 if (nleft >= 0 and i >= nleft) else 0, cx, cy), []).append(i)
    res = np.zeros((len(q), 7), np.int32)
    nmatches = 0
    skip_right_half = False
    left_window_empty = False
    for k in range(len(q)):
        x, y, r, xr = f32(q["x"][k]), f32(q["y"][k]), f32(q["radius"][k]), f32(q["xr"][k])
        minl, maxl, flags = int(q["min_level"][k]), int(q["max_level"][k]), int(q["flags"][k])
        res[k] = (-1, 256, -1, 256, -1, 0, 0)
        drop = (bool(flags & 8) and skip_right_half) or (bool(flags & 32) and left_window_empty)
        skip_right_half = False
        left_window_empty = False
        if flags & 4:
            res[k, 6] = 4
            continue
        if drop:
            res[k, 6] = 6
            continue
        side = 1 if flags & 1 else 0
        cand = []
        x0 = max(0, int(math.floor((x - minx - r) * invw)))
        x1 = min(63, int(math.ceil((x - minx + r) * invw)))
        y0 = max(0, int(math.floor((y - miny - r) * invh)))
        y1 = min(47, int(math.ceil((y - miny + r) * invh)))
        if x0 < 64 and x1 >= 0 and y0 < 48 and y1 >= 0:
            for ix in range(x0, x1 + 1):
                for iy in range(y0, y1 + 1):
                    for i in cells.get((side, ix, iy), []):
                        o = int(kps["octave"][i])
                        if (minl > 0 or maxl >= 0) and (o < minl or (maxl >= 0 and o > maxl)):
                            continue
                        if abs(f32(kps["x"][i]) - x) < r and abs(f32(kps["y"][i]) - y) < r:
                            cand.append(i)
        if not cand:
            res[k, 6] = 8
            left_window_empty = not side
            continue
        best, blvl, best2, blvl2, bidx = 256, -1, 256, -1, -1
        for i in cand:
            if taken[i]:
                continue
            if (flags & 64) and sig is not None:                 # Fuse: chi-square gate on the reprojection error (ORBmatcher.cc:1714-1740)
                ex, ey = x - f32(kps["x"][i]), y - f32(kps["y"][i])
                lim = 5.99
                e2 = f32(f32(ex * ex) + f32(ey * ey))
                if ur is not None and ur[i] >= 0:
                    er = xr - f32(ur[i])
                    e2 = f32(e2 + f32(er * er))
                    lim = 7.8
                if float(f32(e2 * f32(sig[int(kps["octave"][i])]))) > lim:
                    continue
            elif nleft < 0 and ur is not None and ur[i] > 0 and abs(xr - f32(ur[i])) > r:
                continue
            d = int(popc[desc[i] ^ qd[k]].sum())
            if d < best:
                best2, blvl2 = best, blvl
                best, blvl, bidx = d, int(kps["octave"][i]), i
            elif d < best2:
                best2, blvl2 = d, int(kps["octave"][i])
        res[k, :5] = (bidx, best, blvl, best2, blvl2)
        if bidx < 0:
            continue
        if best > th_high:
            res[k, 6] = 2
            continue
        if not (flags & 16) and blvl == blvl2 and f32(best) > ratio * f32(best2):
            res[k, 6] = 3
            skip_right_half = not side
            continue
        res[k, 5], res[k, 6] = 1, 1
        nmatches += 1
        if flags & 2:
            taken[bidx] = True
        if nleft >= 0 and partner is not None and partner[bidx] >= 0:
            nmatches += 1
            if flags & 2:
                taken[partner[bidx]] = True
        if check_ori:
            rot = f32(q["angle"][k]) - f32(kps["angle"][bidx])
            if rot < 0:
                rot = f32(rot + f32(360))
            v = f32(rot * f32(f32(1) / f32(30)))
            b = int(math.floor(abs(float(v)) + 0.5))              # C round(): half away from zero (v >= 0 here)
            if b == 30:
                b = 0
            rot_hist[b].append(k)
    if check_ori:
        sizes = [len(h) for h in rot_hist]
        ind, mx = [-1, -1, -1], [0, 0, 0]
        for i, sz in enumerate(sizes):                            # ComputeThreeMaxima
            if sz > mx[0]:
                mx = [sz, mx[0], mx[1]]; ind = [i, ind[0], ind[1]]
            elif sz > mx[1]:
                mx = [mx[0], sz, mx[1]]; ind = [ind[0], i, ind[1]]
            elif sz > mx[2]:
                mx[2] = sz; ind[2] = i
        if f32(mx[1]) < f32(0.1) * f32(mx[0]):
            ind[1] = ind[2] = -1
        elif f32(mx[2]) < f32(0.1) * f32(mx[0]):
            ind[2] = -1
        for i in range(30):
            if i not in ind:
                for k in rot_hist[i]:
                    res[k, 5], res[k, 6] = 0, 7
                    taken[res[k, 0]] = False
                    nmatches -= 2 if (nleft >= 0 and partner is not None and partner[res[k, 0]] >= 0) else 1
    return res, taken.astype(np.uint8), nmatches


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, kind, W, H, seed, params, laps, stereo, knn in CASES:
        L, R = special_image(kind, W, H, seed)
        ex = CvExtractor(*params)
        rec = {"params": np.array(params, np.float64), "wh": np.array([W, H]), "seed": np.array([seed]),
               "kind": np.array([kind]), "quota": np.array(ex.quota), "umax": np.array(ex.umax),
               "scale": np.array(ex.scale, f32)}
        if kind != "synth":
            rec["imgL"], rec["imgR"] = L, R
        else:
            rec["imgL_sha"], rec["imgR_sha"] = np.array([sha(L)]), np.array([sha(R)])
        for li, lap in enumerate(laps):
            mono, kps, desc = ex(L, lap)
            rec["lap%d" % li] = np.array(lap)
            rec["mono%d" % li] = np.array([mono])
            rec["kps%d" % li] = kps.astype(f32)
            rec["desc%d" % li] = desc
        rec["pyr_sha"] = np.array([sha(p) for p in ex.padded])
        rec["blur_sha"] = np.array([sha(b) if b is not None else "" for b in ex.blurred])
        rec["level_wh"] = np.array([[p.shape[1], p.shape[0]] for p in ex.pyr])
        for l in range(ex.nlevels):
            rec["cand%d" % l] = np.array(ex.cand[l], np.int32).reshape(-1, 3)
            rec["lkp%d" % l] = np.array(ex.level_kp[l], f32).reshape(-1, 6)
        print(name, "n =", len(rec["kps0"]), "cands", [len(c) for c in ex.cand])
        if stereo or knn:
            exR = CvExtractor(*params)
            monoR, kR, dR = exR(R, laps[-1])
            mono, kL, dL = ex(L, laps[-1])
            rec["kpsR"], rec["descR"], rec["monoR"] = kR.astype(f32), dR, np.array([monoR])
        if stereo:
            uR, depth = stereo_matches(ex, exR, kL, dL, kR, dR, stereo[0], stereo[1])
            rec["stereo_params"] = np.array(stereo, np.float64)
            rec["uRight"], rec["depth"] = uR, depth
            print("   stereo matches:", int((uR >= 0).sum()))
        if knn:
            bf = cv2.BFMatcher(cv2.NORM_HAMMING)
            ms = bf.knnMatch(dL[mono:], dR[monoR:], 2)
            m = np.full(len(ms), -1, np.int32)
            d01 = np.full((len(ms), 2), -1, np.int32)
            for i, it in enumerate(ms):
                if len(it) >= 2:
                    d01[i] = (int(it[0].distance), int(it[1].distance))
                    if it[0].distance < it[1].distance * 0.7:
                        m[i] = it[0].trainIdx
            rec["knn_match"], rec["knn_d01"] = m, d01
            print("   knn accepted:", int((m >= 0).sum()), "of", len(m))
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **rec)

    # primitive-level pins on random inputs (cv2 outputs stored)
    rng = np.random.RandomState(7)
    prim = {}
    src = rng.randint(0, 256, (97, 131)).astype(np.uint8)
    prim["resize_src"] = src
    for k, (dw, dh) in enumerate([(109, 81), (66, 49), (131, 97), (130, 96)]):
        prim["resize_wh%d" % k] = np.array([dw, dh])
        prim["resize_out%d" % k] = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
    prim["blur_out"] = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    roi = (rng.randint(0, 256, (47, 57)) // 3 * 3).astype(np.uint8)
    sm = cv2.GaussianBlur(roi, (5, 5), 1.2)
    prim["fast_roi"] = sm
    for t in (20, 7):
        det = cv2.FastFeatureDetector_create(threshold=t, nonmaxSuppression=True, type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        prim["fast_out%d" % t] = np.array([(k.pt[0], k.pt[1], k.response) for k in det.detect(sm)], np.int32).reshape(-1, 3)
    yx = rng.randint(-40000, 40000, (4000, 2)).astype(f32)
    yx[:16] = [[0, 0], [0, 1], [1, 0], [0, -1], [-1, 0], [1, 1], [-1, -1], [1, -1], [-1, 1], [5, 5], [3, 4], [-3, 4], [4, -3], [1e-3, 1], [1, 1e-3], [7, 7]]
    prim["atan_yx"] = yx
    prim["atan_out"] = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in yx], f32)
    np.savez_compressed(os.path.join(out_dir, "primitives.npz"), **prim)

    # input conditioning: cv2.remap (float maps, INTER_LINEAR) and cv2.cvtColor to gray
    cond = {}
    H, W = 120, 163
    src = rng.randint(0, 256, (H, W)).astype(np.uint8)
    yy, xx = np.mgrid[0:H, 0:W].astype(f32)
    k1, k2 = f32(-0.28), f32(0.07)                        # a radial-distortion-like map plus jitter; parts fall outside the image
    xn, yn = (xx - f32(W / 2)) / f32(140), (yy - f32(H / 2)) / f32(140)
    r2 = xn * xn + yn * yn
    fac = f32(1) + k1 * r2 + k2 * r2 * r2
    cond["src"] = src
    cond["mapx"] = (xn * fac * f32(158) + f32(W / 2) + f32(7.3) + rng.uniform(-0.4, 0.4, (H, W))).astype(f32)
    cond["mapy"] = (yn * fac * f32(158) + f32(H / 2) - f32(5.7) + rng.uniform(-0.4, 0.4, (H, W))).astype(f32)
    cond["remap_out"] = cv2.remap(src, cond["mapx"], cond["mapy"], cv2.INTER_LINEAR)
    col = rng.randint(0, 256, (61, 83, 4)).astype(np.uint8)
    cond["color"] = col
    cond["gray_rgb"] = cv2.cvtColor(np.ascontiguousarray(col[..., :3]), cv2.COLOR_RGB2GRAY)
    cond["gray_bgr"] = cv2.cvtColor(np.ascontiguousarray(col[..., :3]), cv2.COLOR_BGR2GRAY)
    cond["gray_rgba"] = cv2.cvtColor(col, cv2.COLOR_RGBA2GRAY)
    cond["gray_bgra"] = cv2.cvtColor(col, cv2.COLOR_BGRA2GRAY)
    # CLAHE exactly as the TUM-VI drivers build it: cv::createCLAHE(3.0, cv::Size(8, 8))
    clahe = cv2.createCLAHE(clipLimit=3.0, tileGridSize=(8, 8))
    L, _ = synth.stereo_pair(1, 512, 512)                 # the tumvi case's left image (dark-ish version: CLAHE has work to do)
    dark = (L // 3 + 20).astype(np.uint8)
    cond["clahe_tumvi_src_sha"] = np.array([sha(dark)])
    cond["clahe_tumvi_sha"] = np.array([sha(clahe.apply(dark))])
    E, _ = synth.stereo_pair(0, 752, 480)
    cond["clahe_euroc_sha"] = np.array([sha(clahe.apply(E))])
    for tag, (h, w) in (("a", (70, 100)), ("b", (64, 100)), ("c", (61, 88))):    # sizes that do not divide by 8 (one or both ways)
        im = np.clip(rng.normal(120, 30, (h, w)), 0, 255).astype(np.uint8)
        im[5:25, 10:50] = np.clip(im[5:25, 10:50].astype(np.int32) + 90, 0, 255)
        im[40:, 60:] //= 4
        cond["clahe_src_" + tag] = im
        cond["clahe_out_" + tag] = clahe.apply(im)
    flat = np.full((64, 64), 77, np.uint8)                # a constant image: every histogram is one clipped bin
    cond["clahe_flat_out"] = clahe.apply(flat)
    c2 = cv2.createCLAHE(clipLimit=40.0, tileGridSize=(4, 6))
    cond["clahe_out_d"] = c2.apply(cond["clahe_src_a"])   # other parameters: clip 40, 4 x 6 tiles (cv::Size(4, 6): 4 across)
    np.savez_compressed(os.path.join(out_dir, "conditioning.npz"), **cond)
    # projection-gated search: inputs from synth/projection_cases.py, expected output from the Python restatement
    from orb_slam3_detailed_comments_kor_b200.synth import projection_cases
    proj = {}
    for name, kw in projection_cases.CASES.items():
        c = projection_cases.make_case(**kw)
        res, taken, nm = py_search_by_projection(c)
        for k, v in c.items():
            proj[name + "/" + k] = v
        proj[name + "/expect"] = res
        proj[name + "/taken_after"] = taken
        proj[name + "/nmatches"] = np.array([nm], np.int32)
        print("projection", name, "nmatches", nm, "accepted", int(res[:, 5].sum()))
    np.savez_compressed(os.path.join(out_dir, "projection.npz"), **proj)
    print("wrote goldens to", out_dir, "cv2", cv2.__version__)


if __name__ == "__main__":
    main()
