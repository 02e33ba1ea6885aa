/* match_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Matching side: restates the Hamming core of /root/reference/src/ORBmatcher.cc and the stereo
 * matchers of /root/reference/src/Frame.cc. */
#include "orb_oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <utility>
#include <vector>

extern "C" const uint8_t* orc__pyr_data(const orc_extractor* e, int level, int* w, int* h, int* step);
extern "C" float orc__inv_scale(const orc_extractor* e, int level);

namespace {
const int TH_HIGH = 100; /* ORBmatcher.cc:36 */
const int TH_LOW = 50;   /* ORBmatcher.cc:37 */

/* ORBmatcher.cc:2591-2607 — the SWAR popcount is kept as written. */
inline int descriptor_distance(const uint8_t* a, const uint8_t* b) {
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t pa, pb;
        memcpy(&pa, a + 4 * i, 4);
        memcpy(&pb, b + 4 * i, 4);
        unsigned int v = pa ^ pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}
}  // namespace

extern "C" {

int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

/* ORBmatcher.cc:311-375 / 870-908: bestDist1/bestDist2/bestIdx update rule and acceptance tests. */
void orc_hamming_top2(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int th, float ratio,
                      int strict_lt, orc_match* out) {
    for (int i = 0; i < nq; i++) {
        const uint8_t* dq = q + (size_t)i * 32;
        int best1 = 256, best2 = 256, idx = -1;
        for (int64_t j = 0; j < ndb; j++) {
            const int dist = descriptor_distance(dq, db + (size_t)j * 32);
            if (dist < best1) {
                best2 = best1;
                best1 = dist;
                idx = (int)j;
            } else if (dist < best2) {
                best2 = dist;
            }
        }
        const bool thOk = strict_lt ? (best1 < th) : (best1 <= th);
        const bool ok = thOk && (static_cast<float>(best1) < ratio * static_cast<float>(best2));
        out[i].best1 = best1;
        out[i].best2 = best2;
        out[i].idx = idx;
        out[i].accepted = ok ? 1 : 0;
    }
}

/* Frame.cc:797-967 */
int orc_stereo_match(const orc_extractor* exL, const orc_extractor* exR,
                     const orc_keypoint* kpsL, const uint8_t* descL, int N,
                     const orc_keypoint* kpsR, const uint8_t* descR, int Nr,
                     float mbf, float minD, float maxD, float* uRight, float* depth) {
    for (int i = 0; i < N; i++) uRight[i] = depth[i] = -1.0f;
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    int w0, nRows, st0;
    orc__pyr_data(exL, 0, &w0, &nRows, &st0);
    const int nlevels = orc_num_levels(exL);

    std::vector<std::vector<size_t>> rowIdx(nRows);
    for (int iR = 0; iR < Nr; iR++) {
        const float kpY = kpsR[iR].y;
        const float r = 2.0f * orc_scale_factor(exR, kpsR[iR].octave);
        const int maxr = (int)std::ceil(kpY + r);
        const int minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR); /* the reference indexes unchecked */
    }

    std::vector<std::pair<int, int>> distIdx;
    distIdx.reserve(N);

    for (int iL = 0; iL < N; iL++) {
        const orc_keypoint& kpL = kpsL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        if (!(vL >= 0) || (size_t)vL >= (size_t)nRows) continue;
        const std::vector<size_t>& cands = rowIdx[(size_t)vL];
        if (cands.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;

        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        const uint8_t* dL = descL + (size_t)iL * 32;
        for (size_t iC = 0; iC < cands.size(); iC++) {
            const size_t iR = cands[iC];
            const orc_keypoint& kpR = kpsR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL, descR + iR * 32);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdxR = iR;
                }
            }
        }

        if (bestDist < thOrbDist) {
            if (levelL < 0 || levelL >= nlevels) continue;
            const float uR0 = kpsR[bestIdxR].x;
            const float scaleFactor = orc__inv_scale(exL, levelL);
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5;
            int wl, hl, stl, wr, hr, str;
            const uint8_t* pl = orc__pyr_data(exL, levelL, &wl, &hl, &stl);
            const uint8_t* pr = orc__pyr_data(exR, levelL, &wr, &hr, &str);
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);

            int bestDistS = INT_MAX;
            int bestincR = 0;
            const int L = 5;
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= wr) continue;

            for (int incR = -L; incR <= +L; incR++) {
                const int xr0 = (int)(scaleduR0 + incR - w);
                int sad = 0;
                for (int dy = 0; dy < 2 * w + 1; dy++) {
                    const uint8_t* a = pl + (ptrdiff_t)(y0 + dy) * stl + xl0;
                    const uint8_t* b = pr + (ptrdiff_t)(y0 + dy) * str + xr0;
                    for (int dx = 0; dx < 2 * w + 1; dx++) sad += std::abs((int)a[dx] - (int)b[dx]);
                }
                const float dist = (float)sad; /* cv::norm(IL,IR,NORM_L1) -> double -> float */
                if (dist < bestDistS) {
                    bestDistS = (int)dist;
                    bestincR = incR;
                }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;

            const float dist1 = vDists[L + bestincR - 1];
            const float dist2 = vDists[L + bestincR];
            const float dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;

            float bestuR = orc_scale_factor(exL, levelL) * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) {
                    disparity = 0.01;
                    bestuR = uL - 0.01;
                }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                distIdx.push_back(std::pair<int, int>(bestDistS, iL));
            }
        }
    }

    if (distIdx.empty()) return 0; /* defined deviation (iii) */
    std::sort(distIdx.begin(), distIdx.end());
    const float median = (float)distIdx[distIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)distIdx.size();
    for (int i = (int)distIdx.size() - 1; i >= 0; i--) {
        if (distIdx[i].first < thDist) break;
        uRight[distIdx[i].second] = -1;
        depth[distIdx[i].second] = -1;
        kept--;
    }
    return kept;
}

/* Frame.cc:1137-1144: knnMatch(k=2) (two smallest (distance, trainIdx) lexicographically) + ratio in double. */
void orc_knn2_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio,
                    int32_t* match, int32_t* d0o, int32_t* d1o) {
    for (int i = 0; i < nq; i++) {
        int d0 = 256 + 1, d1 = 256 + 1, i0 = -1;
        for (int j = 0; j < nt; j++) {
            const int d = descriptor_distance(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < d0) { d1 = d0; d0 = d; i0 = j; }
            else if (d < d1) { d1 = d; }
        }
        const bool two = nt >= 2;
        const bool ok = two && ((double)(float)d0 < (double)(float)d1 * ratio);
        match[i] = ok ? i0 : -1;
        if (d0o) d0o[i] = (nt >= 1) ? d0 : -1;
        if (d1o) d1o[i] = two ? d1 : -1;
    }
}

/* MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:386-417, for npoints map points (rows offsets[p]..offsets[p+1]). */
void orc_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best_idx, int32_t* best_median) {
    for (int p = 0; p < npoints; p++) {
        const uint8_t* d = desc + (size_t)offsets[p] * 32;
        const size_t N = (size_t)(offsets[p + 1] - offsets[p]);
        best_idx[p] = -1;
        if (best_median) best_median[p] = INT_MAX;
        if (N == 0) continue;
        std::vector<float> Distances(N * N);
        for (size_t i = 0; i < N; i++) {
            Distances[i * N + i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                const int distij = descriptor_distance(d + i * 32, d + j * 32);
                Distances[i * N + j] = distij;
                Distances[j * N + i] = distij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + (i + 1) * N);
            std::sort(vDists.begin(), vDists.end());
            const int median = vDists[0.5 * (N - 1)];
            if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
        }
        best_idx[p] = BestIdx;
        if (best_median) best_median[p] = BestMedian;
    }
}

}  /* extern "C" */
