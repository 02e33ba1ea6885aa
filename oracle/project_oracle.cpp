/* project_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Projection-gated search of Tracking::SearchLocalPoints, restating
 *   /root/reference/src/Frame.cc:380-411    Frame::AssignFeaturesToGrid
 *   /root/reference/src/Frame.cc:643-709    Frame::GetFeaturesInArea
 *   /root/reference/src/Frame.cc:711-721    Frame::PosInGrid
 *   /root/reference/src/ORBmatcher.cc:44-214 ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, ...)
 *   /root/reference/src/ORBmatcher.cc:2193-2419 ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)
 *   /root/reference/src/ORBmatcher.cc:2545-2586 ORBmatcher::ComputeThreeMaxima
 *   /root/reference/src/ORBmatcher.cc:1643-1841 ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight): the search (PQ_ERROR_GATE queries)
 * on plain arrays, in the reference's serial order (a map point that was matched blocks its keypoint for the map points
 * after it).  No third-party arithmetic is involved.  PINNED against the reference's own bodies, cut out of src/ORBmatcher.cc, src/Frame.cc
 * and src/KeyFrame.cc and compiled into oracle/_ref/liborb_ref.so (tests/test_ref_pinned.py, tests/test_initialization.py): the
 * SearchLocalPoints, frame-to-frame, relocalisation and both Sim3 SearchByProjection overloads, both Fuse functions, SearchBySim3 and
 * SearchForInitialization;
 * an independent Python restatement in tools/make_goldens.py (tests/golden/projection.npz) besides. */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "orb_oracle.h"

namespace {
const int GRID_COLS = 64, GRID_ROWS = 48; /* include/Frame.h:37-38 */

struct Grid {
    std::vector<int> cell[2][GRID_COLS][GRID_ROWS]; /* [0] = mGrid, [1] = mGridRight; entries are rows of the kps array */
};
}  // namespace

extern "C" {

/* Layout-identical to orbfe_frame_grid / orbfe_proj_query / orbfe_proj_match (include/orbfe.h). */
typedef struct { float min_x, min_y, inv_w, inv_h; } orc_frame_grid;
typedef struct { float x, y, radius, xr; int32_t min_level, max_level, flags; float angle; } orc_proj_query;
typedef struct { int32_t best_idx, best_dist, best_level, second_dist, second_level, accepted, status; } orc_proj_match;

enum { PQ_RIGHT = 1, PQ_BLOCKS = 2, PQ_SKIP = 4, PQ_AFTER_LEFT = 8, PQ_BEST_ONLY = 16, PQ_NEEDS_LEFT_WINDOW = 32, PQ_ERROR_GATE = 64 };
enum { PS_NO_CANDIDATES = 0, PS_ACCEPTED = 1, PS_OVER_THRESHOLD = 2, PS_RATIO_REJECTED = 3, PS_SKIPPED = 4, PS_DROPPED = 6, PS_ROTATION = 7, PS_EMPTY_WINDOW = 8 };
const int HISTO_LENGTH = 30; /* ORBmatcher.cc:38 */

/* kps: n rows (rows >= n_left belong to the right camera when n_left >= 0); taken (in/out): rows whose map point has
 * observations; partner: row matched in the other camera or -1 (may be NULL); u_right may be NULL.
 * Returns nmatches as the reference counts them (a matched row with a partner counts twice). */
int orc_search_by_projection(const orc_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right, const int32_t* partner,
                             uint8_t* taken, const orc_frame_grid* g, const orc_proj_query* q, const uint8_t* qdesc, int nq, int th_high,
                             float nnratio, int check_orientation, const float* inv_level_sigma2, orc_proj_match* out) {
    Grid* grid = new Grid();
    /* AssignFeaturesToGrid + PosInGrid */
    for (int i = 0; i < n; i++) {
        const int posX = (int)std::round((kps[i].x - g->min_x) * g->inv_w), posY = (int)std::round((kps[i].y - g->min_y) * g->inv_h);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid->cell[(n_left >= 0 && i >= n_left) ? 1 : 0][posX][posY].push_back(i);
    }
    int nmatches = 0;
    bool prevRatioRejected = false, prevWindowEmpty = false;
    std::vector<int> rotHist[HISTO_LENGTH];            /* query indices; the reference stores the keypoint row */
    const float factor = 1.0f / HISTO_LENGTH;
    for (int iq = 0; iq < nq; iq++) {
        const orc_proj_query& p = q[iq];
        orc_proj_match& o = out[iq];
        o.best_idx = -1; o.best_dist = 256; o.best_level = -1; o.second_dist = 256; o.second_level = -1; o.accepted = 0;
        const bool dropped = ((p.flags & PQ_AFTER_LEFT) && prevRatioRejected) || ((p.flags & PQ_NEEDS_LEFT_WINDOW) && prevWindowEmpty);
        prevRatioRejected = false;
        prevWindowEmpty = false;
        if (p.flags & PQ_SKIP) { o.status = PS_SKIPPED; continue; }
        if (dropped) { o.status = PS_DROPPED; continue; }
        const bool right = (p.flags & PQ_RIGHT) != 0;
        /* GetFeaturesInArea */
        std::vector<int> vIndices;
        const float x = p.x, y = p.y, r = p.radius;
        do {
            const int nMinCellX = std::max(0, (int)std::floor((x - g->min_x - r) * g->inv_w));
            if (nMinCellX >= GRID_COLS) break;
            const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - g->min_x + r) * g->inv_w));
            if (nMaxCellX < 0) break;
            const int nMinCellY = std::max(0, (int)std::floor((y - g->min_y - r) * g->inv_h));
            if (nMinCellY >= GRID_ROWS) break;
            const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - g->min_y + r) * g->inv_h));
            if (nMaxCellY < 0) break;
            const bool bCheckLevels = (p.min_level > 0) || (p.max_level >= 0);
            for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                    for (int idx : grid->cell[right ? 1 : 0][ix][iy]) {
                        if (bCheckLevels) {
                            if (kps[idx].octave < p.min_level) continue;
                            if (p.max_level >= 0 && kps[idx].octave > p.max_level) continue;
                        }
                        const float distx = kps[idx].x - x, disty = kps[idx].y - y;
                        if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back(idx);
                    }
        } while (false);
        if (vIndices.empty()) { o.status = PS_EMPTY_WINDOW; prevWindowEmpty = !right; continue; }
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (taken[idx]) continue;
            if ((p.flags & PQ_ERROR_GATE) && inv_level_sigma2) {          /* ORBmatcher::Fuse, ORBmatcher.cc:1714-1740 */
                const float kpx = kps[idx].x, kpy = kps[idx].y;
                const float ex = p.x - kpx, ey = p.y - kpy;
                if (u_right && u_right[idx] >= 0) {
                    const float er = p.xr - u_right[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * inv_level_sigma2[kps[idx].octave] > 7.8) continue;
                } else {
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * inv_level_sigma2[kps[idx].octave] > 5.99) continue;
                }
            } else if (n_left < 0 && u_right && u_right[idx] > 0) {
                const float er = std::fabs(p.xr - u_right[idx]);
                if (er > r) continue;
            }
            const int dist = orc_descriptor_distance(qdesc + (size_t)iq * 32, desc + (size_t)idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kps[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = kps[idx].octave; bestDist2 = dist;
            }
        }
        o.best_idx = bestIdx; o.best_dist = bestDist; o.best_level = bestLevel; o.second_dist = bestDist2; o.second_level = bestLevel2;
        if (bestIdx < 0) { o.status = PS_NO_CANDIDATES; continue; } /* every candidate was taken or failed the uRight gate */
        if (bestDist > th_high) { o.status = PS_OVER_THRESHOLD; continue; }
        if (!(p.flags & PQ_BEST_ONLY) && bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) {
            o.status = PS_RATIO_REJECTED;
            prevRatioRejected = !right; /* the `continue` at ORBmatcher.cc:117-118 also skips the right-camera half of this map point */
            continue;
        }
        o.status = PS_ACCEPTED; o.accepted = 1;
        nmatches++;
        if (p.flags & PQ_BLOCKS) taken[bestIdx] = 1;
        if (n_left >= 0 && partner && partner[bestIdx] >= 0) {
            nmatches++;
            if (p.flags & PQ_BLOCKS) taken[partner[bestIdx]] = 1;
        }
        if (check_orientation) {                               /* ORBmatcher.cc:2305-2322 */
            float rot = p.angle - kps[bestIdx].angle;
            if (rot < 0.0) rot += 360.0f;
            int bin = (int)std::round(rot * factor);
            if (bin == HISTO_LENGTH) bin = 0;
            if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(iq);
        }
    }
    if (check_orientation) {                                   /* ORBmatcher.cc:2391-2412 with ComputeThreeMaxima (:2545-2586) */
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int iq : rotHist[i]) {
                    out[iq].accepted = 0; out[iq].status = PS_ROTATION;
                    taken[out[iq].best_idx] = 0;               /* mvpMapPoints[row] = NULL */
                    nmatches -= 1 + ((n_left >= 0 && partner && partner[out[iq].best_idx] >= 0) ? 1 : 0);   /* what it had added */
                }
    }
    delete grid;
    return nmatches;
}

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), /root/reference/src/ORBmatcher.cc:706-821
 * (the matcher of the monocular initialiser, Tracking.cc:2462), on plain arrays.  kps1 / kps2 are F1.mvKeysUn / F2.mvKeysUn.
 * The loop is serial: a level-0 keypoint of F1 takes the best level-0 keypoint of F2 inside the window around its previous
 * match, a keypoint of F2 that is already matched at a distance <= the new one is invisible (:745), a better match displaces
 * the old one (:764-768); the rotation histogram counts every acceptance, displaced ones included (:783).
 * prev_matched [n1][2] is updated in place (:816-818); matches12 [n1] receives vnMatches12.  Returns nmatches. */
int orc_search_for_initialization(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const orc_keypoint* kps2, const uint8_t* desc2, int n2,
                                  const orc_frame_grid* g, float* prev_matched, int window, float nnratio, int check_orientation, int th_low,
                                  int32_t* matches12) {
    Grid* grid = new Grid();
    for (int i = 0; i < n2; i++) {
        const int posX = (int)std::round((kps2[i].x - g->min_x) * g->inv_w), posY = (int)std::round((kps2[i].y - g->min_y) * g->inv_h);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid->cell[0][posX][posY].push_back(i);
    }
    int nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vMatchedDistance(n2, INT32_MAX), vnMatches21(n2, -1);
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = kps1[i1].octave;
        if (level1 > 0) continue;
        std::vector<int> vIndices2;
        const float x = prev_matched[2 * i1], y = prev_matched[2 * i1 + 1], r = (float)window;
        do {   /* GetFeaturesInArea(x, y, windowSize, level1, level1) */
            const int nMinCellX = std::max(0, (int)std::floor((x - g->min_x - r) * g->inv_w));
            if (nMinCellX >= GRID_COLS) break;
            const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - g->min_x + r) * g->inv_w));
            if (nMaxCellX < 0) break;
            const int nMinCellY = std::max(0, (int)std::floor((y - g->min_y - r) * g->inv_h));
            if (nMinCellY >= GRID_ROWS) break;
            const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - g->min_y + r) * g->inv_h));
            if (nMaxCellY < 0) break;
            const bool bCheckLevels = (level1 > 0) || (level1 >= 0);
            for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                    for (int idx : grid->cell[0][ix][iy]) {
                        if (bCheckLevels) {
                            if (kps2[idx].octave < level1) continue;
                            if (level1 >= 0 && kps2[idx].octave > level1) continue;
                        }
                        const float distx = kps2[idx].x - x, disty = kps2[idx].y - y;
                        if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices2.push_back(idx);
                    }
        } while (false);
        if (vIndices2.empty()) continue;
        int bestDist = INT32_MAX, bestDist2 = INT32_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = orc_descriptor_distance(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= th_low) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (check_orientation) {
                    float rot = kps1[i1].angle - kps2[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (matches12[i1] >= 0) { prev_matched[2 * i1] = kps2[matches12[i1]].x; prev_matched[2 * i1 + 1] = kps2[matches12[i1]].y; }
    delete grid;
    return nmatches;
}

}  /* extern "C" */
