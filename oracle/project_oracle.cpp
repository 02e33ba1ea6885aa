/* project_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Projection-gated search of Tracking::SearchLocalPoints, restating
 *   /root/reference/src/Frame.cc:380-411    Frame::AssignFeaturesToGrid
 *   /root/reference/src/Frame.cc:643-709    Frame::GetFeaturesInArea
 *   /root/reference/src/Frame.cc:711-721    Frame::PosInGrid
 *   /root/reference/src/ORBmatcher.cc:44-214 ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, ...)
 * on plain arrays, in the reference's serial order (a map point that was matched blocks its keypoint for the map points
 * after it).  No third-party arithmetic is involved; parity unpinned beyond the independent Python restatement in
 * tools/make_goldens.py (tests/golden/projection.npz). */
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
typedef struct { float x, y, radius, xr; int32_t min_level, max_level, flags; } orc_proj_query;
typedef struct { int32_t best_idx, best_dist, best_level, second_dist, second_level, accepted, status; } orc_proj_match;

enum { PQ_RIGHT = 1, PQ_BLOCKS = 2, PQ_SKIP = 4, PQ_AFTER_LEFT = 8 };
enum { PS_NO_CANDIDATES = 0, PS_ACCEPTED = 1, PS_OVER_THRESHOLD = 2, PS_RATIO_REJECTED = 3, PS_SKIPPED = 4, PS_DROPPED = 6 };

/* kps: n rows (rows >= n_left belong to the right camera when n_left >= 0); taken (in/out): rows whose map point has
 * observations; partner: row matched in the other camera or -1 (may be NULL); u_right may be NULL.
 * Returns nmatches as the reference counts them (a matched row with a partner counts twice). */
int orc_search_by_projection(const orc_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right, const int32_t* partner,
                             uint8_t* taken, const orc_frame_grid* g, const orc_proj_query* q, const uint8_t* qdesc, int nq, int th_high,
                             float nnratio, orc_proj_match* out) {
    Grid* grid = new Grid();
    /* AssignFeaturesToGrid + PosInGrid */
    for (int i = 0; i < n; i++) {
        const int posX = (int)std::round((kps[i].x - g->min_x) * g->inv_w), posY = (int)std::round((kps[i].y - g->min_y) * g->inv_h);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid->cell[(n_left >= 0 && i >= n_left) ? 1 : 0][posX][posY].push_back(i);
    }
    int nmatches = 0;
    bool prevRatioRejected = false;
    for (int iq = 0; iq < nq; iq++) {
        const orc_proj_query& p = q[iq];
        orc_proj_match& o = out[iq];
        o.best_idx = -1; o.best_dist = 256; o.best_level = -1; o.second_dist = 256; o.second_level = -1; o.accepted = 0;
        const bool dropped = (p.flags & PQ_AFTER_LEFT) && prevRatioRejected;
        prevRatioRejected = false;
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
        if (vIndices.empty()) { o.status = PS_NO_CANDIDATES; continue; }
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (taken[idx]) continue;
            if (n_left < 0 && u_right && u_right[idx] > 0) {
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
        if (bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) {
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
    }
    delete grid;
    return nmatches;
}

}  /* extern "C" */
