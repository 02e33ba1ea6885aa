/* bow_oracle.cpp — CPU ORACLE (test infrastructure only; see orb_oracle.h for the rules).
 * Bag-of-words assignment of Frame::ComputeBoW, restating
 *   /root/reference/src/Frame.cc:724-731                                     mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)
 *   /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1125-1193   transform(features, BowVector&, FeatureVector&, levelsup)
 *   /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1217-1260   transform(feature, word_id, weight, nid, levelsup)
 *   /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1426   loadFromTextFile (node / word numbering)
 *   /root/reference/Thirdparty/DBoW2/DBoW2/BowVector.cpp:34-84, FeatureVector.cpp:31-45, FORB.cpp:81-101, ScoringObject.h:73-90
 * The vocabulary file (Vocabulary/ORBvoc.txt) is not in /root/reference, so the tests run on synthetic trees.  PINNED: the transform against
 * the reference's vendored DBoW2 compiled into oracle/_ref/libdbow_ref.so (bit-identical doubles), SearchByBoW x2 and SearchForTriangulation
 * against the reference's own bodies compiled into oracle/_ref/liborb_ref.so (tests/test_ref_pinned.py, tests/test_ref_triangulation.py);
 * an independent Python restatement in tests/test_bow.py besides. */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "orb_oracle.h"

namespace {
struct Node {
    std::vector<int> children;
    uint8_t descriptor[32];
    double weight = 0;
    int word_id = 0;
    bool isLeaf() const { return children.empty(); }
};
}  // namespace

struct orc_vocabulary {
    int k, L, scoring, weighting, nwords;
    std::vector<Node> nodes;
};

extern "C" {

/* Rows as in the text file: node id = row + 1 (the root is node 0), `parent`, `is_leaf`, 32 descriptor bytes, weight; words
 * are numbered in the order their rows appear. */
orc_vocabulary* orc_vocabulary_create(int k, int L, int scoring, int weighting, int nrows, const int32_t* parent, const uint8_t* is_leaf,
                                      const uint8_t* desc, const double* weight) {
    orc_vocabulary* v = new orc_vocabulary();
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting; v->nwords = 0;
    v->nodes.resize(1);
    for (int r = 0; r < nrows; r++) {
        const int nid = (int)v->nodes.size();
        v->nodes.resize(nid + 1);
        if (parent[r] < 0 || parent[r] >= nid) { delete v; return nullptr; }
        v->nodes[parent[r]].children.push_back(nid);
        memcpy(v->nodes[nid].descriptor, desc + (size_t)r * 32, 32);
        v->nodes[nid].weight = weight[r];
        if (is_leaf[r]) v->nodes[nid].word_id = v->nwords++;
    }
    return v;
}

void orc_vocabulary_destroy(orc_vocabulary* v) { delete v; }
int orc_vocabulary_words(const orc_vocabulary* v) { return v->nwords; }

/* word_ids / node_ids / weights: per feature (also for stopped words); bow_*: the BowVector in key order; fv_*: the FeatureVector
 * in key order, node e owns fv_feats[fv_starts[e] .. fv_starts[e+1]). */
int orc_bow_transform(const orc_vocabulary* voc, const uint8_t* features, int n, int levelsup, int32_t* word_ids, int32_t* node_ids,
                      int32_t* bow_words, double* bow_values, int32_t* nbow, int32_t* fv_nodes, int32_t* fv_starts, int32_t* fv_feats,
                      int32_t* nfv) {
    std::map<unsigned, double> v;
    std::map<unsigned, std::vector<unsigned>> fv;
    *nbow = 0; *nfv = 0;
    if (voc->nodes.size() <= 1) return 0;   /* empty() */
    const bool must = voc->scoring != 5;    /* DotProductScoring is the only one that does not normalise */
    const bool l2 = voc->scoring == 1;
    for (int i = 0; i < n; i++) {
        const uint8_t* f = features + (size_t)i * 32;
        /* transform(feature, word_id, weight, nid, levelsup) */
        const int nid_level = voc->L - levelsup;
        unsigned nid = 0;
        int final_id = 0, current_level = 0;
        do {
            ++current_level;
            const std::vector<int>& nodes = voc->nodes[final_id].children;
            final_id = nodes[0];
            double best_d = orc_descriptor_distance(f, voc->nodes[final_id].descriptor);
            for (size_t c = 1; c < nodes.size(); c++) {
                const double d = orc_descriptor_distance(f, voc->nodes[nodes[c]].descriptor);
                if (d < best_d) { best_d = d; final_id = nodes[c]; }
            }
            if (current_level == nid_level) nid = final_id;
        } while (!voc->nodes[final_id].isLeaf());
        const unsigned id = voc->nodes[final_id].word_id;
        const double w = voc->nodes[final_id].weight;
        if (word_ids) word_ids[i] = (int32_t)id;
        if (node_ids) node_ids[i] = (int32_t)nid;
        if (w > 0) {
            if (voc->weighting == 0 || voc->weighting == 1) {   /* TF_IDF, TF: addWeight */
                auto it = v.lower_bound(id);
                if (it != v.end() && it->first == id) it->second += w;
                else v.insert(it, std::make_pair(id, w));
            } else {                                            /* IDF, BINARY: addIfNotExist */
                if (v.find(id) == v.end()) v[id] = w;
            }
            fv[nid].push_back((unsigned)i);
        }
    }
    if ((voc->weighting == 0 || voc->weighting == 1) && !v.empty() && !must) {
        const double nd = (double)v.size();
        for (auto& e : v) e.second /= nd;
    }
    if (must) {
        double norm = 0.0;
        if (!l2) { for (auto& e : v) norm += std::fabs(e.second); }
        else { for (auto& e : v) norm += e.second * e.second; norm = std::sqrt(norm); }
        if (norm > 0.0) for (auto& e : v) e.second /= norm;
    }
    int k = 0;
    for (auto& e : v) { bow_words[k] = (int32_t)e.first; bow_values[k] = e.second; k++; }
    *nbow = k;
    int e = 0, pos = 0;
    for (auto& nd : fv) {
        fv_nodes[e] = (int32_t)nd.first;
        fv_starts[e] = pos;
        for (unsigned i : nd.second) fv_feats[pos++] = (int32_t)i;
        e++;
    }
    fv_starts[e] = pos;
    *nfv = e;
    return 0;
}

/* ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*>&), src/ORBmatcher.cc:269-470, the F.Nleft == -1 path, with
 * ComputeThreeMaxima (:2545-2586); kf_kf selects SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (:823-963, NLeft == -1): candidates
 * need a map point too and the threshold is strict.  The feature vectors are the (nodes, starts, feats) arrays of orc_bow_transform.
 * match_f[i] = keyframe row whose map point lands in vpMapPointMatches[i], or -1; kf_out (optional): per keyframe row the scan's
 * best1/best2/idx and accepted (1 kept, 2 removed by the rotation histogram).  Returns nmatches. */
int orc_search_by_bow(const uint8_t* desc_kf, const orc_keypoint* kps_kf, int n_kf, const int32_t* nodes_kf, const int32_t* starts_kf, const int32_t* feats_kf,
                      int nfv_kf, const uint8_t* has_map_point, const uint8_t* desc_f, const orc_keypoint* kps_f, int n_f, const int32_t* nodes_f,
                      const int32_t* starts_f, const int32_t* feats_f, int nfv_f, const uint8_t* has_map_point_f, int n_left_f, int th_low,
                      float nnratio, int check_orientation, int kf_kf, int32_t* match_f, orc_match* kf_out) {
    const int HISTO_LENGTH = 30;
    for (int i = 0; i < n_f; i++) match_f[i] = -1;
    if (kf_out) for (int i = 0; i < n_kf; i++) { kf_out[i].best1 = 256; kf_out[i].best2 = 256; kf_out[i].idx = -1; kf_out[i].accepted = 0; }
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int KFit = 0, Fit = 0;
    while (KFit != nfv_kf && Fit != nfv_f) {
        if (nodes_kf[KFit] == nodes_f[Fit]) {
            for (int iKF = starts_kf[KFit]; iKF < starts_kf[KFit + 1]; iKF++) {
                const int realIdxKF = feats_kf[iKF];
                if (has_map_point && !has_map_point[realIdxKF]) continue;
                const uint8_t* dKF = desc_kf + (size_t)realIdxKF * 32;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
                for (int iF = starts_f[Fit]; iF < starts_f[Fit + 1]; iF++) {
                    const int realIdxF = feats_f[iF];
                    if (match_f[realIdxF] >= 0) continue;
                    if (kf_kf && has_map_point_f && !has_map_point_f[realIdxF]) continue;   /* :886-892 (keyframe-keyframe overload) */
                    const int dist = orc_descriptor_distance(dKF, desc_f + (size_t)realIdxF * 32);
                    if (n_left_f < 0 || kf_kf) {                                            /* F.Nleft == -1, :314-329 */
                        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                        else if (dist < bestDist2) { bestDist2 = dist; }
                    } else {                                                                /* :330-358 */
                        if (realIdxF < n_left_f && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                        else if (realIdxF < n_left_f && dist < bestDist2) { bestDist2 = dist; }
                        if (realIdxF >= n_left_f && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = realIdxF; }
                        else if (realIdxF >= n_left_f && dist < bestDist2R) { bestDist2R = dist; }
                    }
                }
                if (!kf_kf && n_left_f >= 0 && bestDist1 <= th_low && bestDist1R <= th_low) {   /* :388-416, `|| true` */
                    match_f[bestIdxFR] = realIdxKF;
                    if (check_orientation) {
                        float rot = kps_kf[realIdxKF].angle - kps_f[bestIdxFR].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(bestIdxFR);
                    }
                    nmatches++;
                }
                bool ok = false;
                if ((kf_kf ? bestDist1 < th_low : bestDist1 <= th_low) && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
                    ok = true;
                    match_f[bestIdxF] = realIdxKF;
                    if (check_orientation) {
                        float rot = kps_kf[realIdxKF].angle - kps_f[bestIdxF].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(bestIdxF);
                    }
                    nmatches++;
                }
                if (kf_out) { kf_out[realIdxKF].best1 = bestDist1; kf_out[realIdxKF].best2 = bestDist2; kf_out[realIdxKF].idx = bestIdxF; kf_out[realIdxKF].accepted = ok; }
            }
            KFit++; Fit++;
        } else if (nodes_kf[KFit] < nodes_f[Fit]) {
            while (KFit != nfv_kf && nodes_kf[KFit] < nodes_f[Fit]) KFit++;   /* lower_bound */
        } else {
            while (Fit != nfv_f && nodes_f[Fit] < nodes_kf[KFit]) Fit++;
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int c = (int)rotHist[i].size();
            if (c > max1) { max3 = max2; max2 = max1; max1 = c; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (c > max2) { max3 = max2; max2 = c; ind3 = ind2; ind2 = i; }
            else if (c > max3) { max3 = c; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idxF : rotHist[i]) {
                if (kf_out) kf_out[match_f[idxF]].accepted = 2;
                match_f[idxF] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

/* layout of orbfe_triangulation_rig (include/orbfe.h): cameras 1 left / right, 2 left / right; poses ll, lr, rl, rr */
typedef struct { float cam[4][8]; float R[4][9], t[4][3]; int32_t n_left_1, n_left_2; } orc_triangulation_rig;
void orc_kb8_triangulate_matches(const float* cam1, const float* cam2, const float* R12, const float* t12, const float* uv1, const float* uv2,
                                 const float* sigma1, const float* sigma2, int n, float* depth, float* p3d);   /* triangulate_oracle.cpp */

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, bCoarse), src/ORBmatcher.cc:965-1206, for pinhole
 * keyframes without a second camera, with Pinhole::epipolarConstrain (src/CameraModels/Pinhole.cpp:135-157) evaluated on the given
 * F12 (row-major 3x3).  match_12[i] = vMatches12[i].  Returns nmatches. */
int orc_search_for_triangulation(const uint8_t* desc_1, const orc_keypoint* kps_1, const float* u_right_1, const uint8_t* has_mp_1, int n_1,
                                 const int32_t* nodes_1, const int32_t* starts_1, const int32_t* feats_1, int nfv_1, const uint8_t* desc_2,
                                 const orc_keypoint* kps_2, const float* u_right_2, const uint8_t* has_mp_2, int n_2, const int32_t* nodes_2,
                                 const int32_t* starts_2, const int32_t* feats_2, int nfv_2, const float* F12, const float* ep,
                                 const float* scale_factors, const float* level_sigma2, int th_low, int only_stereo, int coarse,
                                 int check_orientation, int32_t* match_12, const orc_triangulation_rig* rig) {
    const int HISTO_LENGTH = 30;
    std::vector<bool> vbMatched2(n_2, false);   /* declared and tested by the reference, never set */
    for (int i = 0; i < n_1; i++) match_12[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int f1it = 0, f2it = 0;
    while (f1it != nfv_1 && f2it != nfv_2) {
        if (nodes_1[f1it] == nodes_2[f2it]) {
            for (int i1 = starts_1[f1it]; i1 < starts_1[f1it + 1]; i1++) {
                const int idx1 = feats_1[i1];
                if (has_mp_1 && has_mp_1[idx1]) continue;
                const bool bStereo1 = !rig && u_right_1 && u_right_1[idx1] >= 0;   /* !pKF1->mpCamera2 && mvuRight[idx1] >= 0 */
                if (only_stereo && !bStereo1) continue;
                const orc_keypoint& kp1 = kps_1[idx1];
                const uint8_t* d1 = desc_1 + (size_t)idx1 * 32;
                int bestDist = th_low, bestIdx2 = -1;
                for (int i2 = starts_2[f2it]; i2 < starts_2[f2it + 1]; i2++) {
                    const int idx2 = feats_2[i2];
                    if (vbMatched2[idx2] || (has_mp_2 && has_mp_2[idx2])) continue;
                    const bool bStereo2 = !rig && u_right_2 && u_right_2[idx2] >= 0;
                    if (only_stereo && !bStereo2) continue;
                    const int dist = orc_descriptor_distance(d1, desc_2 + (size_t)idx2 * 32);
                    if (dist > th_low || dist > bestDist) continue;
                    const orc_keypoint& kp2 = kps_2[idx2];
                    if (!bStereo1 && !bStereo2 && !rig) {
                        const float distex = ep[0] - kp2.x, distey = ep[1] - kp2.y;
                        if (distex * distex + distey * distey < 100 * scale_factors[kp2.octave]) continue;
                    }
                    bool ok = coarse != 0;
                    if (!ok && rig) {   /* two-camera rig (:1102-1132): KannalaBrandt8::epipolarConstrain = TriangulateMatches(...) > 0.0001f */
                        const int r1 = idx1 >= rig->n_left_1 ? 1 : 0, r2 = idx2 >= rig->n_left_2 ? 1 : 0, combo = 2 * r1 + r2;
                        const float uv1[2] = {kp1.x, kp1.y}, uv2[2] = {kp2.x, kp2.y};
                        float depth = -1.f, p3d[3];
                        orc_kb8_triangulate_matches(rig->cam[r1], rig->cam[2 + r2], rig->R[combo], rig->t[combo], uv1, uv2, &level_sigma2[kp1.octave],
                                                    &level_sigma2[kp2.octave], 1, &depth, p3d);
                        ok = depth > 0.0001f;
                    } else if (!ok) {   /* Pinhole::epipolarConstrain */
                        const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                        const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                        const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                        const float num = a * kp2.x + b * kp2.y + c;
                        const float den = a * a + b * b;
                        if (den != 0) {
                            const float dsqr = num * num / den;
                            ok = dsqr < 3.84 * level_sigma2[kp2.octave];
                        }
                    }
                    if (ok) { bestIdx2 = idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    match_12[idx1] = bestIdx2;
                    nmatches++;
                    if (check_orientation) {
                        float rot = kp1.angle - kps_2[bestIdx2].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(idx1);
                    }
                }
            }
            f1it++; f2it++;
        } else if (nodes_1[f1it] < nodes_2[f2it]) {
            while (f1it != nfv_1 && nodes_1[f1it] < nodes_2[f2it]) f1it++;
        } else {
            while (f2it != nfv_2 && nodes_2[f2it] < nodes_1[f1it]) f2it++;
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int c = (int)rotHist[i].size();
            if (c > max1) { max3 = max2; max2 = max1; max1 = c; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (c > max2) { max3 = max2; max2 = c; ind3 = ind2; ind2 = i; }
            else if (c > max3) { max3 = c; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) { match_12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}

}  /* extern "C" */
