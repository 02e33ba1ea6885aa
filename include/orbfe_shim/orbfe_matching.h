/* orbfe_matching.h (orbfe shim) — the GPU bodies of the descriptor-matching members of ORB_SLAM3::ORBmatcher and Frame, as free
 * functions in namespace ORB_SLAM3::orbfe that take what those members read (/root/reference/include/ORBmatcher.h:35-111,
 * include/Frame.h:114,304).
 *
 * This header deliberately does NOT declare a class named ORBmatcher and is not called ORBmatcher.h: the reference's own
 * include/ORBmatcher.h (with SearchByProjection x5, SearchByBoW x2, Fuse x2, SearchBySim3, SearchForTriangulation_, ... that
 * Tracking, LocalMapping, LoopClosing, KeyFrame, MapPoint, Sim3Solver and Initializer call) stays in the build untouched; the
 * bodies of a few of its members, and of two Frame members, become one-line calls into the functions below (INTEGRATION.md
 * §3-§5 gives each replacement).  Only include/orbfe_shim/ORBextractor.h replaces a reference header, because it reproduces
 * that header's whole public surface.
 * The full SearchByBoW / SearchForTriangulation / Fuse searches are C-ABI entry points (include/orbfe.h). */
#ifndef ORBFE_SHIM_MATCHING_H
#define ORBFE_SHIM_MATCHING_H

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "ORBextractor.h"

namespace ORB_SLAM3 {
namespace orbfe {

/* Body of ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2591-2607.  One 256-bit pair: eight popcounts on the host — launching
 * a kernel for 32 bytes would cost more than the answer; the batched form is orbfe_descriptor_distances(). */
inline int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
    const uint32_t* pa = a.ptr<uint32_t>();
    const uint32_t* pb = b.ptr<uint32_t>();
    int dist = 0;
    for (int i = 0; i < 8; i++) dist += __builtin_popcount(pa[i] ^ pb[i]);
    return dist;
}

/* The top-2 core of SearchByBoW over whole descriptor matrices (rows x 32, CV_8U): for every query row the two smallest
 * distances, the lowest best row, and the thLow (ORBmatcher::TH_LOW) + nnratio (mfNNratio) verdict.  kfkf selects
 * `best1 < TH_LOW` (src/ORBmatcher.cc:906) instead of `<=` (:373). */
inline int SearchTop2(const cv::Mat& queries, const cv::Mat& database, std::vector<orbfe_match>& out, float nnratio, int thLow = 50, bool kfkf = false,
                      int device = 0) {
    out.resize(queries.rows);
    std::vector<unsigned char> q((size_t)queries.rows * 32), db((size_t)database.rows * 32);
    for (int i = 0; i < queries.rows; i++) std::memcpy(&q[(size_t)i * 32], queries.ptr<unsigned char>(i), 32);
    for (int i = 0; i < database.rows; i++) std::memcpy(&db[(size_t)i * 32], database.ptr<unsigned char>(i), 32);
    if (orbfe_hamming_top2(device, q.data(), queries.rows, db.data(), database.rows, thLow, nnratio, kfkf ? 1 : 0, out.data()) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_hamming_top2: ") + orbfe_last_error());
    int n = 0;
    for (const orbfe_match& m : out) n += m.accepted;
    return n;
}

/* Body of Frame::ComputeStereoMatches (src/Frame.cc:797-967) for the frames the two extractors processed last.
 * The reference's maxD is mbf/mb with mb = mbf/fx (src/Frame.cc:827-829, :173); pass that float as maxD (it equals fx for the
 * EuRoC and KITTI settings).  mvuRight / mvDepth are resized to N and filled (-1 = no match). */
inline void ComputeStereoMatches(ORBextractor* left, ORBextractor* right, int N, float mbf, float maxD,
                                 std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
    mvuRight.assign(N, -1.0f);
    mvDepth.assign(N, -1.0f);
    if (N == 0) return;
    if (orbfe_stereo_match_batch_device(left->Impl(), right->Impl(), 1, mbf, 0.0f, maxD, nullptr) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_stereo_match: ") + orbfe_last_error());
    int n = 0;
    if (orbfe_stereo_result(left->Impl(), 0, mvuRight.data(), mvDepth.data(), N, &n) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_stereo_result: ") + orbfe_last_error());
}

/* Matching step of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1137-1144): knnMatch(k=2) + `d0 < d1*0.7`.
 * match[i] = row of stereoDescRight or -1; triangulation of the survivors stays in KannalaBrandt8 (host). */
inline void FishEyeKnnMatches(const cv::Mat& stereoDescLeft, const cv::Mat& stereoDescRight, std::vector<int>& match, int device = 0) {
    match.assign(stereoDescLeft.rows, -1);
    if (stereoDescLeft.rows == 0) return;
    std::vector<unsigned char> q((size_t)stereoDescLeft.rows * 32), t((size_t)stereoDescRight.rows * 32);
    for (int i = 0; i < stereoDescLeft.rows; i++) std::memcpy(&q[(size_t)i * 32], stereoDescLeft.ptr<unsigned char>(i), 32);
    for (int i = 0; i < stereoDescRight.rows; i++) std::memcpy(&t[(size_t)i * 32], stereoDescRight.ptr<unsigned char>(i), 32);
    if (orbfe_knn2_ratio(device, q.data(), stereoDescLeft.rows, t.data(), stereoDescRight.rows, 0.7, match.data(), nullptr, nullptr) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_knn2_ratio: ") + orbfe_last_error());
}

/* Body of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159) as a whole, for the frame the two extractors have just processed with
 * the lapping interval: knnMatch(k=2) + `d0 < d1*0.7`, KannalaBrandt8::TriangulateMatches per survivor, `depth > 0.0001f`.
 *   cam1 / cam2: mpCamera->mvParameters / mpCamera2->mvParameters (fx, fy, cx, cy, k0..k3); Rlr row-major 3x3 and tlr: mRlr, mtlr.
 * Inside the member the body becomes
 *     ORB_SLAM3::orbfe::ComputeStereoFishEyeMatches(mpORBextractorLeft, mpORBextractorRight, Nleft, Nright, cam1, cam2, Rlr, tlr,
 *                                                   mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, mvStereo3Dpoints);
 * mvStereo3Dpoints receives x, y, z per left keypoint (the reference stores cv::Mat; the caller wraps the triple). */
inline void ComputeStereoFishEyeMatches(ORBextractor* left, ORBextractor* right, int Nleft, int Nright, const float cam1[8], const float cam2[8], const float Rlr[9],
                                        const float tlr[3], std::vector<int>& mvLeftToRightMatch, std::vector<int>& mvRightToLeftMatch,
                                        std::vector<float>& mvDepth, std::vector<float>& mvStereo3Dpoints) {
    mvLeftToRightMatch.assign(Nleft, -1);
    mvRightToLeftMatch.assign(Nright, -1);
    mvDepth.assign(Nleft, -1.0f);
    mvStereo3Dpoints.assign((size_t)Nleft * 3, 0.0f);
    if (Nleft == 0 || Nright == 0) return;
    orbfe_fisheye_rig rig;
    std::memcpy(rig.cam1, cam1, sizeof rig.cam1); std::memcpy(rig.cam2, cam2, sizeof rig.cam2);
    std::memcpy(rig.R12, Rlr, sizeof rig.R12); std::memcpy(rig.t12, tlr, sizeof rig.t12);
    const int cap = orbfe_keypoint_capacity(left->Impl());
    std::vector<int> lr(cap), rl(cap);
    std::vector<float> d(cap), p((size_t)cap * 3);
    if (orbfe_fisheye_stereo_frames(left->Impl(), right->Impl(), 1, &rig, 0.7, lr.data(), rl.data(), d.data(), p.data(), cap) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_fisheye_stereo_frames: ") + orbfe_last_error());
    std::copy(lr.begin(), lr.begin() + Nleft, mvLeftToRightMatch.begin());
    std::copy(rl.begin(), rl.begin() + Nright, mvRightToLeftMatch.begin());
    for (int i = 0; i < Nleft; i++) {
        if (lr[i] < 0) continue;
        mvDepth[i] = d[i];
        for (int k = 0; k < 3; k++) mvStereo3Dpoints[(size_t)i * 3 + k] = p[(size_t)i * 3 + k];
    }
}

/* The loop of ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th, bFarPoints, thFarPoints)
 * (src/ORBmatcher.cc:44-214) — or, with ORBFE_PQ_BEST_ONLY queries and checkOrientation = mbCheckOrientation, of
 * SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono) (:2193-2419) — on the frame's arrays; the grid of
 * Frame::AssignFeaturesToGrid is rebuilt on the GPU.
 *   keys / descriptors: F.mvKeysUn + F.mDescriptors (Nleft == -1) or mvKeys followed by mvKeysRight (fisheye rig);
 *   uRight: &F.mvuRight or NULL;   partner: per row the row matched in the other camera or -1 (see orbfe.h), or NULL;
 *   taken[row] != 0 iff F.mvpMapPoints[row] has observations — updated in place;
 *   queries[k] / queryDescriptors.row(k): one entry per (map point, camera) search the reference loop would start, in its order
 *   (fields in orbfe.h cite the members they come from).
 * The caller then assigns F.mvpMapPoints[out[k].best_idx] = pMP (and the partner row) for every out[k].accepted, in order.
 * Returns nmatches as the reference counts them. */
inline int SearchByProjection(const std::vector<cv::KeyPoint>& keys, const cv::Mat& descriptors, int Nleft, const std::vector<float>* uRight,
                              const std::vector<int>* partner, std::vector<unsigned char>& taken, float mnMinX, float mnMinY,
                              float mfGridElementWidthInv, float mfGridElementHeightInv, const std::vector<orbfe_proj_query>& queries,
                              const cv::Mat& queryDescriptors, float mfNNratio, std::vector<orbfe_proj_match>& out, bool checkOrientation = false,
                              int device = 0, const float* invLevelSigma2 = nullptr /* mvInvLevelSigma2 (16 entries) for ORBFE_PQ_ERROR_GATE queries (Fuse) */,
                              int thHigh = 100 /* TH_HIGH; Fuse passes TH_LOW */) {
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint layout");
    const int n = (int)keys.size(), nq = (int)queries.size();
    out.resize(nq);
    taken.resize(n, 0);
    std::vector<unsigned char> d((size_t)n * 32), qd((size_t)nq * 32);
    for (int i = 0; i < n; i++) std::memcpy(&d[(size_t)i * 32], descriptors.ptr<unsigned char>(i), 32);
    for (int i = 0; i < nq; i++) std::memcpy(&qd[(size_t)i * 32], queryDescriptors.ptr<unsigned char>(i), 32);
    const orbfe_frame_grid grid = {mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv};
    int nmatches = 0;
    if (orbfe_search_by_projection(device, reinterpret_cast<const orbfe_keypoint*>(keys.data()), d.data(), n, Nleft, uRight ? uRight->data() : nullptr,
                                   partner ? partner->data() : nullptr, taken.data(), &grid, queries.data(), qd.data(), nq, thHigh,
                                   mfNNratio, checkOrientation ? 1 : 0, invLevelSigma2, out.data(), &nmatches) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_search_by_projection: ") + orbfe_last_error());
    return nmatches;
}

/* Body of ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12,
 * int windowSize) (src/ORBmatcher.cc:706-821), a drop-in: keys1 / descriptors1 = F1.mvKeysUn / F1.mDescriptors, likewise for F2;
 * the four grid numbers are Frame's static members.  vbPrevMatched is updated as the reference does (:816-818).
 * Inside ORBmatcher::SearchForInitialization the whole body becomes
 *     return ORB_SLAM3::orbfe::SearchForInitialization(F1.mvKeysUn, F1.mDescriptors, F2.mvKeysUn, F2.mDescriptors, Frame::mnMinX, Frame::mnMinY,
 *                Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv, vbPrevMatched, vnMatches12, windowSize, mfNNratio, mbCheckOrientation, TH_LOW); */
inline int SearchForInitialization(const std::vector<cv::KeyPoint>& keys1, const cv::Mat& descriptors1, const std::vector<cv::KeyPoint>& keys2,
                                   const cv::Mat& descriptors2, float mnMinX, float mnMinY, float mfGridElementWidthInv, float mfGridElementHeightInv,
                                   std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize, float mfNNratio,
                                   bool checkOrientation, int thLow = 50, int device = 0) {
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint layout");
    static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "cv::Point2f layout");
    const int n1 = (int)keys1.size(), n2 = (int)keys2.size();
    vnMatches12.assign(n1, -1);
    std::vector<unsigned char> d1((size_t)n1 * 32), d2((size_t)n2 * 32);
    for (int i = 0; i < n1; i++) std::memcpy(&d1[(size_t)i * 32], descriptors1.ptr<unsigned char>(i), 32);
    for (int i = 0; i < n2; i++) std::memcpy(&d2[(size_t)i * 32], descriptors2.ptr<unsigned char>(i), 32);
    const orbfe_frame_grid grid = {mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv};
    int nmatches = 0;
    if (orbfe_search_for_initialization(device, reinterpret_cast<const orbfe_keypoint*>(keys1.data()), d1.data(), n1,
                                        reinterpret_cast<const orbfe_keypoint*>(keys2.data()), d2.data(), n2, &grid,
                                        reinterpret_cast<float*>(vbPrevMatched.data()), windowSize, mfNNratio, checkOrientation ? 1 : 0, thLow,
                                        vnMatches12.data(), &nmatches) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_search_for_initialization: ") + orbfe_last_error());
    return nmatches;
}

/* DBoW2::FeatureVector (a std::map<NodeId, std::vector<unsigned int>>, Thirdparty/DBoW2/DBoW2/FeatureVector.h) — or any ordered map from node id to
 * feature rows — as the three arrays the C ABI takes: node ids ascending, offsets, rows in the map's order. */
template <class FeatureVector>
inline void FlattenFeatureVector(const FeatureVector& fv, std::vector<int32_t>& nodes, std::vector<int32_t>& starts, std::vector<int32_t>& feats) {
    nodes.clear(); feats.clear();
    starts.assign(1, 0);
    for (const auto& kv : fv) {
        nodes.push_back((int32_t)kv.first);
        for (const auto r : kv.second) feats.push_back((int32_t)r);
        starts.push_back((int32_t)feats.size());
    }
}

inline void PackDescriptors(const cv::Mat& descriptors, int n, std::vector<unsigned char>& out) {
    out.resize((size_t)n * 32);
    for (int i = 0; i < n; i++) std::memcpy(&out[(size_t)i * 32], descriptors.ptr<unsigned char>(i), 32);
}

/* The search of ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (src/ORBmatcher.cc:269-470): keysKF /
 * descriptorsKF / fvKF = pKF->mvKeysUn, mDescriptors, mFeatVec; hasMapPointKF[r] = vpMapPointsKF[r] && !vpMapPointsKF[r]->isBad(); the frame
 * likewise with F.mvKeys (followed by mvKeysRight on a fisheye rig, NleftF = F.Nleft).  matchF[i] = keyframe row whose map point row i of the
 * frame receives, or -1; the return value is nmatches.  kfkf = the keyframe-keyframe overload (:823-963; thLow, hasMapPointF as there). */
template <class FeatureVector>
inline int SearchByBoW(const cv::Mat& descriptorsKF, const std::vector<cv::KeyPoint>& keysKF, const FeatureVector& fvKF, const std::vector<unsigned char>& hasMapPointKF,
                       const cv::Mat& descriptorsF, const std::vector<cv::KeyPoint>& keysF, const FeatureVector& fvF, int NleftF, float mfNNratio,
                       bool checkOrientation, std::vector<int>& matchF, int thLow = 50, bool kfkf = false, const std::vector<unsigned char>* hasMapPointF = nullptr,
                       int device = 0) {
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint layout");
    const int nk = (int)keysKF.size(), nf = (int)keysF.size();
    std::vector<unsigned char> dk, df;
    PackDescriptors(descriptorsKF, nk, dk);
    PackDescriptors(descriptorsF, nf, df);
    std::vector<int32_t> nA, sA, fA, nB, sB, fB;
    FlattenFeatureVector(fvKF, nA, sA, fA);
    FlattenFeatureVector(fvF, nB, sB, fB);
    matchF.assign(nf > 0 ? nf : 1, -1);
    int nmatches = 0;
    if (orbfe_search_by_bow(device, dk.data(), reinterpret_cast<const orbfe_keypoint*>(keysKF.data()), nk, nA.data(), sA.data(), fA.data(), (int)nA.size(),
                            hasMapPointKF.empty() ? nullptr : hasMapPointKF.data(), df.data(), reinterpret_cast<const orbfe_keypoint*>(keysF.data()), nf, nB.data(),
                            sB.data(), fB.data(), (int)nB.size(), hasMapPointF && !hasMapPointF->empty() ? hasMapPointF->data() : nullptr, NleftF, thLow, mfNNratio,
                            checkOrientation ? 1 : 0, kfkf ? 1 : 0, matchF.data(), nullptr, &nmatches) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_search_by_bow: ") + orbfe_last_error());
    matchF.resize(nf);
    return nmatches;
}

/* The search of ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, bCoarse) (src/ORBmatcher.cc:965-1206): F12 row-major
 * (Pinhole.cpp:137-140), ep = pKF2->mpCamera->project(C2) (:974-977), scaleFactors / levelSigma2 = pKF2->mvScaleFactors / mvLevelSigma2,
 * hasMapPoint1/2[r] = pKF->GetMapPoint(r) != NULL, uRight1/2 = mvuRight (NULL: monocular).  vMatches12[r] = row of keyframe 2 or -1; the
 * caller builds vMatchedPairs from it (:1196-1203).  rig: the two-camera case (see orbfe.h), NULL for pinhole keyframes. */
template <class FeatureVector>
inline int SearchForTriangulation(const cv::Mat& descriptors1, const std::vector<cv::KeyPoint>& keys1, const std::vector<float>* uRight1,
                                  const std::vector<unsigned char>& hasMapPoint1, const FeatureVector& fv1, const cv::Mat& descriptors2,
                                  const std::vector<cv::KeyPoint>& keys2, const std::vector<float>* uRight2, const std::vector<unsigned char>& hasMapPoint2,
                                  const FeatureVector& fv2, const float F12[9], const float ep[2], const std::vector<float>& scaleFactors,
                                  const std::vector<float>& levelSigma2, bool bOnlyStereo, bool bCoarse, bool checkOrientation, std::vector<int>& vMatches12,
                                  int thLow = 50, const orbfe_triangulation_rig* rig = nullptr, int device = 0) {
    const int n1 = (int)keys1.size(), n2 = (int)keys2.size();
    std::vector<unsigned char> d1, d2;
    PackDescriptors(descriptors1, n1, d1);
    PackDescriptors(descriptors2, n2, d2);
    std::vector<int32_t> nA, sA, fA, nB, sB, fB;
    FlattenFeatureVector(fv1, nA, sA, fA);
    FlattenFeatureVector(fv2, nB, sB, fB);
    vMatches12.assign(n1 > 0 ? n1 : 1, -1);
    int nmatches = 0;
    if (orbfe_search_for_triangulation(device, d1.data(), reinterpret_cast<const orbfe_keypoint*>(keys1.data()), uRight1 ? uRight1->data() : nullptr,
                                       hasMapPoint1.empty() ? nullptr : hasMapPoint1.data(), n1, nA.data(), sA.data(), fA.data(), (int)nA.size(), d2.data(),
                                       reinterpret_cast<const orbfe_keypoint*>(keys2.data()), uRight2 ? uRight2->data() : nullptr,
                                       hasMapPoint2.empty() ? nullptr : hasMapPoint2.data(), n2, nB.data(), sB.data(), fB.data(), (int)nB.size(), F12, ep,
                                       scaleFactors.data(), levelSigma2.data(), (int)std::min(scaleFactors.size(), levelSigma2.size()), thLow, bOnlyStereo ? 1 : 0,
                                       bCoarse ? 1 : 0, checkOrientation ? 1 : 0, vMatches12.data(), &nmatches, rig) != ORBFE_OK)
        throw std::runtime_error(std::string("orbfe_search_for_triangulation: ") + orbfe_last_error());
    vMatches12.resize(n1);
    return nmatches;
}

}  // namespace orbfe
}  // namespace ORB_SLAM3
#endif
