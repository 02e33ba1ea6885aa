/* ref_shim.cpp — C entry points around the REFERENCE'S OWN code, compiled where it lies.  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile (target _ref/liborb_ref.so) compiles /root/reference/src/ORBextractor.cc unmodified against
 * oracle/refstub/ (see cvstub.hpp), and cuts function bodies out of files that cannot be compiled whole
 * (they pull in Eigen, Sophus, g2o, ...; src/ORBmatcher.cc also stops at a syntax error on line 2226):
 *     src/Frame.cc       Frame::ComputeStereoMatches, GetFeaturesInArea, PosInGrid, AssignFeaturesToGrid
 *     src/ORBmatcher.cc  DescriptorDistance, TH_LOW / TH_HIGH, SearchByBoW (both overloads), ComputeThreeMaxima,
 *                        SearchByProjection(Frame&, vpMapPoints, th, ...) with RadiusByViewingCos
 *     src/MapPoint.cc    MapPoint::ComputeDistinctiveDescriptors
 * into _ref/*.inc, which live only under oracle/_ref/ (git-ignored); no reference source is stored in this repository.
 * The `Frame`, `KeyFrame`, `MapPoint` and `ORBmatcher` classes below declare just the members those bodies name
 * (include/Frame.h, include/KeyFrame.h, include/MapPoint.h, include/ORBmatcher.h); DBoW2::FeatureVector is the
 * reference's own (Thirdparty/DBoW2).
 *
 * Allocation: every `new` inside a ref_* call is served from a per-thread bump arena and `delete` is a no-op, so heap
 * addresses grow with creation order.  DistributeOctTree sorts pair<size, ExtractorNode*> (src/ORBextractor.cc:682):
 * equal sizes are ordered by node address, i.e. by whatever the allocator returned; under a monotonic allocator that
 * is creation order, which is the tie rule the oracle and the kernels define.  ref_set_arena(0) switches to plain
 * malloc/free to see what glibc's allocator does with the same inputs.
 */
#include <sys/mman.h>

#include <atomic>
#include <map>
#include <set>
#include <mutex>
#include <new>
#include <tuple>

#include "refstub/cvstub.hpp"
#include "ref_camera.h" /* GeometricCamera, Pinhole, KannalaBrandt8: member bodies from the reference, compiled in ref_camera_shim.cpp */
#include "FeatureVector.h" /* the reference's vendored DBoW2 (Thirdparty/DBoW2/DBoW2): the real FeatureVector type */
#include "ORBextractor.h"  /* the reference's header: -I/root/reference/include */

/* ---- bump arena ------------------------------------------------------------------------------------------------ */
namespace {
constexpr size_t kSlotBytes = size_t(2) << 30; /* virtual, committed lazily */
constexpr int kSlots = 64;
char* g_region = nullptr;
std::atomic<int> g_nextSlot{0};
std::atomic<int> g_useArena{1};
struct Arena {
    char* base = nullptr;
    size_t off = 0, peak = 0;
    bool active = false;
};
thread_local Arena t_arena;

char* region() {
    static char* r = [] {
        void* p = mmap(nullptr, kSlotBytes * kSlots, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        return p == MAP_FAILED ? nullptr : (char*)p;
    }();
    return r;
}

struct ArenaScope {
    bool on;
    ArenaScope() : on(false) {
        if (!g_useArena.load() || t_arena.active) return;
        if (!t_arena.base) {
            g_region = region();
            int s = g_nextSlot.fetch_add(1);
            if (!g_region || s >= kSlots) return; /* more threads than slots: that thread uses malloc */
            t_arena.base = g_region + kSlotBytes * s;
        }
        t_arena.off = 0;
        t_arena.active = on = true;
    }
    ~ArenaScope() {
        if (!on) return;
        t_arena.active = false;
        /* give the pages back so a long bench run does not keep every thread's high-water mark resident */
        if (t_arena.off > (size_t(64) << 20)) madvise(t_arena.base, t_arena.off, MADV_DONTNEED);
    }
};

inline void* arena_new(size_t n) {
    Arena& a = t_arena;
    if (a.active) {
        size_t o = (a.off + 15) & ~size_t(15);
        if (o + n > kSlotBytes) throw std::bad_alloc();
        a.off = o + n;
        if (a.off > a.peak) a.peak = a.off;
        return a.base + o;
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
inline void arena_delete(void* p) noexcept {
    if (g_region && (char*)p >= g_region && (char*)p < g_region + kSlotBytes * kSlots) return;
    std::free(p);
}
}  // namespace

/* Local to this library (the linker script exports ref_* only): the host process keeps its own allocator. */
void* operator new(size_t n) { return arena_new(n); }
void* operator new[](size_t n) { return arena_new(n); }
void operator delete(void* p) noexcept { arena_delete(p); }
void operator delete[](void* p) noexcept { arena_delete(p); }
void operator delete(void* p, size_t) noexcept { arena_delete(p); }
void operator delete[](void* p, size_t) noexcept { arena_delete(p); }

/* ---- the function bodies taken from Frame.cc / ORBmatcher.cc / MapPoint.cc ----------------------------------------- */
using namespace std;
namespace ORB_SLAM3 {

/* only what the cut-out bodies name: include/MapPoint.h (isBad), include/KeyFrame.h and include/Frame.h (the members below) */
class KeyFrame;
class Frame;
class MapPoint {
public:
    bool bad = false;
    bool isBad() { return bad; }
    /* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423) and the members it reads (include/MapPoint.h) */
    void ComputeDistinctiveDescriptors();
    std::map<KeyFrame*, std::tuple<int, int>> mObservations;
    std::mutex mMutexFeatures;
    bool mbBad = false;
    /* the tracking fields SearchByProjection reads (include/MapPoint.h:126-137) */
    bool mbTrackInView = false, mbTrackInViewR = false;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 0, mTrackProjXR = 0, mTrackProjYR = 0, mTrackViewCos = 0, mTrackViewCosR = 0;
    int mnTrackScaleLevel = 0, mnTrackScaleLevelR = 0;
    int nObs = 0;
    cv::Mat mDescriptor;
    int Observations() { return nObs; }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    cv::Mat mWorldPos;
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    /* what ORBmatcher::Fuse reads and calls (include/MapPoint.h); the bookkeeping calls only record what happened */
    bool inKF = false;
    float mfMinDistance = 0, mfMaxDistance = 0;
    cv::Mat mNormalVector;
    int predictedLevel = 0, queryIndex = -1, rowIndex = -1, fusedRow = -1, fusedHow = 0;
    bool IsInKeyFrame(KeyFrame*) { return inKF; }
    std::tuple<int, int> GetIndexInKeyFrame(KeyFrame*) { return std::tuple<int, int>(indexInOther, -1); }   /* SearchBySim3 asks for the index in pKF2 */
    int indexInOther = -1;
    float GetMinDistanceInvariance() { return mfMinDistance; }
    float GetMaxDistanceInvariance() { return mfMaxDistance; }
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    int PredictScale(const float&, KeyFrame*) { return predictedLevel; }
    int PredictScale(const float&, Frame*) { return predictedLevel; }
    void AddObservation(KeyFrame*, int idx) { fusedRow = idx; fusedHow = 1; }
    void Replace(MapPoint* other) {   /* one of the two is the candidate map point, the other the one the keypoint row holds (rowIndex >= 0) */
        MapPoint* q = rowIndex < 0 ? this : other;
        MapPoint* r = rowIndex < 0 ? other : this;
        q->fusedRow = r->rowIndex; q->fusedHow = rowIndex < 0 ? 2 : 3;
    }
};

}  // namespace ORB_SLAM3 (reopened below)
namespace cv {
/* cv::BFMatcher(cv::NORM_HAMMING)::knnMatch as Frame::ComputeStereoFishEyeMatches uses it — OpenCV's code, restated: per query row the k smallest
 * Hamming distances in ascending order, the lower train row first among equals (what orc_knn2_ratio keeps too; pinned against cv2.BFMatcher by
 * the knn goldens, tests/test_oracle_golden.py) */
struct DMatch { int queryIdx, trainIdx, imgIdx; float distance; };
enum { NORM_HAMMING = 6 };
class BFMatcher {
public:
    explicit BFMatcher(int = NORM_HAMMING) {}
    void knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch>>& matches, int k) const {
        matches.assign(q.rows, std::vector<DMatch>());
        for (int i = 0; i < q.rows; i++) {
            std::vector<std::pair<int, int>> best;   /* (distance, train row), ascending, at most k */
            for (int j = 0; j < t.rows; j++) {
                int d = 0;
                const unsigned char *a = q.ptr<unsigned char>(i), *b = t.ptr<unsigned char>(j);
                for (int c = 0; c < 32; c++) d += __builtin_popcount((unsigned)(a[c] ^ b[c]));
                size_t pos = best.size();
                while (pos > 0 && d < best[pos - 1].first) pos--;
                if ((int)pos < k) { best.insert(best.begin() + pos, std::make_pair(d, j)); if ((int)best.size() > k) best.pop_back(); }
            }
            for (const auto& e : best) matches[i].push_back(DMatch{i, e.second, 0, (float)e.first});
        }
    }
};
}  // namespace cv
namespace ORB_SLAM3 {

#define FRAME_GRID_ROWS 48 /* include/Frame.h:37-38 */
#define FRAME_GRID_COLS 64

class Frame {
public:
    void ComputeStereoMatches();
    int N;
    int Nleft = -1;
    float mbf, mb;
    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<float> mvuRight, mvDepth;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    DBoW2::FeatureVector mFeatVec;
    GeometricCamera* mpCamera2 = nullptr;
    /* what ComputeStereoFishEyeMatches reads and fills on top (include/Frame.h) */
    void ComputeStereoFishEyeMatches();
    int monoLeft = 0, monoRight = 0, Nright = 0, mnCloseMPs = 0;
    std::vector<cv::Mat> mvStereo3Dpoints;
    std::vector<float> mvLevelSigma2;
    cv::Mat mRlr, mtlr;
    cv::BFMatcher BFmatcher;
    /* the grid of include/Frame.h:187-189,232-235,296 and what SearchByProjection touches */
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1,
                                          const bool bRight = false) const;
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    void AssignFeaturesToGrid();
    static float mfGridElementWidthInv, mfGridElementHeightInv, mnMinX, mnMinY;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    std::vector<std::size_t> mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
    /* what SearchByProjection(CurrentFrame, LastFrame, ...) reads on top (include/Frame.h) */
    cv::Mat mTcw, mTrl;
    std::vector<bool> mvbOutlier;
    GeometricCamera* mpCamera = nullptr;
    static float mnMaxX, mnMaxY;
};
float Frame::mnMaxX, Frame::mnMaxY;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv, Frame::mnMinX, Frame::mnMinY;

class KeyFrame {
public:
    /* what SearchForTriangulation reads (include/KeyFrame.h): poses as cv::Mat, the cameras, per-keypoint stereo coordinate, level tables */
    cv::Mat mRcw, mtcw, mOw, mRrw, mtrw;
    cv::Mat GetCameraCenter() { return mOw.clone(); }
    cv::Mat GetRotation() { return mRcw.clone(); }
    cv::Mat GetTranslation() { return mtcw.clone(); }
    cv::Mat GetRightRotation() { return mRrw.clone(); }
    cv::Mat GetRightTranslation() { return mtrw.clone(); }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    GeometricCamera* mpCamera = nullptr;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2;
    int N = 0;
    /* what ORBmatcher::Fuse and KeyFrame::GetFeaturesInArea / IsInImage read on top (include/KeyFrame.h) */
    cv::Mat mOwr;
    cv::Mat GetRightCameraCenter() { return mOwr.clone(); }
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0;
    std::vector<float> mvInvLevelSigma2;
    int mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0, mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    std::vector<std::vector<std::vector<size_t>>> mGrid, mGridRight;
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight = false) const;
    bool IsInImage(const float& x, const float& y) const;
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; pMP->rowIndex = (int)idx; }
    std::set<MapPoint*> GetMapPoints() { std::set<MapPoint*> r; for (MapPoint* m : mvpMapPoints) if (m && !m->isBad()) r.insert(m); return r; }   /* src/KeyFrame.cc:392-405 */
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    std::vector<MapPoint*> mvpMapPoints;
    DBoW2::FeatureVector mFeatVec;
    cv::Mat mDescriptors;
    int NLeft = -1;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    GeometricCamera* mpCamera2 = nullptr;
    bool mbBadKF = false;
    bool isBad() { return mbBadKF; }
};

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3, const bool bFarPoints = false, const float thFarPoints = 50.0f);
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize = 10);
    int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t>>& vMatchedPairs, const bool bOnlyStereo, const bool bCoarse = false);
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
    int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th, const int ORBdist);
    int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0, const bool bRight = false);
    int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched, int th, float ratioHamming = 1.0);
    int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, const std::vector<KeyFrame*>& vpPointsKFs,
                           std::vector<MapPoint*>& vpMatched, std::vector<KeyFrame*>& vpMatchedKF, int th, float ratioHamming = 1.0);
    int Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, std::vector<MapPoint*>& vpReplacePoint);
    int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12, const cv::Mat& t12, const float th);
    float RadiusByViewingCos(const float& viewCos);
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;
    float mfNNratio;
    bool mbCheckOrientation;
};
const int ORBmatcher::HISTO_LENGTH = 30;   /* src/ORBmatcher.cc:38 */

#include "_ref/ORBmatcher_constants.inc"
#include "_ref/ORBmatcher_DescriptorDistance.inc"
#include "_ref/Frame_ComputeStereoMatches.inc"
#include "_ref/Frame_ComputeStereoFishEyeMatches.inc"
#include "_ref/ORBmatcher_ComputeThreeMaxima.inc"
#include "_ref/ORBmatcher_SearchByBoW.inc"
#include "_ref/Frame_grid.inc"
#include "_ref/MapPoint_ComputeDistinctiveDescriptors.inc"
#include "_ref/ORBmatcher_SearchByProjection.inc"
#include "_ref/ORBmatcher_SearchForInitialization.inc"
#include "_ref/ORBmatcher_SearchForTriangulation.inc"
#include "_ref/ORBmatcher_SearchByProjectionLastFrame.inc"
#include "_ref/KeyFrame_grid.inc"
#include "_ref/ORBmatcher_SearchByProjectionReloc.inc"
#include "_ref/ORBmatcher_Fuse.inc"
#include "_ref/ORBmatcher_SearchByProjectionSim3.inc"
#include "_ref/ORBmatcher_SearchByProjectionSim3KFs.inc"
#include "_ref/ORBmatcher_FuseSim3.inc"
#include "_ref/ORBmatcher_SearchBySim3.inc"

/* protected members of the extractor, for the stage-level comparisons */
struct Probe : ORBextractor {
    using ORBextractor::ORBextractor;
    using ORBextractor::ComputeKeyPointsOctTree;
    using ORBextractor::ComputePyramid;
    using ORBextractor::mnFeaturesPerLevel;
    using ORBextractor::umax;
};

}  // namespace ORB_SLAM3

using ORB_SLAM3::Probe;

namespace {
int copy_out(const std::vector<cv::KeyPoint>& kps, const cv::Mat& desc, orc_keypoint* okps, uint8_t* odesc, int cap, int* n) {
    *n = (int)kps.size();
    if (*n > cap) return -2;
    if (*n) std::memcpy(okps, kps.data(), sizeof(orc_keypoint) * kps.size());
    for (int i = 0; i < *n; ++i) std::memcpy(odesc + 32 * i, desc.ptr(i), 32);
    return 0;
}
}  // namespace

extern "C" {

void ref_set_arena(int on) { g_useArena.store(on); }
size_t ref_arena_peak(void) { return t_arena.peak; }

/* ORBextractor ctor tables (src/ORBextractor.cc:408-468): quotas[nlevels], umax[16], scale[nlevels]. */
void ref_tables(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int* quotas, int* umax, float* scale) {
    ArenaScope scope;
    Probe ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    for (int l = 0; l < nlevels; ++l) quotas[l] = ex.mnFeaturesPerLevel[l];
    for (int v = 0; v < 16; ++v) umax[v] = ex.umax[v];
    std::vector<float> s = ex.GetScaleFactors();
    for (int l = 0; l < nlevels; ++l) scale[l] = s[l];
}

/* ORBextractor::operator() (src/ORBextractor.cc:1068-1150) on one image.  Returns monoIndex (-1: empty image,
 * -2: cap too small).  pyramid (optional): the padded levels, dense (h+38) x (w+38), back to back. */
int ref_extract(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const uint8_t* img, int rows, int cols,
                int step, int lap0, int lap1, orc_keypoint* kps, uint8_t* desc, int cap, int* n, uint8_t* pyramid) {
    ArenaScope scope;
    Probe ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    cv::Mat image = img ? cv::Mat(rows, cols, CV_8UC1, (void*)img, (size_t)step) : cv::Mat();
    std::vector<cv::KeyPoint> keys;
    cv::Mat descriptors;
    std::vector<int> lap{lap0, lap1};
    int mono = ex(image, cv::Mat(), keys, descriptors, lap);
    if (mono < 0) { *n = 0; return mono; }
    if (copy_out(keys, descriptors, kps, desc, cap, n)) return -2;
    if (pyramid)
        for (int l = 0; l < nlevels; ++l) {
            cv::Mat p = ex.mvImagePyramid[l];
            p.adjustROI(19, 19, 19, 19);
            for (int y = 0; y < p.rows; ++y, pyramid += p.cols) std::memcpy(pyramid, p.ptr(y), p.cols);
        }
    return mono;
}

/* ComputePyramid + ComputeKeyPointsOctTree (src/ORBextractor.cc:763-878): the keypoints of every level after the
 * quadtree and the orientation step, level coordinates, in the reference's order; counts[nlevels]. */
int ref_level_keypoints(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const uint8_t* img, int rows,
                        int cols, int step, orc_keypoint* out, int cap, int* counts) {
    ArenaScope scope;
    Probe ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    ex.ComputePyramid(cv::Mat(rows, cols, CV_8UC1, (void*)img, (size_t)step));
    std::vector<std::vector<cv::KeyPoint>> all;
    ex.ComputeKeyPointsOctTree(all);
    int total = 0;
    for (int l = 0; l < nlevels; ++l) {
        counts[l] = (int)all[l].size();
        if (total + counts[l] > cap) return -2;
        if (counts[l]) std::memcpy(out + total, all[l].data(), sizeof(orc_keypoint) * all[l].size());
        total += counts[l];
    }
    return total;
}

/* Both extractions of a stereo frame followed by Frame::ComputeStereoMatches (src/Frame.cc:797-967).
 * `mb` is the member the reference divides mbf by to get maxD (:827).  Returns the number of left keypoints. */
int ref_stereo(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const uint8_t* imgL, const uint8_t* imgR,
               int rows, int cols, int step, float mbf, float mb, orc_keypoint* kpsL, uint8_t* descL, int* nL,
               orc_keypoint* kpsR, uint8_t* descR, int* nR, int cap, float* uRight, float* depth) {
    ArenaScope scope;
    Probe exL(nfeatures, scaleFactor, nlevels, iniTh, minTh), exR(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    ORB_SLAM3::Frame F;
    std::vector<int> lap{0, 0};
    exL(cv::Mat(rows, cols, CV_8UC1, (void*)imgL, (size_t)step), cv::Mat(), F.mvKeys, F.mDescriptors, lap);
    exR(cv::Mat(rows, cols, CV_8UC1, (void*)imgR, (size_t)step), cv::Mat(), F.mvKeysRight, F.mDescriptorsRight, lap);
    F.N = (int)F.mvKeys.size();
    F.mbf = mbf;
    F.mb = mb;
    F.mpORBextractorLeft = &exL;
    F.mpORBextractorRight = &exR;
    F.mvScaleFactors = exL.GetScaleFactors();
    F.mvInvScaleFactors = exL.GetInverseScaleFactors();
    if (copy_out(F.mvKeys, F.mDescriptors, kpsL, descL, cap, nL)) return -2;
    if (copy_out(F.mvKeysRight, F.mDescriptorsRight, kpsR, descR, cap, nR)) return -2;
    F.ComputeStereoMatches();
    for (int i = 0; i < F.N; ++i) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
    return F.N;
}

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2591-2607) on two 32-byte rows. */
int ref_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    alignas(4) uint8_t ra[32], rb[32];
    std::memcpy(ra, a, 32);
    std::memcpy(rb, b, 32);
    return ORB_SLAM3::ORBmatcher::DescriptorDistance(cv::Mat(1, 32, CV_8UC1, ra, 32), cv::Mat(1, 32, CV_8UC1, rb, 32));
}

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:269-470) and, with kf_kf, SearchByBoW(KeyFrame*,
 * KeyFrame*, vpMatches12) (:823-963) on plain arrays — the argument list of orc_search_by_bow.  has_map_point[i]: 0 = NULL, 1 = a
 * good map point, 2 = a bad one (isBad()).  match_out: kf_kf = 0: per frame row the keyframe row whose map point was assigned, or -1;
 * kf_kf = 1: per keyframe-1 row the keyframe-2 row, or -1.  Returns nmatches. */
int ref_search_by_bow(const uint8_t* desc_kf, const orc_keypoint* kps_kf, int n_kf, const int32_t* nodes_kf, const int32_t* starts_kf, const int32_t* feats_kf,
                      int nfv_kf, const uint8_t* has_map_point, const uint8_t* desc_f, const orc_keypoint* kps_f, int n_f, const int32_t* nodes_f,
                      const int32_t* starts_f, const int32_t* feats_f, int nfv_f, const uint8_t* has_map_point_f, int n_left_f, float nnratio,
                      int check_orientation, int kf_kf, int32_t* match_out) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    auto fill_fv = [](DBoW2::FeatureVector& fv, const int32_t* nodes, const int32_t* starts, const int32_t* feats, int n) {
        for (int e = 0; e < n; e++) {
            std::vector<unsigned int> v(feats + starts[e], feats + starts[e + 1]);
            fv.insert(fv.end(), std::make_pair((DBoW2::NodeId)nodes[e], v));
        }
    };
    auto keys_of = [](const orc_keypoint* k, int n) {
        std::vector<cv::KeyPoint> v(n);
        if (n) std::memcpy((void*)v.data(), k, sizeof(orc_keypoint) * n);
        return v;
    };
    std::vector<MapPoint> store(n_kf + n_f);
    auto points_of = [&](const uint8_t* has, int n, int base) {
        std::vector<MapPoint*> v(n, nullptr);
        for (int i = 0; i < n; i++)
            if (has && has[i]) { store[base + i].bad = has[i] == 2; v[i] = &store[base + i]; }
        return v;
    };
    KeyFrame kf;
    kf.mDescriptors = cv::Mat(n_kf, 32, CV_8UC1, (void*)desc_kf, 32);
    kf.mvKeysUn = kf.mvKeys = keys_of(kps_kf, n_kf);
    kf.mvpMapPoints = points_of(has_map_point, n_kf, 0);
    fill_fv(kf.mFeatVec, nodes_kf, starts_kf, feats_kf, nfv_kf);
    ORBmatcher matcher(nnratio, check_orientation != 0);
    std::vector<MapPoint*> out;
    int nm;
    if (!kf_kf) {
        Frame F;
        F.N = n_f;
        F.Nleft = n_left_f;
        static int rig_marker;
        if (n_left_f >= 0) F.mpCamera2 = reinterpret_cast<GeometricCamera*>(&rig_marker);   /* a fisheye rig: the frame has a second camera */
        F.mDescriptors = cv::Mat(n_f, 32, CV_8UC1, (void*)desc_f, 32);
        if (n_left_f < 0) F.mvKeysUn = F.mvKeys = keys_of(kps_f, n_f);
        else { F.mvKeys = keys_of(kps_f, n_left_f); F.mvKeysRight = keys_of(kps_f + n_left_f, n_f - n_left_f); }
        fill_fv(F.mFeatVec, nodes_f, starts_f, feats_f, nfv_f);
        nm = matcher.SearchByBoW(&kf, F, out);
        for (int i = 0; i < n_f; i++) match_out[i] = out[i] ? (int32_t)(out[i] - &store[0]) : -1;
    } else {
        KeyFrame kf2;
        kf2.mDescriptors = cv::Mat(n_f, 32, CV_8UC1, (void*)desc_f, 32);
        kf2.mvKeysUn = kf2.mvKeys = keys_of(kps_f, n_f);
        kf2.mvpMapPoints = points_of(has_map_point_f, n_f, n_kf);
        fill_fv(kf2.mFeatVec, nodes_f, starts_f, feats_f, nfv_f);
        nm = matcher.SearchByBoW(&kf, &kf2, out);
        for (int i = 0; i < n_kf; i++) match_out[i] = out[i] ? (int32_t)(out[i] - &store[n_kf]) : -1;
    }
    return nm;
}

/* ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th, bFarPoints = false) (src/ORBmatcher.cc:44-214) on a frame given as arrays, with the
 * grid built by the reference's own AssignFeaturesToGrid.  kps / desc: n rows (rows >= n_left: the right camera, n_left = -1: none);
 * l2r [n_left] / r2l [n - n_left]: mvLeftToRightMatch / mvRightToLeftMatch; taken[i] != 0: the row starts with a map point that has
 * observations.  Map points as arrays of nmp entries (the MapPoint fields of the same names; desc_mp nmp x 32).  grid4 = mnMinX, mnMinY,
 * mfGridElementWidthInv, mfGridElementHeightInv.  owner[i] = index of the map point row i holds afterwards, -2 = the initial one, -1 = none. */
int ref_search_by_projection(const orc_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right, const int32_t* l2r, const int32_t* r2l,
                             const uint8_t* taken, const float* grid4, const float* scale_factors, int nlevels, int nmp, const uint8_t* in_view,
                             const uint8_t* in_view_r, const int32_t* level, const int32_t* level_r, const float* view_cos, const float* view_cos_r,
                             const float* proj_x, const float* proj_y, const float* proj_xr, const float* proj_yr, const int32_t* obs, const uint8_t* desc_mp,
                             float th, float nnratio, int32_t* owner) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F;
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    F.N = n;
    F.Nleft = n_left;
    F.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    std::vector<cv::KeyPoint> all(n);
    if (n) std::memcpy((void*)all.data(), kps, sizeof(orc_keypoint) * n);
    if (n_left < 0) F.mvKeysUn = F.mvKeys = all;
    else { F.mvKeys.assign(all.begin(), all.begin() + n_left); F.mvKeysRight.assign(all.begin() + n_left, all.end()); }
    F.mvuRight.assign(n, -1.f);
    if (u_right) F.mvuRight.assign(u_right, u_right + n);
    if (n_left >= 0) { F.mvLeftToRightMatch.assign(l2r, l2r + n_left); F.mvRightToLeftMatch.assign(r2l, r2l + (n - n_left)); }
    F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    F.AssignFeaturesToGrid();
    MapPoint initial;
    initial.nObs = 1;
    F.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; i++)
        if (taken && taken[i]) F.mvpMapPoints[i] = &initial;
    std::vector<MapPoint> store(nmp);
    std::vector<MapPoint*> vp(nmp);
    for (int i = 0; i < nmp; i++) {
        MapPoint& m = store[i];
        m.mbTrackInView = in_view[i]; m.mbTrackInViewR = in_view_r[i];
        m.mnTrackScaleLevel = level[i]; m.mnTrackScaleLevelR = level_r[i];
        m.mTrackViewCos = view_cos[i]; m.mTrackViewCosR = view_cos_r[i];
        m.mTrackProjX = proj_x[i]; m.mTrackProjY = proj_y[i]; m.mTrackProjXR = proj_xr[i]; m.mTrackProjYR = proj_yr[i];
        m.nObs = obs[i];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc_mp + (size_t)i * 32), 32);
        vp[i] = &m;
    }
    ORBmatcher matcher(nnratio, true);
    const int nm = matcher.SearchByProjection(F, vp, th, false, 50.0f);
    for (int i = 0; i < n; i++) owner[i] = !F.mvpMapPoints[i] ? -1 : F.mvpMapPoints[i] == &initial ? -2 : (int32_t)(F.mvpMapPoints[i] - store.data());
    return nm;
}

/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:706-821) on two frames given as arrays (kps = mvKeysUn), F2's grid built by the
 * reference's own AssignFeaturesToGrid.  prev_matched [n1][2] in/out, matches12 [n1] out.  Returns nmatches. */
int ref_search_for_initialization(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const orc_keypoint* kps2, const uint8_t* desc2, int n2,
                                  const float* grid4, float* prev_matched, int window, float nnratio, int check_orientation, int32_t* matches12) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F1, F2;
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    F1.N = n1; F2.N = n2;
    F1.mDescriptors = cv::Mat(n1, 32, CV_8UC1, (void*)desc1, 32);
    F2.mDescriptors = cv::Mat(n2, 32, CV_8UC1, (void*)desc2, 32);
    F1.mvKeysUn.resize(n1); F2.mvKeysUn.resize(n2);
    if (n1) std::memcpy((void*)F1.mvKeysUn.data(), kps1, sizeof(orc_keypoint) * n1);
    if (n2) std::memcpy((void*)F2.mvKeysUn.data(), kps2, sizeof(orc_keypoint) * n2);
    F1.mvKeys = F1.mvKeysUn; F2.mvKeys = F2.mvKeysUn;
    F2.AssignFeaturesToGrid();
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; i++) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int nm = matcher.SearchForInitialization(F1, F2, prev, m12, window);
    for (int i = 0; i < n1; i++) { matches12[i] = m12[i]; prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y; }
    return nm;
}

namespace {
cv::Mat mat_from(const float* v, int r, int c) {
    cv::Mat m(r, c, CV_32F);
    for (int i = 0; i < r * c; i++) m.at<float>(i / c, i % c) = v[i];
    return m;
}
void mat_to(const cv::Mat& m, float* out) {
    for (int i = 0; i < m.rows; i++)
        for (int j = 0; j < m.cols; j++) out[i * m.cols + j] = m.at<float>(i, j);
}
}  // namespace

/* KannalaBrandt8::TriangulateMatches (src/CameraModels/KannalaBrandt8.cpp:338-407) for n keypoint pairs: depth[i] = its return value,
 * p3d[3i..] = p3D when it was set (0 otherwise).  cam1 / cam2: mvParameters (8 floats). */
void ref_kb8_triangulate_matches(const float* cam1, const float* cam2, const float* R12, const float* t12, const float* uv1, const float* uv2,
                                 const float* sigma1, const float* sigma2, int n, float* depth, float* p3d) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    KannalaBrandt8 c1(std::vector<float>(cam1, cam1 + 8)), c2(std::vector<float>(cam2, cam2 + 8));
    const cv::Mat R = mat_from(R12, 3, 3), t = mat_from(t12, 3, 1);
    for (int i = 0; i < n; i++) {
        cv::KeyPoint k1, k2;
        k1.pt = cv::Point2f(uv1[2 * i], uv1[2 * i + 1]);
        k2.pt = cv::Point2f(uv2[2 * i], uv2[2 * i + 1]);
        cv::Mat p;
        depth[i] = c1.TriangulateMatches(&c2, k1, k2, R, t, sigma1[i], sigma2[i], p);
        for (int k = 0; k < 3; k++) p3d[3 * i + k] = p.empty() ? 0.f : p.at<float>(k);
    }
}

/* Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159) on the two images' keypoints and descriptors (the extractor's output order: the
 * monocular part first, then the rows inside the lapping interval): left_to_right [nl], right_to_left [nr], depth [nl], p3d [nl][3] (0 where
 * mvStereo3Dpoints stays empty).  cam1 / cam2 = mvParameters of the two KannalaBrandt8 cameras, Rlr / tlr = mRlr / mtlr.  Returns the matches kept. */
int ref_fisheye_stereo(const orc_keypoint* kps_l, const uint8_t* desc_l, int nl, int mono_l, const orc_keypoint* kps_r, const uint8_t* desc_r, int nr, int mono_r,
                       const float* cam1, const float* cam2, const float* Rlr, const float* tlr, const float* level_sigma2, int nlevels, int32_t* left_to_right,
                       int32_t* right_to_left, float* depth, float* p3d) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F;
    KannalaBrandt8 c1(std::vector<float>(cam1, cam1 + 8)), c2(std::vector<float>(cam2, cam2 + 8));
    F.mpCamera = &c1; F.mpCamera2 = &c2;
    F.Nleft = nl; F.Nright = nr; F.monoLeft = mono_l; F.monoRight = mono_r;
    F.mvKeys.resize(nl); F.mvKeysRight.resize(nr);
    if (nl) std::memcpy((void*)F.mvKeys.data(), kps_l, sizeof(orc_keypoint) * nl);
    if (nr) std::memcpy((void*)F.mvKeysRight.data(), kps_r, sizeof(orc_keypoint) * nr);
    F.mDescriptors = cv::Mat(nl, 32, CV_8UC1, (void*)desc_l, 32);
    F.mDescriptorsRight = cv::Mat(nr, 32, CV_8UC1, (void*)desc_r, 32);
    F.mvLevelSigma2.assign(level_sigma2, level_sigma2 + nlevels);
    F.mRlr = mat_from(Rlr, 3, 3); F.mtlr = mat_from(tlr, 3, 1);
    F.ComputeStereoFishEyeMatches();
    int kept = 0;
    for (int i = 0; i < nl; i++) {
        left_to_right[i] = F.mvLeftToRightMatch[i];
        depth[i] = F.mvDepth[i];
        const cv::Mat& p = F.mvStereo3Dpoints[i];
        for (int k = 0; k < 3; k++) p3d[3 * i + k] = p.empty() ? 0.f : p.at<float>(k);
        kept += F.mvLeftToRightMatch[i] >= 0;
    }
    for (int i = 0; i < nr; i++) right_to_left[i] = F.mvRightToLeftMatch[i];
    return kept;
}

/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:2193-2419) on two frames given as arrays.
 *   current frame: kps / desc [n] (rows >= n_left: the right camera; n_left = -1: one camera), u_right [n] or NULL, obs_cur [n] = -1 no map
 *   point, else Observations() of the one the row holds; grid4 + bounds4 (mnMinX, mnMinY, inv widths; mnMinX, mnMaxX, mnMinY, mnMaxY); Tcw [16],
 *   Trl [12] (3x4, rig only); cam: Pinhole fx, fy, cx, cy;
 *   last frame: nl rows with octave / angle in kps_last (n_left_last like above), has_mp / outlier [nl], world [nl][3], obs_mp [nl], desc_mp
 *   [nl][32], Tlw [16].
 * owner[row] = index of the last-frame row whose map point the row holds afterwards, -2 = the initial one, -1 = none. */
int ref_search_by_projection_last_frame(const orc_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right, const int32_t* obs_cur,
                                        const float* grid4, const float* bounds4, const float* scale_factors, int nlevels, float mbf, float mb, const float* Tcw,
                                        const float* Trl, const float* cam4, const orc_keypoint* kps_last, int nl, int n_left_last, const uint8_t* has_mp,
                                        const uint8_t* outlier, const float* world, const int32_t* obs_mp, const uint8_t* desc_mp, const float* Tlw, float th,
                                        int mono, int check_orientation, int32_t* owner) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F, Last;
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    Frame::mnMaxX = bounds4[1]; Frame::mnMaxY = bounds4[3];
    Pinhole camera(std::vector<float>(cam4, cam4 + 4));
    F.N = n; F.Nleft = n_left; F.mbf = mbf; F.mb = mb; F.mpCamera = &camera;
    F.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    std::vector<cv::KeyPoint> all(n);
    if (n) std::memcpy((void*)all.data(), kps, sizeof(orc_keypoint) * n);
    if (n_left < 0) F.mvKeysUn = F.mvKeys = all;
    else { F.mvKeys.assign(all.begin(), all.begin() + n_left); F.mvKeysRight.assign(all.begin() + n_left, all.end()); }
    F.mvuRight.assign(n, -1.f);
    if (u_right) F.mvuRight.assign(u_right, u_right + n);
    F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    F.mTcw = mat_from(Tcw, 4, 4);
    if (n_left >= 0) F.mTrl = mat_from(Trl, 3, 4);
    F.AssignFeaturesToGrid();
    std::vector<MapPoint> initial(n);
    F.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; i++)
        if (obs_cur[i] >= 0) { initial[i].nObs = obs_cur[i]; F.mvpMapPoints[i] = &initial[i]; }
    Last.N = nl; Last.Nleft = n_left_last;
    Last.mTcw = mat_from(Tlw, 4, 4);
    std::vector<cv::KeyPoint> lall(nl);
    if (nl) std::memcpy((void*)lall.data(), kps_last, sizeof(orc_keypoint) * nl);
    if (n_left_last < 0) Last.mvKeysUn = Last.mvKeys = lall;
    else { Last.mvKeys.assign(lall.begin(), lall.begin() + n_left_last); Last.mvKeysRight.assign(lall.begin() + n_left_last, lall.end()); }
    std::vector<MapPoint> store(nl);
    Last.mvpMapPoints.assign(nl, nullptr);
    Last.mvbOutlier.assign(nl, false);
    for (int i = 0; i < nl; i++) {
        store[i].nObs = obs_mp[i];
        store[i].mWorldPos = mat_from(world + 3 * i, 3, 1);
        store[i].mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc_mp + (size_t)i * 32), 32);
        if (has_mp[i]) Last.mvpMapPoints[i] = &store[i];
        Last.mvbOutlier[i] = outlier[i] != 0;
    }
    ORBmatcher matcher(0.9f, check_orientation != 0);
    const int nm = matcher.SearchByProjection(F, Last, th, mono != 0);
    for (int i = 0; i < n; i++) {
        MapPoint* p = F.mvpMapPoints[i];
        owner[i] = !p ? -1 : (p >= initial.data() && p < initial.data() + n) ? -2 : (int32_t)(p - store.data());
    }
    return nm;
}

/* ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2421-2543, Tracking::Relocalization) on arrays.
 *   current frame as in ref_search_by_projection_last_frame (pinhole, no second camera): kps / desc [n], obs_cur [n] (-1: the row holds no map point),
 *   grid4, bounds4, scale_factors, Tcw [16], cam4;
 *   keyframe: nk rows with the angle in kps_kf; per row has_mp, bad, found (the point is in sAlreadyFound), world [3], min_dist / max_dist (the
 *   distance-invariance interval), level (what PredictScale returns), desc_mp [32].
 * owner[row] = index of the keyframe row whose map point the row holds afterwards, -2 = the initial one, -1 = none.  Returns nmatches. */
int ref_search_by_projection_reloc(const orc_keypoint* kps, const uint8_t* desc, int n, const int32_t* obs_cur, const float* grid4, const float* bounds4,
                                   const float* scale_factors, int nlevels, const float* Tcw, const float* cam4, const orc_keypoint* kps_kf, int nk,
                                   const uint8_t* has_mp, const uint8_t* bad, const uint8_t* found, const float* world, const float* min_dist, const float* max_dist,
                                   const int32_t* level, const uint8_t* desc_mp, float th, int orb_dist, int check_orientation, int32_t* owner) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F;
    KeyFrame kf;
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    Frame::mnMaxX = bounds4[1]; Frame::mnMaxY = bounds4[3];
    Pinhole camera(std::vector<float>(cam4, cam4 + 4));
    F.N = n; F.Nleft = -1; F.mpCamera = &camera;
    F.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    F.mvKeysUn.resize(n);
    if (n) std::memcpy((void*)F.mvKeysUn.data(), kps, sizeof(orc_keypoint) * n);
    F.mvKeys = F.mvKeysUn;
    F.mvuRight.assign(n, -1.f);
    F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    F.mTcw = mat_from(Tcw, 4, 4);
    F.AssignFeaturesToGrid();
    std::vector<MapPoint> initial(n);
    F.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; i++)
        if (obs_cur[i] >= 0) { initial[i].nObs = obs_cur[i]; F.mvpMapPoints[i] = &initial[i]; }
    kf.N = nk;
    kf.mvKeysUn.resize(nk);
    if (nk) std::memcpy((void*)kf.mvKeysUn.data(), kps_kf, sizeof(orc_keypoint) * nk);
    std::vector<MapPoint> store(nk);
    kf.mvpMapPoints.assign(nk, nullptr);
    std::set<MapPoint*> already;
    for (int i = 0; i < nk; i++) {
        store[i].bad = bad[i] != 0;
        store[i].mWorldPos = mat_from(world + 3 * i, 3, 1);
        store[i].mfMinDistance = min_dist[i]; store[i].mfMaxDistance = max_dist[i];
        store[i].predictedLevel = level[i];
        store[i].mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc_mp + (size_t)i * 32), 32);
        if (has_mp[i]) kf.mvpMapPoints[i] = &store[i];
        if (found[i]) already.insert(&store[i]);
    }
    ORBmatcher matcher(0.9f, check_orientation != 0);
    const int nm = matcher.SearchByProjection(F, &kf, already, th, orb_dist);
    for (int i = 0; i < n; i++) {
        MapPoint* p = F.mvpMapPoints[i];
        owner[i] = !p ? -1 : (p >= initial.data() && p < initial.data() + n) ? -2 : (int32_t)(p - store.data());
    }
    return nm;
}

/* ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) (src/ORBmatcher.cc:1643-1841) on a keyframe and map points given as arrays.
 *   keyframe: kps / desc [n] (rows >= n_left: the right camera), u_right [n], row_obs [n] = -1 no map point else its Observations(),
 *   row_bad [n]; grid4, bounds4 (mnMinX, mnMaxX, mnMinY, mnMaxY), scale_factors / inv_level_sigma2 [nlevels], bf, cam4 (Pinhole), pose = Rcw[9]
 *   tcw[3] of the camera that is searched;
 *   map points [nmp]: present (0 = NULL pointer), bad, in_kf, world [3], normal [3], min_dist, max_dist, level (what PredictScale returns),
 *   obs, desc_mp [32].
 * fused_row[i] = the keypoint row map point i was fused with (-1: none), fused_how[i] = 1 AddObservation + AddMapPoint, 2 pMP->Replace(pMPinKF),
 * 3 pMPinKF->Replace(pMP), 4 the row's point was bad (counted, nothing done).  Returns nFused. */
int ref_fuse(const orc_keypoint* kps, const uint8_t* desc, int n, int n_left, const float* u_right, const int32_t* row_obs, const uint8_t* row_bad,
             const float* grid4, const float* bounds4, const float* scale_factors, const float* inv_level_sigma2, int nlevels, float bf, const float* cam4,
             const float* pose, int nmp, const uint8_t* present, const uint8_t* bad, const uint8_t* in_kf, const float* world, const float* normal,
             const float* min_dist, const float* max_dist, const int32_t* level, const int32_t* obs, const uint8_t* desc_mp, float th, int right,
             int32_t* fused_row, int32_t* fused_how) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F;   /* the grid is built by Frame::AssignFeaturesToGrid and copied, as KeyFrame's constructor does (src/KeyFrame.cc:74-90) */
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    F.N = n; F.Nleft = n_left;
    std::vector<cv::KeyPoint> all(n);
    if (n) std::memcpy((void*)all.data(), kps, sizeof(orc_keypoint) * n);
    if (n_left < 0) F.mvKeysUn = F.mvKeys = all;
    else { F.mvKeys.assign(all.begin(), all.begin() + n_left); F.mvKeysRight.assign(all.begin() + n_left, all.end()); }
    F.AssignFeaturesToGrid();
    Pinhole camera(std::vector<float>(cam4, cam4 + 4));
    KeyFrame kf;
    kf.N = n; kf.NLeft = n_left;
    kf.mvKeys = F.mvKeys; kf.mvKeysUn = F.mvKeysUn; kf.mvKeysRight = F.mvKeysRight;
    kf.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    kf.mvuRight.assign(n, -1.f);
    if (u_right) kf.mvuRight.assign(u_right, u_right + n);
    kf.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    kf.mvInvLevelSigma2.assign(inv_level_sigma2, inv_level_sigma2 + nlevels);
    kf.mnMinX = bounds4[0]; kf.mnMaxX = bounds4[1]; kf.mnMinY = bounds4[2]; kf.mnMaxY = bounds4[3];
    kf.mfGridElementWidthInv = grid4[2]; kf.mfGridElementHeightInv = grid4[3];
    kf.fx = cam4[0]; kf.fy = cam4[1]; kf.cx = cam4[2]; kf.cy = cam4[3]; kf.mbf = bf;
    kf.mGrid.assign(FRAME_GRID_COLS, std::vector<std::vector<size_t>>(FRAME_GRID_ROWS));
    kf.mGridRight = kf.mGrid;
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) { kf.mGrid[i][j] = F.mGrid[i][j]; kf.mGridRight[i][j] = F.mGridRight[i][j]; }
    const cv::Mat Rcw = mat_from(pose, 3, 3), tcw = mat_from(pose + 9, 3, 1);
    cv::Mat Ow = -Rcw.t() * tcw;
    kf.mRcw = kf.mRrw = Rcw; kf.mtcw = kf.mtrw = tcw; kf.mOw = kf.mOwr = Ow;
    kf.mpCamera = &camera; kf.mpCamera2 = &camera;
    std::vector<MapPoint> rowPoint(n);
    kf.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; i++)
        if (row_obs[i] >= 0) { rowPoint[i].nObs = row_obs[i]; rowPoint[i].bad = row_bad && row_bad[i]; rowPoint[i].rowIndex = i; kf.mvpMapPoints[i] = &rowPoint[i]; }
    std::vector<MapPoint> store(nmp);
    std::vector<MapPoint*> vp(nmp, nullptr);
    for (int i = 0; i < nmp; i++) {
        MapPoint& m = store[i];
        m.queryIndex = i; m.bad = bad[i] != 0; m.inKF = in_kf[i] != 0; m.nObs = obs[i];
        m.mWorldPos = mat_from(world + 3 * i, 3, 1); m.mNormalVector = mat_from(normal + 3 * i, 3, 1);
        m.mfMinDistance = min_dist[i]; m.mfMaxDistance = max_dist[i]; m.predictedLevel = level[i];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc_mp + (size_t)i * 32), 32);
        if (present[i]) vp[i] = &m;
    }
    ORBmatcher matcher(0.6f, true);
    const int nFused = matcher.Fuse(&kf, vp, th, right != 0);
    for (int i = 0; i < nmp; i++) { fused_row[i] = store[i].fusedRow; fused_how[i] = store[i].fusedHow; }
    return nFused;
}

/* ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) (src/ORBmatcher.cc:473-586) on arrays.  matched_in [n]: -1 the
 * row has no match yet, -2 some other map point, k >= 0 candidate k itself (so that it is "already found").  Scw [16].  owner as elsewhere. */
int ref_search_by_projection_sim3(const orc_keypoint* kps, const uint8_t* desc, int n, const int32_t* matched_in, const float* grid4, const float* bounds4,
                                  const float* scale_factors, int nlevels, const float* cam4, const float* Scw, int nmp, const uint8_t* bad, const float* world,
                                  const float* normal, const float* min_dist, const float* max_dist, const int32_t* level, const uint8_t* desc_mp, int th,
                                  float ratio_hamming, int32_t* owner, const int32_t* point_kf, int32_t* owner_kf) {
    /* point_kf != NULL: the second overload (src/ORBmatcher.cc:588-704), vpPointsKFs[i] = keyframe number point_kf[i]; owner_kf[row] = the number
     * vpMatchedKF[row] ends up with (-1: untouched) */
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F;
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    F.N = n; F.Nleft = -1;
    F.mvKeysUn.resize(n);
    if (n) std::memcpy((void*)F.mvKeysUn.data(), kps, sizeof(orc_keypoint) * n);
    F.mvKeys = F.mvKeysUn;
    F.AssignFeaturesToGrid();
    Pinhole camera(std::vector<float>(cam4, cam4 + 4));
    KeyFrame kf;
    kf.N = n; kf.NLeft = -1;
    kf.mvKeys = kf.mvKeysUn = F.mvKeysUn;
    kf.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    kf.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    kf.mnMinX = bounds4[0]; kf.mnMaxX = bounds4[1]; kf.mnMinY = bounds4[2]; kf.mnMaxY = bounds4[3];
    kf.mfGridElementWidthInv = grid4[2]; kf.mfGridElementHeightInv = grid4[3];
    kf.fx = cam4[0]; kf.fy = cam4[1]; kf.cx = cam4[2]; kf.cy = cam4[3];
    kf.mGrid.assign(FRAME_GRID_COLS, std::vector<std::vector<size_t>>(FRAME_GRID_ROWS));
    kf.mGridRight = kf.mGrid;
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) kf.mGrid[i][j] = F.mGrid[i][j];
    kf.mpCamera = &camera;
    std::vector<MapPoint> store(nmp);
    std::vector<MapPoint*> vp(nmp);
    for (int i = 0; i < nmp; i++) {
        MapPoint& m = store[i];
        m.bad = bad[i] != 0;
        m.mWorldPos = mat_from(world + 3 * i, 3, 1); m.mNormalVector = mat_from(normal + 3 * i, 3, 1);
        m.mfMinDistance = min_dist[i]; m.mfMaxDistance = max_dist[i]; m.predictedLevel = level[i];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc_mp + (size_t)i * 32), 32);
        vp[i] = &m;
    }
    MapPoint other;
    std::vector<MapPoint*> matched(n, nullptr);
    for (int i = 0; i < n; i++) matched[i] = matched_in[i] == -1 ? nullptr : matched_in[i] == -2 ? &other : &store[matched_in[i]];
    ORBmatcher matcher(0.75f, true);
    int nm;
    if (point_kf) {
        int nkf = 1;
        for (int i = 0; i < nmp; i++) nkf = std::max(nkf, point_kf[i] + 1);
        std::vector<KeyFrame> kfs(nkf);
        std::vector<KeyFrame*> pointKFs(nmp), matchedKF(n, nullptr);
        for (int i = 0; i < nmp; i++) pointKFs[i] = &kfs[point_kf[i]];
        nm = matcher.SearchByProjection(&kf, mat_from(Scw, 4, 4), vp, pointKFs, matched, matchedKF, th, ratio_hamming);
        for (int i = 0; i < n; i++) owner_kf[i] = matchedKF[i] ? (int32_t)(matchedKF[i] - kfs.data()) : -1;
    } else
        nm = matcher.SearchByProjection(&kf, mat_from(Scw, 4, 4), vp, matched, th, ratio_hamming);
    for (int i = 0; i < n; i++) owner[i] = !matched[i] ? -1 : matched[i] == &other ? -2 : (int32_t)(matched[i] - store.data());
    return nm;
}

/* ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:1843-1965) on arrays.  row_state [n]: -1 the row has no map point,
 * -2 another map point, -3 another, bad one, k >= 0 candidate k itself (so that it is in pKF->GetMapPoints()).
 * fused_row[i] = the row candidate i was fused with (added, or to be replaced by the row's point), -1 none; replaced[i] = 1 iff vpReplacePoint[i]
 * was set.  Returns nFused. */
int ref_fuse_sim3(const orc_keypoint* kps, const uint8_t* desc, int n, const int32_t* row_state, const float* grid4, const float* bounds4,
                  const float* scale_factors, int nlevels, const float* cam4, const float* Scw, int nmp, const uint8_t* bad, const float* world,
                  const float* normal, const float* min_dist, const float* max_dist, const int32_t* level, const uint8_t* desc_mp, float th,
                  int32_t* fused_row, int32_t* replaced) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Frame F;
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    F.N = n; F.Nleft = -1;
    F.mvKeysUn.resize(n);
    if (n) std::memcpy((void*)F.mvKeysUn.data(), kps, sizeof(orc_keypoint) * n);
    F.mvKeys = F.mvKeysUn;
    F.AssignFeaturesToGrid();
    Pinhole camera(std::vector<float>(cam4, cam4 + 4));
    KeyFrame kf;
    kf.N = n; kf.NLeft = -1;
    kf.mvKeys = kf.mvKeysUn = F.mvKeysUn;
    kf.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    kf.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    kf.mnMinX = bounds4[0]; kf.mnMaxX = bounds4[1]; kf.mnMinY = bounds4[2]; kf.mnMaxY = bounds4[3];
    kf.mfGridElementWidthInv = grid4[2]; kf.mfGridElementHeightInv = grid4[3];
    kf.fx = cam4[0]; kf.fy = cam4[1]; kf.cx = cam4[2]; kf.cy = cam4[3];
    kf.mGrid.assign(FRAME_GRID_COLS, std::vector<std::vector<size_t>>(FRAME_GRID_ROWS));
    kf.mGridRight = kf.mGrid;
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) kf.mGrid[i][j] = F.mGrid[i][j];
    kf.mpCamera = &camera;
    std::vector<MapPoint> store(nmp), rowPoint(n);
    std::vector<MapPoint*> vp(nmp), repl(nmp, nullptr);
    for (int i = 0; i < nmp; i++) {
        MapPoint& m = store[i];
        m.queryIndex = i; m.bad = bad[i] != 0;
        m.mWorldPos = mat_from(world + 3 * i, 3, 1); m.mNormalVector = mat_from(normal + 3 * i, 3, 1);
        m.mfMinDistance = min_dist[i]; m.mfMaxDistance = max_dist[i]; m.predictedLevel = level[i];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc_mp + (size_t)i * 32), 32);
        vp[i] = &m;
    }
    kf.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; i++) {
        if (row_state[i] == -1) continue;
        if (row_state[i] >= 0) { kf.mvpMapPoints[i] = &store[row_state[i]]; store[row_state[i]].rowIndex = i; continue; }
        rowPoint[i].rowIndex = i; rowPoint[i].bad = row_state[i] == -3;
        kf.mvpMapPoints[i] = &rowPoint[i];
    }
    ORBmatcher matcher(0.8f, true);
    const int nFused = matcher.Fuse(&kf, mat_from(Scw, 4, 4), vp, th, repl);
    for (int i = 0; i < nmp; i++) {
        replaced[i] = repl[i] != nullptr;
        fused_row[i] = repl[i] ? repl[i]->rowIndex : store[i].fusedRow;
    }
    return nFused;
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1967-2191) on two keyframes given as arrays (identity
 * poses: a map point's world position is its position in both cameras' frames up to the similarity).  Per keyframe k: kps / desc [n], has_mp, bad,
 * world [n][3], min/max distance, level (PredictScale's answer in the OTHER keyframe); pre12 [n1]: -1 or the row of keyframe 2 whose map point
 * vpMatches12[i] already holds.  match12[i] = row of keyframe 2 whose map point vpMatches12[i] holds afterwards (-1 none).  Returns nFound. */
int ref_search_by_sim3(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has1, const uint8_t* bad1, const float* world1,
                       const float* min1, const float* max1, const int32_t* level1, const orc_keypoint* kps2, const uint8_t* desc2, int n2,
                       const uint8_t* has2, const uint8_t* bad2, const float* world2, const float* min2, const float* max2, const int32_t* level2,
                       const int32_t* pre12, const float* grid4, const float* bounds4, const float* scale_factors, int nlevels, const float* cam4,
                       float s12, const float* R12, const float* t12, float th, int32_t* match12) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    Pinhole camera(std::vector<float>(cam4, cam4 + 4));
    Frame::mnMinX = grid4[0]; Frame::mnMinY = grid4[1]; Frame::mfGridElementWidthInv = grid4[2]; Frame::mfGridElementHeightInv = grid4[3];
    KeyFrame kf[2];
    std::vector<MapPoint> pts[2];
    const orc_keypoint* kps[2] = {kps1, kps2};
    const uint8_t* desc[2] = {desc1, desc2};
    const int nn[2] = {n1, n2};
    const uint8_t* has[2] = {has1, has2};
    const uint8_t* bad[2] = {bad1, bad2};
    const float* world[2] = {world1, world2};
    const float* mn[2] = {min1, min2};
    const float* mx[2] = {max1, max2};
    const int32_t* lvl[2] = {level1, level2};
    const float eye[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};
    for (int k = 0; k < 2; k++) {
        const int n = nn[k];
        Frame F;
        F.N = n; F.Nleft = -1;
        F.mvKeysUn.resize(n);
        if (n) std::memcpy((void*)F.mvKeysUn.data(), kps[k], sizeof(orc_keypoint) * n);
        F.mvKeys = F.mvKeysUn;
        F.AssignFeaturesToGrid();
        KeyFrame& K = kf[k];
        K.N = n; K.NLeft = -1;
        K.mvKeys = K.mvKeysUn = F.mvKeysUn;
        K.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc[k], 32);
        K.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
        K.mnMinX = bounds4[0]; K.mnMaxX = bounds4[1]; K.mnMinY = bounds4[2]; K.mnMaxY = bounds4[3];
        K.mfGridElementWidthInv = grid4[2]; K.mfGridElementHeightInv = grid4[3];
        K.fx = cam4[0]; K.fy = cam4[1]; K.cx = cam4[2]; K.cy = cam4[3];
        K.mGrid.assign(FRAME_GRID_COLS, std::vector<std::vector<size_t>>(FRAME_GRID_ROWS));
        K.mGridRight = K.mGrid;
        for (int i = 0; i < FRAME_GRID_COLS; i++)
            for (int j = 0; j < FRAME_GRID_ROWS; j++) K.mGrid[i][j] = F.mGrid[i][j];
        K.mpCamera = &camera;
        K.mRcw = mat_from(eye, 3, 3); K.mtcw = mat_from(eye + 9, 3, 1);
        { std::vector<MapPoint> fresh(n); pts[k].swap(fresh); }   /* MapPoint holds a mutex: no resize() */
        K.mvpMapPoints.assign(n, nullptr);
        for (int i = 0; i < n; i++) {
            MapPoint& m = pts[k][i];
            m.bad = bad[k][i] != 0; m.rowIndex = i;
            m.mWorldPos = mat_from(world[k] + 3 * i, 3, 1);
            m.mfMinDistance = mn[k][i]; m.mfMaxDistance = mx[k][i]; m.predictedLevel = lvl[k][i];
            m.mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(desc[k] + (size_t)i * 32), 32);   /* a map point's descriptor: its keypoint's */
            if (has[k][i]) K.mvpMapPoints[i] = &m;
        }
    }
    std::vector<MapPoint*> matches(n1, nullptr);
    for (int i = 0; i < n1; i++)
        if (pre12[i] >= 0) { matches[i] = &pts[1][pre12[i]]; pts[1][pre12[i]].indexInOther = pre12[i]; }
    ORBmatcher matcher(0.75f, true);
    const int nFound = matcher.SearchBySim3(&kf[0], &kf[1], matches, s12, mat_from(R12, 3, 3), mat_from(t12, 3, 1), th);
    for (int i = 0; i < n1; i++) match12[i] = matches[i] ? matches[i]->rowIndex : -1;
    return nFound;
}

/* KannalaBrandt8::project / unproject (src/CameraModels/KannalaBrandt8.cpp:25-41, 96-125) on n points */
void ref_kb8_project(const float* cam, const float* p3, int n, float* uv) {
    ORB_SLAM3::KannalaBrandt8 c(std::vector<float>(cam, cam + 8));
    for (int i = 0; i < n; i++) { const cv::Point2f r = c.project(cv::Point3f(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2])); uv[2 * i] = r.x; uv[2 * i + 1] = r.y; }
}
void ref_kb8_unproject(const float* cam, const float* uv, int n, float* ray) {
    ORB_SLAM3::KannalaBrandt8 c(std::vector<float>(cam, cam + 8));
    for (int i = 0; i < n; i++) { const cv::Point3f r = c.unproject(cv::Point2f(uv[2 * i], uv[2 * i + 1])); ray[3 * i] = r.x; ray[3 * i + 1] = r.y; ray[3 * i + 2] = r.z; }
}

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:965-1206) on two keyframes given as arrays.
 *   kps / desc / has_mp [n]: keypoints (mvKeysUn, or mvKeys followed by mvKeysRight when n_left >= 0), descriptors, "has a map point";
 *   u_right [n] or NULL (all -1); feature vector as (nodes [nfv], starts [nfv+1], feats);
 *   cams [4][8]: keyframe 1 left, right, keyframe 2 left, right (a pinhole keyframe uses the first four numbers of its left entry);
 *   rig != 0: both keyframes have a second camera (KannalaBrandt8), else both are Pinhole;
 *   poses: Rcw[9], tcw[3] of keyframe 1, of keyframe 2, then (rig) Rrw[9], trw[3] = GetRightRotation / GetRightTranslation of 1 and of 2.
 * geom_out: pinhole — F12[9] as Pinhole::epipolarConstrain forms it and the epipole ep[2] (:972-977); rig — R[4][9] then t[4][3] for the
 * side pairs ll, lr, rl, rr (:995-1003): what the library's entry point takes as inputs, evaluated by the same expressions. */
int ref_search_for_triangulation(const uint8_t* desc1, const orc_keypoint* kps1, const float* u_right1, const uint8_t* has_mp1, int n1, int n_left1,
                                 const int32_t* nodes1, const int32_t* starts1, const int32_t* feats1, int nfv1, const uint8_t* desc2,
                                 const orc_keypoint* kps2, const float* u_right2, const uint8_t* has_mp2, int n2, int n_left2, const int32_t* nodes2,
                                 const int32_t* starts2, const int32_t* feats2, int nfv2, const float* cams, int rig, const float* poses,
                                 const float* scale_factors, const float* level_sigma2, int nlevels, int only_stereo, int coarse, int check_orientation,
                                 int32_t* match12, float* geom_out) {
    using namespace ORB_SLAM3;
    ArenaScope scope;
    std::vector<GeometricCamera*> cam(4, nullptr);
    for (int c = 0; c < 4; c++) {
        if (rig) cam[c] = new KannalaBrandt8(std::vector<float>(cams + 8 * c, cams + 8 * c + 8));
        else if (c % 2 == 0) cam[c] = new Pinhole(std::vector<float>(cams + 8 * c, cams + 8 * c + 4));
    }
    MapPoint some;
    auto fill = [&](KeyFrame& kf, const uint8_t* desc, const orc_keypoint* kps, const float* ur, const uint8_t* has, int n, int nl, const int32_t* nodes,
                    const int32_t* starts, const int32_t* feats, int nfv, const float* pose, const float* poseR, GeometricCamera* cl, GeometricCamera* cr) {
        kf.N = n; kf.NLeft = rig ? nl : -1;
        kf.mDescriptors = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
        std::vector<cv::KeyPoint> all(n);
        if (n) std::memcpy((void*)all.data(), kps, sizeof(orc_keypoint) * n);
        if (!rig) kf.mvKeysUn = kf.mvKeys = all;
        else { kf.mvKeys.assign(all.begin(), all.begin() + nl); kf.mvKeysRight.assign(all.begin() + nl, all.end()); }
        kf.mvuRight.assign(n, -1.f);
        if (ur) kf.mvuRight.assign(ur, ur + n);
        kf.mvpMapPoints.assign(n, nullptr);
        for (int i = 0; i < n; i++) if (has && has[i]) kf.mvpMapPoints[i] = &some;
        for (int k = 0; k < nfv; k++)
            for (int j = starts[k]; j < starts[k + 1]; j++) kf.mFeatVec.addFeature((DBoW2::NodeId)nodes[k], (unsigned)feats[j]);
        kf.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
        kf.mvLevelSigma2.assign(level_sigma2, level_sigma2 + nlevels);
        kf.mRcw = mat_from(pose, 3, 3); kf.mtcw = mat_from(pose + 9, 3, 1);
        kf.mOw = -kf.mRcw.t() * kf.mtcw;                                    /* KeyFrame::SetPose: Ow = -Rwc * tcw */
        if (rig) { kf.mRrw = mat_from(poseR, 3, 3); kf.mtrw = mat_from(poseR + 9, 3, 1); }
        kf.mpCamera = cl; kf.mpCamera2 = rig ? cr : nullptr;
    };
    KeyFrame kf1, kf2;
    fill(kf1, desc1, kps1, u_right1, has_mp1, n1, n_left1, nodes1, starts1, feats1, nfv1, poses, poses + 24, cam[0], cam[1]);
    fill(kf2, desc2, kps2, u_right2, has_mp2, n2, n_left2, nodes2, starts2, feats2, nfv2, poses + 12, poses + 36, cam[2], cam[3]);
    ORBmatcher matcher(0.6f, check_orientation != 0);
    std::vector<std::pair<size_t, size_t>> pairs;
    const int nm = matcher.SearchForTriangulation(&kf1, &kf2, cv::Mat(), pairs, only_stereo != 0, coarse != 0);
    for (int i = 0; i < n1; i++) match12[i] = -1;
    for (const auto& pr : pairs) match12[pr.first] = (int32_t)pr.second;
    if (geom_out) {
        if (!rig) {   /* the same expressions as :972-977, :990-993 and Pinhole.cpp:137-140 */
            cv::Mat Cw = kf1.GetCameraCenter(), R2w = kf2.GetRotation(), t2w = kf2.GetTranslation();
            cv::Mat C2 = R2w * Cw + t2w;
            const cv::Point2f ep = kf2.mpCamera->project(C2);
            cv::Mat R1w = kf1.GetRotation(), t1w = kf1.GetTranslation();
            cv::Mat R12 = R1w * R2w.t();
            cv::Mat t12 = -R1w * R2w.t() * t2w + t1w;
            Pinhole* p1 = static_cast<Pinhole*>(kf1.mpCamera);
            cv::Mat t12x = p1->SkewSymmetricMatrix(t12);
            cv::Mat K1 = p1->toK(), K2 = kf2.mpCamera->toK();
            cv::Mat F12 = K1.t().inv() * t12x * R12 * K2.inv();
            mat_to(F12, geom_out);
            geom_out[9] = ep.x; geom_out[10] = ep.y;
        } else {
            cv::Mat R[4], t[4];
            R[0] = kf1.GetRotation() * kf2.GetRotation().t();
            R[1] = kf1.GetRotation() * kf2.GetRightRotation().t();
            R[2] = kf1.GetRightRotation() * kf2.GetRotation().t();
            R[3] = kf1.GetRightRotation() * kf2.GetRightRotation().t();
            t[0] = kf1.GetRotation() * (-kf2.GetRotation().t() * kf2.GetTranslation()) + kf1.GetTranslation();
            t[1] = kf1.GetRotation() * (-kf2.GetRightRotation().t() * kf2.GetRightTranslation()) + kf1.GetTranslation();
            t[2] = kf1.GetRightRotation() * (-kf2.GetRotation().t() * kf2.GetTranslation()) + kf1.GetRightTranslation();
            t[3] = kf1.GetRightRotation() * (-kf2.GetRightRotation().t() * kf2.GetRightTranslation()) + kf1.GetRightTranslation();
            for (int c = 0; c < 4; c++) { mat_to(R[c], geom_out + 9 * c); mat_to(t[c], geom_out + 36 + 3 * c); }
        }
    }
    return nm;
}

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423) for one map point observed in n keyframes (one descriptor row each;
 * bad_kf[i] != 0: that keyframe isBad()).  The observation map is ordered by keyframe address; the keyframes sit in one array, so that
 * order is the row order.  Returns the row of `desc` whose copy becomes mDescriptor, or -1 when nothing is selected. */
int ref_distinctive_descriptor(const uint8_t* desc, int n, const uint8_t* bad_kf) {
    using namespace ORB_SLAM3;
    std::vector<KeyFrame> kfs(n);
    MapPoint mp;
    for (int i = 0; i < n; i++) {
        kfs[i].mDescriptors = cv::Mat(1, 32, CV_8UC1, (void*)(desc + (size_t)i * 32), 32);
        kfs[i].mbBadKF = bad_kf && bad_kf[i];
        mp.mObservations[&kfs[i]] = std::make_tuple(0, -1);
    }
    mp.ComputeDistinctiveDescriptors();
    if (mp.mDescriptor.empty()) return -1;
    for (int i = 0; i < n; i++)
        if (!(bad_kf && bad_kf[i]) && std::memcmp(mp.mDescriptor.data, desc + (size_t)i * 32, 32) == 0) return i;   /* first equal row */
    return -1;
}

int ref_th_low(void) { return ORB_SLAM3::ORBmatcher::TH_LOW; }
int ref_th_high(void) { return ORB_SLAM3::ORBmatcher::TH_HIGH; }

}  // extern "C"
