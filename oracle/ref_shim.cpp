/* ref_shim.cpp — C entry points around the REFERENCE'S OWN code, compiled where it lies.  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile (target _ref/liborb_ref.so) compiles /root/reference/src/ORBextractor.cc unmodified against
 * oracle/refstub/ (see cvstub.hpp), and cuts two function bodies out of files that cannot be compiled whole
 * (they pull in Eigen, Sophus, g2o, DBoW2, ...):
 *     /root/reference/src/Frame.cc       Frame::ComputeStereoMatches          -> _ref/Frame_ComputeStereoMatches.inc
 *     /root/reference/src/ORBmatcher.cc  ORBmatcher::DescriptorDistance, TH_* -> _ref/ORBmatcher_*.inc
 * Those generated files live only under oracle/_ref/ (git-ignored); no reference source is stored in this repository.
 * The `Frame` and `ORBmatcher` classes below declare just the members those bodies name (include/Frame.h,
 * include/ORBmatcher.h:42,95-96).
 *
 * Allocation: every `new` inside a ref_* call is served from a per-thread bump arena and `delete` is a no-op, so heap
 * addresses grow with creation order.  DistributeOctTree sorts pair<size, ExtractorNode*> (src/ORBextractor.cc:682):
 * equal sizes are ordered by node address, i.e. by whatever the allocator returned; under a monotonic allocator that
 * is creation order, which is the tie rule the oracle and the kernels define.  ref_set_arena(0) switches to plain
 * malloc/free to see what glibc's allocator does with the same inputs.
 */
#include <sys/mman.h>

#include <atomic>
#include <new>

#include "refstub/cvstub.hpp"
#include "ORBextractor.h" /* the reference's header: -I/root/reference/include */

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

/* ---- the two function bodies taken from Frame.cc / ORBmatcher.cc ------------------------------------------------- */
using namespace std;
namespace ORB_SLAM3 {

class ORBmatcher {
public:
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    static const int TH_LOW;
    static const int TH_HIGH;
};

class Frame {
public:
    void ComputeStereoMatches();
    int N;
    float mbf, mb;
    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<float> mvuRight, mvDepth;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
};

#include "_ref/ORBmatcher_constants.inc"
#include "_ref/ORBmatcher_DescriptorDistance.inc"
#include "_ref/Frame_ComputeStereoMatches.inc"

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

int ref_th_low(void) { return ORB_SLAM3::ORBmatcher::TH_LOW; }
int ref_th_high(void) { return ORB_SLAM3::ORBmatcher::TH_HIGH; }

}  // extern "C"
