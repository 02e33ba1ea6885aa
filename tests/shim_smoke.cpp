// Compiles the C++ shim (include/orbfe_shim) the way Frame.cc uses the reference classes and runs one stereo frame.
// Built and run by tests/test_shim.py (GPU) / syntax-checked on CPU.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <map>
#include <vector>
#include "orbfe_shim/ORBextractor.h"     // replaces the reference's include/ORBextractor.h (same public surface)
#include "orbfe_shim/orbfe_matching.h"   // next to the reference's own ORBmatcher.h / Frame.h, which stay as they are
extern "C" int orbfe_synth_stereo(uint64_t frame_index, int W, int H, uint8_t* left, uint8_t* right);
using namespace ORB_SLAM3;

// Stand-ins for the reference's own declarations (include/ORBmatcher.h:42,95-97; include/Frame.h:114 and the members the body
// reads), followed by the member bodies exactly as INTEGRATION.md section 3 and 5 give them.
namespace ORB_SLAM3 {
class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    static const int TH_LOW;
    static const int TH_HIGH;
    float mfNNratio;
    bool mbCheckOrientation;
};
const int ORBmatcher::TH_HIGH = 100;
const int ORBmatcher::TH_LOW = 50;
class Frame {
public:
    void ComputeStereoMatches();
    int N;
    float mbf, mb;
    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
    std::vector<float> mvuRight, mvDepth;
};
// src/ORBmatcher.cc:2591-2607
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbfe::DescriptorDistance(a, b); }
// src/Frame.cc:797-967
void Frame::ComputeStereoMatches() { orbfe::ComputeStereoMatches(mpORBextractorLeft, mpORBextractorRight, N, mbf, mbf / mb, mvuRight, mvDepth); }
}  // namespace ORB_SLAM3
#include <chrono>
int main(int argc, char** argv) {
    const int W = 752, H = 480;
    if (argc > 2 && std::string(argv[1]) == "--latency") {
        // single-frame latency of the drop-in as shipped: ORBextractor::operator() on a pageable host image, keypoints, descriptors and
        // (unless built with -DORBFE_SHIM_LAZY_PYRAMID) the eight mvImagePyramid mirrors back on the host
        const int iters = atoi(argv[2]);
        cv::Mat im(H, W, CV_8UC1), im2(H, W, CV_8UC1), mask;
        orbfe_synth_stereo(3, W, H, im.ptr<unsigned char>(0), im2.ptr<unsigned char>(0));
        ORBextractor ex(1200, 1.2f, 8, 20, 7);
        std::vector<cv::KeyPoint> k;
        cv::Mat d;
        std::vector<int> lap = {0, 0};
        for (int i = 0; i < 10; i++) ex(im, mask, k, d, lap);
        const auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < iters; i++) ex(i & 1 ? im2 : im, mask, k, d, lap);
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / iters;
#ifdef ORBFE_SHIM_LAZY_PYRAMID
        printf("shim_latency_us=%.1f pyramid=lazy n=%zu\n", us, k.size());
#else
        printf("shim_latency_us=%.1f pyramid=mirrored n=%zu\n", us, k.size());
#endif
        return 0;
    }
    cv::Mat imL(H, W, CV_8UC1), imR(H, W, CV_8UC1), mask;
    orbfe_synth_stereo(0, W, H, imL.ptr<unsigned char>(0), imR.ptr<unsigned char>(0));
    ORBextractor left(1200, 1.2f, 8, 20, 7), right(1200, 1.2f, 8, 20, 7);   // src/Tracking.cc:1151-1154
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR;
    std::vector<int> lap = {0, 0};
    const int monoL = left(imL, mask, kL, dL, lap);                           // src/Frame.cc:413-420
    const int monoR = right(imR, mask, kR, dR, lap);
    Frame F;
    F.N = (int)kL.size(); F.mbf = 47.90639f; F.mb = F.mbf / 435.2047f;           // src/Frame.cc:173
    F.mpORBextractorLeft = &left; F.mpORBextractorRight = &right;
    F.ComputeStereoMatches();
    const std::vector<float>& uR = F.mvuRight;
    int matched = 0;
    for (float v : uR) matched += v >= 0;
    const int d01 = ORBmatcher::DescriptorDistance(dL.row(0), dL.row(1));
    ORBmatcher m(0.9f, true);
    std::vector<orbfe_match> res;
    const int acc = orbfe::SearchTop2(dL, dR, res, m.mfNNratio, ORBmatcher::TH_LOW);
    // Tracking::SearchLocalPoints-like: every left keypoint re-projected half a pixel off as a map point (src/ORBmatcher.cc:44-214)
    std::vector<orbfe_proj_query> pq(kL.size());
    for (size_t i = 0; i < kL.size(); i++) {
        pq[i].x = kL[i].pt.x + 0.5f; pq[i].y = kL[i].pt.y - 0.5f;
        pq[i].radius = 4.0f * left.GetScaleFactors()[kL[i].octave];
        pq[i].xr = uR[i]; pq[i].min_level = kL[i].octave - 1; pq[i].max_level = kL[i].octave; pq[i].flags = ORBFE_PQ_BLOCKS; pq[i].angle = 0.f;
    }
    std::vector<unsigned char> taken;
    std::vector<orbfe_proj_match> pm;
    const int proj = orbfe::SearchByProjection(kL, dL, -1, &uR, nullptr, taken, 0.f, 0.f, 64.f / W, 48.f / H, pq, dL, 0.8f, pm);
    int projSelf = 0;
    for (size_t i = 0; i < pm.size(); i++) projSelf += pm[i].accepted && pm[i].best_idx == (int)i;
    // Tracking::MonocularInitialization-like (src/Tracking.cc:2462): the left keypoints against the right ones, previous matches = own positions
    std::vector<cv::Point2f> prevMatched(kL.size());
    for (size_t i = 0; i < kL.size(); i++) prevMatched[i] = kL[i].pt;
    std::vector<int> ini12;
    const int ini = orbfe::SearchForInitialization(kL, dL, kR, dR, 0.f, 0.f, 64.f / W, 48.f / H, prevMatched, ini12, 100, 0.9f, true, ORBmatcher::TH_LOW);
    int iniCount = 0;
    for (int v : ini12) iniCount += v >= 0;
    // SearchByBoW / SearchForTriangulation over std::map feature vectors (the shape of DBoW2::FeatureVector); the "vocabulary" here just groups
    // features by the top six bits of their first descriptor byte, which is all the matchers need of one
    std::map<unsigned, std::vector<unsigned>> fvL, fvR;
    for (size_t i = 0; i < kL.size(); i++) fvL[dL.ptr<unsigned char>((int)i)[0] >> 2].push_back((unsigned)i);
    for (size_t i = 0; i < kR.size(); i++) fvR[dR.ptr<unsigned char>((int)i)[0] >> 2].push_back((unsigned)i);
    std::vector<unsigned char> allMp(kL.size(), 1), noMp;
    std::vector<int> bowMatch, tri12;
    const int bow = orbfe::SearchByBoW(dL, kL, fvL, allMp, dR, kR, fvR, -1, 0.7f, true, bowMatch);
    long bowSum = 0;
    for (size_t i = 0; i < bowMatch.size(); i++) bowSum += (long)(i + 1) * (bowMatch[i] + 1);
    const float F12[9] = {0, 0, 0, 0, 0, -1, 0, 1, 0}, ep[2] = {-3000.f, 240.f};   // rectified pair: the epipolar line of a keypoint is its own row
    std::vector<float> sigma2(left.GetScaleFactors().size());
    for (size_t i = 0; i < sigma2.size(); i++) sigma2[i] = left.GetScaleFactors()[i] * left.GetScaleFactors()[i];
    const int tri = orbfe::SearchForTriangulation(dL, kL, nullptr, noMp, fvL, dR, kR, nullptr, noMp, fvR, F12, ep, left.GetScaleFactors(), sigma2, false, false, true, tri12);
    long triSum = 0;
    for (size_t i = 0; i < tri12.size(); i++) triSum += (long)(i + 1) * (tri12[i] + 1);
    // a fisheye rig: both images with the lapping interval, then the whole Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159)
    int fishMatches = 0;
    long fishSum = 0;
    {
        const int FW = 512, FH = 512;
        cv::Mat fL(FH, FW, CV_8UC1), fR(FH, FW, CV_8UC1);
        orbfe_synth_stereo(800, FW, FH, fL.ptr<unsigned char>(0), fR.ptr<unsigned char>(0));
        ORBextractor exl(1500, 1.2f, 8, 20, 7), exr(1500, 1.2f, 8, 20, 7);
        std::vector<cv::KeyPoint> fkL, fkR;
        cv::Mat fdL, fdR;
        std::vector<int> flap = {0, 511};
        exl(fL, mask, fkL, fdL, flap);
        exr(fR, mask, fkR, fdR, flap);
        const float cam[8] = {190.f, 190.f, 256.f, 256.f, 0.f, 0.f, 0.f, 0.f};
        const float Rlr[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, tlr[3] = {0.1f, 0.f, 0.f};
        std::vector<int> l2r, r2l;
        std::vector<float> fdepth, fp3d;
        orbfe::ComputeStereoFishEyeMatches(&exl, &exr, (int)fkL.size(), (int)fkR.size(), cam, cam, Rlr, tlr, l2r, r2l, fdepth, fp3d);
        for (size_t i = 0; i < l2r.size(); i++)
            if (l2r[i] >= 0) { fishMatches++; fishSum += (long)(i + 1) * (l2r[i] + 1); }
    }
    cv::Mat empty;
    const int rcEmpty = left(empty, mask, kR, dR, lap);
    printf("shim n=%zu mono=%d monoR=%d levels=%d pyr0=%dx%d stereo=%d d01=%d top2acc=%d proj=%d projself=%d init=%d initcount=%d bow=%d bowsum=%ld tri=%d trisum=%ld fish=%d fishsum=%ld empty=%d first=(%.1f,%.1f,%.3f,%d)\n",
           kL.size(), monoL, monoR, left.GetLevels(), left.mvImagePyramid[0].cols, left.mvImagePyramid[0].rows, matched, d01, acc, proj, projSelf, ini, iniCount, bow, bowSum, tri, triSum, fishMatches, fishSum, rcEmpty,
           kL[0].pt.x, kL[0].pt.y, kL[0].angle, kL[0].octave);
    return 0;
}
