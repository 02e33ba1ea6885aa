// Compiles the C++ shim (include/orbfe_shim) the way Frame.cc uses the reference classes and runs one stereo frame.
// Built and run by tests/test_shim.py (GPU) / syntax-checked on CPU.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "orbfe_shim/ORBmatcher.h"
extern "C" int orbfe_synth_stereo(uint64_t frame_index, int W, int H, uint8_t* left, uint8_t* right);
using namespace ORB_SLAM3;
int main(int argc, char** argv) {
    const int W = 752, H = 480;
    cv::Mat imL(H, W, CV_8UC1), imR(H, W, CV_8UC1), mask;
    orbfe_synth_stereo(0, W, H, imL.ptr<unsigned char>(0), imR.ptr<unsigned char>(0));
    ORBextractor left(1200, 1.2f, 8, 20, 7), right(1200, 1.2f, 8, 20, 7);   // src/Tracking.cc:1151-1154
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR;
    std::vector<int> lap = {0, 0};
    const int monoL = left(imL, mask, kL, dL, lap);                           // src/Frame.cc:413-420
    const int monoR = right(imR, mask, kR, dR, lap);
    std::vector<float> uR, depth;
    ComputeStereoMatches(&left, &right, (int)kL.size(), 47.90639f, 435.2047f, uR, depth);
    int matched = 0;
    for (float v : uR) matched += v >= 0;
    const int d01 = ORBmatcher::DescriptorDistance(dL.row(0), dL.row(1));
    ORBmatcher m(0.9f, true);
    std::vector<orbfe_match> res;
    const int acc = m.SearchTop2(dL, dR, res);
    // Tracking::SearchLocalPoints-like: every left keypoint re-projected half a pixel off as a map point (src/ORBmatcher.cc:44-214)
    std::vector<orbfe_proj_query> pq(kL.size());
    for (size_t i = 0; i < kL.size(); i++) {
        pq[i].x = kL[i].pt.x + 0.5f; pq[i].y = kL[i].pt.y - 0.5f;
        pq[i].radius = 4.0f * left.GetScaleFactors()[kL[i].octave];
        pq[i].xr = uR[i]; pq[i].min_level = kL[i].octave - 1; pq[i].max_level = kL[i].octave; pq[i].flags = ORBFE_PQ_BLOCKS; pq[i].angle = 0.f;
    }
    std::vector<unsigned char> taken;
    std::vector<orbfe_proj_match> pm;
    const int proj = SearchByProjection(kL, dL, -1, &uR, nullptr, taken, 0.f, 0.f, 64.f / W, 48.f / H, pq, dL, 0.8f, pm);
    int projSelf = 0;
    for (size_t i = 0; i < pm.size(); i++) projSelf += pm[i].accepted && pm[i].best_idx == (int)i;
    cv::Mat empty;
    const int rcEmpty = left(empty, mask, kR, dR, lap);
    printf("shim n=%zu mono=%d monoR=%d levels=%d pyr0=%dx%d stereo=%d d01=%d top2acc=%d proj=%d projself=%d empty=%d first=(%.1f,%.1f,%.3f,%d)\n",
           kL.size(), monoL, monoR, left.GetLevels(), left.mvImagePyramid[0].cols, left.mvImagePyramid[0].rows, matched, d01, acc, proj, projSelf, rcEmpty,
           kL[0].pt.x, kL[0].pt.y, kL[0].angle, kL[0].octave);
    return 0;
}
