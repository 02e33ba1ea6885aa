/* ORBextractor.h (orbfe shim) — ORB_SLAM3::ORBextractor with the reference's exact public surface
 * (/root/reference/include/ORBextractor.h:43-109), forwarding to the C ABI of liborbfe.so (include/orbfe.h).
 *
 * Drop-in use: put this directory before the reference's include/ on the include path and link liborbfe.so;
 * Frame.cc / Tracking.cc compile unchanged (constructor, operator(), the six getters, the public member
 * mvImagePyramid).  Differences a maintainer should know:
 *   - the GPU object is shape-specialised, so it is created lazily on the first operator() call (the reference
 *     constructor does not know the image size) and re-created if the image size changes;
 *   - mvImagePyramid is filled after each call from the device-resident pyramid (Frame::ComputeStereoMatches
 *     reads it, src/Frame.cc:804,894-909); define ORBFE_SHIM_LAZY_PYRAMID to skip the copy when stereo matching
 *     also runs through the shim (orbfe_stereo_match_batch_device needs no host pyramid);
 *   - no CPU fallback: if the device call fails, operator() throws std::runtime_error. */
#ifndef ORBFE_SHIM_ORBEXTRACTOR_H
#define ORBFE_SHIM_ORBEXTRACTOR_H

#include <list>
#include <stdexcept>
#include <string>
#include <vector>

#if defined(__has_include)
#if __has_include(<opencv2/core.hpp>) && !defined(ORBFE_SHIM_NO_OPENCV)
#include <opencv2/core.hpp>
#define ORBFE_SHIM_REAL_OPENCV 1
#endif
#endif
#ifndef ORBFE_SHIM_REAL_OPENCV
#include "opencv_min.hpp"
#endif

#include "../orbfe.h"

namespace ORB_SLAM3 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
        : nfeatures(nfeatures), scaleFactor(scaleFactor), nlevels(nlevels), iniThFAST(iniThFAST), minThFAST(minThFAST) {
        /* scale tables exactly as src/ORBextractor.cc:413-429 (double member times float accumulator) */
        mvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
        mvScaleFactor[0] = 1.0f; mvLevelSigma2[0] = 1.0f;
        for (int i = 1; i < nlevels; i++) {
            mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * this->scaleFactor);
            mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
        }
        mvInvScaleFactor.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        for (int i = 0; i < nlevels; i++) {
            mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
            mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
        }
        mvImagePyramid.resize(nlevels);
    }
    ~ORBextractor() {
        for (unsigned char* p : mvPinned) if (p) orbfe_unpin_host(p);
        if (mpImpl) orbfe_extractor_destroy(mpImpl);
    }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    /* src/ORBextractor.cc:1068-1150.  Mask is ignored, as in the reference. */
    int operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        (void)_mask;
        if (_image.empty()) return -1;
        cv::Mat image = _image.getMat();
        if (image.type() != CV_8UC1) throw std::runtime_error("ORBextractor: CV_8UC1 image expected");
        Ensure(image.cols, image.rows);
        const int cap = orbfe_keypoint_capacity(mpImpl);
        mvKeyBuf.resize(cap);
        mvDescBuf.resize((size_t)cap * 32);
        int n = 0, mono = 0;
        const int lap0 = vLappingArea.size() > 0 ? vLappingArea[0] : 0, lap1 = vLappingArea.size() > 1 ? vLappingArea[1] : 0;
        const int rc = orbfe_extract(mpImpl, image.ptr<unsigned char>(0), image.rows, image.cols, image.step, lap0, lap1,
                                     mvKeyBuf.data(), mvDescBuf.data(), cap, &n, &mono);
        if (rc == ORBFE_E_EMPTY) return -1;
        if (rc != ORBFE_OK) throw std::runtime_error(std::string("orbfe_extract: ") + orbfe_last_error());
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint and orbfe_keypoint must share a layout");
        _keypoints = std::vector<cv::KeyPoint>(n);
        if (n) std::memcpy((void*)_keypoints.data(), mvKeyBuf.data(), (size_t)n * sizeof(orbfe_keypoint));
        if (n == 0) _descriptors.release();
        else {
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; i++) std::memcpy(d.ptr<unsigned char>(i), mvDescBuf.data() + (size_t)i * 32, 32);
        }
#ifndef ORBFE_SHIM_LAZY_PYRAMID
        FetchPyramid();
#endif
        return mono;
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    std::vector<cv::Mat> mvImagePyramid;

    /* orbfe additions (not in the reference): the device object, for the stereo shim; and the pyramid fetch */
    orbfe_extractor* Impl() { return mpImpl; }
    void FetchPyramid() {
        /* like the reference, each level is an ROI of a bordered buffer (19 px), so reads slightly outside stay valid */
        mvPadded.resize(nlevels);
        std::vector<unsigned char*> dst(nlevels);
        std::vector<int> ws(nlevels), hs(nlevels);
        mvPinned.resize(nlevels, nullptr);
        for (int l = 0; l < nlevels; l++) {
            orbfe_level_size(mpImpl, l, &ws[l], &hs[l]);
            mvPadded[l].create(hs[l] + 38, ws[l] + 38, CV_8U);
            dst[l] = mvPadded[l].ptr<unsigned char>(0);
            if (mvPinned[l] != dst[l]) {   /* the mirrors are reused frame after frame: page-lock them once, the downloads run at link speed */
                if (mvPinned[l]) orbfe_unpin_host(mvPinned[l]);
                mvPinned[l] = orbfe_pin_host(dst[l], (size_t)(hs[l] + 38) * (ws[l] + 38)) == ORBFE_OK ? dst[l] : nullptr;
            }
        }
        if (orbfe_pyramid_levels(mpImpl, 0, 1, dst.data()) != ORBFE_OK)
            throw std::runtime_error(std::string("orbfe_pyramid_levels: ") + orbfe_last_error());
        for (int l = 0; l < nlevels; l++) mvImagePyramid[l] = cv::Mat(hs[l], ws[l], CV_8U, mvPadded[l].ptr<unsigned char>(19) + 19, mvPadded[l].step);
    }

protected:
    void Ensure(int w, int h) {
        if (mpImpl && w == mW && h == mH) return;
        if (mpImpl) { orbfe_extractor_destroy(mpImpl); mpImpl = nullptr; }
        orbfe_params p = {nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST};
        if (orbfe_extractor_create(&p, mDevice, w, h, 1, &mpImpl) != ORBFE_OK)
            throw std::runtime_error(std::string("orbfe_extractor_create: ") + orbfe_last_error());
        mW = w; mH = h;
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

    orbfe_extractor* mpImpl = nullptr;
    int mW = 0, mH = 0, mDevice = 0;
    std::vector<orbfe_keypoint> mvKeyBuf;
    std::vector<unsigned char> mvDescBuf;
    std::vector<cv::Mat> mvPadded;
    std::vector<unsigned char*> mvPinned;
};

}  // namespace ORB_SLAM3
#endif
