/* ref_camera.h — declarations of the camera-model classes whose member bodies oracle/Makefile cuts out of the reference's
 * src/CameraModels/Pinhole.cpp and KannalaBrandt8.cpp (the cv::Mat members that SearchForTriangulation and ComputeStereoFishEyeMatches reach).
 * Only what those bodies name (include/CameraModels/GeometricCamera.h, Pinhole.h, KannalaBrandt8.h).  TEST INFRASTRUCTURE ONLY. */
#ifndef ORB_REF_CAMERA_H
#define ORB_REF_CAMERA_H
#include <vector>

#include "refstub/cvstub.hpp"

namespace ORB_SLAM3 {

class GeometricCamera {
public:
    GeometricCamera() {}
    explicit GeometricCamera(const std::vector<float>& p) : mvParameters(p) {}
    virtual ~GeometricCamera() {}
    virtual cv::Point2f project(const cv::Point3f& p3D) = 0;
    virtual cv::Point2f project(const cv::Mat& m3D) = 0;
    virtual cv::Point3f unproject(const cv::Point2f& p2D) = 0;
    virtual cv::Mat unprojectMat(const cv::Point2f& p2D) = 0;
    virtual cv::Mat toK() = 0;
    virtual bool epipolarConstrain(GeometricCamera* otherCamera, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12,
                                   const float sigmaLevel, const float unc) = 0;
    std::vector<float> mvParameters;
};

class Pinhole : public GeometricCamera {
public:
    explicit Pinhole(const std::vector<float>& p) : GeometricCamera(p) {}
    cv::Point2f project(const cv::Point3f& p3D);
    cv::Point2f project(const cv::Mat& m3D);
    cv::Point3f unproject(const cv::Point2f&) { std::abort(); }      /* not reached by the bodies taken here */
    cv::Mat unprojectMat(const cv::Point2f&) { std::abort(); }
    cv::Mat toK();
    bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12, const float sigmaLevel,
                           const float unc);
    cv::Mat SkewSymmetricMatrix(const cv::Mat& v);
};

class KannalaBrandt8 : public GeometricCamera {
public:
    explicit KannalaBrandt8(const std::vector<float>& p) : GeometricCamera(p), precision(1e-6) {}   /* include/CameraModels/KannalaBrandt8.h:44 */
    cv::Point2f project(const cv::Point3f& p3D);
    cv::Point2f project(const cv::Mat& m3D);
    cv::Point3f unproject(const cv::Point2f& p2D);
    cv::Mat unprojectMat(const cv::Point2f& p2D);
    cv::Mat toK();
    bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12, const float sigmaLevel,
                           const float unc);
    float TriangulateMatches(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12, const float sigmaLevel,
                             const float unc, cv::Mat& p3D);
    void Triangulate(const cv::Point2f& p1, const cv::Point2f& p2, const cv::Mat& Tcw1, const cv::Mat& Tcw2, cv::Mat& x3D);
    const float precision;
};

}  // namespace ORB_SLAM3
#endif
