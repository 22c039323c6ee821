/*
 * oracle/refshim_frame/ref_frame_stubs.h — TEST INFRASTRUCTURE ONLY.
 *
 * Force-included (-include) in front of the UNMODIFIED /root/reference/src/Frame.cc so that it compiles on its own: the REAL
 * include/Frame.h, ORBextractor.h, ORBVocabulary.h (vendored DBoW2) and Config.h are used; the headers below are replaced by
 * doubles that hold exactly what Frame.cc touches (their include guards are defined here, so the real files become empty):
 *   ImuTypes.h (IMU::Bias / Calib / Preintegrated), G2oTypes.h (ConstraintPoseImu), Converter.h (toDescriptorVector),
 *   MapPoint.h (the mTrack* fields isInFrustum writes + the getters it reads), KeyFrame.h (opaque),
 *   CameraModels/{GeometricCamera,Pinhole,KannalaBrandt8}.h (need Eigen), ORBmatcher.h (thresholds + DescriptorDistance).
 * The reference's own headers leak `using namespace std;` into Frame.h (it writes `vector<...>` unqualified); so does this file.
 */
#ifndef ORB_REF_FRAME_STUBS_H
#define ORB_REF_FRAME_STUBS_H

#define IMUTYPES_H
#define G2OTYPES_H
#define CONVERTER_H
#define MAPPOINT_H
#define KEYFRAME_H
#define CAMERAMODELS_GEOMETRICCAMERA_H
#define CAMERAMODELS_PINHOLE_H
#define CAMERAMODELS_KANNALABRANDT8_H
#define ORBMATCHER_H

#include <cmath>
#include <iostream>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>
#include <opencv2/opencv.hpp>

using namespace std;

extern "C" int orc_hamming256(const void* a, const void* b);

namespace ORB_SLAM3 {

class Frame;
class KeyFrame {};                 /* Frame.cc only stores pointers to keyframes */
class ConstraintPoseImu {};

namespace IMU {
struct Bias { float bax = 0, bay = 0, baz = 0, bwx = 0, bwy = 0, bwz = 0; };
struct Calib {
    cv::Mat Tcb, Tbc;              /* include/ImuTypes.h:99-100 */
    Calib() { Tcb = cv::Mat::eye(4, 4, CV_32F); Tbc = cv::Mat::eye(4, 4, CV_32F); }
};
struct Preintegrated { void SetNewBias(const Bias&) {} };
}

struct Converter {
    /* src/Converter.cc:28-36 */
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors)
    {
        std::vector<cv::Mat> vDesc;
        vDesc.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
        return vDesc;
    }
};

/* include/CameraModels/GeometricCamera.h; Pinhole::project is src/CameraModels/Pinhole.cpp:33-53 (fx * x / z + cx) */
class GeometricCamera {
public:
    std::vector<float> mvParameters;       /* fx fy cx cy */
    virtual ~GeometricCamera() {}
    virtual cv::Point2f project(const cv::Point3f& p3D) = 0;
    virtual cv::Point2f project(const cv::Matx31f& m3D) = 0;
    virtual cv::Point2f project(const cv::Mat& m3D) = 0;
    virtual cv::Mat toK() = 0;
};
class Pinhole : public GeometricCamera {
public:
    Pinhole(float fx, float fy, float cx, float cy) { mvParameters = {fx, fy, cx, cy}; }
    cv::Point2f project(const cv::Point3f& p3D) override
    {
        return cv::Point2f(mvParameters[0] * p3D.x / p3D.z + mvParameters[2], mvParameters[1] * p3D.y / p3D.z + mvParameters[3]);
    }
    cv::Point2f project(const cv::Matx31f& m3D) override { return this->project(cv::Point3f(m3D(0), m3D(1), m3D(2))); }
    cv::Point2f project(const cv::Mat& m3D) override
    {
        const float* p = m3D.ptr<float>();
        return this->project(cv::Point3f(p[0], p[1], p[2]));
    }
    cv::Mat toK() override
    {
        cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
        K.at<float>(0, 0) = mvParameters[0]; K.at<float>(1, 1) = mvParameters[1];
        K.at<float>(0, 2) = mvParameters[2]; K.at<float>(1, 2) = mvParameters[3];
        return K;
    }
};
/* the fisheye model: only the lapping area and the triangulation hook of ComputeStereoFishEyeMatches (src/Frame.cc:1470) */
class KannalaBrandt8 : public Pinhole {
public:
    KannalaBrandt8(float fx, float fy, float cx, float cy) : Pinhole(fx, fy, cx, cy), mvLappingArea(2, 0) {}
    std::vector<int> mvLappingArea;
    float precision = 1e-6f;
    /* test hook: the depth the double reports for a (left, right) keypoint pair */
    float (*triangulate)(void* user, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2) = nullptr;
    void* user = nullptr;
    float TriangulateMatches(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat&, const cv::Mat&,
                             const float, const float, cv::Mat& p3D)
    {
        const float z = triangulate ? triangulate(user, kp1, kp2) : 1.f;
        p3D = cv::Mat(3, 1, CV_32F);
        p3D.at<float>(0) = kp1.pt.x; p3D.at<float>(1) = kp1.pt.y; p3D.at<float>(2) = z;
        return z;
    }
};

/* include/MapPoint.h: the tracking fields (:113-130) and the getters Frame::isInFrustum reads; PredictScale is
 * src/MapPoint.cc:725-741 */
class MapPoint {
public:
    bool mbTrackInView = false, mbTrackInViewR = false;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 0, mTrackDepthR = 0, mTrackProjXR = 0, mTrackProjYR = 0;
    int mnTrackScaleLevel = 0, mnTrackScaleLevelR = 0;
    float mTrackViewCos = 0, mTrackViewCosR = 0;
    cv::Mat mWorldPos, mNormalVector;
    cv::Matx31f mWorldPosx, mNormalVectorx;
    float mfMinDistance = 0, mfMaxDistance = 0;
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Matx31f GetWorldPos2() { return mWorldPosx; }
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    cv::Matx31f GetNormal2() { return mNormalVectorx; }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
    int PredictScale(const float& currentDist, Frame* pF);
};

/* include/ORBmatcher.h:39-45,98-99 and src/ORBmatcher.cc:33-35; DescriptorDistance (:2911-2931) is pinned separately against
 * the compiled ORBmatcher.cc (tests/test_oracle_ref_match.py::test_descriptor_distance_and_three_maxima) */
class ORBmatcher {
public:
    static const int TH_LOW = 50;
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orc_hamming256(a.ptr<uint8_t>(), b.ptr<uint8_t>()); }
};

} // namespace ORB_SLAM3

#endif
