/*
 * oracle/refshim/ref_stubs.h — TEST INFRASTRUCTURE ONLY.
 *
 * Test doubles for the SLAM object graph that the reference's src/ORBmatcher.cc walks (Frame, KeyFrame, MapPoint,
 * GeometricCamera), so that the UNMODIFIED ORBmatcher.cc compiles and runs here without Eigen, DBoW2's boost
 * dependency, g2o or Pangolin.  oracle/build_ref.sh force-includes this file (-include) in front of
 * /root/reference/src/ORBmatcher.cc; it defines the include guards of include/MapPoint.h, KeyFrame.h and Frame.h, so
 * those headers (which ORBmatcher.h pulls in) expand to nothing and the classes below stand in for them.
 *
 * The doubles are plain data holders: every member is what the matcher reads, filled by oracle/ref_match_wrap.cc from
 * the SoA arrays of a test case.  Three pieces of behaviour live here, none of it matcher logic:
 *   - Frame / KeyFrame::GetFeaturesInArea forward to the oracle's grid (orc_grid_query, restating src/Frame.cc:784-871
 *     and src/KeyFrame.cc:916-962);
 *   - the camera double projects a point (x, y, z) to (x, y): test cases hand the projections in as "world points" with
 *     identity poses, so whatever pose arithmetic the matcher does (cv::Mat products, :2482-2518 etc.) is executed by
 *     the reference's own code and lands on the intended pixel;
 *   - epipolarConstrain / matchAndtriangulate answer through a callback the test installs.
 */
#pragma once
#define MAPPOINT_H
#define KEYFRAME_H
#define FRAME_H
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

#include <map>
#include <mutex>
#include <set>
#include <vector>

#include <opencv2/opencv.hpp>

#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

/* the reference's headers leak this into every file that includes them (e.g. include/ORBVocabulary.h, Converter.h);
 * include/ORBmatcher.h depends on it (unqualified `vector`, `pair`) */
using namespace std;

namespace ORB_SLAM3 {

class KeyFrame;
class Frame;

typedef int (*ref_verdict_fn)(void* user, int idx1, int idx2);

class GeometricCamera {
public:
    /* pinhole parameters for toK(); the projection itself is the identity on (x, y) - see the header comment */
    float fx = 1, fy = 1, cx = 0, cy = 0;
    ref_verdict_fn verdict = nullptr;      /* epipolarConstrain / matchAndtriangulate answer */
    void* user = nullptr;
    const std::vector<cv::KeyPoint>* keys1 = nullptr;   /* to recover idx1 / idx2 from the KeyPoint references handed in */
    const std::vector<cv::KeyPoint>* keys1b = nullptr;
    const std::vector<cv::KeyPoint>* keys2 = nullptr;
    const std::vector<cv::KeyPoint>* keys2b = nullptr;
    int nleft1 = 0, nleft2 = 0;

    cv::Point2f project(const cv::Point3f& p) { return cv::Point2f(p.x, p.y); }
    cv::Point2f project(const cv::Matx31f& m) { return cv::Point2f(m(0), m(1)); }
    cv::Point2f project(const cv::Mat& m) { return cv::Point2f(m.at<float>(0), m.at<float>(1)); }
    float uncertainty2(const cv::Point2f&) { return 1.f; }

    int index_of(const cv::KeyPoint& kp, const std::vector<cv::KeyPoint>* a, const std::vector<cv::KeyPoint>* b, int nleft) const
    {
        if (a && !a->empty() && &kp >= a->data() && &kp < a->data() + a->size()) return (int)(&kp - a->data());
        if (b && !b->empty() && &kp >= b->data() && &kp < b->data() + b->size()) return nleft + (int)(&kp - b->data());
        cv::cvshim_fail("camera double: keypoint reference outside the keyframe's vectors");
    }
    bool ask(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2)
    {
        if (!verdict) cv::cvshim_fail("camera double: no verdict callback installed");
        return verdict(user, index_of(kp1, keys1, keys1b, nleft1), index_of(kp2, keys2, keys2b, nleft2)) != 0;
    }
    bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat&, const cv::Mat&, float, float) { return ask(kp1, kp2); }
    bool epipolarConstrain_(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Matx33f&, const cv::Matx31f&, float, float) { return ask(kp1, kp2); }
    bool matchAndtriangulate(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, GeometricCamera*, cv::Mat&, cv::Mat&, float, float, cv::Mat& x3D)
    {
        const bool ok = ask(kp1, kp2);
        if (ok) x3D = cv::Mat::zeros(3, 1, CV_32F);
        return ok;
    }
};

class MapPoint {
public:
    /* what the matcher reads */
    cv::Mat mWorldPos;                  /* 3x1 CV_32F */
    cv::Mat mDescriptor;                /* 1x32 CV_8U */
    cv::Mat mNormal;
    bool mbBad = false;
    int nObs = 1;
    float mfMinDistance = 0.f, mfMaxDistance = 1e30f;
    int predicted_level = 0;            /* PredictScale answer (the caller's geometry) */
    /* Frame::isInFrustum leaves these (src/Frame.cc:603-711) */
    float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 0, mTrackDepthR = 0, mTrackProjXR = 0, mTrackProjYR = 0, mTrackViewCos = 1, mTrackViewCosR = 1;
    bool mbTrackInView = false, mbTrackInViewR = false;
    int mnTrackScaleLevel = 0, mnTrackScaleLevelR = 0;
    long unsigned int mnLastFrameSeen = 0;
    long unsigned int mnFuseCandidateForKF = 0;
    long unsigned int mnId = 0;
    int index = -1;                     /* row of the test case this point came from */
    std::map<KeyFrame*, std::pair<int, int> > obs;       /* IsInKeyFrame / GetIndexInKeyFrame */
    MapPoint* replaced = nullptr;

    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Matx31f GetWorldPos2() { return cv::Matx31f(mWorldPos.at<float>(0), mWorldPos.at<float>(1), mWorldPos.at<float>(2)); }
    cv::Mat GetNormal() { return mNormal.clone(); }
    cv::Matx31f GetNormal2() { return cv::Matx31f(mNormal.at<float>(0), mNormal.at<float>(1), mNormal.at<float>(2)); }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    bool isBad() { return mbBad; }
    int Observations() { return nObs; }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
    int PredictScale(const float&, KeyFrame*) { return predicted_level; }
    int PredictScale(const float&, Frame*) { return predicted_level; }
    bool IsInKeyFrame(KeyFrame* pKF) { return obs.count(pKF) != 0; }
    std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* pKF)
    {
        auto it = obs.find(pKF);
        return it == obs.end() ? std::tuple<int, int>(-1, -1) : std::tuple<int, int>(it->second.first, it->second.second);
    }
    /* Fuse mutates the map (:2000-2030): record what it did */
    void Replace(MapPoint* pMP) { replaced = pMP; }
    void AddObservation(KeyFrame* pKF, int idx) { obs[pKF] = std::make_pair(idx, -1); nObs++; }
};

/* fields shared by the Frame and KeyFrame doubles */
struct RefViewBase {
    int N = 0, Nleft = -1, NLeft = -1, NRight = -1, Nright = -1;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    int mnScaleLevels = 8;
    float mfScaleFactor = 1.2f, mfLogScaleFactor = 0.18232156f;
    float fx = 1, fy = 1, cx = 0, cy = 0, invfx = 1, invfy = 1, mbf = 0, mb = 0, mThDepth = 0;
    GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
    /* the oracle's grids (left / mono keypoints, right keypoints) */
    orc_grid* grid = nullptr; orc_grid* gridR = nullptr;

    std::vector<size_t> area(float x, float y, float r, int minLevel, int maxLevel, bool bRight) const
    {
        const bool rig = Nleft != -1 || NLeft != -1;
        const std::vector<cv::KeyPoint>& k = !rig ? mvKeysUn : (bRight ? mvKeysRight : mvKeys);
        const orc_grid* g = bRight ? gridR : grid;
        std::vector<int32_t> buf(k.size() + 1);
        const int n = g ? orc_grid_query(g, reinterpret_cast<const orc_keypoint*>(k.data()), x, y, r, minLevel, maxLevel, buf.data(), (int)k.size()) : 0;
        return std::vector<size_t>(buf.begin(), buf.begin() + n);
    }
};

class Frame : public RefViewBase {
public:
    cv::Mat mTcw, mTrl, mTlr, mRcw, mtcw, mOw;
    std::vector<bool> mvbOutlier;
    long unsigned int mnId = 0;
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1,
                                          const bool bRight = false) const
    {
        return area(x, y, r, minLevel, maxLevel, bRight);
    }
};

class KeyFrame : public RefViewBase {
public:
    cv::Mat Tcw, Trl;                   /* 4x4 CV_32F */
    long unsigned int mnId = 0;
    std::mutex mMutexFeatures;
    bool mbBad = false;

    cv::Mat GetPose() { return Tcw.clone(); }
    cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetCameraCenter() { return -GetRotation().t() * GetTranslation(); }
    cv::Matx33f GetRotation_() { cv::Matx33f R; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R(i, j) = Tcw.at<float>(i, j); return R; }
    cv::Matx31f GetTranslation_() { return cv::Matx31f(Tcw.at<float>(0, 3), Tcw.at<float>(1, 3), Tcw.at<float>(2, 3)); }
    cv::Matx31f GetCameraCenter_() { return -(GetRotation_().t() * GetTranslation_()); }
    cv::Mat GetRightPose() { return Trl * Tcw; }
    cv::Mat GetRightRotation() { return GetRightPose().rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetRightTranslation() { return GetRightPose().rowRange(0, 3).col(3).clone(); }
    cv::Mat GetRightCameraCenter() { return -GetRightRotation().t() * GetRightTranslation(); }
    cv::Matx33f GetRightRotation_() { cv::Mat R = GetRightRotation(); cv::Matx33f M; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) M(i, j) = R.at<float>(i, j); return M; }
    cv::Matx31f GetRightTranslation_() { cv::Mat t = GetRightTranslation(); return cv::Matx31f(t.at<float>(0), t.at<float>(1), t.at<float>(2)); }
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    std::set<MapPoint*> GetMapPoints() { std::set<MapPoint*> s; for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p); return s; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }   /* src/KeyFrame.cc:964 */
    bool isBad() { return mbBad; }
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight = false) const
    {
        return area(x, y, r, -1, -1, bRight);     /* src/KeyFrame.cc:916-962: no level filter */
    }
};

} // namespace ORB_SLAM3
