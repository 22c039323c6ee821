/*
 * oracle/refshim_mp/ref_mp_stubs.h — TEST INFRASTRUCTURE ONLY.
 *
 * Force-included in front of the UNMODIFIED /root/reference/src/MapPoint.cc: the REAL include/MapPoint.h is used; KeyFrame.h,
 * Frame.h, Map.h and ORBmatcher.h are replaced by doubles holding exactly what MapPoint.cc touches (their include guards are
 * defined here).  The reference's headers leak `using namespace std;`; so does this file.
 */
#ifndef ORB_REF_MP_STUBS_H
#define ORB_REF_MP_STUBS_H

#define KEYFRAME_H
#define FRAME_H
#define MAP_H
#define ORBMATCHER_H

#include <cmath>
#include <iostream>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <tuple>
#include <vector>
#include <opencv2/opencv.hpp>

using namespace std;

extern "C" int orc_hamming256(const void* a, const void* b);

namespace ORB_SLAM3 {

class MapPoint;
class GeometricCamera;

/* fields shared by the Frame and KeyFrame doubles */
struct RefMpView {
    long unsigned int mnId = 0, mnFrameId = 0;
    int N = 0, Nleft = -1, NLeft = -1;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    std::vector<float> mvScaleFactors;
    int mnScaleLevels = 8;
    float mfLogScaleFactor = 0.18232156f;
    GeometricCamera* mpCamera2 = nullptr;
    cv::Mat Ow = cv::Mat::zeros(3, 1, CV_32F), OwR = cv::Mat::zeros(3, 1, CV_32F);
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    cv::Mat GetRightCameraCenter() { return OwR.clone(); }
};
class Frame : public RefMpView {
public:
    cv::Mat mRwc = cv::Mat::eye(3, 3, CV_32F), mTlr = cv::Mat::eye(4, 4, CV_32F), mOw = cv::Mat::zeros(3, 1, CV_32F);   /* :104-107, rig branch */
};
class KeyFrame : public RefMpView {
public:
    bool mbBad = false;
    bool isBad() { return mbBad; }
    std::vector<std::pair<int, MapPoint*> > log;        /* what MapPoint told the keyframe */
    void EraseMapPointMatch(const int& idx) { log.push_back(std::make_pair(idx, (MapPoint*)nullptr)); }
    void ReplaceMapPointMatch(const int& idx, MapPoint* pMP) { log.push_back(std::make_pair(idx, pMP)); }
};
class Map {
public:
    std::mutex mMutexPointCreation;
    std::vector<MapPoint*> erased;
    void EraseMapPoint(MapPoint* pMP) { erased.push_back(pMP); }
    long unsigned int GetId() { return 0; }
};

/* include/ORBmatcher.h:39-45; DescriptorDistance pinned separately against the compiled ORBmatcher.cc */
class ORBmatcher {
public:
    static const int TH_LOW = 50;
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orc_hamming256(a.ptr<uint8_t>(), b.ptr<uint8_t>()); }
};

} // namespace ORB_SLAM3

#endif
