// ORBmatcher.h — the matcher-side shims over the C ABI: what src/ORBmatcher.cc / src/Frame.cc would call
// instead of their scalar loops.  Object-graph bookkeeping (MapPoint*, KeyFrame*) stays with the caller;
// the device sees SoA PODs.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "ORBextractor.h"

namespace ORB_SLAM3 {

struct ORBmatcherGPU {
    static const int TH_LOW = 50;       // src/ORBmatcher.cc:37
    static const int TH_HIGH = 100;     // src/ORBmatcher.cc:36
    static const int HISTO_LENGTH = 30; // src/ORBmatcher.cc:38

    // static int ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&), include/ORBmatcher.h:42
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbx_hamming256(a.data, b.data); }

    // Frame::ComputeStereoMatches(), include/Frame.h:114, src/Frame.cc:1027-1276: run right after the two
    // ExtractORB calls of the stereo Frame constructor (src/Frame.cc:132-137) on the same extractor objects.
    static void ComputeStereoMatches(ORBextractor& left, ORBextractor& right, int N, float mb, float mbf,
                                     std::vector<float>& mvuRight, std::vector<float>& mvDepth)
    {
        mvuRight.assign((size_t)N, -1.0f);
        mvDepth.assign((size_t)N, -1.0f);
        if (N == 0) return;
        if (orbx_stereo_match(left.handle(), right.handle(), 0, mb, mbf, mvuRight.data(), mvDepth.data(), N) != ORBX_OK)
            throw std::runtime_error(std::string("ComputeStereoMatches: ") + orbx_last_error());
    }

    // cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, matches, 2) of ComputeStereoFishEyeMatches, src/Frame.cc:1458
    static void KnnMatch2(const cv::Mat& q, const cv::Mat& t, std::vector<int>& idx, std::vector<int>& dist, int device = 0)
    {
        idx.assign((size_t)q.rows * 2, -1); dist.assign((size_t)q.rows * 2, -1);
        if (orbx_knn2_hamming(device, q.data, q.rows, t.data, t.rows, idx.data(), dist.data()) != ORBX_OK)
            throw std::runtime_error(std::string("KnnMatch2: ") + orbx_last_error());
    }

    // SoA view of the map points SearchByProjection visits (fields filled by Frame::isInFrustum,
    // src/Frame.cc:603-711).
    struct MapPointsSoA {
        std::vector<uint8_t> desc;      // n x 32
        std::vector<float> projX, projY, projXR, viewCos, depth;
        std::vector<int32_t> level;
        std::vector<uint8_t> inView;
    };

    // int ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)
    // include/ORBmatcher.h:46, src/ORBmatcher.cc:45-168 (F.Nleft == -1).  assign[i] >= 0 on entry marks a
    // keypoint whose F.mvpMapPoints[i] already has observations; on return it holds the map point index.
    static int SearchByProjection(const std::vector<cv::KeyPoint>& mvKeysUn, const cv::Mat& mDescriptors,
                                  const std::vector<float>* mvuRight, float mnMinX, float mnMaxX, float mnMinY, float mnMaxY,
                                  const std::vector<float>& mvScaleFactors, const MapPointsSoA& mp,
                                  float th, float mfNNratio, bool bFarPoints, float thFarPoints,
                                  std::vector<int32_t>& assign, int device = 0)
    {
        const int n = (int)mvKeysUn.size();
        assign.resize((size_t)n, -1);
        const int r = orbx_search_by_projection(
            device, reinterpret_cast<const orbx_keypoint*>(mvKeysUn.data()), mDescriptors.data, n,
            mvuRight ? mvuRight->data() : nullptr, mnMinX, mnMaxX, mnMinY, mnMaxY, mvScaleFactors.data(),
            (int)mvScaleFactors.size(), (int)mp.projX.size(), mp.desc.data(), mp.projX.data(), mp.projY.data(),
            mp.projXR.data(), mp.level.data(), mp.viewCos.data(), mp.depth.data(), mp.inView.data(), th, mfNNratio,
            bFarPoints ? 1 : 0, thFarPoints, assign.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByProjection: ") + orbx_last_error());
        return r;
    }
};

}  // namespace ORB_SLAM3
