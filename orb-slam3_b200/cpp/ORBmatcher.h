// ORBmatcher.h — the matcher-side shims over the C ABI: what src/ORBmatcher.cc / src/Frame.cc would call
// instead of their scalar loops.  Object-graph bookkeeping (MapPoint*, KeyFrame*) stays with the caller;
// the device sees SoA PODs.
#pragma once
#include <stdexcept>
#include <type_traits>
#include <string>
#include <utility>
#include <vector>

#include "ORBextractor.h"

namespace ORB_SLAM3 {

struct ORBmatcherGPU {
    static const int TH_LOW = 50;       // src/ORBmatcher.cc:37
    static const int TH_HIGH = 100;     // src/ORBmatcher.cc:36
    static const int HISTO_LENGTH = 30; // src/ORBmatcher.cc:38

    // ORBmatcher::ORBmatcher(float nnratio = 0.6, bool checkOri = true), include/ORBmatcher.h:39, src/ORBmatcher.cc:40
    explicit ORBmatcherGPU(float nnratio = 0.6f, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
    float mfNNratio;
    bool mbCheckOrientation;

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

    // The stereo Frame constructor's hot part as one call (src/Frame.cc:132-137: threadLeft / threadRight ExtractORB with
    // vLappingArea {0,0}; :143 N = mvKeys.size(); :154 ComputeStereoMatches): both eyes are extracted concurrently on the
    // device and the match is queued behind them - one host wait instead of two thread joins and a third call.
    static void ExtractStereo(ORBextractor& left, ORBextractor& right, const cv::Mat& imLeft, const cv::Mat& imRight,
                              std::vector<cv::KeyPoint>& mvKeys, cv::Mat& mDescriptors,
                              std::vector<cv::KeyPoint>& mvKeysRight, cv::Mat& mDescriptorsRight,
                              float mb, float mbf, std::vector<float>& mvuRight, std::vector<float>& mvDepth)
    {
        const int cap = orbx_max_keypoints(left.handle(), imLeft.rows, imLeft.cols);
        if (cap < 0) throw std::runtime_error(std::string("ExtractStereo: ") + orbx_last_error());
        mvKeys.resize((size_t)cap); mvKeysRight.resize((size_t)cap);
        std::vector<uint8_t> dl((size_t)cap * 32), dr((size_t)cap * 32);
        mvuRight.assign((size_t)cap, -1.0f); mvDepth.assign((size_t)cap, -1.0f);
        int nL = 0, nR = 0;
        if (orbx_extract_stereo(left.handle(), right.handle(), imLeft.data, imLeft.step, imRight.data, imRight.step,
                                imLeft.rows, imLeft.cols, reinterpret_cast<orbx_keypoint*>(mvKeys.data()), dl.data(), &nL,
                                reinterpret_cast<orbx_keypoint*>(mvKeysRight.data()), dr.data(), &nR, cap, mb, mbf,
                                mvuRight.data(), mvDepth.data()) != ORBX_OK)
            throw std::runtime_error(std::string("ExtractStereo: ") + orbx_last_error());
        mvKeys.resize((size_t)nL); mvKeysRight.resize((size_t)nR);
        mvuRight.resize((size_t)nL); mvDepth.resize((size_t)nL);
        auto fill = [](cv::Mat& D, const std::vector<uint8_t>& src, int n) {
            if (n == 0) { D.release(); return; }
            ORBX_MAT_CREATE(D, n, 32);
            for (int i = 0; i < n; i++) std::memcpy(D.ptr(i), src.data() + (size_t)i * 32, 32);
        };
        fill(mDescriptors, dl, nL); fill(mDescriptorsRight, dr, nR);
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

    // View of the frame fields the guided searches read (mono / rectified stereo, Nleft == -1).
    struct FrameView {
        const std::vector<cv::KeyPoint>* mvKeysUn; const cv::Mat* mDescriptors;
        const std::vector<float>* mvuRight;            // nullptr = monocular
        float mnMinX, mnMaxX, mnMinY, mnMaxY;          // Frame::ComputeImageBounds, src/Frame.cc:978-1009
        const std::vector<float>* mvScaleFactors;
        int N() const { return (int)mvKeysUn->size(); }
        const orbx_keypoint* kps() const { return reinterpret_cast<const orbx_keypoint*>(mvKeysUn->data()); }
    };

    // Rows of the last frame as SearchByProjection(Frame&, const Frame&, ...) visits them: map point
    // descriptor, projection into the current camera (src/ORBmatcher.cc:2508-2518), octave, angle, flags.
    struct LastFrameSoA {
        std::vector<uint8_t> desc;                     // n x 32
        std::vector<float> u, v, invz, angle;
        std::vector<int32_t> octave;
        std::vector<uint8_t> valid, obs;               // pMP && !mvbOutlier[i] ; pMP->Observations() > 0
    };

    // int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono)
    // include/ORBmatcher.h:50, src/ORBmatcher.cc:2469-2711.  assign[i] = row of `last` now held by keypoint i.
    int SearchByProjection(const FrameView& CurrentFrame, const LastFrameSoA& last, float th, float mbf,
                           bool bForward, bool bBackward, const std::vector<uint8_t>* curHeld,
                           std::vector<int32_t>& assign, int device = 0) const
    {
        const int n = CurrentFrame.N();
        assign.assign((size_t)n, -1);
        const int r = orbx_search_by_projection_frame(
            device, CurrentFrame.kps(), CurrentFrame.mDescriptors->data, n,
            CurrentFrame.mvuRight ? CurrentFrame.mvuRight->data() : nullptr,
            CurrentFrame.mnMinX, CurrentFrame.mnMaxX, CurrentFrame.mnMinY, CurrentFrame.mnMaxY,
            CurrentFrame.mvScaleFactors->data(), (int)CurrentFrame.mvScaleFactors->size(),
            (int)last.u.size(), last.desc.data(), last.u.data(), last.v.data(), last.invz.data(), last.octave.data(),
            last.angle.data(), last.valid.data(), last.obs.data(), th, mbf, bForward ? 1 : 0, bBackward ? 1 : 0,
            mbCheckOrientation ? 1 : 0, curHeld ? curHeld->data() : nullptr, assign.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByProjection(Frame,Frame): ") + orbx_last_error());
        return r;
    }

    // Rows of a keyframe as SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) visits them.
    struct KeyFramePointsSoA {
        std::vector<uint8_t> desc;                     // n x 32: pMP->GetDescriptor()
        std::vector<float> u, v, angle;                // projection into the current camera, pKF->mvKeysUn[i].angle
        std::vector<int32_t> level;                    // pMP->PredictScale(dist3D, &CurrentFrame)
        std::vector<uint8_t> valid;                    // pMP && !isBad() && !sAlreadyFound.count(pMP) && distance gate
    };

    // int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound,
    // const float th, const int ORBdist)  include/ORBmatcher.h:54, src/ORBmatcher.cc:2723-2862
    int SearchByProjection(const FrameView& CurrentFrame, const KeyFramePointsSoA& kf, float th, int ORBdist,
                           const std::vector<uint8_t>* curHeld, std::vector<int32_t>& assign, int device = 0) const
    {
        const int n = CurrentFrame.N();
        assign.assign((size_t)n, -1);
        const int r = orbx_search_by_projection_keyframe(
            device, CurrentFrame.kps(), CurrentFrame.mDescriptors->data, n,
            CurrentFrame.mnMinX, CurrentFrame.mnMaxX, CurrentFrame.mnMinY, CurrentFrame.mnMaxY,
            CurrentFrame.mvScaleFactors->data(), (int)CurrentFrame.mvScaleFactors->size(),
            (int)kf.u.size(), kf.desc.data(), kf.u.data(), kf.v.data(), kf.level.data(), kf.angle.data(), kf.valid.data(),
            th, ORBdist, mbCheckOrientation ? 1 : 0, curHeld ? curHeld->data() : nullptr, assign.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByProjection(Frame,KeyFrame): ") + orbx_last_error());
        return r;
    }

    // int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched,
    // vector<int>& vnMatches12, int windowSize = 10)  include/ORBmatcher.h:71, src/ORBmatcher.cc:799-943
    int SearchForInitialization(const FrameView& F1, const FrameView& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10, int device = 0) const
    {
        static_assert(sizeof(cv::Point2f) == 8, "Point2f is two floats");
        vnMatches12.assign((size_t)F1.N(), -1);
        const int r = orbx_search_for_initialization(
            device, F1.kps(), F1.mDescriptors->data, F1.N(), F2.kps(), F2.mDescriptors->data, F2.N(),
            F2.mnMinX, F2.mnMaxX, F2.mnMinY, F2.mnMaxY, reinterpret_cast<float*>(vbPrevMatched.data()), windowSize,
            mfNNratio, mbCheckOrientation ? 1 : 0, vnMatches12.data());
        if (r < 0) throw std::runtime_error(std::string("SearchForInitialization: ") + orbx_last_error());
        return r;
    }

    // View of a keyframe for SearchForTriangulation: the DBoW2::FeatureVector flattened to CSR.
    struct KeyFrameView {
        const std::vector<cv::KeyPoint>* mvKeysUn; const cv::Mat* mDescriptors;
        const std::vector<float>* mvuRight;            // nullptr = monocular
        std::vector<uint8_t> hasMapPoint;              // GetMapPoint(idx) != NULL
        std::vector<int32_t> fvNode, fvStart, fvIdx;   // mFeatVec: ascending node ids, CSR offsets, keypoint indices
        const std::vector<float>* mvScaleFactors; const std::vector<float>* mvLevelSigma2;
        // flatten any ordered map<node id, vector<unsigned>> (DBoW2::FeatureVector)
        template <class FeatVec> void SetFeatVec(const FeatVec& fv)
        {
            fvNode.clear(); fvIdx.clear(); fvStart.assign(1, 0);
            for (const auto& kv : fv) {
                fvNode.push_back((int32_t)kv.first);
                for (auto i : kv.second) fvIdx.push_back((int32_t)i);
                fvStart.push_back((int32_t)fvIdx.size());
            }
        }
    };

    // int ORBmatcher::SearchForTriangulation(KeyFrame*, KeyFrame*, cv::Mat F12, vector<pair<size_t,size_t>>&,
    // bOnlyStereo, bCoarse) include/ORBmatcher.h:74, src/ORBmatcher.cc:1107-1386; underscore = true selects
    // SearchForTriangulation_ (include/ORBmatcher.h:77, :1388-1630).  ep = epipole of KF1 in KF2 (:1120),
    // F12 = the matrix Pinhole::epipolarConstrain forms (src/CameraModels/Pinhole.cpp:243), row-major.
    int SearchForTriangulation(const KeyFrameView& KF1, const KeyFrameView& KF2, const float F12[9], const float ep[2],
                               std::vector<std::pair<size_t, size_t>>& vMatchedPairs, bool bOnlyStereo, bool bCoarse = false,
                               bool underscore = false, int device = 0) const
    {
        const int n1 = (int)KF1.mvKeysUn->size(), n2 = (int)KF2.mvKeysUn->size();
        std::vector<int32_t> m12((size_t)n1, -1);
        const int r = orbx_search_for_triangulation(
            device, reinterpret_cast<const orbx_keypoint*>(KF1.mvKeysUn->data()), KF1.mDescriptors->data, n1,
            KF1.hasMapPoint.data(), KF1.mvuRight ? KF1.mvuRight->data() : nullptr,
            KF1.fvNode.data(), KF1.fvStart.data(), KF1.fvIdx.data(), (int)KF1.fvNode.size(),
            reinterpret_cast<const orbx_keypoint*>(KF2.mvKeysUn->data()), KF2.mDescriptors->data, n2,
            KF2.hasMapPoint.data(), KF2.mvuRight ? KF2.mvuRight->data() : nullptr,
            KF2.fvNode.data(), KF2.fvStart.data(), KF2.fvIdx.data(), (int)KF2.fvNode.size(),
            KF2.mvScaleFactors->data(), KF2.mvLevelSigma2->data(), (int)KF2.mvScaleFactors->size(), ep, F12,
            bOnlyStereo ? 1 : 0, bCoarse ? 1 : 0, mbCheckOrientation ? 1 : 0, underscore ? 1 : 0, m12.data());
        if (r < 0) throw std::runtime_error(std::string("SearchForTriangulation: ") + orbx_last_error());
        vMatchedPairs.clear();                                              // :1358-1367
        vMatchedPairs.reserve((size_t)r);
        for (int i = 0; i < n1; i++)
            if (m12[(size_t)i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[(size_t)i]));
        return r;
    }

    // int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) include/ORBmatcher.h:67,
    // src/ORBmatcher.cc:314-547 (bKeyFrames = false) and SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)
    // include/ORBmatcher.h:68, :955-1105 (bKeyFrames = true).  For side 1 (and side 2 of the KF-KF form)
    // hasMapPoint must mean "map point present and !isBad()".  matches: KF-Frame -> per Frame keypoint the index of
    // the KF keypoint whose map point it receives; KF-KF -> per KF1 keypoint the matched KF2 keypoint.
    int SearchByBoW(const KeyFrameView& KF1, const KeyFrameView& KF2, bool bKeyFrames, std::vector<int32_t>& matches, int device = 0) const
    {
        const int n1 = (int)KF1.mvKeysUn->size(), n2 = (int)KF2.mvKeysUn->size();
        matches.assign((size_t)(bKeyFrames ? n1 : n2), -1);
        const int r = orbx_search_by_bow(
            device, bKeyFrames ? 1 : 0,
            reinterpret_cast<const orbx_keypoint*>(KF1.mvKeysUn->data()), KF1.mDescriptors->data, n1, KF1.hasMapPoint.data(),
            KF1.fvNode.data(), KF1.fvStart.data(), KF1.fvIdx.data(), (int)KF1.fvNode.size(),
            reinterpret_cast<const orbx_keypoint*>(KF2.mvKeysUn->data()), KF2.mDescriptors->data, n2,
            bKeyFrames ? KF2.hasMapPoint.data() : nullptr,
            KF2.fvNode.data(), KF2.fvStart.data(), KF2.fvIdx.data(), (int)KF2.fvNode.size(),
            mfNNratio, mbCheckOrientation ? 1 : 0, matches.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByBoW: ") + orbx_last_error());
        return r;
    }

    // The two-camera rig form (F.Nleft != -1) of SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints,
    // thFarPoints), src/ORBmatcher.cc:45-239: the fields Frame::isInFrustum leaves per camera, SoA.
    struct RigMapPointsSoA {
        std::vector<uint8_t> desc;                                   // n x 32
        std::vector<float> depth;                                    // mTrackDepth
        std::vector<uint8_t> inView, inViewR;                        // mbTrackInView, mbTrackInViewR
        std::vector<float> projX, projY, viewCos, projXR, projYR, viewCosR;
        std::vector<int32_t> level, levelR;                          // mnTrackScaleLevel, mnTrackScaleLevelR (-1 allowed)
    };
    // mvKeys / mvKeysRight / mDescriptors (left rows first) / mvLeftToRightMatch / mvRightToLeftMatch of the frame;
    // assign = F.mvpMapPoints as rows of `mp` (>= 0 on entry: holds a map point with observations), right keypoints at Nleft + i.
    int SearchByProjection(const std::vector<cv::KeyPoint>& mvKeys, const std::vector<cv::KeyPoint>& mvKeysRight, const cv::Mat& mDescriptors,
                           const std::vector<int>& mvLeftToRightMatch, const std::vector<int>& mvRightToLeftMatch,
                           float mnMinX, float mnMaxX, float mnMinY, float mnMaxY, const std::vector<float>& mvScaleFactors,
                           const RigMapPointsSoA& mp, float th, bool bFarPoints, float thFarPoints,
                           std::vector<int32_t>& assign, int device = 0) const
    {
        static_assert(sizeof(int) == sizeof(int32_t), "int is 32 bits");
        const int nL = (int)mvKeys.size(), nR = (int)mvKeysRight.size();
        assign.resize((size_t)(nL + nR), -1);
        const int r = orbx_search_by_projection_rig(
            device, reinterpret_cast<const orbx_keypoint*>(mvKeys.data()), nL, reinterpret_cast<const orbx_keypoint*>(mvKeysRight.data()), nR,
            mDescriptors.data, mvLeftToRightMatch.data(), mvRightToLeftMatch.data(), mnMinX, mnMaxX, mnMinY, mnMaxY,
            mvScaleFactors.data(), (int)mvScaleFactors.size(), (int)mp.depth.size(), mp.desc.data(), mp.depth.data(),
            mp.inView.data(), mp.projX.data(), mp.projY.data(), mp.level.data(), mp.viewCos.data(),
            mp.inViewR.data(), mp.projXR.data(), mp.projYR.data(), mp.levelR.data(), mp.viewCosR.data(),
            th, mfNNratio, bFarPoints ? 1 : 0, thFarPoints, assign.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByProjection(rig): ") + orbx_last_error());
        return r;
    }

    // ---- two-camera rig (Nleft != -1) forms of the other guided searches ----

    // Rows of the last frame for the rig form of SearchByProjection(Frame&, const Frame&, th, bMono): as LastFrameSoA plus
    // the projection of Trl * x3Dc into the right camera (src/ORBmatcher.cc:2613-2616).
    struct LastFrameRigSoA : LastFrameSoA { std::vector<float> ur, vr; };

    // int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), CurrentFrame.Nleft != -1
    // (src/ORBmatcher.cc:2469-2711; right camera :2610-2684).  mDescriptors holds the left rows first (as Frame does, :2647);
    // assign[i] / assign[Nleft + j] = row of `last` now held by left keypoint i / right keypoint j.
    int SearchByProjection(const std::vector<cv::KeyPoint>& mvKeys, const std::vector<cv::KeyPoint>& mvKeysRight, const cv::Mat& mDescriptors,
                           float mnMinX, float mnMaxX, float mnMinY, float mnMaxY, const std::vector<float>& mvScaleFactors,
                           const LastFrameRigSoA& last, float th, bool bForward, bool bBackward,
                           const std::vector<uint8_t>* curHeld, std::vector<int32_t>& assign, int device = 0) const
    {
        const int nL = (int)mvKeys.size(), nR = (int)mvKeysRight.size();
        assign.assign((size_t)(nL + nR), -1);
        const int r = orbx_search_by_projection_frame_rig(
            device, reinterpret_cast<const orbx_keypoint*>(mvKeys.data()), nL, reinterpret_cast<const orbx_keypoint*>(mvKeysRight.data()), nR,
            mDescriptors.data, mnMinX, mnMaxX, mnMinY, mnMaxY, mvScaleFactors.data(), (int)mvScaleFactors.size(),
            (int)last.u.size(), last.desc.data(), last.u.data(), last.v.data(), last.ur.data(), last.vr.data(), last.invz.data(),
            last.octave.data(), last.angle.data(), last.valid.data(), last.obs.data(), th, bForward ? 1 : 0, bBackward ? 1 : 0,
            mbCheckOrientation ? 1 : 0, curHeld ? curHeld->data() : nullptr, assign.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByProjection(Frame,Frame) rig: ") + orbx_last_error());
        return r;
    }

    // int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>&) with F.Nleft != -1 (src/ORBmatcher.cc:386-430,
    // :455-486).  KF2.mvKeysUn must point at the Frame's keypoints in mvpMapPoints order: mvKeys rows first, then mvKeysRight
    // (the angles :442-448 reads); Nleft = F.Nleft.
    int SearchByBoWRig(const KeyFrameView& KF1, const KeyFrameView& F, int Nleft, std::vector<int32_t>& matches, int device = 0) const
    {
        const int n1 = (int)KF1.mvKeysUn->size(), n2 = (int)F.mvKeysUn->size();
        matches.assign((size_t)n2, -1);
        const int r = orbx_search_by_bow_rig(
            device, reinterpret_cast<const orbx_keypoint*>(KF1.mvKeysUn->data()), KF1.mDescriptors->data, n1, KF1.hasMapPoint.data(),
            KF1.fvNode.data(), KF1.fvStart.data(), KF1.fvIdx.data(), (int)KF1.fvNode.size(),
            reinterpret_cast<const orbx_keypoint*>(F.mvKeysUn->data()), F.mDescriptors->data, n2, Nleft,
            F.fvNode.data(), F.fvStart.data(), F.fvIdx.data(), (int)F.fvNode.size(), mfNNratio, mbCheckOrientation ? 1 : 0, matches.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByBoW rig: ") + orbx_last_error());
        return r;
    }

    // The triangulation searches whose geometric test belongs to the camera model (a KannalaBrandt8 rig):
    //   overload 0: SearchForTriangulation (include/ORBmatcher.h:74, src/ORBmatcher.cc:1107) with pKF1->mpCamera2 && pKF2->mpCamera2:
    //               verdict(idx1, idx2) = pCamera1->epipolarConstrain(pCamera2, kp1, kp2, R12, t12, sigma1, sigma2) with the
    //               camera / pose pair of :1252-1279 (bRight1 = idx1 >= NLeft, bRight2 = idx2 >= NLeft)
    //   overload 1: SearchForTriangulation_ (:1388), epipolarConstrain_
    //   overload 2: SearchForTriangulation(..., vector<cv::Mat>& vMatchedPoints) (:1632): verdict = matchAndtriangulate(...);
    //               let it remember x3D per idx1 when it accepts to rebuild vMatchedPoints (:1733-1735, :1815)
    // Verdict is any callable bool(int idx1, int idx2); it is invoked in the reference's call order from this thread.
    template <class Verdict>
    int SearchForTriangulation(int overload, const KeyFrameView& KF1, const KeyFrameView& KF2, Verdict&& verdict,
                               std::vector<std::pair<size_t, size_t>>& vMatchedPairs, bool bOnlyStereo, bool bCoarse = false, int device = 0) const
    {
        const int n1 = (int)KF1.mvKeysUn->size(), n2 = (int)KF2.mvKeysUn->size();
        std::vector<int32_t> m12((size_t)n1, -1);
        auto thunk = [](void* u, int i1, int i2) -> int { return (*static_cast<typename std::remove_reference<Verdict>::type*>(u))(i1, i2) ? 1 : 0; };
        const int r = orbx_search_for_triangulation_cb(
            device, overload, reinterpret_cast<const orbx_keypoint*>(KF1.mvKeysUn->data()), KF1.mDescriptors->data, n1, KF1.hasMapPoint.data(),
            KF1.fvNode.data(), KF1.fvStart.data(), KF1.fvIdx.data(), (int)KF1.fvNode.size(),
            reinterpret_cast<const orbx_keypoint*>(KF2.mvKeysUn->data()), KF2.mDescriptors->data, n2, KF2.hasMapPoint.data(),
            KF2.fvNode.data(), KF2.fvStart.data(), KF2.fvIdx.data(), (int)KF2.fvNode.size(),
            bOnlyStereo ? 1 : 0, bCoarse ? 1 : 0, mbCheckOrientation ? 1 : 0, thunk, &verdict, m12.data());
        if (r < 0) throw std::runtime_error(std::string("SearchForTriangulation(rig / matchAndtriangulate): ") + orbx_last_error());
        vMatchedPairs.clear();
        vMatchedPairs.reserve((size_t)r);
        for (int i = 0; i < n1; i++)
            if (m12[(size_t)i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[(size_t)i]));
        return r;
    }

    // void Frame::ComputeStereoFishEyeMatches(), include/Frame.h:304, src/Frame.cc:1440-1480, up to the triangulation gate, on
    // the descriptors the two ExtractORB calls (src/Frame.cc:1363-1366) left on the device: left / right are the two
    // extractor objects, Nleft / Nright the keypoint counts, monoLeft / monoRight the values operator() returned.
    // Fills the CANDIDATE mvLeftToRightMatch / mvRightToLeftMatch (the caller keeps a pair iff TriangulateMatches > 0.0001, :1471).
    static int ComputeStereoFishEyeMatches(ORBextractor& left, ORBextractor& right, int Nleft, int monoLeft, int Nright, int monoRight,
                                           std::vector<int>& mvLeftToRightMatch, std::vector<int>& mvRightToLeftMatch)
    {
        mvLeftToRightMatch.assign((size_t)Nleft, -1); mvRightToLeftMatch.assign((size_t)Nright, -1);
        const int r = orbx_stereo_fisheye_match(left.handle(), right.handle(), 0, Nleft, monoLeft, Nright, monoRight,
                                                mvLeftToRightMatch.data(), mvRightToLeftMatch.data(), nullptr);
        if (r < 0) throw std::runtime_error(std::string("ComputeStereoFishEyeMatches: ") + orbx_last_error());
        return r;
    }

    // void MapPoint::ComputeDistinctiveDescriptors() src/MapPoint.cc:450-531 for a batch of map points: start is the CSR
    // offset of each point's observation descriptors in `desc` (rows of 32 bytes); bestIdx[p] = row within group p.
    static void ComputeDistinctiveDescriptors(const std::vector<int32_t>& start, const std::vector<uint8_t>& desc,
                                              std::vector<int32_t>& bestIdx, int device = 0)
    {
        const int np = start.empty() ? 0 : (int)start.size() - 1;
        bestIdx.assign((size_t)np, -1);
        if (orbx_distinctive_descriptors(device, np, start.data(), desc.data(), bestIdx.data(), nullptr) != ORBX_OK)
            throw std::runtime_error(std::string("ComputeDistinctiveDescriptors: ") + orbx_last_error());
    }

    // The local map as Tracking::SearchLocalPoints (src/Tracking.cc:3808-3893) walks it: per map point the world
    // position, mean viewing direction, distance-invariance range, descriptor, and candidate = !isBad() &&
    // mnLastFrameSeen != mCurrentFrame.mnId.
    struct LocalMapSoA {
        std::vector<float> pos, normal;                // n x 3 each: GetWorldPos2(), GetNormal2()
        std::vector<float> minDistance, maxDistance;   // mfMinDistance, mfMaxDistance as stored (= GetMin/MaxDistanceInvariance() / 0.8f, / 1.2f)
        std::vector<uint8_t> desc, candidate;          // n x 32 ; n
    };
    struct TrackFields {                               // what Frame::isInFrustum leaves in each MapPoint
        std::vector<uint8_t> mbTrackInView;
        std::vector<float> mTrackProjX, mTrackProjY, mTrackProjXR, mTrackDepth, mTrackViewCos;
        std::vector<int32_t> mnTrackScaleLevel;
        void resize(size_t n)
        {
            mbTrackInView.resize(n); mTrackProjX.resize(n); mTrackProjY.resize(n); mTrackProjXR.resize(n);
            mTrackDepth.resize(n); mTrackViewCos.resize(n); mnTrackScaleLevel.resize(n);
        }
    };

    // Frame::isInFrustum (src/Frame.cc:603-711) for the whole local map + SearchByProjection(Frame&, vector<MapPoint*>&,
    // th, bFarPoints, thFarPoints) (src/ORBmatcher.cc:45-168) in one device round trip.  Rcw row-major 3x3.
    // assign[i] in: >= 0 when keypoint i already holds a map point with observations; out: row matched.
    int SearchLocalPoints(const FrameView& F, const float* Rcw, const float* tcw, const float* Ow,
                          float fx, float fy, float cx, float cy, float mbf, float mfLogScaleFactor,
                          const LocalMapSoA& map, float viewingCosLimit, float th, bool bFarPoints, float thFarPoints,
                          std::vector<int32_t>& assign, TrackFields& out, int device = 0) const
    {
        const size_t nmp = map.minDistance.size();
        out.resize(nmp);
        assign.resize((size_t)F.N(), -1);
        const int r = orbx_search_local_points(
            device, F.kps(), F.mDescriptors->data, F.N(), F.mvuRight ? F.mvuRight->data() : nullptr,
            F.mnMinX, F.mnMaxX, F.mnMinY, F.mnMaxY, F.mvScaleFactors->data(), (int)F.mvScaleFactors->size(),
            Rcw, tcw, Ow, fx, fy, cx, cy, mbf, mfLogScaleFactor,
            (int)nmp, map.pos.data(), map.normal.data(), map.minDistance.data(), map.maxDistance.data(), map.desc.data(),
            map.candidate.empty() ? nullptr : map.candidate.data(), viewingCosLimit, th, mfNNratio, bFarPoints ? 1 : 0, thFarPoints,
            assign.data(), out.mbTrackInView.data(), out.mTrackProjX.data(), out.mTrackProjY.data(), out.mTrackProjXR.data(),
            out.mTrackDepth.data(), out.mTrackViewCos.data(), out.mnTrackScaleLevel.data());
        if (r < 0) throw std::runtime_error(std::string("SearchLocalPoints: ") + orbx_last_error());
        return r;
    }

    // Rows of candidate map points projected into a keyframe, as the loop-closing / local-mapping searches visit
    // them (Scw or the keyframe pose applied on the host): descriptor, projection, predicted level, validity.
    struct ProjectedPointsSoA {
        std::vector<uint8_t> desc;                     // n x 32: pMP->GetDescriptor()
        std::vector<float> u, v;                       // pKF->mpCamera->project(Rcw * p3Dw + tcw)
        std::vector<float> ur;                         // u - bf * invz (first Fuse form only; empty otherwise)
        std::vector<int32_t> level;                    // pMP->PredictScale(dist3D, pKF)
        std::vector<uint8_t> valid;                    // every gate before KeyFrame::IsInImage
    };

    // int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints,
    // vector<MapPoint*>& vpMatched, int th, float ratioHamming)  src/ORBmatcher.cc:549-678 (and :681-797).
    // matched[idx] = vpMatched[idx] != NULL; assign[idx] = row newly matched (the caller stores vpPoints[row] /
    // vpPointsKFs[row] into vpMatched / vpMatchedKF).
    int SearchByProjection(const FrameView& KF, const ProjectedPointsSoA& pts, int th, float ratioHamming,
                           const std::vector<uint8_t>* matched, std::vector<int32_t>& assign, int device = 0) const
    {
        const int n = KF.N();
        assign.assign((size_t)n, -1);
        const int r = orbx_search_by_projection_sim3(
            device, KF.kps(), KF.mDescriptors->data, n, KF.mnMinX, KF.mnMaxX, KF.mnMinY, KF.mnMaxY,
            KF.mvScaleFactors->data(), (int)KF.mvScaleFactors->size(),
            (int)pts.u.size(), pts.desc.data(), pts.u.data(), pts.v.data(), pts.level.data(), pts.valid.data(),
            (float)th, ratioHamming, matched ? matched->data() : nullptr, assign.data());
        if (r < 0) throw std::runtime_error(std::string("SearchByProjection(KeyFrame,Sim3): ") + orbx_last_error());
        return r;
    }

    // The search of int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th, const bool
    // bRight) src/ORBmatcher.cc:1823-2035 (pts.ur filled, mvInvLevelSigma2 given) and of Fuse(KeyFrame* pKF, cv::Mat Scw,
    // const vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint) :2051-2185 (pts.ur empty).
    // bestIdx[row] = keypoint the map point fuses with or -1; the caller then walks the rows in order and does
    // Replace / AddObservation / AddMapPoint / vpReplacePoint exactly as :2001-2026 / :2165-2180.
    int Fuse(const FrameView& KF, const std::vector<float>* mvInvLevelSigma2, const ProjectedPointsSoA& pts, float th,
             std::vector<int32_t>& bestIdx, std::vector<int32_t>& bestDist, int device = 0) const
    {
        const size_t nmp = pts.u.size();
        bestIdx.assign(nmp, -1); bestDist.assign(nmp, 256);
        const bool chi2 = !pts.ur.empty();
        const int r = orbx_fuse(
            device, KF.kps(), KF.mDescriptors->data, KF.N(), KF.mvuRight ? KF.mvuRight->data() : nullptr,
            KF.mnMinX, KF.mnMaxX, KF.mnMinY, KF.mnMaxY, KF.mvScaleFactors->data(),
            mvInvLevelSigma2 ? mvInvLevelSigma2->data() : nullptr, (int)KF.mvScaleFactors->size(),
            (int)nmp, pts.desc.data(), pts.u.data(), pts.v.data(), chi2 ? pts.ur.data() : nullptr, pts.level.data(),
            pts.valid.data(), th, bestIdx.data(), bestDist.data());
        if (r < 0) throw std::runtime_error(std::string("Fuse: ") + orbx_last_error());
        return r;
    }

    // int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12,
    // const cv::Mat& R12, const cv::Mat& t12, const float th)  src/ORBmatcher.cc:2201-2434.
    // pts1: KF1's map points projected into KF2 (one row per KF1 keypoint), pts2 the reverse; match12[i1] = idx2 of
    // every new cross-checked match (the caller sets vpMatches12[i1] = vpMapPoints2[idx2]).
    int SearchBySim3(const FrameView& KF1, const FrameView& KF2, const ProjectedPointsSoA& pts1, const ProjectedPointsSoA& pts2,
                     float th, std::vector<int32_t>& match12, int device = 0) const
    {
        match12.assign((size_t)KF1.N(), -1);
        const int r = orbx_search_by_sim3(
            device, KF1.kps(), KF1.mDescriptors->data, KF1.N(), KF2.kps(), KF2.mDescriptors->data, KF2.N(),
            KF2.mnMinX, KF2.mnMaxX, KF2.mnMinY, KF2.mnMaxY, KF2.mvScaleFactors->data(), (int)KF2.mvScaleFactors->size(),
            pts1.desc.data(), pts1.u.data(), pts1.v.data(), pts1.level.data(), pts1.valid.data(),
            pts2.desc.data(), pts2.u.data(), pts2.v.data(), pts2.level.data(), pts2.valid.data(), th, match12.data());
        if (r < 0) throw std::runtime_error(std::string("SearchBySim3: ") + orbx_last_error());
        return r;
    }

    // void ORBmatcher::ComputeThreeMaxima(vector<int>* histo, const int L, int&, int&, int&)
    // include/ORBmatcher.h:107, src/ORBmatcher.cc:2863-2905
    static void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3)
    {
        std::vector<int32_t> sz((size_t)L);
        for (int i = 0; i < L; i++) sz[(size_t)i] = (int32_t)histo[i].size();
        int32_t a, b, c;
        orbx_three_maxima(sz.data(), L, &a, &b, &c);
        ind1 = a; ind2 = b; ind3 = c;
    }
};

}  // namespace ORB_SLAM3
