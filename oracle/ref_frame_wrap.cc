/*
 * oracle/ref_frame_wrap.cc — TEST INFRASTRUCTURE ONLY.
 *
 * C entry points around the UNMODIFIED reference /root/reference/src/Frame.cc (compiled where it lies by oracle/build_ref.sh over
 * oracle/cvshim, oracle/refshim_frame/ref_frame_stubs.h and the real include/Frame.h, with the real ORBextractor.cc and the
 * vendored DBoW2 linked in) -> oracle/_ref/libref_frame.so.  Each function fills a real ORB_SLAM3::Frame with the inputs of the
 * restatement it pins and calls the reference's own member function:
 *   ref_frame_stereo_matches      Frame::ComputeStereoMatches        src/Frame.cc:1027-1276   pins orc_stereo_match
 *   ref_frame_grid_*              AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea  :446-480, :784-899   pins orc_grid_*
 *   ref_frame_is_in_frustum       Frame::isInFrustum (Nleft == -1)   :603-711                 pins orc_is_in_frustum
 *   ref_frame_stereo_fisheye      Frame::ComputeStereoFishEyeMatches :1439-1480               pins orc_stereo_fisheye_match
 *   ref_frame_stereo_ctor         Frame::Frame(imLeft, imRight, ...) :103-225  the whole rectified-stereo front end
 */
#define private public
#define protected public
#include "Frame.h"
#undef private
#undef protected
#include "ORBextractor.h"

#include <cstring>
#include <memory>
#include <new>
#include <type_traits>

extern "C" {
#include "orc.h"
}

using namespace ORB_SLAM3;

/* MapPoint::PredictScale(const float&, Frame*), src/MapPoint.cc:725-741 (MapPoint.cc itself is not compiled: it drags in Map,
 * KeyFrame and the optimiser types).  `log` on a float with namespace std imported = std::log(float) = logf. */
int MapPoint::PredictScale(const float& currentDist, Frame* pF)
{
    float ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
}

namespace {

std::vector<cv::KeyPoint> keys(const orc_keypoint* k, int n)
{
    std::vector<cv::KeyPoint> v((size_t)n);
    if (n) std::memcpy(v.data(), k, sizeof(orc_keypoint) * (size_t)n);
    return v;
}

cv::Mat desc_matrix(const uint8_t* d, int n)
{
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) std::memcpy(m.data, d, (size_t)n * 32);
    return n > 0 ? m : m.rowRange(0, 0);
}

void set_levels(Frame& F, const float* sf, int nlevels)
{
    F.mnScaleLevels = nlevels;
    F.mfScaleFactor = nlevels > 1 ? sf[1] : 1.2f;
    F.mfLogScaleFactor = log(F.mfScaleFactor);
    F.mvScaleFactors.assign(sf, sf + nlevels);
    F.mvInvScaleFactors.resize((size_t)nlevels); F.mvLevelSigma2.resize((size_t)nlevels); F.mvInvLevelSigma2.resize((size_t)nlevels);
    for (int i = 0; i < nlevels; i++) {
        F.mvInvScaleFactors[i] = 1.0f / sf[i]; F.mvLevelSigma2[i] = sf[i] * sf[i]; F.mvInvLevelSigma2[i] = 1.0f / F.mvLevelSigma2[i];
    }
}

/* an ORBextractor whose mvImagePyramid views the caller's padded level planes (plane l: (rows[l] + 38) x (cols[l] + 38), the
 * level image starts at (19, 19) exactly as ComputePyramid leaves it, src/ORBextractor.cc:1672-1676) */
std::unique_ptr<ORBextractor> extractor_over(const uint8_t* const* planes, const int* rows, const int* cols, const size_t* steps, int nlevels)
{
    std::unique_ptr<ORBextractor> e(new ORBextractor(100, 1.2f, nlevels, 20, 7));
    e->mvImagePyramid.resize((size_t)nlevels);
    for (int l = 0; l < nlevels; l++) {
        cv::Mat padded(rows[l] + 38, cols[l] + 38, CV_8U, const_cast<uint8_t*>(planes[l]), steps[l]);
        e->mvImagePyramid[(size_t)l] = padded(cv::Rect(19, 19, cols[l], rows[l]));
    }
    return e;
}

} // namespace

extern "C" {

/* Frame::ComputeStereoMatches on caller-provided keypoints, descriptors and pyramids */
int ref_frame_stereo_matches(const uint8_t* const* planesL, const uint8_t* const* planesR, const int* rows, const int* cols,
                             const size_t* stepsL, const size_t* stepsR, int nlevels, const float* sf,
                             const orc_keypoint* kpL, int nL, const uint8_t* descL, const orc_keypoint* kpR, int nR, const uint8_t* descR,
                             float mb, float mbf, float* uRight, float* depth)
{
    Frame F;
    std::unique_ptr<ORBextractor> eL = extractor_over(planesL, rows, cols, stepsL, nlevels), eR = extractor_over(planesR, rows, cols, stepsR, nlevels);
    F.mpORBextractorLeft = eL.get(); F.mpORBextractorRight = eR.get();
    F.N = nL; F.mvKeys = keys(kpL, nL); F.mvKeysRight = keys(kpR, nR);
    F.mDescriptors = desc_matrix(descL, nL); F.mDescriptorsRight = desc_matrix(descR, nR);
    set_levels(F, sf, nlevels);
    F.mb = mb; F.mbf = mbf;
    F.ComputeStereoMatches();
    for (int i = 0; i < nL; i++) { uRight[i] = F.mvuRight[(size_t)i]; depth[i] = F.mvDepth[(size_t)i]; }
    return 0;
}

/* the frame grid: a Frame whose mvKeysUn are the caller's keypoints, bounds set as ComputeImageBounds would, grid filled by the
 * reference's AssignFeaturesToGrid; queries through the reference's GetFeaturesInArea */
void* ref_frame_grid_build(const orc_keypoint* kps, int n, float minX, float maxX, float minY, float maxY)
{
    Frame* F = new Frame();
    F->N = n; F->Nleft = -1; F->Nright = -1;
    F->mvKeys = keys(kps, n); F->mvKeysUn = F->mvKeys;
    Frame::mnMinX = minX; Frame::mnMaxX = maxX; Frame::mnMinY = minY; Frame::mnMaxY = maxY;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(maxX - minX);      /* src/Frame.cc:193-194 */
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(maxY - minY);
    F->AssignFeaturesToGrid();
    return F;
}
void ref_frame_grid_destroy(void* f) { delete static_cast<Frame*>(f); }
int ref_frame_grid_query(void* f, float x, float y, float r, int minLevel, int maxLevel, int32_t* out, int cap)
{
    const std::vector<size_t> v = static_cast<Frame*>(f)->GetFeaturesInArea(x, y, r, minLevel, maxLevel, false);
    int n = 0;
    for (size_t i = 0; i < v.size() && n < cap; i++) out[n++] = (int32_t)v[i];
    return (int)v.size();
}

/* Frame::isInFrustum over a list of map points; arguments as orc_is_in_frustum (Ow is recomputed by the reference's own
 * UpdatePoseMatrices from Rcw / tcw and must agree with the caller's) */
int ref_frame_is_in_frustum(const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx, float cy, float mbf,
                            float minX, float maxX, float minY, float maxY, float logScaleFactor, int nlevels,
                            int nmp, const float* pos, const float* normal, const float* min_dist, const float* max_dist,
                            const uint8_t* candidate, float viewingCosLimit,
                            uint8_t* inview, float* projx, float* projy, float* projxr, float* depth, float* viewcos, int32_t* level)
{
    Frame F;
    Pinhole cam(fx, fy, cx, cy);
    F.mpCamera = &cam; F.mpCamera2 = nullptr; F.Nleft = -1; F.Nright = -1;
    F.mbf = mbf; F.mnScaleLevels = nlevels; F.mfLogScaleFactor = logScaleFactor;
    Frame::mnMinX = minX; Frame::mnMaxX = maxX; Frame::mnMinY = minY; Frame::mnMaxY = maxY;
    cv::Mat Tcw = cv::Mat::eye(4, 4, CV_32F);
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) Tcw.at<float>(i, j) = Rcw[3 * i + j];
        Tcw.at<float>(i, 3) = tcw[i];
    }
    F.SetPose(Tcw);
    int bad_ow = 0;
    for (int i = 0; i < 3; i++) if (F.mOw.at<float>(i) != Ow[i]) bad_ow = 1;
    int nin = 0;
    for (int i = 0; i < nmp; i++) {
        inview[i] = 0; projx[i] = projy[i] = projxr[i] = depth[i] = viewcos[i] = 0.f; level[i] = 0;
        if (!candidate[i]) continue;
        MapPoint P;
        P.mWorldPosx = cv::Matx31f(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        P.mNormalVectorx = cv::Matx31f(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
        P.mfMinDistance = min_dist[i]; P.mfMaxDistance = max_dist[i];
        P.mTrackProjXR = 0; P.mTrackDepth = 0; P.mTrackViewCos = 0; P.mnTrackScaleLevel = 0;
        const bool in = F.isInFrustum(&P, viewingCosLimit);
        inview[i] = in ? 1 : 0; nin += in ? 1 : 0;
        projx[i] = P.mTrackProjX; projy[i] = P.mTrackProjY; projxr[i] = P.mTrackProjXR; depth[i] = P.mTrackDepth;
        viewcos[i] = P.mTrackViewCos; level[i] = P.mnTrackScaleLevel;
    }
    return bad_ow ? -1000 - nin : nin;
}

/* Frame::ComputeStereoFishEyeMatches with a camera double whose TriangulateMatches reports depth 1 for every pair (so the
 * depth gate :1471 keeps all of them, as orc_stereo_fisheye_match's "candidate" outputs do); arguments as that function, dist_out
 * is not observable (-1). */
int ref_frame_stereo_fisheye(const uint8_t* descL, int nL, int monoLeft, const uint8_t* descR, int nR, int monoRight,
                             int32_t* l2r, int32_t* r2l, int32_t* dist_out)
{
    Frame F;
    KannalaBrandt8 camL(1, 1, 0, 0), camR(1, 1, 0, 0);
    F.mpCamera = &camL; F.mpCamera2 = &camR;
    F.Nleft = nL; F.Nright = nR; F.N = nL + nR; F.monoLeft = monoLeft; F.monoRight = monoRight;
    F.mvKeys.assign((size_t)nL, cv::KeyPoint()); F.mvKeysRight.assign((size_t)nR, cv::KeyPoint());
    for (auto& k : F.mvKeys) k.octave = 0;
    for (auto& k : F.mvKeysRight) k.octave = 0;
    F.mDescriptors = desc_matrix(descL, nL); F.mDescriptorsRight = desc_matrix(descR, nR);
    const float sf[8] = {1.f, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985984f, 3.5831808f};
    set_levels(F, sf, 8);
    F.mRlr = cv::Mat::eye(3, 3, CV_32F); F.mtlr = cv::Mat::zeros(3, 1, CV_32F);
    F.ComputeStereoFishEyeMatches();
    int good = 0;
    for (int i = 0; i < nL; i++) { l2r[i] = F.mvLeftToRightMatch[(size_t)i]; dist_out[i] = -1; good += l2r[i] >= 0; }
    for (int i = 0; i < nR; i++) r2l[i] = F.mvRightToLeftMatch[(size_t)i];
    return good;
}

/* The rectified-stereo constructor, src/Frame.cc:103-225: two ExtractORB threads, UndistortKeyPoints (no distortion),
 * ComputeStereoMatches, AssignFeaturesToGrid.  Outputs: N, keypoints / descriptors of both eyes (caller's arrays sized cap),
 * mvuRight, mvDepth.  The quadtree tie-break of the two extractions is the reference's literal (heap-address) one here.
 * NOTE a reference quirk: this constructor calls ComputeStereoMatches() (:137) BEFORE it assigns mb = mbf / fx (:163), so the
 * disparity bound maxD = mbf / mb (:1084) is computed from an UNINITIALISED member.  The object is built in zero-filled storage
 * here, i.e. mb = 0 and maxD = +inf (no upper disparity bound) - which is also what a fresh stack page gives the reference. */
int ref_frame_stereo_ctor(const uint8_t* imL, const uint8_t* imR, int rows, int cols, size_t step,
                          int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                          float fx, float fy, float cx, float cy, float bf, float thDepth,
                          orc_keypoint* kpL, uint8_t* descL, int* nL, orc_keypoint* kpR, uint8_t* descR, int* nR, int cap,
                          float* uRight, float* depth)
{
    ORBextractor eL(nfeatures, scaleFactor, nlevels, iniTh, minTh), eR(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    cv::Mat L(rows, cols, CV_8U, const_cast<uint8_t*>(imL), step), R(rows, cols, CV_8U, const_cast<uint8_t*>(imR), step);
    Pinhole cam(fx, fy, cx, cy);
    cv::Mat K = cam.toK(), dist = cv::Mat::zeros(4, 1, CV_32F);
    Frame::mbInitialComputations = true;
    static std::aligned_storage<sizeof(Frame), alignof(Frame)>::type storage;
    std::memset(&storage, 0, sizeof(storage));
    Frame* pF = new (&storage) Frame(L, R, 0.0, &eL, &eR, nullptr, K, dist, bf, thDepth, &cam);
    struct Guard { Frame* p; ~Guard() { p->~Frame(); } } guard{pF};
    Frame& F = *pF;
    *nL = F.N; *nR = (int)F.mvKeysRight.size();
    if (F.N > cap || *nR > cap) return -2;
    for (int i = 0; i < F.N; i++) {
        std::memcpy(&kpL[i], &F.mvKeys[(size_t)i], sizeof(orc_keypoint));
        std::memcpy(descL + 32 * (size_t)i, F.mDescriptors.ptr(i), 32);
        uRight[i] = F.mvuRight[(size_t)i]; depth[i] = F.mvDepth[(size_t)i];
    }
    for (int i = 0; i < *nR; i++) {
        std::memcpy(&kpR[i], &F.mvKeysRight[(size_t)i], sizeof(orc_keypoint));
        std::memcpy(descR + 32 * (size_t)i, F.mDescriptorsRight.ptr(i), 32);
    }
    return 0;
}

} // extern "C"
