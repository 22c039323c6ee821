/*
 * oracle/ref_match_wrap.cc — TEST INFRASTRUCTURE ONLY.
 *
 * C entry points around the UNMODIFIED reference matcher: /root/reference/src/ORBmatcher.cc is compiled where it lies
 * (oracle/build_ref.sh) against oracle/cvshim (OpenCV stand-in) and oracle/refshim/ref_stubs.h (data-holder doubles of
 * Frame / KeyFrame / MapPoint / GeometricCamera) and linked with this file into oracle/_ref/libref_match.so.
 * Every ref_* function takes the SAME arguments as the orc_* restatement of oracle/orc_match.c it pins, builds the object
 * graph the reference expects from those SoA arrays, calls the reference's own member function and flattens the result
 * back - so tests/test_oracle_ref_match.py can require orc_* == ref_* on identical inputs.
 *
 * Geometry convention (see ref_stubs.h): poses are the identity and the camera double projects (x, y, z) -> (x, y), so
 * a projection the test wants at pixel (u, v) with inverse depth invz is handed in as the world point (u, v, 1 / invz);
 * the reference's own pose arithmetic then runs on it and reproduces (u, v, invz).
 */
#define protected public          /* ComputeThreeMaxima is a protected member (include/ORBmatcher.h:100-107) */
#include "ORBmatcher.h"
#undef protected

#include <memory>

using namespace ORB_SLAM3;

namespace {

cv::Mat desc_row(const uint8_t* d)
{
    cv::Mat m(1, 32, CV_8U);
    std::memcpy(m.data, d, 32);
    return m;
}

cv::Mat desc_matrix(const uint8_t* d, int n)
{
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) std::memcpy(m.data, d, (size_t)n * 32);
    return m;
}

std::vector<cv::KeyPoint> keys(const orc_keypoint* k, int n)
{
    std::vector<cv::KeyPoint> v((size_t)n);
    if (n) std::memcpy(v.data(), k, sizeof(orc_keypoint) * (size_t)n);
    return v;
}

cv::Mat point3(float x, float y, float z)
{
    cv::Mat m(3, 1, CV_32F);
    m.at<float>(0) = x; m.at<float>(1) = y; m.at<float>(2) = z;
    return m;
}

void set_levels(RefViewBase& F, const float* sf, int nlevels = 8)
{
    F.mvScaleFactors.assign(sf, sf + nlevels);
    F.mvLevelSigma2.resize((size_t)nlevels); F.mvInvScaleFactors.resize((size_t)nlevels); F.mvInvLevelSigma2.resize((size_t)nlevels);
    for (int i = 0; i < nlevels; i++) {
        F.mvLevelSigma2[i] = sf[i] * sf[i]; F.mvInvScaleFactors[i] = 1.0f / sf[i]; F.mvInvLevelSigma2[i] = 1.0f / F.mvLevelSigma2[i];
    }
}

void set_bounds(RefViewBase& F, const orc_grid* g)
{
    float b[4];
    orc_grid_bounds(g, b);
    F.mnMinX = b[0]; F.mnMaxX = b[1]; F.mnMinY = b[2]; F.mnMaxY = b[3];
}

void set_featvec(RefViewBase& F, const int32_t* node, const int32_t* start, const int32_t* idx, int nn)
{
    for (int a = 0; a < nn; a++)
        for (int p = start[a]; p < start[a + 1]; p++) F.mFeatVec.addFeature((DBoW2::NodeId)node[a], (unsigned)idx[p]);
}

/* keeps the MapPoint doubles of one call alive */
struct Pool {
    std::vector<std::unique_ptr<MapPoint> > all;
    MapPoint* make(int index, const uint8_t* desc, int nobs = 1)
    {
        all.emplace_back(new MapPoint());
        MapPoint* p = all.back().get();
        p->index = index; p->nObs = nobs;
        if (desc) p->mDescriptor = desc_row(desc);
        return p;
    }
};

} // namespace

extern "C" {

/* ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2911-2931 */
int ref_hamming256(const void* a, const void* b)
{
    return ORBmatcher::DescriptorDistance(desc_row((const uint8_t*)a), desc_row((const uint8_t*)b));
}

/* ORBmatcher::ComputeThreeMaxima, :2863-2905 */
void ref_three_maxima(const int* sizes, int L, int* ind1, int* ind2, int* ind3)
{
    std::vector<std::vector<int> > histo((size_t)L);
    for (int i = 0; i < L; i++) histo[(size_t)i].assign((size_t)sizes[i], 0);
    ORBmatcher m;
    m.ComputeThreeMaxima(histo.data(), L, *ind1, *ind2, *ind3);
}

/* SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints), :45-239, F.Nleft == -1; args as orc_search_by_projection */
int ref_search_by_projection(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n,
                             const float* uRight, const float* scaleFactors,
                             int nmp, const uint8_t* mp_desc, const float* mp_projx, const float* mp_projy,
                             const float* mp_projxr, const int32_t* mp_level, const float* mp_viewcos,
                             const float* mp_depth, const uint8_t* mp_inview,
                             float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    Frame F; Pool pool;
    F.N = n; F.Nleft = -1; F.mvKeysUn = keys(kps, n); F.mvKeys = F.mvKeysUn; F.mDescriptors = desc_matrix(desc, n);
    F.mvuRight.assign((size_t)n, -1.f);
    if (uRight) F.mvuRight.assign(uRight, uRight + n);
    set_levels(F, scaleFactors); set_bounds(F, g); F.grid = const_cast<orc_grid*>(g);
    F.mvpMapPoints.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (assign[i] >= 0) F.mvpMapPoints[i] = pool.make(-2 - assign[i], nullptr, 1);   // already held, with observations
    std::vector<MapPoint*> mps((size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        MapPoint* p = pool.make(i, mp_desc + 32 * (size_t)i);
        p->mbTrackInView = mp_inview[i] != 0; p->mbTrackInViewR = false;
        p->mTrackProjX = mp_projx[i]; p->mTrackProjY = mp_projy[i]; p->mTrackProjXR = mp_projxr ? mp_projxr[i] : 0.f;
        p->mnTrackScaleLevel = mp_level[i]; p->mTrackViewCos = mp_viewcos[i]; p->mTrackDepth = mp_depth[i];
        mps[(size_t)i] = p;
    }
    ORBmatcher m(nnratio, true);
    const int r = m.SearchByProjection(F, mps, th, bFarPoints != 0, thFarPoints);
    for (int i = 0; i < n; i++) {
        MapPoint* p = F.mvpMapPoints[i];
        assign[i] = !p ? -1 : (p->index >= 0 ? p->index : -2 - p->index);
    }
    return r;
}

/* the same with F.Nleft != -1 (:170-236 added); args as orc_search_by_projection_rig */
int ref_search_by_projection_rig(const orc_grid* gL, const orc_keypoint* kpsL, int nL,
                                 const orc_grid* gR, const orc_keypoint* kpsR, int nR,
                                 const uint8_t* desc, const int32_t* l2r, const int32_t* r2l, const float* sf,
                                 int nmp, const uint8_t* mp_desc, const float* mp_depth,
                                 const uint8_t* inviewL, const float* projxL, const float* projyL, const int32_t* levelL, const float* viewcosL,
                                 const uint8_t* inviewR, const float* projxR, const float* projyR, const int32_t* levelR, const float* viewcosR,
                                 float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    Frame F; Pool pool;
    const int n = nL + nR;
    F.N = n; F.Nleft = nL; F.Nright = nR; F.mvKeys = keys(kpsL, nL); F.mvKeysRight = keys(kpsR, nR); F.mDescriptors = desc_matrix(desc, n);
    F.mvuRight.assign((size_t)n, -1.f);
    F.mvLeftToRightMatch.assign(l2r, l2r + nL); F.mvRightToLeftMatch.assign(r2l, r2l + nR);
    set_levels(F, sf); set_bounds(F, gL); F.grid = const_cast<orc_grid*>(gL); F.gridR = const_cast<orc_grid*>(gR);
    F.mvpMapPoints.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (assign[i] >= 0) F.mvpMapPoints[i] = pool.make(-2 - assign[i], nullptr, 1);
    std::vector<MapPoint*> mps((size_t)nmp);
    for (int i = 0; i < nmp; i++) {
        MapPoint* p = pool.make(i, mp_desc + 32 * (size_t)i);
        p->mbTrackInView = inviewL[i] != 0; p->mbTrackInViewR = inviewR[i] != 0;
        p->mTrackProjX = projxL[i]; p->mTrackProjY = projyL[i]; p->mnTrackScaleLevel = levelL[i]; p->mTrackViewCos = viewcosL[i];
        p->mTrackProjXR = projxR[i]; p->mTrackProjYR = projyR[i]; p->mnTrackScaleLevelR = levelR[i]; p->mTrackViewCosR = viewcosR[i];
        p->mTrackDepth = mp_depth[i];
        mps[(size_t)i] = p;
    }
    ORBmatcher m(nnratio, true);
    const int r = m.SearchByProjection(F, mps, th, bFarPoints != 0, thFarPoints);
    for (int i = 0; i < n; i++) {
        MapPoint* p = F.mvpMapPoints[i];
        assign[i] = !p ? -1 : (p->index >= 0 ? p->index : -2 - p->index);
    }
    return r;
}

/* fills a Frame double for the frame-to-frame searches: identity pose; bForward / bBackward come out of the LAST frame's
 * translation along z (:2491-2496: tlc = Rlw * twc + tlw with twc = 0) */
static void frame_pair_poses(Frame& Cur, Frame& Last, int bForward, int bBackward, float mb)
{
    Cur.mTcw = cv::Mat::eye(4, 4, CV_32F); Last.mTcw = cv::Mat::eye(4, 4, CV_32F);
    Cur.mb = mb;
    Last.mTcw.at<float>(2, 3) = bForward ? 2.f * mb + 1.f : (bBackward ? -(2.f * mb + 1.f) : 0.f);
}

/* SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), :2469-2711, Nleft == -1; args as orc_search_by_projection_frame */
int ref_search_by_projection_frame(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n,
                                   const float* uRight, const float* sf,
                                   int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                   const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                   const uint8_t* last_valid, const uint8_t* last_obs,
                                   float th, float mbf, int bForward, int bBackward, int checkOrientation,
                                   const uint8_t* cur_held, int32_t* assign)
{
    Frame Cur, Last; Pool pool; GeometricCamera cam;
    Cur.N = n; Cur.Nleft = -1; Cur.mvKeysUn = keys(kps, n); Cur.mvKeys = Cur.mvKeysUn; Cur.mDescriptors = desc_matrix(desc, n);
    Cur.mvuRight.assign((size_t)n, -1.f);
    if (uRight) Cur.mvuRight.assign(uRight, uRight + n);
    Cur.mbf = mbf; Cur.mpCamera = &cam;
    set_levels(Cur, sf); set_bounds(Cur, g); Cur.grid = const_cast<orc_grid*>(g);
    Cur.mvpMapPoints.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (cur_held && cur_held[i]) Cur.mvpMapPoints[i] = pool.make(-2, nullptr, 1);
    frame_pair_poses(Cur, Last, bForward, bBackward, 0.5f);
    Last.N = nlast; Last.Nleft = -1;
    Last.mvKeys.resize((size_t)nlast); Last.mvKeysUn.resize((size_t)nlast);
    Last.mvpMapPoints.assign((size_t)nlast, nullptr); Last.mvbOutlier.assign((size_t)nlast, false);
    for (int i = 0; i < nlast; i++) {
        Last.mvKeys[i].octave = last_octave[i]; Last.mvKeys[i].angle = last_angle[i];
        Last.mvKeysUn[i] = Last.mvKeys[i];
        if (!last_valid[i]) continue;                                   /* pMP == NULL (or an outlier, :2503-2506) */
        MapPoint* p = pool.make(i, last_desc + 32 * (size_t)i, last_obs[i] ? 1 : 0);
        p->mWorldPos = point3(last_u[i], last_v[i], 1.0f / last_invz[i]);
        Last.mvpMapPoints[i] = p;
    }
    ORBmatcher m(0.9f, checkOrientation != 0);
    const int r = m.SearchByProjection(Cur, Last, th, /*bMono=*/!(bForward || bBackward) && uRight == nullptr);
    for (int i = 0; i < n; i++) {
        MapPoint* p = Cur.mvpMapPoints[i];
        assign[i] = (p && p->index >= 0) ? p->index : -1;
    }
    return r;
}

/* the same with CurrentFrame.Nleft != -1 (:2610-2684); args as orc_search_by_projection_frame_rig.  The reference projects
 * Trl.R * x3Dc + Trl.t into the right camera: with the identity rotation that is one constant pixel offset, so the rows must
 * satisfy (last_ur, last_vr) == (last_u + dx, last_v + dy) in float arithmetic for one (dx, dy); -1000 otherwise. */
int ref_search_by_projection_frame_rig(const orc_grid* gL, const orc_keypoint* kpsL, int nL,
                                       const orc_grid* gR, const orc_keypoint* kpsR, int nR,
                                       const uint8_t* desc, const float* sf,
                                       int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                       const float* last_ur, const float* last_vr,
                                       const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                       const uint8_t* last_valid, const uint8_t* last_obs,
                                       float th, int bForward, int bBackward, int checkOrientation,
                                       const uint8_t* cur_held, int32_t* assign)
{
    const int n = nL + nR;
    float dx = 0.f, dy = 0.f;
    if (nlast > 0) { dx = last_ur[0] - last_u[0]; dy = last_vr[0] - last_v[0]; }
    for (int i = 0; i < nlast; i++)
        if (last_u[i] + dx != last_ur[i] || last_v[i] + dy != last_vr[i]) return -1000;
    Frame Cur, Last; Pool pool; GeometricCamera cam;
    Cur.N = n; Cur.Nleft = nL; Cur.Nright = nR; Cur.mvKeys = keys(kpsL, nL); Cur.mvKeysRight = keys(kpsR, nR);
    Cur.mDescriptors = desc_matrix(desc, n); Cur.mvuRight.assign((size_t)n, -1.f); Cur.mpCamera = &cam;
    set_levels(Cur, sf); set_bounds(Cur, gL); Cur.grid = const_cast<orc_grid*>(gL); Cur.gridR = const_cast<orc_grid*>(gR);
    Cur.mvpMapPoints.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (cur_held && cur_held[i]) Cur.mvpMapPoints[i] = pool.make(-2, nullptr, 1);
    frame_pair_poses(Cur, Last, bForward, bBackward, 0.5f);
    Cur.mTrl = cv::Mat::eye(3, 4, CV_32F);
    Cur.mTrl.at<float>(0, 3) = dx; Cur.mTrl.at<float>(1, 3) = dy;
    Last.N = nlast; Last.Nleft = -1;
    Last.mvKeys.resize((size_t)nlast); Last.mvKeysUn.resize((size_t)nlast);
    Last.mvpMapPoints.assign((size_t)nlast, nullptr); Last.mvbOutlier.assign((size_t)nlast, false);
    for (int i = 0; i < nlast; i++) {
        Last.mvKeys[i].octave = last_octave[i]; Last.mvKeys[i].angle = last_angle[i];
        Last.mvKeysUn[i] = Last.mvKeys[i];
        if (!last_valid[i]) continue;
        MapPoint* p = pool.make(i, last_desc + 32 * (size_t)i, last_obs[i] ? 1 : 0);
        p->mWorldPos = point3(last_u[i], last_v[i], 1.0f / last_invz[i]);
        Last.mvpMapPoints[i] = p;
    }
    ORBmatcher m(0.9f, checkOrientation != 0);
    const int r = m.SearchByProjection(Cur, Last, th, /*bMono=*/!(bForward || bBackward));
    for (int i = 0; i < n; i++) {
        MapPoint* p = Cur.mvpMapPoints[i];
        assign[i] = (p && p->index >= 0) ? p->index : -1;
    }
    return r;
}

/* a KeyFrame double from SoA arrays; nleft < 0: mono / rectified stereo (mvKeysUn), else rig (mvKeys | mvKeysRight) */
static void fill_keyframe(KeyFrame& K, Pool& pool, const orc_keypoint* kps, const uint8_t* desc, int n, int nleft, const uint8_t* has_mp,
                          const float* uRight, const int32_t* fn, const int32_t* fs, const int32_t* fi, int nn, GeometricCamera* cam, GeometricCamera* cam2)
{
    K.N = n; K.Tcw = cv::Mat::eye(4, 4, CV_32F); K.Trl = cv::Mat::eye(4, 4, CV_32F);
    K.mDescriptors = desc_matrix(desc, n);
    if (nleft < 0) { K.NLeft = K.Nleft = -1; K.mvKeysUn = keys(kps, n); K.mvKeys = K.mvKeysUn; }
    else { K.NLeft = K.Nleft = nleft; K.NRight = K.Nright = n - nleft; K.mvKeys = keys(kps, nleft); K.mvKeysRight = keys(kps + nleft, n - nleft); }
    K.mvuRight.assign((size_t)n, -1.f);
    if (uRight) K.mvuRight.assign(uRight, uRight + n);
    K.mvpMapPoints.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (has_mp && has_mp[i]) K.mvpMapPoints[i] = pool.make(i, desc + 32 * (size_t)i);
    set_featvec(K, fn, fs, fi, nn);
    K.mpCamera = cam; K.mpCamera2 = cam2;
}

/* SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) :314-547 (mode 0) and SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)
 * :955-1105 (mode 1); args as orc_search_by_bow */
int ref_search_by_bow(int mode,
                      const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                      const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                      const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* valid2,
                      const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                      float nnratio, int checkOrientation, int32_t* matches)
{
    Pool pool; KeyFrame K1;
    fill_keyframe(K1, pool, kps1, desc1, n1, -1, valid1, nullptr, fv1_node, fv1_start, fv1_idx, nn1, nullptr, nullptr);
    ORBmatcher m(nnratio, checkOrientation != 0);
    std::vector<MapPoint*> out;
    int r;
    if (mode == 0) {
        Frame F;
        F.N = n2; F.Nleft = -1; F.mvKeys = keys(kps2, n2); F.mvKeysUn = F.mvKeys; F.mDescriptors = desc_matrix(desc2, n2);
        set_featvec(F, fv2_node, fv2_start, fv2_idx, nn2);
        r = m.SearchByBoW(&K1, F, out);
        for (int i = 0; i < n2; i++) matches[i] = out[(size_t)i] ? out[(size_t)i]->index : -1;
    } else {
        KeyFrame K2;
        fill_keyframe(K2, pool, kps2, desc2, n2, -1, valid2, nullptr, fv2_node, fv2_start, fv2_idx, nn2, nullptr, nullptr);
        r = m.SearchByBoW(&K1, &K2, out);
        for (int i = 0; i < n1; i++) matches[i] = out[(size_t)i] ? out[(size_t)i]->index : -1;     /* vpMatches12[idx1] = KF2's map point */
    }
    return r;
}

/* SearchByBoW(KeyFrame*, Frame&, ...) with F.Nleft != -1; args as orc_search_by_bow_rig.  The keyframe is a rig too
 * (pKF->mpCamera2 != NULL, NLeft = n1: all its keypoints count as left) so that :437-448 reads mvKeys / mvKeysRight. */
int ref_search_by_bow_rig(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                          const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                          const orc_keypoint* kps2, const uint8_t* desc2, int n2, int nleft2,
                          const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                          float nnratio, int checkOrientation, int32_t* matches)
{
    Pool pool; KeyFrame K1; GeometricCamera cam, cam2;
    fill_keyframe(K1, pool, kps1, desc1, n1, n1, valid1, nullptr, fv1_node, fv1_start, fv1_idx, nn1, &cam, &cam2);
    Frame F;
    F.N = n2; F.Nleft = nleft2; F.Nright = n2 - nleft2; F.mvKeys = keys(kps2, nleft2); F.mvKeysRight = keys(kps2 + nleft2, n2 - nleft2);
    F.mDescriptors = desc_matrix(desc2, n2); F.mpCamera = &cam; F.mpCamera2 = &cam2;
    set_featvec(F, fv2_node, fv2_start, fv2_idx, nn2);
    ORBmatcher m(nnratio, checkOrientation != 0);
    std::vector<MapPoint*> out;
    const int r = m.SearchByBoW(&K1, F, out);
    for (int i = 0; i < n2; i++) matches[i] = out[(size_t)i] ? out[(size_t)i]->index : -1;
    return r;
}

/* SearchForInitialization(Frame& F1, Frame& F2, vbPrevMatched, vnMatches12, windowSize), :799-943; args as orc_search_for_initialization */
int ref_search_for_initialization(const orc_keypoint* kps1, const uint8_t* desc1, int n1,
                                  const orc_grid* g2, const orc_keypoint* kps2, const uint8_t* desc2, int n2,
                                  float* prev_matched, int windowSize, float nnratio, int checkOrientation, int32_t* matches12)
{
    Frame F1, F2;
    F1.N = n1; F1.Nleft = -1; F1.mvKeysUn = keys(kps1, n1); F1.mvKeys = F1.mvKeysUn; F1.mDescriptors = desc_matrix(desc1, n1);
    F2.N = n2; F2.Nleft = -1; F2.mvKeysUn = keys(kps2, n2); F2.mvKeys = F2.mvKeysUn; F2.mDescriptors = desc_matrix(desc2, n2);
    set_bounds(F2, g2); F2.grid = const_cast<orc_grid*>(g2);
    std::vector<cv::Point2f> prev((size_t)n1);
    for (int i = 0; i < n1; i++) prev[(size_t)i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher m(nnratio, checkOrientation != 0);
    const int r = m.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < n1; i++) { matches12[i] = m12[(size_t)i]; prev_matched[2 * i] = prev[(size_t)i].x; prev_matched[2 * i + 1] = prev[(size_t)i].y; }
    return r;
}

/* verdict used for the PINHOLE triangulation searches: Pinhole::epipolarConstrain, src/CameraModels/Pinhole.cpp:237-260, with
 * the F12 the test hands in (the camera double has no intrinsics of its own) */
struct PinholeVerdict { const orc_keypoint *k1, *k2; const float* F; const float* sigma2_2; };
static int pinhole_verdict(void* u, int i1, int i2)
{
    const PinholeVerdict* v = (const PinholeVerdict*)u;
    const orc_keypoint& kp1 = v->k1[i1]; const orc_keypoint& kp2 = v->k2[i2];
    const float* F12 = v->F;
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return 0;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * v->sigma2_2[kp2.octave];
}

static int run_triangulation(int which, KeyFrame& K1, KeyFrame& K2, int n1, int bOnlyStereo, int bCoarse, int checkOrientation, int32_t* matches12)
{
    ORBmatcher m(0.6f, checkOrientation != 0);
    std::vector<std::pair<size_t, size_t> > pairs;
    int r;
    if (which == 0) r = m.SearchForTriangulation(&K1, &K2, cv::Mat::eye(3, 3, CV_32F), pairs, bOnlyStereo != 0, bCoarse != 0);
    else if (which == 1) r = m.SearchForTriangulation_(&K1, &K2, cv::Matx33f(), pairs, bOnlyStereo != 0, bCoarse != 0);
    else { std::vector<cv::Mat> pts; r = m.SearchForTriangulation(&K1, &K2, cv::Mat::eye(3, 3, CV_32F), pairs, bOnlyStereo != 0, pts); }
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    for (const auto& p : pairs) matches12[p.first] = (int32_t)p.second;
    return r;
}

/* SearchForTriangulation (:1107-1386, variant 0) / SearchForTriangulation_ (:1388-1630, variant 1), pinhole / Nleft == -1;
 * args as orc_search_for_triangulation.  KF2's translation puts the epipole where the test wants it (:1115-1120). */
int ref_search_for_triangulation(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1, const float* uRight1,
                                 const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                 const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2, const float* uRight2,
                                 const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                 const float* sf2, const float* sigma2_2, const float* ep, const float* F12,
                                 int bOnlyStereo, int bCoarse, int checkOrientation, int variant, int32_t* matches12)
{
    Pool pool; KeyFrame K1, K2; GeometricCamera cam;
    PinholeVerdict pv = {kps1, kps2, F12, sigma2_2};
    cam.verdict = pinhole_verdict; cam.user = &pv;
    fill_keyframe(K1, pool, kps1, desc1, n1, -1, has_mp1, uRight1, fv1_node, fv1_start, fv1_idx, nn1, &cam, nullptr);
    fill_keyframe(K2, pool, kps2, desc2, n2, -1, has_mp2, uRight2, fv2_node, fv2_start, fv2_idx, nn2, &cam, nullptr);
    cam.keys1 = &K1.mvKeysUn; cam.keys2 = &K2.mvKeysUn;
    set_levels(K1, sf2); set_levels(K2, sf2);
    K2.mvLevelSigma2.assign(sigma2_2, sigma2_2 + 8);
    K2.Tcw.at<float>(0, 3) = ep[0]; K2.Tcw.at<float>(1, 3) = ep[1]; K2.Tcw.at<float>(2, 3) = 1.f;
    return run_triangulation(variant ? 1 : 0, K1, K2, n1, bOnlyStereo, bCoarse, checkOrientation, matches12);
}

/* the rig branches (overload 0 / 1) and the matchAndtriangulate overload (2); args as orc_search_for_triangulation_cb.
 * Rig: both keyframes carry mpCamera2 and NLeft = N (every keypoint is a left one; which camera pair the reference picks at
 * :1252-1279 does not matter to the double, whose verdict is the callback). */
int ref_search_for_triangulation_cb(int overload,
                                    const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                    const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                    const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                    const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                    int bOnlyStereo, int bCoarse, int checkOrientation,
                                    orc_pair_verdict_fn verdict, void* user, int32_t* matches12)
{
    Pool pool; KeyFrame K1, K2; GeometricCamera cam, cam2;
    cam.verdict = cam2.verdict = verdict; cam.user = cam2.user = user;
    /* overload 2 reads NLeft only to pick the keypoint vector; a rig there too keeps one code path in this wrapper */
    fill_keyframe(K1, pool, kps1, desc1, n1, n1, has_mp1, nullptr, fv1_node, fv1_start, fv1_idx, nn1, &cam, &cam2);
    fill_keyframe(K2, pool, kps2, desc2, n2, n2, has_mp2, nullptr, fv2_node, fv2_start, fv2_idx, nn2, &cam, &cam2);
    for (GeometricCamera* c : {&cam, &cam2}) { c->keys1 = &K1.mvKeys; c->keys1b = &K1.mvKeysRight; c->nleft1 = n1; c->keys2 = &K2.mvKeys; c->keys2b = &K2.mvKeysRight; c->nleft2 = n2; }
    std::vector<float> sf(8, 1.f);
    for (int i = 1; i < 8; i++) sf[(size_t)i] = sf[(size_t)i - 1] * 1.2f;
    set_levels(K1, sf.data()); set_levels(K2, sf.data());
    return run_triangulation(overload, K1, K2, n1, bOnlyStereo, bCoarse, checkOrientation, matches12);
}

} // extern "C"

namespace {

cv::Mat identity4()
{
    cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
    return T;
}

/* a map point the reference's pose arithmetic (identity pose, camera double (x, y, z) -> (x, y)) projects to pixel (u, v)
 * at depth z, passing the distance (:610-617) and viewing-angle (:621-624) gates: normal = the viewing ray itself */
MapPoint* projected_point(Pool& pool, int index, const uint8_t* desc, float u, float v, float z, int level, int nobs = 1)
{
    MapPoint* p = pool.make(index, desc, nobs);
    p->mWorldPos = point3(u, v, z);
    p->mNormal = point3(u, v, z);
    p->mfMinDistance = 0.f; p->mfMaxDistance = 3.0e38f;
    p->predicted_level = level;
    return p;
}

void fill_keyframe(KeyFrame& K, const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf, GeometricCamera* cam)
{
    K.N = n; K.Nleft = -1; K.NLeft = -1;
    K.mvKeysUn = keys(kps, n); K.mvKeys = K.mvKeysUn; K.mDescriptors = desc_matrix(desc, n);
    K.mvuRight.assign((size_t)n, -1.f);
    K.mpCamera = cam; K.Tcw = identity4(); K.Trl = identity4();
    set_levels(K, sf); set_bounds(K, g); K.grid = const_cast<orc_grid*>(g);
    K.mvpMapPoints.assign((size_t)n, nullptr);
}

} // namespace

extern "C" {

/* ORBmatcher::SearchByProjection(Frame&, KeyFrame*, const set<MapPoint*>&, th, ORBdist), src/ORBmatcher.cc:2723-2862;
 * arguments as orc_search_by_projection_keyframe */
int ref_search_by_projection_keyframe(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                                      int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                      const int32_t* mp_level, const float* mp_angle, const uint8_t* mp_valid,
                                      float th, int ORBdist, int checkOrientation, const uint8_t* cur_held, int32_t* assign)
{
    Frame Cur; KeyFrame K; Pool pool; GeometricCamera cam;
    Cur.N = n; Cur.Nleft = -1; Cur.mvKeysUn = keys(kps, n); Cur.mvKeys = Cur.mvKeysUn; Cur.mDescriptors = desc_matrix(desc, n);
    Cur.mvuRight.assign((size_t)n, -1.f); Cur.mpCamera = &cam; Cur.mTcw = identity4();
    set_levels(Cur, sf); set_bounds(Cur, g); Cur.grid = const_cast<orc_grid*>(g);
    Cur.mvpMapPoints.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (cur_held && cur_held[i]) Cur.mvpMapPoints[i] = pool.make(-2, nullptr, 1);
    K.N = nmp; K.mvKeysUn.resize((size_t)nmp); K.mvpMapPoints.assign((size_t)nmp, nullptr);
    std::set<MapPoint*> already;
    for (int i = 0; i < nmp; i++) {
        K.mvKeysUn[i].angle = mp_angle[i];
        /* invalid rows alternate between the three ways the reference skips a keyframe keypoint (:2746-2750) */
        if (!mp_valid[i] && i % 3 == 0) continue;                                        /* pMP == NULL */
        MapPoint* p = projected_point(pool, i, mp_desc + 32 * (size_t)i, mp_u[i], mp_v[i], 1.f, mp_level[i]);
        if (!mp_valid[i]) { if (i % 3 == 1) p->mbBad = true; else already.insert(p); }
        K.mvpMapPoints[i] = p;
    }
    K.mvKeys = K.mvKeysUn;
    ORBmatcher m(0.9f, checkOrientation != 0);
    const int r = m.SearchByProjection(Cur, &K, already, th, ORBdist);
    for (int i = 0; i < n; i++) {
        MapPoint* p = Cur.mvpMapPoints[i];
        assign[i] = (p && p->index >= 0) ? p->index : -1;
    }
    return r;
}

/* ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:549-678, and (every
 * second call) its twin with vpPointsKFs / vpMatchedKF, :681-797; arguments as orc_search_by_projection_sim3 */
int ref_search_by_projection_sim3(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                                  int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                  const int32_t* mp_level, const uint8_t* mp_valid,
                                  float th, float ratioHamming, const uint8_t* matched, int32_t* assign)
{
    static int flip = 0;
    KeyFrame K, other; Pool pool; GeometricCamera cam;
    fill_keyframe(K, g, kps, desc, n, sf, &cam);
    std::vector<MapPoint*> vpPoints((size_t)nmp), vpMatched((size_t)n, nullptr);
    std::vector<KeyFrame*> vpPointsKFs((size_t)nmp, &other), vpMatchedKF((size_t)n, nullptr);
    for (int i = 0; i < n; i++) if (matched && matched[i]) vpMatched[i] = pool.make(-2, nullptr, 1);
    for (int i = 0; i < nmp; i++) {
        MapPoint* p = projected_point(pool, i, mp_desc + 32 * (size_t)i, mp_u[i], mp_v[i], 1.f, mp_level[i]);
        if (!mp_valid[i]) p->mbBad = true;                                               /* :582 */
        vpPoints[i] = p;
    }
    ORBmatcher m(0.9f, true);
    cv::Mat Scw = identity4();
    const int r = (flip++ & 1) ? m.SearchByProjection(&K, Scw, vpPoints, vpPointsKFs, vpMatched, vpMatchedKF, (int)th, ratioHamming)
                               : m.SearchByProjection(&K, Scw, vpPoints, vpMatched, (int)th, ratioHamming);
    for (int i = 0; i < n; i++) {
        MapPoint* p = vpMatched[i];
        assign[i] = (p && p->index >= 0) ? p->index : -1;
    }
    return r;
}

/* ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th, bRight = false), src/ORBmatcher.cc:1823-2035 (mp_ur != NULL), and
 * Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint), :2051-2185 (mp_ur == NULL); arguments as orc_fuse.
 * The first form computes ur = uv.x - bf * invz itself (:1912): the double uses bf = 1 and the depth z = 1 / (mp_u - mp_ur), so
 * every valid row must satisfy mp_u - fl(1 / z) == mp_ur in float arithmetic (disparities that are powers of two do); -1000
 * otherwise.  best_dist is not observable through the reference's interface: rows come back as -1.
 * best_idx is read from what the reference did to the map: AddObservation(pKF, bestIdx) for the first point that lands on a
 * keypoint, Replace(the point already there) for the later ones (the doubles' observation counts make that the branch taken). */
int ref_fuse(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
             const float* sf, const float* inv_sigma2,
             int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v, const float* mp_ur,
             const int32_t* mp_level, const uint8_t* mp_valid, float th, int32_t* best_idx, int32_t* best_dist)
{
    KeyFrame K; Pool pool; GeometricCamera cam;
    fill_keyframe(K, g, kps, desc, n, sf, &cam);
    if (uRight) K.mvuRight.assign(uRight, uRight + n);
    if (inv_sigma2) K.mvInvLevelSigma2.assign(inv_sigma2, inv_sigma2 + 8);
    K.mbf = 1.f;
    std::vector<MapPoint*> vp((size_t)nmp, nullptr);
    for (int i = 0; i < nmp; i++) {
        best_idx[i] = -1; best_dist[i] = -1;
        if (!mp_valid[i] && mp_ur && i % 3 == 0) continue;                               /* !pMP, :1864 (first form only) */
        float z = 1.f;
        if (mp_ur && mp_valid[i]) {
            const float d = mp_u[i] - mp_ur[i];
            if (!(d > 0.f)) return -1000;
            z = 1.f / d;
            const float invz = 1 / z;
            const float ur = mp_u[i] - 1.f * invz;
            if (ur != mp_ur[i]) return -1000;
        }
        MapPoint* p = projected_point(pool, i, mp_desc + 32 * (size_t)i, mp_u[i], mp_v[i], z, mp_level[i], /*nobs=*/1);
        if (!mp_valid[i]) { if (i % 3 == 1 || !mp_ur) p->mbBad = true; else p->obs[&K] = std::make_pair(0, -1); }   /* isBad / IsInKeyFrame */
        vp[i] = p;
    }
    ORBmatcher m(0.9f, true);
    int r;
    std::vector<MapPoint*> vpReplace((size_t)nmp, nullptr);
    if (mp_ur) r = m.Fuse(&K, vp, th, false);
    else { cv::Mat Scw = identity4(); r = m.Fuse(&K, Scw, vp, th, vpReplace); }
    for (int i = 0; i < nmp; i++) {
        MapPoint* p = vp[i];
        if (!p || !mp_valid[i]) continue;
        MapPoint* there = mp_ur ? p->replaced : vpReplace[i];
        if (there) {                                                                     /* landed on an occupied keypoint */
            auto it = there->obs.find(&K);
            best_idx[i] = it == there->obs.end() ? -2 : it->second.first;
        } else {
            auto it = p->obs.find(&K);
            if (it != p->obs.end()) best_idx[i] = it->second.first;
        }
    }
    return r;
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th), src/ORBmatcher.cc:2201-2434; arguments as orc_search_by_sim3 */
int ref_search_by_sim3(const orc_grid* g1, const orc_keypoint* kps1, const uint8_t* desc1, int n1,
                       const orc_grid* g2, const orc_keypoint* kps2, const uint8_t* desc2, int n2, const float* sf,
                       const uint8_t* mp1_desc, const float* u1in2, const float* v1in2, const int32_t* level1, const uint8_t* valid1,
                       const uint8_t* mp2_desc, const float* u2in1, const float* v2in1, const int32_t* level2, const uint8_t* valid2,
                       float th, int32_t* match12)
{
    KeyFrame K1, K2; Pool pool; GeometricCamera cam;
    fill_keyframe(K1, g1, kps1, desc1, n1, sf, &cam);
    fill_keyframe(K2, g2, kps2, desc2, n2, sf, &cam);
    std::vector<MapPoint*> vpMatches12((size_t)n1, nullptr);
    for (int i = 0; i < n1; i++) {
        if (!valid1[i] && i % 3 == 0) continue;                                          /* !pMP, :2262 */
        MapPoint* p = projected_point(pool, i, mp1_desc + 32 * (size_t)i, u1in2[i], v1in2[i], 1.f, level1[i]);
        if (!valid1[i]) { if (i % 3 == 1) p->mbBad = true; else vpMatches12[i] = pool.make(-2, nullptr, 1); }   /* isBad / vbAlreadyMatched1 */
        K1.mvpMapPoints[i] = p;
    }
    for (int i = 0; i < n2; i++) {
        if (!valid2[i] && i % 2 == 0) continue;
        MapPoint* p = projected_point(pool, i, mp2_desc + 32 * (size_t)i, u2in1[i], v2in1[i], 1.f, level2[i]);
        if (!valid2[i]) p->mbBad = true;
        K2.mvpMapPoints[i] = p;
    }
    ORBmatcher m(0.9f, true);
    const float s12 = 1.f;
    cv::Mat R12 = cv::Mat::eye(3, 3, CV_32F), t12 = cv::Mat::zeros(3, 1, CV_32F);
    const int r = m.SearchBySim3(&K1, &K2, vpMatches12, s12, R12, t12, th);
    for (int i = 0; i < n1; i++) {
        MapPoint* p = vpMatches12[i];
        match12[i] = (p && p->index >= 0) ? p->index : -1;
    }
    return r;
}

} // extern "C"
