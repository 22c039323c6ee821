/*
 * oracle/ref_mp_wrap.cc — TEST INFRASTRUCTURE ONLY.
 *
 * C entry points around the UNMODIFIED reference /root/reference/src/MapPoint.cc (compiled where it lies by oracle/build_ref.sh with
 * the real include/MapPoint.h over oracle/cvshim and oracle/refshim_mp/ref_mp_stubs.h) -> oracle/_ref/libref_mp.so:
 *   ref_mp_distinctive     MapPoint::ComputeDistinctiveDescriptors   src/MapPoint.cc:450-531   pins orc_distinctive_descriptors
 *   ref_mp_predict_scale   MapPoint::PredictScale (Frame* and KeyFrame* forms)  :698-741        pins the level formula of
 *                          orc_is_in_frustum and of the MapPoint double in oracle/refshim_frame
 * (MapPoint::UpdateNormalAndDepth, :561-652, locks mMutexFeatures twice in this fork - :568-569 - and is not called here.)
 */
#define private public
#define protected public
#include "MapPoint.h"
#undef private
#undef protected

#include <cstdint>
#include <cstring>

using namespace ORB_SLAM3;

extern "C" {

/* groups of observation descriptors in CSR form (start[npoints + 1], desc rows of 32 bytes).  Each descriptor becomes the single
 * feature of its own keyframe double; `pairs` != 0 makes consecutive descriptors (2k, 2k+1) of a group the LEFT and RIGHT
 * observation of one rig keyframe (NLeft != -1, :479-486).  The keyframes of a group live in one array, so that the
 * std::map<KeyFrame*, ...> the reference iterates (pointer order) visits them in input order.
 * out_desc[32 * g] = mDescriptor after the call, out_has[g] = 0 when the reference returned early (no observation). */
int ref_mp_distinctive(int npoints, const int32_t* start, const uint8_t* desc, int pairs, uint8_t* out_desc, uint8_t* out_has)
{
    Map map;
    for (int g = 0; g < npoints; g++) {
        const int m = start[g + 1] - start[g];
        const uint8_t* d = desc + 32 * (size_t)start[g];
        const int nkf = pairs ? (m + 1) / 2 : m;
        std::vector<KeyFrame> kfs((size_t)(nkf > 0 ? nkf : 1));
        for (int k = 0; k < nkf; k++) {
            KeyFrame& K = kfs[(size_t)k];
            const int rows = pairs ? ((2 * k + 1 < m) ? 2 : 1) : 1;
            K.mDescriptors = cv::Mat(rows, 32, CV_8U);
            std::memcpy(K.mDescriptors.ptr(0), d + 32 * (size_t)(pairs ? 2 * k : k), 32);
            if (rows == 2) std::memcpy(K.mDescriptors.ptr(1), d + 32 * (size_t)(2 * k + 1), 32);
            K.mvuRight.assign((size_t)rows, -1.f);
            K.NLeft = pairs ? 1 : -1;                 /* index 0 = left feature, index 1 (>= NLeft) = right feature */
            K.mnId = (unsigned long)k;
        }
        cv::Mat pos = cv::Mat::zeros(3, 1, CV_32F);
        MapPoint mp(pos, &kfs[0], &map);
        for (int k = 0; k < nkf; k++) {
            mp.AddObservation(&kfs[(size_t)k], 0);
            if (pairs && 2 * k + 1 < m) mp.AddObservation(&kfs[(size_t)k], 1);
        }
        mp.ComputeDistinctiveDescriptors();
        cv::Mat best = mp.GetDescriptor();
        out_has[g] = best.empty() ? 0 : 1;
        if (!best.empty()) std::memcpy(out_desc + 32 * (size_t)g, best.ptr(0), 32);
        else std::memset(out_desc + 32 * (size_t)g, 0, 32);
    }
    return 0;
}

/* MapPoint::PredictScale for n (maxDistance, currentDist) pairs through both overloads; returns the number of disagreements
 * between the Frame* and the KeyFrame* form (they are the same function body) */
int ref_mp_predict_scale(int n, const float* max_dist, const float* dist, float logScaleFactor, int nlevels, int32_t* level)
{
    Map map; KeyFrame K; Frame F;
    K.mfLogScaleFactor = logScaleFactor; K.mnScaleLevels = nlevels;
    F.mfLogScaleFactor = logScaleFactor; F.mnScaleLevels = nlevels;
    cv::Mat pos = cv::Mat::zeros(3, 1, CV_32F);
    MapPoint mp(pos, &K, &map);
    int diff = 0;
    for (int i = 0; i < n; i++) {
        mp.mfMaxDistance = max_dist[i];
        level[i] = mp.PredictScale(dist[i], &F);
        if (mp.PredictScale(dist[i], &K) != level[i]) diff++;
    }
    return diff;
}

} // extern "C"
