// example_stereo.cc — what the reference's stereo Frame constructor (src/Frame.cc:103-225) does with the
// extractor and matcher, written against the shims (no OpenCV needed).  Build + run on a GPU box:
//   g++ -O2 -std=c++17 orb-slam3_b200/cpp/example_stereo.cc -Lorb-slam3_b200 -lorbx -Wl,-rpath,$PWD/orb-slam3_b200 -o example_stereo
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cmath>
#include <memory>
#include <thread>

#include "ORBmatcher.h"
#include "ORBVocabulary.h"

static cv::Mat synth(int rows, int cols, unsigned seed, int shift)
{
    cv::Mat m(rows, cols);
    std::vector<uint8_t> base((size_t)rows * (cols + 64));
    srand(seed);
    for (auto& v : base) v = (uint8_t)(rand() & 0xFF);
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            // 3x3 box of the noise = band-limited texture; the right image is the left shifted by `shift`
            int s = 0;
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++) {
                    int yy = y + dy < 0 ? 0 : (y + dy >= rows ? rows - 1 : y + dy);
                    s += base[(size_t)yy * (cols + 64) + x + shift + 1 + dx];
                }
            m.ptr(y)[x] = (uint8_t)(s / 9);
        }
    return m;
}

static ORB_SLAM3::ORBmatcherGPU::KeyFrameView K1nomp(ORB_SLAM3::ORBmatcherGPU::KeyFrameView k)
{
    k.hasMapPoint.assign(k.hasMapPoint.size(), 0);
    return k;
}

int main()
{
    using namespace ORB_SLAM3;
    const int rows = 480, cols = 752;
    try {
        // src/Tracking.cc:1106-1110: mpORBextractorLeft / mpORBextractorRight
        ORBextractor left(1200, 1.2f, 8, 20, 7), right(1200, 1.2f, 8, 20, 7);
        cv::Mat imLeft = synth(rows, cols, 1, 0), imRight = synth(rows, cols, 1, 20);   // uR = uL - 20
        std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
        cv::Mat mDescriptors, mDescriptorsRight;
        std::vector<int> vLapping = {0, 0};
        // src/Frame.cc:132-137: two threads, one extractor each
        int monoLeft = 0, monoRight = 0;
        std::thread tl([&] { monoLeft = left(imLeft, cv::Mat(), mvKeys, mDescriptors, vLapping); });
        std::thread tr([&] { monoRight = right(imRight, cv::Mat(), mvKeysRight, mDescriptorsRight, vLapping); });
        tl.join(); tr.join();
        // src/Frame.cc:159: ComputeStereoMatches (EuRoC: bf = 47.90639, fx = 435.2047)
        std::vector<float> mvuRight, mvDepth;
        const float mbf = 47.90639f, mb = mbf / 435.2047f;
        ORBmatcherGPU::ComputeStereoMatches(left, right, (int)mvKeys.size(), mb, mbf, mvuRight, mvDepth);
        int matched = 0;
        for (float u : mvuRight) matched += u >= 0;
        std::printf("left %zu kps (mono %d), right %zu kps (mono %d), stereo matches %d, d(0,1) = %d\n", mvKeys.size(), monoLeft,
                    mvKeysRight.size(), monoRight, matched,
                    mvKeys.size() > 1 ? ORBmatcherGPU::DescriptorDistance(cv::Mat(1, 32, mDescriptors.ptr(0), 32), cv::Mat(1, 32, mDescriptors.ptr(1), 32)) : -1);
        // --- the tracking-side searches on a second view (the scene moved 6 px to the left) ---
        cv::Mat imNext = synth(rows, cols, 1, 6);
        std::vector<cv::KeyPoint> mvKeys2;
        cv::Mat mDescriptors2;
        left(imNext, cv::Mat(), mvKeys2, mDescriptors2, vLapping);
        std::vector<float> sf = left.GetScaleFactors(), sigma2 = left.GetScaleSigmaSquares();
        ORBmatcherGPU::FrameView F1 = {&mvKeys, &mDescriptors, &mvuRight, 0.f, (float)cols, 0.f, (float)rows, &sf};
        ORBmatcherGPU::FrameView F2 = {&mvKeys2, &mDescriptors2, nullptr, 0.f, (float)cols, 0.f, (float)rows, &sf};
        // Tracking::MonocularInitialization, src/Tracking.cc: matcher(0.9, true).SearchForInitialization(..., 100)
        ORBmatcherGPU matcherIni(0.9f, true);
        std::vector<cv::Point2f> vbPrevMatched(mvKeys.size());
        for (size_t i = 0; i < mvKeys.size(); i++) vbPrevMatched[i] = mvKeys[i].pt;
        std::vector<int> vnMatches12;
        const int nIni = matcherIni.SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, 100);
        // Tracking::TrackWithMotionModel: matcher(0.9, true).SearchByProjection(mCurrentFrame, mLastFrame, th, bMono)
        ORBmatcherGPU::LastFrameSoA last;
        last.desc.assign(mDescriptors.data, mDescriptors.data + (size_t)mDescriptors.rows * 32);
        for (const cv::KeyPoint& kp : mvKeys) {
            last.u.push_back(kp.pt.x - 6.f); last.v.push_back(kp.pt.y); last.invz.push_back(0.2f);
            last.angle.push_back(kp.angle); last.octave.push_back(kp.octave); last.valid.push_back(1); last.obs.push_back(1);
        }
        std::vector<int32_t> assign;
        const int nProj = matcherIni.SearchByProjection(F2, last, 15.f, 0.f, false, false, nullptr, assign);
        // LocalMapping::CreateNewMapPoints: matcher(0.6, false).SearchForTriangulation(pKF1, pKF2, F12, pairs, false)
        ORBmatcherGPU matcherTri(0.6f, false);
        ORBmatcherGPU::KeyFrameView K1 = {&mvKeys, &mDescriptors, nullptr, std::vector<uint8_t>(mvKeys.size(), 0), {}, {}, {}, &sf, &sigma2};
        ORBmatcherGPU::KeyFrameView K2 = {&mvKeys2, &mDescriptors2, nullptr, std::vector<uint8_t>(mvKeys2.size(), 0), {}, {}, {}, &sf, &sigma2};
        for (int k = 0; k < 2; k++) {           // one vocabulary node holding every keypoint
            ORBmatcherGPU::KeyFrameView& K = k ? K2 : K1;
            K.fvNode = {0}; K.fvStart = {0, (int32_t)K.mvKeysUn->size()};
            for (size_t i = 0; i < K.mvKeysUn->size(); i++) K.fvIdx.push_back((int32_t)i);
        }
        const float F12[9] = {0, 0, 0, 0, 0, -1, 0, 1, 0};   // sideways motion: epipolar lines are the image rows
        const float ep[2] = {-1e6f, 0.f};
        std::vector<std::pair<size_t, size_t>> vMatchedPairs;
        const int nTri = matcherTri.SearchForTriangulation(K1, K2, F12, ep, vMatchedPairs, false);
        // Tracking::TrackReferenceKeyFrame: matcher(0.7, true).SearchByBoW(mpReferenceKF, mCurrentFrame, vpMapPointMatches)
        ORBmatcherGPU matcherBoW(0.7f, true);
        K1.hasMapPoint.assign(mvKeys.size(), 1);
        std::vector<int32_t> bowMatches;
        const int nBoW = matcherBoW.SearchByBoW(K1, K2, false, bowMatches);
        std::printf("second view: %zu kps; SearchForInitialization %d, SearchByProjection(Frame,Frame) %d, SearchForTriangulation %d, "
                    "SearchByBoW %d\n", mvKeys2.size(), nIni, nProj, nTri, nBoW);
        // --- two-camera rig (KannalaBrandt stereo, Frame::Nleft != -1): the fisheye Frame constructor's sequence ---
        {
            std::vector<int> lapRig = {100, 600};                   // vLappingArea of the two fisheye cameras
            std::vector<cv::KeyPoint> kL, kR;
            cv::Mat dL, dR;
            const int monoLeft = left(imLeft, cv::Mat(), kL, dL, lapRig);        // src/Frame.cc:1363-1366 (two threads there)
            const int monoRight = right(imRight, cv::Mat(), kR, dR, lapRig);
            std::vector<int> l2r, r2l;
            const int nFish = ORBmatcherGPU::ComputeStereoFishEyeMatches(left, right, (int)kL.size(), monoLeft, (int)kR.size(), monoRight, l2r, r2l);
            // Frame keypoints in mvpMapPoints order (left rows, then right rows) for the rig searches
            std::vector<cv::KeyPoint> kAll(kL); kAll.insert(kAll.end(), kR.begin(), kR.end());
            cv::Mat dAll; ORBX_MAT_CREATE(dAll, (int)kAll.size(), 32);
            for (int i = 0; i < dL.rows; i++) std::memcpy(dAll.ptr(i), dL.ptr(i), 32);
            for (int i = 0; i < dR.rows; i++) std::memcpy(dAll.ptr(dL.rows + i), dR.ptr(i), 32);
            ORBmatcherGPU::KeyFrameView FR = {&kAll, &dAll, nullptr, std::vector<uint8_t>(kAll.size(), 0), {0}, {0, (int32_t)kAll.size()}, {}, &sf, &sigma2};
            for (size_t i = 0; i < kAll.size(); i++) FR.fvIdx.push_back((int32_t)i);
            std::vector<int32_t> bowRig;
            const int nBoWRig = matcherBoW.SearchByBoWRig(K1, FR, (int)kL.size(), bowRig);
            ORBmatcherGPU::LastFrameRigSoA lastRig;
            static_cast<ORBmatcherGPU::LastFrameSoA&>(lastRig) = last;
            lastRig.ur = last.u; lastRig.vr = last.v;
            for (float& u : lastRig.ur) u -= 20.f;
            std::vector<int32_t> assignRig;
            const int nProjRig = matcherIni.SearchByProjection(kL, kR, dAll, 0.f, (float)cols, 0.f, (float)rows, sf, lastRig, 15.f, false, false,
                                                               nullptr, assignRig);
            // SearchForTriangulation with the camera model's own test as a callable (here: rows within 2 px)
            int asked = 0;
            auto verdict = [&](int i1, int i2) { asked++; return std::fabs(mvKeys[(size_t)i1].pt.y - kAll[(size_t)i2].pt.y) < 2.f; };
            std::vector<std::pair<size_t, size_t>> pairsRig;
            const int nTriRig = matcherTri.SearchForTriangulation(0, K1nomp(K1), FR, verdict, pairsRig, false);
            std::printf("rig: fisheye stereo candidates %d (monoLeft %d, monoRight %d), SearchByBoW %d, SearchByProjection(Frame,Frame) %d, "
                        "SearchForTriangulation %d (%d verdict calls)\n", nFish, monoLeft, monoRight, nBoWRig, nProjRig, nTriRig, asked);
            if (nFish <= 0 || nBoWRig <= 0 || nProjRig <= 0 || nTriRig <= 0 || asked < nTriRig) return 3;
            left(imLeft, cv::Mat(), mvKeys, mDescriptors, vLapping);             // restore the resident state used below
            right(imRight, cv::Mat(), mvKeysRight, mDescriptorsRight, vLapping);
        }
        // --- one host batch over several extractor handles (one per GPU on a multi-GPU box): results in frame order ---
        {
            int ndevs = orbx_device_count();
            std::vector<std::unique_ptr<ORBextractor>> ex;
            std::vector<orbx_extractor*> hs;
            const int nh = ndevs > 1 ? ndevs : 2;                   // a single-GPU box: two handles on device 0
            for (int d = 0; d < nh; d++) { ex.emplace_back(new ORBextractor(1200, 1.2f, 8, 20, 7, d % ndevs)); hs.push_back(ex.back()->handle()); }
            const int nfr = 2 * nh + 1;
            std::vector<uint8_t> batch((size_t)nfr * rows * cols);
            for (int f = 0; f < nfr; f++) {
                const cv::Mat& src = (f & 1) ? imRight : imLeft;
                for (int y = 0; y < rows; y++) std::memcpy(batch.data() + ((size_t)f * rows + y) * cols, src.ptr(y), (size_t)cols);
            }
            const int cap = orbx_max_keypoints(hs[0], rows, cols);
            std::vector<orbx_keypoint> bk((size_t)nfr * cap);
            std::vector<uint8_t> bd((size_t)nfr * cap * 32);
            std::vector<int> bn((size_t)nfr), bm((size_t)nfr);
            if (orbx_extract_batch_multi(hs.data(), nh, nfr, batch.data(), rows, cols, (size_t)cols, (size_t)rows * cols, 0, 0,
                                         bk.data(), bd.data(), cap, bn.data(), bm.data()) != ORBX_OK)
                throw std::runtime_error(std::string("extract_batch_multi: ") + orbx_last_error());
            bool ok = true;
            for (int f = 0; f < nfr && ok; f++) {
                const std::vector<cv::KeyPoint>& rk = (f & 1) ? mvKeysRight : mvKeys;
                const cv::Mat& rd = (f & 1) ? mDescriptorsRight : mDescriptors;
                ok = bn[(size_t)f] == (int)rk.size() && std::memcmp(&bk[(size_t)f * cap], rk.data(), rk.size() * sizeof(cv::KeyPoint)) == 0;
                for (int i = 0; ok && i < rd.rows; i++) ok = std::memcmp(&bd[((size_t)f * cap + i) * 32], rd.ptr(i), 32) == 0;
            }
            std::printf("multi-device batch: %d frames over %d handles on %d device(s): %s\n", nfr, nh, ndevs, ok ? "identical to operator()" : "MISMATCH");
            if (!ok) return 4;
        }
        // --- the same stereo frame through the one-call entry: identical results ---
        std::vector<cv::KeyPoint> k1, k2;
        cv::Mat d1, d2;
        std::vector<float> u1, z1;
        ORBmatcherGPU::ExtractStereo(left, right, imLeft, imRight, k1, d1, k2, d2, mb, mbf, u1, z1);
        bool same = k1.size() == mvKeys.size() && k2.size() == mvKeysRight.size() && u1 == mvuRight && z1 == mvDepth;
        for (size_t i = 0; same && i < k1.size(); i++)
            same = std::memcmp(&k1[i], &mvKeys[i], sizeof(cv::KeyPoint)) == 0 && std::memcmp(d1.ptr((int)i), mDescriptors.ptr((int)i), 32) == 0;
        // --- what the C++ caller waits per frame (host cv::Mat in, host vectors out) ---
        auto ms = [](auto&& fn, int reps) {
            for (int i = 0; i < 10; i++) fn();
            const auto t0 = std::chrono::steady_clock::now();
            for (int i = 0; i < reps; i++) fn();
            return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / reps;
        };
        const double tMono = ms([&] { left(imLeft, cv::Mat(), mvKeys, mDescriptors, vLapping); }, 200);
        const double tThreads = ms([&] {
            std::thread a([&] { left(imLeft, cv::Mat(), mvKeys, mDescriptors, vLapping); });
            std::thread b([&] { right(imRight, cv::Mat(), mvKeysRight, mDescriptorsRight, vLapping); });
            a.join(); b.join();
            ORBmatcherGPU::ComputeStereoMatches(left, right, (int)mvKeys.size(), mb, mbf, mvuRight, mvDepth);
        }, 200);
        const double tOne = ms([&] { ORBmatcherGPU::ExtractStereo(left, right, imLeft, imRight, k1, d1, k2, d2, mb, mbf, u1, z1); }, 200);
        std::printf("one-call stereo identical: %s; latency per frame: mono operator() %.3f ms, stereo (two threads + "
                    "ComputeStereoMatches) %.3f ms, stereo (ExtractStereo) %.3f ms\n", same ? "yes" : "NO", tMono, tThreads, tOne);
        return (matched > 0 && nIni > 0 && nProj > 0 && nTri > 0 && nBoW > 0 && (int)vMatchedPairs.size() == nTri && same) ? 0 : 2;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
