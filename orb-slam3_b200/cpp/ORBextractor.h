// ORBextractor.h — drop-in replacement of the reference's include/ORBextractor.h:28-114 on top of the
// C ABI (include/orbx.h).  Same namespace, class name, constructor, operator() and getters, so that
// src/Frame.cc / src/Tracking.cc compile unchanged against it.
//
// With OpenCV headers available define ORBX_WITH_OPENCV (the real cv::Mat / cv::KeyPoint are used).
// Without them (this build image has no OpenCV C++ headers) a minimal layout-compatible stand-in
// (`cv::Mat`, `cv::KeyPoint`, 28 bytes) is provided so the shim and its example still compile.
#pragma once
#include <cassert>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/orbx.h"

#ifdef ORBX_WITH_OPENCV
#include <opencv2/opencv.hpp>       // as the reference's include/ORBextractor.h:24
#else
namespace cv {
struct Point2f { float x, y; };
struct KeyPoint {                   // memory layout of cv::KeyPoint (28 bytes)
    Point2f pt; float size, angle, response; int octave, class_id;
};
struct Mat {                        // CV_8UC1 only: what the hot path exchanges
    int rows = 0, cols = 0; size_t step = 0; uint8_t* data = nullptr;
    std::vector<uint8_t> storage;
    Mat() {}
    Mat(int r, int c) { create(r, c); }
    Mat(int r, int c, uint8_t* ext, size_t st) : rows(r), cols(c), step(st), data(ext) {}
    void create(int r, int c) { rows = r; cols = c; step = (size_t)c; storage.assign((size_t)r * c, 0); data = storage.data(); }
    void release() { rows = cols = 0; step = 0; data = nullptr; storage.clear(); }
    void shrinkRows(int r) { rows = r; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    uint8_t* ptr(int r) { return data + (size_t)r * step; }
    const uint8_t* ptr(int r) const { return data + (size_t)r * step; }
};
}  // namespace cv
#endif

#ifdef ORBX_WITH_OPENCV
#define ORBX_MAT_CREATE(m, r, c) (m).create((r), (c), CV_8U)
#define ORBX_MAT_SHRINK(m, r) ((m) = (m).rowRange(0, (r)))      // a view of the first r rows: no copy
#else
#define ORBX_MAT_CREATE(m, r, c) (m).create((r), (c))
#define ORBX_MAT_SHRINK(m, r) (m).shrinkRows(r)
#endif

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");

namespace ORB_SLAM3 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // include/ORBextractor.h:49-50
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
    {
        if (orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, &h_) != ORBX_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        const int n = orbx_levels(h_);
        mvScaleFactor.assign(orbx_scale_factors(h_), orbx_scale_factors(h_) + n);
        mvInvScaleFactor.assign(orbx_inv_scale_factors(h_), orbx_inv_scale_factors(h_) + n);
        mvLevelSigma2.assign(orbx_level_sigma2(h_), orbx_level_sigma2(h_) + n);
        mvInvLevelSigma2.assign(orbx_inv_level_sigma2(h_), orbx_inv_level_sigma2(h_) + n);
    }
    ~ORBextractor() { orbx_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // include/ORBextractor.h:57-59, src/ORBextractor.cc:1534-1659.  The mask is ignored (as in the
    // reference); outputs are resized by the callee; returns monoIndex, -1 for an empty image.
    // The descriptor matrix is created with the capacity, filled by the library in place and cut to the
    // keypoint count as a row-range view: no second host pass over the descriptors.
#ifdef ORBX_WITH_OPENCV
    int operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea)
    {
        (void)_mask;
        if (_image.empty()) return -1;                                      // :1538
        cv::Mat image = _image.getMat();
        assert(image.type() == CV_8UC1);                                    // :1544
        cv::Mat descriptors;
        const int mono = extract(image, _keypoints, descriptors, vLappingArea);
        if (descriptors.empty()) _descriptors.release();                    // :1570-1572
        else descriptors.copyTo(_descriptors);                              // shares nothing with the capacity buffer the view came from
        return mono;
    }
    // the overload Frame::ExtractORB reaches when it passes plain cv::Mat objects (src/Frame.cc:482-492): fills the
    // caller's Mat in place
    int operator()(const cv::Mat& image, const cv::Mat& /*mask*/, std::vector<cv::KeyPoint>& keypoints,
                   cv::Mat& descriptors, std::vector<int>& vLappingArea)
    {
        if (image.empty()) return -1;
        assert(image.type() == CV_8UC1);
        return extract(image, keypoints, descriptors, vLappingArea);
    }
#else
    int operator()(const cv::Mat& image, const cv::Mat& /*mask*/, std::vector<cv::KeyPoint>& keypoints,
                   cv::Mat& descriptors, std::vector<int>& vLappingArea)
    {
        if (image.empty()) return -1;                                       // :1538
        return extract(image, keypoints, descriptors, vLappingArea);
    }
#endif

    // include/ORBextractor.h:61-81
    int GetLevels() { return orbx_levels(h_); }
    float GetScaleFactor() { return orbx_scale_factor(h_); }
    std::vector<float> GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // mvImagePyramid (include/ORBextractor.h:83) lives on the device; copy a level out on demand.
    cv::Mat GetPyramidLevel(int level, bool padded = false)
    {
        int r = 0, c = 0;
        if (orbx_get_pyramid_level(h_, 0, level, padded, nullptr, 0, &r, &c) != ORBX_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        cv::Mat m; ORBX_MAT_CREATE(m, r, c);
        orbx_get_pyramid_level(h_, 0, level, padded, m.data, m.step, &r, &c);
        return m;
    }

    // ---- frame preparation on the device (no reference member: replaces the caller-side cv::remap of
    // Examples/Stereo/stereo_euroc.cc:168-169 and the cvtColor of src/Tracking.cc:1264-1290) ----
    // M1 / M2 of cv::initUndistortRectifyMap(..., CV_32F, M1, M2) (stereo_euroc.cc:117-118) for this camera.
    void SetRectifyMaps(const float* mapx, const float* mapy, int rows, int cols, size_t step_bytes)
    {
        if (orbx_set_rectify_maps(h_, mapx, mapy, rows, cols, step_bytes) != ORBX_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        rect_rows_ = mapx ? rows : 0; rect_cols_ = mapx ? cols : 0;
    }
#ifdef ORBX_WITH_OPENCV
    void SetRectifyMaps(const cv::Mat& M1, const cv::Mat& M2)
    {
        assert(M1.type() == CV_32FC1 && M2.type() == CV_32FC1 && M1.size() == M2.size() && M1.step == M2.step);
        SetRectifyMaps(M1.ptr<float>(), M2.ptr<float>(), M1.rows, M1.cols, M1.step);
    }
    // raw camera image (CV_8UC1 / 8UC3 / 8UC4, unrectified) -> remap -> gray -> operator()
    int ExtractRaw(const cv::Mat& im, bool bRGB, std::vector<cv::KeyPoint>& keypoints, cv::Mat& descriptors, std::vector<int>& vLappingArea)
    {
        if (im.empty()) return -1;
        return ExtractRaw(im.data, im.rows, im.cols, im.channels(), im.step, bRGB, keypoints, descriptors, vLappingArea);
    }
#endif
    int ExtractRaw(const uint8_t* data, int rows, int cols, int channels, size_t step, bool bRGB,
                   std::vector<cv::KeyPoint>& keypoints, cv::Mat& descriptors, std::vector<int>& vLappingArea)
    {
        if (!data || rows <= 0 || cols <= 0) return -1;
        assert(vLappingArea.size() >= 2);
        const int prow = rect_rows_ ? rect_rows_ : rows, pcol = rect_cols_ ? rect_cols_ : cols;
        const int cap = orbx_max_keypoints(h_, prow, pcol);
        if (cap < 0) throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        keypoints.resize((size_t)cap);
        scratch_.resize((size_t)cap * 32);
        int n = 0;
        const int mono = orbx_prepare_extract(h_, data, rows, cols, channels, bRGB ? 1 : 0, step, vLappingArea[0], vLappingArea[1],
                                              reinterpret_cast<orbx_keypoint*>(keypoints.data()), scratch_.data(), cap, &n);
        if (mono < 0) throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        keypoints.resize((size_t)n);
        if (n == 0) descriptors.release();
        else {
            ORBX_MAT_CREATE(descriptors, n, 32);
            for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), scratch_.data() + (size_t)i * 32, 32);
        }
        prep_rows_ = prow; prep_cols_ = pcol;
        return mono;
    }
    // imLeftRect / mImGray: the frame the last ExtractRaw call extracted from (for the viewer, Tracking::mImGray)
    cv::Mat GetPreparedImage()
    {
        cv::Mat m; ORBX_MAT_CREATE(m, prep_rows_, prep_cols_);
        if (orbx_get_prepared(h_, 0, m.data, m.step) != ORBX_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        return m;
    }

    orbx_extractor* handle() { return h_; }

    // include/ORBextractor.h:83: `std::vector<cv::Mat> mvImagePyramid`, which only Frame::ComputeStereoMatches reads
    // (src/Frame.cc:1051,1168,1194) - replaced by ORBmatcherGPU::ComputeStereoMatches on the device-resident pyramid.
    // Callers that still want the member: SetKeepHostPyramid(true) makes every operator() call copy the 8 levels out.
    std::vector<cv::Mat> mvImagePyramid;
    void SetKeepHostPyramid(bool on) { keep_host_pyramid_ = on; }

protected:
    int extract(const cv::Mat& image, std::vector<cv::KeyPoint>& keypoints, cv::Mat& descriptors, std::vector<int>& vLappingArea)
    {
        assert(vLappingArea.size() >= 2);
        const int cap = orbx_max_keypoints(h_, image.rows, image.cols);
        if (cap < 0) throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        keypoints.resize((size_t)cap);
        ORBX_MAT_CREATE(descriptors, cap, 32);                              // :1577 (CV_8U), at capacity
        int n = 0;
        const int mono = orbx_extract(h_, image.data, image.rows, image.cols, image.step, vLappingArea[0], vLappingArea[1],
                                      reinterpret_cast<orbx_keypoint*>(keypoints.data()), descriptors.data, cap, &n);
        if (mono < 0) throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error());
        keypoints.resize((size_t)n);                                        // :1587
        if (n == 0) descriptors.release();                                  // :1570-1572
        else ORBX_MAT_SHRINK(descriptors, n);
        if (keep_host_pyramid_) {
            mvImagePyramid.resize((size_t)orbx_levels(h_));
            for (int l = 0; l < (int)mvImagePyramid.size(); l++) mvImagePyramid[(size_t)l] = GetPyramidLevel(l);
        }
        return mono;
    }
    bool keep_host_pyramid_ = false;
    orbx_extractor* h_ = nullptr;
    std::vector<uint8_t> scratch_;
    int rect_rows_ = 0, rect_cols_ = 0, prep_rows_ = 0, prep_cols_ = 0;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
};

}  // namespace ORB_SLAM3
