/*
 * orc_prepare.c — CPU ORACLE (test infrastructure) for the frame preparation in front of
 * ORBextractor::operator(): stereo rectification and grayscale conversion.
 *
 * Reference call sites:
 *   Examples/Stereo/stereo_euroc.cc:117-118  cv::initUndistortRectifyMap(..., CV_32F, M1, M2)   (maps, once)
 *   Examples/Stereo/stereo_euroc.cc:168-169  cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR)    (every frame)
 *   src/Tracking.cc:1264-1290                cvtColor(..., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY)      (every frame)
 * Both primitives live in OpenCV (un-vendored, unpinned by the reference); they are restated here from
 * the published fixed-point algorithms and pinned against cv2 4.13.0 (tests/test_oracle_vs_cv2.py):
 *
 *   remap, CV_32FC1 maps, u8, INTER_LINEAR, BORDER_CONSTANT(0):
 *     sx = cvRound(mapx * 32), sy = cvRound(mapy * 32); ix = sx >> 5, iy = sy >> 5 (saturated to short),
 *     fx = sx & 31, fy = sy & 31; weights (32-fx)(32-fy)*32, fx(32-fy)*32, (32-fx)fy*32, fx*fy*32 (sum 2^15);
 *     dst = (sum_k tap_k * w_k + 2^14) >> 15, taps outside the source read 0.
 *   cvtColor to gray, u8: (R*9798 + G*19235 + B*3735 + 2^14) >> 15.
 */
#include "orc.h"
#include <math.h>

static int sat_short(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

void orc_remap_linear_u8(const uint8_t* src, int srows, int scols, size_t sstep, int channels,
                         const float* mapx, const float* mapy, size_t mstep_floats,
                         uint8_t* dst, int drows, int dcols, size_t dstep)
{
    for (int y = 0; y < drows; y++) {
        const float* mx = mapx + (size_t)y * mstep_floats;
        const float* my = mapy + (size_t)y * mstep_floats;
        uint8_t* d = dst + (size_t)y * dstep;
        for (int x = 0; x < dcols; x++) {
            const int sx = orc_cvround_f(mx[x] * 32.0f), sy = orc_cvround_f(my[x] * 32.0f);
            const int ix = sat_short(sx >> 5), iy = sat_short(sy >> 5);
            const int fx = sx & 31, fy = sy & 31;
            const int w[4] = {(32 - fx) * (32 - fy) * 32, fx * (32 - fy) * 32, (32 - fx) * fy * 32, fx * fy * 32};
            for (int c = 0; c < channels; c++) {
                int acc = 1 << 14;
                for (int k = 0; k < 4; k++) {
                    const int tx = ix + (k & 1), ty = iy + (k >> 1);
                    const int v = (tx >= 0 && tx < scols && ty >= 0 && ty < srows)
                                      ? src[(size_t)ty * sstep + (size_t)tx * channels + c] : 0;
                    acc += v * w[k];
                }
                d[(size_t)x * channels + c] = (uint8_t)(acc >> 15);
            }
        }
    }
}

/* channels 3 or 4; rgb != 0: first channel is R (mbRGB, src/Tracking.cc:1266), else B. */
void orc_cvt_gray_u8(const uint8_t* src, int rows, int cols, size_t sstep, int channels, int rgb,
                     uint8_t* dst, size_t dstep)
{
    for (int y = 0; y < rows; y++) {
        const uint8_t* s = src + (size_t)y * sstep;
        uint8_t* d = dst + (size_t)y * dstep;
        for (int x = 0; x < cols; x++) {
            const int c0 = s[(size_t)x * channels], c1 = s[(size_t)x * channels + 1], c2 = s[(size_t)x * channels + 2];
            const int r = rgb ? c0 : c2, b = rgb ? c2 : c0;
            d[x] = (uint8_t)((r * 9798 + c1 * 19235 + b * 3735 + (1 << 14)) >> 15);
        }
    }
}

/* cv::undistortPoints(src, dst, K, distCoef, cv::Mat(), K) as Frame::UndistortKeyPoints calls it (src/Frame.cc:924-960)
 * and Frame::ComputeImageBounds (:978-1009): OpenCV's cvUndistortPointsInternal restated (double arithmetic, 5 fixed
 * iterations = TermCriteria(MAX_ITER, 5, 0.01), R = identity, P = K), pinned against cv2 4.13
 * (tests/test_oracle_vs_cv2.py).  dist: k1, k2, p1, p2, k3 (mDistCoef has 4 or 5 entries; pass k3 = 0 for 4). */
void orc_undistort_points(const float* xy_in, int n, float fx, float fy, float cx, float cy, const float* dist, float* xy_out)
{
    const double dfx = fx, dfy = fy, dcx = cx, dcy = cy;
    const double ifx = 1. / dfx, ify = 1. / dfy;
    double k[12] = {0};
    k[0] = dist[0]; k[1] = dist[1]; k[2] = dist[2]; k[3] = dist[3]; k[4] = dist[4];
    for (int i = 0; i < n; i++) {
        double x = xy_in[2 * i], y = xy_in[2 * i + 1];
        const double u = x, v = y;
        x = (x - dcx) * ifx;
        y = (y - dcy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - dcx) * ifx; y = (v - dcy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = dfx * x + 0.0 * y + dcx, yy = 0.0 * x + dfy * y + dcy;
        const double ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        xy_out[2 * i] = (float)(xx * ww);
        xy_out[2 * i + 1] = (float)(yy * ww);
    }
}
