/*
 * orc_prims.c — ORACLE (test infrastructure): restatements of the OpenCV primitives the
 * reference's hot path calls.  OpenCV is an un-vendored, unpinned dependency of the reference
 * (CMakeLists.txt:37-43); these follow OpenCV 4.x's published algorithms and are pinned against
 * cv2 4.13.0 by tests/test_oracle_vs_cv2.py and the fixtures in tests/golden/.
 * Call sites in the reference: resize src/ORBextractor.cc:1679; copyMakeBorder :1689,1711;
 * FAST :1112,1121; GaussianBlur :1609; fastAtan2 :137; cvRound :97,160,168,519,559,1669;
 * cv::norm(L1) src/Frame.cc:1197; BFMatcher::knnMatch src/Frame.cc:1458.
 */
#include "orc.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* cvRound: SSE cvtss2si / cvtsd2si => round to nearest, ties to even (default MXCSR). */
int orc_cvround_f(float v) { return (int)lrintf(v); }
int orc_cvround_d(double v) { return (int)lrint(v); }

/* cv::fastAtan2 (scalar atan_f32): 7th-order odd polynomial in degrees, float32, no FMA. */
float orc_fast_atan2(float y, float x)
{
    const float scale = (float)(180.0 / 3.1415926535897932384626433832795);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/*
 * glibc >= 2.28 sinf/cosf (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf_poly.h — the ARM
 * optimized-routines algorithm): double-precision range reduction by pi/2 and degree-7/8 minimax
 * polynomials.  Restated for |x| < 120 (the descriptor only ever passes [0, 2*pi]); evaluated
 * WITHOUT FMA contraction (= glibc's __sinf_sse2/__cosf_sse2 variants).  The reference calls
 * cos/sin on a float (src/ORBextractor.cc:155), i.e. these functions on a glibc host.
 */
static const double SC_HPI_INV = 0x1.45F306DC9C883p+23; /* 2/pi * 2^24 */
static const double SC_HPI = 0x1.921FB54442D18p0;
static const double SC_C0 = 0x1p0, SC_C1 = -0x1.ffffffd0c621cp-2, SC_C2 = 0x1.55553e1068f19p-5,
                    SC_C3 = -0x1.6c087e89a359dp-10, SC_C4 = 0x1.99343027bf8c3p-16;
static const double SC_S1 = -0x1.555545995a603p-3, SC_S2 = 0x1.1107605230bc4p-7,
                    SC_S3 = -0x1.994eb3774cf24p-13;

static inline uint32_t abstop12(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    return (u >> 20) & 0x7ff;
}

/* sinf_poly: n even -> sine polynomial of x, n odd -> cosine polynomial; neg selects table[1]. */
static inline float sc_poly(double x, double x2, int neg, int n)
{
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = SC_S2 + x2 * SC_S3;
        double x7 = x3 * x2;
        double s = x + x3 * SC_S1;
        return (float)(s + x7 * s1);
    } else {
        double sg = neg ? -1.0 : 1.0;
        double x4 = x2 * x2;
        double c2 = sg * SC_C3 + x2 * (sg * SC_C4);
        double c1 = sg * SC_C0 + x2 * (sg * SC_C1);
        double x6 = x4 * x2;
        double c = c1 + x4 * (sg * SC_C2);
        return (float)(c + x6 * c2);
    }
}

static inline double sc_reduce(double x, int* np)
{
    double r = x * SC_HPI_INV;
    int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return x - n * SC_HPI;
}

static const double SC_SIGN[4] = {1.0, -1.0, -1.0, 1.0};

float orc_sinf(float y)
{
    double x = y;
    int n;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) { /* |y| < pi/4 */
        double s = x * x;
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return sc_poly(x, s, 0, 0);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        x = sc_reduce(x, &n);
        double s = SC_SIGN[n & 3];
        return sc_poly(x * s, x * x, (n & 2) != 0, n);
    }
    return sinf(y); /* out of the restated range; never reached by the hot path */
}

float orc_cosf(float y)
{
    double x = y;
    int n;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double s = x * x;
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return sc_poly(x, s, 0, 1);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        x = sc_reduce(x, &n);
        double s = SC_SIGN[n & 3];
        return sc_poly(x * s, x * x, (n & 2) != 0, n ^ 1);
    }
    return cosf(y);
}

/*
 * cv::resize(..., INTER_LINEAR) for CV_8UC1 (imgproc/src/resize.cpp: resizeGeneric_ with
 * HResizeLinear<uchar,int,short,2048> and VResizeLinear<uchar,int,short,FixedPtCast<...,22>>).
 * Horizontal taps are reset (fx=0) at the borders; vertical source rows are only clipped.
 */
void orc_resize_linear_u8(const uint8_t* src, int srows, int scols, size_t sstep,
                          uint8_t* dst, int drows, int dcols, size_t dstep)
{
    const double scale_x = (double)scols / dcols, scale_y = (double)srows / drows;
    int* xofs = (int*)malloc(sizeof(int) * dcols);
    short* ialpha = (short*)malloc(sizeof(short) * 2 * dcols);
    int* row0 = (int*)malloc(sizeof(int) * dcols);
    int* row1 = (int*)malloc(sizeof(int) * dcols);
    for (int dx = 0; dx < dcols; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= scols - 1) { fx = 0; sx = scols - 1; }
        xofs[dx] = sx;
        ialpha[2 * dx] = (short)orc_cvround_f((1.f - fx) * 2048.f);
        ialpha[2 * dx + 1] = (short)orc_cvround_f(fx * 2048.f);
    }
    int prev_sy0 = -1000000, prev_sy1 = -1000000;
    for (int dy = 0; dy < drows; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)floorf(fy);
        fy -= sy;
        short b0 = (short)orc_cvround_f((1.f - fy) * 2048.f);
        short b1 = (short)orc_cvround_f(fy * 2048.f);
        int sy0 = sy < 0 ? 0 : (sy > srows - 1 ? srows - 1 : sy);
        int sy1 = sy + 1 < 0 ? 0 : (sy + 1 > srows - 1 ? srows - 1 : sy + 1);
        if (sy0 != prev_sy0 || sy1 != prev_sy1) {
            const uint8_t* S0 = src + (size_t)sy0 * sstep;
            const uint8_t* S1 = src + (size_t)sy1 * sstep;
            for (int dx = 0; dx < dcols; dx++) {
                int sx = xofs[dx];
                int sx1 = sx + 1 < scols ? sx + 1 : scols - 1;
                int a0 = ialpha[2 * dx], a1 = ialpha[2 * dx + 1];
                row0[dx] = S0[sx] * a0 + S0[sx1] * a1;
                row1[dx] = S1[sx] * a0 + S1[sx1] * a1;
            }
            prev_sy0 = sy0; prev_sy1 = sy1;
        }
        uint8_t* D = dst + (size_t)dy * dstep;
        for (int dx = 0; dx < dcols; dx++)
            D[dx] = (uint8_t)((((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2);
    }
    free(xofs); free(ialpha); free(row0); free(row1);
}

/* copyMakeBorder(..., BORDER_REFLECT_101) in place: `plane` points at the padded buffer origin,
 * the image occupies [border, border+rows) x [border, border+cols).  Index map -k -> k,
 * n-1+k -> n-1-k. */
void orc_border_reflect101(uint8_t* plane, int rows, int cols, size_t step, int border)
{
    for (int y = 0; y < rows; y++) {
        uint8_t* r = plane + (size_t)(y + border) * step + border;
        for (int k = 1; k <= border; k++) {
            r[-k] = r[k];
            r[cols - 1 + k] = r[cols - 1 - k];
        }
    }
    const size_t w = (size_t)cols + 2 * (size_t)border;
    for (int k = 1; k <= border; k++) {
        memcpy(plane + (size_t)(border - k) * step, plane + (size_t)(border + k) * step, w);
        memcpy(plane + (size_t)(border + rows - 1 + k) * step, plane + (size_t)(border + rows - 1 - k) * step, w);
    }
}

/* FAST-9/16 (features2d/src/fast.cpp FAST_t<16>, fast_score.cpp cornerScore<16>). */
static const int FAST_DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int FAST_DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* returns 0 for a non-corner at `threshold`, else the OpenCV corner score (>= threshold). */
static int fast_score_px(const uint8_t* p, const int* off, int threshold)
{
    int d[25];
    int v = p[0];
    {   /* exact early-out (same idea as OpenCV's pair tests): a 9-arc contains one pixel of
           every opposite pair (k, k+8), so each pair must hold a brighter (or darker) pixel. */
        const int vb = v + threshold, vd = v - threshold;
        int okb = 1, okd = 1;
        for (int k = 0; k < 8; k++) {
            int a = p[off[k]], b = p[off[k + 8]];
            okb &= (a > vb) | (b > vb);
            okd &= (a < vd) | (b < vd);
            if (!(okb | okd)) return 0;
        }
    }
    for (int k = 0; k < 16; k++) d[k] = v - p[off[k]];
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int A = -256, B = 256; /* A = max over arcs of min d ; B = min over arcs of max d */
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; j++) {
            if (d[k + j] < mn) mn = d[k + j];
            if (d[k + j] > mx) mx = d[k + j];
        }
        if (mn > A) A = mn;
        if (mx < B) B = mx;
    }
    if (!(A > threshold || B < -threshold)) return 0;
    int s = threshold;
    if (A > s) s = A;
    if (-B > s) s = -B;
    return s - 1;
}

int orc_fast9_16(const uint8_t* img, int rows, int cols, size_t step, int threshold,
                 int nonmax, orc_cand* out, int cap)
{
    if (rows < 7 || cols < 7) return 0;
    int off[16];
    for (int k = 0; k < 16; k++) off[k] = FAST_DY[k] * (int)step + FAST_DX[k];
    int* sc = (int*)calloc((size_t)rows * cols, sizeof(int));
    for (int y = 3; y < rows - 3; y++)
        for (int x = 3; x < cols - 3; x++)
            sc[(size_t)y * cols + x] = fast_score_px(img + (size_t)y * step + x, off, threshold);
    int n = 0;
    for (int y = 3; y < rows - 3; y++)
        for (int x = 3; x < cols - 3; x++) {
            int s = sc[(size_t)y * cols + x];
            if (s == 0) continue;
            if (nonmax) {
                const int* c = sc + (size_t)y * cols + x;
                if (!(s > c[-1] && s > c[1] && s > c[-cols - 1] && s > c[-cols] && s > c[-cols + 1] &&
                      s > c[cols - 1] && s > c[cols] && s > c[cols + 1]))
                    continue;
            }
            if (n < cap) { out[n].x = x; out[n].y = y; out[n].score = s; }
            n++;
        }
    free(sc);
    return n;
}

/*
 * cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 in OpenCV 4.x:
 * fixed-point path (imgproc/src/smooth.dispatch.cpp / fixedpoint: ufixedpoint16 8.8 kernel
 * [18,34,48,56,48,34,18]/256, horizontal pass in 8.8, vertical in 16.16, round by +0.5).
 */
static inline int reflect101(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

void orc_gaussian7x7_s2(const uint8_t* src, int rows, int cols, size_t sstep, uint8_t* dst, size_t dstep)
{
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    uint16_t* H = (uint16_t*)malloc(sizeof(uint16_t) * (size_t)rows * cols);
    for (int y = 0; y < rows; y++) {
        const uint8_t* S = src + (size_t)y * sstep;
        uint16_t* h = H + (size_t)y * cols;
        for (int x = 0; x < cols; x++) {
            int acc = 0;
            if (x >= 3 && x < cols - 3)
                for (int k = 0; k < 7; k++) acc += K[k] * S[x + k - 3];
            else
                for (int k = 0; k < 7; k++) acc += K[k] * S[reflect101(x + k - 3, cols)];
            h[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < rows; y++) {
        const uint16_t* r[7];
        for (int k = 0; k < 7; k++) r[k] = H + (size_t)reflect101(y + k - 3, rows) * cols;
        uint8_t* D = dst + (size_t)y * dstep;
        for (int x = 0; x < cols; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)K[k] * r[k][x];
            acc = (acc + 32768u) >> 16;
            D[x] = (uint8_t)(acc > 255 ? 255 : acc);
        }
    }
    free(H);
}

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2911-2931): 8 x 32-bit SWAR popcount. */
int orc_hamming256(const void* a, const void* b)
{
    uint32_t pa[8], pb[8];
    memcpy(pa, a, 32);
    memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
    }
    return dist;
}

/* cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2) as used at src/Frame.cc:1458: the two smallest
 * distances per query, ties resolved by ascending train index. */
void orc_bf_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* out_idx, int32_t* out_dist)
{
    for (int i = 0; i < nq; i++) {
        int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
        for (int j = 0; j < nt; j++) {
            int d = orc_hamming256(q + 32 * (size_t)i, t + 32 * (size_t)j);
            if (d < b0) { b1 = b0; i1 = i0; b0 = d; i0 = j; }
            else if (d < b1) { b1 = d; i1 = j; }
        }
        out_idx[2 * i] = i0; out_idx[2 * i + 1] = i1;
        out_dist[2 * i] = i0 >= 0 ? b0 : -1; out_dist[2 * i + 1] = i1 >= 0 ? b1 : -1;
    }
}
