/*
 * orc_detect.c — ORACLE (test infrastructure): ComputeKeyPointsOctTree, IC_Angle,
 * computeOrbDescriptor and ORBextractor::operator() of reference src/ORBextractor.cc.
 * Build with -ffp-contract=off.
 */
#include "orc_internal.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* IC_Angle, src/ORBextractor.cc:91-138.  `center` points at the keypoint pixel of the level image. */
float orc_ic_angle(const uint8_t* center, size_t step_, const int* u_max)
{
    int m_01 = 0, m_10 = 0;
    const int step = (int)step_;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];   /* :103-105 */
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {                                        /* :110-133 */
        int v_sum = 0;
        int d = u_max[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return orc_fast_atan2((float)m_01, (float)m_10);                                    /* :137 */
}

/* computeOrbDescriptor, src/ORBextractor.cc:150-203.  `center` points into the BLURRED level image. */
void orc_orb_descriptor(const uint8_t* center, size_t step_, float angle_deg, int trig_mode, uint8_t* desc)
{
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);          /* :141 */
    float angle = (float)angle_deg * factorPI;                                          /* :153 */
    float a, b;                                                                         /* :155 */
    if (trig_mode == ORC_TRIG_LIBM) { a = (float)cosf(angle); b = (float)sinf(angle); }
    else { a = orc_cosf(angle); b = orc_sinf(angle); }
    const int step = (int)step_;
    const signed char* pat = ORC_PATTERN;
#define GET_VALUE(idx) \
    center[orc_cvround_f(pat[2 * (idx)] * b + pat[2 * (idx) + 1] * a) * step + \
           orc_cvround_f(pat[2 * (idx)] * a - pat[2 * (idx) + 1] * b)]
    for (int i = 0; i < 32; ++i, pat += 32) {                                           /* :176-198 */
        int val = 0;
        for (int k = 0; k < 8; k++) {
            int t0 = GET_VALUE(2 * k), t1 = GET_VALUE(2 * k + 1);
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
#undef GET_VALUE
}

/* ComputeKeyPointsOctTree cell loop, src/ORBextractor.cc:1038-1143: fills e->cand[level]. */
static int detect_level(orc_extractor* e, int level)
{
    const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;                  /* :1053-1054 */
    const int maxBorderX = e->cols[level] - EDGE_THRESHOLD + 3;
    const int maxBorderY = e->rows[level] - EDGE_THRESHOLD + 3;
    const float W = 35;
    const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);                       /* :1064-1065 */
    if (nCols < 1 || nRows < 1) return -1;  /* reference divides by zero here */
    const int wCell = (int)ceilf(width / nCols), hCell = (int)ceilf(height / nRows);     /* :1067-1068 */
    size_t cap = 4096, n = 0;
    orc_cand* out = (orc_cand*)malloc(sizeof(orc_cand) * cap);
    const int cellcap = (wCell + 6) * (hCell + 6);
    orc_cand* cell = (orc_cand*)malloc(sizeof(orc_cand) * (size_t)cellcap);
    const uint8_t* img = e->plane[level] + EDGE_THRESHOLD * e->step[level] + EDGE_THRESHOLD;
    for (int i = 0; i < nRows; i++) {                                                    /* :1071 */
        const float iniY = (float)(minBorderY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBorderY - 3) continue;                                            /* :1085 */
        if (maxY > maxBorderY) maxY = (float)maxBorderY;
        for (int j = 0; j < nCols; j++) {                                                /* :1093 */
            const float iniX = (float)(minBorderX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBorderX - 6) continue;                                        /* :1102 */
            if (maxX > maxBorderX) maxX = (float)maxBorderX;
            const int y0 = (int)iniY, y1 = (int)maxY, x0 = (int)iniX, x1 = (int)maxX;
            const uint8_t* sub = img + (size_t)y0 * e->step[level] + x0;
            int nc = orc_fast9_16(sub, y1 - y0, x1 - x0, e->step[level], e->iniTh, 1, cell, cellcap);  /* :1112 */
            if (nc == 0)
                nc = orc_fast9_16(sub, y1 - y0, x1 - x0, e->step[level], e->minTh, 1, cell, cellcap);  /* :1121 */
            for (int k = 0; k < nc; k++) {                                               /* :1131-1140 */
                if (n == cap) { cap *= 2; out = (orc_cand*)realloc(out, sizeof(orc_cand) * cap); }
                out[n].x = cell[k].x + j * wCell;
                out[n].y = cell[k].y + i * hCell;
                out[n].score = cell[k].score;
                n++;
            }
        }
    }
    free(cell);
    e->cand[level] = out;
    e->ncand[level] = (int)n;
    return 0;
}

/* ORBextractor::operator(), src/ORBextractor.cc:1534-1659 (with ComputeKeyPointsOctTree :1038-1185). */
int orc_extract(orc_extractor* e, const uint8_t* img, int rows, int cols, size_t step,
                int lap0, int lap1, orc_keypoint* kps, uint8_t* desc, int cap, int* nkp)
{
    if (nkp) *nkp = 0;
    if (!img || rows <= 0 || cols <= 0) return -1;                                       /* :1538 */
    orc__free_levels(e);
    orc__compute_pyramid(e, img, rows, cols, step);                                      /* :1548 */
    const int minBorder = EDGE_THRESHOLD - 3;
    int total = 0;
    for (int level = 0; level < e->nlevels; ++level) {                                   /* :1048-1177 */
        if (detect_level(e, level) != 0) return -3;
        const int maxBorderX = e->cols[level] - EDGE_THRESHOLD + 3;
        const int maxBorderY = e->rows[level] - EDGE_THRESHOLD + 3;
        int selcap = e->ncand[level] + 8;
        e->sel[level] = (orc_cand*)malloc(sizeof(orc_cand) * (size_t)selcap);
        int ns = orc_distribute_octree(e->cand[level], e->ncand[level], minBorder, maxBorderX, minBorder,
                                       maxBorderY, e->nper[level], e->tie_mode, e->sel[level], selcap); /* :1150 */
        if (ns < 0) return -3;
        e->nsel[level] = ns;
        e->ang[level] = (float*)malloc(sizeof(float) * (size_t)(ns > 0 ? ns : 1));
        total += ns;
    }
    for (int level = 0; level < e->nlevels; ++level) {                                   /* :1181-1184 */
        const uint8_t* roi = e->plane[level] + EDGE_THRESHOLD * e->step[level] + EDGE_THRESHOLD;
        for (int i = 0; i < e->nsel[level]; i++) {
            /* pt += minBorder (:1167-1168); integral, so cvRound(pt) is the identity (:97) */
            int x = e->sel[level][i].x + minBorder, y = e->sel[level][i].y + minBorder;
            e->ang[level][i] = orc_ic_angle(roi + (size_t)y * e->step[level] + x, e->step[level], e->umax);
        }
    }
    if (nkp) *nkp = total;
    if (total > cap) return -2;
    int monoIndex = 0, stereoIndex = total - 1;                                          /* :1589 */
    uint8_t d[32];
    for (int level = 0; level < e->nlevels; ++level) {
        const int n = e->nsel[level];
        if (n == 0) continue;                                                            /* :1595 */
        const int h = e->rows[level], w = e->cols[level];
        const uint8_t* roi = e->plane[level] + EDGE_THRESHOLD * e->step[level] + EDGE_THRESHOLD;
        e->blurred[level] = (uint8_t*)malloc((size_t)h * w);
        orc_gaussian7x7_s2(roi, h, w, e->step[level], e->blurred[level], (size_t)w);    /* :1606-1614 */
        const float scale = e->sf[level];                                                /* :1628 */
        const int scaledPatchSize = (int)(PATCH_SIZE * e->sf[level]);                    /* :1161 */
        for (int i = 0; i < n; i++) {
            int xl = e->sel[level][i].x + minBorder, yl = e->sel[level][i].y + minBorder;
            orc_orb_descriptor(e->blurred[level] + (size_t)yl * w + xl, (size_t)w, e->ang[level][i], e->trig_mode, d); /* :1621 */
            orc_keypoint kp;
            kp.x = (float)xl; kp.y = (float)yl;
            kp.size = (float)scaledPatchSize;
            kp.angle = e->ang[level][i];
            kp.response = (float)e->sel[level][i].score;
            kp.octave = level;
            kp.class_id = -1;
            if (level != 0) { kp.x *= scale; kp.y *= scale; }                            /* :1639-1642 */
            int dst;
            if (kp.x >= lap0 && kp.x <= lap1) dst = stereoIndex--;                       /* :1644-1648 */
            else dst = monoIndex++;                                                      /* :1649-1653 */
            kps[dst] = kp;
            memcpy(desc + 32 * (size_t)dst, d, 32);
        }
    }
    return monoIndex;
}
