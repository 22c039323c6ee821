/*
 * orc_extractor.c — ORACLE (test infrastructure): plain-C restatement of ORB_SLAM3::ORBextractor
 * (reference src/ORBextractor.cc, include/ORBextractor.h).  Every function cites the lines it follows.
 * Build with -ffp-contract=off (canonical float semantics, see orc.h).
 */
#include "orc.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "orc_internal.h"

const signed char ORC_PATTERN[1024] = {
#include "../include/orb_pattern_31.inc"
};

/* ORBextractor::ORBextractor, src/ORBextractor.cc:468-571 */
orc_extractor* orc_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
{
    if (nlevels < 1 || nlevels > ORC_MAX_LEVELS) return NULL;
    orc_extractor* e = (orc_extractor*)calloc(1, sizeof(*e));
    e->nfeatures = nfeatures; e->nlevels = nlevels; e->iniTh = iniThFAST; e->minTh = minThFAST;
    e->scaleFactor = (double)scaleFactor;
    e->sf[0] = 1.0f; e->sig2[0] = 1.0f;                                   /* :485-486 */
    for (int i = 1; i < nlevels; i++) {
        e->sf[i] = (float)((double)e->sf[i - 1] * e->scaleFactor);         /* :488 float*double */
        e->sig2[i] = e->sf[i] * e->sf[i];                                  /* :490 */
    }
    for (int i = 0; i < nlevels; i++) {                                    /* :494-500 */
        e->isf[i] = 1.0f / e->sf[i];
        e->isig2[i] = 1.0f / e->sig2[i];
    }
    float factor = (float)(1.0f / e->scaleFactor);                         /* :509 */
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels)); /* :511 */
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {                    /* :516-524 */
        e->nper[level] = orc_cvround_f(nDesired);
        sum += e->nper[level];
        nDesired *= factor;
    }
    e->nper[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0;      /* :526 */
    /* umax, :542-570 */
    int vmax = (int)floor(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
    int vmin = (int)ceil(HALF_PATCH_SIZE * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (int v = 0; v <= vmax; ++v) e->umax[v] = orc_cvround_d(sqrt(hp2 - v * v));
    for (int v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
    return e;
}

void orc__free_levels(orc_extractor* e)
{
    for (int l = 0; l < ORC_MAX_LEVELS; l++) {
        free(e->plane[l]); free(e->blurred[l]); free(e->cand[l]); free(e->sel[l]); free(e->ang[l]);
        e->plane[l] = e->blurred[l] = NULL; e->cand[l] = e->sel[l] = NULL; e->ang[l] = NULL;
        e->ncand[l] = e->nsel[l] = 0;
    }
}

void orc_destroy(orc_extractor* e) { if (e) { orc__free_levels(e); free(e); } }
void orc_set_modes(orc_extractor* e, int tie_mode, int trig_mode) { e->tie_mode = tie_mode; e->trig_mode = trig_mode; }
int orc_levels(const orc_extractor* e) { return e->nlevels; }
const float* orc_scale_factors(const orc_extractor* e) { return e->sf; }
const float* orc_inv_scale_factors(const orc_extractor* e) { return e->isf; }
const float* orc_level_sigma2(const orc_extractor* e) { return e->sig2; }
const float* orc_inv_level_sigma2(const orc_extractor* e) { return e->isig2; }
const int* orc_features_per_level(const orc_extractor* e) { return e->nper; }
const int* orc_umax(const orc_extractor* e) { return e->umax; }
void orc_level_dims(const orc_extractor* e, int l, int* rows, int* cols) { *rows = e->rows[l]; *cols = e->cols[l]; }
const uint8_t* orc_level_plane(const orc_extractor* e, int l, size_t* step) { *step = e->step[l]; return e->plane[l]; }
const uint8_t* orc_level_blurred(const orc_extractor* e, int l, size_t* step) { *step = (size_t)e->cols[l]; return e->blurred[l]; }
int orc_level_candidates(const orc_extractor* e, int l, const orc_cand** out) { *out = e->cand[l]; return e->ncand[l]; }
int orc_level_selected(const orc_extractor* e, int l, const orc_cand** out, const float** angles)
{
    *out = e->sel[l];
    if (angles) *angles = e->ang[l];
    return e->nsel[l];
}

/* ORBextractor::ComputePyramid, src/ORBextractor.cc:1664-1715 */
void orc__compute_pyramid(orc_extractor* e, const uint8_t* img, int rows, int cols, size_t step)
{
    for (int level = 0; level < e->nlevels; ++level) {
        float scale = e->isf[level];
        int w = orc_cvround_f((float)cols * scale), h = orc_cvround_f((float)rows * scale);  /* :1669 */
        e->rows[level] = h; e->cols[level] = w;
        e->step[level] = (size_t)(w + 2 * EDGE_THRESHOLD);
        e->plane[level] = (uint8_t*)malloc(e->step[level] * (size_t)(h + 2 * EDGE_THRESHOLD));
        uint8_t* roi = e->plane[level] + EDGE_THRESHOLD * e->step[level] + EDGE_THRESHOLD;
        if (level != 0) {                                                  /* :1679-1693 */
            const uint8_t* prev = e->plane[level - 1] + EDGE_THRESHOLD * e->step[level - 1] + EDGE_THRESHOLD;
            orc_resize_linear_u8(prev, e->rows[level - 1], e->cols[level - 1], e->step[level - 1],
                                 roi, h, w, e->step[level]);
        } else {                                                           /* :1711-1713 */
            for (int y = 0; y < h; y++) memcpy(roi + (size_t)y * e->step[0], img + (size_t)y * step, (size_t)w);
        }
        orc_border_reflect101(e->plane[level], h, w, e->step[level], EDGE_THRESHOLD);
    }
}
