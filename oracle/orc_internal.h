/* orc_internal.h — ORACLE (test infrastructure): shared state of the extractor restatement. */
#ifndef ORC_INTERNAL_H
#define ORC_INTERNAL_H
#include "orc.h"
#define PATCH_SIZE 31       /* src/ORBextractor.cc:76 */
#define HALF_PATCH_SIZE 15  /* :77 */
#define EDGE_THRESHOLD 19   /* :78 */
#define ORC_MAX_LEVELS 32

extern const signed char ORC_PATTERN[1024];

struct orc_extractor {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor; /* include/ORBextractor.h:96: member is double, ctor argument float */
    float sf[ORC_MAX_LEVELS], isf[ORC_MAX_LEVELS], sig2[ORC_MAX_LEVELS], isig2[ORC_MAX_LEVELS];
    int nper[ORC_MAX_LEVELS];
    int umax[HALF_PATCH_SIZE + 1];
    int tie_mode, trig_mode;
    /* per-level state of the last orc_extract call (mvImagePyramid + stage dumps) */
    int rows[ORC_MAX_LEVELS], cols[ORC_MAX_LEVELS];
    uint8_t* plane[ORC_MAX_LEVELS];
    size_t step[ORC_MAX_LEVELS];
    uint8_t* blurred[ORC_MAX_LEVELS];
    orc_cand* cand[ORC_MAX_LEVELS];
    int ncand[ORC_MAX_LEVELS];
    orc_cand* sel[ORC_MAX_LEVELS];
    float* ang[ORC_MAX_LEVELS];
    int nsel[ORC_MAX_LEVELS];
};


void orc__compute_pyramid(orc_extractor* e, const uint8_t* img, int rows, int cols, size_t step);
void orc__free_levels(orc_extractor* e);
#endif
