/*
 * orc_match.c — ORACLE (test infrastructure): Frame::ComputeStereoMatches (src/Frame.cc:1027-1276),
 * the 64x48 frame grid (include/Frame.h:37-38, src/Frame.cc:446-480, 784-899) and
 * ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) (src/ORBmatcher.cc:45-249).
 * Build with -ffp-contract=off.
 */
#include "orc_internal.h"
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define TH_HIGH 100 /* src/ORBmatcher.cc:36 */
#define TH_LOW 50   /* :37 */

typedef struct { int dist, idx; } distidx;
static int cmp_distidx(const void* a, const void* b)
{
    const distidx *p = (const distidx*)a, *q = (const distidx*)b;
    if (p->dist != q->dist) return p->dist < q->dist ? -1 : 1;
    return p->idx < q->idx ? -1 : (p->idx > q->idx ? 1 : 0);
}

void orc_stereo_match(const orc_extractor* eL, const orc_extractor* eR,
                      const orc_keypoint* kpL, int N, const uint8_t* descL,
                      const orc_keypoint* kpR, int Nr, const uint8_t* descR,
                      float mb, float mbf, float* uRight, float* depth,
                      int32_t* best_dist_out, int32_t* best_sad_out)
{
    for (int i = 0; i < N; i++) {                                             /* :1041-1042 */
        uRight[i] = -1.0f; depth[i] = -1.0f;
        if (best_dist_out) best_dist_out[i] = -1;
        if (best_sad_out) best_sad_out[i] = -1;
    }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;                             /* :1045 */
    const int nRows = eL->rows[0];                                            /* :1048 */
    const float* sf = eL->sf; const float* isf = eL->isf;
    /* row buckets :1051-1078 (counting sort into CSR, insertion order kept) */
    int* cnt = (int*)calloc((size_t)nRows + 1, sizeof(int));
    for (int pass = 0; pass < 1; pass++)
        for (int iR = 0; iR < Nr; iR++) {
            const float kpY = kpR[iR].y;
            const float r = 2.0f * sf[kpR[iR].octave];
            const int maxr = (int)ceilf(kpY + r), minr = (int)floorf(kpY - r);
            for (int yi = minr; yi <= maxr; yi++) if (yi >= 0 && yi < nRows) cnt[yi + 1]++;
        }
    for (int y = 0; y < nRows; y++) cnt[y + 1] += cnt[y];
    int* fill = (int*)malloc(sizeof(int) * (size_t)(nRows + 1));
    memcpy(fill, cnt, sizeof(int) * (size_t)(nRows + 1));
    int* rowidx = (int*)malloc(sizeof(int) * (size_t)(cnt[nRows] > 0 ? cnt[nRows] : 1));
    for (int iR = 0; iR < Nr; iR++) {
        const float kpY = kpR[iR].y;
        const float r = 2.0f * sf[kpR[iR].octave];
        const int maxr = (int)ceilf(kpY + r), minr = (int)floorf(kpY - r);
        for (int yi = minr; yi <= maxr; yi++) if (yi >= 0 && yi < nRows) rowidx[fill[yi]++] = iR;
    }
    const float minZ = mb, minD = 0;                                          /* :1081-1083 */
    const float maxD = mbf / minZ;
    distidx* vDistIdx = (distidx*)malloc(sizeof(distidx) * (size_t)(N > 0 ? N : 1));
    int nd = 0;
    for (int iL = 0; iL < N; iL++) {                                          /* :1090 */
        const int levelL = kpL[iL].octave;
        const float vL = kpL[iL].y, uL = kpL[iL].x;
        const int row = (int)vL;                                              /* :1097 index by float */
        if (row < 0 || row >= nRows) continue;
        const int c0 = cnt[row], c1 = cnt[row + 1];
        if (c0 == c1) continue;                                               /* :1099 */
        const float minU = uL - maxD, maxU = uL - minD;                       /* :1102-1103 */
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;                                               /* :1108 */
        int bestIdxR = 0;
        const uint8_t* dL = descL + 32 * (size_t)iL;
        for (int c = c0; c < c1; c++) {                                       /* :1114-1135 */
            const int iR = rowidx[c];
            if (kpR[iR].octave < levelL - 1 || kpR[iR].octave > levelL + 1) continue;
            const float uR = kpR[iR].x;
            if (uR >= minU && uR <= maxU) {
                const int dist = orc_hamming256(dL, descR + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (best_dist_out) best_dist_out[iL] = bestDist;
        if (bestDist < thOrbDist) {                                           /* :1139 */
            const float uR0 = kpR[bestIdxR].x;
            const float scaleFactor = isf[levelL];
            const float scaleduL = roundf(kpL[iL].x * scaleFactor);           /* :1144-1146 */
            const float scaledvL = roundf(kpL[iL].y * scaleFactor);
            const float scaleduR0 = roundf(uR0 * scaleFactor);
            const int w = 5, L = 5;
            const uint8_t* PL = eL->plane[levelL] + EDGE_THRESHOLD * eL->step[levelL] + EDGE_THRESHOLD;
            const uint8_t* PR = eR->plane[levelL] + EDGE_THRESHOLD * eR->step[levelL] + EDGE_THRESHOLD;
            const size_t sL = eL->step[levelL], sR = eR->step[levelL];
            int bestSad = INT_MAX, bestincR = 0;
            float vDists[11];
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1; /* :1162-1163 */
            if (iniu < 0 || endu >= eR->cols[levelL]) continue;               /* :1164 */
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
            for (int incR = -L; incR <= +L; incR++) {                         /* :1167-1180 */
                const int xr0 = (int)(scaleduR0 + incR - w);
                int sad = 0;
                for (int yy = 0; yy < 2 * w + 1; yy++) {
                    const uint8_t* a = PL + (ptrdiff_t)(y0 + yy) * (ptrdiff_t)sL + xl0;
                    const uint8_t* b = PR + (ptrdiff_t)(y0 + yy) * (ptrdiff_t)sR + xr0;
                    for (int xx = 0; xx < 2 * w + 1; xx++) sad += abs((int)a[xx] - (int)b[xx]);
                }
                float dist = (float)sad;                                      /* cv::norm(IL,IR,NORM_L1) */
                if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (best_sad_out) best_sad_out[iL] = bestSad;
            if (bestincR == -L || bestincR == L) continue;                    /* :1182 */
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2)); /* :1190 */
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = sf[levelL] * ((float)scaleduR0 + (float)bestincR + deltaR);      /* :1196 */
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {                      /* :1200 */
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                vDistIdx[nd].dist = bestSad; vDistIdx[nd].idx = iL; nd++;
            }
        }
    }
    if (nd > 0) {                                                             /* :1261-1275 (guarded: reference UB when empty) */
        qsort(vDistIdx, (size_t)nd, sizeof(distidx), cmp_distidx);
        const float median = (float)vDistIdx[nd / 2].dist;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = nd - 1; i >= 0; i--) {
            if (vDistIdx[i].dist < thDist) break;
            uRight[vDistIdx[i].idx] = -1; depth[vDistIdx[i].idx] = -1;
        }
    }
    free(vDistIdx); free(cnt); free(fill); free(rowidx);
}

/* ---- frame grid ---- */
#define GRID_COLS 64 /* include/Frame.h:38 */
#define GRID_ROWS 48 /* include/Frame.h:37 */
struct orc_grid {
    float minX, minY, winv, hinv;
    int start[GRID_COLS * GRID_ROWS + 1]; /* cell id = ix*GRID_ROWS + iy (mGrid[ix][iy]) */
    int* idx;
};

/* Frame::PosInGrid, src/Frame.cc:883-899 */
static int pos_in_grid(const orc_grid* g, const orc_keypoint* kp, int* px, int* py)
{
    *px = (int)roundf((kp->x - g->minX) * g->winv);
    *py = (int)roundf((kp->y - g->minY) * g->hinv);
    if (*px < 0 || *px >= GRID_COLS || *py < 0 || *py >= GRID_ROWS) return 0;
    return 1;
}

/* Frame::AssignFeaturesToGrid, src/Frame.cc:446-480 (mono / rectified stereo: Nleft == -1) */
orc_grid* orc_grid_build(const orc_keypoint* kps, int n, float minX, float maxX, float minY, float maxY)
{
    orc_grid* g = (orc_grid*)calloc(1, sizeof(orc_grid));
    g->minX = minX; g->minY = minY;
    g->winv = (float)GRID_COLS / (float)(maxX - minX);                        /* src/Frame.cc:292 */
    g->hinv = (float)GRID_ROWS / (float)(maxY - minY);                        /* :294 */
    g->idx = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    int px, py;
    for (int i = 0; i < n; i++) if (pos_in_grid(g, &kps[i], &px, &py)) g->start[px * GRID_ROWS + py + 1]++;
    for (int c = 0; c < GRID_COLS * GRID_ROWS; c++) g->start[c + 1] += g->start[c];
    int* fill = (int*)malloc(sizeof(int) * GRID_COLS * GRID_ROWS);
    memcpy(fill, g->start, sizeof(int) * GRID_COLS * GRID_ROWS);
    for (int i = 0; i < n; i++) if (pos_in_grid(g, &kps[i], &px, &py)) g->idx[fill[px * GRID_ROWS + py]++] = i;
    free(fill);
    return g;
}
void orc_grid_destroy(orc_grid* g) { if (g) { free(g->idx); free(g); } }

/* Frame::GetFeaturesInArea, src/Frame.cc:784-871 */
int orc_grid_query(const orc_grid* g, const orc_keypoint* kps, float x, float y, float r,
                   int minLevel, int maxLevel, int32_t* out, int cap)
{
    int n = 0;
    const float factorX = r, factorY = r;
    int nMinCellX = (int)floorf((x - g->minX - factorX) * g->winv); if (nMinCellX < 0) nMinCellX = 0;
    if (nMinCellX >= GRID_COLS) return 0;
    int nMaxCellX = (int)ceilf((x - g->minX + factorX) * g->winv); if (nMaxCellX > GRID_COLS - 1) nMaxCellX = GRID_COLS - 1;
    if (nMaxCellX < 0) return 0;
    int nMinCellY = (int)floorf((y - g->minY - factorY) * g->hinv); if (nMinCellY < 0) nMinCellY = 0;
    if (nMinCellY >= GRID_ROWS) return 0;
    int nMaxCellY = (int)ceilf((y - g->minY + factorY) * g->hinv); if (nMaxCellY > GRID_ROWS - 1) nMaxCellY = GRID_ROWS - 1;
    if (nMaxCellY < 0) return 0;
    const int bCheckLevels = (minLevel > 0) || (maxLevel >= 0);               /* :828 */
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const int c = ix * GRID_ROWS + iy;
            for (int j = g->start[c]; j < g->start[c + 1]; j++) {
                const orc_keypoint* kp = &kps[g->idx[j]];
                if (bCheckLevels) {
                    if (kp->octave < minLevel) continue;
                    if (maxLevel >= 0 && kp->octave > maxLevel) continue;
                }
                const float distx = kp->x - x, disty = kp->y - y;
                if (fabsf(distx) < factorX && fabsf(disty) < factorY) {
                    if (n < cap) out[n] = g->idx[j];
                    n++;
                }
            }
        }
    return n;
}

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints),
 * src/ORBmatcher.cc:45-168, branch F.Nleft == -1 (mono / rectified stereo). */
int orc_search_by_projection(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n,
                             const float* uRight, const float* sf,
                             int nmp, const uint8_t* mp_desc, const float* mp_projx, const float* mp_projy,
                             const float* mp_projxr, const int32_t* mp_level, const float* mp_viewcos,
                             const float* mp_depth, const uint8_t* mp_inview,
                             float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    int nmatches = 0;
    const int bFactor = th != 1.0;                                            /* :49 */
    int32_t* vIndices = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int iMP = 0; iMP < nmp; iMP++) {
        if (!mp_inview[iMP]) continue;                                        /* :54 */
        if (bFarPoints && mp_depth[iMP] > thFarPoints) continue;              /* :57 */
        const int nPredictedLevel = mp_level[iMP];
        float r = mp_viewcos[iMP] > 0.998 ? 2.5f : 4.0f;                      /* RadiusByViewingCos :242-249 */
        if (bFactor) r *= th;
        const int nc = orc_grid_query(g, kps, mp_projx[iMP], mp_projy[iMP], r * sf[nPredictedLevel],
                                      nPredictedLevel - 1, nPredictedLevel, vIndices, n); /* :84-87 */
        if (nc == 0) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int k = 0; k < nc; k++) {                                        /* :100-141 */
            const int idx = vIndices[k];
            if (assign[idx] >= 0) continue;                                   /* :102-104 */
            if (uRight && uRight[idx] > 0) {                                  /* :107-117 */
                const float er = fabsf(mp_projxr[iMP] - uRight[idx]);
                if (er > r * sf[nPredictedLevel]) continue;
            }
            const int dist = orc_hamming256(mp_desc + 32 * (size_t)iMP, desc + 32 * (size_t)idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist;
                bestLevel2 = bestLevel; bestLevel = kps[idx].octave;
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = kps[idx].octave; bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {                                            /* :147-165 */
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= nnratio * bestDist2) {
                assign[bestIdx] = iMP;
                nmatches++;
            }
        }
    }
    free(vIndices);
    return nmatches;
}
