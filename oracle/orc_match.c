/*
 * orc_match.c — ORACLE (test infrastructure): Frame::ComputeStereoMatches (src/Frame.cc:1027-1276),
 * the 64x48 frame grid (include/Frame.h:37-38, src/Frame.cc:446-480, 784-899) and
 * ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) (src/ORBmatcher.cc:45-249),
 * SearchByProjection(Frame&, const Frame&, th, bMono) (:2469-2711), SearchForInitialization (:799-943),
 * SearchForTriangulation / SearchForTriangulation_ (:1107-1630) and ComputeThreeMaxima (:2863-2905).
 * Build with -ffp-contract=off.
 */
#include "orc_internal.h"
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define TH_HIGH 100 /* src/ORBmatcher.cc:36 */
#define TH_LOW 50   /* :37 */
#define HISTO_LENGTH 30 /* :38 */

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
    float minX, minY, maxX, maxY, winv, hinv;
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
    g->minX = minX; g->minY = minY; g->maxX = maxX; g->maxY = maxY;
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

/* ---- rotation histogram ---- */
typedef struct { int* v[HISTO_LENGTH]; int n[HISTO_LENGTH], cap[HISTO_LENGTH]; } rot_hist;
static void hist_init(rot_hist* h) { memset(h, 0, sizeof(*h)); }
static void hist_free(rot_hist* h) { for (int i = 0; i < HISTO_LENGTH; i++) free(h->v[i]); }
static void hist_push(rot_hist* h, int bin, int val)
{
    if (h->n[bin] == h->cap[bin]) {
        h->cap[bin] = h->cap[bin] ? 2 * h->cap[bin] : 64;
        h->v[bin] = (int*)realloc(h->v[bin], sizeof(int) * (size_t)h->cap[bin]);
    }
    h->v[bin][h->n[bin]++] = val;
}
/* the bin of one accepted pair, e.g. src/ORBmatcher.cc:895-906 (factor is each caller's own) */
static int hist_bin(float angle1, float angle2, float factor)
{
    float rot = angle1 - angle2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)roundf(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

/* ORBmatcher::ComputeThreeMaxima, src/ORBmatcher.cc:2863-2905 */
void orc_three_maxima(const int* sizes, int L, int* ind1, int* ind2, int* ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = sizes[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

/* ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono),
 * src/ORBmatcher.cc:2469-2711, branch CurrentFrame.Nleft == -1.  The projection of the last frame's map
 * points into the current camera (:2501-2520) is the caller's: last_u/last_v = uv, last_invz = invzc.
 * last_valid = pMP && !mvbOutlier; last_obs = pMP->Observations() > 0.  cur_held (may be NULL) marks
 * current keypoints that already hold a map point with observations; assign[n] out = index of the last
 * frame point each current keypoint ends with (-1 = none).  bForward/bBackward as computed at :2495-2496. */
int orc_search_by_projection_frame(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n,
                                   const float* uRight, const float* sf,
                                   int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                   const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                   const uint8_t* last_valid, const uint8_t* last_obs,
                                   float th, float mbf, int bForward, int bBackward, int checkOrientation,
                                   const uint8_t* cur_held, int32_t* assign)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = HISTO_LENGTH / 360.0f;                               /* :2480 */
    uint8_t* held = (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1);
    int32_t* vIndices2 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { assign[i] = -1; if (cur_held) held[i] = cur_held[i] != 0; }
    for (int i = 0; i < nlast; i++) {
        if (!last_valid[i]) continue;                                         /* :2503-2506 */
        const float invzc = last_invz[i];
        if (invzc < 0) continue;                                              /* :2515 */
        const float u = last_u[i], v = last_v[i];
        if (u < g->minX || u > g->maxX) continue;                             /* :2520-2523 */
        if (v < g->minY || v > g->maxY) continue;
        const int nLastOctave = last_octave[i];
        const float radius = th * sf[nLastOctave];                            /* :2529 */
        int nc;
        if (bForward) nc = orc_grid_query(g, kps, u, v, radius, nLastOctave, -1, vIndices2, n);
        else if (bBackward) nc = orc_grid_query(g, kps, u, v, radius, 0, nLastOctave, vIndices2, n);
        else nc = orc_grid_query(g, kps, u, v, radius, nLastOctave - 1, nLastOctave + 1, vIndices2, n);
        if (nc == 0) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {                                        /* :2548-2574 */
            const int i2 = vIndices2[k];
            if (held[i2]) continue;
            if (uRight && uRight[i2] > 0) {
                const float ur = u - mbf * invzc;
                const float er = fabsf(ur - uRight[i2]);
                if (er > radius) continue;
            }
            const int dist = orc_hamming256(last_desc + 32 * (size_t)i, desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {                                            /* :2576-2601 */
            assign[bestIdx2] = i;
            held[bestIdx2] = last_obs[i] != 0;
            nmatches++;
            if (checkOrientation) hist_push(&H, hist_bin(last_angle[i], kps[bestIdx2].angle, factor), bestIdx2);
        }
    }
    if (checkOrientation) {                                                   /* :2686-2708 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int b = 0; b < HISTO_LENGTH; b++)
            if (b != ind1 && b != ind2 && b != ind3)
                for (int j = 0; j < H.n[b]; j++) { assign[H.v[b][j]] = -1; nmatches--; }
    }
    hist_free(&H); free(held); free(vIndices2);
    return nmatches;
}

/* ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:799-943.  g2 is the grid of frame 2.
 * prev_matched[2*n1] (x,y) in/out (vbPrevMatched), matches12[n1] out.  Returns nmatches. */
int orc_search_for_initialization(const orc_keypoint* kps1, const uint8_t* desc1, int n1,
                                  const orc_grid* g2, const orc_keypoint* kps2, const uint8_t* desc2, int n2,
                                  float* prev_matched, int windowSize, float nnratio, int checkOrientation,
                                  int32_t* matches12)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = HISTO_LENGTH / 360.0f;                               /* :812 */
    const size_t m2 = (size_t)(n2 > 0 ? n2 : 1);
    int* vMatchedDistance = (int*)malloc(sizeof(int) * m2);
    int* vnMatches21 = (int*)malloc(sizeof(int) * m2);
    int32_t* vIndices2 = (int32_t*)malloc(sizeof(int32_t) * m2);
    for (int i = 0; i < n2; i++) { vMatchedDistance[i] = INT_MAX; vnMatches21[i] = -1; }
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = kps1[i1].octave;
        if (level1 > 0) continue;                                             /* :825 */
        const int nc = orc_grid_query(g2, kps2, prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)windowSize,
                                      level1, level1, vIndices2, n2);         /* :831 */
        if (nc == 0) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {                                        /* :845-866 */
            const int i2 = vIndices2[k];
            const int dist = orc_hamming256(desc1 + 32 * (size_t)i1, desc2 + 32 * (size_t)i2);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {                                             /* :870-908 */
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOrientation) hist_push(&H, hist_bin(kps1[i1].angle, kps2[bestIdx2].angle, factor), i1);
            }
        }
    }
    if (checkOrientation) {                                                   /* :914-935 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == ind1 || b == ind2 || b == ind3) continue;
            for (int j = 0; j < H.n[b]; j++) {
                const int idx1 = H.v[b][j];
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
            }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)                                           /* :939-941 */
        if (matches12[i1] >= 0) { prev_matched[2 * i1] = kps2[matches12[i1]].x; prev_matched[2 * i1 + 1] = kps2[matches12[i1]].y; }
    hist_free(&H); free(vMatchedDistance); free(vnMatches21); free(vIndices2);
    return nmatches;
}

/* Pinhole::epipolarConstrain, src/CameraModels/Pinhole.cpp:237-260, with F12 (row-major 3x3) formed by the
 * caller (K1^-T [t12]x R12 K2^-1, :243) */
static int epipolar_ok(const float* F12, float x1, float y1, float x2, float y2, float unc)
{
    const float a = x1 * F12[0] + y1 * F12[3] + F12[6];
    const float b = x1 * F12[1] + y1 * F12[4] + F12[7];
    const float c = x1 * F12[2] + y1 * F12[5] + F12[8];
    const float num = a * x2 + b * y2 + c;
    const float den = a * a + b * b;
    if (den == 0) return 0;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * unc;
}

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:1107-1386, variant 0) and SearchForTriangulation_
 * (:1388-1630, variant 1: factor 1/HISTO_LENGTH :1441 and no vbMatched2 bookkeeping), pinhole branch
 * (mpCamera2 == NULL).  FeatureVectors as CSR: fvN_node[nnN] ascending node ids, fvN_start[nnN+1],
 * fvN_idx[].  has_mpN = GetMapPoint(idx) != NULL; uRightN may be NULL (all < 0 = mono).  ep = epipole in
 * image 2 (:1120), F12 row-major.  matches12[n1] out (vMatchedPairs = the >=0 entries, ascending idx1). */
int orc_search_for_triangulation(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                 const float* uRight1,
                                 const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                 const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                 const float* uRight2,
                                 const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                 const float* sf2, const float* sigma2_2, const float* ep, const float* F12,
                                 int bOnlyStereo, int bCoarse, int checkOrientation, int variant,
                                 int32_t* matches12)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = variant ? 1.0f / HISTO_LENGTH : HISTO_LENGTH / 360.0f; /* :1166 / :1441 */
    uint8_t* vbMatched2 = (uint8_t*)calloc((size_t)(n2 > 0 ? n2 : 1), 1);
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {                                              /* :1176-1330 */
        if (fv1_node[a] < fv2_node[b]) { a++; continue; }                     /* lower_bound on sorted keys */
        if (fv1_node[a] > fv2_node[b]) { b++; continue; }
        for (int p1 = fv1_start[a]; p1 < fv1_start[a + 1]; p1++) {
            const int idx1 = fv1_idx[p1];
            if (has_mp1[idx1]) continue;                                      /* :1188-1192 */
            const int bStereo1 = uRight1 && uRight1[idx1] >= 0;               /* :1194 */
            if (bOnlyStereo && !bStereo1) continue;
            const orc_keypoint* kp1 = &kps1[idx1];
            int bestDist = TH_LOW, bestIdx2 = -1;
            for (int p2 = fv2_start[b]; p2 < fv2_start[b + 1]; p2++) {
                const int idx2 = fv2_idx[p2];
                if (vbMatched2[idx2] || has_mp2[idx2]) continue;              /* :1218 */
                const int bStereo2 = uRight2 && uRight2[idx2] >= 0;
                if (bOnlyStereo && !bStereo2) continue;
                const int dist = orc_hamming256(desc1 + 32 * (size_t)idx1, desc2 + 32 * (size_t)idx2);
                if (dist > TH_LOW || dist > bestDist) continue;               /* :1231 */
                const orc_keypoint* kp2 = &kps2[idx2];
                if (!bStereo1 && !bStereo2) {                                 /* :1241-1250 */
                    const float distex = ep[0] - kp2->x, distey = ep[1] - kp2->y;
                    if (distex * distex + distey * distey < 100 * sf2[kp2->octave]) continue;
                }
                if (epipolar_ok(F12, kp1->x, kp1->y, kp2->x, kp2->y, sigma2_2[kp2->octave]) || bCoarse) { /* :1282 */
                    bestIdx2 = idx2; bestDist = dist;
                }
            }
            if (bestIdx2 >= 0) {                                              /* :1289-1313 */
                matches12[idx1] = bestIdx2;
                if (!variant) vbMatched2[bestIdx2] = 1;
                nmatches++;
                if (checkOrientation) hist_push(&H, hist_bin(kp1->angle, kps2[bestIdx2].angle, factor), idx1);
            }
        }
        a++; b++;
    }
    if (checkOrientation) {                                                   /* :1332-1355 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int bi = 0; bi < HISTO_LENGTH; bi++) {
            if (bi == ind1 || bi == ind2 || bi == ind3) continue;
            for (int j = 0; j < H.n[bi]; j++) { matches12[H.v[bi][j]] = -1; nmatches--; }
        }
    }
    hist_free(&H); free(vbMatched2);
    return nmatches;
}

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (src/ORBmatcher.cc:314-547, F.Nleft == -1;
 * mode 0) and SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&) (:955-1105; mode 1).  valid1 / valid2 =
 * map point present and not bad (valid2 may be NULL in mode 0: a Frame keypoint only needs to be unmatched).
 * matches out: mode 0 -> [n2] side-1 index per Frame keypoint, mode 1 -> [n1] side-2 index per KF1 keypoint. */
int orc_search_by_bow(int mode,
                      const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                      const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                      const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* valid2,
                      const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                      float nnratio, int checkOrientation, int32_t* matches)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = HISTO_LENGTH / 360.0f;                               /* :334 / :978 */
    uint8_t* matched2 = (uint8_t*)calloc((size_t)(n2 > 0 ? n2 : 1), 1);
    const int nout = mode == 0 ? n2 : n1;
    for (int i = 0; i < nout; i++) matches[i] = -1;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (fv1_node[a] < fv2_node[b]) { a++; continue; }
        if (fv1_node[a] > fv2_node[b]) { b++; continue; }
        for (int p1 = fv1_start[a]; p1 < fv1_start[a + 1]; p1++) {
            const int idx1 = fv1_idx[p1];
            if (!valid1[idx1]) continue;                                      /* :343-349 / :1006-1011 */
            int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
            for (int p2 = fv2_start[b]; p2 < fv2_start[b + 1]; p2++) {
                const int idx2 = fv2_idx[p2];
                if (matched2[idx2]) continue;                                 /* :365 / :1038 */
                if (valid2 && !valid2[idx2]) continue;                        /* :1038-1045 */
                const int dist = orc_hamming256(desc1 + 32 * (size_t)idx1, desc2 + 32 * (size_t)idx2);
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                else if (dist < bestDist2) bestDist2 = dist;
            }
            const int low = mode == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;   /* :431 / :1064 */
            if (low && (float)bestDist1 < nnratio * (float)bestDist2) {       /* :433 / :1066 */
                matched2[bestIdx2] = 1;
                if (mode == 0) matches[bestIdx2] = idx1; else matches[idx1] = bestIdx2;
                if (checkOrientation)
                    hist_push(&H, hist_bin(kps1[idx1].angle, kps2[bestIdx2].angle, factor), mode == 0 ? bestIdx2 : idx1);
                nmatches++;
            }
        }
        a++; b++;
    }
    if (checkOrientation) {                                                   /* :520-544 / :1090-1110 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int bi = 0; bi < HISTO_LENGTH; bi++) {
            if (bi == ind1 || bi == ind2 || bi == ind3) continue;
            for (int j = 0; j < H.n[bi]; j++) { matches[H.v[bi][j]] = -1; nmatches--; }
        }
    }
    hist_free(&H); free(matched2);
    return nmatches;
}

/* ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th,
 * ORBdist), src/ORBmatcher.cc:2723-2862.  Rows = keyframe keypoints: mp_valid folds pMP && !isBad() &&
 * !sAlreadyFound.count(pMP) and the distance-invariance gate (:2768-2776); mp_level = PredictScale (:2778). */
int orc_search_by_projection_keyframe(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                                      int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                      const int32_t* mp_level, const float* mp_angle, const uint8_t* mp_valid,
                                      float th, int ORBdist, int checkOrientation, const uint8_t* cur_held, int32_t* assign)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = HISTO_LENGTH / 360.0f;                               /* :2735 */
    uint8_t* held = (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1);
    int32_t* vIndices2 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { assign[i] = -1; if (cur_held) held[i] = cur_held[i] != 0; }
    for (int i = 0; i < nmp; i++) {
        if (!mp_valid[i]) continue;
        const float u = mp_u[i], v = mp_v[i];
        if (u < g->minX || u > g->maxX) continue;                             /* :2761-2764 */
        if (v < g->minY || v > g->maxY) continue;
        const int lvl = mp_level[i];
        const float radius = th * sf[lvl];                                    /* :2781 */
        const int nc = orc_grid_query(g, kps, u, v, radius, lvl - 1, lvl + 1, vIndices2, n);   /* :2784 */
        if (nc == 0) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {                                        /* :2795-2811 */
            const int i2 = vIndices2[k];
            if (held[i2]) continue;
            const int dist = orc_hamming256(mp_desc + 32 * (size_t)i, desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= ORBdist) {                                            /* :2813-2831 */
            assign[bestIdx2] = i; held[bestIdx2] = 1;
            nmatches++;
            if (checkOrientation) hist_push(&H, hist_bin(mp_angle[i], kps[bestIdx2].angle, factor), bestIdx2);
        }
    }
    if (checkOrientation) {                                                   /* :2834-2857 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int b = 0; b < HISTO_LENGTH; b++)
            if (b != ind1 && b != ind2 && b != ind3)
                for (int j = 0; j < H.n[b]; j++) { assign[H.v[b][j]] = -1; nmatches--; }
    }
    hist_free(&H); free(held); free(vIndices2);
    return nmatches;
}

/* ---- the remaining windowed searches of SURVEY 8a row 16 (loop closing / local mapping) ----
 * All of them: project a map point into a keyframe (caller geometry: mp_u, mp_v, mp_level = PredictScale,
 * mp_valid = every gate before KeyFrame::IsInImage), KeyFrame::GetFeaturesInArea(u, v, th*scale[level])
 * (src/KeyFrame.cc:916-962: no level filter, same grid as the Frame), keypoints at levels
 * [level-1, level], best Hamming distance, first best wins. */

/* KeyFrame::IsInImage, src/KeyFrame.cc:965-968 */
static int kf_in_image(const orc_grid* g, float x, float y) { return x >= g->minX && x < g->maxX && y >= g->minY && y < g->maxY; }

/* ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:549-678,
 * and its twin with vpPointsKFs / vpMatchedKF, :681-797 (same search; the extra vectors are caller bookkeeping).
 * matched: vpMatched[idx] != NULL on entry (may be NULL); assign[idx] = row of the map point newly matched. */
int orc_search_by_projection_sim3(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                                  int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                  const int32_t* mp_level, const uint8_t* mp_valid,
                                  float th, float ratioHamming, const uint8_t* matched, int32_t* assign)
{
    int nmatches = 0;
    uint8_t* held = (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1);
    int32_t* vIndices = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { assign[i] = -1; if (matched) held[i] = matched[i] != 0; }
    for (int i = 0; i < nmp; i++) {
        if (!mp_valid[i]) continue;                                           /* :582-594, 610-625 */
        const float u = mp_u[i], v = mp_v[i];
        if (!kf_in_image(g, u, v)) continue;                                  /* :603 */
        const int lvl = mp_level[i];
        const float radius = th * sf[lvl];                                    /* :631 */
        const int nc = orc_grid_query(g, kps, u, v, radius, -1, -1, vIndices, n);   /* :634 */
        if (nc == 0) continue;
        int bestDist = 256, bestIdx = -1;
        for (int k = 0; k < nc; k++) {                                        /* :645-664 */
            const int idx = vIndices[k];
            if (held[idx]) continue;
            const int kpLevel = kps[idx].octave;
            if (kpLevel < lvl - 1 || kpLevel > lvl) continue;
            const int dist = orc_hamming256(mp_desc + 32 * (size_t)i, desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestIdx >= 0 && bestDist <= TH_LOW * ratioHamming) {              /* :667 (bestIdx = -1 would index vpMatched[-1] there) */
            assign[bestIdx] = i; held[bestIdx] = 1;
            nmatches++;
        }
    }
    free(held); free(vIndices);
    return nmatches;
}

/* The search of ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th, bRight), src/ORBmatcher.cc:1823-2035
 * (mp_ur != NULL: reprojection gates :1958-1985 with uRight (NULL = mono keyframe, all -1) and invLevelSigma2),
 * and of Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint), :2051-2185 (mp_ur == NULL: no such gates).
 * best_idx[i] = keypoint the map point fuses with (-1: none, bestDist > TH_LOW); Replace / AddObservation stay
 * with the caller.  Returns the number of rows with a fuse target. */
int orc_fuse(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
             const float* sf, const float* inv_sigma2,
             int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v, const float* mp_ur,
             const int32_t* mp_level, const uint8_t* mp_valid, float th, int32_t* best_idx, int32_t* best_dist)
{
    int nFused = 0;
    int32_t* vIndices = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < nmp; i++) {
        best_idx[i] = -1; best_dist[i] = 256;
        if (!mp_valid[i]) continue;
        const float u = mp_u[i], v = mp_v[i];
        if (!kf_in_image(g, u, v)) continue;                                  /* :1905 / :2096 */
        const int lvl = mp_level[i];
        const float radius = th * sf[lvl];                                    /* :1938 / :2124 */
        const int nc = orc_grid_query(g, kps, u, v, radius, -1, -1, vIndices, n);
        if (nc == 0) continue;
        int bestDist = 256, bestIdx = -1;
        for (int k = 0; k < nc; k++) {
            const int idx = vIndices[k];
            const int kpLevel = kps[idx].octave;
            if (kpLevel < lvl - 1 || kpLevel > lvl) continue;                 /* :1955 / :2143 */
            if (mp_ur) {
                const float ex = u - kps[idx].x, ey = v - kps[idx].y;
                if (uRight && uRight[idx] >= 0) {                             /* :1958-1972 */
                    const float er = mp_ur[i] - uRight[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * inv_sigma2[kpLevel] > 7.8) continue;
                } else {                                                      /* :1974-1985 */
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * inv_sigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = orc_hamming256(mp_desc + 32 * (size_t)i, desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        best_dist[i] = bestDist;
        if (bestDist <= TH_LOW) { best_idx[i] = bestIdx; nFused++; }          /* :1999 / :2163 */
    }
    free(vIndices);
    return nFused;
}

/* one direction of SearchBySim3 (:2248-2330 / :2333-2412): best keypoint per projected map point, <= TH_HIGH */
static void sim3_direction(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                           int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                           const int32_t* mp_level, const uint8_t* mp_valid, float th, int32_t* vnMatch)
{
    int32_t* vIndices = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < nmp; i++) {
        vnMatch[i] = -1;
        if (!mp_valid[i]) continue;
        const float u = mp_u[i], v = mp_v[i];
        if (!kf_in_image(g, u, v)) continue;                                  /* :2280 */
        const int lvl = mp_level[i];
        const float radius = th * sf[lvl];                                    /* :2296 */
        const int nc = orc_grid_query(g, kps, u, v, radius, -1, -1, vIndices, n);
        if (nc == 0) continue;
        int bestDist = 0x7fffffff, bestIdx = -1;
        for (int k = 0; k < nc; k++) {
            const int idx = vIndices[k];
            if (kps[idx].octave < lvl - 1 || kps[idx].octave > lvl) continue;
            const int dist = orc_hamming256(mp_desc + 32 * (size_t)i, desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= TH_HIGH) vnMatch[i] = bestIdx;                        /* :2326 */
    }
    free(vIndices);
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th), src/ORBmatcher.cc:2201-2434.
 * Rows 1 = keypoints of KF1 (their map points projected into KF2), rows 2 the reverse; valid folds the
 * map point, vbAlreadyMatched, isBad, depth and distance gates.  match12[i1] = idx2 of a cross-checked
 * new match (:2415-2431) or -1.  Returns nFound. */
int orc_search_by_sim3(const orc_grid* g1, const orc_keypoint* kps1, const uint8_t* desc1, int n1,
                       const orc_grid* g2, const orc_keypoint* kps2, const uint8_t* desc2, int n2, const float* sf,
                       const uint8_t* mp1_desc, const float* u1in2, const float* v1in2, const int32_t* level1, const uint8_t* valid1,
                       const uint8_t* mp2_desc, const float* u2in1, const float* v2in1, const int32_t* level2, const uint8_t* valid2,
                       float th, int32_t* match12)
{
    int32_t* m1 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n1 > 0 ? n1 : 1));
    int32_t* m2 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n2 > 0 ? n2 : 1));
    sim3_direction(g2, kps2, desc2, n2, sf, n1, mp1_desc, u1in2, v1in2, level1, valid1, th, m1);
    sim3_direction(g1, kps1, desc1, n1, sf, n2, mp2_desc, u2in1, v2in1, level2, valid2, th, m2);
    int nFound = 0;
    for (int i1 = 0; i1 < n1; i1++) {
        match12[i1] = -1;
        const int idx2 = m1[i1];
        if (idx2 >= 0 && m2[idx2] == i1) { match12[i1] = idx2; nFound++; }
    }
    free(m1); free(m2);
    return nFound;
}

/* ---- Frame::isInFrustum (src/Frame.cc:603-711, branch Nleft == -1) for a list of map points ----
 * Float arithmetic follows the reference operation by operation: cv::Matx33f * Matx31f accumulates
 * ((R0*P0 + R1*P1) + R2*P2) in float then adds t; cv::norm(Matx31f) sums the squares in DOUBLE
 * (normL2Sqr<float, double>) and the square root is rounded to float on assignment; Pinhole::project
 * (src/CameraModels/Pinhole.cpp:33-37) is fx*x/z + cx; Matx::dot accumulates in float;
 * MapPoint::PredictScale (src/MapPoint.cc:725-741) is ceil(logf(maxDistance/dist) / mfLogScaleFactor)
 * clamped to [0, nLevels-1] (std::log(float): TemplatedVocabulary.h imports namespace std).
 * candidate[i] = the point is tested at all (src/Tracking.cc:3835-3839).  Outputs are the mTrack* fields;
 * inview[i] = the return value.  Returns the number of points in view (nToMatch). */
int orc_is_in_frustum(const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx, float cy, float mbf,
                      float minX, float maxX, float minY, float maxY, float logScaleFactor, int nlevels,
                      int nmp, const float* pos, const float* normal, const float* min_dist, const float* max_dist,
                      const uint8_t* candidate, float viewingCosLimit,
                      uint8_t* inview, float* projx, float* projy, float* projxr, float* depth, float* viewcos, int32_t* level)
{
    int nin = 0;
    for (int i = 0; i < nmp; i++) {
        inview[i] = 0; projx[i] = -1; projy[i] = -1; projxr[i] = 0; depth[i] = 0; viewcos[i] = 0; level[i] = 0;   /* :607-609 */
        if (candidate && !candidate[i]) continue;
        const float* P = pos + 3 * (size_t)i;
        float Pc[3];
        for (int r = 0; r < 3; r++) {
            float s = 0;
            for (int k = 0; k < 3; k++) s += Rcw[3 * r + k] * P[k];
            Pc[r] = s + tcw[r];                                               /* :615 */
        }
        double d2 = 0;
        for (int k = 0; k < 3; k++) { const double v = Pc[k]; d2 += v * v; }
        const float Pc_dist = (float)sqrt(d2);                                /* :616 */
        const float PcZ = Pc[2];
        const float invz = 1.0f / PcZ;
        if (PcZ < 0.0f) continue;                                             /* :621 */
        const float u = fx * Pc[0] / PcZ + cx, v = fy * Pc[1] / PcZ + cy;     /* :624 */
        if (u < minX || u > maxX) continue;
        if (v < minY || v > maxY) continue;
        projx[i] = u; projy[i] = v;                                           /* :631-632 */
        float PO[3];
        for (int k = 0; k < 3; k++) PO[k] = P[k] - Ow[k];                     /* :637 */
        double p2 = 0;
        for (int k = 0; k < 3; k++) { const double w = PO[k]; p2 += w * w; }
        const float dist = (float)sqrt(p2);
        if (dist < 0.8f * min_dist[i] || dist > 1.2f * max_dist[i]) continue; /* :647-656; GetMin/MaxDistanceInvariance, src/MapPoint.cc:668-681 */
        float dot = 0;
        for (int k = 0; k < 3; k++) dot += PO[k] * normal[3 * (size_t)i + k];
        const float vc = dot / dist;                                          /* :646 */
        if (vc < viewingCosLimit) continue;
        const float ratio = max_dist[i] / dist;                               /* PredictScale: mfMaxDistance itself, src/MapPoint.cc:730 */
        int nScale = (int)ceilf(logf(ratio) / logScaleFactor);
        if (nScale < 0) nScale = 0; else if (nScale >= nlevels) nScale = nlevels - 1;
        inview[i] = 1;
        projxr[i] = u - mbf * invz;                                           /* :657 */
        depth[i] = Pc_dist; level[i] = nScale; viewcos[i] = vc;
        nin++;
    }
    return nin;
}

/* ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:450-531) for a batch of map points ----
 * Group p = descriptor rows [start[p], start[p+1]) (the observations' descriptors in std::map order, as collected
 * at :463-481).  best_idx[p] = row (within the group) with the least median distance to the others (:505-524:
 * sorted row incl. the zero self-distance, element (size_t)(0.5*(N-1)), first least wins), -1 for an empty group. */
static int cmp_int(const void* a, const void* b) { return *(const int*)a - *(const int*)b; }
void orc_distinctive_descriptors(int npoints, const int32_t* start, const uint8_t* desc, int32_t* best_idx, int32_t* best_median)
{
    for (int p = 0; p < npoints; p++) {
        const int N = start[p + 1] - start[p];
        best_idx[p] = -1; best_median[p] = 0x7fffffff;
        if (N <= 0) continue;
        const uint8_t* D = desc + 32 * (size_t)start[p];
        int* v = (int*)malloc(sizeof(int) * (size_t)N);
        int BestMedian = 0x7fffffff, BestIdx = 0;
        for (int i = 0; i < N; i++) {
            for (int j = 0; j < N; j++) v[j] = i == j ? 0 : orc_hamming256(D + 32 * (size_t)i, D + 32 * (size_t)j);
            qsort(v, (size_t)N, sizeof(int), cmp_int);
            const int median = v[(size_t)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        free(v);
        best_idx[p] = BestIdx; best_median[p] = BestMedian;
    }
}

/* ---- DBoW2 vocabulary-tree transform behind Frame::ComputeBoW (src/Frame.cc:905-922) ----
 * TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1258:
 * from the root, move to the child with the least FORB::distance (Hamming, FORB.cpp:81-100; first least wins) until a
 * leaf; nid = the node passed at level L - levelsup (0 = root when that level is <= 0).  Tree as CSR: children of node
 * i = child_idx[child_start[i] .. child_start[i+1]) in push_back order (loadFromTextFile :1338-1415). */
void orc_bow_transform(const int32_t* child_start, const int32_t* child_idx, const uint8_t* node_desc,
                       const double* node_weight, const int32_t* node_word, int L, int levelsup,
                       int n, const uint8_t* desc, int32_t* word, double* weight, int32_t* node)
{
    const int nid_level = L - levelsup;
    for (int f = 0; f < n; f++) {
        const uint8_t* feat = desc + 32 * (size_t)f;
        int nid = 0, final_id = 0, current_level = 0;
        do {
            ++current_level;
            const int b = child_start[final_id], e = child_start[final_id + 1];
            final_id = child_idx[b];
            double best_d = orc_hamming256(feat, node_desc + 32 * (size_t)final_id);
            for (int c = b + 1; c < e; c++) {
                const int id = child_idx[c];
                const double d = orc_hamming256(feat, node_desc + 32 * (size_t)id);
                if (d < best_d) { best_d = d; final_id = id; }
            }
            if (current_level == nid_level) nid = final_id;
        } while (child_start[final_id] != child_start[final_id + 1]);
        word[f] = node_word[final_id]; weight[f] = node_weight[final_id]; node[f] = nid;
    }
}

/* transform(features, BowVector&, FeatureVector&, levelsup) :1127-1196 for TF_IDF weighting and L1 scoring (ORBVocabulary):
 * v.addWeight(id, w) in feature order for w > 0 (BowVector.cpp:34-46), fv.addFeature(nid, i), then v.normalize(L1)
 * (BowVector.cpp:62-84, sum in ascending word id).  Flat outputs: (bow_id, bow_val) ascending, FeatureVector as CSR
 * (fv_node ascending, indices in insertion order).  Returns the number of words; *nfv = number of feature-vector nodes. */
typedef struct { int32_t key; int32_t i; double w; } bow_item;
static int cmp_bow(const void* a, const void* b)
{
    const bow_item* x = (const bow_item*)a; const bow_item* y = (const bow_item*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->i - y->i;
}
int orc_bow_vectors(int n, const int32_t* word, const double* weight, const int32_t* node,
                    int32_t* bow_id, double* bow_val, int32_t* fv_node, int32_t* fv_start, int32_t* fv_idx, int* nfv)
{
    bow_item* it = (bow_item*)malloc(sizeof(bow_item) * (size_t)(n > 0 ? n : 1));
    int m = 0;
    for (int i = 0; i < n; i++) if (weight[i] > 0) { it[m].key = word[i]; it[m].i = i; it[m].w = weight[i]; m++; }
    qsort(it, (size_t)m, sizeof(bow_item), cmp_bow);
    int nw = 0;
    for (int k = 0; k < m; k++) {
        if (nw > 0 && bow_id[nw - 1] == it[k].key) bow_val[nw - 1] += it[k].w;    /* feature order inside a word */
        else { bow_id[nw] = it[k].key; bow_val[nw] = it[k].w; nw++; }
    }
    double norm = 0.0;
    for (int k = 0; k < nw; k++) norm += fabs(bow_val[k]);
    if (norm > 0.0) for (int k = 0; k < nw; k++) bow_val[k] /= norm;
    m = 0;
    for (int i = 0; i < n; i++) if (weight[i] > 0) { it[m].key = node[i]; it[m].i = i; it[m].w = 0; m++; }
    qsort(it, (size_t)m, sizeof(bow_item), cmp_bow);
    int nn = 0;
    for (int k = 0; k < m; k++) {
        if (nn == 0 || fv_node[nn - 1] != it[k].key) { fv_node[nn] = it[k].key; fv_start[nn] = k; nn++; }
        fv_idx[k] = it[k].i;
    }
    fv_start[nn] = m;
    *nfv = nn;
    free(it);
    return nw;
}

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints), src/ORBmatcher.cc:45-239,
 * branch F.Nleft != -1 (two-camera rig, e.g. the TUM-VI stereo-fisheye configuration): left keypoints mvKeys / mGrid,
 * right keypoints mvKeysRight / mGridRight (right-local indices), descriptor rows and mvpMapPoints indexed left first
 * then right (+ Nleft), mvLeftToRightMatch / mvRightToLeftMatch from ComputeStereoFishEyeMatches.  Per map point the
 * left fields (mbTrackInView, mTrackProjX/Y, mnTrackScaleLevel, mTrackViewCos) and the right ones (...R).
 * Quirks kept: a failed ratio test on the left side `continue`s past the right side (:148-149); the right radius is
 * not multiplied by th (:173 has no bFactor); a stereo partner is overwritten without the "already held" test
 * (:155-159, :222-226).  assign[nL + nR] in/out (>= 0 on entry: holds a map point with observations). */
int orc_search_by_projection_rig(const orc_grid* gL, const orc_keypoint* kpsL, int nL,
                                 const orc_grid* gR, const orc_keypoint* kpsR, int nR,
                                 const uint8_t* desc, const int32_t* l2r, const int32_t* r2l, const float* sf,
                                 int nmp, const uint8_t* mp_desc, const float* mp_depth,
                                 const uint8_t* inviewL, const float* projxL, const float* projyL, const int32_t* levelL, const float* viewcosL,
                                 const uint8_t* inviewR, const float* projxR, const float* projyR, const int32_t* levelR, const float* viewcosR,
                                 float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    int nmatches = 0;
    const int bFactor = th != 1.0;
    const int cap = (nL > nR ? nL : nR) + 1;
    int32_t* vIndices = (int32_t*)malloc(sizeof(int32_t) * (size_t)cap);
    for (int iMP = 0; iMP < nmp; iMP++) {
        if (!inviewL[iMP] && !inviewR[iMP]) continue;                         /* :54 */
        if (bFarPoints && mp_depth[iMP] > thFarPoints) continue;              /* :57 */
        const uint8_t* MPd = mp_desc + 32 * (size_t)iMP;
        if (inviewL[iMP]) {                                                   /* :63-167 */
            const int lvl = levelL[iMP];
            float r = viewcosL[iMP] > 0.998 ? 2.5f : 4.0f;
            if (bFactor) r *= th;
            const int nc = orc_grid_query(gL, kpsL, projxL[iMP], projyL[iMP], r * sf[lvl], lvl - 1, lvl, vIndices, cap);
            if (nc > 0) {
                int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
                for (int k = 0; k < nc; k++) {
                    const int idx = vIndices[k];
                    if (assign[idx] >= 0) continue;
                    const int dist = orc_hamming256(MPd, desc + 32 * (size_t)idx);
                    if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kpsL[idx].octave; bestIdx = idx; }
                    else if (dist < bestDist2) { bestLevel2 = kpsL[idx].octave; bestDist2 = dist; }
                }
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;       /* skips the right side too */
                    if (bestLevel != bestLevel2 || bestDist <= nnratio * bestDist2) {
                        assign[bestIdx] = iMP;
                        if (l2r[bestIdx] != -1) { assign[l2r[bestIdx] + nL] = iMP; nmatches++; }    /* :155-159 */
                        nmatches++;
                    }
                }
            }
        }
        if (inviewR[iMP]) {                                                   /* :170-236 */
            const int lvl = levelR[iMP];
            if (lvl != -1) {
                const float r = viewcosR[iMP] > 0.998 ? 2.5f : 4.0f;
                const int nc = orc_grid_query(gR, kpsR, projxR[iMP], projyR[iMP], r * sf[lvl], lvl - 1, lvl, vIndices, cap);
                if (nc == 0) continue;
                int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
                for (int k = 0; k < nc; k++) {
                    const int idx = vIndices[k];
                    if (assign[idx + nL] >= 0) continue;
                    const int dist = orc_hamming256(MPd, desc + 32 * (size_t)(idx + nL));
                    if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kpsR[idx].octave; bestIdx = idx; }
                    else if (dist < bestDist2) { bestLevel2 = kpsR[idx].octave; bestDist2 = dist; }
                }
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
                    if (r2l[bestIdx] != -1) { assign[r2l[bestIdx]] = iMP; nmatches++; }             /* :222-226 */
                    assign[bestIdx + nL] = iMP;
                    nmatches++;
                }
            }
        }
    }
    free(vIndices);
    return nmatches;
}

/* ===== two-camera rig (Nleft != -1) branches and the callback-driven triangulation searches ===== */

/* ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), src/ORBmatcher.cc:2469-2711,
 * branch CurrentFrame.Nleft != -1 (KannalaBrandt stereo rig, config C3).  Keypoints: kpsL = mvKeys (nL = Nleft, grid gL =
 * mGrid), kpsR = mvKeysRight (nR, grid gR = mGridRight); desc / cur_held / assign are indexed like mvpMapPoints:
 * left keypoint i at i, right keypoint j at nL + j.  Per last-frame point the caller supplies both projections:
 * (last_u, last_v) into the left camera (:2516) and (last_ur, last_vr) of Trl * x3Dc (:2613-2616), plus invzc, octave
 * (:2530, picks mvKeys / mvKeysRight of the LAST frame), the last keypoint's angle, valid = pMP && !outlier, obs.
 * Quirks kept: a point whose LEFT window is empty skips its right half as well (`continue`, :2541); there is no stereo
 * gate (Nleft != -1, :2556) and no image-bounds test on the right projection. */
int orc_search_by_projection_frame_rig(const orc_grid* gL, const orc_keypoint* kpsL, int nL,
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
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = HISTO_LENGTH / 360.0f;                               /* :2480 */
    uint8_t* held = (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1);
    int32_t* vIndices2 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { assign[i] = -1; if (cur_held) held[i] = cur_held[i] != 0; }
    for (int i = 0; i < nlast; i++) {
        if (!last_valid[i]) continue;                                         /* :2503-2506 */
        if (last_invz[i] < 0) continue;                                       /* :2515 */
        const float u = last_u[i], v = last_v[i];
        if (u < gL->minX || u > gL->maxX) continue;                           /* :2520-2523 */
        if (v < gL->minY || v > gL->maxY) continue;
        const int nLastOctave = last_octave[i];
        const float radius = th * sf[nLastOctave];                            /* :2529 */
        int nc;
        if (bForward) nc = orc_grid_query(gL, kpsL, u, v, radius, nLastOctave, -1, vIndices2, n);
        else if (bBackward) nc = orc_grid_query(gL, kpsL, u, v, radius, 0, nLastOctave, vIndices2, n);
        else nc = orc_grid_query(gL, kpsL, u, v, radius, nLastOctave - 1, nLastOctave + 1, vIndices2, n);
        if (nc == 0) continue;                                                /* :2541: also skips the right camera */
        int bestDist = 256, bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {                                        /* :2548-2574 */
            const int i2 = vIndices2[k];
            if (held[i2]) continue;
            const int dist = orc_hamming256(last_desc + 32 * (size_t)i, desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {                                            /* :2576-2601 */
            assign[bestIdx2] = i;
            held[bestIdx2] = last_obs[i] != 0;
            nmatches++;
            if (checkOrientation) hist_push(&H, hist_bin(last_angle[i], kpsL[bestIdx2].angle, factor), bestIdx2);
        }
        /* right camera, :2610-2684 */
        const float ur = last_ur[i], vr = last_vr[i];
        if (bForward) nc = orc_grid_query(gR, kpsR, ur, vr, radius, nLastOctave, -1, vIndices2, n);
        else if (bBackward) nc = orc_grid_query(gR, kpsR, ur, vr, radius, 0, nLastOctave, vIndices2, n);
        else nc = orc_grid_query(gR, kpsR, ur, vr, radius, nLastOctave - 1, nLastOctave + 1, vIndices2, n);
        bestDist = 256; bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {
            const int i2 = vIndices2[k];
            if (held[i2 + nL]) continue;                                      /* :2640-2642 */
            const int dist = orc_hamming256(last_desc + 32 * (size_t)i, desc + 32 * (size_t)(i2 + nL));
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {                                            /* :2658-2683 */
            assign[bestIdx2 + nL] = i;
            held[bestIdx2 + nL] = last_obs[i] != 0;
            nmatches++;
            if (checkOrientation) hist_push(&H, hist_bin(last_angle[i], kpsR[bestIdx2].angle, factor), bestIdx2 + nL);
        }
    }
    if (checkOrientation) {                                                   /* :2686-2708 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int b = 0; b < HISTO_LENGTH; b++)
            if (b != ind1 && b != ind2 && b != ind3)
                for (int j = 0; j < H.n[b]; j++) { assign[H.v[b][j]] = -1; nmatches--; }
    }
    hist_free(&H); free(held); free(vIndices2);
    return nmatches;
}

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&), src/ORBmatcher.cc:314-547, branch F.Nleft != -1:
 * Frame rows < nleft2 are left-camera keypoints, the others right-camera ones; best / second are kept per camera
 * (:380-430).  Quirks kept: the right-camera match is only considered when the LEFT best passed TH_LOW (:431, the block
 * at :455 is nested in it) and ignores its own ratio test (`|| true`, :457).  kps1 / kps2 carry the angles the reference
 * picks (:437-448, :459-470: mvKeys / mvKeysRight by index).  matches[n2] out: KF keypoint per Frame keypoint. */
int orc_search_by_bow_rig(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                          const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                          const orc_keypoint* kps2, const uint8_t* desc2, int n2, int nleft2,
                          const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                          float nnratio, int checkOrientation, int32_t* matches)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = HISTO_LENGTH / 360.0f;                               /* :334 */
    uint8_t* matched2 = (uint8_t*)calloc((size_t)(n2 > 0 ? n2 : 1), 1);
    for (int i = 0; i < n2; i++) matches[i] = -1;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (fv1_node[a] < fv2_node[b]) { a++; continue; }
        if (fv1_node[a] > fv2_node[b]) { b++; continue; }
        for (int p1 = fv1_start[a]; p1 < fv1_start[a + 1]; p1++) {
            const int idx1 = fv1_idx[p1];
            if (!valid1[idx1]) continue;                                      /* :343-349 */
            int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
            int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
            for (int p2 = fv2_start[b]; p2 < fv2_start[b + 1]; p2++) {
                const int idx2 = fv2_idx[p2];
                if (matched2[idx2]) continue;                                 /* :390 */
                const int dist = orc_hamming256(desc1 + 32 * (size_t)idx1, desc2 + 32 * (size_t)idx2);
                if (idx2 < nleft2 && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = idx2; }
                else if (idx2 < nleft2 && dist < bestDist2) bestDist2 = dist;
                if (idx2 >= nleft2 && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = idx2; }
                else if (idx2 >= nleft2 && dist < bestDist2R) bestDist2R = dist;
            }
            if (bestDist1 <= TH_LOW) {                                        /* :431 */
                if ((float)bestDist1 < nnratio * (float)bestDist2) {          /* :433 */
                    matches[bestIdxF] = idx1; matched2[bestIdxF] = 1;
                    if (checkOrientation) hist_push(&H, hist_bin(kps1[idx1].angle, kps2[bestIdxF].angle, factor), bestIdxF);
                    nmatches++;
                }
                if (bestDist1R <= TH_LOW) {                                   /* :455-457: ratio test disabled by `|| true` */
                    matches[bestIdxFR] = idx1; matched2[bestIdxFR] = 1;
                    if (checkOrientation) hist_push(&H, hist_bin(kps1[idx1].angle, kps2[bestIdxFR].angle, factor), bestIdxFR);
                    nmatches++;
                }
            }
        }
        a++; b++;
    }
    if (checkOrientation) {                                                   /* :520-544 */
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int bi = 0; bi < HISTO_LENGTH; bi++) {
            if (bi == ind1 || bi == ind2 || bi == ind3) continue;
            for (int j = 0; j < H.n[bi]; j++) { matches[H.v[bi][j]] = -1; nmatches--; }
        }
    }
    hist_free(&H); free(matched2);
    return nmatches;
}

/* The triangulation searches whose geometric verdict is the camera model's own (caller-side) code:
 *   overload 0  SearchForTriangulation  src/ORBmatcher.cc:1107-1386, branch pKF1->mpCamera2 && pKF2->mpCamera2
 *               (verdict = pCamera1->epipolarConstrain(...) with the Rll/Rlr/Rrl/Rrr pose picked at :1252-1279)
 *   overload 1  SearchForTriangulation_ :1388-1630, same branch (epipolarConstrain_)
 *   overload 2  SearchForTriangulation(..., vector<cv::Mat>& vMatchedPoints) :1632-1821
 *               (verdict = pCamera1->matchAndtriangulate(...), which also yields the 3-D point)
 * verdict(user, idx1, idx2) is called exactly where the reference calls the camera model: for the candidates of idx1 in
 * FeatureVector order that pass `!vbMatched2 && !pMP2`, `dist <= TH_LOW && dist <= bestDist`.  With a rig bStereo1 /
 * bStereo2 are false (:1194, :1222), so bOnlyStereo leaves no candidate and the epipole gate (:1241) is off.
 * Differences kept: overload 0 marks vbMatched2 (:1297) and un-marks it in the histogram filter (:1362), 1 and 2 never
 * mark it; the histogram factor is 30/360 in overload 0 and 1/30 in 1 and 2 (:1166 / :1441 / :1666); with bCoarse
 * (overloads 0, 1) the camera model is still asked - `A || bCoarse` evaluates A first - and its answer ignored. */
int orc_search_for_triangulation_cb(int overload,
                                    const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                    const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                    const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                    const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                    int bOnlyStereo, int bCoarse, int checkOrientation,
                                    orc_pair_verdict_fn verdict, void* user, int32_t* matches12)
{
    int nmatches = 0;
    rot_hist H; hist_init(&H);
    const float factor = overload == 0 ? HISTO_LENGTH / 360.0f : 1.0f / HISTO_LENGTH;
    uint8_t* vbMatched2 = (uint8_t*)calloc((size_t)(n2 > 0 ? n2 : 1), 1);
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    int a = 0, b = 0;
    while (a < nn1 && b < nn2) {
        if (fv1_node[a] < fv2_node[b]) { a++; continue; }
        if (fv1_node[a] > fv2_node[b]) { b++; continue; }
        for (int p1 = fv1_start[a]; p1 < fv1_start[a + 1]; p1++) {
            const int idx1 = fv1_idx[p1];
            if (has_mp1[idx1]) continue;                                      /* :1188-1192 / :1682-1684 */
            if (overload != 2 && bOnlyStereo) continue;                       /* :1194-1198: bStereo1 is false with a rig */
            int bestDist = TH_LOW, bestIdx2 = -1;
            for (int p2 = fv2_start[b]; p2 < fv2_start[b + 1]; p2++) {
                const int idx2 = fv2_idx[p2];
                if (vbMatched2[idx2] || has_mp2[idx2]) continue;              /* :1218 / :1701 */
                const int dist = orc_hamming256(desc1 + 32 * (size_t)idx1, desc2 + 32 * (size_t)idx2);
                if (dist > TH_LOW || dist > bestDist) continue;               /* :1231 / :1707 */
                /* :1282 `epipolarConstrain(...) || bCoarse`: the camera model is asked FIRST, bCoarse only overrides its answer */
                int ok = verdict ? verdict(user, idx1, idx2) != 0 : 0;
                if (overload != 2 && bCoarse) ok = 1;
                if (ok) { bestIdx2 = idx2; bestDist = dist; }
            }
            if (bestIdx2 >= 0) {
                matches12[idx1] = bestIdx2;
                if (overload == 0) vbMatched2[bestIdx2] = 1;                  /* :1297 */
                nmatches++;
                if (checkOrientation) hist_push(&H, hist_bin(kps1[idx1].angle, kps2[bestIdx2].angle, factor), idx1);
            }
        }
        a++; b++;
    }
    if (checkOrientation) {
        int ind1, ind2, ind3;
        orc_three_maxima(H.n, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int bi = 0; bi < HISTO_LENGTH; bi++) {
            if (bi == ind1 || bi == ind2 || bi == ind3) continue;
            for (int j = 0; j < H.n[bi]; j++) {
                if (overload == 0) vbMatched2[matches12[H.v[bi][j]]] = 0;     /* :1362 */
                matches12[H.v[bi][j]] = -1; nmatches--;
            }
        }
    }
    hist_free(&H); free(vbMatched2);
    return nmatches;
}

/* Frame::ComputeStereoFishEyeMatches, src/Frame.cc:1440-1480, up to the triangulation: the descriptors of the lapping
 * area are the rows [monoLeft, nL) / [monoRight, nR) of the two extractors' outputs (:1445-1446); BFMatcher knnMatch k=2
 * (:1458); Lowe ratio 0.7 (:1465: size >= 2 && d0 < d1 * 0.7).  l2r[nL] / r2l[nR] out = mvLeftToRightMatch /
 * mvRightToLeftMatch candidates in FULL-frame indices (:1477-1478 before the triangulation gate), -1 elsewhere.
 * Later matches overwrite earlier ones exactly as the loop at :1463 does.  Returns the number of ratio-test survivors. */
int orc_stereo_fisheye_match(const uint8_t* descL, int nL, int monoLeft, const uint8_t* descR, int nR, int monoRight,
                             int32_t* l2r, int32_t* r2l, int32_t* dist_out)
{
    for (int i = 0; i < nL; i++) { l2r[i] = -1; if (dist_out) dist_out[i] = -1; }
    for (int i = 0; i < nR; i++) r2l[i] = -1;
    const int nq = nL - monoLeft, nt = nR - monoRight;
    if (nq <= 0 || nt <= 0) return 0;
    int32_t* idx = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)nq);
    int32_t* dst = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)nq);
    orc_bf_knn2(descL + 32 * (size_t)monoLeft, nq, descR + 32 * (size_t)monoRight, nt, idx, dst);
    int good = 0;
    for (int q = 0; q < nq; q++) {
        if (idx[2 * q] < 0 || idx[2 * q + 1] < 0) continue;                   /* (*it).size() >= 2 */
        if ((double)dst[2 * q] < (double)dst[2 * q + 1] * 0.7) {              /* float distance * double 0.7, :1465 */
            const int iL = q + monoLeft, iR = idx[2 * q] + monoRight;
            l2r[iL] = iR; r2l[iR] = iL;
            if (dist_out) dist_out[iL] = dst[2 * q];
            good++;
        }
    }
    free(idx); free(dst);
    return good;
}

/* bounds of a grid (mnMinX, mnMaxX, mnMinY, mnMaxY) - for oracle/ref_match_wrap.cc, which fills the Frame doubles */
void orc_grid_bounds(const orc_grid* g, float* out4) { out4[0] = g->minX; out4[1] = g->maxX; out4[2] = g->minY; out4[3] = g->maxY; }
