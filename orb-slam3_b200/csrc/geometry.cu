// geometry.cu — host side: ORBextractor constructor tables (reference src/ORBextractor.cc:468-571),
// per-level geometry (ComputePyramid :1668-1672, cell grid :1053-1072, quadtree roots :695-698) and
// the fixed-point resize tap tables (OpenCV resize INTER_LINEAR, see oracle/orc_prims.c).
#include "orbx_host.h"
#include <cmath>
#include <cstring>
#include <algorithm>
#include <vector>

static inline int cv_round_f(float v) { return (int)lrintf(v); }   // cvRound: round-half-even
static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }
static inline int reflect101(int i, int n) { if (i < 0) i = -i; if (i >= n) i = 2 * (n - 1) - i; return i; }

void orbx_init_tables(orbx_extractor* e)
{
    const int nlevels = e->nlevels;
    e->scaleFactorD = (double)e->scaleFactorF;                       // double member, float argument
    e->sf[0] = 1.0f; e->sig2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        e->sf[i] = (float)((double)e->sf[i - 1] * e->scaleFactorD);   // :488
        e->sig2[i] = e->sf[i] * e->sf[i];
    }
    for (int i = 0; i < nlevels; i++) { e->isf[i] = 1.0f / e->sf[i]; e->isig2[i] = 1.0f / e->sig2[i]; }
    float factor = (float)(1.0f / e->scaleFactorD);                  // :509
    volatile float num = (float)e->nfeatures * (1 - factor);
    volatile float den = 1 - (float)pow((double)factor, (double)nlevels);
    float nDesired = num / den;                                      // :511
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        e->nper[l] = cv_round_f(nDesired);
        sum += e->nper[l];
        nDesired *= factor;
    }
    e->nper[nlevels - 1] = e->nfeatures - sum > 0 ? e->nfeatures - sum : 0;
}

// Fills G (without device pointers) for an image of rows x cols.  Returns 0 or ORBX_ERR_GEOMETRY.
// Fast-path word taps of a level from its byte-indexed tap table; false if some word's taps do not fit
// the 8-byte window (scale factors > ~2) -> the level uses the generic kernel.
bool orbx_build_word_taps(const std::vector<ResizeTapX>& TX, int w, int pitch, std::vector<ResizeWord>* out)
{
    out->assign(pitch / 4, ResizeWord{0, 0, 0, {0, 0, 0, 0}, 0});
    // interior words plus one halo word on each side; TX holds the reflect-101 taps for the bytes outside
    // the image, so the halo words (and the pixels of the last word past column w-1) come out mirrored
    for (int wi = ORBX_XOFF / 4 - 1; wi <= (ORBX_XOFF + w + 3) / 4; wi++) {
        ResizeWord rw{};
        int base = 1 << 30;
        for (int k = 0; k < 4; k++) base = std::min(base, (int)TX[wi * 4 + k].s0);
        rw.wb = base & ~3; rw.shift = (uint32_t)(base & 3) * 8;
        for (int k = 0; k < 4; k++) {
            const ResizeTapX& t = TX[wi * 4 + k];
            // window byte n = source column (base & ~3) + shift/8 + n
            const int n0 = t.s0 - base, n1 = t.s1 - base;
            if (n0 < 0 || n0 > 7 || n1 < 0 || n1 > 7) return false;
            rw.sels |= (uint32_t)(n0 | (n1 << 4)) << (8 * k);
            rw.a01[k] = (uint32_t)(uint16_t)t.a0 | ((uint32_t)(uint16_t)t.a1 << 16);
        }
        (*out)[wi] = rw;
    }
    return true;
}

int orbx_compute_geometry(const orbx_extractor* e, int rows, int cols, int batch, FrameGeom* G,
                          std::vector<std::vector<ResizeTapX>>* tapx, std::vector<std::vector<ResizeTapY>>* tapy)
{
    memset(G, 0, sizeof(*G));
    G->nlevels = e->nlevels; G->rows = rows; G->cols = cols; G->batch = batch;
    G->iniTh = e->iniTh; G->minTh = e->minTh;
    int cells = 0, sel = 0;
    long long keys = 0;
    tapx->assign(e->nlevels, {}); tapy->assign(e->nlevels, {});
    for (int l = 0; l < e->nlevels; l++) {
        LevelGeom& L = G->L[l];
        const float scale = e->isf[l];
        L.w = cv_round_f((float)cols * scale);                        // :1669
        L.h = cv_round_f((float)rows * scale);
        L.pitch = align_up(ORBX_XOFF + L.w + ORBX_EDGE, 128);
        L.prow = L.h + 2 * ORBX_EDGE;
        L.plane_bytes = (long long)L.pitch * L.prow;
        L.maxBX = L.w - ORBX_EDGE + 3; L.maxBY = L.h - ORBX_EDGE + 3;   // :1055-1056
        const float width = (float)(L.maxBX - ORBX_MINB), height = (float)(L.maxBY - ORBX_MINB);
        if (width < 35.f || height < 35.f) return ORBX_ERR_GEOMETRY;   // nCols/nRows would be 0 (:1064-1068 divide by zero)
        if (L.maxBX - ORBX_MINB > 32767 || L.maxBY - ORBX_MINB > 32767) return ORBX_ERR_GEOMETRY;   // 16-bit node bounds / resize taps
        L.nCols = (int)(width / 35.f); L.nRows = (int)(height / 35.f);
        L.wCell = (int)ceilf(width / L.nCols); L.hCell = (int)ceilf(height / L.nRows);
        L.cell_base = cells;
        L.cell_cap = ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2);       // NMS keeps <= 1 corner per 2x2 block
        L.key_base = keys;
        const int ncell = L.nCols * L.nRows;
        cells += ncell;
        L.cand_cap = ncell * L.cell_cap;
        L.cand_base = keys;
        keys += (long long)ncell * L.cell_cap;
        L.N = e->nper[l];
        L.nIni = (int)roundf((float)(L.maxBX - ORBX_MINB) / (float)(L.maxBY - ORBX_MINB));   // :695
        if (L.nIni < 1) return ORBX_ERR_GEOMETRY;                     // reference: hX = x/0
        L.hX = (float)(L.maxBX - ORBX_MINB) / (float)L.nIni;          // :698
        const int mx = (L.N + 3 > 4 * L.nIni ? L.N + 3 : 4 * L.nIni);
        L.sel_cap = mx;
        L.node_cap = align_up(mx + 1, 4);
        L.sel_base = sel;
        sel += L.sel_cap;
        L.scale = e->sf[l];
        L.patch_size = (float)(int)(ORBX_PATCH * e->sf[l]);           // :1161
        if (l > 0) {
            const LevelGeom& P = G->L[l - 1];
            const double sx = (double)P.w / L.w, sy = (double)P.h / L.h;
            auto& TX = (*tapx)[l]; auto& TY = (*tapy)[l];
            // TX is indexed by the BYTE of the padded row (padded column X = byte - (XOFF - EDGE)); bytes that
            // are not image columns get an all-zero tap, which makes the kernel write 0 there.
            TX.assign(L.pitch, ResizeTapX{0, 0, 0, 0}); TY.resize(L.h + 2 * ORBX_EDGE);
            for (int X = 0; X < L.w + 2 * ORBX_EDGE; X++) {
                const int dx = reflect101(X - ORBX_EDGE, L.w);
                float fx = (float)((dx + 0.5) * sx - 0.5);
                int s = (int)floorf(fx);
                fx -= s;
                if (s < 0) { fx = 0; s = 0; }
                if (s >= P.w - 1) { fx = 0; s = P.w - 1; }
                ResizeTapX t;
                t.s0 = (short)s; t.s1 = (short)(s + 1 < P.w ? s + 1 : P.w - 1);
                t.a0 = (short)cv_round_f((1.f - fx) * 2048.f); t.a1 = (short)cv_round_f(fx * 2048.f);
                TX[X + ORBX_XOFF - ORBX_EDGE] = t;
            }
            for (int Y = 0; Y < L.h + 2 * ORBX_EDGE; Y++) {
                const int dy = reflect101(Y - ORBX_EDGE, L.h);
                float fy = (float)((dy + 0.5) * sy - 0.5);
                int s = (int)floorf(fy);
                fy -= s;
                ResizeTapY t;
                t.s0 = (short)(s < 0 ? 0 : (s > P.h - 1 ? P.h - 1 : s));
                t.s1 = (short)(s + 1 < 0 ? 0 : (s + 1 > P.h - 1 ? P.h - 1 : s + 1));
                t.b0 = (short)cv_round_f((1.f - fy) * 2048.f); t.b1 = (short)cv_round_f(fy * 2048.f);
                TY[Y] = t;
            }
        }
    }
    G->cells_per_frame = cells;
    G->keys_per_frame = keys;
    G->cand_per_frame = keys;
    G->sel_per_frame = sel;
    G->out_cap = sel;
    return 0;
}
