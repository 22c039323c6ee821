// fast.cu — per-cell FAST-9/16 with the iniThFAST/minThFAST retry
// (ORBextractor::ComputeKeyPointsOctTree cell loop, reference src/ORBextractor.cc:1038-1143, and the
// cv::FAST / cornerScore<16> / 3x3 non-max suppression it calls, restated in oracle/orc_prims.c).
//
// Work unit ("band"): the interiors of up to 768 px worth of cells of one cell row.  The interiors of
// the reference's overlapping cell sub-images tile the level without overlap, so every pixel is scored
// once.  Three phases per CTA:
//   A  threshold-free score S = max(A, Bn) - 1 of every pixel (A = v - min_arcs max_arc ring,
//      Bn = max_arcs min_arc ring - v; pixel is a corner at threshold t  <=>  S >= t, and its OpenCV
//      score is S).  Each thread owns 4 adjacent pixels and slides down the band with a 7-row register
//      window; the ring is evaluated two pixels per instruction on u16x2 lanes (VIMNMX3.U16x2): a lane
//      is the raw byte pair (garbage, pixel) so the byte-shifted row words are used as they are —
//      lexicographic min/max keeps the exact pixel value in the high byte.
//   B1 3x3 non-max suppression on the shared-memory score tile, again on u16x2 lanes; neighbours
//      outside the pixel's own cell interior count as 0 (the reference runs cv::FAST per cell).  The
//      strict-maximum test does not depend on the threshold, so one pass serves both thresholds;
//      per cell the survivors >= iniThFAST and those in [minThFAST, iniThFAST) are counted.
//   B2 survivors >= iniThFAST are emitted; a cell without any emits its >= minThFAST survivors
//      (the retry of :1118-1125).  Emission order is arbitrary: the quadtree breaks ties by the
//      candidate's canonical rank (cell row, cell col, y, x), i.e. the reference's vector order.
// Integer work on the ALU pipe (~55 lane-ops per pixel); HBM traffic = one read of the level.
#include "orbx_internal.h"
#include <cstdlib>
#include <algorithm>
#include <vector>

#define FAST_WARPS 4
#define FAST_THREADS (FAST_WARPS * 32)
#define FAST_MAXSPAN 128               // pixels per work item: one warp, 4 px per lane
#define FAST_MAXCELLS 4

// ring offsets of cv::FAST 9_16 (oracle/orc_prims.c)
#define RING_DX {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1}
#define RING_DY {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3}

// Score of two pixels.  r[k]: u16x2 lanes (low byte garbage, high byte ring pixel k), c: centre lanes.
// Returns the two scores (0..254) in the low bytes of the two lanes.
__device__ __forceinline__ uint32_t fast_score_pair(const uint32_t (&r)[16], uint32_t c)
{
    // arcs k and k+3 share two of their three triples:  min(arc_k, arc_k+3) =
    // max(t3[k+3], t3[k+6], min(t3[k], t3[k+9]));  the 16 arcs pair up as (k, k+3) for these 8 starts
    constexpr int PK[8] = {0, 6, 12, 2, 8, 14, 4, 10};
    uint32_t t3[16], w[8];
    uint32_t P, Q;
    // P: min over the 16 arcs of (max of the arc's 9 ring pixels)
#pragma unroll
    for (int k = 0; k < 16; k++) t3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int k = PK[i];
        w[i] = __vimax3_u16x2(t3[(k + 3) & 15], t3[(k + 6) & 15], __vminu2(t3[k], t3[(k + 9) & 15]));
    }
    P = __vimin3_u16x2(w[0], w[1], w[2]);
    P = __vimin3_u16x2(P, w[3], w[4]);
    P = __vimin3_u16x2(P, w[5], w[6]);
    P = __vminu2(P, w[7]);
    // Q: max over the arcs of (min of the arc)
#pragma unroll
    for (int k = 0; k < 16; k++) t3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int k = PK[i];
        w[i] = __vimin3_u16x2(t3[(k + 3) & 15], t3[(k + 6) & 15], __vmaxu2(t3[k], t3[(k + 9) & 15]));
    }
    Q = __vimax3_u16x2(w[0], w[1], w[2]);
    Q = __vimax3_u16x2(Q, w[3], w[4]);
    Q = __vimax3_u16x2(Q, w[5], w[6]);
    Q = __vmaxu2(Q, w[7]);
    // clean lanes (high byte -> value), biased by 0x100 so that plain 32-bit adds never borrow across lanes
    const uint32_t Pc = __byte_perm(P, 0u, 0x4341), Qc = __byte_perm(Q, 0u, 0x4341), Vc = __byte_perm(c, 0u, 0x4341);
    const uint32_t Ab = Vc + 0x01000100u - Pc;        // 0x100 + (v - P)  per lane, in [1, 0x1FF]
    const uint32_t Bb = Qc + 0x01000100u - Vc;        // 0x100 + (Q - v)
    const uint32_t Mb = __vmaxu2(Ab, Bb);
    return __viaddmax_s16x2(Mb, 0xFEFFFEFFu, 0u);     // max(M - 1, 0): add -0x101 per lane
}

// window row: the 8 byte-shifted words starting at x-4 .. x+3 (x = first pixel of the thread's word)
// ORBX_FAST_IMAD_SHIFT (build-time experiment, measured SLOWER: FAST 1.997 vs 1.943 ms): the six funnel shifts (ALU pipe) as IMAD.HI + IMAD with multipliers the compiler
// cannot see (constant bank), i.e. on the FMA pipe: (l >> s) | (m << (32 - s)) = umulhi(l, 2^(32-s)) + m * 2^(32-s)
__constant__ uint32_t c_fast_pow[3] = {1u << 24, 1u << 16, 1u << 8};
__device__ __forceinline__ uint32_t fast_fsr(uint32_t lo, uint32_t hi, int i)
{
    const uint32_t k = c_fast_pow[i];
    return __umulhi(lo, k) + hi * k;
}
__device__ __forceinline__ void fast_fill_row(uint32_t (&w)[8], uint32_t l, uint32_t m, uint32_t r)
{
#ifdef ORBX_FAST_IMAD_SHIFT
    w[0] = l; w[1] = fast_fsr(l, m, 0); w[2] = fast_fsr(l, m, 1); w[3] = fast_fsr(l, m, 2);
    w[4] = m; w[5] = fast_fsr(m, r, 0); w[6] = fast_fsr(m, r, 1); w[7] = fast_fsr(m, r, 2);
    return;
#endif
    w[0] = l;
    w[1] = __funnelshift_r(l, m, 8);
    w[2] = __funnelshift_r(l, m, 16);
    w[3] = __funnelshift_r(l, m, 24);
    w[4] = m;
    w[5] = __funnelshift_r(m, r, 8);
    w[6] = __funnelshift_r(m, r, 16);
    w[7] = __funnelshift_r(m, r, 24);
}

// scores of the thread's 4 pixels of the row whose window slots are (PH + dy + 3) % 7, dy = -3..3
template <int PH>
__device__ __forceinline__ uint32_t fast_score_word(const uint32_t (&W)[7][8])
{
    constexpr int DX[16] = RING_DX;
    constexpr int DY[16] = RING_DY;
    uint32_t re[16], ro[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        re[k] = W[(PH + DY[k] + 3) % 7][DX[k] + 3];     // lanes: pixels x+dx, x+dx+2   (ring of px 0 and 2)
        ro[k] = W[(PH + DY[k] + 3) % 7][DX[k] + 4];     // lanes: pixels x+dx+1, x+dx+3 (ring of px 1 and 3)
    }
    const uint32_t se = fast_score_pair(re, W[(PH + 3) % 7][3]);
    const uint32_t so = fast_score_pair(ro, W[(PH + 3) % 7][4]);
    return se + (so << 8);                               // bytes S0 S1 S2 S3
}

// 3x3 NMS row step.  R[slot][0..3]: score words starting at x-2, x-1, x, x+1 of rows r-1, r, r+1 at slots
// (PH, PH+1, PH+2) % 3.  Returns non-zero lanes for the survivors: te = px 0,2 ; to = px 1,3.
template <int PH>
__device__ __forceinline__ void nms_row(const uint32_t (&R)[3][4], uint32_t thr, uint32_t ML_e, uint32_t MR_e,
                                        uint32_t ML_o, uint32_t MR_o, uint32_t V_e, uint32_t V_o,
                                        uint32_t& te, uint32_t& to)
{
    const uint32_t (&U)[4] = R[PH % 3];
    const uint32_t (&C)[4] = R[(PH + 1) % 3];
    const uint32_t (&D)[4] = R[(PH + 2) % 3];
    // even pixels (x, x+2): left neighbours = lanes of word x-2, centre column = word x-1, right = word x
    uint32_t cl = __vimax3_u16x2(U[0], C[0], D[0]) & ML_e;
    uint32_t cr = __vimax3_u16x2(U[2], C[2], D[2]) & MR_e;
    uint32_t cm = __vimax3_u16x2(U[1], D[1], thr);
    uint32_t nb = __vimax3_u16x2(cl, cr, cm) | 0x00FF00FFu;
    te = __vmaxu2(C[1] & V_e, nb) ^ nb;
    // odd pixels (x+1, x+3): left = word x-1, centre = word x, right = word x+1
    cl = __vimax3_u16x2(U[1], C[1], D[1]) & ML_o;
    cr = __vimax3_u16x2(U[3], C[3], D[3]) & MR_o;
    cm = __vimax3_u16x2(U[2], D[2], thr);
    nb = __vimax3_u16x2(cl, cr, cm) | 0x00FF00FFu;
    to = __vmaxu2(C[2] & V_o, nb) ^ nb;
}

// append the 32-bit entry e to the lane's list at shared address wa when flag != 0: predicated store + bump,
// no divergent region.  Entry i of lane l lives at word i * 32 + l of the list area, so a push never
// meets a bank conflict.
__device__ __forceinline__ void list_push(uint32_t& wa, uint32_t flag, uint32_t e)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p st.shared.u32 [%0], %2;\n\t@p add.u32 %0, %0, 128;\n\t}"
                 : "+r"(wa) : "r"(flag), "r"(e) : "memory");
}

__device__ __forceinline__ void nms_fill(uint32_t (&w)[4], uint32_t l, uint32_t m, uint32_t r)
{
    w[0] = __funnelshift_r(l, m, 16);
    w[1] = __funnelshift_r(l, m, 24);
    w[2] = m;
    w[3] = __funnelshift_r(m, r, 8);
}

#ifndef FAST_MINBLOCKS
#define FAST_MINBLOCKS 4
#endif
__global__ void __launch_bounds__(FAST_THREADS, FAST_MINBLOCKS)
fast_band_kernel(FrameGeom G, DevBuffers D)
{
    extern __shared__ uint32_t s_dyn[];               // per warp: S words [ihmax][32], then 32 lane-private u16 survivor lists
    orbx_pdl_prologue();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int item = blockIdx.x * FAST_WARPS + warp;
    if (item >= G.bands_per_frame) return;            // warps are independent: no block-level barrier below
    const BandDesc bd = D.bands[item];
    const int f = blockIdx.y;
    const LevelGeom& L = G.L[bd.lvl];
    const int Wr = L.maxBX - ORBX_MINB, Hr = L.maxBY - ORBX_MINB;     // region size, :1053-1059
    // interior rows / columns of the band in level coordinates (cell sub-image minus its 3-px FAST border)
    const int ys = ORBX_MINB + bd.ci * L.hCell + 3;
    const int ye = ORBX_MINB + min(bd.ci * L.hCell + L.hCell + 6, Hr) - 3;
    const int xs = ORBX_MINB + bd.cj0 * L.wCell + 3;
    const int xe = ORBX_MINB + min((bd.cj1 - 1) * L.wCell + L.wCell + 6, Wr) - 3;
    const int ih = ye - ys;
    const int xa = xs & ~3;                           // word aligned (XOFF and the pitch are multiples of 4)
    const int nw = (xe - xa + 3) >> 2;                // words per row, <= 32
    const bool active = lane < nw;
    const int wi = active ? lane : nw - 1;
    const int x = xa + 4 * wi;
    uint32_t* s_tile = s_dyn + warp * G.fast_warp_words;

    // ---------------- phase A: score tile ----------------
    {
        const uint32_t* p0 = reinterpret_cast<const uint32_t*>(
            L.pyr + (long long)f * L.plane_bytes + (long long)(ORBX_YOFF + ys - 3) * L.pitch + ORBX_XOFF + x);
        const int pw = L.pitch >> 2;
        const int klast = ih + 5;                     // last window row (ye + 2) relative to ys - 3
        uint32_t W[7][8];
#pragma unroll
        for (int i = 0; i < 6; i++) { const uint32_t* p = p0 + i * pw; fast_fill_row(W[i], __ldg(p - 1), __ldg(p), __ldg(p + 1)); }
        // three rows in flight ahead of the window, in statically indexed slots (no register moves that
        // would wait for the loads): slot k % 7 holds window row k = r + 6 .. r + 8 at step r
        uint32_t pf[7][3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const uint32_t* pn = p0 + min(6 + i, klast) * pw;
            pf[(6 + i) % 7][0] = __ldg(pn - 1); pf[(6 + i) % 7][1] = __ldg(pn); pf[(6 + i) % 7][2] = __ldg(pn + 1);
        }
        uint32_t* out = s_tile + wi;
#define FAST_STEP(PH)                                                                              \
        if (rb + PH < ih) {                                                                        \
            const int r = rb + PH;                                                                 \
            fast_fill_row(W[(PH + 6) % 7], pf[(PH + 6) % 7][0], pf[(PH + 6) % 7][1], pf[(PH + 6) % 7][2]); \
            { const uint32_t* pn = p0 + min(r + 9, klast) * pw;                                    \
              pf[(PH + 9) % 7][0] = __ldg(pn - 1); pf[(PH + 9) % 7][1] = __ldg(pn); pf[(PH + 9) % 7][2] = __ldg(pn + 1); } \
            const uint32_t sw = fast_score_word<PH>(W);                                            \
            if (active) out[r * 32] = sw;                                                          \
        }
#pragma unroll 1
        for (int rb = 0; rb < ih; rb += 7) {
            FAST_STEP(0) FAST_STEP(1) FAST_STEP(2) FAST_STEP(3) FAST_STEP(4) FAST_STEP(5) FAST_STEP(6)
        }
#undef FAST_STEP
    }
    __syncwarp();

    // two zero rows under the tile: the NMS loop below runs whole 3-row steps without a row guard
    s_tile[ih * 32 + lane] = 0;
    s_tile[(ih + 1) * 32 + lane] = 0;
    __syncwarp();

    // ---------------- phase B1: 3x3 NMS inside each cell, survivors appended to the list ----------------
    uint32_t sh[4];                                   // 8 * (cell of each of the thread's 4 pixels, 0.. inside the item)
    uint32_t ML_e = 0, MR_e = 0, ML_o = 0, MR_o = 0, V_e = 0, V_o = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int xp = x + q;
        const bool valid = active && xp >= xs && xp < xe;
        const int X3 = xp - ORBX_MINB - 3;
        const int j = valid ? X3 / L.wCell : bd.cj0;
        const int xin = X3 - j * L.wCell;
        const int cend = ORBX_MINB + min(j * L.wCell + L.wCell + 6, Wr) - 3;      // end of cell j's interior
        const bool lok = valid && xin > 0, rok = valid && xp + 1 < cend;
        sh[q] = 8u * (uint32_t)(j - bd.cj0);
        const uint32_t ln = (q & 2) ? 0xFFFF0000u : 0x0000FFFFu;
        if (q & 1) { if (lok) ML_o |= ln; if (rok) MR_o |= ln; if (valid) V_o |= ln & 0xFF00FF00u; }
        else       { if (lok) ML_e |= ln; if (rok) MR_e |= ln; if (valid) V_e |= ln & 0xFF00FF00u; }
    }
    const int iniTh = G.iniTh, minTh = min(G.minTh, G.iniTh);
    const uint32_t thr = (uint32_t)(((minTh - 1) << 8) | 0xFF) * 0x00010001u;
    // one entry per (lane, row) holding a survivor: bits 0,1,16,17 = px 0,1,2,3 survive, bits 8.. = row
    const uint32_t* mylist = s_tile + (ih + 2) * 32 + lane;
    int cnt = 0;                                      // entries of this lane
    {
        // the first word's left and the last word's right neighbour lie outside every cell interior and
        // are masked by ML/MR: those lanes re-read their own word instead of a guarded load
        const uint32_t* pm = s_tile + wi;
        const uint32_t* pl = pm - (wi > 0 ? 1 : 0);
        const uint32_t* pr = pm + (wi + 1 < nw ? 1 : 0);
        uint32_t R[3][4];
        const uint32_t wa0 = (uint32_t)__cvta_generic_to_shared(mylist);
        uint32_t wa = wa0, rr = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) R[0][k] = 0;
        nms_fill(R[1], pl[0], pm[0], pr[0]);
#define NMS_STEP(PH)                                                                                        \
        {                                                                                                   \
            nms_fill(R[(PH + 2) % 3], pl[(PH + 1) * 32], pm[(PH + 1) * 32], pr[(PH + 1) * 32]);             \
            uint32_t te, to;                                                                                \
            nms_row<PH>(R, thr, ML_e, MR_e, ML_o, MR_o, V_e, V_o, te, to);                                  \
            const uint32_t fl = __vminu2(te, 0x00010001u) + 2u * __vminu2(to, 0x00010001u);                 \
            list_push(wa, fl, fl | (rr + PH * 256u));                                                       \
        }
#pragma unroll 1
        for (int rb = 0; rb < ih; rb += 3) {          // rows >= ih are zero: no survivors there
            NMS_STEP(0) NMS_STEP(1) NMS_STEP(2)
            pl += 96; pm += 96; pr += 96; rr += 768u;
        }
#undef NMS_STEP
        cnt = (int)((wa - wa0) >> 7);
    }
    __syncwarp();

    // ---------------- phase B2: per-cell retry rule (:1118-1125) and emission ----------------
    // every lane walks its own entries; a pixel's cell is a per-lane constant (sh[q]), so the per-cell
    // counts are byte fields of one register per class, summed over the warp afterwards
    const int xrel = xa - ORBX_MINB + 4 * wi, yrel = ys - ORBX_MINB;
    const uint32_t* srow = s_tile + wi;
    uint32_t hic = 0, loc = 0;                         // per-cell byte counters: >= iniTh / in [minTh, iniTh)
    for (int i = 0; i < cnt; i++) {
        const uint32_t e = mylist[i * 32];
        const uint32_t sw = srow[((e >> 8) & 127u) * 32];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t bit = (e >> ((q & 1) + (q & 2) * 8)) & 1u;
            const uint32_t hi = ((sw >> (8 * q)) & 255u) >= (uint32_t)iniTh ? bit : 0u;
            hic += hi << sh[q];
            loc += (bit ^ hi) << sh[q];
        }
    }
    // a lane holds <= 4 * ih < 256 survivors per cell; the warp sums need 16-bit fields
    const uint32_t h02 = __reduce_add_sync(0xffffffffu, hic & 0x00FF00FFu);
    const uint32_t h13 = __reduce_add_sync(0xffffffffu, (hic >> 8) & 0x00FF00FFu);
    uint32_t zm = 0;                                   // byte c = 0xFF when cell c has no survivor >= iniTh
    if ((h02 & 0xFFFFu) == 0) zm |= 0x000000FFu;
    if ((h13 & 0xFFFFu) == 0) zm |= 0x0000FF00u;
    if ((h02 >> 16) == 0)     zm |= 0x00FF0000u;
    if ((h13 >> 16) == 0)     zm |= 0xFF000000u;
    const int mine = __dp4a(hic + (loc & zm), 0x01010101u, 0u);      // emitted by this lane
    const int total = __reduce_add_sync(0xffffffffu, mine);
    if (total == 0) return;
    int base = 0;
    if (lane == 0) base = atomicAdd(&D.candcnt[f * G.nlevels + bd.lvl], total);
    base = __shfl_sync(0xffffffffu, base, 0);
    // entry i of every lane is written side by side (neighbouring candidates then fall into different
    // quadtree nodes more often than a lane-major order would give).  Up to three pixels of a 4-px row can be
    // strict 3x3 maxima of their cells (two on either side of a cell boundary plus one two pixels away):
    // three ballots rank the lanes.
    orbx_key_t* out = D.cand + (long long)f * G.cand_per_frame + L.cand_base + base;
    const uint32_t lt = (1u << lane) - 1u;
    const int maxcnt = __reduce_max_sync(0xffffffffu, cnt);
    for (int i = 0; i < maxcnt; i++) {
        const uint32_t e = i < cnt ? mylist[i * 32] : 0u;
        const uint32_t r = (e >> 8) & 127u;
        const uint32_t sw = srow[r * 32];
        bool em[4];
        int c = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t sc = (sw >> (8 * q)) & 255u;
            const bool bit = (e >> ((q & 1) + (q & 2) * 8)) & 1u;
            em[q] = bit && (sc >= (uint32_t)iniTh || ((zm >> sh[q]) & 1u));
            c += em[q];
        }
        const uint32_t b1 = __ballot_sync(0xffffffffu, c >= 1), b2 = __ballot_sync(0xffffffffu, c >= 2),
                       b3 = __ballot_sync(0xffffffffu, c >= 3);
        int pos = __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
#pragma unroll
        for (int q = 0; q < 4; q++)
            if (em[q]) out[pos++] = MAKE_KEY(xrel + q, yrel + r, (sw >> (8 * q)) & 255u);
        out += __popc(b1) + __popc(b2) + __popc(b3);
    }
}

// ------------------------------------------------------------------------------------------------
// Strip variant.  fast_band_kernel gives a warp a run of WHOLE cells (<= 128 px): at cell widths of 35-42 px that is
// 3 cells = 27-32 of the 32 lanes, and some 2-cell items (18 lanes) - 74 % lane utilisation on 752x480.  Here a GROUP
// of whole cells spans up to FAST_WARPS * 128 px and is cut into 128-px strips that ignore the cell boundaries, one
// warp each, side by side in one CTA (88 % utilisation).  What crosses a strip edge goes through shared memory behind
// two block barriers: the 3x3 NMS reads the edge column of the neighbouring warp's score tile, and the per-cell
// "any survivor >= iniThFAST" flags of the retry rule (:1118-1125) are per CTA (one array per group, indexed by the
// cell column), set by every lane for its own pixels.  Phases A / B1 are those of fast_band_kernel.
// ------------------------------------------------------------------------------------------------
#define FS_MAXCOLS 128                 // cell columns per level the flag arrays hold

// HALVES = 2 (launches of a few frames, where one warp walking a whole cell row IS the kernel's duration): two warps
// per strip, warp w and w + FAST_WARPS share strip slot w - rows [0, r0) and [r0, ih) of the same score tile, r0 a
// multiple of 3 (the NMS loop's step), each with its own half of the survivor list.
template <int HALVES>
__global__ void __launch_bounds__(FAST_THREADS * HALVES, HALVES == 1 ? FAST_MINBLOCKS : 2)
fast_strip_kernel(FrameGeom G, DevBuffers D, int cta0)
{
    extern __shared__ uint32_t s_dyn[];               // per warp: S words [ihmax + 2][32], then the lane-private survivor lists
    __shared__ uint32_t s_ini[FAST_WARPS][FS_MAXCOLS];
    orbx_pdl_prologue();
    const int lane = threadIdx.x & 31, warp = (threadIdx.x >> 5) & (FAST_WARPS - 1), half = threadIdx.x >> 7;
    const StripDesc sd = D.strips[(cta0 + blockIdx.x) * FAST_WARPS + warp];
    const bool live = sd.lvl >= 0;
    const int f = blockIdx.y;
    const LevelGeom& L = G.L[live ? sd.lvl : 0];
    const int Hr = L.maxBY - ORBX_MINB, Wr = L.maxBX - ORBX_MINB;
    const int ys = ORBX_MINB + sd.ci * L.hCell + 3;
    const int ye = ORBX_MINB + min(sd.ci * L.hCell + L.hCell + 6, Hr) - 3;
    const int xs = sd.gxs, xe = sd.gxe;               // the group's scored columns
    const int ih = live ? ye - ys : 0;
    const int xa = (xs & ~3) + 128 * sd.k;            // word aligned (XOFF and the pitch are multiples of 4)
    const int nw = live ? min(32, (xe - xa + 3) >> 2) : 1;
    const bool active = live && lane < nw;
    const int wi = lane < nw ? lane : nw - 1;
    const int x = xa + 4 * wi;
    uint32_t* s_tile = s_dyn + warp * G.fast_warp_words;
    // this warp's rows [ra, rb_end) of the tile
    const int r0 = HALVES == 2 ? 3 * ((ih + 5) / 6) : ih;
    const int ra = half ? r0 : 0, re = half ? ih : min(r0, ih);
    if (half == 0) {
#pragma unroll
        for (int c = lane; c < FS_MAXCOLS; c += 32) s_ini[warp][c] = 0;
    }

    // ---------------- phase A: score tile ----------------
    if (live) {
        const int pw = L.pitch >> 2;
        // window rows are counted from the first row this warp scores
        const uint32_t* p0 = reinterpret_cast<const uint32_t*>(
            L.pyr + (long long)f * L.plane_bytes + (long long)(ORBX_YOFF + ys - 3 + ra) * L.pitch + ORBX_XOFF + x);
        const int nrow = re - ra;
        const int klast = nrow + 5;
        uint32_t W[7][8];
#pragma unroll
        for (int i = 0; i < 6; i++) { const uint32_t* p = p0 + i * pw; fast_fill_row(W[i], __ldg(p - 1), __ldg(p), __ldg(p + 1)); }
        uint32_t pf[7][3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const uint32_t* pn = p0 + min(6 + i, klast) * pw;
            pf[(6 + i) % 7][0] = __ldg(pn - 1); pf[(6 + i) % 7][1] = __ldg(pn); pf[(6 + i) % 7][2] = __ldg(pn + 1);
        }
        uint32_t* out = s_tile + ra * 32 + wi;
#define FAST_STEP(PH)                                                                              \
        if (rb + PH < nrow) {                                                                        \
            const int r = rb + PH;                                                                 \
            fast_fill_row(W[(PH + 6) % 7], pf[(PH + 6) % 7][0], pf[(PH + 6) % 7][1], pf[(PH + 6) % 7][2]); \
            { const uint32_t* pn = p0 + min(r + 9, klast) * pw;                                    \
              pf[(PH + 9) % 7][0] = __ldg(pn - 1); pf[(PH + 9) % 7][1] = __ldg(pn); pf[(PH + 9) % 7][2] = __ldg(pn + 1); } \
            const uint32_t sw = fast_score_word<PH>(W);                                            \
            if (active) out[r * 32] = sw;                                                          \
        }
#pragma unroll 1
        for (int rb = 0; rb < nrow; rb += 7) {
            FAST_STEP(0) FAST_STEP(1) FAST_STEP(2) FAST_STEP(3) FAST_STEP(4) FAST_STEP(5) FAST_STEP(6)
        }
#undef FAST_STEP
        // two zero rows under the tile: the NMS loop below runs whole 3-row steps without a row guard
        if (half == HALVES - 1) {
            s_tile[ih * 32 + lane] = 0;
            s_tile[(ih + 1) * 32 + lane] = 0;
        }
    }
    __syncthreads();                                  // every tile of the CTA is complete, the flags are cleared

    // ---------------- phase B1: 3x3 NMS inside each cell, survivors appended to the list ----------------
    int cq[4];                                        // cell column of each of the thread's 4 pixels
    uint32_t ML_e = 0, MR_e = 0, ML_o = 0, MR_o = 0, V_e = 0, V_o = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int xp = x + q;
        const bool valid = active && xp >= xs && xp < xe;
        const int X3 = xp - ORBX_MINB - 3;
        const int j = valid ? X3 / L.wCell : 0;
        const int xin = X3 - j * L.wCell;
        const int cend = ORBX_MINB + min(j * L.wCell + L.wCell + 6, Wr) - 3;      // end of cell j's interior
        const bool lok = valid && xin > 0, rok = valid && xp + 1 < cend;
        cq[q] = j;
        const uint32_t ln = (q & 2) ? 0xFFFF0000u : 0x0000FFFFu;
        if (q & 1) { if (lok) ML_o |= ln; if (rok) MR_o |= ln; if (valid) V_o |= ln & 0xFF00FF00u; }
        else       { if (lok) ML_e |= ln; if (rok) MR_e |= ln; if (valid) V_e |= ln & 0xFF00FF00u; }
    }
    const int iniTh = G.iniTh, minTh = min(G.minTh, G.iniTh);
    const uint32_t thr = (uint32_t)(((minTh - 1) << 8) | 0xFF) * 0x00010001u;
    const uint32_t* mylist = s_tile + (ih + 2 + ra) * 32 + lane;     // the second half appends behind the first half's rows
    int cnt = 0;
    if (live && ra < re) {
        // across a strip edge the neighbour column is word 31 / word 0 of the adjacent warp's tile (same cell row, same
        // ih); at the group's ends it lies outside every cell interior and is masked: those lanes re-read their own word
        const uint32_t* pm = s_tile + ra * 32 + wi;
        const uint32_t* pl = wi > 0 ? pm - 1 : (sd.k > 0 ? pm - G.fast_warp_words + 31 : pm);
        const uint32_t* pr = wi + 1 < nw ? pm + 1 : ((wi == 31 && sd.k + 1 < sd.nstrips) ? pm + G.fast_warp_words - 31 : pm);
        uint32_t R[3][4];
        const uint32_t wa0 = (uint32_t)__cvta_generic_to_shared(mylist);
        uint32_t wa = wa0, rr = (uint32_t)ra * 256u;
#pragma unroll
        for (int k = 0; k < 4; k++) R[0][k] = 0;
        if (ra > 0) nms_fill(R[0], pl[-32], pm[-32], pr[-32]);         // the row above belongs to the first half
        nms_fill(R[1], pl[0], pm[0], pr[0]);
#define NMS_STEP(PH)                                                                                        \
        {                                                                                                   \
            nms_fill(R[(PH + 2) % 3], pl[(PH + 1) * 32], pm[(PH + 1) * 32], pr[(PH + 1) * 32]);             \
            uint32_t te, to;                                                                                \
            nms_row<PH>(R, thr, ML_e, MR_e, ML_o, MR_o, V_e, V_o, te, to);                                  \
            const uint32_t fl = __vminu2(te, 0x00010001u) + 2u * __vminu2(to, 0x00010001u);                 \
            list_push(wa, fl, fl | (rr + PH * 256u));                                                       \
        }
#pragma unroll 1
        for (int rb = ra; rb < re; rb += 3) {
            NMS_STEP(0) NMS_STEP(1) NMS_STEP(2)
            pl += 96; pm += 96; pr += 96; rr += 768u;
        }
#undef NMS_STEP
        cnt = (int)((wa - wa0) >> 7);
    }
    __syncwarp();

    // ---------------- phase B2: per-cell retry rule (:1118-1125) and emission ----------------
    const int xrel = xa - ORBX_MINB + 4 * wi, yrel = ys - ORBX_MINB;
    const uint32_t* srow = s_tile + wi;        // list entries carry absolute tile rows
    uint32_t* ini = s_ini[live ? sd.first : 0];       // the group's flags: cell column -> a survivor >= iniTh exists
    for (int i = 0; i < cnt; i++) {
        const uint32_t e = mylist[i * 32];
        const uint32_t sw = srow[((e >> 8) & 127u) * 32];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t bit = (e >> ((q & 1) + (q & 2) * 8)) & 1u;
            if (bit && ((sw >> (8 * q)) & 255u) >= (uint32_t)iniTh) ini[cq[q]] = 1u;
        }
    }
    __syncthreads();                                  // the flags of every group of the CTA are final
    if (!live) return;
    uint32_t zq = 0;                                  // bit q: the cell of pixel q has no survivor >= iniTh
#pragma unroll
    for (int q = 0; q < 4; q++) if (ini[cq[q]] == 0u) zq |= 1u << q;
    int mine = 0;
    for (int i = 0; i < cnt; i++) {
        const uint32_t e = mylist[i * 32];
        const uint32_t sw = srow[((e >> 8) & 127u) * 32];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t bit = (e >> ((q & 1) + (q & 2) * 8)) & 1u;
            mine += (int)(bit & ((((sw >> (8 * q)) & 255u) >= (uint32_t)iniTh ? 1u : 0u) | (zq >> q)));
        }
    }
    const int total = __reduce_add_sync(0xffffffffu, mine);
    if (total == 0) return;
    int base = 0;
    if (lane == 0) base = atomicAdd(&D.candcnt[f * G.nlevels + sd.lvl], total);
    base = __shfl_sync(0xffffffffu, base, 0);
    orbx_key_t* out = D.cand + (long long)f * G.cand_per_frame + L.cand_base + base;
    const uint32_t lt = (1u << lane) - 1u;
    const int maxcnt = __reduce_max_sync(0xffffffffu, cnt);
    for (int i = 0; i < maxcnt; i++) {
        const uint32_t e = i < cnt ? mylist[i * 32] : 0u;
        const uint32_t r = (e >> 8) & 127u;
        const uint32_t sw = srow[r * 32];
        bool em[4];
        int c = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t sc = (sw >> (8 * q)) & 255u;
            const bool bit = (e >> ((q & 1) + (q & 2) * 8)) & 1u;
            em[q] = bit && (sc >= (uint32_t)iniTh || ((zq >> q) & 1u));
            c += em[q];
        }
        const uint32_t b1 = __ballot_sync(0xffffffffu, c >= 1), b2 = __ballot_sync(0xffffffffu, c >= 2),
                       b3 = __ballot_sync(0xffffffffu, c >= 3);
        int pos = __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
#pragma unroll
        for (int q = 0; q < 4; q++)
            if (em[q]) out[pos++] = MAKE_KEY(xrel + q, yrel + r, (sw >> (8 * q)) & 255u);
        out += __popc(b1) + __popc(b2) + __popc(b3);
    }
}

// Work items of the strip kernel: per level and cell row, greedy runs of whole valid cells spanning <= FAST_WARPS * 128
// px ("groups"), each cut into 128-px strips; the groups are packed into CTAs of FAST_WARPS warps (largest first, first
// fit).  Returns the CTAs per frame; 0 when the geometry does not fit (the band kernel serves it then).
int fast_build_strips(const FrameGeom& G, std::vector<StripDesc>* out, int* lvl_cta)
{
    struct Group { int lvl, ci, xs, xe, n; };
    std::vector<Group> groups;
    out->clear();
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        const int Wr = L.maxBX - ORBX_MINB, Hr = L.maxBY - ORBX_MINB;
        if (L.nCols > FS_MAXCOLS || L.maxBX > 32000) return 0;
        int ncv = 0, nrv = 0;
        while (ncv < L.nCols && ncv * L.wCell <= Wr - 7) ncv++;
        while (nrv < L.nRows && nrv * L.hCell <= Hr - 7) nrv++;
        auto xs_of = [&](int j0) { return ORBX_MINB + j0 * L.wCell + 3; };
        auto xe_of = [&](int j1) { return ORBX_MINB + ((j1 - 1) * L.wCell + L.wCell + 6 < Wr ? (j1 - 1) * L.wCell + L.wCell + 6 : Wr) - 3; };
        auto span = [&](int j0, int j1) { return xe_of(j1) - (xs_of(j0) & ~3); };
        for (int i = 0; i < nrv; i++) {
            int j0 = 0;
            while (j0 < ncv) {
                int j1 = j0 + 1;
                while (j1 < ncv && span(j0, j1 + 1) <= FAST_WARPS * 128) j1++;
                const int sp = span(j0, j1);
                if (sp > FAST_WARPS * 128 || sp <= 0) return 0;
                groups.push_back({l, i, xs_of(j0), xe_of(j1), (sp + 127) / 128});
                j0 = j1;
            }
        }
    }
    // first fit of the groups of [lo, hi) (largest first) into CTAs appended to *out; returns the CTAs added
    auto pack = [&](std::vector<Group> gs) {
        std::stable_sort(gs.begin(), gs.end(), [](const Group& a, const Group& b) { return a.n > b.n; });
        std::vector<int> used;                         // warps taken per CTA
        std::vector<std::vector<StripDesc>> ctas;
        for (const Group& g : gs) {
            size_t c = 0;
            while (c < used.size() && used[c] + g.n > FAST_WARPS) c++;
            if (c == used.size()) { used.push_back(0); ctas.emplace_back(); }
            for (int k = 0; k < g.n; k++) {
                StripDesc sd; sd.lvl = (short)g.lvl; sd.ci = (short)g.ci; sd.gxs = (short)g.xs; sd.gxe = (short)g.xe;
                sd.k = (short)k; sd.first = (short)used[c]; sd.nstrips = (short)g.n; sd.pad = 0;
                ctas[c].push_back(sd);
            }
            used[c] += g.n;
        }
        for (auto& v : ctas) {
            StripDesc none = {}; none.lvl = -1;
            while ((int)v.size() < FAST_WARPS) v.push_back(none);
            out->insert(out->end(), v.begin(), v.end());
        }
        return (int)ctas.size();
    };
    const int total = pack(groups);                    // plan 1: all levels packed together (batches)
    int at = total;                                    // plan 2: level by level (the per-level launches of a single-frame graph)
    for (int l = 0; l < G.nlevels; l++) {
        std::vector<Group> gl;
        for (const Group& g : groups) if (g.lvl == l) gl.push_back(g);
        lvl_cta[l] = at;
        at += pack(gl);
    }
    lvl_cta[G.nlevels] = at;
    return total;
}

// FAST work items of one frame: per level and cell row, runs of whole cells spanning <= 128 px (one warp,
// 4 px per lane).  Cells the reference skips (:1085,1102: iniX >= maxBorderX-6, iniY >= maxBorderY-3, or
// narrower than the 7-px FAST footprint) are left out.  Returns the kernel's dynamic shared memory and
// sets *warp_words (32-bit words of shared memory per warp).
int fast_build_bands(const FrameGeom& G, std::vector<BandDesc>* out, int* warp_words)
{
    int ihmax = 1;
    out->clear();
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        const int Wr = L.maxBX - ORBX_MINB, Hr = L.maxBY - ORBX_MINB;
        int ncv = 0, nrv = 0;
        while (ncv < L.nCols && ncv * L.wCell <= Wr - 7) ncv++;
        while (nrv < L.nRows && nrv * L.hCell <= Hr - 7) nrv++;
        int per = (FAST_MAXSPAN - 3) / L.wCell;
        if (per > FAST_MAXCELLS) per = FAST_MAXCELLS;
        if (per < 1 || ncv < 1) continue;
        const int nseg = (ncv + per - 1) / per, q = ncv / nseg, rem = ncv % nseg;   // balanced split
        for (int i = 0; i < nrv; i++) {
            const int ih = (L.hCell + 6 < Hr - i * L.hCell ? L.hCell + 6 : Hr - i * L.hCell) - 6;
            if (ih > ihmax) ihmax = ih;
            int j0 = 0;
            for (int sgm = 0; sgm < nseg; sgm++) {
                const int j1 = j0 + q + (sgm < rem ? 1 : 0);
                BandDesc b; b.lvl = (short)l; b.ci = (short)i; b.cj0 = (short)j0; b.cj1 = (short)j1;
                out->push_back(b);
                j0 = j1;
            }
        }
    }
    if (ihmax > 127) return -1;                    // row field of a list entry, byte counters of phase B2
    *warp_words = (ihmax + 2) * 32 + ihmax * 32;   // S tile + 2 zero rows, then one 32-bit list entry per (lane, row)
    return FAST_WARPS * *warp_words * 4;
}

static bool fast_strips_enabled()
{
    static const bool strips = [] { const char* e = getenv("ORBX_FAST_STRIPS"); return !e || atoi(e) != 0; }();
    return strips;
}

// per-level launches (launch_fast(..., level)) exist for the strip kernel only
bool fast_per_level(const FrameGeom& G) { return fast_strips_enabled() && G.strip_ctas > 0; }

int launch_fast(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s, int level)
{
    // D.candcnt was zeroed by pyr_level0_kernel
    int smem = G.fast_smem;
    // experiment hook: ORBX_FAST_SMEM pads the CTA's shared memory, i.e. caps the FAST CTAs per SM
    static const int pad = [] { const char* e = getenv("ORBX_FAST_SMEM"); return e ? atoi(e) : 0; }();
    if (pad > smem) smem = pad;
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(fast_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (fast_per_level(G)) {
        const bool two = nframes * G.nlevels <= 64;        // few frames: two warps per strip
        if (smem + (int)sizeof(uint32_t) * FAST_WARPS * FS_MAXCOLS > 48 * 1024)      // static flags + dynamic tiles
            cudaFuncSetAttribute(two ? fast_strip_kernel<2> : fast_strip_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (level >= 0) {
            const int c0 = G.strip_lvl_cta[level], nc = G.strip_lvl_cta[level + 1] - c0;
            if (nc <= 0) return 0;
            if (two) orbx_launch(fast_strip_kernel<2>, dim3(nc, nframes), dim3(2 * FAST_THREADS), (size_t)smem, s, G, D, c0);
            else orbx_launch(fast_strip_kernel<1>, dim3(nc, nframes), dim3(FAST_THREADS), (size_t)smem, s, G, D, c0);
            return 1;
        }
        if (two) orbx_launch(fast_strip_kernel<2>, dim3(G.strip_ctas, nframes), dim3(2 * FAST_THREADS), (size_t)smem, s, G, D, 0);
        else orbx_launch(fast_strip_kernel<1>, dim3(G.strip_ctas, nframes), dim3(FAST_THREADS), (size_t)smem, s, G, D, 0);
        return 1;
    }
    const dim3 grid((G.bands_per_frame + FAST_WARPS - 1) / FAST_WARPS, nframes);
    orbx_launch(fast_band_kernel, grid, dim3(FAST_THREADS), (size_t)smem, s, G, D);
    return 1;
}

// scalar restatement kept for the unit-test entry below
__device__ __forceinline__ int fast_score(const uint8_t* p, int pitch, int t)
{
    const int v = p[0];
    int r[16];
    r[0] = p[3 * pitch];       r[1] = p[3 * pitch + 1];   r[2] = p[2 * pitch + 2];   r[3] = p[pitch + 3];
    r[4] = p[3];               r[5] = p[-pitch + 3];      r[6] = p[-2 * pitch + 2];  r[7] = p[-3 * pitch + 1];
    r[8] = p[-3 * pitch];      r[9] = p[-3 * pitch - 1];  r[10] = p[-2 * pitch - 2]; r[11] = p[-pitch - 3];
    r[12] = p[-3];             r[13] = p[pitch - 3];      r[14] = p[2 * pitch - 2];  r[15] = p[3 * pitch - 1];
    const int vb = v + t, vd = v - t;
    bool okb = true, okd = true;
#pragma unroll
    for (int k = 0; k < 8; k++) {   // a 9-arc holds one pixel of every opposite pair
        okb &= (r[k] > vb) | (r[k + 8] > vb);
        okd &= (r[k] < vd) | (r[k + 8] < vd);
    }
    if (!(okb | okd)) return 0;
    // d = v - ring (darker arcs), e = ring - v (brighter arcs).  A = max over the 16 arcs of min(d),
    // Bn = max over arcs of min(e) = -(OpenCV's b0).  Written negation-free on purpose: ptxas 12.9 for
    // sm_100a drops the negation of `max(max(A, -B), t)` when it fuses it into VIMNMX3
    // (tools/scratch/ptxas_vimnmx3_neg_repro.cu), which silently turned every score into max(A, B, t)-1.
    int d[16], e[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { d[k] = v - r[k]; e[k] = r[k] - v; }
    // circular sliding min over windows of 9 by doubling (2,4,8,+1)
    int dn2[16], en2[16], dn4[16], en4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { dn2[k] = min(d[k], d[(k + 1) & 15]); en2[k] = min(e[k], e[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; k++) { dn4[k] = min(dn2[k], dn2[(k + 2) & 15]); en4[k] = min(en2[k], en2[(k + 2) & 15]); }
    int A = -256, Bn = -256;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        A = max(A, min(min(dn4[k], dn4[(k + 4) & 15]), d[(k + 8) & 15]));
        Bn = max(Bn, min(min(en4[k], en4[(k + 4) & 15]), e[(k + 8) & 15]));
    }
    if (!(A > t || Bn > t)) return 0;
    return max(max(A, Bn), t) - 1;
}

// ---- unit-test entry: plain per-pixel score map of a host image (no cells, no NMS) ----
__global__ void fast_score_map_kernel(const uint8_t* img, int rows, int cols, int pitch, int th, uint8_t* out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= cols || y >= rows) return;
    int s = 0;
    if (x >= 3 && y >= 3 && x < cols - 3 && y < rows - 3) s = fast_score(img + (long long)y * pitch + x, pitch, th);
    out[(long long)y * cols + x] = (uint8_t)s;
}

extern "C" int orbx_debug_fast_score_map(int device, const uint8_t* img, int rows, int cols, size_t step, int th, uint8_t* out)
{
    uint8_t *d_in = nullptr, *d_out = nullptr;
    if (cudaSetDevice(device) != cudaSuccess) return ORBX_ERR_CUDA;
    if (cudaMalloc(&d_in, (size_t)rows * cols) != cudaSuccess || cudaMalloc(&d_out, (size_t)rows * cols) != cudaSuccess) return ORBX_ERR_CUDA;
    cudaMemcpy2D(d_in, cols, img, step, cols, rows, cudaMemcpyHostToDevice);
    dim3 b(32, 8), g((cols + 31) / 32, (rows + 7) / 8);
    fast_score_map_kernel<<<g, b>>>(d_in, rows, cols, cols, th, d_out);
    cudaError_t e = cudaMemcpy(out, d_out, (size_t)rows * cols, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out);
    return e == cudaSuccess ? ORBX_OK : ORBX_ERR_CUDA;
}
