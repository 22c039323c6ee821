// octree.cu — deterministic GPU rewrite of ORBextractor::DistributeOctTree
// (reference src/ORBextractor.cc:688-1034, ExtractorNode::DivideNode :602-674).
//
// The reference walks a std::list of nodes holding std::vector<KeyPoint> copies.  Here one CTA owns
// one (frame, level): node = list POSITION, every candidate carries `4*node + child` (the child of
// its node it would fall into).  One round of the reference's loop =
//   node step   : from the per-child counts decide which nodes split (all expandable ones, or — in
//                 the ordered final phase :909-992 — the largest first until the list reaches N),
//                 compute by scans where children / surviving nodes land in the new list (children
//                 are push_front'ed n1..n4, so the last split node's n4 ends up first), build the new
//                 node table and a (node, child) -> new position map;
//   candidate step (ONE pass per round): relabel through that map, find the child inside the new
//                 node and add it to the next round's per-child counts.
// The ordered phase sorts (count desc, creation desc == position asc): the canonical tie-break of
// the oracle (the reference compares heap pointers, :927).  Winner per node = max response, ties to
// the first keypoint in the reference's vector order = smallest canonical rank (cell row, cell col,
// y, x) since the FAST kernel emits candidates unordered.
// Latency-bound integer work (no roofline); reported as a share of the step.
#include "orbx_internal.h"
#include <cstdlib>

#define OCT_TABLES 25     // int arrays of node_cap entries in the per-CTA node scratch
// "Fast rounds": the subdivision GEOMETRY of the reference's quadtree does not depend on the data (a node's children
// are its four quadrants, ceil-halved) - only which nodes exist and when the splitting stops does.  So ONE candidate
// pass bins every candidate by its path through the implicit tree down to depth OCT_HDEPTH; the per-child counts a
// round needs are then table lookups, and the per-round candidate pass disappears as long as no node deeper than
// OCT_HDEPTH has to split (dense frames: never - 4^5 nodes per root exceed every per-level quota).  A level that does
// go deeper (few, clustered keypoints) materialises the per-candidate node labels once and continues with the
// relabel-per-round scheme.  Shared-memory variant only, <= OCT_HROOTS root nodes.
#define OCT_HDEPTH 5
#define OCT_HROOTS 4
#define OCT_HNODES 1365   // (4^(OCT_HDEPTH+1) - 1) / 3 implicit nodes per root

// exclusive scan of in[0..m) into out[0..m) (out may alias in); returns the total to all threads.
template <int NT>
__device__ int block_scan_excl_t(const int* in, int* out, int m, int* s_warp)
{
    const int T = NT;
    const int chunk = (m + T - 1) / T;
    const int b = min(m, (int)threadIdx.x * chunk), e = min(m, b + chunk);
    int sum = 0;
    for (int i = b; i < e; i++) sum += in[i];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += v;
    }
    __syncthreads();      // protect s_warp reuse
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < NT / 32; w++) {
        const int v = s_warp[w];
        if (w < (int)(threadIdx.x >> 5)) base += v;
        total += v;
    }
    int run = base + incl - sum;
    for (int i = b; i < e; i++) { const int v = in[i]; out[i] = run; run += v; }
    __syncthreads();
    return total;
}

// node bounds are packed lo | hi << 16 (coordinates relative to minBorder, < 32768)
__device__ __forceinline__ int child_of(int x, int y, int bx, int by)
{
    const int x0 = bx & 0xFFFF, x1 = bx >> 16, y0 = by & 0xFFFF, y1 = by >> 16;
    const int mx = x0 + ((x1 - x0 + 1) >> 1);     // ceil(float(w)/2), :608
    const int my = y0 + ((y1 - y0 + 1) >> 1);     // :609
    return (x < mx) ? ((y < my) ? 0 : 2) : ((y < my) ? 1 : 3);   // :651-661
}

// position of a candidate in the reference's vToDistributeKeys order (cell row-major, raster inside a cell)
// The two divisions by the cell size are multiplications by ceil(2^32 / d) (mw, mh; one division per thread at kernel start): exact for
// x < 2^16 and 2 <= d <= 128, since x * (ceil(2^32 / d) * d - 2^32) < 2^16 * 128 < 2^32.
__device__ __forceinline__ int cand_rank(orbx_key_t key, const LevelGeom& L, uint32_t mw, uint32_t mh)
{
    const int X3 = (int)KEY_X(key) - 3, Y3 = (int)KEY_Y(key) - 3;
    const int cj = (int)__umulhi((uint32_t)X3, mw), ci = (int)__umulhi((uint32_t)Y3, mh);
    return ((ci * L.nCols + cj) << 14) | ((Y3 - ci * L.hCell) << 7) | (X3 - cj * L.wCell);
}

// Candidate passes are chains of dependent lookups (global label -> shared tables -> shared atomic); four
// candidates per thread are fetched before any of them is processed so that their global loads overlap.
#define OCT_UNROLL 4
template <int NT, class F>
__device__ __forceinline__ void for_cands_t(const orbx_key_t* __restrict__ cand, const uint32_t* kcode, int n, F body)
{
    for (int i0 = threadIdx.x; i0 < n; i0 += OCT_UNROLL * NT) {
        orbx_key_t key[OCT_UNROLL]; uint32_t code[OCT_UNROLL];
#pragma unroll
        for (int u = 0; u < OCT_UNROLL; u++) {
            const int i = i0 + u * NT;
            key[u] = i < n ? cand[i] : (orbx_key_t)0;
            code[u] = (kcode && i < n) ? kcode[i] : 0u;
        }
#pragma unroll
        for (int u = 0; u < OCT_UNROLL; u++) {
            const int i = i0 + u * NT;
            if (i < n) body(i, key[u], (int)code[u]);
        }
    }
}

// levels served by one launch (levels of similar table size share a launch: dynamic shared memory is per launch)
struct OctLevels { int n; int lvl[ORBX_MAX_LEVELS]; };

template <bool SMEM, int NT>
__global__ void __launch_bounds__(NT)
octree_kernel(FrameGeom G, DevBuffers D, OctLevels LV)
{
#define OCT_THREADS NT
#define block_scan_excl block_scan_excl_t<NT>
#define for_cands for_cands_t<NT>
    extern __shared__ int dyn_smem[];
    __shared__ int s_warp[NT / 32 + 1];
    __shared__ int s_val[4];
    __shared__ int s_thr[OCT_HROOTS + 1], s_rx[OCT_HROOTS + 1];

    orbx_pdl_prologue();
    // blockIdx.x = frame (dispatched first), blockIdx.y = level in the group's order (largest level first): the long level-0 CTAs of
    // ALL frames start before any small level does, so the launch ends with short CTAs instead of waiting for a late level-0 one
    const int lvl = LV.lvl[blockIdx.y], f = blockIdx.x;
    const LevelGeom& L = G.L[lvl];
    const int tid = threadIdx.x;
    const int cap = L.node_cap;

    const orbx_key_t* __restrict__ cand = D.cand + (long long)f * G.cand_per_frame + L.cand_base;
    uint32_t* kcode = D.candnode + (long long)f * G.cand_per_frame + L.cand_base;     // 4*node + child
    int* T = SMEM ? dyn_smem : (D.nodetab + ((long long)f * G.nlevels + lvl) * D.nodetab_per_frame);
    int* BX[2] = {T, T + 3 * cap};               // x0 | x1 << 16
    int* BY[2] = {T + cap, T + 4 * cap};         // y0 | y1 << 16
    int* CN[2] = {T + 2 * cap, T + 5 * cap};     // keypoints in the node
    int* CC[2] = {T + 6 * cap, T + 10 * cap};    // keypoints per child (4 per node)
    int* NP4 = T + 14 * cap;                     // (old node, child) -> new list position
    int* aA = T + 18 * cap; int* aB = T + 19 * cap; int* aC = T + 20 * cap; int* aD = T + 21 * cap; int* aE = T + 22 * cap;
    int* PD[2] = {T + 23 * cap, T + 24 * cap};   // depth << 24 | path of the node in the implicit tree (fast rounds)
    int* Hh = T + OCT_TABLES * cap;              // per implicit node: candidates inside (later: list-position markers)
    bool fast = SMEM && L.nIni <= OCT_HROOTS;
    const int nI = L.nIni;
    // first bin of depth d: nI * (4^d - 1) / 3
#define HOFF(d) (nI * (((1 << (2 * (d))) - 1) / 3))

    const int n = min(D.candcnt[f * G.nlevels + lvl], L.cand_cap);   // compacted (unordered) by fast_band_kernel
    const int N = L.N;
    const int H = L.maxBY - ORBX_MINB;

    // ---- roots, :695-763 ----
    for (int i = tid; i < L.nIni; i += OCT_THREADS) {
        const int x0 = (int)__fmul_rn(L.hX, (float)i), x1 = (int)__fmul_rn(L.hX, (float)(i + 1));
        BX[0][i] = x0 | (x1 << 16); BY[0][i] = H << 16; CN[0][i] = 0;
    }
    __syncthreads();
    if (fast) {
        for (int i = tid; i < nI * OCT_HNODES; i += OCT_THREADS) Hh[i] = 0;
        // root of a keypoint = min((int)(x / hX), nIni - 1) (:740) without a division per candidate: x is an integer and the quotient
        // is monotone, so root r <=> s_thr[r] <= x < s_thr[r + 1] with s_thr[i] = the smallest x whose float quotient reaches i
        if (tid <= nI) {
            s_rx[tid] = (int)__fmul_rn(L.hX, (float)tid);                   // root bounds, :698-731
            int c = (int)ceilf(__fmul_rn(L.hX, (float)tid));
            while (c > 0 && (int)__fdiv_rn((float)(c - 1), L.hX) >= tid) c--;
            while ((int)__fdiv_rn((float)c, L.hX) < tid) c++;
            s_thr[tid] = (tid == 0) ? 0 : (tid == nI ? 0x7FFFFFFF : c);
        }
        __syncthreads();
        int* Hd = Hh + HOFF(OCT_HDEPTH);
        for_cands(cand, nullptr, n, [&](int i, orbx_key_t k, int) {
            const int x = KEY_X(k), y = KEY_Y(k);
            int r = 0;
#pragma unroll
            for (int q = 1; q < OCT_HROOTS; q++) r += (q < nI) && (x >= s_thr[q]);
            int x0 = s_rx[r], x1 = s_rx[r + 1], y0 = 0, y1 = H;
            int path = r;
#pragma unroll
            for (int d = 0; d < OCT_HDEPTH; d++) {                   // DivideNode geometry, :602-661
                const int mx = x0 + ((x1 - x0 + 1) >> 1), my = y0 + ((y1 - y0 + 1) >> 1);
                const int cx = x >= mx, cy = y >= my;
                path = 4 * path + 2 * cy + cx;
                x0 = cx ? mx : x0; x1 = cx ? x1 : mx; y0 = cy ? my : y0; y1 = cy ? y1 : my;
            }
            kcode[i] = path;
            atomicAdd(&Hd[path], 1);
        });
        __syncthreads();
        for (int d = OCT_HDEPTH - 1; d >= 0; d--) {
            int* Hp = Hh + HOFF(d); const int* Hc = Hh + HOFF(d + 1);
            for (int q = tid; q < (nI << (2 * d)); q += OCT_THREADS) Hp[q] = Hc[4 * q] + Hc[4 * q + 1] + Hc[4 * q + 2] + Hc[4 * q + 3];
            __syncthreads();
        }
        for (int i = tid; i < nI; i += OCT_THREADS) CN[0][i] = Hh[i];
    } else {
        for_cands(cand, nullptr, n, [&](int i, orbx_key_t k, int) {
            int r = (int)__fdiv_rn((float)KEY_X(k), L.hX);           // vpIniNodes[kp.pt.x/hX], :740
            r = min(r, L.nIni - 1);
            kcode[i] = 4 * r;
            atomicAdd(&CN[0][r], 1);
        });
    }
    __syncthreads();
    for (int i = tid; i < L.nIni; i += OCT_THREADS) aA[i] = CN[0][i] > 0;
    __syncthreads();
    int size = block_scan_excl(aA, aB, L.nIni, s_warp);              // empty roots are erased, :746-763
    for (int i = tid; i < L.nIni; i += OCT_THREADS) {
        const int p = aB[i];
        if (CN[0][i] > 0) {
            BX[1][p] = BX[0][i]; BY[1][p] = BY[0][i]; CN[1][p] = CN[0][i]; PD[1][p] = i;     // depth 0, path = root index
            *reinterpret_cast<int4*>(CC[1] + 4 * p) = make_int4(0, 0, 0, 0);
        }
        *reinterpret_cast<int4*>(NP4 + 4 * i) = make_int4(p, p, p, p);
    }
    int cur = 1;
    bool final_phase = false;
    bool finish = (n == 0);        // empty list: first iteration leaves size == prevSize, :889
    bool more = !finish;
    __syncthreads();

#pragma unroll 1
    while (more) {
        bool materialise = false;
        if (fast) {
            if (finish) break;                                             // no labels to refresh: the winner pass looks nodes up
            // ---- fast round: per-child counts of the expandable nodes straight from the implicit-tree bins ----
            if (tid == 0) s_val[2] = 0;
            __syncthreads();
            const int* cn = CN[cur]; int* cc = CC[cur]; const int* pd = PD[cur];
            int deep = 0;
            for (int i = tid; i < size; i += OCT_THREADS) {
                if (cn[i] > 1) {
                    const int d = pd[i] >> 24, q = pd[i] & 0xFFFFFF;
                    if (d >= OCT_HDEPTH) deep = 1;
                    else {
                        const int* Hc = Hh + HOFF(d + 1) + 4 * q;
                        *reinterpret_cast<int4*>(cc + 4 * i) = make_int4(Hc[0], Hc[1], Hc[2], Hc[3]);
                    }
                }
            }
            if (deep) s_val[2] = 1;
            __syncthreads();
            if (s_val[2]) {
                // a node below the binned depth has to split: mark every list node in the bins (-1 - position), let each
                // candidate find its node, and continue with per-round relabelling
                fast = false; materialise = true;
                for (int i = tid; i < size; i += OCT_THREADS) {
                    Hh[HOFF(pd[i] >> 24) + (pd[i] & 0xFFFFFF)] = -1 - i;
                    *reinterpret_cast<int4*>(cc + 4 * i) = make_int4(0, 0, 0, 0);
                }
                __syncthreads();
            }
        }
        if (!fast) {
            // ---- candidate step: relabel, child inside the new node, next round's per-child counts ----
            const int* bx = BX[cur]; const int* by = BY[cur]; const int* cn = CN[cur]; int* cc = CC[cur];
            for_cands(cand, kcode, n, [&](int i, orbx_key_t k, int code) {
                int p;
                if (materialise) {
                    p = 0;
#pragma unroll
                    for (int d = OCT_HDEPTH; d >= 0; d--) {
                        const int v = Hh[HOFF(d) + (code >> (2 * (OCT_HDEPTH - d)))];
                        if (v < 0) { p = -1 - v; break; }
                    }
                } else p = NP4[code];
                int c = 0;
                if (!finish && cn[p] > 1) {
                    c = child_of(KEY_X(k), KEY_Y(k), bx[p], by[p]);
                    atomicAdd(&cc[4 * p + c], 1);
                }
                kcode[i] = 4 * p + c;
            });
            __syncthreads();
            if (finish) break;
        }
        // ---- node step ----
        const int prevSize = size;
        const int nxt = cur ^ 1;
        const int* cn = CN[cur]; const int* cc = CC[cur];
        int nsplit, totalNe;
        // on exit of either branch: aB[i] = first child slot of a node that splits (else -1),
        //                           aC[i] = number of non-splitting nodes before i
        if (!final_phase) {
            // every expandable node splits, in list order (:800-886); e << 16 | ne scanned in one go
            for (int i = tid; i < size; i += OCT_THREADS) {
                int v = 0;
                if (cn[i] > 1) {
                    const int4 c4 = *reinterpret_cast<const int4*>(cc + 4 * i);
                    v = 0x10000 | ((c4.x > 0) + (c4.y > 0) + (c4.z > 0) + (c4.w > 0));
                }
                aA[i] = v;
            }
            __syncthreads();
            const int tot = block_scan_excl(aA, aD, size, s_warp);
            nsplit = tot >> 16; totalNe = tot & 0xFFFF;
            for (int i = tid; i < size; i += OCT_THREADS) {
                const int v = aA[i], pre = aD[i];
                // later-processed nodes sit in front: slot = sum of ne over the nodes after i
                aB[i] = (v >> 16) ? totalNe - (pre & 0xFFFF) - (v & 0xFFFF) : -1;
                aC[i] = i - (pre >> 16);
            }
            __syncthreads();
        } else {
            // ordered phase (:909-992): expandable nodes by (size desc, creation desc == position asc).  The compacted
            // list e = 0..E-1 of expandable nodes is in list order, so "position asc" is "e asc": the sort key is the bare
            // count, ties go to the smaller e - no bit field limits the list length or the count.  NP4 is free between the
            // candidate step and the table build: it maps e back to the list position.
            for (int i = tid; i < size; i += OCT_THREADS) { aA[i] = cn[i] > 1; aB[i] = -1; }
            __syncthreads();
            const int E = block_scan_excl(aA, aD, size, s_warp);
            for (int i = tid; i < size; i += OCT_THREADS)
                if (cn[i] > 1) { aC[aD[i]] = cn[i]; NP4[aD[i]] = i; }
            __syncthreads();
            for (int e = tid; e < E; e += OCT_THREADS) {
                const int my = aC[e];
                int r = 0, j = 0;
                const int e4 = e & ~3;
                for (; j < e4; j += 4) {                                     // in front of e: ties rank first
                    const int4 k4 = *reinterpret_cast<const int4*>(aC + j);
                    r += (k4.x >= my) + (k4.y >= my) + (k4.z >= my) + (k4.w >= my);
                }
                for (; j < e; j++) r += aC[j] >= my;
                for (j = e + 1; j < E && (j & 3); j++) r += aC[j] > my;      // behind e: only larger counts
                for (; j + 4 <= E; j += 4) {
                    const int4 k4 = *reinterpret_cast<const int4*>(aC + j);
                    r += (k4.x > my) + (k4.y > my) + (k4.z > my) + (k4.w > my);
                }
                for (; j < E; j++) r += aC[j] > my;
                const int i = NP4[e];
                const int4 c4 = *reinterpret_cast<const int4*>(cc + 4 * i);
                aD[r] = i;                                                   // node processed r-th
                aE[r] = (c4.x > 0) + (c4.y > 0) + (c4.z > 0) + (c4.w > 0);  // its non-empty children
            }
            if (tid == 0) s_val[0] = E;                                      // default: split everything
            __syncthreads();
            block_scan_excl(aE, aA, E, s_warp);                              // aA[r] = sum of ne before rank r
            for (int r = tid; r < E; r += OCT_THREADS) {
                // list size after splitting ranks 0..r : prevSize + sum(ne) - (r+1)
                const int after = prevSize + aA[r] + aE[r] - (r + 1);
                const int before = prevSize + aA[r] - r;
                if (after >= N && before < N) s_val[0] = r + 1;              // break right after this split, :983
            }
            __syncthreads();
            nsplit = s_val[0];
            totalNe = (nsplit > 0) ? aA[nsplit - 1] + aE[nsplit - 1] : 0;
            __syncthreads();
            for (int r = tid; r < nsplit; r += OCT_THREADS) aB[aD[r]] = totalNe - aA[r] - aE[r];
            __syncthreads();
            for (int i = tid; i < size; i += OCT_THREADS) aC[i] = aB[i] < 0;
            __syncthreads();
            block_scan_excl(aC, aC, size, s_warp);
        }
        const int newSize = totalNe + size - nsplit;
        // ---- build the next node table and the relabel map ----
        if (tid == 0) s_val[1] = 0;
        __syncthreads();
        {
            const int* bxo = BX[cur]; const int* byo = BY[cur]; const int* pdo = PD[cur];
            int* bxn = BX[nxt]; int* byn = BY[nxt]; int* cnn = CN[nxt]; int* ccn = CC[nxt]; int* pdn = PD[nxt];
            int myexp = 0;
            for (int i = tid; i < size; i += OCT_THREADS) {
                const int bx = bxo[i], by = byo[i];
                int p = aB[i];
                if (p >= 0) {
                    const int x0 = bx & 0xFFFF, x1 = bx >> 16, y0 = by & 0xFFFF, y1 = by >> 16;
                    const int mx = x0 + ((x1 - x0 + 1) >> 1), my = y0 + ((y1 - y0 + 1) >> 1);
                    const int4 c4 = *reinterpret_cast<const int4*>(cc + 4 * i);
                    int4 np = make_int4(0, 0, 0, 0);
                    const int pdc = pdo[i] + (1 << 24) + 3 * (pdo[i] & 0xFFFFFF);      // depth + 1, path * 4 (+ child below)
                    if (c4.w > 0) pdn[p] = pdc + 3;
                    if (c4.z > 0) pdn[p + (c4.w > 0)] = pdc + 2;
                    if (c4.y > 0) pdn[p + (c4.w > 0) + (c4.z > 0)] = pdc + 1;
                    if (c4.x > 0) pdn[p + (c4.w > 0) + (c4.z > 0) + (c4.y > 0)] = pdc;
                    if (c4.w > 0) { bxn[p] = mx | (x1 << 16); byn[p] = my | (y1 << 16); cnn[p] = c4.w; np.w = p;
                                    *reinterpret_cast<int4*>(ccn + 4 * p) = make_int4(0, 0, 0, 0); p++; myexp += c4.w > 1; }
                    if (c4.z > 0) { bxn[p] = x0 | (mx << 16); byn[p] = my | (y1 << 16); cnn[p] = c4.z; np.z = p;
                                    *reinterpret_cast<int4*>(ccn + 4 * p) = make_int4(0, 0, 0, 0); p++; myexp += c4.z > 1; }
                    if (c4.y > 0) { bxn[p] = mx | (x1 << 16); byn[p] = y0 | (my << 16); cnn[p] = c4.y; np.y = p;
                                    *reinterpret_cast<int4*>(ccn + 4 * p) = make_int4(0, 0, 0, 0); p++; myexp += c4.y > 1; }
                    if (c4.x > 0) { bxn[p] = x0 | (mx << 16); byn[p] = y0 | (my << 16); cnn[p] = c4.x; np.x = p;
                                    *reinterpret_cast<int4*>(ccn + 4 * p) = make_int4(0, 0, 0, 0); p++; myexp += c4.x > 1; }
                    *reinterpret_cast<int4*>(NP4 + 4 * i) = np;
                } else {
                    p = totalNe + aC[i];
                    bxn[p] = bx; byn[p] = by; cnn[p] = cn[i]; pdn[p] = pdo[i];
                    *reinterpret_cast<int4*>(ccn + 4 * p) = make_int4(0, 0, 0, 0);
                    *reinterpret_cast<int4*>(NP4 + 4 * i) = make_int4(p, p, p, p);
                }
            }
            if (myexp) atomicAdd(&s_val[1], myexp);
        }
        __syncthreads();
        const int nToExpand = s_val[1];
        cur = nxt;
        size = newSize;
        if (size >= N || size == prevSize) finish = true;                  // :889 / :990
        else if (!final_phase && size + nToExpand * 3 > N) final_phase = true;   // :909
        __syncthreads();
    }

    // ---- best response per node; ties go to the first keypoint in the reference's vector order
    // (:1005-1030), i.e. the smallest canonical rank (cell row, cell col, y, x).  The rank determines the
    // keypoint (cand_rank is invertible), so the winner is rebuilt from (score, rank) by a node pass: ONE
    // candidate pass when score and rank pack into 32 bits (<= 1024 cells: score << 24 | inverted rank),
    // two otherwise ----
    unsigned* best = reinterpret_cast<unsigned*>(aA);
    const uint32_t mw = 0xFFFFFFFFu / (uint32_t)L.wCell + 1u, mh = 0xFFFFFFFFu / (uint32_t)L.hCell + 1u;   // ceil(2^32 / cell size); cells are 35..127 px
    int* bestrank = aB;
    orbx_key_t* sel = D.sel + (long long)f * G.sel_per_frame + L.sel_base;
    const int nout = min(size, L.sel_cap);
    const bool packed = L.nCols * L.nRows <= 1024;
    for (int i = tid; i < size; i += OCT_THREADS) {
        best[i] = 0u; bestrank[i] = 0x7FFFFFFF;
        if (fast) Hh[HOFF(PD[cur][i] >> 24) + (PD[cur][i] & 0xFFFFFF)] = -1 - i;      // list-position markers
    }
    __syncthreads();
    auto node_of = [&](int code) -> int {
        if (!fast) return code >> 2;
        int p = 0;
#pragma unroll
        for (int d = OCT_HDEPTH; d >= 0; d--) {       // deepest first: the final nodes sit near the bottom of the binned tree, and
            const int v = Hh[HOFF(d) + (code >> (2 * (OCT_HDEPTH - d)))];      // exactly one node on a candidate's path holds a marker
            if (v < 0) { p = -1 - v; break; }
        }
        return p;
    };
    if (packed) {
        for_cands(cand, kcode, n, [&](int, orbx_key_t k, int code) {
            atomicMax(&best[node_of(code)], (KEY_S(k) << 24) | (0xFFFFFFu - (uint32_t)cand_rank(k, L, mw, mh)));
        });
    } else {
        for_cands(cand, kcode, n, [&](int, orbx_key_t k, int code) { atomicMax(&best[node_of(code)], KEY_S(k)); });
        __syncthreads();
        for_cands(cand, kcode, n, [&](int, orbx_key_t k, int code) {
            const int nd = node_of(code);
            if (KEY_S(k) == best[nd]) atomicMin(&bestrank[nd], cand_rank(k, L, mw, mh));
        });
    }
    __syncthreads();
    for (int nd = tid; nd < nout; nd += OCT_THREADS) {
        const uint32_t score = packed ? best[nd] >> 24 : best[nd];
        const int rank = packed ? (int)(0xFFFFFFu - (best[nd] & 0xFFFFFFu)) : bestrank[nd];
        const int cell = rank >> 14, ci = cell / L.nCols, cj = cell - ci * L.nCols;
        const int X3 = cj * L.wCell + (rank & 127), Y3 = ci * L.hCell + ((rank >> 7) & 127);
        sel[nd] = MAKE_KEY(X3 + 3, Y3 + 3, score);
    }
    if (tid == 0) D.selcnt[f * G.nlevels + lvl] = nout;
#undef OCT_THREADS
#undef block_scan_excl
#undef for_cands
#undef HOFF
}

size_t octree_nodetab_ints(const FrameGeom& G)
{
    int mx = 0;
    for (int l = 0; l < G.nlevels; l++) mx = G.L[l].node_cap > mx ? G.L[l].node_cap : mx;
    return (size_t)mx * OCT_TABLES;
}

// Levels whose node tables fit 64 KB of shared memory share one launch of 256-thread CTAs (3 per SM).  Larger
// levels (8000-feature frames: level 0 needs 160 KB) share a launch of 1024-thread CTAs with their tables in shared
// memory up to 220 KB (one CTA per SM), and levels beyond that a launch with the tables in global memory: their
// candidate passes (250 k candidates per frame at 4K) are what the step waits for.
#define OCT_NT_FEW 512
template <bool SMEM, int NT>
static void launch_group(const FrameGeom& G, const DevBuffers& D, const OctLevels& LV, int nframes, size_t smem, cudaStream_t s)
{
    if (smem > 48 * 1024) cudaFuncSetAttribute(octree_kernel<SMEM, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    orbx_launch(octree_kernel<SMEM, NT>, dim3(nframes, LV.n), dim3(NT), smem, s, G, D, LV);
}

int launch_octree(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s, int level)
{
    OctLevels small, big, huge;
    small.n = big.n = huge.n = 0;
    size_t small_need = 0, big_need = 0;
    for (int l = 0; l < G.nlevels; l++) {
        if (level >= 0 && l != level) continue;
        const size_t need = ((size_t)G.L[l].node_cap * OCT_TABLES + (G.L[l].nIni <= OCT_HROOTS ? G.L[l].nIni * OCT_HNODES : 0)) * sizeof(int);
        if (need <= 64 * 1024) { small.lvl[small.n++] = l; small_need = need > small_need ? need : small_need; }
        else if (need <= 220 * 1024) { big.lvl[big.n++] = l; big_need = need > big_need ? need : big_need; }
        else huge.lvl[huge.n++] = l;
    }
    int launches = 0;
    // the long-running launches first: CTAs of the later launches fill the SMs they leave idle
    if (huge.n) { launch_group<false, 1024>(G, D, huge, nframes, 0, s); launches++; }
    if (big.n) { launch_group<true, 1024>(G, D, big, nframes, big_need, s); launches++; }
    if (small.n) {
        // few CTAs in flight (single frames): the candidate passes of level 0 are the critical path, wider CTAs shorten them
        static const int nt_env = [] { const char* e = getenv("ORBX_OCT_NT"); return e ? atoi(e) : 0; }();
        const int nt = nt_env ? nt_env : (small.n * nframes <= 148 ? OCT_NT_FEW : 256);
        if (nt == 1024) launch_group<true, 1024>(G, D, small, nframes, small_need, s);
        else if (nt == 512) launch_group<true, 512>(G, D, small, nframes, small_need, s);
        else launch_group<true, 256>(G, D, small, nframes, small_need, s);
        launches++;
    }
    return launches;
}
