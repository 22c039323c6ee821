// octree.cu — deterministic GPU rewrite of ORBextractor::DistributeOctTree
// (reference src/ORBextractor.cc:688-1034, ExtractorNode::DivideNode :602-674).
//
// The reference walks a std::list of nodes holding std::vector<KeyPoint> copies.  Here one CTA owns
// one (frame, level): every candidate carries the list POSITION of its node, and a round of the
// reference's loop becomes  (1) a key pass that histograms candidates into the 4 children of every
// node about to be split,  (2) scans over the node table that compute where each child / surviving
// node lands in the new list (children are push_front'ed in the order n1..n4, so the last split node's
// n4 ends up first),  (3) a key pass that relabels candidates.  The ordered final phase (:909-992:
// sort by size, split largest first, stop the moment the list reaches N) is a rank-by-counting sort
// plus a prefix cut.  Tie-break among equal-size nodes is the canonical (size, creation order) rule
// of the oracle (the reference compares heap pointers, :927): among the just-created nodes, creation
// order descending == list position ascending.  Winner per node = max response, first in candidate
// order (:1005-1030), via a packed atomicMax.
// Latency-bound integer work (no roofline); reported as a share of the step.
#include "orbx_internal.h"

#define OCT_THREADS 512
#define OCT_TABLES 19     // int arrays of node_cap entries in the per-CTA node scratch

struct Scan {
    int* warp;            // shared: OCT_THREADS/32 + 1 ints
};

// exclusive scan of in[0..m) into out[0..m) (out may alias in); returns the total to all threads.
__device__ int block_scan_excl(const int* in, int* out, int m, int* s_warp)
{
    const int T = OCT_THREADS;
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
    for (int w = 0; w < OCT_THREADS / 32; w++) {
        const int v = s_warp[w];
        if (w < (int)(threadIdx.x >> 5)) base += v;
        total += v;
    }
    int run = base + incl - sum;
    for (int i = b; i < e; i++) { const int v = in[i]; out[i] = run; run += v; }
    __syncthreads();
    return total;
}

__device__ __forceinline__ int child_of(int x, int y, int x0, int y0, int x1, int y1)
{
    const int mx = x0 + ((x1 - x0 + 1) >> 1);     // ceil(float(w)/2), :608
    const int my = y0 + ((y1 - y0 + 1) >> 1);     // :609
    return (x < mx) ? ((y < my) ? 0 : 2) : ((y < my) ? 1 : 3);   // :651-661
}

// position of a candidate in the reference's vToDistributeKeys order (cell row-major, raster inside a cell)
__device__ __forceinline__ int cand_rank(uint32_t key, const LevelGeom& L)
{
    const int X3 = (int)KEY_X(key) - 3, Y3 = (int)KEY_Y(key) - 3;
    const int cj = X3 / L.wCell, ci = Y3 / L.hCell;
    return ((ci * L.nCols + cj) << 14) | ((Y3 - ci * L.hCell) << 7) | (X3 - cj * L.wCell);
}

__global__ void __launch_bounds__(OCT_THREADS)
octree_kernel(FrameGeom G, DevBuffers D, int use_smem)
{
    extern __shared__ int dyn_smem[];
    __shared__ int s_warp[OCT_THREADS / 32 + 1];
    __shared__ int s_val[4];

    const int lvl = blockIdx.x, f = blockIdx.y;
    const LevelGeom& L = G.L[lvl];
    const int tid = threadIdx.x;
    const int cap = L.node_cap;

    uint32_t* cand = D.cand + (long long)f * G.cand_per_frame + L.cand_base;
    uint32_t* knode = D.candnode + (long long)f * G.cand_per_frame + L.cand_base;
    int* T = use_smem ? dyn_smem : (D.nodetab + ((long long)f * G.nlevels + lvl) * D.nodetab_per_frame);
    // node tables (double-buffered bounds+count), per-round temporaries
    int* bx0[2] = {T + 0 * cap, T + 5 * cap};
    int* by0[2] = {T + 1 * cap, T + 6 * cap};
    int* bx1[2] = {T + 2 * cap, T + 7 * cap};
    int* by1[2] = {T + 3 * cap, T + 8 * cap};
    int* ncnt[2] = {T + 4 * cap, T + 9 * cap};
    int* childcnt = T + 10 * cap;     // 4*cap
    int* tmpA = T + 14 * cap;         // flags / ne / gains
    int* tmpB = T + 15 * cap;         // scan results
    int* rnk = T + 16 * cap;          // processing rank of a node (or -1)
    int* headoff = T + 17 * cap;      // new list position of a split node's first child slot
    int* tailpos = T + 18 * cap;      // new list position of an unsplit node

    // candidates were compacted (in arbitrary order) by fast_band_kernel
    const int n = min(D.candcnt[f * G.nlevels + lvl], L.cand_cap);
    const int N = L.N;
    const int H = L.maxBY - ORBX_MINB;

    // ---- roots, :695-763 ----
    int cur = 0;
    for (int i = tid; i < L.nIni; i += OCT_THREADS) {
        bx0[0][i] = (int)__fmul_rn(L.hX, (float)i);
        bx1[0][i] = (int)__fmul_rn(L.hX, (float)(i + 1));
        by0[0][i] = 0; by1[0][i] = H;
        ncnt[0][i] = 0;
    }
    __syncthreads();
    for (int i = tid; i < n; i += OCT_THREADS) {
        int r = (int)__fdiv_rn((float)KEY_X(cand[i]), L.hX);     // vpIniNodes[kp.pt.x/hX], :740
        r = min(r, L.nIni - 1);
        knode[i] = r;
        atomicAdd(&ncnt[0][r], 1);
    }
    __syncthreads();
    for (int i = tid; i < L.nIni; i += OCT_THREADS) tmpA[i] = ncnt[0][i] > 0;
    __syncthreads();
    int size = block_scan_excl(tmpA, tmpB, L.nIni, s_warp);
    for (int i = tid; i < L.nIni; i += OCT_THREADS) {
        if (ncnt[0][i] > 0) {
            const int p = tmpB[i];
            bx0[1][p] = bx0[0][i]; by0[1][p] = by0[0][i]; bx1[1][p] = bx1[0][i]; by1[1][p] = by1[0][i];
            ncnt[1][p] = ncnt[0][i];
        }
    }
    for (int i = tid; i < n; i += OCT_THREADS) knode[i] = tmpB[knode[i]];
    cur = 1;
    __syncthreads();

    bool final_phase = false;
    bool finish = (n == 0);        // empty list: first iteration leaves size == prevSize, :889
#pragma unroll 1
    while (!finish) {
        const int prevSize = size;
        int *X0 = bx0[cur], *Y0 = by0[cur], *X1 = bx1[cur], *Y1 = by1[cur], *CN = ncnt[cur];
        int *NX0 = bx0[cur ^ 1], *NY0 = by0[cur ^ 1], *NX1 = bx1[cur ^ 1], *NY1 = by1[cur ^ 1], *NCN = ncnt[cur ^ 1];
        // (1) histogram candidates of expandable nodes into their four children
        for (int i = tid; i < 4 * size; i += OCT_THREADS) childcnt[i] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += OCT_THREADS) {
            const int nd = knode[i];
            if (CN[nd] > 1) {
                const uint32_t k = cand[i];
                const int c = child_of(KEY_X(k), KEY_Y(k), X0[nd], Y0[nd], X1[nd], Y1[nd]);
                atomicAdd(&childcnt[4 * nd + c], 1);
            }
        }
        __syncthreads();
        // (2) processing order of the nodes to split
        int nsplit, totalNe;
        if (!final_phase) {
            // all expandable nodes, in list order (:800-886)
            for (int i = tid; i < size; i += OCT_THREADS) tmpA[i] = CN[i] > 1;
            __syncthreads();
            nsplit = block_scan_excl(tmpA, rnk, size, s_warp);       // rnk = #expandable nodes before i
            for (int i = tid; i < size; i += OCT_THREADS) {
                int ne = 0;
                tailpos[i] = i - rnk[i];                                // #unsplit nodes before i
                if (CN[i] > 1) {
                    ne = (childcnt[4 * i] > 0) + (childcnt[4 * i + 1] > 0) + (childcnt[4 * i + 2] > 0) + (childcnt[4 * i + 3] > 0);
                } else rnk[i] = -1;
                tmpA[i] = ne;
            }
            __syncthreads();
            totalNe = block_scan_excl(tmpA, tmpB, size, s_warp);       // tmpB = exclusive prefix of ne in list order
            for (int i = tid; i < size; i += OCT_THREADS) {
                // later-processed nodes sit in front: offset = sum of ne over the nodes after i
                headoff[i] = totalNe - tmpB[i] - tmpA[i];
                tailpos[i] += totalNe;
            }
            __syncthreads();
        } else {
            // ordered phase (:909-992): sort expandable nodes by (size desc, creation desc == position asc)
            for (int i = tid; i < size; i += OCT_THREADS) {
                int r = -1;
                const int ci = CN[i];
                if (ci > 1) {
                    r = 0;
                    for (int j = 0; j < size; j++) {
                        const int cj = CN[j];
                        r += (cj > 1) && (cj > ci || (cj == ci && j < i));
                    }
                }
                rnk[i] = r;
            }
            __syncthreads();
            int* order = tmpB;     // order[r] = node
            int E;
            for (int i = tid; i < size; i += OCT_THREADS) tmpA[i] = CN[i] > 1;
            __syncthreads();
            E = block_scan_excl(tmpA, headoff, size, s_warp);   // headoff used as throwaway here
            for (int i = tid; i < size; i += OCT_THREADS) if (rnk[i] >= 0) order[rnk[i]] = i;
            __syncthreads();
            // ne in rank order -> tmpA ; inclusive prefix -> tailpos (temporary)
            for (int r = tid; r < E; r += OCT_THREADS) {
                const int i = order[r];
                tmpA[r] = (childcnt[4 * i] > 0) + (childcnt[4 * i + 1] > 0) + (childcnt[4 * i + 2] > 0) + (childcnt[4 * i + 3] > 0);
            }
            __syncthreads();
            block_scan_excl(tmpA, tailpos, E, s_warp);          // tailpos[r] = exclusive prefix of ne (rank order)
            if (tid == 0) s_val[0] = E;                          // default: split everything
            __syncthreads();
            for (int r = tid; r < E; r += OCT_THREADS) {
                // list size after splitting ranks 0..r : prevSize + sum(ne) - (r+1)
                const int after = prevSize + tailpos[r] + tmpA[r] - (r + 1);
                const int before = prevSize + tailpos[r] - r;
                if (after >= N && before < N) s_val[0] = r + 1;  // break right after this split, :983
            }
            __syncthreads();
            nsplit = s_val[0];
            totalNe = (nsplit > 0) ? tailpos[nsplit - 1] + tmpA[nsplit - 1] : 0;
            __syncthreads();
            for (int r = tid; r < E; r += OCT_THREADS) {
                const int i = order[r];
                if (r < nsplit) headoff[i] = totalNe - tailpos[r] - tmpA[r];
                else rnk[i] = -1;                                // not split in this round
            }
            __syncthreads();
            for (int i = tid; i < size; i += OCT_THREADS) tmpA[i] = rnk[i] >= 0;
            __syncthreads();
            block_scan_excl(tmpA, tmpB, size, s_warp);
            for (int i = tid; i < size; i += OCT_THREADS) tailpos[i] = totalNe + i - tmpB[i];
            __syncthreads();
        }
        const int newSize = totalNe + size - nsplit;
        // (3) build the next node table
        if (tid == 0) s_val[1] = 0;
        __syncthreads();
        int myexp = 0;
        for (int i = tid; i < size; i += OCT_THREADS) {
            if (rnk[i] >= 0) {
                const int x0 = X0[i], y0 = Y0[i], x1 = X1[i], y1 = Y1[i];
                const int mx = x0 + ((x1 - x0 + 1) >> 1), my = y0 + ((y1 - y0 + 1) >> 1);
                const int c0 = childcnt[4 * i], c1 = childcnt[4 * i + 1], c2 = childcnt[4 * i + 2], c3 = childcnt[4 * i + 3];
                int p = headoff[i];
                if (c3 > 0) { NX0[p] = mx; NY0[p] = my; NX1[p] = x1; NY1[p] = y1; NCN[p] = c3; p++; myexp += c3 > 1; }
                if (c2 > 0) { NX0[p] = x0; NY0[p] = my; NX1[p] = mx; NY1[p] = y1; NCN[p] = c2; p++; myexp += c2 > 1; }
                if (c1 > 0) { NX0[p] = mx; NY0[p] = y0; NX1[p] = x1; NY1[p] = my; NCN[p] = c1; p++; myexp += c1 > 1; }
                if (c0 > 0) { NX0[p] = x0; NY0[p] = y0; NX1[p] = mx; NY1[p] = my; NCN[p] = c0; p++; myexp += c0 > 1; }
            } else {
                const int p = tailpos[i];
                NX0[p] = X0[i]; NY0[p] = Y0[i]; NX1[p] = X1[i]; NY1[p] = Y1[i]; NCN[p] = CN[i];
            }
        }
        if (myexp) atomicAdd(&s_val[1], myexp);
        // (4) relabel candidates
        for (int i = tid; i < n; i += OCT_THREADS) {
            const int nd = knode[i];
            if (rnk[nd] >= 0) {
                const uint32_t k = cand[i];
                const int c = child_of(KEY_X(k), KEY_Y(k), X0[nd], Y0[nd], X1[nd], Y1[nd]);
                int p = headoff[nd];
                for (int cc = 3; cc > c; cc--) p += childcnt[4 * nd + cc] > 0;
                knode[i] = p;
            } else {
                knode[i] = tailpos[nd];
            }
        }
        __syncthreads();
        const int nToExpand = s_val[1];
        cur ^= 1;
        size = newSize;
        if (size >= N || size == prevSize) finish = true;                  // :889 / :990
        else if (!final_phase && size + nToExpand * 3 > N) final_phase = true;   // :909
        __syncthreads();
    }

    // ---- best response per node; ties go to the first keypoint in the reference's vector order
    // (:1005-1030), i.e. the smallest canonical rank (cell row, cell col, y, x) ----
    int* best = tmpA;
    int* bestrank = tmpB;
    for (int i = tid; i < size; i += OCT_THREADS) { best[i] = 0; bestrank[i] = 0x7FFFFFFF; }
    __syncthreads();
    for (int i = tid; i < n; i += OCT_THREADS) atomicMax(&best[knode[i]], (int)KEY_S(cand[i]));
    __syncthreads();
    for (int i = tid; i < n; i += OCT_THREADS) {
        const uint32_t k = cand[i];
        const int nd = knode[i];
        if ((int)KEY_S(k) == best[nd]) atomicMin(&bestrank[nd], cand_rank(k, L));
    }
    __syncthreads();
    uint32_t* sel = D.sel + (long long)f * G.sel_per_frame + L.sel_base;
    const int nout = min(size, L.sel_cap);
    for (int i = tid; i < n; i += OCT_THREADS) {
        const uint32_t k = cand[i];
        const int nd = knode[i];
        if (nd < nout && (int)KEY_S(k) == best[nd] && cand_rank(k, L) == bestrank[nd]) sel[nd] = k;
    }
    if (tid == 0) D.selcnt[f * G.nlevels + lvl] = nout;
}

size_t octree_nodetab_ints(const FrameGeom& G)
{
    int mx = 0;
    for (int l = 0; l < G.nlevels; l++) mx = G.L[l].node_cap > mx ? G.L[l].node_cap : mx;
    return (size_t)mx * OCT_TABLES;
}

int launch_octree(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    const size_t need = octree_nodetab_ints(G) * sizeof(int);
    const int use_smem = need <= 96 * 1024;
    if (use_smem && need > 48 * 1024)
        cudaFuncSetAttribute(octree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    const dim3 grid(G.nlevels, nframes);
    octree_kernel<<<grid, OCT_THREADS, use_smem ? need : 0, s>>>(G, D, use_smem);
    return 1;
}
