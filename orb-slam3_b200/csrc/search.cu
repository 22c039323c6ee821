// search.cu — the guided searches of src/ORBmatcher.cc that share one inner loop:
//   SearchByProjection(Frame&, const Frame&, th, bMono)   :2469-2711
//   SearchForInitialization                               :799-943
//   SearchForTriangulation / SearchForTriangulation_      :1107-1630
// Device side: candidate generation (GetFeaturesInArea windows over the CSR grid, or the per-BoW-node
// all-pairs sweep), every gate that does not depend on earlier acceptances, and the 256-bit Hamming
// distances (XOR + POPC, integer pipe).  Host side (this file, below the kernels): the reference's
// sequential acceptance bookkeeping and the 30-bin rotation histogram, replayed over the candidate lists
// in the reference's visiting order — that part is a few thousand dependent scalar steps per call.
// Each calling thread owns a grow-only workspace per device (device arena + pinned staging + stream), so
// a call is one H2D copy, two or three launches and one or two D2H copies, without cudaMalloc.
#include "match_internal.h"
#include <climits>
#include <cmath>
#include <cstring>
#include <vector>

#define CUDA_TRYS(expr)                                                                        \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            orbx_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return ORBX_ERR_CUDA;                                                              \
        }                                                                                      \
    } while (0)

// ---- per-thread workspace (declared in match_internal.h, shared with match.cu) ----
OrbxWorkspace& orbx_workspace() { static thread_local OrbxWorkspace w; return w; }
namespace {
typedef OrbxWorkspace Workspace;
typedef OrbxCarver Carver;
}
#define t_ws orbx_workspace()

// ---- windows: GetFeaturesInArea (src/Frame.cc:784-871) + DescriptorDistance for nq queries ----
struct WinArgs {
    const orbx_keypoint* kps; const uint4* desc; int n; const float* uRight;
    float minX, minY, winv, hinv;
    const int* cell_start; const int* cell_idx;
    int nq; const uint4* q_desc; const float *q_x, *q_y, *q_r, *q_ur; const int *q_min, *q_max;
    int* q_base; int* q_cnt;    // per query: where its list starts in cand, and its length
    int2* cand; int cap;        // (idx, dist)
    int* total;                 // bump cursor; > cap after the launch means the lists did not fit
};

#define WIN_WARPS 8
// One warp per query, lanes sweep the cells' CSR ranges 32 keypoints at a time.  Sweep 1 counts the
// hits, lane 0 reserves the list with one atomicAdd, sweep 2 writes (index, distance) in the
// reference's visiting order (cell column-major, insertion order inside a cell).
__global__ void __launch_bounds__(WIN_WARPS * 32)
win_candidates_kernel(WinArgs A)
{
    const int q = blockIdx.x * WIN_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= A.nq) return;
    const float r = A.q_r[q];
    int cx0 = 1, cx1 = 0, cy0 = 1, cy1 = 0;
    const float x = A.q_x[q], y = A.q_y[q];
    const int minLevel = A.q_min[q], maxLevel = A.q_max[q];
    if (r >= 0.f) {
        cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), r), A.winv)));                  // :800
        cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), r), A.winv)));       // :807
        cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), r), A.hinv)));                  // :814
        cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), r), A.hinv)));       // :820
        if (cx0 >= GRID_COLS || cx1 < 0 || cy0 >= GRID_ROWS || cy1 < 0) { cx0 = 1; cx1 = 0; }
    }
    const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);                                            // :828
    const float ur = A.q_ur ? A.q_ur[q] : 0.f;
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    int base = 0, cnt = 0;
    for (int pass = 0; pass < 2; pass++) {
        int pos = 0;
        for (int ix = cx0; ix <= cx1; ix++)
            for (int iy = cy0; iy <= cy1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                const int b = A.cell_start[c], e = A.cell_start[c + 1];
                for (int j0 = b; j0 < e; j0 += 32) {
                    const int j = j0 + lane;
                    bool ok = j < e;
                    int idx = 0;
                    if (ok) {
                        idx = A.cell_idx[j];
                        const orbx_keypoint kp = A.kps[idx];
                        if (checkLevels) {                                                                 // :848-857
                            if (kp.octave < minLevel) ok = false;
                            if (maxLevel >= 0 && kp.octave > maxLevel) ok = false;
                        }
                        const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                        if (!(fabsf(dx) < r && fabsf(dy) < r)) ok = false;                                 // :864
                        if (ok && A.uRight && A.uRight[idx] > 0) {                                         // src/ORBmatcher.cc:2556-2562
                            const float er = fabsf(__fsub_rn(ur, A.uRight[idx]));
                            if (er > r) ok = false;
                        }
                    }
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    if (pass && ok) {
                        const int p = base + pos + __popc(bal & ((1u << lane) - 1));
                        if (p < A.cap) A.cand[p] = make_int2(idx, ham256(a0, a1, A.desc[2 * idx], A.desc[2 * idx + 1]));
                    }
                    pos += __popc(bal);
                }
            }
        if (!pass) {
            cnt = pos;
            if (lane == 0) { base = cnt ? atomicAdd(A.total, cnt) : 0; A.q_base[q] = base; A.q_cnt[q] = cnt; }
            base = __shfl_sync(0xffffffffu, base, 0);
            if (cnt == 0) return;
            a0 = A.q_desc[2 * q]; a1 = A.q_desc[2 * q + 1];
        }
    }
}

// ---- per-BoW-node all-pairs sweep of SearchForTriangulation (src/ORBmatcher.cc:1176-1330) ----
struct TriArgs {
    const orbx_keypoint* kps1; const uint4* desc1; const float* uRight1;
    const orbx_keypoint* kps2; const uint4* desc2; const float* uRight2; const uint8_t* has_mp2;
    const int* fv2_idx;
    int nitems; const int4* items;   // (idx1, first, last) into fv2_idx; .w unused
    float sf2[ORBX_MAX_LEVELS], sigma2_2[ORBX_MAX_LEVELS];
    float epx, epy, F[9];
    int bOnlyStereo, bCoarse;
    int mode;                        // 0: SearchForTriangulation gates, 1: SearchByBoW (ratio-bounded distances),
                                     // 2: dist <= TH_LOW only (rig / callback triangulation: geometry is the caller's)
    float nnratio;
    int* it_base; int* it_cnt;
    int2* cand; int cap;             // (idx2 | epi_ok << 30, dist)
    int* total;
};

#define TRI_WARPS 8
// One warp per keypoint of KF1 that may still be matched (no map point yet, stereo gate passed on the
// host); lanes sweep the keypoints of KF2 filed under the same vocabulary node.  Kept: pairs with
// dist <= TH_LOW that pass the gates that do not depend on earlier matches (:1218-1250); the epipolar
// verdict (:1282, Pinhole::epipolarConstrain src/CameraModels/Pinhole.cpp:246-259) rides in bit 30.
__global__ void __launch_bounds__(TRI_WARPS * 32)
tri_candidates_kernel(TriArgs A)
{
    const int it = blockIdx.x * TRI_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (it >= A.nitems) return;
    const int4 item = A.items[it];
    const int idx1 = item.x;
    const orbx_keypoint kp1 = A.kps1[idx1];
    const bool bStereo1 = A.uRight1 && A.uRight1[idx1] >= 0;
    const uint4 a0 = A.desc1[2 * idx1], a1 = A.desc1[2 * idx1 + 1];
    // epipolar line of kp1 in image 2
    const float la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, A.F[0]), __fmul_rn(kp1.y, A.F[3])), A.F[6]);
    const float lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, A.F[1]), __fmul_rn(kp1.y, A.F[4])), A.F[7]);
    const float lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, A.F[2]), __fmul_rn(kp1.y, A.F[5])), A.F[8]);
    const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
    int base = 0;
    for (int pass = 0; pass < 2; pass++) {
        int pos = 0;
        for (int j0 = item.y; j0 < item.z; j0 += 32) {
            const int j = j0 + lane;
            bool ok = j < item.z;
            int idx2 = 0, dist = 0, epi = 0;
            if (ok) {
                idx2 = A.fv2_idx[j];
                if (A.has_mp2[idx2]) ok = false;                                             // :1218 (BoW: keypoint 2 not usable)
                const bool bStereo2 = A.uRight2 && A.uRight2[idx2] >= 0;
                if (A.bOnlyStereo && !bStereo2) ok = false;                                  // :1224-1226
                if (ok) {
                    dist = ham256(a0, a1, A.desc2[2 * idx2], A.desc2[2 * idx2 + 1]);
                    if (A.mode == 0 || A.mode == 2) { if (dist > TH_LOW) ok = false; }       // :1231 (mode 2: the verdict is the caller's)
                    // SearchByBoW accepts best1 <= TH_LOW when best1 < ratio * best2 (:431-433): a distance with
                    // ratio * d > TH_LOW can neither win nor make that test fail, so it need not be reported
                    else { epi = 1; if (__fmul_rn(A.nnratio, (float)dist) > (float)TH_LOW) ok = false; }
                }
                if (ok && A.mode == 0) {
                    const orbx_keypoint kp2 = A.kps2[idx2];
                    if (!bStereo1 && !bStereo2) {                                            // :1241-1250
                        const float ex = __fsub_rn(A.epx, kp2.x), ey = __fsub_rn(A.epy, kp2.y);
                        if (__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)) < __fmul_rn(100.f, A.sf2[kp2.octave])) ok = false;
                    }
                    const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, kp2.x), __fmul_rn(lb, kp2.y)), lc);
                    if (den != 0.f) {
                        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
                        epi = (double)dsqr < __dmul_rn(3.84, (double)A.sigma2_2[kp2.octave]);
                    }
                    if (A.bCoarse) epi = 1;
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (pass && ok) {
                const int p = base + pos + __popc(bal & ((1u << lane) - 1));
                if (p < A.cap) A.cand[p] = make_int2(idx2 | (epi << 30), dist);
            }
            pos += __popc(bal);
        }
        if (!pass) {
            if (lane == 0) { base = pos ? atomicAdd(A.total, pos) : 0; A.it_base[it] = base; A.it_cnt[it] = pos; }
            base = __shfl_sync(0xffffffffu, base, 0);
            if (pos == 0) return;
        }
    }
}

// ---- host: run the window kernel over caller (pageable) arrays through the workspace ----
namespace {
struct WinHost {
    const orbx_keypoint* kps; const uint8_t* desc; int n; const float* uRight;
    float minX, maxX, minY, maxY;
    int nq; const uint8_t* q_desc; const float *q_x, *q_y, *q_r, *q_ur; const int32_t *q_min, *q_max;
};
struct WinResult { const int* base; const int* cnt; const int2* cand; int total; };

thread_local int t_win_cap = 1 << 16, t_tri_cap = 1 << 16;

int run_windows(int device, const WinHost& in, WinResult* out)
{
    Workspace& W = t_ws;
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    const size_t n = (size_t)in.n, nq = (size_t)in.nq;
    int cap = t_win_cap;
    for (int attempt = 0; attempt < 2; attempt++) {
        Carver c;
        const size_t o_kps = c.take(n * sizeof(orbx_keypoint)), o_desc = c.take(n * 32), o_ur = c.take(in.uRight ? n * 4 : 0);
        const size_t o_qd = c.take(nq * 32), o_qx = c.take(nq * 4), o_qy = c.take(nq * 4), o_qr = c.take(nq * 4);
        const size_t o_qur = c.take(in.q_ur ? nq * 4 : 0), o_qmin = c.take(nq * 4), o_qmax = c.take(nq * 4);
        const size_t o_total = c.take(4);
        const size_t in_bytes = c.off;
        const size_t o_base = c.take(nq * 4), o_cnt = c.take(nq * 4);
        const size_t small_end = c.off;
        const size_t o_cs = c.take((GRID_CELLS + 1) * 4), o_ci = c.take(n * 4);
        const size_t o_cand = c.take((size_t)cap * 8);
        rc = W.reserve(c.off, small_end + (size_t)cap * 8 + 256);
        if (rc != ORBX_OK) return rc;
        char* h = W.h; char* d = W.d;
        memcpy(h + o_kps, in.kps, n * sizeof(orbx_keypoint)); memcpy(h + o_desc, in.desc, n * 32);
        if (in.uRight) memcpy(h + o_ur, in.uRight, n * 4);
        memcpy(h + o_qd, in.q_desc, nq * 32); memcpy(h + o_qx, in.q_x, nq * 4); memcpy(h + o_qy, in.q_y, nq * 4);
        memcpy(h + o_qr, in.q_r, nq * 4); memcpy(h + o_qmin, in.q_min, nq * 4); memcpy(h + o_qmax, in.q_max, nq * 4);
        if (in.q_ur) memcpy(h + o_qur, in.q_ur, nq * 4);
        *(int*)(h + o_total) = 0;
        CUDA_TRYS(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, W.s));
        GridArgs g;
        g.kps = (const orbx_keypoint*)(d + o_kps); g.n = in.n; g.minX = in.minX; g.minY = in.minY;
        g.winv = (float)GRID_COLS / (float)(in.maxX - in.minX);                   // src/Frame.cc:292
        g.hinv = (float)GRID_ROWS / (float)(in.maxY - in.minY);                   // :294
        g.cell_start = (int*)(d + o_cs); g.cell_idx = (int*)(d + o_ci);
        orbx_launch_grid_build(g, W.s);
        WinArgs A;
        A.kps = g.kps; A.desc = (const uint4*)(d + o_desc); A.n = in.n; A.uRight = in.uRight ? (const float*)(d + o_ur) : nullptr;
        A.minX = g.minX; A.minY = g.minY; A.winv = g.winv; A.hinv = g.hinv; A.cell_start = g.cell_start; A.cell_idx = g.cell_idx;
        A.nq = in.nq; A.q_desc = (const uint4*)(d + o_qd); A.q_x = (const float*)(d + o_qx); A.q_y = (const float*)(d + o_qy);
        A.q_r = (const float*)(d + o_qr); A.q_ur = in.q_ur ? (const float*)(d + o_qur) : nullptr;
        A.q_min = (const int*)(d + o_qmin); A.q_max = (const int*)(d + o_qmax);
        A.q_base = (int*)(d + o_base); A.q_cnt = (int*)(d + o_cnt); A.cand = (int2*)(d + o_cand); A.cap = cap; A.total = (int*)(d + o_total);
        win_candidates_kernel<<<(in.nq + WIN_WARPS - 1) / WIN_WARPS, WIN_WARPS * 32, 0, W.s>>>(A);
        CUDA_TRYS(cudaGetLastError());
        CUDA_TRYS(cudaMemcpyAsync(h + o_total, d + o_total, small_end - o_total, cudaMemcpyDeviceToHost, W.s));
        CUDA_TRYS(cudaStreamSynchronize(W.s));
        const int total = *(const int*)(h + o_total);
        if (total > cap) { cap = total + total / 4; t_win_cap = cap; continue; }
        if (total > 0) {
            CUDA_TRYS(cudaMemcpyAsync(h + small_end, d + o_cand, (size_t)total * 8, cudaMemcpyDeviceToHost, W.s));
            CUDA_TRYS(cudaStreamSynchronize(W.s));
        }
        out->base = (const int*)(h + o_base); out->cnt = (const int*)(h + o_cnt);
        out->cand = (const int2*)(h + small_end); out->total = total;
        return ORBX_OK;
    }
    orbx_set_error("window candidate lists did not fit after regrowing");
    return ORBX_ERR_CAPACITY;
}

bool bad_frame(const orbx_keypoint* kps, const uint8_t* desc, int n, float minX, float maxX, float minY, float maxY)
{
    return n < 0 || (n > 0 && (!kps || !desc)) || !(maxX > minX) || !(maxY > minY);
}

// the 30-bin rotation histogram every search keeps (e.g. src/ORBmatcher.cc:2477-2480, 2584-2600)
struct RotHist {
    std::vector<int> bin[HISTO_LENGTH];
    float factor;
    explicit RotHist(float f) : factor(f) { for (auto& b : bin) b.reserve(500); }
    void push(float angle1, float angle2, int v)
    {
        float rot = angle1 - angle2;
        if (rot < 0.0) rot += 360.0f;
        int b = (int)roundf(rot * factor);
        if (b == HISTO_LENGTH) b = 0;
        if (b >= 0 && b < HISTO_LENGTH) bin[b].push_back(v);
    }
    void top3(int* i1, int* i2, int* i3) const
    {
        int32_t sz[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sz[i] = (int32_t)bin[i].size();
        orbx_three_maxima(sz, HISTO_LENGTH, i1, i2, i3);
    }
};
}  // namespace

extern "C" void orbx_three_maxima(const int32_t* sizes, int L, int32_t* ind1, int32_t* ind2, int32_t* ind3)
{
    int max1 = 0, max2 = 0, max3 = 0, a = -1, b = -1, c = -1;
    for (int i = 0; i < L; i++) {
        const int s = sizes[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; c = b; b = a; a = i; }
        else if (s > max2) { max3 = max2; max2 = s; c = b; b = i; }
        else if (s > max3) { max3 = s; c = i; }
    }
    if (max2 < 0.1f * (float)max1) { b = -1; c = -1; }
    else if (max3 < 0.1f * (float)max1) c = -1;
    *ind1 = a; *ind2 = b; *ind3 = c;
}

extern "C" int orbx_window_candidates(int device, const orbx_keypoint* kps, const uint8_t* desc, int n,
                                      float minX, float maxX, float minY, float maxY,
                                      int nq, const uint8_t* q_desc, const float* q_x, const float* q_y, const float* q_r,
                                      const int32_t* q_minLevel, const int32_t* q_maxLevel,
                                      int32_t* out_start, int32_t* out_idx, int32_t* out_dist, int cap)
{
    if (bad_frame(kps, desc, n, minX, maxX, minY, maxY) || nq < 0 || cap < 0 || !out_start ||
        (nq > 0 && (!q_desc || !q_x || !q_y || !q_r || !q_minLevel || !q_maxLevel)) || (cap > 0 && (!out_idx || !out_dist))) {
        orbx_set_error("bad window_candidates arguments"); return ORBX_ERR_ARG;
    }
    out_start[0] = 0;
    if (nq == 0) return 0;
    WinHost in = {kps, desc, n, nullptr, minX, maxX, minY, maxY, nq, q_desc, q_x, q_y, q_r, nullptr, q_minLevel, q_maxLevel};
    WinResult R;
    const int rc = run_windows(device, in, &R);
    if (rc != ORBX_OK) return rc;
    int run = 0;
    for (int q = 0; q < nq; q++) {
        const int2* c = R.cand + R.base[q];
        for (int k = 0; k < R.cnt[q]; k++, run++)
            if (run < cap) { out_idx[run] = c[k].x; out_dist[run] = c[k].y; }
        out_start[q + 1] = run;
    }
    if (run > cap) { orbx_set_error("window_candidates: %d candidates, capacity %d", run, cap); return ORBX_ERR_CAPACITY; }
    return run;
}


// Windows + sequential acceptance shared by SearchByProjection(Frame,Frame) (src/ORBmatcher.cc:2545-2601, 2686-2708)
// and SearchByProjection(Frame,KeyFrame,...) (:2790-2857): best unheld candidate per row, accepted when its
// distance is <= maxDist; row_obs == NULL means every accepted keypoint becomes held.
static int projection_replay(int device, const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                             float minX, float maxX, float minY, float maxY,
                             int nrows, const uint8_t* row_desc, const float* row_u, const float* row_v,
                             const float* q_r, const float* q_ur, const int32_t* q_min, const int32_t* q_max,
                             const float* row_angle, const uint8_t* row_obs, int maxDist, int checkOrientation,
                             const uint8_t* cur_held, int32_t* assign)
{
    WinHost in = {kps, desc, n, uRight, minX, maxX, minY, maxY, nrows, row_desc, row_u, row_v, q_r, q_ur, q_min, q_max};
    WinResult R;
    const int rc = run_windows(device, in, &R);
    if (rc != ORBX_OK) return rc;
    int nmatches = 0;
    RotHist H(HISTO_LENGTH / 360.0f);                                             // :2480 / :2735
    std::vector<uint8_t> held(n, 0);
    if (cur_held) for (int i = 0; i < n; i++) held[i] = cur_held[i] != 0;
    for (int i = 0; i < nrows; i++) {
        const int nc = R.cnt[i];
        if (nc == 0) continue;
        const int2* c = R.cand + R.base[i];
        int bestDist = 256, bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {
            if (held[c[k].x]) continue;                                           // :2552-2554 / :2800
            if (c[k].y < bestDist) { bestDist = c[k].y; bestIdx2 = c[k].x; }
        }
        if (bestDist <= maxDist) {
            assign[bestIdx2] = i;
            held[bestIdx2] = row_obs ? row_obs[i] != 0 : 1;
            nmatches++;
            if (checkOrientation) H.push(row_angle[i], kps[bestIdx2].angle, bestIdx2);
        }
    }
    if (checkOrientation) {                                                       // :2686-2708 / :2834-2857
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++)
            if (b != i1 && b != i2 && b != i3)
                for (int v : H.bin[b]) { assign[v] = -1; nmatches--; }
    }
    return nmatches;
}

extern "C" int orbx_search_by_projection_frame(int device,
                                               const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                                               float minX, float maxX, float minY, float maxY,
                                               const float* scaleFactors, int nlevels,
                                               int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                               const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                               const uint8_t* last_valid, const uint8_t* last_obs,
                                               float th, float mbf, int bForward, int bBackward, int checkOrientation,
                                               const uint8_t* cur_held, int32_t* assign)
{
    if (bad_frame(kps, desc, n, minX, maxX, minY, maxY) || nlast < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors ||
        (n > 0 && !assign) ||
        (nlast > 0 && (!last_desc || !last_u || !last_v || !last_invz || !last_octave || !last_angle || !last_valid || !last_obs))) {
        orbx_set_error("bad search_by_projection_frame arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n; i++) assign[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n == 0 || nlast == 0) return 0;
    // per last-frame point: the window of :2529-2538 (or none)
    std::vector<float> q_r(nlast), q_ur(nlast);
    std::vector<int32_t> q_min(nlast), q_max(nlast);
    for (int i = 0; i < nlast; i++) {
        q_r[i] = -1.f; q_ur[i] = 0.f; q_min[i] = -1; q_max[i] = -1;
        if (!last_valid[i]) continue;                                             // :2503-2506
        const float invzc = last_invz[i];
        if (invzc < 0) continue;                                                  // :2515
        if (last_u[i] < minX || last_u[i] > maxX) continue;                       // :2520-2523
        if (last_v[i] < minY || last_v[i] > maxY) continue;
        const int oct = last_octave[i];
        if (oct < 0 || oct >= nlevels) { orbx_set_error("last-frame point %d: octave %d out of range", i, oct); return ORBX_ERR_ARG; }
        q_r[i] = th * scaleFactors[oct];                                          // :2529
        if (!(q_r[i] >= 0.f)) { q_r[i] = -1.f; continue; }
        q_ur[i] = last_u[i] - mbf * invzc;                                        // :2558
        if (bForward) { q_min[i] = oct; q_max[i] = -1; }                          // :2533
        else if (bBackward) { q_min[i] = 0; q_max[i] = oct; }                     // :2535
        else { q_min[i] = oct - 1; q_max[i] = oct + 1; }                          // :2537
    }
    return projection_replay(device, kps, desc, n, uRight, minX, maxX, minY, maxY, nlast, last_desc, last_u, last_v,
                             q_r.data(), uRight ? q_ur.data() : nullptr, q_min.data(), q_max.data(), last_angle, last_obs,
                             TH_HIGH, checkOrientation, cur_held, assign);
}

extern "C" int orbx_search_by_projection_keyframe(int device,
                                                  const orbx_keypoint* kps, const uint8_t* desc, int n,
                                                  float minX, float maxX, float minY, float maxY,
                                                  const float* scaleFactors, int nlevels,
                                                  int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                                  const int32_t* mp_level, const float* mp_angle, const uint8_t* mp_valid,
                                                  float th, int ORBdist, int checkOrientation,
                                                  const uint8_t* cur_held, int32_t* assign)
{
    if (bad_frame(kps, desc, n, minX, maxX, minY, maxY) || nmp < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors ||
        (n > 0 && !assign) || (nmp > 0 && (!mp_desc || !mp_u || !mp_v || !mp_level || !mp_angle || !mp_valid))) {
        orbx_set_error("bad search_by_projection_keyframe arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n; i++) assign[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n == 0 || nmp == 0) return 0;
    std::vector<float> q_r(nmp);
    std::vector<int32_t> q_min(nmp), q_max(nmp);
    for (int i = 0; i < nmp; i++) {
        q_r[i] = -1.f; q_min[i] = -1; q_max[i] = -1;
        if (!mp_valid[i]) continue;                                               // :2749-2753, 2772-2776 (caller's gates)
        if (mp_u[i] < minX || mp_u[i] > maxX) continue;                           // :2761-2764
        if (mp_v[i] < minY || mp_v[i] > maxY) continue;
        const int lvl = mp_level[i];
        if (lvl < 0 || lvl >= nlevels) { orbx_set_error("map point %d: level %d out of range", i, lvl); return ORBX_ERR_ARG; }
        q_r[i] = th * scaleFactors[lvl];                                          // :2781
        if (!(q_r[i] >= 0.f)) { q_r[i] = -1.f; continue; }
        q_min[i] = lvl - 1; q_max[i] = lvl + 1;                                   // :2784
    }
    return projection_replay(device, kps, desc, n, nullptr, minX, maxX, minY, maxY, nmp, mp_desc, mp_u, mp_v,
                             q_r.data(), nullptr, q_min.data(), q_max.data(), mp_angle, nullptr,
                             ORBdist, checkOrientation, cur_held, assign);
}

extern "C" int orbx_search_for_initialization(int device,
                                              const orbx_keypoint* kps1, const uint8_t* desc1, int n1,
                                              const orbx_keypoint* kps2, const uint8_t* desc2, int n2,
                                              float minX, float maxX, float minY, float maxY,
                                              float* prev_matched, int windowSize, float nnratio, int checkOrientation,
                                              int32_t* matches12)
{
    if (bad_frame(kps2, desc2, n2, minX, maxX, minY, maxY) || n1 < 0 || (n1 > 0 && (!kps1 || !desc1 || !prev_matched || !matches12)) || windowSize < 0) {
        orbx_set_error("bad search_for_initialization arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<float> q_x(n1), q_y(n1), q_r(n1);
    std::vector<int32_t> q_lvl(n1);
    for (int i = 0; i < n1; i++) {
        q_x[i] = prev_matched[2 * i]; q_y[i] = prev_matched[2 * i + 1];
        q_lvl[i] = kps1[i].octave;
        q_r[i] = kps1[i].octave > 0 ? -1.f : (float)windowSize;                   // :825, :831
    }
    WinHost in = {kps2, desc2, n2, nullptr, minX, maxX, minY, maxY, n1, desc1, q_x.data(), q_y.data(), q_r.data(), nullptr,
                  q_lvl.data(), q_lvl.data()};
    WinResult R;
    const int rc = run_windows(device, in, &R);
    if (rc != ORBX_OK) return rc;
    int nmatches = 0;
    RotHist H(HISTO_LENGTH / 360.0f);                                             // :812
    std::vector<int> vMatchedDistance(n2, INT_MAX), vnMatches21(n2, -1);
    for (int i1 = 0; i1 < n1; i1++) {
        const int nc = R.cnt[i1];
        if (nc == 0) continue;
        const int2* c = R.cand + R.base[i1];
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int k = 0; k < nc; k++) {                                            // :845-866
            const int i2 = c[k].x, dist = c[k].y;
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW && bestDist < (float)bestDist2 * nnratio) {        // :870-873
            if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
            matches12[i1] = bestIdx2;
            vnMatches21[bestIdx2] = i1;
            vMatchedDistance[bestIdx2] = bestDist;
            nmatches++;
            if (checkOrientation) H.push(kps1[i1].angle, kps2[bestIdx2].angle, i1);
        }
    }
    if (checkOrientation) {                                                       // :914-935
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (int idx1 : H.bin[b])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)                                               // :939-941
        if (matches12[i1] >= 0) { prev_matched[2 * i1] = kps2[matches12[i1]].x; prev_matched[2 * i1 + 1] = kps2[matches12[i1]].y; }
    return nmatches;
}


// ---- host: run the pair kernel over caller (pageable) arrays through the workspace ----
namespace {
struct PairHost {
    const orbx_keypoint* kps1; const uint8_t* desc1; int n1; const float* uRight1;
    const orbx_keypoint* kps2; const uint8_t* desc2; int n2; const float* uRight2; const uint8_t* skip2;
    const int32_t* fv2_idx; int len2; const int4* items; int nitems;
    const float* sf2; const float* sigma2_2; int nlevels; const float* ep; const float* F12;
    int bOnlyStereo, bCoarse, mode; float nnratio;
};
struct PairResult { const int* base; const int* cnt; const int2* cand; int total; };

int run_pairs(int device, const PairHost& in, PairResult* out)
{
    Workspace& W = t_ws;
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    const size_t m1 = (size_t)in.n1, m2 = (size_t)in.n2;
    const int nitems = in.nitems, len2 = in.len2;
    int cap = t_tri_cap;
    for (int attempt = 0; attempt < 2; attempt++) {
        Carver c;
        const size_t o_k1 = c.take(m1 * sizeof(orbx_keypoint)), o_d1 = c.take(m1 * 32), o_u1 = c.take(in.uRight1 ? m1 * 4 : 0);
        const size_t o_k2 = c.take(m2 * sizeof(orbx_keypoint)), o_d2 = c.take(m2 * 32), o_u2 = c.take(in.uRight2 ? m2 * 4 : 0);
        const size_t o_h2 = c.take(m2), o_f2 = c.take((size_t)len2 * 4), o_it = c.take((size_t)nitems * 16);
        const size_t o_total = c.take(4);
        const size_t in_bytes = c.off;
        const size_t o_base = c.take((size_t)nitems * 4), o_cnt = c.take((size_t)nitems * 4);
        const size_t small_end = c.off;
        const size_t o_cand = c.take((size_t)cap * 8);
        rc = W.reserve(c.off, small_end + (size_t)cap * 8 + 256);
        if (rc != ORBX_OK) return rc;
        char* h = W.h; char* d = W.d;
        memcpy(h + o_k1, in.kps1, m1 * sizeof(orbx_keypoint)); memcpy(h + o_d1, in.desc1, m1 * 32);
        memcpy(h + o_k2, in.kps2, m2 * sizeof(orbx_keypoint)); memcpy(h + o_d2, in.desc2, m2 * 32);
        if (in.uRight1) memcpy(h + o_u1, in.uRight1, m1 * 4);
        if (in.uRight2) memcpy(h + o_u2, in.uRight2, m2 * 4);
        memcpy(h + o_h2, in.skip2, m2); memcpy(h + o_f2, in.fv2_idx, (size_t)len2 * 4); memcpy(h + o_it, in.items, (size_t)nitems * 16);
        *(int*)(h + o_total) = 0;
        CUDA_TRYS(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, W.s));
        TriArgs A;
        A.kps1 = (const orbx_keypoint*)(d + o_k1); A.desc1 = (const uint4*)(d + o_d1); A.uRight1 = in.uRight1 ? (const float*)(d + o_u1) : nullptr;
        A.kps2 = (const orbx_keypoint*)(d + o_k2); A.desc2 = (const uint4*)(d + o_d2); A.uRight2 = in.uRight2 ? (const float*)(d + o_u2) : nullptr;
        A.has_mp2 = (const uint8_t*)(d + o_h2); A.fv2_idx = (const int*)(d + o_f2);
        A.nitems = nitems; A.items = (const int4*)(d + o_it);
        for (int l = 0; l < ORBX_MAX_LEVELS; l++) {
            A.sf2[l] = (in.sf2 && l < in.nlevels) ? in.sf2[l] : 0.f;
            A.sigma2_2[l] = (in.sigma2_2 && l < in.nlevels) ? in.sigma2_2[l] : 0.f;
        }
        A.epx = in.ep ? in.ep[0] : 0.f; A.epy = in.ep ? in.ep[1] : 0.f;
        for (int k = 0; k < 9; k++) A.F[k] = in.F12 ? in.F12[k] : 0.f;
        A.bOnlyStereo = in.bOnlyStereo; A.bCoarse = in.bCoarse; A.mode = in.mode; A.nnratio = in.nnratio;
        A.it_base = (int*)(d + o_base); A.it_cnt = (int*)(d + o_cnt); A.cand = (int2*)(d + o_cand); A.cap = cap; A.total = (int*)(d + o_total);
        tri_candidates_kernel<<<(nitems + TRI_WARPS - 1) / TRI_WARPS, TRI_WARPS * 32, 0, W.s>>>(A);
        CUDA_TRYS(cudaGetLastError());
        CUDA_TRYS(cudaMemcpyAsync(h + o_total, d + o_total, small_end - o_total, cudaMemcpyDeviceToHost, W.s));
        CUDA_TRYS(cudaStreamSynchronize(W.s));
        const int total = *(const int*)(h + o_total);
        if (total > cap) { cap = total + total / 4; t_tri_cap = cap; continue; }
        if (total > 0) {
            CUDA_TRYS(cudaMemcpyAsync(h + small_end, d + o_cand, (size_t)total * 8, cudaMemcpyDeviceToHost, W.s));
            CUDA_TRYS(cudaStreamSynchronize(W.s));
        }
        out->base = (const int*)(h + o_base); out->cnt = (const int*)(h + o_cnt); out->cand = (const int2*)(h + small_end); out->total = total;
        return ORBX_OK;
    }
    orbx_set_error("pair candidate lists did not fit after regrowing");
    return ORBX_ERR_CAPACITY;
}

// merge of two sorted FeatureVectors (e.g. src/ORBmatcher.cc:1176-1178, 1316-1329): one work item per keypoint
// of the first keyframe that passes `take`, in the reference's visiting order
template <class Take>
void merge_feature_vectors(const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                           const int32_t* fv2_node, const int32_t* fv2_start, int nn2, Take take, std::vector<int4>* items)
{
    for (int a = 0, b = 0; a < nn1 && b < nn2;) {
        if (fv1_node[a] < fv2_node[b]) { a++; continue; }
        if (fv1_node[a] > fv2_node[b]) { b++; continue; }
        for (int p1 = fv1_start[a]; p1 < fv1_start[a + 1]; p1++)
            if (take(fv1_idx[p1])) items->push_back(make_int4(fv1_idx[p1], fv2_start[b], fv2_start[b + 1], 0));
        a++; b++;
    }
}
}  // namespace

extern "C" int orbx_search_for_triangulation(int device,
                                             const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                             const float* uRight1,
                                             const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                             const orbx_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                             const float* uRight2,
                                             const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                             const float* sf2, const float* sigma2_2, int nlevels, const float* ep, const float* F12,
                                             int bOnlyStereo, int bCoarse, int checkOrientation, int variant,
                                             int32_t* matches12)
{
    if (n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !sf2 || !sigma2_2 || !ep || !F12 ||
        (n1 > 0 && (!kps1 || !desc1 || !has_mp1 || !matches12)) || (n2 > 0 && (!kps2 || !desc2 || !has_mp2)) ||
        (nn1 > 0 && (!fv1_node || !fv1_start || !fv1_idx)) || (nn2 > 0 && (!fv2_node || !fv2_start || !fv2_idx))) {
        orbx_set_error("bad search_for_triangulation arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n1 == 0 || n2 == 0 || nn1 == 0 || nn2 == 0) return 0;
    const int len1 = fv1_start[nn1], len2 = fv2_start[nn2];
    for (int i = 0; i < len1; i++) if (fv1_idx[i] < 0 || fv1_idx[i] >= n1) { orbx_set_error("fv1_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    for (int i = 0; i < len2; i++) if (fv2_idx[i] < 0 || fv2_idx[i] >= n2) { orbx_set_error("fv2_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    for (int i = 0; i < n2; i++) if (kps2[i].octave < 0 || kps2[i].octave >= nlevels) { orbx_set_error("kps2[%d]: octave out of range", i); return ORBX_ERR_ARG; }
    // merge of the two sorted FeatureVectors (:1176-1178, 1316-1329) -> one work item per eligible KF1 keypoint
    std::vector<int4> items;
    items.reserve(len1);
    merge_feature_vectors(fv1_node, fv1_start, fv1_idx, nn1, fv2_node, fv2_start, nn2, [&](int idx1) {
        if (has_mp1[idx1]) return false;                                          // :1188-1192
        return !(bOnlyStereo && !(uRight1 && uRight1[idx1] >= 0));                // :1194-1198
    }, &items);
    const int nitems = (int)items.size();
    if (nitems == 0) return 0;

    PairHost in = {kps1, desc1, n1, uRight1, kps2, desc2, n2, uRight2, has_mp2, fv2_idx, len2, items.data(), nitems,
                   sf2, sigma2_2, nlevels, ep, F12, bOnlyStereo, bCoarse, 0, 0.f};
    PairResult R;
    {
        const int rc = run_pairs(device, in, &R);
        if (rc != ORBX_OK) return rc;
    }
    const int* h_base = R.base; const int* h_cnt = R.cnt; const int2* h_cand = R.cand;
    // sequential part, :1210-1313
    int nmatches = 0;
    RotHist H(variant ? 1.0f / HISTO_LENGTH : HISTO_LENGTH / 360.0f);             // :1166 / :1441
    std::vector<uint8_t> vbMatched2(n2, 0);
    for (int it = 0; it < nitems; it++) {
        const int idx1 = items[it].x;
        const int2* c = h_cand + h_base[it];
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int k = 0; k < h_cnt[it]; k++) {
            const int idx2 = c[k].x & 0x3FFFFFFF, dist = c[k].y;
            if (vbMatched2[idx2]) continue;                                       // :1218
            if (dist > bestDist) continue;                                        // :1231
            if (c[k].x >> 30) { bestIdx2 = idx2; bestDist = dist; }               // :1282-1286
        }
        if (bestIdx2 >= 0) {
            matches12[idx1] = bestIdx2;
            if (!variant) vbMatched2[bestIdx2] = 1;                               // :1297 (absent in the _ variant)
            nmatches++;
            if (checkOrientation) H.push(kps1[idx1].angle, kps2[bestIdx2].angle, idx1);
        }
    }
    if (checkOrientation) {                                                       // :1332-1355
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (int idx1 : H.bin[b]) { matches12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}

extern "C" int orbx_search_by_bow(int device, int mode,
                                  const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                                  const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                  const orbx_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* valid2,
                                  const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                  float nnratio, int checkOrientation, int32_t* matches)
{
    if ((mode != 0 && mode != 1) || n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || !matches ||
        (n1 > 0 && (!kps1 || !desc1 || !valid1)) || (n2 > 0 && (!kps2 || !desc2)) || (mode == 1 && n2 > 0 && !valid2) ||
        (nn1 > 0 && (!fv1_node || !fv1_start || !fv1_idx)) || (nn2 > 0 && (!fv2_node || !fv2_start || !fv2_idx))) {
        orbx_set_error("bad search_by_bow arguments"); return ORBX_ERR_ARG;
    }
    const int nout = mode == 0 ? n2 : n1;            // per Frame keypoint (KF,F) / per KF1 keypoint (KF,KF)
    for (int i = 0; i < nout; i++) matches[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n1 == 0 || n2 == 0 || nn1 == 0 || nn2 == 0) return 0;
    const int len1 = fv1_start[nn1], len2 = fv2_start[nn2];
    for (int i = 0; i < len1; i++) if (fv1_idx[i] < 0 || fv1_idx[i] >= n1) { orbx_set_error("fv1_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    for (int i = 0; i < len2; i++) if (fv2_idx[i] < 0 || fv2_idx[i] >= n2) { orbx_set_error("fv2_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    std::vector<int4> items;
    items.reserve(len1);
    merge_feature_vectors(fv1_node, fv1_start, fv1_idx, nn1, fv2_node, fv2_start, nn2,
                          [&](int idx1) { return valid1[idx1] != 0; }, &items);   // :343-349 / :1006-1011: pMP && !pMP->isBad()
    const int nitems = (int)items.size();
    if (nitems == 0) return 0;
    std::vector<uint8_t> skip2((size_t)n2, 0);
    if (valid2) for (int i = 0; i < n2; i++) skip2[i] = valid2[i] == 0;           // :1038-1045
    PairHost in = {kps1, desc1, n1, nullptr, kps2, desc2, n2, nullptr, skip2.data(), fv2_idx, len2, items.data(), nitems,
                   nullptr, nullptr, 0, nullptr, nullptr, 0, 0, 1, nnratio};
    PairResult R;
    {
        const int rc = run_pairs(device, in, &R);
        if (rc != ORBX_OK) return rc;
    }
    int nmatches = 0;
    RotHist H(HISTO_LENGTH / 360.0f);                                             // :334 / :978
    std::vector<uint8_t> taken2((size_t)n2, 0);                                   // vpMapPointMatches[idxF] != NULL / vbMatched2
    for (int it = 0; it < nitems; it++) {
        const int idx1 = items[it].x;
        const int2* c = R.cand + R.base[it];
        int bestDist1 = 256, bestDist2 = 256, bestIdx2 = -1;
        for (int k = 0; k < R.cnt[it]; k++) {                                     // :361-384 / :1028-1062
            const int idx2 = c[k].x & 0x3FFFFFFF, dist = c[k].y;
            if (taken2[idx2]) continue;
            if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        const bool low = mode == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;    // :431 (<=) / :1064 (<)
        if (low && (float)bestDist1 < nnratio * (float)bestDist2) {               // :433 / :1066
            taken2[bestIdx2] = 1;
            if (mode == 0) matches[bestIdx2] = idx1; else matches[idx1] = bestIdx2;
            if (checkOrientation) H.push(kps1[idx1].angle, kps2[bestIdx2].angle, mode == 0 ? bestIdx2 : idx1);
            nmatches++;
        }
    }
    if (checkOrientation) {                                                       // :520-544 / :1090-1110
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (int v : H.bin[b]) { matches[v] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ---- the loop-closing / local-mapping searches of SURVEY 8a row 16: best keypoint in a window ----
//   Fuse(KeyFrame*, const vector<MapPoint*>&, th, bRight)            src/ORBmatcher.cc:1823-2035
//   Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)               :2051-2185
//   SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)         :2201-2434
// Every map point's search is independent of the others (nothing the loop mutates is read by the search
// itself), so the whole decision runs on the device: one warp per projected map point sweeps the window's
// cells (KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:916-962), keeps keypoints at levels [level-1, level]
// that pass the optional reprojection gates, and reduces (distance, visiting position) to the first best.
struct BestArgs {
    const orbx_keypoint* kps; const uint4* desc; int n; const float* uRight;
    float minX, minY, winv, hinv;
    const int* cell_start; const int* cell_idx;
    int nq; const uint4* q_desc; const float *q_x, *q_y, *q_r, *q_ur; const int* q_lvl;
    int chi2;                              // reprojection gates of the first Fuse form (:1958-1985)
    float isig2[ORBX_MAX_LEVELS];
    int2* best;                            // (idx, dist); idx = -1: no candidate
};

__global__ void __launch_bounds__(WIN_WARPS * 32)
best_in_window_kernel(BestArgs A)
{
    const int q = blockIdx.x * WIN_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= A.nq) return;
    const float r = A.q_r[q];
    int bestDist = INT_MAX, bestIdx = -1, bestPos = INT_MAX;
    if (r >= 0.f) {
        const float x = A.q_x[q], y = A.q_y[q];
        const int lvl = A.q_lvl[q];
        int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), r), A.winv)));              // src/KeyFrame.cc:924
        int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), r), A.winv)));   // :929
        const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), r), A.hinv)));        // :933
        const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), r), A.hinv)));   // :937
        if (cx0 >= GRID_COLS || cx1 < 0 || cy0 >= GRID_ROWS || cy1 < 0) { cx0 = 1; cx1 = 0; }
        const uint4 a0 = A.q_desc[2 * q], a1 = A.q_desc[2 * q + 1];
        const float ur = (A.chi2 && A.q_ur) ? A.q_ur[q] : 0.f;
        int pos = lane;                                    // visiting position: warp-uniform base + lane
        for (int ix = cx0; ix <= cx1; ix++)
            for (int iy = cy0; iy <= cy1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                const int b = A.cell_start[c], e = A.cell_start[c + 1];
                for (int j0 = b; j0 < e; j0 += 32, pos += 32) {
                    const int j = j0 + lane;
                    if (j >= e) continue;
                    const int idx = A.cell_idx[j];
                    const orbx_keypoint kp = A.kps[idx];
                    const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                    if (!(fabsf(dx) < r && fabsf(dy) < r)) continue;                                       // :955
                    if (kp.octave < lvl - 1 || kp.octave > lvl) continue;                                  // src/ORBmatcher.cc:1955
                    if (A.chi2) {
                        const float ex = __fsub_rn(x, kp.x), ey = __fsub_rn(y, kp.y);
                        const float kr = A.uRight ? A.uRight[idx] : -1.f;
                        if (kr >= 0) {                                                                     // :1958-1972
                            const float er = __fsub_rn(ur, kr);
                            const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
                            if ((double)__fmul_rn(e2, A.isig2[kp.octave]) > 7.8) continue;
                        } else {                                                                           // :1974-1985
                            const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                            if ((double)__fmul_rn(e2, A.isig2[kp.octave]) > 5.99) continue;
                        }
                    }
                    const int dist = ham256(a0, a1, A.desc[2 * idx], A.desc[2 * idx + 1]);
                    if (dist < bestDist) { bestDist = dist; bestIdx = idx; bestPos = pos; }
                }
            }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const int od = __shfl_xor_sync(0xffffffffu, bestDist, o), oi = __shfl_xor_sync(0xffffffffu, bestIdx, o);
            const int op = __shfl_xor_sync(0xffffffffu, bestPos, o);
            if (od < bestDist || (od == bestDist && op < bestPos)) { bestDist = od; bestIdx = oi; bestPos = op; }
        }
    }
    if (lane == 0) A.best[q] = make_int2(bestIdx, bestDist);
}

namespace {
struct BestHost {
    const orbx_keypoint* kps; const uint8_t* desc; int n; const float* uRight;
    float minX, maxX, minY, maxY;
    int nq; const uint8_t* q_desc; const float *q_x, *q_y, *q_r, *q_ur; const int32_t* q_lvl;
    int chi2; const float* isig2; int nlevels;
};

// one H2D copy, grid build + best_in_window_kernel, one D2H copy of nq (idx, dist) pairs (pinned; valid until the
// calling thread's next search call)
int run_best(int device, const BestHost& in, const int2** out)
{
    Workspace& W = t_ws;
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    const size_t n = (size_t)in.n, nq = (size_t)in.nq;
    Carver c;
    const size_t o_kps = c.take(n * sizeof(orbx_keypoint)), o_desc = c.take(n * 32), o_ur = c.take(in.uRight ? n * 4 : 0);
    const size_t o_qd = c.take(nq * 32), o_qx = c.take(nq * 4), o_qy = c.take(nq * 4), o_qr = c.take(nq * 4);
    const size_t o_qur = c.take(in.q_ur ? nq * 4 : 0), o_ql = c.take(nq * 4);
    const size_t in_bytes = c.off;
    const size_t o_best = c.take(nq * 8);
    const size_t out_end = c.off;
    const size_t o_cs = c.take((GRID_CELLS + 1) * 4), o_ci = c.take(n * 4);
    rc = W.reserve(c.off, out_end);
    if (rc != ORBX_OK) return rc;
    char* h = W.h; char* d = W.d;
    memcpy(h + o_kps, in.kps, n * sizeof(orbx_keypoint)); memcpy(h + o_desc, in.desc, n * 32);
    if (in.uRight) memcpy(h + o_ur, in.uRight, n * 4);
    memcpy(h + o_qd, in.q_desc, nq * 32); memcpy(h + o_qx, in.q_x, nq * 4); memcpy(h + o_qy, in.q_y, nq * 4);
    memcpy(h + o_qr, in.q_r, nq * 4); memcpy(h + o_ql, in.q_lvl, nq * 4);
    if (in.q_ur) memcpy(h + o_qur, in.q_ur, nq * 4);
    CUDA_TRYS(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, W.s));
    GridArgs g;
    g.kps = (const orbx_keypoint*)(d + o_kps); g.n = in.n; g.minX = in.minX; g.minY = in.minY;
    g.winv = (float)GRID_COLS / (float)(in.maxX - in.minX);
    g.hinv = (float)GRID_ROWS / (float)(in.maxY - in.minY);
    g.cell_start = (int*)(d + o_cs); g.cell_idx = (int*)(d + o_ci);
    orbx_launch_grid_build(g, W.s);
    BestArgs A;
    A.kps = g.kps; A.desc = (const uint4*)(d + o_desc); A.n = in.n; A.uRight = in.uRight ? (const float*)(d + o_ur) : nullptr;
    A.minX = g.minX; A.minY = g.minY; A.winv = g.winv; A.hinv = g.hinv; A.cell_start = g.cell_start; A.cell_idx = g.cell_idx;
    A.nq = in.nq; A.q_desc = (const uint4*)(d + o_qd); A.q_x = (const float*)(d + o_qx); A.q_y = (const float*)(d + o_qy);
    A.q_r = (const float*)(d + o_qr); A.q_ur = in.q_ur ? (const float*)(d + o_qur) : nullptr; A.q_lvl = (const int*)(d + o_ql);
    A.chi2 = in.chi2;
    for (int l = 0; l < ORBX_MAX_LEVELS; l++) A.isig2[l] = (in.isig2 && l < in.nlevels) ? in.isig2[l] : 0.f;
    A.best = (int2*)(d + o_best);
    best_in_window_kernel<<<(in.nq + WIN_WARPS - 1) / WIN_WARPS, WIN_WARPS * 32, 0, W.s>>>(A);
    CUDA_TRYS(cudaGetLastError());
    CUDA_TRYS(cudaMemcpyAsync(h + o_best, d + o_best, nq * 8, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYS(cudaStreamSynchronize(W.s));
    *out = (const int2*)(h + o_best);
    return ORBX_OK;
}

// window of a projected map point: KeyFrame::IsInImage (src/KeyFrame.cc:965-968), radius th * scale[level]; -1 = no search
int keyframe_windows(int nmp, const float* mp_u, const float* mp_v, const int32_t* mp_level, const uint8_t* mp_valid,
                     float minX, float maxX, float minY, float maxY, const float* scaleFactors, int nlevels, float th,
                     std::vector<float>* q_r)
{
    q_r->assign((size_t)nmp, -1.f);
    for (int i = 0; i < nmp; i++) {
        if (!mp_valid[i]) continue;
        if (!(mp_u[i] >= minX && mp_u[i] < maxX && mp_v[i] >= minY && mp_v[i] < maxY)) continue;
        const int lvl = mp_level[i];
        if (lvl < 0 || lvl >= nlevels) { orbx_set_error("map point %d: level %d out of range", i, lvl); return ORBX_ERR_ARG; }
        const float r = th * scaleFactors[lvl];
        if (r >= 0.f) (*q_r)[i] = r;
    }
    return ORBX_OK;
}
}  // namespace

extern "C" int orbx_search_by_projection_sim3(int device,
                                              const orbx_keypoint* kps, const uint8_t* desc, int n,
                                              float minX, float maxX, float minY, float maxY,
                                              const float* scaleFactors, int nlevels,
                                              int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                              const int32_t* mp_level, const uint8_t* mp_valid,
                                              float th, float ratioHamming, const uint8_t* matched, int32_t* assign)
{
    if (bad_frame(kps, desc, n, minX, maxX, minY, maxY) || nmp < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors ||
        (n > 0 && !assign) || (nmp > 0 && (!mp_desc || !mp_u || !mp_v || !mp_level || !mp_valid))) {
        orbx_set_error("bad search_by_projection_sim3 arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n; i++) assign[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n == 0 || nmp == 0) return 0;
    std::vector<float> q_r;
    int rc = keyframe_windows(nmp, mp_u, mp_v, mp_level, mp_valid, minX, maxX, minY, maxY, scaleFactors, nlevels, th, &q_r);
    if (rc != ORBX_OK) return rc;
    std::vector<int32_t> q_min(nmp), q_max(nmp);
    for (int i = 0; i < nmp; i++) { q_min[i] = mp_level[i] - 1; q_max[i] = mp_level[i]; }    // :653
    // bestDist <= TH_LOW * ratioHamming (:667, int against float) <=> bestDist <= floor of the float product
    const float lim = (float)TH_LOW * ratioHamming;
    const int maxDist = lim >= 255.f ? 255 : (lim < 0.f ? -1 : (int)floorf(lim));   // 256 = "no candidate"
    return projection_replay(device, kps, desc, n, nullptr, minX, maxX, minY, maxY, nmp, mp_desc, mp_u, mp_v,
                             q_r.data(), nullptr, q_min.data(), q_max.data(), nullptr, nullptr, maxDist, 0, matched, assign);
}

extern "C" int orbx_fuse(int device,
                         const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                         float minX, float maxX, float minY, float maxY,
                         const float* scaleFactors, const float* invLevelSigma2, int nlevels,
                         int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v, const float* mp_ur,
                         const int32_t* mp_level, const uint8_t* mp_valid, float th,
                         int32_t* best_idx, int32_t* best_dist)
{
    if (bad_frame(kps, desc, n, minX, maxX, minY, maxY) || nmp < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors ||
        (mp_ur && !invLevelSigma2) ||
        (nmp > 0 && (!mp_desc || !mp_u || !mp_v || !mp_level || !mp_valid || !best_idx || !best_dist))) {
        orbx_set_error("bad fuse arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < nmp; i++) { best_idx[i] = -1; best_dist[i] = 256; }
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n == 0 || nmp == 0) return 0;
    std::vector<float> q_r;
    int rc = keyframe_windows(nmp, mp_u, mp_v, mp_level, mp_valid, minX, maxX, minY, maxY, scaleFactors, nlevels, th, &q_r);
    if (rc != ORBX_OK) return rc;
    BestHost in = {kps, desc, n, uRight, minX, maxX, minY, maxY, nmp, mp_desc, mp_u, mp_v, q_r.data(), mp_ur, mp_level,
                   mp_ur != nullptr, invLevelSigma2, nlevels};
    const int2* best = nullptr;
    rc = run_best(device, in, &best);
    if (rc != ORBX_OK) return rc;
    int nFused = 0;
    for (int i = 0; i < nmp; i++) {
        if (best[i].x < 0) continue;
        best_dist[i] = best[i].y;
        if (best[i].y <= TH_LOW) { best_idx[i] = best[i].x; nFused++; }                // :1999 / :2163
    }
    return nFused;
}

extern "C" int orbx_search_by_sim3(int device,
                                   const orbx_keypoint* kps1, const uint8_t* desc1, int n1,
                                   const orbx_keypoint* kps2, const uint8_t* desc2, int n2,
                                   float minX, float maxX, float minY, float maxY,
                                   const float* scaleFactors, int nlevels,
                                   const uint8_t* mp1_desc, const float* u1in2, const float* v1in2, const int32_t* level1, const uint8_t* valid1,
                                   const uint8_t* mp2_desc, const float* u2in1, const float* v2in1, const int32_t* level2, const uint8_t* valid2,
                                   float th, int32_t* match12)
{
    if (bad_frame(kps1, desc1, n1, minX, maxX, minY, maxY) || bad_frame(kps2, desc2, n2, minX, maxX, minY, maxY) ||
        nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors || (n1 > 0 && !match12) ||
        (n1 > 0 && (!mp1_desc || !u1in2 || !v1in2 || !level1 || !valid1)) || (n2 > 0 && (!mp2_desc || !u2in1 || !v2in1 || !level2 || !valid2))) {
        orbx_set_error("bad search_by_sim3 arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n1; i++) match12[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<float> q_r;
    // KF1's map points in KF2 (:2248-2330)
    int rc = keyframe_windows(n1, u1in2, v1in2, level1, valid1, minX, maxX, minY, maxY, scaleFactors, nlevels, th, &q_r);
    if (rc != ORBX_OK) return rc;
    BestHost a = {kps2, desc2, n2, nullptr, minX, maxX, minY, maxY, n1, mp1_desc, u1in2, v1in2, q_r.data(), nullptr, level1, 0, nullptr, nlevels};
    const int2* best = nullptr;
    rc = run_best(device, a, &best);
    if (rc != ORBX_OK) return rc;
    std::vector<int> vnMatch1((size_t)n1, -1), vnMatch2((size_t)n2, -1);
    for (int i = 0; i < n1; i++) if (best[i].x >= 0 && best[i].y <= TH_HIGH) vnMatch1[i] = best[i].x;   // :2326
    // KF2's map points in KF1 (:2333-2412)
    rc = keyframe_windows(n2, u2in1, v2in1, level2, valid2, minX, maxX, minY, maxY, scaleFactors, nlevels, th, &q_r);
    if (rc != ORBX_OK) return rc;
    BestHost b = {kps1, desc1, n1, nullptr, minX, maxX, minY, maxY, n2, mp2_desc, u2in1, v2in1, q_r.data(), nullptr, level2, 0, nullptr, nlevels};
    rc = run_best(device, b, &best);
    if (rc != ORBX_OK) return rc;
    for (int i = 0; i < n2; i++) if (best[i].x >= 0 && best[i].y <= TH_HIGH) vnMatch2[i] = best[i].x;
    int nFound = 0;
    for (int i1 = 0; i1 < n1; i1++) {                                                   // :2415-2431
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && vnMatch2[idx2] == i1) { match12[i1] = idx2; nFound++; }
    }
    return nFound;
}

// ---- two-camera rig (Nleft != -1) branches: config C3's KannalaBrandt stereo set-up ----

// SearchByProjection(Frame& Current, const Frame& Last, th, bMono), src/ORBmatcher.cc:2469-2711, CurrentFrame.Nleft != -1.
// Two window sweeps (left grid over mvKeys, right grid over mvKeysRight with the right projections) and one host replay
// over rows 0..nlast-1 in order: left half (:2540-2601), then right half (:2610-2684) - skipped together with the left
// one when the left WINDOW is empty (`continue`, :2541).
extern "C" int orbx_search_by_projection_frame_rig(int device,
                                                   const orbx_keypoint* kpsL, int nL, const orbx_keypoint* kpsR, int nR, const uint8_t* desc,
                                                   float minX, float maxX, float minY, float maxY,
                                                   const float* scaleFactors, int nlevels,
                                                   int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                                   const float* last_ur, const float* last_vr,
                                                   const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                                   const uint8_t* last_valid, const uint8_t* last_obs,
                                                   float th, int bForward, int bBackward, int checkOrientation,
                                                   const uint8_t* cur_held, int32_t* assign)
{
    const int n = nL + nR;
    if (nL < 0 || nR < 0 || (nL > 0 && !kpsL) || (nR > 0 && !kpsR) || (n > 0 && (!desc || !assign)) || !(maxX > minX) || !(maxY > minY) ||
        nlast < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors ||
        (nlast > 0 && (!last_desc || !last_u || !last_v || !last_ur || !last_vr || !last_invz || !last_octave || !last_angle || !last_valid || !last_obs))) {
        orbx_set_error("bad search_by_projection_frame_rig arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n; i++) assign[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n == 0 || nlast == 0) return 0;
    std::vector<float> q_r(nlast);
    std::vector<int32_t> q_min(nlast), q_max(nlast);
    for (int i = 0; i < nlast; i++) {
        q_r[i] = -1.f; q_min[i] = -1; q_max[i] = -1;
        if (!last_valid[i]) continue;                                             // :2503-2506
        if (last_invz[i] < 0) continue;                                           // :2515
        if (last_u[i] < minX || last_u[i] > maxX) continue;                       // :2520-2523 (left projection only)
        if (last_v[i] < minY || last_v[i] > maxY) continue;
        const int oct = last_octave[i];
        if (oct < 0 || oct >= nlevels) { orbx_set_error("last-frame point %d: octave %d out of range", i, oct); return ORBX_ERR_ARG; }
        q_r[i] = th * scaleFactors[oct];                                          // :2529 / :2620
        if (!(q_r[i] >= 0.f)) { q_r[i] = -1.f; continue; }
        if (bForward) { q_min[i] = oct; q_max[i] = -1; }
        else if (bBackward) { q_min[i] = 0; q_max[i] = oct; }
        else { q_min[i] = oct - 1; q_max[i] = oct + 1; }
    }
    // the results of run_windows live in the calling thread's pinned workspace: keep the left sweep before the right one
    std::vector<int> cntL(nlast, 0), baseL(nlast, 0);
    std::vector<int2> candL;
    if (nL > 0) {
        WinHost in = {kpsL, desc, nL, nullptr, minX, maxX, minY, maxY, nlast, last_desc, last_u, last_v, q_r.data(), nullptr, q_min.data(), q_max.data()};
        WinResult R;
        const int rc = run_windows(device, in, &R);
        if (rc != ORBX_OK) return rc;
        cntL.assign(R.cnt, R.cnt + nlast); baseL.assign(R.base, R.base + nlast);
        candL.assign(R.cand, R.cand + R.total);
    }
    WinResult RR = {nullptr, nullptr, nullptr, 0};
    if (nR > 0) {
        WinHost in = {kpsR, desc + 32 * (size_t)nL, nR, nullptr, minX, maxX, minY, maxY, nlast, last_desc, last_ur, last_vr, q_r.data(), nullptr,
                      q_min.data(), q_max.data()};
        const int rc = run_windows(device, in, &RR);
        if (rc != ORBX_OK) return rc;
    }
    int nmatches = 0;
    RotHist H(HISTO_LENGTH / 360.0f);                                             // :2480
    std::vector<uint8_t> held(n, 0);
    if (cur_held) for (int i = 0; i < n; i++) held[i] = cur_held[i] != 0;
    for (int i = 0; i < nlast; i++) {
        if (cntL[i] == 0) continue;                                               // :2541
        {
            const int2* c = candL.data() + baseL[i];
            int bestDist = 256, bestIdx2 = -1;
            for (int k = 0; k < cntL[i]; k++) {
                if (held[c[k].x]) continue;                                       // :2552-2554
                if (c[k].y < bestDist) { bestDist = c[k].y; bestIdx2 = c[k].x; }
            }
            if (bestDist <= TH_HIGH) {
                assign[bestIdx2] = i; held[bestIdx2] = last_obs[i] != 0; nmatches++;
                if (checkOrientation) H.push(last_angle[i], kpsL[bestIdx2].angle, bestIdx2);
            }
        }
        if (nR > 0 && RR.cnt[i] > 0) {
            const int2* c = RR.cand + RR.base[i];
            int bestDist = 256, bestIdx2 = -1;
            for (int k = 0; k < RR.cnt[i]; k++) {
                if (held[c[k].x + nL]) continue;                                  // :2640-2642
                if (c[k].y < bestDist) { bestDist = c[k].y; bestIdx2 = c[k].x; }
            }
            if (bestDist <= TH_HIGH) {
                assign[bestIdx2 + nL] = i; held[bestIdx2 + nL] = last_obs[i] != 0; nmatches++;
                if (checkOrientation) H.push(last_angle[i], kpsR[bestIdx2].angle, bestIdx2 + nL);
            }
        }
    }
    if (checkOrientation) {                                                       // :2686-2708
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++)
            if (b != i1 && b != i2 && b != i3)
                for (int v : H.bin[b]) { assign[v] = -1; nmatches--; }
    }
    return nmatches;
}

// SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&), src/ORBmatcher.cc:314-547, F.Nleft != -1: the pair sweep is that
// of orbx_search_by_bow(mode 0); the replay keeps best / second per camera (:380-430), accepts the right-camera match
// only inside the left one's TH_LOW test and without a ratio test of its own (:431, :455-457).
extern "C" int orbx_search_by_bow_rig(int device,
                                      const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                                      const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                      const orbx_keypoint* kps2, const uint8_t* desc2, int n2, int nleft2,
                                      const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                      float nnratio, int checkOrientation, int32_t* matches)
{
    if (n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || nleft2 < 0 || nleft2 > n2 || (n2 > 0 && !matches) ||
        (n1 > 0 && (!kps1 || !desc1 || !valid1)) || (n2 > 0 && (!kps2 || !desc2)) ||
        (nn1 > 0 && (!fv1_node || !fv1_start || !fv1_idx)) || (nn2 > 0 && (!fv2_node || !fv2_start || !fv2_idx))) {
        orbx_set_error("bad search_by_bow_rig arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n2; i++) matches[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n1 == 0 || n2 == 0 || nn1 == 0 || nn2 == 0) return 0;
    const int len1 = fv1_start[nn1], len2 = fv2_start[nn2];
    for (int i = 0; i < len1; i++) if (fv1_idx[i] < 0 || fv1_idx[i] >= n1) { orbx_set_error("fv1_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    for (int i = 0; i < len2; i++) if (fv2_idx[i] < 0 || fv2_idx[i] >= n2) { orbx_set_error("fv2_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    std::vector<int4> items;
    items.reserve(len1);
    merge_feature_vectors(fv1_node, fv1_start, fv1_idx, nn1, fv2_node, fv2_start, nn2, [&](int idx1) { return valid1[idx1] != 0; }, &items);
    const int nitems = (int)items.size();
    if (nitems == 0) return 0;
    std::vector<uint8_t> skip2((size_t)n2, 0);
    // a distance with ratio * d > TH_LOW can neither be accepted (left: best1 <= TH_LOW and best1 < ratio * best2; right:
    // best1R <= TH_LOW) nor make the left ratio test fail, so the pair sweep need not report it.  For ratio > 1 that bound
    // would drop right-camera candidates <= TH_LOW: report everything then.
    PairHost in = {kps1, desc1, n1, nullptr, kps2, desc2, n2, nullptr, skip2.data(), fv2_idx, len2, items.data(), nitems,
                   nullptr, nullptr, 0, nullptr, nullptr, 0, 0, 1, nnratio > 1.f ? 1.f : nnratio};
    PairResult R;
    {
        const int rc = run_pairs(device, in, &R);
        if (rc != ORBX_OK) return rc;
    }
    int nmatches = 0;
    RotHist H(HISTO_LENGTH / 360.0f);                                             // :334
    std::vector<uint8_t> taken2((size_t)n2, 0);
    for (int it = 0; it < nitems; it++) {
        const int idx1 = items[it].x;
        const int2* c = R.cand + R.base[it];
        int bestDist1 = 256, bestDist2 = 256, bestIdxF = -1, bestDist1R = 256, bestDist2R = 256, bestIdxFR = -1;
        for (int k = 0; k < R.cnt[it]; k++) {                                     // :386-430
            const int idx2 = c[k].x & 0x3FFFFFFF, dist = c[k].y;
            if (taken2[idx2]) continue;
            if (idx2 < nleft2) {
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = idx2; }
                else if (dist < bestDist2) bestDist2 = dist;
            } else {
                if (dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = idx2; }
                else if (dist < bestDist2R) bestDist2R = dist;
            }
        }
        (void)bestDist2R;
        if (bestDist1 <= TH_LOW) {                                                // :431
            if ((float)bestDist1 < nnratio * (float)bestDist2) {                  // :433
                matches[bestIdxF] = idx1; taken2[bestIdxF] = 1;
                if (checkOrientation) H.push(kps1[idx1].angle, kps2[bestIdxF].angle, bestIdxF);
                nmatches++;
            }
            if (bestDist1R <= TH_LOW) {                                           // :455-457 (`|| true`)
                matches[bestIdxFR] = idx1; taken2[bestIdxFR] = 1;
                if (checkOrientation) H.push(kps1[idx1].angle, kps2[bestIdxFR].angle, bestIdxFR);
                nmatches++;
            }
        }
    }
    if (checkOrientation) {                                                       // :520-544
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (int v : H.bin[b]) { matches[v] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// The triangulation searches whose geometric verdict is the camera model's own (caller-side) code - see the
// declaration in include/orbx.h for the three overloads.  The device sweeps every (KF1 keypoint, KF2 keypoint) pair
// under a common vocabulary node and keeps those with !pMP2 and dist <= TH_LOW, in FeatureVector order; the replay
// asks `verdict` exactly where the reference calls epipolarConstrain / matchAndtriangulate (dist <= bestDist so far).
extern "C" int orbx_search_for_triangulation_cb(int device, int overload,
                                                const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                                const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                                const orbx_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                                const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                                int bOnlyStereo, int bCoarse, int checkOrientation,
                                                orbx_pair_verdict_fn verdict, void* user, int32_t* matches12)
{
    if (overload < 0 || overload > 2 || n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || (n1 > 0 && (!kps1 || !desc1 || !has_mp1 || !matches12)) ||
        (n2 > 0 && (!kps2 || !desc2 || !has_mp2)) || (nn1 > 0 && (!fv1_node || !fv1_start || !fv1_idx)) ||
        (nn2 > 0 && (!fv2_node || !fv2_start || !fv2_idx)) || (!verdict && !(bCoarse && overload != 2))) {
        orbx_set_error("bad search_for_triangulation_cb arguments"); return ORBX_ERR_ARG;
    }
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    if (n1 == 0 || n2 == 0 || nn1 == 0 || nn2 == 0) return 0;
    if (overload != 2 && bOnlyStereo) return 0;                                   // bStereo1 is false with a rig: every idx1 is skipped (:1194-1198)
    const int len1 = fv1_start[nn1], len2 = fv2_start[nn2];
    for (int i = 0; i < len1; i++) if (fv1_idx[i] < 0 || fv1_idx[i] >= n1) { orbx_set_error("fv1_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    for (int i = 0; i < len2; i++) if (fv2_idx[i] < 0 || fv2_idx[i] >= n2) { orbx_set_error("fv2_idx[%d] out of range", i); return ORBX_ERR_ARG; }
    std::vector<int4> items;
    items.reserve(len1);
    merge_feature_vectors(fv1_node, fv1_start, fv1_idx, nn1, fv2_node, fv2_start, nn2, [&](int idx1) { return !has_mp1[idx1]; }, &items);
    const int nitems = (int)items.size();
    if (nitems == 0) return 0;
    PairHost in = {kps1, desc1, n1, nullptr, kps2, desc2, n2, nullptr, has_mp2, fv2_idx, len2, items.data(), nitems,
                   nullptr, nullptr, 0, nullptr, nullptr, 0, 0, 2, 0.f};
    PairResult R;
    {
        const int rc = run_pairs(device, in, &R);
        if (rc != ORBX_OK) return rc;
    }
    // the candidate lists live in the calling thread's workspace, which `verdict` may reuse (another orbx call): copy them
    const std::vector<int> base(R.base, R.base + nitems), cnt(R.cnt, R.cnt + nitems);
    const std::vector<int2> cand(R.cand, R.cand + R.total);
    int nmatches = 0;
    RotHist H(overload == 0 ? HISTO_LENGTH / 360.0f : 1.0f / HISTO_LENGTH);       // :1166 / :1441 / :1666
    std::vector<uint8_t> vbMatched2(n2, 0);
    for (int it = 0; it < nitems; it++) {
        const int idx1 = items[it].x;
        const int2* c = cand.data() + base[it];
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int k = 0; k < cnt[it]; k++) {
            const int idx2 = c[k].x & 0x3FFFFFFF, dist = c[k].y;
            if (vbMatched2[idx2]) continue;                                       // :1218
            if (dist > bestDist) continue;                                        // :1231 / :1707
            // :1282 `epipolarConstrain(...) || bCoarse` evaluates the camera model first: it is asked even when bCoarse overrides it
            bool ok = verdict ? verdict(user, idx1, idx2) != 0 : false;                           // :1282 / :1729
            if (overload != 2 && bCoarse) ok = true;
            if (ok) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
            matches12[idx1] = bestIdx2;
            if (overload == 0) vbMatched2[bestIdx2] = 1;                          // :1297 only
            nmatches++;
            if (checkOrientation) H.push(kps1[idx1].angle, kps2[bestIdx2].angle, idx1);
        }
    }
    if (checkOrientation) {
        int i1, i2, i3;
        H.top3(&i1, &i2, &i3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (int v : H.bin[b]) { matches12[v] = -1; nmatches--; }
        }
    }
    return nmatches;
}

