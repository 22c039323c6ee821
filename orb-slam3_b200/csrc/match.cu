// match.cu — Hamming matchers: brute-force 2-NN (cv::BFMatcher(NORM_HAMMING).knnMatch k=2 as used by
// Frame::ComputeStereoFishEyeMatches, reference src/Frame.cc:1440-1480) and
// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) (src/ORBmatcher.cc:45-168) over the
// 64x48 frame grid (Frame::AssignFeaturesToGrid / GetFeaturesInArea, src/Frame.cc:446-480, 784-899).
// Integer pipe work (XOR + POPC); no tensor cores: nothing here is a dense contraction.
#include "match_internal.h"
#include <climits>
#include <vector>

#define CUDA_TRYM(expr)                                                                        \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            orbx_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            rc = ORBX_ERR_CUDA; goto done;                                                     \
        }                                                                                      \
    } while (0)

// ---- brute-force 2-NN: one thread per query, train rows broadcast from shared memory ----
#define KNN_THREADS 128
#define KNN_TILE 128

__global__ void __launch_bounds__(KNN_THREADS)
knn2_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t, int nt, int* __restrict__ out_idx, int* __restrict__ out_dist)
{
    __shared__ uint4 s_t[KNN_TILE * 2];
    const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (i < nq) { a0 = q[2 * i]; a1 = q[2 * i + 1]; }
    int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
    for (int j0 = 0; j0 < nt; j0 += KNN_TILE) {
        const int m = min(KNN_TILE, nt - j0);
        __syncthreads();
        for (int k = threadIdx.x; k < 2 * m; k += KNN_THREADS) s_t[k] = t[2 * j0 + k];
        __syncthreads();
        for (int j = 0; j < m; j++) {
            const int d = ham256(a0, a1, s_t[2 * j], s_t[2 * j + 1]);
            if (d < b0) { b1 = b0; i1 = i0; b0 = d; i0 = j0 + j; }     // ties keep the lower train index
            else if (d < b1) { b1 = d; i1 = j0 + j; }
        }
    }
    if (i < nq) {
        out_idx[2 * i] = i0; out_idx[2 * i + 1] = i1;
        out_dist[2 * i] = i0 >= 0 ? b0 : -1; out_dist[2 * i + 1] = i1 >= 0 ? b1 : -1;
    }
}

extern "C" int orbx_knn2_hamming(int device, const uint8_t* q, int nq, const uint8_t* t, int nt,
                                 int32_t* out_idx, int32_t* out_dist)
{
    if (nq < 0 || nt < 0 || (nq > 0 && (!q || !out_idx || !out_dist)) || (nt > 0 && !t)) { orbx_set_error("bad knn2 arguments"); return ORBX_ERR_ARG; }
    if (nq == 0) return ORBX_OK;
    int rc = ORBX_OK;
    uint8_t *dq = nullptr, *dt = nullptr; int *di = nullptr, *dd = nullptr;
    cudaStream_t s = nullptr;
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    CUDA_TRYM(cudaSetDevice(device));
    CUDA_TRYM(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    CUDA_TRYM(cudaMalloc(&dq, (size_t)nq * 32)); CUDA_TRYM(cudaMalloc(&dt, (size_t)(nt > 0 ? nt : 1) * 32));
    CUDA_TRYM(cudaMalloc(&di, (size_t)nq * 8)); CUDA_TRYM(cudaMalloc(&dd, (size_t)nq * 8));
    CUDA_TRYM(cudaMemcpyAsync(dq, q, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
    if (nt > 0) CUDA_TRYM(cudaMemcpyAsync(dt, t, (size_t)nt * 32, cudaMemcpyHostToDevice, s));
    knn2_kernel<<<(nq + KNN_THREADS - 1) / KNN_THREADS, KNN_THREADS, 0, s>>>((const uint4*)dq, nq, (const uint4*)dt, nt, di, dd);
    CUDA_TRYM(cudaGetLastError());
    CUDA_TRYM(cudaMemcpyAsync(out_idx, di, (size_t)nq * 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRYM(cudaMemcpyAsync(out_dist, dd, (size_t)nq * 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRYM(cudaStreamSynchronize(s));
done:
    cudaFree(dq); cudaFree(dt); cudaFree(di); cudaFree(dd);
    if (s) cudaStreamDestroy(s);
    return rc;
}

// ---- frame grid + SearchByProjection ----
#define PROJ_KMAX 256

struct ProjArgs {
    const orbx_keypoint* kps; const uint4* desc; int n; const float* uRight;
    float minX, minY, winv, hinv;
    float sf[ORBX_MAX_LEVELS];
    int nmp; const uint4* mp_desc; const float *projx, *projy, *projxr; const int* level; const float *viewcos, *depth;
    const uint8_t* inview;
    float th, nnratio; int bFar; float thFar;
    int* cell_start;      // GRID_CELLS + 1
    int* cell_idx;        // n
    int* cand_cnt;        // nmp
    int* cand;            // nmp * PROJ_KMAX * 2 : (idx | octave << 24, dist)
    int* assign;          // n (in/out)
    int* claim0; int* claim1;   // n
    int* choice;          // nmp
    int* result;          // [0] nmatches, [1] overflow flag
};

// Frame::AssignFeaturesToGrid (src/Frame.cc:446-480) + PosInGrid (:883-899): CSR, insertion order kept
__global__ void __launch_bounds__(1024)
grid_build_kernel(GridArgs A)
{
    __shared__ int s_cnt[GRID_CELLS + 1];
    __shared__ int s_warp[32];
    const int tid = threadIdx.x;
    for (int c = tid; c <= GRID_CELLS; c += 1024) s_cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < A.n; i += 1024) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(A.kps[i].x, A.minX), A.winv));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(A.kps[i].y, A.minY), A.hinv));
        if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) atomicAdd(&s_cnt[px * GRID_ROWS + py], 1);
    }
    __syncthreads();
    // exclusive scan of 3072 counts: 3 per thread
    int v[3], sum = 0;
    for (int k = 0; k < 3; k++) { v[k] = s_cnt[3 * tid + k]; sum += v[k]; }
    int incl = sum;
    for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= o) incl += u; }
    if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < (tid >> 5); w++) base += s_warp[w];
    int run = base + incl - sum;
    __syncthreads();
    for (int k = 0; k < 3; k++) { s_cnt[3 * tid + k] = run; A.cell_start[3 * tid + k] = run; run += v[k]; }
    if (tid == 1023) A.cell_start[GRID_CELLS] = run;
    __syncthreads();
    for (int i = tid; i < A.n; i += 1024) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(A.kps[i].x, A.minX), A.winv));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(A.kps[i].y, A.minY), A.hinv));
        if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) A.cell_idx[atomicAdd(&s_cnt[px * GRID_ROWS + py], 1)] = i;
    }
    __syncthreads();
    // restore push_back order (ascending keypoint index) inside every cell
    for (int c = tid; c < GRID_CELLS; c += 1024) {
        const int b = A.cell_start[c], e = s_cnt[c];
        for (int i = b + 1; i < e; i++) {
            const int x = A.cell_idx[i];
            int j = i - 1;
            while (j >= b && A.cell_idx[j] > x) { A.cell_idx[j + 1] = A.cell_idx[j]; j--; }
            A.cell_idx[j + 1] = x;
        }
    }
}

void orbx_launch_grid_build(const GridArgs& g, cudaStream_t stream) { grid_build_kernel<<<1, 1024, 0, stream>>>(g); }

// candidate lists in the reference's visiting order (GetFeaturesInArea, src/Frame.cc:784-871) with the
// gates that do not depend on earlier assignments (src/ORBmatcher.cc:57-117) and the Hamming distance.
#define PROJ_WARPS 8
__global__ void __launch_bounds__(PROJ_WARPS * 32)
proj_candidates_kernel(ProjArgs A)
{
    const int iMP = blockIdx.x * PROJ_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (iMP >= A.nmp) return;
    int cnt = 0;
    bool live = A.inview[iMP] != 0;                                       // :54
    if (A.bFar && A.depth[iMP] > A.thFar) live = false;                    // :57
    if (live) {
        const int lvl = A.level[iMP];
        float r = ((double)A.viewcos[iMP] > 0.998) ? 2.5f : 4.0f;          // RadiusByViewingCos :242-249
        if ((double)A.th != 1.0) r = __fmul_rn(r, A.th);                   // :49,:79
        const float rr = __fmul_rn(r, A.sf[lvl]);
        const float x = A.projx[iMP], y = A.projy[iMP];
        const int minLevel = lvl - 1, maxLevel = lvl;
        int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), rr), A.winv)));
        int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), rr), A.winv)));
        int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), rr), A.hinv)));
        int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), rr), A.hinv)));
        if (cx0 >= GRID_COLS || cx1 < 0 || cy0 >= GRID_ROWS || cy1 < 0) { cx0 = 1; cx1 = 0; }
        const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);        // :828
        const uint4 a0 = A.mp_desc[2 * iMP], a1 = A.mp_desc[2 * iMP + 1];
        int* out = A.cand + (long long)iMP * PROJ_KMAX * 2;
        for (int ix = cx0; ix <= cx1; ix++)
            for (int iy = cy0; iy <= cy1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                const int b = A.cell_start[c], e = A.cell_start[c + 1];
                for (int j0 = b; j0 < e; j0 += 32) {
                    const int j = j0 + lane;
                    bool ok = j < e;
                    int idx = 0, oct = 0, dist = 0;
                    if (ok) {
                        idx = A.cell_idx[j];
                        const orbx_keypoint kp = A.kps[idx];
                        oct = kp.octave;
                        if (checkLevels) {
                            if (oct < minLevel) ok = false;
                            if (maxLevel >= 0 && oct > maxLevel) ok = false;
                        }
                        const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                        if (!(fabsf(dx) < rr && fabsf(dy) < rr)) ok = false;
                        if (ok && A.uRight && A.uRight[idx] > 0) {         // :107-117
                            const float er = fabsf(__fsub_rn(A.projxr[iMP], A.uRight[idx]));
                            if (er > rr) ok = false;
                        }
                        if (ok) dist = ham256(a0, a1, A.desc[2 * idx], A.desc[2 * idx + 1]);
                    }
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    if (ok) {
                        const int p = cnt + __popc(bal & ((1u << lane) - 1));
                        if (p < PROJ_KMAX) { out[2 * p] = idx | (oct << 24); out[2 * p + 1] = dist; }
                    }
                    cnt += __popc(bal);
                }
            }
    }
    if (lane == 0) {
        if (cnt > PROJ_KMAX) { A.result[1] = 1; cnt = PROJ_KMAX; }
        A.cand_cnt[iMP] = cnt;
    }
}

// Replays the reference's sequential acceptance (a keypoint taken by an earlier map point is skipped,
// :102-104) as a fixpoint: iteration t decides every map point against the claims of iteration t-1; a
// map point's decision only depends on EARLIER map points, so the first t map points are final after t
// iterations and equality of two consecutive iterations proves the whole vector final.
__global__ void __launch_bounds__(1024)
proj_resolve_kernel(ProjArgs A)
{
    __shared__ int s_changed, s_matches;
    const int tid = threadIdx.x;
    int* prev = A.claim0; int* next = A.claim1;
    for (int i = tid; i < A.n; i += 1024) prev[i] = A.assign[i] >= 0 ? -1 : INT_MAX;
    for (int i = tid; i < A.nmp; i += 1024) A.choice[i] = -2;
    __syncthreads();
    for (int iter = 0; iter <= A.nmp + 1; iter++) {
        if (tid == 0) { s_changed = 0; s_matches = 0; }
        for (int i = tid; i < A.n; i += 1024) next[i] = A.assign[i] >= 0 ? -1 : INT_MAX;
        __syncthreads();
        int changed = 0, matches = 0;
        for (int i = tid; i < A.nmp; i += 1024) {
            const int* c = A.cand + (long long)i * PROJ_KMAX * 2;
            const int nc = A.cand_cnt[i];
            int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
            for (int k = 0; k < nc; k++) {                                 // :100-141
                const int idx = c[2 * k] & 0xFFFFFF, oct = c[2 * k] >> 24, dist = c[2 * k + 1];
                if (prev[idx] < i) continue;                               // already holds a map point
                if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = idx; }
                else if (dist < bestDist2) { bestLevel2 = oct; bestDist2 = dist; }
            }
            int ch = -1;
            if (bestDist <= TH_HIGH) {                                     // :147-165
                const float lim = __fmul_rn(A.nnratio, (float)bestDist2);
                if (!(bestLevel == bestLevel2 && (float)bestDist > lim)) ch = bestIdx;
            }
            if (ch >= 0) { atomicMin(&next[ch], i); matches++; }
            if (ch != A.choice[i]) { changed = 1; A.choice[i] = ch; }
        }
        if (changed) atomicOr(&s_changed, 1);
        if (matches) atomicAdd(&s_matches, matches);
        __syncthreads();
        int* t = prev; prev = next; next = t;
        const int ch = s_changed;
        __syncthreads();
        if (!ch) break;
    }
    for (int i = tid; i < A.n; i += 1024) if (A.assign[i] < 0 && prev[i] != INT_MAX) A.assign[i] = prev[i];
    if (tid == 0) A.result[0] = s_matches;
}

extern "C" int orbx_search_by_projection(int device,
                                         const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                                         float minX, float maxX, float minY, float maxY,
                                         const float* scaleFactors, int nlevels,
                                         int nmp, const uint8_t* mp_desc, const float* mp_projx, const float* mp_projy,
                                         const float* mp_projxr, const int32_t* mp_level, const float* mp_viewcos,
                                         const float* mp_depth, const uint8_t* mp_inview,
                                         float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    if (n < 0 || nmp < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors || !(maxX > minX) || !(maxY > minY) ||
        (n > 0 && (!kps || !desc || !assign)) ||
        (nmp > 0 && (!mp_desc || !mp_projx || !mp_projy || !mp_level || !mp_viewcos || !mp_inview)) ||
        (uRight && !mp_projxr) || (bFarPoints && !mp_depth)) {
        orbx_set_error("bad search_by_projection arguments"); return ORBX_ERR_ARG;
    }
    if (n >= (1 << 24)) { orbx_set_error("too many keypoints"); return ORBX_ERR_ARG; }
    if (n == 0 || nmp == 0) return 0;
    for (int i = 0; i < nmp; i++)
        if (mp_level[i] < 0 || mp_level[i] >= nlevels) { orbx_set_error("map point %d: level %d out of range", i, mp_level[i]); return ORBX_ERR_ARG; }
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    int rc = ORBX_OK;
    int result[2] = {0, 0};
    ProjArgs A;
    memset(&A, 0, sizeof(A));
    std::vector<void*> allocs;
    cudaStream_t s = nullptr;
    auto up = [&](const void* h, size_t bytes, cudaError_t* ce) -> void* {
        void* d = nullptr;
        *ce = cudaMalloc(&d, bytes ? bytes : 4);
        if (*ce != cudaSuccess) return nullptr;
        allocs.push_back(d);
        if (h) *ce = cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s);
        return d;
    };
    cudaError_t ce = cudaSuccess;
    CUDA_TRYM(cudaSetDevice(device));
    CUDA_TRYM(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
#define UP(dst, type, src, bytes) do { dst = (type)up(src, bytes, &ce); CUDA_TRYM(ce); } while (0)
    UP(A.kps, const orbx_keypoint*, kps, sizeof(orbx_keypoint) * (size_t)n);
    UP(A.desc, const uint4*, desc, 32 * (size_t)n);
    if (uRight) UP(A.uRight, const float*, uRight, 4 * (size_t)n);
    UP(A.mp_desc, const uint4*, mp_desc, 32 * (size_t)nmp);
    UP(A.projx, const float*, mp_projx, 4 * (size_t)nmp);
    UP(A.projy, const float*, mp_projy, 4 * (size_t)nmp);
    if (mp_projxr) UP(A.projxr, const float*, mp_projxr, 4 * (size_t)nmp);
    UP(A.level, const int*, mp_level, 4 * (size_t)nmp);
    UP(A.viewcos, const float*, mp_viewcos, 4 * (size_t)nmp);
    if (mp_depth) UP(A.depth, const float*, mp_depth, 4 * (size_t)nmp);
    UP(A.inview, const uint8_t*, mp_inview, (size_t)nmp);
    UP(A.assign, int*, assign, 4 * (size_t)n);
    UP(A.cell_start, int*, nullptr, 4 * (size_t)(GRID_CELLS + 1));
    UP(A.cell_idx, int*, nullptr, 4 * (size_t)n);
    UP(A.cand_cnt, int*, nullptr, 4 * (size_t)nmp);
    UP(A.cand, int*, nullptr, 8 * (size_t)nmp * PROJ_KMAX);
    UP(A.claim0, int*, nullptr, 4 * (size_t)n);
    UP(A.claim1, int*, nullptr, 4 * (size_t)n);
    UP(A.choice, int*, nullptr, 4 * (size_t)nmp);
    UP(A.result, int*, result, 8);
#undef UP
    A.n = n; A.nmp = nmp;
    A.minX = minX; A.minY = minY;
    A.winv = (float)GRID_COLS / (float)(maxX - minX);                     // src/Frame.cc:292
    A.hinv = (float)GRID_ROWS / (float)(maxY - minY);                     // :294
    for (int l = 0; l < nlevels; l++) A.sf[l] = scaleFactors[l];
    A.th = th; A.nnratio = nnratio; A.bFar = bFarPoints; A.thFar = thFarPoints;
    {
        GridArgs g;
        g.kps = A.kps; g.n = n; g.minX = A.minX; g.minY = A.minY; g.winv = A.winv; g.hinv = A.hinv;
        g.cell_start = A.cell_start; g.cell_idx = A.cell_idx;
        orbx_launch_grid_build(g, s);
    }
    proj_candidates_kernel<<<(nmp + PROJ_WARPS - 1) / PROJ_WARPS, PROJ_WARPS * 32, 0, s>>>(A);
    proj_resolve_kernel<<<1, 1024, 0, s>>>(A);
    CUDA_TRYM(cudaGetLastError());
    CUDA_TRYM(cudaMemcpyAsync(assign, A.assign, 4 * (size_t)n, cudaMemcpyDeviceToHost, s));
    CUDA_TRYM(cudaMemcpyAsync(result, A.result, 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRYM(cudaStreamSynchronize(s));
    if (result[1]) { orbx_set_error("a map point has more than %d candidates", PROJ_KMAX); rc = ORBX_ERR_CAPACITY; }
    else rc = result[0];
done:
    for (void* p : allocs) cudaFree(p);
    if (s) cudaStreamDestroy(s);
    return rc;
}
