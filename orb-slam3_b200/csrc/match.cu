// match.cu — Hamming matchers: brute-force 2-NN (cv::BFMatcher(NORM_HAMMING).knnMatch k=2 as used by
// Frame::ComputeStereoFishEyeMatches, reference src/Frame.cc:1440-1480) and
// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) (src/ORBmatcher.cc:45-168) over the
// 64x48 frame grid (Frame::AssignFeaturesToGrid / GetFeaturesInArea, src/Frame.cc:446-480, 784-899).
// Integer pipe work (XOR + POPC); no tensor cores: nothing here is a dense contraction.
#include "match_internal.h"
#include <climits>
#include <cstring>
#include <mutex>
#include <vector>

#define CUDA_TRYM(expr)                                                                        \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            orbx_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            rc = ORBX_ERR_CUDA; goto done;                                                     \
        }                                                                                      \
    } while (0)

// ---- brute-force 2-NN ----
// Grid = (query tiles of 128, train splits): one thread per query, the split's train rows broadcast from
// shared memory; every (split, query) leaves its partial best / second best, knn2_merge_kernel folds the
// splits in ascending train order (strict '<' keeps the lower train index on ties, as BFMatcher does).
// The split count is chosen so that the grid gives every SM about 8 CTAs when the train set allows (>= 64 rows per split).
#define KNN_THREADS 128
#define KNN_TILE 128

__global__ void __launch_bounds__(KNN_THREADS)
knn2_partial_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t, int nt, int per_split, int4* __restrict__ part)
{
    __shared__ uint4 s_t[KNN_TILE * 2];
    const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
    const int jb = blockIdx.y * per_split, je = min(nt, jb + per_split);
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (i < nq) { a0 = q[2 * i]; a1 = q[2 * i + 1]; }
    int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
    for (int j0 = jb; j0 < je; j0 += KNN_TILE) {
        const int m = min(KNN_TILE, je - j0);
        __syncthreads();
        for (int k = threadIdx.x; k < 2 * m; k += KNN_THREADS) s_t[k] = t[2 * j0 + k];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < m; j++) {
            const int d = ham256_csa(a0, a1, s_t[2 * j], s_t[2 * j + 1]);
            if (d < b0) { b1 = b0; i1 = i0; b0 = d; i0 = j0 + j; }     // ties keep the lower train index
            else if (d < b1) { b1 = d; i1 = j0 + j; }
        }
    }
    if (i < nq) part[(size_t)blockIdx.y * nq + i] = make_int4(b0, i0, b1, i1);
}

__global__ void __launch_bounds__(KNN_THREADS)
knn2_merge_kernel(const int4* __restrict__ part, int nq, int nsplit, int* __restrict__ out_idx, int* __restrict__ out_dist)
{
    const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
    if (i >= nq) return;
    int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
    for (int s = 0; s < nsplit; s++) {
        const int4 p = part[(size_t)s * nq + i];
        // a split's pair is in (distance, index) order; splits ascend in index
        if (p.y >= 0) {
            if (p.x < b0) { b1 = b0; i1 = i0; b0 = p.x; i0 = p.y; }
            else if (p.x < b1) { b1 = p.x; i1 = p.y; }
        }
        if (p.w >= 0) {
            if (p.z < b0) { b1 = b0; i1 = i0; b0 = p.z; i0 = p.w; }
            else if (p.z < b1) { b1 = p.z; i1 = p.w; }
        }
    }
    out_idx[2 * i] = i0; out_idx[2 * i + 1] = i1;
    out_dist[2 * i] = i0 >= 0 ? b0 : -1; out_dist[2 * i + 1] = i1 >= 0 ? b1 : -1;
}

extern "C" int orbx_knn2_hamming(int device, const uint8_t* q, int nq, const uint8_t* t, int nt,
                                 int32_t* out_idx, int32_t* out_dist)
{
    if (nq < 0 || nt < 0 || (nq > 0 && (!q || !out_idx || !out_dist)) || (nt > 0 && !t)) { orbx_set_error("bad knn2 arguments"); return ORBX_ERR_ARG; }
    if (nq == 0) return ORBX_OK;
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    const int tiles = (nq + KNN_THREADS - 1) / KNN_THREADS;
    int nsplit = (8 * 148 + tiles - 1) / tiles;                       // ~8 CTAs of 128 threads per SM
    const int max_split = nt > 0 ? (nt + 63) / 64 : 1;                // at least 64 train rows per split
    if (nsplit > max_split) nsplit = max_split;
    if (nsplit < 1) nsplit = 1;
    const int per_split = nt > 0 ? (nt + nsplit - 1) / nsplit : 1;
    nsplit = nt > 0 ? (nt + per_split - 1) / per_split : 1;
    OrbxCarver c;
    const size_t o_q = c.take((size_t)nq * 32), o_t = c.take((size_t)nt * 32);
    const size_t in_bytes = c.off;
    const size_t o_idx = c.take((size_t)nq * 8), o_dist = c.take((size_t)nq * 8);
    const size_t out_end = c.off;
    const size_t o_part = c.take((size_t)nsplit * nq * sizeof(int4));
    rc = W.reserve(c.off, out_end);
    if (rc != ORBX_OK) return rc;
    memcpy(W.h + o_q, q, (size_t)nq * 32);
    if (nt > 0) memcpy(W.h + o_t, t, (size_t)nt * 32);
    CUDA_TRYW(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.s));
    knn2_partial_kernel<<<dim3(tiles, nsplit), KNN_THREADS, 0, W.s>>>((const uint4*)(W.d + o_q), nq, (const uint4*)(W.d + o_t), nt, per_split,
                                                                     (int4*)(W.d + o_part));
    knn2_merge_kernel<<<tiles, KNN_THREADS, 0, W.s>>>((const int4*)(W.d + o_part), nq, nsplit, (int*)(W.d + o_idx), (int*)(W.d + o_dist));
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(W.h + o_idx, W.d + o_idx, out_end - o_idx, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    memcpy(out_idx, W.h + o_idx, (size_t)nq * 8);
    memcpy(out_dist, W.h + o_dist, (size_t)nq * 8);
    return ORBX_OK;
}

// ---- Frame::ComputeStereoFishEyeMatches on the descriptors the two extractions left on the device ----
// src/Frame.cc:1440-1480 up to the triangulation: rows [monoLeft, nL) x [monoRight, nR) of the two handles' resident
// descriptor matrices (:1445-1446), knnMatch k = 2 (:1458), Lowe ratio (:1465; float distance * DOUBLE 0.7).
__global__ void __launch_bounds__(KNN_THREADS)
fisheye_ratio_kernel(const int4* __restrict__ part, int nq, int nsplit, int monoLeft, int monoRight, int* __restrict__ l2r, int* __restrict__ r2l,
                     int* __restrict__ dist_out, int* __restrict__ good)
{
    const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
    int ok = 0;
    if (i < nq) {
        int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
        for (int s = 0; s < nsplit; s++) {                             // as knn2_merge_kernel
            const int4 p = part[(size_t)s * nq + i];
            if (p.y >= 0) { if (p.x < b0) { b1 = b0; i1 = i0; b0 = p.x; i0 = p.y; } else if (p.x < b1) { b1 = p.x; i1 = p.y; } }
            if (p.w >= 0) { if (p.z < b0) { b1 = b0; i1 = i0; b0 = p.z; i0 = p.w; } else if (p.z < b1) { b1 = p.z; i1 = p.w; } }
        }
        if (i0 >= 0 && i1 >= 0 && (double)b0 < __dmul_rn((double)b1, 0.7)) {   // (*it).size() >= 2 && d0 < d1 * 0.7
            l2r[i + monoLeft] = i0 + monoRight;
            dist_out[i + monoLeft] = b0;
            atomicMax(&r2l[i0 + monoRight], i + monoLeft);              // the loop at :1463 runs in query order: the last writer wins
            ok = 1;
        }
    }
    const int n = __syncthreads_count(ok);
    if (threadIdx.x == 0 && n) atomicAdd(good, n);
}

__global__ void fill_minus1_kernel(int* p, int n) { const int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = -1; }

extern "C" int orbx_stereo_fisheye_match(orbx_extractor* eL, orbx_extractor* eR, int frame, int nL, int monoLeft, int nR, int monoRight,
                                         int32_t* l2r, int32_t* r2l, int32_t* best_dist)
{
    if (!eL || !eR || !eL->reserved || !eR->reserved || eL->last_nframes <= 0 || eL->last_nframes != eR->last_nframes) {
        orbx_set_error("fisheye stereo match needs a preceding extract of the same number of frames on both handles"); return ORBX_ERR_STATE;
    }
    if (eL->device != eR->device) { orbx_set_error("left/right extractors live on different devices"); return ORBX_ERR_ARG; }
    if (frame < 0 || frame >= eL->last_nframes || nL < 0 || nR < 0 || monoLeft < 0 || monoRight < 0 || monoLeft > nL || monoRight > nR ||
        nL > eL->G.out_cap || nR > eR->G.out_cap || (nL > 0 && !l2r) || (nR > 0 && !r2l)) {
        orbx_set_error("bad fisheye stereo arguments (frame %d, nL %d, monoLeft %d, nR %d, monoRight %d)", frame, nL, monoLeft, nR, monoRight); return ORBX_ERR_ARG;
    }
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(eL->device);
    if (rc != ORBX_OK) return rc;
    const int nq = nL - monoLeft, nt = nR - monoRight;
    for (int i = 0; i < nL; i++) { l2r[i] = -1; if (best_dist) best_dist[i] = -1; }
    for (int i = 0; i < nR; i++) r2l[i] = -1;
    if (nq <= 0 || nt <= 0) return 0;
    const int tiles = (nq + KNN_THREADS - 1) / KNN_THREADS;
    int nsplit = (8 * 148 + tiles - 1) / tiles;
    const int max_split = (nt + 63) / 64;
    if (nsplit > max_split) nsplit = max_split;
    if (nsplit < 1) nsplit = 1;
    const int per_split = (nt + nsplit - 1) / nsplit;
    nsplit = (nt + per_split - 1) / per_split;
    OrbxCarver c;
    const size_t o_l2r = c.take((size_t)nL * 4), o_dist = c.take((size_t)nL * 4), o_r2l = c.take((size_t)nR * 4), o_good = c.take(4);
    const size_t out_end = c.off;
    const size_t o_part = c.take((size_t)nsplit * nq * sizeof(int4));
    rc = W.reserve(c.off, out_end);
    if (rc != ORBX_OK) return rc;
    // order behind both extractions (each handle has its own stream)
    CUDA_TRYW(cudaEventRecord(eL->join_ev, eL->stream)); CUDA_TRYW(cudaStreamWaitEvent(W.s, eL->join_ev, 0));
    CUDA_TRYW(cudaEventRecord(eR->join_ev, eR->stream)); CUDA_TRYW(cudaStreamWaitEvent(W.s, eR->join_ev, 0));
    const int nfill = (int)(out_end / 4);
    fill_minus1_kernel<<<(nfill + 255) / 256, 256, 0, W.s>>>((int*)W.d, nfill);
    CUDA_TRYW(cudaMemsetAsync(W.d + o_good, 0, 4, W.s));
    const uint8_t* dL = eL->D.desc + ((size_t)frame * eL->G.out_cap + monoLeft) * 32;
    const uint8_t* dR = eR->D.desc + ((size_t)frame * eR->G.out_cap + monoRight) * 32;
    knn2_partial_kernel<<<dim3(tiles, nsplit), KNN_THREADS, 0, W.s>>>((const uint4*)dL, nq, (const uint4*)dR, nt, per_split, (int4*)(W.d + o_part));
    fisheye_ratio_kernel<<<tiles, KNN_THREADS, 0, W.s>>>((const int4*)(W.d + o_part), nq, nsplit, monoLeft, monoRight,
                                                          (int*)(W.d + o_l2r), (int*)(W.d + o_r2l), (int*)(W.d + o_dist), (int*)(W.d + o_good));
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(W.h, W.d, out_end, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    memcpy(l2r, W.h + o_l2r, (size_t)nL * 4);
    memcpy(r2l, W.h + o_r2l, (size_t)nR * 4);
    if (best_dist) memcpy(best_dist, W.h + o_dist, (size_t)nL * 4);
    return *(const int*)(W.h + o_good);
}

// ---- brute-force 2-NN without any allocation (no counterpart in the reference) ----
// Same two kernels as orbx_knn2_hamming, but the (split, query) partial table lives in a file-scope device array
// instead of the grow-only workspace arena, and descriptors / results are the CALLER's device buffers: the entry
// point calls no cudaMalloc / cudaFree / cudaMallocHost and creates no stream.  For callers that cannot allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid it.
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
// zero: knn2_merge_kernel reads exactly the nsplit x nq entries that the knn2_partial_kernel launch in front of it
// (same stream) has written in this call.  Calls that overlap on the GPU (two streams, two host threads) RACE on it.
#define KNN2_NOALLOC_MAX_QUERIES 16384                       // = ORBX_KNN2_NOALLOC_MAX_QUERIES (include/orbx.h)
#define KNN2_NOALLOC_PARTIALS (16 * KNN2_NOALLOC_MAX_QUERIES) // int4 entries: 4 MiB per GPU that loads the library
static __device__ __align__(256) int4 g_knn2_noalloc_scratch[KNN2_NOALLOC_PARTIALS];

#define KNN2_NOALLOC_MAX_DEVICES 64
static int4* g_knn2_noalloc_ptr[KNN2_NOALLOC_MAX_DEVICES];   // device address of the array, per device (host code
static std::mutex g_knn2_noalloc_mutex;                      // must not pass the symbol itself to a kernel)

// set-up, once per device, with that device current
static int knn2_noalloc_scratch(int device, int4** out)
{
    std::lock_guard<std::mutex> lock(g_knn2_noalloc_mutex);
    if (!g_knn2_noalloc_ptr[device]) {
        int4* p = nullptr;
        CUDA_TRYW(cudaGetSymbolAddress((void**)&p, g_knn2_noalloc_scratch));
        if (!p) { orbx_set_error("cudaGetSymbolAddress returned a null scratch address on device %d", device); return ORBX_ERR_CUDA; }
        g_knn2_noalloc_ptr[device] = p;
    }
    *out = g_knn2_noalloc_ptr[device];
    return ORBX_OK;
}

extern "C" int orbx_knn2_hamming_noalloc(int device, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt,
                                         int32_t* d_out_idx, int32_t* d_out_dist, void* stream)
{
    if (nq < 0 || nt < 0 || (nq > 0 && (!d_q || !d_out_idx || !d_out_dist)) || (nt > 0 && !d_t)) { orbx_set_error("bad knn2 arguments"); return ORBX_ERR_ARG; }
    if (nq > KNN2_NOALLOC_MAX_QUERIES) {
        orbx_set_error("orbx_knn2_hamming_noalloc: %d query descriptors exceed the fixed scratch (max %d); use orbx_knn2_hamming", nq, KNN2_NOALLOC_MAX_QUERIES);
        return ORBX_ERR_CAPACITY;
    }
    if (((uintptr_t)d_q | (uintptr_t)d_t) & 15) { orbx_set_error("orbx_knn2_hamming_noalloc: descriptor arrays must be 16-byte aligned"); return ORBX_ERR_ARG; }
    if (device < 0 || device >= orbx_device_count() || device >= KNN2_NOALLOC_MAX_DEVICES) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    CUDA_TRYW(cudaSetDevice(device));
    int4* part = nullptr;
    const int rc = knn2_noalloc_scratch(device, &part);
    if (rc != ORBX_OK) return rc;
    if (nq == 0) return ORBX_OK;                                      // set-up only (call once outside a capture)
    const int tiles = (nq + KNN_THREADS - 1) / KNN_THREADS;
    int nsplit = (8 * 148 + tiles - 1) / tiles;                       // same split rule as orbx_knn2_hamming ...
    const int max_split = nt > 0 ? (nt + 63) / 64 : 1;
    if (nsplit > max_split) nsplit = max_split;
    if (nsplit > KNN2_NOALLOC_PARTIALS / nq) nsplit = KNN2_NOALLOC_PARTIALS / nq;   // ... capped by the scratch (>= 16)
    if (nsplit < 1) nsplit = 1;
    const int per_split = nt > 0 ? (nt + nsplit - 1) / nsplit : 1;
    nsplit = nt > 0 ? (nt + per_split - 1) / per_split : 1;
    cudaStream_t s = (cudaStream_t)stream;
    knn2_partial_kernel<<<dim3(tiles, nsplit), KNN_THREADS, 0, s>>>((const uint4*)d_q, nq, (const uint4*)d_t, nt, per_split, part);
    knn2_merge_kernel<<<tiles, KNN_THREADS, 0, s>>>(part, nq, nsplit, d_out_idx, d_out_dist);
    CUDA_TRYW(cudaGetLastError());
    return ORBX_OK;
}

// ---- frame grid + SearchByProjection ----
#define PROJ_KMAX0 256            // first capacity of a map point's candidate list; grown on demand (t_proj_kmax)
static thread_local int t_proj_kmax = PROJ_KMAX0;
#define ORBX_RETRY_KMAX (-1000)   // private: a candidate list overflowed, t_proj_kmax was raised, run again

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
    int* cand;            // nmp * kmax * 2 : (idx | octave << 24, dist)
    int kmax;             // capacity of one list
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
    if (live && A.level[iMP] < 0) live = false;                            // mnTrackScaleLevelR == -1, :172
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
        int* out = A.cand + (long long)iMP * A.kmax * 2;
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
                        if (p < A.kmax) { out[2 * p] = idx | (oct << 24); out[2 * p + 1] = dist; }
                    }
                    cnt += __popc(bal);
                }
            }
    }
    if (lane == 0) {
        if (cnt > A.kmax) { atomicMax(&A.result[1], cnt); cnt = A.kmax; }      // the host re-runs with room for the longest list
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
            const int* c = A.cand + (long long)i * A.kmax * 2;
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

static int orbx_search_by_projection_once(int device,
                                         const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                                         float minX, float maxX, float minY, float maxY,
                                         const float* scaleFactors, int nlevels,
                                         int nmp, const uint8_t* mp_desc, const float* mp_projx, const float* mp_projy,
                                         const float* mp_projxr, const int32_t* mp_level, const float* mp_viewcos,
                                         const float* mp_depth, const uint8_t* mp_inview,
                                         float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    const int kmax = t_proj_kmax;
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
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    ProjArgs A;
    memset(&A, 0, sizeof(A));
    const size_t N = (size_t)n, M = (size_t)nmp;
    // inputs (one H2D copy), then the small outputs (one D2H copy), then device-only scratch
    OrbxCarver c;
    const size_t o_kps = c.take(N * sizeof(orbx_keypoint)), o_desc = c.take(N * 32), o_ur = c.take(uRight ? N * 4 : 0);
    const size_t o_md = c.take(M * 32), o_px = c.take(M * 4), o_py = c.take(M * 4), o_pxr = c.take(mp_projxr ? M * 4 : 0);
    const size_t o_lv = c.take(M * 4), o_vc = c.take(M * 4), o_dp = c.take(mp_depth ? M * 4 : 0), o_iv = c.take(M);
    const size_t o_assign = c.take(N * 4), o_result = c.take(8);
    const size_t in_bytes = c.off;
    const size_t o_cs = c.take(4 * (size_t)(GRID_CELLS + 1)), o_ci = c.take(N * 4), o_cc = c.take(M * 4);
    const size_t o_cand = c.take(8 * M * (size_t)kmax), o_c0 = c.take(N * 4), o_c1 = c.take(N * 4), o_ch = c.take(M * 4);
    rc = W.reserve(c.off, in_bytes);
    if (rc != ORBX_OK) return rc;
    char* h = W.h; char* d = W.d;
    memcpy(h + o_kps, kps, N * sizeof(orbx_keypoint)); memcpy(h + o_desc, desc, N * 32);
    if (uRight) memcpy(h + o_ur, uRight, N * 4);
    memcpy(h + o_md, mp_desc, M * 32); memcpy(h + o_px, mp_projx, M * 4); memcpy(h + o_py, mp_projy, M * 4);
    if (mp_projxr) memcpy(h + o_pxr, mp_projxr, M * 4);
    memcpy(h + o_lv, mp_level, M * 4); memcpy(h + o_vc, mp_viewcos, M * 4);
    if (mp_depth) memcpy(h + o_dp, mp_depth, M * 4);
    memcpy(h + o_iv, mp_inview, M); memcpy(h + o_assign, assign, N * 4);
    memset(h + o_result, 0, 8);
    CUDA_TRYW(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, W.s));
    A.kps = (const orbx_keypoint*)(d + o_kps); A.desc = (const uint4*)(d + o_desc);
    A.uRight = uRight ? (const float*)(d + o_ur) : nullptr;
    A.mp_desc = (const uint4*)(d + o_md); A.projx = (const float*)(d + o_px); A.projy = (const float*)(d + o_py);
    A.projxr = mp_projxr ? (const float*)(d + o_pxr) : nullptr;
    A.level = (const int*)(d + o_lv); A.viewcos = (const float*)(d + o_vc);
    A.depth = mp_depth ? (const float*)(d + o_dp) : nullptr; A.inview = (const uint8_t*)(d + o_iv);
    A.assign = (int*)(d + o_assign); A.result = (int*)(d + o_result); A.kmax = kmax;
    A.cell_start = (int*)(d + o_cs); A.cell_idx = (int*)(d + o_ci); A.cand_cnt = (int*)(d + o_cc); A.cand = (int*)(d + o_cand);
    A.claim0 = (int*)(d + o_c0); A.claim1 = (int*)(d + o_c1); A.choice = (int*)(d + o_ch);
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
        orbx_launch_grid_build(g, W.s);
    }
    proj_candidates_kernel<<<(nmp + PROJ_WARPS - 1) / PROJ_WARPS, PROJ_WARPS * 32, 0, W.s>>>(A);
    proj_resolve_kernel<<<1, 1024, 0, W.s>>>(A);
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(h + o_assign, d + o_assign, in_bytes - o_assign, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    const int* result = (const int*)(h + o_result);
    if (result[1] > kmax) { t_proj_kmax = (result[1] + 63) / 64 * 64; return ORBX_RETRY_KMAX; }   // nothing of the caller's was touched yet
    memcpy(assign, h + o_assign, N * 4);
    return result[0];
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
    // a map point whose window holds more keypoints than the candidate lists have room for makes the pass report the
    // longest list; the lists are regrown (per calling thread, kept) and the pass repeated - the reference has no such limit
    for (int attempt = 0; attempt < 4; attempt++) {
        const int r = orbx_search_by_projection_once(device, kps, desc, n, uRight, minX, maxX, minY, maxY, scaleFactors, nlevels, nmp, mp_desc, mp_projx, mp_projy, mp_projxr, mp_level, mp_viewcos, mp_depth, mp_inview, th, nnratio, bFarPoints, thFarPoints, assign);
        if (r != ORBX_RETRY_KMAX) return r;
    }
    orbx_set_error("candidate lists kept overflowing");
    return ORBX_ERR_CAPACITY;
}


// ---- Tracking::SearchLocalPoints (src/Tracking.cc:3808-3893) in one call: Frame::isInFrustum for every local map
// point on the device (src/Frame.cc:603-711, Nleft == -1), its mTrack* fields feeding SearchByProjection directly ----

// glibc's logf (sysdeps/ieee754/flt-32/e_logf.c: 16-entry table, degree-3 polynomial in double) restated; checked
// against this image's libm over all 2 139 095 039 positive finite floats: 0 mismatches, with and without FMA
// contraction (tools/scratch/logf_glibc_check.c).  MapPoint::PredictScale needs it bit for bit: the level is
// ceil(logf(ratio) / logScaleFactor).
__constant__ double c_logf_tab[32] = {
    0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2,
    0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2, 0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3,
    0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3,
    0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4,
    0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5, 0x1p+0, 0x0p+0,
    0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5, 0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4,
    0x1.b2036576afce6p-1, 0x1.526e57720db08p-3, 0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3,
    0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2, 0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2};

__device__ float glibc_logf(float x)
{
    uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return NAN;
        ix = __float_as_uint(__fmul_rn(x, 8388608.0f)); ix -= 23u << 23;       // subnormal
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (tmp >> 19) & 15;
    const int k = (int)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = c_logf_tab[2 * i], logc = c_logf_tab[2 * i + 1];
    const double z = (double)__uint_as_float(iz);
    const double r = __dadd_rn(__dmul_rn(z, invc), -1.0);
    const double y0 = __dadd_rn(logc, __dmul_rn((double)k, 0x1.62e42fefa39efp-1));
    const double r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(0x1.5575b0be00b6ap-2, r), -0x1.ffffef20a4123p-2);
    y = __dadd_rn(__dmul_rn(-0x1.00ea348b88334p-2, r2), y);
    y = __dadd_rn(__dmul_rn(y, r2), __dadd_rn(y0, r));
    return __double2float_rn(y);
}

struct FrustumArgs {
    float R[9], t[3], Ow[3];
    float fx, fy, cx, cy, mbf;
    float minX, maxX, minY, maxY, logsf; int nlevels;
    int nmp; const float* pos; const float* normal; const float* dmin; const float* dmax; const uint8_t* candidate;
    float cosLimit;
    uint8_t* inview; float *projx, *projy, *projxr, *depth, *viewcos; int* level;
};

__global__ void __launch_bounds__(256)
frustum_kernel(FrustumArgs A)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= A.nmp) return;
    uint8_t in = 0; float px = -1.f, py = -1.f, pxr = 0.f, dep = 0.f, vcos = 0.f; int lvl = 0;        // :607-609
    if (!A.candidate || A.candidate[i]) {
        const float P0 = A.pos[3 * i], P1 = A.pos[3 * i + 1], P2 = A.pos[3 * i + 2];
        float Pc[3];
#pragma unroll
        for (int r = 0; r < 3; r++)                                                                     // cv::Matx33f * Matx31f + t, :615
            Pc[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(A.R[3 * r], P0), __fmul_rn(A.R[3 * r + 1], P1)), __fmul_rn(A.R[3 * r + 2], P2)), A.t[r]);
        const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(Pc[0], Pc[0]), __dmul_rn(Pc[1], Pc[1])), __dmul_rn(Pc[2], Pc[2]));
        const float Pc_dist = __double2float_rn(__dsqrt_rn(d2));                                        // cv::norm, :616
        const float z = Pc[2];
        do {
            if (z < 0.0f) break;                                                                        // :621
            const float invz = __fdiv_rn(1.0f, z);
            const float u = __fadd_rn(__fdiv_rn(__fmul_rn(A.fx, Pc[0]), z), A.cx);                      // Pinhole::project
            const float v = __fadd_rn(__fdiv_rn(__fmul_rn(A.fy, Pc[1]), z), A.cy);
            if (u < A.minX || u > A.maxX) break;
            if (v < A.minY || v > A.maxY) break;
            px = u; py = v;                                                                             // :631-632
            const float O0 = __fsub_rn(P0, A.Ow[0]), O1 = __fsub_rn(P1, A.Ow[1]), O2 = __fsub_rn(P2, A.Ow[2]);
            const double p2 = __dadd_rn(__dadd_rn(__dmul_rn(O0, O0), __dmul_rn(O1, O1)), __dmul_rn(O2, O2));
            const float dist = __double2float_rn(__dsqrt_rn(p2));
            if (dist < __fmul_rn(0.8f, A.dmin[i]) || dist > __fmul_rn(1.2f, A.dmax[i])) break;          // :647-656, src/MapPoint.cc:668-681
            const float dot = __fadd_rn(__fadd_rn(__fmul_rn(O0, A.normal[3 * i]), __fmul_rn(O1, A.normal[3 * i + 1])), __fmul_rn(O2, A.normal[3 * i + 2]));
            const float vc = __fdiv_rn(dot, dist);                                                      // :646
            if (vc < A.cosLimit) break;
            const float ratio = __fdiv_rn(A.dmax[i], dist);                                             // MapPoint::PredictScale: the RAW mfMaxDistance
            int nScale = (int)ceilf(__fdiv_rn(glibc_logf(ratio), A.logsf));
            if (nScale < 0) nScale = 0; else if (nScale >= A.nlevels) nScale = A.nlevels - 1;
            in = 1; pxr = __fsub_rn(u, __fmul_rn(A.mbf, invz)); dep = Pc_dist; lvl = nScale; vcos = vc; // :654-662
        } while (0);
    }
    A.inview[i] = in; A.projx[i] = px; A.projy[i] = py; A.projxr[i] = pxr; A.depth[i] = dep; A.viewcos[i] = vcos; A.level[i] = lvl;
}

static int orbx_search_local_points_once(int device,
                                        const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                                        float minX, float maxX, float minY, float maxY,
                                        const float* scaleFactors, int nlevels,
                                        const float* Rcw, const float* tcw, const float* Ow,
                                        float fx, float fy, float cx, float cy, float mbf, float logScaleFactor,
                                        int nmp, const float* mp_pos, const float* mp_normal, const float* mp_min_dist,
                                        const float* mp_max_dist, const uint8_t* mp_desc, const uint8_t* mp_candidate,
                                        float viewingCosLimit, float th, float nnratio, int bFarPoints, float thFarPoints,
                                        int32_t* assign,
                                        uint8_t* out_inview, float* out_projx, float* out_projy, float* out_projxr,
                                        float* out_depth, float* out_viewcos, int32_t* out_level)
{
    const int kmax = t_proj_kmax;
    if (n < 0 || nmp < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors || !(maxX > minX) || !(maxY > minY) ||
        !Rcw || !tcw || !Ow || (n > 0 && (!kps || !desc || !assign)) ||
        (nmp > 0 && (!mp_pos || !mp_normal || !mp_min_dist || !mp_max_dist || (n > 0 && !mp_desc)))) {
        orbx_set_error("bad search_local_points arguments"); return ORBX_ERR_ARG;
    }
    if (n >= (1 << 24)) { orbx_set_error("too many keypoints"); return ORBX_ERR_ARG; }
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    if (nmp == 0) return 0;
    const size_t N = (size_t)n, M = (size_t)nmp;
    const bool search = n > 0;
    OrbxCarver c;
    // inputs
    const size_t o_kps = c.take(N * sizeof(orbx_keypoint)), o_desc = c.take(N * 32), o_ur = c.take(uRight ? N * 4 : 0);
    const size_t o_pos = c.take(M * 12), o_nrm = c.take(M * 12), o_dmin = c.take(M * 4), o_dmax = c.take(M * 4);
    const size_t o_md = c.take(search ? M * 32 : 0), o_cand = c.take(mp_candidate ? M : 0);
    const size_t in_bytes = c.off;
    // outputs (one D2H copy): assign, result, then the mTrack* fields
    const size_t o_assign = c.take(N * 4 + 4), o_result = c.take(8);
    const size_t o_iv = c.take(M), o_px = c.take(M * 4), o_py = c.take(M * 4), o_pxr = c.take(M * 4), o_dp = c.take(M * 4), o_vc = c.take(M * 4), o_lv = c.take(M * 4);
    const size_t out_end = c.off;
    const size_t o_cs = c.take(4 * (size_t)(GRID_CELLS + 1)), o_ci = c.take(N * 4 + 4), o_cc = c.take(M * 4);
    const size_t o_cl = c.take(search ? 8 * M * (size_t)kmax : 0), o_c0 = c.take(N * 4 + 4), o_c1 = c.take(N * 4 + 4), o_ch = c.take(M * 4);
    rc = W.reserve(c.off, out_end);
    if (rc != ORBX_OK) return rc;
    char* h = W.h; char* d = W.d;
    if (search) { memcpy(h + o_kps, kps, N * sizeof(orbx_keypoint)); memcpy(h + o_desc, desc, N * 32); memcpy(h + o_md, mp_desc, M * 32); }
    if (uRight && search) memcpy(h + o_ur, uRight, N * 4);
    memcpy(h + o_pos, mp_pos, M * 12); memcpy(h + o_nrm, mp_normal, M * 12);
    memcpy(h + o_dmin, mp_min_dist, M * 4); memcpy(h + o_dmax, mp_max_dist, M * 4);
    if (mp_candidate) memcpy(h + o_cand, mp_candidate, M);
    CUDA_TRYW(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, W.s));
    if (search) {
        memcpy(h + o_assign, assign, N * 4); memset(h + o_result, 0, 8);
        CUDA_TRYW(cudaMemcpyAsync(d + o_assign, h + o_assign, o_iv - o_assign, cudaMemcpyHostToDevice, W.s));
    }
    FrustumArgs F;
    for (int i = 0; i < 9; i++) F.R[i] = Rcw[i];
    for (int i = 0; i < 3; i++) { F.t[i] = tcw[i]; F.Ow[i] = Ow[i]; }
    F.fx = fx; F.fy = fy; F.cx = cx; F.cy = cy; F.mbf = mbf;
    F.minX = minX; F.maxX = maxX; F.minY = minY; F.maxY = maxY; F.logsf = logScaleFactor; F.nlevels = nlevels;
    F.nmp = nmp; F.pos = (const float*)(d + o_pos); F.normal = (const float*)(d + o_nrm);
    F.dmin = (const float*)(d + o_dmin); F.dmax = (const float*)(d + o_dmax);
    F.candidate = mp_candidate ? (const uint8_t*)(d + o_cand) : nullptr;
    F.cosLimit = viewingCosLimit;
    F.inview = (uint8_t*)(d + o_iv); F.projx = (float*)(d + o_px); F.projy = (float*)(d + o_py); F.projxr = (float*)(d + o_pxr);
    F.depth = (float*)(d + o_dp); F.viewcos = (float*)(d + o_vc); F.level = (int*)(d + o_lv);
    frustum_kernel<<<(nmp + 255) / 256, 256, 0, W.s>>>(F);
    if (search) {
        ProjArgs A;
        memset(&A, 0, sizeof(A));
        A.kps = (const orbx_keypoint*)(d + o_kps); A.desc = (const uint4*)(d + o_desc);
        A.uRight = uRight ? (const float*)(d + o_ur) : nullptr;
        A.mp_desc = (const uint4*)(d + o_md); A.projx = F.projx; A.projy = F.projy; A.projxr = F.projxr;
        A.level = F.level; A.viewcos = F.viewcos; A.depth = F.depth; A.inview = F.inview;
        A.assign = (int*)(d + o_assign); A.result = (int*)(d + o_result); A.kmax = kmax;
        A.cell_start = (int*)(d + o_cs); A.cell_idx = (int*)(d + o_ci); A.cand_cnt = (int*)(d + o_cc); A.cand = (int*)(d + o_cl);
        A.claim0 = (int*)(d + o_c0); A.claim1 = (int*)(d + o_c1); A.choice = (int*)(d + o_ch);
        A.n = n; A.nmp = nmp;
        A.minX = minX; A.minY = minY;
        A.winv = (float)GRID_COLS / (float)(maxX - minX);
        A.hinv = (float)GRID_ROWS / (float)(maxY - minY);
        for (int l = 0; l < nlevels; l++) A.sf[l] = scaleFactors[l];
        A.th = th; A.nnratio = nnratio; A.bFar = bFarPoints; A.thFar = thFarPoints;
        GridArgs g;
        g.kps = A.kps; g.n = n; g.minX = A.minX; g.minY = A.minY; g.winv = A.winv; g.hinv = A.hinv;
        g.cell_start = A.cell_start; g.cell_idx = A.cell_idx;
        orbx_launch_grid_build(g, W.s);
        proj_candidates_kernel<<<(nmp + PROJ_WARPS - 1) / PROJ_WARPS, PROJ_WARPS * 32, 0, W.s>>>(A);
        proj_resolve_kernel<<<1, 1024, 0, W.s>>>(A);
    }
    CUDA_TRYW(cudaGetLastError());
    const size_t d2h0 = search ? o_assign : o_iv;
    CUDA_TRYW(cudaMemcpyAsync(h + d2h0, d + d2h0, out_end - d2h0, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    if (search && ((const int*)(h + o_result))[1] > kmax) { t_proj_kmax = (((const int*)(h + o_result))[1] + 63) / 64 * 64; return ORBX_RETRY_KMAX; }
    if (out_inview) memcpy(out_inview, h + o_iv, M);
    if (out_projx) memcpy(out_projx, h + o_px, M * 4);
    if (out_projy) memcpy(out_projy, h + o_py, M * 4);
    if (out_projxr) memcpy(out_projxr, h + o_pxr, M * 4);
    if (out_depth) memcpy(out_depth, h + o_dp, M * 4);
    if (out_viewcos) memcpy(out_viewcos, h + o_vc, M * 4);
    if (out_level) memcpy(out_level, h + o_lv, M * 4);
    if (!search) return 0;
    const int* result = (const int*)(h + o_result);
    if (result[1] > kmax) { t_proj_kmax = (result[1] + 63) / 64 * 64; return ORBX_RETRY_KMAX; }   // nothing of the caller's was touched yet
    memcpy(assign, h + o_assign, N * 4);
    return result[0];
}

extern "C" int orbx_search_local_points(int device,
                                        const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                                        float minX, float maxX, float minY, float maxY,
                                        const float* scaleFactors, int nlevels,
                                        const float* Rcw, const float* tcw, const float* Ow,
                                        float fx, float fy, float cx, float cy, float mbf, float logScaleFactor,
                                        int nmp, const float* mp_pos, const float* mp_normal, const float* mp_min_dist,
                                        const float* mp_max_dist, const uint8_t* mp_desc, const uint8_t* mp_candidate,
                                        float viewingCosLimit, float th, float nnratio, int bFarPoints, float thFarPoints,
                                        int32_t* assign,
                                        uint8_t* out_inview, float* out_projx, float* out_projy, float* out_projxr,
                                        float* out_depth, float* out_viewcos, int32_t* out_level)
{
    // a map point whose window holds more keypoints than the candidate lists have room for makes the pass report the
    // longest list; the lists are regrown (per calling thread, kept) and the pass repeated - the reference has no such limit
    for (int attempt = 0; attempt < 4; attempt++) {
        const int r = orbx_search_local_points_once(device, kps, desc, n, uRight, minX, maxX, minY, maxY, scaleFactors, nlevels, Rcw, tcw, Ow, fx, fy, cx, cy, mbf, logScaleFactor, nmp, mp_pos, mp_normal, mp_min_dist, mp_max_dist, mp_desc, mp_candidate, viewingCosLimit, th, nnratio, bFarPoints, thFarPoints, assign, out_inview, out_projx, out_projy, out_projxr, out_depth, out_viewcos, out_level);
        if (r != ORBX_RETRY_KMAX) return r;
    }
    orbx_set_error("candidate lists kept overflowing");
    return ORBX_ERR_CAPACITY;
}


// ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:450-531), batched ----
// One warp per map point.  For every row i of its N observation descriptors the lanes compute the N distances
// d(i, .) (XOR + POPC), drop them into a 257-bin histogram in shared memory, and a warp scan finds the element
// floor(0.5*(N-1)) of the sorted row - its median as the reference takes it; the first row with the least median
// wins.  Any N; integer pipe + shared-memory atomics.
#define DD_WARPS 8
#define DD_BINS 288              // 257 used; 9 bins per lane

__global__ void __launch_bounds__(DD_WARPS * 32)
distinctive_kernel(int npoints, const int* __restrict__ start, const uint4* __restrict__ desc, int* __restrict__ best_idx, int* __restrict__ best_median)
{
    __shared__ int s_hist[DD_WARPS][DD_BINS];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int p = blockIdx.x * DD_WARPS + w;
    if (p >= npoints) return;
    const int b = start[p], N = start[p + 1] - b;
    if (N <= 0) { if (lane == 0) { best_idx[p] = -1; best_median[p] = INT_MAX; } return; }
    const int k = (N - 1) >> 1;                                               // (size_t)(0.5*(N-1)), :516
    int* hist = s_hist[w];
    int BestMedian = INT_MAX, BestIdx = 0;
    for (int i = 0; i < N; i++) {
#pragma unroll
        for (int q = 0; q < 9; q++) hist[lane * 9 + q] = 0;
        __syncwarp();
        const uint4 a0 = desc[2 * (size_t)(b + i)], a1 = desc[2 * (size_t)(b + i) + 1];
        for (int j = lane; j < N; j += 32) {
            const int d = j == i ? 0 : ham256(a0, a1, desc[2 * (size_t)(b + j)], desc[2 * (size_t)(b + j) + 1]);
            atomicAdd(&hist[d], 1);
        }
        __syncwarp();
        int c[9], tot = 0;
#pragma unroll
        for (int q = 0; q < 9; q++) { c[q] = hist[lane * 9 + q]; tot += c[q]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        int before = inc - tot, median = -1;
        if (before <= k && k < inc) {
#pragma unroll
            for (int q = 0; q < 9; q++) { if (median < 0 && k < before + c[q]) median = lane * 9 + q; before += c[q]; }
        }
        const unsigned who = __ballot_sync(0xffffffffu, median >= 0);
        median = __shfl_sync(0xffffffffu, median, __ffs(who) - 1);
        if (median < BestMedian) { BestMedian = median; BestIdx = i; }        // :518-522
        __syncwarp();
    }
    if (lane == 0) { best_idx[p] = BestIdx; best_median[p] = BestMedian; }
}

extern "C" int orbx_distinctive_descriptors(int device, int npoints, const int32_t* start, const uint8_t* desc,
                                            int32_t* best_idx, int32_t* best_median)
{
    if (npoints < 0 || (npoints > 0 && (!start || !best_idx))) { orbx_set_error("bad distinctive_descriptors arguments"); return ORBX_ERR_ARG; }
    if (npoints == 0) return ORBX_OK;
    if (start[0] != 0) { orbx_set_error("distinctive_descriptors: start[0] must be 0"); return ORBX_ERR_ARG; }
    for (int p = 0; p < npoints; p++)
        if (start[p + 1] < start[p]) { orbx_set_error("distinctive_descriptors: start must not decrease"); return ORBX_ERR_ARG; }
    const size_t total = (size_t)start[npoints];
    if (total > 0 && !desc) { orbx_set_error("bad distinctive_descriptors arguments"); return ORBX_ERR_ARG; }
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    OrbxCarver c;
    const size_t o_start = c.take(4 * ((size_t)npoints + 1)), o_desc = c.take(total * 32);
    const size_t in_bytes = c.off;
    const size_t o_bi = c.take(4 * (size_t)npoints), o_bm = c.take(4 * (size_t)npoints);
    rc = W.reserve(c.off, c.off);
    if (rc != ORBX_OK) return rc;
    memcpy(W.h + o_start, start, 4 * ((size_t)npoints + 1));
    if (total) memcpy(W.h + o_desc, desc, total * 32);
    CUDA_TRYW(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.s));
    distinctive_kernel<<<(npoints + DD_WARPS - 1) / DD_WARPS, DD_WARPS * 32, 0, W.s>>>(
        npoints, (const int*)(W.d + o_start), (const uint4*)(W.d + o_desc), (int*)(W.d + o_bi), (int*)(W.d + o_bm));
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(W.h + o_bi, W.d + o_bi, c.off - o_bi, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    memcpy(best_idx, W.h + o_bi, 4 * (size_t)npoints);
    if (best_median) memcpy(best_median, W.h + o_bm, 4 * (size_t)npoints);
    return ORBX_OK;
}


// ---- Frame::UndistortKeyPoints (src/Frame.cc:924-960): cv::undistortPoints(mat, mat, K, mDistCoef, cv::Mat(), mK) ----
// OpenCV's cvUndistortPointsInternal restated (double arithmetic, 5 fixed iterations = TermCriteria(MAX_ITER, 5, 0.01),
// R = identity, P = K; k1 k2 p1 p2 k3), non-contracted so that the float results equal cv2's bit for bit
// (oracle/orc_prepare.c is pinned against cv2 4.13 on 2.4 M values).  One thread per keypoint.
struct UndistArgs { int n; const orbx_keypoint* in; orbx_keypoint* out; double fx, fy, cx, cy, k[5]; };

__global__ void __launch_bounds__(256)
undistort_kernel(UndistArgs A)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= A.n) return;
    orbx_keypoint kp = A.in[i];
    const double ifx = __ddiv_rn(1.0, A.fx), ify = __ddiv_rn(1.0, A.fy);
    const double u = (double)kp.x, v = (double)kp.y;
    double x = __dmul_rn(__dsub_rn(u, A.cx), ifx), y = __dmul_rn(__dsub_rn(v, A.cy), ify);
    const double x0 = x, y0 = y;
    const double k0 = A.k[0], k1 = A.k[1], k2 = A.k[2], k3 = A.k[3], k4 = A.k[4];
#pragma unroll 1
    for (int j = 0; j < 5; j++) {
        const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
        // (1 + ((k[7]*r2 + k[6])*r2 + k[5])*r2) with k5..k7 = 0  ==  1 + 0 exactly
        const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(0.0, r2), 0.0), r2), 0.0), r2));
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k4, r2), k1), r2), k0), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) { x = __dmul_rn(__dsub_rn(u, A.cx), ifx); y = __dmul_rn(__dsub_rn(v, A.cy), ify); break; }
        // deltaX = 2*k[2]*x*y + k[3]*(r2 + 2*x*x) + k[8]*r2 + k[9]*r2*r2   (k8..k11 = 0)
        const double dX = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, k2), x), y),
                                                        __dmul_rn(k3, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x)))),
                                              __dmul_rn(0.0, r2)), __dmul_rn(__dmul_rn(0.0, r2), r2));
        const double dY = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(k2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                                        __dmul_rn(__dmul_rn(__dmul_rn(2.0, k3), x), y)),
                                              __dmul_rn(0.0, r2)), __dmul_rn(__dmul_rn(0.0, r2), r2));
        x = __dmul_rn(__dsub_rn(x0, dX), icdist);
        y = __dmul_rn(__dsub_rn(y0, dY), icdist);
    }
    const double xx = __dadd_rn(__dadd_rn(__dmul_rn(A.fx, x), __dmul_rn(0.0, y)), A.cx);
    const double yy = __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(A.fy, y)), A.cy);
    const double ww = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(0.0, y)), 1.0));
    kp.x = __double2float_rn(__dmul_rn(xx, ww));
    kp.y = __double2float_rn(__dmul_rn(yy, ww));
    A.out[i] = kp;
}

extern "C" int orbx_undistort_keypoints(int device, const orbx_keypoint* kps, int n, float fx, float fy, float cx, float cy,
                                        const float* distCoef, int ncoef, orbx_keypoint* kps_un)
{
    if (n < 0 || (n > 0 && (!kps || !kps_un)) || !distCoef || (ncoef != 4 && ncoef != 5)) { orbx_set_error("bad undistort_keypoints arguments"); return ORBX_ERR_ARG; }
    if (n == 0) return ORBX_OK;
    if (distCoef[0] == 0.0f) {                                   // src/Frame.cc:926-931: mvKeysUn = mvKeys
        if (kps_un != kps) memcpy(kps_un, kps, sizeof(orbx_keypoint) * (size_t)n);
        return ORBX_OK;
    }
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    OrbxCarver c;
    const size_t o_in = c.take(sizeof(orbx_keypoint) * (size_t)n), o_out = c.take(sizeof(orbx_keypoint) * (size_t)n);
    rc = W.reserve(c.off, c.off);
    if (rc != ORBX_OK) return rc;
    memcpy(W.h + o_in, kps, sizeof(orbx_keypoint) * (size_t)n);
    CUDA_TRYW(cudaMemcpyAsync(W.d + o_in, W.h + o_in, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyHostToDevice, W.s));
    UndistArgs A;
    A.n = n; A.in = (const orbx_keypoint*)(W.d + o_in); A.out = (orbx_keypoint*)(W.d + o_out);
    A.fx = fx; A.fy = fy; A.cx = cx; A.cy = cy;
    for (int i = 0; i < 5; i++) A.k[i] = i < ncoef ? (double)distCoef[i] : 0.0;
    undistort_kernel<<<(n + 255) / 256, 256, 0, W.s>>>(A);
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(W.h + o_out, W.d + o_out, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    memcpy(kps_un, W.h + o_out, sizeof(orbx_keypoint) * (size_t)n);
    return ORBX_OK;
}


// ---- SearchByProjection(Frame&, vector<MapPoint*>&, ...) for a two-camera rig (F.Nleft != -1), src/ORBmatcher.cc:45-239 ----
// Candidate lists per side by proj_candidates_kernel (left: mvKeys / mGrid, right: mvKeysRight / mGridRight with
// right-local indices; the right radius is not scaled by th, :173).  rig_resolve_kernel replays the sequential
// bookkeeping as the same fixpoint as proj_resolve_kernel over rows (2*iMP: left side, 2*iMP+1: right side): a row
// skips keypoints written by an earlier row; an accepted row writes its keypoint AND overwrites the keypoint's stereo
// partner (mvLeftToRightMatch / mvRightToLeftMatch, :155-159, :222-226); a failed left ratio test skips the right side
// (:148-149).  A keypoint ends up with its LAST writer.
struct RigArgs {
    int nL, nR, nmp;
    const int* candL; const int* cntL; const int* candR; const int* cntR;
    const int* l2r; const int* r2l;
    float nnratio; int kmax;
    int* assign;                 // nL + nR, in/out
    int* claim0; int* claim1;    // first writer row per keypoint (iteration t-1 / t)
    int* last;                   // last writer row per keypoint
    int* choice;                 // 2 per map point
    int* result;
};

__device__ __forceinline__ int rig_pick(const int* c, int nc, const int* prev, int off, int row, int skipIdx, float nnratio, bool* ratioFail)
{
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int k = 0; k < nc; k++) {
        const int idx = c[2 * k] & 0xFFFFFF, oct = c[2 * k] >> 24, dist = c[2 * k + 1];
        if (prev[idx + off] < row || idx == skipIdx) continue;             // already holds a map point
        if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = idx; }
        else if (dist < bestDist2) { bestLevel2 = oct; bestDist2 = dist; }
    }
    *ratioFail = false;
    if (bestDist <= TH_HIGH) {
        if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) { *ratioFail = true; return -1; }
        return bestIdx;
    }
    return -1;
}

__global__ void __launch_bounds__(1024)
rig_resolve_kernel(RigArgs A)
{
    __shared__ int s_changed, s_matches;
    const int tid = threadIdx.x;
    const int n = A.nL + A.nR;
    int* prev = A.claim0; int* next = A.claim1;
    for (int i = tid; i < n; i += 1024) prev[i] = A.assign[i] >= 0 ? -1 : INT_MAX;
    for (int i = tid; i < 2 * A.nmp; i += 1024) A.choice[i] = -2;
    __syncthreads();
    for (int iter = 0; iter <= A.nmp + 1; iter++) {
        if (tid == 0) { s_changed = 0; s_matches = 0; }
        for (int i = tid; i < n; i += 1024) { next[i] = A.assign[i] >= 0 ? -1 : INT_MAX; A.last[i] = -1; }
        __syncthreads();
        int changed = 0, matches = 0;
        for (int i = tid; i < A.nmp; i += 1024) {
            bool fail = false;
            const int chL = rig_pick(A.candL + (long long)i * A.kmax * 2, A.cntL[i], prev, 0, 2 * i, -1, A.nnratio, &fail);
            int partner = -1;
            if (chL >= 0) {
                atomicMin(&next[chL], 2 * i); atomicMax(&A.last[chL], 2 * i);
                partner = A.l2r[chL];
                if (partner != -1) { atomicMin(&next[partner + A.nL], 2 * i); atomicMax(&A.last[partner + A.nL], 2 * i); matches++; }
                matches++;
            }
            int chR = -1;
            if (!fail) {                                                   // a failed left ratio test `continue`s past the right side
                bool failR = false;
                chR = rig_pick(A.candR + (long long)i * A.kmax * 2, A.cntR[i], prev, A.nL, 2 * i + 1, partner, A.nnratio, &failR);
                if (chR >= 0) {
                    const int p = A.r2l[chR];
                    if (p != -1) { atomicMin(&next[p], 2 * i + 1); atomicMax(&A.last[p], 2 * i + 1); matches++; }
                    atomicMin(&next[chR + A.nL], 2 * i + 1); atomicMax(&A.last[chR + A.nL], 2 * i + 1);
                    matches++;
                }
            }
            if (chL != A.choice[2 * i] || chR != A.choice[2 * i + 1]) { changed = 1; A.choice[2 * i] = chL; A.choice[2 * i + 1] = chR; }
        }
        if (changed) atomicOr(&s_changed, 1);
        if (matches) atomicAdd(&s_matches, matches);
        __syncthreads();
        int* t = prev; prev = next; next = t;
        const int ch = s_changed;
        __syncthreads();
        if (!ch) break;
    }
    for (int i = tid; i < n; i += 1024) if (A.last[i] >= 0) A.assign[i] = A.last[i] >> 1;
    if (tid == 0) A.result[0] = s_matches;
}

static int orbx_search_by_projection_rig_once(int device,
                                             const orbx_keypoint* kpsL, int nL, const orbx_keypoint* kpsR, int nR, const uint8_t* desc,
                                             const int32_t* leftToRight, const int32_t* rightToLeft,
                                             float minX, float maxX, float minY, float maxY,
                                             const float* scaleFactors, int nlevels,
                                             int nmp, const uint8_t* mp_desc, const float* mp_depth,
                                             const uint8_t* inviewL, const float* projxL, const float* projyL, const int32_t* levelL, const float* viewcosL,
                                             const uint8_t* inviewR, const float* projxR, const float* projyR, const int32_t* levelR, const float* viewcosR,
                                             float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    const int kmax = t_proj_kmax;
    if (nL < 0 || nR < 0 || nmp < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !scaleFactors || !(maxX > minX) || !(maxY > minY) ||
        (nL > 0 && (!kpsL || !leftToRight)) || (nR > 0 && (!kpsR || !rightToLeft)) || (nL + nR > 0 && (!desc || !assign)) ||
        (nmp > 0 && (!mp_desc || !inviewL || !projxL || !projyL || !levelL || !viewcosL || !inviewR || !projxR || !projyR || !levelR || !viewcosR)) ||
        (bFarPoints && !mp_depth)) {
        orbx_set_error("bad search_by_projection_rig arguments"); return ORBX_ERR_ARG;
    }
    if (nL + nR >= (1 << 24)) { orbx_set_error("too many keypoints"); return ORBX_ERR_ARG; }
    if (nL + nR == 0 || nmp == 0) return 0;
    for (int i = 0; i < nmp; i++) {
        if (inviewL[i] && (levelL[i] < 0 || levelL[i] >= nlevels)) { orbx_set_error("map point %d: level %d out of range", i, levelL[i]); return ORBX_ERR_ARG; }
        if (inviewR[i] && (levelR[i] < -1 || levelR[i] >= nlevels)) { orbx_set_error("map point %d: right level %d out of range", i, levelR[i]); return ORBX_ERR_ARG; }
    }
    for (int i = 0; i < nL; i++) if (leftToRight[i] < -1 || leftToRight[i] >= nR) { orbx_set_error("mvLeftToRightMatch[%d] out of range", i); return ORBX_ERR_ARG; }
    for (int i = 0; i < nR; i++) if (rightToLeft[i] < -1 || rightToLeft[i] >= nL) { orbx_set_error("mvRightToLeftMatch[%d] out of range", i); return ORBX_ERR_ARG; }
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(device);
    if (rc != ORBX_OK) return rc;
    const size_t NL = (size_t)nL, NR = (size_t)nR, N = NL + NR, M = (size_t)nmp;
    OrbxCarver c;
    const size_t o_kl = c.take(NL * sizeof(orbx_keypoint)), o_kr = c.take(NR * sizeof(orbx_keypoint)), o_desc = c.take(N * 32);
    const size_t o_l2r = c.take(NL * 4), o_r2l = c.take(NR * 4);
    const size_t o_md = c.take(M * 32), o_dp = c.take(mp_depth ? M * 4 : 0);
    size_t o_iv[2], o_px[2], o_py[2], o_lv[2], o_vc[2];
    for (int sd = 0; sd < 2; sd++) { o_iv[sd] = c.take(M); o_px[sd] = c.take(M * 4); o_py[sd] = c.take(M * 4); o_lv[sd] = c.take(M * 4); o_vc[sd] = c.take(M * 4); }
    const size_t o_assign = c.take(N * 4), o_result = c.take(8);
    const size_t in_bytes = c.off;
    size_t o_cs[2], o_ci[2], o_cc[2], o_cl[2];
    for (int sd = 0; sd < 2; sd++) { o_cs[sd] = c.take(4 * (size_t)(GRID_CELLS + 1)); o_ci[sd] = c.take((sd ? NR : NL) * 4 + 4); o_cc[sd] = c.take(M * 4); o_cl[sd] = c.take(8 * M * (size_t)kmax); }
    const size_t o_c0 = c.take(N * 4), o_c1 = c.take(N * 4), o_last = c.take(N * 4), o_ch = c.take(M * 8);
    rc = W.reserve(c.off, in_bytes);
    if (rc != ORBX_OK) return rc;
    char* h = W.h; char* d = W.d;
    if (nL) { memcpy(h + o_kl, kpsL, NL * sizeof(orbx_keypoint)); memcpy(h + o_l2r, leftToRight, NL * 4); }
    if (nR) { memcpy(h + o_kr, kpsR, NR * sizeof(orbx_keypoint)); memcpy(h + o_r2l, rightToLeft, NR * 4); }
    memcpy(h + o_desc, desc, N * 32); memcpy(h + o_md, mp_desc, M * 32);
    if (mp_depth) memcpy(h + o_dp, mp_depth, M * 4);
    const uint8_t* iv[2] = {inviewL, inviewR}; const float* px[2] = {projxL, projxR}; const float* py[2] = {projyL, projyR};
    const int32_t* lv[2] = {levelL, levelR}; const float* vc[2] = {viewcosL, viewcosR};
    for (int sd = 0; sd < 2; sd++) {
        memcpy(h + o_iv[sd], iv[sd], M); memcpy(h + o_px[sd], px[sd], M * 4); memcpy(h + o_py[sd], py[sd], M * 4);
        memcpy(h + o_lv[sd], lv[sd], M * 4); memcpy(h + o_vc[sd], vc[sd], M * 4);
    }
    memcpy(h + o_assign, assign, N * 4); memset(h + o_result, 0, 8);
    CUDA_TRYW(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, W.s));
    for (int sd = 0; sd < 2; sd++) {
        const int n = sd ? nR : nL;
        ProjArgs A;
        memset(&A, 0, sizeof(A));
        A.kps = (const orbx_keypoint*)(d + (sd ? o_kr : o_kl)); A.desc = (const uint4*)(d + o_desc + (sd ? NL * 32 : 0));
        A.uRight = nullptr;                                                   // the stereo gate is for Nleft == -1 only, :105
        A.mp_desc = (const uint4*)(d + o_md); A.projx = (const float*)(d + o_px[sd]); A.projy = (const float*)(d + o_py[sd]);
        A.level = (const int*)(d + o_lv[sd]); A.viewcos = (const float*)(d + o_vc[sd]);
        A.depth = mp_depth ? (const float*)(d + o_dp) : nullptr; A.inview = (const uint8_t*)(d + o_iv[sd]);
        A.result = (int*)(d + o_result); A.kmax = kmax;
        A.cell_start = (int*)(d + o_cs[sd]); A.cell_idx = (int*)(d + o_ci[sd]); A.cand_cnt = (int*)(d + o_cc[sd]); A.cand = (int*)(d + o_cl[sd]);
        A.n = n; A.nmp = nmp;
        A.minX = minX; A.minY = minY;
        A.winv = (float)GRID_COLS / (float)(maxX - minX);
        A.hinv = (float)GRID_ROWS / (float)(maxY - minY);
        for (int l = 0; l < nlevels; l++) A.sf[l] = scaleFactors[l];
        A.th = sd ? 1.0f : th;                                                // :173: no bFactor on the right side
        A.nnratio = nnratio; A.bFar = bFarPoints; A.thFar = thFarPoints;
        GridArgs g;
        g.kps = A.kps; g.n = n; g.minX = A.minX; g.minY = A.minY; g.winv = A.winv; g.hinv = A.hinv;
        g.cell_start = A.cell_start; g.cell_idx = A.cell_idx;
        orbx_launch_grid_build(g, W.s);
        proj_candidates_kernel<<<(nmp + PROJ_WARPS - 1) / PROJ_WARPS, PROJ_WARPS * 32, 0, W.s>>>(A);
    }
    RigArgs R;
    R.nL = nL; R.nR = nR; R.nmp = nmp;
    R.candL = (const int*)(d + o_cl[0]); R.cntL = (const int*)(d + o_cc[0]); R.candR = (const int*)(d + o_cl[1]); R.cntR = (const int*)(d + o_cc[1]);
    R.l2r = (const int*)(d + o_l2r); R.r2l = (const int*)(d + o_r2l); R.nnratio = nnratio; R.kmax = kmax;
    R.assign = (int*)(d + o_assign); R.claim0 = (int*)(d + o_c0); R.claim1 = (int*)(d + o_c1); R.last = (int*)(d + o_last);
    R.choice = (int*)(d + o_ch); R.result = (int*)(d + o_result);
    rig_resolve_kernel<<<1, 1024, 0, W.s>>>(R);
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(h + o_assign, d + o_assign, in_bytes - o_assign, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    const int* result = (const int*)(h + o_result);
    if (result[1] > kmax) { t_proj_kmax = (result[1] + 63) / 64 * 64; return ORBX_RETRY_KMAX; }   // nothing of the caller's was touched yet
    memcpy(assign, h + o_assign, N * 4);
    return result[0];
}

extern "C" int orbx_search_by_projection_rig(int device,
                                             const orbx_keypoint* kpsL, int nL, const orbx_keypoint* kpsR, int nR, const uint8_t* desc,
                                             const int32_t* leftToRight, const int32_t* rightToLeft,
                                             float minX, float maxX, float minY, float maxY,
                                             const float* scaleFactors, int nlevels,
                                             int nmp, const uint8_t* mp_desc, const float* mp_depth,
                                             const uint8_t* inviewL, const float* projxL, const float* projyL, const int32_t* levelL, const float* viewcosL,
                                             const uint8_t* inviewR, const float* projxR, const float* projyR, const int32_t* levelR, const float* viewcosR,
                                             float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign)
{
    // a map point whose window holds more keypoints than the candidate lists have room for makes the pass report the
    // longest list; the lists are regrown (per calling thread, kept) and the pass repeated - the reference has no such limit
    for (int attempt = 0; attempt < 4; attempt++) {
        const int r = orbx_search_by_projection_rig_once(device, kpsL, nL, kpsR, nR, desc, leftToRight, rightToLeft, minX, maxX, minY, maxY, scaleFactors, nlevels, nmp, mp_desc, mp_depth, inviewL, projxL, projyL, levelL, viewcosL, inviewR, projxR, projyR, levelR, viewcosR, th, nnratio, bFarPoints, thFarPoints, assign);
        if (r != ORBX_RETRY_KMAX) return r;
    }
    orbx_set_error("candidate lists kept overflowing");
    return ORBX_ERR_CAPACITY;
}
