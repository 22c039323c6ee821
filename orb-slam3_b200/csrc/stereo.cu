// stereo.cu — Frame::ComputeStereoMatches (reference src/Frame.cc:1027-1276) on the device-resident
// keypoints, descriptors and pyramids of a left and a right extractor handle.
//
// stereo_rows_kernel (one CTA per frame) is vRowIndices (:1064-1078) at 16-row granularity: every right
// keypoint's row band [floor(y-r), ceil(y+r)], r = 2*scale[octave], is computed ONCE, and for every
// 16-row band of the image the right keypoints touching it are listed in ascending iR (ballot
// compaction keeps the reference's push_back order) as 12-byte entries {minr|maxr|octave, x, iR}.
// stereo_match_kernel: one warp per left keypoint.  Lanes stride the entries of band (int)vL >> 4
// (~5 % of the right keypoints), apply the exact row / octave +-1 / uR in [uL-maxD, uL] gates
// (:1114-1125) and the 256-bit Hamming distance by 8 x __popc; warp arg-min with the reference's
// "first candidate wins" tie-break (lowest iR).  If the best distance is < 75 the same warp does the
// 11 x (11x11) SAD search in the left keypoint's pyramid level (:1153-1207, no mean subtraction in this
// fork) and lane 0 the parabola / disparity tail in non-contracted FP32.
// stereo_median_kernel: median of the accepted SADs (element nd/2 of the ascending list, :1261-1263) by a
// two-pass radix select on the 15-bit SAD values, then the 1.5*1.4*median reject (:1266-1275).
// Integer pipe bound (POPC/LOP3/compare); reported in G int-op/s, not against HBM.
#include "orbx_host.h"

#define TH_HIGH 100
#define TH_LOW 50
#define ST_BAND_SHIFT 4                    // 16-row bands
#define ST_ENTRY_WORDS 3

struct StereoArgs {
    const orbx_keypoint* kpL; const uint8_t* descL; const int* countsL;
    const orbx_keypoint* kpR; const uint8_t* descR; const int* countsR;
    int out_cap;
    float* uRight; float* depth; int* sad;     // [frames][out_cap]
    uint32_t* rows;                            // [frames][rows_stride]: band starts (nbands+1), then entries
    long long rows_stride; int nbands; int start_words; int entry_cap;
    float mb, mbf;
    int nRows;
    int nlevels;
    float sf[ORBX_MAX_LEVELS], isf[ORBX_MAX_LEVELS];
    const uint8_t* pyrL[ORBX_MAX_LEVELS]; const uint8_t* pyrR[ORBX_MAX_LEVELS];
    int pitch[ORBX_MAX_LEVELS]; long long plane_bytes[ORBX_MAX_LEVELS]; int cols[ORBX_MAX_LEVELS];
};

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4* b)
{
    const uint4 b0 = b[0], b1 = b[1];
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// ---------------------------------------------------------------------------------------------------
// vRowIndices at band granularity
#define SR_THREADS 512
#define SR_TILE 4096                        // right keypoints staged per pass (32 KB of shared memory)

__global__ void __launch_bounds__(SR_THREADS)
stereo_rows_kernel(StereoArgs A)
{
    __shared__ uint32_t s_word[SR_TILE];    // minr | maxr << 13 | octave << 26
    __shared__ float s_x[SR_TILE];
    __shared__ int s_start[1024 + 1];       // nbands <= 1024 (images up to 16 k rows)
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int Nr = A.countsR[2 * f];
    const long long fo = (long long)f * A.out_cap;
    uint32_t* start = A.rows + (long long)f * A.rows_stride;
    uint32_t* ent = start + A.start_words;
    const int nb = A.nbands;
    // pass 0 counts, pass 1 fills; each pass walks the right keypoints tile by tile
    for (int pass = 0; pass < 2; pass++) {
        if (pass == 0) for (int b = tid; b <= nb; b += SR_THREADS) s_start[b] = 0;
        __syncthreads();
        for (int t0 = 0; t0 < Nr; t0 += SR_TILE) {
            const int nt = Nr - t0 < SR_TILE ? Nr - t0 : SR_TILE;
            for (int i = tid; i < nt; i += SR_THREADS) {
                const orbx_keypoint kr = A.kpR[fo + t0 + i];
                const float r = __fmul_rn(2.0f, A.sf[kr.octave]);                       // :1071
                int maxr = (int)ceilf(__fadd_rn(kr.y, r)), minr = (int)floorf(__fsub_rn(kr.y, r));
                if (minr < 0) minr = 0;
                if (maxr > A.nRows - 1) maxr = A.nRows - 1;                             // rows that exist (the reference would index out of range)
                s_word[i] = (maxr < minr) ? 0xFFFFFFFFu : ((uint32_t)minr | ((uint32_t)maxr << 13) | ((uint32_t)kr.octave << 26));
                s_x[i] = kr.x;
            }
            __syncthreads();
            for (int b = warp; b < nb; b += SR_THREADS / 32) {
                int base = pass ? s_start[b] : 0;
                for (int i0 = 0; i0 < nt; i0 += 32) {
                    const int i = i0 + lane;
                    uint32_t w = 0xFFFFFFFFu;
                    if (i < nt) w = s_word[i];
                    const bool in = w != 0xFFFFFFFFu && (int)((w & 0x1FFFu) >> ST_BAND_SHIFT) <= b && (int)(((w >> 13) & 0x1FFFu) >> ST_BAND_SHIFT) >= b;
                    const uint32_t m = __ballot_sync(0xffffffffu, in);
                    if (pass && in) {
                        const int pos = base + __popc(m & ((1u << lane) - 1u));
                        if (pos < A.entry_cap) {
                            uint32_t* e = ent + (long long)pos * ST_ENTRY_WORDS;
                            e[0] = w; e[1] = __float_as_uint(s_x[i]); e[2] = (uint32_t)(t0 + i);
                        }
                    }
                    base += __popc(m);
                }
                if (lane == 0) { if (pass) s_start[b] = base; else s_start[b] += base; }
            }
            __syncthreads();
        }
        if (pass == 0) {
            // exclusive scan of the band counts (nb is small: one warp, chunked)
            if (warp == 0) {
                int run = 0;
                for (int b0 = 0; b0 <= nb; b0 += 32) {
                    const int b = b0 + lane;
                    const int c = b < nb ? s_start[b] : 0;
                    int inc = c;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
                    if (b <= nb) { s_start[b] = run + inc - c; start[b] = (uint32_t)(run + inc - c); }
                    run += __shfl_sync(0xffffffffu, inc, 31);
                }
            }
            __syncthreads();
        }
    }
}

#define ST_WARPS 8

__global__ void __launch_bounds__(ST_WARPS * 32)
stereo_match_kernel(StereoArgs A)
{
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * ST_WARPS + (threadIdx.x >> 5);
    const int N = A.countsL[2 * f];
    if (iL >= N) return;
    const long long fo = (long long)f * A.out_cap;
    const orbx_keypoint kl = A.kpL[fo + iL];
    float uRight = -1.0f, depth = -1.0f;             // :1041-1042
    int sadOut = -1;
    const int levelL = kl.octave;
    const float vL = kl.y, uL = kl.x;
    const int row = (int)vL;                          // vRowIndices[vL], :1097
    const float minD = 0.f, maxD = __fdiv_rn(A.mbf, A.mb);     // :1081-1083
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    int bestDist = TH_HIGH, bestIdxR = 0x7fffffff;
    if (maxU >= 0 && row >= 0 && row < A.nRows) {
        const uint4* dl = reinterpret_cast<const uint4*>(A.descL + (fo + iL) * 32);
        const uint4 a0 = dl[0], a1 = dl[1];
        const uint32_t* start = A.rows + (long long)f * A.rows_stride;
        const uint32_t* ent = start + A.start_words;
        const int b = row >> ST_BAND_SHIFT;
        int e0 = (int)start[b], e1 = (int)start[b + 1];
        if (e1 > A.entry_cap) e1 = A.entry_cap;
        for (int e = e0 + lane; e < e1; e += 32) {
            const uint32_t w = ent[(long long)e * ST_ENTRY_WORDS];
            const int minr = (int)(w & 0x1FFFu), maxr = (int)((w >> 13) & 0x1FFFu), oct = (int)(w >> 26);
            if (row < minr || row > maxr) continue;
            if (oct < levelL - 1 || oct > levelL + 1) continue;                         // :1118
            const float uR = __uint_as_float(ent[(long long)e * ST_ENTRY_WORDS + 1]);
            if (uR >= minU && uR <= maxU) {                                             // :1123
                const int iR = (int)ent[(long long)e * ST_ENTRY_WORDS + 2];
                const int dist = hamming256(a0, a1, reinterpret_cast<const uint4*>(A.descR + (fo + iR) * 32));
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }                // entries ascend in iR: first wins
            }
        }
        // warp arg-min, ties to the lowest iR (= first in the reference's candidate order)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const int od = __shfl_xor_sync(0xffffffffu, bestDist, o), oi = __shfl_xor_sync(0xffffffffu, bestIdxR, o);
            if (od < bestDist || (od == bestDist && oi < bestIdxR)) { bestDist = od; bestIdxR = oi; }
        }
    }
    if (bestDist < (TH_HIGH + TH_LOW) / 2 && bestIdxR != 0x7fffffff) {            // :1139
        const float uR0 = A.kpR[fo + bestIdxR].x;
        const float scaleFactor = A.isf[levelL];
        const float scaleduL = roundf(__fmul_rn(kl.x, scaleFactor));              // :1144-1146
        const float scaledvL = roundf(__fmul_rn(kl.y, scaleFactor));
        const float scaleduR0 = roundf(__fmul_rn(uR0, scaleFactor));
        const int w = 5, Lr = 5;
        const float iniu = scaleduR0 + Lr - w, endu = scaleduR0 + Lr + w + 1;     // :1162-1163
        if (!(iniu < 0 || endu >= (float)A.cols[levelL])) {                       // :1164
            const int pitch = A.pitch[levelL];
            const uint8_t* PL = A.pyrL[levelL] + (long long)f * A.plane_bytes[levelL] + (long long)ORBX_YOFF * pitch + ORBX_XOFF;
            const uint8_t* PR = A.pyrR[levelL] + (long long)f * A.plane_bytes[levelL] + (long long)ORBX_YOFF * pitch + ORBX_XOFF;
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w), xr00 = (int)(scaleduR0 - w);
            // each lane owns up to 4 of the 121 patch pixels
            int il[4], off[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int p = lane + 32 * k;
                if (p < 121) {
                    const int yy = p / 11, xx = p - yy * 11;
                    il[k] = PL[(long long)(y0 + yy) * pitch + xl0 + xx];
                    off[k] = (y0 + yy) * pitch + xr00 + xx;
                } else { il[k] = 0; off[k] = -1; }
            }
            int bestSad = 0x7fffffff, bestincR = 0;
            float vDists[11];
#pragma unroll
            for (int inc = -5; inc <= 5; inc++) {                                  // :1167-1180
                int s = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) if (off[k] >= 0) s += abs(il[k] - (int)PR[off[k] + inc]);
                s = __reduce_add_sync(0xffffffffu, s);
                const float dist = (float)s;
                if (dist < (float)bestSad) { bestSad = s; bestincR = inc; }
                vDists[inc + 5] = dist;
            }
            if (!(bestincR == -Lr || bestincR == Lr)) {                            // :1182
                float dist1 = 0.f, dist2 = 0.f, dist3 = 0.f;
#pragma unroll
                for (int k = 1; k < 10; k++) if (k == bestincR + 5) { dist1 = vDists[k - 1]; dist2 = vDists[k]; dist3 = vDists[k + 1]; }
                const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                               __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));   // :1190
                if (!(deltaR < -1 || deltaR > 1)) {
                    float bestuR = __fmul_rn(A.sf[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));      // :1196
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= minD && disparity < maxD) {                  // :1200
                        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
                        depth = __fdiv_rn(A.mbf, disparity);
                        uRight = bestuR;
                        sadOut = bestSad;
                    }
                }
            }
        }
    }
    if (lane == 0) { A.uRight[fo + iL] = uRight; A.depth[fo + iL] = depth; A.sad[fo + iL] = sadOut; }
}

#define MED_THREADS 256

// median of the accepted SADs = element nd/2 of the ascending list (:1261-1263), then reject (:1266-1275).
// SAD <= 121*255 = 30855 < 2^15: radix select, high 8 bits then low 7 bits.
__global__ void __launch_bounds__(MED_THREADS)
stereo_median_kernel(StereoArgs A)
{
    __shared__ int s_hist[256];
    __shared__ int s_bin, s_rank, s_median;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int N = A.countsL[2 * f];
    const long long fo = (long long)f * A.out_cap;
    const int* sad = A.sad + fo;
    s_hist[tid] = 0;
    if (tid == 0) { s_bin = -1; s_median = -1; }
    __syncthreads();
    for (int i = tid; i < N; i += MED_THREADS) { const int v = sad[i]; if (v >= 0) atomicAdd(&s_hist[v >> 7], 1); }
    __syncthreads();
    // bin holding rank k = nd/2: warp 0 scans the 256 bins
    if (tid < 32) {
        int c[8], tot = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) { c[j] = s_hist[lane * 8 + j]; tot += c[j]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        const int nd = __shfl_sync(0xffffffffu, inc, 31);
        const int k = nd / 2;
        int before = inc - tot;
        if (nd > 0 && before <= k && k < inc) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (k >= before && k < before + c[j]) { s_bin = lane * 8 + j; s_rank = k - before; }
                before += c[j];
            }
        }
    }
    __syncthreads();
    const int bin = s_bin;
    if (bin < 0) return;                              // no match at all: the reference indexes an empty vector here (UB); guarded
    const int rank = s_rank;
    __syncthreads();
    s_hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += MED_THREADS) { const int v = sad[i]; if (v >= 0 && (v >> 7) == bin) atomicAdd(&s_hist[v & 127], 1); }
    __syncthreads();
    if (tid < 32) {
        int c[4], tot = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) { c[j] = s_hist[lane * 4 + j]; tot += c[j]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        int before = inc - tot;
        if (before <= rank && rank < inc) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (rank >= before && rank < before + c[j]) s_median = (bin << 7) | (lane * 4 + j);
                before += c[j];
            }
        }
    }
    __syncthreads();
    const float median = (float)s_median;
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), median);
    for (int i = tid; i < N; i += MED_THREADS) {
        const int v = sad[i];
        if (v >= 0 && !((float)v < thDist)) { A.uRight[fo + i] = -1.f; A.depth[fo + i] = -1.f; }
    }
}

static int fill_args(orbx_extractor* eL, orbx_extractor* eR, float mb, float mbf, StereoArgs* A)
{
    if (!eL || !eR || !eL->reserved || !eR->reserved || eL->last_nframes <= 0 || eL->last_nframes != eR->last_nframes) {
        orbx_set_error("stereo match needs a preceding extract of the same number of frames on both handles");
        return ORBX_ERR_STATE;
    }
    if (eL->device != eR->device || eL->G.rows != eR->G.rows || eL->G.cols != eR->G.cols || eL->nlevels != eR->nlevels ||
        eL->G.out_cap != eR->G.out_cap || eL->scaleFactorF != eR->scaleFactorF) {
        orbx_set_error("left/right extractors differ in device, image size or parameters");
        return ORBX_ERR_ARG;
    }
    // mb == 0 is accepted: maxD = mbf / 0 = +inf, no upper disparity bound - what the reference's rectified-stereo constructor
    // effectively runs with (it calls ComputeStereoMatches, src/Frame.cc:137, before it assigns mb, :163)
    if (!(mb >= 0.f)) { orbx_set_error("mb must be >= 0"); return ORBX_ERR_ARG; }
    if (!eL->d_st_scratch) {
        void* q = nullptr;
        if (cudaMalloc(&q, sizeof(int) * (size_t)eL->G.batch * eL->G.out_cap) != cudaSuccess) {
            orbx_set_error("cudaMalloc failed (stereo scratch)"); return ORBX_ERR_CUDA;
        }
        eL->allocs.push_back(q);
        eL->d_st_scratch = (int*)q;
    }
    // row-band index of the right keypoints: band starts + entries; a keypoint's band [floor(y-r), ceil(y+r)]
    // spans at most 2*ceil(r)+2 rows, r = 2*scale[top level]
    const int nRows = eL->G.L[0].h;
    if (nRows > 8191) { orbx_set_error("stereo match: images taller than 8191 rows are not supported"); return ORBX_ERR_GEOMETRY; }
    const int span = 2 * (int)ceilf(2.0f * eL->sf[eL->nlevels - 1]) + 2;
    const int bands_per_kp = (span + (1 << ST_BAND_SHIFT) - 2) / (1 << ST_BAND_SHIFT) + 1;
    A->nbands = ((nRows - 1) >> ST_BAND_SHIFT) + 1;
    A->start_words = (A->nbands + 1 + 3) / 4 * 4;
    A->entry_cap = bands_per_kp * eL->G.out_cap;
    A->rows_stride = A->start_words + (long long)A->entry_cap * ST_ENTRY_WORDS;
    if (!eL->d_st_rows) {
        void* q = nullptr;
        if (cudaMalloc(&q, sizeof(uint32_t) * (size_t)eL->G.batch * (size_t)A->rows_stride) != cudaSuccess) {
            orbx_set_error("cudaMalloc failed (stereo row index)"); return ORBX_ERR_CUDA;
        }
        eL->allocs.push_back(q);
        eL->d_st_rows = (uint32_t*)q;
    }
    A->rows = eL->d_st_rows;
    A->kpL = eL->D.kps; A->descL = eL->D.desc; A->countsL = eL->D.counts;
    A->kpR = eR->D.kps; A->descR = eR->D.desc; A->countsR = eR->D.counts;
    A->out_cap = eL->G.out_cap;
    A->uRight = eL->d_uright; A->depth = eL->d_depth; A->sad = eL->d_st_scratch;
    A->mb = mb; A->mbf = mbf;
    A->nRows = eL->G.L[0].h;                                                      // mvImagePyramid[0].rows, :1048
    A->nlevels = eL->nlevels;
    for (int l = 0; l < eL->nlevels; l++) {
        A->sf[l] = eL->sf[l]; A->isf[l] = eL->isf[l];
        A->pyrL[l] = eL->G.L[l].pyr; A->pyrR[l] = eR->G.L[l].pyr;
        A->pitch[l] = eL->G.L[l].pitch; A->plane_bytes[l] = eL->G.L[l].plane_bytes; A->cols[l] = eL->G.L[l].w;
    }
    return 0;
}

extern "C" int orbx_stereo_match_resident(orbx_extractor* eL, orbx_extractor* eR, float mb, float mbf)
{
    StereoArgs A;
    int r = fill_args(eL, eR, mb, mbf, &A);
    if (r) return r;
    if (cudaSetDevice(eL->device) != cudaSuccess) { orbx_set_error("cudaSetDevice failed"); return ORBX_ERR_CUDA; }
    // order after both extractions: the right handle has its own stream
    cudaEvent_t ev = eR->join_ev;                 // not a profiling event: eR's stage times stay intact
    cudaEventRecord(ev, eR->stream);
    cudaStreamWaitEvent(eL->stream, ev, 0);
    const int nf = eL->last_nframes;
    const dim3 grid((eL->G.out_cap + ST_WARPS - 1) / ST_WARPS, nf);
    stereo_rows_kernel<<<nf, SR_THREADS, 0, eL->stream>>>(A);
    stereo_match_kernel<<<grid, ST_WARPS * 32, 0, eL->stream>>>(A);
    stereo_median_kernel<<<nf, MED_THREADS, 0, eL->stream>>>(A);
    eL->launches += 3;
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) { orbx_set_error("stereo launch failed: %s", cudaGetErrorString(ce)); return ORBX_ERR_CUDA; }
    return ORBX_OK;
}

static int stereo_download(orbx_extractor* eL, int frame0, int nframes, float* uRight, float* depth, int cap)
{
    const int oc = eL->G.out_cap, n = cap < oc ? cap : oc;
    cudaError_t ce = cudaMemcpy2DAsync(uRight, sizeof(float) * cap, eL->d_uright + (size_t)frame0 * oc, sizeof(float) * oc,
                                       sizeof(float) * n, nframes, cudaMemcpyDeviceToHost, eL->stream);
    if (ce == cudaSuccess)
        ce = cudaMemcpy2DAsync(depth, sizeof(float) * cap, eL->d_depth + (size_t)frame0 * oc, sizeof(float) * oc,
                               sizeof(float) * n, nframes, cudaMemcpyDeviceToHost, eL->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(eL->stream);
    if (ce != cudaSuccess) { orbx_set_error("stereo download failed: %s", cudaGetErrorString(ce)); return ORBX_ERR_CUDA; }
    return ORBX_OK;
}

extern "C" int orbx_stereo_match_batch(orbx_extractor* eL, orbx_extractor* eR, float mb, float mbf,
                                       float* uRight, float* depth, int cap)
{
    if (!uRight || !depth || cap <= 0) { orbx_set_error("bad stereo output buffers"); return ORBX_ERR_ARG; }
    int r = orbx_stereo_match_resident(eL, eR, mb, mbf);
    if (r) return r;
    return stereo_download(eL, 0, eL->last_nframes, uRight, depth, cap);
}

extern "C" int orbx_stereo_match(orbx_extractor* eL, orbx_extractor* eR, int frame, float mb, float mbf,
                                 float* uRight, float* depth, int cap)
{
    if (!uRight || !depth || cap <= 0) { orbx_set_error("bad stereo output buffers"); return ORBX_ERR_ARG; }
    int r = orbx_stereo_match_resident(eL, eR, mb, mbf);
    if (r) return r;
    if (frame < 0 || frame >= eL->last_nframes) { orbx_set_error("frame out of range"); return ORBX_ERR_ARG; }
    return stereo_download(eL, frame, 1, uRight, depth, cap);
}
