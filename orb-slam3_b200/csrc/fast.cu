// fast.cu — per-cell FAST-9/16 with the iniThFAST/minThFAST retry
// (ORBextractor::ComputeKeyPointsOctTree cell loop, reference src/ORBextractor.cc:1038-1143, and the
// cv::FAST / cornerScore<16> / 3x3 non-max suppression it calls, restated in oracle/orc_prims.c).
//
// One CTA per 35-px cell: the (wCell+6)x(hCell+6) sub-image is staged in shared memory once, the
// score map of its interior is computed at iniThFAST; only if NMS leaves nothing the cell is redone
// at minThFAST (exactly the reference's per-cell semantics: NMS sees zero scores outside the cell).
// Survivors are compacted in raster order into the cell's slot range; the quadtree kernel later
// concatenates cells in row-major order, which reproduces vToDistributeKeys' order.
// Integer/byte work, HBM-bound in the limit: algorithmic bytes per frame = read sum(L0..L7).
#include "orbx_internal.h"

#define FAST_THREADS 128
#define TILE_MAXW 80      // >= wCell+6 (wCell < 70)
#define TILE_MAXH 76
#define SC_MAXW 72        // interior + 1-px zero ring

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

__global__ void __launch_bounds__(FAST_THREADS)
fast_cells_kernel(FrameGeom G, DevBuffers D)
{
    __shared__ uint8_t tile[TILE_MAXH * TILE_MAXW];
    __shared__ uint8_t sc[(SC_MAXW) * (SC_MAXW)];
    __shared__ uint8_t keep[70 * 70];
    __shared__ int warp_tot[FAST_THREADS / 32];
    __shared__ int s_total;

    const int f = blockIdx.y;
    int c = blockIdx.x;
    int lvl = 0;
#pragma unroll 1
    for (int l = 1; l < G.nlevels; l++) if (c >= G.L[l].cell_base) lvl = l;
    const LevelGeom& L = G.L[lvl];
    c -= L.cell_base;
    const int ci = c / L.nCols, cj = c - ci * L.nCols;
    uint32_t* cnt_out = D.cellcnt + (long long)f * G.cells_per_frame + L.cell_base + c;
    uint32_t* key_out = D.cellkey + (long long)f * G.keys_per_frame + L.key_base + (long long)c * L.cell_cap;

    // cell geometry, src/ORBextractor.cc:1078-1106 (all values are small integers, float math is exact)
    const int y0 = ORBX_MINB + ci * L.hCell, x0 = ORBX_MINB + cj * L.wCell;
    const int y1 = min(y0 + L.hCell + 6, L.maxBY), x1 = min(x0 + L.wCell + 6, L.maxBX);
    const int sw = x1 - x0, sh = y1 - y0;
    if (y0 >= L.maxBY - 3 || x0 >= L.maxBX - 6 || sw < 7 || sh < 7) {
        if (threadIdx.x == 0) *cnt_out = 0;
        return;
    }
    const int iw = sw - 6, ih = sh - 6;      // tested interior
    const int scp = iw + 2;

    const uint8_t* src = L.pyr + (long long)f * L.plane_bytes + (long long)(ORBX_YOFF + y0) * L.pitch + ORBX_XOFF + x0;
    for (int i = threadIdx.x; i < sw * sh; i += FAST_THREADS) {
        const int r = i / sw, cc = i - r * sw;
        tile[r * TILE_MAXW + cc] = __ldg(src + (long long)r * L.pitch + cc);
    }
    for (int i = threadIdx.x; i < scp * (ih + 2); i += FAST_THREADS) sc[i] = 0;
    __syncthreads();

    const int P = iw * ih;
    const int chunk = (P + FAST_THREADS - 1) / FAST_THREADS;
    const int p0 = min(P, (int)threadIdx.x * chunk), p1 = min(P, p0 + chunk);
    int th = G.iniTh;
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        for (int i = threadIdx.x; i < P; i += FAST_THREADS) {
            const int y = i / iw, x = i - y * iw;
            sc[(y + 1) * scp + x + 1] = (uint8_t)fast_score(tile + (y + 3) * TILE_MAXW + x + 3, TILE_MAXW, th);
        }
        __syncthreads();
        int mycount = 0;
        for (int i = p0; i < p1; i++) {
            const int y = i / iw, x = i - y * iw;
            const uint8_t* q = sc + (y + 1) * scp + x + 1;
            const int s = q[0];
            const bool k = s != 0 && s > q[-1] && s > q[1] && s > q[-scp - 1] && s > q[-scp] && s > q[-scp + 1] &&
                           s > q[scp - 1] && s > q[scp] && s > q[scp + 1];
            keep[i] = k;
            mycount += k;
        }
        // block exclusive scan of mycount
        int incl = mycount;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += n;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
        __syncthreads();
        int base = 0, total = 0;
#pragma unroll
        for (int wv = 0; wv < FAST_THREADS / 32; wv++) {
            if (wv < (int)(threadIdx.x >> 5)) base += warp_tot[wv];
            total += warp_tot[wv];
        }
        if (total > 0 || pass == 1) {
            int o = base + incl - mycount;
            for (int i = p0; i < p1; i++) {
                if (keep[i]) {
                    const int y = i / iw, x = i - y * iw;
                    key_out[o++] = MAKE_KEY(x + 3 + cj * L.wCell, y + 3 + ci * L.hCell, sc[(y + 1) * scp + x + 1]);
                }
            }
            if (threadIdx.x == 0) *cnt_out = (uint32_t)total;
            return;
        }
        th = G.minTh;          // cell empty at iniThFAST -> retry, src/ORBextractor.cc:1118-1125
        __syncthreads();       // everyone done with sc/keep/warp_tot before the second pass
    }
    (void)s_total;
}

int launch_fast(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    const dim3 grid(G.cells_per_frame, nframes);
    fast_cells_kernel<<<grid, FAST_THREADS, 0, s>>>(G, D);
    return 1;
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
