// pyramid.cu — image pyramid (ORBextractor::ComputePyramid, reference src/ORBextractor.cc:1664-1715).
//
// Level 0 = copyMakeBorder(image, 19, BORDER_REFLECT_101); level l>0 = cv::resize(level l-1,
// INTER_LINEAR) in OpenCV's 11-bit fixed point followed by the same reflect-101 border.  Planes keep
// room for the border (rows are 128-byte pitched, interior starts on a 32-byte sector boundary) but the
// hot path only writes the interior; the border is produced on demand (orbx_get_pyramid_level).
// HBM-bound: algorithmic bytes per frame = read sum(L0..L6) + write sum(L1..L7).
#include "orbx_internal.h"
#include <cstdlib>

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

// level 0: input frame -> interior of the level-0 plane, 16 bytes per thread, plus the thin reflect-101
// halo the blur reads (PYR_HALO rows above / below, one 4-px word left / right; the pixels of the last
// word past column w-1 are mirrored too).  The full 19-px border of mvImagePyramid is NOT materialised
// on the hot path; orbx_get_pyramid_level(padded) fills it on demand with the two kernels below.
#define PYR_HALO 3
// ALIGNED: source rows start on 16-byte boundaries (uint4 loads); otherwise five aligned words + funnel shifts
template <bool ALIGNED>
__global__ void __launch_bounds__(256)
pyr_level0_kernel(const uint8_t* __restrict__ in, int in_pitch, long long in_frame_bytes,
                  uint8_t* __restrict__ out, int pitch, long long plane_bytes, int w, int h,
                  int* __restrict__ candcnt, int nlevels)
{
    orbx_pdl_prologue();
    // the per-level candidate counters FAST appends to start at zero (one thread per frame; saves a memset node
    // between the pyramid and FAST launches)
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.y == 0 && threadIdx.x < nlevels)
        candcnt[blockIdx.z * nlevels + threadIdx.x] = 0;
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;     // 16-byte group of the image row
    const int y = blockIdx.y * blockDim.y + threadIdx.y - PYR_HALO;
    if (y >= h + PYR_HALO || qi * 16 >= w) return;
    const int f = blockIdx.z;
    const uint8_t* srow = in + (long long)f * in_frame_bytes + (long long)reflect101(y, h) * in_pitch;
    uint8_t* drow = out + (long long)f * plane_bytes + (long long)(ORBX_YOFF + y) * pitch + ORBX_XOFF;
    uint4 v;
    if (ALIGNED) v = __ldg(reinterpret_cast<const uint4*>(srow) + qi);
    else {
        const uint8_t* sp = srow + 16 * qi;
        const uint32_t sh = ((uint32_t)(uintptr_t)sp & 3u) * 8u;
        const uint32_t* ap = reinterpret_cast<const uint32_t*>(sp - (sh >> 3));
        const uint32_t w0 = __ldg(ap), w1 = __ldg(ap + 1), w2 = __ldg(ap + 2), w3 = __ldg(ap + 3), w4 = __ldg(ap + 4);
        v = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
    }
    reinterpret_cast<uint4*>(drow)[qi] = v;
    if (qi == 0) *reinterpret_cast<uint32_t*>(drow - 4) = __byte_perm(v.x, v.y, 0x1234);      // columns 4,3,2,1
    if (qi * 16 + 16 >= w) {                                  // last group: words nw-1 and nw
        const int p0 = ((w + 3) / 4 - 1) * 4;
        uint32_t t[2] = {0u, 0u};
#pragma unroll
        for (int q = 0; q < 8; q++) t[q >> 2] |= (uint32_t)srow[reflect101(p0 + q, w)] << (8 * (q & 3));
        *reinterpret_cast<uint32_t*>(drow + p0) = t[0];
        *reinterpret_cast<uint32_t*>(drow + p0 + 4) = t[1];
    }
}

// on-demand border, step 1: rows above / below the image = reflect-101 rows (interior columns)
__global__ void __launch_bounds__(256)
pyr_border_rows_kernel(uint8_t* __restrict__ plane, int pitch, int w, int h)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;                                 // 0..37: 19 rows above, 19 below
    if (x >= w) return;
    const int Y = j < ORBX_EDGE ? j : h + j;                  // padded row
    const int sy = j < ORBX_EDGE ? ORBX_EDGE - j : h - 2 - (j - ORBX_EDGE);
    plane[(long long)Y * pitch + ORBX_XOFF + x] = plane[(long long)(ORBX_YOFF + sy) * pitch + ORBX_XOFF + x];
}

// on-demand border, step 2: columns left / right of every padded row: out[-k] = out[k],
// out[w-1+k] = out[w-1-k], k = 1..19 (copyMakeBorder BORDER_REFLECT_101, :1689-1693,1711-1713)
__global__ void __launch_bounds__(256)
pyr_border_cols_kernel(uint8_t* __restrict__ plane, int pitch, int w, int prow)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int Y = idx / (2 * ORBX_EDGE), j = idx - Y * (2 * ORBX_EDGE);
    if (Y >= prow) return;
    uint8_t* row = plane + (long long)Y * pitch + ORBX_XOFF;
    if (j < ORBX_EDGE) row[-(j + 1)] = row[j + 1];
    else { const int k = j - ORBX_EDGE + 1; row[w - 1 + k] = row[w - 1 - k]; }
}

void pyramid_fill_border(const LevelGeom& L, int frame, cudaStream_t s)
{
    uint8_t* plane = L.pyr + (long long)frame * L.plane_bytes;
    pyr_border_rows_kernel<<<dim3((L.w + 255) / 256, 2 * ORBX_EDGE), 256, 0, s>>>(plane, L.pitch, L.w, L.h);
    pyr_border_cols_kernel<<<(L.prow * 2 * ORBX_EDGE + 255) / 256, 256, 0, s>>>(plane, L.pitch, L.w, L.prow);
}

// fast path of the resize: interior words and the thin halo (PYR_HALO rows, one word per side), taps from the 8 source bytes starting at the first
// pixel's left tap (3 aligned word loads + 2 funnel shifts per source row, PRMT + IDP.2A per pixel)
struct HRow4 { int h[4]; };
__device__ __forceinline__ HRow4 pyr_hrow_fast(const uint8_t* __restrict__ S, const ResizeWord& t, const uint32_t (&sel)[4])
{
    const uint32_t* p = reinterpret_cast<const uint32_t*>(S + t.wb);
    const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
    const uint32_t lo = __funnelshift_r(w0, w1, t.shift), hi = __funnelshift_r(w1, w2, t.shift);
    HRow4 r;
#pragma unroll
    for (int k = 0; k < 4; k++)
        r.h[k] = (int)(__dp2a_lo(t.a01[k], __byte_perm(lo, hi, sel[k]), 0u) >> 4);
    return r;
}

// PYRF_ROWS output rows per thread: 8 for batches (one horizontal pass serves ~1.2 output rows), 2 when a launch holds
// only a few frames - there the serial chain of dependent row loads per thread IS the launch's duration (8 rows: 8-10 us
// per level for a single 752x480 frame, 2 rows: a third of that), and the redundant horizontal passes cost nothing
// PF: the warp first asks L2 for the ~11 source rows its 8 output rows will read (prefetch.global.L2, 6 sectors per row):
// the demand loads of the row loop then pay L2 instead of DRAM latency without holding registers for the rows in flight
// (pyramid stage 0.66 -> 0.58 ms per 512 frames).  Tried and dropped: all source rows of the 8 output rows requested up
// front into registers and consumed in source order (73 registers, 3 CTAs per SM): 0.69 ms - the kernel is as much
// issue- as latency-bound; a 4-deep unroll of the row loop: no change.
template <int PYRF_ROWS, bool PF>
__global__ void __launch_bounds__(256)
pyr_resize_fast_kernel(const uint8_t* __restrict__ prev, int ppitch, long long pplane_bytes,
                       uint8_t* __restrict__ out, int pitch, long long plane_bytes, int w, int h,
                       const ResizeWord* __restrict__ tapw, const ResizeTapY* __restrict__ tapy)
{
    orbx_pdl_prologue();
    const int wi = ORBX_XOFF / 4 - 1 + blockIdx.x * blockDim.x + threadIdx.x; // interior words + one halo word per side
    if (wi > (ORBX_XOFF + w + 3) / 4) return;
    const int Y0 = (blockIdx.y * blockDim.y + threadIdx.y) * PYRF_ROWS - PYR_HALO;   // image row (interior + halo rows)
    if (Y0 >= h + PYR_HALO) return;
    const int f = blockIdx.z;
    ResizeWord t;
    {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(tapw + wi)), b = __ldg(reinterpret_cast<const uint4*>(tapw + wi) + 1);
        t.wb = (int)a.x; t.shift = a.y; t.sels = a.z; t.a01[0] = a.w; t.a01[1] = b.x; t.a01[2] = b.y; t.a01[3] = b.z;
    }
    uint32_t sel[4];
#pragma unroll
    for (int k = 0; k < 4; k++) sel[k] = (t.sels >> (8 * k)) & 0xFFu;
    const uint8_t* base = prev + (long long)f * pplane_bytes + (long long)ORBX_YOFF * ppitch + ORBX_XOFF;
    uint32_t* dst = reinterpret_cast<uint32_t*>(out + (long long)f * plane_bytes + (long long)(ORBX_YOFF + Y0) * pitch) + wi;
    const int pw = pitch >> 2;
    if (PF) {
        const int sb = tapy[ORBX_EDGE + max(Y0, 0)].s0;                    // first source row of the interior rows
        const int wb0 = __shfl_sync(__activemask(), t.wb, 0) & ~31;        // lane 0 holds the warp's leftmost word
        const uint8_t* pbase = base + (long long)sb * ppitch + wb0;
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const int idx = lane + 32 * j, row = idx / 6, sec = idx - row * 6;
            if (row < 11) asm volatile("prefetch.global.L2 [%0];" :: "l"(pbase + (long long)row * ppitch + sec * 32));
        }
    }
    int have = -1;
    HRow4 hl;
#pragma unroll
    for (int k = 0; k < 4; k++) hl.h[k] = 0;
    const int nr = min(PYRF_ROWS, h + PYR_HALO - Y0);
#pragma unroll 2
    for (int rr = 0; rr < nr; rr++) {
        // the tap table is indexed by padded row; one 8-byte load (source rows and coefficients are non-negative)
        const uint2 tv = __ldg(reinterpret_cast<const uint2*>(tapy + ORBX_EDGE + Y0) + rr);
        const int s0 = (int)(tv.x & 0xFFFFu), s1 = (int)(tv.x >> 16);
        const HRow4 h0 = (s0 == have) ? hl : pyr_hrow_fast(base + (uint32_t)(s0 * ppitch), t, sel);
        const HRow4 h1 = (s1 == s0) ? h0 : pyr_hrow_fast(base + (uint32_t)(s1 * ppitch), t, sel);
        const uint32_t b0 = tv.y & 0xFFFFu, b1 = tv.y >> 16;
        // ((b0*h0)>>16) + ((b1*h1)>>16) + 2 for two pixels per register: PRMT gathers the upper halves of
        // the products (each < 2^27), the sums stay below 2^11 per 16-bit lane; result byte = bits 2..9
        const uint32_t A01 = __byte_perm(b0 * (uint32_t)h0.h[0], b0 * (uint32_t)h0.h[1], 0x7632);
        const uint32_t A23 = __byte_perm(b0 * (uint32_t)h0.h[2], b0 * (uint32_t)h0.h[3], 0x7632);
        const uint32_t B01 = __byte_perm(b1 * (uint32_t)h1.h[0], b1 * (uint32_t)h1.h[1], 0x7632);
        const uint32_t B23 = __byte_perm(b1 * (uint32_t)h1.h[2], b1 * (uint32_t)h1.h[3], 0x7632);
        const uint32_t S01 = A01 + B01 + 0x00020002u, S23 = A23 + B23 + 0x00020002u;
        dst[rr * pw] = __byte_perm(S01 >> 2, S23 >> 2, 0x6420);
        hl = h1; have = s1;
    }
}

// ---- TMA variant of the fast resize (default; ORBX_PYR_TMA=0 selects the global-memory kernel above) ----
// Same thread mapping and arithmetic as pyr_resize_fast_kernel<8, *>, but the source rows of the CTA's 64 output rows x 128
// pixels come from ONE cp.async.bulk.tensor.2d tile (SASS: UTMALDG) of PT_ROWS x PT_BYTES behind an mbarrier: the row loop
// reads shared memory (3 LDS per source row) instead of global memory, the L2 prefetch and the 64-bit address arithmetic go.
#include <cuda.h>
// PT_ROWS / PT_BYTES (orbx_internal.h): the tile holds the sources of 64 rows x 32 words at scale factors up to ~1.25; the host
// checks every level's tap tables against them (LevelGeom::pyr_tma_ok) and falls back to the global-memory kernel otherwise
struct PyrMaps { CUtensorMap map[ORBX_MAX_LEVELS]; };

__device__ __forceinline__ HRow4 pyr_hrow_smem(const uint8_t* S, uint32_t shift, const uint32_t (&a01)[4], const uint32_t (&sel)[4])
{
    const uint32_t* p = reinterpret_cast<const uint32_t*>(S);
    const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
    const uint32_t lo = __funnelshift_r(w0, w1, shift), hi = __funnelshift_r(w1, w2, shift);
    HRow4 r;
#pragma unroll
    for (int k = 0; k < 4; k++)
        r.h[k] = (int)(__dp2a_lo(a01[k], __byte_perm(lo, hi, sel[k]), 0u) >> 4);
    return r;
}

__global__ void __launch_bounds__(256)
pyr_resize_tma_kernel(const __grid_constant__ PyrMaps M, int lvl_src, int slot0, int src_prow,
                      uint8_t* __restrict__ out, int pitch, long long plane_bytes, int w, int h,
                      const ResizeWord* __restrict__ tapw, const ResizeTapY* __restrict__ tapy)
{
    __shared__ __align__(128) uint8_t tile[PT_ROWS * PT_BYTES];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ int s_org[2];
    orbx_pdl_prologue();
    const int lane = threadIdx.x, ty = threadIdx.y;
    const int wi = ORBX_XOFF / 4 - 1 + blockIdx.x * 32 + lane;             // interior words + one halo word per side
    const bool colvalid = wi <= (ORBX_XOFF + w + 3) / 4;
    const int Y0c = blockIdx.y * 64 - PYR_HALO, Y0 = Y0c + ty * 8;
    const int f = blockIdx.z;
    ResizeWord t;
    {
        const int wl = colvalid ? wi : ORBX_XOFF / 4;                        // any valid table row
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(tapw + wl)), b = __ldg(reinterpret_cast<const uint4*>(tapw + wl) + 1);
        t.wb = (int)a.x; t.shift = a.y; t.sels = a.z; t.a01[0] = a.w; t.a01[1] = b.x; t.a01[2] = b.y; t.a01[3] = b.z;
    }
    if (ty == 0) {
        // tile origin: leftmost tap word of the warp (16-byte aligned down: TMA box start), first source row of the CTA's rows
        const int wbmin = __reduce_min_sync(0xffffffffu, colvalid ? t.wb : 0x7fffffff);
        if (lane == 0) {
            const int xs = (ORBX_XOFF + wbmin) & ~15;                       // byte column inside the padded plane row
            // first source row: the CTA's first image row, or - bottom CTA - the reflected halo row h + 2 -> h - 4 (the top halo reflects downwards)
            const int ys = min((int)tapy[ORBX_EDGE + max(Y0c, 0)].s0, (int)tapy[ORBX_EDGE + min(Y0c + 63, h + PYR_HALO - 1)].s0);
            s_org[0] = xs - ORBX_XOFF; s_org[1] = ys;
            const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mb));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"((uint32_t)(PT_ROWS * PT_BYTES)) : "memory");
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         :: "r"((uint32_t)__cvta_generic_to_shared(tile)), "l"(reinterpret_cast<uint64_t>(&M.map[lvl_src])),
                            "r"(xs), "r"((slot0 + f) * src_prow + ORBX_YOFF + ys), "r"(mb) : "memory");
        }
    }
    __syncthreads();                                                        // barrier initialised, origin published
    if (!colvalid || Y0 >= h + PYR_HALO) return;
    uint32_t sel[4];
#pragma unroll
    for (int k = 0; k < 4; k++) sel[k] = (t.sels >> (8 * k)) & 0xFFu;
    const uint8_t* base = tile + (t.wb - s_org[0]) - s_org[1] * PT_BYTES;   // source row s -> base + s * PT_BYTES
    uint32_t* dst = reinterpret_cast<uint32_t*>(out + (long long)f * plane_bytes + (long long)(ORBX_YOFF + Y0) * pitch) + wi;
    const int pw = pitch >> 2;
    {
        const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(mb), "r"(0u) : "memory");
    }
    int have = -1;
    HRow4 hl;
#pragma unroll
    for (int k = 0; k < 4; k++) hl.h[k] = 0;
    const int nr = min(8, h + PYR_HALO - Y0);
#pragma unroll 2
    for (int rr = 0; rr < nr; rr++) {
        const uint2 tv = __ldg(reinterpret_cast<const uint2*>(tapy + ORBX_EDGE + Y0) + rr);
        const int s0 = (int)(tv.x & 0xFFFFu), s1 = (int)(tv.x >> 16);
        const HRow4 h0 = (s0 == have) ? hl : pyr_hrow_smem(base + s0 * PT_BYTES, t.shift, t.a01, sel);
        const HRow4 h1 = (s1 == s0) ? h0 : pyr_hrow_smem(base + s1 * PT_BYTES, t.shift, t.a01, sel);
        const uint32_t b0 = tv.y & 0xFFFFu, b1 = tv.y >> 16;
        const uint32_t A01 = __byte_perm(b0 * (uint32_t)h0.h[0], b0 * (uint32_t)h0.h[1], 0x7632);
        const uint32_t A23 = __byte_perm(b0 * (uint32_t)h0.h[2], b0 * (uint32_t)h0.h[3], 0x7632);
        const uint32_t B01 = __byte_perm(b1 * (uint32_t)h1.h[0], b1 * (uint32_t)h1.h[1], 0x7632);
        const uint32_t B23 = __byte_perm(b1 * (uint32_t)h1.h[2], b1 * (uint32_t)h1.h[3], 0x7632);
        const uint32_t S01 = A01 + B01 + 0x00020002u, S23 = A23 + B23 + 0x00020002u;
        dst[rr * pw] = __byte_perm(S01 >> 2, S23 >> 2, 0x6420);
        hl = h1; have = s1;
    }
}

typedef CUresult (*pyr_tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                       const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                       CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

const void* pyr_tma_create(const FrameGeom& G)
{
    static const bool want = [] { const char* e = getenv("ORBX_PYR_TMA"); return !e || atoi(e) != 0; }();   // default on: 0.544 -> 0.463 ms per 512 C1 frames
    if (!want) return nullptr;
    static pyr_tmap_encode_fn enc = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
        return (pyr_tmap_encode_fn)fn;
    }();
    if (!enc) return nullptr;
    PyrMaps* M = new PyrMaps();
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        const cuuint64_t dims[2] = {(cuuint64_t)L.pitch, (cuuint64_t)L.prow * (cuuint64_t)G.batch};
        const cuuint64_t strides[1] = {(cuuint64_t)L.pitch};
        const cuuint32_t box[2] = {(cuuint32_t)PT_BYTES, (cuuint32_t)PT_ROWS};
        const cuuint32_t estr[2] = {1u, 1u};
        if (enc(&M->map[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)L.pyr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            delete M;
            return nullptr;
        }
    }
    return M;
}

void pyr_tma_destroy(const void* maps) { delete static_cast<const PyrMaps*>(maps); }

// level l>0 from level l-1.  OpenCV HResizeLinear/VResizeLinear u8 arithmetic (see oracle/orc_prims.c):
//   H = S[s0]*a0 + S[s1]*a1 ;  dst = (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2
// A thread owns one 32-bit word (4 px) of PYR_ROWS consecutive padded rows: the 4 horizontal taps are
// loaded once, and a horizontal row result is reused when the next output row's upper source row is
// this row's lower one (scale 1.2: most of the time), i.e. ~1.2 horizontal passes per output row.
#define PYR_ROWS 8
#define PYR_BX 64
#define PYR_BY 4

struct HRow { int h[4]; };     // (S[s0]*a0 + S[s1]*a1) >> 4 of the thread's 4 pixels

__device__ __forceinline__ HRow pyr_hrow(const uint8_t* __restrict__ S, const int (&s0)[4], const int (&s1)[4],
                                         const int (&a0)[4], const int (&a1)[4])
{
    HRow r;
#pragma unroll
    for (int k = 0; k < 4; k++) r.h[k] = ((int)__ldg(S + s0[k]) * a0[k] + (int)__ldg(S + s1[k]) * a1[k]) >> 4;
    return r;
}

__global__ void __launch_bounds__(PYR_BX * PYR_BY)
pyr_resize_kernel(const uint8_t* __restrict__ prev, int ppitch, long long pplane_bytes,
                  uint8_t* __restrict__ out, int pitch, long long plane_bytes, int prow,
                  const ResizeTapX* __restrict__ tapx, const ResizeTapY* __restrict__ tapy)
{
    const int wi = blockIdx.x * PYR_BX + threadIdx.x;
    if (wi * 4 >= pitch) return;
    const int Y0 = (blockIdx.y * PYR_BY + threadIdx.y) * PYR_ROWS;
    if (Y0 >= prow) return;
    const int f = blockIdx.z;
    int s0[4], s1[4], a0[4], a1[4];
    {
        const uint4 t01 = __ldg(reinterpret_cast<const uint4*>(tapx + wi * 4));
        const uint4 t23 = __ldg(reinterpret_cast<const uint4*>(tapx + wi * 4) + 1);
        const uint32_t w[8] = {t01.x, t01.y, t01.z, t01.w, t23.x, t23.y, t23.z, t23.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            s0[k] = (int)(w[2 * k] & 0xFFFF); s1[k] = (int)(w[2 * k] >> 16);
            a0[k] = (int)(w[2 * k + 1] & 0xFFFF); a1[k] = (int)(w[2 * k + 1] >> 16);
        }
    }
    const uint8_t* base = prev + (long long)f * pplane_bytes + (long long)ORBX_YOFF * ppitch + ORBX_XOFF;
    uint8_t* dst = out + (long long)f * plane_bytes + (long long)Y0 * pitch + wi * 4;
    int have = -1;              // source row whose horizontal pass is held in hl
    HRow hl;
#pragma unroll
    for (int k = 0; k < 4; k++) hl.h[k] = 0;
    const int nr = min(PYR_ROWS, prow - Y0);
#pragma unroll 2
    for (int rr = 0; rr < nr; rr++) {
        const ResizeTapY ty = tapy[Y0 + rr];
        const HRow h0 = (ty.s0 == have) ? hl : pyr_hrow(base + (long long)ty.s0 * ppitch, s0, s1, a0, a1);
        const HRow h1 = (ty.s1 == ty.s0) ? h0 : pyr_hrow(base + (long long)ty.s1 * ppitch, s0, s1, a0, a1);
        const int b0 = ty.b0, b1 = ty.b1;
        uint32_t word = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int v = (((b0 * h0.h[k]) >> 16) + ((b1 * h1.h[k]) >> 16) + 2) >> 2;
            word |= (uint32_t)v << (8 * k);        // v <= 255 by construction
        }
        *reinterpret_cast<uint32_t*>(dst + (long long)rr * pitch) = word;
        hl = h1; have = ty.s1;
    }
}

// experiment hook: ORBX_PYR_ROWS=2 / 8 forces one variant for every batch size
static int pyr_rows_override()
{
    static const int v = [] { const char* e = getenv("ORBX_PYR_ROWS"); return e ? atoi(e) : 0; }();
    return v;
}

int launch_pyramid_level(const FrameGeom& G, const DevBuffers& D, int nframes, int l, cudaStream_t s)
{
    {
        const LevelGeom& L = G.L[l];
        if (l == 0) {
            const dim3 block(64, 4);
            const dim3 grid(((L.w + 15) / 16 + block.x - 1) / block.x, (L.h + 2 * PYR_HALO + block.y - 1) / block.y, nframes);
            const uint8_t* in = D.zc_input ? D.zc_input : D.input;
            const int ip = D.zc_input ? D.zc_pitch : D.in_pitch;
            const long long ifb = D.zc_input ? D.zc_frame_bytes : (long long)D.in_pitch * G.rows;
            if ((ip & 15) == 0 && (ifb & 15) == 0 && ((uintptr_t)in & 15) == 0)
                orbx_launch(pyr_level0_kernel<true>, grid, block, 0, s, in, ip, ifb,
                            L.pyr, L.pitch, L.plane_bytes, L.w, L.h, D.candcnt, G.nlevels);
            else
                orbx_launch(pyr_level0_kernel<false>, grid, block, 0, s, in, ip, ifb,
                            L.pyr, L.pitch, L.plane_bytes, L.w, L.h, D.candcnt, G.nlevels);
        } else if (L.tapw) {
            const LevelGeom& P = G.L[l - 1];
            const dim3 block(32, 8);        // one warp per 32 words: the partial warp at the right edge wastes less than a 64-wide block
            const int gx = ((L.w + 3) / 4 + 2 + block.x - 1) / block.x;
            const int gy8 = (L.h + 2 * PYR_HALO + block.y * 8 - 1) / (block.y * 8);
            if (D.pyr_maps && L.pyr_tma_ok && (long long)gx * gy8 * nframes >= 4 * 148) {
                orbx_launch(pyr_resize_tma_kernel, dim3(gx, gy8, nframes), block, 0, s, *static_cast<const PyrMaps*>(D.pyr_maps), l - 1, D.slot0, P.prow,
                            L.pyr, L.pitch, L.plane_bytes, L.w, L.h, L.tapw, L.tapy);
            } else if (pyr_rows_override() ? pyr_rows_override() == 8 : (long long)gx * gy8 * nframes >= 4 * 148) {
                static const bool pf = [] { const char* e = getenv("ORBX_PYR_PF"); return !e || atoi(e) != 0; }();
                orbx_launch(pf ? pyr_resize_fast_kernel<8, true> : pyr_resize_fast_kernel<8, false>, dim3(gx, gy8, nframes), block, 0, s, P.pyr, P.pitch, P.plane_bytes, L.pyr, L.pitch,
                                                                                  L.plane_bytes, L.w, L.h, L.tapw, L.tapy);
            } else {
                const int gy2 = (L.h + 2 * PYR_HALO + block.y * 2 - 1) / (block.y * 2);
                orbx_launch(pyr_resize_fast_kernel<2, false>, dim3(gx, gy2, nframes), block, 0, s, P.pyr, P.pitch, P.plane_bytes, L.pyr, L.pitch,
                                                                                  L.plane_bytes, L.w, L.h, L.tapw, L.tapy);
            }
        } else {
            // generic path (taps of a word do not fit 8 source bytes); it also writes the border
            const LevelGeom& P = G.L[l - 1];
            const dim3 rblock(PYR_BX, PYR_BY);
            const dim3 rgrid((L.pitch / 4 + PYR_BX - 1) / PYR_BX, (L.prow + PYR_BY * PYR_ROWS - 1) / (PYR_BY * PYR_ROWS), nframes);
            pyr_resize_kernel<<<rgrid, rblock, 0, s>>>(P.pyr, P.pitch, P.plane_bytes, L.pyr, L.pitch, L.plane_bytes,
                                                       L.prow, L.tapx, L.tapy);
        }
    }
    return 1;
}

int launch_pyramid(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    int launches = 0;
    for (int l = 0; l < G.nlevels; l++) launches += launch_pyramid_level(G, D, nframes, l, s);
    return launches;
}
