// describe.cu — output assembly, 7x7 Gaussian blur, IC_Angle orientation and rBRIEF-256 descriptors.
// Reference: src/ORBextractor.cc  operator() :1534-1659, computeOrientation/IC_Angle :91-138,580-591,
// GaussianBlur call :1606-1614, computeOrbDescriptor :150-203.
#include "orbx_internal.h"

// umax of the 31x31 circular patch (src/ORBextractor.cc:542-570); depends only on HALF_PATCH_SIZE=15
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

// ------------------------------------------------------------------------------------------------
// assemble: final row of every selected keypoint (mono ascending / lapping-area descending,
// src/ORBextractor.cc:1589,1633-1655) + all KeyPoint fields except the angle.
// ------------------------------------------------------------------------------------------------
#define ASM_THREADS 512

// One CTA per frame walks the selected keypoints of all levels as ONE list (level-major = the reference's order) in
// chunks of ASM_THREADS: a frame of 1000 keypoints is two chunks (two dependent global loads deep), not one chunk per level.
__global__ void __launch_bounds__(ASM_THREADS)
assemble_kernel(FrameGeom G, DevBuffers D, float lap0, float lap1)
{
    __shared__ int s_warp[ASM_THREADS / 32];
    __shared__ int s_pref[ORBX_MAX_LEVELS + 1];
    orbx_pdl_prologue();
    const int f = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) {
        const int* selcnt = D.selcnt + f * G.nlevels;
        int a = 0;
        for (int l = 0; l < G.nlevels; l++) { s_pref[l] = a; a += selcnt[l]; }
        s_pref[G.nlevels] = a;
    }
    __syncthreads();
    const int total = s_pref[G.nlevels];
    int nLap = 0, nMono = 0;       // running counts in reference order
    orbx_keypoint* kps = D.kps + (long long)f * G.out_cap;
    for (int base = 0; base < total; base += ASM_THREADS) {
        const int g = base + tid;
        const int valid = g < total;
        float x = 0.f, y = 0.f;
        orbx_key_t key = 0;
        int flag = 0, l = 0, i = 0;
        if (valid) {
            while (g >= s_pref[l + 1]) l++;
            i = g - s_pref[l];
            const LevelGeom& L = G.L[l];
            key = D.sel[(long long)f * G.sel_per_frame + L.sel_base + i];
            x = (float)(KEY_X(key) + ORBX_MINB);                 // :1167
            y = (float)(KEY_Y(key) + ORBX_MINB);
            if (l != 0) { x = __fmul_rn(x, L.scale); y = __fmul_rn(y, L.scale); }   // :1639-1642
            flag = (x >= lap0 && x <= lap1);                      // :1644
        }
        // block exclusive scan of flag
        const unsigned bal = __ballot_sync(0xffffffffu, flag);
        const int lane = tid & 31, warp = tid >> 5;
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = __popc(bal & ((1u << lane) - 1)), tot = 0;
#pragma unroll
        for (int w = 0; w < ASM_THREADS / 32; w++) { const int v = s_warp[w]; if (w < warp) before += v; tot += v; }
        const int nvalid = min(ASM_THREADS, total - base);
        if (valid) {
            const LevelGeom& L = G.L[l];
            const int lapBefore = nLap + before;
            const int monoBefore = nMono + tid - before;
            const int dst = flag ? (total - 1 - lapBefore) : monoBefore;
            D.outidx[(long long)f * G.sel_per_frame + L.sel_base + i] = dst;
            orbx_keypoint kp;
            kp.x = x; kp.y = y; kp.size = L.patch_size; kp.angle = -1.f;
            kp.response = (float)KEY_S(key); kp.octave = l; kp.class_id = -1;
            kps[dst] = kp;
        }
        nLap += tot; nMono += nvalid - tot;
        __syncthreads();
    }
    if (tid == 0) { D.counts[2 * f] = total; D.counts[2 * f + 1] = nMono; }
}

int launch_assemble(const FrameGeom& G, const DevBuffers& D, int nframes, int lap0, int lap1, cudaStream_t s)
{
    orbx_launch(assemble_kernel, dim3(nframes), dim3(ASM_THREADS), 0, s, G, D, (float)lap0, (float)lap1);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// GaussianBlur(7x7, sigma 2) in OpenCV 4.x fixed point: kernel [18,34,48,56,48,34,18]/256, horizontal
// pass in 8.8 (u16), vertical in 16.16, +0.5 rounding (oracle/orc_prims.c orc_gaussian7x7_s2).
// Reads the pyramid plane with its thin reflect-101 halo (written by the pyramid kernels), writes the
// interior of the blur plane.  HBM-bound: algorithmic bytes per frame = read + write sum(L0..L7).
// ------------------------------------------------------------------------------------------------
// The two passes are exact integer sums without intermediate rounding, so they commute: the VERTICAL
// pass runs first, on raw bytes.  A thread owns one 32-bit word (4 px) of a 32-row strip; the word is
// split into its even / odd pixels as u16x2 lanes (E = px 0,2; O = px 1,3) and the 7-tap column sum
// (<= 255*256, fits 16 bits) is 7 IMAD per lane PAIR (SWAR).  The horizontal pass then needs the column
// sums of the neighbouring words (4 SHFL) and is 18 IDP.2A (u16 x u8 dot products) per word; byte 2 of
// each 32-bit accumulator (started at 32768) is the rounded result.  A warp produces 30 words per row:
// lanes 0 and 31 only supply the halo column sums.
#define BL_RT 32
#define BL_BY 4
#define BL_OUTW 30            // output words per warp row
#ifndef BL_MINB
#define BL_MINB 8              // CTAs per SM the register allocation is held to (64 registers)
#endif
#define HALO_ROWS_BLUR 3      // rows of reflect-101 halo the pyramid planes carry (written by the pyramid kernels)

template <int PH>
__device__ __forceinline__ void blur_vcol(const uint32_t (&WE)[7], const uint32_t (&WO)[7], uint32_t& VE, uint32_t& VO)
{
    VE = 18u * WE[PH % 7] + 34u * WE[(PH + 1) % 7] + 48u * WE[(PH + 2) % 7] + 56u * WE[(PH + 3) % 7] +
         48u * WE[(PH + 4) % 7] + 34u * WE[(PH + 5) % 7] + 18u * WE[(PH + 6) % 7];
    VO = 18u * WO[PH % 7] + 34u * WO[(PH + 1) % 7] + 48u * WO[(PH + 2) % 7] + 56u * WO[(PH + 3) % 7] +
         48u * WO[(PH + 4) % 7] + 34u * WO[(PH + 5) % 7] + 18u * WO[(PH + 6) % 7];
}

// E = (V[x], V[x+2]), O = (V[x+1], V[x+3]) of the thread's word, l / r suffix: the words left / right of it.
__device__ __forceinline__ uint32_t blur_hword(uint32_t El, uint32_t Ol, uint32_t E, uint32_t O, uint32_t Er, uint32_t Or)
{
#define KK(lo, hi) ((uint32_t)(lo) | ((uint32_t)(hi) << 8))
    const uint32_t rnd = 32768u;
    const uint32_t a0 = __dp2a_lo(Ol, KK(18, 48), __dp2a_lo(El, KK(0, 34), __dp2a_lo(E, KK(56, 34), __dp2a_lo(O, KK(48, 18), rnd))));
    const uint32_t a1 = __dp2a_lo(El, KK(0, 18), __dp2a_lo(Ol, KK(0, 34), __dp2a_lo(E, KK(48, 48), __dp2a_lo(O, KK(56, 34), __dp2a_lo(Er, KK(18, 0), rnd)))));
    const uint32_t a2 = __dp2a_lo(Ol, KK(0, 18), __dp2a_lo(E, KK(34, 56), __dp2a_lo(O, KK(48, 48), __dp2a_lo(Er, KK(34, 0), __dp2a_lo(Or, KK(18, 0), rnd)))));
    const uint32_t a3 = __dp2a_lo(E, KK(18, 48), __dp2a_lo(O, KK(34, 56), __dp2a_lo(Er, KK(48, 18), __dp2a_lo(Or, KK(34, 0), rnd))));
#undef KK
    // (acc >> 16) <= 255 is byte 2 of each accumulator
    return __byte_perm(__byte_perm(a0, a1, 0x0062), __byte_perm(a2, a3, 0x0062), 0x5410);
}

// One warp per (120-px column block, 32-row strip); no shared memory, no barriers, one straight code
// path: BORDER_REFLECT_101 is already in the plane (the 3-row / one-word halo written by
// the pyramid kernels), so the first / last strips and the edge column blocks read like any other.
// Rows are fetched 7 stream rows ahead into statically indexed registers (slot = stream row % 7, the
// same rotation as the E/O window).  A slot's old value is consumed (unpacked) BEFORE the load that
// refills it is issued — the refill's address carries a fake dependency on the unpacked value — so the
// load lands in the slot's own register and nothing waits for it until 7 rows later.
template <bool PF>
__global__ void __launch_bounds__(32 * BL_BY, BL_MINB)
blur_kernel(FrameGeom G)
{
    orbx_pdl_prologue();
    const int lane = threadIdx.x, warp = threadIdx.y;
    // all levels in one launch: blockIdx.x enumerates (level, row-strip group, column block)
    int lvl = 0, bid = blockIdx.x;
    while (bid >= G.blur_end[lvl]) lvl++;
    if (lvl > 0) bid -= G.blur_end[lvl - 1];
    const LevelGeom& L = G.L[lvl];
    const int pitch = L.pitch, w = L.w, h = L.h;
    const int nbx = (w + 4 * BL_OUTW - 1) / (4 * BL_OUTW);
    const int by = bid / nbx, bx = bid - by * nbx;
    const int wi = bx * BL_OUTW - 1 + lane;                    // lane 0 / 31: halo words
    const int y0 = (by * BL_BY + warp) * BL_RT;
    if (y0 >= h) return;                                       // warp-uniform
    const int pw = pitch >> 2;
    const int nr = min(BL_RT, h - y0);
    const long long plane = (long long)blockIdx.y * L.plane_bytes + (long long)ORBX_YOFF * pitch + ORBX_XOFF;
    // lanes past the right halo word read a valid word and drop the result
    const uint32_t* nxt = reinterpret_cast<const uint32_t*>(L.pyr + plane) + (long long)(y0 - HALO_ROWS_BLUR) * pw + min(wi, (w + 3) >> 2);
    uint32_t* q = reinterpret_cast<uint32_t*>(L.blur + plane) + (long long)y0 * pw + wi;
    const bool colvalid = lane >= 1 && lane <= BL_OUTW && wi * 4 < w;
    uint32_t RA[7], WE[7], WO[7];
    if (PF) {
        // L2 prefetch of rows 7.. of the strip (the first 7 are demanded right below): lane = row, 4 sectors of the
        // warp's 128-byte row; the register ring then refills from L2 (blur stage 0.452 -> 0.442 ms per 512 frames)
        const uint32_t* rowp = nxt - min(wi, (w + 3) >> 2) + (bx * BL_OUTW - 1) + (long long)(7 + lane) * pw;
        if (7 + lane < nr + 2 * HALO_ROWS_BLUR) {
#pragma unroll
            for (int sct = 0; sct < 4; sct++) asm volatile("prefetch.global.L2 [%0];" :: "l"(rowp + 8 * sct));
        }
    }
#pragma unroll
    for (int i = 0; i < 7; i++) { RA[i] = __ldg(nxt); nxt += pw; }     // may run a few rows past the strip, inside the plane
#define BLUR_ROW(SLOT)                                                                    \
    {   uint32_t e, o;                                                                    \
        asm("prmt.b32 %0, %2, 0, 0x4240;\n\tprmt.b32 %1, %2, 0, 0x4341;" : "=r"(e), "=r"(o) : "r"(RA[SLOT])); \
        WE[SLOT] = e; WO[SLOT] = o;                                                       \
        RA[SLOT] = __ldg(nxt + (o >> 24)); nxt += pw; }
    BLUR_ROW(0) BLUR_ROW(1) BLUR_ROW(2) BLUR_ROW(3) BLUR_ROW(4) BLUR_ROW(5)
#define BLUR_STEP(PH)                                                   \
    {   /* unguarded: rows past nr are computed and dropped, so the body is straight-line code */ \
        BLUR_ROW((PH + 6) % 7)                                          \
        uint32_t VE, VO;                                                \
        blur_vcol<PH>(WE, WO, VE, VO);                                  \
        const uint32_t El = __shfl_up_sync(0xffffffffu, VE, 1), Ol = __shfl_up_sync(0xffffffffu, VO, 1);     \
        const uint32_t Er = __shfl_down_sync(0xffffffffu, VE, 1), Or = __shfl_down_sync(0xffffffffu, VO, 1); \
        const uint32_t vw = blur_hword(El, Ol, VE, VO, Er, Or);         \
        if (colvalid && rb + PH < nr) *q = vw;                          \
        q += pw;                                                        \
    }
#pragma unroll 1
    for (int rb = 0; rb < nr; rb += 7) {
        BLUR_STEP(0) BLUR_STEP(1) BLUR_STEP(2) BLUR_STEP(3) BLUR_STEP(4) BLUR_STEP(5) BLUR_STEP(6)
    }
#undef BLUR_STEP
#undef BLUR_ROW
}

// ------------------------------------------------------------------------------------------------
// TMA variant of the blur (north_star: "staged through shared memory/TMA tiles"): the same arithmetic, but the
// CTA's input tile - 128 output rows + 6 halo rows of the 128-byte column block - arrives in shared memory by ONE
// cp.async.bulk.tensor.2d issued by one thread (SASS: UTMALDG) behind an mbarrier, instead of 32 lanes x 38 LDG per warp
// through a 7-deep register ring.  What it removes: the ring registers, the L2 prefetches and the per-row global
// address arithmetic; what stays: one LDS per word-row where there was one LDG, and the whole ALU work.
// Opt-in (ORBX_BLUR_TMA=1): measured on a B200 in DESIGN.md section 5.
// ------------------------------------------------------------------------------------------------
#include <cuda.h>
#define BT_ROWS (BL_BY * BL_RT + 2 * HALO_ROWS_BLUR)      // 134 tile rows
#define BT_WORDS 36                                       // 144-byte tile rows: 128 bytes + up to 12 of alignment slack
struct BlurMaps { CUtensorMap map[ORBX_MAX_LEVELS]; };

// tile t of a launch = (frame, level, row group, column block); input box coordinates and the level
__device__ __forceinline__ int blur_tma_tile(const FrameGeom& G, int t, int per_frame, int slot0, int* f, int* bx, int* y0c, int* cx, int* cy)
{
    int bid = t, lvl = 0;
    if (per_frame > 0) { *f = t / per_frame; bid = t - *f * per_frame; }      // persistent form: linear tile index
    else *f = blockIdx.y;                                                      // one tile per CTA: grid = (tiles per frame, frames)
    while (bid >= G.blur_tma_end[lvl]) lvl++;
    if (lvl > 0) bid -= G.blur_tma_end[lvl - 1];
    const LevelGeom& L = G.L[lvl];
    const int nbx = (L.w + 4 * BL_OUTW - 1) / (4 * BL_OUTW);
    const int by = bid / nbx;
    *bx = bid - by * nbx;
    *y0c = by * BL_BY * BL_RT;
    // the box must start on a 16-byte address (an unaligned start traps with "illegal instruction",
    // tools/scratch/tma_probe2.cu): start at the aligned address below the left halo word, 144 bytes wide
    *cx = (ORBX_XOFF + *bx * 4 * BL_OUTW - 4) & ~15;
    *cy = (slot0 + *f) * L.prow + ORBX_YOFF + *y0c - HALO_ROWS_BLUR;          // row in the stack of planes
    return lvl;
}

// STAGES = 1: one tile per CTA (grid = tiles).  STAGES = 2: persistent CTAs walk the tiles with stride gridDim.x and a
// two-deep shared-memory ring - thread 0 issues the copy of tile k+1 before the CTA computes tile k.
template <int STAGES>
__global__ void __launch_bounds__(32 * BL_BY, STAGES == 1 ? BL_MINB : 5)
blur_tma_kernel(FrameGeom G, const __grid_constant__ BlurMaps M, int slot0, int per_frame, int total)
{
    __shared__ __align__(128) uint32_t tile[STAGES][BT_ROWS + 2][BT_WORDS];   // 136 x 144 B: every stage starts on a 128-byte address (TMA destination)
    __shared__ __align__(8) uint64_t mbar[STAGES];
    orbx_pdl_prologue();
    const int lane = threadIdx.x, warp = threadIdx.y;
    const bool leader = lane == 0 && warp == 0;
    if (leader) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((uint32_t)__cvta_generic_to_shared(&mbar[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int t, int stage) {
        int f, bx, y0c, cx, cy;
        const int lvl = blur_tma_tile(G, t, per_frame, slot0, &f, &bx, &y0c, &cx, &cy);
        const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar[stage]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"((uint32_t)(BT_ROWS * BT_WORDS * 4)) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     :: "r"((uint32_t)__cvta_generic_to_shared(&tile[stage][0][0])), "l"(reinterpret_cast<uint64_t>(&M.map[lvl])), "r"(cx), "r"(cy), "r"(mb)
                     : "memory");
    };
    if (STAGES == 1) per_frame = 0;
    int t = blockIdx.x;
    if (leader && t < total) issue(t, 0);
#pragma unroll 1
    for (int k = 0; t < total; k++, t += gridDim.x) {
        const int stage = STAGES == 1 ? 0 : (k & 1);
        if (STAGES == 2 && leader && t + (int)gridDim.x < total) issue(t + gridDim.x, stage ^ 1);   // that buffer was released by the barrier below
        int f, bx, y0c, cx, cy;
        const int lvl = blur_tma_tile(G, t, per_frame, slot0, &f, &bx, &y0c, &cx, &cy);
        const LevelGeom& L = G.L[lvl];
        {   // every thread waits for the tile; the stage's barrier completes once per use: phase = use count & 1
            const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar[stage]);
            const uint32_t phase = STAGES == 1 ? 0u : (uint32_t)((k >> 1) & 1);
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(mb), "r"(phase) : "memory");
        }
        const int pitch = L.pitch, w = L.w, h = L.h;
        const int y0 = y0c + warp * BL_RT;
        if (y0 < h) {                                          // warp-uniform
            const int wi = bx * BL_OUTW - 1 + lane;
            const int pw = pitch >> 2;
            const int nr = min(BL_RT, h - y0);
            const long long plane = (long long)f * L.plane_bytes + (long long)ORBX_YOFF * pitch + ORBX_XOFF;
            uint32_t* q = reinterpret_cast<uint32_t*>(L.blur + plane) + (long long)y0 * pw + wi;
            const bool colvalid = lane >= 1 && lane <= BL_OUTW && wi * 4 < w;
            const int woff = ((ORBX_XOFF + bx * 4 * BL_OUTW - 4) & 15) >> 2;   // words between the box start and the left halo word (1 or 3)
            const uint32_t* src = &tile[stage][warp * BL_RT][woff + lane];      // tile row 0 of this warp = image row y0 - 3
            uint32_t WE[7], WO[7];
#define BT_ROW(SLOT, R)                                                                   \
            {   const uint32_t raw = src[(R) * BT_WORDS];                                 \
                WE[SLOT] = __byte_perm(raw, 0u, 0x4240); WO[SLOT] = __byte_perm(raw, 0u, 0x4341); }
            BT_ROW(0, 0) BT_ROW(1, 1) BT_ROW(2, 2) BT_ROW(3, 3) BT_ROW(4, 4) BT_ROW(5, 5)
#define BT_STEP(PH)                                                     \
            {   BT_ROW((PH + 6) % 7, rb + PH + 6)                       \
                uint32_t VE, VO;                                        \
                blur_vcol<PH>(WE, WO, VE, VO);                          \
                const uint32_t El = __shfl_up_sync(0xffffffffu, VE, 1), Ol = __shfl_up_sync(0xffffffffu, VO, 1);     \
                const uint32_t Er = __shfl_down_sync(0xffffffffu, VE, 1), Or = __shfl_down_sync(0xffffffffu, VO, 1); \
                const uint32_t vw = blur_hword(El, Ol, VE, VO, Er, Or); \
                if (colvalid && rb + PH < nr) *q = vw;                  \
                q += pw;                                                \
            }
            // rows past the strip stay inside the tile only for rb + PH + 6 < 38: the last (partial) step group is guarded
#pragma unroll 1
            for (int rb = 0; rb < nr; rb += 7) {
                if (rb + 12 < BL_RT + 2 * HALO_ROWS_BLUR) {
                    BT_STEP(0) BT_STEP(1) BT_STEP(2) BT_STEP(3) BT_STEP(4) BT_STEP(5) BT_STEP(6)
                } else {                                       // rb = 28: rows 28..31 only (tile rows up to 37)
                    BT_STEP(0) BT_STEP(1) BT_STEP(2) BT_STEP(3)
                }
            }
#undef BT_STEP
#undef BT_ROW
        }
        if (STAGES == 2) __syncthreads();                      // every warp is done with this stage's buffer
    }
}

typedef CUresult (*orbx_tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

const void* blur_tma_create(const FrameGeom& G)
{
    static orbx_tmap_encode_fn enc = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
        return (orbx_tmap_encode_fn)fn;
    }();
    if (!enc) return nullptr;
    BlurMaps* M = new BlurMaps();
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        // the planes of all frame slots stacked: rows of `pitch` bytes, prow rows per slot
        const cuuint64_t dims[2] = {(cuuint64_t)L.pitch, (cuuint64_t)L.prow * (cuuint64_t)G.batch};
        const cuuint64_t strides[1] = {(cuuint64_t)L.pitch};
        const cuuint32_t box[2] = {(cuuint32_t)(BT_WORDS * 4), (cuuint32_t)BT_ROWS};
        const cuuint32_t estr[2] = {1u, 1u};
        if (enc(&M->map[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)L.pyr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            delete M;
            return nullptr;
        }
    }
    return M;
}

void blur_tma_destroy(const void* maps) { delete static_cast<const BlurMaps*>(maps); }

void blur_plan(FrameGeom& G)
{
    int run = 0, run_tma = 0;
    for (int l = 0; l < G.nlevels; l++) {
        const int nbx = (G.L[l].w + 4 * BL_OUTW - 1) / (4 * BL_OUTW);
        run += nbx * ((G.L[l].h + BL_BY * BL_RT - 1) / (BL_BY * BL_RT));
        G.blur_end[l] = run;
        run_tma += nbx * ((G.L[l].h + BL_BY * BL_RT - 1) / (BL_BY * BL_RT));
        G.blur_tma_end[l] = run_tma;
    }
}

int launch_blur(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    static const bool pf = [] { const char* e = getenv("ORBX_BLUR_PF"); return !e || atoi(e) != 0; }();
    // ORBX_BLUR_TMA: 0 = register-ring kernel, 1 = one TMA tile per CTA, 2 = persistent CTAs with a two-deep TMA ring
    static const int tma = [] { const char* e = getenv("ORBX_BLUR_TMA"); return e ? atoi(e) : 1; }();
    if (tma && D.blur_maps) {
        const int per_frame = G.blur_tma_end[G.nlevels - 1], total = per_frame * nframes;
        if (tma == 2) {
            const int grid = total < 148 * 5 ? total : 148 * 5;
            orbx_launch(blur_tma_kernel<2>, dim3(grid), dim3(32, BL_BY), 0, s, G, *static_cast<const BlurMaps*>(D.blur_maps), D.slot0, per_frame, total);
        } else {
            orbx_launch(blur_tma_kernel<1>, dim3(per_frame, nframes), dim3(32, BL_BY), 0, s, G, *static_cast<const BlurMaps*>(D.blur_maps), D.slot0, per_frame, per_frame);
        }
        return 1;
    }
    orbx_launch(pf ? blur_kernel<true> : blur_kernel<false>, dim3(G.blur_end[G.nlevels - 1], nframes), dim3(32, BL_BY), 0, s, G);
    (void)D;
    return 1;
}

// ------------------------------------------------------------------------------------------------
// orientation + descriptor: one warp per selected keypoint.
// ------------------------------------------------------------------------------------------------
// cv::fastAtan2 (degrees), float32, no FMA — see oracle/orc_prims.c orc_fast_atan2
__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float scale = 57.29577951308232f;                     // (float)(180/CV_PI)
    const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale);
    const float p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
    const float eps = 2.2204460492503131e-16f;                  // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// glibc sinf/cosf (generic C algorithm, double polynomial, no FMA) — see oracle/orc_prims.c orc_sinf.
__device__ __forceinline__ float sc_poly(double x, double x2, bool neg, int n)
{
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __dadd_rn(S2, __dmul_rn(x2, S3));
        const double x7 = __dmul_rn(x3, x2);
        const double s = __dadd_rn(x, __dmul_rn(x3, S1));
        return __double2float_rn(__dadd_rn(s, __dmul_rn(x7, s1)));
    } else {
        const double sg = neg ? -1.0 : 1.0;
        const double x4 = __dmul_rn(x2, x2);
        const double c2 = __dadd_rn(sg * C3, __dmul_rn(x2, sg * C4));
        const double c1 = __dadd_rn(sg * C0, __dmul_rn(x2, sg * C1));
        const double x6 = __dmul_rn(x4, x2);
        const double c = __dadd_rn(c1, __dmul_rn(x4, sg * C2));
        return __double2float_rn(__dadd_rn(c, __dmul_rn(x6, c2)));
    }
}

__device__ __forceinline__ void sincosf_glibc(float y, float* sn, float* cs)
{
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
    double x = (double)y;
    if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ff)) {       // |y| < pi/4
        const double x2 = __dmul_rn(x, x);
        if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ff)) { *sn = y; *cs = 1.0f; return; }
        *sn = sc_poly(x, x2, false, 0);
        *cs = sc_poly(x, x2, false, 1);
        return;
    }
    // reduce_fast: |y| < 120 always holds here (angle in [0, 2*pi])
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const bool neg = (n & 2) != 0;
    const double xs = x * sgn, x2 = __dmul_rn(x, x);
    *sn = sc_poly(xs, x2, neg, n);
    *cs = sc_poly(xs, x2, neg, n ^ 1);
}

#define DESC_WARPS 8
#ifndef DESC_KPW
#define DESC_KPW 1
#endif
#ifndef DESC_MINB
#define DESC_MINB 6
#endif
//#define DESC_KPW_                 // keypoint slots per warp (pattern registers and masks are set up once)
#define DESC_R 18                  // max |rotated tap| : the pattern lies inside radius 18.4
#define DESC_ROWS (2 * DESC_R + 1)
#define DESC_PW 11                 // words per patch row: 37 bytes at any alignment

__device__ __align__(16) signed char g_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};

__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c)      // sum of u8(a) * s8(b) + c
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// One warp per keypoint, DESC_KPW keypoints per warp.
//  IC_Angle (:91-138): lane = patch row v = lane - 15; the row's 31 bytes are 8 aligned-by-funnel-shift
//    words, masked to |u| <= umax[|v|], and reduced with two IDP.4A per word (sum of u*I with constant
//    s8 weights, sum of I with ones); m_10 / m_01 by warp reductions.
//  computeOrbDescriptor (:150-203): the lane's 16 tap pairs live in registers as floats; the 37x37
//    neighbourhood of the blurred level is staged in shared memory; cvRound = add 1.5 * 2^23 (round half
//    even, exact for |v| < 2^22); lane i builds byte i.
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_MINB)
describe_kernel(FrameGeom G, DevBuffers D)
{
    __shared__ uint32_t s_patch[DESC_WARPS][DESC_ROWS * DESC_PW];
    __shared__ uint32_t s_raw[DESC_WARPS][ORBX_PATCH * 9];
    orbx_pdl_prologue();
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot0 = (blockIdx.x * DESC_WARPS + warp) * DESC_KPW;
    if (slot0 >= G.sel_per_frame) return;

    // per-lane constants: tap pattern (floats) and the disc mask of this lane's patch row
    float pat[32];
    {
        const uint4* gp = reinterpret_cast<const uint4*>(g_pattern) + 2 * lane;
        const uint4 a = gp[0], b = gp[1];
        const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 32; k++) pat[k] = (float)(signed char)((w[k >> 2] >> (8 * (k & 3))) & 0xFF);
    }
    const int v = lane - ORBX_HALF_PATCH;
    uint32_t rmask[8];
    {
        const int d = lane < 31 ? c_umax[abs(v)] : -1;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            uint32_t m = 0;
#pragma unroll
            for (int b = 0; b < 4; b++) { const int u = -15 + 4 * j + b; if (abs(u) <= d) m |= 0xFFu << (8 * b); }
            rmask[j] = m;
        }
    }
    uint32_t* patch = s_patch[warp];
    uint32_t* raw = s_raw[warp];

#pragma unroll 1
    for (int kk = 0; kk < DESC_KPW; kk++) {
        const int slot = slot0 + kk;
        if (slot >= G.sel_per_frame) break;
        int lvl = 0;
#pragma unroll 1
        for (int l = 1; l < G.nlevels; l++) if (slot >= G.L[l].sel_base) lvl = l;
        const LevelGeom& L = G.L[lvl];
        if (slot - L.sel_base >= D.selcnt[f * G.nlevels + lvl]) continue;
        const orbx_key_t key = D.sel[(long long)f * G.sel_per_frame + slot];
        const int x = KEY_X(key) + ORBX_MINB, y = KEY_Y(key) + ORBX_MINB;
        const int pw = L.pitch >> 2;

        // ---- stage the 31x31 patch of the level (rows of 9 words) and the 37x37 patch of the blurred level
        // (rows of 11 words) in shared memory with row-major coalesced loads: the L1 wavefront count of a
        // keypoint is then ~1-2 per patch row instead of one per lane ----
        const int xr = ORBX_XOFF + x - ORBX_HALF_PATCH;           // byte of u = -15 in the row
        __syncwarp();                                             // previous keypoint's reads are done
        {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(
                L.pyr + (long long)f * L.plane_bytes + (long long)(ORBX_YOFF + y - ORBX_HALF_PATCH) * L.pitch + (xr & ~3));
            const int g = lane / 9, c = lane - g * 9;             // lanes 0..26: 3 rows of 9 words per iteration
            if (lane < 27) {
                const uint32_t* sp = src + g * pw + c;
                uint32_t* dp = raw + g * 9 + c;
#pragma unroll
                for (int t = 0; t < 11; t++) if (3 * t + g < ORBX_PATCH) dp[3 * t * 9] = __ldg(sp + 3 * t * pw);
            }
        }
        {
            const int xb = ORBX_XOFF + x - DESC_R;
            const uint32_t* src = reinterpret_cast<const uint32_t*>(
                L.blur + (long long)f * L.plane_bytes + (long long)(ORBX_YOFF + y - DESC_R) * L.pitch + (xb & ~3));
            // lanes 0..10 -> even rows, 11..21 -> odd rows; 19 iterations cover the 37 rows
            const int g = lane >= DESC_PW ? 1 : 0, c = lane - g * DESC_PW;
            if (lane < 2 * DESC_PW) {
                const uint32_t* sp = src + g * pw + c;
                uint32_t* dp = patch + g * DESC_PW + c;
#pragma unroll
                for (int t = 0; t < (DESC_ROWS + 1) / 2; t++) {
                    if (2 * t + g < DESC_ROWS) dp[2 * t * DESC_PW] = __ldg(sp + 2 * t * pw);
                }
            }
        }
        __syncwarp();
        // this lane's patch row: 9 conflict-free word reads (row pitch 9 words)
        uint32_t rw[9];
        {
            const uint32_t* rp = raw + (lane < 31 ? lane : 0) * 9;
#pragma unroll
            for (int j = 0; j < 9; j++) rw[j] = rp[j];
        }
        // ---- IC_Angle ----
        const int sh = (xr & 3) * 8;
        int sI = 0, sUI = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t a = __funnelshift_r(rw[j], rw[j + 1], sh) & rmask[j];     // bytes u = -15+4j .. -12+4j
            const uint32_t uw = (uint32_t)((-15 + 4 * j) & 0xFF) | ((uint32_t)((-14 + 4 * j) & 0xFF) << 8) |
                                ((uint32_t)((-13 + 4 * j) & 0xFF) << 16) | ((uint32_t)((-12 + 4 * j) & 0xFF) << 24);
            sUI = dp4a_us(a, uw, sUI);
            sI = (int)__dp4a(a, 0x01010101u, (uint32_t)sI);
        }
        const int m10 = __reduce_add_sync(0xffffffffu, sUI);
        const int m01 = __reduce_add_sync(0xffffffffu, v * sI);
        const float angle = fast_atan2_deg((float)m01, (float)m10);
        float a, b;
        sincosf_glibc(__fmul_rn(angle, 0.017453292519943295f), &b, &a);   // factorPI = (float)(CV_PI/180.f)
        // ---- descriptor byte of this lane ----
        const int xoff = (ORBX_XOFF + x - DESC_R) & 3;
        const uint8_t* bc = reinterpret_cast<const uint8_t*>(patch) + DESC_R * (DESC_PW * 4) + DESC_R + xoff;
        const float MAGIC = 12582912.f;                           // 1.5 * 2^23
        // tap address = ry * 44 + rx with r = bits(v + MAGIC) - bits(MAGIC)
        const uint32_t abias = 0x4B400000u * (uint32_t)(DESC_PW * 4 + 1);    // wraps; the difference below is exact
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            int t[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const float px = pat[4 * k + 2 * e], py = pat[4 * k + 2 * e + 1];
                const float fy = __fadd_rn(__fadd_rn(__fmul_rn(px, b), __fmul_rn(py, a)), MAGIC);
                const float fx = __fadd_rn(__fsub_rn(__fmul_rn(px, a), __fmul_rn(py, b)), MAGIC);
                t[e] = bc[(int)(__float_as_uint(fy) * (uint32_t)(DESC_PW * 4) + __float_as_uint(fx) - abias)];
            }
            val |= (t[0] < t[1]) << k;
        }
        const int dst = D.outidx[(long long)f * G.sel_per_frame + slot];
        D.desc[((long long)f * G.out_cap + dst) * 32 + lane] = (uint8_t)val;
        if (lane == 0) {
            D.kps[(long long)f * G.out_cap + dst].angle = angle;
            D.angle[(long long)f * G.sel_per_frame + slot] = angle;
        }
    }
}

int launch_describe(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    const dim3 grid((G.sel_per_frame + DESC_WARPS * DESC_KPW - 1) / (DESC_WARPS * DESC_KPW), nframes);
    orbx_launch(describe_kernel, grid, dim3(DESC_WARPS * 32), 0, s, G, D);
    return 1;
}
