// orbx_internal.h — device data layout and kernel launchers of the B200 ORB front-end.
//
// HBM layout (all per extractor handle, sized for `batch` frame slots):
//   input    : batch frames, rows x in_pitch (the caller's row stride when its frames are packed back to back,
//              else cols rounded up to 16).
//   pyr[l]   : batch planes of level l, each `prow x pitch` bytes; interior pixel (0,0) at byte
//              ORBX_YOFF*pitch + ORBX_XOFF (XOFF=32 keeps interior rows 32B-sector aligned, pitch is a
//              multiple of 128).  The hot path writes the interior plus a thin reflect-101 halo (3 rows
//              above / below, one 4-px word left / right, written by the pyramid kernels) for the 7x7 blur; the full
//              19-px border of mvImagePyramid is filled on demand (pyramid_fill_border) when a padded
//              level is read back.  FAST, IC_Angle and the descriptor taps never leave the interior.
//   blur[l]  : same geometry, GaussianBlur(7x7, s=2) of the level (interior only).
//   bands    : FAST work items (BandDesc): runs of whole cells of one cell row, <= 128 px wide.
//   cand     : per frame, per level FAST keypoints that survive the cell's NMS, packed keys
//              (y<<20 | x<<8 | score), UNORDERED (one atomic bump per work item); candcnt = counts.
//              The reference's order (vToDistributeKeys) is recovered where it matters by a canonical
//              rank computed from the coordinates (cand_rank in octree.cu).
//   sel      : per frame, per level quadtree selection in std::list order (+ count), packed keys.
//   nodetab  : quadtree node tables for levels whose tables do not fit 64 KB of shared memory.
//   out      : per frame `out_cap` cv::KeyPoint records + 32-byte descriptors + {nkp, mono}.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/orbx.h"

#define ORBX_EDGE 19
#define ORBX_XOFF 32
#define ORBX_YOFF 19
#define ORBX_MINB 16          // EDGE_THRESHOLD-3, src/ORBextractor.cc:1053
#define ORBX_HALF_PATCH 15
#define ORBX_PATCH 31

// candidate / selected-keypoint key: y (16 bits) | x (16 bits) | FAST score (8 bits), coordinates relative to minBorder.
// 64-bit words: 12-bit coordinates in a 32-bit key stopped at level planes of 4111 px.
typedef uint64_t orbx_key_t;
#define KEY_X(k) ((uint32_t)((k) >> 8) & 0xFFFFu)
#define KEY_Y(k) ((uint32_t)((k) >> 24) & 0xFFFFu)
#define KEY_S(k) ((uint32_t)(k) & 0xFFu)
#define MAKE_KEY(x, y, s) (((orbx_key_t)(uint32_t)(y) << 24) | ((orbx_key_t)(uint32_t)(x) << 8) | (orbx_key_t)(uint32_t)(s))

struct ResizeTapX { short s0, s1, a0, a1; };   // source cols (clipped) and 11-bit coefficients
struct ResizeTapY { short s0, s1, b0, b1; };

// horizontal taps of one output word (4 px) of the fast resize path: the 8 source bytes starting at
// the first pixel's left tap hold every tap of the 4 pixels
struct ResizeWord {
    int wb;                   // byte offset (multiple of 4) of the aligned source word holding the first tap
    uint32_t shift;           // bit shift that brings the first tap to byte 0 of the 8-byte window
    uint32_t sels;            // per pixel one byte: PRMT selector (left tap | right tap << 4) into the window
    uint32_t a01[4];          // per pixel the 11-bit coefficients a0 | a1 << 16
    uint32_t pad;
};

#define PT_ROWS 82            // TMA resize tile: 64 output rows * 1.2 + taps + the reflected halo rows
#define PT_BYTES 192          // 32 words * 4 px * 1.2 + 12 tap bytes + 15 alignment bytes, multiple of 16

struct LevelGeom {
    int pyr_tma_ok;           // the TMA resize tile (PT_ROWS x PT_BYTES) covers every CTA's source rows / tap bytes of this level
    int w, h;                 // level image size (cols, rows)
    int pitch, prow;          // plane row pitch (bytes) and padded row count (h + 38)
    long long plane_bytes;    // pitch * prow
    uint8_t* pyr;             // level planes of all frame slots
    uint8_t* blur;
    // FAST cell grid, src/ORBextractor.cc:1053-1072
    int maxBX, maxBY;         // maxBorderX/Y (minBorder = 16)
    int nCols, nRows, wCell, hCell;
    int cell_base;            // first cell of this level in a frame's flattened cell list
    int cell_cap;             // key slots per cell
    long long key_base;       // first slot of this level in a frame's cellkey array
    // quadtree
    int N;                    // mnFeaturesPerLevel[level]
    int nIni;                 // root nodes, :695
    float hX;                 // :698
    int cand_cap;             // capacity of compact candidate list
    long long cand_base;      // offset in a frame's cand array
    int node_cap;             // capacity of node tables
    int sel_cap;              // capacity of selection list
    int sel_base;             // offset in a frame's sel arrays
    // descriptor stage
    float scale;              // mvScaleFactor[level]
    float patch_size;         // (float)(int)(31*scale), :1161
    // resize taps producing THIS level from level-1, indexed by padded coordinate (border folded in)
    const ResizeTapX* tapx;   // pitch entries, indexed by byte of the padded row
    const ResizeTapY* tapy;   // h + 38 entries
    const ResizeWord* tapw;   // pitch/4 entries (interior words only are meaningful); NULL: generic path
};

// one FAST work item (one warp): the interiors of cells [cj0, cj1) of cell row ci of level lvl (<= 128 px wide)
struct BandDesc { short lvl, ci, cj0, cj1; };

// one FAST work item of the strip kernel (one warp): 128-px strip k of a GROUP = the interiors of a run of whole cells
// of cell row ci, [gxs, gxe) in level coordinates; the nstrips warps of a group sit side by side in one CTA starting
// at warp `first`.  lvl < 0: empty slot.
struct StripDesc { short lvl, ci, gxs, gxe, k, first, nstrips, pad; };

struct FrameGeom {
    int nlevels;
    int strip_ctas;           // CTAs per frame of fast_strip_kernel (0: use fast_band_kernel)
    int strip_lvl_cta[ORBX_MAX_LEVELS + 1];   // second, level-major plan behind the first: CTAs [lvl_cta[l], lvl_cta[l+1]) serve level l
    int bands_per_frame;      // FAST band segments per frame (DevBuffers::bands)
    int fast_smem;            // dynamic shared memory of fast_band_kernel (bytes)
    int fast_warp_words;      // ... of which 32-bit words per warp
    int rows, cols;           // input image size
    int batch;                // frame slots
    int cells_per_frame;
    long long keys_per_frame; // cellkey slots per frame
    long long cand_per_frame;
    int sel_per_frame;        // == out_cap
    int out_cap;
    int iniTh, minTh;
    int blur_end[ORBX_MAX_LEVELS];   // cumulative blur CTAs per frame up to and including level l
    int blur_tma_end[ORBX_MAX_LEVELS];   // same for the TMA variant (128-row tiles)
    LevelGeom L[ORBX_MAX_LEVELS];
};

struct DevBuffers {
    uint8_t* input;           // batch x rows x in_pitch staging of the uploaded frames
    int in_pitch;             // row stride of the frames currently resident (the caller's stride when packed)
    int in_pitch_cap;         // largest row stride the staging buffer is sized for
    const BandDesc* bands;    // [bands_per_frame] FAST work units (same for every frame)
    const StripDesc* strips;  // [strip_ctas * 4] work units of the strip kernel
    orbx_key_t* cand;         // [batch][cand_per_frame]
    uint32_t* candnode;       // [batch][cand_per_frame]  quadtree scratch: node of each candidate
    int* nodetab;             // quadtree scratch: node tables
    long long nodetab_per_frame;
    orbx_key_t* sel;          // [batch][sel_per_frame]
    int* selcnt;              // [batch][nlevels]
    int* candcnt;             // [batch][nlevels]
    float* angle;             // [batch][sel_per_frame]
    int* outidx;              // [batch][sel_per_frame] final row of each selected keypoint
    orbx_keypoint* kps;       // [batch][out_cap]
    uint8_t* desc;            // [batch][out_cap][32]
    int* counts;              // [batch][2] = {nkp, mono}
    const void* blur_maps;    // HOST pointer: BlurMaps (TMA tensor maps of the pyramid planes, describe.cu) or NULL
    const void* pyr_maps;     // HOST pointer: PyrMaps (pyramid.cu; ORBX_PYR_TMA=0 disables) or NULL
    int slot0;                // first frame slot of this view (TMA coordinates are absolute)
    const uint8_t* zc_input;  // != NULL: level 0 reads the frames from here (mapped pinned host memory) instead of `input`
    int zc_pitch; long long zc_frame_bytes;
};

// kernel launchers (each returns the number of kernel launches it issued)
int launch_pyramid(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s);
void blur_plan(FrameGeom& G);           // fills blur_end[] (and blur_tma_end[])
const void* blur_tma_create(const FrameGeom& G);   // tensor maps over the level planes (NULL: TMA path unavailable)
void blur_tma_destroy(const void* maps);
const void* pyr_tma_create(const FrameGeom& G);
void pyr_tma_destroy(const void* maps);
void pyramid_fill_border(const LevelGeom& L, int frame, cudaStream_t s);   // mvImagePyramid border on demand
bool fast_per_level(const FrameGeom& G);
int launch_pyramid_level(const FrameGeom& G, const DevBuffers& D, int nframes, int l, cudaStream_t s);
int launch_fast(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s, int level = -1);     // level >= 0: that level only
int launch_octree(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s, int level = -1);
int launch_assemble(const FrameGeom& G, const DevBuffers& D, int nframes, int lap0, int lap1, cudaStream_t s);
int launch_blur(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s);
int launch_describe(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s);
#ifdef __cplusplus
#include <vector>
int fast_build_bands(const FrameGeom& G, std::vector<BandDesc>* out, int* warp_words);   // returns the kernel's dynamic smem bytes
int fast_build_strips(const FrameGeom& G, std::vector<StripDesc>* out, int* lvl_cta);     // returns the CTAs per frame, 0 if unsupported
#endif
size_t octree_nodetab_ints(const FrameGeom& G);   // ints of node scratch per frame
void octree_assign_node_offsets(FrameGeom& G);

// ---- programmatic dependent launch (sm_90+) ----------------------------------------------------
// The kernels of one extraction form a chain on one stream.  Each of them starts with orbx_pdl_prologue():
// `launch_dependents` lets the NEXT launch of the stream be set up and its CTAs made resident while this grid is still
// running, `wait` then holds every thread until the preceding grid has completed and flushed.  The data dependences are
// exactly those of plain stream order (nothing is read or written before the wait); what should disappear is the launch
// latency between dependent kernels.  MEASURED (B200, one 752x480 frame per call, 14 launches): best case 229 vs 234 us
// per call, but the median got WORSE (251 vs 237 us, bimodal) and batch throughput did not move - so it is opt-in
// (ORBX_PDL=1), off by default.  Launched without the attribute the two instructions are no-ops.
#ifdef __CUDACC__
#include <cstdlib>
__device__ __forceinline__ void orbx_pdl_prologue()
{
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

inline bool orbx_pdl_enabled()
{
    static const bool v = [] { const char* c = getenv("ORBX_PDL"); return c && atoi(c) != 0; }();
    return v;
}

template <class... KArgs, class... Args>
inline void orbx_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = orbx_pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif
