// orbx_host.h — host-side extractor handle.
#pragma once
#include "orbx_internal.h"
#include <vector>

#define ORBX_NSTAGES 6
#define ORBX_NCSTREAMS 4

struct orbx_extractor {
    // ORBextractor members (include/ORBextractor.h:93-112)
    int nfeatures, nlevels, iniTh, minTh;
    float scaleFactorF;
    double scaleFactorD;
    float sf[ORBX_MAX_LEVELS], isf[ORBX_MAX_LEVELS], sig2[ORBX_MAX_LEVELS], isig2[ORBX_MAX_LEVELS];
    int nper[ORBX_MAX_LEVELS];
    // device state
    int device;
    cudaStream_t stream;
    cudaStream_t cstream[ORBX_NCSTREAMS];  // compute streams: sub-batches run concurrently so kernel tails overlap
    cudaEvent_t ev_fork, ev_join[ORBX_NCSTREAMS];
    // side streams: the blur of a sub-batch runs beside its FAST / quadtree / assemble (it only needs the pyramid); one per compute
    // stream plus one for the handle's own stream, created on first use
    cudaStream_t bstream[ORBX_NCSTREAMS + 1] = {};
    cudaEvent_t ev_pyr[ORBX_NCSTREAMS + 1] = {}, ev_blur[ORBX_NCSTREAMS + 1] = {};
    cudaStream_t stream_h2d, stream_d2h;   // copy engines of the chunked extract_batch pipeline
    std::vector<cudaEvent_t> ev_up, ev_done, ev_down, ev_piece;   // ev_piece: upload / kernel events of the ramp pieces of a call's first and last chunk
    int* h_counts; size_t h_counts_cap;    // pinned {nkp, mono} staging (a pageable target would make the D2H copy synchronous)
    // single-chunk calls (one frame / a few frames): pinned staging of the input frames and a pinned landing zone for
    // {keypoints, descriptors}: with pageable user buffers every cudaMemcpyAsync is a synchronous staged copy
    // the pipeline of a single-chunk call as an instantiated CUDA graph (one graph launch instead of 14 kernel launches);
    // valid for one {frames, lapping area, input pitch} and the current device buffers
    cudaGraphExec_t graph_exec; int graph_key[4]; int graph_launches;
    // per-level branches of that graph: FAST and the quadtree of level l start as soon as pyramid level l exists
    cudaStream_t dag_stream[ORBX_MAX_LEVELS]; cudaEvent_t dag_ev[ORBX_MAX_LEVELS], dag_done[ORBX_MAX_LEVELS]; bool dag_ready;
    uint8_t* h_in; size_t h_in_cap;
    uint8_t* h_land; size_t h_land_cap;
    bool reserved;
    FrameGeom G;
    DevBuffers D;
    std::vector<void*> allocs;           // every cudaMalloc of this handle
    cudaEvent_t ev[ORBX_NSTAGES + 1];
    cudaEvent_t tev[2];                  // stopwatch (orbx_timer_start/stop)
    cudaEvent_t join_ev;                 // stereo: orders the left handle's stream behind this (right) handle's
    bool profiling;
    float stage_ms[ORBX_NSTAGES + 1];    // [ORBX_NSTAGES] = total
    long long launches;
    int last_nframes;                    // frames processed by the last run (valid pyramids / outputs)
    int batch_limit = 512;               // most frame slots ensure_reserved may give this geometry (8 GB bound)
    // stereo-match outputs (device), sized out_cap per frame slot
    float* d_uright; float* d_depth;
    int* d_st_scratch;
    uint32_t* d_st_rows;                 // stereo: row-band index of the right keypoints
    // frame preparation (prepare.cu): packed rectification maps of this camera and raw frame slots.
    // Both live outside `allocs`: they survive a change of the extractor geometry.
    int2* d_map; int map_rows, map_cols;
    uint8_t* d_raw; size_t raw_bytes;    // raw (unrectified / colour) frames, one slot per frame slot
    int raw_rows, raw_cols, raw_ch, raw_rgb, raw_pitch;
};

void orbx_init_tables(orbx_extractor* e);
int orbx_compute_geometry(const orbx_extractor* e, int rows, int cols, int batch, FrameGeom* G,
                          std::vector<std::vector<ResizeTapX>>* tapx, std::vector<std::vector<ResizeTapY>>* tapy);
bool orbx_build_word_taps(const std::vector<ResizeTapX>& TX, int w, int pitch, std::vector<ResizeWord>* out);
void orbx_set_error(const char* fmt, ...);
int launch_prepare(orbx_extractor* e, int slot0, int nframes, cudaStream_t s);
int launch_rectify_pack(const float* mx, const float* my, int rows, int cols, int mstep_floats, int2* out, cudaStream_t s);
