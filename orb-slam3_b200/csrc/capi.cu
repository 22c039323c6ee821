// capi.cu — C ABI of the extractor (include/orbx.h): handle life cycle, device buffers, the
// extraction pipeline and host<->device transfers.  No CPU fallback: every compute entry point needs
// a CUDA device and fails with ORBX_ERR_CUDA otherwise.
#include "orbx_host.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <string>
#include <thread>

static thread_local char g_err[512] = "";

void orbx_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            orbx_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return ORBX_ERR_CUDA;                                                              \
        }                                                                                      \
    } while (0)

extern "C" const char* orbx_last_error(void) { return g_err; }
extern "C" int orbx_version(void) { return 100; }
extern "C" int orbx_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int orbx_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                           int device, orbx_extractor** out)
{
    if (!out) { orbx_set_error("out is NULL"); return ORBX_ERR_ARG; }
    *out = nullptr;
    if (nlevels < 1 || nlevels > ORBX_MAX_LEVELS || nfeatures < 0 || !(scaleFactor > 1.0f)) {
        orbx_set_error("bad parameters: nfeatures=%d scaleFactor=%g nlevels=%d (1..%d)", nfeatures, scaleFactor, nlevels, ORBX_MAX_LEVELS);
        return ORBX_ERR_ARG;
    }
    int ndev = orbx_device_count();
    if (device < 0 || device >= ndev) {
        orbx_set_error("CUDA device %d not available (%d visible); this library has no CPU path", device, ndev);
        return ORBX_ERR_CUDA;
    }
    orbx_extractor* e = new orbx_extractor();
    e->nfeatures = nfeatures; e->nlevels = nlevels; e->iniTh = iniThFAST; e->minTh = minThFAST;
    e->scaleFactorF = scaleFactor;
    e->device = device; e->reserved = false; e->profiling = false; e->launches = 0; e->last_nframes = 0;
    e->d_uright = e->d_depth = nullptr; e->d_st_scratch = nullptr; e->d_st_rows = nullptr;
    e->h_counts = nullptr; e->h_counts_cap = 0;
    e->h_in = e->h_land = nullptr; e->h_in_cap = e->h_land_cap = 0;
    e->graph_exec = nullptr; e->graph_launches = 0; e->dag_ready = false;
    e->d_map = nullptr; e->map_rows = e->map_cols = 0;
    e->d_raw = nullptr; e->raw_bytes = 0; e->raw_rows = e->raw_cols = e->raw_ch = e->raw_rgb = e->raw_pitch = 0;
    memset(&e->G, 0, sizeof(e->G)); memset(&e->D, 0, sizeof(e->D));
    memset(e->stage_ms, 0, sizeof(e->stage_ms));
    orbx_init_tables(e);
    cudaError_t ce = cudaSetDevice(device);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    for (int i = 0; i < ORBX_NCSTREAMS && ce == cudaSuccess; i++) {
        ce = cudaStreamCreateWithFlags(&e->cstream[i], cudaStreamNonBlocking);
        if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ev_join[i], cudaEventDisableTiming);
    }
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->stream_h2d, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->stream_d2h, cudaStreamNonBlocking);
    for (int i = 0; i <= ORBX_NSTAGES && ce == cudaSuccess; i++) ce = cudaEventCreate(&e->ev[i]);
    for (int i = 0; i < 2 && ce == cudaSuccess; i++) ce = cudaEventCreate(&e->tev[i]);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->join_ev, cudaEventDisableTiming);
    if (ce != cudaSuccess) {
        orbx_set_error("CUDA init failed: %s", cudaGetErrorString(ce));
        delete e;
        return ORBX_ERR_CUDA;
    }
    *out = e;
    return ORBX_OK;
}

static void free_device(orbx_extractor* e)
{
    cudaSetDevice(e->device);
    if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }   // it holds the old pointers
    for (void* p : e->allocs) cudaFree(p);
    e->allocs.clear();
    if (e->reserved && e->D.blur_maps) { blur_tma_destroy(e->D.blur_maps); e->D.blur_maps = nullptr; }
    if (e->reserved && e->D.pyr_maps) { pyr_tma_destroy(e->D.pyr_maps); e->D.pyr_maps = nullptr; }
    e->reserved = false;
    e->last_nframes = 0;
    e->d_uright = e->d_depth = nullptr; e->d_st_scratch = nullptr; e->d_st_rows = nullptr;
}

extern "C" void orbx_destroy(orbx_extractor* e)
{
    if (!e) return;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    free_device(e);
    if (e->d_map) cudaFree(e->d_map);
    if (e->d_raw) cudaFree(e->d_raw);
    for (int i = 0; i <= ORBX_NSTAGES; i++) cudaEventDestroy(e->ev[i]);
    for (int i = 0; i < 2; i++) cudaEventDestroy(e->tev[i]);
    cudaEventDestroy(e->join_ev);
    for (cudaEvent_t ev : e->ev_up) cudaEventDestroy(ev);
    for (cudaEvent_t ev : e->ev_done) cudaEventDestroy(ev);
    for (cudaEvent_t ev : e->ev_down) cudaEventDestroy(ev);
    for (cudaEvent_t ev : e->ev_piece) cudaEventDestroy(ev);
    for (int i = 0; i < ORBX_NCSTREAMS; i++) { cudaStreamDestroy(e->cstream[i]); cudaEventDestroy(e->ev_join[i]); }
    for (int i = 0; i <= ORBX_NCSTREAMS; i++) if (e->bstream[i]) { cudaStreamDestroy(e->bstream[i]); cudaEventDestroy(e->ev_pyr[i]); cudaEventDestroy(e->ev_blur[i]); }
    cudaEventDestroy(e->ev_fork);
    if (e->dag_ready)
        for (int l = 0; l < ORBX_MAX_LEVELS; l++) { cudaStreamDestroy(e->dag_stream[l]); cudaEventDestroy(e->dag_ev[l]); cudaEventDestroy(e->dag_done[l]); }
    if (e->h_counts) cudaFreeHost(e->h_counts);
    if (e->h_in) cudaFreeHost(e->h_in);
    if (e->h_land) cudaFreeHost(e->h_land);
    cudaStreamDestroy(e->stream_h2d); cudaStreamDestroy(e->stream_d2h);
    cudaStreamDestroy(e->stream);
    delete e;
}

extern "C" int orbx_levels(const orbx_extractor* e) { return e->nlevels; }
extern "C" float orbx_scale_factor(const orbx_extractor* e) { return (float)e->scaleFactorD; }
extern "C" const float* orbx_scale_factors(const orbx_extractor* e) { return e->sf; }
extern "C" const float* orbx_inv_scale_factors(const orbx_extractor* e) { return e->isf; }
extern "C" const float* orbx_level_sigma2(const orbx_extractor* e) { return e->sig2; }
extern "C" const float* orbx_inv_level_sigma2(const orbx_extractor* e) { return e->isig2; }
extern "C" const int* orbx_features_per_level(const orbx_extractor* e) { return e->nper; }

extern "C" int orbx_max_keypoints(const orbx_extractor* e, int rows, int cols)
{
    FrameGeom G;
    std::vector<std::vector<ResizeTapX>> tx;
    std::vector<std::vector<ResizeTapY>> ty;
    int r = orbx_compute_geometry(e, rows, cols, 1, &G, &tx, &ty);
    if (r != 0) { orbx_set_error("image %dx%d too small for %d pyramid levels", cols, rows, e->nlevels); return r; }
    return G.out_cap;
}

template <typename T>
static int dev_alloc(orbx_extractor* e, T** p, size_t count)
{
    void* q = nullptr;
    CUDA_TRY(cudaMalloc(&q, count * sizeof(T) + 256));
    e->allocs.push_back(q);
    *p = (T*)q;
    return 0;
}

extern "C" int orbx_reserve(orbx_extractor* e, int rows, int cols, int max_batch)
{
    if (!e || rows <= 0 || cols <= 0 || max_batch <= 0) { orbx_set_error("bad reserve arguments"); return ORBX_ERR_ARG; }
    if (e->reserved && e->G.rows == rows && e->G.cols == cols && e->G.batch >= max_batch) return ORBX_OK;
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    free_device(e);
    FrameGeom G;
    std::vector<std::vector<ResizeTapX>> tx;
    std::vector<std::vector<ResizeTapY>> ty;
    int r = orbx_compute_geometry(e, rows, cols, max_batch, &G, &tx, &ty);
    if (r != 0) { orbx_set_error("image %dx%d too small for %d pyramid levels (35-px cell grid)", cols, rows, e->nlevels); return r; }
    DevBuffers D;
    memset(&D, 0, sizeof(D));
    const size_t B = (size_t)max_batch;
    // device rows keep the caller's stride when the frames are packed back to back (one linear H2D copy, also
    // for odd widths such as KITTI's 1241); otherwise rows are 16-byte aligned.  Room for strides up to +64 bytes.
    D.in_pitch = (cols + 15) / 16 * 16;
    D.in_pitch_cap = D.in_pitch + 64;
    int rc = 0;
    bool geometry_error = false;
    rc |= dev_alloc(e, &D.input, B * rows * D.in_pitch_cap + 64);
    for (int l = 0; l < G.nlevels && !rc; l++) {
        rc |= dev_alloc(e, &G.L[l].pyr, B * G.L[l].plane_bytes);
        rc |= dev_alloc(e, &G.L[l].blur, B * G.L[l].plane_bytes);
        if (!rc) {
            // zero once: bytes outside the padded image are never written by the kernels
            if (cudaMemsetAsync(G.L[l].pyr, 0, B * G.L[l].plane_bytes, e->stream) != cudaSuccess) rc = ORBX_ERR_CUDA;
            if (cudaMemsetAsync(G.L[l].blur, 0, B * G.L[l].plane_bytes, e->stream) != cudaSuccess) rc = ORBX_ERR_CUDA;
        }
        if (l > 0 && !rc) {
            ResizeTapX* dx; ResizeTapY* dy;
            rc |= dev_alloc(e, &dx, tx[l].size());
            rc |= dev_alloc(e, &dy, ty[l].size());
            if (!rc) {
                CUDA_TRY(cudaMemcpy(dx, tx[l].data(), tx[l].size() * sizeof(ResizeTapX), cudaMemcpyHostToDevice));
                CUDA_TRY(cudaMemcpy(dy, ty[l].data(), ty[l].size() * sizeof(ResizeTapY), cudaMemcpyHostToDevice));
                G.L[l].tapx = dx; G.L[l].tapy = dy;
                std::vector<ResizeWord> tw;
                if (orbx_build_word_taps(tx[l], G.L[l].w, G.L[l].pitch, &tw)) {
                    ResizeWord* dw;
                    rc |= dev_alloc(e, &dw, tw.size());
                    if (!rc) CUDA_TRY(cudaMemcpy(dw, tw.data(), tw.size() * sizeof(ResizeWord), cudaMemcpyHostToDevice));
                    G.L[l].tapw = dw;
                    // does the TMA resize tile hold the sources of every CTA (32 words x 64 rows) of this level?
                    const int wlast = (ORBX_XOFF + G.L[l].w + 3) / 4, HALO = 3;
                    bool ok = true;
                    for (int w0 = ORBX_XOFF / 4 - 1; w0 <= wlast && ok; w0 += 32) {
                        int lo = 1 << 30, hi = 0;
                        for (int wi = w0; wi < w0 + 32 && wi <= wlast; wi++) { lo = std::min(lo, tw[wi].wb); hi = std::max(hi, tw[wi].wb + 12); }
                        const int start = ((ORBX_XOFF + lo) & ~15) - ORBX_XOFF;
                        if (lo < -ORBX_XOFF || hi - start > PT_BYTES) ok = false;
                    }
                    for (int y0 = -HALO; y0 < G.L[l].h + HALO && ok; y0 += 64) {
                        int lo = 1 << 30, hi = 0;
                        for (int y = y0; y < y0 + 64 && y < G.L[l].h + HALO; y++) {
                            const ResizeTapY& t = ty[l][ORBX_EDGE + y];
                            lo = std::min(lo, (int)t.s0); hi = std::max(hi, (int)t.s1);
                        }
                        const int first = std::min((int)ty[l][ORBX_EDGE + std::max(y0, 0)].s0, (int)ty[l][ORBX_EDGE + std::min(y0 + 63, G.L[l].h + HALO - 1)].s0);
                        if (lo < first || hi - first + 1 > PT_ROWS) ok = false;
                    }
                    G.L[l].pyr_tma_ok = ok ? 1 : 0;
                }
            }
        }
    }
    D.nodetab_per_frame = (long long)octree_nodetab_ints(G);
    blur_plan(G);
    {
        std::vector<BandDesc> bands;
        G.fast_smem = fast_build_bands(G, &bands, &G.fast_warp_words);
        if (G.fast_smem < 0) { orbx_set_error("cell rows taller than 127 px are not supported"); geometry_error = true; rc |= ORBX_ERR_GEOMETRY; G.fast_smem = 0; }
        G.bands_per_frame = (int)bands.size();
        BandDesc* db = nullptr;
        if (!rc) rc |= dev_alloc(e, &db, bands.size());
        if (!rc) CUDA_TRY(cudaMemcpy(db, bands.data(), bands.size() * sizeof(BandDesc), cudaMemcpyHostToDevice));
        D.bands = db;
        std::vector<StripDesc> strips;
        G.strip_ctas = rc ? 0 : fast_build_strips(G, &strips, G.strip_lvl_cta);
        D.strips = nullptr;
        if (G.strip_ctas > 0) {
            StripDesc* ds = nullptr;
            rc |= dev_alloc(e, &ds, strips.size());
            if (!rc) CUDA_TRY(cudaMemcpy(ds, strips.data(), strips.size() * sizeof(StripDesc), cudaMemcpyHostToDevice));
            D.strips = ds;
        }
    }
    if (!rc) rc |= dev_alloc(e, &D.cand, B * G.cand_per_frame);
    if (!rc) rc |= dev_alloc(e, &D.candnode, B * G.cand_per_frame);
    if (!rc) rc |= dev_alloc(e, &D.nodetab, B * G.nlevels * D.nodetab_per_frame);
    if (!rc) rc |= dev_alloc(e, &D.sel, B * G.sel_per_frame);
    if (!rc) rc |= dev_alloc(e, &D.selcnt, B * G.nlevels);
    if (!rc) rc |= dev_alloc(e, &D.candcnt, B * G.nlevels);
    if (!rc) rc |= dev_alloc(e, &D.angle, B * G.sel_per_frame);
    if (!rc) rc |= dev_alloc(e, &D.outidx, B * G.sel_per_frame);
    if (!rc) rc |= dev_alloc(e, &D.kps, B * G.out_cap);
    if (!rc) rc |= dev_alloc(e, &D.desc, B * G.out_cap * 32);
    if (!rc) rc |= dev_alloc(e, &D.counts, B * 2);
    if (!rc) rc |= dev_alloc(e, &e->d_uright, B * G.out_cap);
    if (!rc) rc |= dev_alloc(e, &e->d_depth, B * G.out_cap);
    if (rc) { free_device(e); return geometry_error ? ORBX_ERR_GEOMETRY : ORBX_ERR_CUDA; }   // keep the cause
    // a slot that no call has written yet reads as "no keypoints" (stereo / getters on a partly used batch)
    CUDA_TRY(cudaMemsetAsync(D.counts, 0, sizeof(*D.counts) * B * 2, e->stream));
    CUDA_TRY(cudaMemsetAsync(D.selcnt, 0, sizeof(*D.selcnt) * B * G.nlevels, e->stream));
    CUDA_TRY(cudaMemsetAsync(D.candcnt, 0, sizeof(*D.candcnt) * B * G.nlevels, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    D.blur_maps = blur_tma_create(G); D.slot0 = 0;
    D.pyr_maps = pyr_tma_create(G);
    e->G = G; e->D = D; e->reserved = true;
    return ORBX_OK;
}

static int upload_frames(orbx_extractor* e, int slot0, int nframes, const uint8_t* imgs, size_t step, size_t frame_stride,
                         cudaStream_t st)
{
    const FrameGeom& G = e->G;
    const bool packed = frame_stride == step * (size_t)G.rows && step <= (size_t)e->D.in_pitch_cap;
    // all frames resident at one time share one row stride: the caller's when packed, else the aligned one
    e->D.in_pitch = packed ? (int)step : (G.cols + 15) / 16 * 16;
    uint8_t* dst = e->D.input + (size_t)slot0 * G.rows * e->D.in_pitch;
    if (packed) {
        // frames back to back: one linear copy, whatever the alignment of the rows (the level-0 kernel copes)
        CUDA_TRY(cudaMemcpyAsync(dst, imgs, (size_t)nframes * G.rows * step, cudaMemcpyHostToDevice, st));
    } else if (frame_stride == step * (size_t)G.rows) {
        // frames are back to back: one 2-D copy of nframes*rows rows
        CUDA_TRY(cudaMemcpy2DAsync(dst, e->D.in_pitch, imgs, step, G.cols, (size_t)nframes * G.rows, cudaMemcpyHostToDevice, st));
    } else {
        for (int f = 0; f < nframes; f++)
            CUDA_TRY(cudaMemcpy2DAsync(dst + (size_t)f * G.rows * e->D.in_pitch, e->D.in_pitch,
                                       imgs + (size_t)f * frame_stride, step, G.cols, G.rows, cudaMemcpyHostToDevice, st));
    }
    return 0;
}

// views of the geometry / buffers that start at frame slot s0 (kernels index frames from 0)
static void slot_view(const orbx_extractor* e, int s0, FrameGeom* Go, DevBuffers* Do)
{
    *Go = e->G; *Do = e->D;
    Do->slot0 = s0; Do->zc_input = nullptr; Do->zc_pitch = 0; Do->zc_frame_bytes = 0;
    const FrameGeom& G = e->G;
    const size_t s = (size_t)s0;
    for (int l = 0; l < G.nlevels; l++) { Go->L[l].pyr += s * G.L[l].plane_bytes; Go->L[l].blur += s * G.L[l].plane_bytes; }
    Do->input += s * G.rows * e->D.in_pitch;
    Do->cand += s * G.cand_per_frame; Do->candnode += s * G.cand_per_frame;
    Do->nodetab += s * G.nlevels * e->D.nodetab_per_frame;
    Do->sel += s * G.sel_per_frame; Do->selcnt += s * G.nlevels; Do->candcnt += s * G.nlevels;
    Do->angle += s * G.sel_per_frame; Do->outidx += s * G.sel_per_frame;
    Do->kps += s * G.out_cap; Do->desc += s * G.out_cap * 32; Do->counts += 2 * s;
}

// the whole extraction pipeline on frame slots [0, nframes), asynchronous on the handle's stream
static int run_pipeline(orbx_extractor* e, int slot0, int nframes, int lap0, int lap1, cudaStream_t s,
                        const uint8_t* zc_input = nullptr, size_t zc_step = 0, size_t zc_frame_stride = 0)
{
    FrameGeom G; DevBuffers D;
    slot_view(e, slot0, &G, &D);
    D.zc_input = zc_input; D.zc_pitch = (int)zc_step; D.zc_frame_bytes = (long long)zc_frame_stride;
    const bool prof = e->profiling && s == e->stream;
    // blur beside FAST on a side stream (ORBX_SIDE_BLUR=1; launches of >= 16 frames outside profiling / graph capture).  Measured on
    // B200, C1: 142.6 k vs 142.4 k frames/s, and 141.1 k vs 141.4 k with FAST capped at 3 CTAs per SM - nothing to gain, off
    static const bool side_on = [] { const char* v = getenv("ORBX_SIDE_BLUR"); return v && atoi(v) != 0; }();
    int side = -1;
    if (side_on && !e->profiling && nframes >= 16) {
        cudaStreamCaptureStatus cst = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(s, &cst) == cudaSuccess && cst == cudaStreamCaptureStatusNone) {
            side = ORBX_NCSTREAMS;
            for (int i = 0; i < ORBX_NCSTREAMS; i++) if (e->cstream[i] == s) side = i;
            if (side == ORBX_NCSTREAMS && s != e->stream) side = -1;
            if (side >= 0 && !e->bstream[side]) {
                if (cudaStreamCreateWithFlags(&e->bstream[side], cudaStreamNonBlocking) != cudaSuccess ||
                    cudaEventCreateWithFlags(&e->ev_pyr[side], cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&e->ev_blur[side], cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); side = -1; }
            }
        }
    }
    if (side >= 0) {
        e->launches += launch_pyramid(G, D, nframes, s);
        CUDA_TRY(cudaEventRecord(e->ev_pyr[side], s));
        CUDA_TRY(cudaStreamWaitEvent(e->bstream[side], e->ev_pyr[side], 0));
        e->launches += launch_blur(G, D, nframes, e->bstream[side]);
        CUDA_TRY(cudaEventRecord(e->ev_blur[side], e->bstream[side]));
        e->launches += launch_fast(G, D, nframes, s);
        e->launches += launch_octree(G, D, nframes, s);
        e->launches += launch_assemble(G, D, nframes, lap0, lap1, s);
        CUDA_TRY(cudaStreamWaitEvent(s, e->ev_blur[side], 0));
        e->launches += launch_describe(G, D, nframes, s);
        CUDA_TRY(cudaGetLastError());
        return 0;
    }
    if (prof) cudaEventRecord(e->ev[0], s);
    e->launches += launch_pyramid(G, D, nframes, s);
    if (prof) cudaEventRecord(e->ev[1], s);
    e->launches += launch_fast(G, D, nframes, s);
    if (prof) cudaEventRecord(e->ev[2], s);
    e->launches += launch_octree(G, D, nframes, s);
    if (prof) cudaEventRecord(e->ev[3], s);
    e->launches += launch_assemble(G, D, nframes, lap0, lap1, s);
    if (prof) cudaEventRecord(e->ev[4], s);
    e->launches += launch_blur(G, D, nframes, s);
    if (prof) cudaEventRecord(e->ev[5], s);
    e->launches += launch_describe(G, D, nframes, s);
    if (prof) cudaEventRecord(e->ev[6], s);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static void collect_stage_times(orbx_extractor* e)
{
    if (!e->profiling) return;
    for (int i = 0; i < ORBX_NSTAGES; i++) cudaEventElapsedTime(&e->stage_ms[i], e->ev[i], e->ev[i + 1]);
    cudaEventElapsedTime(&e->stage_ms[ORBX_NSTAGES], e->ev[0], e->ev[ORBX_NSTAGES]);
}

extern "C" int orbx_upload(orbx_extractor* e, int nframes, const uint8_t* imgs, int rows, int cols,
                           size_t step, size_t frame_stride)
{
    if (!e || !imgs || nframes <= 0) { orbx_set_error("bad upload arguments"); return ORBX_ERR_ARG; }
    int r = orbx_reserve(e, rows, cols, nframes);
    if (r) return r;
    CUDA_TRY(cudaSetDevice(e->device));
    r = upload_frames(e, 0, nframes, imgs, step, frame_stride, e->stream);
    if (r) return r;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->last_nframes = -nframes;     // uploaded, not yet run
    return ORBX_OK;
}

extern "C" int orbx_run_resident(orbx_extractor* e, int lap0, int lap1)
{
    if (!e || !e->reserved || e->last_nframes == 0) { orbx_set_error("run_resident before upload"); return ORBX_ERR_STATE; }
    CUDA_TRY(cudaSetDevice(e->device));
    const int n = e->last_nframes < 0 ? -e->last_nframes : e->last_nframes;
    e->last_nframes = n;
    int ns = ORBX_NCSTREAMS;
    if (const char* cs = getenv("ORBX_STREAMS")) { const int v = atoi(cs); if (v >= 1 && v <= ORBX_NCSTREAMS) ns = v; }
    if (e->profiling || ns == 1 || n < 4 * ns) return run_pipeline(e, 0, n, lap0, lap1, e->stream);
    // fork: sub-batches on the compute streams (ordered after everything queued on the handle's stream), join back
    CUDA_TRY(cudaEventRecord(e->ev_fork, e->stream));
    const int per = (n + ns - 1) / ns;
    for (int i = 0; i < ns; i++) {
        const int s0 = i * per, m = n - s0 < per ? n - s0 : per;
        if (m <= 0) break;
        CUDA_TRY(cudaStreamWaitEvent(e->cstream[i], e->ev_fork, 0));
        int r = run_pipeline(e, s0, m, lap0, lap1, e->cstream[i]);
        if (r) return r;
        CUDA_TRY(cudaEventRecord(e->ev_join[i], e->cstream[i]));
        CUDA_TRY(cudaStreamWaitEvent(e->stream, e->ev_join[i], 0));
    }
    return ORBX_OK;
}

extern "C" int orbx_sync(orbx_extractor* e)
{
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    collect_stage_times(e);
    return ORBX_OK;
}

static int download_results(orbx_extractor* e, int slot0, int nframes, orbx_keypoint* kps, uint8_t* desc, int cap,
                            int* counts_host /* 2*nframes */, cudaStream_t st)
{
    const FrameGeom& G = e->G;
    const int ncopy = cap < G.out_cap ? cap : G.out_cap;
    CUDA_TRY(cudaMemcpyAsync(counts_host, e->D.counts + 2 * slot0, sizeof(int) * 2 * nframes, cudaMemcpyDeviceToHost, st));
    if (cap == G.out_cap) {       // caller's per-frame capacity equals the device layout: two linear copies
        CUDA_TRY(cudaMemcpyAsync(kps, e->D.kps + (size_t)slot0 * G.out_cap, (size_t)nframes * cap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(desc, e->D.desc + (size_t)slot0 * G.out_cap * 32, (size_t)nframes * cap * 32, cudaMemcpyDeviceToHost, st));
        return 0;
    }
    CUDA_TRY(cudaMemcpy2DAsync(kps, (size_t)cap * sizeof(orbx_keypoint), e->D.kps + (size_t)slot0 * G.out_cap,
                               (size_t)G.out_cap * sizeof(orbx_keypoint), (size_t)ncopy * sizeof(orbx_keypoint), nframes,
                               cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpy2DAsync(desc, (size_t)cap * 32, e->D.desc + (size_t)slot0 * G.out_cap * 32, (size_t)G.out_cap * 32,
                               (size_t)ncopy * 32, nframes, cudaMemcpyDeviceToHost, st));
    return 0;
}

extern "C" int orbx_download(orbx_extractor* e, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono)
{
    if (!e || !e->reserved || e->last_nframes <= 0) { orbx_set_error("download before run"); return ORBX_ERR_STATE; }
    CUDA_TRY(cudaSetDevice(e->device));
    const int n = e->last_nframes;
    std::vector<int> counts(2 * (size_t)n);
    int r = download_results(e, 0, n, kps, desc, cap, counts.data(), e->stream);
    if (r) return r;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    collect_stage_times(e);
    int rc = ORBX_OK;
    for (int f = 0; f < n; f++) {
        if (nkp) nkp[f] = counts[2 * f];
        if (mono) mono[f] = counts[2 * f + 1];
        if (counts[2 * f] > cap) rc = ORBX_ERR_CAPACITY;
    }
    if (rc) orbx_set_error("keypoint capacity %d too small", cap);
    return rc;
}

// raw frame slots for the prepare path: room for `batch` frames of raw_rows x raw_pitch bytes
static int reserve_raw(orbx_extractor* e, int batch)
{
    const size_t need = (size_t)batch * e->raw_rows * e->raw_pitch + 64;
    if (e->d_raw && e->raw_bytes >= need) return 0;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (e->d_raw) cudaFree(e->d_raw);
    e->d_raw = nullptr; e->raw_bytes = 0;
    CUDA_TRY(cudaMalloc((void**)&e->d_raw, need));
    e->raw_bytes = need;
    return 0;
}

static int upload_raw(orbx_extractor* e, int slot0, int nframes, const uint8_t* imgs, size_t step, size_t frame_stride,
                      cudaStream_t st)
{
    const size_t fbytes = (size_t)e->raw_rows * e->raw_pitch;
    uint8_t* dst = e->d_raw + (size_t)slot0 * fbytes;
    if (frame_stride == step * (size_t)e->raw_rows && step == (size_t)e->raw_pitch) {
        CUDA_TRY(cudaMemcpyAsync(dst, imgs, (size_t)nframes * fbytes, cudaMemcpyHostToDevice, st));
    } else {
        for (int f = 0; f < nframes; f++)
            CUDA_TRY(cudaMemcpy2DAsync(dst + (size_t)f * fbytes, e->raw_pitch, imgs + (size_t)f * frame_stride, step,
                                       (size_t)e->raw_cols * e->raw_ch, e->raw_rows, cudaMemcpyHostToDevice, st));
    }
    return 0;
}

// validates a prepare call and fixes the raw layout in the handle; *rows/*cols = size of the prepared (gray,
// rectified) frame = the frame operator() sees.  Returns 1 when the frames need no preparation at all.
static int prepare_begin(orbx_extractor* e, const uint8_t* imgs, int src_rows, int src_cols, int channels, int rgb,
                         size_t step, size_t frame_stride, int nframes, int* rows, int* cols)
{
    if (!e) { orbx_set_error("null handle"); return ORBX_ERR_ARG; }
    if (!imgs || src_rows <= 0 || src_cols <= 0 || nframes <= 0) { orbx_set_error("empty image"); return ORBX_EMPTY_IMAGE; }
    if (channels != 1 && channels != 3 && channels != 4) {          // src/Tracking.cc:1264-1290 handles 3 and 4
        orbx_set_error("prepare: %d channels (1, 3 or 4 expected)", channels); return ORBX_ERR_ARG;
    }
    if (step < (size_t)src_cols * channels) { orbx_set_error("prepare: step smaller than a row"); return ORBX_ERR_ARG; }
    *rows = e->d_map ? e->map_rows : src_rows;
    *cols = e->d_map ? e->map_cols : src_cols;
    if (!e->d_map && channels == 1) return 1;
    e->raw_rows = src_rows; e->raw_cols = src_cols; e->raw_ch = channels; e->raw_rgb = rgb != 0;
    const bool packed = frame_stride == step * (size_t)src_rows || nframes == 1;
    e->raw_pitch = packed ? (int)step : (src_cols * channels + 15) / 16 * 16;
    return 0;
}

// One chunk, few frames (the per-frame call of a SLAM front end): everything on the handle's stream, host side kept off
// the critical path.  Frames go through a pinned staging buffer (one host memcpy, then a truly asynchronous H2D copy
// that the first kernels queue behind), results land in a pinned zone by two asynchronous D2H copies, ONE stream
// synchronise, then only the nkp valid rows are copied into the caller's (pageable) arrays.
#define ORBX_LAND_MAX (4u << 20)
static bool orbx_no_land()
{
    static const bool v = [] { const char* c = getenv("ORBX_NO_LAND"); return c && atoi(c) != 0; }();
    return v;
}

static int pinned_reserve(uint8_t** p, size_t* cap, size_t need)
{
    if (need <= *cap) return 0;
    if (*p) cudaFreeHost(*p);
    *p = nullptr; *cap = 0;
    CUDA_TRY(cudaHostAlloc((void**)p, need, cudaHostAllocDefault));
    *cap = need;
    return 0;
}

// run_pipeline on frame slots [0, nframes) of the handle's stream through a cached CUDA graph
static bool orbx_no_graph()
{
    static const bool v = [] { const char* c = getenv("ORBX_NO_GRAPH"); return c && atoi(c) != 0; }();
    return v;
}

// The pipeline of frame slots [0, nframes) as a DAG (only ever CAPTURED into a graph): the levels' FAST + quadtree
// kernels hang off the pyramid chain level by level on their own streams instead of waiting for the whole pyramid, the
// blur runs beside them.  Critical path of one 752x480 frame: level 0 -> FAST(0) -> quadtree(0) -> assemble -> describe.
static int run_pipeline_dag(orbx_extractor* e, int nframes, int lap0, int lap1)
{
    FrameGeom G; DevBuffers D;
    slot_view(e, 0, &G, &D);
    cudaStream_t s = e->stream;
    for (int l = 0; l < G.nlevels; l++) {
        e->launches += launch_pyramid_level(G, D, nframes, l, s);
        CUDA_TRY(cudaEventRecord(e->dag_ev[l], s));
        cudaStream_t b = e->dag_stream[l];
        CUDA_TRY(cudaStreamWaitEvent(b, e->dag_ev[l], 0));
        e->launches += launch_fast(G, D, nframes, b, l);
        e->launches += launch_octree(G, D, nframes, b, l);
        CUDA_TRY(cudaEventRecord(e->dag_done[l], b));
    }
    e->launches += launch_blur(G, D, nframes, s);
    for (int l = 0; l < G.nlevels; l++) CUDA_TRY(cudaStreamWaitEvent(s, e->dag_done[l], 0));
    e->launches += launch_assemble(G, D, nframes, lap0, lap1, s);
    e->launches += launch_describe(G, D, nframes, s);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static bool orbx_no_dag()
{
    static const bool v = [] { const char* c = getenv("ORBX_NO_DAG"); return c && atoi(c) != 0; }();
    return v;
}

static int run_pipeline_graph(orbx_extractor* e, int nframes, int lap0, int lap1)
{
    if (e->profiling || orbx_no_graph()) return run_pipeline(e, 0, nframes, lap0, lap1, e->stream);
    const int key[4] = {nframes, lap0, lap1, e->D.in_pitch};
    if (!e->graph_exec || memcmp(key, e->graph_key, sizeof(key)) != 0) {
        if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
        cudaGraph_t g = nullptr;
        const bool dag = fast_per_level(e->G) && !orbx_no_dag() && nframes * e->G.nlevels <= 64;
        if (dag && !e->dag_ready) {                       // branch streams / events exist before the capture starts
            for (int l = 0; l < ORBX_MAX_LEVELS; l++) {
                CUDA_TRY(cudaStreamCreateWithFlags(&e->dag_stream[l], cudaStreamNonBlocking));
                CUDA_TRY(cudaEventCreateWithFlags(&e->dag_ev[l], cudaEventDisableTiming));
                CUDA_TRY(cudaEventCreateWithFlags(&e->dag_done[l], cudaEventDisableTiming));
            }
            e->dag_ready = true;
        }
        CUDA_TRY(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
        const long long l0 = e->launches;
        const int r = dag ? run_pipeline_dag(e, nframes, lap0, lap1) : run_pipeline(e, 0, nframes, lap0, lap1, e->stream);
        const cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
        e->graph_launches = (int)(e->launches - l0);
        e->launches = l0;
        if (r) { if (g) cudaGraphDestroy(g); return r; }
        CUDA_TRY(ce);
        const cudaError_t ci = cudaGraphInstantiate(&e->graph_exec, g, 0);
        cudaGraphDestroy(g);
        CUDA_TRY(ci);
        memcpy(e->graph_key, key, sizeof(key));
    }
    CUDA_TRY(cudaGraphLaunch(e->graph_exec, e->stream));
    e->launches += e->graph_launches;
    return 0;
}

// stage the frames, upload, run the pipeline: all asynchronous on the handle's stream
static int small_enqueue(orbx_extractor* e, int nframes, const uint8_t* imgs, size_t step, size_t frame_stride, int lap0, int lap1)
{
    const FrameGeom& G = e->G;
    const size_t pitch = (size_t)(G.cols + 15) / 16 * 16, fbytes = pitch * G.rows;
    const uint8_t* src_all = imgs;
    size_t sstep = step, sstride = frame_stride;
    cudaPointerAttributes pa;
    const bool pinned = cudaPointerGetAttributes(&pa, imgs) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    if (!pinned) {                                       // a caller that keeps its frames in pinned memory skips the staging copy
        cudaGetLastError();
        int r = pinned_reserve(&e->h_in, &e->h_in_cap, fbytes * nframes);
        if (r) return r;
        for (int f = 0; f < nframes; f++) {
            const uint8_t* src = imgs + (size_t)f * frame_stride;
            uint8_t* dst = e->h_in + (size_t)f * fbytes;
            if (step == pitch) memcpy(dst, src, fbytes);
            else for (int y = 0; y < G.rows; y++) memcpy(dst + (size_t)y * pitch, src + (size_t)y * step, G.cols);
        }
        src_all = e->h_in; sstep = pitch; sstride = fbytes;
    }
    int r = upload_frames(e, 0, nframes, src_all, sstep, sstride, e->stream);
    if (r) return r;
    r = run_pipeline_graph(e, nframes, lap0, lap1);
    if (r) return r;
    e->last_nframes = nframes;
    return 0;
}

// results -> the pinned landing zone {keypoints | descriptors | extra}, asynchronous on the handle's stream
static int small_download(orbx_extractor* e, int nframes, size_t extra_bytes)
{
    const int oc = e->G.out_cap;
    const size_t kb = (size_t)nframes * oc * sizeof(orbx_keypoint), db = (size_t)nframes * oc * 32;
    if (e->h_counts_cap < 2 * (size_t)nframes) {
        if (e->h_counts) cudaFreeHost(e->h_counts);
        e->h_counts = nullptr; e->h_counts_cap = 0;
        CUDA_TRY(cudaHostAlloc((void**)&e->h_counts, sizeof(int) * 2 * (size_t)nframes, cudaHostAllocDefault));
        e->h_counts_cap = 2 * (size_t)nframes;
    }
    int r = pinned_reserve(&e->h_land, &e->h_land_cap, kb + db + extra_bytes);
    if (r) return r;
    cudaStream_t st = e->stream;
    CUDA_TRY(cudaMemcpyAsync(e->h_counts, e->D.counts, sizeof(int) * 2 * nframes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(e->h_land, e->D.kps, kb, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(e->h_land + kb, e->D.desc, db, cudaMemcpyDeviceToHost, st));
    return 0;
}

// after the stream synchronise: counts and the valid rows into the caller's arrays
static int small_finish(orbx_extractor* e, int nframes, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono)
{
    const int oc = e->G.out_cap;
    const size_t kb = (size_t)nframes * oc * sizeof(orbx_keypoint);
    const int* counts = e->h_counts;
    int rc = ORBX_OK;
    for (int f = 0; f < nframes; f++) {
        const int n = counts[2 * f];
        if (nkp) nkp[f] = n;
        if (mono) mono[f] = counts[2 * f + 1];
        if (n > cap) rc = ORBX_ERR_CAPACITY;
        const int m = n < cap ? n : cap;
        memcpy(kps + (size_t)f * cap, e->h_land + (size_t)f * oc * sizeof(orbx_keypoint), (size_t)m * sizeof(orbx_keypoint));
        memcpy(desc + (size_t)f * cap * 32, e->h_land + kb + (size_t)f * oc * 32, (size_t)m * 32);
    }
    if (rc) orbx_set_error("keypoint capacity %d too small", cap);
    return rc;
}

static int extract_small(orbx_extractor* e, int nframes, const uint8_t* imgs, size_t step, size_t frame_stride,
                         int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono)
{
    int r = small_enqueue(e, nframes, imgs, step, frame_stride, lap0, lap1);
    if (r) return r;
    r = small_download(e, nframes, 0);
    if (r) return r;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    collect_stage_times(e);
    return small_finish(e, nframes, kps, desc, cap, nkp, mono);
}

// lazy sizing of the handle for rows x cols frames: bound the device footprint to ~8 GB
static int ensure_reserved(orbx_extractor* e, int rows, int cols, int nframes)
{
    if (e->reserved && e->G.rows == rows && e->G.cols == cols && (e->G.batch >= nframes || e->G.batch >= e->batch_limit)) return 0;
    FrameGeom G;
    std::vector<std::vector<ResizeTapX>> tx; std::vector<std::vector<ResizeTapY>> ty;
    int r = orbx_compute_geometry(e, rows, cols, 1, &G, &tx, &ty);
    if (r) { orbx_set_error("image %dx%d too small for %d pyramid levels (35-px cell grid)", cols, rows, e->nlevels); return r; }
    long long per = (long long)rows * ((cols + 127) / 128 * 128) + G.keys_per_frame * 12;
    for (int l = 0; l < G.nlevels; l++) per += 2 * G.L[l].plane_bytes;
    long long maxslots = 512, budget = 8LL << 30;
    if (const char* sl = getenv("ORBX_SLOTS")) { const long long v = atoll(sl); if (v >= 1 && v <= 8192) { maxslots = v; budget = 64LL << 30; } }
    long long b = budget / per;
    if (b < 1) b = 1;
    if (b > maxslots) b = maxslots;
    e->batch_limit = (int)b;                                 // a later, larger batch grows the slots up to here
    if (b > nframes) b = nframes;
    if (e->reserved && e->G.rows == rows && e->G.cols == cols && b < e->G.batch) b = e->G.batch;
    return orbx_reserve(e, rows, cols, (int)b);
}

static int extract_batch_impl(orbx_extractor* e, int nframes, const uint8_t* imgs, int rows, int cols,
                              size_t step, size_t frame_stride, int lap0, int lap1,
                              orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono, bool raw)
{
    if (!e) { orbx_set_error("null handle"); return ORBX_ERR_ARG; }
    if (!imgs || rows <= 0 || cols <= 0 || nframes <= 0) { orbx_set_error("empty image"); return ORBX_EMPTY_IMAGE; }
    if (!kps || !desc || cap <= 0 || (!raw && step < (size_t)cols)) { orbx_set_error("bad output buffers / step"); return ORBX_ERR_ARG; }
    int r = ensure_reserved(e, rows, cols, nframes);
    if (r) return r;
    CUDA_TRY(cudaSetDevice(e->device));
    const int B = e->G.batch;
    int rc = ORBX_OK;
    if (raw) {
        r = reserve_raw(e, B);
        if (r) return r;
        e->D.in_pitch = (cols + 15) / 16 * 16;               // the prepare kernel writes whole words
    }
    // Chunked three-stream ring pipeline: the frames are cut into chunks that cycle through the R chunk
    // slots of the device buffers; upload of chunk k+1, kernels of chunk k and download of chunk k-1
    // overlap.  Hazards of slot reuse: upload(k) waits for kernels(k-R) (input consumed), kernels(k) wait
    // for upload(k) and download(k-R) (outputs of the slot copied out), download(k) waits for kernels(k).
    int chunk = 128;
    if (const char* cs = getenv("ORBX_CHUNK")) { const int v = atoi(cs); if (v > 0) chunk = v; }
    // at least 4 chunks: upload / kernels / download of neighbouring chunks overlap also for small batches of big frames
    const bool laps = nframes > B;                           // more frames than device slots: the slots are reused
    const int span = laps ? B : nframes;
    if (chunk > (span + 3) / 4) chunk = (span + 3) / 4;
    if (chunk < 1) chunk = 1;
    // A batch that fits keeps frame f in slot f: chunk k sits at slot offset k*chunk and the last one may be short, so
    // ceil(span/chunk) chunk slots never reach past slot B.  A lapping batch cycles through the floor(B/chunk) whole ones.
    const int R = laps ? (B / chunk > 0 ? B / chunk : 1) : (span + chunk - 1) / chunk;
    int ncs = ORBX_NCSTREAMS;
    if (const char* cs_ = getenv("ORBX_STREAMS")) { const int v = atoi(cs_); if (v >= 1 && v <= ORBX_NCSTREAMS) ncs = v; }
    while (R % ncs) ncs--;                                   // a slot must always be served by the same compute stream
    const int nchunks = (nframes + chunk - 1) / chunk;
    const bool multi = nchunks > 1;
    while ((int)e->ev_up.size() < R) {
        cudaEvent_t a, b, c;
        CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&c, cudaEventDisableTiming));
        e->ev_up.push_back(a); e->ev_done.push_back(b); e->ev_down.push_back(c);
    }
    if (e->h_counts_cap < 2 * (size_t)nframes) {
        if (e->h_counts) cudaFreeHost(e->h_counts);
        e->h_counts = nullptr; e->h_counts_cap = 0;
        CUDA_TRY(cudaHostAlloc((void**)&e->h_counts, sizeof(int) * 2 * (size_t)nframes, cudaHostAllocDefault));
        e->h_counts_cap = 2 * (size_t)nframes;
    }
    int* counts = e->h_counts;
    if (!multi && !raw && (size_t)nframes * e->G.out_cap * (sizeof(orbx_keypoint) + 32) <= ORBX_LAND_MAX && !orbx_no_land())
        return extract_small(e, nframes, imgs, step, frame_stride, lap0, lap1, kps, desc, cap, nkp, mono);
    cudaStream_t up = multi ? e->stream_h2d : e->stream, dn = multi ? e->stream_d2h : e->stream;
    // Ramp: the upload of the first chunk and the kernels + download of the last one have nothing to overlap with, so those two
    // chunks are cut into pieces of 1/8, 1/8, 1/4, 1/2 (mirrored at the end) that move through the same slot one behind the
    // other: only 1/8 of a chunk is exposed at either end of a call (C1: 0.9 ms -> 0.1 ms of a 16 ms call of 2048 frames).
    // Zero-copy experiment (ORBX_ZEROCOPY=1): frames in mapped pinned host memory are read by the level-0 kernel itself, no copy engine
    bool zerocopy = false; const uint8_t* zc_base = nullptr;
    if (const char* zs = getenv("ORBX_ZEROCOPY")) {
        cudaPointerAttributes pa;
        if (atoi(zs) != 0 && !raw && multi && cudaPointerGetAttributes(&pa, imgs) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer) {
            zerocopy = true; zc_base = (const uint8_t*)pa.devicePointer;
        }
        cudaGetLastError();
    }
    bool ramp = false;                                       // measured on B200: 124.5 k vs 125.9 k frames/s without (C1) - off
    if (const char* rs = getenv("ORBX_RAMP")) ramp = multi && !e->profiling && atoi(rs) != 0;
    while (ramp && (int)e->ev_piece.size() < 8) {
        cudaEvent_t a;
        CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        e->ev_piece.push_back(a);
    }
    for (int k = 0; k < nchunks; k++) {
        const int fr0 = k * chunk, n = nframes - fr0 < chunk ? nframes - fr0 : chunk;
        const int slot = k % R, s0 = slot * chunk;
        cudaStream_t cs = (multi && !e->profiling) ? e->cstream[k % ncs] : e->stream;   // chunk slot k%R is always served by the same stream
        int pc[4] = {n, 0, 0, 0}, np = 1;
        if (ramp && n >= 32 && (k == 0 || k == nchunks - 1)) {
            const int h = n / 2, q = n / 4, o = n / 8, rest = n - h - q - o;
            if (k == 0) { pc[0] = rest; pc[1] = o; pc[2] = q; pc[3] = h; }
            else { pc[0] = h; pc[1] = q; pc[2] = o; pc[3] = rest; }
            np = 4;
        }
        if (multi && k >= R) CUDA_TRY(cudaStreamWaitEvent(up, e->ev_done[slot], 0));
        if (multi && k >= R) CUDA_TRY(cudaStreamWaitEvent(cs, e->ev_down[slot], 0));
        for (int j = 0, off = 0; j < np; off += pc[j], j++) {
            const int pn = pc[j], pf = fr0 + off, ps = s0 + off;
            static const bool skip_h2d = getenv("ORBX_DIAG_NO_H2D") != nullptr, skip_d2h = getenv("ORBX_DIAG_NO_D2H") != nullptr;   // diagnostics
            r = (skip_h2d || zerocopy) ? 0 : raw ? upload_raw(e, ps, pn, imgs + (size_t)pf * frame_stride, step, frame_stride, up)
                    : upload_frames(e, ps, pn, imgs + (size_t)pf * frame_stride, step, frame_stride, up);
            if (r) return r;
            if (multi) {
                cudaEvent_t evu = np > 1 ? e->ev_piece[(size_t)(2 * j)] : e->ev_up[slot];
                CUDA_TRY(cudaEventRecord(evu, up));
                CUDA_TRY(cudaStreamWaitEvent(cs, evu, 0));
            }
            if (raw) e->launches += launch_prepare(e, ps, pn, cs);   // remap + gray into the slot's frames
            static const bool skip_kernels = getenv("ORBX_DIAG_COPIES_ONLY") != nullptr;   // diagnostic: the ring's copy rate alone
            r = skip_kernels ? 0 : zerocopy ? run_pipeline(e, ps, pn, lap0, lap1, cs, zc_base + (size_t)pf * frame_stride, step, frame_stride)
                                            : run_pipeline(e, ps, pn, lap0, lap1, cs);
            if (r) return r;
            if (multi) {
                cudaEvent_t evd = np > 1 ? e->ev_piece[(size_t)(2 * j + 1)] : e->ev_done[slot];
                CUDA_TRY(cudaEventRecord(evd, cs));
                CUDA_TRY(cudaStreamWaitEvent(dn, evd, 0));
            }
            r = skip_d2h ? (int)cudaMemcpyAsync(counts + 2 * (size_t)pf, e->D.counts + 2 * ps, sizeof(int) * 2 * pn, cudaMemcpyDeviceToHost, dn)
                         : download_results(e, ps, pn, kps + (size_t)pf * cap, desc + (size_t)pf * cap * 32, cap, counts + 2 * (size_t)pf, dn);
            if (r) return r;
        }
        if (multi) {
            if (np > 1) {                                     // the slot's own events mark the end of its last piece
                CUDA_TRY(cudaEventRecord(e->ev_up[slot], up));
                CUDA_TRY(cudaEventRecord(e->ev_done[slot], cs));
            }
            CUDA_TRY(cudaEventRecord(e->ev_down[slot], dn));
        }
    }
    if (multi) {
        // callers (timer, stereo match) order against the handle's stream: join the copy streams into it
        CUDA_TRY(cudaStreamWaitEvent(e->stream, e->ev_down[(nchunks - 1) % R], 0));
        CUDA_TRY(cudaStreamSynchronize(e->stream_d2h));
    }
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    collect_stage_times(e);
    // resident pyramids / outputs (orbx_stereo_match_batch, orbx_get_pyramid_level, the debug getters): every frame of a
    // batch that fits; a lapping batch leaves a mix of passes in the slots, so nothing is reported resident
    e->last_nframes = laps ? 0 : nframes;
    for (int f = 0; f < nframes; f++) {
        if (nkp) nkp[f] = counts[2 * f];
        if (mono) mono[f] = counts[2 * f + 1];
        if (counts[2 * f] > cap) rc = ORBX_ERR_CAPACITY;
    }
    if (rc) orbx_set_error("keypoint capacity %d too small", cap);
    return rc;
}

extern "C" int orbx_extract_batch(orbx_extractor* e, int nframes, const uint8_t* imgs, int rows, int cols,
                                  size_t step, size_t frame_stride, int lap0, int lap1,
                                  orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono)
{
    return extract_batch_impl(e, nframes, imgs, rows, cols, step, frame_stride, lap0, lap1, kps, desc, cap, nkp, mono, false);
}

// ---- one host batch over several devices (north_star: "partitioning frame batches across the 8 GPUs of one box with
// per-GPU CUDA streams"; the reference's only analogue is the two extraction threads of src/Frame.cc:132-137) ----
// handles[d] may live on any device (several on one device are allowed).  Frames are split into contiguous ranges, the
// first nframes % ndev handles take one frame more (the rule of orb-slam3_b200/shard.py); every range is driven by its
// own host thread through orbx_extract_batch, results land in the caller's arrays in frame order.  No collective, no
// device-to-device traffic: frames are independent.
extern "C" int orbx_extract_batch_multi(orbx_extractor* const* handles, int ndev, int nframes, const uint8_t* imgs, int rows, int cols,
                                        size_t step, size_t frame_stride, int lap0, int lap1,
                                        orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono)
{
    if (!handles || ndev <= 0) { orbx_set_error("no extractor handles"); return ORBX_ERR_ARG; }
    for (int d = 0; d < ndev; d++) {
        if (!handles[d]) { orbx_set_error("handle %d is null", d); return ORBX_ERR_ARG; }
        for (int k = 0; k < d; k++) if (handles[k] == handles[d]) { orbx_set_error("handle %d is used twice: a handle is not re-entrant", d); return ORBX_ERR_ARG; }
    }
    if (!imgs || rows <= 0 || cols <= 0 || nframes <= 0) { orbx_set_error("empty image"); return ORBX_EMPTY_IMAGE; }
    if (!kps || !desc || cap <= 0 || !nkp || !mono || step < (size_t)cols) { orbx_set_error("bad output buffers / step"); return ORBX_ERR_ARG; }
    std::vector<int> rc((size_t)ndev, ORBX_OK);
    std::vector<std::string> msg((size_t)ndev);
    std::vector<std::thread> th;
    const int q = nframes / ndev, r = nframes % ndev;
    int f0 = 0;
    for (int d = 0; d < ndev; d++) {
        const int n = q + (d < r ? 1 : 0);
        if (n == 0) continue;
        const int first = f0;
        th.emplace_back([=, &rc, &msg] {
            rc[(size_t)d] = orbx_extract_batch(handles[d], n, imgs + (size_t)first * frame_stride, rows, cols, step, frame_stride, lap0, lap1,
                                              kps + (size_t)first * cap, desc + (size_t)first * cap * 32, cap, nkp + first, mono + first);
            if (rc[(size_t)d] != ORBX_OK) msg[(size_t)d] = orbx_last_error();       // the message is thread-local: carry it over
        });
        f0 += n;
    }
    for (std::thread& t : th) t.join();
    for (int d = 0; d < ndev; d++)
        if (rc[(size_t)d] != ORBX_OK) { orbx_set_error("device range %d: %s", d, msg[(size_t)d].c_str()); return rc[(size_t)d]; }
    return ORBX_OK;
}

// ---- frame preparation in front of operator(): remap + cvtColor on the device (prepare.cu) ----

extern "C" int orbx_set_rectify_maps(orbx_extractor* e, const float* mapx, const float* mapy, int rows, int cols, size_t map_step)
{
    if (!e) { orbx_set_error("null handle"); return ORBX_ERR_ARG; }
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (e->d_map) { cudaFree(e->d_map); e->d_map = nullptr; e->map_rows = e->map_cols = 0; }
    if (!mapx && !mapy) return ORBX_OK;                       // cleared: frames are used as they come
    if (!mapx || !mapy || rows <= 0 || cols <= 0 || map_step < sizeof(float) * (size_t)cols || map_step % sizeof(float)) {
        orbx_set_error("bad rectification maps (CV_32FC1 mapx and mapy of the rectified size expected)"); return ORBX_ERR_ARG;
    }
    float* tmp = nullptr;
    const size_t plane = map_step * (size_t)rows;
    CUDA_TRY(cudaMalloc((void**)&tmp, 2 * plane));
    cudaError_t ce = cudaMemcpyAsync(tmp, mapx, plane, cudaMemcpyHostToDevice, e->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync((char*)tmp + plane, mapy, plane, cudaMemcpyHostToDevice, e->stream);
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&e->d_map, sizeof(int2) * (size_t)rows * cols);
    if (ce == cudaSuccess) {
        e->launches += launch_rectify_pack(tmp, (const float*)((char*)tmp + plane), rows, cols, (int)(map_step / sizeof(float)), e->d_map, e->stream);
        ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
    cudaFree(tmp);
    if (ce != cudaSuccess) {
        if (e->d_map) { cudaFree(e->d_map); e->d_map = nullptr; }
        orbx_set_error("rectification map upload failed: %s", cudaGetErrorString(ce)); return ORBX_ERR_CUDA;
    }
    e->map_rows = rows; e->map_cols = cols;
    return ORBX_OK;
}

extern "C" int orbx_prepare_extract_batch(orbx_extractor* e, int nframes, const uint8_t* imgs, int src_rows, int src_cols,
                                          int channels, int rgb, size_t step, size_t frame_stride, int lap0, int lap1,
                                          orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono)
{
    int rows = 0, cols = 0;
    const int p = prepare_begin(e, imgs, src_rows, src_cols, channels, rgb, step, frame_stride, nframes, &rows, &cols);
    if (p < 0) return p;
    return extract_batch_impl(e, nframes, imgs, rows, cols, step, frame_stride, lap0, lap1, kps, desc, cap, nkp, mono, p == 0);
}

extern "C" int orbx_prepare_extract(orbx_extractor* e, const uint8_t* img, int src_rows, int src_cols, int channels, int rgb,
                                    size_t step, int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp)
{
    int n = 0, mono = 0;
    if (nkp) *nkp = 0;
    int r = orbx_prepare_extract_batch(e, 1, img, src_rows, src_cols, channels, rgb, step, step * (size_t)(src_rows > 0 ? src_rows : 0),
                                       lap0, lap1, kps, desc, cap, &n, &mono);
    if (nkp) *nkp = n;
    if (r < 0) return r;
    return mono;
}

// resident twin of orbx_upload: raw frames -> raw slots -> prepare kernel -> frame slots (then orbx_run_resident)
extern "C" int orbx_prepare_upload(orbx_extractor* e, int nframes, const uint8_t* imgs, int src_rows, int src_cols,
                                   int channels, int rgb, size_t step, size_t frame_stride)
{
    int rows = 0, cols = 0;
    const int p = prepare_begin(e, imgs, src_rows, src_cols, channels, rgb, step, frame_stride, nframes, &rows, &cols);
    if (p < 0) return p;
    if (p == 1) return orbx_upload(e, nframes, imgs, rows, cols, step, frame_stride);
    int r = orbx_reserve(e, rows, cols, nframes);
    if (r) return r;
    CUDA_TRY(cudaSetDevice(e->device));
    r = reserve_raw(e, e->G.batch);
    if (r) return r;
    e->D.in_pitch = (cols + 15) / 16 * 16;
    r = upload_raw(e, 0, nframes, imgs, step, frame_stride, e->stream);
    if (r) return r;
    e->launches += launch_prepare(e, 0, nframes, e->stream);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->last_nframes = -nframes;
    return ORBX_OK;
}

// the prepare kernel alone over the resident raw frames (kernel timing)
extern "C" int orbx_prepare_resident(orbx_extractor* e)
{
    if (!e || !e->reserved || !e->d_raw || e->last_nframes == 0) { orbx_set_error("prepare_resident before prepare_upload"); return ORBX_ERR_STATE; }
    CUDA_TRY(cudaSetDevice(e->device));
    const int n = e->last_nframes < 0 ? -e->last_nframes : e->last_nframes;
    e->launches += launch_prepare(e, 0, n, e->stream);
    CUDA_TRY(cudaGetLastError());
    return ORBX_OK;
}

extern "C" int orbx_get_prepared(orbx_extractor* e, int frame, uint8_t* dst, size_t dst_step)
{
    if (!e || !e->reserved || e->last_nframes == 0) { orbx_set_error("no frame resident"); return ORBX_ERR_STATE; }
    const int n = e->last_nframes < 0 ? -e->last_nframes : e->last_nframes;
    if (frame < 0 || frame >= n || !dst || dst_step < (size_t)e->G.cols) { orbx_set_error("bad frame / destination"); return ORBX_ERR_ARG; }
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaMemcpy2DAsync(dst, dst_step, e->D.input + (size_t)frame * e->G.rows * e->D.in_pitch, e->D.in_pitch,
                               e->G.cols, e->G.rows, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return ORBX_OK;
}

extern "C" int orbx_extract(orbx_extractor* e, const uint8_t* img, int rows, int cols, size_t step,
                            int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp)
{
    int n = 0, mono = 0;
    if (nkp) *nkp = 0;
    int r = orbx_extract_batch(e, 1, img, rows, cols, step, step * (size_t)(rows > 0 ? rows : 0), lap0, lap1, kps, desc, cap, &n, &mono);
    if (nkp) *nkp = n;
    if (r < 0) return r;
    return mono;
}

// The stereo Frame constructor's hot part in ONE call (src/Frame.cc:132-137 two ExtractORB threads, :154
// ComputeStereoMatches): both extractions run concurrently on the two handles' streams, the match is queued behind them,
// the right eye's results travel back while the match kernels run; the host blocks once.
extern "C" int orbx_extract_stereo(orbx_extractor* eL, orbx_extractor* eR, const uint8_t* imgL, size_t stepL,
                                   const uint8_t* imgR, size_t stepR, int rows, int cols,
                                   orbx_keypoint* kpsL, uint8_t* descL, int* nL,
                                   orbx_keypoint* kpsR, uint8_t* descR, int* nR, int cap,
                                   float mb, float mbf, float* uRight, float* depth)
{
    if (nL) *nL = 0;
    if (nR) *nR = 0;
    if (!eL || !eR || eL == eR) { orbx_set_error("two distinct extractor handles needed"); return ORBX_ERR_ARG; }
    if (!imgL || !imgR || rows <= 0 || cols <= 0) { orbx_set_error("empty image"); return ORBX_EMPTY_IMAGE; }
    if (!kpsL || !descL || !kpsR || !descR || !uRight || !depth || cap <= 0 || stepL < (size_t)cols || stepR < (size_t)cols) {
        orbx_set_error("bad output buffers / step"); return ORBX_ERR_ARG;
    }
    if (eL->device != eR->device) { orbx_set_error("left/right extractors live on different devices"); return ORBX_ERR_ARG; }
    int r = ensure_reserved(eL, rows, cols, 1);
    if (!r) r = ensure_reserved(eR, rows, cols, 1);
    if (r) return r;
    CUDA_TRY(cudaSetDevice(eL->device));
    r = small_enqueue(eL, 1, imgL, stepL, stepL * (size_t)rows, 0, 0);
    if (!r) r = small_enqueue(eR, 1, imgR, stepR, stepR * (size_t)rows, 0, 0);
    if (!r) r = orbx_stereo_match_resident(eL, eR, mb, mbf);       // orders eL's stream behind eR's pipeline
    if (r) return r;
    const int oc = eL->G.out_cap;
    r = small_download(eR, 1, 0);
    if (!r) r = small_download(eL, 1, 2 * sizeof(float) * (size_t)oc);
    if (r) return r;
    float* su = reinterpret_cast<float*>(eL->h_land + (size_t)oc * (sizeof(orbx_keypoint) + 32));
    CUDA_TRY(cudaMemcpyAsync(su, eL->d_uright, sizeof(float) * oc, cudaMemcpyDeviceToHost, eL->stream));
    CUDA_TRY(cudaMemcpyAsync(su + oc, eL->d_depth, sizeof(float) * oc, cudaMemcpyDeviceToHost, eL->stream));
    CUDA_TRY(cudaStreamSynchronize(eR->stream));
    int monoR = 0, monoL = 0, cR = 0, cL = 0;
    const int rcR = small_finish(eR, 1, kpsR, descR, cap, &cR, &monoR);
    CUDA_TRY(cudaStreamSynchronize(eL->stream));
    collect_stage_times(eL);
    const int rcL = small_finish(eL, 1, kpsL, descL, cap, &cL, &monoL);
    if (nL) *nL = cL;
    if (nR) *nR = cR;
    const int m = cL < cap ? cL : cap;
    memcpy(uRight, su, sizeof(float) * (size_t)m);
    memcpy(depth, su + oc, sizeof(float) * (size_t)m);
    return rcL ? rcL : rcR;
}

extern "C" int orbx_get_pyramid_level(orbx_extractor* e, int frame, int level, int padded,
                                      uint8_t* dst, size_t dst_step, int* rows, int* cols)
{
    if (!e || !e->reserved || e->last_nframes <= 0 || frame < 0 || frame >= e->last_nframes || level < 0 || level >= e->nlevels) {
        orbx_set_error("pyramid level not available"); return ORBX_ERR_STATE;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const LevelGeom& L = e->G.L[level];
    const int r = padded ? L.h + 2 * ORBX_EDGE : L.h, c = padded ? L.w + 2 * ORBX_EDGE : L.w;
    if (rows) *rows = r;
    if (cols) *cols = c;
    if (!dst) return ORBX_OK;
    if (padded) pyramid_fill_border(L, frame, e->stream);     // the hot path leaves the reflect-101 border unwritten
    const uint8_t* src = L.pyr + (size_t)frame * L.plane_bytes +
                         (padded ? (size_t)(ORBX_XOFF - ORBX_EDGE) : (size_t)ORBX_YOFF * L.pitch + ORBX_XOFF);
    CUDA_TRY(cudaMemcpy2DAsync(dst, dst_step, src, L.pitch, c, r, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return ORBX_OK;
}

extern "C" int orbx_debug_blurred_level(orbx_extractor* e, int frame, int level, uint8_t* dst, size_t dst_step)
{
    if (!e || !e->reserved || e->last_nframes <= 0 || frame < 0 || frame >= e->last_nframes || level < 0 || level >= e->nlevels) {
        orbx_set_error("blurred level not available"); return ORBX_ERR_STATE;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const LevelGeom& L = e->G.L[level];
    const uint8_t* src = L.blur + (size_t)frame * L.plane_bytes + (size_t)ORBX_YOFF * L.pitch + ORBX_XOFF;
    CUDA_TRY(cudaMemcpy2DAsync(dst, dst_step, src, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return ORBX_OK;
}

extern "C" int orbx_debug_level_keys(orbx_extractor* e, int frame, int level, int stage,
                                     int32_t* out, float* angles, int cap)
{
    if (!e || !e->reserved || e->last_nframes <= 0 || frame < 0 || frame >= e->last_nframes || level < 0 || level >= e->nlevels) {
        orbx_set_error("level keys not available"); return ORBX_ERR_STATE;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const FrameGeom& G = e->G;
    const LevelGeom& L = G.L[level];
    int n = 0;
    const int* cntp = (stage == 0 ? e->D.candcnt : e->D.selcnt) + frame * G.nlevels + level;
    CUDA_TRY(cudaMemcpyAsync(&n, cntp, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (!out) return n;
    const int m = n < cap ? n : cap;
    std::vector<orbx_key_t> keys((size_t)m);
    const orbx_key_t* src = stage == 0 ? e->D.cand + (size_t)frame * G.cand_per_frame + L.cand_base
                                     : e->D.sel + (size_t)frame * G.sel_per_frame + L.sel_base;
    if (m > 0) CUDA_TRY(cudaMemcpyAsync(keys.data(), src, sizeof(orbx_key_t) * m, cudaMemcpyDeviceToHost, e->stream));
    if (m > 0 && stage == 1 && angles)
        CUDA_TRY(cudaMemcpyAsync(angles, e->D.angle + (size_t)frame * G.sel_per_frame + L.sel_base, sizeof(float) * m,
                                 cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (stage == 0) {
        // the device list is unordered; present it in the reference's vToDistributeKeys order
        // (cell row-major, raster inside a cell) = ascending canonical rank, as the quadtree sees it
        auto rank = [&L](orbx_key_t k) {
            const int X3 = (int)KEY_X(k) - 3, Y3 = (int)KEY_Y(k) - 3;
            const int cj = X3 / L.wCell, ci = Y3 / L.hCell;
            return ((ci * L.nCols + cj) << 14) | ((Y3 - ci * L.hCell) << 7) | (X3 - cj * L.wCell);
        };
        std::sort(keys.begin(), keys.end(), [&](orbx_key_t a, orbx_key_t b) { return rank(a) < rank(b); });
    }
    for (int i = 0; i < m; i++) {
        out[3 * i] = (int32_t)KEY_X(keys[i]); out[3 * i + 1] = (int32_t)KEY_Y(keys[i]); out[3 * i + 2] = (int32_t)KEY_S(keys[i]);
    }
    return n;
}

extern "C" float orbx_last_stage_ms(orbx_extractor* e, int stage)
{
    if (!e || stage >= ORBX_NSTAGES) return -1.f;
    return e->stage_ms[stage < 0 ? ORBX_NSTAGES : stage];
}
// CUDA-event stopwatch on the handle's stream (bench.py times K steps with it)
extern "C" int orbx_timer_start(orbx_extractor* e)
{
    if (!e) return ORBX_ERR_ARG;
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaEventRecord(e->tev[0], e->stream));
    return ORBX_OK;
}
extern "C" float orbx_timer_stop(orbx_extractor* e)
{
    if (!e) return -1.f;
    float ms = -1.f;
    if (cudaSetDevice(e->device) != cudaSuccess) return -1.f;
    if (cudaEventRecord(e->tev[1], e->stream) != cudaSuccess) return -1.f;
    if (cudaEventSynchronize(e->tev[1]) != cudaSuccess) return -1.f;
    if (cudaEventElapsedTime(&ms, e->tev[0], e->tev[1]) != cudaSuccess) return -1.f;
    return ms;
}
extern "C" void orbx_set_profiling(orbx_extractor* e, int on) { if (e) e->profiling = on != 0; }
extern "C" long long orbx_launch_count(const orbx_extractor* e) { return e ? e->launches : 0; }

extern "C" int orbx_hamming256(const void* a, const void* b)
{
    uint64_t x[4], y[4];
    memcpy(x, a, 32); memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
           __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}
