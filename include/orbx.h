/*
 * orbx.h — C ABI of the B200-native ORB front-end (liborbx.so, CUDA sm_100a).
 *
 * This is the drop-in boundary for the reference's hot path (Wangkkkkkkkk/ORB-SLAM3).  The
 * reference has no FFI: its "operator API" is the C++ class surface of ORBextractor / ORBmatcher /
 * Frame inside libORB_SLAM3.so.  Each entry point below names the reference interface it replaces
 * (file:line relative to the reference tree); orb-slam3_b200/cpp/ holds header-only C++ shims with
 * the reference's own class names and signatures on top of this ABI, and INTEGRATION.md shows the
 * change a maintainer makes.
 *
 * Conventions: plain pointers and sizes only; no exceptions cross the boundary; functions return
 * >= 0 on success and a negative orbx_status on failure, with a thread-local message available from
 * orbx_last_error().  There is NO CPU fallback: without a CUDA device every compute entry point
 * fails with ORBX_ERR_CUDA.
 *
 * Threading (same contract as the reference): one extractor handle = one CUDA
 * stream + buffers; different handles may be used from different threads concurrently; calls on the
 * same handle must be serialised by the caller (the reference never shares an ORBextractor).
 */
#ifndef ORBX_H
#define ORBX_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_MAX_LEVELS 16

typedef enum {
    ORBX_OK = 0,
    ORBX_EMPTY_IMAGE = -1,   /* operator() returns -1 for an empty image, src/ORBextractor.cc:1538 */
    ORBX_ERR_CAPACITY = -2,  /* caller's keypoint/descriptor buffers too small; *nkp holds the need */
    ORBX_ERR_GEOMETRY = -3,  /* image too small for the 35-px cell grid (reference divides by zero) */
    ORBX_ERR_ARG = -4,
    ORBX_ERR_CUDA = -5,
    ORBX_ERR_STATE = -6      /* e.g. stereo match before both extractions */
} orbx_status;

/* cv::KeyPoint memory layout (28 bytes): the shim reinterprets std::vector<cv::KeyPoint> storage. */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbx_keypoint;

typedef struct orbx_extractor orbx_extractor;

const char* orbx_last_error(void);
int orbx_version(void);
/* number of visible CUDA devices (0 when none / no driver); never fails */
int orbx_device_count(void);

/* ---- ORBextractor -------------------------------------------------------------------------- */

/* ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * include/ORBextractor.h:49-50, src/ORBextractor.cc:468-571.  `device` = CUDA ordinal. */
int orbx_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                int device, orbx_extractor** out);
void orbx_destroy(orbx_extractor* e);

/* GetLevels / GetScaleFactor / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares, include/ORBextractor.h:61-81.  Returned arrays have nlevels entries
 * and live as long as the handle. */
int          orbx_levels(const orbx_extractor* e);
float        orbx_scale_factor(const orbx_extractor* e);
const float* orbx_scale_factors(const orbx_extractor* e);
const float* orbx_inv_scale_factors(const orbx_extractor* e);
const float* orbx_level_sigma2(const orbx_extractor* e);
const float* orbx_inv_level_sigma2(const orbx_extractor* e);
const int*   orbx_features_per_level(const orbx_extractor* e);   /* mnFeaturesPerLevel */

/* int ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea)
 * include/ORBextractor.h:57-59, src/ORBextractor.cc:1534-1659.  The mask is ignored by the
 * reference (include/ORBextractor.h:56) and has no parameter here.  `img` is a HOST CV_8UC1 image
 * (rows x cols, `step` bytes per row).  kps/desc are caller-owned with room for `cap` keypoints
 * (desc: cap x 32 bytes); *nkp receives the keypoint count.  Returns monoIndex (>= 0), or
 * ORBX_EMPTY_IMAGE, or an error.  The image pyramid stays resident on the device (mvImagePyramid,
 * include/ORBextractor.h:83) for orbx_stereo_match / orbx_get_pyramid_level. */
int orbx_extract(orbx_extractor* e, const uint8_t* img, int rows, int cols, size_t step,
                 int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp);

/* Throughput twin of operator() (no reference analogue; identical per-frame results): nframes
 * HOST images of equal size, frame f at imgs + f*frame_stride.  Frame f's outputs are written at
 * kps + f*cap, desc + f*cap*32; nkp[f], mono[f] receive the count and the monoIndex. */
int orbx_extract_batch(orbx_extractor* e, int nframes, const uint8_t* imgs, int rows, int cols,
                       size_t step, size_t frame_stride, int lap0, int lap1,
                       orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono);

/* One HOST batch spread over several devices (no reference analogue beyond the two extraction threads of
 * src/Frame.cc:132-137; BASELINE north_star: frame batches partitioned across the GPUs of one box, no collective).
 * handles[0..ndev): extractors created with the same parameters on the devices to use (distinct handles; several may
 * share a device).  Frames are split into contiguous ranges - the first nframes % ndev handles take one more - each
 * driven by its own host thread through orbx_extract_batch; outputs are written in frame order exactly as
 * orbx_extract_batch writes them.  Returns the first failing range's status. */
int orbx_extract_batch_multi(orbx_extractor* const* handles, int ndev, int nframes, const uint8_t* imgs, int rows, int cols,
                             size_t step, size_t frame_stride, int lap0, int lap1,
                             orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono);

/* ---- frame preparation in front of operator() ------------------------- */

/* cv::initUndistortRectifyMap(..., CV_32F, M1, M2) maps of this extractor's camera
 * (Examples/Stereo/stereo_euroc.cc:117-118): HOST CV_32FC1 mapx / mapy of the RECTIFIED size
 * rows x cols, `map_step` bytes per row.  Converted once to remap's internal fixed point and kept on
 * the device.  mapx == mapy == NULL clears the maps (frames are used as they come). */
int orbx_set_rectify_maps(orbx_extractor* e, const float* mapx, const float* mapy, int rows, int cols, size_t map_step);

/* cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) (Examples/Stereo/stereo_euroc.cc:168-169; skipped when
 * no maps are set), then cvtColor(.., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY) (src/Tracking.cc:1264-1290; skipped for
 * channels == 1; rgb != 0 <=> mbRGB, the first channel is R), then operator() on the result, all on the
 * device in one call.  `img`: HOST u8 image src_rows x src_cols x channels (1, 3 or 4), `step` bytes per
 * row.  Outputs and return value as orbx_extract; the prepared frame has the maps' size (or the source's). */
int orbx_prepare_extract(orbx_extractor* e, const uint8_t* img, int src_rows, int src_cols, int channels, int rgb,
                         size_t step, int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp);
/* batch twin (cf. orbx_extract_batch): frame f at imgs + f*frame_stride */
int orbx_prepare_extract_batch(orbx_extractor* e, int nframes, const uint8_t* imgs, int src_rows, int src_cols,
                               int channels, int rgb, size_t step, size_t frame_stride, int lap0, int lap1,
                               orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono);
/* resident twins (kernel timing): raw frames -> device, prepared into the frame slots (then
 * orbx_run_resident); orbx_prepare_resident re-runs the prepare kernel alone on the resident raw frames */
int orbx_prepare_upload(orbx_extractor* e, int nframes, const uint8_t* imgs, int src_rows, int src_cols,
                        int channels, int rgb, size_t step, size_t frame_stride);
int orbx_prepare_resident(orbx_extractor* e);
/* imLeftRect / mImGray of frame slot `frame`: the frame operator() saw (rows x cols of the extractor geometry) */
int orbx_get_prepared(orbx_extractor* e, int frame, uint8_t* dst, size_t dst_step);

/* Upper bound on the keypoints one frame can yield (quota + quadtree overshoot per level). */
int orbx_max_keypoints(const orbx_extractor* e, int rows, int cols);

/* Size the device buffers for `max_batch` frames of rows x cols ahead of time (otherwise done
 * lazily by the first extract call). */
int orbx_reserve(orbx_extractor* e, int rows, int cols, int max_batch);

/* Device-resident staging for kernel-only timing: upload once, run many times, download. */
int orbx_upload(orbx_extractor* e, int nframes, const uint8_t* imgs, int rows, int cols,
                size_t step, size_t frame_stride);
int orbx_run_resident(orbx_extractor* e, int lap0, int lap1);          /* async on the handle's stream */
int orbx_sync(orbx_extractor* e);
int orbx_download(orbx_extractor* e, orbx_keypoint* kps, uint8_t* desc, int cap, int* nkp, int* mono);

/* mvImagePyramid[level] of frame `frame` of the last extract (include/ORBextractor.h:83): copies
 * the unpadded rows x cols level image (padded=0) or the (rows+38) x (cols+38) buffer with its
 * reflect-101 border (padded=1) to dst (dst_step bytes per row).  rows / cols receive the copied size. */
int orbx_get_pyramid_level(orbx_extractor* e, int frame, int level, int padded,
                           uint8_t* dst, size_t dst_step, int* rows, int* cols);

/* Stage dumps for parity tests (no reference analogue).  stage: 0 = FAST candidates of `level` in
 * reference order (vToDistributeKeys, src/ORBextractor.cc:1131-1140), 1 = quadtree selection in
 * list order (:1150).  out receives (x, y, score) int32 triplets, coordinates relative to
 * minBorder=16; angles (stage 1, may be NULL) the IC_Angle of each selected keypoint. */
int orbx_debug_level_keys(orbx_extractor* e, int frame, int level, int stage,
                          int32_t* out, float* angles, int cap);
/* blurred level image (GaussianBlur 7x7 sigma 2 of the level, src/ORBextractor.cc:1606-1614) */
int orbx_debug_blurred_level(orbx_extractor* e, int frame, int level, uint8_t* dst, size_t dst_step);

/* timing of the most recent orbx_run_resident / extract on this handle, CUDA events on the handle's
 * stream, milliseconds; stage < 0 = whole pipeline.  Stages: 0 pyramid, 1 FAST, 2 quadtree,
 * 3 assemble, 4 blur, 5 orientation+descriptor. */
float orbx_last_stage_ms(orbx_extractor* e, int stage);
/* CUDA-event stopwatch on the handle's stream: start records an event, stop records a second one,
 * waits for it and returns the elapsed milliseconds (everything queued on the handle in between). */
int   orbx_timer_start(orbx_extractor* e);
float orbx_timer_stop(orbx_extractor* e);
/* unit-test entry: FAST-9/16 score map (no NMS) of a HOST image at threshold th -> out[rows*cols] */
int   orbx_debug_fast_score_map(int device, const uint8_t* img, int rows, int cols, size_t step, int th, uint8_t* out);
/* enable per-stage event recording (off by default: events serialise nothing but cost launches) */
void orbx_set_profiling(orbx_extractor* e, int on);
/* number of kernel launches issued by this handle since creation */
long long orbx_launch_count(const orbx_extractor* e);

/* ---- ORBmatcher / Frame -------------------------------------------------------------------- */

/* static int ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)
 * include/ORBmatcher.h:42, src/ORBmatcher.cc:2911-2931: 256-bit Hamming distance (host inline). */
int orbx_hamming256(const void* a, const void* b);

/* void Frame::ComputeStereoMatches()  include/Frame.h:114, src/Frame.cc:1027-1276.
 * Uses the keypoints, descriptors and pyramids left on the device by the preceding extract calls on
 * eL and eR (frame slot `frame` of each).  uRight/depth: N_left floats (mvuRight, mvDepth).
 * mb = Frame::mb (baseline in metres; minZ, so maxD = mbf / mb, :1081-1084), mb >= 0.  NOTE: the reference's rectified-stereo
 * constructor calls ComputeStereoMatches (:137) BEFORE it assigns mb = mbf / fx (:163), i.e. with an uninitialised member; pass
 * mb = mbf / fx for the intended bound, or mb = 0 (maxD = +inf, no upper disparity bound) for what a zero-filled Frame computes. */
int orbx_stereo_match(orbx_extractor* eL, orbx_extractor* eR, int frame, float mb, float mbf,
                      float* uRight, float* depth, int cap);
/* batch twin: all frames of the last extract; outputs at uRight + f*cap, depth + f*cap.  The frames of a batch stay
 * resident (frame f in device slot f) as long as the batch fits the handle's frame slots (up to 512, bounded by an
 * 8 GB footprint); a larger batch reuses the slots chunk by chunk and leaves NOTHING resident: this call, the
 * per-frame variant, orbx_get_pyramid_level and the debug getters then return ORBX_ERR_STATE. */
int orbx_stereo_match_batch(orbx_extractor* eL, orbx_extractor* eR, float mb, float mbf,
                            float* uRight, float* depth, int cap);
/* resident variant (results stay on device; for kernel timing) */
int orbx_stereo_match_resident(orbx_extractor* eL, orbx_extractor* eR, float mb, float mbf);

/* The stereo Frame constructor's extraction + matching in ONE call: the two ExtractORB threads (src/Frame.cc:132-137,
 * both with vLappingArea {0,0}) and ComputeStereoMatches (:154, src/Frame.cc:1027-1276).  Both extractions run
 * concurrently on the two handles' streams, the match is queued behind them, one host wait.  HOST images in (CV_8UC1,
 * row steps stepL / stepR), HOST results out: cap keypoints / descriptors per eye, uRight / depth for the left
 * keypoints (mvuRight, mvDepth).  nL / nR: keypoints found.  Pyramids and results stay resident as after orbx_extract. */
int orbx_extract_stereo(orbx_extractor* eL, orbx_extractor* eR, const uint8_t* imgL, size_t stepL,
                        const uint8_t* imgR, size_t stepR, int rows, int cols,
                        orbx_keypoint* kpsL, uint8_t* descL, int* nL,
                        orbx_keypoint* kpsR, uint8_t* descR, int* nR, int cap,
                        float mb, float mbf, float* uRight, float* depth);

/* cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2) as used by Frame::ComputeStereoFishEyeMatches
 * src/Frame.cc:1440-1480 (matcher at src/Frame.cc:48).  HOST descriptors in, out_idx/out_dist
 * [nq][2] out (missing neighbour = -1); ties by ascending train index. */
int orbx_knn2_hamming(int device, const uint8_t* q, int nq, const uint8_t* t, int nt,
                      int32_t* out_idx, int32_t* out_dist);

/* orbx_knn2_hamming for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same result as orbx_knn2_hamming, but d_q / d_t / d_out_idx / d_out_dist are the caller's DEVICE buffers
 * (descriptors 16-byte aligned; outputs [nq][2] int32), the kernels are queued on `stream` (a cudaStream_t; NULL =
 * the default stream) and the call returns WITHOUT synchronising.  It calls no cudaMalloc / cudaFree / allocating
 * routine: the per-split partial table lives in a static 4 MiB device array that every GPU loading the library holds
 * until exit.  nq > ORBX_KNN2_NOALLOC_MAX_QUERIES -> ORBX_ERR_CAPACITY (nt is not limited).  The first call on a device
 * resolves the array's device address (cudaGetSymbolAddress): make it outside a capture, nq = 0 does only that.
 * All calls on a device share the array; it keeps the last call's contents, none of which a later call reads before
 * writing.  Calls that OVERLAP on the GPU (different streams or host threads) race on it: serialise them. */
#define ORBX_KNN2_NOALLOC_MAX_QUERIES 16384
int orbx_knn2_hamming_noalloc(int device, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt,
                              int32_t* d_out_idx, int32_t* d_out_dist, void* stream);

/* int ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)
 * include/ORBmatcher.h:46, src/ORBmatcher.cc:45-168 (F.Nleft == -1 branch: mono / rectified stereo).
 * Frame side: n keypoints (undistorted, mvKeysUn), descriptors, optional uRight (NULL for mono),
 * image bounds [minX,maxX]x[minY,maxY] (Frame::ComputeImageBounds src/Frame.cc:978-1009), scale
 * factors.  Map-point side (SoA, filled by Frame::isInFrustum src/Frame.cc:603-711): descriptor,
 * mTrackProjX/Y/XR, mnTrackScaleLevel, mTrackViewCos, mTrackDepth, mbTrackInView.
 * assign[n] in/out: -1 = keypoint free, >= 0 = already holds a map point (F.mvpMapPoints[idx] with
 * Observations() > 0, src/ORBmatcher.cc:102-104); on return newly matched keypoints hold the index
 * of their map point.  Returns nmatches. */
int orbx_search_by_projection(int device,
                              const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                              float minX, float maxX, float minY, float maxY,
                              const float* scaleFactors, int nlevels,
                              int nmp, const uint8_t* mp_desc, const float* mp_projx, const float* mp_projy,
                              const float* mp_projxr, const int32_t* mp_level, const float* mp_viewcos,
                              const float* mp_depth, const uint8_t* mp_inview,
                              float th, float nnratio, int bFarPoints, float thFarPoints,
                              int32_t* assign);

/* std::vector<size_t> Frame::GetFeaturesInArea(x, y, r, minLevel, maxLevel) include/Frame.h:101,
 * src/Frame.cc:784-871 for nq windows at once, with ORBmatcher::DescriptorDistance
 * (src/ORBmatcher.cc:2911-2931) of every hit against the query's descriptor: the candidate core shared
 * by the windowed searches of src/ORBmatcher.cc (:831, :2533-2538).  q_r[i] < 0 skips query i.
 * out_start[nq+1] (CSR offsets), out_idx/out_dist[cap] in the reference's visiting order.
 * Returns the total number of candidates (ORBX_ERR_CAPACITY if it exceeds cap). */
int orbx_window_candidates(int device, const orbx_keypoint* kps, const uint8_t* desc, int n,
                           float minX, float maxX, float minY, float maxY,
                           int nq, const uint8_t* q_desc, const float* q_x, const float* q_y, const float* q_r,
                           const int32_t* q_minLevel, const int32_t* q_maxLevel,
                           int32_t* out_start, int32_t* out_idx, int32_t* out_dist, int cap);

/* int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th,
 * const bool bMono) include/ORBmatcher.h:50, src/ORBmatcher.cc:2469-2711 (CurrentFrame.Nleft == -1).
 * Current frame: kps/desc/uRight (NULL = mono), bounds, scale factors.  Last frame, one row per keypoint
 * i (SoA): descriptor of its map point, the projection uv and invzc of that point in the current camera
 * (:2508-2518, host geometry), mvKeys[i].octave, the angle used at :2584, last_valid = pMP &&
 * !mvbOutlier[i], last_obs = pMP->Observations() > 0.  bForward/bBackward as at :2495-2496.
 * cur_held (may be NULL): current keypoints already holding a map point with observations (:2552-2554).
 * assign[n] out: index of the last-frame row each current keypoint is given, -1 = none.  Returns nmatches. */
int orbx_search_by_projection_frame(int device,
                                    const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
                                    float minX, float maxX, float minY, float maxY,
                                    const float* scaleFactors, int nlevels,
                                    int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                    const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                    const uint8_t* last_valid, const uint8_t* last_obs,
                                    float th, float mbf, int bForward, int bBackward, int checkOrientation,
                                    const uint8_t* cur_held, int32_t* assign);

/* int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound,
 * const float th, const int ORBdist) include/ORBmatcher.h:54, src/ORBmatcher.cc:2723-2862 (relocalisation).
 * One row per keyframe keypoint i: descriptor of its map point, projection uv into the current camera
 * (:2756-2758), predicted level (MapPoint::PredictScale, :2778), pKF->mvKeysUn[i].angle, and mp_valid = pMP &&
 * !isBad() && !sAlreadyFound.count(pMP) && the distance-invariance gate (:2768-2776).  cur_held (may be NULL):
 * CurrentFrame.mvpMapPoints[i2] != NULL.  assign[n] out as above.  Returns nmatches. */
int orbx_search_by_projection_keyframe(int device,
                                       const orbx_keypoint* kps, const uint8_t* desc, int n,
                                       float minX, float maxX, float minY, float maxY,
                                       const float* scaleFactors, int nlevels,
                                       int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                       const int32_t* mp_level, const float* mp_angle, const uint8_t* mp_valid,
                                       float th, int ORBdist, int checkOrientation,
                                       const uint8_t* cur_held, int32_t* assign);

/* int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched,
 * vector<int>& vnMatches12, int windowSize) include/ORBmatcher.h:71, src/ORBmatcher.cc:799-943.
 * prev_matched[2*n1] in/out, matches12[n1] out.  Bounds are frame 2's.  Returns nmatches. */
int orbx_search_for_initialization(int device,
                                   const orbx_keypoint* kps1, const uint8_t* desc1, int n1,
                                   const orbx_keypoint* kps2, const uint8_t* desc2, int n2,
                                   float minX, float maxX, float minY, float maxY,
                                   float* prev_matched, int windowSize, float nnratio, int checkOrientation,
                                   int32_t* matches12);

/* int ORBmatcher::SearchForTriangulation(KeyFrame*, KeyFrame*, cv::Mat F12, vector<pair<size_t,size_t>>&,
 * bOnlyStereo, bCoarse) include/ORBmatcher.h:74, src/ORBmatcher.cc:1107-1386 (variant 0) and
 * SearchForTriangulation_ (include/ORBmatcher.h:77, src/ORBmatcher.cc:1388-1630, variant 1: histogram factor 1/HISTO_LENGTH and no vbMatched2
 * marking), pinhole branch (mpCamera2 == NULL).  DBoW2 FeatureVectors as CSR (fv_node ascending ids,
 * fv_start[nn+1], fv_idx).  has_mp = GetMapPoint(idx) != NULL; uRight NULL = mono.  sf2 / sigma2_2 are
 * pKF2's mvScaleFactors / mvLevelSigma2; ep the epipole (:1120); F12 row-major 3x3 as formed inside
 * Pinhole::epipolarConstrain (src/CameraModels/Pinhole.cpp:243).  matches12[n1] out.  Returns nmatches. */
int orbx_search_for_triangulation(int device,
                                  const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                  const float* uRight1,
                                  const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                  const orbx_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                  const float* uRight2,
                                  const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                  const float* sf2, const float* sigma2_2, int nlevels, const float* ep, const float* F12,
                                  int bOnlyStereo, int bCoarse, int checkOrientation, int variant,
                                  int32_t* matches12);

/* int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches)
 * include/ORBmatcher.h:67, src/ORBmatcher.cc:314-547 (F.Nleft == -1 branch; mode 0) and
 * int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12)
 * include/ORBmatcher.h:68, src/ORBmatcher.cc:955-1105 (mode 1).  Side 1 = the keyframe whose map points are
 * matched: valid1 = map point present and !isBad().  Side 2 = the Frame (mode 0: valid2 may be NULL) or the
 * second keyframe (mode 1: valid2 = map point present and !isBad()).  FeatureVectors as CSR (see
 * orbx_search_for_triangulation).  matches out: mode 0 -> [n2], index of the side-1 keypoint whose map point
 * each Frame keypoint receives; mode 1 -> [n1], index of the matched side-2 keypoint.  Returns nmatches. */
int orbx_search_by_bow(int device, int mode,
                       const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                       const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                       const orbx_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* valid2,
                       const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                       float nnratio, int checkOrientation, int32_t* matches);

/* int ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th, const bool
 * bFarPoints, const float thFarPoints) src/ORBmatcher.cc:45-239 for a two-camera rig (F.Nleft != -1; the TUM-VI
 * stereo-fisheye configuration of BASELINE config C3).  Left keypoints mvKeys[0..nL), right keypoints mvKeysRight
 * [0..nR); desc = F.mDescriptors ((nL + nR) x 32, left rows first); leftToRight / rightToLeft = mvLeftToRightMatch /
 * mvRightToLeftMatch (-1: none); per map point the fields Frame::isInFrustum leaves for each camera
 * (mbTrackInView[R], mTrackProjX/Y[R], mnTrackScaleLevel[R] (right: -1 allowed), mTrackViewCos[R]) and mTrackDepth.
 * assign[nL + nR] in/out = F.mvpMapPoints as map-point rows (right keypoints at nL + i); >= 0 on entry: the keypoint
 * holds a map point with observations.  Returns nmatches (left + right increments, :155-165, :222-231). */
int orbx_search_by_projection_rig(int device,
                                  const orbx_keypoint* kpsL, int nL, const orbx_keypoint* kpsR, int nR, const uint8_t* desc,
                                  const int32_t* leftToRight, const int32_t* rightToLeft,
                                  float minX, float maxX, float minY, float maxY,
                                  const float* scaleFactors, int nlevels,
                                  int nmp, const uint8_t* mp_desc, const float* mp_depth,
                                  const uint8_t* inviewL, const float* projxL, const float* projyL, const int32_t* levelL, const float* viewcosL,
                                  const uint8_t* inviewR, const float* projxR, const float* projyR, const int32_t* levelR, const float* viewcosR,
                                  float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign);

/* void Frame::UndistortKeyPoints() src/Frame.cc:924-960: cv::undistortPoints(mat, mat, K, mDistCoef, cv::Mat(), mK)
 * for the keypoint positions (every other KeyPoint field is copied).  distCoef: k1 k2 p1 p2 [k3] (ncoef = 4 or 5);
 * distCoef[0] == 0 copies the keypoints (:926-931).  HOST arrays in and out; kps_un may alias kps.  The four image
 * corners through the same call give Frame::ComputeImageBounds (:978-1009). */
int orbx_undistort_keypoints(int device, const orbx_keypoint* kps, int n, float fx, float fy, float cx, float cy,
                             const float* distCoef, int ncoef, orbx_keypoint* kps_un);

/* void Tracking::SearchLocalPoints() src/Tracking.cc:3808-3893 in one call: bool
 * Frame::isInFrustum(MapPoint*, float viewingCosLimit) include/Frame.h, src/Frame.cc:603-711 (Nleft == -1) for every
 * local map point ON THE DEVICE - pose Rcw (row-major 3x3), tcw, Ow; pinhole fx, fy, cx, cy (src/CameraModels/
 * Pinhole.cpp:33-37); mbf; logScaleFactor = Frame::mfLogScaleFactor; per point world position, normal, the map point's
 * mfMinDistance / mfMaxDistance AS STORED (the 0.8 / 1.2 of GetMin/MaxDistanceInvariance, src/MapPoint.cc:668-681, are applied here;
 * MapPoint::PredictScale :725-741 divides the stored mfMaxDistance), descriptor, candidate = !isBad() && mnLastFrameSeen != frame id (may be NULL = all) - followed
 * by ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints) src/ORBmatcher.cc:45-168
 * on the mTrack* fields, which never leave the device in between.  One upload, one download.
 * assign[n] in/out as orbx_search_by_projection.  out_* (each may be NULL): mbTrackInView, mTrackProjX/Y/XR,
 * mTrackDepth, mTrackViewCos, mnTrackScaleLevel.  n == 0: isInFrustum only.  Returns nmatches. */
int orbx_search_local_points(int device,
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
                             float* out_depth, float* out_viewcos, int32_t* out_level);

/* ---- DBoW2 vocabulary transform ---------------------------------------- */
typedef struct orbx_vocabulary orbx_vocabulary;

/* ORBVocabulary (= DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>, include/ORBVocabulary.h) built from the rows
 * of its text file exactly as loadFromTextFile does (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1415): header
 * k, L; row r (node id r+1): parent id, leaf flag, 32 descriptor bytes, weight.  Children keep file order, word
 * ids number the leaves in file order.  The tree lives on `device`.  TF_IDF weighting / L1 scoring (the ORBvoc.txt
 * header "10 6 0 0") is what orbx_compute_bow implements. */
int orbx_vocabulary_create(int device, int k, int L, int nrows, const int32_t* parent, const uint8_t* is_leaf,
                           const uint8_t* desc, const double* weight, orbx_vocabulary** out);
void orbx_vocabulary_destroy(orbx_vocabulary* v);

/* TemplatedVocabulary::transform(const TDescriptor& feature, WordId&, WordValue&, NodeId* nid, int levelsup)
 * TemplatedVocabulary.h:1218-1258 for n features (HOST descriptors, n x 32): word id, word weight and the node at
 * level L - levelsup of every feature. */
int orbx_bow_transform(const orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup,
                       int32_t* word, double* weight, int32_t* node);

/* void Frame::ComputeBoW() src/Frame.cc:905-922 (KeyFrame::ComputeBoW alike) = transform(features, mBowVec, mFeatVec,
 * levelsup = 4) TemplatedVocabulary.h:1127-1196.  BowVector as ascending (bow_id, bow_val) pairs, L1-normalised
 * (capacity n each, *nbow entries); FeatureVector as CSR: fv_node[*nfv] ascending, features of node k =
 * fv_idx[fv_start[k] .. fv_start[k+1]) in insertion order (capacities n, n+1, n) - the layout orbx_search_by_bow and
 * orbx_search_for_triangulation take. */
int orbx_compute_bow(const orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup,
                     int32_t* bow_id, double* bow_val, int* nbow,
                     int32_t* fv_node, int32_t* fv_start, int32_t* fv_idx, int* nfv);

/* void MapPoint::ComputeDistinctiveDescriptors() include/MapPoint.h, src/MapPoint.cc:450-531, for a batch of map
 * points.  Group p = descriptor rows [start[p], start[p+1]) of `desc` (the observations'
 * descriptors in the order :463-481 collects them; start[0] == 0).  best_idx[p]: row within the group whose median
 * distance to the others is least (:505-524), -1 for an empty group; best_median (may be NULL): that median.
 * The caller copies desc row start[p] + best_idx[p] into mDescriptor (:526-530). */
int orbx_distinctive_descriptors(int device, int npoints, const int32_t* start, const uint8_t* desc,
                                 int32_t* best_idx, int32_t* best_median);

/* int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints,
 * vector<MapPoint*>& vpMatched, int th, float ratioHamming) include/ORBmatcher.h, src/ORBmatcher.cc:549-678, and the
 * overload with vpPointsKFs / vpMatchedKF, :681-797 (same search; the extra vectors are caller bookkeeping).
 * Keyframe: kps (mvKeysUn) / desc, bounds, scale factors.  One row per candidate map point: descriptor, projection
 * uv under Scw (:584-600, host geometry), predicted level (:628), mp_valid = !isBad() && !spAlreadyFound.count(pMP)
 * && depth > 0 && the distance and viewing-angle gates (:582-625).  KeyFrame::IsInImage (:603) is applied here.
 * matched (may be NULL): vpMatched[idx] != NULL on entry.  assign[n] out: row of the map point each keypoint newly
 * receives, -1 = none.  Returns nmatches. */
int orbx_search_by_projection_sim3(int device,
                                   const orbx_keypoint* kps, const uint8_t* desc, int n,
                                   float minX, float maxX, float minY, float maxY,
                                   const float* scaleFactors, int nlevels,
                                   int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                   const int32_t* mp_level, const uint8_t* mp_valid,
                                   float th, float ratioHamming, const uint8_t* matched, int32_t* assign);

/* The search of int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th, const bool bRight)
 * src/ORBmatcher.cc:1823-2035 (mp_ur != NULL: reprojection gates :1958-1985 against uRight (NULL: mono keyframe) and
 * invLevelSigma2) and of Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th,
 * vector<MapPoint*>& vpReplacePoint) :2051-2185 (mp_ur == NULL).  Rows as above; mp_valid folds every gate before
 * IsInImage (:1861-1903, 1916-1932 / :2078-2117).  best_idx[i]: keypoint row i fuses with (-1: none or distance >
 * TH_LOW), best_dist[i]: its distance (256: no candidate).  Replace / AddObservation / AddMapPoint (:2001-2026,
 * 2165-2180) stay with the caller, in row order.  Returns the number of rows with a fuse target. */
int orbx_fuse(int device,
              const orbx_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
              float minX, float maxX, float minY, float maxY,
              const float* scaleFactors, const float* invLevelSigma2, int nlevels,
              int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v, const float* mp_ur,
              const int32_t* mp_level, const uint8_t* mp_valid, float th,
              int32_t* best_idx, int32_t* best_dist);

/* int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12,
 * const cv::Mat& R12, const cv::Mat& t12, const float th) src/ORBmatcher.cc:2201-2434.  Rows 1: per keypoint of KF1
 * its map point's descriptor, projection into KF2 (:2259-2277), predicted level in KF2 (:2293), valid1 = pMP &&
 * !vbAlreadyMatched1[i1] && !isBad() && depth > 0 && distance gate; rows 2 the reverse direction.  match12[n1] out:
 * idx2 of every cross-checked new match (:2415-2431), -1 otherwise.  Returns nFound. */
int orbx_search_by_sim3(int device,
                        const orbx_keypoint* kps1, const uint8_t* desc1, int n1,
                        const orbx_keypoint* kps2, const uint8_t* desc2, int n2,
                        float minX, float maxX, float minY, float maxY,
                        const float* scaleFactors, int nlevels,
                        const uint8_t* mp1_desc, const float* u1in2, const float* v1in2, const int32_t* level1, const uint8_t* valid1,
                        const uint8_t* mp2_desc, const float* u2in1, const float* v2in1, const int32_t* level2, const uint8_t* valid2,
                        float th, int32_t* match12);

/* void ORBmatcher::ComputeThreeMaxima(vector<int>* histo, const int L, int&, int&, int&)
 * include/ORBmatcher.h:107, src/ORBmatcher.cc:2863-2905 on the bin sizes (host, no device work). */
void orbx_three_maxima(const int32_t* sizes, int L, int32_t* ind1, int32_t* ind2, int32_t* ind3);

/* ---- two-camera rig (Frame::Nleft != -1: KannalaBrandt stereo, e.g. TUM-VI = config C3) ----------------------- */

/* int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono)
 * include/ORBmatcher.h:50, src/ORBmatcher.cc:2469-2711, branch CurrentFrame.Nleft != -1 (left half :2540-2601, right
 * half :2610-2684).  kpsL = mvKeys (nL = Nleft), kpsR = mvKeysRight (nR); desc / cur_held / assign are indexed like
 * mvpMapPoints and mDescriptors: left keypoint i at i, right keypoint j at nL + j.  Per last-frame point the caller
 * supplies the projection into the left camera (last_u, last_v; :2516) and that of Trl * x3Dc (last_ur, last_vr;
 * :2613-2616), invzc, the octave the reference picks at :2530, the last keypoint's angle (:2583-2585), valid = pMP &&
 * !mvbOutlier, obs = Observations() > 0.  Returns nmatches. */
int orbx_search_by_projection_frame_rig(int device,
                                        const orbx_keypoint* kpsL, int nL, const orbx_keypoint* kpsR, int nR, const uint8_t* desc,
                                        float minX, float maxX, float minY, float maxY,
                                        const float* scaleFactors, int nlevels,
                                        int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                        const float* last_ur, const float* last_vr,
                                        const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                        const uint8_t* last_valid, const uint8_t* last_obs,
                                        float th, int bForward, int bBackward, int checkOrientation,
                                        const uint8_t* cur_held, int32_t* assign);

/* int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>&), src/ORBmatcher.cc:314-547, branch
 * F.Nleft != -1 (:386-430, :455-486): Frame rows < nleft2 are left-camera keypoints, the rest right-camera ones
 * (kps2 carries mvKeys / mvKeysRight angles accordingly, kps1 the keyframe's, :437-448).  Otherwise as
 * orbx_search_by_bow(mode 0).  matches[n2] out. */
int orbx_search_by_bow_rig(int device,
                           const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                           const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                           const orbx_keypoint* kps2, const uint8_t* desc2, int n2, int nleft2,
                           const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                           float nnratio, int checkOrientation, int32_t* matches);

/* The triangulation searches whose geometric test is the camera model's own code (KannalaBrandt8: an SVD
 * triangulation, src/CameraModels/KannalaBrandt8.cpp:350-457), which stays the caller's:
 *   overload 0  SearchForTriangulation   include/ORBmatcher.h:74, src/ORBmatcher.cc:1107-1386, branch pKF1->mpCamera2 &&
 *               pKF2->mpCamera2 (:1252-1279; verdict = pCamera1->epipolarConstrain(pCamera2, kp1, kp2, R12, t12, ...))
 *   overload 1  SearchForTriangulation_  include/ORBmatcher.h:77, :1388-1630, same branch (epipolarConstrain_)
 *   overload 2  SearchForTriangulation(..., vector<cv::Mat>& vMatchedPoints)  include/ORBmatcher.h:80, :1632-1821
 *               (verdict = pCamera1->matchAndtriangulate(...), :1729; caller src/Tracking.cc:4739)
 * `verdict(user, idx1, idx2)` (non-zero = accepted) is called from the calling thread in exactly the reference's call
 * order: per KF1 keypoint in FeatureVector order, for the KF2 keypoints under the same node with no map point, not
 * matched before (overload 0 only), dist <= TH_LOW and dist <= the best accepted so far.  A callback that stores its
 * 3-D point per idx1 on acceptance reproduces vMatchedPoints.  bOnlyStereo (overloads 0, 1) yields no match with a rig
 * (:1194); with bCoarse the callback is still invoked (`A || bCoarse` evaluates A first) and its answer ignored - pass
 * verdict = NULL to skip it.  has_mp1 / has_mp2: the keypoint already holds a map point.  matches12[n1]. */
typedef int (*orbx_pair_verdict_fn)(void* user, int idx1, int idx2);
int orbx_search_for_triangulation_cb(int device, int overload,
                                     const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                     const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                     const orbx_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                     const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                     int bOnlyStereo, int bCoarse, int checkOrientation,
                                     orbx_pair_verdict_fn verdict, void* user, int32_t* matches12);

/* void Frame::ComputeStereoFishEyeMatches() include/Frame.h:304, src/Frame.cc:1440-1480, up to the triangulation gate:
 * knnMatch(k = 2) between rows [monoLeft, nL) of eL's and rows [monoRight, nR) of eR's descriptors of frame slot
 * `frame` - the matrices the two preceding orbx_extract calls left ON THE DEVICE (nL / nR = their keypoint counts,
 * monoLeft / monoRight = their return values, :1445-1446) - and Lowe's ratio test d0 < d1 * 0.7 (:1465).
 * l2r[nL] / r2l[nR] out: candidate mvLeftToRightMatch / mvRightToLeftMatch in full-frame indices (-1 = none; the
 * caller applies TriangulateMatches > 0.0001, :1471); best_dist[nL] (may be NULL) the winning distance.
 * Returns the number of pairs that pass the ratio test (descMatches). */
int orbx_stereo_fisheye_match(orbx_extractor* eL, orbx_extractor* eR, int frame, int nL, int monoLeft, int nR, int monoRight,
                              int32_t* l2r, int32_t* r2l, int32_t* best_dist);


#ifdef __cplusplus
}
#endif
#endif /* ORBX_H */
