/*
 * orc.h — CPU ORACLE for the ORB front-end hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's algorithm (Wangkkkkkkkk/ORB-SLAM3, CPU, OpenCV based):
 *   src/ORBextractor.cc, src/ORBmatcher.cc, src/Frame.cc  (file:line cited at every function).
 * The OpenCV primitives the reference calls (resize, copyMakeBorder, FAST, GaussianBlur,
 * fastAtan2, cvRound, norm L1, BFMatcher) are NOT in /root/reference (un-vendored, unpinned
 * dependency, CMakeLists.txt:37-43); they are restated here from their published algorithms and
 * PINNED against cv2 4.13.0 (tests/test_oracle_vs_cv2.py, tests/golden/ built by
 * oracle/make_golden.py which runs the real cv2 primitives).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * link or call this.  The product (orb-slam3_b200/) never does.
 *
 * Canonicalisations (documented in DESIGN.md):
 *   - quadtree tie-break: (size, creation sequence) instead of (size, heap pointer)
 *     (src/ORBextractor.cc:927 sorts pair<int,ExtractorNode*>);
 *   - cosf/sinf: glibc's generic C algorithm evaluated without FMA contraction (orc_sinf/orc_cosf);
 *   - tap rotation x*b+y*a: no FMA contraction (build uses -ffp-contract=off).
 */
#ifndef ORC_H
#define ORC_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* cv::KeyPoint memory layout (28 bytes). */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orc_keypoint;

/* candidate / selected keypoint inside one level (coordinates relative to minBorder=16). */
typedef struct {
    int32_t x, y, score;
} orc_cand;

typedef struct orc_extractor orc_extractor;

enum { ORC_TIE_CANONICAL = 0, ORC_TIE_REVERSED = 1 };
enum { ORC_TRIG_CANONICAL = 0, ORC_TRIG_LIBM = 1 };

/* ---- OpenCV primitive restatements (orc_prims.c) ---- */
int   orc_cvround_f(float v);                       /* cvRound(float): round-half-even      */
int   orc_cvround_d(double v);                      /* cvRound(double)                      */
float orc_fast_atan2(float y, float x);             /* cv::fastAtan2, degrees [0,360)       */
float orc_sinf(float x);                            /* glibc generic sinf, no FMA           */
float orc_cosf(float x);
void  orc_resize_linear_u8(const uint8_t* src, int srows, int scols, size_t sstep,
                           uint8_t* dst, int drows, int dcols, size_t dstep);
void  orc_border_reflect101(uint8_t* plane, int rows, int cols, size_t step, int border);
/* cv::FAST(img, kps, threshold, nms=true) TYPE_9_16; returns count, writes raster-ordered (x,y,score) */
int   orc_fast9_16(const uint8_t* img, int rows, int cols, size_t step, int threshold,
                   int nonmax, orc_cand* out, int cap);
void  orc_gaussian7x7_s2(const uint8_t* src, int rows, int cols, size_t sstep,
                         uint8_t* dst, size_t dstep);    /* BORDER_REFLECT_101, fixed point */
int   orc_hamming256(const void* a, const void* b);      /* ORBmatcher::DescriptorDistance  */
/* BFMatcher(NORM_HAMMING).knnMatch(k=2): out_idx/out_dist are [nq][2]; missing = -1 */
void  orc_bf_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* out_idx, int32_t* out_dist);

/* ---- frame preparation in front of operator() (orc_prepare.c) ---- */
/* cv::remap(src, dst, mapx, mapy, INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT 0; 1, 3 or 4 channels */
void  orc_remap_linear_u8(const uint8_t* src, int srows, int scols, size_t sstep, int channels,
                          const float* mapx, const float* mapy, size_t mstep_floats,
                          uint8_t* dst, int drows, int dcols, size_t dstep);
/* cvtColor(COLOR_{RGB,BGR,RGBA,BGRA}2GRAY): channels 3 or 4, rgb != 0 when the first channel is R */
void  orc_cvt_gray_u8(const uint8_t* src, int rows, int cols, size_t sstep, int channels, int rgb,
                      uint8_t* dst, size_t dstep);
/* cv::undistortPoints(src, dst, K, distCoef, Mat(), K) (Frame::UndistortKeyPoints, src/Frame.cc:924-960); dist = k1 k2 p1 p2 k3 */
void  orc_undistort_points(const float* xy_in, int n, float fx, float fy, float cx, float cy, const float* dist, float* xy_out);


/* ---- ORBextractor (orc_extractor.c) ---- */
orc_extractor* orc_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
void  orc_destroy(orc_extractor* e);
void  orc_set_modes(orc_extractor* e, int tie_mode, int trig_mode);
/* ORBextractor::operator(): returns monoIndex, -1 on empty image, -2 if cap too small. */
int   orc_extract(orc_extractor* e, const uint8_t* img, int rows, int cols, size_t step,
                  int lap0, int lap1, orc_keypoint* kps, uint8_t* desc, int cap, int* nkp);
/* getters */
int   orc_levels(const orc_extractor* e);
const float* orc_scale_factors(const orc_extractor* e);
const float* orc_inv_scale_factors(const orc_extractor* e);
const float* orc_level_sigma2(const orc_extractor* e);
const float* orc_inv_level_sigma2(const orc_extractor* e);
const int*   orc_features_per_level(const orc_extractor* e);
const int*   orc_umax(const orc_extractor* e);
/* stage accessors (valid after orc_extract) */
void  orc_level_dims(const orc_extractor* e, int level, int* rows, int* cols);
const uint8_t* orc_level_plane(const orc_extractor* e, int level, size_t* step);   /* padded origin */
const uint8_t* orc_level_blurred(const orc_extractor* e, int level, size_t* step); /* unpadded      */
int   orc_level_candidates(const orc_extractor* e, int level, const orc_cand** out);
int   orc_level_selected(const orc_extractor* e, int level, const orc_cand** out, const float** angles);
/* stand-alone pieces for unit parity */
int   orc_distribute_octree(const orc_cand* keys, int nkeys, int minX, int maxX, int minY, int maxY,
                            int N, int tie_mode, orc_cand* out, int cap);
float orc_ic_angle(const uint8_t* center, size_t step, const int* umax);
void  orc_orb_descriptor(const uint8_t* center, size_t step, float angle_deg, int trig_mode, uint8_t* desc);

/* ---- matching (orc_match.c) ---- */
/* Frame::ComputeStereoMatches (src/Frame.cc:1027-1276). best_dist/best_sad optional (may be NULL). */
void  orc_stereo_match(const orc_extractor* eL, const orc_extractor* eR,
                       const orc_keypoint* kpL, int nL, const uint8_t* descL,
                       const orc_keypoint* kpR, int nR, const uint8_t* descR,
                       float mb, float mbf, float* uRight, float* depth,
                       int32_t* best_dist, int32_t* best_sad);

/* Frame grid (include/Frame.h:37-38, src/Frame.cc:446-480, 784-899) */
typedef struct orc_grid orc_grid;
orc_grid* orc_grid_build(const orc_keypoint* kps, int n, float minX, float maxX, float minY, float maxY);
void  orc_grid_destroy(orc_grid* g);
void  orc_grid_bounds(const orc_grid* g, float* out4);
int   orc_grid_query(const orc_grid* g, const orc_keypoint* kps, float x, float y, float r,
                     int minLevel, int maxLevel, int32_t* out, int cap);

/* ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFar, thFar) mono/rectified-stereo
 * branch (F.Nleft == -1), src/ORBmatcher.cc:45-168.  SoA map-point inputs; assign[n] in/out:
 * -1 = free, >=0 = index of the map point already held (treated as "Observations()>0").
 * uRight may be NULL (mono).  Returns nmatches. */
int   orc_search_by_projection(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n,
                               const float* uRight, const float* scaleFactors,
                               int nmp, const uint8_t* mp_desc, const float* mp_projx, const float* mp_projy,
                               const float* mp_projxr, const int32_t* mp_level, const float* mp_viewcos,
                               const float* mp_depth, const uint8_t* mp_inview,
                               float th, float nnratio, int bFarPoints, float thFarPoints,
                               int32_t* assign);

/* Frame::isInFrustum (src/Frame.cc:603-711, Nleft == -1) over a list of map points; see orc_match.c */
int   orc_is_in_frustum(const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx, float cy, float mbf,
                        float minX, float maxX, float minY, float maxY, float logScaleFactor, int nlevels,
                        int nmp, const float* pos, const float* normal, const float* min_dist, const float* max_dist,
                        const uint8_t* candidate, float viewingCosLimit,
                        uint8_t* inview, float* projx, float* projy, float* projxr, float* depth, float* viewcos, int32_t* level);

/* MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:450-531, batched over CSR groups of descriptors */
void  orc_distinctive_descriptors(int npoints, const int32_t* start, const uint8_t* desc, int32_t* best_idx, int32_t* best_median);

/* DBoW2 TemplatedVocabulary::transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1258 and :1127-1196) behind
 * Frame::ComputeBoW (src/Frame.cc:905-922); see orc_match.c */
void  orc_bow_transform(const int32_t* child_start, const int32_t* child_idx, const uint8_t* node_desc,
                        const double* node_weight, const int32_t* node_word, int L, int levelsup,
                        int n, const uint8_t* desc, int32_t* word, double* weight, int32_t* node);
int   orc_bow_vectors(int n, const int32_t* word, const double* weight, const int32_t* node,
                      int32_t* bow_id, double* bow_val, int32_t* fv_node, int32_t* fv_start, int32_t* fv_idx, int* nfv);

/* ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) src/ORBmatcher.cc:45-239, branch Nleft != -1 (two-camera rig) */
int   orc_search_by_projection_rig(const orc_grid* gL, const orc_keypoint* kpsL, int nL,
                                   const orc_grid* gR, const orc_keypoint* kpsR, int nR,
                                   const uint8_t* desc, const int32_t* l2r, const int32_t* r2l, const float* sf,
                                   int nmp, const uint8_t* mp_desc, const float* mp_depth,
                                   const uint8_t* inviewL, const float* projxL, const float* projyL, const int32_t* levelL, const float* viewcosL,
                                   const uint8_t* inviewR, const float* projxR, const float* projyR, const int32_t* levelR, const float* viewcosR,
                                   float th, float nnratio, int bFarPoints, float thFarPoints, int32_t* assign);

/* ORBmatcher::ComputeThreeMaxima, src/ORBmatcher.cc:2863-2905 (sizes[i] = histo[i].size()) */
void  orc_three_maxima(const int* sizes, int L, int* ind1, int* ind2, int* ind3);

/* ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono), src/ORBmatcher.cc:2469-2711,
 * Nleft == -1 branch; see orc_match.c for the argument meaning.  Returns nmatches. */
int   orc_search_by_projection_frame(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n,
                                     const float* uRight, const float* scaleFactors,
                                     int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                     const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                     const uint8_t* last_valid, const uint8_t* last_obs,
                                     float th, float mbf, int bForward, int bBackward, int checkOrientation,
                                     const uint8_t* cur_held, int32_t* assign);

/* ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:799-943 */
int   orc_search_for_initialization(const orc_keypoint* kps1, const uint8_t* desc1, int n1,
                                    const orc_grid* g2, const orc_keypoint* kps2, const uint8_t* desc2, int n2,
                                    float* prev_matched, int windowSize, float nnratio, int checkOrientation,
                                    int32_t* matches12);

/* ORBmatcher::SearchForTriangulation (:1107-1386, variant 0) / SearchForTriangulation_ (:1388-1630, variant 1) */
int   orc_search_for_triangulation(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                   const float* uRight1,
                                   const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                   const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                   const float* uRight2,
                                   const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                   const float* sf2, const float* sigma2_2, const float* ep, const float* F12,
                                   int bOnlyStereo, int bCoarse, int checkOrientation, int variant,
                                   int32_t* matches12);

/* ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:2723-2862 */
int   orc_search_by_projection_keyframe(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                                        int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                        const int32_t* mp_level, const float* mp_angle, const uint8_t* mp_valid,
                                        float th, int ORBdist, int checkOrientation, const uint8_t* cur_held, int32_t* assign);

/* ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:549-678 (and :681-797) */
int   orc_search_by_projection_sim3(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* sf,
                                    int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v,
                                    const int32_t* mp_level, const uint8_t* mp_valid,
                                    float th, float ratioHamming, const uint8_t* matched, int32_t* assign);
/* the search of ORBmatcher::Fuse, src/ORBmatcher.cc:1823-2035 (mp_ur != NULL) and :2051-2185 (mp_ur == NULL) */
int   orc_fuse(const orc_grid* g, const orc_keypoint* kps, const uint8_t* desc, int n, const float* uRight,
               const float* sf, const float* inv_sigma2,
               int nmp, const uint8_t* mp_desc, const float* mp_u, const float* mp_v, const float* mp_ur,
               const int32_t* mp_level, const uint8_t* mp_valid, float th, int32_t* best_idx, int32_t* best_dist);
/* ORBmatcher::SearchBySim3, src/ORBmatcher.cc:2201-2434 */
int   orc_search_by_sim3(const orc_grid* g1, const orc_keypoint* kps1, const uint8_t* desc1, int n1,
                         const orc_grid* g2, const orc_keypoint* kps2, const uint8_t* desc2, int n2, const float* sf,
                         const uint8_t* mp1_desc, const float* u1in2, const float* v1in2, const int32_t* level1, const uint8_t* valid1,
                         const uint8_t* mp2_desc, const float* u2in1, const float* v2in1, const int32_t* level2, const uint8_t* valid2,
                         float th, int32_t* match12);

/* ORBmatcher::SearchByBoW (KF,Frame: mode 0, src/ORBmatcher.cc:314-547; KF,KF: mode 1, :955-1105) */
int   orc_search_by_bow(int mode,
                        const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                        const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                        const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* valid2,
                        const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                        float nnratio, int checkOrientation, int32_t* matches);

/* ---- two-camera rig branches, callback-driven triangulation searches, fisheye stereo (orc_match.c, bottom) ---- */
typedef int (*orc_pair_verdict_fn)(void* user, int idx1, int idx2);
int   orc_search_by_projection_frame_rig(const orc_grid* gL, const orc_keypoint* kpsL, int nL,
                                         const orc_grid* gR, const orc_keypoint* kpsR, int nR,
                                         const uint8_t* desc, const float* sf,
                                         int nlast, const uint8_t* last_desc, const float* last_u, const float* last_v,
                                         const float* last_ur, const float* last_vr,
                                         const float* last_invz, const int32_t* last_octave, const float* last_angle,
                                         const uint8_t* last_valid, const uint8_t* last_obs,
                                         float th, int bForward, int bBackward, int checkOrientation,
                                         const uint8_t* cur_held, int32_t* assign);
int   orc_search_by_bow_rig(const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* valid1,
                            const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                            const orc_keypoint* kps2, const uint8_t* desc2, int n2, int nleft2,
                            const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                            float nnratio, int checkOrientation, int32_t* matches);
int   orc_search_for_triangulation_cb(int overload,
                                      const orc_keypoint* kps1, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                      const int32_t* fv1_node, const int32_t* fv1_start, const int32_t* fv1_idx, int nn1,
                                      const orc_keypoint* kps2, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                      const int32_t* fv2_node, const int32_t* fv2_start, const int32_t* fv2_idx, int nn2,
                                      int bOnlyStereo, int bCoarse, int checkOrientation,
                                      orc_pair_verdict_fn verdict, void* user, int32_t* matches12);
int   orc_stereo_fisheye_match(const uint8_t* descL, int nL, int monoLeft, const uint8_t* descR, int nR, int monoRight,
                               int32_t* l2r, int32_t* r2l, int32_t* dist_out);

#ifdef __cplusplus
}
#endif
#endif
