/*
 * oracle/ref_wrap.cc — TEST INFRASTRUCTURE ONLY.
 *
 * C entry points around the UNMODIFIED reference extractor: /root/reference/src/ORBextractor.cc and
 * /root/reference/include/ORBextractor.h are compiled where they lie (oracle/build_ref.sh) against
 * oracle/cvshim/ and linked with this file into oracle/_ref/libref*.so.  The reference's own code then
 * decides everything the hot path's parity is about — DistributeOctTree (src/ORBextractor.cc:688-1034),
 * the cell loop with the iniThFAST / minThFAST retry (:1038-1143), IC_Angle (:91-138), computeOrbDescriptor
 * (:150-203), the lapping-area ordering of operator() (:1534-1659) — and only the OpenCV pixel
 * primitives underneath come from orc_prims.c (pinned to cv2 4.13).
 *
 * Builds:
 *   libref.so         system allocator: `sort(vector<pair<int,ExtractorNode*>>)` (:927) breaks size ties by
 *                     whatever heap addresses malloc hands out — the reference's literal behaviour.
 *   libref_seq.so     -DREF_SEQ_ALLOC: operator new is a per-call bump arena, so addresses grow in creation
 *                     order and the pointer tie-break becomes the canonical (size, creation sequence) rule
 *                     that oracle/orc_octree.c and the CUDA quadtree implement.  Same reference code, a
 *                     deterministic allocator underneath.
 *   libref_native.so  libref.so with -march=native, as the reference's CMakeLists.txt builds it.
 */
#include <mutex>
#include <new>
#include <vector>

#include "ORBextractor.h"

#include "orc.h"

#ifdef REF_SEQ_ALLOC
#include <sys/mman.h>
namespace {
const size_t ARENA_BYTES = (size_t)8 << 30; /* virtual; pages are touched on demand */
unsigned char* g_arena = nullptr;
size_t g_top = 0;
std::mutex g_call_mutex;
void* arena_alloc(size_t n)
{
    if (!g_arena) {
        g_arena = (unsigned char*)mmap(nullptr, ARENA_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (g_arena == (unsigned char*)MAP_FAILED) std::abort();
    }
    n = (n + 15) & ~(size_t)15;
    if (g_top + n > ARENA_BYTES) std::abort();
    void* p = g_arena + g_top;
    g_top += n;
    return p;
}
} // namespace
void* operator new(size_t n) { return arena_alloc(n ? n : 1); }
void* operator new[](size_t n) { return arena_alloc(n ? n : 1); }
void operator delete(void*) noexcept {}
void operator delete[](void*) noexcept {}
void operator delete(void*, size_t) noexcept {}
void operator delete[](void*, size_t) noexcept {}
#endif

using ORB_SLAM3::ORBextractor;

extern "C" {

void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
{
    return new ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}

void ref_destroy(void* h) { delete (ORBextractor*)h; }

int ref_is_seq_alloc(void)
{
#ifdef REF_SEQ_ALLOC
    return 1;
#else
    return 0;
#endif
}

/* ORBextractor::operator() through the reference's own signature.  Same convention as orc_extract:
 * returns monoIndex (-1 for an empty image), *nkp = keypoints found, -2 if cap is too small. */
int ref_extract(void* h, const uint8_t* img, int rows, int cols, size_t step, int lap0, int lap1,
                orc_keypoint* kps, uint8_t* desc, int cap, int* nkp)
{
    ORBextractor* e = (ORBextractor*)h;
    int mono;
#ifdef REF_SEQ_ALLOC
    std::lock_guard<std::mutex> lock(g_call_mutex);
    const size_t mark = g_top;
#endif
    {
        cv::Mat image = (rows > 0 && cols > 0) ? cv::Mat(rows, cols, CV_8UC1, (void*)img, step) : cv::Mat();
        std::vector<cv::KeyPoint> keypoints;
        cv::Mat descriptors;
        std::vector<int> lapping = {lap0, lap1};
        mono = (*e)(image, cv::Mat(), keypoints, descriptors, lapping);
        const int n = (int)keypoints.size();
        if (nkp) *nkp = mono < 0 ? 0 : n;
        if (mono >= 0) {
            if (n > cap) {
                mono = -2;
            } else {
                static_assert(sizeof(cv::KeyPoint) == sizeof(orc_keypoint), "KeyPoint layout");
                if (n) std::memcpy(kps, keypoints.data(), sizeof(orc_keypoint) * (size_t)n);
                for (int i = 0; i < n; i++) std::memcpy(desc + (size_t)32 * i, descriptors.ptr(i), 32);
            }
        }
    }
#ifdef REF_SEQ_ALLOC
    g_top = mark; /* every operator-new allocation of the call has been destroyed by now */
#endif
    return mono;
}

int ref_levels(void* h) { return ((ORBextractor*)h)->GetLevels(); }
float ref_scale_factor(void* h) { return ((ORBextractor*)h)->GetScaleFactor(); }
/* which: 0 scale factors, 1 inverse scale factors, 2 sigma^2, 3 inverse sigma^2 */
void ref_tables(void* h, int which, float* out)
{
    ORBextractor* e = (ORBextractor*)h;
    std::vector<float> v = which == 0 ? e->GetScaleFactors() : which == 1 ? e->GetInverseScaleFactors()
                         : which == 2 ? e->GetScaleSigmaSquares() : e->GetInverseScaleSigmaSquares();
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
}

/* mvImagePyramid[level] of the last operator() call (include/ORBextractor.h:83).  padded != 0 copies the
 * whole (rows+38) x (cols+38) allocation the level is a view of. */
void ref_level_dims(void* h, int level, int* rows, int* cols)
{
    const cv::Mat& m = ((ORBextractor*)h)->mvImagePyramid[level];
    *rows = m.rows;
    *cols = m.cols;
}
void ref_level_copy(void* h, int level, int padded, uint8_t* dst, size_t dstep)
{
    const cv::Mat& m = ((ORBextractor*)h)->mvImagePyramid[level];
    const int b = padded ? 19 : 0;
    for (int y = -b; y < m.rows + b; y++) std::memcpy(dst + (size_t)(y + b) * dstep, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step.v - b, (size_t)(m.cols + 2 * b));
}

} // extern "C"
