// match_internal.h — pieces shared by match.cu and search.cu: the 64x48 frame grid
// (include/Frame.h:37-38 of the reference) in CSR form and the 256-bit Hamming distance.
#pragma once
#include "orbx_host.h"

#define GRID_COLS 64
#define GRID_ROWS 48
#define GRID_CELLS (GRID_COLS * GRID_ROWS)
#define TH_HIGH 100       // src/ORBmatcher.cc:36
#define TH_LOW 50         // :37
#define HISTO_LENGTH 30   // :38

struct GridArgs {
    const orbx_keypoint* kps; int n;
    float minX, minY, winv, hinv;
    int* cell_start;      // GRID_CELLS + 1
    int* cell_idx;        // n
};

// Frame::AssignFeaturesToGrid (src/Frame.cc:446-480): one block builds the CSR grid on `stream`
void orbx_launch_grid_build(const GridArgs& g, cudaStream_t stream);

__device__ __forceinline__ int ham256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// Same distance with 6 instead of 8 POPC (quarter-rate on this part): two carry-save adders fold words 0-2 and 3-5
// into (sum, carry) pairs first - popc(x0)+popc(x1)+popc(x2) = popc(x0^x1^x2) + 2*popc(maj(x0,x1,x2)).  12 LOP3 + 6 POPC
// against 8 LOP3 + 8 POPC: the POPC pipe, not the ALU, bounds the brute-force matcher.
__device__ __forceinline__ int ham256_csa(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const uint32_t x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const uint32_t x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x0 & x2) | (x1 & x2);
    const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x3 & x5) | (x4 & x5);
    return __popc(s1) + __popc(s2) + __popc(x6) + __popc(x7) + 2 * (__popc(c1) + __popc(c2));
}

// ---- per-thread workspace: grow-only device arena + pinned mirror + stream (no cudaMalloc after warm-up) ----
#define CUDA_TRYW(expr)                                                                        \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            orbx_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return ORBX_ERR_CUDA;                                                              \
        }                                                                                      \
    } while (0)

extern "C" int orbx_device_count(void);

struct OrbxWorkspace {
    int device = -1;
    cudaStream_t s = nullptr;
    char* d = nullptr; size_t dcap = 0;     // device arena
    char* h = nullptr; size_t hcap = 0;     // pinned mirror of the arena's input part / output landing zone
    ~OrbxWorkspace() { release(); }
    void release()
    {
        if (device < 0) return;
        if (cudaSetDevice(device) == cudaSuccess) {
            if (d) cudaFree(d);
            if (h) cudaFreeHost(h);
            if (s) cudaStreamDestroy(s);
        }
        d = nullptr; h = nullptr; s = nullptr; dcap = hcap = 0; device = -1;
    }
    int bind(int dev)
    {
        if (dev < 0 || dev >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", dev); return ORBX_ERR_CUDA; }
        if (device != dev) release();
        CUDA_TRYW(cudaSetDevice(dev));
        if (!s) CUDA_TRYW(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        device = dev;
        return ORBX_OK;
    }
    int reserve(size_t dbytes, size_t hbytes)
    {
        if (dbytes > dcap) {
            if (d) { CUDA_TRYW(cudaStreamSynchronize(s)); CUDA_TRYW(cudaFree(d)); d = nullptr; dcap = 0; }
            const size_t want = dbytes + dbytes / 2 + (1u << 20);
            CUDA_TRYW(cudaMalloc((void**)&d, want)); dcap = want;
        }
        if (hbytes > hcap) {
            if (h) { CUDA_TRYW(cudaStreamSynchronize(s)); CUDA_TRYW(cudaFreeHost(h)); h = nullptr; hcap = 0; }
            const size_t want = hbytes + hbytes / 2 + (1u << 20);
            CUDA_TRYW(cudaMallocHost((void**)&h, want)); hcap = want;
        }
        return ORBX_OK;
    }
};
OrbxWorkspace& orbx_workspace();            // the calling thread's workspace (search.cu)

// lays consecutive 256-byte aligned regions out of one arena
struct OrbxCarver {
    size_t off = 0;
    size_t take(size_t bytes) { const size_t o = off; off = (off + bytes + 255) & ~(size_t)255; return o; }
};
