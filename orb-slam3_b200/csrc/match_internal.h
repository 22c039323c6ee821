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
