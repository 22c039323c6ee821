// pyramid.cu — image pyramid (ORBextractor::ComputePyramid, reference src/ORBextractor.cc:1664-1715).
//
// Level 0 = copyMakeBorder(image, 19, BORDER_REFLECT_101); level l>0 = cv::resize(level l-1,
// INTER_LINEAR) in OpenCV's 11-bit fixed point followed by the same reflect-101 border.  The border
// is folded into the tap tables (indexed by PADDED coordinate), so each level is one streaming pass:
// every thread produces 4 consecutive bytes of one padded row and stores them as one 32-bit word
// (rows are 128-byte pitched, interior starts on a 32-byte sector boundary).
// HBM-bound: algorithmic bytes per frame = read sum(L0..L6) + write sum(L1..L7).
#include "orbx_internal.h"

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

// level 0: input frame -> padded plane
__global__ void __launch_bounds__(256)
pyr_level0_kernel(const uint8_t* __restrict__ in, int in_pitch, long long in_frame_bytes,
                  uint8_t* __restrict__ out, int pitch, long long plane_bytes, int w, int h)
{
    const int wi = blockIdx.x * blockDim.x + threadIdx.x;     // 32-bit word of the row
    const int Y = blockIdx.y * blockDim.y + threadIdx.y;      // padded row
    if (Y >= h + 2 * ORBX_EDGE || wi * 4 >= pitch) return;
    const int f = blockIdx.z;
    const uint8_t* src = in + (long long)f * in_frame_bytes + (long long)reflect101(Y - ORBX_EDGE, h) * in_pitch;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int X = wi * 4 + k - (ORBX_XOFF - ORBX_EDGE);   // padded column
        if (X >= 0 && X < w + 2 * ORBX_EDGE)
            word |= (uint32_t)__ldg(src + reflect101(X - ORBX_EDGE, w)) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(out + (long long)f * plane_bytes + (long long)Y * pitch + wi * 4) = word;
}

// level l>0 from level l-1.  OpenCV HResizeLinear/VResizeLinear u8 arithmetic (see oracle/orc_prims.c):
//   H = S[s0]*a0 + S[s1]*a1 ;  dst = (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2
__global__ void __launch_bounds__(256)
pyr_resize_kernel(const uint8_t* __restrict__ prev, int ppitch, long long pplane_bytes,
                  uint8_t* __restrict__ out, int pitch, long long plane_bytes, int w, int h,
                  const ResizeTapX* __restrict__ tapx, const ResizeTapY* __restrict__ tapy)
{
    const int wi = blockIdx.x * blockDim.x + threadIdx.x;
    const int Y = blockIdx.y * blockDim.y + threadIdx.y;
    if (Y >= h + 2 * ORBX_EDGE || wi * 4 >= pitch) return;
    const int f = blockIdx.z;
    const ResizeTapY ty = tapy[Y];
    const uint8_t* base = prev + (long long)f * pplane_bytes + ORBX_XOFF;
    const uint8_t* S0 = base + (long long)(ORBX_YOFF + ty.s0) * ppitch;
    const uint8_t* S1 = base + (long long)(ORBX_YOFF + ty.s1) * ppitch;
    const int b0 = ty.b0, b1 = ty.b1;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int X = wi * 4 + k - (ORBX_XOFF - ORBX_EDGE);
        if (X >= 0 && X < w + 2 * ORBX_EDGE) {
            const ResizeTapX tx = tapx[X];
            const int h0 = (int)__ldg(S0 + tx.s0) * tx.a0 + (int)__ldg(S0 + tx.s1) * tx.a1;
            const int h1 = (int)__ldg(S1 + tx.s0) * tx.a0 + (int)__ldg(S1 + tx.s1) * tx.a1;
            const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
            word |= (uint32_t)(v & 0xFF) << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(out + (long long)f * plane_bytes + (long long)Y * pitch + wi * 4) = word;
}

int launch_pyramid(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    int launches = 0;
    const dim3 block(64, 4);
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        const dim3 grid((L.pitch / 4 + block.x - 1) / block.x, (L.prow + block.y - 1) / block.y, nframes);
        if (l == 0) {
            pyr_level0_kernel<<<grid, block, 0, s>>>(D.input, D.in_pitch, (long long)D.in_pitch * G.rows,
                                                     L.pyr, L.pitch, L.plane_bytes, L.w, L.h);
        } else {
            const LevelGeom& P = G.L[l - 1];
            pyr_resize_kernel<<<grid, block, 0, s>>>(P.pyr, P.pitch, P.plane_bytes, L.pyr, L.pitch, L.plane_bytes,
                                                     L.w, L.h, L.tapx, L.tapy);
        }
        launches++;
    }
    return launches;
}
