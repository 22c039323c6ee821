// prepare.cu — the frame preparation in front of ORBextractor::operator(), on the device:
//   cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR)          Examples/Stereo/stereo_euroc.cc:168-169
//     (maps from cv::initUndistortRectifyMap(..., CV_32F, M1, M2), :117-118, once per camera)
//   cvtColor(mImGray, mImGray, COLOR_{RGB,BGR,RGBA,BGRA}2GRAY) src/Tracking.cc:1264-1290
// fused into ONE pass that writes the extractor's frame slot: per output pixel the four source taps of
// every colour channel are gathered, blended in OpenCV's fixed point (5 fractional bits per axis,
// weights summing to 2^15, round by +2^14) and the per-channel results are folded to gray
// ((R*9798 + G*19235 + B*3735 + 2^14) >> 15) in registers.  Rounding happens per channel BEFORE the gray
// fold, exactly as remap-then-cvtColor does on the CPU.
//
// rectify_pack_kernel (once per camera) turns the float maps into the fixed-point form remap uses
// internally: sx = cvRound(mapx*32), ix = sat_short(sx >> 5), fx = sx & 31 (same for y), stored as
// int2 {ix | iy << 16, fx | fy << 5}: the per-frame kernel does no float work at all.  The maps of one
// camera are shared by every frame of a batch (2.9 MB at 752x480), i.e. L2-resident; compulsory HBM
// traffic per frame is raw frame in + gray frame out.
//
// One thread per output word (4 px): map entries by two 16-byte loads, 16 x CH byte gathers (neighbouring
// output pixels read neighbouring source pixels: the gathers of a warp fall into 1-2 sectors per
// request), one 32-bit store.  Bound: L1 gather wavefronts, then HBM.
#include "orbx_host.h"

struct PrepArgs {
    const uint8_t* raw; int srows, scols, spitch; long long sframe;
    const int2* map;
    uint8_t* dst; int rows, cols, dpitch; long long dframe;
    int rgb;
};

__global__ void rectify_pack_kernel(const float* __restrict__ mx, const float* __restrict__ my, int rows, int cols,
                                    int mstep /* floats */, int2* __restrict__ out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    const int sx = __float2int_rn(__fmul_rn(mx[(size_t)y * mstep + x], 32.0f));
    const int sy = __float2int_rn(__fmul_rn(my[(size_t)y * mstep + x], 32.0f));
    int ix = sx >> 5, iy = sy >> 5;
    ix = max(-32768, min(32767, ix)); iy = max(-32768, min(32767, iy));
    out[(size_t)y * cols + x] = make_int2((ix & 0xFFFF) | (iy << 16), (sx & 31) | ((sy & 31) << 5));
}

template <int CH, bool REMAP>
__device__ __forceinline__ uint32_t prepare_pixel(const PrepArgs& A, const uint8_t* __restrict__ S, int x, int y, int2 m)
{
    constexpr int NC = CH == 1 ? 1 : 3;               // alpha is never read
    int v[NC];
    if (REMAP) {
        const int ix = (int)(short)(m.x & 0xFFFF), iy = m.x >> 16;
        const int fx = m.y & 31, fy = (m.y >> 5) & 31;
        // (sum_k tap_k * w_k * 32 + 2^14) >> 15 == (sum_k tap_k * w_k + 2^9) >> 10 with w = (32-fx | fx) * (32-fy | fy):
        // two horizontal lerps and one vertical, all exact in int32
        const uint8_t* p = S + (long long)iy * A.spitch + (long long)ix * CH;
        if ((unsigned)ix < (unsigned)(A.scols - 1) && (unsigned)iy < (unsigned)(A.srows - 1)) {
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const int top = p[c] * (32 - fx) + p[CH + c] * fx, bot = p[A.spitch + c] * (32 - fx) + p[A.spitch + CH + c] * fx;
                v[c] = (top * (32 - fy) + bot * fy + 512) >> 10;
            }
        } else {                                          // some tap reads the constant border (0)
            const bool x0 = ix >= 0 && ix < A.scols, x1 = ix + 1 >= 0 && ix + 1 < A.scols;
            const bool y0 = iy >= 0 && iy < A.srows, y1 = iy + 1 >= 0 && iy + 1 < A.srows;
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const int t00 = (x0 && y0) ? p[c] : 0, t01 = (x1 && y0) ? p[CH + c] : 0;
                const int t10 = (x0 && y1) ? p[A.spitch + c] : 0, t11 = (x1 && y1) ? p[A.spitch + CH + c] : 0;
                const int top = t00 * (32 - fx) + t01 * fx, bot = t10 * (32 - fx) + t11 * fx;
                v[c] = (top * (32 - fy) + bot * fy + 512) >> 10;
            }
        }
    } else {
        const uint8_t* p = S + (long long)y * A.spitch + (long long)x * CH;
#pragma unroll
        for (int c = 0; c < NC; c++) v[c] = p[c];
    }
    if (CH == 1) return (uint32_t)v[0];
    const int r = A.rgb ? v[0] : v[NC - 1], b = A.rgb ? v[NC - 1] : v[0];
    return (uint32_t)((r * 9798 + v[NC == 3 ? 1 : 0] * 19235 + b * 3735 + (1 << 14)) >> 15);
}

template <int CH, bool REMAP>
__global__ void __launch_bounds__(256)
prepare_kernel(PrepArgs A)
{
    const int wx = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, f = blockIdx.z;
    const int x0 = wx * 4;
    if (x0 >= A.cols || y >= A.rows) return;
    const uint8_t* S = A.raw + (long long)f * A.sframe;
    int2 m[4] = {};
    if (REMAP) {
        const int2* mp = A.map + (size_t)y * A.cols + x0;
        if ((A.cols & 3) == 0) {
            const int4 a = reinterpret_cast<const int4*>(mp)[0], b = reinterpret_cast<const int4*>(mp)[1];
            m[0] = make_int2(a.x, a.y); m[1] = make_int2(a.z, a.w); m[2] = make_int2(b.x, b.y); m[3] = make_int2(b.z, b.w);
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) if (x0 + k < A.cols) m[k] = mp[k];
        }
    }
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (x0 + k < A.cols) out |= prepare_pixel<CH, REMAP>(A, S, x0 + k, y, m[k]) << (8 * k);
    *reinterpret_cast<uint32_t*>(A.dst + (long long)f * A.dframe + (long long)y * A.dpitch + x0) = out;
}

// raw frames of frame slots [slot0, slot0 + nframes) -> gray frame slots.  Returns the number of launches.
int launch_prepare(orbx_extractor* e, int slot0, int nframes, cudaStream_t s)
{
    PrepArgs A;
    A.srows = e->raw_rows; A.scols = e->raw_cols; A.spitch = e->raw_pitch;
    A.sframe = (long long)e->raw_rows * e->raw_pitch;
    A.raw = e->d_raw + (size_t)slot0 * A.sframe;
    A.map = e->d_map;
    A.rows = e->G.rows; A.cols = e->G.cols; A.dpitch = e->D.in_pitch;
    A.dframe = (long long)e->G.rows * e->D.in_pitch;
    A.dst = e->D.input + (size_t)slot0 * A.dframe;
    A.rgb = e->raw_rgb;
    const dim3 block(32, 8), grid(((A.cols + 3) / 4 + 31) / 32, (A.rows + 7) / 8, nframes);
    const bool remap = e->d_map != nullptr;
    switch (e->raw_ch * 2 + (remap ? 1 : 0)) {
    case 3: prepare_kernel<1, true><<<grid, block, 0, s>>>(A); break;
    case 6: prepare_kernel<3, false><<<grid, block, 0, s>>>(A); break;
    case 7: prepare_kernel<3, true><<<grid, block, 0, s>>>(A); break;
    case 8: prepare_kernel<4, false><<<grid, block, 0, s>>>(A); break;
    case 9: prepare_kernel<4, true><<<grid, block, 0, s>>>(A); break;
    default: return 0;            // 1 channel, no maps: the frame IS the gray frame (copied straight into the slot)
    }
    return 1;
}

int launch_rectify_pack(const float* mx, const float* my, int rows, int cols, int mstep_floats, int2* out, cudaStream_t s)
{
    rectify_pack_kernel<<<dim3((cols + 255) / 256, rows), 256, 0, s>>>(mx, my, rows, cols, mstep_floats, out);
    return 1;
}
