// describe.cu — output assembly, 7x7 Gaussian blur, IC_Angle orientation and rBRIEF-256 descriptors.
// Reference: src/ORBextractor.cc  operator() :1534-1659, computeOrientation/IC_Angle :91-138,580-591,
// GaussianBlur call :1606-1614, computeOrbDescriptor :150-203.
#include "orbx_internal.h"

__constant__ signed char c_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};
// umax of the 31x31 circular patch (src/ORBextractor.cc:542-570); depends only on HALF_PATCH_SIZE=15
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

// ------------------------------------------------------------------------------------------------
// assemble: final row of every selected keypoint (mono ascending / lapping-area descending,
// src/ORBextractor.cc:1589,1633-1655) + all KeyPoint fields except the angle.
// ------------------------------------------------------------------------------------------------
#define ASM_THREADS 256

__global__ void __launch_bounds__(ASM_THREADS)
assemble_kernel(FrameGeom G, DevBuffers D, float lap0, float lap1)
{
    __shared__ int s_warp[ASM_THREADS / 32];
    const int f = blockIdx.x, tid = threadIdx.x;
    const int* selcnt = D.selcnt + f * G.nlevels;
    int total = 0;
    for (int l = 0; l < G.nlevels; l++) total += selcnt[l];
    int nLap = 0, nMono = 0;       // running counts in reference order
    orbx_keypoint* kps = D.kps + (long long)f * G.out_cap;
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        const uint32_t* sel = D.sel + (long long)f * G.sel_per_frame + L.sel_base;
        int* outidx = D.outidx + (long long)f * G.sel_per_frame + L.sel_base;
        const int n = selcnt[l];
        for (int base = 0; base < n; base += ASM_THREADS) {
            const int i = base + tid;
            float x = 0.f, y = 0.f;
            uint32_t key = 0;
            int flag = 0, valid = i < n;
            if (valid) {
                key = sel[i];
                x = (float)(KEY_X(key) + ORBX_MINB);                 // :1167
                y = (float)(KEY_Y(key) + ORBX_MINB);
                if (l != 0) { x = __fmul_rn(x, L.scale); y = __fmul_rn(y, L.scale); }   // :1639-1642
                flag = (x >= lap0 && x <= lap1);                      // :1644
            }
            // block exclusive scan of flag
            const unsigned bal = __ballot_sync(0xffffffffu, flag);
            const int lane = tid & 31, warp = tid >> 5;
            if (lane == 0) s_warp[warp] = __popc(bal);
            __syncthreads();
            int before = __popc(bal & ((1u << lane) - 1)), tot = 0;
            for (int w = 0; w < ASM_THREADS / 32; w++) { if (w < warp) before += s_warp[w]; tot += s_warp[w]; }
            const int nvalid = min(ASM_THREADS, n - base);
            if (valid) {
                const int lapBefore = nLap + before;
                const int monoBefore = nMono + (i - base) - before;
                const int dst = flag ? (total - 1 - lapBefore) : monoBefore;
                outidx[i] = dst;
                orbx_keypoint kp;
                kp.x = x; kp.y = y; kp.size = L.patch_size; kp.angle = -1.f;
                kp.response = (float)KEY_S(key); kp.octave = l; kp.class_id = -1;
                kps[dst] = kp;
            }
            nLap += tot; nMono += nvalid - tot;
            __syncthreads();
        }
    }
    if (tid == 0) { D.counts[2 * f] = total; D.counts[2 * f + 1] = nMono; }
}

int launch_assemble(const FrameGeom& G, const DevBuffers& D, int nframes, int lap0, int lap1, cudaStream_t s)
{
    assemble_kernel<<<nframes, ASM_THREADS, 0, s>>>(G, D, (float)lap0, (float)lap1);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// GaussianBlur(7x7, sigma 2) in OpenCV 4.x fixed point: kernel [18,34,48,56,48,34,18]/256, horizontal
// pass in 8.8 (u16), vertical in 16.16, +0.5 rounding (oracle/orc_prims.c orc_gaussian7x7_s2).
// Reads the padded pyramid plane (its reflect-101 border IS the blur's BORDER_REFLECT_101), writes the
// interior of the blur plane.  HBM-bound: algorithmic bytes per frame = read + write sum(L0..L7).
// ------------------------------------------------------------------------------------------------
#define BL_TW 128
#define BL_TH 16
#define BL_THREADS 256

__global__ void __launch_bounds__(BL_THREADS)
blur_kernel(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur, int pitch, long long plane_bytes, int w, int h)
{
    __shared__ uint32_t s_in[(BL_TH + 6)][(BL_TW + 8) / 4];      // bytes x0-4 .. x0+131
    __shared__ uint16_t s_h[(BL_TH + 6)][BL_TW];
    const int f = blockIdx.z;
    const int x0 = blockIdx.x * BL_TW, y0 = blockIdx.y * BL_TH;
    const uint8_t* src = pyr + (long long)f * plane_bytes;
    const int tid = threadIdx.x;
    // stage rows y0-3 .. y0+TH+2, byte columns XOFF+x0-4 .. (word aligned)
    for (int i = tid; i < (BL_TH + 6) * ((BL_TW + 8) / 4); i += BL_THREADS) {
        const int r = i / ((BL_TW + 8) / 4), c = i - r * ((BL_TW + 8) / 4);
        const int yy = ORBX_YOFF + y0 - 3 + r;                    // padded row
        const int bx = ORBX_XOFF + x0 - 4 + 4 * c;                // byte in row
        uint32_t v = 0;
        if (yy < h + 2 * ORBX_EDGE && bx + 4 <= pitch)
            v = __ldg(reinterpret_cast<const uint32_t*>(src + (long long)yy * pitch + bx));
        s_in[r][c] = v;
    }
    __syncthreads();
    // horizontal pass: 4 outputs per work item
    for (int i = tid; i < (BL_TH + 6) * (BL_TW / 4); i += BL_THREADS) {
        const int r = i / (BL_TW / 4), c = i - r * (BL_TW / 4);
        const uint32_t w0 = s_in[r][c], w1 = s_in[r][c + 1], w2 = s_in[r][c + 2];
        // bytes b[0..11] = columns x-4 .. x+7 for x = x0+4c ; output j uses b[j+1 .. j+7]
        int b[12];
#pragma unroll
        for (int k = 0; k < 4; k++) { b[k] = (w0 >> (8 * k)) & 0xFF; b[4 + k] = (w1 >> (8 * k)) & 0xFF; b[8 + k] = (w2 >> (8 * k)) & 0xFF; }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int acc = 18 * (b[j + 1] + b[j + 7]) + 34 * (b[j + 2] + b[j + 6]) + 48 * (b[j + 3] + b[j + 5]) + 56 * b[j + 4];
            s_h[r][4 * c + j] = (uint16_t)acc;
        }
    }
    __syncthreads();
    // vertical pass: thread -> 4 columns x 2 rows
    const int tx = tid & 31, ty = tid >> 5;
#pragma unroll
    for (int rr = 0; rr < 2; rr++) {
        const int r = ty + 8 * rr, y = y0 + r, x = x0 + 4 * tx;
        if (y < h && x < w) {
            uint32_t word = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int cx = 4 * tx + j;
                const uint32_t acc = 18u * (s_h[r][cx] + s_h[r + 6][cx]) + 34u * (s_h[r + 1][cx] + s_h[r + 5][cx]) +
                                     48u * (s_h[r + 2][cx] + s_h[r + 4][cx]) + 56u * s_h[r + 3][cx];
                word |= (((acc + 32768u) >> 16) & 0xFFu) << (8 * j);
            }
            // columns beyond w inside the last word land in the (unused) right margin of the blur plane
            *reinterpret_cast<uint32_t*>(blur + (long long)f * plane_bytes + (long long)(ORBX_YOFF + y) * pitch + ORBX_XOFF + x) = word;
        }
    }
}

int launch_blur(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    int launches = 0;
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom& L = G.L[l];
        const dim3 grid((L.w + BL_TW - 1) / BL_TW, (L.h + BL_TH - 1) / BL_TH, nframes);
        blur_kernel<<<grid, BL_THREADS, 0, s>>>(L.pyr, L.blur, L.pitch, L.plane_bytes, L.w, L.h);
        launches++;
    }
    return launches;
}

// ------------------------------------------------------------------------------------------------
// orientation + descriptor: one warp per selected keypoint.
// ------------------------------------------------------------------------------------------------
// cv::fastAtan2 (degrees), float32, no FMA — see oracle/orc_prims.c orc_fast_atan2
__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float scale = 57.29577951308232f;                     // (float)(180/CV_PI)
    const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale);
    const float p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
    const float eps = 2.2204460492503131e-16f;                  // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// glibc sinf/cosf (generic C algorithm, double polynomial, no FMA) — see oracle/orc_prims.c orc_sinf.
__device__ __forceinline__ float sc_poly(double x, double x2, bool neg, int n)
{
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __dadd_rn(S2, __dmul_rn(x2, S3));
        const double x7 = __dmul_rn(x3, x2);
        const double s = __dadd_rn(x, __dmul_rn(x3, S1));
        return __double2float_rn(__dadd_rn(s, __dmul_rn(x7, s1)));
    } else {
        const double sg = neg ? -1.0 : 1.0;
        const double x4 = __dmul_rn(x2, x2);
        const double c2 = __dadd_rn(sg * C3, __dmul_rn(x2, sg * C4));
        const double c1 = __dadd_rn(sg * C0, __dmul_rn(x2, sg * C1));
        const double x6 = __dmul_rn(x4, x2);
        const double c = __dadd_rn(c1, __dmul_rn(x4, sg * C2));
        return __double2float_rn(__dadd_rn(c, __dmul_rn(x6, c2)));
    }
}

__device__ __forceinline__ void sincosf_glibc(float y, float* sn, float* cs)
{
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
    double x = (double)y;
    if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ff)) {       // |y| < pi/4
        const double x2 = __dmul_rn(x, x);
        if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ff)) { *sn = y; *cs = 1.0f; return; }
        *sn = sc_poly(x, x2, false, 0);
        *cs = sc_poly(x, x2, false, 1);
        return;
    }
    // reduce_fast: |y| < 120 always holds here (angle in [0, 2*pi])
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const bool neg = (n & 2) != 0;
    const double xs = x * sgn, x2 = __dmul_rn(x, x);
    *sn = sc_poly(xs, x2, neg, n);
    *cs = sc_poly(xs, x2, neg, n ^ 1);
}

#define DESC_WARPS 8

__global__ void __launch_bounds__(DESC_WARPS * 32)
describe_kernel(FrameGeom G, DevBuffers D)
{
    __shared__ signed char s_pat[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_pat[i] = c_pattern[i];
    __syncthreads();
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5);
    if (slot >= G.sel_per_frame) return;
    int lvl = 0;
#pragma unroll 1
    for (int l = 1; l < G.nlevels; l++) if (slot >= G.L[l].sel_base) lvl = l;
    const LevelGeom& L = G.L[lvl];
    const int i = slot - L.sel_base;
    if (i >= D.selcnt[f * G.nlevels + lvl]) return;
    const uint32_t key = D.sel[(long long)f * G.sel_per_frame + slot];
    const int x = KEY_X(key) + ORBX_MINB, y = KEY_Y(key) + ORBX_MINB;
    const long long off = (long long)f * L.plane_bytes + (long long)(ORBX_YOFF + y) * L.pitch + ORBX_XOFF + x;

    // IC_Angle (:91-138): lane = u + 15
    const uint8_t* center = L.pyr + off;
    int m10 = 0, m01 = 0;
    const int u = lane - ORBX_HALF_PATCH;
    if (lane < 31) {
        m10 = u * (int)center[u];
#pragma unroll
        for (int v = 1; v <= ORBX_HALF_PATCH; v++) {
            if (abs(u) <= c_umax[v]) {
                const int vp = center[u + v * L.pitch], vm = center[u - v * L.pitch];
                m01 += v * (vp - vm);
                m10 += u * (vp + vm);
            }
        }
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    m01 = __reduce_add_sync(0xffffffffu, m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // computeOrbDescriptor (:150-203)
    float a = 0.f, b = 0.f;
    if (lane == 0) sincosf_glibc(__fmul_rn(angle, 0.017453292519943295f), &b, &a);   // factorPI = (float)(CV_PI/180.f)
    a = __shfl_sync(0xffffffffu, a, 0);
    b = __shfl_sync(0xffffffffu, b, 0);
    const uint8_t* bc = L.blur + off;
    const signed char* pat = s_pat + 32 * lane;
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        int t[2];
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const float px = (float)pat[4 * k + 2 * e], py = (float)pat[4 * k + 2 * e + 1];
            const int ry = __float2int_rn(__fadd_rn(__fmul_rn(px, b), __fmul_rn(py, a)));
            const int rx = __float2int_rn(__fsub_rn(__fmul_rn(px, a), __fmul_rn(py, b)));
            t[e] = bc[ry * L.pitch + rx];
        }
        val |= (t[0] < t[1]) << k;
    }
    const int dst = D.outidx[(long long)f * G.sel_per_frame + slot];
    D.desc[((long long)f * G.out_cap + dst) * 32 + lane] = (uint8_t)val;
    if (lane == 0) {
        D.kps[(long long)f * G.out_cap + dst].angle = angle;
        D.angle[(long long)f * G.sel_per_frame + slot] = angle;
    }
}

int launch_describe(const FrameGeom& G, const DevBuffers& D, int nframes, cudaStream_t s)
{
    const dim3 grid((G.sel_per_frame + DESC_WARPS - 1) / DESC_WARPS, nframes);
    describe_kernel<<<grid, DESC_WARPS * 32, 0, s>>>(G, D);
    return 1;
}
