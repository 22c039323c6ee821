#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <algorithm>
#include <cuda_runtime.h>
template <int VARIANT>
__host__ __device__ inline int fast_score_v(const uint8_t* p, int pitch, int t)
{
    const int v = p[0];
    int r[16];
    r[0] = p[3 * pitch];       r[1] = p[3 * pitch + 1];   r[2] = p[2 * pitch + 2];   r[3] = p[pitch + 3];
    r[4] = p[3];               r[5] = p[-pitch + 3];      r[6] = p[-2 * pitch + 2];  r[7] = p[-3 * pitch + 1];
    r[8] = p[-3 * pitch];      r[9] = p[-3 * pitch - 1];  r[10] = p[-2 * pitch - 2]; r[11] = p[-pitch - 3];
    r[12] = p[-3];             r[13] = p[pitch - 3];      r[14] = p[2 * pitch - 2];  r[15] = p[3 * pitch - 1];
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = v - r[k];
    int A = -256, B = 256;
    if (VARIANT == 0) {
        int mn2[16], mx2[16], mn4[16], mx4[16];
#pragma unroll
        for (int k = 0; k < 16; k++) { mn2[k] = min(d[k], d[(k + 1) & 15]); mx2[k] = max(d[k], d[(k + 1) & 15]); }
#pragma unroll
        for (int k = 0; k < 16; k++) { mn4[k] = min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = max(mx2[k], mx2[(k + 2) & 15]); }
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int mn9 = min(min(mn4[k], mn4[(k + 4) & 15]), d[(k + 8) & 15]);
            const int mx9 = max(max(mx4[k], mx4[(k + 4) & 15]), d[(k + 8) & 15]);
            A = max(A, mn9); B = min(B, mx9);
        }
    } else if (VARIANT == 2) {
        // negation-free: B' = max over arcs of min(-d)
        int e[16], mn2[16], nn2[16], mn4[16], nn4[16];
#pragma unroll
        for (int k = 0; k < 16; k++) e[k] = r[k] - v;
#pragma unroll
        for (int k = 0; k < 16; k++) { mn2[k] = min(d[k], d[(k + 1) & 15]); nn2[k] = min(e[k], e[(k + 1) & 15]); }
#pragma unroll
        for (int k = 0; k < 16; k++) { mn4[k] = min(mn2[k], mn2[(k + 2) & 15]); nn4[k] = min(nn2[k], nn2[(k + 2) & 15]); }
        int Bn = -256;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            A = max(A, min(min(mn4[k], mn4[(k + 4) & 15]), d[(k + 8) & 15]));
            Bn = max(Bn, min(min(nn4[k], nn4[(k + 4) & 15]), e[(k + 8) & 15]));
        }
        if (!(A > t || Bn > t)) return 0;
        return max(max(A, Bn), t) - 1;
    } else {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            int mn = d[k], mx = d[k];
#pragma unroll
            for (int j = 1; j < 9; j++) { mn = min(mn, d[(k + j) & 15]); mx = max(mx, d[(k + j) & 15]); }
            A = max(A, mn); B = min(B, mx);
        }
    }
    if (!(A > t || B < -t)) return 0;
    return max(max(A, -B), t) - 1;
}
template <int V> __global__ void k(const uint8_t* img, int rows, int cols, int th, int* out)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= cols || y >= rows) return;
    int s = 0;
    if (x >= 3 && y >= 3 && x < cols - 3 && y < rows - 3) s = fast_score_v<V>(img + y * cols + x, cols, th);
    out[y * cols + x] = s;
}
using std::min; using std::max;
int main()
{
    const int W = 256, H = 256;
    static uint8_t img[H * W]; static int o0[H * W], o1[H * W];
    srand(3);
    for (int i = 0; i < H * W; i++) img[i] = (rand() % 3 == 0) ? rand() % 256 : 100 + rand() % 40;
    uint8_t* d; int* dout;
    cudaMalloc(&d, H * W); cudaMalloc(&dout, H * W * 4);
    cudaMemcpy(d, img, H * W, cudaMemcpyHostToDevice);
    for (int th : {20, 7}) {
        k<0><<<dim3(W / 32, H / 8), dim3(32, 8)>>>(d, H, W, th, dout); cudaMemcpy(o0, dout, H * W * 4, cudaMemcpyDeviceToHost);
        k<2><<<dim3(W / 32, H / 8), dim3(32, 8)>>>(d, H, W, th, dout); cudaMemcpy(o1, dout, H * W * 4, cudaMemcpyDeviceToHost);
        int b0 = 0, b1 = 0, bh = 0, nc = 0;
        for (int y = 3; y < H - 3; y++) for (int x = 3; x < W - 3; x++) {
            int h0 = fast_score_v<0>(img + y * W + x, W, th), h1 = fast_score_v<1>(img + y * W + x, W, th);
            nc += h1 != 0; bh += h0 != h1; b0 += o0[y * W + x] != h1; b1 += o1[y * W + x] != h1;
        }
        printf("th %d corners %d host v0!=v1 %d  dev v0 bad %d  dev v2 bad %d  err %s\n", th, nc, bh, b0, b1, cudaGetErrorString(cudaGetLastError()));
    }
}
