// stream_patterns.cu — achievable HBM bandwidth of the access patterns used by the pyramid / blur kernels:
// a plane set of F frames (pitch 768 x 480 rows), each thread owns a column of W bytes and walks ROWS rows.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
template <typename T, int ROWS, int PF>
__global__ void colwalk(const T* __restrict__ in, T* __restrict__ out, int pw /*elements per row*/, int h, long long plane, int wcols)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int y0 = (blockIdx.y * blockDim.y + threadIdx.y) * ROWS;
    if (c >= wcols || y0 >= h) return;
    const T* p = in + blockIdx.z * plane + (long long)y0 * pw + c;
    T* q = out + blockIdx.z * plane + (long long)y0 * pw + c;
    T buf[PF];
#pragma unroll
    for (int i = 0; i < PF; i++) buf[i] = __ldg(p + (long long)min(i, ROWS - 1) * pw);
#pragma unroll
    for (int r = 0; r < ROWS; r++) {
        const T v = buf[r % PF];
        if (r + PF < ROWS) buf[r % PF] = __ldg(p + (long long)(r + PF) * pw);
        q[(long long)r * pw] = v;
    }
}
__global__ void rowmajor(const uint4* __restrict__ in, uint4* __restrict__ out, long long n)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = __ldg(in + i);
}
template <typename F> float timeit(F f) { cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); f(); cudaDeviceSynchronize(); cudaEventRecord(a); for (int i = 0; i < 5; i++) f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); return ms / 5; }
int main()
{
    const int F = 512, pitch = 768, h = 480, w = 752;
    const long long plane = (long long)pitch * h, bytes = plane * F;
    uint8_t *in, *out; cudaMalloc(&in, bytes); cudaMalloc(&out, bytes); cudaMemset(in, 1, bytes);
    const double gb = 2.0 * (double)w * h * F / 1e9;
#define RUN(T, ROWS, PF, BX, BY) { const int wc = w / (int)sizeof(T); dim3 b(BX, BY), g((wc + BX - 1) / BX, (h + BY * ROWS - 1) / (BY * ROWS), F); \
    float ms = timeit([&] { colwalk<T, ROWS, PF><<<g, b>>>((const T*)in, (T*)out, pitch / (int)sizeof(T), h, plane / (long long)sizeof(T), wc); }); \
    printf("colwalk %-8s rows %2d prefetch %d block %dx%d : %7.3f ms %7.1f GB/s\n", #T, ROWS, PF, BX, BY, ms, gb / ms * 1e3); }
    RUN(uint32_t, 32, 1, 64, 2) RUN(uint32_t, 32, 2, 64, 2) RUN(uint32_t, 32, 4, 64, 2) RUN(uint32_t, 32, 8, 64, 2)
    RUN(uint32_t, 8, 1, 64, 4) RUN(uint32_t, 8, 4, 64, 4) RUN(uint32_t, 8, 8, 64, 4)
    RUN(uint32_t, 32, 4, 192, 1) RUN(uint32_t, 16, 4, 192, 1)
    RUN(uint2, 32, 1, 32, 4) RUN(uint2, 32, 4, 32, 4) RUN(uint4, 32, 1, 16, 8) RUN(uint4, 32, 4, 16, 8) RUN(uint4, 8, 8, 16, 8) RUN(uint4, 32, 4, 48, 2)
    { float ms = timeit([&] { rowmajor<<<148 * 16, 256>>>((const uint4*)in, (uint4*)out, bytes / 16); }); printf("rowmajor uint4 grid-stride: %7.3f ms %7.1f GB/s (full pitch)\n", ms, 2.0 * bytes / 1e9 / ms * 1e3); }
    return 0;
}
