// pipes.cu — issue-rate microbenchmark for the integer / half2 min-max instructions the FAST and
// blur kernels are built from (B200, sm_100a).  Prints warp-instructions per cycle per SM.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define ITERS 4096
#define NACC 8
template <int OP> __device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c)
{
    if (OP == 0) return __vimax3_u16x2(a, b, c);
    if (OP == 1) return __vmaxu2(a, b);
    if (OP == 2) { __half2 r = __hmax2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }
    if (OP == 3) return __funnelshift_r(a, b, 8);
    if (OP == 4) return __byte_perm(a, b, 0x4321);
    if (OP == 5) return a * b + c;            // IMAD
    if (OP == 6) return (a & b) | c;          // LOP3
    if (OP == 7) return a + b + c;            // IADD3
    if (OP == 8) return max(max((int)a, (int)b), (int)c);   // VIMNMX3 s32
    if (OP == 9) return __vabsdiffu4(a, b);
    if (OP == 10) return __dp4a(a, b, c);
    if (OP == 11) return __dp2a_lo(a, b, c);
    if (OP == 12) return __viaddmax_s16x2(a, b, c);
    if (OP == 13) return __vadd2(a, b);
    if (OP == 14) return __popc(a) + b;
    if (OP == 15) return __umulhi(a, b);
    return 0;
}
template <int OP> __global__ void k(uint32_t* out, uint32_t s0, uint32_t s1)
{
    uint32_t acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 7 + i + s0;
    const uint32_t b = s1 + threadIdx.x, c = s0 ^ 0x12345;
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < NACC; i++) acc[i] = op<OP>(acc[i], acc[(i + 3) % NACC], c + r);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// mixed: half of the chains on VIMNMX3.U16x2, half on HMNMX2
__global__ void kmix(uint32_t* out, uint32_t s0, uint32_t s1)
{
    uint32_t acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 7 + i + s0;
    const uint32_t b = s1 + threadIdx.x, c = s0 ^ 0x12345;
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < NACC; i++) acc[i] = (i & 1) ? op<0>(acc[i], acc[(i + 2) % NACC], c + r) : op<2>(acc[i], acc[(i + 2) % NACC], c);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void kmix2(uint32_t* out, uint32_t s0, uint32_t s1)   // VIMNMX3 + IMAD
{
    uint32_t acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 7 + i + s0;
    const uint32_t b = s1 + threadIdx.x, c = s0 ^ 0x12345;
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < NACC; i++) acc[i] = (i & 1) ? op<0>(acc[i], acc[(i + 2) % NACC], c + r) : op<5>(acc[i], acc[(i + 2) % NACC], c);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F> static void run(const char* name, F launch, int sms, double mhz)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double winstr = (double)sms * 8 * (1024 / 32) * ITERS * 4.0 * NACC;   // 8 blocks of 1024/SM... see launch
    printf("%-28s %8.3f ms  %6.2f warp-instr/clk/SM (at %.0f MHz)\n", name, ms, winstr / (ms * 1e-3) / (mhz * 1e6) / sms, mhz);
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double mhz = khz / 1000.0;
    uint32_t* out; cudaMalloc(&out, (size_t)sms * 8 * 1024 * 4);
    dim3 g(sms * 8), b(1024);
#define RUN(OP, NAME) run(NAME, [&] { k<OP><<<g, b>>>(out, 1, 2); }, sms, mhz)
    RUN(0, "VIMNMX3.U16x2"); RUN(1, "VIMNMX.U16x2"); RUN(2, "HMNMX2"); RUN(3, "SHF.R (funnel)"); RUN(4, "PRMT");
    RUN(5, "IMAD"); RUN(6, "LOP3"); RUN(7, "IADD3"); RUN(8, "VIMNMX3.S32"); RUN(9, "VABSDIFF4"); RUN(10, "IDP.4A");
    RUN(11, "IDP.2A"); RUN(12, "VIADDMNMX.S16x2"); RUN(13, "VIADD.16x2"); RUN(14, "POPC+IADD"); RUN(15, "IMAD.HI");
    run("mix VIMNMX3.U16x2 + HMNMX2", [&] { kmix<<<g, b>>>(out, 1, 2); }, sms, mhz);
    run("mix VIMNMX3.U16x2 + IMAD", [&] { kmix2<<<g, b>>>(out, 1, 2); }, sms, mhz);
    printf("SMs %d clock %.0f MHz\n", sms, mhz);
    return 0;
}
