// tma_probe.cu - which way of handing a CUtensorMap to cp.async.bulk.tensor.2d works on this toolchain (sm_100a):
//   1: one map as a __grid_constant__ parameter   2: array of maps in a __grid_constant__ struct, dynamic index
//   3: array of maps in global memory
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_probe tma_probe.cu ; run: ./tma_probe <variant>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#define ROWS 134
struct Maps { CUtensorMap map[16]; };
__device__ __forceinline__ void load_tile(uint32_t (*tile)[32], uint64_t* mbar, const CUtensorMap* m, int cx, int cy)
{
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(mbar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mb));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"((uint32_t)(ROWS * 128)) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     :: "r"((uint32_t)__cvta_generic_to_shared(&tile[0][0])), "l"(reinterpret_cast<uint64_t>(m)), "r"(cx), "r"(cy), "r"(mb) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(mb) : "memory");
}
#define BODY(MAPPTR)                                                                   \
    __shared__ __align__(128) uint32_t tile[ROWS][32];                                 \
    __shared__ __align__(8) uint64_t mbar;                                             \
    load_tile(tile, &mbar, MAPPTR, cx, cy);                                            \
    uint32_t s = 0;                                                                    \
    for (int r = 0; r < ROWS; r++) s += tile[r][threadIdx.x & 31];                     \
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
__global__ void k1(const __grid_constant__ CUtensorMap M, int cx, int cy, uint32_t* out) { BODY(&M) }
__global__ void k2(const __grid_constant__ Maps M, int lvl, int cx, int cy, uint32_t* out) { BODY(&M.map[lvl]) }
__global__ void k3(const CUtensorMap* M, int lvl, int cx, int cy, uint32_t* out) { BODY(M + lvl) }
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                           const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv)
{
    const int variant = argc > 1 ? atoi(argv[1]) : 1;
    const int pitch = 896, prow = 518, batch = 4;
    uint8_t* d; cudaMalloc(&d, (size_t)pitch * prow * batch);
    std::vector<uint8_t> h((size_t)pitch * prow * batch);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)(i * 7 + (i >> 9));
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) { printf("no entry point\n"); return 2; }
    Maps M;
    const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)prow * batch}; const cuuint64_t strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {128u, ROWS}; const cuuint32_t estr[2] = {1u, 1u};
    for (int l = 0; l < 16; l++) {
        CUresult r = ((enc_fn)fn)(&M.map[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 3; }
    }
    uint32_t* out; cudaMalloc(&out, 128 * 4);
    const int cx = 28 + 120, cy = 1 * prow + 19 + 128 - 3;
    if (variant == 1) k1<<<1, 128>>>(M.map[3], cx, cy, out);
    else if (variant == 2) k2<<<1, 128>>>(M, 3, cx, cy, out);
    else { CUtensorMap* dm; cudaMalloc(&dm, sizeof(M)); cudaMemcpy(dm, &M, sizeof(M), cudaMemcpyHostToDevice); k3<<<1, 128>>>(dm, 3, cx, cy, out); }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("variant %d: %s\n", variant, cudaGetErrorString(e)); return 1; }
    uint32_t ho[128]; cudaMemcpy(ho, out, sizeof(ho), cudaMemcpyDeviceToHost);
    uint32_t want = 0;
    for (int r = 0; r < ROWS; r++) { const uint8_t* p = &h[(size_t)(cy + r) * pitch + cx + 4 * 5]; want += (uint32_t)p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
    printf("variant %d: ok, lane 5 sum %u (want %u) %s\n", variant, ho[5], want, ho[5] == want ? "MATCH" : "MISMATCH");
    return 0;
}
