#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cstdlib>
struct alignas(64) Maps { CUtensorMap m[16]; };
#ifndef PADB
#define PADB (2560 - 16)
#endif
struct Pad { char b[PADB]; };
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int MODE>
__global__ void k(const __grid_constant__ Pad pad, int lvl, int x, int y, int z, int bw, uint8_t* out, const __grid_constant__ Maps tm) {
    extern __shared__ __align__(128) uint8_t tile[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        uint32_t b = smem_addr(&bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(1) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bw * 40) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(smem_addr(tile)), "l"(&tm.m[lvl]), "r"(x), "r"(y), "r"(z), "r"(b) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_addr(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < bw * 40; i += blockDim.x) out[i] = tile[i] + pad.b[0];
}
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    const int X = argc > 1 ? atoi(argv[1]) : 13; const int BW = argc > 2 ? atoi(argv[2]) : 48;
    const int pitch = 640, h = 400, frames = 4;
    std::vector<uint8_t> img((size_t)pitch * h * frames);
    for (size_t i = 0; i < img.size(); i++) img[i] = (uint8_t)(i * 7 + i / pitch);
    uint8_t *d, *o; cudaMalloc(&d, img.size()); cudaMalloc(&o, 128 * 40);
    cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    printf("entry %d %d %p\n", (int)e, (int)q, fn);
    Maps M = {}; Pad P = {};
    for (int l = 0; l < 16; l++) {
        cuuint64_t dims[3] = {pitch, h, frames}, str[2] = {pitch, (cuuint64_t)pitch * h};
        cuuint32_t box[3] = {(cuuint32_t)BW, 40, 1}, es[3] = {1, 1, 1};
        CUresult r = ((enc_fn)fn)(&M.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (l == 0) printf("encode %d\n", (int)r);
    }
    for (int lvl : {0, 3, 7, 13}) {
        k<0><<<1, 128, 128 * 40>>>(P, lvl, X, 29, 2, BW, o, M);
        e = cudaDeviceSynchronize();
        std::vector<uint8_t> got(BW * 40);
        cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int r = 0; r < 40; r++) for (int c = 0; c < BW; c++) bad += got[r * BW + c] != img[(size_t)2 * pitch * h + (size_t)(29 + r) * pitch + X + c];
        printf("lvl %d: %s, mismatches %d\n", lvl, cudaGetErrorString(e), bad);
        if (e != cudaSuccess) break;
    }
    return 0;
}
