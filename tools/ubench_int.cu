// Integer-pipe microbenchmark for the Hamming roofline denominator: measured issue rates of
// POPC, LOP3 (xor / full-adder), IADD3 and the two candidate Hamming inner loops on this GPU.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_int tools/ubench_int.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define ITERS 4096

template <int OP> __global__ void k(uint32_t* out, uint32_t seed) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9e3779b9u;
    uint32_t acc = 0;
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) a[i] = __popc(a[i]) + seed;                       // POPC + IADD
            if (OP == 1) a[i] = (a[i] ^ a[(i + 1) & 7]) + 1;               // LOP3/IADD chain (alu)
            if (OP == 2) a[i] = a[i] + a[(i + 3) & 7] + seed;              // IADD3
            if (OP == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]), "r"(seed));   // LOP3 only
            if (OP == 4) a[i] = __popc(a[i] ^ seed);                       // XOR+POPC dependent pair, no add
        }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) acc ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// Hamming inner loops: Q query descriptors in registers vs targets streamed from shared memory.
// CSA = 0: the textbook loop (8 XOR + 8 POPC + adds per pair). CSA = 1: the loop of the product kernel k_allpairs_best2
// (motioncheck_ccm_slam_b200/csrc/match_allpairs.cu) restated instruction for instruction - carry-save tree as explicit LOP3
// (0x96 = xor3, 0xe8 = majority), 4 POPC, packed (distance << 23 | index) keys with a branch-free best / second-best
// update, next tile prefetched into registers while the current one is used - on a database that stays in L2, so the
// figure is the ceiling of THAT loop and the product kernel (which also streams its shard from HBM) cannot beat it.
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
template <int Q, int CSA> __global__ void __launch_bounds__(256) hamming(const uint4* __restrict__ tg, int nt, uint32_t* out, uint32_t seed) {
    __shared__ uint4 st[2 * 256];
    uint32_t q[Q][8];
#pragma unroll
    for (int j = 0; j < Q; j++)
#pragma unroll
        for (int i = 0; i < 8; i++) q[j][i] = seed * (threadIdx.x * Q + j + 1) + i * 0x9e3779b9u;
    uint32_t b1[Q], b2[Q];
#pragma unroll
    for (int j = 0; j < Q; j++) b1[j] = b2[j] = 0xffffffffu;
    uint4 p0 = tg[threadIdx.x], p1 = tg[threadIdx.x + 256];
    for (int base = 0; base < nt; base += 256) {
        __syncthreads();
        st[threadIdx.x] = p0; st[threadIdx.x + 256] = p1;
        __syncthreads();
        if (base + 256 < nt) { p0 = tg[(size_t)(base + 256) * 2 + threadIdx.x]; p1 = tg[(size_t)(base + 256) * 2 + threadIdx.x + 256]; }
#pragma unroll 2
        for (int t = 0; t < 256; t++) {
            const uint4 lo = st[2 * t], hi = st[2 * t + 1];
            const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
            for (int j = 0; j < Q; j++) {
                uint32_t d;
                if (!CSA) {
                    d = 0;
#pragma unroll
                    for (int i = 0; i < 8; i++) d += __popc(q[j][i] ^ w[i]);
                } else {
                    uint32_t x[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) x[i] = q[j][i] ^ w[i];
                    const uint32_t o0 = xor3(x[0], x[1], x[2]), t0 = maj3(x[0], x[1], x[2]);
                    const uint32_t o1 = xor3(x[3], x[4], x[5]), t1 = maj3(x[3], x[4], x[5]);
                    const uint32_t o2 = xor3(o0, o1, x[6]), t2 = maj3(o0, o1, x[6]);
                    const uint32_t tw = xor3(t0, t1, t2), fo = maj3(t0, t1, t2);
                    d = __popc(o2) + __popc(x[7]) + 2 * __popc(tw) + 4 * __popc(fo);
                }
                const uint32_t key = (d << 23) + (uint32_t)(base + t);
                const uint32_t m = max(b1[j], key);
                b1[j] = min(b1[j], key);
                b2[j] = min(b2[j], m);
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < Q; j++) acc ^= b1[j] ^ b2[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <typename F> float timeit(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 8, threads = 256;
    uint32_t* out; cudaMalloc(&out, sizeof(uint32_t) * blocks * threads * 4);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d", p.name, sms, clk);
    const char* names[5] = {"popc_iadd", "xor_iadd", "iadd3", "lop3", "xor_popc"};
    float ms[5];
    ms[0] = timeit([&] { k<0><<<blocks, threads>>>(out, 3); });
    ms[1] = timeit([&] { k<1><<<blocks, threads>>>(out, 3); });
    ms[2] = timeit([&] { k<2><<<blocks, threads>>>(out, 3); });
    ms[3] = timeit([&] { k<3><<<blocks, threads>>>(out, 3); });
    ms[4] = timeit([&] { k<4><<<blocks, threads>>>(out, 3); });
    for (int i = 0; i < 5; i++) {
        const double ops = (double)blocks * threads * ITERS * 8;      // loop bodies (each 1-2 instructions)
        printf(", \"%s_Gbody_per_s\": %.1f, \"%s_body_per_clk_per_sm\": %.2f", names[i], ops / ms[i] / 1e6, names[i],
               ops / (ms[i] * 1e-3) / sms / (clk * 1e3));
    }
    const int nt = 1 << 16;
    uint4* tg; cudaMalloc(&tg, (size_t)nt * 32);
    {   // random descriptors: constant data toggles few bits and flatters the loop
        uint32_t* h = (uint32_t*)malloc((size_t)nt * 32);
        uint32_t x = 0x12345678u;
        for (size_t i = 0; i < (size_t)nt * 8; i++) { x ^= x << 13; x ^= x >> 17; x ^= x << 5; h[i] = x; }
        cudaMemcpy(tg, h, (size_t)nt * 32, cudaMemcpyHostToDevice);
        free(h);
    }
    auto rep = [&](const char* name, float m, int Q) {
        const double pairs = (double)blocks * threads * Q * nt;
        printf(", \"%s_Gpairs_per_s\": %.1f", name, pairs / m / 1e6);
    };
    rep("hamming_q4_popc", timeit([&] { hamming<4, 0><<<blocks, threads>>>(tg, nt, out, 7); }), 4);
    rep("hamming_q4_csa", timeit([&] { hamming<4, 1><<<blocks, threads>>>(tg, nt, out, 7); }), 4);
    {   // the product kernel's geometry: one wave, three CTAs per SM (73 registers)
        const int b3 = sms * 3;
        const float m3 = timeit([&] { hamming<4, 1><<<b3, threads>>>(tg, nt, out, 7); });
        printf(", \"hamming_q4_csa_one_wave_Gpairs_per_s\": %.1f", (double)b3 * threads * 4 * nt / m3 / 1e6);
    }
    rep("hamming_q8_popc", timeit([&] { hamming<8, 0><<<blocks, threads>>>(tg, nt, out, 7); }), 8);
    rep("hamming_q8_csa", timeit([&] { hamming<8, 1><<<blocks, threads>>>(tg, nt, out, 7); }), 8);
    printf("}\n");
    return 0;
}
