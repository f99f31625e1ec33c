// Integer-pipe microbenchmark for the Hamming roofline denominator: measured issue rates of
// POPC, LOP3 (xor / full-adder), IADD3 and the two candidate Hamming inner loops on this GPU.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_int tools/ubench_int.cu
#include <cstdio>
#include <cstdint>
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

// Hamming inner loops: Q query descriptors in registers vs targets streamed from shared memory
template <int Q, int CSA> __global__ void hamming(const uint4* __restrict__ tg, int nt, uint32_t* out, uint32_t seed) {
    __shared__ uint4 st[2 * 256];
    uint32_t q[Q][8];
#pragma unroll
    for (int j = 0; j < Q; j++)
#pragma unroll
        for (int i = 0; i < 8; i++) q[j][i] = seed * (threadIdx.x * Q + j + 1) + i * 0x9e3779b9u;
    uint32_t best[Q];
#pragma unroll
    for (int j = 0; j < Q; j++) best[j] = 0xffffffffu;
    for (int base = 0; base < nt; base += 256) {
        __syncthreads();
        for (int i = threadIdx.x; i < 512; i += blockDim.x) st[i] = tg[(size_t)base * 2 + i];
        __syncthreads();
#pragma unroll 4
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
                    // carry-save adders: 7 words -> ones/twos/fours, plus the eighth word
                    const uint32_t o0 = x[0] ^ x[1] ^ x[2], t0 = (x[0] & x[1]) | (x[2] & (x[0] | x[1]));
                    const uint32_t o1 = x[3] ^ x[4] ^ x[5], t1 = (x[3] & x[4]) | (x[5] & (x[3] | x[4]));
                    const uint32_t o2 = o0 ^ o1 ^ x[6], t2 = (o0 & o1) | (x[6] & (o0 | o1));
                    const uint32_t tw = t0 ^ t1 ^ t2, fo = (t0 & t1) | (t2 & (t0 | t1));
                    d = __popc(o2) + __popc(x[7]) + 2 * __popc(tw) + 4 * __popc(fo);
                }
                const uint32_t key = (d << 22) | (uint32_t)(base + t);
                best[j] = min(best[j], key);
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < Q; j++) acc ^= best[j];
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
    uint4* tg; cudaMalloc(&tg, (size_t)nt * 32); cudaMemset(tg, 0x5a, (size_t)nt * 32);
    auto rep = [&](const char* name, float m, int Q) {
        const double pairs = (double)blocks * threads * Q * nt;
        printf(", \"%s_Gpairs_per_s\": %.1f", name, pairs / m / 1e6);
    };
    rep("hamming_q4_popc", timeit([&] { hamming<4, 0><<<blocks, threads>>>(tg, nt, out, 7); }), 4);
    rep("hamming_q4_csa", timeit([&] { hamming<4, 1><<<blocks, threads>>>(tg, nt, out, 7); }), 4);
    rep("hamming_q8_popc", timeit([&] { hamming<8, 0><<<blocks, threads>>>(tg, nt, out, 7); }), 8);
    rep("hamming_q8_csa", timeit([&] { hamming<8, 1><<<blocks, threads>>>(tg, nt, out, 7); }), 8);
    printf("}\n");
    return 0;
}
