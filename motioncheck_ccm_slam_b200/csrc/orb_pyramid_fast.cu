// K1 (pyramid level) and K2+K3 (per-cell FAST-9/16 + NMS + threshold fallback + compaction), sm_100a.
// Integer/byte work bounded by HBM/L2 bandwidth and the ALU pipe; no tensor cores by design.
#include <cuda.h>   // CUtensorMap (types only; the encoder is fetched with cudaGetDriverEntryPoint, libcuda is not linked)
#include "ccm_common.cuh"

// ---------------------------------------------------------------------------------------------
// K1: one pyramid level from the previous one. Replaces cv::resize(INTER_LINEAR) at
// cslam/src/ORBextractor.cpp:1293 (ComputePyramid :1280-1304). Bit-exact 8U fixed-point bilinear:
// horizontal taps scaled by 2048 into int32, vertical (((b0*(r0>>4))>>16)+((b1*(r1>>4))>>16)+2)>>2.
// Each thread produces 4 horizontally adjacent pixels and stores one 32-bit word; a warp writes 128 B.
// For scale factors <= 2 the source pixels of the 4 outputs span at most 8 bytes: the thread loads three aligned
// words per source row, funnel-shifts them to the first tap, picks each output's two taps with one PRMT (selector
// from the host-built group table) and forms a0*p0 + a1*p1 with one IDP2A. k_resize_level_generic (byte loads) covers
// larger factors.
// ---------------------------------------------------------------------------------------------
// A thread walks RESIZE_ROWS consecutive output rows of its 4-pixel column: the horizontal pass of a source row is kept in
// registers and reused when the next output row starts on it (at scale 1.2 an output row's lower source row is the next
// row's upper one four times out of five), and the column's table entries are loaded once.
#define RESIZE_ROWS 4
// vertical pass of 4 outputs: v = (((b0 * h0) >> 16) + ((b1 * h1) >> 16) + 2) >> 2, v <= 255 (b0 + b1 = 2048, h <= 32640).
// With b pre-shifted by 16 each product-shift is one IMAD.HI (the +2 and the second product ride on its addend); two
// sums share a register (s0 + s1 << 16: both < 1024) for the final shift, and one PRMT gathers the four bytes.
__device__ __forceinline__ uint32_t resize_vpack(uint32_t b0s, uint32_t b1s, const uint32_t (&h0)[4], const uint32_t (&h1)[4]) {
    uint32_t t[4];
#pragma unroll
    for (int k = 0; k < 4; k++) t[k] = __umulhi(b1s, h1[k]) + (__umulhi(b0s, h0[k]) + 2u);
    const uint32_t q01 = (t[0] + (t[1] << 16)) >> 2, q23 = (t[2] + (t[3] << 16)) >> 2;
    return __byte_perm(q01, q23, 0x6420);
}
__device__ __forceinline__ void resize_hrow(const uint8_t* __restrict__ rp, bool in1, bool in2, uint32_t shift, uint32_t sel4,
                                            const uint32_t (&coef)[4], uint32_t (&r)[4]) {
    const uint32_t u0 = *reinterpret_cast<const uint32_t*>(rp);
    const uint32_t u1 = in1 ? *reinterpret_cast<const uint32_t*>(rp + 4) : 0u;
    const uint32_t u2 = in2 ? *reinterpret_cast<const uint32_t*>(rp + 8) : 0u;
    const uint32_t lo = __funnelshift_r(u0, u1, shift), hi = __funnelshift_r(u1, u2, shift);
#pragma unroll
    for (int i = 0; i < 4; i++) r[i] = __dp2a_lo(coef[i], __byte_perm(lo, hi, sel4 >> (8 * i)), 0u) >> 4;   // IDP2A reads result bytes 0,1 only: selector nibbles 2,3 are don't-care (all < 8)
}

__global__ void __launch_bounds__(256) k_resize_level(const __grid_constant__ LevelTable T, int lvl) {
    const LevelDesc& D = T.l[lvl];
    const LevelDesc& S = T.l[lvl - 1];
    const int f = blockIdx.z;
    const int g = blockIdx.x * 32 + threadIdx.x, x4 = g * 4;
    const int y0 = (blockIdx.y * 8 + threadIdx.y) * RESIZE_ROWS;
    pdl_launch();
    if (y0 >= D.h || x4 >= D.w) { pdl_wait(); return; }      // no CTA leaves before its predecessor is complete (the chain is transitive)
    // group table (resize_group_table, orb_api.cu): {first tap's aligned byte offset, bit shift to the first tap,
    // 4 x (tap0 | tap1 << 4) byte selectors, -} {4 x (a0 | a1 << 16)}
    const uint4 ga = __ldg(D.xgrp + 2 * g), gc = __ldg(D.xgrp + 2 * g + 1);
    const int base = (int)ga.x, wlim = S.w;                   // words starting at or beyond the row's last pixel are not read
    const bool in1 = base + 4 < wlim, in2 = base + 8 < wlim;
    const uint32_t coef[4] = {gc.x, gc.y, gc.z, gc.w};
    const uint8_t* __restrict__ src = S.img + (size_t)f * S.img_stride + base;
    uint8_t* __restrict__ dst = D.img + (size_t)f * D.img_stride + (size_t)y0 * D.img_pitch + x4;
    const int pitch = S.img_pitch, last = S.h - 1, ny = min(RESIZE_ROWS, D.h - y0);
    int cached = -1;
    uint32_t hc[4] = {0u, 0u, 0u, 0u};                        // horizontal pass of source row `cached`
    const int sy_first = __ldg(D.yofs + y0);                  // host tables: loaded before the previous level is waited for
    (void)sy_first;
    pdl_wait();                                               // level lvl-1 is complete
#pragma unroll
    for (int j = 0; j < RESIZE_ROWS; j++) {
        if (j < ny) {                                         // warp-uniform (a warp is one row strip)
            const int y = y0 + j;
            const int sy0 = __ldg(D.yofs + y), sy1 = min(sy0 + 1, last);
            const uint32_t b0 = (uint32_t)__ldg(D.ycoef + 2 * y), b1 = (uint32_t)__ldg(D.ycoef + 2 * y + 1);
            uint32_t h0[4], h1[4];
            if (sy0 == cached) {
#pragma unroll
                for (int i = 0; i < 4; i++) h0[i] = hc[i];
            } else resize_hrow(src + (size_t)sy0 * pitch, in1, in2, ga.y, ga.z, coef, h0);
            if (sy1 == sy0) {
#pragma unroll
                for (int i = 0; i < 4; i++) h1[i] = h0[i];
            } else resize_hrow(src + (size_t)sy1 * pitch, in1, in2, ga.y, ga.z, coef, h1);
            const uint32_t packed = resize_vpack(b0 << 16, b1 << 16, h0, h1);
#pragma unroll
            for (int i = 0; i < 4; i++) hc[i] = h1[i];
            cached = sy1;
            *reinterpret_cast<uint32_t*>(dst) = packed;       // pitch is a multiple of 32: in-bounds
            dst += D.img_pitch;
        }
    }
}

// Loads-first variant for scale factors where RESIZE_ROWS output rows tap at most RESIZE_SRC_ROWS source rows (the host
// checks the y table; 6 rows at ORB-SLAM's 1.2). k_resize_level interleaves each source row's loads with its arithmetic,
// so a thread has 3-6 loads in flight and the kernel waits on HBM/L2 latency (ncu: long-scoreboard 7.4 per issue at IPC
// 2.0); here the strip's RESIZE_SRC_ROWS x 3 words are all requested before the first is used.
#define RESIZE_SRC_ROWS 6
#ifndef RESIZE_MLP_CTAS
#define RESIZE_MLP_CTAS 4
#endif
__global__ void __launch_bounds__(256, RESIZE_MLP_CTAS) k_resize_level_mlp(const __grid_constant__ LevelTable T, int lvl) {
    const LevelDesc& D = T.l[lvl];
    const LevelDesc& S = T.l[lvl - 1];
    const int f = blockIdx.z;
    const int g = blockIdx.x * 32 + threadIdx.x, x4 = g * 4;
    const int y0 = (blockIdx.y * 8 + threadIdx.y) * RESIZE_ROWS;
    pdl_launch();
    if (y0 >= D.h || x4 >= D.w) { pdl_wait(); return; }
    const uint4 ga = __ldg(D.xgrp + 2 * g), gc = __ldg(D.xgrp + 2 * g + 1);
    const int base = (int)ga.x, wlim = S.w;
    const bool in1 = base + 4 < wlim, in2 = base + 8 < wlim;
    const uint32_t coef[4] = {gc.x, gc.y, gc.z, gc.w};
    const int pitch = S.img_pitch, last = S.h - 1, ny = min(RESIZE_ROWS, D.h - y0);
    int off[RESIZE_ROWS]; uint32_t bb[RESIZE_ROWS];
    const int s_first = __ldg(D.yofs + y0);
#pragma unroll
    for (int j = 0; j < RESIZE_ROWS; j++) {
        const int y = min(y0 + j, D.h - 1);
        off[j] = __ldg(D.yofs + y) - s_first;                                          // j or j + 1 ... checked on the host: off[j] + 1 < RESIZE_SRC_ROWS
        bb[j] = __ldg(reinterpret_cast<const uint32_t*>(D.ycoef) + y);                 // b0 | b1 << 16
    }
    const uint8_t* src = S.img + (size_t)f * S.img_stride + base;
    uint8_t* dst = D.img + (size_t)f * D.img_stride + (size_t)y0 * D.img_pitch + x4;
    pdl_wait();                                               // level lvl-1 is complete
    CCM_TS(blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0 && threadIdx.y == 0, lvl);
    uint32_t u[RESIZE_SRC_ROWS][3];
#pragma unroll
    for (int k = 0; k < RESIZE_SRC_ROWS; k++) {
        const uint8_t* rp = src + (size_t)min(s_first + k, last) * pitch;
        u[k][0] = *reinterpret_cast<const uint32_t*>(rp);
        u[k][1] = in1 ? *reinterpret_cast<const uint32_t*>(rp + 4) : 0u;
        u[k][2] = in2 ? *reinterpret_cast<const uint32_t*>(rp + 8) : 0u;
    }
    uint32_t H[RESIZE_SRC_ROWS][4];
#pragma unroll
    for (int k = 0; k < RESIZE_SRC_ROWS; k++) {
        const uint32_t lo = __funnelshift_r(u[k][0], u[k][1], ga.y), hi = __funnelshift_r(u[k][1], u[k][2], ga.y);
#pragma unroll
        for (int i = 0; i < 4; i++) H[k][i] = __dp2a_lo(coef[i], __byte_perm(lo, hi, ga.z >> (8 * i)), 0u) >> 4;
    }
#pragma unroll
    for (int j = 0; j < RESIZE_ROWS; j++) {
        if (j < ny) {
            // row j starts at source row s_first + off[j], off[j] in [j, j + 2) for scale factors below 4/3: a warp-uniform branch
            uint32_t packed;
            if (off[j] != j) packed = resize_vpack(bb[j] << 16, bb[j] & 0xffff0000u, H[j + 1], H[j + 2 < RESIZE_SRC_ROWS ? j + 2 : RESIZE_SRC_ROWS - 1]);
            else packed = resize_vpack(bb[j] << 16, bb[j] & 0xffff0000u, H[j], H[j + 1]);
            *reinterpret_cast<uint32_t*>(dst) = packed;
            dst += D.img_pitch;
        }
    }
}

__global__ void __launch_bounds__(256) k_resize_level_generic(const __grid_constant__ LevelTable T, int lvl) {
    const LevelDesc& D = T.l[lvl];
    const LevelDesc& S = T.l[lvl - 1];
    const int f = blockIdx.z;
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (y >= D.h || x4 >= D.w) return;
    const uint8_t* __restrict__ src = S.img + (size_t)f * S.img_stride;
    uint8_t* __restrict__ dst = D.img + (size_t)f * D.img_stride;
    const int sy0 = __ldg(D.yofs + y);
    const int sy1 = min(sy0 + 1, S.h - 1);
    const int b0 = __ldg(D.ycoef + 2 * y), b1 = __ldg(D.ycoef + 2 * y + 1);
    const uint8_t* r0p = src + (size_t)sy0 * S.img_pitch;
    const uint8_t* r1p = src + (size_t)sy1 * S.img_pitch;
    uint32_t packed = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int x = min(x4 + i, D.w - 1);
        const int sx0 = __ldg(D.xofs + x);
        const int sx1 = min(sx0 + 1, S.w - 1);
        const int a0 = __ldg(D.xcoef + 2 * x), a1 = __ldg(D.xcoef + 2 * x + 1);
        const int r0 = r0p[sx0] * a0 + r0p[sx1] * a1;
        const int r1 = r1p[sx0] * a0 + r1p[sx1] * a1;
        const int v = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2;
        packed |= (uint32_t)(v & 255) << (8 * i);
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * D.img_pitch + x4) = packed;   // pitch is a multiple of 32: in-bounds
}

// ---------------------------------------------------------------------------------------------
// K2+K3 building blocks: the FAST-9/16 corner score and the TMA / mbarrier helpers of the tile load. The kernel itself is
// k_fast_chunks (orb_fast_chunks.cu). A cell's detection range is its (wCell x hCell) block of [19,w-19)x[19,h-19);
// NMS neighbours outside that block count as 0 exactly as in a per-cell cv::FAST call. Survivors are
// appended to the (frame, level) corner list; list order is not significant (the quadtree's only order dependence is
// re-derived from coordinates, see orb_quadtree.cu).
// ---------------------------------------------------------------------------------------------
// cornerScore<16> (OpenCV fast_score.cpp): max over the 16 arcs of 9 ring pixels of min(d) and of min(-d), minus 1, with
// d = v - ring pixel. d + 256 and 256 - d (both in [1, 511]) travel as the two halves of one register - one IMAD per ring
// pixel, r * 65535 + ((v + 256) * (1 - 65536) + (512 << 16)), no borrow between the halves - so each step of the arc
// reduction is one VIMNMX.S16x2 / VIMNMX3.S16x2 for both polarities: min over 2, over 4, then min3(4-run, 4-run, 9th
// pixel) and a max3 over pairs of arcs: 56 packed min/max instead of 160 scalar ones. (The scalar form also had to keep -d
// behind an opaque asm: ptxas 12.9 for sm_100a miscompiles max(max(x,-y),max(z,-w)) into a VIMNMX3 with negated operands,
// DESIGN.md "toolchain notes"; nothing is negated inside a min/max here.)
// A pixel is a FAST-9 corner at threshold t exactly when this score is >= t (some arc has every |d| > t, same sign), so
// the segment test and the score are one computation.
__device__ __forceinline__ int fast_score16(int v, const uint32_t (&r)[16]) {
    uint32_t bias = (uint32_t)(v + 256) * 0xFFFF0001u + (512u << 16);
    asm("" : "+r"(bias));      // opaque: otherwise the compiler rewrites r * 65535 + bias as (r - v) * 65535 + const, two instructions per ring pixel
    uint32_t p[16], m2[16], m4[16], a9[16];
#pragma unroll
    for (int k = 0; k < 16; k++) p[k] = r[k] * 65535u + bias;            // lo = d + 256, hi = 256 - d
#pragma unroll
    for (int k = 0; k < 16; k++) m2[k] = __vmins2(p[k], p[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) m4[k] = __vmins2(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) a9[k] = __vimin3_s16x2(m4[k], m4[(k + 4) & 15], p[(k + 8) & 15]);
    uint32_t best = __vmaxs2(a9[0], a9[1]);
#pragma unroll
    for (int k = 2; k < 16; k += 2) best = __vimax3_s16x2(best, a9[k], a9[k + 1]);
    return (int)max(best & 0xffffu, best >> 16) - 257;
}

// 32-bit shared-memory address of a generic pointer (PTX operands of the mbarrier / TMA / st.shared helpers)
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// TMA tile load (cp.async.bulk.tensor, sm_90+): one elected thread asks the copy engine for the (cell + 3-px ring) box of
// its level image; the bytes land in shared memory and complete an mbarrier transaction count. One descriptor per level
// (u8 tensor [frames][h][pitch], box = [1][tile_rows][TS]), passed by value as a __grid_constant__ parameter.
struct alignas(64) FastTma { CUtensorMap m[CCM_MAX_LEVELS]; };
__device__ __forceinline__ void mbar_init(uint32_t bar_s, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar_s), "r"(count) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");            // the init must be visible to the async (TMA) proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar_s, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst_s, const CUtensorMap* map, int x, int y, int z, uint32_t bar_s) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst_s), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar_s) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar_s, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
                 ::"r"(bar_s), "r"(parity) : "memory");
}
