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
// K2+K3: FAST-9/16 per 30-px cell. Replaces the cell loop of ComputeKeyPointsOctTree,
// cslam/src/ORBextractor.cpp:957-998 (cv::FAST(cell, iniTh, nms) and the minTh rerun when the cell is empty).
// One CTA per (cell, frame). The cell's detection range is its (wCell x hCell) block of [19,w-19)x[19,h-19);
// NMS neighbours outside that block count as 0 exactly as in a per-cell cv::FAST call.
// Survivors are appended to the (frame, level) corner list with a warp-aggregated atomic; list order is not
// significant (the quadtree's only order dependence is re-derived from coordinates, see orb_quadtree.cu).
// ---------------------------------------------------------------------------------------------
#define FAST_MAX_CELL 64                       // wCell,hCell <= 60 (width/nCols with nCols=floor(width/30))
#define FAST_TILE (FAST_MAX_CELL + 6)

__device__ __forceinline__ bool has_run9(uint32_t m16) {
    uint32_t x = m16 | (m16 << 16);
    x &= x >> 1;   // runs >= 2
    x &= x >> 2;   // runs >= 4
    x &= x >> 4;   // runs >= 8
    x &= x >> 1;   // runs >= 9
    return x != 0;
}

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

// Work is staged through shared-memory lists so that the expensive steps only run on pixels that need them, with all
// lanes busy: (B) 4-point compass test on every pixel (a 9-arc always holds two adjacent compass points) -> candidate
// list; (C) full 16-ring FAST-9 test on candidates -> corner list; (D) cornerScore on corners -> score map;
// (E) 3x3 NMS on corners only -> warp-aggregated append to the (frame, level) corner list in HBM.
#define FAST_TS 76                             // tile row stride: 64 + 6 halo + up to 3 bytes of word-alignment shift

// Shared-memory list plumbing written as PTX on 32-bit shared addresses: one ATOMS per warp (nvcc otherwise wraps every
// atomicAdd in its own warp-aggregation sequence) and predicated 16-bit stores instead of a branch per store.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t atoms_add(uint32_t saddr, uint32_t v) {
    uint32_t r;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(r) : "r"(saddr), "r"(v) : "memory");
    return r;
}
__device__ __forceinline__ void sts_u16_if(uint32_t pred, uint32_t saddr, uint32_t val) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.shared.u16 [%1], %2;\n\t}" ::"r"(pred), "r"(saddr), "h"((unsigned short)val) : "memory");
}
__device__ __forceinline__ uint32_t lanemask_lt() { uint32_t m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

// append val to a shared u16 list (byte address list_s, counter at count_s) for the lanes with pred set
__device__ __forceinline__ void list_append(bool pred, uint32_t val, uint32_t list_s, uint32_t count_s) {
    const uint32_t ballot = __ballot_sync(0xffffffffu, pred);
    if (!ballot) return;
    uint32_t base = 0;
    if ((threadIdx.x & 31) == 0) base = atoms_add(count_s, __popc(ballot));
    base = __shfl_sync(0xffffffffu, base, 0);
    sts_u16_if(pred, list_s + 2 * (base + __popc(ballot & lanemask_lt())), val);
}

// CTA size NT is a template parameter (96 / 128 / 160 threads): a cell is ngx x ch 4-pixel groups (~290 at the reference's 30-px
// cells) and the host picks the NT that wastes the fewest lanes in the sweeps over them (fast_pick_threads, orb_api.cu)
// Dynamic shared memory, sized by the host for the largest cell of the configured shape (fast_smem_layout):
// tile [max_ch + 6][ts] | score [max_ch + 2][ss] | cand [max_px] u16 | corn [max_px] u16
struct FastSmem { int score_off, cand_off, corn_off, total, ts, tile_rows, ss; };
static inline FastSmem fast_smem_layout(int max_cw, int max_ch, int ts = FAST_TS) {
    FastSmem m;
    const int tile_bytes = ((max_ch + 6) * ts + 127) / 128 * 128;
    m.ss = (max_cw + 2 + 3) / 4 * 4;                                 // score row stride: widest cell + its 1-px zero ring
    const int score_bytes = ((max_ch + 2) * m.ss + 15) / 16 * 16;
    const int list_bytes = (max_cw * max_ch * 2 + 15) / 16 * 16;
    m.score_off = tile_bytes; m.cand_off = m.score_off + score_bytes; m.corn_off = m.cand_off + list_bytes;
    m.total = m.corn_off + list_bytes;
    m.ts = ts; m.tile_rows = max_ch + 6;
    return m;
}

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

#ifndef FAST_RESIDENT_THREADS
#define FAST_RESIDENT_THREADS 2048          // threads per SM the register budget is sized for (32 registers per thread)
#endif
template <int TS, bool TMA, int NT>
__global__ void __launch_bounds__(NT, FAST_RESIDENT_THREADS / NT) k_fast_cells(const __grid_constant__ LevelTable T,
                                                    const uint4* __restrict__ cell_info,  // per cell 2 x uint4, see configure_shape
                                                    uint32_t* __restrict__ corners,      // [frames][corners_per_frame]
                                                    int* __restrict__ corner_count,      // [frames][CCM_MAX_LEVELS]
                                                    const FastSmem M, int frame0,        // arena slot of frame 0 (TMA z coordinate)
                                                    const __grid_constant__ FastTma tm) {
    extern __shared__ __align__(128) uint8_t fast_smem[];
    uint8_t* const tile = fast_smem;                                   // [ch + 6][TS]
    uint8_t* const score = fast_smem + M.score_off;                    // [cell + 1-px zero ring][SS]
    const int SS = M.ss;
    uint16_t* const cand = reinterpret_cast<uint16_t*>(fast_smem + M.cand_off);
    uint16_t* const corn = reinterpret_cast<uint16_t*>(fast_smem + M.corn_off);
    __shared__ int n_cand, n_corn;
    __shared__ __align__(8) unsigned long long tile_bar;

    const int f = blockIdx.y;
    // detection block of this cell in level coordinates and the index-arithmetic constants of its tile
    // (host-built table: configure_shape, orb_api.cu)
    const uint4 ci = __ldg(cell_info + 2 * blockIdx.x), cj = __ldg(cell_info + 2 * blockIdx.x + 1);
    const int x0 = (int)(ci.x & 0xffff), y0 = (int)(ci.x >> 16), cw = (int)(ci.y & 0xffff), ch = (int)(ci.y >> 16), lvl = (int)ci.z;
    pdl_launch();
    pdl_wait();                              // the pyramid (previous kernels of the chain) is complete
    if (cw <= 0 || ch <= 0) return;          // clipped-away cell (the :962/:971 `continue`s)
    const LevelDesc& L = T.l[lvl];

    const uint8_t* __restrict__ img = L.img + (size_t)f * L.img_stride;
    const int tid = threadIdx.x;
    // (A) tile = cell + 3-px ring. TMA: one thread asks for the box whose corner is (gxa, y0-3) of frame f — the innermost start
    //     coordinate of a tiled TMA load must be a multiple of 16 bytes (measured: tools/tma_probe.cu; x = 13 raises an illegal
    //     instruction), so gxa = (x0-3) & ~15 and pixel x0-3 is tile column `shift`; otherwise aligned 32-bit words (rows are
    //     32-B aligned, pitch a multiple of 32).
    const int shift = (int)(ci.w & 0xff), wpr = (int)(ci.w >> 8) & 0xff;    // x0-3 = aligned gxa + shift (word-aligned: 0..3, TMA: 16-byte aligned, 0..15); words per tile row
    const int g0 = (int)(ci.w >> 16) & 0xff, ngx = (int)(ci.w >> 24);       // aligned 4-px words covering the cell: [g0, g0 + ngx)
    const uint32_t wmagic = cj.x, cmagic = cj.y, gmagic = cj.z;             // i / wpr, i / cw, i / ngx for i < 4096
    const int gxa = x0 - 3 - shift, th = ch + 6;
    if (TMA) {
        if (tid == 0) {
            const uint32_t bar_s = smem_addr(&tile_bar);
            mbar_init(bar_s, 1);
            mbar_expect_tx(bar_s, (uint32_t)(M.tile_rows * TS));
            tma_load_3d(smem_addr(tile), &tm.m[lvl], gxa, y0 - 3, frame0 + f, bar_s);
        }
    } else {
        // NT / wpr tile rows per sweep, one word per thread, pointers stepped by whole sweeps
        const int rps = (int)((NT * wmagic) >> 20);
        const int ty0 = (int)(((uint32_t)tid * wmagic) >> 20), tx = tid - ty0 * wpr;
        if (ty0 < rps) {
            const uint8_t* src = img + (size_t)(y0 - 3 + ty0) * L.img_pitch + gxa + 4 * tx;
            uint8_t* dst = tile + ty0 * TS + 4 * tx;
            const size_t sstep = (size_t)rps * L.img_pitch;
            const int dstep = rps * TS;
            for (int ty = ty0; ty < th; ty += rps, src += sstep, dst += dstep)
                *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(src);
        }
    }
    for (int i = tid; i < ((ch + 2) * SS + 15) / 16; i += NT) reinterpret_cast<uint4*>(score)[i] = make_uint4(0, 0, 0, 0);
    if (tid == 0) { n_cand = 0; n_corn = 0; }
    __syncthreads();
    if (TMA) mbar_wait(smem_addr(&tile_bar), 0);     // every thread observes the completed transaction: the tile is visible
    const uint32_t cand_s = smem_addr(cand), corn_s = smem_addr(corn), ncand_s = smem_addr(&n_cand), ncorn_s = smem_addr(&n_corn);

    const int S = TS;
    const int ngrp = ngx * ch;
    int threshold = T.ini_th;
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        // (B) opposite-pair test on 4 pixels per thread (a 9-arc of the ring contains one pixel of every diametral pair,
        //     so |v-p0|<=t && |v-p8|<=t, or the same for p4/p12, rules a pixel out): aligned 32-bit tile words, byte SIMD
        const uint32_t t4 = (uint32_t)threshold * 0x01010101u;
        for (int base = 0; base < ngrp; base += NT) {
            const int gi = base + tid;
            uint32_t nib = 0;
            int py = 0, pxb = 0;
            if (gi < ngrp) {
                py = (int)(((uint32_t)gi * gmagic) >> 20);
                const int g = g0 + gi - py * ngx;
                pxb = 4 * g - 3 - shift;                                      // pixel of byte 0 of this word
                const uint32_t* rowc = reinterpret_cast<const uint32_t*>(tile + (py + 3) * TS) + g;
                const uint32_t cen = rowc[0], lft = rowc[-1], rgt = rowc[1];
                const uint32_t up = rowc[-3 * (TS / 4)], dn = rowc[3 * (TS / 4)];
                const uint32_t p4 = __funnelshift_r(cen, rgt, 24), p12 = __funnelshift_r(lft, cen, 8);
                const uint32_t m = (__vcmpgtu4(__vabsdiffu4(cen, dn), t4) | __vcmpgtu4(__vabsdiffu4(cen, up), t4)) &
                                   (__vcmpgtu4(__vabsdiffu4(cen, p4), t4) | __vcmpgtu4(__vabsdiffu4(cen, p12), t4));
                nib = (((m & 0x01010101u) * 0x01020408u) >> 24) & 15u;
                // keep the bytes that are pixels of this cell: 0 <= pxb + k < cw
                const uint32_t lo_ok = pxb >= 0 ? 15u : (15u << (-pxb)) & 15u;
                const uint32_t hi_ok = pxb + 3 < cw ? 15u : 15u >> (pxb + 4 - cw);
                nib &= lo_ok & hi_ok;
            }
            const uint32_t b0 = __ballot_sync(0xffffffffu, nib & 1u), b1 = __ballot_sync(0xffffffffu, nib & 2u);
            const uint32_t b2 = __ballot_sync(0xffffffffu, nib & 4u), b3 = __ballot_sync(0xffffffffu, nib & 8u);
            if (b0 | b1 | b2 | b3) {
                const uint32_t lt = lanemask_lt();
                const uint32_t c0 = __popc(b0), c1 = __popc(b1), c2 = __popc(b2), c3 = __popc(b3);
                uint32_t slot = 0;
                if ((tid & 31) == 0) slot = atoms_add(ncand_s, c0 + c1 + c2 + c3);
                slot = __shfl_sync(0xffffffffu, slot, 0);
                const uint32_t a = cand_s + 2 * slot;
                const uint32_t pi = (uint32_t)(py * cw + pxb);
                sts_u16_if(nib & 1u, a + 2 * __popc(b0 & lt), pi);
                sts_u16_if(nib & 2u, a + 2 * (c0 + __popc(b1 & lt)), pi + 1);
                sts_u16_if(nib & 4u, a + 2 * (c0 + c1 + __popc(b2 & lt)), pi + 2);
                sts_u16_if(nib & 8u, a + 2 * (c0 + c1 + c2 + __popc(b3 & lt)), pi + 3);
            }
        }
        __syncthreads();
        // (C) FAST-9/16 on the candidates: the corner score decides (score >= threshold) and goes into the score map
        const int nc = n_cand;
        for (int base = 0; base < nc; base += NT) {
            const int j = base + tid;
            bool is_corner = false;
            int i = 0;
            if (j < nc) {
                i = cand[j];
                const int py = (int)(((uint32_t)i * cmagic) >> 20), px = i - py * cw;
                const uint8_t* c = tile + (py + 3) * TS + px + 3 + shift;
                uint32_t r[16];
                r[0] = c[3 * S];       r[1] = c[3 * S + 1];   r[2] = c[2 * S + 2];   r[3] = c[S + 3];
                r[4] = c[3];           r[5] = c[-S + 3];      r[6] = c[-2 * S + 2];  r[7] = c[-3 * S + 1];
                r[8] = c[-3 * S];      r[9] = c[-3 * S - 1];  r[10] = c[-2 * S - 2]; r[11] = c[-S - 3];
                r[12] = c[-3];         r[13] = c[S - 3];      r[14] = c[2 * S - 2];  r[15] = c[3 * S - 1];
                const int sc = fast_score16((int)c[0], r);
                is_corner = sc >= threshold;                  // threshold >= 1 (enforced at create): a corner's score is never 0
                if (is_corner) score[(py + 1) * SS + px + 1] = (uint8_t)sc;
            }
            list_append(is_corner, (uint32_t)i, corn_s, ncorn_s);
        }
        __syncthreads();
        const int nk = n_corn;
        // (E) non-maximum suppression over the corners, append survivors
        int found = 0;
        for (int base = 0; base < nk; base += NT) {     // uniform trip count: ballots below need whole warps
            const int j = base + tid;
            bool keep = false;
            int px = 0, py = 0, s = 0;
            if (j < nk) {
                const int i = corn[j];
                py = (int)(((uint32_t)i * cmagic) >> 20); px = i - py * cw;
                const uint8_t* sc = &score[(py + 1) * SS + px + 1];
                s = sc[0];
                const uint8_t *su = sc - SS, *sd = sc + SS;
                keep = s > su[-1] && s > su[0] && s > su[1] && s > sc[-1] && s > sc[1] && s > sd[-1] && s > sd[0] && s > sd[1];
            }
            const uint32_t ballot = __ballot_sync(0xffffffffu, keep);
            if (ballot) {
                const int lane = tid & 31;
                int slot = 0;
                if (lane == 0) slot = atomicAdd(&corner_count[f * CCM_MAX_LEVELS + lvl], __popc(ballot));
                slot = __shfl_sync(0xffffffffu, slot, 0);
                if (keep) {
                    const int idx = slot + __popc(ballot & ((1u << lane) - 1));
                    if (idx < L.corner_cap)
                        corners[(size_t)f * T.corners_per_frame + L.corner_off + idx] =
                            (uint32_t)(x0 + px) | ((uint32_t)(y0 + py) << 12) | ((uint32_t)s << 24);
                }
                found = 1;
            }
        }
        if (__syncthreads_or(found) || threshold == T.min_th) break;
        threshold = T.min_th;     // cell empty at iniThFAST -> rerun at minThFAST (:981-985)
        if (tid == 0) { n_cand = 0; n_corn = 0; }
        __syncthreads();
    }
}
