// K6 (7x7 Gaussian blur), K5 (intensity-centroid orientation) + K7 (steered rBRIEF-256), sm_100a.
#include "ccm_common.cuh"

// ---------------------------------------------------------------------------------------------
// K6: cv::GaussianBlur(level.clone(), 7x7, sigma 2, BORDER_REFLECT_101), cslam/src/ORBextractor.cpp:1258-1259.
// OpenCV 4.x CV_8U fixed-point path: 8.8 kernel {18,34,48,56,48,34,18}/256, exact integer row and column
// accumulation, single rounding (sum + 2^15) >> 16.
// All levels of all frames in ONE launch. A warp owns a 120-px x BLUR_RPW-row strip (128-px windows overlapping by one word): each lane loads one aligned 32-bit
// word per source row (a warp reads 128 contiguous bytes), takes its left/right neighbour words by shuffle, forms the
// four horizontal sums of its 4 pixels, and keeps the last seven row results in registers as a sliding window for the
// vertical pass: no shared memory, every source byte is loaded once per strip (BLUR_RPW + 6 rows for BLUR_RPW outputs), the output
// is one coalesced 32-bit store per lane. Reflect-101 borders take a per-byte path only in the lanes that touch them.
// ---------------------------------------------------------------------------------------------
#ifndef BLUR_OVERLAP
#define BLUR_OVERLAP 1                  // warps overlap by one word on each side: lanes 0 / 31 only feed their neighbours, no edge loads
#endif
#define BLUR_TW (BLUR_OVERLAP ? 120 : 128)   // output pixels per warp row
#ifndef BLUR_PF
#define BLUR_PF 3                        // rows of prefetch (divides 6): 0.0903 / 0.0811 / 0.0804 ms per 64 frames at 1 / 2 / 3
#endif
#ifndef BLUR_ROLLED
#define BLUR_ROLLED 1
#endif
#ifndef BLUR_RPW
#define BLUR_RPW 30                      // output rows per warp (a multiple of 6 in the rolled form)
#endif
#define BLUR_TH (8 * BLUR_RPW)           // rows per CTA (8 warps)

// own word of this lane for one source row (0 beyond the row); issued one row ahead of its use
__device__ __forceinline__ uint32_t blur_load(const uint8_t* __restrict__ row, int x, int w) {
    return (unsigned)x < (unsigned)w ? *reinterpret_cast<const uint32_t*>(row + x) : 0u;   // x = -4 for lane 0 of the first tile (overlap mode)
}

// four horizontal 7-tap sums of pixels x..x+3 from the words holding pixels x-4..x+7: eight unaligned 4-byte windows
// (PRMT) and eight IDP4A with the packed 8.8 coefficients {18,34,48,56 | 48,34,18,0}.
// Right border (reflect-101): the taps reach at most pixels w..w+2 = pixels w-2, w-3, w-4. With wa = (w-1) & ~3 the row's
// last aligned word and e = w - wa its valid pixels, the 8 bytes "pixels wa..w-1, then w-2, w-3, w-4" are a byte permutation
// of the two words at wa-4 and wa, which the two lanes concerned already hold: two PRMT with selectors that depend only
// on e (blur_tail_selectors), no extra loads. The lane that owns word wa takes (own, next) from them, the lane before
// takes next. Levels narrower than 16 px never get here (blur_small_level).
__device__ __forceinline__ void blur_tail_selectors(int w, uint32_t& sel_lo, uint32_t& sel_hi) {
    const int e = w - ((w - 1) & ~3);                       // 1..4
    // byte j of the tail is pixel wa + j for j < e (index 4 + j of the word pair) and pixel w - 2 - (j - e) after it (index 2e + 2 - j)
    sel_lo = (uint32_t)(0x7654565434541234ull >> (16 * (e - 1))) & 0xffffu;     // e = 1: 0x1234, 2: 0x3454, 3: 0x5654, 4: 0x7654
    sel_hi = (uint32_t)(0x3456123400120000ull >> (16 * (e - 1))) & 0xffffu;     // e = 1: 0x0000, 2: 0x0012, 3: 0x1234, 4: 0x3456
}

// Tap sets. OCV24 = false: OpenCV >= 3.4.1 / 4.x fixed-point GaussianBlur, 8.8 kernel {18,34,48,56,48,34,18} (sums to 256),
// pinned to cv2 4.13.0 golden vectors. OCV24 = true: the integer path OpenCV 2.4.13's FilterEngine takes for CV_8U with a
// symmetric smoothing kernel (createSeparableLinearFilter: getGaussianKernel(7, 2) in float, each tap times 256 and rounded on
// its own -> {18,34,49,55,49,34,18}, which sums to 257; same exact integer passes, (sum + 2^15) >> 16, saturated to 255).
// The reference's CMakeLists pins 2.4.13; its source is a missing blob, so that variant is a restatement from knowledge of
// the 2.4 sources and cannot be checked against a build here. Selected with CCM_ORB_BLUR=opencv2.
template <bool OCV24> struct BlurTaps {
    static constexpr uint32_t k0 = 18u, k1 = 34u, k2 = OCV24 ? 49u : 48u, k3 = OCV24 ? 55u : 56u;
};

template <bool OCV24>
__device__ __forceinline__ void blur_hrow(const uint8_t* __restrict__ row, uint32_t own, int x, int x0, int w, int lane,
                                          uint32_t sel_lo, uint32_t sel_hi, uint32_t (&hs)[4]) {
    typedef BlurTaps<OCV24> K;
    uint32_t prev = __shfl_up_sync(0xffffffffu, own, 1);
    uint32_t next = __shfl_down_sync(0xffffffffu, own, 1);
#if !BLUR_OVERLAP
    if (lane == 0 && x0 > 0) prev = *reinterpret_cast<const uint32_t*>(row + x - 4);
    if (lane == 31 && x + 4 < w) next = *reinterpret_cast<const uint32_t*>(row + x + 4);
#endif
    const int wa = (w - 1) & ~3;
    if (x == 0) prev = __byte_perm(own, 0u, 0x1230);       // reflect-101 on the left: pixels -3,-2,-1 are pixels 3,2,1 of this word
    if (x == wa) {                                          // the two lanes whose windows reach past the last pixel
        const uint32_t lo = __byte_perm(prev, own, sel_lo), hi = __byte_perm(prev, own, sel_hi);
        own = lo; next = hi;
    } else if (x == wa - 4) next = __byte_perm(own, next, sel_lo);
    const uint32_t C1 = K::k0 | (K::k1 << 8) | (K::k2 << 16) | (K::k3 << 24), C2 = K::k2 | (K::k1 << 8) | (K::k0 << 16);
    const uint32_t wm3 = __byte_perm(prev, own, 0x4321), wm2 = __byte_perm(prev, own, 0x5432), wm1 = __byte_perm(prev, own, 0x6543);
    const uint32_t wp1 = __byte_perm(own, next, 0x4321), wp2 = __byte_perm(own, next, 0x5432), wp3 = __byte_perm(own, next, 0x6543);
    hs[0] = __dp4a(wm3, C1, __dp4a(wp1, C2, 0u));     // <= 255 * 257 = 65535
    hs[1] = __dp4a(wm2, C1, __dp4a(wp2, C2, 0u));
    hs[2] = __dp4a(wm1, C1, __dp4a(wp3, C2, 0u));
    hs[3] = __dp4a(own, C1, __dp4a(next, C2, 0u));
}

// Levels narrower than 16 px (the last levels of very small images): the word tricks of blur_hrow do not apply, and such a
// level is a few hundred pixels - each lane evaluates its 4 outputs tap by tap with reflect-101 (same integers: the
// horizontal sums are exact, the vertical pass rounds once). Out of line so that the unrolled main loop stays small.
template <bool OCV24>
__device__ __noinline__ void blur_small_level(const uint8_t* __restrict__ img, int pitch, uint8_t* __restrict__ out, int opitch,
                                              int w, int h, int x, int wy0) {
    typedef BlurTaps<OCV24> K;
    const int cf[7] = {(int)K::k0, (int)K::k1, (int)K::k2, (int)K::k3, (int)K::k2, (int)K::k1, (int)K::k0};
#pragma unroll 1
    for (int oy = wy0; oy < min(wy0 + BLUR_RPW, h); oy++) {
        if (x >= w) continue;
        uint32_t packed = 0;
#pragma unroll 1
        for (int i = 0; i < 4; i++) {
            const int px = min(x + i, w - 1);
            uint32_t acc = 0;
#pragma unroll 1
            for (int dy = -3; dy <= 3; dy++) {
                const uint8_t* row = img + (size_t)ccm_reflect101(oy + dy, h) * pitch;
                uint32_t hs = 0;
#pragma unroll 1
                for (int dx = -3; dx <= 3; dx++) hs += (uint32_t)cf[dx + 3] * row[ccm_reflect101(px + dx, w)];
                acc += (uint32_t)cf[dy + 3] * hs;
            }
            packed |= min((acc + 32768u) >> 16, 255u) << (8 * i);
        }
        *reinterpret_cast<uint32_t*>(out + (size_t)oy * opitch + x) = packed;
    }
}

template <bool OCV24>
__global__ void __launch_bounds__(256, 5) k_blur_levels(const __grid_constant__ LevelTable T) {
    typedef BlurTaps<OCV24> K;
    int lvl = 0;
#pragma unroll 1
    while (lvl + 1 < T.nlevels && (int)blockIdx.x >= T.l[lvl + 1].blur_tile_base) lvl++;
    const LevelDesc& L = T.l[lvl];
    const int tile = blockIdx.x - L.blur_tile_base;
    const int tyi = tile / L.blur_tiles_x, txi = tile - tyi * L.blur_tiles_x;
    const int f = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = txi * BLUR_TW - (BLUR_OVERLAP ? 4 : 0), x = x0 + 4 * lane;
    const bool out_lane = !BLUR_OVERLAP || (lane >= 1 && lane <= 30);
    const int wy0 = tyi * BLUR_TH + warp * BLUR_RPW;
    const int w = L.w, h = L.h;
    CCM_TS(blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0, 65);
    if (wy0 >= h) return;                                    // warp-uniform
    const uint8_t* __restrict__ img = L.img + (size_t)f * L.img_stride;
    const int pitch = L.img_pitch;
    uint8_t* __restrict__ outp = L.blur + (size_t)f * L.blur_stride + (size_t)wy0 * L.blur_pitch + x;
    if (w < 16) { blur_small_level<OCV24>(img, pitch, L.blur + (size_t)f * L.blur_stride, L.blur_pitch, w, h, out_lane ? x : w, wy0); return; }   // warp-uniform
    // Vertical pass on row PAIRS: pr[j][k] = hsum(row j-1) | hsum(row j) << 16 of pixel k (the sums are <= 65535), so that one
    // IDP2A applies two taps: out(c) = dp2a(pr[c-2], k0|k1) + dp2a(pr[c], k2|k3) + dp2a(pr[c+2], k2|k1) + dp2a(pr[c+3], 0|k0)
    // = 4 instructions per pixel (+ one PRMT to form the pair) instead of 3 adds + 4 multiply-adds.
    uint32_t pr[6][4];                 // pairs ending at the last six rows
    uint32_t last[4] = {0u, 0u, 0u, 0u};
    uint32_t sel_lo, sel_hi;
    blur_tail_selectors(w, sel_lo, sel_hi);
    // source row of window row r is wy0 + r - 3 (rows past the bottom only feed outputs that are never stored).
    // Strips that touch neither the top nor the bottom border step the row pointer; the others reflect per row.
    const bool inner = wy0 >= 3 && wy0 + BLUR_RPW + 3 <= h;   // warp-uniform
    const uint8_t* row = img + (size_t)ccm_reflect101(min(wy0 - 3, h + 2), h) * pitch;
#if BLUR_ROLLED
    // the lane's words of the next BLUR_PF rows are in flight while a row is worked on (q[t % BLUR_PF] = window row t)
    uint32_t q[BLUR_PF];
    q[0] = blur_load(row, x, w);
#pragma unroll
    for (int d = 1; d < BLUR_PF; d++) {
        row = inner ? row + pitch : img + (size_t)ccm_reflect101(min(wy0 + d - 3, h + 2), h) * pitch;
        q[d] = blur_load(row, x, w);
    }
#else
    uint32_t own = blur_load(row, x, w);
#endif
    const uint32_t W01 = K::k0 | (K::k1 << 8), W23 = K::k2 | (K::k3 << 8), W21 = K::k2 | (K::k1 << 8), W_0 = K::k0 << 8;
#if BLUR_ROLLED
    // rows in blocks of six (the pair ring's period): the first block only fills the ring, every later block is the same code
    // with r = 6 * blk + j, so the loop body is 6 rows long instead of BLUR_RPW + 6 (instruction cache) and BLUR_RPW can grow
    // (6 halo rows per BLUR_RPW outputs).
    static_assert(BLUR_RPW % 6 == 0, "BLUR_RPW must be a multiple of 6");
#pragma unroll 1
    for (int blk = 0; blk <= BLUR_RPW / 6; blk++) {
#pragma unroll
        for (int j = 0; j < 6; j++) {
            const int r = 6 * blk + j;
            const uint32_t own = q[j % BLUR_PF];
            row = inner ? row + pitch : img + (size_t)ccm_reflect101(min(wy0 + r + BLUR_PF - 3, h + 2), h) * pitch;   // window row r + BLUR_PF
            q[j % BLUR_PF] = (r + BLUR_PF < BLUR_RPW + 6) ? blur_load(row, x, w) : 0u;
            uint32_t hs[4];
            blur_hrow<OCV24>(row, own, x, x0, w, lane, sel_lo, sel_hi, hs);
#pragma unroll
            for (int k = 0; k < 4; k++) { pr[j][k] = __byte_perm(last[k], hs[k], 0x5410); last[k] = hs[k]; }   // (row r-1, row r)
            if (blk > 0) {
                const int oy = wy0 + r - 6;
                uint32_t packed = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t acc = __dp2a_lo(pr[(j + 1) % 6][k], W01, __dp2a_lo(pr[(j + 3) % 6][k], W23, __dp2a_lo(pr[(j + 5) % 6][k], W21,
                                                   __dp2a_lo(pr[j][k], W_0, 32768u))));
                    const uint32_t v = acc >> 16;
                    packed |= (OCV24 ? min(v, 255u) : v) << (8 * k);
                }
                if (out_lane && oy < h && x < w) *reinterpret_cast<uint32_t*>(outp) = packed;   // pitch multiple of 32: in-bounds
                outp += L.blur_pitch;
            }
        }
    }
#else
#pragma unroll
    for (int r = 0; r < BLUR_RPW + 6; r++) {
        const uint8_t* row_n = inner ? row + pitch : img + (size_t)ccm_reflect101(min(wy0 + r - 2, h + 2), h) * pitch;
        const uint32_t own_n = (r + 1 < BLUR_RPW + 6) ? blur_load(row_n, x, w) : 0u;
        uint32_t hs[4];
        blur_hrow<OCV24>(row, own, x, x0, w, lane, sel_lo, sel_hi, hs);
        row = row_n; own = own_n;
#pragma unroll
        for (int k = 0; k < 4; k++) { pr[r % 6][k] = __byte_perm(last[k], hs[k], 0x5410); last[k] = hs[k]; }   // (row r-1, row r)
        if (r >= 6) {
            const int oy = wy0 + r - 6;               // centre row c = r - 3: pairs ending at c-2, c, c+2, c+3 = r-5, r-3, r-1, r
            uint32_t packed = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t acc = __dp2a_lo(pr[(r - 5) % 6][k], W01, __dp2a_lo(pr[(r - 3) % 6][k], W23, __dp2a_lo(pr[(r - 1) % 6][k], W21,
                                               __dp2a_lo(pr[r % 6][k], W_0, 32768u))));
                const uint32_t v = acc >> 16;
                packed |= (OCV24 ? min(v, 255u) : v) << (8 * k);
            }
            if (out_lane && oy < h && x < w) *reinterpret_cast<uint32_t*>(outp) = packed;   // pitch multiple of 32: in-bounds
            outp += L.blur_pitch;
        }
    }
#endif
}

// ---------------------------------------------------------------------------------------------
// K5 + K7, one warp per selected keypoint.
//   IC_Angle (cslam/src/ORBextractor.cpp:68-95): integer moments over the 749-px disc of the UNBLURRED level,
//   then cv::fastAtan2 restated with separately rounded float ops (bit-exact).
//   computeOrbDescriptor (:100-316): 256 steered comparisons on the BLURRED level; lane l produces byte l.
//   cos/sin are evaluated in double and rounded to float (glibc cosf/sinf are not reproducible bit-for-bit on a
//   GPU; this differs from them by 1 ulp for ~1.3 % of angles, which flips a tap only at an exact .5 boundary —
//   the north-star tolerance >= 99.9 % identical descriptor bits covers it; measured in tests).
//   Output: ccm_keypoint with pt scaled by mvScaleFactor[level] for level != 0 (:1268-1274), descriptor as two
//   16-byte stores per keypoint.
// ---------------------------------------------------------------------------------------------
// byte at base + pitch * v (v a compile-time row offset) / at base + off (off a signed 32-bit byte offset), global memory
__device__ __forceinline__ uint32_t ldg_u8_row(const uint8_t* base, int pitch, int v) {
    uint32_t r;
    asm("{\n\t.reg .s64 a;\n\tmad.wide.s32 a, %2, %3, %1;\n\tld.global.u8 %0, [a];\n\t}" : "=r"(r) : "l"(base), "r"(pitch), "r"(v));
    return r;
}
__device__ __forceinline__ uint32_t ldg_u8_off(const uint8_t* base, uint32_t off) {
    uint32_t r;
    asm("{\n\t.reg .s64 a;\n\tmad.wide.s32 a, %2, 1, %1;\n\tld.global.u8 %0, [a];\n\t}" : "=r"(r) : "l"(base), "r"(off));
    return r;
}
#ifndef DESC_SINCOS
#define DESC_SINCOS 1
#endif
// rBRIEF test pattern (bit_pattern_31_, ORBextractor.cpp:320-577) as it sits in the source; the kernel reads the float4 table
// orb_pattern_table() builds from it once per handle: entry [k][lane] = (x1, y1, x2, y2) of test 8 * lane + k
static const signed char h_brief_pattern[256][4] = {
#include "brief_pattern.inc"
};
static void orb_pattern_table(float4 (&out)[256]) {
    for (int t = 0; t < 256; t++)
        out[(t & 7) * 32 + (t >> 3)] = make_float4((float)h_brief_pattern[t][0], (float)h_brief_pattern[t][1], (float)h_brief_pattern[t][2], (float)h_brief_pattern[t][3]);
}
// umax[v] of the 15-px orientation patch, ORBextractor.cpp:623-638: {15,15,15,15,14,14,14,13,13,12,11,10,9,8,6,3}
__host__ __device__ constexpr int umax_of(int v) {
    return v <= 3 ? 15 : v <= 6 ? 14 : v <= 8 ? 13 : v == 9 ? 12 : v == 10 ? 11 : v == 11 ? 10 : v == 12 ? 9 : v == 13 ? 8 : v == 14 ? 6 : 3;
}

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    // OpenCV atan2_p1..p7: float(coef) * float(180/pi), evaluated in float
    const float p1 = 0x1.ca44dep+5f, p3 = -0x1.2aaddcp+4f, p5 = 0x1.1d3f7ep+3f, p7 = -0x1.4515b2p+1f;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, 0x1p-52f));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, 0x1p-52f));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__global__ void __launch_bounds__(256) k_orient_describe(const __grid_constant__ LevelTable T,
                                                         const float4* __restrict__ pat_table,       // [8][32], see below (host-built)
                                                         const uint8_t* __restrict__ block_level,    // [total_desc_blocks] level of each CTA
                                                         const uint32_t* __restrict__ level_kps,     // [frames][kps_per_frame]
                                                         const int* __restrict__ level_kp_count,     // [frames][CCM_MAX_LEVELS]
                                                         ccm_keypoint_dev* __restrict__ out_kp,      // [frames][capacity]
                                                         uint8_t* __restrict__ out_desc,             // [frames][capacity][32]
                                                         int* __restrict__ out_count,                // [frames]
                                                         int capacity) {
    __shared__ float4 pat[8][32];     // pat[k][lane] = (x1, y1, x2, y2) of test 8*lane+k as floats: one LDS.128, no conversions in the loop
    __shared__ int s_cnt[CCM_MAX_LEVELS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int f = blockIdx.y;
    reinterpret_cast<float4*>(&pat[0][0])[tid] = __ldg(pat_table + tid);
    if (tid < CCM_MAX_LEVELS) s_cnt[tid] = tid < T.nlevels ? level_kp_count[f * CCM_MAX_LEVELS + tid] : 0;
    // A CTA's 8 slots belong to one level (host grid: desc_block_base), so the level, its image bases and pitches are
    // CTA-uniform: they live in uniform registers and every pixel load below is [uniform base + 32-bit offset].
    const int lvl = __ldg(block_level + blockIdx.x);          // host table (configure_shape): level of every CTA of the grid
    CCM_TS(blockIdx.x == 0 && blockIdx.y == 0 && tid == 0, 64);
    const LevelDesc& L = T.l[lvl];
    __syncthreads();
    // keypoints of the lower levels come first in the output (:1239-1276): lane l holds level l's count
    const int c = lane < CCM_MAX_LEVELS ? s_cnt[lane] : 0;
    const int total = __reduce_add_sync(0xffffffffu, c);
    const int before = __reduce_add_sync(0xffffffffu, lane < lvl ? c : 0);
    const int cnt_lvl = s_cnt[lvl];
    if (blockIdx.x == 0 && tid == 0) out_count[f] = total;
    const int pos = ((int)blockIdx.x - L.desc_block_base) * 8 + warp;
    if (pos >= cnt_lvl) return;
    const int oi = before + pos;
    if (oi >= capacity) return;
    const int slot = L.kp_off + pos;

    const uint32_t pk = level_kps[(size_t)f * T.kps_per_frame + slot];
    const int kx = (int)(pk & 4095u), ky = (int)((pk >> 12) & 4095u), score = (int)(pk >> 24);

    // --- orientation ---
    const uint8_t* __restrict__ imgf = L.img + (size_t)f * L.img_stride;      // uniform
    const int ipitch = L.img_pitch;
    // rows +v and -v share |u| <= umax[v] (:80-92); fully unrolled so that all 31 row loads of a lane are in flight at once
    int m10 = 0, m01 = 0;
    const int u = lane - CCM_HALF_PATCH;
    const int au = abs(u);
    const int o0 = ky * ipitch + kx + u;
    // one IMAD.WIDE forms the 64-bit address of row +-v (pitch * v + centre) and the load follows: two instructions per
    // pixel instead of the IMAD + IADD3 + IADD3.X + LDG ptxas builds from `imgf[o0 + v * ipitch]` (0.0830 -> 0.0795 ms)
    const uint8_t* pc = imgf + o0;
    if (lane < 31) m10 = u * (int)ldg_u8_row(pc, ipitch, 0);
#pragma unroll
    for (int v = 1; v <= CCM_HALF_PATCH; v++) {
        if (lane < 31 && au <= umax_of(v)) {
            const int vp = (int)ldg_u8_row(pc, ipitch, v), vm = (int)ldg_u8_row(pc, ipitch, -v);
            m10 += u * (vp + vm);
            m01 += v * (vp - vm);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { m10 += __shfl_xor_sync(0xffffffffu, m10, o); m01 += __shfl_xor_sync(0xffffffffu, m01, o); }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // --- descriptor ---
    const float rad = __fmul_rn(angle, 0x1.1df46ap-6f);      // factorPI = (float)(CV_PI/180.f)
#if DESC_SINCOS
    double sd, cd;
    sincos((double)rad, &sd, &cd);
    const float a = (float)cd, b = (float)sd;
#else
    const float a = (float)cos((double)rad), b = (float)sin((double)rad);
#endif
    const uint8_t* __restrict__ blf = L.blur + (size_t)f * L.blur_stride;     // uniform
    // cvRound of the steered tap (:107, :111-124) without the quarter-rate F2I: v + 1.5 * 2^23 rounds v to the nearest-even integer
    // in the low mantissa bits (|v| < 2^22), so int(v) = bits - 0x4B400000. The two constants of row * pitch + col are folded
    // into one (32-bit modular arithmetic: the true offset is small).
    const float MAGIC = 12582912.0f;
    const uint32_t pitch = (uint32_t)L.blur_pitch, koff = (uint32_t)ky * pitch + (uint32_t)kx - 0x4B400000u * (pitch + 1u);
    uint32_t byte = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float4 p = pat[k][lane];
        const uint32_t r1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)), MAGIC));
        const uint32_t c1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)), MAGIC));
        const uint32_t r2 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)), MAGIC));
        const uint32_t c2 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)), MAGIC));
        const int t0 = (int)ldg_u8_off(blf, r1 * pitch + c1 + koff), t1 = (int)ldg_u8_off(blf, r2 * pitch + c2 + koff);
        byte |= (uint32_t)(t0 < t1) << k;
    }
    uint32_t w = byte << (8 * (lane & 3));
    w |= __shfl_xor_sync(0xffffffffu, w, 1);
    w |= __shfl_xor_sync(0xffffffffu, w, 2);                  // every lane of a quad now holds descriptor word lane/4
    const int half = lane & 1;
    uint4 q;
    q.x = __shfl_sync(0xffffffffu, w, 16 * half + 0);
    q.y = __shfl_sync(0xffffffffu, w, 16 * half + 4);
    q.z = __shfl_sync(0xffffffffu, w, 16 * half + 8);
    q.w = __shfl_sync(0xffffffffu, w, 16 * half + 12);
    uint8_t* d = out_desc + ((size_t)f * capacity + oi) * 32;
    if (lane < 2) *reinterpret_cast<uint4*>(d + 16 * lane) = q;
    if (lane == 0) {
        ccm_keypoint_dev kp;
        kp.x = (float)kx; kp.y = (float)ky;
        if (lvl != 0) { kp.x = __fmul_rn(kp.x, L.scale); kp.y = __fmul_rn(kp.y, L.scale); }
        kp.size = L.kp_size; kp.angle = angle; kp.response = (float)score; kp.octave = lvl;
        out_kp[(size_t)f * capacity + oi] = kp;
    }
}
