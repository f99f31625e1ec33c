// K1 (pyramid level) and K2+K3 (per-cell FAST-9/16 + NMS + threshold fallback + compaction), sm_100a.
// Integer/byte work bounded by HBM/L2 bandwidth and the ALU pipe; no tensor cores by design.
#include "ccm_common.cuh"

// ---------------------------------------------------------------------------------------------
// K1: one pyramid level from the previous one. Replaces cv::resize(INTER_LINEAR) at
// cslam/src/ORBextractor.cpp:1293 (ComputePyramid :1280-1304). Bit-exact 8U fixed-point bilinear:
// horizontal taps scaled by 2048 into int32, vertical (((b0*(r0>>4))>>16)+((b1*(r1>>4))>>16)+2)>>2.
// Each thread produces 4 horizontally adjacent pixels and stores one 32-bit word; a warp writes 128 B.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_resize_level(const __grid_constant__ LevelTable T, int lvl) {
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

// cornerScore<16> (OpenCV fast_score.cpp): max over the 16 arcs of 9 ring pixels of min(d) and of min(-d), minus 1.
// Written with min-only arc reductions over d and over nd = -d (materialised through an opaque asm so the negation
// cannot be folded into a 3-input min/max): ptxas 12.9 -O1+ for sm_100a miscompiles max(max(x,-y),max(z,-w))
// (VIMNMX3 with negated operands; reproduced on B200, see DESIGN.md "toolchain notes").
__device__ __forceinline__ int fast_score16(const int (&d)[16]) {
    int nd[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { nd[k] = -d[k]; asm volatile("" : "+r"(nd[k])); }
    int a2[16], b2[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { a2[k] = min(d[k], d[(k + 1) & 15]); b2[k] = min(nd[k], nd[(k + 1) & 15]); }
    int a4[16], b4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { a4[k] = min(a2[k], a2[(k + 2) & 15]); b4[k] = min(b2[k], b2[(k + 2) & 15]); }
    int best = -1000;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const int a9 = min(min(a4[k], a4[(k + 4) & 15]), d[(k + 8) & 15]);
        const int b9 = min(min(b4[k], b4[(k + 4) & 15]), nd[(k + 8) & 15]);
        best = max(best, max(a9, b9));
    }
    return best - 1;
}

__global__ void __launch_bounds__(256) k_fast_cells(const __grid_constant__ LevelTable T,
                                                    uint32_t* __restrict__ corners,      // [frames][corners_per_frame]
                                                    int* __restrict__ corner_count) {    // [frames][CCM_MAX_LEVELS]
    __shared__ uint8_t tile[FAST_TILE][FAST_TILE + 2];
    __shared__ uint8_t score[FAST_MAX_CELL + 2][FAST_MAX_CELL + 4];   // 1-px zero ring around the cell

    const int f = blockIdx.y;
    int lvl = 0;
#pragma unroll 1
    while (lvl + 1 < T.nlevels && (int)blockIdx.x >= T.l[lvl + 1].cell_base) lvl++;
    const LevelDesc& L = T.l[lvl];
    const int cell = blockIdx.x - L.cell_base;
    const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
    // detection block of this cell in level coordinates
    const int x0 = CCM_EDGE + cj * L.wCell, y0 = CCM_EDGE + ci * L.hCell;
    const int x1 = min(x0 + L.wCell, L.w - CCM_EDGE), y1 = min(y0 + L.hCell, L.h - CCM_EDGE);
    const int cw = x1 - x0, ch = y1 - y0;
    if (cw <= 0 || ch <= 0) return;          // clipped-away cell (the :962/:971 `continue`s)

    const uint8_t* __restrict__ img = L.img + (size_t)f * L.img_stride;
    const int tid = threadIdx.x;
    const int tw = cw + 6, th = ch + 6;
    for (int i = tid; i < tw * th; i += 256) {
        const int ty = i / tw, tx = i - ty * tw;
        tile[ty][tx] = img[(size_t)(y0 - 3 + ty) * L.img_pitch + (x0 - 3 + tx)];
    }
    for (int i = tid; i < (FAST_MAX_CELL + 2) * (FAST_MAX_CELL + 4); i += 256) (&score[0][0])[i] = 0;
    __syncthreads();

    int threshold = T.ini_th;
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        for (int i = tid; i < cw * ch; i += 256) {
            const int py = i / cw, px = i - py * cw;
            const uint8_t* c = &tile[py + 3][px + 3];
            const int v = c[0];
            const int S = FAST_TILE + 2;
            int d[16];
            d[0] = v - c[3 * S];       d[1] = v - c[3 * S + 1];   d[2] = v - c[2 * S + 2];   d[3] = v - c[S + 3];
            d[4] = v - c[3];           d[5] = v - c[-S + 3];      d[6] = v - c[-2 * S + 2];  d[7] = v - c[-3 * S + 1];
            d[8] = v - c[-3 * S];      d[9] = v - c[-3 * S - 1];  d[10] = v - c[-2 * S - 2]; d[11] = v - c[-S - 3];
            d[12] = v - c[-3];         d[13] = v - c[S - 3];      d[14] = v - c[2 * S - 2];  d[15] = v - c[3 * S - 1];
            uint32_t dark = 0, bright = 0;      // ring darker than v-t  /  brighter than v+t
#pragma unroll
            for (int k = 0; k < 16; k++) { dark |= (uint32_t)(d[k] > threshold) << k; bright |= (uint32_t)(d[k] < -threshold) << k; }
            int s = 0;
            if (has_run9(dark) || has_run9(bright)) s = fast_score16(d);
            score[py + 1][px + 1] = (uint8_t)s;   // s >= threshold > 0 for corners (threshold >= 1 enforced at create)
        }
        __syncthreads();
        int found = 0;
        for (int base = 0; base < cw * ch; base += 256) {     // uniform trip count: ballots below need whole warps
            const int i = base + tid;
            bool keep = false;
            int px = 0, py = 0, s = 0;
            if (i < cw * ch) {
                py = i / cw; px = i - py * cw;
                s = score[py + 1][px + 1];
                if (s > 0) {
                    keep = s > score[py][px] && s > score[py][px + 1] && s > score[py][px + 2] &&
                           s > score[py + 1][px] && s > score[py + 1][px + 2] &&
                           s > score[py + 2][px] && s > score[py + 2][px + 1] && s > score[py + 2][px + 2];
                }
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
    }
}
