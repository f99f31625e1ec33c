// K2+K3, second form: FAST-9/16 + per-cell NMS + per-cell threshold fallback with TWO WARPS per chunk of whole cells.
// Replaces the cell loop of ComputeKeyPointsOctTree, cslam/src/ORBextractor.cpp:957-998 (cv::FAST(cell, iniThFAST, nms) and
// the minThFAST rerun when the cell came out empty, :978-985). Successor of the round-1 kernel
// (one CTA per cell: ~300 warp instructions per 128 pixels, five barriers per pass, a 4-word prefilter that recomputed
// every difference twice, four ballots per 128 pixels for the candidate list; 167 M warp instructions per 64 frames).
//
// A chunk = up to four horizontally adjacent cells of one cell row (<= 121 px wide): lane l owns the aligned 4-pixel word
// xw0 + 4l of every row of the chunk's tile (128 px + ring; one TMA box or one word-copy loop per CTA). The two warps of
// the CTA take the upper and the lower half of the rows; they share the tile and the score map and meet at two barriers
// per pass (scores complete -> NMS; survivors known -> fallback decision). Everything else is warp-private.
//  (B) prefilter, streamed down the rows. A FAST-9 arc holds one pixel of every diametral pair of the ring, so a corner at
//      threshold t needs (|v-up3|>t or |v-dn3|>t) and (|v-lf3|>t or |v-rt3|>t). The vertical difference V(x,y) =
//      |I(x,y) - I(x,y+3)| > t serves pixel (x,y) as "down" and pixel (x,y+3) as "up"; the horizontal H(x,y) =
//      |I(x,y) - I(x+3,y)| > t serves (x,y) and (x+3,y): each is computed ONCE per pixel (VABSDIFF4 + a 3-instruction
//      SWAR compare whose result is bit 7 of each byte), kept in a 3-row register ring (V) or passed to the right
//      neighbour by one shuffle (H). One LDS, two shuffles, ~20 instructions per 4 pixels x 32 lanes.
//  (C) candidate bits of 8 rows are gathered in one register per lane (byte = pixel, bit = row). Lanes whose register is
//      not zero queue it (one ballot); when 32 registers are queued every lane unpacks one of them into the candidate
//      list (u16 codes: row block, lane, bit), so the unpack loop runs with all lanes busy.
//  (D) as soon as 32 candidates are pending the warp scores them (fast_score16: segment test and cornerScore in one),
//      writes scores >= t into the chunk's score map and keeps the corners' codes at the front of the list; cells are
//      separated by a zero column in the score map, so
//  (E) the 3x3 NMS over the corners needs no cell-border logic (neighbours outside the cell's detection block count as 0
//      exactly as in a per-cell cv::FAST call); survivors are compacted in place and leave with ONE global atomic per
//      warp. Cells without a survivor rerun B-E at minThFAST with the valid mask restricted to their pixels.
// A warp whose candidates outgrow its list (noise, checkerboards) scores what is pending, and its NMS walks the score map
// instead of the list (8 rows at a time through the same list): slower, same result.
#include "ccm_common.cuh"

#define FC_TW 144                               // tile row stride = TMA box width: 128 px of lane words + up to 12 px of 16-byte alignment
#define FC_SS 128                               // score-map row stride: <= 121 px + 1 zero column per cell + 1
#define FC_LIST 1024                            // candidate list entries per warp (>= 32 queued registers x 32 bits)
#define FC_QUEUE 64                             // queued (candidate register, row block | lane) pairs per warp
#define FC_MAX_SPAN 123                         // last valid pixel <= xw0 + 123 (lane 31 only lends its word to lane 30)
#ifndef FC_WARPS
#define FC_WARPS 2
#endif

struct FastChunkSmem { int tile_rows, score_off, list_off, queue_off, total; };
static inline FastChunkSmem fast_chunk_smem(int max_ch) {
    FastChunkSmem m;
    m.tile_rows = max_ch + 6;
    m.score_off = (m.tile_rows * FC_TW + 127) / 128 * 128;
    m.list_off = m.score_off + (max_ch + 2) * FC_SS;
    m.queue_off = m.list_off + FC_WARPS * FC_LIST * 2;
    m.total = m.queue_off + FC_WARPS * FC_QUEUE * 8;
    return m;
}

__device__ __forceinline__ void sts_u16(uint32_t saddr, uint32_t val) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(saddr), "h"((unsigned short)val) : "memory");
}
__device__ __forceinline__ uint32_t bfind_u32(uint32_t v) { uint32_t p; asm("bfind.u32 %0, %1;" : "=r"(p) : "r"(v)); return p; }

// |a - b| > t per byte, result in bit 7 of each byte (other bits undefined). t < 128: d > t <=> (d & 0x7f) + (127 - t)
// carries into bit 7, or d >= 128; t >= 128: d > t <=> d >= 128 and (d & 0x7f) + (255 - t) carries.
template <bool HI> __device__ __forceinline__ uint32_t fc_gt(uint32_t a, uint32_t b, uint32_t k4) {
    const uint32_t d = __vabsdiffu4(a, b);
    const uint32_t s = (d & 0x7f7f7f7fu) + k4;
    return HI ? (s & d) : (s | d);
}

struct FcChunk {
    int lane, f, lvl;
    int x_c0, y0, wc, ch;                       // detection block of the chunk (level coordinates)
    int r0, r1;                                 // detection rows [r0, r1) of this warp
    int xw0, shift, xoff;                       // lane 0's word; its byte offset in a tile row; x_c0 - xw0
    uint32_t cmagic;                            // (i * cmagic) >> 20 == i / wCell
    const uint8_t* tile; uint8_t* score;
    uint16_t* list; uint2* queue;               // this warp's candidate list and register queue
    uint32_t list_s;
    int th;                                     // threshold of the current pass
    uint32_t vm;                                // 0x80 in the bytes of this lane's word that are pixels of the pass
    int qn;                                     // queued registers
    int total, scored, ncorn;                   // list: corners [0, ncorn), pending candidates [scored, total)
    bool drained;                               // the list no longer holds every corner of the pass
    uint32_t found;                             // bit j: cell j has an NMS survivor (this lane's view)
};

// code = row block << 10 | lane << 5 | (pixel-in-word << 3 | 7 - row-in-block)
__device__ __forceinline__ void fc_decode(const FcChunk& C, uint32_t code, int& tc, int& tr, int& xrel) {
    const int p = (int)(code & 31u), ln = (int)((code >> 5) & 31u), rb = (int)(code >> 10);
    const int k = p >> 3, j = 7 - (p & 7);
    tc = C.shift + 4 * ln + k;
    tr = C.r0 + 8 * rb + j;                     // detection row within the chunk
    xrel = 4 * ln + k - C.xoff;
}

// (D) score list entries [base, base + 32) (those < end); corners move to the front of the list
__device__ __forceinline__ void fc_score_batch(FcChunk& C, int base, int end) {
    const int j = base + C.lane;
    bool corner = false;
    uint32_t code = 0;
    if (j < end) {
        code = C.list[j];
        int tc, tr, xrel;
        fc_decode(C, code, tc, tr, xrel);
        const uint8_t* c = C.tile + (tr + 3) * FC_TW + tc;
        constexpr int S = FC_TW;
        uint32_t r[16];
        r[0] = c[3 * S];       r[1] = c[3 * S + 1];   r[2] = c[2 * S + 2];   r[3] = c[S + 3];
        r[4] = c[3];           r[5] = c[-S + 3];      r[6] = c[-2 * S + 2];  r[7] = c[-3 * S + 1];
        r[8] = c[-3 * S];      r[9] = c[-3 * S - 1];  r[10] = c[-2 * S - 2]; r[11] = c[-S - 3];
        r[12] = c[-3];         r[13] = c[S - 3];      r[14] = c[2 * S - 2];  r[15] = c[3 * S - 1];
        const int sc = fast_score16((int)c[0], r);
        corner = sc >= C.th;                    // th >= 1: a corner's score is never 0
        if (corner) {
            const int cj = (int)(((uint32_t)xrel * C.cmagic) >> 20);
            C.score[(tr + 1) * FC_SS + xrel + cj + 1] = (uint8_t)sc;
        }
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, corner);      // every lane has read its entry; ncorn <= base
    if (corner) C.list[C.ncorn + __popc(ballot & ((1u << C.lane) - 1u))] = (uint16_t)code;
    C.ncorn += __popc(ballot);
    __syncwarp();
}

// unpack one candidate register per lane (w, meta = row block << 5 | lane) into the list
__device__ __forceinline__ void fc_unpack(FcChunk& C, uint32_t w, uint32_t meta) {
    const uint32_t full = 0xffffffffu;
    const uint32_t cnt = __popc(w);
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(full, inc, d);
        if (C.lane >= d) inc += t;
    }
    const int tot = (int)__shfl_sync(full, inc, 31);
    if (C.total + tot > FC_LIST) {              // drain: score what is pending; the list restarts
        for (; C.scored < C.total; C.scored += 32) fc_score_batch(C, C.scored, C.total);
        C.total = 0; C.scored = 0; C.ncorn = 0; C.drained = true;
    }
    uint32_t addr = C.list_s + 2u * (uint32_t)(C.total + (int)(inc - cnt));
    const uint32_t code0 = meta << 5;
    while (w) {
        const uint32_t p = bfind_u32(w);
        w ^= 1u << p;
        sts_u16(addr, code0 + p);
        addr += 2;
    }
    C.total += tot;
    __syncwarp();
}

// (C) queue the candidate bits of row block rb; unpack and score whenever 32 registers / 32 candidates are available
__device__ __forceinline__ void fc_emit(FcChunk& C, uint32_t acc, int rb) {
    const uint32_t full = 0xffffffffu;
    const uint32_t m = __ballot_sync(full, acc != 0);
    if (!m) return;
    if (acc) C.queue[C.qn + __popc(m & ((1u << C.lane) - 1u))] = make_uint2(acc, ((uint32_t)rb << 5) | (uint32_t)C.lane);
    C.qn += __popc(m);
    __syncwarp();
    if (C.qn < 32) return;
    const uint2 e = C.queue[C.lane];
    const uint2 tail = C.queue[32 + C.lane];    // (garbage beyond qn - 32: not stored back)
    __syncwarp();
    if (C.lane < C.qn - 32) C.queue[C.lane] = tail;
    C.qn -= 32;
    fc_unpack(C, e.x, e.y);
    for (; C.total - C.scored >= 32; C.scored += 32) fc_score_batch(C, C.scored, C.total);
}

// (B) one pass of the streamed prefilter at threshold C.th over the pixels of C.vm, rows [r0, r1) of this warp
template <bool HI> __device__ __forceinline__ void fc_prefilter_pass(FcChunk& C) {
    const uint32_t full = 0xffffffffu;
    const uint32_t k4 = (uint32_t)(HI ? 255 - C.th : 127 - C.th) * 0x01010101u;
    constexpr int WS = FC_TW / 4;               // words per tile row
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(C.tile + C.shift) + C.lane + C.r0 * WS;
    // detection row r is tile row r + 3; V(q) relates tile rows q and q + 3
    uint32_t c0 = wp[3 * WS], c1 = wp[4 * WS], c2 = wp[5 * WS];
    uint32_t va = fc_gt<HI>(wp[0], c0, k4), vb = fc_gt<HI>(wp[WS], c1, k4), vc = fc_gt<HI>(wp[2 * WS], c2, k4);
    wp += 6 * WS;                               // next row to load: tile row 6 + r
    const uint32_t vm = C.vm;
    // one detection row: c0 = its pixels, n = the row three below, va = V of the row three above
#define FC_ROW(j)                                                                              \
    {                                                                                          \
        const uint32_t n = wp[(j) * WS];                                                       \
        const uint32_t vr = fc_gt<HI>(c0, n, k4);                                              \
        const uint32_t rgt = __shfl_down_sync(full, c0, 1);                                    \
        const uint32_t h = fc_gt<HI>(c0, __funnelshift_r(c0, rgt, 24), k4);                    \
        const uint32_t hl = __shfl_up_sync(full, h, 1);                                        \
        const uint32_t cand = (vr | va) & (h | __funnelshift_r(hl, h, 8)) & vm;                \
        acc |= cand >> (j);                                                                    \
        c0 = c1; c1 = c2; c2 = n; va = vb; vb = vc; vc = vr;                                   \
    }
    const int rows = C.r1 - C.r0, nfull = rows >> 3, rem = rows & 7;
    for (int rb = 0; rb < nfull; rb++) {
        uint32_t acc = 0;
        FC_ROW(0) FC_ROW(1) FC_ROW(2) FC_ROW(3) FC_ROW(4) FC_ROW(5) FC_ROW(6) FC_ROW(7)
        wp += 8 * WS;
        fc_emit(C, acc, rb);
    }
    if (rem) {
        uint32_t acc = 0;
#pragma unroll 1
        for (int j = 0; j < rem; j++) {
            uint32_t a1 = 0;
            { uint32_t acc = 0; FC_ROW(0) a1 = acc; }
            acc |= a1 >> j;
            wp += WS;
        }
        fc_emit(C, acc, nfull);
    }
#undef FC_ROW
    // what is left in the queue and in the list
    if (__any_sync(full, C.qn > 0)) {
        uint2 e = make_uint2(0u, 0u);
        if (C.lane < C.qn) e = C.queue[C.lane];
        C.qn = 0;
        __syncwarp();
        fc_unpack(C, e.x, e.y);
    }
    for (; C.scored < C.total; C.scored += 32) fc_score_batch(C, C.scored, C.total);
}

// (E) NMS over the corners list[0, n): survivors compacted in place, then appended to the (frame, level) corner list
__device__ __forceinline__ void fc_nms_output(FcChunk& C, int n, const LevelDesc& L, uint32_t* __restrict__ corners,
                                              int* __restrict__ corner_count, int corners_per_frame) {
    const uint32_t full = 0xffffffffu;
    int nsurv = 0;
    for (int base = 0; base < n; base += 32) {
        const int j = base + C.lane;
        bool keep = false;
        uint32_t code = 0;
        int cj = 0;
        if (j < n) {
            code = C.list[j];
            int tc, tr, xrel;
            fc_decode(C, code, tc, tr, xrel);
            cj = (int)(((uint32_t)xrel * C.cmagic) >> 20);
            const uint8_t* sc = C.score + (tr + 1) * FC_SS + xrel + cj + 1;
            const int s = sc[0];
            const int m0 = max(max((int)sc[-FC_SS - 1], (int)sc[-FC_SS]), max((int)sc[-FC_SS + 1], (int)sc[-1]));
            const int m1 = max(max((int)sc[1], (int)sc[FC_SS - 1]), max((int)sc[FC_SS], (int)sc[FC_SS + 1]));
            keep = s > max(m0, m1);
        }
        const uint32_t ballot = __ballot_sync(full, keep);      // every lane has read its entry: the writes below cannot overtake a read
        if (keep) {
            C.found |= 1u << cj;
            C.list[nsurv + __popc(ballot & ((1u << C.lane) - 1u))] = (uint16_t)code;
        }
        nsurv += __popc(ballot);
        __syncwarp();
    }
    if (nsurv == 0) return;
    int slot = 0;
    if (C.lane == 0) slot = atomicAdd(&corner_count[C.f * CCM_MAX_LEVELS + C.lvl], nsurv);
    slot = __shfl_sync(full, slot, 0);
    for (int i = C.lane; i < nsurv; i += 32) {
        int tc, tr, xrel;
        fc_decode(C, C.list[i], tc, tr, xrel);
        const int cj = (int)(((uint32_t)xrel * C.cmagic) >> 20);
        const uint32_t s = C.score[(tr + 1) * FC_SS + xrel + cj + 1];
        if (slot + i < L.corner_cap)
            corners[(size_t)C.f * corners_per_frame + L.corner_off + slot + i] = (uint32_t)(C.x_c0 + xrel) | ((uint32_t)(C.y0 + tr) << 12) | (s << 24);
    }
    __syncwarp();
}

// NMS of a drained pass: the corners of this warp's rows are re-read from the score map, 8 rows at a time through the list
__device__ __forceinline__ void fc_nms_from_score_map(FcChunk& C, const LevelDesc& L, uint32_t* __restrict__ corners,
                                                      int* __restrict__ corner_count, int corners_per_frame) {
    const int rows = C.r1 - C.r0, nblk = (rows + 7) >> 3;
    const int x0 = 4 * C.lane - C.xoff;                           // xrel of byte 0 of this lane's word
    for (int rb = 0; rb < nblk; rb++) {
        uint32_t acc = 0;
        for (int j = 0; j < 8 && 8 * rb + j < rows; j++)
            for (int k = 0; k < 4; k++)
                if (C.vm & (0x80u << (8 * k))) {
                    const int xrel = x0 + k, cj = (int)(((uint32_t)xrel * C.cmagic) >> 20);
                    if (C.score[(C.r0 + 8 * rb + j + 1) * FC_SS + xrel + cj + 1]) acc |= 1u << (8 * k + 7 - j);
                }
        C.total = 0; C.scored = 0; C.ncorn = 0;
        fc_unpack(C, acc, ((uint32_t)rb << 5) | (uint32_t)C.lane);     // <= 31 x 32 entries: never drains
        fc_nms_output(C, C.total, L, corners, corner_count, corners_per_frame);
    }
}

template <bool TMA>
__global__ void __launch_bounds__(32 * FC_WARPS) k_fast_chunks(const __grid_constant__ LevelTable T,
                                                    const uint4* __restrict__ items,     // per chunk 2 x uint4, see fast_chunk_table (orb_api.cu)
                                                    uint32_t* __restrict__ corners,      // [frames][corners_per_frame]
                                                    int* __restrict__ corner_count,      // [frames][CCM_MAX_LEVELS]
                                                    const FastChunkSmem M, int frame0,   // arena slot of frame 0 (TMA z coordinate)
                                                    const __grid_constant__ FastTma tm) {
    extern __shared__ __align__(128) uint8_t fc_smem[];
    __shared__ __align__(8) unsigned long long tile_bar;
    __shared__ uint32_t found_s;
    const int warp = threadIdx.x >> 5;

    FcChunk C;
    C.lane = threadIdx.x & 31; C.f = blockIdx.y;
    const uint4 a = __ldg(items + 2 * blockIdx.x), b = __ldg(items + 2 * blockIdx.x + 1);
    C.x_c0 = (int)(a.x & 0xffffu); C.y0 = (int)(a.x >> 16);
    C.wc = (int)(a.y & 0xffffu); C.ch = (int)(a.y >> 16);
    C.lvl = (int)(a.z & 0xffu);
    const int ncells = (int)((a.z >> 8) & 0xffu);
    C.xw0 = (int)a.w; C.cmagic = b.x;
    const int gxa = C.xw0 & ~15;
    C.shift = C.xw0 - gxa; C.xoff = C.x_c0 - C.xw0;
    const int half = (C.ch + FC_WARPS - 1) / FC_WARPS;
    C.r0 = min(C.ch, warp * half); C.r1 = min(C.ch, C.r0 + half);
    C.tile = fc_smem; C.score = fc_smem + M.score_off;
    C.list = reinterpret_cast<uint16_t*>(fc_smem + M.list_off) + warp * FC_LIST;
    C.queue = reinterpret_cast<uint2*>(fc_smem + M.queue_off) + warp * FC_QUEUE;
    C.list_s = smem_addr(C.list);
    pdl_launch();
    pdl_wait();                                 // the pyramid (previous kernels of the chain) is complete
    CCM_TS(blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0, 16 + C.lvl);
    const LevelDesc& L = T.l[C.lvl];

    // (A) tile rows 0 .. ch+5 = image rows y0-3 .. y0+ch+2, columns gxa .. gxa+143
    if (TMA) {
        if (threadIdx.x == 0) {
            const uint32_t bar_s = smem_addr(&tile_bar);
            mbar_init(bar_s, 1);
            mbar_expect_tx(bar_s, (uint32_t)(M.tile_rows * FC_TW));
            tma_load_3d(smem_addr(fc_smem), &tm.m[C.lvl], gxa, C.y0 - 3, frame0 + C.f, bar_s);
        }
    } else {
        const uint8_t* img = L.img + (size_t)C.f * L.img_stride + (size_t)(C.y0 - 3) * L.img_pitch + gxa;
        const int nw = min(FC_TW / 4, (L.img_pitch - gxa) >> 2);      // whole words of the row (pitch is a multiple of 4)
        for (int i = threadIdx.x; i < (C.ch + 6) * (FC_TW / 4); i += 32 * FC_WARPS) {
            const int r = i / (FC_TW / 4), w = i - r * (FC_TW / 4);
            reinterpret_cast<uint32_t*>(fc_smem)[i] = w < nw ? reinterpret_cast<const uint32_t*>(img + (size_t)r * L.img_pitch)[w] : 0u;
        }
    }
    for (int i = threadIdx.x; i < ((C.ch + 2) * FC_SS) / 16; i += 32 * FC_WARPS) reinterpret_cast<uint4*>(C.score)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) found_s = 0;
    __syncthreads();
    if (TMA) mbar_wait(smem_addr(&tile_bar), 0);

    // pixels of this lane's word that belong to the chunk
    uint32_t vm_all = 0;
    const int xr0 = 4 * C.lane - C.xoff;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (xr0 + k >= 0 && xr0 + k < C.wc) vm_all |= 0x80u << (8 * k);
    const uint32_t cells_all = (1u << ncells) - 1u;

    C.th = T.ini_th; C.vm = vm_all; C.found = 0;
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        C.total = 0; C.scored = 0; C.ncorn = 0; C.qn = 0; C.drained = false;
        if (C.th >= 128) fc_prefilter_pass<true>(C);
        else fc_prefilter_pass<false>(C);
        __syncthreads();                        // every score of the pass is in the map
        if (C.drained) fc_nms_from_score_map(C, L, corners, corner_count, T.corners_per_frame);
        else fc_nms_output(C, C.ncorn, L, corners, corner_count, T.corners_per_frame);
        if (C.th == T.min_th) break;
        const uint32_t fw = __reduce_or_sync(0xffffffffu, C.found);
        if (C.lane == 0 && fw) atomicOr(&found_s, fw);
        __syncthreads();
        const uint32_t empty = cells_all & ~found_s;
        if (!empty) break;
        // cells without a survivor at iniThFAST rerun at minThFAST (:981-985): only their pixels stay valid
        C.th = T.min_th;
        uint32_t vm2 = 0;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (vm_all & (0x80u << (8 * k))) {
                const int cj = (int)(((uint32_t)(xr0 + k) * C.cmagic) >> 20);
                if ((empty >> cj) & 1u) vm2 |= 0x80u << (8 * k);
            }
        C.vm = vm2;
    }
}
