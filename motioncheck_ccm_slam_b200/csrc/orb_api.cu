// Host side of the extractor C ABI (include/ccm_orb.h): scale/quota tables exactly as the reference constructor
// computes them, per-shape level geometry, the device arena, and the kernel sequence of one batch.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ccm_orb.h"
#include "ccm_common.cuh"

// kernels: this file is compiled as part of the single translation unit ccm_orb_cuda.cu, after
// orb_pyramid_fast.cu, orb_quadtree.cu and orb_describe.cu.

#define CCM_MAX_CHUNKS 8
#define CCM_MAX_TICKETS 2
static thread_local std::string g_create_error = "";

struct ccm_orb {
    ccm_orb_config cfg;
    // reference constructor tables (ORBextractor.cpp:579-615)
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> quota;
    // current shape
    int rows = 0, cols = 0;
    LevelTable T;
    int qt_cap = 0;           // quadtree node capacity (power of two)
    int qt_ccap = 0;          // corners per (frame, level) the quadtree keeps in shared memory
    size_t qt_smem = 0;
    int max_kps = 0;          // == T.kps_per_frame
    // device arena
    std::vector<void*> allocs;
    void* encode_tiled = nullptr;                      // cuTensorMapEncodeTiled, fetched through the runtime
    // k_fast_chunks: two warps per chunk of whole cells (orb_fast_chunks.cu)
    uint4* d_chunks = nullptr; int n_chunks = 0;
    int chunk_off[CCM_MAX_LEVELS + 1] = {};   // chunks are in level order: level l owns [chunk_off[l], chunk_off[l + 1])
    // single-frame graph: one branch per level (FAST + quadtree of level l start when level l exists) and one for the blur
    cudaStream_t lvl_s[CCM_MAX_LEVELS + 1] = {}; cudaEvent_t lvl_ready[CCM_MAX_LEVELS + 1] = {}, lvl_done[CCM_MAX_LEVELS + 1] = {};
    FastChunkSmem fc_smem = {};
    bool fc_tma = false;                               // tensor maps of the arena levels are encoded (box FC_TW x tile_rows)
    FastTma fc_tmaps = {};
    float4* d_pat = nullptr;                           // k_orient_describe's test pattern table
    uint8_t* d_desc_block_level = nullptr;             // level of every CTA of k_orient_describe's grid
    uint32_t* d_corners = nullptr; int* d_corner_count = nullptr; uint16_t* d_node_of = nullptr;
    uint32_t* d_level_kps = nullptr; int* d_level_kp_count = nullptr;
    ccm_keypoint* d_kp = nullptr; uint8_t* d_desc = nullptr; int* d_n = nullptr;
    // pinned staging
    ccm_keypoint* h_kp = nullptr; uint8_t* h_desc = nullptr; int* h_n = nullptr;
    cudaStream_t stream = nullptr;
    // host entry point: copy-in / compute / copy-out overlap over sub-batches ("chunks") of one call
    cudaStream_t s_in = nullptr, s_out = nullptr;
    // two pipeline lanes: lane 0 runs on `stream`, lane 1 on its own stream; each has a side stream for the blur. Half
    // batches on the two lanes fill each other's latency-bound stages (quadtree, small pyramid levels).
    struct Lane { cudaStream_t s = nullptr, aux = nullptr; cudaEvent_t fork = nullptr, join = nullptr; } lane[2];
    uint8_t* d_stage = nullptr; size_t stage_bytes = 0;   // host entry point: staging for rows that are not 4-byte multiples
    cudaEvent_t ev_in[CCM_MAX_TICKETS][CCM_MAX_CHUNKS] = {}, ev_done[CCM_MAX_TICKETS][CCM_MAX_CHUNKS] = {}, ev_out[CCM_MAX_TICKETS][CCM_MAX_CHUNKS] = {};
    // batches in flight through the asynchronous host entry point (ccm_orb_extract_batch_submit / _wait)
    struct Ticket {
        bool active = false, empty = false, direct = false, trace = false;
        int n = 0, slot0 = 0, n_chunks = 0, capacity = 0;
        int first[CCM_MAX_CHUNKS + 1] = {};
        ccm_keypoint* out_kp = nullptr; uint8_t* out_desc = nullptr; int* out_n = nullptr;
        double t_call = 0, t_enq = 0;
        cudaEvent_t tev[1 + 3 * CCM_MAX_CHUNKS] = {};
    } tk[CCM_MAX_TICKETS];
    int next_slot = 0, next_lane = 0;
    long long launches = 0;
    // single-frame call (ccm_orb_extract, the call Frame::ExtractORB makes): the whole kernel sequence of one frame in arena
    // slot 0 plus the copy-out into the pinned staging is ONE CUDA graph per (shape, level-0 pitch)
    cudaGraphExec_t graph1 = nullptr;
    size_t graph1_pitch = 0;
    long long graph1_launches = 0;
    bool graph1_failed = false;
    bool blur_ocv24 = false;                           // CCM_ORB_BLUR=opencv2: OpenCV 2.4.13's 8-bit Gaussian taps (orb_describe.cu)
    // optional per-stage CUDA-event timing (ccm_orb_set_stage_timing): events bracket each stage of run_pipeline
    bool stage_timing = false;
    cudaEvent_t ev[CCM_N_STAGES + 1] = {};
    double stage_ms[CCM_N_STAGES] = {};
    long long stage_calls = 0;
    bool stage_pending = false;
    int last_frames = 0;
    const uint8_t* last_l0 = nullptr; size_t last_l0_pitch = 0, last_l0_stride = 0;
    std::string err;
};

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                 \
            return CCM_E_CUDA;                                                                           \
        }                                                                                                \
    } while (0)

// launch check that names the kernel: a failed launch is reported where it happened, not at the end of the batch
#define CKL(name, expr)                                                                                  \
    do {                                                                                                 \
        cudaError_t e_ = (expr);                                                                         \
        if (e_ == cudaSuccess) e_ = cudaGetLastError();                                                  \
        if (e_ != cudaSuccess) {                                                                         \
            h->err = std::string("launch of " name ": ") + cudaGetErrorString(e_);                       \
            return CCM_E_CUDA;                                                                           \
        }                                                                                                \
    } while (0)

static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

template <typename Tp> static int dev_alloc(ccm_orb* h, Tp** p, size_t count) {
    void* q = nullptr;
    CK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(Tp)));
    h->allocs.push_back(q);
    *p = (Tp*)q;
    return CCM_OK;
}

static void free_arena(ccm_orb* h) {
    for (void* p : h->allocs) cudaFree(p);
    h->allocs.clear();
    if (h->h_kp) cudaFreeHost(h->h_kp);
    if (h->h_desc) cudaFreeHost(h->h_desc);
    if (h->h_n) cudaFreeHost(h->h_n);
    h->h_kp = nullptr; h->h_desc = nullptr; h->h_n = nullptr;
}

// cv::resize INTER_LINEAR coefficient table (OpenCV resize.cpp), see oracle/cv_primitives.cpp linear_table
static void linear_table(int ssize, int dsize, std::vector<int>& ofs, std::vector<short>& coef) {
    const double scale = 1. / ((double)dsize / ssize);
    ofs.resize(dsize); coef.resize(2 * (size_t)dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (s < 0) { f = 0; s = 0; }
        if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        ofs[d] = s;
        coef[2 * d] = (short)lrintf((1.f - f) * 2048.f);
        coef[2 * d + 1] = (short)lrintf(f * 2048.f);
    }
}

// k_resize_level's table for groups of 4 adjacent outputs: the taps of a group start at source pixel s0 = ofs[4g]; entry
// = {s0 & ~3, 8 * (s0 & 3), 4 selector bytes (tap0 - s0) | (tap1 - s0) << 4, 0} {4 x (a0 | a1 << 16)}. Outputs past
// the row end repeat the last pixel. Returns false when some group's taps do not fit in 8 bytes (scale factor > 2).
static bool resize_group_table(int ssize, int dsize, const std::vector<int>& ofs, const std::vector<short>& coef, std::vector<uint4>& out) {
    const int groups = (dsize + 3) / 4;
    out.resize(2 * (size_t)groups);
    for (int g = 0; g < groups; g++) {
        const int s0 = ofs[4 * g];
        uint32_t sel = 0, c[4];
        for (int i = 0; i < 4; i++) {
            const int x = std::min(4 * g + i, dsize - 1);
            const int t0 = ofs[x] - s0, t1 = std::min(ofs[x] + 1, ssize - 1) - s0;
            if (t0 < 0 || t1 < 0 || t0 > 7 || t1 > 7) return false;
            sel |= (uint32_t)(t0 | (t1 << 4)) << (8 * i);
            c[i] = (uint32_t)(uint16_t)coef[2 * x] | ((uint32_t)(uint16_t)coef[2 * x + 1] << 16);
        }
        out[2 * g] = make_uint4((uint32_t)(s0 & ~3), (uint32_t)(8 * (s0 & 3)), sel, 0u);
        out[2 * g + 1] = make_uint4(c[0], c[1], c[2], c[3]);
    }
    return true;
}

// ---- TMA descriptors for k_fast_chunks<true> ----
typedef CUresult (*ccm_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// u8 tensor [frames][rows][pitch] at base, box [1][box_rows][ts]; out-of-range parts of a box are zero-filled (never read)
static bool encode_level_map(ccm_orb* h, CUtensorMap* out, const void* base, size_t pitch, int rows, size_t frame_stride, int frames,
                             int ts, int box_rows) {
    if (!h->encode_tiled || ((uintptr_t)base & 15) || (pitch & 15) || (frame_stride & 15) || pitch == 0 || frames < 1) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frame_stride};
    const cuuint32_t box[3] = {(cuuint32_t)ts, (cuuint32_t)box_rows, 1u}, estr[3] = {1u, 1u, 1u};
    return ((ccm_encode_tiled_fn)h->encode_tiled)(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// launch with the programmatic-stream-serialization attribute (see pdl_wait / pdl_launch, ccm_common.cuh).
// tl_pdl: run_pipeline turns the attribute off for a batch submitted while another ticket is in flight (the asynchronous
// pair in steady state): early-resident CTAs of the next kernel then only take slots from the other lane's kernels
// (measured: 124.8 k -> 128.5 k frames/s through the asynchronous pair; alone on the GPU the attribute is worth 129 k -> 135 k).
static thread_local bool tl_pdl = true;
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    static const bool off = std::getenv("CCM_ORB_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = (off || !tl_pdl) ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// k_fast_chunks' work list: every cell row of every level cut into chunks of whole cells that one warp covers with its 32
// aligned 4-pixel words (lane 0's word holds x_c0 - 3, the last valid pixel lies in lane 30's word or before).
// Entry = {x_c0 | y0 << 16, wc | ch << 16, level | ncells << 8 | wCell << 16, xw0} {magic(wCell), 0, 0, 0}.
static void fast_chunk_table(const LevelTable& T, std::vector<uint4>& out, int* max_ch) {
    auto magic = [](int d) { return (0x100000u + (uint32_t)d - 1) / (uint32_t)d; };
    *max_ch = 1;
    for (int l = 0; l < T.nlevels; l++) {
        const LevelDesc& D = T.l[l];
        for (int i = 0; i < D.nRows; i++) {
            const int y0 = CCM_EDGE + i * D.hCell, ch = std::min(y0 + D.hCell, D.h - CCM_EDGE) - y0;
            if (ch <= 0) continue;
            *max_ch = std::max(*max_ch, ch);
            for (int j = 0; j < D.nCols;) {
                const int x_c0 = CCM_EDGE + j * D.wCell, xw0 = (x_c0 - 3) & ~3;
                int wc = 0, n = 0;
                while (j + n < D.nCols) {
                    const int x = x_c0 + wc, cw = std::min(x + D.wCell, D.w - CCM_EDGE) - x;
                    if (cw <= 0 || (n > 0 && x + cw - 1 > xw0 + FC_MAX_SPAN)) break;
                    wc += cw; n++;
                }
                if (n == 0) break;                               // the rest of the row is clipped away (:971)
                out.push_back(make_uint4((uint32_t)x_c0 | ((uint32_t)y0 << 16), (uint32_t)wc | ((uint32_t)ch << 16),
                                         (uint32_t)l | ((uint32_t)n << 8) | ((uint32_t)D.wCell << 16), (uint32_t)xw0));
                out.push_back(make_uint4(magic(D.wCell), 0u, 0u, 0u));
                j += n;
            }
        }
    }
}

// Level geometry for one image shape, following ComputePyramid (:1284-1286), ComputeKeyPointsOctTree (:941-955)
// and DistributeOctTree (:711-713). Allocates the arena for cfg.max_batch frames of this shape.
static int configure_shape(ccm_orb* h, int rows, int cols) {
    if (rows == h->rows && cols == h->cols) return CCM_OK;
    if (rows > h->cfg.max_height || cols > h->cfg.max_width) { h->err = "image larger than max_width x max_height of the handle"; return CCM_E_CAPACITY; }
    CK(cudaStreamSynchronize(h->stream));
    free_arena(h);
    if (h->graph1) { cudaGraphExecDestroy(h->graph1); h->graph1 = nullptr; }
    h->rows = h->cols = 0;
    LevelTable& T = h->T;
    std::memset(&T, 0, sizeof(T));
    const int L = h->cfg.nlevels, B = h->cfg.max_batch;
    T.nlevels = L; T.ini_th = h->cfg.ini_th_fast; T.min_th = h->cfg.min_th_fast;
    int cells = 0, corners = 0, kps = 0, max_node_cap = 0, blur_tiles = 0, desc_blocks = 0;
    for (int l = 0; l < L; l++) {
        LevelDesc& D = T.l[l];
        D.w = (int)lrintf((float)cols * h->isf[l]);
        D.h = (int)lrintf((float)rows * h->isf[l]);
        const float width = (float)(D.w - 2 * CCM_MIN_BORDER), height = (float)(D.h - 2 * CCM_MIN_BORDER);
        D.nCols = (int)(width / 30.f); D.nRows = (int)(height / 30.f);
        if (D.nCols <= 0 || D.nRows <= 0) { h->err = "image too small for this many pyramid levels (a level has no 30-px cell)"; return CCM_E_BADARG; }
        D.wCell = (int)std::ceil(width / D.nCols); D.hCell = (int)std::ceil(height / D.nRows);
        if (D.wCell > 64 || D.hCell > 64) { h->err = "internal: cell larger than 64 px"; return CCM_E_BADARG; }
        D.cell_base = cells; cells += D.nCols * D.nRows;
        D.blur_tile_base = blur_tiles; D.blur_tiles_x = (D.w + BLUR_TW - 1) / BLUR_TW;
        blur_tiles += D.blur_tiles_x * ((D.h + BLUR_TH - 1) / BLUR_TH);
        D.quota = h->quota[l];
        D.nIni = (int)std::round((float)(D.w - 32) / (D.h - 32));
        if (D.nIni < 1) { h->err = "unsupported aspect ratio (height > 2*width): the reference divides by zero here"; return CCM_E_BADARG; }
        D.hX = (float)(D.w - 32) / D.nIni;
        D.node_cap = std::max(D.quota + 3, 4 * D.nIni);
        max_node_cap = std::max(max_node_cap, D.node_cap);
        int cc = 0;
        for (int i = 0; i < D.nRows; i++)
            for (int j = 0; j < D.nCols; j++) {
                const int x0 = CCM_EDGE + j * D.wCell, y0 = CCM_EDGE + i * D.hCell;
                const int cw = std::min(x0 + D.wCell, D.w - CCM_EDGE) - x0, ch = std::min(y0 + D.hCell, D.h - CCM_EDGE) - y0;
                if (cw > 0 && ch > 0) cc += ((cw + 1) / 2) * ((ch + 1) / 2);
            }
        D.corner_cap = std::max(cc, 1); D.corner_off = corners; corners += align_up(D.corner_cap, 4);
        if (D.corner_cap > 0x7FFFF) { h->err = "image too large (corner capacity per level exceeds 2^19)"; return CCM_E_BADARG; }
        D.kp_off = kps; kps += D.node_cap;
        D.desc_block_base = desc_blocks; desc_blocks += (D.node_cap + 7) / 8;
        D.scale = h->sf[l];
        D.kp_size = (float)(int)(31 * h->sf[l]);
        D.img_pitch = D.blur_pitch = align_up(D.w, 32);
        D.img_stride = D.blur_stride = (size_t)D.img_pitch * D.h;
    }
    if (max_node_cap > 2048) { h->err = "nfeatures too large: more than 2045 features on one level"; return CCM_E_BADARG; }
    T.total_cells = cells; T.total_blur_tiles = blur_tiles; T.total_desc_blocks = desc_blocks; T.corners_per_frame = corners; T.kps_per_frame = kps;
    h->max_kps = kps;
    int cap = 32; while (cap < max_node_cap) cap <<= 1;
    h->qt_cap = cap;
    h->qt_ccap = 4096;                                   // corners per (frame, level) cached in shared memory: 24 KB
    h->qt_smem = (size_t)cap * 80 + 32 * sizeof(int) + (size_t)h->qt_ccap * 6;
    CK(cudaFuncSetAttribute(k_quadtree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->qt_smem));

    for (int l = 0; l < L; l++) {
        LevelDesc& D = T.l[l];
        int rc;
        if ((rc = dev_alloc(h, &D.img, D.img_stride * B + 64))) return rc;     // +64: 4-pixel stores may touch the pitch padding
        if ((rc = dev_alloc(h, &D.blur, D.blur_stride * B + 64))) return rc;
        if (l > 0) {
            std::vector<int> xo, yo; std::vector<short> xa, ya;
            linear_table(T.l[l - 1].w, D.w, xo, xa);
            linear_table(T.l[l - 1].h, D.h, yo, ya);
            int* dxo; int* dyo; short* dxa; short* dya;
            if ((rc = dev_alloc(h, &dxo, xo.size()))) return rc;
            if ((rc = dev_alloc(h, &dyo, yo.size()))) return rc;
            if ((rc = dev_alloc(h, &dxa, xa.size()))) return rc;
            if ((rc = dev_alloc(h, &dya, ya.size()))) return rc;
            CK(cudaMemcpy(dxo, xo.data(), xo.size() * sizeof(int), cudaMemcpyHostToDevice));
            CK(cudaMemcpy(dyo, yo.data(), yo.size() * sizeof(int), cudaMemcpyHostToDevice));
            CK(cudaMemcpy(dxa, xa.data(), xa.size() * sizeof(short), cudaMemcpyHostToDevice));
            CK(cudaMemcpy(dya, ya.data(), ya.size() * sizeof(short), cudaMemcpyHostToDevice));
            D.xofs = dxo; D.yofs = dyo; D.xcoef = dxa; D.ycoef = dya;
            std::vector<uint4> grp;
            D.xgrp = nullptr;
            if (resize_group_table(T.l[l - 1].w, D.w, xo, xa, grp)) {
                uint4* dg;
                if ((rc = dev_alloc(h, &dg, grp.size()))) return rc;
                CK(cudaMemcpy(dg, grp.data(), grp.size() * sizeof(uint4), cudaMemcpyHostToDevice));
                D.xgrp = dg;
            }
            D.resize_mlp = D.xgrp != nullptr && !std::getenv("CCM_ORB_NO_RESIZE_MLP");
            for (int y0 = 0; y0 < D.h && D.resize_mlp; y0 += RESIZE_ROWS)
                for (int j = 0; j < RESIZE_ROWS && y0 + j < D.h; j++) {
                    const int o = yo[y0 + j] - yo[y0];
                    if (o != j && o != j + 1) D.resize_mlp = 0;
                }
        }
    }
    int rc;
    {
        float4 pt[256];
        orb_pattern_table(pt);
        if ((rc = dev_alloc(h, &h->d_pat, 256))) return rc;
        CK(cudaMemcpy(h->d_pat, pt, sizeof(pt), cudaMemcpyHostToDevice));
        std::vector<uint8_t> bl((size_t)std::max(desc_blocks, 1), 0);
        for (int l = 0; l < L; l++)
            for (int b = T.l[l].desc_block_base; b < (l + 1 < L ? T.l[l + 1].desc_block_base : desc_blocks); b++) bl[b] = (uint8_t)l;
        if ((rc = dev_alloc(h, &h->d_desc_block_level, bl.size()))) return rc;
        CK(cudaMemcpy(h->d_desc_block_level, bl.data(), bl.size(), cudaMemcpyHostToDevice));
    }
    if (!h->encode_tiled && !std::getenv("CCM_ORB_NO_TMA")) {
        cudaDriverEntryPointQueryResult q;
        void* fn = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) h->encode_tiled = fn;
        else cudaGetLastError();
    }
    {
        std::vector<uint4> chunks;
        int fc_max_ch = 1;
        fast_chunk_table(T, chunks, &fc_max_ch);
        h->n_chunks = (int)(chunks.size() / 2);
        for (int l = 0; l <= CCM_MAX_LEVELS; l++) h->chunk_off[l] = 0;
        for (size_t c = 0; c < chunks.size(); c += 2) h->chunk_off[(chunks[c].z & 0xffu) + 1]++;
        for (int l = 0; l < CCM_MAX_LEVELS; l++) h->chunk_off[l + 1] += h->chunk_off[l];
        h->fc_smem = fast_chunk_smem(fc_max_ch);
        if ((rc = dev_alloc(h, &h->d_chunks, chunks.size()))) return rc;
        CK(cudaMemcpy(h->d_chunks, chunks.data(), chunks.size() * sizeof(uint4), cudaMemcpyHostToDevice));
        CK(cudaFuncSetAttribute(k_fast_chunks<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CK(cudaFuncSetAttribute(k_fast_chunks<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CK(cudaFuncSetAttribute(k_fast_chunks<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fc_smem.total));
        CK(cudaFuncSetAttribute(k_fast_chunks<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fc_smem.total));
        // TMA tile load: one descriptor per level, box = FC_TW x tile_rows; the arena levels are encoded once per shape,
        // level 0 (the caller's buffer) per call (run_pipeline). CCM_ORB_NO_TMA=1 forces the word-load variant.
        h->fc_tma = h->encode_tiled != nullptr;
        for (int l = 1; l < L && h->fc_tma; l++)
            h->fc_tma = encode_level_map(h, &h->fc_tmaps.m[l], T.l[l].img, (size_t)T.l[l].img_pitch, T.l[l].h, T.l[l].img_stride, B, FC_TW, h->fc_smem.tile_rows);
    }
    if ((rc = dev_alloc(h, &h->d_corners, (size_t)corners * B))) return rc;
    if ((rc = dev_alloc(h, &h->d_node_of, (size_t)corners * B))) return rc;
    if ((rc = dev_alloc(h, &h->d_corner_count, (size_t)CCM_MAX_LEVELS * B))) return rc;
    if ((rc = dev_alloc(h, &h->d_level_kp_count, (size_t)CCM_MAX_LEVELS * B))) return rc;
    if ((rc = dev_alloc(h, &h->d_level_kps, (size_t)kps * B))) return rc;
    if ((rc = dev_alloc(h, &h->d_kp, (size_t)kps * B))) return rc;
    if ((rc = dev_alloc(h, &h->d_desc, (size_t)kps * B * 32))) return rc;
    if ((rc = dev_alloc(h, &h->d_n, (size_t)B))) return rc;
    CK(cudaMallocHost((void**)&h->h_kp, (size_t)kps * B * sizeof(ccm_keypoint)));
    CK(cudaMallocHost((void**)&h->h_desc, (size_t)kps * B * 32));
    CK(cudaMallocHost((void**)&h->h_n, (size_t)B * sizeof(int)));
    h->rows = rows; h->cols = cols;
    return CCM_OK;
}

// Fold the event intervals of the previous timed batch into the per-stage sums (waits for that batch).
static int collect_stage_times(ccm_orb* h) {
    if (!h->stage_pending) return CCM_OK;
    CK(cudaEventSynchronize(h->ev[CCM_N_STAGES]));
    for (int i = 0; i < CCM_N_STAGES; i++) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
        h->stage_ms[i] += ms;
    }
    h->stage_calls++;
    h->stage_pending = false;
    return CCM_OK;
}

// One batch through K1..K7 on the handle's stream. Level 0 is read in place from (l0, pitch, stride).
// Frames [f0, f0+n) of the arena: every per-frame array is addressed through pointers advanced by f0 frames, so a call
// can be cut into sub-batches that overlap with their own copies (ccm_orb_extract_batch).
static int run_pipeline(ccm_orb* h, int n, const uint8_t* l0, size_t l0_pitch, size_t l0_stride,
                        ccm_keypoint* d_kp, uint8_t* d_desc, int* d_n, int capacity, int f0 = 0, bool allow_timing = true, int lane_id = 0,
                        bool pdl = true) {
    tl_pdl = pdl;
    h->T.l[0].img = const_cast<uint8_t*>(l0);
    h->T.l[0].img_pitch = (int)l0_pitch;
    h->T.l[0].img_stride = l0_stride;
    if (f0 == 0) { h->last_l0 = l0; h->last_l0_pitch = l0_pitch; h->last_l0_stride = l0_stride; }
    h->last_frames = f0 + n;
    LevelTable T = h->T;
    for (int l = 0; l < T.nlevels; l++) {
        T.l[l].img += (size_t)f0 * T.l[l].img_stride;
        T.l[l].blur += (size_t)f0 * T.l[l].blur_stride;
    }
    struct Off {
        uint32_t* corners; int* corner_count; uint16_t* node_of; uint32_t* level_kps; int* level_kp_count;
    } A = {h->d_corners + (size_t)f0 * T.corners_per_frame, h->d_corner_count + (size_t)f0 * CCM_MAX_LEVELS,
           h->d_node_of + (size_t)f0 * T.corners_per_frame, h->d_level_kps + (size_t)f0 * T.kps_per_frame,
           h->d_level_kp_count + (size_t)f0 * CCM_MAX_LEVELS};
    ccm_orb::Lane& LN = h->lane[lane_id];
    cudaStream_t s = LN.s;
    const bool st = h->stage_timing && allow_timing && lane_id == 0;
    if (st) { int rc = collect_stage_times(h); if (rc != CCM_OK) return rc; }
    // The blur depends only on the pyramid; the quadtree is one small CTA per (frame, level) and latency-bound. Outside
    // stage-timing mode the blur runs on a side stream next to the quadtree and joins before the descriptor stage.
    const bool fork = !st;
    {
        CK(cudaMemsetAsync(A.corner_count, 0, sizeof(int) * CCM_MAX_LEVELS * n, s));
        if (st) CK(cudaEventRecord(h->ev[0], s));
        for (int l = 1; l < T.nlevels; l++) {
            dim3 g((T.l[l].w + 127) / 128, (T.l[l].h + 7) / 8, n);
            if (T.l[l].resize_mlp) CKL("k_resize_level_mlp", launch_pdl(k_resize_level_mlp, dim3(g.x, (T.l[l].h + 8 * RESIZE_ROWS - 1) / (8 * RESIZE_ROWS), n), dim3(32, 8), 0, s, T, l));
            else if (T.l[l].xgrp) CKL("k_resize_level", launch_pdl(k_resize_level, dim3(g.x, (T.l[l].h + 8 * RESIZE_ROWS - 1) / (8 * RESIZE_ROWS), n), dim3(32, 8), 0, s, T, l));
            else { k_resize_level_generic<<<g, dim3(32, 8), 0, s>>>(T, l); CKL("k_resize_level_generic", cudaSuccess); }
            h->launches++;
        }
        if (st) CK(cudaEventRecord(h->ev[1], s));
        {
            // level 0 is the caller's buffer: its descriptor is encoded per call; a base / pitch / stride that TMA cannot address
            // (not 16-byte multiples) takes the word-load variant of the kernel for this call
            FastTma tm = h->fc_tmaps;
            const bool tma = h->fc_tma && encode_level_map(h, &tm.m[0], l0, l0_pitch, T.l[0].h, l0_stride, f0 + n, FC_TW, h->fc_smem.tile_rows);
            const dim3 grid(h->n_chunks, n);
            if (tma) CKL("k_fast_chunks<TMA>", launch_pdl(k_fast_chunks<true>, grid, dim3(32 * FC_WARPS), (size_t)h->fc_smem.total, s, T, h->d_chunks, A.corners, A.corner_count, h->fc_smem, f0, tm));
            else CKL("k_fast_chunks<words>", launch_pdl(k_fast_chunks<false>, grid, dim3(32 * FC_WARPS), (size_t)h->fc_smem.total, s, T, h->d_chunks, A.corners, A.corner_count, h->fc_smem, f0, tm));
            h->launches++;
        }
        if (st) CK(cudaEventRecord(h->ev[2], s));
        if (fork) CK(cudaEventRecord(LN.fork, s));
        if (!fork) {
            if (h->blur_ocv24) k_blur_levels<true><<<dim3(T.total_blur_tiles, n), 256, 0, s>>>(T); else k_blur_levels<false><<<dim3(T.total_blur_tiles, n), 256, 0, s>>>(T);
            CKL("k_blur_levels", cudaSuccess);
            h->launches++;
        }
        if (st) CK(cudaEventRecord(h->ev[3], s));
    }
    {
        CKL("k_quadtree", launch_pdl(k_quadtree, dim3(T.nlevels, n), dim3(QT_THREADS), h->qt_smem, s, T, h->qt_cap, h->qt_ccap, 0, (const uint32_t*)A.corners,
                                     (const int*)A.corner_count, A.node_of, A.level_kps, A.level_kp_count));
        h->launches++;
        if (fork) {
            CK(cudaStreamWaitEvent(LN.aux, LN.fork, 0));
            if (h->blur_ocv24) k_blur_levels<true><<<dim3(T.total_blur_tiles, n), 256, 0, LN.aux>>>(T); else k_blur_levels<false><<<dim3(T.total_blur_tiles, n), 256, 0, LN.aux>>>(T);
            CKL("k_blur_levels", cudaSuccess);
            h->launches++;
            CK(cudaEventRecord(LN.join, LN.aux));
            CK(cudaStreamWaitEvent(s, LN.join, 0));
        }
        if (st) CK(cudaEventRecord(h->ev[4], s));
        k_orient_describe<<<dim3(T.total_desc_blocks, n), 256, 0, s>>>(T, h->d_pat, h->d_desc_block_level, A.level_kps, A.level_kp_count,
                                                                            (ccm_keypoint_dev*)d_kp, d_desc, d_n, capacity);
        CKL("k_orient_describe", cudaSuccess);
        h->launches++;
        if (st) { CK(cudaEventRecord(h->ev[5], s)); h->stage_pending = true; }
    }
    CK(cudaGetLastError());
    return CCM_OK;
}

extern "C" {

const char* ccm_last_error(const ccm_orb* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int ccm_orb_create(const ccm_orb_config* cfg, ccm_orb** out) {
    if (!cfg || !out) { g_create_error = "null argument"; return CCM_E_BADARG; }
    *out = nullptr;
    if (cfg->nfeatures < 1 || cfg->nlevels < 1 || cfg->nlevels > CCM_MAX_LEVELS || !(cfg->scale_factor > 1.0f) ||
        cfg->ini_th_fast < 1 || cfg->ini_th_fast > 255 || cfg->min_th_fast < 1 || cfg->min_th_fast > 255 ||
        cfg->max_width < 1 || cfg->max_width > 4095 || cfg->max_height < 1 || cfg->max_height > 4095 || cfg->max_batch < 1 ||
        cfg->max_batch > 65535) {
        g_create_error = "bad ccm_orb_config (need nfeatures>=1, 1<=nlevels<=16, scale_factor>1, FAST thresholds in 1..255, size<=4095, 1<=max_batch<=65535)";
        return CCM_E_BADARG;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) {
        g_create_error = std::string("no usable CUDA device ") + std::to_string(cfg->device) + ": " + cudaGetErrorString(e) + " (there is no CPU fallback)";
        return CCM_E_CUDA;
    }
    ccm_orb* h = new ccm_orb();
    h->cfg = *cfg;
    if (const char* e = std::getenv("CCM_ORB_BLUR")) h->blur_ocv24 = std::strcmp(e, "opencv2") == 0;
    // ORBextractor::ORBextractor, cslam/src/ORBextractor.cpp:579-615 — float/double mix kept as written
    const int L = cfg->nlevels;
    const double scaleFactor = cfg->scale_factor;
    h->sf.assign(L, 1.0f); h->sig2.assign(L, 1.0f); h->isf.resize(L); h->isig2.resize(L); h->quota.assign(L, 0);
    for (int i = 1; i < L; i++) { h->sf[i] = (float)(h->sf[i - 1] * scaleFactor); h->sig2[i] = h->sf[i] * h->sf[i]; }
    for (int i = 0; i < L; i++) { h->isf[i] = 1.0f / h->sf[i]; h->isig2[i] = 1.0f / h->sig2[i]; }
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = cfg->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l < L - 1; l++) { h->quota[l] = (int)lrintf(nDesired); sum += h->quota[l]; nDesired *= factor; }
    h->quota[L - 1] = std::max(cfg->nfeatures - sum, 0);

    auto fail = [&](int rc) { g_create_error = h->err; free_arena(h); if (h->stream) cudaStreamDestroy(h->stream); delete h; return rc; };
    if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) { h->err = cudaGetErrorString(e); return fail(CCM_E_CUDA); }
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) { h->err = cudaGetErrorString(e); return fail(CCM_E_CUDA); }
    if ((e = cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&h->lane[1].s, cudaStreamNonBlocking)) != cudaSuccess) { h->err = cudaGetErrorString(e); return fail(CCM_E_CUDA); }
    h->lane[0].s = h->stream;
    for (int i = 0; i < 2; i++)
        if ((e = cudaStreamCreateWithFlags(&h->lane[i].aux, cudaStreamNonBlocking)) != cudaSuccess ||
            (e = cudaEventCreateWithFlags(&h->lane[i].fork, cudaEventDisableTiming)) != cudaSuccess ||
            (e = cudaEventCreateWithFlags(&h->lane[i].join, cudaEventDisableTiming)) != cudaSuccess) { h->err = cudaGetErrorString(e); return fail(CCM_E_CUDA); }
    for (int t = 0; t < CCM_MAX_TICKETS; t++)
        for (int i = 0; i < CCM_MAX_CHUNKS; i++)
            if ((e = cudaEventCreateWithFlags(&h->ev_in[t][i], cudaEventDisableTiming)) != cudaSuccess ||
                (e = cudaEventCreateWithFlags(&h->ev_done[t][i], cudaEventDisableTiming)) != cudaSuccess ||
                (e = cudaEventCreateWithFlags(&h->ev_out[t][i], cudaEventDisableTiming)) != cudaSuccess) { h->err = cudaGetErrorString(e); return fail(CCM_E_CUDA); }
    int rc = configure_shape(h, cfg->max_height, cfg->max_width);
    if (rc != CCM_OK) return fail(rc);
    *out = h;
    return CCM_OK;
}

void ccm_orb_destroy(ccm_orb* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    for (cudaStream_t st : {h->s_in, h->stream, h->lane[1].s, h->lane[0].aux, h->lane[1].aux, h->s_out}) if (st) cudaStreamSynchronize(st);
    free_arena(h);
    if (h->graph1) cudaGraphExecDestroy(h->graph1);
    for (int l = 0; l <= CCM_MAX_LEVELS; l++) {
        if (h->lvl_s[l]) cudaStreamDestroy(h->lvl_s[l]);
        if (h->lvl_ready[l]) cudaEventDestroy(h->lvl_ready[l]);
        if (h->lvl_done[l]) cudaEventDestroy(h->lvl_done[l]);
    }
    for (int i = 0; i <= CCM_N_STAGES; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (int t = 0; t < CCM_MAX_TICKETS; t++) {
        for (int i = 0; i < CCM_MAX_CHUNKS; i++) {
            if (h->ev_in[t][i]) cudaEventDestroy(h->ev_in[t][i]);
            if (h->ev_done[t][i]) cudaEventDestroy(h->ev_done[t][i]);
            if (h->ev_out[t][i]) cudaEventDestroy(h->ev_out[t][i]);
        }
        for (auto& e : h->tk[t].tev) if (e) cudaEventDestroy(e);
    }
    if (h->d_stage) cudaFree(h->d_stage);
    for (int i = 0; i < 2; i++) {
        if (h->lane[i].fork) cudaEventDestroy(h->lane[i].fork);
        if (h->lane[i].join) cudaEventDestroy(h->lane[i].join);
        if (h->lane[i].aux) cudaStreamDestroy(h->lane[i].aux);
    }
    if (h->lane[1].s) cudaStreamDestroy(h->lane[1].s);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int ccm_orb_get_tables(const ccm_orb* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* fpl) {
    if (!h) return CCM_E_BADARG;
    for (int l = 0; l < h->cfg.nlevels; l++) {
        if (scale) scale[l] = h->sf[l];
        if (inv_scale) inv_scale[l] = h->isf[l];
        if (sigma2) sigma2[l] = h->sig2[l];
        if (inv_sigma2) inv_sigma2[l] = h->isig2[l];
        if (fpl) fpl[l] = h->quota[l];
    }
    return CCM_OK;
}

int ccm_orb_max_keypoints(const ccm_orb* h) { return h ? h->max_kps : CCM_E_BADARG; }
void* ccm_orb_stream(ccm_orb* h) { return h ? (void*)h->stream : nullptr; }
long long ccm_orb_launch_count(const ccm_orb* h) { return h ? h->launches : 0; }

int ccm_orb_set_stage_timing(ccm_orb* h, int enable) {
    if (!h) return CCM_E_BADARG;
    CK(cudaSetDevice(h->cfg.device));
    if (enable && !h->ev[0])
        for (int i = 0; i <= CCM_N_STAGES; i++) CK(cudaEventCreate(&h->ev[i]));
    if (!enable) { int rc = collect_stage_times(h); if (rc != CCM_OK) return rc; }
    h->stage_timing = enable != 0;
    return CCM_OK;
}

int ccm_orb_get_stage_ms(ccm_orb* h, double* sum_ms, long long* calls, int reset) {
    if (!h) return CCM_E_BADARG;
    CK(cudaSetDevice(h->cfg.device));
    int rc = collect_stage_times(h);
    if (rc != CCM_OK) return rc;
    for (int i = 0; i < CCM_N_STAGES; i++) { if (sum_ms) sum_ms[i] = h->stage_ms[i]; if (reset) h->stage_ms[i] = 0; }
    if (calls) *calls = h->stage_calls;
    if (reset) h->stage_calls = 0;
    return CCM_OK;
}

int ccm_orb_sync(ccm_orb* h) {
    if (!h) return CCM_E_BADARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    return CCM_OK;
}

int ccm_orb_extract_batch_device(ccm_orb* h, const uint8_t* d_imgs, int n, int rows, int cols, size_t pitch, size_t frame_stride,
                                 ccm_keypoint* d_kp, uint8_t* d_desc, int capacity, int* d_n) {
    if (!h) return CCM_E_BADARG;
    if (!d_imgs || !d_kp || !d_desc || !d_n || n < 1 || rows < 1 || cols < 1 || capacity < 1 || pitch < (size_t)cols) { h->err = "bad argument"; return CCM_E_BADARG; }
    if (((uintptr_t)d_imgs | pitch | frame_stride) & 3) { h->err = "device images: base, pitch and frame stride must be multiples of 4 bytes"; return CCM_E_BADARG; }
    if (n > h->cfg.max_batch) { h->err = "batch larger than max_batch"; return CCM_E_CAPACITY; }
    for (int i = 0; i < CCM_MAX_TICKETS; i++)
        if (h->tk[i].active) { h->err = "ccm_orb_extract_batch_device while a submitted batch is in flight: wait for it first (both use the arena's frame slots)"; return CCM_E_BADARG; }
    CK(cudaSetDevice(h->cfg.device));
    int rc = configure_shape(h, rows, cols);
    if (rc != CCM_OK) return rc;
    // one lane: splitting a resident batch over both lanes measured slower (75.3k vs 77.0k frames/s at 64 frames)
    return run_pipeline(h, n, d_imgs, pitch, frame_stride, d_kp, d_desc, d_n, capacity);
}

// Host entry point, asynchronous pair. submit enqueues copy-in, kernels and copy-out of one batch and returns; wait blocks
// until that batch's results are in the caller's arrays. Every batch takes a contiguous range of the arena's frame slots
// (ring allocation over max_batch), so a second batch can be submitted while the first is still on the device: its
// copy-in overlaps the first one's kernels and the link and the SMs are busy at the same time across calls.
static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static bool is_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

int ccm_orb_extract_batch_submit(ccm_orb* h, const uint8_t* const* imgs, int n, int rows, int cols, size_t step,
                                 ccm_keypoint* out_kp, uint8_t* out_desc, int capacity, int* out_n, int* ticket) {
    if (!h) return CCM_E_BADARG;
    if (!imgs || !out_n || !ticket || n < 1 || capacity < 0 || (capacity > 0 && (!out_kp || !out_desc))) { h->err = "bad argument"; return CCM_E_BADARG; }
    if (rows < 0 || cols < 0 || (rows > 0 && cols > 0 && step < (size_t)cols)) { h->err = "bad image geometry"; return CCM_E_BADARG; }
    if (n > h->cfg.max_batch) { h->err = "batch larger than max_batch"; return CCM_E_CAPACITY; }
    int tk = -1, in_flight = 0;
    for (int i = 0; i < CCM_MAX_TICKETS; i++) { if (h->tk[i].active) in_flight++; else if (tk < 0) tk = i; }
    if (tk < 0) { h->err = "two batches are already in flight: wait for one first"; return CCM_E_CAPACITY; }
    ccm_orb::Ticket& K = h->tk[tk];
    {
        cudaEvent_t keep[1 + 3 * CCM_MAX_CHUNKS];
        std::memcpy(keep, K.tev, sizeof(keep));
        K = ccm_orb::Ticket();
        std::memcpy(K.tev, keep, sizeof(keep));
    }
    K.n = n; K.out_kp = out_kp; K.out_desc = out_desc; K.out_n = out_n; K.capacity = capacity; K.t_call = now_ms();
    if (rows == 0 || cols == 0) { K.active = true; K.empty = true; *ticket = tk; return CCM_OK; }          // ORBextractor.cpp:1219-1220
    for (int f = 0; f < n; f++)
        if (!imgs[f]) { h->err = "null image in batch"; return CCM_E_BADARG; }
    CK(cudaSetDevice(h->cfg.device));
    if ((rows != h->rows || cols != h->cols) && in_flight) { h->err = "image shape changed while a batch is in flight"; return CCM_E_BADARG; }
    int rc = configure_shape(h, rows, cols);
    if (rc != CCM_OK) return rc;
    // frame slots: a synchronous caller (nothing in flight) always starts at slot 0, which the debug taps rely on
    if (!in_flight) { h->next_slot = 0; h->next_lane = 0; }
    int s0 = h->next_slot;
    if (s0 + n > h->cfg.max_batch) s0 = 0;
    for (int i = 0; i < CCM_MAX_TICKETS; i++)
        if (h->tk[i].active && !h->tk[i].empty && s0 < h->tk[i].slot0 + h->tk[i].n && h->tk[i].slot0 < s0 + n) {
            h->err = "frames in flight exceed max_batch: create the handle with max_batch >= the two batches together";
            return CCM_E_CAPACITY;
        }
    K.slot0 = s0;
    h->next_slot = s0 + n;
    // Level 0 lives in the handle's own level-0 buffer. When the caller's rows are 4-byte multiples that fit the
    // buffer's row budget the device pitch equals the host step and a frame is ONE contiguous copy; otherwise a
    // pitched 2-D copy. The call is cut into up to CCM_MAX_CHUNKS sub-batches: copy-in (s_in), kernels (lanes) and
    // copy-out (s_out) of different sub-batches overlap.
    uint8_t* l0 = (uint8_t*)h->allocs[0];
    const size_t budget = (size_t)align_up(cols, 32);
    const bool flat = (step % 4 == 0) && step <= budget;
    const size_t frame_bytes = step * (size_t)rows;
    // frames that follow each other in host memory (a [n][rows][step] array) keep that layout on the device, so a whole
    // sub-batch is ONE copy instead of one per frame
    bool contiguous = n > 1;
    for (int f = 1; contiguous && f < n; f++) contiguous = imgs[f] == imgs[f - 1] + frame_bytes;
    const bool packed = flat && contiguous;
    const size_t pitch = flat ? step : budget, stride = packed ? frame_bytes : budget * rows;
    if (!flat && h->stage_bytes < frame_bytes * h->cfg.max_batch) {
        if (in_flight) { h->err = "staging buffer must grow while a batch is in flight"; return CCM_E_BADARG; }
        CK(cudaStreamSynchronize(h->s_in));
        if (h->d_stage) cudaFree(h->d_stage);
        h->d_stage = nullptr; h->stage_bytes = 0;
        CK(cudaMalloc((void**)&h->d_stage, frame_bytes * h->cfg.max_batch));
        h->stage_bytes = frame_bytes * h->cfg.max_batch;
    }
    const int cap = h->max_kps;
    // sub-batch plan: first[c] .. first[c+1]. With another batch already on the device the copy-in of this one is hidden
    // behind that batch's kernels, so fewer, larger sub-batches (better kernel efficiency) are the better cut.
    int* first = K.first;
    int n_chunks = n <= 8 ? 1 : std::min(CCM_MAX_CHUNKS, std::min(in_flight ? 2 : 4, n / 4));
    static const char* const env_chunks = std::getenv("CCM_ORB_CHUNKS");
    static const char* const env_plan = std::getenv("CCM_ORB_CHUNK_PLAN");
    if (const char* e = env_chunks) n_chunks = std::max(1, std::min(std::min(CCM_MAX_CHUNKS, n), std::atoi(e)));   // tuning knob
    {
        const int per = (n + n_chunks - 1) / n_chunks;
        n_chunks = (n + per - 1) / per;
        for (int c = 0; c <= n_chunks; c++) first[c] = std::min(c * per, n);
    }
    if (const char* e = env_plan) {                                      // tuning knob: comma-separated sub-batch sizes
        int c = 0, at = 0;
        first[0] = 0;
        while (*e && c < CCM_MAX_CHUNKS - 1 && at < n) {
            const int v = std::atoi(e);
            if (v <= 0) break;
            at = std::min(n, at + v); first[++c] = at;
            while (*e && *e != ',') e++;
            if (*e == ',') e++;
        }
        if (at < n) first[++c] = n;
        n_chunks = c;
    }
    K.n_chunks = n_chunks;
    // results go straight into the caller's arrays when those are pinned and laid out like the arena's ([n][max_kps]);
    // otherwise through the handle's pinned staging and a host-side scatter in wait
    K.direct = capacity == cap && is_pinned(out_kp) && is_pinned(out_desc);
    static const bool trace = std::getenv("CCM_ORB_TRACE") != nullptr;
    K.trace = trace;
    if (trace) {
        for (auto& e : K.tev) if (!e) CK(cudaEventCreate(&e));
        CK(cudaEventRecord(K.tev[0], h->s_in));
    }
    auto enqueue = [&]() -> int {
    for (int c = 0; c < n_chunks; c++) {
        const int f0 = first[c], m = first[c + 1] - f0, g0 = s0 + f0;        // g0: first arena slot of this sub-batch
        if (packed) CK(cudaMemcpyAsync(l0 + g0 * stride, imgs[f0], frame_bytes * (size_t)(m - 1) + step * (size_t)(rows - 1) + cols, cudaMemcpyHostToDevice, h->s_in));
        else if (contiguous) {
            // rows that are not 4-byte multiples, frames adjacent in host memory: the sub-batch goes up as ONE copy into the
            // staging area and is re-pitched by ONE device-side 2-D copy (the frames are m * rows consecutive rows on both sides)
            uint8_t* st = h->d_stage + (size_t)g0 * frame_bytes;
            CK(cudaMemcpyAsync(st, imgs[f0], frame_bytes * (size_t)(m - 1) + step * (size_t)(rows - 1) + cols, cudaMemcpyHostToDevice, h->s_in));
            CK(cudaMemcpy2DAsync(l0 + g0 * stride, pitch, st, step, cols, (size_t)rows * m, cudaMemcpyDeviceToDevice, h->s_in));
        }
        for (int f = f0; f < f0 + m && !packed && !contiguous; f++) {
            if (flat) CK(cudaMemcpyAsync(l0 + (s0 + f) * stride, imgs[f], step * (size_t)(rows - 1) + cols, cudaMemcpyHostToDevice, h->s_in));
            else {
                // rows that are not 4-byte multiples: one contiguous host->device copy into a staging area, then a
                // device-side pitched copy (a pitched host->device copy runs at a fraction of the link rate)
                uint8_t* st = h->d_stage + (size_t)(s0 + f) * frame_bytes;
                CK(cudaMemcpyAsync(st, imgs[f], step * (size_t)(rows - 1) + cols, cudaMemcpyHostToDevice, h->s_in));
                CK(cudaMemcpy2DAsync(l0 + (s0 + f) * stride, pitch, st, step, cols, rows, cudaMemcpyDeviceToDevice, h->s_in));
            }
        }
        cudaEvent_t ev_in = h->ev_in[tk][c], ev_done = h->ev_done[tk][c], ev_out = h->ev_out[tk][c];
        CK(cudaEventRecord(ev_in, h->s_in));
        if (trace) CK(cudaEventRecord(K.tev[1 + 3 * c], h->s_in));
        const int ln = h->next_lane;
        h->next_lane ^= 1;
        CK(cudaStreamWaitEvent(h->lane[ln].s, ev_in, 0));
        rc = run_pipeline(h, m, l0, pitch, stride, h->d_kp + (size_t)g0 * cap, h->d_desc + (size_t)g0 * cap * 32, h->d_n + g0, cap, g0,
                          n_chunks == 1 && !in_flight, ln, !in_flight);
        if (rc != CCM_OK) return rc;
        CK(cudaEventRecord(ev_done, h->lane[ln].s));
        if (trace) CK(cudaEventRecord(K.tev[2 + 3 * c], h->lane[ln].s));
        CK(cudaStreamWaitEvent(h->s_out, ev_done, 0));
        CK(cudaMemcpyAsync(h->h_n + g0, h->d_n + g0, sizeof(int) * m, cudaMemcpyDeviceToHost, h->s_out));
        ccm_keypoint* dst_kp = K.direct ? out_kp + (size_t)f0 * cap : h->h_kp + (size_t)g0 * cap;
        uint8_t* dst_desc = K.direct ? out_desc + (size_t)f0 * cap * 32 : h->h_desc + (size_t)g0 * cap * 32;
        CK(cudaMemcpyAsync(dst_kp, h->d_kp + (size_t)g0 * cap, sizeof(ccm_keypoint) * (size_t)cap * m, cudaMemcpyDeviceToHost, h->s_out));
        CK(cudaMemcpyAsync(dst_desc, h->d_desc + (size_t)g0 * cap * 32, (size_t)32 * cap * m, cudaMemcpyDeviceToHost, h->s_out));
        CK(cudaEventRecord(ev_out, h->s_out));
        if (trace) CK(cudaEventRecord(K.tev[3 + 3 * c], h->s_out));
    }
    return CCM_OK;
    };
    rc = enqueue();
    if (rc != CCM_OK) {
        // part of the batch may be enqueued and still writing into the caller's arrays: nothing of it may be in flight
        // when the error returns, because the ticket stays inactive and cannot be waited on
        const std::string err = h->err;
        cudaStreamSynchronize(h->s_in);
        for (int i = 0; i < 2; i++) { cudaStreamSynchronize(h->lane[i].s); cudaStreamSynchronize(h->lane[i].aux); }
        cudaStreamSynchronize(h->s_out);
        cudaGetLastError();
        h->err = err;
        return rc;
    }
    K.t_enq = now_ms();
    K.active = true;
    *ticket = tk;
    return CCM_OK;
}

int ccm_orb_extract_batch_wait(ccm_orb* h, int ticket) {
    if (!h) return CCM_E_BADARG;
    if (ticket < 0 || ticket >= CCM_MAX_TICKETS || !h->tk[ticket].active) { h->err = "no such batch in flight"; return CCM_E_BADARG; }
    ccm_orb::Ticket& K = h->tk[ticket];
    if (K.empty) { for (int f = 0; f < K.n; f++) K.out_n[f] = 0; K.active = false; return CCM_OK; }
    CK(cudaSetDevice(h->cfg.device));
    const int cap = h->max_kps;
    const double t0 = K.trace ? now_ms() : 0.0;
    double t_wait = 0;
    for (int c = 0; c < K.n_chunks; c++) {
        const int f0 = K.first[c], m = K.first[c + 1] - f0;
        const double ta = K.trace ? now_ms() : 0.0;
        cudaError_t e = cudaEventSynchronize(h->ev_out[ticket][c]);
        if (e != cudaSuccess) { K.active = false; h->err = std::string("cudaEventSynchronize: ") + cudaGetErrorString(e); return CCM_E_CUDA; }
        if (K.trace) t_wait += now_ms() - ta;
        for (int f = f0; f < f0 + m; f++) {
            const int g = K.slot0 + f, cnt = h->h_n[g];
            K.out_n[f] = cnt;
            const int k = std::min(std::min(cnt, K.capacity), cap);
            if (k > 0 && !K.direct) {
                std::memcpy(K.out_kp + (size_t)f * K.capacity, h->h_kp + (size_t)g * cap, sizeof(ccm_keypoint) * k);
                std::memcpy(K.out_desc + (size_t)f * K.capacity * 32, h->h_desc + (size_t)g * cap * 32, (size_t)32 * k);
            }
        }
    }
    K.active = false;
    if (K.trace) {
        const double t_end = now_ms();
        fprintf(stderr, "[ccm_orb trace] ticket %d slots %d..%d%s: enqueue %.3f ms, submit->wait %.3f ms, wait %.3f ms, scatter %.3f ms | device (ms after first copy starts):",
                ticket, K.slot0, K.slot0 + K.n, K.direct ? " direct" : "", K.t_enq - K.t_call, t0 - K.t_enq, t_wait, t_end - t0 - t_wait);
        for (int c = 0; c < K.n_chunks; c++) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, K.tev[0], K.tev[1 + 3 * c]); cudaEventElapsedTime(&b, K.tev[0], K.tev[2 + 3 * c]);
            cudaEventElapsedTime(&d, K.tev[0], K.tev[3 + 3 * c]);
            fprintf(stderr, " [in %.3f done %.3f out %.3f]", a, b, d);
        }
        fprintf(stderr, "\n");
    }
    return CCM_OK;
}

int ccm_orb_extract_batch(ccm_orb* h, const uint8_t* const* imgs, int n, int rows, int cols, size_t step,
                          ccm_keypoint* out_kp, uint8_t* out_desc, int capacity, int* out_n) {
    int ticket = -1;
    const int rc = ccm_orb_extract_batch_submit(h, imgs, n, rows, cols, step, out_kp, out_desc, capacity, out_n, &ticket);
    if (rc != CCM_OK) return rc;
    return ccm_orb_extract_batch_wait(h, ticket);
}

// One frame, enqueued for capture into the single-frame graph: the frame fills ~7 % of the machine, so its latency is the
// length of the dependency chain, not the work. FAST and the quadtree of level l need nothing but level l: each level gets
// its own branch that starts when its resize has finished (level 0 at once), the blur is one more branch behind the last
// resize, and k_orient_describe joins them all and writes keypoints, descriptors and the count straight into the handle's
// pinned host mirrors (mapped into the device's address space: no device-to-host copy node at all).
static int run_pipeline_single_levels(ccm_orb* h, const uint8_t* l0, size_t l0_pitch) {
    const size_t l0_stride = l0_pitch * (size_t)h->rows;
    tl_pdl = true;
    h->T.l[0].img = const_cast<uint8_t*>(l0);
    h->T.l[0].img_pitch = (int)l0_pitch;
    h->T.l[0].img_stride = l0_stride;
    h->last_l0 = l0; h->last_l0_pitch = l0_pitch; h->last_l0_stride = l0_stride; h->last_frames = 1;
    const LevelTable& T = h->T;
    const int L = T.nlevels;
    cudaStream_t s = h->stream;
    for (int l = 0; l <= L; l++) {
        if (!h->lvl_s[l]) CK(cudaStreamCreateWithFlags(&h->lvl_s[l], cudaStreamNonBlocking));
        if (!h->lvl_ready[l]) CK(cudaEventCreateWithFlags(&h->lvl_ready[l], cudaEventDisableTiming));
        if (!h->lvl_done[l]) CK(cudaEventCreateWithFlags(&h->lvl_done[l], cudaEventDisableTiming));
    }
    FastTma tm = h->fc_tmaps;
    const bool tma = h->fc_tma && encode_level_map(h, &tm.m[0], l0, l0_pitch, T.l[0].h, l0_stride, 1, FC_TW, h->fc_smem.tile_rows);
    static const int qt_single = std::getenv("CCM_ORB_QT_SINGLE") ? std::atoi(std::getenv("CCM_ORB_QT_SINGLE")) : QT_THREADS_SINGLE;   // tuning knob
    CK(cudaMemsetAsync(h->d_corner_count, 0, sizeof(int) * CCM_MAX_LEVELS, s));
    for (int l = 0; l < L; l++) {
        if (l > 0) {
            dim3 g((T.l[l].w + 127) / 128, (T.l[l].h + 7) / 8, 1);
            if (T.l[l].resize_mlp) CKL("k_resize_level_mlp", launch_pdl(k_resize_level_mlp, dim3(g.x, (T.l[l].h + 8 * RESIZE_ROWS - 1) / (8 * RESIZE_ROWS), 1), dim3(32, 8), 0, s, T, l));
            else if (T.l[l].xgrp) CKL("k_resize_level", launch_pdl(k_resize_level, dim3(g.x, (T.l[l].h + 8 * RESIZE_ROWS - 1) / (8 * RESIZE_ROWS), 1), dim3(32, 8), 0, s, T, l));
            else { k_resize_level_generic<<<g, dim3(32, 8), 0, s>>>(T, l); CKL("k_resize_level_generic", cudaSuccess); }
            h->launches++;
        }
        CK(cudaEventRecord(h->lvl_ready[l], s));
        cudaStream_t x = h->lvl_s[l];
        CK(cudaStreamWaitEvent(x, h->lvl_ready[l], 0));
        const int c0 = h->chunk_off[l], nc = h->chunk_off[l + 1] - c0;
        if (nc > 0) {
            const uint4* items = h->d_chunks + 2 * (size_t)c0;
            if (tma) k_fast_chunks<true><<<dim3(nc, 1), dim3(32 * FC_WARPS), (size_t)h->fc_smem.total, x>>>(T, items, h->d_corners, h->d_corner_count, h->fc_smem, 0, tm);
            else k_fast_chunks<false><<<dim3(nc, 1), dim3(32 * FC_WARPS), (size_t)h->fc_smem.total, x>>>(T, items, h->d_corners, h->d_corner_count, h->fc_smem, 0, tm);
            CKL("k_fast_chunks (one level)", cudaSuccess);
            h->launches++;
        }
        k_quadtree<<<dim3(1, 1), dim3(qt_single), h->qt_smem, x>>>(T, h->qt_cap, h->qt_ccap, l, (const uint32_t*)h->d_corners, (const int*)h->d_corner_count,
                                                                 h->d_node_of, h->d_level_kps, h->d_level_kp_count);
        CKL("k_quadtree (one level)", cudaSuccess);
        h->launches++;
        CK(cudaEventRecord(h->lvl_done[l], x));
    }
    {
        cudaStream_t x = h->lvl_s[L];
        CK(cudaStreamWaitEvent(x, h->lvl_ready[L - 1], 0));
        if (h->blur_ocv24) k_blur_levels<true><<<dim3(T.total_blur_tiles, 1), 256, 0, x>>>(T); else k_blur_levels<false><<<dim3(T.total_blur_tiles, 1), 256, 0, x>>>(T);
        CKL("k_blur_levels", cudaSuccess);
        h->launches++;
        CK(cudaEventRecord(h->lvl_done[L], x));
    }
    for (int l = 0; l <= L; l++) CK(cudaStreamWaitEvent(s, h->lvl_done[l], 0));
    k_orient_describe<<<dim3(T.total_desc_blocks, 1), 256, 0, s>>>(T, h->d_pat, h->d_desc_block_level, h->d_level_kps, h->d_level_kp_count,
                                                                     (ccm_keypoint_dev*)h->h_kp, h->h_desc, h->h_n, h->max_kps);
    CKL("k_orient_describe", cudaSuccess);
    h->launches++;
    return CCM_OK;
}

// One frame, synchronously, as Frame::ExtractORB calls the extractor (cslam/src/Frame.cpp:120-123): copy-in, ONE graph launch
// (memset, 7 resize levels, FAST, quadtree next to the blur, orientation + descriptors, copy-out of slot 0 into the pinned
// staging), wait, hand the results over. The graph is captured from run_pipeline itself on first use, so it is the same
// kernel sequence with the same arguments; every pointer in it is an arena address, which only changes with the shape.
// Returns 1 when the fast path cannot be used (a batch in flight, capture unavailable): the caller takes the general path.
static int extract_single_graph(ccm_orb* h, const uint8_t* img, int rows, int cols, size_t step, ccm_keypoint* out_kp, uint8_t* out_desc,
                                int capacity, int* out_n) {
    static const bool off = std::getenv("CCM_ORB_NO_GRAPH") != nullptr;
    if (off || h->graph1_failed || h->stage_timing) return 1;
    for (int i = 0; i < CCM_MAX_TICKETS; i++) if (h->tk[i].active) return 1;
    if (capacity < 0 || (capacity > 0 && (!out_kp || !out_desc)) || step < (size_t)cols) { h->err = "bad argument"; return CCM_E_BADARG; }
    CK(cudaSetDevice(h->cfg.device));
    int rc = configure_shape(h, rows, cols);
    if (rc != CCM_OK) return rc;
    uint8_t* l0 = (uint8_t*)h->allocs[0];
    const size_t budget = (size_t)align_up(cols, 32);
    const bool flat = (step % 4 == 0) && step <= budget;
    const size_t pitch = flat ? step : budget;
    cudaStream_t s = h->stream;
    const int cap = h->max_kps;
    if (!h->graph1 || h->graph1_pitch != pitch) {
        if (h->graph1) { cudaGraphExecDestroy(h->graph1); h->graph1 = nullptr; }
        CK(cudaStreamSynchronize(s));
        const long long l_before = h->launches;
        cudaGraph_t g = nullptr;
        cudaError_t e = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        if (e == cudaSuccess) {
            static const bool flat_graph = std::getenv("CCM_ORB_GRAPH_FLAT") != nullptr;     // A/B knob: the batch pipeline's launch order + three copies
            if (flat_graph) {
                rc = run_pipeline(h, 1, l0, pitch, pitch * (size_t)rows, h->d_kp, h->d_desc, h->d_n, cap, 0, false, 0, true);
                if (rc == CCM_OK) {
                    cudaMemcpyAsync(h->h_n, h->d_n, sizeof(int), cudaMemcpyDeviceToHost, s);
                    cudaMemcpyAsync(h->h_kp, h->d_kp, sizeof(ccm_keypoint) * (size_t)cap, cudaMemcpyDeviceToHost, s);
                    cudaMemcpyAsync(h->h_desc, h->d_desc, (size_t)32 * cap, cudaMemcpyDeviceToHost, s);
                }
            } else {
                rc = run_pipeline_single_levels(h, l0, pitch);
            }
            e = cudaStreamEndCapture(s, &g);
            if (e == cudaSuccess && rc == CCM_OK) e = cudaGraphInstantiate(&h->graph1, g, 0);
            if (g) cudaGraphDestroy(g);
        }
        if (e != cudaSuccess || rc != CCM_OK || !h->graph1) {       // no graph on this driver / configuration: general path from now on
            cudaGetLastError();
            h->graph1 = nullptr; h->graph1_failed = true; h->launches = l_before;
            return 1;
        }
        h->graph1_pitch = pitch;
        h->graph1_launches = h->launches - l_before;
        h->launches = l_before;
    }
    if (flat) CK(cudaMemcpyAsync(l0, img, step * (size_t)(rows - 1) + cols, cudaMemcpyHostToDevice, s));
    else CK(cudaMemcpy2DAsync(l0, pitch, img, step, cols, rows, cudaMemcpyHostToDevice, s));
    CK(cudaGraphLaunch(h->graph1, s));
    h->launches += h->graph1_launches;
    h->last_l0 = l0; h->last_l0_pitch = pitch; h->last_l0_stride = pitch * (size_t)rows; h->last_frames = 1;
    h->T.l[0].img = l0; h->T.l[0].img_pitch = (int)pitch; h->T.l[0].img_stride = pitch * (size_t)rows;
    h->next_slot = 0; h->next_lane = 0;
    CK(cudaStreamSynchronize(s));
    const int cnt = h->h_n[0];
    *out_n = cnt;
    const int k = std::min(std::min(cnt, capacity), cap);
    if (k > 0) {
        std::memcpy(out_kp, h->h_kp, sizeof(ccm_keypoint) * (size_t)k);
        std::memcpy(out_desc, h->h_desc, (size_t)32 * k);
    }
    return CCM_OK;
}

int ccm_orb_extract(ccm_orb* h, const uint8_t* img, int rows, int cols, size_t step, ccm_keypoint* out_kp, uint8_t* out_desc,
                    int capacity, int* out_n) {
    if (!h) return CCM_E_BADARG;
    if (!out_n) { h->err = "bad argument"; return CCM_E_BADARG; }
    if (!img || rows == 0 || cols == 0) { *out_n = 0; return CCM_OK; }   // ORBextractor.cpp:1219-1220
    if (rows > 0 && cols > 0) {
        const int rc = extract_single_graph(h, img, rows, cols, step, out_kp, out_desc, capacity, out_n);
        if (rc != 1) return rc;
    }
    const uint8_t* one[1] = {img};
    return ccm_orb_extract_batch(h, one, 1, rows, cols, step, out_kp, out_desc, capacity, out_n);
}

}  // extern "C"

// ---- readback of intermediate stages of the last call ----
static int read_level(ccm_orb* h, int frame, int level, bool blurred, int border, uint8_t* out, size_t out_step, int* ow, int* oh) {
    if (!h) return CCM_E_BADARG;
    if (!out || frame < 0 || frame >= h->last_frames || level < 0 || level >= h->cfg.nlevels) { h->err = "bad argument"; return CCM_E_BADARG; }
    CK(cudaSetDevice(h->cfg.device));
    const LevelDesc& D = h->T.l[level];
    const int W = D.w + 2 * border, H = D.h + 2 * border;
    if (out_step < (size_t)W) { h->err = "out_step too small"; return CCM_E_CAPACITY; }
    std::vector<uint8_t> tmp((size_t)D.w * D.h);
    const uint8_t* src = blurred ? D.blur + (size_t)frame * D.blur_stride : D.img + (size_t)frame * D.img_stride;
    const size_t pitch = blurred ? D.blur_pitch : D.img_pitch;
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy2D(tmp.data(), D.w, src, pitch, D.w, D.h, cudaMemcpyDeviceToHost));
    auto refl = [](int p, int len) { if (len == 1) return 0; while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p; return p; };
    for (int y = 0; y < H; y++) {
        const uint8_t* s = &tmp[(size_t)refl(y - border, D.h) * D.w];
        uint8_t* d = out + (size_t)y * out_step;
        if (border == 0) std::memcpy(d, s, D.w);
        else for (int x = 0; x < W; x++) d[x] = s[refl(x - border, D.w)];
    }
    if (ow) *ow = W;
    if (oh) *oh = H;
    return CCM_OK;
}

extern "C" {

int ccm_orb_get_pyramid_level(ccm_orb* h, int frame, int level, int with_border, uint8_t* out, size_t out_step, int* ow, int* oh) {
    return read_level(h, frame, level, false, with_border ? CCM_EDGE : 0, out, out_step, ow, oh);
}
int ccm_orb_debug_blurred_level(ccm_orb* h, int frame, int level, uint8_t* out, size_t out_step, int* ow, int* oh) {
    return read_level(h, frame, level, true, 0, out, out_step, ow, oh);
}

int ccm_orb_debug_level_corners(ccm_orb* h, int frame, int level, int* xys, int capacity, int* out_n) {
    if (!h) return CCM_E_BADARG;
    if (!out_n || frame < 0 || frame >= h->last_frames || level < 0 || level >= h->cfg.nlevels || (capacity > 0 && !xys)) { h->err = "bad argument"; return CCM_E_BADARG; }
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    const LevelDesc& D = h->T.l[level];
    int n = 0;
    CK(cudaMemcpy(&n, h->d_corner_count + frame * CCM_MAX_LEVELS + level, sizeof(int), cudaMemcpyDeviceToHost));
    n = std::min(n, D.corner_cap);
    std::vector<uint32_t> pk(n);
    if (n) CK(cudaMemcpy(pk.data(), h->d_corners + (size_t)frame * h->T.corners_per_frame + D.corner_off, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
    // reference order: cells row-major, raster inside a cell (ORBextractor.cpp:957-997)
    auto ord = [&](uint32_t c) {
        const int x = (int)(c & 4095u), y = (int)((c >> 12) & 4095u);
        const int cj = (x - CCM_EDGE) / D.wCell, ci = (y - CCM_EDGE) / D.hCell;
        return (long long)(((ci * D.nCols + cj) * D.hCell + (y - CCM_EDGE - ci * D.hCell))) * D.wCell + (x - CCM_EDGE - cj * D.wCell);
    };
    std::sort(pk.begin(), pk.end(), [&](uint32_t a, uint32_t b) { return ord(a) < ord(b); });
    for (int i = 0; i < n && i < capacity; i++) { xys[3 * i] = (int)(pk[i] & 4095u); xys[3 * i + 1] = (int)((pk[i] >> 12) & 4095u); xys[3 * i + 2] = (int)(pk[i] >> 24); }
    *out_n = n;
    return CCM_OK;
}

}  // extern "C"

#ifdef CCM_DEBUG_TS
extern "C" int ccm_debug_ts(unsigned long long* out) {            // debug builds only: kernel start stamps, see ccm_common.cuh
    return cudaMemcpyFromSymbol(out, g_dbg_ts, sizeof(unsigned long long) * 128) == cudaSuccess ? 0 : -1;
}
#endif
