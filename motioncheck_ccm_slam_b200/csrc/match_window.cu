// K8a — windowed Hamming searches of ORBmatcher: every SearchByProjection overload, the SearchBySim3 passes, the
// distance phase of Fuse, and SearchForInitialization (cslam/src/ORBmatcher.cpp:71-148, 308-446, 448-563, 929-962,
// 1082-1103, 1216-1252, 1296-1327, 1350-1476), over the feature grid of Frame / KeyFrame
// (cslam/src/Frame.cpp:103-118, 200-265; cslam/src/KeyFrame.cpp:1162-1201).
//
// One CTA per job (one query list against one frame pack), one warp per query, lanes striding the candidates.
// The grid is CSR in (ix*rows + iy) cell order, so the cells of one grid column that a window touches are one
// contiguous item range: a query walks (window columns) ranges in exactly the order GetFeaturesInArea emits them.
// Candidates are ranked by the packed key (distance << 32 | position-in-that-order << 16 | feature index); the two
// smallest keys of a warp-shuffle reduction are the reference's best / second-best under its strict-`<` first-wins rule.
//
// The reference loops are greedy and sequential: an accepted match writes F.mvpMapPoints[idx] / vpMatched[idx] and a
// LATER query skips that feature. Query i's outcome depends only on queries j < i, so the sequential result is the
// unique fixpoint of "evaluate every query against the claims of lower-numbered queries"; the kernel iterates that
// map in parallel until the claim vector stops changing (after k rounds the first k queries are final, in practice
// 2-3 rounds because conflicts are rare). SearchForInitialization's steal-back rule (:487-513) compares distances of
// earlier claims and is run in query order by one warp instead (it is called a handful of times per session).
//
// Two kernels per search: k_window_eval walks the grid ONCE per query, fully parallel over (job, query) warps, under the
// filters that do not depend on other queries, and keeps the WIN_TOPK best candidate keys plus the candidate count;
// k_window_resolve (one CTA per job) runs the fixpoint on those keys alone, one thread per query. Only a query that
// finds too few unblocked keys among a truncated list re-walks the grid (exact, rare).
#include "match_common.cuh"

#define WIN_THREADS 1024
#define WIN_MODE_MAPPOINTS 0
#define WIN_MODE_LASTFRAME 1
#define WIN_MODE_KF_SIM3 2
#define WIN_MODE_BEST 3
#define WIN_MODE_INIT 4
#define WQ_VALID 1u
#define WQ_BLOCKS 2u
#define KEY_NONE 0xffffffffffffffffull
#define WIN_TOPK 8

// ---- frame pack construction ----------------------------------------------------------------------------------
// ccm_keypoint (AoS, extractor output) -> pack SoA, one CTA per pack; xy_un overrides the positions (mvKeysUn of a
// distorted camera). Pack j reads frame j of the extractor's batch output (kp_stride keypoints apart).
struct PackFill { float* x; float* y; float* angle; int* octave; int* n; int cap; };
__global__ void __launch_bounds__(256) k_pack_from_keypoints(const PackFill* __restrict__ fills, const ccm_keypoint_dev* __restrict__ kp,
                                                             const float* __restrict__ xy_un, const int* __restrict__ n_ptr, int kp_stride,
                                                             const CameraDev cam) {
    const PackFill F = fills[blockIdx.x];
    kp += (size_t)blockIdx.x * kp_stride;
    if (xy_un) xy_un += (size_t)blockIdx.x * kp_stride * 2;
    const int n = min(n_ptr[blockIdx.x], min(F.cap, kp_stride));
    if (threadIdx.x == 0) *F.n = n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const ccm_keypoint_dev k = kp[i];
        const float2 un = xy_un ? make_float2(xy_un[2 * i], xy_un[2 * i + 1]) : undistort_point(cam, k.x, k.y);   // mvKeysUn[i].pt
        F.x[i] = un.x;
        F.y[i] = un.y;
        F.angle[i] = k.angle;
        F.octave[i] = k.octave;
    }
}

// Frame::UndistortKeyPoints (Frame.cpp:277-307) for flat point arrays and for an extractor batch (frame = blockIdx.y)
__global__ void __launch_bounds__(256) k_undistort_points(const CameraDev cam, int n, const float2* __restrict__ in, float2* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const float2 p = in[i]; out[i] = undistort_point(cam, p.x, p.y); }
}
__global__ void __launch_bounds__(256) k_undistort_keypoints(const CameraDev cam, const ccm_keypoint_dev* __restrict__ kp, const int* __restrict__ n_ptr,
                                                             int kp_stride, float2* __restrict__ out) {
    const int f = blockIdx.y, n = min(n_ptr[f], kp_stride);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const ccm_keypoint_dev k = kp[(size_t)f * kp_stride + i]; out[(size_t)f * kp_stride + i] = undistort_point(cam, k.x, k.y); }
}

// Frame::AssignFeaturesToGrid + PosInGrid (Frame.cpp:103-118, 255-265; note round(), not floor()): one CTA per pack.
// Everything between reading x / y and writing the CSR happens in shared memory: a counting pass that also hands every
// feature its (cell, arrival rank), a block scan, a scatter without atomics, an insertion sort inside each cell that
// restores ascending feature index (what push_back in index order produces), and one coalesced store of items[].
// Dynamic shared memory: ncell ints | cap u32 (cell << 16 | rank) | cap u16 (items); packs too large for that
// (cap > 65536 or more than the launch reserved) take k_build_grid_global below.
struct GridJob { PackDev pack; int* cell_start; int* items; };
__device__ __forceinline__ int grid_cell_of(const PackDev& P, int i) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(P.x[i], P.min_x), P.inv_w));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(P.y[i], P.min_y), P.inv_h));
    if (px < 0 || px >= P.cols || py < 0 || py >= P.rows) return -1;
    return px * P.rows + py;
}
// exclusive block scan of cnt[0..ncell) in place (each thread owns a contiguous run of cells); returns the total to all threads
__device__ __forceinline__ int grid_scan(int* cnt, int ncell, int* cell_start, int* wsum) {
    const int per = (ncell + blockDim.x - 1) / blockDim.x;
    const int c0 = min(threadIdx.x * per, ncell), c1 = min(c0 + per, ncell);
    int s = 0;
    for (int c = c0; c < c1; c++) s += cnt[c];
    int incl = s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int v = wsum[lane], w = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
        wsum[lane] = w - v;
        if (lane == 31) wsum[32] = w;
    }
    __syncthreads();
    int run = wsum[wid] + incl - s;
    for (int c = c0; c < c1; c++) { const int k = cnt[c]; cell_start[c] = run; cnt[c] = run; run += k; }
    const int total = wsum[32];
    if (threadIdx.x == 0) cell_start[ncell] = total;
    return total;
}

__global__ void __launch_bounds__(1024) k_build_grid(const GridJob* __restrict__ jobs, int smem_cap) {
    const PackDev P = jobs[blockIdx.x].pack;
    int* cell_start = jobs[blockIdx.x].cell_start;
    int* items = jobs[blockIdx.x].items;
    extern __shared__ int g_sm[];
    __shared__ int wsum[33];
    const int ncell = P.cols * P.rows;
    const int n = min(*P.n, P.cap);
    int* cnt = g_sm;
    uint32_t* cr = reinterpret_cast<uint32_t*>(cnt + ncell);
    uint16_t* its = reinterpret_cast<uint16_t*>(cr + smem_cap);
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) cnt[c] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int cell = grid_cell_of(P, i);
        cr[i] = cell < 0 ? 0xffffffffu : ((uint32_t)cell << 16) | (uint32_t)atomicAdd(&cnt[cell], 1);
    }
    __syncthreads();
    const int total = grid_scan(cnt, ncell, cell_start, wsum);
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const uint32_t v = cr[i];
        if (v != 0xffffffffu) its[cnt[v >> 16] + (v & 0xffffu)] = (uint16_t)i;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
        const int a = cnt[c], b = c + 1 < ncell ? cnt[c + 1] : total;
        for (int i = a + 1; i < b; i++) {
            const uint16_t v = its[i];
            int j = i - 1;
            while (j >= a && its[j] > v) { its[j + 1] = its[j]; j--; }
            its[j + 1] = v;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += blockDim.x) items[i] = its[i];
}

// the same with items[] sorted in global memory: packs whose working set does not fit in shared memory
__global__ void __launch_bounds__(1024) k_build_grid_global(const GridJob* __restrict__ jobs) {
    const PackDev P = jobs[blockIdx.x].pack;
    int* cell_start = jobs[blockIdx.x].cell_start;
    int* items = jobs[blockIdx.x].items;
    __shared__ int cnt[CCM_MAX_GRID_CELLS];
    __shared__ int wsum[33];
    const int ncell = P.cols * P.rows;
    const int n = min(*P.n, P.cap);
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) cnt[c] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int cell = grid_cell_of(P, i);
        if (cell >= 0) atomicAdd(&cnt[cell], 1);
    }
    __syncthreads();
    grid_scan(cnt, ncell, cell_start, wsum);
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int cell = grid_cell_of(P, i);
        if (cell >= 0) items[atomicAdd(&cnt[cell], 1)] = i;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
        const int a = cell_start[c], b = cnt[c];   // cnt[c] is now the end of cell c
        for (int i = a + 1; i < b; i++) {
            const int v = items[i];
            int j = i - 1;
            while (j >= a && items[j] > v) { items[j + 1] = items[j]; j--; }
            items[j + 1] = v;
        }
    }
}

// ---- one query against the grid ---------------------------------------------------------------------------------
struct WinEval { unsigned long long b1, b2; };

// blocked(k): the feature already holds a map point that makes the reference `continue` (:113-115, :393, :1417-1419)
template <int MODE>
__device__ __forceinline__ WinEval eval_query(const PackDev& P, const WindowParamsDev& prm, const WindowQueryDev& q,
                                              const uint4 qd0, const uint4 qd1, int qi, const int* __restrict__ holder,
                                              const int* __restrict__ minclaim, const int* __restrict__ mdist, int lane) {
    WinEval e; e.b1 = KEY_NONE; e.b2 = KEY_NONE;
    // GetFeaturesInArea cell window (Frame.cpp:205-221)
    const float fx = __fsub_rn(q.u, P.min_x), fy = __fsub_rn(q.v, P.min_y);
    const int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(fx, q.r), P.inv_w)));
    if (minCX >= P.cols) return e;
    const int maxCX = min(P.cols - 1, (int)ceilf(__fmul_rn(__fadd_rn(fx, q.r), P.inv_w)));
    if (maxCX < 0) return e;
    const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(fy, q.r), P.inv_h)));
    if (minCY >= P.rows) return e;
    const int maxCY = min(P.rows - 1, (int)ceilf(__fmul_rn(__fadd_rn(fy, q.r), P.inv_h)));
    if (maxCY < 0) return e;
    const uint4* __restrict__ desc = reinterpret_cast<const uint4*>(P.desc);
    int pos_base = 0;
    for (int ix = minCX; ix <= maxCX; ix++) {
        const int s = P.cell_start[ix * P.rows + minCY], t = P.cell_start[ix * P.rows + maxCY + 1];
        for (int p = s + lane; p < t; p += 32) {
            const int k = P.items[p];
            const int oct = P.octave[k];
            if (oct < q.min_level) continue;
            if (q.max_level >= 0 && oct > q.max_level) continue;
            const float dx = __fsub_rn(P.x[k], q.u), dy = __fsub_rn(P.y[k], q.v);
            if (!(fabsf(dx) < q.r && fabsf(dy) < q.r)) continue;
            if (MODE == WIN_MODE_BEST && prm.use_gate) {       // Fuse chi2 gate, ORBmatcher.cpp:944-951
                const float ex = __fsub_rn(q.u, P.x[k]), ey = __fsub_rn(q.v, P.y[k]);
                const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                if ((double)__fmul_rn(e2, prm.inv_sigma2[oct]) > 5.99) continue;
            }
            if (MODE == WIN_MODE_MAPPOINTS || MODE == WIN_MODE_LASTFRAME) {
                if (holder[k] == -2 || minclaim[k] < qi) continue;
            } else if (MODE == WIN_MODE_KF_SIM3) {
                if (holder[k] != -1 || minclaim[k] < qi) continue;
            }
            const int dist = hamming256(qd0, qd1, desc[2 * (size_t)k], desc[2 * (size_t)k + 1]);
            if (MODE == WIN_MODE_INIT) { if (mdist[k] <= dist) continue; }   // vMatchedDistance, :487-488
            const unsigned long long key = ((unsigned long long)dist << 32) | ((unsigned long long)(pos_base + p - s) << 16) | (unsigned)k;
            push_top2(e.b1, e.b2, key);
        }
        pos_base += t - s;
    }
    warp_top2(e.b1, e.b2);
    return e;
}

// static pass: candidates under the query-independent filters (window, levels, gate, the holder state the call
// started with), best WIN_TOPK keys in keys[] (ascending, KEY_NONE padded) and the number of candidates that passed
// While a window is walked its keys are 32 bits: distance << 16 | CSR position of the candidate in the pack's grid (the
// position is swapped for the feature index when the list is stored). The CSR is in
// (ix * rows + iy) cell order with ascending feature indices inside a cell, i.e. positions inside one window ascend exactly in
// the order GetFeaturesInArea emits the features, so the position is both the reference's first-wins tie-break and (through
// items[]) the feature index. One REDUX per warp minimum instead of five 64-bit shuffle steps.
#define KEY32_NONE 0xffffffffu
__device__ __forceinline__ unsigned long long key32_to_key64(uint32_t k) {      // stored key (distance << 16 | feature) -> WinEval key
    return ((unsigned long long)(k >> 16) << 32) | (unsigned long long)(k & 0xffffu);
}

// G lanes per query (8, 16 or 32; the launch picks by the number of queries in flight): a window holds a handful of candidates, so thirty-two lanes were mostly idle and the
// kernel was bound by the number of resident warps times the length of the dependent load chain. The two halves run in
// lock step (loop trip counts are the maximum over the warp, shuffles are G wide), `valid` = this group has a query at all.
// Few queries in flight (one frame) are latency-bound instead and take G = 32.
template <int MODE, int G>
__device__ __forceinline__ void eval_query_topk(const PackDev& P, const WindowParamsDev& prm, const WindowQueryDev& q,
                                                const uint4 qd0, const uint4 qd1, const int* __restrict__ holder, int lane,
                                                bool valid, uint32_t& mine, int& total) {
    const unsigned FULL = 0xffffffffu;
    const int sub = lane & (G - 1);
    const unsigned gmask = (0xffffffffu >> (32 - G)) << (lane & (32 - G));
    uint32_t k0 = KEY32_NONE, k1 = KEY32_NONE, k2 = KEY32_NONE, k3 = KEY32_NONE;
    int cnt = 0;
    const float fx = __fsub_rn(q.u, P.min_x), fy = __fsub_rn(q.v, P.min_y);
    int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(fx, q.r), P.inv_w)));
    int maxCX = min(P.cols - 1, (int)ceilf(__fmul_rn(__fadd_rn(fx, q.r), P.inv_w)));
    const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(fy, q.r), P.inv_h)));
    const int maxCY = min(P.rows - 1, (int)ceilf(__fmul_rn(__fadd_rn(fy, q.r), P.inv_h)));
    if (!(valid && minCX < P.cols && maxCX >= 0 && minCY < P.rows && maxCY >= 0)) { minCX = 0; maxCX = -1; }
    const int ncol = max(0, maxCX - minCX + 1);
    int ncol_max = ncol;
#pragma unroll
    for (int o = G; o < 32; o <<= 1) ncol_max = max(ncol_max, __shfl_xor_sync(FULL, ncol_max, o));
    const uint4* __restrict__ desc = reinterpret_cast<const uint4*>(P.desc);
    // The window's grid columns are walked as ONE flattened candidate list (GetFeaturesInArea order: column by column,
    // cells top to bottom, features in index order): lane j fetches the CSR range of column cb + j, an inclusive scan gives
    // every column's first position, and the lanes then stride over all candidates of up to G columns at once. The
    // dependent load chain (range -> item -> fields -> descriptor) is paid once per G candidates, not once per column.
    for (int cb = 0; cb < ncol_max; cb += G) {
        const int ix = minCX + cb + sub;
        int s = 0, n = 0;
        if (ix <= maxCX) { s = P.cell_start[ix * P.rows + minCY]; n = P.cell_start[ix * P.rows + maxCY + 1] - s; }
        int inc = n;
#pragma unroll
        for (int o = 1; o < G; o <<= 1) { const int v = __shfl_up_sync(FULL, inc, o, G); if (sub >= o) inc += v; }
        const int tot = __shfl_sync(FULL, inc, G - 1, G);
        const int exc = inc - n;                           // position of this column's first candidate
        int tot_max = tot;
#pragma unroll
        for (int o = G; o < 32; o <<= 1) tot_max = max(tot_max, __shfl_xor_sync(FULL, tot_max, o));
        for (int c0 = 0; c0 < tot_max; c0 += G) {         // warp-uniform trip count: the shuffles below need all lanes
            const int c = c0 + sub;
            // column of candidate c = the first lane whose inclusive count exceeds c (binary search over the lanes by shuffle)
            int lo = 0;
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) { const int v = __shfl_sync(FULL, inc, lo + o - 1, G); if (v <= c) lo += o; }
            const int s_j = __shfl_sync(FULL, s, lo, G), exc_j = __shfl_sync(FULL, exc, lo, G);
            if (c >= tot) continue;
            const int p = s_j + (c - exc_j);
            const int k = P.items[p];
            const int oct = P.octave[k];
            if (oct < q.min_level) continue;
            if (q.max_level >= 0 && oct > q.max_level) continue;
            const float dx = __fsub_rn(P.x[k], q.u), dy = __fsub_rn(P.y[k], q.v);
            if (!(fabsf(dx) < q.r && fabsf(dy) < q.r)) continue;
            if (MODE == WIN_MODE_BEST && prm.use_gate) {
                const float ex = __fsub_rn(q.u, P.x[k]), ey = __fsub_rn(q.v, P.y[k]);
                const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                if ((double)__fmul_rn(e2, prm.inv_sigma2[oct]) > 5.99) continue;
            }
            if (MODE == WIN_MODE_MAPPOINTS || MODE == WIN_MODE_LASTFRAME) { if (holder[k] == -2) continue; }
            else if (MODE == WIN_MODE_KF_SIM3) { if (holder[k] != -1) continue; }
            const int dist = hamming256(qd0, qd1, desc[2 * (size_t)k], desc[2 * (size_t)k + 1]);
            uint32_t key = ((uint32_t)dist << 16) | (uint32_t)p;
            cnt++;
            if (key < k3) {
                k3 = key;
                if (k3 < k2) { const uint32_t t2 = k2; k2 = k3; k3 = t2; }
                if (k2 < k1) { const uint32_t t1 = k1; k1 = k2; k2 = t1; }
                if (k1 < k0) { const uint32_t t0 = k0; k0 = k1; k1 = t0; }
            }
        }
    }
    // A lane keeps its own four best. The query's WIN_TOPK best are popped off its lanes in order; once a lane that saw more
    // than four candidates has given all four away, its fifth (unknown) might be the next one, so the list ends there: the
    // keys stored are always an exact PREFIX of the window's ranking, KEY32_NONE after it, and the resolve pass compares
    // `total` with the number of keys it was given to know whether the list is complete.
    int left = min(cnt, 4);
    bool open = true;
#pragma unroll
    for (int r = 0; r < WIN_TOPK; r++) {
        uint32_t m = __reduce_min_sync(gmask, k0);         // the two halves reduce side by side, each over its own lanes
        if (!open) m = KEY32_NONE;
        if (m == KEY32_NONE) open = false;
        else if (k0 == m) { k0 = k1; k1 = k2; k2 = k3; k3 = KEY32_NONE; --left; }          // keys are unique: one lane pops
        if (sub == r) mine = m;                            // lane r of the half keeps the r-th best
        if (__ballot_sync(FULL, left == 0 && cnt > 4) & gmask) open = false;
        if (!__any_sync(FULL, open)) break;
    }
    // the resolve pass wants the feature, not its CSR position (the list order already carries the first-wins tie-break)
    if (mine != KEY32_NONE) mine = (mine & 0xffff0000u) | (uint32_t)P.items[mine & 0xffffu];
    total = __reduce_add_sync(gmask, cnt);
}

template <int MODE>
__device__ __forceinline__ bool accept_rule(const PackDev& P, const WindowParamsDev& prm, const WinEval& e, int& bestIdx,
                                            int& bestDist, int& bestDist2) {
    bestIdx = -1; bestDist = 256; bestDist2 = 256;
    if (e.b1 == KEY_NONE) return false;
    bestIdx = (int)(e.b1 & 0xffff); bestDist = (int)(e.b1 >> 32);
    const bool has2 = e.b2 != KEY_NONE;
    if (has2) bestDist2 = (int)(e.b2 >> 32);
    if (MODE == WIN_MODE_MAPPOINTS) {
        if (bestDist > prm.th_dist) return false;
        const int bestLevel = P.octave[bestIdx];
        const int bestLevel2 = has2 ? P.octave[(int)(e.b2 & 0xffff)] : -1;
        // `bestDist > mfNNratio*bestDist2` with int operands promoted to float (:139)
        if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(prm.nnratio, (float)bestDist2)) return false;
        return true;
    }
    if (MODE == WIN_MODE_INIT) {
        if (bestDist > prm.th_dist) return false;
        // `bestDist < (float)bestDist2*mfNNratio` (:498); with no second candidate bestDist2 is INT_MAX there
        if (!has2) return true;
        return (float)bestDist < __fmul_rn((float)bestDist2, prm.nnratio);
    }
    return bestDist <= prm.th_dist;
}

template <int MODE>
__device__ void window_job(const WindowJob& J, const WindowParamsDev& prm) {
    __shared__ int s_nfb[3], s_nm, s_hist[32], s_keep[3];
    const PackDev& P = J.pack;
    const int nq = J.nq_ptr ? min(*J.nq_ptr, J.nq) : J.nq;
    const int n = min(*P.n, P.cap);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const uint4* __restrict__ qdesc = reinterpret_cast<const uint4*>(J.q_desc);
    const bool stateful = (MODE != WIN_MODE_BEST);
    // Three "lowest blocking claimant per target" arrays rotate through the rounds: round `it` READS mc[it % 3] (built while
    // round it-1 ran, from that round's claims), BUILDS mc[(it+1) % 3] with an atomicMin per blocking claim as the claims are
    // decided, and CLEARS mc[(it+2) % 3] for the round after. A round is therefore one pass over the queries and ONE barrier
    // (it used to be clear / rebuild / evaluate with five); the iteration itself is unchanged: every query is evaluated against
    // the claims of the previous round (Jacobi), until no claim changes.
    // The three arrays live in shared memory when the host reserved it for this launch (3 * cap ints), else in the job's scratch.
    extern __shared__ int s_mc[];
    int* mcbase = prm.mc_smem_ints >= 3 * P.cap ? s_mc : J.head;
    int* mc[3] = {mcbase, mcbase + P.cap, mcbase + 2 * P.cap};
    // the keys of a thread's first query stay in registers over the rounds (every query when nq <= blockDim.x): 64-bit form,
    // i.e. the feature index is fetched from items[] once, not once per round
    uint32_t rk[WIN_TOPK];
    uint32_t rflags = 0; int rtotal = 0;
    if ((int)threadIdx.x < nq) {
        rflags = J.q[threadIdx.x].flags; rtotal = J.total[threadIdx.x];
#pragma unroll
        for (int r = 0; r < WIN_TOPK; r++) {
            const uint32_t k = J.keys[(size_t)WIN_TOPK * threadIdx.x + r];
            rk[r] = r < min(rtotal, WIN_TOPK) ? k : KEY32_NONE;
        }
    }

    if (threadIdx.x == 0) { s_nm = 0; s_nfb[0] = 0; s_nfb[1] = 0; s_nfb[2] = 0; }
    if (threadIdx.x < 32) s_hist[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < nq; i += blockDim.x) J.claim[0][i] = -1;           // claims of the previous round
    if (stateful)
        for (int i = threadIdx.x; i < n; i += blockDim.x) { mc[0][i] = 0x7fffffff; mc[1][i] = 0x7fffffff; }
    __syncthreads();

    for (int it = 0; it <= nq; it++) {
        const int* __restrict__ mcur = mc[it % 3];
        int* mnext = mc[(it + 1) % 3];
        int* mclear = mc[(it + 2) % 3];
        int* nfb = &s_nfb[it % 3];                           // this round's fallback count; next round's is reset now, the previous
        int changed = 0;                                     // round's may still be read by a slow thread and is left alone
        if (threadIdx.x == 0) s_nfb[(it + 1) % 3] = 0;
        if (stateful)
            for (int i = threadIdx.x; i < n; i += blockDim.x) mclear[i] = 0x7fffffff;
        // one thread per query, from the keys k_window_eval stored
        for (int i = threadIdx.x; i < nq; i += blockDim.x) {
            const bool first = i == (int)threadIdx.x;
            const uint32_t flags = first ? rflags : J.q[i].flags;
            WinEval e; e.b1 = KEY_NONE; e.b2 = KEY_NONE;
            bool fallback = false;
            if (flags & WQ_VALID) {
                const int total = first ? rtotal : J.total[i];
                const int have = min(total, WIN_TOPK);
                int found = 0, given = 0;                  // given: keys seen (counts the whole list whenever found < need)
#pragma unroll
                for (int r = 0; r < WIN_TOPK; r++) {
                    if (r >= have || found == 2) break;
                    const uint32_t k32 = first ? rk[r] : J.keys[(size_t)WIN_TOPK * i + r];
                    if (k32 == KEY32_NONE) break;
                    const unsigned long long key = key32_to_key64(k32);
                    ++given;
                    if (stateful && mcur[(int)(key & 0xffff)] < i) continue;
                    if (found == 0) e.b1 = key; else { e.b2 = key; }
                    ++found;
                }
                const int need = (MODE == WIN_MODE_MAPPOINTS) ? 2 : 1;
                fallback = found < need && total > given;
            }
            if (fallback) { J.fb[atomicAdd(nfb, 1)] = i; continue; }
            int bestIdx, bestDist, bestDist2;
            const bool acc = accept_rule<MODE>(P, prm, e, bestIdx, bestDist, bestDist2);
            const int c = acc ? bestIdx : -1;
            if (c != J.claim[0][i]) { changed = 1; J.claim[0][i] = c; }
            if (stateful && c >= 0 && (flags & WQ_BLOCKS)) atomicMin(&mnext[c], i);
            J.qbest[3 * i] = bestIdx; J.qbest[3 * i + 1] = bestDist;
            J.qbest[3 * i + 2] = (MODE == WIN_MODE_MAPPOINTS) ? bestDist2 : 256;
        }
        int any = __syncthreads_or(changed);               // the round's barrier: claims, mnext and the fallback list are complete
        // truncated key lists that ran dry: exact re-walk of the grid, one warp per query, against the same mcur (rare)
        const int cnt = *nfb;
        if (cnt > 0) {
            changed = 0;
            for (int j = wid; j < cnt; j += nw) {
                const int i = J.fb[j];
                const WindowQueryDev q = J.q[i];
                int bestIdx = -1, bestDist = 256, bestDist2 = 256;
                const WinEval e = eval_query<MODE>(P, prm, q, qdesc[2 * (size_t)i], qdesc[2 * (size_t)i + 1], i, J.holder, mcur, nullptr, lane);
                const bool acc = accept_rule<MODE>(P, prm, e, bestIdx, bestDist, bestDist2);
                if (lane == 0) {
                    const int c = acc ? bestIdx : -1;
                    if (c != J.claim[0][i]) { changed = 1; J.claim[0][i] = c; }
                    if (stateful && c >= 0 && (q.flags & WQ_BLOCKS)) atomicMin(&mnext[c], i);
                    J.qbest[3 * i] = bestIdx; J.qbest[3 * i + 1] = bestDist;
                    J.qbest[3 * i + 2] = (MODE == WIN_MODE_MAPPOINTS) ? bestDist2 : 256;
                }
            }
            any |= __syncthreads_or(changed);
        }
#ifdef CCM_DEBUG_ROUNDS
        if (threadIdx.x == 0 && blockIdx.x < 3) printf("mode %d job %d round %d nfb %d any %d nq %d n %d\n", MODE, blockIdx.x, it, cnt, any, nq, n);
#endif
        if (!stateful || !any) break;
    }
    __syncthreads();
    const int cur = 0;

    // ---- outcome: res[], holder[], nmatches, rotation histogram (LASTFRAME, :1436-1465) ----
    const bool ori = (MODE == WIN_MODE_LASTFRAME) && prm.check_orientation;
    int my_nm = 0;
    for (int i = threadIdx.x; i < nq; i += blockDim.x) {
        const int t = J.claim[cur][i];
        if (J.res) {
            MatchResultDev r; r.best_idx = J.qbest[3 * i]; r.best_dist = J.qbest[3 * i + 1]; r.best_dist2 = J.qbest[3 * i + 2];
            r.accepted = t >= 0;
            J.res[i] = r;
        }
        if (t < 0) continue;
        const bool writes = (MODE == WIN_MODE_MAPPOINTS || MODE == WIN_MODE_LASTFRAME) ||
                            (MODE == WIN_MODE_KF_SIM3 && (J.q[i].flags & WQ_BLOCKS));
        if (writes) atomicMax(&J.holder[t], i);       // later queries overwrite earlier non-blocking claims
        if (MODE != WIN_MODE_KF_SIM3 || writes) my_nm++;
        if (ori) atomicAdd(&s_hist[rot_bin(J.q[i].angle, P.angle[t])], 1);
    }
    if (my_nm) atomicAdd(&s_nm, my_nm);
    __syncthreads();
    if (ori) {
        if (threadIdx.x == 0) three_maxima(s_hist, 30, s_keep[0], s_keep[1], s_keep[2]);
        __syncthreads();
        int drop = 0;
        for (int i = threadIdx.x; i < nq; i += blockDim.x) {
            const int t = J.claim[cur][i];
            if (t < 0) continue;
            const int b = rot_bin(J.q[i].angle, P.angle[t]);
            if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) { J.holder[t] = -1; drop++; }
        }
        if (drop) atomicSub(&s_nm, drop);
        __syncthreads();
    }
    if (threadIdx.x == 0) *J.nmatches = s_nm;
}

// SearchForInitialization (ORBmatcher.cpp:448-563): holder = vnMatches21 (output), head = vMatchedDistance, claim[0] =
// vnMatches12, claim[1] = rotation bin of the accept event. The steal-back rule (:487-513) makes a query's candidate set
// depend on the DISTANCES of earlier accepts (`if(vMatchedDistance[i2]<=dist) continue`), so the queries are resolved in
// order by one warp; but the distances themselves do not depend on that state, so k_window_eval has already walked every
// window in parallel and left each query's WIN_TOPK best keys. The ordered pass only filters those keys against
// vMatchedDistance (lanes 0..3 = the four keys; 32 queries' keys are staged in shared memory at a time so that the loop never
// waits on global memory) and re-walks the window exactly only when fewer than two keys of a truncated list survive.
__device__ void window_job_init(const WindowJob& J, const WindowParamsDev& prm) {
    __shared__ int s_hist[32], s_keep[3], s_nm;
    __shared__ uint32_t s_key[32][WIN_TOPK];
    __shared__ int s_idx[32][WIN_TOPK], s_total[32];
    __shared__ uint32_t s_flags[32];
    const PackDev& P = J.pack;
    const int nq = J.nq_ptr ? min(*J.nq_ptr, J.nq) : J.nq;
    const int n = min(*P.n, P.cap);
    const int lane = threadIdx.x & 31;
    const uint4* __restrict__ qdesc = reinterpret_cast<const uint4*>(J.q_desc);
    int* mdist = J.head;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { mdist[i] = 0x7fffffff; J.holder[i] = -1; }
    for (int i = threadIdx.x; i < nq; i += blockDim.x) { J.claim[0][i] = -1; J.claim[1][i] = -1; }
    if (threadIdx.x < 32) s_hist[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_nm = 0;
    __syncthreads();
    if (threadIdx.x < 32) {
        int nm = 0;
        for (int i0 = 0; i0 < nq; i0 += 32) {
            {   // stage the keys of queries i0 .. i0+31: lane = query
                const int i = i0 + lane;
                uint32_t fl = 0; int tot = 0;
                if (i < nq) { fl = J.q[i].flags; tot = J.total[i]; }
#pragma unroll
                for (int r = 0; r < WIN_TOPK; r++) {
                    const uint32_t k = i < nq ? J.keys[(size_t)WIN_TOPK * i + r] : KEY32_NONE;
                    s_key[lane][r] = k;
                    s_idx[lane][r] = k != KEY32_NONE ? (int)(k & 0xffffu) : 0;
                }
                s_flags[lane] = fl; s_total[lane] = tot;
            }
            __syncwarp();
            const int jn = min(32, nq - i0);
            for (int j = 0; j < jn; j++) {
                const int i = i0 + j;
                int bestIdx = -1, bestDist = 256, bestDist2 = 256;
                bool acc = false;
                if (s_flags[j] & WQ_VALID) {
                    const int total = s_total[j];
                    uint32_t key = KEY32_NONE; int idx = 0;
                    if (lane < WIN_TOPK) { key = s_key[j][lane]; idx = s_idx[j][lane]; }
                    const int dist = (int)(key >> 16);
                    const bool ok = key != KEY32_NONE && mdist[idx] > dist;              // vMatchedDistance, :487-488
                    const uint32_t b = __ballot_sync(0xffffffffu, ok);
                    const int given = __popc(__ballot_sync(0xffffffffu, key != KEY32_NONE));
                    if (__popc(b) < 2 && total > given) {                              // truncated list ran dry: exact re-walk
                        const WinEval e = eval_query<WIN_MODE_INIT>(P, prm, J.q[i], qdesc[2 * (size_t)i], qdesc[2 * (size_t)i + 1], i,
                                                                    J.holder, nullptr, mdist, lane);
                        acc = accept_rule<WIN_MODE_INIT>(P, prm, e, bestIdx, bestDist, bestDist2);
                    } else if (b) {
                        const int first = __ffs(b) - 1;
                        const uint32_t rest = b & (b - 1);
                        WinEval e;
                        const unsigned long long k64 = ((unsigned long long)dist << 32) | ((unsigned long long)(key & 0xffffu) << 16) | (unsigned)idx;
                        e.b1 = __shfl_sync(0xffffffffu, k64, first);
                        e.b2 = rest ? __shfl_sync(0xffffffffu, k64, __ffs(rest) - 1) : KEY_NONE;
                        acc = accept_rule<WIN_MODE_INIT>(P, prm, e, bestIdx, bestDist, bestDist2);
                    }
                }
                if (lane == 0) {
                    J.qbest[3 * i] = bestIdx; J.qbest[3 * i + 1] = bestDist; J.qbest[3 * i + 2] = bestDist2;
                    if (acc) {
                        const int old = J.holder[bestIdx];
                        if (old >= 0) { J.claim[0][old] = -1; nm--; }          // steal-back, :506-510
                        J.claim[0][i] = bestIdx; J.holder[bestIdx] = i; mdist[bestIdx] = bestDist; nm++;
                        if (prm.check_orientation) {
                            const int bb = rot_bin(J.q[i].angle, P.angle[bestIdx]);
                            J.claim[1][i] = bb; s_hist[bb]++;
                        }
                    }
                }
                __syncwarp();
            }
        }
        if (lane == 0) s_nm = nm;
    }
    __syncthreads();
    if (prm.check_orientation) {
        if (threadIdx.x == 0) three_maxima(s_hist, 30, s_keep[0], s_keep[1], s_keep[2]);
        __syncthreads();
        int drop = 0;
        for (int i = threadIdx.x; i < nq; i += blockDim.x) {
            const int b = J.claim[1][i], t = J.claim[0][i];
            if (b < 0 || t < 0) continue;                                   // `if(vnMatches12[idx1]>=0)`, :548
            if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) { J.claim[0][i] = -1; J.holder[t] = -1; drop++; }
        }
        if (drop) atomicSub(&s_nm, drop);
        __syncthreads();
    }
    for (int i = threadIdx.x; i < nq; i += blockDim.x)
        if (J.res) {
            MatchResultDev r; r.best_idx = J.qbest[3 * i]; r.best_dist = J.qbest[3 * i + 1]; r.best_dist2 = J.qbest[3 * i + 2];
            r.accepted = J.claim[0][i] >= 0;
            J.res[i] = r;
        }
    if (threadIdx.x == 0) *J.nmatches = s_nm;
}

template <int MODE, int G>
__device__ __forceinline__ void window_eval_job(const WindowJob& J, const WindowParamsDev& prm, int i, int nq, int lane) {
    const bool have = i < nq;
    WindowQueryDev q = J.q[have ? i : 0];
    const bool valid = have && (q.flags & WQ_VALID);
    const uint4* __restrict__ qdesc = reinterpret_cast<const uint4*>(J.q_desc);
    const size_t iq = have ? i : 0;
    uint32_t mine = KEY32_NONE;
    int total = 0;
    eval_query_topk<MODE, G>(J.pack, prm, q, qdesc[2 * iq], qdesc[2 * iq + 1], J.holder, lane, valid, mine, total);
    if (!have) return;
    const int sub = lane & (G - 1);
    if (sub < WIN_TOPK) J.keys[(size_t)WIN_TOPK * i + sub] = mine;
    if (sub == 0) J.total[i] = total;
}

// grid (query groups of 256 / G, jobs); G lanes = one query of one job (CCM_WIN_INIT takes the filter-free instantiation).
#ifndef WIN_EVAL_MIN_CTAS
#define WIN_EVAL_MIN_CTAS 6
#endif
template <int G>
__global__ void __launch_bounds__(256, WIN_EVAL_MIN_CTAS) k_window_eval(const WindowJob* __restrict__ jobs, WindowParamsDev prm) {
    const WindowJob& J = jobs[blockIdx.y];
    const int nq = J.nq_ptr ? min(*J.nq_ptr, J.nq) : J.nq;
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (256 / G) + threadIdx.x / G;
    if (i - lane / G >= nq) return;                        // the warp's first query is past the end: the whole warp leaves
    switch (prm.mode) {
        case WIN_MODE_MAPPOINTS: window_eval_job<WIN_MODE_MAPPOINTS, G>(J, prm, i, nq, lane); break;
        case WIN_MODE_LASTFRAME: window_eval_job<WIN_MODE_LASTFRAME, G>(J, prm, i, nq, lane); break;
        case WIN_MODE_KF_SIM3: window_eval_job<WIN_MODE_KF_SIM3, G>(J, prm, i, nq, lane); break;
        default: window_eval_job<WIN_MODE_BEST, G>(J, prm, i, nq, lane); break;
    }
}

__global__ void __launch_bounds__(WIN_THREADS) k_window_resolve(const WindowJob* __restrict__ jobs, WindowParamsDev prm) {
    const WindowJob J = jobs[blockIdx.x];
    switch (prm.mode) {
        case WIN_MODE_MAPPOINTS: window_job<WIN_MODE_MAPPOINTS>(J, prm); break;
        case WIN_MODE_LASTFRAME: window_job<WIN_MODE_LASTFRAME>(J, prm); break;
        case WIN_MODE_KF_SIM3: window_job<WIN_MODE_KF_SIM3>(J, prm); break;
        case WIN_MODE_BEST: window_job<WIN_MODE_BEST>(J, prm); break;
        default: window_job_init(J, prm); break;
    }
}
