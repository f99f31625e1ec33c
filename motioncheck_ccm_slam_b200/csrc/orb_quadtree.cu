// K4: quadtree keypoint distribution, one CTA per (frame, level). Replaces ORBextractor::DistributeOctTree
// (cslam/src/ORBextractor.cpp:707-931) and ExtractorNode::DivideNode (:650-705).
//
// The reference walks a std::list with push_front/erase; here the list is an array indexed by list position
// (front = 0) that is rebuilt level-synchronously each round:
//   new list = reverse(children of the nodes split this round, in split order, n1..n4)  ++  untouched nodes in old order
// which is exactly what repeated push_front + erase produces. Every corner carries the list position of its node.
// Round kinds:
//   "sweep"  (:767-830)  every node holding >1 corner is split, in list order;
//   "greedy" (:841-905)  entered when size + 3*nToExpand > N: nodes holding >1 corner are split largest first
//                        until the list reaches N.  Ties between equal sizes: the reference compares heap
//                        addresses (:852); the rule fixed here is "later-created first" == smaller list position
//                        first, i.e. the reference under a monotone allocator (quadtree tie rule).
// Final pick (:910-928): per node the corner with the largest response, first in vToDistributeKeys order on
// ties; that order (cells row-major, raster inside a cell) is recomputed from the coordinates, so the corner
// list itself may be in any order.
#include "ccm_common.cuh"

namespace {

struct QNode { short ulx, uly, brx, bry; int cnt; };

// exclusive prefix sum of v[0..n) in place (block-wide, any multiple of 32 threads up to 1024); returns the total. scratch: 32 ints.
// Each thread owns a contiguous run of ceil(n / threads) elements; the per-thread sums are scanned inside every warp
// at once and the warp totals by every warp redundantly, so the call costs two barriers and ten shuffle steps.
#define QT_THREADS 256               // batch launch: CTA = (level, frame); 0.0399 / 0.0364 / 0.0355 ms per 64 frames at 512 / 384 / 256
#define QT_THREADS_SINGLE 384        // one frame, one launch per level: latency, not throughput
__device__ int block_exclusive_scan(int* v, int n, int* scratch) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = blockDim.x, NW = NT >> 5;
    const int per = (n + NT - 1) / NT;
    const int b = min(tid * per, n), e = min(b + per, n);
    int s = 0;
    for (int i = b; i < e; i++) s += v[i];
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
    if (lane == 31) scratch[warp] = inc;
    __syncthreads();
    const int wt = lane < NW ? scratch[lane] : 0;
    int winc = wt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += y; }
    const int total = __shfl_sync(0xffffffffu, winc, 31);
    int run = __shfl_sync(0xffffffffu, winc - wt, warp) + inc - s;
    for (int i = b; i < e; i++) { const int x = v[i]; v[i] = run; run += x; }
    __syncthreads();
    return total;
}

__device__ __forceinline__ void child_bounds(const QNode& p, int q, QNode& c) {
    const int mx = p.ulx + ((p.brx - p.ulx + 1) >> 1);   // ceil(float(UR.x-UL.x)/2), :652
    const int my = p.uly + ((p.bry - p.uly + 1) >> 1);
    c.ulx = (q & 1) ? mx : p.ulx;  c.brx = (q & 1) ? p.brx : mx;
    c.uly = (q & 2) ? my : p.uly;  c.bry = (q & 2) ? p.bry : my;
}
__device__ __forceinline__ int quadrant(const QNode& p, int x, int y) {
    const int mx = p.ulx + ((p.brx - p.ulx + 1) >> 1);
    const int my = p.uly + ((p.bry - p.uly + 1) >> 1);
    return (x < mx ? 0 : 1) | (y < my ? 0 : 2);           // n1=0 n2=1 n3=2 n4=3, :680-695
}

}  // namespace

__device__ __forceinline__ int n_children(const int* cc, int i) {
    return (cc[4 * i] > 0) + (cc[4 * i + 1] > 0) + (cc[4 * i + 2] > 0) + (cc[4 * i + 3] > 0);
}
// One node of the old list goes into the new list B: a node that is split leaves its non-empty children (creation index
// `created`.., push_front order: later-created sits closer to the front, so child c lands on C - 1 - c), any other node
// keeps its place behind the C new children, moved up by the `before` split nodes in front of it.
__device__ __forceinline__ void place_node(const QNode* A, QNode* B, const int* cc, int* newpos, short* childpos, int* s_expand,
                                           int i, bool split, int created, int before, int C) {
    const QNode nd = A[i];
    if (split) {
        int expand = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int cnt = cc[4 * i + q];
            if (cnt > 0) {
                const int pos = C - 1 - created;
                QNode c; child_bounds(nd, q, c); c.cnt = cnt;
                B[pos] = c;
                childpos[4 * i + q] = (short)pos;
                created++;
                expand += cnt > 1;
            }
        }
        newpos[i] = -1;
        if (expand) atomicAdd(s_expand, expand);
    } else {
        const int pos = C + i - before;
        B[pos] = nd;
        newpos[i] = pos;
    }
}

// dynamic smem layout (cap = node capacity rounded up to a power of two, >= 32):
//   QNode A[cap], B[cap]; int cc0[4*cap], cc1[4*cap]; int key[cap]; int aux[cap]; int aux2[cap]; int newpos[cap];
//   short childpos[4*cap]; int scratch[32]; uint32_t corner_cache[ccap]; uint16_t node_cache[ccap]     (80 B per node)
// A (frame, level) with at most ccap corners keeps its corner list and the corner -> node map in shared memory for the
// whole run (every round walks both once); larger ones work on the global arrays.
// A round is few barriers on purpose (one frame's latency is this kernel's barrier count): the node lists A / B and the
// child-occupancy tables ping-pong; the pass that moves every corner to its new node also counts the occupancy of that
// node's children for the next round; in a sweep round the creation index of a node's children and the number of split
// nodes in front of it come out of ONE scan of a packed value. Sweep round: 4 barriers, greedy round: 10 (was 14 / 20).
__global__ void __launch_bounds__(512, 4) k_quadtree(const __grid_constant__ LevelTable T, int cap, int ccap, int lvl0,   // CTA x handles level lvl0 + x
                                                  const uint32_t* corners, const int* corner_count,   // written by the PDL predecessor: no __restrict__, so these are coherent loads the compiler cannot hoist above pdl_wait()
                                                  uint16_t* __restrict__ node_of,          // [frames][corners_per_frame]
                                                  uint32_t* __restrict__ level_kps,        // [frames][kps_per_frame] packed x|y<<12|score<<24
                                                  int* __restrict__ level_kp_count) {      // [frames][CCM_MAX_LEVELS]
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lvl = lvl0 + blockIdx.x, f = blockIdx.y;
    const LevelDesc& L = T.l[lvl];
    const int tid = threadIdx.x, NT = blockDim.x;
    pdl_launch();
    pdl_wait();                              // the corner lists of k_fast_chunks are complete
    CCM_TS(blockIdx.y == 0 && threadIdx.x == 0, 32 + lvl);

    QNode* A = reinterpret_cast<QNode*>(smem_raw);
    QNode* B = A + cap;
    int* cc = reinterpret_cast<int*>(B + cap);          // child occupancy of the current list
    int* cc_next = cc + 4 * cap;                        // ... of the list being built
    int* key = cc_next + 4 * cap;
    int* aux = key + cap;
    int* aux2 = aux + cap;
    int* newpos = aux2 + cap;
    short* childpos = reinterpret_cast<short*>(newpos + cap);
    int* scratch = reinterpret_cast<int*>(childpos + 4 * cap);
    __shared__ int s_n, s_expand, s_P, s_C;

    const int nC = min(corner_count[f * CCM_MAX_LEVELS + lvl], L.corner_cap);
    const uint32_t* ck = corners + (size_t)f * T.corners_per_frame + L.corner_off;
    uint16_t* nof = node_of + (size_t)f * T.corners_per_frame + L.corner_off;
    if (nC <= ccap) {
        uint32_t* sck = reinterpret_cast<uint32_t*>(scratch + 32);
        for (int k = tid; k < nC; k += NT) sck[k] = ck[k];
        ck = sck;
        nof = reinterpret_cast<uint16_t*>(sck + ccap);
    }
    const int N = L.quota;
    const int H = L.h - 2 * CCM_MIN_BORDER;   // maxY-minY

    // ---- initial nodes (:711-741) ----
    if (tid < L.nIni) {
        QNode nd;
        nd.ulx = (short)(int)(L.hX * (float)tid);
        nd.brx = (short)(int)(L.hX * (float)(tid + 1));
        nd.uly = 0; nd.bry = (short)H; nd.cnt = 0;
        A[tid] = nd;
    }
    for (int i = tid; i < 4 * L.nIni; i += NT) cc_next[i] = 0;
    __syncthreads();
    for (int k = tid; k < nC; k += NT) {
        const uint32_t c = ck[k];
        const int x = (int)(c & 4095u) - CCM_MIN_BORDER;
        const int nd = (int)(__fdiv_rn((float)x, L.hX));
        nof[k] = (uint16_t)nd;
        atomicAdd(&A[nd].cnt, 1);
    }
    __syncthreads();
    // drop empty initial nodes (:743-757), keeping order
    if (tid == 0) {
        int n = 0;
        for (int i = 0; i < L.nIni; i++) { newpos[i] = n; if (A[i].cnt > 0) { B[n] = A[i]; n++; } }
        s_n = n;
    }
    __syncthreads();
    int n = s_n;
    for (int k = tid; k < nC; k += NT) {
        const int np = newpos[nof[k]];
        nof[k] = (uint16_t)np;
        const QNode nn = B[np];
        if (nn.cnt > 1) {
            const uint32_t c = ck[k];
            atomicAdd(&cc_next[4 * np + quadrant(nn, (int)(c & 4095u) - CCM_MIN_BORDER, (int)((c >> 12) & 4095u) - CCM_MIN_BORDER)], 1);
        }
    }
    __syncthreads();
    { QNode* t = A; A = B; B = t; int* u = cc; cc = cc_next; cc_next = u; }

    bool greedy = false;
#pragma unroll 1
    while (true) {
        int P, C;
        if (!greedy) {
            // sweep (:767-830): every node holding more than one corner is split, in list order. One scan of
            // (children | 1 << 16) per split node gives a node both the creation index of its first child and the number
            // of split nodes in front of it.
            if (tid == 0) s_expand = 0;
            for (int i = tid; i < n; i += NT) aux[i] = A[i].cnt > 1 ? (n_children(cc, i) | 0x10000) : 0;
            __syncthreads();
            const int tot = block_exclusive_scan(aux, n, scratch);
            C = tot & 0xffff; P = tot >> 16;
            for (int i = tid; i < 4 * (C + n - P); i += NT) cc_next[i] = 0;
            for (int i = tid; i < n; i += NT)
                place_node(A, B, cc, newpos, childpos, &s_expand, i, A[i].cnt > 1, aux[i] & 0xffff, aux[i] >> 16, C);
        } else {
            // greedy (:841-905): candidates sorted descending by (count, earlier list position first). Keys are distinct
            // (the position is part of the key), so a candidate's place is the number of larger keys: one pass of broadcast
            // reads, no barrier ladder
            if (tid == 0) { s_expand = 0; s_P = 0; s_C = 0; }
            for (int i = tid; i < n; i += NT) {
                key[i] = A[i].cnt > 1 ? ((min(A[i].cnt, 0x7FFFF) << 12) | (4095 - i)) : 0;
                newpos[i] = 0;                                // sorted keys
                aux2[i] = 0;                                  // 1 = split this round
            }
            __syncthreads();
            for (int i = tid; i < n; i += NT) {
                const int k = key[i];
                if (k) {
                    int r = 0;
                    for (int j = 0; j < n; j++) r += key[j] > k;
                    newpos[r] = k;
                }
            }
            __syncthreads();
            // aux[r] = growth of the list when the r-th candidate is split (children - 1); prefix -> cut (:895-896)
            for (int r = tid; r < n; r += NT) {
                const int k = newpos[r];
                aux[r] = k ? n_children(cc, 4095 - (k & 4095)) - 1 : 0;
                key[r] = -1;                                  // from here: creation index of a split node's first child
            }
            __syncthreads();
            block_exclusive_scan(aux, n, scratch);            // aux[r] = growth before splitting candidate r
            for (int r = tid; r < n; r += NT) {
                const int k = newpos[r];
                if (k != 0 && n + aux[r] < N) {               // candidates split form a prefix: growth is monotone
                    const int i = 4095 - (k & 4095);
                    key[i] = aux[r] + r;                      // children created before it = growth before it + one per candidate
                    aux2[i] = 1;
                    atomicAdd(&s_P, 1);
                    atomicAdd(&s_C, n_children(cc, i));
                }
            }
            __syncthreads();
            P = s_P; C = s_C;
            block_exclusive_scan(aux2, n, scratch);           // aux2[i] = number of split nodes before list position i
            for (int i = tid; i < 4 * (C + n - P); i += NT) cc_next[i] = 0;
            for (int i = tid; i < n; i += NT)
                place_node(A, B, cc, newpos, childpos, &s_expand, i, key[i] >= 0, key[i], aux2[i], C);
        }
        __syncthreads();
        const int n_new = C + n - P;
        const int nToExpand = s_expand;
        const bool last = n_new >= N || n_new == n;              // :833-836 / :901-902
        // every corner moves to its new node and, unless this was the last round, is counted in that node's child occupancy
        for (int k = tid; k < nC; k += NT) {
            const int nd = nof[k];
            int np = newpos[nd];
            const uint32_t c = ck[k];
            const int x = (int)(c & 4095u) - CCM_MIN_BORDER, y = (int)((c >> 12) & 4095u) - CCM_MIN_BORDER;
            if (np < 0) np = childpos[4 * nd + quadrant(A[nd], x, y)];
            nof[k] = (uint16_t)np;
            if (!last) {
                const QNode nn = B[np];
                if (nn.cnt > 1) atomicAdd(&cc_next[4 * np + quadrant(nn, x, y)], 1);
            }
        }
        __syncthreads();
        { QNode* t = A; A = B; B = t; int* u = cc; cc = cc_next; cc_next = u; }
        n = n_new;
        if (last) break;
        if (!greedy && n + nToExpand * 3 > N) greedy = true;     // :838
    }

    // ---- best corner per node (:910-928): max response, earliest in (cell row, cell col, y, x) order ----
    unsigned long long* best = reinterpret_cast<unsigned long long*>(cc);   // cap*8 B <= 4*cap*4 B
    for (int i = tid; i < n; i += NT) best[i] = 0ull;
    __syncthreads();
    for (int k = tid; k < nC; k += NT) {
        const uint32_t c = ck[k];
        const int x = (int)(c & 4095u), y = (int)((c >> 12) & 4095u);
        const int cj = (x - CCM_EDGE) / L.wCell, ci = (y - CCM_EDGE) / L.hCell;
        const uint32_t ord = (uint32_t)(((ci * L.nCols + cj) * L.hCell + (y - CCM_EDGE - ci * L.hCell)) * L.wCell + (x - CCM_EDGE - cj * L.wCell));
        const unsigned long long v = ((unsigned long long)(c >> 24) << 56) | ((unsigned long long)(0xFFFFFFFFu - ord) << 24) | (unsigned long long)(c & 0xFFFFFFu);
        atomicMax(&best[nof[k]], v);
    }
    __syncthreads();
    uint32_t* out = level_kps + (size_t)f * T.kps_per_frame + L.kp_off;
    for (int i = tid; i < n; i += NT) {
        const unsigned long long v = best[i];
        out[i] = (uint32_t)(v & 0xFFFFFFu) | ((uint32_t)(v >> 56) << 24);
    }
    if (tid == 0) level_kp_count[f * CCM_MAX_LEVELS + lvl] = n;
    CCM_TS(blockIdx.y == 0 && threadIdx.x == 0, 48 + lvl);
}
