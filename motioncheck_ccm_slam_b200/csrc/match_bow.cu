// K8b — BoW-guided Hamming searches: SearchByBoW(KeyFrame, Frame) and SearchByBoW(KeyFrame, KeyFrame)
// (cslam/src/ORBmatcher.cpp:178-306, 565-698).
//
// The reference merge-joins two DBoW2 FeatureVectors on node id and, inside a shared node, runs a greedy loop:
// every A-side feature (ascending) against every still-unmatched B-side feature of the node, best / second-best,
// threshold + ratio test, and the accepted B feature is removed from play. A feature belongs to exactly one node, so
// the greedy state never crosses nodes: nodes are independent jobs. One warp per shared node walks the node's A
// features in order (that order is the data dependence), lanes striding its B features, warp-shuffle top-2.
// The host side hands the two vectors over as CSR (order[] sorted by (node id, feature index) + per-shared-node
// ranges), which is the FeatureVector's own iteration order (thirdparty/DBoW2/DBoW2/FeatureVector.cpp:31-45).
#include "match_common.cuh"

#define BOW_MODE_KF_FRAME 0
#define BOW_MODE_KF_KF 1
#define BOW_MODE_TRIANGULATION 2

// SearchForTriangulation's geometric predicates (ORBmatcher.cpp:159-176, :773-778), thresholds folded per level by the host
struct TriGeom {
    const float* x_a; const float* y_a;                    // mvKeysUn of pKF1
    const float* x_b; const float* y_b; const int* oct_b;  // mvKeysUn of pKF2
    float f[9];                                            // F12 row-major
    float ex, ey;                                          // epipole in pKF2 (:707-714)
    float epipole_th[16];                                  // 100 * mvScaleFactors[l]                         (:777)
    float line_th[16];                                     // smallest float >= 3.84 * mvLevelSigma2[l] (double) (:175)
};

struct BowNode { int a0, a1, b0, b1; };   // ranges into order_a / order_b

struct BowJob {
    const uint8_t* desc_a; const uint8_t* desc_b;
    const uint8_t* valid_a; const uint8_t* valid_b;
    const float* angle_a; const float* angle_b;
    const int* order_a; const int* order_b;
    const BowNode* nodes; int n_nodes;
    int n_a, n_b;
    int* match_a; int* match_b;       // [n_a], [n_b], initialised to -1 by the caller (cudaMemset 0xff)
    MatchResultDev* res_a;            // [n_a] or null, initialised by k_bow_finish's counterpart below
    int* nmatches;
    int mode; float nnratio; int th_low; int check_orientation;
};

__global__ void k_bow_init(BowJob J) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < max(J.n_a, J.n_b); i += gridDim.x * blockDim.x) {
        if (i < J.n_a) {
            J.match_a[i] = -1;
            if (J.res_a) { MatchResultDev r; r.best_idx = -1; r.best_dist = 256; r.best_dist2 = 256; r.accepted = 0; J.res_a[i] = r; }
        }
        if (i < J.n_b) J.match_b[i] = -1;
    }
}

__device__ __forceinline__ uint4 shfl_u4(uint4 v, int src) {
    return make_uint4(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src),
                      __shfl_sync(0xffffffffu, v.w, src));
}

// One warp per shared vocabulary node; the A-side features are taken in FeatureVector order, each against the node's B-side
// features that no earlier A-feature has taken (:221-251 / :605-645). A feature belongs to one node, so that state is private to
// the warp: for nodes of up to 32 B-features (every node of a real vocabulary at 1000-2000 features) lane l keeps B-feature l -
// index, descriptor, "taken by" - in registers, the A-features are loaded 32 at a time (lane = feature) and handed round by
// shuffle, and an A-feature costs two REDUX minima over (distance << 16 | position) instead of a chain of four dependent
// global loads. Nodes of up to 128 B-features stage the B side in shared memory, larger ones take the streaming loop at the end.
#define BOW_SMEM_B 128               // B-features of a node staged in shared memory (per warp: 4 KB of descriptors + 1 KB)
__global__ void __launch_bounds__(256) k_bow_search(BowJob J) {
    __shared__ uint4 s_bdesc[8][2 * BOW_SMEM_B];
    __shared__ int s_bidx[8][BOW_SMEM_B], s_btaken[8][BOW_SMEM_B];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int node = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (node >= J.n_nodes) return;
    const BowNode N = J.nodes[node];
    const uint4* __restrict__ da = reinterpret_cast<const uint4*>(J.desc_a);
    const uint4* __restrict__ db = reinterpret_cast<const uint4*>(J.desc_b);
    const int nb = N.b1 - N.b0;
    if (nb <= 32) {
        int b = -1, taken_by = -1;
        bool ok_b = false;
        uint4 d0 = make_uint4(0, 0, 0, 0), d1 = d0;
        if (lane < nb) {
            b = J.order_b[N.b0 + lane];
            ok_b = J.mode != BOW_MODE_KF_KF || J.valid_b[b];                // :619-623
            d0 = db[2 * (size_t)b]; d1 = db[2 * (size_t)b + 1];
        }
        for (int ia0 = N.a0; ia0 < N.a1; ia0 += 32) {
            int a = -1, va = 0, my_match = -1;
            uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
            if (ia0 + lane < N.a1) {
                a = J.order_a[ia0 + lane];
                va = J.valid_a[a] ? 1 : 0;                                   // no good map point, :213-219 / :597-603
                q0 = da[2 * (size_t)a]; q1 = da[2 * (size_t)a + 1];
            }
            MatchResultDev r; r.best_idx = -1; r.best_dist = 256; r.best_dist2 = 256; r.accepted = 0;
            bool has_r = false;
            const int cnt = min(32, N.a1 - ia0);
            for (int j = 0; j < cnt; j++) {
                if (!__shfl_sync(full, va, j)) continue;                     // warp-uniform
                const uint4 x0 = shfl_u4(q0, j), x1 = shfl_u4(q1, j);
                uint32_t key = 0xffffffffu;
                if (ok_b && taken_by < 0) key = ((uint32_t)hamming256(x0, x1, d0, d1) << 16) | (uint32_t)lane;   // vpMapPointMatches[realIdxF] / vbMatched2[idx2]
                const uint32_t k1 = __reduce_min_sync(full, key);
                if (k1 == 0xffffffffu) continue;
                const uint32_t k2 = __reduce_min_sync(full, key == k1 ? 0xffffffffu : key);
                const int best_lane = (int)(k1 & 0xffffu), bestDist = (int)(k1 >> 16);
                const int bestDist2 = k2 == 0xffffffffu ? 256 : (int)(k2 >> 16);
                const bool under = J.mode == BOW_MODE_KF_KF ? bestDist < J.th_low : bestDist <= J.th_low;   // :641 vs :241
                const bool acc = under && (float)bestDist < __fmul_rn(J.nnratio, (float)bestDist2);
                const int bestIdx = __shfl_sync(full, b, best_lane), aj = __shfl_sync(full, a, j);
                if (acc && lane == best_lane) taken_by = aj;
                if (lane == j) {
                    if (acc) my_match = bestIdx;
                    r.best_idx = bestIdx; r.best_dist = bestDist; r.best_dist2 = bestDist2; r.accepted = acc;
                    has_r = true;
                }
            }
            if (a >= 0) {
                if (my_match >= 0) J.match_a[a] = my_match;
                if (has_r && J.res_a) J.res_a[a] = r;
            }
        }
        if (taken_by >= 0) J.match_b[b] = taken_by;
        return;
    }
    if (nb <= BOW_SMEM_B) {
        // up to 128 B-features: descriptors, indices and "taken by" staged in the warp's slice of shared memory, "taken" also as
        // a bit mask in (warp-uniform) registers with the B-features that may not be matched at all (KF-KF: no map point) set
        // from the start; an A-feature is ceil(nb / 32) shared-memory steps and two REDUX minima
        const int warp = threadIdx.x >> 5;
        uint4* sd = s_bdesc[warp];
        int* sb = s_bidx[warp];
        int* st = s_btaken[warp];
        uint32_t taken[BOW_SMEM_B / 32];
#pragma unroll
        for (int c = 0; c < BOW_SMEM_B / 32; c++) {
            const int bl = 32 * c + lane;
            bool blocked = true;
            if (bl < nb) {
                const int b = J.order_b[N.b0 + bl];
                sb[bl] = b; st[bl] = -1;
                sd[2 * bl] = db[2 * (size_t)b]; sd[2 * bl + 1] = db[2 * (size_t)b + 1];
                blocked = J.mode == BOW_MODE_KF_KF && !J.valid_b[b];        // :619-623
            }
            taken[c] = __ballot_sync(full, blocked);
        }
        __syncwarp();
        for (int ia0 = N.a0; ia0 < N.a1; ia0 += 32) {
            int a = -1, va = 0, my_match = -1;
            uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
            if (ia0 + lane < N.a1) {
                a = J.order_a[ia0 + lane];
                va = J.valid_a[a] ? 1 : 0;
                q0 = da[2 * (size_t)a]; q1 = da[2 * (size_t)a + 1];
            }
            MatchResultDev r; r.best_idx = -1; r.best_dist = 256; r.best_dist2 = 256; r.accepted = 0;
            bool has_r = false;
            const int cnt = min(32, N.a1 - ia0);
            for (int j = 0; j < cnt; j++) {
                if (!__shfl_sync(full, va, j)) continue;                     // warp-uniform
                const uint4 x0 = shfl_u4(q0, j), x1 = shfl_u4(q1, j);
                uint32_t ka = 0xffffffffu, kb = 0xffffffffu;                 // this lane's best two
#pragma unroll
                for (int c = 0; c < BOW_SMEM_B / 32; c++) {
                    const int bl = 32 * c + lane;
                    if (32 * c < nb && !((taken[c] >> lane) & 1u)) {
                        const uint32_t key = ((uint32_t)hamming256(x0, x1, sd[2 * bl], sd[2 * bl + 1]) << 16) | (uint32_t)bl;
                        if (key < ka) { kb = ka; ka = key; } else if (key < kb) kb = key;
                    }
                }
                const uint32_t k1 = __reduce_min_sync(full, ka);
                if (k1 == 0xffffffffu) continue;
                const uint32_t k2 = __reduce_min_sync(full, ka == k1 ? kb : ka);
                const int best_bl = (int)(k1 & 0xffffu), bestDist = (int)(k1 >> 16);
                const int bestDist2 = k2 == 0xffffffffu ? 256 : (int)(k2 >> 16);
                const bool under = J.mode == BOW_MODE_KF_KF ? bestDist < J.th_low : bestDist <= J.th_low;   // :641 vs :241
                const bool acc = under && (float)bestDist < __fmul_rn(J.nnratio, (float)bestDist2);
                const int bestIdx = sb[best_bl], aj = __shfl_sync(full, a, j);
                if (acc) {
#pragma unroll
                    for (int c = 0; c < BOW_SMEM_B / 32; c++) if ((best_bl >> 5) == c) taken[c] |= 1u << (best_bl & 31);
                    if (lane == 0) st[best_bl] = aj;
                }
                if (lane == j) {
                    if (acc) my_match = bestIdx;
                    r.best_idx = bestIdx; r.best_dist = bestDist; r.best_dist2 = bestDist2; r.accepted = acc;
                    has_r = true;
                }
            }
            if (a >= 0) {
                if (my_match >= 0) J.match_a[a] = my_match;
                if (has_r && J.res_a) J.res_a[a] = r;
            }
        }
        __syncwarp();
        for (int bl = lane; bl < nb; bl += 32) if (st[bl] >= 0) J.match_b[sb[bl]] = st[bl];
        return;
    }
    for (int ia = N.a0; ia < N.a1; ia++) {
        const int a = J.order_a[ia];
        if (!J.valid_a[a]) continue;                                   // no good map point, :213-219 / :597-603
        const uint4 q0 = da[2 * (size_t)a], q1 = da[2 * (size_t)a + 1];
        unsigned long long b1 = 0xffffffffffffffffull, b2 = b1;
        for (int ib = N.b0 + lane; ib < N.b1; ib += 32) {
            const int b = J.order_b[ib];
            if (J.match_b[b] >= 0) continue;                           // vpMapPointMatches[realIdxF] / vbMatched2[idx2]
            if (J.mode == BOW_MODE_KF_KF && !J.valid_b[b]) continue;   // :619-623
            const int dist = hamming256(q0, q1, db[2 * (size_t)b], db[2 * (size_t)b + 1]);
            push_top2(b1, b2, ((unsigned long long)dist << 32) | ((unsigned long long)(ib - N.b0) << 16) | (unsigned)b);
        }
        warp_top2(b1, b2);
        if (lane == 0 && b1 != 0xffffffffffffffffull) {
            const int bestIdx = (int)(b1 & 0xffff), bestDist = (int)(b1 >> 32);
            const int bestDist2 = b2 == 0xffffffffffffffffull ? 256 : (int)(b2 >> 32);
            const bool under = J.mode == BOW_MODE_KF_KF ? bestDist < J.th_low : bestDist <= J.th_low;   // :641 vs :241
            const bool acc = under && (float)bestDist < __fmul_rn(J.nnratio, (float)bestDist2);
            if (acc) { J.match_b[bestIdx] = a; J.match_a[a] = bestIdx; }
            if (J.res_a) { MatchResultDev r; r.best_idx = bestIdx; r.best_dist = bestDist; r.best_dist2 = bestDist2; r.accepted = acc; J.res_a[a] = r; }
        }
        __syncwarp();
    }
}

// rotation-histogram pruning (:256-262, 284-303 / :648-659, 676-695) and the return value
__global__ void __launch_bounds__(1024) k_bow_finish(BowJob J) {
    __shared__ int s_hist[32], s_keep[3], s_nm;
    if (threadIdx.x < 32) s_hist[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_nm = 0;
    __syncthreads();
    int nm = 0;
    for (int a = threadIdx.x; a < J.n_a; a += blockDim.x) {
        const int b = J.match_a[a];
        if (b < 0) continue;
        nm++;
        if (J.check_orientation) atomicAdd(&s_hist[rot_bin(J.angle_a[a], J.angle_b[b])], 1);
    }
    if (nm) atomicAdd(&s_nm, nm);
    __syncthreads();
    if (J.check_orientation) {
        if (threadIdx.x == 0) three_maxima(s_hist, 30, s_keep[0], s_keep[1], s_keep[2]);
        __syncthreads();
        int drop = 0;
        for (int a = threadIdx.x; a < J.n_a; a += blockDim.x) {
            const int b = J.match_a[a];
            if (b < 0) continue;
            const int bin = rot_bin(J.angle_a[a], J.angle_b[b]);
            if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) { J.match_a[a] = -1; J.match_b[b] = -1; drop++; }
        }
        if (drop) atomicSub(&s_nm, drop);
        __syncthreads();
    }
    if (threadIdx.x == 0) *J.nmatches = s_nm;
}

// SearchForTriangulation (ORBmatcher.cpp:700-852). vbMatched2 is never set upstream, so the pKF1 features are
// independent: one warp per (shared node, pKF1 feature), lanes over the node's pKF2 features. The running
// `dist > bestDist` test (:768) only ever compares against candidates that passed both geometric gates, so the loop
// returns the gated candidate of minimum distance, the LAST one among equals: key = dist << 16 | (0xffff - position).
// Float expressions keep the reference's operation order with explicit round-to-nearest mul/add (no FMA contraction).
__global__ void __launch_bounds__(256) k_tri_search(BowJob J, TriGeom G, const int* __restrict__ a_node) {
    const int lane = threadIdx.x & 31;
    const int ia = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);      // position in order_a
    if (ia >= J.n_a) return;
    const int ni = a_node[ia];
    if (ni < 0) return;                                                     // node not shared with pKF2
    const int a = J.order_a[ia];
    if (!J.valid_a[a]) return;                                              // pKF1 feature holds a map point, :741-744
    const BowNode N = J.nodes[ni];
    const uint4* __restrict__ da = reinterpret_cast<const uint4*>(J.desc_a);
    const uint4* __restrict__ db = reinterpret_cast<const uint4*>(J.desc_b);
    const uint4 q0 = da[2 * (size_t)a], q1 = da[2 * (size_t)a + 1];
    const float x1 = G.x_a[a], y1 = G.y_a[a];
    const float la = __fadd_rn(__fadd_rn(__fmul_rn(x1, G.f[0]), __fmul_rn(y1, G.f[3])), G.f[6]);
    const float lb = __fadd_rn(__fadd_rn(__fmul_rn(x1, G.f[1]), __fmul_rn(y1, G.f[4])), G.f[7]);
    const float lc = __fadd_rn(__fadd_rn(__fmul_rn(x1, G.f[2]), __fmul_rn(y1, G.f[5])), G.f[8]);
    const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
    uint32_t best = 0xffffffffu;
    if (den != 0.f) {
        for (int ib = N.b0 + lane; ib < N.b1; ib += 32) {
            const int b = J.order_b[ib];
            if (!J.valid_b[b]) continue;                                    // pKF2 feature holds a map point, :758-762
            const int dist = hamming256(q0, q1, db[2 * (size_t)b], db[2 * (size_t)b + 1]);
            if (dist > J.th_low) continue;
            const float x2 = G.x_b[b], y2 = G.y_b[b];
            const int o = G.oct_b[b];
            const float dex = __fsub_rn(G.ex, x2), dey = __fsub_rn(G.ey, y2);
            if (__fadd_rn(__fmul_rn(dex, dex), __fmul_rn(dey, dey)) < G.epipole_th[o]) continue;
            const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, x2), __fmul_rn(lb, y2)), lc);
            const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
            if (!(dsqr < G.line_th[o])) continue;
            best = min(best, ((uint32_t)dist << 16) | (uint32_t)(0xffff - (ib - N.b0)));
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0 && best != 0xffffffffu) {
        const int b = J.order_b[N.b0 + (0xffff - (int)(best & 0xffff))];
        J.match_a[a] = b;
        if (J.res_a) { MatchResultDev r; r.best_idx = b; r.best_dist = (int)(best >> 16); r.best_dist2 = 256; r.accepted = 1; J.res_a[a] = r; }
    }
}

// Vocabulary descent of Frame::ComputeBoW (cslam/src/Frame.cpp:268-275 -> DBoW2 TemplatedVocabulary::transform,
// thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1261): per feature, from the root to a leaf through the child at the
// smallest Hamming distance (first child wins ties), remembering the node passed at level nid_level = L - levelsup.
// One warp per feature; the children of a node are stored contiguously (CSR order), lane c takes child c, so a level is
// one coalesced read of k x 32 B and one warp min over (distance << 16 | child position). A k = 10, L = 6 vocabulary is
// 35 MB: resident in the 126 MB L2 after the first batch.
struct VocabDev {
    const int* child_start;      // [n_nodes + 1]
    const int* child_id;         // [n_nodes - 1] node id per CSR position
    const uint4* child_desc;     // [n_nodes - 1][2] descriptor per CSR position
    const int* node_word;        // [n_nodes] word id of a leaf, -1 for inner nodes
};

__global__ void __launch_bounds__(256) k_vocab_transform(VocabDev V, int nid_level, int n, const uint4* __restrict__ desc,
                                                         int* __restrict__ word, int* __restrict__ node_out) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= n) return;
    const uint4 q0 = __ldg(desc + 2 * (size_t)i), q1 = __ldg(desc + 2 * (size_t)i + 1);
    int node = 0, nid = 0, level = 0;
    for (;;) {
        const int c0 = __ldg(V.child_start + node), c1 = __ldg(V.child_start + node + 1);
        if (c1 == c0) break;                                                   // a node without children is a word (:1256)
        ++level;
        uint32_t best = 0xffffffffu;
        for (int p = c0 + lane; p < c1; p += 32) {
            const int d = hamming256(q0, q1, __ldg(V.child_desc + 2 * (size_t)p), __ldg(V.child_desc + 2 * (size_t)p + 1));
            best = min(best, ((uint32_t)d << 16) | (uint32_t)min(p - c0, 0xffff));
        }
        best = __reduce_min_sync(0xffffffffu, best);
        node = __ldg(V.child_id + c0 + (int)(best & 0xffffu));
        if (level == nid_level) nid = node;
    }
    if (lane == 0) { word[i] = __ldg(V.node_word + node); node_out[i] = nid; }
}
