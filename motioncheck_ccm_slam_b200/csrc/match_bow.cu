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

__global__ void __launch_bounds__(256) k_bow_search(BowJob J) {
    const int lane = threadIdx.x & 31;
    const int node = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (node >= J.n_nodes) return;
    const BowNode N = J.nodes[node];
    const uint4* __restrict__ da = reinterpret_cast<const uint4*>(J.desc_a);
    const uint4* __restrict__ db = reinterpret_cast<const uint4*>(J.desc_b);
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
