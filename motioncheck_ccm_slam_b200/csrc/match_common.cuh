// Shared declarations of the Hamming-search kernels (K8a windowed, K8b BoW-grouped, K9 all-pairs), sm_100a.
// Integer-pipe work: XOR/LOP3 + POPC; no tensor cores by design (not a dense floating-point contraction).
//
// HBM layout of a frame pack (one Frame / KeyFrame as the matcher sees it), structure-of-arrays so that a warp
// walking candidate features reads each field as a contiguous stream:
//   desc   [cap][32] u8, 32-B aligned rows (two 16-B loads per descriptor)
//   x, y   [cap] f32   undistorted keypoint position (mvKeysUn[i].pt)
//   angle  [cap] f32   degrees
//   octave [cap] i32
//   cell_start [cols*rows+1] i32, items [cap] u16-in-i32: the feature grid as CSR, cells in (ix*rows + iy) order,
//   features inside a cell in ascending index (what push_back in index order produces, Frame.cpp:103-118)
//   n      i32 (device-resident feature count: the extractor writes it, nobody reads it back)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define CCM_MAX_GRID_CELLS 8192
#define CCM_MATCH_MAX_LEVELS 16

struct PackDev {
    const uint8_t* desc;
    const float* x;
    const float* y;
    const float* angle;
    const int* octave;
    const int* cell_start;
    const int* items;
    const int* n;
    int cap;
    float min_x, min_y, inv_w, inv_h;
    int cols, rows;
};

struct WindowQueryDev {   // == ccm_window_query (include/ccm_match.h), 24 bytes
    float u, v, r, angle;
    int16_t min_level, max_level;
    uint32_t flags;
};

struct MatchResultDev { int best_idx, best_dist, best_dist2, accepted; };   // == ccm_match_result

struct Best2Dev { uint32_t idx; uint16_t d1, d2; };                          // == ccm_best2

// One windowed-search job: a query list against one pack. Scratch (per job) lives in the matcher's arena.
struct WindowJob {
    PackDev pack;
    const WindowQueryDev* q;
    const uint8_t* q_desc;     // [nq][32]
    const int* nq_ptr;         // device count (batched path) or null
    int nq;                    // used when nq_ptr == null
    int* holder;               // [pack.cap] in/out
    MatchResultDev* res;       // [nq] or null
    int* nmatches;             // [1]
    // scratch (per job, carved from the matcher's arena)
    int* head;                 // [3 * pack.cap] lowest blocking claimant per target, three rotating copies (INIT: vMatchedDistance in the first)
    int* claim[2];             // [nq_cap] x2  claimed target per query (ping-pong), -1 none (INIT: vnMatches12, accept bin)
    int* qbest;                // [nq_cap] x3  best idx, best dist, best dist2 of the last evaluation
    uint32_t* keys;            // [nq_cap] x WIN_TOPK  best candidates of the static pass (k_window_eval), best first: distance << 16 | feature index; an exact prefix of the window's ranking, 0xffffffff after it
    int* total;                // [nq_cap]   candidates that passed the static filters
    int* fb;                   // [nq_cap]   queries that must re-walk the grid this round
};

struct WindowParamsDev {
    int mode;
    float nnratio;
    int th_dist;
    int check_orientation;
    int use_gate;
    int mc_smem_ints;          // dynamic shared memory of k_window_resolve in ints (0: the claimant arrays stay in global scratch)
    float inv_sigma2[CCM_MATCH_MAX_LEVELS];
};

// ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1653-1669): the SWAR popcount there equals POPC on each 32-bit word.
__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// rotation-histogram bin (e.g. ORBmatcher.cpp:256-262): float subtract, +360 if negative, round(rot * (1/30))
__device__ __forceinline__ int rot_bin(float angle_a, float angle_b) {
    const float factor = 1.0f / 30;
    float rot = __fsub_rn(angle_a, angle_b);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, factor));
    if (bin == 30) bin = 0;
    return bin;
}

// ORBmatcher::ComputeThreeMaxima (ORBmatcher.cpp:1607-1648) over bin counts
__device__ inline void three_maxima(const int* cnt, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = cnt[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

// merge two sorted (best, second) key pairs; keys are unique except for the all-ones sentinel
__device__ __forceinline__ void merge_top2(unsigned long long& b1, unsigned long long& b2, unsigned long long o1, unsigned long long o2) {
    const unsigned long long lo = min(b1, o1), hi = max(b1, o1);
    b2 = min(hi, min(b2, o2));
    b1 = lo;
}
__device__ __forceinline__ void push_top2(unsigned long long& b1, unsigned long long& b2, unsigned long long k) {
    if (k < b1) { b2 = b1; b1 = k; }
    else if (k < b2) b2 = k;
}
__device__ __forceinline__ void warp_top2(unsigned long long& b1, unsigned long long& b2) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, b1, o);
        const unsigned long long o2 = __shfl_xor_sync(0xffffffffu, b2, o);
        merge_top2(b1, b2, o1, o2);
    }
}

// cv::undistortPoints for one point (OpenCV 4.13.0 cvUndistortPointsInternal, default 5 iterations, R = I, P = K), the
// arithmetic of Frame::UndistortKeyPoints (Frame.cpp:295). Doubles with explicit round-to-nearest intrinsics so that
// nvcc cannot contract a multiply-add: the host library is built without FMA and the results must be identical.
// k[5..11] are 0 for the reference's 4/5-coefficient models; the terms they multiply are kept where a zero product
// still takes part in a rounding (x + 0 is exact, so they are dropped only where they add an exact zero).
struct CameraDev { double fx, fy, cx, cy, ifx, ify, k0, k1, k2, k3, k4; int distorted; };
__device__ __forceinline__ float2 undistort_point(const CameraDev& C, float xf, float yf) {
    if (!C.distorted) return make_float2(xf, yf);
    const double u = (double)xf, v = (double)yf;
    double x = __dmul_rn(__dsub_rn(u, C.cx), C.ifx), y = __dmul_rn(__dsub_rn(v, C.cy), C.ify);
    const double x0 = x, y0 = y;
    const double two_k2 = __dmul_rn(2.0, C.k2), two_k3 = __dmul_rn(2.0, C.k3);
#pragma unroll 1
    for (int j = 0; j < 5; j++) {
        const double xx = __dmul_rn(x, x), yy = __dmul_rn(y, y);
        const double r2 = __dadd_rn(xx, yy);
        // (1 + ((k7*r2 + k6)*r2 + k5)*r2) with k5..k7 = 0: ((0*r2 + 0)*r2 + 0)*r2 = 0 for finite r2, 1 + 0 = 1
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(C.k4, r2), C.k1), r2), C.k0), r2));
        const double icdist = __ddiv_rn(1.0, den);
        if (icdist < 0) { x = x0; y = y0; break; }
        // deltaX = 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2, left to right; the k8..k11 terms add exact zeros
        const double dx = __dadd_rn(__dmul_rn(__dmul_rn(two_k2, x), y), __dmul_rn(C.k3, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
        const double dy = __dadd_rn(__dmul_rn(C.k2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))), __dmul_rn(__dmul_rn(two_k3, x), y));
        x = __dmul_rn(__dsub_rn(x0, dx), icdist);
        y = __dmul_rn(__dsub_rn(y0, dy), icdist);
    }
    // RR = K: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1/(0*x + 0*y + 1) = 1
    const double ox = __dadd_rn(__dadd_rn(__dmul_rn(C.fx, x), __dmul_rn(0.0, y)), C.cx);
    const double oy = __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(C.fy, y)), C.cy);
    return make_float2(__double2float_rn(ox), __double2float_rn(oy));
}
