// K9 — server keyframe-database search: best and second-best Hamming distance of every query descriptor over a
// (shard of a) target descriptor database. This is the inner loop of SearchByBoW(KF,KF)
// (cslam/src/ORBmatcher.cpp:607-637) with one all-inclusive vocabulary node and no cross-query state
// (BASELINE.json configs[4]); ties resolve to the lowest target index, as the reference's strict `<` over ascending
// indices does.
//
// Integer-pipe bound. Each thread keeps Q query descriptors (8 words each) in registers; the CTA streams its
// chunk of the target database through shared memory in 256-descriptor tiles, every thread reading the same
// target (one broadcast LDS.128 pair per target, amortised over Q queries). Distance = popcount of the 256-bit
// XOR, computed with a carry-save adder tree on LOP3 (7 words -> ones/twos/fours planes) so that a pair costs
// 4 POPC instead of 8: POPC issues at 16 lanes/clk/SM on B200 (tools/ubench_int.cu, profiles/r1_ubench_int.json)
// while LOP3 issues at 64. Top-2 is kept as two packed u32 keys (distance << 23 | index-in-chunk) updated with
// three min/max per pair, so there is no branch and no shuffle in the loop. HBM traffic is nq_tiles * nt * 32 B
// per launch, i.e. ~0.03 B per pair: the roofline is the integer pipe, not memory.
#include "match_common.cuh"

#define AP_THREADS 256
#define AP_TILE 256                 // targets per shared-memory tile (8 KB)
#define AP_IDX_BITS 23              // chunk length <= 2^23 targets
#define AP_MAX_CHUNK (1 << AP_IDX_BITS)

__device__ __forceinline__ uint32_t lop3_xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// |q xor w| over 256 bits == ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1653-1669): 8 XOR, 10 LOP3, 4 POPC
__device__ __forceinline__ uint32_t hamming256_csa(const uint32_t* q, const uint32_t* w) {
    uint32_t x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = q[i] ^ w[i];
    const uint32_t o0 = lop3_xor3(x[0], x[1], x[2]), t0 = lop3_maj(x[0], x[1], x[2]);
    const uint32_t o1 = lop3_xor3(x[3], x[4], x[5]), t1 = lop3_maj(x[3], x[4], x[5]);
    const uint32_t o2 = lop3_xor3(o0, o1, x[6]), t2 = lop3_maj(o0, o1, x[6]);
    const uint32_t tw = lop3_xor3(t0, t1, t2), fo = lop3_maj(t0, t1, t2);
    return __popc(o2) + __popc(x[7]) + 2 * __popc(tw) + 4 * __popc(fo);
}

template <int Q>
__global__ void __launch_bounds__(AP_THREADS)
k_allpairs_best2(const uint4* __restrict__ qd, int nq, const uint4* __restrict__ tg, long long nt, int chunk,
                 uint32_t index_base, Best2Dev* __restrict__ parts) {
    __shared__ uint4 st[2 * AP_TILE];
    const int qbase = blockIdx.x * (AP_THREADS * Q);
    const long long c0 = (long long)blockIdx.y * chunk;
    const long long c1 = min(nt, c0 + chunk);

    uint32_t q[Q][8];
#pragma unroll
    for (int j = 0; j < Q; j++) {
        const int qi = qbase + j * AP_THREADS + threadIdx.x;
        uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
        if (qi < nq) { lo = qd[2 * (size_t)qi]; hi = qd[2 * (size_t)qi + 1]; }
        q[j][0] = lo.x; q[j][1] = lo.y; q[j][2] = lo.z; q[j][3] = lo.w;
        q[j][4] = hi.x; q[j][5] = hi.y; q[j][6] = hi.z; q[j][7] = hi.w;
    }
    uint32_t b1[Q], b2[Q];
#pragma unroll
    for (int j = 0; j < Q; j++) b1[j] = b2[j] = 0xffffffffu;

    // register prefetch of the next tile: two uint4 per thread
    const uint4* tsrc = tg + 2 * c0;
    const long long n_u4 = 2 * (c1 - c0);
    uint4 p0 = make_uint4(0, 0, 0, 0), p1 = p0;
    if (threadIdx.x < n_u4) p0 = tsrc[threadIdx.x];
    if (threadIdx.x + AP_THREADS < n_u4) p1 = tsrc[threadIdx.x + AP_THREADS];

    for (long long base = c0; base < c1; base += AP_TILE) {
        __syncthreads();
        st[threadIdx.x] = p0;
        st[threadIdx.x + AP_THREADS] = p1;
        __syncthreads();
        {   // prefetch tile base + AP_TILE
            const long long o = 2 * (base + AP_TILE - c0);
            if (o + threadIdx.x < n_u4) p0 = tsrc[o + threadIdx.x];
            if (o + threadIdx.x + AP_THREADS < n_u4) p1 = tsrc[o + threadIdx.x + AP_THREADS];
        }
        const int cnt = (int)min((long long)AP_TILE, c1 - base);
        const uint32_t kbase = (uint32_t)(base - c0);
#pragma unroll 2
        for (int t = 0; t < cnt; t++) {
            const uint4 lo = st[2 * t], hi = st[2 * t + 1];
            const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
            for (int j = 0; j < Q; j++) {
                const uint32_t d = hamming256_csa(q[j], w);
                const uint32_t key = (d << AP_IDX_BITS) + (kbase + t);
                const uint32_t m = max(b1[j], key);
                b1[j] = min(b1[j], key);
                b2[j] = min(b2[j], m);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < Q; j++) {
        const int qi = qbase + j * AP_THREADS + threadIdx.x;
        if (qi >= nq) continue;
        Best2Dev r;
        r.d1 = (uint16_t)min(b1[j] >> AP_IDX_BITS, 256u);
        r.d2 = (uint16_t)min(b2[j] >> AP_IDX_BITS, 256u);
        r.idx = b1[j] == 0xffffffffu ? 0xffffffffu
                                     : index_base + (uint32_t)c0 + (b1[j] & (AP_MAX_CHUNK - 1));
        parts[(size_t)blockIdx.y * nq + qi] = r;
    }
}

// Merge of per-chunk (or, after the NCCL allgather, per-shard) best-2 records: global best = minimum distance, ties
// to the lowest global index (the reference's first-wins order); global second = second-smallest distance of the
// union. Then the reference accept rule of SearchByBoW(KF,KF): best < TH_LOW and (float)best < nnratio*(float)best2
// (ORBmatcher.cpp:641-643).
__global__ void k_allpairs_merge(const Best2Dev* __restrict__ parts, int n_parts, int nq, float nnratio, int th_low,
                                 Best2Dev* __restrict__ out, uint8_t* __restrict__ accept) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    uint32_t bd = 256, bi = 0xffffffffu, sd = 256;
    for (int p = 0; p < n_parts; p++) {
        const Best2Dev r = parts[(size_t)p * nq + qi];
        if (r.d1 < bd || (r.d1 == bd && r.idx < bi)) { sd = min(sd, bd); bd = r.d1; bi = r.idx; }
        else sd = min(sd, (uint32_t)r.d1);
        sd = min(sd, (uint32_t)r.d2);
    }
    Best2Dev o; o.idx = bi; o.d1 = (uint16_t)bd; o.d2 = (uint16_t)sd;
    out[qi] = o;
    if (accept) accept[qi] = ((int)bd < th_low && (float)bd < __fmul_rn(nnratio, (float)sd)) ? 1 : 0;
}

// static ORBmatcher::DescriptorDistance for n independent pairs
__global__ void k_descriptor_distance(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = hamming256(a[2 * (size_t)i], a[2 * (size_t)i + 1], b[2 * (size_t)i], b[2 * (size_t)i + 1]);
}

// MapPoint::ComputeDistinctiveDescriptors (cslam/src/MapPoint.cpp:929-994), batched over map points: among the
// descriptors of a point's observations, the one whose sorted row of Hamming distances to all of them (its own 0
// included) has the least element at index int(0.5 * (N - 1)); the first such row wins. One warp per map point, one lane
// per row (strided for N > 32). The k-th smallest of a row is found by bisection on the value range [0, 256]: 9 counting
// passes over the point's descriptors, which every lane reads at the same address (one broadcast transaction), so no
// distance matrix is stored.
__global__ void __launch_bounds__(256) k_distinctive(const int* __restrict__ start, const uint4* __restrict__ desc, int n_points,
                                                     int* __restrict__ best, int* __restrict__ best_median) {
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= n_points) return;
    const int r0 = start[p], N = start[p + 1] - r0;
    if (N <= 0) { if (lane == 0) { best[p] = -1; best_median[p] = 0; } return; }
    const int k = (int)(0.5 * (N - 1));
    uint32_t key = 0xffffffffu;
    for (int i = lane; i < N; i += 32) {
        const uint4 q0 = desc[2 * (size_t)(r0 + i)], q1 = desc[2 * (size_t)(r0 + i) + 1];
        int lo = 0, hi = 256;                                   // smallest v with #{j : d_ij <= v} >= k + 1
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < N; j++) cnt += hamming256(q0, q1, desc[2 * (size_t)(r0 + j)], desc[2 * (size_t)(r0 + j) + 1]) <= mid;
            if (cnt >= k + 1) hi = mid; else lo = mid + 1;
        }
        key = min(key, ((uint32_t)lo << 16) | (uint32_t)min(i, 0xffff));
    }
    key = __reduce_min_sync(0xffffffffu, key);
    if (lane == 0) { best[p] = (int)(key & 0xffffu); best_median[p] = (int)(key >> 16); }
}

// ---------------------------------------------------------------------------------------------
// Fused fold + exchange + merge for the sharded keyframe-database search (BASELINE.json configs[4]): the one path with a
// real exchange step. Instead of fold -> NCCL all_gather -> merge (three launches and a collective whose 8 KB payload is
// pure latency), ONE kernel per rank
//   1. folds the rank's chunk records into its per-query best-2 (as k_allpairs_merge does),
//   2. stores that record straight into slot [rank] of EVERY rank's exchange buffer over NVLink (peer memory, plain
//      8-byte stores: a CTA writes 2 KB per peer, coalesced), fences at system scope and raises the CTA's flag on
//      every peer,
//   3. waits until the same CTA index of every other rank has raised its flag in the local buffer (the records it needs
//      are exactly those that CTA wrote: no grid-wide or cross-GPU barrier), and
//   4. merges the `world` records per query and applies the accept rule (ORBmatcher.cpp:641-643).
// Records and flags are double-buffered by the parity of a per-context epoch: a rank can run at most one call ahead of
// a peer (its call e+1 needs the peer's e+1 flags, which the peer raises only after its own call e has finished), so
// the set of parity e&1 is never overwritten while somebody still merges call e.
// Buffer of one rank: Best2Dev rec[2][world][nq_cap] | uint32_t flag[2][world][ctas_cap].
// ---------------------------------------------------------------------------------------------
#define EX_MAX_WORLD 16
#define EX_THREADS 256
struct ExPeers { unsigned char* p[EX_MAX_WORLD]; };

__device__ __forceinline__ unsigned long long ld_sys_u64(const void* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const void* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(void* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// status (device word of the exchange context): set to 1 by a CTA whose wait ran past timeout_ns. A peer that never
// launches (its call failed or its process died) would otherwise leave this kernel spinning with no way to cancel it.
__global__ void __launch_bounds__(EX_THREADS) k_allpairs_exchange(const Best2Dev* __restrict__ parts, int n_parts, int nq, float nnratio, int th_low,
                                                                  const ExPeers peers, int rank, int world, int nq_cap, int ctas_cap, uint32_t epoch,
                                                                  Best2Dev* __restrict__ out, uint8_t* __restrict__ accept,
                                                                  unsigned long long timeout_ns, uint32_t* __restrict__ status) {
    __shared__ int timed_out;
    if (threadIdx.x == 0) timed_out = 0;
    const int qi = blockIdx.x * EX_THREADS + threadIdx.x;
    const uint32_t par = epoch & 1u;
    const size_t rec_bytes = (size_t)2 * world * nq_cap * sizeof(Best2Dev);
    // 1. fold this rank's chunks
    uint32_t bd = 256, bi = 0xffffffffu, sd = 256;
    if (qi < nq) {
        for (int p = 0; p < n_parts; p++) {
            const Best2Dev r = parts[(size_t)p * nq + qi];
            if (r.d1 < bd || (r.d1 == bd && r.idx < bi)) { sd = min(sd, bd); bd = r.d1; bi = r.idx; }
            else sd = min(sd, (uint32_t)r.d1);
            sd = min(sd, (uint32_t)r.d2);
        }
        // 2. publish into slot [par][rank] of every rank's buffer (own included)
        const unsigned long long packed = (unsigned long long)bi | ((unsigned long long)bd << 32) | ((unsigned long long)sd << 48);   // == Best2Dev {idx, d1, d2}
        for (int t = 0; t < world; t++) {
            unsigned long long* dst = reinterpret_cast<unsigned long long*>(peers.p[t]) + ((size_t)par * world + rank) * nq_cap + qi;
            *dst = packed;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) {
        uint32_t* flag = reinterpret_cast<uint32_t*>(peers.p[threadIdx.x] + rec_bytes) + ((size_t)par * world + rank) * ctas_cap + blockIdx.x;
        st_release_sys_u32(flag, epoch);
    }
    // 3. wait for the same CTA of every rank (flags live in the local buffer)
    if (threadIdx.x < world) {
        const uint32_t* flag = reinterpret_cast<const uint32_t*>(peers.p[rank] + rec_bytes) + ((size_t)par * world + threadIdx.x) * ctas_cap + blockIdx.x;
        const unsigned long long t0 = globaltimer_ns();
        while (ld_acquire_sys_u32(flag) != epoch) {
            __nanosleep(100);
            if (globaltimer_ns() - t0 > timeout_ns) { timed_out = 1; atomicExch(status, 1u); break; }
        }
    }
    __syncthreads();
    // 4. merge across ranks
    if (qi >= nq) return;
    if (timed_out) {                                     // a peer never published: no result, the host reports CCM_E_CUDA
        Best2Dev o; o.idx = 0xffffffffu; o.d1 = 256; o.d2 = 256;
        out[qi] = o;
        if (accept) accept[qi] = 0;
        return;
    }
    bd = 256; bi = 0xffffffffu; sd = 256;
    const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(peers.p[rank]) + (size_t)par * world * nq_cap + qi;
    for (int t = 0; t < world; t++) {
        const unsigned long long v = ld_sys_u64(mine + (size_t)t * nq_cap);
        const uint32_t idx = (uint32_t)v, d1 = (uint32_t)(v >> 32) & 0xffffu, d2 = (uint32_t)(v >> 48);
        if (d1 < bd || (d1 == bd && idx < bi)) { sd = min(sd, bd); bd = d1; bi = idx; }
        else sd = min(sd, d1);
        sd = min(sd, d2);
    }
    Best2Dev o; o.idx = bi; o.d1 = (uint16_t)bd; o.d2 = (uint16_t)sd;
    out[qi] = o;
    if (accept) accept[qi] = ((int)bd < th_low && (float)bd < __fmul_rn(nnratio, (float)sd)) ? 1 : 0;
}
