// Allocation-free variant of the server keyframe-database search (ccm_match_allpairs_best2_device, match_api.cu): the
// same two kernels (k_allpairs_best2 per chunk, k_allpairs_merge), with the per-chunk partial records — the one piece of
// scratch memory that search needs — in a fixed file-scope device array instead of the matcher's grow-only arena.
// For callers that may not call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid
// it. No counterpart in the reference.
//
// This file owns the array, its set-up and the entry point. The entry point calls no cudaMalloc / cudaFree / stream or
// event creation and no allocating library routine; it needs no ccm_matcher handle (creating one makes a stream).
//
// Rules the code below follows:
//  * in HOST code the array's name is only a host-side stand-in for the symbol, never a device address. The device
//    address is fetched with cudaGetSymbolAddress, once per device, with that device current, and that pointer is what
//    the kernels receive;
//  * every GPU that loads the library holds its own copy of the array (NA_SCRATCH_BYTES) until the process exits;
//  * every call on a device shares that device's copy, which keeps the previous call's records. Nothing has to start at
//    zero: k_allpairs_best2 stores every record [chunk][query < nq] that k_allpairs_merge reads, in the same stream
//    order, so stale contents are never read. Calls that overlap on one GPU (two streams, two host threads) race on it.
#include <mutex>

#include "../../include/ccm_match.h"
#include "match_common.cuh"

#define NA_MAX_NQ CCM_NOALLOC_MAX_QUERIES
#define NA_MAX_RECORDS (1 << 19)                               // chunk records, 8 B each
#define NA_SCRATCH_BYTES ((size_t)NA_MAX_RECORDS * 8)          // 4 MiB per device
#define NA_MAX_DEVICES 64

static_assert(sizeof(Best2Dev) == 8, "scratch is sized in 8-byte records");

static __device__ __align__(256) unsigned long long g_noalloc_parts[NA_MAX_RECORDS];

namespace {
struct NoallocDevice { Best2Dev* parts = nullptr; int sms = 0; bool ready = false; };
NoallocDevice g_noalloc_dev[NA_MAX_DEVICES];
std::mutex g_noalloc_mu;

int noalloc_fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    g_match_create_error = e == cudaSuccess ? std::string(what) : std::string(what) + ": " + cudaGetErrorString(e);
    return code;
}

// once per device, with that device current: the symbol's address ON THAT DEVICE
int noalloc_setup(int dev, NoallocDevice** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoallocDevice& d = g_noalloc_dev[dev];
    if (!d.ready) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_parts);
        if (e != cudaSuccess || !p) return noalloc_fail(CCM_E_CUDA, "cudaGetSymbolAddress(g_noalloc_parts)", e);
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess || sms <= 0) return noalloc_fail(CCM_E_CUDA, "cudaDeviceGetAttribute(multiProcessorCount)", e);
        d.parts = (Best2Dev*)p; d.sms = sms; d.ready = true;
    }
    *out = &d;
    return CCM_OK;
}
}  // namespace

extern "C" int ccm_match_allpairs_best2_noalloc_device(void* stream, const uint8_t* d_q, int nq, const uint8_t* d_targets,
                                                       long long nt, uint32_t index_base, ccm_best2* d_out) {
    if (nq < 0 || nt < 0 || !d_out || (nq > 0 && !d_q) || (nt > 0 && !d_targets))
        return noalloc_fail(CCM_E_BADARG, "ccm_match_allpairs_best2_noalloc_device: null pointer or negative size");
    if (nq > NA_MAX_NQ)
        return noalloc_fail(CCM_E_CAPACITY, "ccm_match_allpairs_best2_noalloc_device: nq exceeds CCM_NOALLOC_MAX_QUERIES (4096); "
                                            "use ccm_match_allpairs_best2_device or split the queries");
    if (nq == 0) return CCM_OK;
    if (((uintptr_t)d_q | (uintptr_t)d_targets) & 15)
        return noalloc_fail(CCM_E_BADARG, "ccm_match_allpairs_best2_noalloc_device: descriptor arrays must be 16-byte aligned");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(CCM_E_CUDA, "cudaGetDevice (there is no CPU fallback)", e);
    if (dev < 0 || dev >= NA_MAX_DEVICES) return noalloc_fail(CCM_E_BADARG, "ccm_match_allpairs_best2_noalloc_device: device ordinal >= 64");
    NoallocDevice* D = nullptr;
    int rc = noalloc_setup(dev, &D);
    if (rc) return rc;

    // same partition as allpairs_chunks (match_api.cu), with the chunk count capped by the fixed scratch; the merge is
    // associative and order-independent, so the partition does not change the result
    const int Q = nq >= 896 ? 4 : (nq >= 384 ? 2 : 1);
    const int q_tiles = (nq + AP_THREADS * Q - 1) / (AP_THREADS * Q);
    long long want = std::max<long long>(1, ((long long)D->sms * 3 + q_tiles - 1) / q_tiles);
    want = std::min<long long>(want, NA_MAX_RECORDS / nq);
    long long chunk = std::max<long long>(AP_TILE, (nt + want - 1) / want);
    chunk = (chunk + AP_TILE - 1) / AP_TILE * AP_TILE;
    chunk = std::min<long long>(chunk, AP_MAX_CHUNK);
    const long long n_chunks = std::max<long long>(1, (nt + chunk - 1) / chunk);
    if (n_chunks * nq > NA_MAX_RECORDS || n_chunks > 65535)
        return noalloc_fail(CCM_E_CAPACITY, "ccm_match_allpairs_best2_noalloc_device: nt too large for the fixed scratch "
                                            "(nt <= floor(524288 / nq) * 2^23); use ccm_match_allpairs_best2_device");

    cudaStream_t s = (cudaStream_t)stream;
    Best2Dev* parts = D->parts;                                   // the DEVICE address, not the host-side symbol
    const dim3 grid((unsigned)q_tiles, (unsigned)n_chunks);
    if (Q == 4) k_allpairs_best2<4><<<grid, AP_THREADS, 0, s>>>((const uint4*)d_q, nq, (const uint4*)d_targets, nt, (int)chunk, index_base, parts);
    else if (Q == 2) k_allpairs_best2<2><<<grid, AP_THREADS, 0, s>>>((const uint4*)d_q, nq, (const uint4*)d_targets, nt, (int)chunk, index_base, parts);
    else k_allpairs_best2<1><<<grid, AP_THREADS, 0, s>>>((const uint4*)d_q, nq, (const uint4*)d_targets, nt, (int)chunk, index_base, parts);
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(CCM_E_CUDA, "k_allpairs_best2 launch", e);
    k_allpairs_merge<<<(nq + 255) / 256, 256, 0, s>>>(parts, (int)n_chunks, nq, 0.f, CCM_TH_LOW, (Best2Dev*)d_out, nullptr);
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(CCM_E_CUDA, "k_allpairs_merge launch", e);
    return CCM_OK;
}
