// Keyframe wire format (client -> server): the keypoint and descriptor arrays of ccmslam_msgs/KF as ROS 1 serialises them.
// KeyFrame::ConvertToMessage fills them element by element (cslam/src/KeyFrame.cpp:1462-1472, Converter::toCvKeyPointMsg
// cslam/src/Converter.cc:166-179) and KeyFrame::WriteMembersFromMessage reads them back (KeyFrame.cpp:1689-1702,
// Converter::fromCvKeyPointMsg Converter.cc:181-193). Layout (cslam_msgs/msg/CvKeyPoint.msg, Descriptor.msg; ROS 1
// serialisation = fields in declaration order, little-endian, unpadded, variable-length arrays prefixed by a u32 count):
//   CvKeyPoint[]: u32 n | n x { f32 x, f32 y, u8 size, f32 angle, u8 response, i8 octave }   (15 bytes each)
//   Descriptor[]: u32 n | n x u8[32]
// size and response are truncated to u8 on the wire ((u_int8_t)float), octave to i8: server-side keyframes carry the lossy
// values, exactly as here. Byte work bounded by HBM; one CTA stages 256 records through shared memory so that global
// traffic is aligned 32-bit words although the records are 15 bytes.
#include "match_common.cuh"

#define WIRE_KP_BYTES 15
#define WIRE_CTA 256                                        // records per CTA: 3840 bytes, a multiple of 4

__global__ void __launch_bounds__(WIRE_CTA) k_wire_encode(const ccm_keypoint_dev* __restrict__ kp, const float2* __restrict__ xy_un,
                                                          const uint32_t* __restrict__ desc, int n,
                                                          uint32_t* __restrict__ kp_bytes, uint32_t* __restrict__ desc_bytes) {
    __shared__ __align__(16) uint8_t rec[WIRE_CTA * WIRE_KP_BYTES];
    const int base = blockIdx.x * WIRE_CTA, i = base + threadIdx.x;
    if (blockIdx.x == 0 && threadIdx.x == 0) { kp_bytes[0] = (uint32_t)n; desc_bytes[0] = (uint32_t)n; }
    if (i < n) {
        const ccm_keypoint_dev k = kp[i];
        const float2 p = xy_un ? xy_un[i] : make_float2(k.x, k.y);                 // mvKeysUn[i].pt
        uint8_t* r = rec + threadIdx.x * WIRE_KP_BYTES;
        const uint32_t xb = __float_as_uint(p.x), yb = __float_as_uint(p.y), ab = __float_as_uint(k.angle);
#pragma unroll
        for (int b = 0; b < 4; b++) { r[b] = (uint8_t)(xb >> (8 * b)); r[4 + b] = (uint8_t)(yb >> (8 * b)); r[9 + b] = (uint8_t)(ab >> (8 * b)); }
        r[8] = (uint8_t)(int)k.size;                                                // (u_int8_t)kp.size: truncation toward zero
        r[13] = (uint8_t)(int)k.response;
        r[14] = (uint8_t)(int8_t)k.octave;
    }
    __syncthreads();
    const int cnt = min(n - base, WIRE_CTA);
    const int words = (cnt * WIRE_KP_BYTES + 3) / 4;                                // the buffer is padded to a whole word
    uint32_t* out = kp_bytes + 1 + (size_t)base * WIRE_KP_BYTES / 4;
    for (int w = threadIdx.x; w < words; w += WIRE_CTA) {
        uint32_t v = reinterpret_cast<const uint32_t*>(rec)[w];
        const int valid = cnt * WIRE_KP_BYTES - 4 * w;                              // bytes of this word that belong to records
        if (valid < 4) v &= (1u << (8 * valid)) - 1u;
        out[w] = v;
    }
    // descriptors: 8 words per row, word-for-word
    uint32_t* dout = desc_bytes + 1 + (size_t)base * 8;
    const uint32_t* din = desc + (size_t)base * 8;
    for (int w = threadIdx.x; w < cnt * 8; w += WIRE_CTA) dout[w] = din[w];
}

struct WireDecodeOut { float* x; float* y; float* angle; int* octave; uint32_t* desc; ccm_keypoint_dev* kp; int* n; };
__global__ void __launch_bounds__(WIRE_CTA) k_wire_decode(const uint32_t* __restrict__ kp_bytes, const uint32_t* __restrict__ desc_bytes, int n,
                                                          const WireDecodeOut O) {
    __shared__ __align__(16) uint8_t rec[WIRE_CTA * WIRE_KP_BYTES];
    const int base = blockIdx.x * WIRE_CTA, i = base + threadIdx.x;
    const int cnt = min(n - base, WIRE_CTA);
    const int words = (cnt * WIRE_KP_BYTES + 3) / 4;
    const uint32_t* in = kp_bytes + 1 + (size_t)base * WIRE_KP_BYTES / 4;
    for (int w = threadIdx.x; w < words; w += WIRE_CTA) reinterpret_cast<uint32_t*>(rec)[w] = in[w];
    if (blockIdx.x == 0 && threadIdx.x == 0) *O.n = n;
    __syncthreads();
    if (i < n) {
        const uint8_t* r = rec + threadIdx.x * WIRE_KP_BYTES;
        uint32_t xb = 0, yb = 0, ab = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) { xb |= (uint32_t)r[b] << (8 * b); yb |= (uint32_t)r[4 + b] << (8 * b); ab |= (uint32_t)r[9 + b] << (8 * b); }
        ccm_keypoint_dev k;
        k.x = __uint_as_float(xb); k.y = __uint_as_float(yb); k.angle = __uint_as_float(ab);
        k.size = (float)r[8]; k.response = (float)r[13]; k.octave = (int)(int8_t)r[14];   // Converter.cc:185-190
        O.x[i] = k.x; O.y[i] = k.y; O.angle[i] = k.angle; O.octave[i] = k.octave;
        if (O.kp) O.kp[i] = k;
    }
    const uint32_t* din = desc_bytes + 1 + (size_t)base * 8;
    uint32_t* dout = O.desc + (size_t)base * 8;
    for (int w = threadIdx.x; w < cnt * 8; w += WIRE_CTA) dout[w] = din[w];
}
