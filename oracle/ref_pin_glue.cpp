/* TEST INFRASTRUCTURE — C entry points over oracle/_ref/gen_pin.cpp (line ranges of the reference compiled as they are,
 * see pinshim/pin_prelude.h), so that tests can hold oracle/match_oracle.cpp's mo_distinctive_descriptors and
 * mo_wire_* against the reference's own statements. */
#include <cstring>
#include "pinshim/pin_prelude.h"

using namespace cslam;

extern "C" {

/* MapPoint::ComputeDistinctiveDescriptors for n_points map points; point p observes rows [start[p], start[p+1]) of desc,
 * each from its own keyframe. The observation map is keyed by shared_ptr, i.e. ordered by keyframe ADDRESS: the keyframes
 * of a point live in one array here so that the reference walks them in row order (upstream the order is whatever the
 * heap gave). kf_bad [rows] (may be NULL) marks keyframes with isBad(). The observed row of each keyframe sits at a
 * different index of that keyframe's descriptor matrix, as upstream (mit->second).
 * out_desc [n_points][32] = mDescriptor afterwards, out_set [n_points] = 1 when the function assigned it. */
void pin_distinctive_descriptors(int n_points, const int* start, const uint8_t* desc, const uint8_t* kf_bad, uint8_t* out_desc, int* out_set) {
    for (int p = 0; p < n_points; p++) {
        const int r0 = start[p], N = start[p + 1] - r0;
        std::vector<KeyFrame> kfs(N > 0 ? N : 0);
        MapPoint mp;
        for (int i = 0; i < N; i++) {
            const int idx = (r0 + i) % 5;                              /* the observed feature's index in that keyframe */
            kfs[i].mDescriptors = cv::Mat(idx + 2, 32, CV_8U);
            for (int r = 0; r < idx + 2; r++) std::memset(kfs[i].mDescriptors.ptr(r), 0xA5 ^ r, 32);
            std::memcpy(kfs[i].mDescriptors.ptr(idx), desc + (size_t)32 * (r0 + i), 32);
            kfs[i].bad = kf_bad && kf_bad[r0 + i];
            mp.mObservations[MapPoint::kfptr(&kfs[i], [](KeyFrame*) {})] = (size_t)idx;
        }
        mp.ComputeDistinctiveDescriptors();
        out_set[p] = !mp.mDescriptor.empty();
        if (out_set[p]) std::memcpy(out_desc + (size_t)32 * p, mp.mDescriptor.ptr(0), 32);
    }
}

/* ROS 1 serialisation of CvKeyPoint[] and Descriptor[] (roscpp serialization.h: a variable-length array is a u32 count
 * followed by its elements; a message is its fields in declaration order, little-endian, unpadded; a fixed-size array
 * has no count). The messages are filled by the reference's own KeyFrame::ConvertToMessage lines. */
void pin_wire_encode(int n, const float* x, const float* y, const float* size, const float* angle, const float* response, const int* octave,
                     const uint8_t* desc, uint8_t* kp_bytes, uint8_t* desc_bytes) {
    KeyFrame kf;
    kf.mvKeysUn.resize(n);
    kf.mDescriptors = cv::Mat(n, 32, CV_8U);
    for (int i = 0; i < n; i++) {
        kf.mvKeysUn[i] = cv::KeyPoint(x[i], y[i], size[i], angle[i], response[i], octave[i]);
        std::memcpy(kf.mDescriptors.ptr(i), desc + (size_t)32 * i, 32);
    }
    ccmslam_msgs::KF Msg;
    kf.PinConvertToMessage(Msg);
    uint32_t cnt = (uint32_t)Msg.mvKeysUn.size();
    std::memcpy(kp_bytes, &cnt, 4);
    uint8_t* w = kp_bytes + 4;
    for (const ccmslam_msgs::CvKeyPoint& k : Msg.mvKeysUn) {
        std::memcpy(w, &k.fPoint2f_x, 4); w += 4;
        std::memcpy(w, &k.fPoint2f_y, 4); w += 4;
        std::memcpy(w, &k.size, 1); w += 1;
        std::memcpy(w, &k.angle, 4); w += 4;
        std::memcpy(w, &k.response, 1); w += 1;
        std::memcpy(w, &k.octave, 1); w += 1;
    }
    cnt = (uint32_t)Msg.mDescriptors.size();
    std::memcpy(desc_bytes, &cnt, 4);
    w = desc_bytes + 4;
    for (const ccmslam_msgs::Descriptor& d : Msg.mDescriptors) { std::memcpy(w, d.mDescriptor.data(), 32); w += 32; }
}

/* the inverse: deserialise, then the reference's own KeyFrame::WriteMembersFromMessage lines. Returns n (>= 1: upstream
 * dereferences mDescriptors[0]), or -1 when the counts disagree or n == 0. */
int pin_wire_decode(const uint8_t* kp_bytes, const uint8_t* desc_bytes, int cap, float* x, float* y, float* size, float* angle, float* response,
                    int* octave, uint8_t* desc) {
    uint32_t n, nd;
    std::memcpy(&n, kp_bytes, 4);
    std::memcpy(&nd, desc_bytes, 4);
    if (n != nd || n == 0 || (int)n > cap) return -1;
    ccmslam_msgs::KF Msg;
    Msg.mvKeysUn.resize(n); Msg.mDescriptors.resize(n);
    const uint8_t* r = kp_bytes + 4;
    for (uint32_t i = 0; i < n; i++) {
        ccmslam_msgs::CvKeyPoint& k = Msg.mvKeysUn[i];
        std::memcpy(&k.fPoint2f_x, r, 4); r += 4;
        std::memcpy(&k.fPoint2f_y, r, 4); r += 4;
        std::memcpy(&k.size, r, 1); r += 1;
        std::memcpy(&k.angle, r, 4); r += 4;
        std::memcpy(&k.response, r, 1); r += 1;
        std::memcpy(&k.octave, r, 1); r += 1;
        std::memcpy(Msg.mDescriptors[i].mDescriptor.data(), desc_bytes + 4 + (size_t)32 * i, 32);
    }
    KeyFrame kf;
    kf.PinWriteMembersFromMessage(&Msg);
    for (uint32_t i = 0; i < n; i++) {
        const cv::KeyPoint& k = kf.mvKeysUn[i];
        x[i] = k.pt.x; y[i] = k.pt.y; size[i] = k.size; angle[i] = k.angle; response[i] = k.response; octave[i] = k.octave;
        std::memcpy(desc + (size_t)32 * i, kf.mDescriptors.ptr((int)i), 32);
    }
    return (int)n;
}

}  // extern "C"
