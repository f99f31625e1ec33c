// Drop-in replacement for cslam/src/ORBextractor.cpp of taiyaki-go/motioncheck_ccm_slam: the SAME class declaration
// (this file includes the reference's own, unmodified <cslam/ORBextractor.h>), bodies forward to the CUDA library
// through the C ABI of include/ccm_orb.h. Frame.cpp, Tracking.cpp and everything else compile unchanged against it.
// Build it INSTEAD of cslam/src/ORBextractor.cpp and link libccm_orb_cuda.so.
//
// The reference class has no spare member for a handle and its destructor is inline in the header, and the object is
// copied into every Frame (Frame.cpp:37,56), so the device handle lives in a process-wide registry keyed by the
// constructor arguments + device + image shape: copies of an extractor share one handle, the normal and the "ini" (2x
// features) extractor get one each (they are never used concurrently, Tracking.cpp:75,115), a process that alternates two
// image sizes keeps one arena per size instead of re-allocating per frame, and handles live until process exit.
// A CUDA failure prints the library's message and throws estd::infrastructure_ex (the reference's fatal-error convention,
// estd.h:66-74, as KILLSYS does) — there is no CPU fallback.
#include <cslam/ORBextractor.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>

#include "ccm_orb.h"

namespace {

struct Shared {
    ccm_orb* h = nullptr;
    std::mutex mu;                       // the reference extractor is not re-entrant either (it mutates mvImagePyramid)
    std::vector<ccm_keypoint> kp;
    std::vector<unsigned char> desc;
};

[[noreturn]] void fatal(const char* what, const ccm_orb* h) {
    std::fprintf(stderr, "ORBextractor (CUDA): %s: %s\n", what, ccm_last_error(h));
    throw estd::infrastructure_ex();
}

typedef std::tuple<int, float, int, int, int, int, int, int> Key;

int device_from_env() {                  // agent -> GPU pinning: CCM_ORB_DEVICE, default 0
    const char* e = std::getenv("CCM_ORB_DEVICE");
    return e ? std::atoi(e) : 0;
}

Shared& shared_for(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int rows, int cols) {
    static std::mutex reg_mu;
    static std::map<Key, Shared*> reg;
    std::lock_guard<std::mutex> g(reg_mu);
    Shared*& s = reg[Key(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device_from_env(), rows, cols)];
    if (!s) s = new Shared();
    return *s;
}

// create the handle of this (configuration, shape) when its first image arrives
void ensure_handle(Shared& s, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int rows, int cols) {
    if (s.h) return;
    ccm_orb_config cfg;
    cfg.nfeatures = nfeatures; cfg.scale_factor = scaleFactor; cfg.nlevels = nlevels;
    cfg.ini_th_fast = iniThFAST; cfg.min_th_fast = minThFAST;
    cfg.max_width = cols; cfg.max_height = rows;
    cfg.max_batch = 1; cfg.device = device_from_env();
    if (ccm_orb_create(&cfg, &s.h) != CCM_OK) fatal("ccm_orb_create", nullptr);
    const int cap = ccm_orb_max_keypoints(s.h);
    s.kp.resize(cap);
    s.desc.resize((size_t)cap * 32);
}

}  // namespace

namespace cslam {

// cslam/src/ORBextractor.cpp:579-639. The scale tables are computed exactly as there (they are also what the device
// side uses: ccm_orb_create follows the same lines); pattern/umax are kept for the protected interface.
ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST) {
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;
        mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
        mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    mvImagePyramid.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    float factor = 1.0f / scaleFactor;
    float nDesiredFeaturesPerScale = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sumFeatures = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        mnFeaturesPerLevel[level] = cvRound(nDesiredFeaturesPerScale);
        sumFeatures += mnFeaturesPerLevel[level];
        nDesiredFeaturesPerScale *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);
    // umax of the 15-px orientation patch (:623-638); the device side holds the same table as a constant
    const int HALF_PATCH_SIZE = 15;
    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = cvFloor(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
    int vmin = cvCeil(HALF_PATCH_SIZE * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRound(sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// cslam/src/ORBextractor.cpp:1216-1278
void ORBextractor::operator()(InputArray _image, InputArray /*_mask: ignored upstream too, ORBextractor.h:109-111*/,
                              vector<KeyPoint>& _keypoints, OutputArray _descriptors) {
    if (_image.empty()) return;                                        // :1219-1220, outputs untouched
    Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);                                   // :1223
    Shared& s = shared_for(nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, image.rows, image.cols);
    std::lock_guard<std::mutex> g(s.mu);
    ensure_handle(s, nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, image.rows, image.cols);
    int n = 0;
    const int cap = (int)s.kp.size();
    if (ccm_orb_extract(s.h, image.data, image.rows, image.cols, (size_t)image.step, s.kp.data(), s.desc.data(), cap, &n) != CCM_OK)
        fatal("ccm_orb_extract", s.h);
    if (n > cap) fatal("more keypoints than ccm_orb_max_keypoints (internal)", s.h);      // never truncate silently
    _keypoints.clear();
    _keypoints.reserve(n);
    if (n == 0) {
        _descriptors.release();                                        // :1237-1238
    } else {
        _descriptors.create(n, 32, CV_8U);                             // :1241-1242
        Mat descriptors = _descriptors.getMat();
        for (int i = 0; i < n; i++) {
            const ccm_keypoint& k = s.kp[i];
            _keypoints.push_back(KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave));
            std::memcpy(descriptors.ptr(i), &s.desc[(size_t)32 * i], 32);
        }
    }
    // mvImagePyramid (ORBextractor.h:138) has no reader outside the extractor in the reference; it is materialised
    // (level ROI inside its 19-px BORDER_REFLECT_101 apron, :1286-1300) only on request.
    static const bool fill = std::getenv("CCM_ORB_FILL_PYRAMID") != nullptr;
    if (fill) {
        for (int level = 0; level < nlevels; level++) {
            const float scale = mvInvScaleFactor[level];
            const int w = cvRound((float)image.cols * scale), h = cvRound((float)image.rows * scale);
            Mat temp(h + 38, w + 38, CV_8UC1);
            int ow = 0, oh = 0;
            if (ccm_orb_get_pyramid_level(s.h, 0, level, 1, temp.data, (size_t)temp.step, &ow, &oh) != CCM_OK)
                fatal("ccm_orb_get_pyramid_level", s.h);
            mvImagePyramid[level] = temp(Rect(19, 19, w, h));
        }
    }
}

}  // namespace cslam
