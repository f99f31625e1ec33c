/* TEST INFRASTRUCTURE — C wrapper around the reference's OWN cslam::ORBextractor, compiled UNMODIFIED
 * from /root/reference/cslam/src/ORBextractor.cpp against oracle/cvshim (see oracle/Makefile, target
 * _ref/libref_orb.so). Nothing here is shipped or called by the product path.
 *
 * Two builds:
 *   libref_orb.so       glibc malloc — the reference exactly as it behaves in a process.
 *   libref_orb_mono.so  -DCCM_REF_MONOTONE_NEW: operator new hands out monotonically increasing
 *                       addresses during a call, which turns the heap-address tie-break in
 *                       DistributeOctTree's sort (ORBextractor.cpp:852, pair<int,ExtractorNode*>)
 *                       into "later-created node first" — the deterministic rule the oracle
 *                       restatement and the CUDA kernel implement ("quadtree tie rule").
 */
#include <cslam/ORBextractor.h>

#include <cstdio>
#include <cstdlib>
#include <new>

#ifdef CCM_REF_MONOTONE_NEW
#include <sys/mman.h>
namespace {
const size_t kArenaBytes = (size_t)8 << 30; /* virtual reservation; only touched pages are resident */
char* g_arena = nullptr;
size_t g_top = 0;
thread_local bool g_bump = false;
void arena_reset() {
    if (!g_arena) {
        g_arena = (char*)mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (g_arena == (char*)MAP_FAILED) { std::perror("mmap"); std::abort(); }
    } else if (g_top) {
        madvise(g_arena, (g_top + 4095) & ~(size_t)4095, MADV_DONTNEED);
    }
    g_top = 0;
}
inline bool in_arena(void* p) { return g_arena && (char*)p >= g_arena && (char*)p < g_arena + kArenaBytes; }
}  // namespace
void* operator new(size_t n) {
    if (g_bump) {
        size_t a = (g_top + 15) & ~(size_t)15;
        if (a + n > kArenaBytes) { std::fprintf(stderr, "ref arena exhausted\n"); std::abort(); }
        g_top = a + n;
        return g_arena + a;
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }
#endif

namespace {
struct RefHandle {
    int nfeatures, nlevels, ini, min;
    float scale;
    cslam::ORBextractor* ex; /* persistent extractor (glibc build only) */
    std::vector<std::vector<unsigned char>> pyr;
    std::vector<int> pw, ph;
};
struct RefKp { float x, y, size, angle, response; int octave; };

int run(RefHandle* h, cslam::ORBextractor& ex, const unsigned char* img, int rows, int cols, size_t step,
        RefKp* kp, unsigned char* desc, int cap) {
    cv::Mat image(rows, cols, CV_8UC1, (void*)img, step);
    std::vector<cv::KeyPoint> keys;
    cv::Mat d;
    ex(image, cv::Mat(), keys, d);
    int n = (int)keys.size();
    for (int i = 0; i < n && i < cap; i++) {
        kp[i].x = keys[i].pt.x; kp[i].y = keys[i].pt.y; kp[i].size = keys[i].size;
        kp[i].angle = keys[i].angle; kp[i].response = keys[i].response; kp[i].octave = keys[i].octave;
        std::memcpy(desc + (size_t)32 * i, d.ptr(i), 32);
    }
    h->pyr.resize(h->nlevels); h->pw.resize(h->nlevels); h->ph.resize(h->nlevels);
    for (int l = 0; l < h->nlevels; l++) {
        const cv::Mat& m = ex.mvImagePyramid[l];
        h->pw[l] = m.cols; h->ph[l] = m.rows;
        h->pyr[l].resize((size_t)m.cols * m.rows);
        for (int y = 0; y < m.rows; y++) std::memcpy(h->pyr[l].data() + (size_t)y * m.cols, m.ptr(y), m.cols);
    }
    return n;
}
}  // namespace

extern "C" {

void* ref_orb_create(int nfeatures, float scale, int nlevels, int ini, int min) {
    RefHandle* h = (RefHandle*)std::malloc(sizeof(RefHandle));
    new (h) RefHandle();
    h->nfeatures = nfeatures; h->scale = scale; h->nlevels = nlevels; h->ini = ini; h->min = min;
#ifdef CCM_REF_MONOTONE_NEW
    h->ex = nullptr;
#else
    h->ex = new cslam::ORBextractor(nfeatures, scale, nlevels, ini, min);
#endif
    return h;
}

void ref_orb_destroy(void* p) {
    RefHandle* h = (RefHandle*)p;
    if (!h) return;
    delete h->ex;
    h->~RefHandle();
    std::free(h);
}

/* returns the number of keypoints (may exceed cap; only cap are written) */
int ref_orb_extract(void* p, const unsigned char* img, int rows, int cols, size_t step,
                    void* kp_out, unsigned char* desc_out, int cap) {
    RefHandle* h = (RefHandle*)p;
#ifdef CCM_REF_MONOTONE_NEW
    /* results are copied into caller memory and into h->pyr; h->pyr is pre-sized OUTSIDE bump mode */
    int n;
    {
        cslam::ORBextractor probe(h->nfeatures, h->scale, h->nlevels, h->ini, h->min);
        h->pyr.resize(h->nlevels); h->pw.resize(h->nlevels); h->ph.resize(h->nlevels);
        for (int l = 0; l < h->nlevels; l++) {
            float s = probe.GetInverseScaleFactors()[l];
            h->pyr[l].resize((size_t)cvRound((float)cols * s) * cvRound((float)rows * s));
        }
    }
    arena_reset();
    g_bump = true;
    {
        cslam::ORBextractor ex(h->nfeatures, h->scale, h->nlevels, h->ini, h->min);
        n = run(h, ex, img, rows, cols, step, (RefKp*)kp_out, desc_out, cap);
    }
    g_bump = false;
    return n;
#else
    return run(h, *h->ex, img, rows, cols, step, (RefKp*)kp_out, desc_out, cap);
#endif
}

/* pyramid level of the last extract call (ROI only, contiguous w*h) */
int ref_orb_pyramid_level(void* p, int level, unsigned char* out, int cap, int* w, int* h_out) {
    RefHandle* h = (RefHandle*)p;
    if (level < 0 || level >= (int)h->pyr.size()) return -1;
    *w = h->pw[level]; *h_out = h->ph[level];
    if ((int)h->pyr[level].size() > cap) return -2;
    std::memcpy(out, h->pyr[level].data(), h->pyr[level].size());
    return 0;
}

/* scale tables + per-level quotas as the reference constructor computes them (ORBextractor.cpp:579-615).
 * mnFeaturesPerLevel is protected, so it is read through a derived accessor. */
namespace { struct Peek : cslam::ORBextractor { using cslam::ORBextractor::ORBextractor; const std::vector<int>& q() const { return mnFeaturesPerLevel; } const std::vector<int>& um() const { return umax; } }; }
int ref_orb_tables(int nfeatures, float scale, int nlevels, int ini, int min, float* scale_f, float* inv_scale_f,
                   float* sigma2, float* inv_sigma2, int* quota, int* umax16) {
    Peek ex(nfeatures, scale, nlevels, ini, min);
    for (int l = 0; l < nlevels; l++) {
        scale_f[l] = ex.GetScaleFactors()[l]; inv_scale_f[l] = ex.GetInverseScaleFactors()[l];
        sigma2[l] = ex.GetScaleSigmaSquares()[l]; inv_sigma2[l] = ex.GetInverseScaleSigmaSquares()[l];
        quota[l] = ex.q()[l];
    }
    for (int v = 0; v < 16; v++) umax16[v] = ex.um()[v];
    return 0;
}

}  // extern "C"
