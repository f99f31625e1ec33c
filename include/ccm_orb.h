/* libccm_orb_cuda.so — C ABI of the B200 (sm_100a) ORB front-end for CCM-SLAM.
 *
 * The reference (taiyaki-go/motioncheck_ccm_slam) has no FFI layer for this path: the boundary is the C++ class
 * cslam::ORBextractor (cslam/include/cslam/ORBextractor.h:97-164) called from Frame::ExtractORB
 * (cslam/src/Frame.cpp:120-123), and cslam::ORBmatcher (cslam/include/cslam/ORBmatcher.h:77-132).
 * The drop-in C++ classes in motioncheck_ccm_slam_b200/host/ keep those declarations and forward to the entry
 * points below (INTEGRATION.md shows the binding). Plain pointers and sizes only; no exceptions cross this ABI;
 * there is no CPU fallback: every entry point fails with CCM_E_CUDA when no sm_100 device is usable.
 */
#ifndef CCM_ORB_H_
#define CCM_ORB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCM_OK 0
#define CCM_E_BADARG (-1)     /* null pointer, unsupported size or parameter */
#define CCM_E_CAPACITY (-2)   /* caller buffer or handle capacity too small */
#define CCM_E_CUDA (-3)       /* CUDA runtime failure (message via ccm_last_error) */

/* mirrors cv::KeyPoint minus class_id (always -1 in the reference); 24 bytes */
typedef struct ccm_keypoint {
    float x, y;      /* pt, level-0 coordinates (ORBextractor.cpp:1268-1274) */
    float size;      /* (int)(31 * scale[octave]) as float (:1006,1015) */
    float angle;     /* degrees in [0,360), cv::fastAtan2 of the intensity centroid (:94) */
    float response;  /* FAST-9/16 corner score */
    int octave;      /* pyramid level */
} ccm_keypoint;

/* replaces the arguments of ORBextractor::ORBextractor (ORBextractor.h:103-105, ORBextractor.cpp:579-639)
 * plus the arena sizes the handle allocates once (no allocation on the per-frame path). */
typedef struct ccm_orb_config {
    int nfeatures;
    float scale_factor;
    int nlevels;        /* 1..16 */
    int ini_th_fast;
    int min_th_fast;
    int max_width;      /* largest image the handle accepts (<= 4095) */
    int max_height;
    int max_batch;      /* frames in flight per call */
    int device;         /* CUDA device ordinal (agent -> GPU pinning) */
} ccm_orb_config;

typedef struct ccm_orb ccm_orb;   /* opaque extractor handle: one CUDA stream + one device arena */

/* ORBextractor::ORBextractor */
int ccm_orb_create(const ccm_orb_config* cfg, ccm_orb** out);
/* ORBextractor::~ORBextractor */
void ccm_orb_destroy(ccm_orb* h);
/* last error text of this handle (or of the failed create when h is NULL); never NULL */
const char* ccm_last_error(const ccm_orb* h);

/* GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares / GetInverseScaleSigmaSquares
 * (ORBextractor.h:118-136) and mnFeaturesPerLevel; each array has nlevels entries; any pointer may be NULL. */
int ccm_orb_get_tables(const ccm_orb* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* features_per_level);
/* upper bound on keypoints per frame: sum over levels of max(quota+3, 4*nIni) for the handle's max size */
int ccm_orb_max_keypoints(const ccm_orb* h);

/* ORBextractor::operator() (ORBextractor.cpp:1216-1278) on one host image (CV_8UC1, arbitrary step).
 * Writes min(*out_n, capacity) keypoints and 32-byte descriptors; *out_n is the full count.
 * rows==0 or cols==0 or img==NULL: returns CCM_OK with *out_n = 0 and outputs untouched (:1219-1220). */
int ccm_orb_extract(ccm_orb* h, const uint8_t* img, int rows, int cols, size_t step,
                    ccm_keypoint* out_kp, uint8_t* out_desc, int capacity, int* out_n);

/* Same for n_imgs (<= max_batch) host images of identical size in one submission.
 * out_kp is [n_imgs][capacity], out_desc [n_imgs][capacity][32], out_n [n_imgs]. */
int ccm_orb_extract_batch(ccm_orb* h, const uint8_t* const* imgs, int n_imgs, int rows, int cols, size_t step,
                          ccm_keypoint* out_kp, uint8_t* out_desc, int capacity, int* out_n);

/* The same call as an asynchronous pair, for a stream of batches: submit enqueues copy-in, kernels and copy-out and
 * returns a ticket; wait blocks until that batch's results are in out_kp / out_desc / out_n. Two batches may be in
 * flight (submit k+1, then wait k): the copy-in of one overlaps the kernels of the other, so the link and the SMs
 * work at the same time. The frames in flight together must fit max_batch (each batch occupies its own frame slots
 * of the arena), the image shape must not change while a batch is in flight, imgs and the output arrays must stay
 * valid until wait returns, and batches complete in submission order. When out_kp / out_desc are pinned host memory
 * and capacity == ccm_orb_max_keypoints(h), results are copied straight into them (whole rows; only the first
 * out_n[f] entries of a row are meaningful). ccm_orb_extract_batch == submit + wait. Do not interleave with
 * ccm_orb_extract_batch_device on the same handle while a ticket is open. */
int ccm_orb_extract_batch_submit(ccm_orb* h, const uint8_t* const* imgs, int n_imgs, int rows, int cols, size_t step,
                                 ccm_keypoint* out_kp, uint8_t* out_desc, int capacity, int* out_n, int* ticket);
int ccm_orb_extract_batch_wait(ccm_orb* h, int ticket);

/* Device-resident variant: images already in HBM at d_imgs + f*frame_stride (row pitch `pitch`, both multiples
 * of 4 bytes); results stay in HBM (d_kp [n][capacity], d_desc [n][capacity][32], d_n [n]) so the matcher can
 * consume them without a host round trip. Asynchronous on the handle's stream; pair with ccm_orb_sync. */
int ccm_orb_extract_batch_device(ccm_orb* h, const uint8_t* d_imgs, int n_imgs, int rows, int cols, size_t pitch,
                                 size_t frame_stride, ccm_keypoint* d_kp, uint8_t* d_desc, int capacity, int* d_n);
int ccm_orb_sync(ccm_orb* h);
/* the handle's cudaStream_t, for ordering caller work after/before the extractor's */
void* ccm_orb_stream(ccm_orb* h);
/* kernels launched by this handle since creation (bench.py's gpu_launches) */
long long ccm_orb_launch_count(const ccm_orb* h);

/* Per-stage device time of the extractor pipeline, measured with CUDA events on the handle's stream (bench.py's
 * live roofline figure). Stages: 0 pyramid resize (K1), 1 FAST cells (K2+K3), 2 Gaussian blur (K6),
 * 3 quadtree (K4), 4 orientation + rBRIEF (K5+K7). sum_ms[CCM_N_STAGES] accumulates over *calls batches. */
#define CCM_N_STAGES 5
int ccm_orb_set_stage_timing(ccm_orb* h, int enable);
int ccm_orb_get_stage_ms(ccm_orb* h, double* sum_ms, long long* calls, int reset);

/* mvImagePyramid[level] of frame `frame` of the last call (ORBextractor.h:138). with_border != 0 adds the 19-px
 * BORDER_REFLECT_101 apron (ORBextractor.cpp:1295,1300). out has out_step bytes per row. */
int ccm_orb_get_pyramid_level(ccm_orb* h, int frame, int level, int with_border, uint8_t* out, size_t out_step,
                              int* out_w, int* out_h);

/* ---- stage taps for parity tests (not part of the reference interface) ---- */
/* blurred level (GaussianBlur of the level clone, :1258-1259) */
int ccm_orb_debug_blurred_level(ccm_orb* h, int frame, int level, uint8_t* out, size_t out_step, int* out_w, int* out_h);
/* FAST corners handed to the quadtree (vToDistributeKeys, :947-997), sorted into the reference's order;
 * xys = int[3*capacity] (x, y, score) in level coordinates; *out_n = full count */
int ccm_orb_debug_level_corners(ccm_orb* h, int frame, int level, int* xys, int capacity, int* out_n);

#ifdef __cplusplus
}
#endif
#endif /* CCM_ORB_H_ */
