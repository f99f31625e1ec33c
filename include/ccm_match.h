/* libccm_orb_cuda.so — C ABI of the B200 (sm_100a) Hamming-distance searches of CCM-SLAM's ORBmatcher.
 *
 * The reference has no FFI layer: the boundary is the C++ class cslam::ORBmatcher
 * (cslam/include/cslam/ORBmatcher.h:77-132, cslam/src/ORBmatcher.cpp). The drop-in class keeps that declaration;
 * each Search* body gathers what the reference gathers on the host (projection geometry u, v, radius, predicted
 * level, the query map points' descriptors, "has a good map point" flags) and calls one entry point below, which
 * returns exactly the state the reference loop leaves behind (who matched whom, in the reference's greedy order,
 * after the rotation-histogram pruning). Plain pointers and sizes; no exceptions; no CPU fallback.
 *
 * Memory spaces: every pointer argument of a function whose name does not end in `_device` is HOST memory; the
 * call is synchronous. `_device` variants take device pointers, enqueue on the matcher's stream and return
 * without waiting (pair with ccm_matcher_sync), so the extractor's output feeds the matcher without a host copy.
 */
#ifndef CCM_MATCH_H_
#define CCM_MATCH_H_

#include <stddef.h>
#include <stdint.h>

#include "ccm_orb.h"

#ifdef __cplusplus
extern "C" {
#endif

#define CCM_TH_HIGH 100      /* ORBmatcher::TH_HIGH, ORBmatcher.cpp:63 */
#define CCM_TH_LOW 50        /* ORBmatcher::TH_LOW,  :64 */
#define CCM_HISTO_LENGTH 30  /* ORBmatcher::HISTO_LENGTH, :65 */

typedef struct ccm_matcher ccm_matcher;         /* one device, one CUDA stream, a grow-only scratch arena */
typedef struct ccm_frame_pack ccm_frame_pack;   /* device copy of what the matcher reads of a Frame / KeyFrame */

int ccm_matcher_create(int device, ccm_matcher** out);
void ccm_matcher_destroy(ccm_matcher* m);
const char* ccm_matcher_last_error(const ccm_matcher* m);   /* m == NULL: error of the failed create */
int ccm_matcher_sync(ccm_matcher* m);
void* ccm_matcher_stream(ccm_matcher* m);                   /* cudaStream_t */
long long ccm_matcher_launch_count(const ccm_matcher* m);

/* static ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1653-1669) for n descriptor pairs: out[i] = |a[i] xor b[i]| */
int ccm_descriptor_distance(ccm_matcher* m, const uint8_t* a, const uint8_t* b, int n, int* out);

/* ---- frame pack: descriptors + undistorted keypoints + the feature grid ------------------------------------
 * Frame: mDescriptors, mvKeysUn, mGrid[75][48] (Frame.h:128-149; AssignFeaturesToGrid/PosInGrid, Frame.cpp:103-118,
 * 255-265); KeyFrame: same fields (KeyFrame.cpp:1162-1201). The grid is built on the device. */
typedef struct ccm_grid_params {
    float min_x, min_y;        /* mnMinX, mnMinY */
    float inv_w, inv_h;        /* mfGridElementWidthInv, mfGridElementHeightInv (Frame.cpp:87-88) */
    int cols, rows;            /* FRAME_GRID_COLS 75, FRAME_GRID_ROWS 48 (Frame.h:51-52); cols*rows <= 8192 */
} ccm_grid_params;

int ccm_frame_pack_create(ccm_matcher* m, int capacity /* <= 65535 features */, ccm_frame_pack** out);
/* May be called before or after ccm_matcher_destroy of the matcher the pack was created on (it does not touch it). */
void ccm_frame_pack_destroy(ccm_frame_pack* p);
/* host arrays of n features: desc [n][32], x/y = mvKeysUn[i].pt, octave, angle (degrees) */
int ccm_frame_pack_upload(ccm_frame_pack* p, const ccm_grid_params* g, int n, const uint8_t* desc, const float* x,
                          const float* y, const int* octave, const float* angle);
/* straight from ccm_orb_extract_batch_device output (device pointers; *d_n features); the keypoints are used as
 * mvKeysUn, i.e. a camera without distortion (Frame.cpp:279-283), or pass d_xy_un [capacity][2] to override them */
int ccm_frame_pack_from_extractor_device(ccm_frame_pack* p, const ccm_grid_params* g, const ccm_keypoint* d_kp,
                                         const uint8_t* d_desc, const int* d_n, const float* d_xy_un);
/* same for a whole extractor batch in two launches: packs[j] takes frame j of ccm_orb_extract_batch_device's output
 * (d_kp [n][kp_stride], d_desc [n][kp_stride][32], d_n [n]; d_xy_un [n][kp_stride][2] or NULL) */
int ccm_frame_pack_from_extractor_batch_device(ccm_frame_pack* const* packs, int n_packs, const ccm_grid_params* g,
                                               const ccm_keypoint* d_kp, const uint8_t* d_desc, const int* d_n, int kp_stride,
                                               const float* d_xy_un);
/* test tap: the grid as CSR over cells in (ix*rows + iy) order: cell_start[cols*rows+1], items[n]; *n_items out */
int ccm_frame_pack_get_grid(ccm_frame_pack* p, int* cell_start, int* items, int* n_items);
int ccm_frame_pack_count(ccm_frame_pack* p, int* n);

/* ---- Frame::UndistortKeyPoints / ComputeImageBounds (Frame.cpp:277-343) ---------------------------------------
 * cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK): five fixed-point iterations of the inverse
 * Brown-Conrady model per point, in double, every operation separately rounded (no FMA contraction), result
 * rounded to float — bit-identical to OpenCV 4.13.0 (tests/golden/undistort.npz). */
typedef struct ccm_camera {
    float fx, fy, cx, cy;      /* mK (Tracking.cpp:38-43) */
    float dist[5];             /* mDistCoef = k1 k2 p1 p2 [k3] (Tracking.cpp:47-58) */
    int n_dist;                /* 4 or 5 */
} ccm_camera;

/* host in/out: xy_out[i] = undistorted xy_in[i] ([n][2] floats). As Frame.cpp:279-283, dist[0] == 0 copies the input. */
int ccm_undistort_points(ccm_matcher* m, const ccm_camera* cam, int n, const float* xy_in, float* xy_out);
/* device, batched over an extractor batch: frame j's d_n[j] keypoints (d_kp + j*kp_stride) -> d_xy_un [n_frames][kp_stride][2],
 * the array the pack builders take as d_xy_un; enqueued on the matcher's stream */
int ccm_undistort_keypoints_batch_device(ccm_matcher* m, const ccm_camera* cam, int n_frames, const ccm_keypoint* d_kp,
                                         const int* d_n, int kp_stride, float* d_xy_un);
/* Frame::ComputeImageBounds + the grid constants of the Frame constructor (Frame.cpp:85-88, 309-343): fills
 * g->min_x, min_y, inv_w, inv_h for a cols x rows image (g->cols, g->rows as given); bounds[4] = mnMinX, mnMaxX,
 * mnMinY, mnMaxY (may be NULL) */
int ccm_image_bounds(ccm_matcher* m, const ccm_camera* cam, int img_cols, int img_rows, ccm_grid_params* g, float* bounds);
/* ccm_frame_pack_from_extractor_batch_device with the undistortion fused into the pack fill (no d_xy_un round trip):
 * pack positions = mvKeysUn of a camera with distortion */
int ccm_frame_pack_from_extractor_batch_device_cam(ccm_frame_pack* const* packs, int n_packs, const ccm_grid_params* g,
                                                   const ccm_camera* cam, const ccm_keypoint* d_kp, const uint8_t* d_desc,
                                                   const int* d_n, int kp_stride);

/* ---- windowed searches (GetFeaturesInArea candidates) -------------------------------------------------------
 * One query per map point / feature the reference loop reaches; the loop's host-side early `continue`s are the
 * valid flag. Query order is the reference's loop order: results are those of the sequential greedy loop. */
typedef struct ccm_window_query {
    float u, v;            /* projected position */
    float r;               /* search radius as passed to GetFeaturesInArea */
    float angle;           /* query keypoint angle (rotation histogram), degrees */
    int16_t min_level;     /* level window: candidates with octave < min_level are skipped; */
    int16_t max_level;     /* with max_level >= 0, octave > max_level too (Frame.cpp:236-244; KeyFrame callers :397-400) */
    uint32_t flags;        /* CCM_Q_* */
} ccm_window_query;
#define CCM_Q_VALID 1u     /* the reference reaches the candidate loop for this query */
#define CCM_Q_OBS 2u       /* PROJ_MAPPOINTS/LASTFRAME: pMP->Observations()>0 (a later query is blocked by this claim) */
#define CCM_Q_CLAIMS 2u    /* PROJ_KF_SIM3: pMP->GetIndexInKeyFrame(pKF)==-1, the accept writes vpMatched (:436-440) */

typedef struct ccm_match_result {
    int best_idx;          /* -1: no candidate */
    int best_dist, best_dist2;
    int accepted;
} ccm_match_result;

#define CCM_WIN_PROJ_MAPPOINTS 0   /* SearchByProjection(Frame&, vector<mpptr>&, th)           ORBmatcher.cpp:71-148   */
#define CCM_WIN_PROJ_LASTFRAME 1   /* SearchByProjection(Frame&, const Frame&, th)             :1350-1476              */
#define CCM_WIN_PROJ_KF_SIM3 2     /* SearchByProjection(kfptr, Scw, vpPoints, vpMatched, th)  :308-446                */
#define CCM_WIN_BEST 3             /* SearchBySim3 passes :1216-1252,:1296-1327; Fuse distance phase :929-962,:1082-1103 */
#define CCM_WIN_INIT 4             /* SearchForInitialization                                  :448-563                */

typedef struct ccm_window_params {
    int mode;                  /* CCM_WIN_* */
    float nnratio;             /* mfNNratio */
    int th_dist;               /* TH_HIGH or TH_LOW as the reference function uses */
    int check_orientation;     /* mbCheckOrientation (LASTFRAME, INIT) */
    const float* inv_sigma2;   /* CCM_WIN_BEST only: mvInvLevelSigma2 for Fuse's chi2 gate (:950), NULL = no gate; HOST ptr, <= 16 */
    int n_levels;
} ccm_window_params;

/* holder [n_targets] in/out: -1 no map point, -2 map point with Observations()>0 (blocks), -3 map point with
 * Observations()==0; on return a claimed target holds the claiming query index. INIT mode: holder is output only
 * (vnMatches21). res [nq] per-query outcome (may be NULL). *nmatches = the reference function's return value
 * (for CCM_WIN_BEST: number of accepted queries). */
int ccm_search_window(ccm_matcher* m, ccm_frame_pack* targets, const ccm_window_params* prm, int nq,
                      const ccm_window_query* q, const uint8_t* q_desc, int* holder, ccm_match_result* res, int* nmatches);

/* batched, device-resident: job j searches packs[j] with queries d_q + j*q_stride (d_nq[j] of them, <= q_stride);
 * d_holder [n_jobs][holder_stride]; d_res [n_jobs][q_stride]; d_nmatches [n_jobs]. One CTA per job, one launch. */
int ccm_search_window_batch_device(ccm_matcher* m, ccm_frame_pack* const* packs, int n_jobs, const ccm_window_params* prm,
                                   const int* d_nq, int q_stride, const ccm_window_query* d_q, const uint8_t* d_q_desc,
                                   int* d_holder, int holder_stride, ccm_match_result* d_res, int* d_nmatches);

/* ---- BoW-guided searches -----------------------------------------------------------------------------------
 * A DBoW2::FeatureVector is given as node_id per feature (-1 = not in the vector): nodes iterate ascending,
 * features inside a node ascending (FeatureVector.cpp:31-45). valid = "has a good map point". */
#define CCM_BOW_KF_FRAME 0   /* SearchByBoW(kfptr, Frame&, out)   ORBmatcher.cpp:178-306: accept best <= TH_LOW; valid_b ignored */
#define CCM_BOW_KF_KF 1      /* SearchByBoW(kfptr, kfptr, out)    :565-698: accept best < TH_LOW */
/* match_a [n_a]: matched b index or -1 (vpMatches12); match_b [n_b]: matched a index or -1 (vpMapPointMatches) */
int ccm_search_bow(ccm_matcher* m, int mode, float nnratio, int check_orientation,
                   int n_a, const uint8_t* desc_a, const int* node_a, const uint8_t* valid_a, const float* angle_a,
                   int n_b, const uint8_t* desc_b, const int* node_b, const uint8_t* valid_b, const float* angle_b,
                   int* match_a, int* match_b, ccm_match_result* res_a, int* nmatches);

/* SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs), ORBmatcher.cpp:700-852 with CheckDistEpipolarLine :159-176.
 * Same FeatureVector convention; free_a / free_b = the feature holds NO map point (:741-744, :758-762). The epipole
 * (:707-714) and F12 are host geometry and come in as numbers. match_a [n_a]: pKF2 feature index or -1 (vMatches12;
 * vMatchedPairs = its non-negative entries in index order); *nmatches = the return value. A pKF2 feature can be matched
 * by several pKF1 features (upstream never sets vbMatched2). */
typedef struct ccm_epipolar_geom {
    float f12[9];                  /* F12, row-major */
    float ex, ey;                  /* epipole: pKF1's camera centre projected into pKF2 */
    const float* scale_factors;    /* pKF2->mvScaleFactors  (HOST, n_levels) */
    const float* level_sigma2;     /* pKF2->mvLevelSigma2   (HOST, n_levels) */
    int n_levels;                  /* <= 16 */
} ccm_epipolar_geom;
int ccm_search_triangulation(ccm_matcher* m, int check_orientation, const ccm_epipolar_geom* g,
                             int n_a, const uint8_t* desc_a, const int* node_a, const uint8_t* free_a, const float* x_a,
                             const float* y_a, const float* angle_a,
                             int n_b, const uint8_t* desc_b, const int* node_b, const uint8_t* free_b, const float* x_b,
                             const float* y_b, const int* octave_b, const float* angle_b, int* match_a, int* nmatches);

/* ---- vocabulary transform: the per-feature half of Frame::ComputeBoW (cslam/src/Frame.cpp:268-275) ----------
 * mpORBvocabulary->transform(descriptors, mBowVec, mFeatVec, 4) descends the DBoW2 tree once per descriptor
 * (thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1261: child at the smallest FORB::distance, first child on ties) and
 * files the result in two std::maps. The descent runs here; the maps (double weights, :1143-1168) stay on the host.
 * The tree is given as loadFromTextFile numbers it (:1338-1420): node 0 = root, child_start [n_nodes + 1] / child_ids
 * [n_nodes - 1] = each node's children in file order, node_desc [n_nodes][32], node_word [n_nodes] = word id of a
 * leaf (-1 otherwise), depth_levels = m_L. All HOST arrays, copied. */
typedef struct ccm_vocabulary ccm_vocabulary;
int ccm_vocabulary_create(ccm_matcher* m, int n_nodes, const int* child_start, const int* child_ids, const uint8_t* node_desc,
                          const int* node_word, int depth_levels, ccm_vocabulary** out);
/* May be called before or after ccm_matcher_destroy (it does not touch the matcher). */
void ccm_vocabulary_destroy(ccm_vocabulary* v);
/* word [n] = word id of feature i, node [n] = the node `levelsup` levels above the leaves on its path (0 = root when
 * depth_levels - levelsup <= 0, and when the path ends above that level, which upstream leaves undefined). */
int ccm_vocabulary_transform(ccm_matcher* m, const ccm_vocabulary* v, int n, const uint8_t* desc, int levelsup, int* word, int* node);
/* device-resident and asynchronous on the matcher's stream: d_desc [n][32] (16-byte aligned, e.g. the extractor's
 * descriptor rows of a whole batch), d_word / d_node [n]. d_node is what ccm_search_bow takes as node_a / node_b. */
int ccm_vocabulary_transform_device(ccm_matcher* m, const ccm_vocabulary* v, int n, const uint8_t* d_desc, int levelsup,
                                    int* d_word, int* d_node);

/* ---- MapPoint::ComputeDistinctiveDescriptors (cslam/src/MapPoint.cpp:929-994), batched ----------------------
 * Map point p owns descriptor rows [start[p], start[p+1]) of desc (the descriptors of its observations in the
 * caller's iteration order; at most 65535 per point). best[p] = index within its rows of the descriptor with the
 * least median distance to the others (the sorted row's element at int(0.5*(N-1)), self-distance included, first
 * minimum wins), -1 for a point without rows; best_median[p] = that median (may be NULL). HOST arrays. */
int ccm_distinctive_descriptors(ccm_matcher* m, int n_points, const int* start, const uint8_t* desc, int* best, int* best_median);

/* ---- keyframe wire format: the CvKeyPoint[] / Descriptor[] arrays of ccmslam_msgs/KF ----------------------------
 * ROS 1 serialisation of what KeyFrame::ConvertToMessage fills (KeyFrame.cpp:1462-1472, Converter.cc:166-179):
 *   kp_bytes   = u32 n | n x { f32 x, f32 y, u8 size, f32 angle, u8 response, i8 octave }   (CCM_WIRE_KP_BYTES each)
 *   desc_bytes = u32 n | n x u8[32]
 * size / response truncate to u8 and octave to i8 on the wire, as (u_int8_t)kp.size does. Buffers are little-endian
 * byte strings of CCM_WIRE_KP_LEN(n) / CCM_WIRE_DESC_LEN(n) bytes; allocate them rounded up to a multiple of 4. */
#define CCM_WIRE_KP_BYTES 15
#define CCM_WIRE_KP_LEN(n) (4 + (size_t)CCM_WIRE_KP_BYTES * (size_t)(n))
#define CCM_WIRE_DESC_LEN(n) (4 + (size_t)32 * (size_t)(n))
/* client side, device-resident and asynchronous: n keypoints / descriptor rows of the extractor's output (d_xy_un
 * [n][2] overrides the positions with mvKeysUn, or NULL) -> d_kp_bytes, d_desc_bytes (device, 4-byte aligned) */
int ccm_wire_encode_keyframe_device(ccm_matcher* m, const ccm_keypoint* d_kp, const float* d_xy_un, const uint8_t* d_desc,
                                    int n, uint8_t* d_kp_bytes, uint8_t* d_desc_bytes);
/* host convenience: upload, encode, download; synchronous */
int ccm_wire_encode_keyframe(ccm_matcher* m, const ccm_keypoint* kp, const float* xy_un, const uint8_t* desc, int n,
                             uint8_t* kp_bytes, uint8_t* desc_bytes);
/* server side, KeyFrame::WriteMembersFromMessage (KeyFrame.cpp:1689-1702, Converter.cc:181-193): the two received byte
 * strings (HOST) go up as they are and are decoded on the device straight into a frame pack (positions, octaves,
 * angles, descriptors, feature grid). kp_out [n] (HOST, may be NULL) receives the decoded mvKeysUn. Returns
 * CCM_E_BADARG when the counts and lengths disagree, CCM_E_CAPACITY when the pack is too small. */
int ccm_frame_pack_from_wire(ccm_frame_pack* p, const ccm_grid_params* g, const uint8_t* kp_bytes, size_t kp_len,
                             const uint8_t* desc_bytes, size_t desc_len, ccm_keypoint* kp_out);

/* ---- server keyframe-database search (BASELINE.json configs[4]) --------------------------------------------
 * Best and second-best Hamming distance of every query descriptor over a (shard of a) target descriptor database,
 * first-wins on ties: the inner loop of SearchByBoW(KF,KF) (:607-637) with one all-inclusive node. */
typedef struct ccm_best2 {
    uint32_t idx;          /* global target index of the best match (index_base + local), 0xFFFFFFFF if the shard is empty */
    uint16_t d1, d2;       /* best and second-best distance, 256 = none */
} ccm_best2;

/* device-resident, asynchronous: d_out [nq] = best-2 of d_q [nq][32] over d_targets [nt][32] */
int ccm_match_allpairs_best2_device(ccm_matcher* m, const uint8_t* d_q, int nq, const uint8_t* d_targets, long long nt,
                                    uint32_t index_base, ccm_best2* d_out);
/* cross-shard merge after the allgather of per-shard records: d_parts [n_parts][nq] -> d_out [nq], and the
 * reference accept rule (best < TH_LOW and (float)best < nnratio*(float)best2, :641-643) -> d_accept [nq] (may be NULL) */
int ccm_match_allpairs_merge_device(ccm_matcher* m, const ccm_best2* d_parts, int n_parts, int nq, float nnratio,
                                    ccm_best2* d_out, uint8_t* d_accept);
/* ---- fused fold + exchange + merge over peer memory (NVLink), the multi-GPU form of the two calls above ----------
 * One context per rank owns an exchange buffer in device memory that every other rank of the box maps (CUDA IPC across
 * processes, or plain peer access inside one process). ccm_match_allpairs_best2_exchange_device then does the whole sharded
 * search in two launches: the per-chunk search and ONE kernel that folds the chunks, stores the rank's best-2 records into
 * every rank's buffer, flags them, waits for the other ranks' records of the same queries and merges (accept rule
 * included). No NCCL call, no host synchronisation; every rank must make the same sequence of calls. */
typedef struct ccm_exchange ccm_exchange;
#define CCM_IPC_HANDLE_BYTES 64
#define CCM_EXCHANGE_MAX_WORLD 16
/* handle_out (may be NULL) receives the cudaIpcMemHandle_t of this rank's buffer, to be all-gathered by the caller */
int ccm_exchange_create(ccm_matcher* m, int rank, int world, int nq_cap, ccm_exchange** out, uint8_t* handle_out);
/* handles [world][CCM_IPC_HANDLE_BYTES] in rank order (entry `rank` is ignored): opens the peers' buffers */
int ccm_exchange_connect_ipc(ccm_exchange* x, const uint8_t* handles);
/* same-process variant: device pointers of all ranks' buffers (ccm_exchange_buffer) in rank order; enables peer access */
int ccm_exchange_connect_ptrs(ccm_exchange* x, void* const* peer_buffers);
void* ccm_exchange_buffer(ccm_exchange* x);
/* synchronises the device, closes the peer mappings and frees the buffer. Call it when no peer can still be inside a
 * search on this context (a barrier of the caller's), and before or after ccm_matcher_destroy - it does not touch m. */
void ccm_exchange_destroy(ccm_exchange* x);
int ccm_match_allpairs_best2_exchange_device(ccm_matcher* m, ccm_exchange* x, const uint8_t* d_q, int nq, const uint8_t* d_targets,
                                             long long nt, uint32_t index_base, float nnratio, ccm_best2* d_out, uint8_t* d_accept);
/* All ranks must make the same sequence of exchange calls with the same nq (lockstep). A rank whose peer never launches
 * (its call failed, its process died) does not spin forever: after CCM_EXCHANGE_TIMEOUT_MS (environment, default 2000)
 * the kernel gives up, writes "no match" records and raises a status word. ccm_exchange_check synchronises the matcher's
 * stream and returns CCM_E_CUDA if that happened since the context was created; the context is then out of step with
 * its peers and must be destroyed and recreated on every rank. */
int ccm_exchange_check(ccm_exchange* x);
/* host convenience: upload, search one shard, download */
int ccm_match_allpairs_best2(ccm_matcher* m, const uint8_t* q, int nq, const uint8_t* targets, long long nt,
                             uint32_t index_base, ccm_best2* out);

/* ---- allocation-free database search -----------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same result as
 * ccm_match_allpairs_best2_device (d_out [nq] = best-2 of d_q [nq][32] over d_targets [nt][32], device pointers,
 * 16-byte aligned), but it takes no matcher handle and allocates nothing: the per-chunk partial records live in a
 * fixed 4 MiB `static __device__` array inside the library (every GPU that loads the library holds its copy until
 * the process exits). Runs on the CURRENT device, on `stream` (a cudaStream_t; NULL = the default stream),
 * asynchronously. Limits: nq <= CCM_NOALLOC_MAX_QUERIES and nt <= floor(524288 / nq) * 2^23, else CCM_E_CAPACITY;
 * the message is read with ccm_matcher_last_error(NULL). All calls on one GPU share the array: calls that overlap
 * on the GPU (different streams or host threads) race; calls ordered on one stream are safe. The first call on a
 * device resolves the array's device address (cudaGetSymbolAddress) - make it outside stream capture. */
#define CCM_NOALLOC_MAX_QUERIES 4096
int ccm_match_allpairs_best2_noalloc_device(void* stream, const uint8_t* d_q, int nq, const uint8_t* d_targets, long long nt,
                                            uint32_t index_base, ccm_best2* d_out);

#ifdef __cplusplus
}
#endif
#endif /* CCM_MATCH_H_ */
