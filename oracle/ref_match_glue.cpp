/* TEST INFRASTRUCTURE — C wrapper around the reference's OWN cslam::ORBmatcher, compiled UNMODIFIED from
 * /root/reference/cslam/src/ORBmatcher.cpp against the mock Frame / KeyFrame / MapPoint headers in oracle/matchshim
 * (oracle/Makefile target _ref/libref_match.so). Each rm_* function takes the same plain arrays as the corresponding
 * mo_* function of oracle/match_oracle.cpp, builds the mock objects so that the reference's own host-side geometry
 * reproduces exactly those inputs (identity poses, unit focal length, world point = (u, v, 1)), calls the reference
 * method and translates the object state it leaves behind back into arrays. tests/test_match_oracle_vs_ref.py compares
 * the two. Nothing here is shipped or called by the product path. */
#include <cslam/ORBmatcher.h>

#include <cmath>
#include <cstring>
#include <limits>
#include <set>

extern "C" void mo_frame_params(const void* fp, float* min_x, float* min_y, float* inv_w, float* inv_h, int* cols, int* rows);
extern "C" void mo_frame_view(const void* fp, int* n, const uint8_t** desc, const float** x, const float** y, const int** octave, const float** angle);

namespace cslam {
float Frame::fx = 1.f, Frame::fy = 1.f, Frame::cx = 0.f, Frame::cy = 0.f;
float Frame::mnMinX = -1e30f, Frame::mnMaxX = 1e30f, Frame::mnMinY = -1e30f, Frame::mnMaxY = 1e30f;
float Frame::mfGridElementWidthInv = 0.f, Frame::mfGridElementHeightInv = 0.f;
}

namespace {
using namespace cslam;
typedef boost::shared_ptr<MapPoint> mpptr;
typedef boost::shared_ptr<KeyFrame> kfptr;

cv::Mat desc_mat(const uint8_t* d, int n) {
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) std::memcpy(m.data, d, (size_t)n * 32);
    m.rows = n;
    return m;
}
cv::Mat vec3(float a, float b, float c) { cv::Mat m(3, 1, CV_32F); m.at<float>(0) = a; m.at<float>(1) = b; m.at<float>(2) = c; return m; }

void fill_common(const void* fp, int& N, std::vector<cv::KeyPoint>& keysUn, cv::Mat& desc) {
    const uint8_t* d; const float *x, *y, *ang; const int* oct;
    mo_frame_view(fp, &N, &d, &x, &y, &oct, &ang);
    keysUn.resize(N);
    for (int i = 0; i < N; i++) keysUn[i] = cv::KeyPoint(x[i], y[i], 31.f, ang[i], 0.f, oct[i]);
    desc = desc_mat(d, N);
}

/* Identity of the objects built from one frame handle: the same handle searched again is the same Frame / KeyFrame
 * (the drop-in keys its device packs on mId). rm_set_unique_ids(1) gives every object its own id instead. */
static bool g_unique_ids = false;
static size_t g_next_id = 1;
std::pair<size_t, size_t> object_id(const void* fp, size_t kind) {
    return g_unique_ids ? std::make_pair(g_next_id++, (size_t)99) : std::make_pair((size_t)(uintptr_t)fp, kind);
}

void make_frame(Frame& F, const void* fp, const float* sf, int nlevels) {
    fill_common(fp, F.N, F.mvKeysUn, F.mDescriptors);
    F.mId = object_id(fp, 1);
    F.mvKeys = F.mvKeysUn;
    F.grid = const_cast<void*>(fp);
    {   /* static image bounds / grid pitch as Frame.cpp:80-88 would set them for this grid */
        float mx, my, iw, ih; int gc, gr;
        mo_frame_params(fp, &mx, &my, &iw, &ih, &gc, &gr);
        Frame::mnMinX = mx; Frame::mnMinY = my; Frame::mnMaxX = mx + (float)gc / iw; Frame::mnMaxY = my + (float)gr / ih;
        Frame::mfGridElementWidthInv = iw; Frame::mfGridElementHeightInv = ih;
    }
    F.mvpMapPoints.assign(F.N, mpptr());
    F.mvbOutlier.assign(F.N, false);
    F.mTcw = cv::Mat::eye(4, 4, CV_32F);
    if (sf) F.mvScaleFactors.assign(sf, sf + nlevels);
}

kfptr make_kf(const void* fp, const float* sf, int nlevels) {
    kfptr K(new KeyFrame());
    fill_common(fp, K->N, K->mvKeysUn, K->mDescriptors);
    K->mId = object_id(fp, 2);
    K->grid = const_cast<void*>(fp);
    K->mvpMapPoints.assign(K->N, mpptr());
    {
        float mx, my, iw, ih; int gc, gr;
        mo_frame_params(fp, &mx, &my, &iw, &ih, &gc, &gr);
        K->mnMinX = (int)mx; K->mnMinY = (int)my; K->mnMaxX = (int)(mx + (float)gc / iw); K->mnMaxY = (int)(my + (float)gr / ih);
        K->mnGridCols = gc; K->mnGridRows = gr; K->mfGridElementWidthInv = iw; K->mfGridElementHeightInv = ih;
    }
    K->minX = K->minY = -1e30f; K->maxX = K->maxY = 1e30f;
    K->R = cv::Mat::eye(3, 3, CV_32F); K->t = vec3(0, 0, 0); K->C = vec3(0, 0, 0);
    if (sf) K->mvScaleFactors.assign(sf, sf + nlevels);
    return K;
}

mpptr make_mp(int id, const uint8_t* desc32, int nObs) {
    mpptr p(new MapPoint());
    p->id = id; p->nObs = nObs;
    if (desc32) p->descriptor = desc_mat(desc32, 1);
    return p;
}

/* holder codes of oracle/match_oracle.cpp -> occupants of F.mvpMapPoints / vpMatched, and back */
void holder_to_points(const int* holder, int n, std::vector<mpptr>& v) {
    v.assign(n, mpptr());
    for (int i = 0; i < n; i++) {
        if (holder[i] == -2) v[i] = make_mp(-2, nullptr, 1);
        else if (holder[i] == -3) v[i] = make_mp(-3, nullptr, 0);
    }
}
void points_to_holder(const std::vector<mpptr>& v, int* holder) {
    for (size_t i = 0; i < v.size(); i++) holder[i] = v[i] ? v[i]->id : -1;
}

void set_world(mpptr p, float u, float v) {
    p->worldPos = vec3(u, v, 1.f);
    const float d = std::sqrt(u * u + v * v + 1.f);
    p->normal = vec3(u / d, v / d, 1.f / d);          /* viewing-angle gates pass: PO.Pn ~ dist >= 0.5*dist */
}

DBoW2::FeatureVector feat_vec(int n, const int* node) {
    DBoW2::FeatureVector fv;
    for (int i = 0; i < n; i++) if (node[i] >= 0) fv.addFeature((DBoW2::NodeId)node[i], (unsigned)i);
    return fv;
}
const int NLEV = 8;
}  // namespace

extern "C" {

void rm_set_unique_ids(int on) { g_unique_ids = on != 0; }

int rm_descriptor_distance(const uint8_t* a, const uint8_t* b) { return ORBmatcher::DescriptorDistance(desc_mat(a, 1), desc_mat(b, 1)); }

/* ORBmatcher::SearchByProjection(Frame&, const vector<mpptr>&, th), ORBmatcher.cpp:71-148 */
int rm_search_projection_mappoints(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                                   const int* q_level, const float* q_viewcos, const uint8_t* q_desc, const uint8_t* q_obs,
                                   float th, float nnratio, const float* scale_factors, int* holder) {
    Frame F;
    make_frame(F, fp, scale_factors, NLEV);
    holder_to_points(holder, F.N, F.mvpMapPoints);
    std::vector<mpptr> mps(nq);
    for (int i = 0; i < nq; i++) {
        mps[i] = make_mp(i, q_desc + (size_t)32 * i, q_obs[i] ? 1 : 0);
        mps[i]->mbTrackInView = q_valid[i] != 0;
        mps[i]->mTrackProjX = q_u[i]; mps[i]->mTrackProjY = q_v[i];
        mps[i]->mnTrackScaleLevel = q_level[i]; mps[i]->mTrackViewCos = q_viewcos[i];
    }
    ORBmatcher m(nnratio, true);
    const int nm = m.SearchByProjection(F, mps, th);
    points_to_holder(F.mvpMapPoints, holder);
    return nm;
}

/* ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th), ORBmatcher.cpp:1350-1476 */
int rm_search_projection_lastframe(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                                   const int* q_octave, const float* q_angle, const uint8_t* q_desc, const uint8_t* q_obs,
                                   float th, const float* scale_factors, int check_orientation, int* holder) {
    boost::shared_ptr<Frame> Cur(new Frame());
    make_frame(*Cur, fp, scale_factors, NLEV);
    holder_to_points(holder, Cur->N, Cur->mvpMapPoints);
    Frame Last;
    Last.N = nq;
    Last.mTcw = cv::Mat::eye(4, 4, CV_32F);
    Last.mvKeys.resize(nq); Last.mvKeysUn.resize(nq);
    Last.mvpMapPoints.assign(nq, mpptr());
    Last.mvbOutlier.assign(nq, false);
    for (int i = 0; i < nq; i++) {
        Last.mvKeys[i] = cv::KeyPoint(q_u[i], q_v[i], 31.f, q_angle[i], 0.f, q_octave[i]);
        Last.mvKeysUn[i] = Last.mvKeys[i];
        if (!q_valid[i]) continue;
        Last.mvpMapPoints[i] = make_mp(i, q_desc + (size_t)32 * i, q_obs[i] ? 1 : 0);
        set_world(Last.mvpMapPoints[i], q_u[i], q_v[i]);
    }
    ORBmatcher m(0.9f, check_orientation != 0);
    const int nm = m.SearchByProjection(*Cur, Last, th);
    points_to_holder(Cur->mvpMapPoints, holder);
    return nm;
}

/* ORBmatcher::SearchByProjection(Frame& Current, kfptr pKF, sAlreadyFound, th, ORBdist), ORBmatcher.cpp:1478-1605
 * (relocalisation; no caller upstream). q_state: 0 = pKF has no map point at this feature, 1 = a good point, 2 = a bad
 * point, 3 = a good point that is in sAlreadyFound. */
int rm_search_projection_reloc(const void* fp, int nq, const uint8_t* q_state, const float* q_u, const float* q_v, const int* q_level,
                               const float* q_angle, const uint8_t* q_desc, float th, int orb_dist, const float* scale_factors,
                               int check_orientation, int* holder) {
    boost::shared_ptr<Frame> Cur(new Frame());
    make_frame(*Cur, fp, scale_factors, NLEV);
    holder_to_points(holder, Cur->N, Cur->mvpMapPoints);
    kfptr K(new KeyFrame());
    K->N = nq;
    K->mvKeysUn.resize(nq);
    K->mvpMapPoints.assign(nq, mpptr());
    std::set<mpptr> found;
    for (int i = 0; i < nq; i++) {
        K->mvKeysUn[i] = cv::KeyPoint(q_u[i], q_v[i], 31.f, q_angle[i], 0.f, q_level[i]);
        if (!q_state[i]) continue;
        mpptr p = make_mp(i, q_desc + (size_t)32 * i, 1);
        p->bad = q_state[i] == 2;
        p->predictedLevel = q_level[i];
        set_world(p, q_u[i], q_v[i]);
        if (q_state[i] == 3) found.insert(p);
        K->mvpMapPoints[i] = p;
    }
    ORBmatcher m(0.9f, check_orientation != 0);
    const int nm = m.SearchByProjection(*Cur, K, found, th, orb_dist);
    points_to_holder(Cur->mvpMapPoints, holder);
    return nm;
}

/* ORBmatcher::SearchByProjection(kfptr, Scw, vpPoints, vpMatched, th), ORBmatcher.cpp:308-446 */
int rm_search_projection_kf_sim3(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                                 const int* q_level, const uint8_t* q_desc, const uint8_t* q_claims, int th,
                                 const float* scale_factors, int* holder) {
    kfptr K = make_kf(fp, scale_factors, NLEV);
    std::vector<mpptr> matched;
    holder_to_points(holder, K->N, matched);
    std::vector<mpptr> pts(nq);
    for (int i = 0; i < nq; i++) {
        pts[i] = make_mp(i, q_desc + (size_t)32 * i, 1);
        pts[i]->bad = !q_valid[i];
        pts[i]->predictedLevel = q_level[i];
        set_world(pts[i], q_u[i], q_v[i]);
        if (!q_claims[i]) pts[i]->indexInKF[K.get()] = 0;      /* already observed by pKF: the re-map branch, :414-433 */
    }
    ORBmatcher m(0.75f, true);
    const int nm = m.SearchByProjection(K, cv::Mat::eye(4, 4, CV_32F), pts, matched, th);
    points_to_holder(matched, holder);
    return nm;
}

/* ORBmatcher::SearchByBoW(kfptr, Frame&, vpMapPointMatches), ORBmatcher.cpp:178-306 */
int rm_search_bow_kf_frame(int nKF, const uint8_t* desc_kf, const int* node_kf, const uint8_t* valid_kf, const float* angle_kf,
                           int nF, const uint8_t* desc_f, const int* node_f, const float* angle_f,
                           float nnratio, int check_orientation, int* match_of_f) {
    kfptr K(new KeyFrame());
    K->N = nKF; K->mDescriptors = desc_mat(desc_kf, nKF); K->mFeatVec = feat_vec(nKF, node_kf);
    K->mvKeysUn.resize(nKF); K->mvpMapPoints.assign(nKF, mpptr());
    for (int i = 0; i < nKF; i++) {
        K->mvKeysUn[i] = cv::KeyPoint(0, 0, 31.f, angle_kf[i]);
        if (valid_kf[i]) K->mvpMapPoints[i] = make_mp(i, nullptr, 1);
    }
    Frame F;
    F.N = nF; F.mDescriptors = desc_mat(desc_f, nF); F.mFeatVec = feat_vec(nF, node_f);
    F.mvKeys.resize(nF);
    for (int i = 0; i < nF; i++) F.mvKeys[i] = cv::KeyPoint(0, 0, 31.f, angle_f[i]);
    F.mvKeysUn = F.mvKeys;
    std::vector<mpptr> out;
    ORBmatcher m(nnratio, check_orientation != 0);
    const int nm = m.SearchByBoW(K, F, out);
    for (int i = 0; i < nF; i++) match_of_f[i] = (i < (int)out.size() && out[i]) ? out[i]->id : -1;
    return nm;
}

/* ORBmatcher::SearchByBoW(kfptr, kfptr, vpMatches12), ORBmatcher.cpp:565-698 */
int rm_search_bow_kf_kf(int n1, const uint8_t* desc1, const int* node1, const uint8_t* valid1, const float* angle1,
                        int n2, const uint8_t* desc2, const int* node2, const uint8_t* valid2, const float* angle2,
                        float nnratio, int check_orientation, int* match12) {
    kfptr K[2] = {kfptr(new KeyFrame()), kfptr(new KeyFrame())};
    const int n[2] = {n1, n2}; const uint8_t* d[2] = {desc1, desc2}; const int* nd[2] = {node1, node2};
    const uint8_t* v[2] = {valid1, valid2}; const float* a[2] = {angle1, angle2};
    for (int k = 0; k < 2; k++) {
        K[k]->N = n[k]; K[k]->mDescriptors = desc_mat(d[k], n[k]); K[k]->mFeatVec = feat_vec(n[k], nd[k]);
        K[k]->mvKeysUn.resize(n[k]); K[k]->mvpMapPoints.assign(n[k], mpptr());
        for (int i = 0; i < n[k]; i++) {
            K[k]->mvKeysUn[i] = cv::KeyPoint(0, 0, 31.f, a[k][i]);
            if (v[k][i]) K[k]->mvpMapPoints[i] = make_mp(i, nullptr, 1);
        }
    }
    std::vector<mpptr> out;
    ORBmatcher m(nnratio, check_orientation != 0);
    const int nm = m.SearchByBoW(K[0], K[1], out);
    for (int i = 0; i < n1; i++) match12[i] = (i < (int)out.size() && out[i]) ? out[i]->id : -1;
    return nm;
}

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), ORBmatcher.cpp:448-563 */
int rm_search_for_initialization(int n1, const uint8_t* desc1, const int* octave1, const float* angle1, float* prev_xy,
                                 const void* f2, int windowSize, float nnratio, int check_orientation, int* matches12) {
    Frame F1, F2;
    F1.N = n1; F1.mDescriptors = desc_mat(desc1, n1);
    F1.mvKeysUn.resize(n1);
    for (int i = 0; i < n1; i++) F1.mvKeysUn[i] = cv::KeyPoint(prev_xy[2 * i], prev_xy[2 * i + 1], 31.f, angle1[i], 0.f, octave1[i]);
    F1.mvKeys = F1.mvKeysUn;
    make_frame(F2, f2, nullptr, 0);
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; i++) prev[i] = cv::Point2f(prev_xy[2 * i], prev_xy[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher m(nnratio, check_orientation != 0);
    const int nm = m.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < n1; i++) {
        matches12[i] = i < (int)m12.size() ? m12[i] : -1;
        prev_xy[2 * i] = prev[i].x; prev_xy[2 * i + 1] = prev[i].y;
    }
    return nm;
}

/* Occupants of a keyframe's feature slots for the Fuse cases: occ_obs[k] < 0 = empty, else a map point with id
 * OCC_ID + k, that many observations, observed by K at slot k; occ_flags bit0 = bad, bit1 = mbDoNotReplace. */
static const int OCC_ID = 100000;
static void make_occupants(kfptr K, const int* occ_obs, const uint8_t* occ_flags) {
    for (int k = 0; k < K->N; k++) {
        if (occ_obs[k] < 0) continue;
        mpptr o = make_mp(OCC_ID + k, nullptr, occ_obs[k]);
        o->bad = occ_flags[k] & 1; o->mbDoNotReplace = (occ_flags[k] & 2) != 0;
        o->indexInKF[K.get()] = k;
        K->mvpMapPoints[k] = o;
    }
}
/* query points: q_alias[i] = j < i shares point j (the same map point listed twice), q_alias[i] = -(k+2) is the
 * occupant of slot k (a candidate the keyframe already holds), -1 = a fresh point. q_flags bit1 = mbDoNotReplace. */
static void make_points(kfptr K, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v, const int* q_level,
                        const uint8_t* q_desc, const int* q_obs, const uint8_t* q_flags, const int* q_alias, std::vector<mpptr>& pts) {
    pts.resize(nq);
    for (int i = 0; i < nq; i++) {
        if (q_alias[i] >= 0 && q_alias[i] < i) { pts[i] = pts[q_alias[i]]; continue; }
        if (q_alias[i] <= -2 && K->mvpMapPoints[-q_alias[i] - 2]) pts[i] = K->mvpMapPoints[-q_alias[i] - 2];
        else pts[i] = make_mp(i, nullptr, q_obs[i]);
        pts[i]->descriptor = desc_mat(q_desc + (size_t)32 * i, 1);
        if (!q_valid[i]) pts[i]->bad = true;
        pts[i]->mbDoNotReplace = pts[i]->mbDoNotReplace || (q_flags[i] & 2) != 0;
        pts[i]->predictedLevel = q_level[i];
        set_world(pts[i], q_u[i], q_v[i]);
    }
}
static void export_slots(kfptr K, int* slot_id, int* slot_bad) {
    for (int k = 0; k < K->N; k++) {
        slot_id[k] = K->mvpMapPoints[k] ? K->mvpMapPoints[k]->id : -1;
        slot_bad[k] = K->mvpMapPoints[k] ? (int)K->mvpMapPoints[k]->bad : 0;
    }
}

/* ORBmatcher::Fuse(kfptr, vpMapPoints, th), ORBmatcher.cpp:854-993. Outputs per query: replaced_by (id of the point it was
 * replaced by, -1), added_idx (slot it was added to, -1), bad; per slot: id of the final occupant and whether it is bad. */
int rm_fuse(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v, const int* q_level,
            const uint8_t* q_desc, const int* q_obs, const uint8_t* q_flags, const int* q_alias, float th,
            const float* scale_factors, const float* inv_sigma2, const int* occ_obs, const uint8_t* occ_flags,
            int* replaced_by, int* added_idx, int* q_bad, int* slot_id, int* slot_bad, int* occ_replaced_by) {
    kfptr K = make_kf(fp, scale_factors, NLEV);
    K->mvInvLevelSigma2.assign(inv_sigma2, inv_sigma2 + NLEV);
    make_occupants(K, occ_obs, occ_flags);
    std::vector<mpptr> occ = K->mvpMapPoints, pts;
    make_points(K, nq, q_valid, q_u, q_v, q_level, q_desc, q_obs, q_flags, q_alias, pts);
    ORBmatcher m(0.6f, true);
    const int nFused = m.Fuse(K, pts, th);
    for (int i = 0; i < nq; i++) { replaced_by[i] = pts[i]->replacedBy; added_idx[i] = pts[i]->addedObsIdx; q_bad[i] = pts[i]->bad; }
    export_slots(K, slot_id, slot_bad);
    for (int k = 0; k < K->N; k++) occ_replaced_by[k] = occ[k] ? occ[k]->replacedBy : -1;
    return nFused;
}

/* ORBmatcher::Fuse(kfptr, Scw, vpPoints, th, vpReplacePoint), ORBmatcher.cpp:995-1122 (Scw = identity) */
int rm_fuse_sim3(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v, const int* q_level,
                 const uint8_t* q_desc, const int* q_obs, const uint8_t* q_flags, const int* q_alias, float th,
                 const float* scale_factors, const int* occ_obs, const uint8_t* occ_flags,
                 int* replace_point, int* added_idx, int* slot_id, int* slot_bad) {
    kfptr K = make_kf(fp, scale_factors, NLEV);
    make_occupants(K, occ_obs, occ_flags);
    std::vector<mpptr> pts, rep(nq);
    make_points(K, nq, q_valid, q_u, q_v, q_level, q_desc, q_obs, q_flags, q_alias, pts);
    ORBmatcher m(0.8f, true);
    const int nFused = m.Fuse(K, cv::Mat::eye(4, 4, CV_32F), pts, th, rep);
    for (int i = 0; i < nq; i++) { replace_point[i] = rep[i] ? rep[i]->id : -1; added_idx[i] = pts[i]->addedObsIdx; }
    export_slots(K, slot_id, slot_bad);
    return nFused;
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12 = 1, R12 = I, t12 = 0, th), ORBmatcher.cpp:1124-1348.
 * mpX_state[i]: 0 = no map point at slot i, 1 = map point, 2 = bad map point; (u, v) = where it projects in the OTHER
 * keyframe. prior12[i] >= 0: vpMatches12[i] = map point of KF2's slot prior12[i] on entry; -2 = a point KF2 does not
 * hold. matches12 out: slot of KF2 whose map point vpMatches12[i] is on return, -1 none, -2 foreign. */
int rm_search_by_sim3(const void* fp1, const void* fp2, const uint8_t* mp1_state, const float* mp1_u, const float* mp1_v,
                      const int* mp1_level, const uint8_t* mp1_desc, const uint8_t* mp2_state, const float* mp2_u,
                      const float* mp2_v, const int* mp2_level, const uint8_t* mp2_desc, const int* prior12, float th,
                      const float* scale_factors, int* matches12) {
    kfptr K1 = make_kf(fp1, scale_factors, NLEV), K2 = make_kf(fp2, scale_factors, NLEV);
    const int ID2 = 1000000;
    struct Side { kfptr K; const uint8_t* st; const float *u, *v; const int* lvl; const uint8_t* d; int base; };
    Side side[2] = {{K1, mp1_state, mp1_u, mp1_v, mp1_level, mp1_desc, 0}, {K2, mp2_state, mp2_u, mp2_v, mp2_level, mp2_desc, ID2}};
    for (Side& S : side)
        for (int i = 0; i < S.K->N; i++) {
            if (!S.st[i]) continue;
            mpptr p = make_mp(S.base + i, S.d + (size_t)32 * i, 1);
            p->bad = S.st[i] == 2; p->predictedLevel = S.lvl[i];
            set_world(p, S.u[i], S.v[i]);
            p->indexInKF[S.K.get()] = i;
            S.K->mvpMapPoints[i] = p;
        }
    std::vector<mpptr> m12(K1->N);
    mpptr foreign = make_mp(-2, nullptr, 1);
    for (int i = 0; i < K1->N; i++) {
        if (prior12[i] >= 0 && K2->mvpMapPoints[prior12[i]]) m12[i] = K2->mvpMapPoints[prior12[i]];
        else if (prior12[i] == -2) m12[i] = foreign;
    }
    ORBmatcher m(0.75f, true);
    const int nFound = m.SearchBySim3(K1, K2, m12, 1.0f, cv::Mat::eye(3, 3, CV_32F), vec3(0, 0, 0), th);
    for (int i = 0; i < K1->N; i++) matches12[i] = !m12[i] ? -1 : (m12[i]->id >= ID2 ? m12[i]->id - ID2 : -2);
    return nFound;
}

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs), ORBmatcher.cpp:700-852. Plain keyframes from
 * arrays (no grid needed); KF2 has identity pose and unit intrinsics and KF1's camera centre is (ex, ey, 1), so the
 * reference's own epipole code (:707-714) yields exactly (ex, ey). */
int rm_search_for_triangulation(int n1, const uint8_t* desc1, const int* node1, const uint8_t* free1, const float* x1,
                                const float* y1, const float* angle1, int n2, const uint8_t* desc2, const int* node2,
                                const uint8_t* free2, const float* x2, const float* y2, const int* octave2,
                                const float* angle2, const float* f12, float ex, float ey, const float* scale_factors,
                                const float* level_sigma2, int check_orientation, int* match12) {
    kfptr K1(new KeyFrame()), K2(new KeyFrame());
    struct Side { kfptr K; int n; const uint8_t* d; const int* node; const uint8_t* fr; const float *x, *y, *ang; const int* oct; };
    Side side[2] = {{K1, n1, desc1, node1, free1, x1, y1, angle1, nullptr}, {K2, n2, desc2, node2, free2, x2, y2, angle2, octave2}};
    for (Side& S : side) {
        S.K->N = S.n;
        S.K->mDescriptors = desc_mat(S.d, S.n);
        S.K->mvKeysUn.resize(S.n);
        S.K->mvpMapPoints.assign(S.n, mpptr());
        for (int i = 0; i < S.n; i++) {
            S.K->mvKeysUn[i] = cv::KeyPoint(S.x[i], S.y[i], 31.f, S.ang[i], 0.f, S.oct ? S.oct[i] : 0);
            if (!S.fr[i]) S.K->mvpMapPoints[i] = make_mp(i, nullptr, 1);
        }
        S.K->mFeatVec = feat_vec(S.n, S.node);
        S.K->R = cv::Mat::eye(3, 3, CV_32F); S.K->t = vec3(0, 0, 0); S.K->C = vec3(0, 0, 0);
        S.K->mvScaleFactors.assign(scale_factors, scale_factors + NLEV);
        S.K->mvLevelSigma2.assign(level_sigma2, level_sigma2 + NLEV);
    }
    K1->C = vec3(ex, ey, 1.f);
    cv::Mat F12(3, 3, CV_32F);
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F12.at<float>(r, c) = f12[3 * r + c];
    std::vector<std::pair<size_t, size_t> > pairs;
    ORBmatcher m(0.6f, check_orientation != 0);
    const int nm = m.SearchForTriangulation(K1, K2, F12, pairs);
    for (int i = 0; i < n1; i++) match12[i] = -1;
    for (size_t k = 0; k < pairs.size(); k++) match12[pairs[k].first] = (int)pairs[k].second;
    if ((int)pairs.size() != nm) return -1000 - nm;      /* the pair list and the return value must agree */
    return nm;
}

}  // extern "C"
