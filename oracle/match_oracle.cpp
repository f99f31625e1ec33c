/* TEST INFRASTRUCTURE — CPU oracle for the ORBmatcher Hamming searches. Never linked into the product library;
 * only tests/, __graft_entry__.smoke() and bench.py's CPU legs load it.
 *
 * Restates, over plain arrays, the loops of cslam/src/ORBmatcher.cpp and the grid lookups of cslam/src/Frame.cpp /
 * cslam/src/KeyFrame.cpp that feed them. The originals are welded to shared_ptr<MapPoint/KeyFrame/Frame>; here a
 * "map point" is a query index, "has a map point" is a flag array, and all projection geometry (u, v, radius,
 * predicted level) is an input, exactly as the C ABI receives it. Every function cites the lines it follows.
 * Pinned against oracle/_ref/libref_match.so (the reference's own ORBmatcher.cpp compiled unmodified over mock
 * Frame/KeyFrame/MapPoint headers) by tests/test_match_oracle_vs_ref.py where that library is built.
 * Build with -ffp-contract=off (float expressions must round as the reference's -O3 build without FMA does). */
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   /* ORBmatcher.cpp:63-65 */

/* ORBmatcher::DescriptorDistance, ORBmatcher.cpp:1653-1669 (SWAR popcount over 8 x 32 bits) */
inline int descriptor_distance(const uint8_t* a, const uint8_t* b) {
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4);
        std::memcpy(&y, b + 4 * i, 4);
        unsigned int v = x ^ y;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

/* ORBmatcher::ComputeThreeMaxima, ORBmatcher.cpp:1607-1648 */
void compute_three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

/* rotation-histogram bin, e.g. ORBmatcher.cpp:256-262 */
inline int rot_bin(float angle_a, float angle_b) {
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = angle_a - angle_b;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

/* A Frame's or KeyFrame's matching data: descriptors, undistorted keypoints, the feature grid. */
struct MoFrame {
    int n;
    std::vector<uint8_t> desc;
    std::vector<float> x, y, angle;
    std::vector<int> octave;
    float min_x, min_y, inv_w, inv_h;
    int cols, rows;
    std::vector<std::vector<int>> grid;   /* [ix * rows + iy] */

    /* Frame::PosInGrid + AssignFeaturesToGrid, Frame.cpp:103-118,255-265 (note: round(), not floor()) */
    void assign() {
        grid.assign((size_t)cols * rows, std::vector<int>());
        for (int i = 0; i < n; i++) {
            const int px = (int)std::round((x[i] - min_x) * inv_w);
            const int py = (int)std::round((y[i] - min_y) * inv_h);
            if (px < 0 || px >= cols || py < 0 || py >= rows) continue;
            grid[(size_t)px * rows + py].push_back(i);
        }
    }

    /* Frame::GetFeaturesInArea, Frame.cpp:200-253 (check_levels as computed there);
     * KeyFrame::GetFeaturesInArea, KeyFrame.cpp:1162-1201 when kf_variant (no level filter at all). */
    void features_in_area(float qx, float qy, float r, int minLevel, int maxLevel, bool kf_variant, std::vector<int>& out) const {
        out.clear();
        const int nMinCellX = std::max(0, (int)std::floor((qx - min_x - r) * inv_w));
        if (nMinCellX >= cols) return;
        const int nMaxCellX = std::min(cols - 1, (int)std::ceil((qx - min_x + r) * inv_w));
        if (nMaxCellX < 0) return;
        const int nMinCellY = std::max(0, (int)std::floor((qy - min_y - r) * inv_h));
        if (nMinCellY >= rows) return;
        const int nMaxCellY = std::min(rows - 1, (int)std::ceil((qy - min_y + r) * inv_h));
        if (nMaxCellY < 0) return;
        const bool bCheckLevels = !kf_variant && ((minLevel > 0) || (maxLevel >= 0));
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<int>& cell = grid[(size_t)ix * rows + iy];
                for (size_t j = 0; j < cell.size(); j++) {
                    const int k = cell[j];
                    if (bCheckLevels) {
                        if (octave[k] < minLevel) continue;
                        if (maxLevel >= 0 && octave[k] > maxLevel) continue;
                    }
                    const float distx = x[k] - qx, disty = y[k] - qy;
                    if (std::fabs(distx) < r && std::fabs(disty) < r) out.push_back(k);
                }
            }
    }
};

}  // namespace

extern "C" {

int mo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

void* mo_frame_create(int n, const uint8_t* desc, const float* x, const float* y, const int* octave, const float* angle,
                      float min_x, float min_y, float inv_w, float inv_h, int cols, int rows) {
    MoFrame* f = new MoFrame();
    f->n = n;
    f->desc.assign(desc, desc + (size_t)n * 32);
    f->x.assign(x, x + n); f->y.assign(y, y + n); f->angle.assign(angle, angle + n); f->octave.assign(octave, octave + n);
    f->min_x = min_x; f->min_y = min_y; f->inv_w = inv_w; f->inv_h = inv_h; f->cols = cols; f->rows = rows;
    f->assign();
    return f;
}
void mo_frame_destroy(void* f) { delete (MoFrame*)f; }

void mo_frame_params(const void* fp, float* min_x, float* min_y, float* inv_w, float* inv_h, int* cols, int* rows) {
    const MoFrame* f = (const MoFrame*)fp;
    *min_x = f->min_x; *min_y = f->min_y; *inv_w = f->inv_w; *inv_h = f->inv_h; *cols = f->cols; *rows = f->rows;
}
/* read-only view of a frame's arrays (used by oracle/ref_match_glue.cpp to build the mock Frame / KeyFrame) */
void mo_frame_view(const void* fp, int* n, const uint8_t** desc, const float** x, const float** y, const int** octave, const float** angle) {
    const MoFrame* f = (const MoFrame*)fp;
    *n = f->n; *desc = f->desc.data(); *x = f->x.data(); *y = f->y.data(); *octave = f->octave.data(); *angle = f->angle.data();
}

/* the grid as CSR over cells in (ix * rows + iy) order: cell_start[cols*rows+1], items[<= n]; returns #items */
int mo_frame_grid(const void* fp, int* cell_start, int* items) {
    const MoFrame* f = (const MoFrame*)fp;
    int k = 0;
    for (size_t c = 0; c < f->grid.size(); c++) {
        cell_start[c] = k;
        for (int i : f->grid[c]) items[k++] = i;
    }
    cell_start[f->grid.size()] = k;
    return k;
}

int mo_features_in_area(const void* fp, float x, float y, float r, int minLevel, int maxLevel, int kf_variant, int* out, int cap) {
    std::vector<int> v;
    ((const MoFrame*)fp)->features_in_area(x, y, r, minLevel, maxLevel, kf_variant != 0, v);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
    return (int)v.size();
}

}  // extern "C"

/* ---------------------------------------------------------------------------------------------------------
 * Search functions. Conventions shared with the C ABI (include/ccm_match.h):
 *   holder[idx] (in/out, one int per target feature) stands for F.mvpMapPoints[idx] / vpMatched[idx]:
 *     -1 = no map point, -2 = a map point with Observations()>0 (blocks the feature), -3 = a map point with
 *     Observations()==0 (does not block); on return a claimed feature holds the claiming query's index (>= 0).
 *   q_valid[i] = the host-side early `continue`s of the reference loop did not fire for query i.
 *   q_obs[i]   = pMP->Observations()>0 of query i's map point (what a LATER query sees when it meets the claim).
 *   res[i]     = {best_idx, best_dist, best_dist2, accepted} per query (best_idx -1 when no candidate).
 * --------------------------------------------------------------------------------------------------------- */
struct mo_result { int best_idx, best_dist, best_dist2, accepted; };

static inline bool blocked_by_holder(int hv, const uint8_t* q_obs) {
    if (hv == -1 || hv == -3) return false;
    if (hv == -2) return true;
    return q_obs ? q_obs[hv] != 0 : true;
}

extern "C" {

/* ORBmatcher::SearchByProjection(Frame&, const vector<mpptr>&, th), ORBmatcher.cpp:71-148 */
int mo_search_projection_mappoints(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                                   const int* q_level, const float* q_viewcos, const uint8_t* q_desc, const uint8_t* q_obs,
                                   float th, float nnratio, const float* scale_factors, int* holder, mo_result* res) {
    const MoFrame& F = *(const MoFrame*)fp;
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < nq; iMP++) {
        if (res) res[iMP] = mo_result{-1, 256, 256, 0};
        if (!q_valid[iMP]) continue;                                   /* mbTrackInView / isBad, :78-82 */
        const int nPredictedLevel = q_level[iMP];
        float r = q_viewcos[iMP] > 0.998 ? 2.5 : 4.0;                  /* RadiusByViewingCos, :150-156 */
        if (bFactor) r *= th;
        F.features_in_area(q_u[iMP], q_v[iMP], r * scale_factors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel, false, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* MPdescriptor = q_desc + (size_t)iMP * 32;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (blocked_by_holder(holder[idx], q_obs)) continue;       /* :113-115 */
            const int dist = descriptor_distance(MPdescriptor, &F.desc[(size_t)idx * 32]);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F.octave[idx]; bestIdx = idx; }
            else if (dist < bestDist2) { bestLevel2 = F.octave[idx]; bestDist2 = dist; }
        }
        if (res) { res[iMP].best_idx = bestIdx; res[iMP].best_dist = bestDist; res[iMP].best_dist2 = bestDist2; }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;   /* int > float*int, :139 */
            holder[bestIdx] = iMP;
            nmatches++;
            if (res) res[iMP].accepted = 1;
        }
    }
    return nmatches;
}

/* ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th), ORBmatcher.cpp:1350-1476.
 * q_* describe the last frame's features that own a non-outlier map point projecting inside the image (the
 * projection itself, :1378-1397, is host geometry); q_octave = LastFrame.mvKeys[i].octave, q_angle = its angle. */
int mo_search_projection_lastframe(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                                   const int* q_octave, const float* q_angle, const uint8_t* q_desc, const uint8_t* q_obs,
                                   float th, const float* scale_factors, int check_orientation, int* holder, mo_result* res) {
    const MoFrame& F = *(const MoFrame*)fp;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vIndices2;
    for (int i = 0; i < nq; i++) {
        if (res) res[i] = mo_result{-1, 256, 256, 0};
        if (!q_valid[i]) continue;
        const int nLastOctave = q_octave[i];
        const float radius = th * scale_factors[nLastOctave];
        F.features_in_area(q_u[i], q_v[i], radius, nLastOctave - 1, nLastOctave + 1, false, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = q_desc + (size_t)i * 32;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (blocked_by_holder(holder[i2], q_obs)) continue;        /* :1417-1419 */
            const int dist = descriptor_distance(dMP, &F.desc[(size_t)i2 * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (res) { res[i].best_idx = bestIdx2; res[i].best_dist = bestDist; }
        if (bestDist <= TH_HIGH) {
            holder[bestIdx2] = i;
            nmatches++;
            if (res) res[i].accepted = 1;
            if (check_orientation) rotHist[rot_bin(q_angle[i], F.angle[bestIdx2])].push_back(bestIdx2);
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { holder[idx] = -1; nmatches--; }
    }
    return nmatches;
}

/* ORBmatcher::SearchByProjection(kfptr, Scw, vpPoints, vpMatched, th), ORBmatcher.cpp:308-446 — the search part.
 * q_valid covers :333-376; q_level = PredictScale; q_claims[i] = (pMP->GetIndexInKeyFrame(pKF) == -1), i.e. an
 * accepted match is written into vpMatched (:436-440) and blocks later queries; otherwise the reference remaps an
 * existing observation (:414-433), which the host shim replays from res[]. holder: -1 free, -2 vpMatched[idx] set. */
int mo_search_projection_kf_sim3(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                                 const int* q_level, const uint8_t* q_desc, const uint8_t* q_claims, int th,
                                 const float* scale_factors, int* holder, mo_result* res) {
    const MoFrame& F = *(const MoFrame*)fp;
    int nmatches = 0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < nq; iMP++) {
        if (res) res[iMP] = mo_result{-1, 256, 256, 0};
        if (!q_valid[iMP]) continue;
        const int nPredictedLevel = q_level[iMP];
        const float radius = th * scale_factors[nPredictedLevel];
        F.features_in_area(q_u[iMP], q_v[iMP], radius, -1, -1, true, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = q_desc + (size_t)iMP * 32;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            if (holder[idx] != -1) continue;                           /* vpMatched[idx], :393 */
            const int kpLevel = F.octave[idx];
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = descriptor_distance(dMP, &F.desc[(size_t)idx * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (res) { res[iMP].best_idx = bestIdx; res[iMP].best_dist = bestDist; }
        if (bestDist <= TH_LOW) {
            if (res) res[iMP].accepted = 1;
            if (q_claims[iMP]) { holder[bestIdx] = iMP; nmatches++; }
        }
    }
    return nmatches;
}

/* Stateless best-candidate search shared by SearchBySim3's two passes (ORBmatcher.cpp:1216-1252, :1296-1327; accept
 * <= TH_HIGH, INT_MAX start) and the distance phase of both Fuse overloads (:929-962 with the chi2 gate
 * e2*invSigma2[level] > 5.99, :1082-1103 without). KeyFrame grid, level window [pred-1, pred] applied after the gather.
 * res[i].accepted = bestDist <= th_dist. inv_sigma2 == NULL disables the gate. */
void mo_search_best_in_window(const void* fp, int nq, const uint8_t* q_valid, const float* q_u, const float* q_v,
                              const float* q_radius, const int* q_level, const uint8_t* q_desc, int th_dist,
                              const float* inv_sigma2, mo_result* res) {
    const MoFrame& F = *(const MoFrame*)fp;
    std::vector<int> vIndices;
    for (int i = 0; i < nq; i++) {
        res[i] = mo_result{-1, 256, 256, 0};
        if (!q_valid[i]) continue;
        const int nPredictedLevel = q_level[i];
        F.features_in_area(q_u[i], q_v[i], q_radius[i], -1, -1, true, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = q_desc + (size_t)i * 32;
        int bestDist = 256, bestIdx = -1;     /* INT_MAX in SearchBySim3/Fuse(Scw): same outcome, distances are <= 256 */
        for (int idx : vIndices) {
            const int kpLevel = F.octave[idx];
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (inv_sigma2) {
                const float ex = q_u[i] - F.x[idx], ey = q_v[i] - F.y[idx];
                const float e2 = ex * ex + ey * ey;
                if (e2 * inv_sigma2[kpLevel] > 5.99) continue;
            }
            const int dist = descriptor_distance(dMP, &F.desc[(size_t)idx * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        res[i].best_idx = bestIdx; res[i].best_dist = bestDist;
        res[i].accepted = (bestIdx >= 0 && bestDist <= th_dist) ? 1 : 0;
    }
}

}  // extern "C"

/* ---------------------------------------------------------------------------------------------------------
 * BoW-guided searches. A DBoW2::FeatureVector is map<NodeId, vector<feature idx>> filled by addFeature(nid, i)
 * for i ascending (thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1148-1161, FeatureVector.cpp:31-45), so it is
 * fully described by node_id[i] per feature (-1 = stopped word, absent from the vector): nodes iterate in
 * ascending id, features inside a node in ascending index. The merge-join of the two vectors (:203-278) visits
 * exactly the node ids present on both sides, ascending.
 * --------------------------------------------------------------------------------------------------------- */
namespace {
struct NodeLists { std::vector<int> ids; std::vector<std::vector<int>> items; };
NodeLists build_feature_vector(int n, const int* node) {
    std::vector<std::pair<int, int>> v;
    for (int i = 0; i < n; i++) if (node[i] >= 0) v.push_back({node[i], i});
    std::sort(v.begin(), v.end());
    NodeLists L;
    for (auto& p : v) {
        if (L.ids.empty() || L.ids.back() != p.first) { L.ids.push_back(p.first); L.items.emplace_back(); }
        L.items.back().push_back(p.second);
    }
    return L;
}
}  // namespace

extern "C" {

/* ORBmatcher::SearchByBoW(kfptr pKF, Frame& F, vpMapPointMatches), ORBmatcher.cpp:178-306.
 * valid_kf[i] = pKF's feature i has a good map point (:213-219). match_of_f[nF] out: KF feature index or -1. */
int mo_search_bow_kf_frame(int nKF, const uint8_t* desc_kf, const int* node_kf, const uint8_t* valid_kf, const float* angle_kf,
                           int nF, const uint8_t* desc_f, const int* node_f, const float* angle_f,
                           float nnratio, int check_orientation, int* match_of_f, mo_result* res) {
    for (int i = 0; i < nF; i++) match_of_f[i] = -1;
    if (res) for (int i = 0; i < nKF; i++) res[i] = mo_result{-1, 256, 256, 0};
    const NodeLists A = build_feature_vector(nKF, node_kf), B = build_feature_vector(nF, node_f);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    size_t a = 0, b = 0;
    while (a < A.ids.size() && b < B.ids.size()) {
        if (A.ids[a] == B.ids[b]) {
            const std::vector<int>& vIndicesKF = A.items[a];
            const std::vector<int>& vIndicesF = B.items[b];
            for (int realIdxKF : vIndicesKF) {
                if (!valid_kf[realIdxKF]) continue;
                const uint8_t* dKF = desc_kf + (size_t)realIdxKF * 32;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (int realIdxF : vIndicesF) {
                    if (match_of_f[realIdxF] >= 0) continue;                       /* :228 */
                    const int dist = descriptor_distance(dKF, desc_f + (size_t)realIdxF * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                    else if (dist < bestDist2) { bestDist2 = dist; }
                }
                if (res) { res[realIdxKF].best_idx = bestIdxF; res[realIdxKF].best_dist = bestDist1; res[realIdxKF].best_dist2 = bestDist2; }
                if (bestDist1 <= TH_LOW) {
                    if ((float)bestDist1 < nnratio * (float)bestDist2) {
                        match_of_f[bestIdxF] = realIdxKF;
                        if (res) res[realIdxKF].accepted = 1;
                        if (check_orientation) rotHist[rot_bin(angle_kf[realIdxKF], angle_f[bestIdxF])].push_back(bestIdxF);
                        nmatches++;
                    }
                }
            }
            a++; b++;
        } else if (A.ids[a] < B.ids[b]) a++;      /* lower_bound(), :270-277: same visiting order */
        else b++;
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match_of_f[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

/* ORBmatcher::SearchByBoW(kfptr pKF1, kfptr pKF2, vpMatches12), ORBmatcher.cpp:565-698.
 * valid1/valid2 = feature has a good map point. match12[n1] out: KF2 feature index or -1. Note `< TH_LOW` (strict). */
int mo_search_bow_kf_kf(int n1, const uint8_t* desc1, const int* node1, const uint8_t* valid1, const float* angle1,
                        int n2, const uint8_t* desc2, const int* node2, const uint8_t* valid2, const float* angle2,
                        float nnratio, int check_orientation, int* match12, mo_result* res) {
    for (int i = 0; i < n1; i++) match12[i] = -1;
    if (res) for (int i = 0; i < n1; i++) res[i] = mo_result{-1, 256, 256, 0};
    std::vector<char> vbMatched2(n2, 0);
    const NodeLists A = build_feature_vector(n1, node1), B = build_feature_vector(n2, node2);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    size_t a = 0, b = 0;
    while (a < A.ids.size() && b < B.ids.size()) {
        if (A.ids[a] == B.ids[b]) {
            for (int idx1 : A.items[a]) {
                if (!valid1[idx1]) continue;
                const uint8_t* d1 = desc1 + (size_t)idx1 * 32;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int idx2 : B.items[b]) {
                    if (vbMatched2[idx2] || !valid2[idx2]) continue;               /* :619-623 */
                    const int dist = descriptor_distance(d1, desc2 + (size_t)idx2 * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                    else if (dist < bestDist2) { bestDist2 = dist; }
                }
                if (res) { res[idx1].best_idx = bestIdx2; res[idx1].best_dist = bestDist1; res[idx1].best_dist2 = bestDist2; }
                if (bestDist1 < TH_LOW) {
                    if ((float)bestDist1 < nnratio * (float)bestDist2) {
                        match12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = 1;
                        if (res) res[idx1].accepted = 1;
                        if (check_orientation) rotHist[rot_bin(angle1[idx1], angle2[bestIdx2])].push_back(idx1);
                        nmatches++;
                    }
                }
            }
            a++; b++;
        } else if (A.ids[a] < B.ids[b]) a++;
        else b++;
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match12[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs), ORBmatcher.cpp:700-852, with
 * CheckDistEpipolarLine :159-176. free1/free2 = the feature holds NO map point (:739-744, :758-762); f12 = F12 row-major;
 * (ex, ey) = the epipole the reference computes at :707-714 (host geometry, an input here). vbMatched2 is never set
 * upstream, so a KF2 feature may be matched by several KF1 features. match12[n1] out: KF2 feature index or -1
 * (vMatches12; vMatchedPairs = its non-negative entries in index order). Float expressions keep the reference's
 * operation order; the oracle is built with -ffp-contract=off. */
int mo_search_for_triangulation(int n1, const uint8_t* desc1, const int* node1, const uint8_t* free1, const float* x1,
                                const float* y1, const float* angle1, int n2, const uint8_t* desc2, const int* node2,
                                const uint8_t* free2, const float* x2, const float* y2, const int* octave2,
                                const float* angle2, const float* f12, float ex, float ey, const float* scale_factors,
                                const float* level_sigma2, int check_orientation, int* match12) {
    for (int i = 0; i < n1; i++) match12[i] = -1;
    const NodeLists A = build_feature_vector(n1, node1), B = build_feature_vector(n2, node2);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    size_t a = 0, b = 0;
    while (a < A.ids.size() && b < B.ids.size()) {
        if (A.ids[a] == B.ids[b]) {
            for (int idx1 : A.items[a]) {
                if (!free1[idx1]) continue;
                const uint8_t* d1 = desc1 + (size_t)idx1 * 32;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int idx2 : B.items[b]) {
                    if (!free2[idx2]) continue;
                    const int dist = descriptor_distance(d1, desc2 + (size_t)idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;                   /* :768: ties move to the later one */
                    const float distex = ex - x2[idx2], distey = ey - y2[idx2];
                    if (distex * distex + distey * distey < 100 * scale_factors[octave2[idx2]]) continue;   /* :777 */
                    /* CheckDistEpipolarLine :159-176 */
                    const float la = x1[idx1] * f12[0] + y1[idx1] * f12[3] + f12[6];
                    const float lb = x1[idx1] * f12[1] + y1[idx1] * f12[4] + f12[7];
                    const float lc = x1[idx1] * f12[2] + y1[idx1] * f12[5] + f12[8];
                    const float num = la * x2[idx2] + lb * y2[idx2] + lc;
                    const float den = la * la + lb * lb;
                    if (den == 0) continue;
                    const float dsqr = num * num / den;
                    if (dsqr < 3.84 * level_sigma2[octave2[idx2]]) { bestIdx2 = idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    match12[idx1] = bestIdx2;
                    nmatches++;
                    if (check_orientation) rotHist[rot_bin(angle1[idx1], angle2[bestIdx2])].push_back(idx1);
                }
            }
            a++; b++;
        } else if (A.ids[a] < B.ids[b]) a++;
        else b++;
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match12[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

/* DBoW2 TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), thirdparty/DBoW2/DBoW2/
 * TemplatedVocabulary.h:1218-1261 — the per-feature half of Frame::ComputeBoW (cslam/src/Frame.cpp:268-275, levelsup 4):
 * from the root, move to the child at the smallest FORB::distance (Hamming, FORB.cpp:103-121; the FIRST child wins ties)
 * until a node without children; nid = the node reached at level L - levelsup (0 = root if that level is <= 0). If the
 * descent ends ABOVE that level upstream never assigns nid (the batch caller then reads an uninitialised local, :1150-1154:
 * undefined, and impossible with the complete ORB vocabulary); 0 is returned here. The tree comes as CSR children lists in file order
 * (loadFromTextFile, :1338-1420). */
void mo_vocab_transform(const int* child_start, const int* child_ids, const uint8_t* node_desc, const double* node_weight,
                        const int* node_word, int L, int n, const uint8_t* desc, int levelsup, int* word, double* weight, int* node) {
    const int nid_level = L - levelsup;
    for (int i = 0; i < n; i++) {
        const uint8_t* f = desc + (size_t)32 * i;
        int nid = 0, final_id = 0, current_level = 0;
        do {
            ++current_level;
            const int c0 = child_start[final_id], c1 = child_start[final_id + 1];
            int best = child_ids[c0];
            int best_d = descriptor_distance(f, node_desc + (size_t)32 * best);
            for (int c = c0 + 1; c < c1; c++) {
                const int id = child_ids[c];
                const int d = descriptor_distance(f, node_desc + (size_t)32 * id);
                if (d < best_d) { best_d = d; best = id; }
            }
            final_id = best;
            if (current_level == nid_level) nid = final_id;
        } while (child_start[final_id + 1] > child_start[final_id]);
        word[i] = node_word[final_id]; weight[i] = node_weight[final_id]; node[i] = nid;
    }
}

/* MapPoint::ComputeDistinctiveDescriptors, cslam/src/MapPoint.cpp:929-994, for map points given as CSR row ranges of one
 * descriptor array (rows of a point in the order the reference's observation map would iterate them). PINNED: MapPoint.cpp
 * does not compile outside the ROS workspace (ccmslam_msgs, Map, Communicator), so oracle/Makefile cuts lines 929-994 out
 * of it and compiles them as they are over oracle/pinshim (oracle/_ref/libref_pin.so); tests/test_pin_oracle_vs_ref.py
 * holds this restatement against that, tests/test_match_oracle.py against a numpy median. Keeps the float distance
 * matrix, the sort and the vDists[0.5*(N-1)] index of the original. */
void mo_distinctive_descriptors(int n_points, const int* start, const uint8_t* desc, int* best, int* best_median) {
    for (int p = 0; p < n_points; p++) {
        const int r0 = start[p], N = start[p + 1] - r0;
        if (N <= 0) { best[p] = -1; if (best_median) best_median[p] = 0; continue; }
        std::vector<float> D((size_t)N * N);
        for (int i = 0; i < N; i++) {
            D[(size_t)i * N + i] = 0;
            for (int j = i + 1; j < N; j++) {
                const int d = descriptor_distance(desc + (size_t)32 * (r0 + i), desc + (size_t)32 * (r0 + j));
                D[(size_t)i * N + j] = (float)d; D[(size_t)j * N + i] = (float)d;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (int i = 0; i < N; i++) {
            std::vector<int> vDists(D.begin() + (size_t)i * N, D.begin() + (size_t)(i + 1) * N);
            std::sort(vDists.begin(), vDists.end());
            const int median = vDists[(size_t)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        best[p] = BestIdx; if (best_median) best_median[p] = BestMedian;
    }
}

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), ORBmatcher.cpp:448-563.
 * f2 = grid of F2; F1 is given by its arrays. prev_xy[2*n1] in/out (vbPrevMatched). */
int mo_search_for_initialization(int n1, const uint8_t* desc1, const int* octave1, const float* angle1, float* prev_xy,
                                 const void* f2, int windowSize, float nnratio, int check_orientation, int* matches12) {
    const MoFrame& F2 = *(const MoFrame*)f2;
    int nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(F2.n, INT_MAX), vnMatches21(F2.n, -1);
    std::vector<int> vIndices2;
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = octave1[i1];
        if (level1 > 0) continue;
        F2.features_in_area(prev_xy[2 * i1], prev_xy[2 * i1 + 1], (float)windowSize, level1, level1, false, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = desc1 + (size_t)i1 * 32;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = descriptor_distance(d1, &F2.desc[(size_t)i2 * 32]);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) { bestDist2 = dist; }
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (check_orientation) rotHist[rot_bin(angle1[i1], F2.angle[bestIdx2])].push_back(i1);
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        compute_three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (matches12[i1] >= 0) { prev_xy[2 * i1] = F2.x[matches12[i1]]; prev_xy[2 * i1 + 1] = F2.y[matches12[i1]]; }
    return nmatches;
}

/* Server keyframe-database search (BASELINE.json configs[4]; SURVEY.md §0 mismatch 2, §8a SearchByBoW(KF,KF) row):
 * the inner loop of SearchByBoW(KF,KF), ORBmatcher.cpp:607-637, with one all-inclusive vocabulary node and no
 * cross-query state: for every query descriptor the best and second-best Hamming distance over ALL target
 * descriptors, first-wins on ties (strict <, ascending target index). out: best_idx/best_dist/best_dist2;
 * accepted = best < TH_LOW && (float)best < nnratio*(float)best2 (:641-643). */
void mo_match_allpairs_best2(int nq, const uint8_t* q_desc, long long nt, const uint8_t* t_desc, float nnratio, mo_result* res) {
    for (int i = 0; i < nq; i++) {
        const uint8_t* d1 = q_desc + (size_t)i * 32;
        int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
        for (long long j = 0; j < nt; j++) {
            const int dist = descriptor_distance(d1, t_desc + (size_t)j * 32);
            if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = (int)j; }
            else if (dist < bestDist2) { bestDist2 = dist; }
        }
        res[i].best_idx = bestIdx2; res[i].best_dist = bestDist1; res[i].best_dist2 = bestDist2;
        res[i].accepted = (bestDist1 < TH_LOW && (float)bestDist1 < nnratio * (float)bestDist2) ? 1 : 0;
    }
}

/* Keyframe wire format — PINNED to the reference's own conversion statements: cslam/src/Converter.cc needs Eigen and g2o and
 * ccmslam_msgs needs ROS genmsg, none of which exist in this image, so oracle/Makefile cuts Converter.cc:166-193 and
 * KeyFrame.cpp:1462-1472 / 1688-1704 out of the reference and compiles them as they are over message structs with the .msg
 * fields (oracle/_ref/libref_pin.so; tests/test_pin_oracle_vs_ref.py); the ROS 1 byte layout itself (no generated
 * serialiser in the repository) is also checked against an independent numpy packing (tests/test_match_oracle.py). KeyFrame::ConvertToMessage (cslam/src/KeyFrame.cpp:1462-1472) pushes
 * Converter::toCvKeyPointMsg(mvKeysUn[i]) (Converter.cc:166-179: x, y, angle copied; octave to int8; response and size
 * through (u_int8_t)float) and one Descriptor per row; ROS 1 serialises CvKeyPoint[] / Descriptor[] (cslam_msgs/msg) as a
 * u32 count followed by the unpadded little-endian fields in declaration order:
 * fPoint2f_x, fPoint2f_y, size, angle, response, octave = 15 bytes; uint8[32] = 32 bytes. */
void mo_wire_encode_keyframe(int n, const float* x, const float* y, const float* size, const float* angle, const float* response,
                             const int* octave, const uint8_t* desc, uint8_t* kp_bytes, uint8_t* desc_bytes) {
    const uint32_t cnt = (uint32_t)n;
    std::memcpy(kp_bytes, &cnt, 4);
    std::memcpy(desc_bytes, &cnt, 4);
    for (int i = 0; i < n; i++) {
        uint8_t* r = kp_bytes + 4 + (size_t)15 * i;
        const uint8_t sz = (uint8_t)size[i], rs = (uint8_t)response[i];     /* (u_int8_t)kp.size, (u_int8_t)kp.response */
        const int8_t oc = (int8_t)octave[i];                                /* Msg.octave = kp.octave (int8) */
        std::memcpy(r, &x[i], 4); std::memcpy(r + 4, &y[i], 4); r[8] = sz; std::memcpy(r + 9, &angle[i], 4); r[13] = rs;
        std::memcpy(r + 14, &oc, 1);
        for (int b = 0; b < 32; b++) desc_bytes[4 + (size_t)32 * i + b] = desc[(size_t)32 * i + b];   /* KeyFrame.cpp:1464-1471 */
    }
}

/* KeyFrame::WriteMembersFromMessage, KeyFrame.cpp:1689-1702 + Converter::fromCvKeyPointMsg, Converter.cc:181-193.
 * Returns n, or -1 when the two counts disagree. */
int mo_wire_decode_keyframe(const uint8_t* kp_bytes, const uint8_t* desc_bytes, int cap, float* x, float* y, float* size, float* angle,
                            float* response, int* octave, uint8_t* desc) {
    uint32_t n, nd;
    std::memcpy(&n, kp_bytes, 4);
    std::memcpy(&nd, desc_bytes, 4);
    if (n != nd || (int)n > cap) return -1;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t* r = kp_bytes + 4 + (size_t)15 * i;
        int8_t oc;
        std::memcpy(&x[i], r, 4); std::memcpy(&y[i], r + 4, 4); std::memcpy(&angle[i], r + 9, 4); std::memcpy(&oc, r + 14, 1);
        size[i] = (float)r[8]; response[i] = (float)r[13]; octave[i] = oc;
        std::memcpy(desc + (size_t)32 * i, desc_bytes + 4 + (size_t)32 * i, 32);
    }
    return (int)n;
}

}  // extern "C"
