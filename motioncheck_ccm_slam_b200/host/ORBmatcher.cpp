// Drop-in replacement bodies for cslam/src/ORBmatcher.cpp of taiyaki-go/motioncheck_ccm_slam: the SAME class declaration
// (this file includes the reference's own, unmodified <cslam/ORBmatcher.h>); each search keeps the host-side geometry
// the reference performs before its candidate loop (so u, v, radius and level are bit-identical), packs one
// ccm_window_query per loop iteration that reaches GetFeaturesInArea, and lets the CUDA library (include/ccm_match.h)
// run the candidate loops, including their greedy order and the rotation-histogram pruning. The state the reference
// loop would have left (F.mvpMapPoints, vpMatched, vnMatches12, ...) is written back from the returned arrays.
//
// Converted here: DescriptorDistance, SearchByProjection(Frame, MapPoints), SearchByProjection(Frame, LastFrame),
// SearchByProjection(KeyFrame, Scw), SearchByBoW(KeyFrame, Frame), SearchByBoW(KeyFrame, KeyFrame),
// SearchForInitialization, SearchForTriangulation, SearchBySim3 and both Fuse overloads (candidate search on the device,
// CCM_WIN_BEST; the map mutations are replayed on the host in the reference's order), and the relocalisation
// SearchByProjection (:1478-1605, no caller upstream) on the entry point of the LastFrame search. A CUDA failure throws estd::infrastructure_ex (the
// reference's fatal-error convention, estd.h:74); there is no CPU fallback.
#include <cslam/ORBmatcher.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

#include "ccm_match.h"

namespace {

struct Device {
    ccm_matcher* m = nullptr;
    std::mutex mu;
};

Device& device() {
    static Device d;
    static std::once_flag once;
    std::call_once(once, [] {                                 // a failed create throws and leaves the flag unset: the next call retries
        const char* e = std::getenv("CCM_ORB_DEVICE");       // agent -> GPU pinning, same variable as the extractor
        if (ccm_matcher_create(e ? std::atoi(e) : 0, &d.m) != CCM_OK) throw estd::infrastructure_ex();
    });
    return d;
}

// Device copy of what the matcher reads of a Frame / KeyFrame (descriptors, undistorted keypoints, the feature grid).
// Tracking searches the same Frame two or three times per image and LoopFinder / MapMatcher hit the same KeyFrame many
// times, so packs live in a small LRU side table keyed by the object's identity: Frame::mId / KeyFrame::mId
// (Frame.h:156, KeyFrame.h:282; copies of a Frame share the id and the features), the feature count and a checksum of a
// few descriptor rows and keypoints. A hit costs no upload and no grid build. CCM_ORB_PACK_CACHE = entries (default 16,
// 0 = one pack per call as before); ccm_dropin_pack_cache_clear() drops everything (e.g. from a map reset).
struct PackKey {
    int kind; size_t id0, id1; int n; uint64_t sum;
    bool operator<(const PackKey& o) const {
        return std::tie(kind, id0, id1, n, sum) < std::tie(o.kind, o.id0, o.id1, o.n, o.sum);
    }
};
struct PackEntry { ccm_frame_pack* p = nullptr; unsigned long long used = 0; };
struct PackCache {
    std::map<PackKey, PackEntry> m;
    unsigned long long tick = 0, hits = 0, misses = 0;
    int capacity = -1;
};
PackCache& pack_cache() { static PackCache c; return c; }

uint64_t fnv(uint64_t h, const void* p, size_t n) {
    const unsigned char* b = (const unsigned char*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

struct Pack {
    ccm_frame_pack* p = nullptr;
    bool owned = true;
    // kind: 0 Frame, 1 KeyFrame; id: the object's mId
    Pack(ccm_matcher* m, int kind, const std::pair<size_t, size_t>& id, int n, const cv::Mat& desc, const std::vector<cv::KeyPoint>& keysUn,
         float minX, float minY, float invW, float invH, int cols, int rows) {
        PackCache& C = pack_cache();
        if (C.capacity < 0) { const char* e = std::getenv("CCM_ORB_PACK_CACHE"); C.capacity = e ? std::max(0, std::atoi(e)) : 16; }
        PackKey key = {kind, id.first, id.second, n, 0};
        if (C.capacity > 0 && n > 0) {
            uint64_t h = 1469598103934665603ull;
            for (int r : {0, n / 2, n - 1}) { h = fnv(h, desc.ptr(r), 32); h = fnv(h, &keysUn[r].pt, sizeof(keysUn[r].pt)); h = fnv(h, &keysUn[r].octave, sizeof(int)); }
            const float gp[4] = {minX, minY, invW, invH};
            key.sum = fnv(h, gp, sizeof(gp));
            std::map<PackKey, PackEntry>::iterator it = C.m.find(key);
            if (it != C.m.end()) { it->second.used = ++C.tick; C.hits++; p = it->second.p; owned = false; return; }
        }
        const int cap = n > 0 ? n : 1;
        if (ccm_frame_pack_create(m, cap, &p) != CCM_OK) throw estd::infrastructure_ex();
        std::vector<float> x(cap), y(cap), ang(cap);
        std::vector<int> oct(cap);
        std::vector<uint8_t> d((size_t)cap * 32);
        for (int i = 0; i < n; i++) {
            x[i] = keysUn[i].pt.x; y[i] = keysUn[i].pt.y; ang[i] = keysUn[i].angle; oct[i] = keysUn[i].octave;
            std::memcpy(&d[(size_t)32 * i], desc.ptr(i), 32);
        }
        ccm_grid_params g = {minX, minY, invW, invH, cols, rows};
        if (ccm_frame_pack_upload(p, &g, n, d.data(), x.data(), y.data(), oct.data(), ang.data()) != CCM_OK) {
            ccm_frame_pack_destroy(p);
            throw estd::infrastructure_ex();
        }
        if (C.capacity > 0 && n > 0) {
            C.misses++;
            if ((int)C.m.size() >= C.capacity) {                      // evict the least recently used pack
                std::map<PackKey, PackEntry>::iterator lru = C.m.begin();
                for (std::map<PackKey, PackEntry>::iterator it = C.m.begin(); it != C.m.end(); ++it) if (it->second.used < lru->second.used) lru = it;
                ccm_frame_pack_destroy(lru->second.p);
                C.m.erase(lru);
            }
            PackEntry e; e.p = p; e.used = ++C.tick;
            C.m[key] = e;
            owned = false;
        }
    }
    ~Pack() { if (owned && p) ccm_frame_pack_destroy(p); }
    Pack(const Pack&) = delete;
    Pack& operator=(const Pack&) = delete;
};

inline ccm_window_query no_query() { ccm_window_query q; std::memset(&q, 0, sizeof(q)); return q; }

void run_window(ccm_matcher* m, ccm_frame_pack* pack, int mode, float nnratio, int th_dist, bool ori, std::vector<ccm_window_query>& q,
                std::vector<uint8_t>& qd, int* holder, ccm_match_result* res, int& nmatches, const std::vector<float>* inv_sigma2 = nullptr) {
    ccm_window_params prm;
    prm.mode = mode; prm.nnratio = nnratio; prm.th_dist = th_dist; prm.check_orientation = ori ? 1 : 0;
    prm.inv_sigma2 = inv_sigma2 ? inv_sigma2->data() : nullptr; prm.n_levels = inv_sigma2 ? (int)inv_sigma2->size() : 0;
    nmatches = 0;
    if (q.empty()) return;
    if (ccm_search_window(m, pack, &prm, (int)q.size(), q.data(), qd.data(), holder, res, &nmatches) != CCM_OK)
        throw estd::infrastructure_ex();
}

// FeatureVector (node id -> feature indices) as one node id per feature, -1 = not in the vector
std::vector<int> node_ids(const DBoW2::FeatureVector& fv, int n) {
    std::vector<int> node(n > 0 ? n : 1, -1);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it)
        for (size_t k = 0; k < it->second.size(); k++)
            if ((int)it->second[k] < n) node[it->second[k]] = (int)it->first;
    return node;
}

}  // namespace

namespace cslam {

const int ORBmatcher::TH_HIGH = 100;
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

// ORBmatcher.cpp:150-156
float ORBmatcher::RadiusByViewingCos(const float &viewCos) {
    if (viewCos > 0.998) return 2.5;
    return 4.0;
}

// ORBmatcher.cpp:1653-1669 — single pairs stay on the host (called from MapPoint.cpp:645,968 and Tracking.cpp:491);
// batches go through ccm_descriptor_distance.
int ORBmatcher::DescriptorDistance(const cv::Mat &a, const cv::Mat &b) {
    const uint32_t* pa = a.ptr<uint32_t>();
    const uint32_t* pb = b.ptr<uint32_t>();
    int dist = 0;
    for (int i = 0; i < 8; i++) dist += __builtin_popcount(pa[i] ^ pb[i]);
    return dist;
}

// ORBmatcher.cpp:71-148
int ORBmatcher::SearchByProjection(Frame &F, const vector<mpptr> &vpMapPoints, const float th) {
    const bool bFactor = th != 1.0;
    const int nq = (int)vpMapPoints.size();
    std::vector<ccm_window_query> q(nq, no_query());
    std::vector<uint8_t> qd((size_t)nq * 32 + 32);
    for (int i = 0; i < nq; i++) {
        mpptr pMP = vpMapPoints[i];
        if (!pMP->mbTrackInView) continue;
        if (pMP->isBad()) continue;
        const int lvl = pMP->mnTrackScaleLevel;
        float r = RadiusByViewingCos(pMP->mTrackViewCos);
        if (bFactor) r *= th;
        q[i].u = pMP->mTrackProjX; q[i].v = pMP->mTrackProjY; q[i].r = r * F.mvScaleFactors[lvl];
        q[i].min_level = (int16_t)(lvl - 1); q[i].max_level = (int16_t)lvl;
        q[i].flags = CCM_Q_VALID | (pMP->Observations() > 0 ? CCM_Q_OBS : 0u);
        std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
    }
    std::vector<int> holder(F.N > 0 ? F.N : 1, -1);
    for (int k = 0; k < F.N; k++)
        if (F.mvpMapPoints[k]) holder[k] = F.mvpMapPoints[k]->Observations() > 0 ? -2 : -3;
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    Pack pack(D.m, 0, F.mId, F.N, F.mDescriptors, F.mvKeysUn, Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv,
              Frame::mfGridElementHeightInv, FRAME_GRID_COLS, FRAME_GRID_ROWS);
    int nmatches = 0;
    run_window(D.m, pack.p, CCM_WIN_PROJ_MAPPOINTS, mfNNratio, TH_HIGH, mbCheckOrientation, q, qd, holder.data(), nullptr, nmatches);
    for (int k = 0; k < F.N; k++)
        if (holder[k] >= 0) F.mvpMapPoints[k] = vpMapPoints[holder[k]];
    return nmatches;
}

// ORBmatcher.cpp:1350-1476
int ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th) {
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const int nq = LastFrame.N;
    std::vector<ccm_window_query> q(nq, no_query());
    std::vector<uint8_t> qd((size_t)nq * 32 + 32);
    for (int i = 0; i < nq; i++) {
        mpptr pMP = LastFrame.mvpMapPoints[i];
        if (!pMP || LastFrame.mvbOutlier[i]) continue;
        cv::Mat x3Dw = pMP->GetWorldPos();
        cv::Mat x3Dc = Rcw * x3Dw + tcw;
        const float xc = x3Dc.at<float>(0), yc = x3Dc.at<float>(1);
        const float invzc = 1.0 / x3Dc.at<float>(2);
        if (invzc < 0) continue;
        const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        const int nLastOctave = LastFrame.mvKeys[i].octave;
        q[i].u = u; q[i].v = v; q[i].r = th * CurrentFrame.mvScaleFactors[nLastOctave];
        q[i].angle = LastFrame.mvKeysUn[i].angle;
        q[i].min_level = (int16_t)(nLastOctave - 1); q[i].max_level = (int16_t)(nLastOctave + 1);
        q[i].flags = CCM_Q_VALID | (pMP->Observations() > 0 ? CCM_Q_OBS : 0u);
        std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
    }
    const int N = CurrentFrame.N;
    std::vector<int> holder(N > 0 ? N : 1, -1);
    for (int k = 0; k < N; k++)
        if (CurrentFrame.mvpMapPoints[k]) holder[k] = CurrentFrame.mvpMapPoints[k]->Observations() > 0 ? -2 : -3;
    const std::vector<int> before = holder;
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    Pack pack(D.m, 0, CurrentFrame.mId, N, CurrentFrame.mDescriptors, CurrentFrame.mvKeysUn, Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv,
              Frame::mfGridElementHeightInv, FRAME_GRID_COLS, FRAME_GRID_ROWS);
    int nmatches = 0;
    run_window(D.m, pack.p, CCM_WIN_PROJ_LASTFRAME, mfNNratio, TH_HIGH, mbCheckOrientation, q, qd, holder.data(), nullptr, nmatches);
    for (int k = 0; k < N; k++) {
        if (holder[k] >= 0) CurrentFrame.mvpMapPoints[k] = LastFrame.mvpMapPoints[holder[k]];
        else if (holder[k] == -1 && before[k] != -1) CurrentFrame.mvpMapPoints[k] = mpptr();   // pruned by the rotation histogram (:1461)
    }
    return nmatches;
}

// ORBmatcher.cpp:1478-1605 — relocalisation: the key frame's map points projected with the current pose. The candidate
// loop is the one of the LastFrame search (best only, levels [l-1, l+1], rotation histogram) with two changes that fit
// the same entry point: every occupied feature blocks (:1545, no Observations() test) = holder -2, and the bound is
// ORBdist instead of TH_HIGH (:1559) = th_dist.
int ORBmatcher::SearchByProjection(Frame &CurrentFrame, kfptr pKF, const set<mpptr> &sAlreadyFound, const float th, const int ORBdist) {
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat Ow = -Rcw.t() * tcw;
    const vector<mpptr> vpMPs = pKF->GetMapPointMatches();
    const int nq = (int)vpMPs.size();
    std::vector<ccm_window_query> q(nq, no_query());
    std::vector<uint8_t> qd((size_t)nq * 32 + 32);
    for (int i = 0; i < nq; i++) {
        mpptr pMP = vpMPs[i];
        if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
        cv::Mat x3Dw = pMP->GetWorldPos();
        cv::Mat x3Dc = Rcw * x3Dw + tcw;
        const float xc = x3Dc.at<float>(0), yc = x3Dc.at<float>(1);
        const float invzc = 1.0 / x3Dc.at<float>(2);
        const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        cv::Mat PO = x3Dw - Ow;
        float dist3D = cv::norm(PO);
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int lvl = pMP->PredictScale(dist3D, CurrentFrame.shared_from_this());
        q[i].u = u; q[i].v = v; q[i].r = th * CurrentFrame.mvScaleFactors[lvl];
        q[i].angle = pKF->mvKeysUn[i].angle;
        q[i].min_level = (int16_t)(lvl - 1); q[i].max_level = (int16_t)(lvl + 1);
        q[i].flags = CCM_Q_VALID | CCM_Q_OBS;                    // every accepted match blocks the feature for later points
        std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
    }
    const int N = CurrentFrame.N;
    std::vector<int> holder(N > 0 ? N : 1, -1);
    for (int k = 0; k < N; k++)
        if (CurrentFrame.mvpMapPoints[k]) holder[k] = -2;
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    Pack pack(D.m, 0, CurrentFrame.mId, N, CurrentFrame.mDescriptors, CurrentFrame.mvKeysUn, Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv,
              Frame::mfGridElementHeightInv, FRAME_GRID_COLS, FRAME_GRID_ROWS);
    int nmatches = 0;
    run_window(D.m, pack.p, CCM_WIN_PROJ_LASTFRAME, mfNNratio, ORBdist, mbCheckOrientation, q, qd, holder.data(), nullptr, nmatches);
    for (int k = 0; k < N; k++)
        if (holder[k] >= 0) CurrentFrame.mvpMapPoints[k] = vpMPs[holder[k]];
    return nmatches;
}

// ORBmatcher.cpp:308-446
int ORBmatcher::SearchByProjection(kfptr pKF, cv::Mat Scw, const vector<mpptr> &vpPoints, vector<mpptr> &vpMatched, int th) {
    const float &fx = pKF->fx, &fy = pKF->fy, &cx = pKF->cx, &cy = pKF->cy;
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    set<mpptr> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(nullptr);
    const int nq = (int)vpPoints.size();
    std::vector<ccm_window_query> q(nq, no_query());
    std::vector<uint8_t> qd((size_t)nq * 32 + 32);
    for (int i = 0; i < nq; i++) {
        mpptr pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        cv::Mat p3Dw = pMP->GetWorldPos();
        cv::Mat p3Dc = Rcw * p3Dw + tcw;
        if (p3Dc.at<float>(2) < 0.0) continue;
        const float invz = 1 / p3Dc.at<float>(2);
        const float x = p3Dc.at<float>(0) * invz, y = p3Dc.at<float>(1) * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        cv::Mat PO = p3Dw - Ow;
        const float dist = cv::norm(PO);
        if (dist < minDistance || dist > maxDistance) continue;
        cv::Mat Pn = pMP->GetNormal();
        if (PO.dot(Pn) < 0.5 * dist) continue;
        const int lvl = pMP->PredictScale(dist, pKF);
        q[i].u = u; q[i].v = v; q[i].r = th * pKF->mvScaleFactors[lvl];
        q[i].min_level = (int16_t)(lvl - 1); q[i].max_level = (int16_t)lvl;
        q[i].flags = CCM_Q_VALID | (pMP->GetIndexInKeyFrame(pKF) == -1 ? CCM_Q_CLAIMS : 0u);
        std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
    }
    const int N = pKF->N;
    std::vector<int> holder(N > 0 ? N : 1, -1);
    for (int k = 0; k < N; k++) if (vpMatched[k]) holder[k] = -2;
    std::vector<ccm_match_result> res(nq > 0 ? nq : 1);
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    Pack pack(D.m, 1, pKF->mId, N, pKF->mDescriptors, pKF->mvKeysUn, (float)pKF->mnMinX, (float)pKF->mnMinY, pKF->mfGridElementWidthInv,
              pKF->mfGridElementHeightInv, pKF->mnGridCols, pKF->mnGridRows);
    int nmatches = 0;
    run_window(D.m, pack.p, CCM_WIN_PROJ_KF_SIM3, mfNNratio, TH_LOW, mbCheckOrientation, q, qd, holder.data(), res.data(), nmatches);
    for (int i = 0; i < nq; i++) {
        if (!res[i].accepted) continue;
        mpptr pMP = vpPoints[i];
        const int bestIdx = res[i].best_idx, bestDist = res[i].best_dist;
        const int existing_idx = pMP->GetIndexInKeyFrame(pKF);
        if (existing_idx != -1) {
            // the key frame already observes this point: move the observation unless the feature's own descriptor is
            // closer to the point than the match found (:414-433; dist_newplace is the distance to the same feature)
            bool bDoNotReplace = false;
            if (pKF->GetMapPoint(bestIdx)) {
                const int dist_newplace = DescriptorDistance(pMP->GetDescriptor(), pKF->mDescriptors.row(bestIdx));
                if (dist_newplace < bestDist) bDoNotReplace = true;
            }
            if (!bDoNotReplace) pKF->RemapMapPointMatch(pMP, existing_idx, bestIdx);
        } else {
            vpMatched[bestIdx] = pMP;
        }
    }
    return nmatches;
}

// ORBmatcher.cpp:178-306
int ORBmatcher::SearchByBoW(kfptr pKF, Frame &F, vector<mpptr> &vpMapPointMatches) {
    const vector<mpptr> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = vector<mpptr>(F.N, static_cast<mpptr>(NULL));
    const int nKF = (int)vpMapPointsKF.size(), nF = F.N;
    std::vector<uint8_t> valid(nKF > 0 ? nKF : 1, 0);
    std::vector<float> angKF(nKF > 0 ? nKF : 1), angF(nF > 0 ? nF : 1);
    for (int i = 0; i < nKF; i++) {
        valid[i] = (vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad()) ? 1 : 0;
        angKF[i] = pKF->mvKeysUn[i].angle;
    }
    for (int i = 0; i < nF; i++) angF[i] = F.mvKeys[i].angle;
    const std::vector<int> nodeKF = node_ids(pKF->mFeatVec, nKF), nodeF = node_ids(F.mFeatVec, nF);
    std::vector<int> match_f(nF > 0 ? nF : 1, -1);
    int nmatches = 0;
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    if (ccm_search_bow(D.m, CCM_BOW_KF_FRAME, mfNNratio, mbCheckOrientation ? 1 : 0, nKF, pKF->mDescriptors.ptr(0), nodeKF.data(),
                       valid.data(), angKF.data(), nF, F.mDescriptors.ptr(0), nodeF.data(), nullptr, angF.data(), nullptr,
                       match_f.data(), nullptr, &nmatches) != CCM_OK)
        throw estd::infrastructure_ex();
    for (int i = 0; i < nF; i++)
        if (match_f[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[match_f[i]];
    return nmatches;
}

// ORBmatcher.cpp:565-698
int ORBmatcher::SearchByBoW(kfptr pKF1, kfptr pKF2, vector<mpptr> &vpMatches12) {
    const vector<mpptr> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int n1 = (int)vpMapPoints1.size(), n2 = (int)vpMapPoints2.size();
    vpMatches12 = vector<mpptr>(n1, static_cast<mpptr>(NULL));
    std::vector<uint8_t> v1(n1 > 0 ? n1 : 1, 0), v2(n2 > 0 ? n2 : 1, 0);
    std::vector<float> a1(n1 > 0 ? n1 : 1), a2(n2 > 0 ? n2 : 1);
    for (int i = 0; i < n1; i++) { v1[i] = (vpMapPoints1[i] && !vpMapPoints1[i]->isBad()) ? 1 : 0; a1[i] = pKF1->mvKeysUn[i].angle; }
    for (int i = 0; i < n2; i++) { v2[i] = (vpMapPoints2[i] && !vpMapPoints2[i]->isBad()) ? 1 : 0; a2[i] = pKF2->mvKeysUn[i].angle; }
    const std::vector<int> node1 = node_ids(pKF1->mFeatVec, n1), node2 = node_ids(pKF2->mFeatVec, n2);
    std::vector<int> m12(n1 > 0 ? n1 : 1, -1);
    int nmatches = 0;
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    if (ccm_search_bow(D.m, CCM_BOW_KF_KF, mfNNratio, mbCheckOrientation ? 1 : 0, n1, pKF1->mDescriptors.ptr(0), node1.data(), v1.data(),
                       a1.data(), n2, pKF2->mDescriptors.ptr(0), node2.data(), v2.data(), a2.data(), m12.data(), nullptr, nullptr,
                       &nmatches) != CCM_OK)
        throw estd::infrastructure_ex();
    for (int i = 0; i < n1; i++)
        if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];
    return nmatches;
}

// ORBmatcher.cpp:159-176 (scalar helper; the device evaluates the same expression inside ccm_search_triangulation)
bool ORBmatcher::CheckDistEpipolarLine(const cv::KeyPoint &kp1, const cv::KeyPoint &kp2, const cv::Mat &F12, const kfptr pKF2) {
    const float a = kp1.pt.x * F12.at<float>(0, 0) + kp1.pt.y * F12.at<float>(1, 0) + F12.at<float>(2, 0);
    const float b = kp1.pt.x * F12.at<float>(0, 1) + kp1.pt.y * F12.at<float>(1, 1) + F12.at<float>(2, 1);
    const float c = kp1.pt.x * F12.at<float>(0, 2) + kp1.pt.y * F12.at<float>(1, 2) + F12.at<float>(2, 2);
    const float num = a * kp2.pt.x + b * kp2.pt.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * pKF2->mvLevelSigma2[kp2.octave];
}

// ORBmatcher.cpp:700-852
int ORBmatcher::SearchForTriangulation(kfptr pKF1, kfptr pKF2, cv::Mat F12, vector<pair<size_t, size_t> > &vMatchedPairs) {
    // epipole in the second image, :707-714
    cv::Mat Cw = pKF1->GetCameraCenter();
    cv::Mat R2w = pKF2->GetRotation();
    cv::Mat t2w = pKF2->GetTranslation();
    cv::Mat C2 = R2w * Cw + t2w;
    const float invz = 1.0f / C2.at<float>(2);
    ccm_epipolar_geom geo;
    geo.ex = pKF2->fx * C2.at<float>(0) * invz + pKF2->cx;
    geo.ey = pKF2->fy * C2.at<float>(1) * invz + pKF2->cy;
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) geo.f12[3 * r + c] = F12.at<float>(r, c);
    geo.scale_factors = pKF2->mvScaleFactors.data();
    geo.level_sigma2 = pKF2->mvLevelSigma2.data();
    geo.n_levels = (int)std::min(pKF2->mvScaleFactors.size(), pKF2->mvLevelSigma2.size());
    const int n1 = pKF1->N, n2 = pKF2->N;
    struct Side { std::vector<uint8_t> free_; std::vector<float> x, y, ang; std::vector<int> oct; };
    Side S[2];
    kfptr K[2] = {pKF1, pKF2};
    for (int s = 0; s < 2; s++) {
        const int n = K[s]->N, cap = n > 0 ? n : 1;
        S[s].free_.assign(cap, 0); S[s].x.resize(cap); S[s].y.resize(cap); S[s].ang.resize(cap); S[s].oct.assign(cap, 0);
        for (int i = 0; i < n; i++) {
            const cv::KeyPoint &kp = K[s]->mvKeysUn[i];
            S[s].free_[i] = K[s]->GetMapPoint(i) ? 0 : 1;             // :741-744, :758-762
            S[s].x[i] = kp.pt.x; S[s].y[i] = kp.pt.y; S[s].ang[i] = kp.angle; S[s].oct[i] = kp.octave;
        }
    }
    const std::vector<int> node1 = node_ids(pKF1->mFeatVec, n1), node2 = node_ids(pKF2->mFeatVec, n2);
    std::vector<int> m12(n1 > 0 ? n1 : 1, -1);
    int nmatches = 0;
    {
        Device& D = device();
        std::lock_guard<std::mutex> g(D.mu);
        if (ccm_search_triangulation(D.m, mbCheckOrientation ? 1 : 0, &geo, n1, pKF1->mDescriptors.ptr(0), node1.data(), S[0].free_.data(),
                                     S[0].x.data(), S[0].y.data(), S[0].ang.data(), n2, pKF2->mDescriptors.ptr(0), node2.data(),
                                     S[1].free_.data(), S[1].x.data(), S[1].y.data(), S[1].oct.data(), S[1].ang.data(), m12.data(),
                                     &nmatches) != CCM_OK)
            throw estd::infrastructure_ex();
    }
    vMatchedPairs.clear();
    vMatchedPairs.reserve(nmatches);
    for (int i = 0; i < n1; i++)
        if (m12[i] >= 0) vMatchedPairs.push_back(make_pair((size_t)i, (size_t)m12[i]));
    return nmatches;
}

// ORBmatcher.cpp:448-563
int ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12, int windowSize) {
    const int n1 = (int)F1.mvKeysUn.size(), n2 = (int)F2.mvKeysUn.size();
    vnMatches12 = vector<int>(n1, -1);
    std::vector<ccm_window_query> q(n1, no_query());
    std::vector<uint8_t> qd((size_t)n1 * 32 + 32);
    for (int i1 = 0; i1 < n1; i1++) {
        const cv::KeyPoint &kp1 = F1.mvKeysUn[i1];
        const int level1 = kp1.octave;
        std::memcpy(&qd[(size_t)32 * i1], F1.mDescriptors.ptr(i1), 32);
        if (level1 > 0) continue;
        q[i1].u = vbPrevMatched[i1].x; q[i1].v = vbPrevMatched[i1].y; q[i1].r = (float)windowSize;
        q[i1].angle = kp1.angle;
        q[i1].min_level = (int16_t)level1; q[i1].max_level = (int16_t)level1;
        q[i1].flags = CCM_Q_VALID;
    }
    std::vector<int> holder(n2 > 0 ? n2 : 1, -1);           // vnMatches21, output only
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    Pack pack(D.m, 0, F2.mId, n2, F2.mDescriptors, F2.mvKeysUn, Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv,
              Frame::mfGridElementHeightInv, FRAME_GRID_COLS, FRAME_GRID_ROWS);
    int nmatches = 0;
    run_window(D.m, pack.p, CCM_WIN_INIT, mfNNratio, TH_LOW, mbCheckOrientation, q, qd, holder.data(), nullptr, nmatches);
    for (int i2 = 0; i2 < n2; i2++)
        if (holder[i2] >= 0) vnMatches12[holder[i2]] = i2;
    for (int i1 = 0; i1 < n1; i1++)
        if (vnMatches12[i1] >= 0) vbPrevMatched[i1] = F2.mvKeysUn[vnMatches12[i1]].pt;   // :558-560
    return nmatches;
}

}  // namespace cslam

// ---------------------------------------------------------------------------------------------------------------
// Fuse and SearchBySim3: the candidate search of every map point is independent of the map mutations the reference
// interleaves with it (the distance loops never consult map points), so all searches of a call run as one
// CCM_WIN_BEST batch and the mutations are replayed afterwards in the reference's order, re-checking the conditions a
// mutation can change (isBad, IsInKeyFrame) at replay time.
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct KfPack {
    Pack pack;
    KfPack(ccm_matcher* m, const boost::shared_ptr<cslam::KeyFrame>& k)
        : pack(m, 1, k->mId, k->N, k->mDescriptors, k->mvKeysUn, (float)k->mnMinX, (float)k->mnMinY, k->mfGridElementWidthInv,
               k->mfGridElementHeightInv, k->mnGridCols, k->mnGridRows) {}
};
inline void set_query(ccm_window_query& q, float u, float v, float r, int lvl) {
    q.u = u; q.v = v; q.r = r; q.min_level = (int16_t)(lvl - 1); q.max_level = (int16_t)lvl; q.flags = CCM_Q_VALID;
}
}  // namespace

namespace cslam {

// ORBmatcher.cpp:854-993
int ORBmatcher::Fuse(kfptr pKF, const vector<mpptr> &vpMapPoints, const float th) {
    cv::Mat Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation();
    const float &fx = pKF->fx, &fy = pKF->fy, &cx = pKF->cx, &cy = pKF->cy;
    cv::Mat Ow = pKF->GetCameraCenter();
    const int nMPs = (int)vpMapPoints.size();
    std::vector<ccm_window_query> q(nMPs, no_query());
    std::vector<uint8_t> qd((size_t)nMPs * 32 + 32);
    for (int i = 0; i < nMPs; i++) {
        mpptr pMP = vpMapPoints[i];
        if (!pMP) continue;
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        if (pMP->mbDoNotReplace) continue;
        cv::Mat p3Dw = pMP->GetWorldPos();
        cv::Mat p3Dc = Rcw * p3Dw + tcw;
        if (p3Dc.at<float>(2) < 0.0f) continue;
        const float invz = 1 / p3Dc.at<float>(2);
        const float x = p3Dc.at<float>(0) * invz, y = p3Dc.at<float>(1) * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        cv::Mat PO = p3Dw - Ow;
        const float dist3D = cv::norm(PO);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        cv::Mat Pn = pMP->GetNormal();
        if (PO.dot(Pn) < 0.5 * dist3D) continue;
        const int lvl = pMP->PredictScale(dist3D, pKF);
        set_query(q[i], u, v, th * pKF->mvScaleFactors[lvl], lvl);
        std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
    }
    std::vector<ccm_match_result> res(nMPs > 0 ? nMPs : 1);
    {
        Device& D = device();
        std::lock_guard<std::mutex> g(D.mu);
        KfPack kp(D.m, pKF);
        int n_acc = 0;
        run_window(D.m, kp.pack.p, CCM_WIN_BEST, mfNNratio, TH_LOW, false, q, qd, nullptr, res.data(), n_acc, &pKF->mvInvLevelSigma2);
    }
    int nFused = 0;
    for (int i = 0; i < nMPs; i++) {
        if (!(q[i].flags & CCM_Q_VALID) || !res[i].accepted) continue;
        mpptr pMP = vpMapPoints[i];
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;      // an earlier replacement / observation of this call
        const int bestIdx = res[i].best_idx;
        mpptr pMPinKF = pKF->GetMapPoint(bestIdx);
        if (pMPinKF) {
            if (!pMPinKF->isBad() && !pMPinKF->mbDoNotReplace) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, bestIdx);
            pKF->AddMapPoint(pMP, bestIdx);
        }
        nFused++;
    }
    return nFused;
}

// ORBmatcher.cpp:995-1122
int ORBmatcher::Fuse(kfptr pKF, cv::Mat Scw, const vector<mpptr> &vpPoints, float th, vector<mpptr> &vpReplacePoint) {
    const float &fx = pKF->fx, &fy = pKF->fy, &cx = pKF->cx, &cy = pKF->cy;
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    const set<mpptr> spAlreadyFound = pKF->GetMapPoints();
    const int nPoints = (int)vpPoints.size();
    std::vector<ccm_window_query> q(nPoints, no_query());
    std::vector<uint8_t> qd((size_t)nPoints * 32 + 32);
    for (int i = 0; i < nPoints; i++) {
        mpptr pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        cv::Mat p3Dw = pMP->GetWorldPos();
        cv::Mat p3Dc = Rcw * p3Dw + tcw;
        if (p3Dc.at<float>(2) < 0.0f) continue;
        const float invz = 1.0 / p3Dc.at<float>(2);
        const float x = p3Dc.at<float>(0) * invz, y = p3Dc.at<float>(1) * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        cv::Mat PO = p3Dw - Ow;
        const float dist3D = cv::norm(PO);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        cv::Mat Pn = pMP->GetNormal();
        if (PO.dot(Pn) < 0.5 * dist3D) continue;
        const int lvl = pMP->PredictScale(dist3D, pKF);
        set_query(q[i], u, v, th * pKF->mvScaleFactors[lvl], lvl);
        std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
    }
    std::vector<ccm_match_result> res(nPoints > 0 ? nPoints : 1);
    {
        Device& D = device();
        std::lock_guard<std::mutex> g(D.mu);
        KfPack kp(D.m, pKF);
        int n_acc = 0;
        run_window(D.m, kp.pack.p, CCM_WIN_BEST, mfNNratio, TH_LOW, false, q, qd, nullptr, res.data(), n_acc);
    }
    int nFused = 0;
    for (int i = 0; i < nPoints; i++) {
        if (!(q[i].flags & CCM_Q_VALID) || !res[i].accepted) continue;
        mpptr pMP = vpPoints[i];
        const int bestIdx = res[i].best_idx;
        mpptr pMPinKF = pKF->GetMapPoint(bestIdx);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, bestIdx);
            pKF->AddMapPoint(pMP, bestIdx);
        }
        nFused++;
    }
    return nFused;
}

// ORBmatcher.cpp:1124-1348
int ORBmatcher::SearchBySim3(kfptr pKF1, kfptr pKF2, vector<mpptr> &vpMatches12, const float &s12, const cv::Mat &R12,
                             const cv::Mat &t12, const float th) {
    const float &fx = pKF1->fx, &fy = pKF1->fy, &cx = pKF1->cx, &cy = pKF1->cy;
    cv::Mat R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation();
    cv::Mat R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
    cv::Mat sR12 = s12 * R12;
    cv::Mat sR21 = (1.0 / s12) * R12.t();
    cv::Mat t21 = -sR21 * t12;
    const vector<mpptr> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
        mpptr pMP = vpMatches12[i];
        if (pMP) {
            vbAlreadyMatched1[i] = true;
            const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
    }
    // one projection pass per direction: points of `from` through (Rw, tw) then (sR, t) into `to`
    struct Dir { const vector<mpptr>* pts; const vector<bool>* done; cv::Mat Rw, tw, sR, t; kfptr to; int n; };
    Dir dirs[2] = {{&vpMapPoints1, &vbAlreadyMatched1, R1w, t1w, sR21, t21, pKF2, N1},
                   {&vpMapPoints2, &vbAlreadyMatched2, R2w, t2w, sR12, t12, pKF1, N2}};
    std::vector<int> vnMatch[2];
    for (int d = 0; d < 2; d++) {
        const Dir& X = dirs[d];
        std::vector<ccm_window_query> q(X.n, no_query());
        std::vector<uint8_t> qd((size_t)X.n * 32 + 32);
        for (int i = 0; i < X.n; i++) {
            mpptr pMP = (*X.pts)[i];
            if (!pMP || (*X.done)[i]) continue;
            if (pMP->isBad()) continue;
            cv::Mat p3Dw = pMP->GetWorldPos();
            cv::Mat pa = X.Rw * p3Dw + X.tw;
            cv::Mat pb = X.sR * pa + X.t;
            if (pb.at<float>(2) < 0.0) continue;
            const float invz = 1.0 / pb.at<float>(2);
            const float x = pb.at<float>(0) * invz, y = pb.at<float>(1) * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!X.to->IsInImage(u, v)) continue;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            const float dist3D = cv::norm(pb);
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            const int lvl = pMP->PredictScale(dist3D, X.to);
            set_query(q[i], u, v, th * X.to->mvScaleFactors[lvl], lvl);
            std::memcpy(&qd[(size_t)32 * i], pMP->GetDescriptor().ptr(0), 32);
        }
        std::vector<ccm_match_result> res(X.n > 0 ? X.n : 1);
        Device& D = device();
        std::lock_guard<std::mutex> g(D.mu);
        KfPack kp(D.m, X.to);
        int n_acc = 0;
        run_window(D.m, kp.pack.p, CCM_WIN_BEST, mfNNratio, TH_HIGH, false, q, qd, nullptr, res.data(), n_acc);
        vnMatch[d].assign(X.n, -1);
        for (int i = 0; i < X.n; i++)
            if ((q[i].flags & CCM_Q_VALID) && res[i].accepted) vnMatch[d][i] = res[i].best_idx;
    }
    int nFound = 0;
    for (int i1 = 0; i1 < N1; i1++) {
        const int idx2 = vnMatch[0][i1];
        if (idx2 >= 0 && vnMatch[1][idx2] == i1) { vpMatches12[i1] = vpMapPoints2[idx2]; nFound++; }
    }
    return nFound;
}

}  // namespace cslam

// Drops every cached frame pack (call it on a map reset, or from ~Frame / ~KeyFrame if device memory is tight), and
// reports the cache's hit / miss counters since the last clear.
extern "C" void ccm_dropin_pack_cache_clear(unsigned long long* hits, unsigned long long* misses) {
    Device& D = device();
    std::lock_guard<std::mutex> g(D.mu);
    PackCache& C = pack_cache();
    if (hits) *hits = C.hits;
    if (misses) *misses = C.misses;
    for (std::map<PackKey, PackEntry>::iterator it = C.m.begin(); it != C.m.end(); ++it) ccm_frame_pack_destroy(it->second.p);
    C.m.clear();
    C.hits = C.misses = 0;
}
