/* TEST INFRASTRUCTURE — mock of cslam::KeyFrame with the members cslam/src/ORBmatcher.cpp reads (KeyFrame.h:137-342). */
#pragma once
#include <set>
#include <vector>
#include <opencv2/opencv.hpp>
#include <boost/shared_ptr.hpp>
#include <thirdparty/DBoW2/DBoW2/FeatureVector.h>
#include <cslam/MapPoint.h>
#include <cslam/Frame.h>

namespace cslam {
class KeyFrame {
public:
    typedef boost::shared_ptr<MapPoint> mpptr;
    void* grid = nullptr;
    std::pair<size_t, size_t> mId = {0, 0};  /* idpair mId, KeyFrame.h:282 */
    int N = 0;
    float fx = 1, fy = 1, cx = 0, cy = 0;
    float minX = 0, maxX = 0, minY = 0, maxY = 0;       /* IsInImage bounds of the mock */
    int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;   /* KeyFrame.h:345-348 (int upstream) */
    int mnGridCols = 75, mnGridRows = 48;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;   /* KeyFrame.h:289-290 */
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    float mfLogScaleFactor = 0;
    std::vector<mpptr> mvpMapPoints;
    cv::Mat R, t, C;
    /* recorded mutations */
    std::vector<int> added;                  /* AddMapPoint: map point id per feature, -1 none */
    std::vector<std::pair<int, int>> remapped;

    cv::Mat GetCameraCenter() { return C.clone(); }
    cv::Mat GetRotation() { return R.clone(); }
    cv::Mat GetTranslation() { return t.clone(); }
    void AddMapPoint(mpptr pMP, const size_t& idx, bool = false) {      /* KeyFrame.cpp: mvpMapPoints[idx] = pMP */
        if (added.size() < (size_t)N) added.assign(N, -1);
        added[idx] = pMP->id; mvpMapPoints[idx] = pMP;
    }
    std::set<mpptr> GetMapPoints() { std::set<mpptr> s; for (auto& p : mvpMapPoints) if (p) s.insert(p); return s; }
    std::vector<mpptr> GetMapPointMatches() { return mvpMapPoints; }
    mpptr GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    void RemapMapPointMatch(mpptr, const size_t& a, const size_t& b) { remapped.push_back({(int)a, (int)b}); }
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const {
        std::vector<int> tmp(N > 0 ? N : 1);
        const int k = mo_features_in_area(grid, x, y, r, -1, -1, 1, tmp.data(), (int)tmp.size());
        return std::vector<size_t>(tmp.begin(), tmp.begin() + k);
    }
    bool IsInImage(const float& x, const float& y) const { return (x >= minX && x < maxX && y >= minY && y < maxY); }
};
inline void MapPoint::Replace(mpptr pMP, bool) {
    if (pMP->id == id) return;
    replacedBy = pMP->id; bad = true;
    std::map<KeyFrame*, int> obs; obs.swap(indexInKF);
    for (auto& o : obs) {
        if (pMP->indexInKF.count(o.first)) continue;
        if (o.second >= 0 && o.second < (int)o.first->mvpMapPoints.size()) o.first->mvpMapPoints[o.second] = pMP;
        pMP->indexInKF[o.first] = o.second; pMP->nObs++;
    }
}
}
