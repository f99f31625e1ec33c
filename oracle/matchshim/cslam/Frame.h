/* TEST INFRASTRUCTURE — mock of cslam::Frame with the members cslam/src/ORBmatcher.cpp reads (Frame.h:102-171).
 * GetFeaturesInArea delegates to the restated grid lookup of oracle/match_oracle.cpp (Frame.cpp:200-253), which is
 * checked against a numpy brute force in tests/test_match_oracle.py. */
#pragma once
#include <memory>
#include <vector>
#include <opencv2/opencv.hpp>
#include <boost/shared_ptr.hpp>
#include <boost/enable_shared_from_this.hpp>
#include <thirdparty/DBoW2/DBoW2/FeatureVector.h>
#include <cslam/MapPoint.h>

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 75

extern "C" int mo_features_in_area(const void* fp, float x, float y, float r, int minLevel, int maxLevel, int kf_variant, int* out, int cap);

namespace cslam {
class KeyFrame;
class Frame : public boost::enable_shared_from_this<Frame> {
public:
    typedef boost::shared_ptr<Frame> frameptr;
    typedef boost::shared_ptr<MapPoint> mpptr;
    typedef boost::shared_ptr<KeyFrame> kfptr;
    void* grid = nullptr;                    /* MoFrame of oracle/match_oracle.cpp */
    std::pair<size_t, size_t> mId = {0, 0};  /* idpair mId, Frame.h:156 */
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1) const {
        std::vector<int> tmp(N > 0 ? N : 1);
        const int k = mo_features_in_area(grid, x, y, r, minLevel, maxLevel, 0, tmp.data(), (int)tmp.size());
        return std::vector<size_t>(tmp.begin(), tmp.begin() + k);
    }
    static float fx, fy, cx, cy;
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    DBoW2::FeatureVector mFeatVec;
    cv::Mat mDescriptors;
    std::vector<mpptr> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    cv::Mat mTcw;
    std::vector<float> mvScaleFactors;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static float mfGridElementWidthInv, mfGridElementHeightInv;   /* Frame.h:147-148 */
};
}
