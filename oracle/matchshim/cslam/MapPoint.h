/* TEST INFRASTRUCTURE — mock of cslam::MapPoint with the members cslam/src/ORBmatcher.cpp reads (MapPoint.h:134-228).
 * Plain data set by the test glue; AddObservation and Replace record their argument and leave the state the real
 * class leaves that ORBmatcher.cpp can observe afterwards (MapPoint.cpp:417 AddObservation makes IsInKeyFrame true and
 * counts the observation; MapPoint.cpp:583-612 Replace marks the replaced point bad and moves its observations, and
 * the keyframes' slots, to the survivor — the common branch where the survivor does not observe that keyframe yet). */
#pragma once
#include <map>
#include <memory>
#include <opencv2/opencv.hpp>
#include <boost/shared_ptr.hpp>
namespace cslam {
class KeyFrame;
class Frame;
class MapPoint {
public:
    typedef boost::shared_ptr<KeyFrame> kfptr;
    typedef boost::shared_ptr<Frame> frameptr;
    typedef boost::shared_ptr<MapPoint> mpptr;
    int id = -1;
    bool bad = false;
    int nObs = 1;
    int predictedLevel = 0;
    float minDist = 0.f, maxDist = 1e30f;
    cv::Mat worldPos, normal, descriptor;
    std::map<KeyFrame*, int> indexInKF;       /* GetIndexInKeyFrame */
    /* tracking fields, MapPoint.h:224-228 */
    float mTrackProjX = 0, mTrackProjY = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;
    bool mbDoNotReplace = false;
    /* recorded mutations */
    int replacedBy = -1, addedObsIdx = -1;

    cv::Mat GetWorldPos() { return worldPos.clone(); }
    cv::Mat GetNormal() { return normal.clone(); }
    cv::Mat GetDescriptor() { return descriptor.clone(); }
    int Observations() { return nObs; }
    bool isBad() { return bad; }
    float GetMinDistanceInvariance() { return minDist; }
    float GetMaxDistanceInvariance() { return maxDist; }
    int PredictScale(const float&, kfptr) { return predictedLevel; }
    int PredictScale(const float&, frameptr) { return predictedLevel; }
    int GetIndexInKeyFrame(kfptr pKF, bool = false) { auto it = indexInKF.find(pKF.get()); return it == indexInKF.end() ? -1 : it->second; }
    bool IsInKeyFrame(kfptr pKF) { return indexInKF.count(pKF.get()) != 0; }
    void AddObservation(kfptr pKF, size_t idx, bool = false) { addedObsIdx = (int)idx; indexInKF[pKF.get()] = (int)idx; nObs++; }
    void Replace(mpptr pMP, bool = false);     /* defined in the mock KeyFrame.h (needs the complete KeyFrame) */
};
}
