/* TEST INFRASTRUCTURE — declarations the line ranges cut out of the reference need to compile on their own
 * (oracle/Makefile target _ref/libref_pin.so). The Makefile copies, at build time and never into the repository,
 *   cslam/src/MapPoint.cpp:929-994      MapPoint::ComputeDistinctiveDescriptors
 *   cslam/src/ORBmatcher.cpp:1653-1669  ORBmatcher::DescriptorDistance
 *   cslam/src/Converter.cc:166-193      Converter::toCvKeyPointMsg / fromCvKeyPointMsg
 *   cslam/src/KeyFrame.cpp:1462-1472    the mvKeysUn / mDescriptors part of KeyFrame::ConvertToMessage
 *   cslam/src/KeyFrame.cpp:1688-1704    the same part of KeyFrame::WriteMembersFromMessage
 * into oracle/_ref/gen_pin.cpp, each inside namespace cslam, the two KeyFrame fragments inside a member function of the
 * mock below. The classes here hold only the members those lines name; ccmslam_msgs::CvKeyPoint / Descriptor carry the
 * fields of cslam_msgs/msg/CvKeyPoint.msg and Descriptor.msg in declaration order with the C++ types ROS 1 genmsg maps
 * them to (float32 -> float, uint8 -> uint8_t, int8 -> int8_t, uint8[32] -> boost::array<uint8_t, 32>). */
#pragma once
#include <sys/types.h>
#include <algorithm>
#include <array>
#include <climits>
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <vector>
#include <opencv2/opencv.hpp>      /* oracle/matchshim: the mini cv::Mat / cv::KeyPoint */

using namespace std;

namespace ccmslam_msgs {
struct CvKeyPoint { float fPoint2f_x, fPoint2f_y; uint8_t size; float angle; uint8_t response; int8_t octave; };
struct Descriptor { std::array<uint8_t, 32> mDescriptor; };
struct KF { std::vector<CvKeyPoint> mvKeysUn; std::vector<Descriptor> mDescriptors; };
}

namespace cslam {
class ORBmatcher {
public:
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
};
class Converter {
public:
    static ccmslam_msgs::CvKeyPoint toCvKeyPointMsg(const cv::KeyPoint& kp);
    static cv::KeyPoint fromCvKeyPointMsg(const ccmslam_msgs::CvKeyPoint& Msg);
};
class KeyFrame {
public:
    bool bad = false;
    bool isBad() { return bad; }
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    void PinConvertToMessage(ccmslam_msgs::KF& Msg);                  /* body = KeyFrame.cpp:1462-1472 */
    void PinWriteMembersFromMessage(const ccmslam_msgs::KF* pMsg);    /* body = KeyFrame.cpp:1688-1704 */
};
class MapPoint {
public:
    typedef std::shared_ptr<KeyFrame> kfptr;
    std::mutex mMutexFeatures;
    bool mbBad = false;
    std::map<kfptr, size_t> mObservations;
    cv::Mat mDescriptor;
    void ComputeDistinctiveDescriptors();                             /* body = MapPoint.cpp:929-994 */
};
}
