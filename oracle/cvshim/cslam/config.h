/* TEST INFRASTRUCTURE — empty stand-in for <cslam/config.h> (the real one needs ROS + cv::FileStorage);
 * ORBextractor.cpp uses nothing from it. */
#ifndef CCM_ORACLE_CVSHIM_CSLAM_CONFIG_H_
#define CCM_ORACLE_CVSHIM_CSLAM_CONFIG_H_
#endif
