/* TEST INFRASTRUCTURE — empty stand-in for <cslam/estd.h> (the real one needs ROS + PCL);
 * ORBextractor.cpp uses nothing from it. */
#ifndef CCM_ORACLE_CVSHIM_CSLAM_ESTD_H_
#define CCM_ORACLE_CVSHIM_CSLAM_ESTD_H_
#endif
