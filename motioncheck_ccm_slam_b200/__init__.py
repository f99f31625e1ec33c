"""B200-native (sm_100a CUDA) ORB front-end for CCM-SLAM: drop-in for cslam::ORBextractor / cslam::ORBmatcher's
Hamming searches behind a C ABI (include/ccm_orb.h). See DESIGN.md."""
from .extractor import ORBextractor  # noqa: F401
