/* TEST INFRASTRUCTURE — stand-in for <cslam/estd.h> (the real one needs ROS + PCL). ORBextractor.cpp uses nothing from
 * it; the product's drop-in class throws estd::infrastructure_ex (estd.h:66-74) on a CUDA failure. */
#ifndef CCM_ORACLE_CVSHIM_CSLAM_ESTD_H_
#define CCM_ORACLE_CVSHIM_CSLAM_ESTD_H_
#include <exception>
namespace estd {
class infrastructure_ex : public std::exception {
public:
    virtual const char* what() const throw() { return "EXCEPTION: Bad Infrastructure"; }
};
}
#endif
