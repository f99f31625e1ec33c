/* TEST INFRASTRUCTURE — stand-in for <boost/shared_ptr.hpp> (Boost is not installed here). */
#ifndef CCM_ORACLE_CVSHIM_BOOST_SHARED_PTR_HPP_
#define CCM_ORACLE_CVSHIM_BOOST_SHARED_PTR_HPP_
#include <memory>
namespace boost { template <typename T> using shared_ptr = std::shared_ptr<T>; }
#endif
