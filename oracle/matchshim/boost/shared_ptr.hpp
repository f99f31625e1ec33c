/* TEST INFRASTRUCTURE — boost::shared_ptr as std::shared_ptr for the oracle builds of the reference sources. */
#pragma once
#include <memory>
namespace boost {
template <typename T> using shared_ptr = std::shared_ptr<T>;
template <typename T> using enable_shared_from_this = std::enable_shared_from_this<T>;
using std::static_pointer_cast;
using std::dynamic_pointer_cast;
}
