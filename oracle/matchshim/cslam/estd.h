/* TEST INFRASTRUCTURE — the two names ORBmatcher.cpp uses from cslam/estd.h (namespace estd, infrastructure_ex). */
#pragma once
#include <exception>
#include <utility>
#include <cstddef>
namespace estd {
typedef std::pair<size_t, size_t> idpair;
class infrastructure_ex : public std::exception {
public:
    virtual const char* what() const throw() { return "EXCEPTION: Bad Infrastructure"; }
};
}
