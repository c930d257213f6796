// Stand-in for <boost/lexical_cast.hpp> (src/haploclique.cpp:135-137).
#pragma once
#include <sstream>
#include <stdexcept>
#include <string>
namespace boost {
class bad_lexical_cast : public std::bad_cast {
public:
    const char* what() const noexcept override { return "bad lexical cast"; }
};
template <class T, class S> T lexical_cast(const S& s) {
    std::stringstream ss;
    T out;
    if (!(ss << s) || !(ss >> out) || !(ss >> std::ws).eof()) throw bad_lexical_cast();
    return out;
}
}
