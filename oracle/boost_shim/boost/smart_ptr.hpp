// Minimal stand-in for <boost/smart_ptr.hpp>: test infrastructure only (see oracle/README.md).
// The reference uses boost::shared_ptr<std::string> in src/ShortDnaSequence.h:29.
#pragma once
#include <memory>
namespace boost {
template <class T> using shared_ptr = std::shared_ptr<T>;
}
