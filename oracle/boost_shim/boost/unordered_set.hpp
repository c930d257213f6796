// Stand-in for <boost/unordered_set.hpp> (src/CLEVER.cpp:22, src/BronKerbosch.cpp:2).
#pragma once
#include <unordered_set>
namespace boost {
template <class K, class H = std::hash<K>, class E = std::equal_to<K>>
using unordered_set = std::unordered_set<K, H, E>;
}
