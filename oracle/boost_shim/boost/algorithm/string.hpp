// Stand-in for <boost/algorithm/string.hpp>: split / is_any_of / token_compress_on / trim_right
// (src/haploclique.cpp:267-271).
#pragma once
#include <cctype>
#include <string>
#include <vector>
namespace boost {
enum token_compress_mode_type { token_compress_on, token_compress_off };
struct is_any_of_pred {
    std::string set;
    bool operator()(char c) const { return set.find(c) != std::string::npos; }
};
inline is_any_of_pred is_any_of(const std::string& s) { return is_any_of_pred{s}; }
template <class Seq, class Pred>
Seq& split(Seq& result, const std::string& input, Pred pred, token_compress_mode_type mode = token_compress_off) {
    result.clear();
    std::string cur;
    bool last_was_sep = false;
    for (char c : input) {
        if (pred(c)) {
            if (!(mode == token_compress_on && last_was_sep)) { result.push_back(cur); cur.clear(); }
            last_was_sep = true;
        } else { cur.push_back(c); last_was_sep = false; }
    }
    result.push_back(cur);
    return result;
}
inline void trim_right(std::string& s) {
    while (!s.empty() && std::isspace(static_cast<unsigned char>(s.back()))) s.pop_back();
}
inline void trim_left(std::string& s) {
    std::size_t i = 0;
    while (i < s.size() && std::isspace(static_cast<unsigned char>(s[i]))) ++i;
    s.erase(0, i);
}
inline void trim(std::string& s) { trim_right(s); trim_left(s); }
}
