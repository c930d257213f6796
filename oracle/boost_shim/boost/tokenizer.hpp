// Stand-in for <boost/tokenizer.hpp>: char_separator tokenizer with dropped empty tokens
// (src/haploclique.cpp:122-130 is the only live use).
#pragma once
#include <string>
#include <vector>
namespace boost {
template <class Ch> class char_separator {
public:
    explicit char_separator(const Ch* dropped) : d_(dropped) {}
    bool is_sep(Ch c) const { return d_.find(c) != std::basic_string<Ch>::npos; }
private:
    std::basic_string<Ch> d_;
};
template <class Sep> class tokenizer {
public:
    typedef std::vector<std::string>::const_iterator iterator;
    tokenizer(const std::string& s, const Sep& sep) {
        std::string cur;
        for (char c : s) {
            if (sep.is_sep(c)) { if (!cur.empty()) { t_.push_back(cur); cur.clear(); } }
            else cur.push_back(c);
        }
        if (!cur.empty()) t_.push_back(cur);
    }
    iterator begin() const { return t_.begin(); }
    iterator end() const { return t_.end(); }
private:
    std::vector<std::string> t_;
};
}
