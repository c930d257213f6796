// Stand-in for <boost/dynamic_bitset.hpp> (src/Types.h:22-25): only the members the
// reference's clique finders call. Written from the documented semantics, 64-bit blocks.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
namespace boost {
template <class Block = unsigned long, class Alloc = void>
class dynamic_bitset {
public:
    typedef std::size_t size_type;
    typedef Block block_type;
    static const size_type bits_per_block = sizeof(Block) * 8;
    static const size_type npos = static_cast<size_type>(-1);

    dynamic_bitset() : nbits_(0) {}
    explicit dynamic_bitset(size_type n, unsigned long value = 0) : nbits_(n), w_((n + bits_per_block - 1) / bits_per_block, 0) {
        if (!w_.empty()) { w_[0] = static_cast<Block>(value); trim(); }
    }
    size_type size() const { return nbits_; }
    bool empty() const { return nbits_ == 0; }
    void resize(size_type n, bool v = false) {
        size_type old = nbits_;
        w_.resize((n + bits_per_block - 1) / bits_per_block, v ? ~Block(0) : Block(0));
        nbits_ = n;
        if (v) for (size_type i = old; i < n && i < ((old + bits_per_block - 1) / bits_per_block) * bits_per_block; ++i) set(i);
        trim();
    }
    void clear() { nbits_ = 0; w_.clear(); }
    dynamic_bitset& set(size_type i, bool v = true) {
        if (v) w_[i / bits_per_block] |= (Block(1) << (i % bits_per_block)); else reset(i);
        return *this;
    }
    dynamic_bitset& set() { for (auto& b : w_) b = ~Block(0); trim(); return *this; }
    dynamic_bitset& reset(size_type i) { w_[i / bits_per_block] &= ~(Block(1) << (i % bits_per_block)); return *this; }
    dynamic_bitset& reset() { for (auto& b : w_) b = 0; return *this; }
    dynamic_bitset& flip(size_type i) { w_[i / bits_per_block] ^= (Block(1) << (i % bits_per_block)); return *this; }
    dynamic_bitset& flip() { for (auto& b : w_) b = ~b; trim(); return *this; }
    bool test(size_type i) const { return (w_[i / bits_per_block] >> (i % bits_per_block)) & 1; }
    bool operator[](size_type i) const { return test(i); }
    bool any() const { for (auto b : w_) if (b) return true; return false; }
    bool none() const { return !any(); }
    size_type count() const { size_type c = 0; for (auto b : w_) c += __builtin_popcountll((unsigned long long)b); return c; }
    size_type find_first() const { return scan(0); }
    size_type find_next(size_type pos) const { return (pos == npos || pos + 1 >= nbits_) ? npos : scan(pos + 1); }
    bool is_subset_of(const dynamic_bitset& o) const {
        for (size_type i = 0; i < w_.size(); ++i) if (w_[i] & ~o.w_[i]) return false;
        return true;
    }
    bool intersects(const dynamic_bitset& o) const {
        size_type n = w_.size() < o.w_.size() ? w_.size() : o.w_.size();
        for (size_type i = 0; i < n; ++i) if (w_[i] & o.w_[i]) return true;
        return false;
    }
    dynamic_bitset& operator&=(const dynamic_bitset& o) { for (size_type i = 0; i < w_.size(); ++i) w_[i] &= o.w_[i]; return *this; }
    dynamic_bitset& operator|=(const dynamic_bitset& o) { for (size_type i = 0; i < w_.size(); ++i) w_[i] |= o.w_[i]; return *this; }
    dynamic_bitset& operator^=(const dynamic_bitset& o) { for (size_type i = 0; i < w_.size(); ++i) w_[i] ^= o.w_[i]; return *this; }
    dynamic_bitset& operator-=(const dynamic_bitset& o) { for (size_type i = 0; i < w_.size(); ++i) w_[i] &= ~o.w_[i]; return *this; }
    dynamic_bitset operator~() const { dynamic_bitset r(*this); r.flip(); return r; }
    friend dynamic_bitset operator&(const dynamic_bitset& a, const dynamic_bitset& b) { dynamic_bitset r(a); r &= b; return r; }
    friend dynamic_bitset operator|(const dynamic_bitset& a, const dynamic_bitset& b) { dynamic_bitset r(a); r |= b; return r; }
    friend dynamic_bitset operator^(const dynamic_bitset& a, const dynamic_bitset& b) { dynamic_bitset r(a); r ^= b; return r; }
    friend dynamic_bitset operator-(const dynamic_bitset& a, const dynamic_bitset& b) { dynamic_bitset r(a); r -= b; return r; }
    friend bool operator==(const dynamic_bitset& a, const dynamic_bitset& b) { return a.nbits_ == b.nbits_ && a.w_ == b.w_; }
    friend bool operator!=(const dynamic_bitset& a, const dynamic_bitset& b) { return !(a == b); }
    friend bool operator<(const dynamic_bitset& a, const dynamic_bitset& b) {
        for (size_type i = a.w_.size(); i-- > 0;) { if (a.w_[i] != b.w_[i]) return a.w_[i] < b.w_[i]; }
        return false;
    }
private:
    void trim() {
        size_type extra = nbits_ % bits_per_block;
        if (extra && !w_.empty()) w_.back() &= (Block(1) << extra) - 1;
    }
    size_type scan(size_type from) const {
        if (from >= nbits_) return npos;
        size_type bi = from / bits_per_block;
        Block cur = w_[bi] & (~Block(0) << (from % bits_per_block));
        for (;;) {
            if (cur) return bi * bits_per_block + __builtin_ctzll((unsigned long long)cur);
            if (++bi >= w_.size()) return npos;
            cur = w_[bi];
        }
    }
    size_type nbits_;
    std::vector<Block> w_;
};
template <class B, class A> const typename dynamic_bitset<B, A>::size_type dynamic_bitset<B, A>::bits_per_block;
template <class B, class A> const typename dynamic_bitset<B, A>::size_type dynamic_bitset<B, A>::npos;
}
