// Stand-in for <boost/math/distributions/normal.hpp>. Only src/GaussianEdgeCalculator.cpp:35-40
// (out of scope, never constructed without -I) calls into it; NewEdgeCalculator.cpp includes it unused.
#pragma once
#include <cmath>
namespace boost { namespace math {
class normal {
public:
    normal(double mean = 0.0, double sd = 1.0) : m_(mean), s_(sd) {}
    double mean() const { return m_; }
    double standard_deviation() const { return s_; }
private:
    double m_, s_;
};
template <class D, class R> struct complemented2 { const D& dist; const R& param; };
template <class D, class R> complemented2<D, R> complement(const D& d, const R& r) { return complemented2<D, R>{d, r}; }
inline double cdf(const normal& n, double x) { return 0.5 * std::erfc(-(x - n.mean()) / (n.standard_deviation() * std::sqrt(2.0))); }
template <class R> double cdf(const complemented2<normal, R>& c) {
    return 0.5 * std::erfc((c.param - c.dist.mean()) / (c.dist.standard_deviation() * std::sqrt(2.0)));
}
inline double quantile(const normal& n, double p) {
    double lo = -40.0, hi = 40.0;
    for (int i = 0; i < 200; ++i) { double mid = 0.5 * (lo + hi); if (cdf(normal(), mid) < p) lo = mid; else hi = mid; }
    return n.mean() + n.standard_deviation() * 0.5 * (lo + hi);
}
}}
