// Library-private declarations shared by the host packer (hc_host.cpp) and the CUDA side (hc_gpu.cu).
#pragma once
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/hc_edge.h"

namespace hc {

// Error text of the last failed call that had no context to hold it.
void set_global_error(const std::string& msg);
const char* global_error();

// The two 127x127 match-probability tables, built with the reference's exact mixed float/double
// expressions (src/NewEdgeCalculator.cpp:41-74). Row-major [q1][q2]; rows/cols < 33 are zero.
void build_match_tables(double* eq, double* uneq);

// prob0 after n tail columns when no allele-frequency entry applies: n sequential additions of
// log10(0.25) starting from 0.0 (src/NewEdgeCalculator.cpp:133-143), so F[n] is bit-identical to the
// reference's running sum. Filled for 0..n_max.
void build_prob0_ladder(std::vector<double>& F, int n_max);

// Phred update tables of the merges with the reference's arithmetic (src/AlignmentRecord.cpp:40-54, 68-86): error
// probabilities through a double pow rounded to float, the posterior in float, log10 and round in float; stored as the final
// quality char min(v + 33, 126) (:173, :177, :180). tab[0..94*94) = bases agree, tab[94*94..) = bases differ (first index: the
// larger quality). Indices are quality char - 33.
void build_phred_tables(uint8_t* tab2x94x94);

// Host buffer that is page-locked when a CUDA device is available (hc_upload then runs at full PCIe rate)
// and plain malloc otherwise (packing itself never needs a GPU).
struct HostBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool pinned = false;
    bool alloc(size_t bytes);
    void release();
    ~HostBuf() { release(); }
};

// f(begin, end) over [0, n) on up to 16 host threads (per-batch host work: packing, band layout, work items).
template <class F>
void parallel_ranges(int n, int min_per_thread, F f) {
    int T = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
    T = std::max(1, std::min(T, n / std::max(1, min_per_thread)));
    if (T <= 1) { f(0, n); return; }
    std::vector<std::thread> pool;
    for (int t = 1; t < T; ++t) pool.emplace_back([=, &f] { f((int)((long long)n * t / T), (int)((long long)n * (t + 1) / T)); });
    f(0, (int)((long long)n / T));
    for (auto& th : pool) th.join();
}

struct PackedOwner {
    HostBuf hdr, ops;
};

}  // namespace hc
