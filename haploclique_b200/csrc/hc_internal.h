// Library-private declarations shared by the host packer (hc_host.cpp) and the CUDA side (hc_gpu.cu).
#pragma once
#include <string>
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

// Host buffer that is page-locked when a CUDA device is available (hc_upload then runs at full PCIe rate)
// and plain malloc otherwise (packing itself never needs a GPU).
struct HostBuf {
    void* p = nullptr;
    bool pinned = false;
    bool alloc(size_t bytes);
    void release();
    ~HostBuf() { release(); }
};

struct PackedOwner {
    HostBuf hdr, ops;
};

}  // namespace hc
