// GpuEdgeCalculator.cpp -- see GpuEdgeCalculator.h. Host C++ only; all scoring is behind the C ABI (hc_edge.h).
#include "GpuEdgeCalculator.h"

#include <cassert>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

struct GpuEdgeCalculator::Flat {
    std::vector<uint8_t> single_end, is_clique, cig, seq, qual;
    std::vector<int32_t> ncov, cov_first, cov_last, seg_start, seg_end, cig_len, seq_len;
    std::vector<int64_t> cig_off, seq_off;
    hc_reads view() const {
        hc_reads r;
        r.n = (int32_t)single_end.size();
        r.single_end = single_end.data(); r.is_clique = is_clique.data();
        r.ncov = ncov.data(); r.cov_first = cov_first.data(); r.cov_last = cov_last.data();
        r.seg_start = seg_start.data(); r.seg_end = seg_end.data();
        r.cig_off = cig_off.data(); r.cig_len = cig_len.data(); r.seq_off = seq_off.data(); r.seq_len = seq_len.data();
        r.cig = cig.data(); r.seq = seq.data(); r.qual = qual.data();
        return r;
    }
};

// AlignmentRecord getters (src/AlignmentRecord.cpp:1703-1811) -> the SoA the packer reads. Only what
// NewEdgeCalculator::edgeBetween touches: single_end, the "Clique" name test (src/NewEdgeCalculator.cpp:1111),
// cov_pos size / first / last (:1127, :1252), start/end, unrolled CIGARs, bases, qualities.
void GpuEdgeCalculator::flatten(const std::vector<const AlignmentRecord*>& reads, Flat& f) {
    const size_t n = reads.size();
    f.single_end.resize(n); f.is_clique.resize(n); f.ncov.resize(n); f.cov_first.resize(n); f.cov_last.resize(n);
    f.seg_start.assign(2 * n, 0); f.seg_end.assign(2 * n, 0); f.cig_len.assign(2 * n, 0); f.seq_len.assign(2 * n, 0);
    f.cig_off.assign(2 * n, 0); f.seq_off.assign(2 * n, 0);
    f.cig.clear(); f.seq.clear(); f.qual.clear();
    for (size_t r = 0; r < n; ++r) {
        const AlignmentRecord& ar = *reads[r];
        const bool single = ar.isSingleEnd();
        f.single_end[r] = single ? 1 : 0;
        f.is_clique[r] = ar.getName().find("Clique") != std::string::npos ? 1 : 0;
        const std::vector<AlignmentRecord::mapValue>& cov = ar.getCovmap();
        if (cov.empty()) throw std::runtime_error("GpuEdgeCalculator: read " + ar.getName() + " covers no reference position");
        f.ncov[r] = (int32_t)cov.size();
        f.cov_first[r] = cov.front().ref;
        f.cov_last[r] = cov.back().ref;
        for (int s = 0; s < (single ? 1 : 2); ++s) {
            const std::vector<char>& cg = s ? ar.getCigar2Unrolled() : ar.getCigar1Unrolled();
            const ShortDnaSequence& sq = s ? ar.getSequence2() : ar.getSequence1();
            const std::string& bases = sq.toString();
            const std::string& quals = sq.qualityString();
            const size_t k = 2 * r + s;
            f.seg_start[k] = s ? ar.getStart2() : ar.getStart1();
            f.seg_end[k] = s ? ar.getEnd2() : ar.getEnd1();
            f.cig_off[k] = (int64_t)f.cig.size();
            f.cig_len[k] = (int32_t)cg.size();
            f.seq_off[k] = (int64_t)f.seq.size();
            f.seq_len[k] = (int32_t)bases.size();
            f.cig.insert(f.cig.end(), cg.begin(), cg.end());
            f.seq.insert(f.seq.end(), bases.begin(), bases.end());
            f.qual.insert(f.qual.end(), quals.begin(), quals.end());
        }
        if (single) { f.cig_off[2 * r + 1] = (int64_t)f.cig.size(); f.seq_off[2 * r + 1] = (int64_t)f.seq.size(); }
    }
}

void GpuEdgeCalculator::fail(const char* what, int rc) const {
    throw std::runtime_error(std::string("GpuEdgeCalculator: ") + what + " failed (" + std::to_string(rc) + "): " +
                             hc_multi_last_error(multi_));
}

void GpuEdgeCalculator::init(double edge_quasi_cutoff, double overlap, std::unordered_map<int, double>& simpson_map,
                             double edge_quasi_cutoff_single, double overlap_single, double edge_quasi_cutoff_mixed,
                             unsigned int maxPosition, bool noProb0, const std::vector<int>& devices) {
    std::memset(&par_, 0, sizeof par_);
    std::memset(&stats_, 0, sizeof stats_);
    par_.edge_quasi_cutoff_cliques = edge_quasi_cutoff;
    par_.edge_quasi_cutoff_mixed = edge_quasi_cutoff_mixed;
    par_.edge_quasi_cutoff_single = edge_quasi_cutoff_single;
    par_.min_overlap_cliques = overlap;
    par_.min_overlap_single = overlap_single;
    par_.no_prob0 = noProb0 ? 1 : 0;
    devices_ = devices.empty() ? std::vector<int>(1, 0) : devices;
    if (std::getenv("HC_GPU_NO_ROWS")) row_feeding_ = false;
    par_.device = devices_[0];
    if (!simpson_map.empty()) {                            // SIMPSON_MAP.resize(maxPosition+1) (src/NewEdgeCalculator.cpp:85-88)
        simpson_.assign((size_t)maxPosition + 1, 0.0);
        for (const auto& kv : simpson_map)
            if (kv.first >= 0 && (size_t)kv.first < simpson_.size()) simpson_[kv.first] = kv.second;
        par_.simpson = simpson_.data();
        par_.simpson_len = (int32_t)simpson_.size();
    }
    std::vector<int32_t> dv(devices_.begin(), devices_.end());
    const int rc = hc_multi_create(&par_, dv.data(), (int32_t)dv.size(), &multi_);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuEdgeCalculator: hc_multi_create failed: ") + hc_last_error(nullptr));
}

GpuEdgeCalculator::GpuEdgeCalculator(double, double edge_quasi_cutoff, double overlap, bool, std::unordered_map<int, double>& simpson_map,
                                     double edge_quasi_cutoff_single, double overlap_single, double edge_quasi_cutoff_mixed,
                                     unsigned int maxPosition, bool noProb0, int device)
    : multi_(nullptr), aux_(nullptr), resident_(nullptr), flat_(new Flat()), row_feeding_(true) {
    init(edge_quasi_cutoff, overlap, simpson_map, edge_quasi_cutoff_single, overlap_single, edge_quasi_cutoff_mixed, maxPosition,
         noProb0, std::vector<int>(1, device));
}

GpuEdgeCalculator::GpuEdgeCalculator(double, double edge_quasi_cutoff, double overlap, bool, std::unordered_map<int, double>& simpson_map,
                                     double edge_quasi_cutoff_single, double overlap_single, double edge_quasi_cutoff_mixed,
                                     unsigned int maxPosition, bool noProb0, const std::vector<int>& devices)
    : multi_(nullptr), aux_(nullptr), resident_(nullptr), flat_(new Flat()), row_feeding_(true) {
    init(edge_quasi_cutoff, overlap, simpson_map, edge_quasi_cutoff_single, overlap_single, edge_quasi_cutoff_mixed, maxPosition,
         noProb0, devices);
}

GpuEdgeCalculator::~GpuEdgeCalculator() {
    hc_multi_destroy(multi_);
    hc_destroy(aux_);
    hc_packed_free(resident_);
    delete flat_;
}

void GpuEdgeCalculator::scoreBatch(const std::deque<AlignmentRecord*>& reads) {
    std::vector<const AlignmentRecord*> v(reads.begin(), reads.end());
    scoreBatch(v);
}

void GpuEdgeCalculator::scoreBatch(const std::vector<const AlignmentRecord*>& reads) {
    index_.clear();
    reads_ = reads;
    for (size_t k = 0; k < reads.size(); ++k) index_[reads[k]] = (int)k;
    flatten(reads, *flat_);
    const hc_reads view = flat_->view();
    hc_packed* packed = nullptr;
    int rc = hc_pack(&view, &packed);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuEdgeCalculator: hc_pack failed: ") + hc_last_error(nullptr));
    rc = hc_multi_upload(multi_, packed);
    hc_packed_free(resident_);                             // the previous batch's slices are gone with the upload
    resident_ = packed;
    if (rc != HC_OK) fail("hc_multi_upload", rc);
    const int n = (int)reads.size();
    col0_.assign((size_t)std::max(n, 1), 0);
    word_off_.assign((size_t)n + 1, 0);
    int64_t n_words = 0;
    if ((rc = hc_multi_band_layout(multi_, col0_.data(), word_off_.data(), &n_words)) != HC_OK) fail("hc_multi_band_layout", rc);
    words_.assign((size_t)std::max<int64_t>(n_words, 1), 0u);
    near_.clear();
    if ((rc = scoreResident(words_.data(), n_words, &stats_)) != HC_OK) fail("hc_multi_score_band", rc);
}

int GpuEdgeCalculator::scoreResident(uint32_t* words, int64_t n_words, hc_band_stats* stats) {
    // near-threshold pairs come back decided by std::log10 on the host (the bits in `words` are final); keep the report
    near_.assign(1024, hc_ambiguous());
    const int rc = hc_multi_score_band(multi_, words, n_words, near_.data(), (int32_t)near_.size(), stats, nullptr);
    near_.resize(rc == HC_OK ? (size_t)std::min<int64_t>(stats->ambiguous, (int64_t)near_.size()) : 0);
    return rc;
}

int GpuEdgeCalculator::scorePairOn(hc_ctx* ctx, int i, int j, bool* edge) const {
    hc_pair_result res;
    const int32_t pi = i, pj = j;
    const int rc = hc_score_pairs(ctx, &pi, &pj, 1, &res);
    if (rc == HC_OK) *edge = res.edge != 0;
    return rc;
}

hc_reads GpuEdgeCalculator::batchView() const { return flat_->view(); }

int GpuEdgeCalculator::indexOf(const AlignmentRecord& ap) const {
    const auto it = index_.find(&ap);
    return it == index_.end() ? -1 : it->second;
}

bool GpuEdgeCalculator::edgeBetween(const AlignmentRecord& ap1, const AlignmentRecord& ap2) const {
    const auto it1 = index_.find(&ap1), it2 = index_.find(&ap2);
    if (it1 != index_.end() && it2 != index_.end() && it2->second < it1->second) {
        const int i = it1->second, j = it2->second;
        if (j < col0_[i]) return false;                    // left of the band: the reference's first reject (:1252)
        const int rel = j - col0_[i];
        return (words_[(size_t)word_off_[i] + (rel >> 5)] >> (rel & 31)) & 1u;
    }
    // any other two records: a two-read batch, ap2 first (the earlier arrival), scored on a small second context
    if (aux_ == nullptr) {
        const int rc = hc_create(&par_, &aux_);
        if (rc != HC_OK) throw std::runtime_error(std::string("GpuEdgeCalculator: hc_create failed: ") + hc_last_error(nullptr));
    }
    hc_set_params(aux_, &par_);
    std::vector<const AlignmentRecord*> two;
    two.push_back(&ap2);
    two.push_back(&ap1);
    Flat f;
    flatten(two, f);
    const hc_reads view = f.view();
    hc_packed* pk = nullptr;
    int rc = hc_pack(&view, &pk);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuEdgeCalculator: hc_pack failed: ") + hc_last_error(nullptr));
    rc = hc_upload(aux_, pk);
    bool edge = false;
    if (rc == HC_OK) rc = scorePairOn(aux_, 1, 0, &edge);
    hc_packed_free(pk);
    if (rc != HC_OK)
        throw std::runtime_error(std::string("GpuEdgeCalculator: scoring a pair failed (") + std::to_string(rc) + "): " + hc_last_error(aux_));
    return edge;
}

void GpuEdgeCalculator::neighbours(const AlignmentRecord& ap, std::vector<const AlignmentRecord*>& out) const {
    out.clear();
    const auto it = index_.find(&ap);
    if (it == index_.end()) return;
    const int i = it->second;
    const int64_t w0 = word_off_[i], w1 = word_off_[(size_t)i + 1];
    for (int64_t w = w0; w < w1; ++w) {
        uint32_t bits = words_[(size_t)w];
        while (bits) {
            const int b = __builtin_ctz(bits);
            bits &= bits - 1;
            out.push_back(reads_[(size_t)col0_[i] + (size_t)(w - w0) * 32 + b]);
        }
    }
}

bool GpuEdgeCalculator::neighbourIndices(const AlignmentRecord& ap, int id, std::vector<int>& out) const {
    out.clear();
    if (id < 0 || (size_t)id >= reads_.size() || reads_[(size_t)id] != &ap) return false;
    const int64_t w0 = word_off_[(size_t)id], w1 = word_off_[(size_t)id + 1];
    const int c0 = col0_[(size_t)id];
    for (int64_t w = w0; w < w1; ++w) {
        uint32_t bits = words_[(size_t)w];
        while (bits) {
            const int b = __builtin_ctz(bits);
            bits &= bits - 1;
            out.push_back(c0 + (int)(w - w0) * 32 + b);
        }
    }
    return true;
}

void GpuEdgeCalculator::getPartnerLengthRange(const AlignmentRecord&, unsigned int*, unsigned int*) const { assert(false); }

void GpuEdgeCalculator::setOverlapCliques(double d) {      // NewEdgeCalculator::setOverlapCliques (src/NewEdgeCalculator.cpp:1382)
    par_.min_overlap_cliques = d;
    const int rc = hc_multi_set_params(multi_, &par_);
    if (rc != HC_OK) fail("hc_multi_set_params", rc);
}

// ---- QuasispeciesEdgeCalculator behind the same binding -------------------------------------------------------------
GpuQuasispeciesEdgeCalculator::GpuQuasispeciesEdgeCalculator(double Q, double edge_quasi_cutoff, double overlap, bool frameshift_merge,
                                                             std::unordered_map<int, double>& simpson_map,
                                                             double edge_quasi_cutoff_single, double overlap_single,
                                                             double edge_quasi_cutoff_mixed, unsigned int maxPosition, int device)
    : GpuEdgeCalculator(Q, edge_quasi_cutoff, overlap, frameshift_merge, simpson_map, edge_quasi_cutoff_single, overlap_single,
                        edge_quasi_cutoff_mixed, maxPosition, false, device) {
    std::memset(&qpar_, 0, sizeof qpar_);
    qpar_.edge_quasi_cutoff_cliques = edge_quasi_cutoff;
    qpar_.edge_quasi_cutoff_mixed = edge_quasi_cutoff_mixed;
    qpar_.edge_quasi_cutoff_single = edge_quasi_cutoff_single;
    qpar_.min_overlap_cliques = overlap;
    qpar_.min_overlap_single = overlap_single;
    qpar_.frameshift_merge = frameshift_merge ? 1 : 0;
    qpar_.ambiguity_rel_tol = 1e-12;
}

bool GpuQuasispeciesEdgeCalculator::edgeBetween(const AlignmentRecord& ap1, const AlignmentRecord& ap2) const {
    if (ap1.getName().compare(ap2.getName()) == 0) return true;          // src/QuasispeciesEdgeCalculator.cpp:56
    return GpuEdgeCalculator::edgeBetween(ap1, ap2);
}

int GpuQuasispeciesEdgeCalculator::scoreResident(uint32_t* words, int64_t n_words, hc_band_stats* stats) {
    // the older scorer runs on the first device of the handle (it is dead code in the reference's CLI)
    if (hc_multi_n_devices(handle()) != 1) throw std::runtime_error("GpuQuasispeciesEdgeCalculator: one device only");
    return hc_quasi_score_band(hc_multi_context(handle(), 0), &qpar_, 0, batchSize(), words, n_words, stats);
}

int GpuQuasispeciesEdgeCalculator::scorePairOn(hc_ctx* ctx, int i, int j, bool* edge) const {
    hc_quasi_result res;
    const int32_t pi = i, pj = j;
    const int rc = hc_quasi_score_pairs(ctx, &qpar_, &pi, &pj, 1, &res);
    if (rc == HC_OK) *edge = res.edge != 0;
    return rc;
}

void GpuQuasispeciesEdgeCalculator::neighbours(const AlignmentRecord& ap, std::vector<const AlignmentRecord*>& out) const {
    GpuEdgeCalculator::neighbours(ap, out);
    const int i = indexOf(ap);
    if (i <= 0) return;
    // the same-name shortcut (:56): an earlier read of the same name is adjacent whatever its score
    const std::string name = ap.getName();
    std::vector<const AlignmentRecord*> merged;
    size_t k = 0;
    bool any = false;
    for (int j = 0; j < i; ++j) {
        const AlignmentRecord* r = recordAt(j);
        const bool listed = k < out.size() && out[k] == r;
        if (listed) k++;
        const bool same = !listed && r->getName().compare(name) == 0;
        any |= same;
        if (listed || same) merged.push_back(r);
    }
    if (any) out.swap(merged);
}
