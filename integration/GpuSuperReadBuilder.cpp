// GpuSuperReadBuilder.cpp -- see GpuSuperReadBuilder.h. Host C++ only; the merge itself is behind the C ABI (hc_edge.h).
#include "GpuSuperReadBuilder.h"

#include <stdexcept>

GpuSuperReadBuilder::GpuSuperReadBuilder(const GpuEdgeCalculator& calc, int device)
    : calc_(calc), merger_(nullptr), kernel_ms_(0.0), cliques_(0), folds_plain_(0), folds_walked_(0) {
    const int rc = hc_merger_create(device, &merger_);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuSuperReadBuilder: hc_merger_create failed: ") + hc_merger_last_error(nullptr));
}

GpuSuperReadBuilder::~GpuSuperReadBuilder() { hc_merger_destroy(merger_); }

void GpuSuperReadBuilder::build(const std::vector<std::vector<int> >& members, const std::vector<unsigned int>& clique_ids,
                                const std::vector<std::set<int> >& read_names, std::vector<std::string>* read_name_map,
                                std::vector<AlignmentRecord*>& out) {
    out.clear();
    const int n = (int)members.size();
    if (n == 0) return;
    const hc_reads view = calc_.batchView();                  // the reads of this round, as scoreBatch flattened them
    int rc = hc_merger_set_reads(merger_, &view);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuSuperReadBuilder: hc_merger_set_reads failed: ") + hc_merger_last_error(merger_));
    std::vector<int64_t> member_off((size_t)n + 1, 0);
    std::vector<int32_t> flat;
    for (int c = 0; c < n; ++c) {
        for (size_t k = 0; k < members[c].size(); ++k) flat.push_back(members[c][k]);
        member_off[(size_t)c + 1] = (int64_t)flat.size();
    }
    std::vector<hc_merged_record> info((size_t)n);
    int64_t total_ops = 0, total_seq = 0;
    float ms = 0.f;
    rc = hc_merge_run(merger_, n, member_off.data(), flat.data(), info.data(), &total_ops, &total_seq, &ms);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuSuperReadBuilder: hc_merge_run failed: ") + hc_merger_last_error(merger_));
    std::string cig((size_t)total_ops, '\0'), seq((size_t)total_seq, '\0'), qual((size_t)total_seq, '\0');
    rc = hc_merge_fetch(merger_, (uint8_t*)&cig[0], (uint8_t*)&seq[0], (uint8_t*)&qual[0]);
    if (rc != HC_OK) throw std::runtime_error(std::string("GpuSuperReadBuilder: hc_merge_fetch failed: ") + hc_merger_last_error(merger_));
    kernel_ms_ += ms;
    cliques_ += n;
    int64_t a = 0, b = 0;
    hc_merge_fold_counts(merger_, &a, &b);
    folds_plain_ += (long)a;
    folds_walked_ += (long)b;
    size_t oa = 0, sa = 0;
    out.reserve((size_t)n);
    for (int c = 0; c < n; ++c) {
        const hc_merged_record& r = info[(size_t)c];
        if (r.status != HC_MERGE_OK)
            throw std::runtime_error("GpuSuperReadBuilder: the reference's merge is undefined for clique " + std::to_string(clique_ids[(size_t)c]) +
                                     (r.status == HC_MERGE_UNDEFINED ? " (an op pair overlapMerge does not list)" : " (a cursor runs off its CIGAR)"));
        AlignmentRecord::MergedSegment s[2];
        for (int sg = 0; sg < (r.single_end ? 1 : 2); ++sg) {
            s[sg].start = r.start[sg];
            s[sg].end = r.end[sg];
            s[sg].cigar_unrolled = cig.substr(oa, (size_t)r.n_ops[sg]);
            s[sg].dna = seq.substr(sa, (size_t)r.n_seq[sg]);
            s[sg].qualities = qual.substr(sa, (size_t)r.n_seq[sg]);
            oa += (size_t)r.n_ops[sg];
            sa += (size_t)r.n_seq[sg];
        }
        out.push_back(new AlignmentRecord(read_names[(size_t)c], read_name_map, clique_ids[(size_t)c], r.single_end != 0, s[0], s[1]));
    }
}
