// hc_load.cpp -- alignments in BAM order -> the position-sorted SoA batch the scorer packs.
//
// What it replaces: readBamFile (src/haploclique.cpp:143-201) with the record building it drives --
// AlignmentRecord(BamAlignment) (src/AlignmentRecord.cpp:186-209), pairWith (:251-312) and, for mates that overlap,
// getMergedDnaSequence (:380-472) with its BAM-flavoured noOverlapMerge / overlapMerge (:474-580) and computeEntry
// (:169-184) -- up to the point where the scoring path takes over. The reference builds one heap-heavy AlignmentRecord per
// read (six allocations, plus cov_pos at 32 bytes per covered column); here a record is a few integers and three byte runs
// appended to flat streams, and cov_pos is never built: the scorer needs only its size and its first / last position, which
// are counted off the CIGAR.
//
// The decoder stays outside: the caller hands over decoded alignments one by one (integration/GpuBamLoader.cpp feeds them
// from the reference's vendored BamTools reader). The arrival order is the reference's: completed pairs in file order, then
// the unpaired leftovers in the iteration order of the reference's std::unordered_map<std::string, ...> (the same container
// with the same insertions and erasures gives the same order), then std::sort by start on a std::deque with the same
// comparator (not stable; the same algorithm on the same sequence gives the same permutation).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <deque>
#include <string>
#include <unordered_map>
#include <vector>

#include "hc_internal.h"

namespace {

constexpr int NQ = 94;

struct SegRef { int64_t cig_off = 0, seq_off = 0; int32_t cig_len = 0, seq_len = 0, start = 0, end = 0; };
struct Rec { SegRef s[2]; bool single = true; int32_t name = 0; };

inline uint8_t norm_base(uint8_t c) {           // ShortDnaSequence (src/ShortDnaSequence.cpp:37-50)
    if (c >= 'a' && c <= 'z') c = (uint8_t)(c - 32);
    return (c == 'A' || c == 'C' || c == 'G' || c == 'T') ? c : (uint8_t)'N';
}
inline bool clip(uint8_t c) { return c == 'S' || c == 'H'; }

}  // namespace

struct hc_loader {
    std::vector<uint8_t> cig, seq, qual;          // flat streams; a record's segments point into them
    std::deque<Rec> recs;                         // every record ever started (stable addresses)
    std::unordered_map<std::string, Rec*> open;   // names_to_reads (src/haploclique.cpp:146)
    std::deque<Rec*> order;                       // `reads` (:147)
    std::vector<std::string> names;               // readNames (:171-172)
    uint8_t tab[2 * NQ * NQ];
    std::vector<uint8_t> tmp_cig;                 // the mate's unrolled CIGAR
    // the finished batch
    std::vector<uint8_t> o_single, o_clique, o_cig, o_seq, o_qual;
    std::vector<int32_t> o_ncov, o_first, o_last, o_start, o_end, o_clen, o_slen, o_name;
    std::vector<int64_t> o_coff, o_soff;
    bool finished = false;
    int64_t alignments = 0, skipped = 0, merged_mates = 0;
    std::string err;

    SegRef append(int32_t start, int32_t end, const uint8_t* c, int32_t nc, const uint8_t* s, const uint8_t* q, int32_t ns) {
        SegRef r;
        r.start = start; r.end = end; r.cig_off = (int64_t)cig.size(); r.seq_off = (int64_t)seq.size(); r.cig_len = nc; r.seq_len = ns;
        cig.insert(cig.end(), c, c + nc);
        seq.insert(seq.end(), s, s + ns);
        qual.insert(qual.end(), q, q + ns);
        return r;
    }
    // computeEntry (:169-184)
    void entry(uint8_t b1, uint8_t q1, uint8_t b2, uint8_t q2, uint8_t& base, uint8_t& ql) const {
        const int k1 = std::min(std::max((int)q1 - 33, 0), NQ - 1), k2 = std::min(std::max((int)q2 - 33, 0), NQ - 1);
        if (b1 == b2) { base = b1; ql = tab[k1 * NQ + k2]; }
        else if (q1 >= q2) { base = b1; ql = tab[NQ * NQ + k1 * NQ + k2]; }
        else { base = b2; ql = tab[NQ * NQ + k2 * NQ + k1]; }
        if (q1 < 33 || q2 < 33) ql = 33;
    }
};

namespace {

// getMergedDnaSequence (:380-472): the record's segment 1 and the mate that overlaps it become one single-end segment.
// Returns false where the reference's walk is undefined (an op pair its overlapMerge does not list, a cursor past its end).
bool merge_mates(hc_loader* L, Rec& r, int32_t pos1, int32_t end_pos, const uint8_t* mc, int32_t mnc, const uint8_t* mseq, const uint8_t* mqual,
                 int32_t mns) {
    const SegRef a = r.s[0];
    const uint8_t* ac = L->cig.data() + a.cig_off;
    const uint8_t* as = L->seq.data() + a.seq_off;
    const uint8_t* aq = L->qual.data() + a.seq_off;
    auto lead = [](const uint8_t* c, int n) { int k = 0; while (k < n && clip(c[k])) ++k; return k; };
    auto trail = [](const uint8_t* c, int n) { int k = 0; while (k < n && clip(c[n - 1 - k])) ++k; return k; };
    int s1 = a.start - lead(ac, a.cig_len), e1 = a.end + trail(ac, a.cig_len);
    int s2 = pos1 - lead(mc, mnc), e2 = end_pos + trail(mc, mnc);
    int q1 = 0, q2 = 0, c1 = 0, c2 = 0;
    std::vector<uint8_t> oc, os, oq;
    oc.reserve((size_t)a.cig_len + mnc); os.reserve((size_t)a.seq_len + mns); oq.reserve((size_t)a.seq_len + mns);
    bool ok = true;
    auto put = [&](uint8_t op, uint8_t b, uint8_t q) { oc.push_back(op); os.push_back(b); oq.push_back(q); };
    // noOverlapMerge on this record (:1539-1576) / on the mate (:474-505)
    auto lone = [&](bool mate, int& ref) {
        const uint8_t* c = mate ? mc : ac; const int n = mate ? mnc : a.cig_len;
        const uint8_t* sq = mate ? mseq : as; const uint8_t* ql = mate ? mqual : aq; const int ns = mate ? mns : a.seq_len;
        int& cp = mate ? c2 : c1; int& qp = mate ? q2 : q1;
        if (cp >= n) { ok = false; return; }
        const uint8_t op = c[cp];
        if (op == 'M' || op == 'I') {
            if (qp >= ns) { ok = false; return; }
            put(op, sq[qp], ql[qp]); ++qp; ++cp;
            if (op == 'M') ++ref;
        } else if (op == 'D') { oc.push_back('D'); ++ref; ++cp; }
        else if (op == 'S') { ++ref; ++qp; ++cp; }
        else if (op == 'H') { ++ref; ++cp; }
        else ok = false;
    };
    // overlapMerge with a BAM mate (:507-580): M against D is allowed here (the M wins), any I is taken on its own
    auto both = [&](int& ref) {
        if (c1 >= a.cig_len || c2 >= mnc) { ok = false; return; }
        const uint8_t x = ac[c1], y = mc[c2];
        if ((x == 'M' && y == 'M') || (x == 'S' && y == 'S') || (x == 'I' && y == 'I')) {
            if (x != 'S') {
                if (q1 >= a.seq_len || q2 >= mns) { ok = false; return; }
                uint8_t b, q;
                L->entry(as[q1], aq[q1], mseq[q2], mqual[q2], b, q);
                put(x, b, q);
            }
            if (x != 'I') ++ref;
            ++q1; ++q2; ++c1; ++c2;
        } else if ((x == 'D' && y == 'D') || (x == 'H' && y == 'H') || (x == 'D' && y == 'H') || (x == 'H' && y == 'D') || (x == 'D' && y == 'S') ||
                   (x == 'S' && y == 'D')) {
            ++c1; ++c2; ++ref;
            if (x == 'D' || y == 'D') oc.push_back('D');
            if (x == 'S') ++q1; else if (y == 'S') ++q2;
        } else if ((x == 'M' && (y == 'D' || y == 'H' || y == 'S')) || ((x == 'D' || x == 'H' || x == 'S') && y == 'M') || (x == 'S' && y == 'H') ||
                   (x == 'H' && y == 'S')) {
            if (x == 'M') { if (q1 >= a.seq_len) { ok = false; return; } put('M', as[q1], aq[q1]); ++q1; if (y == 'S') ++q2; }
            else if (y == 'M') { if (q2 >= mns) { ok = false; return; } put('M', mseq[q2], mqual[q2]); ++q2; if (x == 'S') ++q1; }
            else if (x == 'S') ++q1;
            else ++q2;
            ++ref; ++c1; ++c2;
        } else if (x == 'I' || y == 'I') {
            if (x == 'I') { if (q1 >= a.seq_len) { ok = false; return; } put('I', as[q1], aq[q1]); ++c1; ++q1; }
            else { if (q2 >= mns) { ok = false; return; } put('I', mseq[q2], mqual[q2]); ++c2; ++q2; }
        } else ok = false;
    };
    if (s1 <= s2 && e1 <= e2) {
        while (ok && s1 < s2) lone(false, s1);
        while (ok && s1 <= e1) both(s1);
        while (ok && s1 <= e2) lone(true, s1);
    } else if (s1 <= s2 && e1 >= e2) {
        while (ok && s1 < s2) lone(false, s1);
        while (ok && s1 <= e2) both(s1);
        while (ok && s1 <= e1) lone(false, s1);
    } else if (s1 >= s2 && e1 <= e2) {
        while (ok && s2 < s1) lone(true, s2);
        while (ok && s2 <= e1) both(s2);
        while (ok && s2 <= e2) lone(true, s2);
    } else {
        while (ok && s2 < s1) lone(true, s2);
        while (ok && s2 <= e2) both(s2);
        while (ok && s2 <= e1) lone(false, s2);
    }
    if (!ok) return false;
    for (uint8_t& b : os) b = norm_base(b);                     // sequence1 = ShortDnaSequence(dna, qualities) (:458)
    r.s[0] = L->append(std::min(a.start, pos1), std::max(end_pos, a.end), oc.data(), (int32_t)oc.size(), os.data(), oq.data(), (int32_t)os.size());
    r.single = true;
    L->merged_mates++;
    return true;
}

}  // namespace

extern "C" {

int hc_loader_create(hc_loader** out) {
    if (!out) { hc::set_global_error("hc_loader_create: null argument"); return HC_E_INVALID; }
    hc_loader* L = new hc_loader();
    hc::build_phred_tables(L->tab);
    *out = L;
    return HC_OK;
}

void hc_loader_destroy(hc_loader* L) { delete L; }

const char* hc_loader_last_error(const hc_loader* L) { return L ? L->err.c_str() : hc::global_error(); }

int hc_loader_add(hc_loader* L, const char* name, int32_t position0, const char* cigar_ops, const uint32_t* cigar_lens, int32_t n_cigar,
                  const char* bases, const char* quals, int32_t n_bases) {
    if (!L || !name || n_cigar < 0 || n_bases < 0 || (n_cigar > 0 && (!cigar_ops || !cigar_lens)) || (n_bases > 0 && (!bases || !quals))) {
        if (L) L->err = "hc_loader_add: bad argument";
        return HC_E_INVALID;
    }
    if (L->finished) { L->err = "hc_loader_add: the batch is finished"; return HC_E_STATE; }
    L->alignments++;
    // readBamFile keeps an alignment only if its CIGAR holds an M op of positive length (:160-165)
    bool valid = false;
    int64_t unrolled = 0;
    for (int k = 0; k < n_cigar; ++k) { valid |= (cigar_ops[k] == 'M' && cigar_lens[k] > 0); unrolled += cigar_lens[k]; }
    if (n_cigar == 0 || !valid) { L->skipped++; return HC_OK; }
    if (unrolled > 65534) { L->err = "hc_loader_add: a CIGAR longer than 65534 ops"; return HC_E_INVALID; }
    // the unrolled CIGAR (:197-200) and GetEndPosition (external/BamTools/src/api/BamAlignment.cpp:478-517: D M X N = advance)
    std::vector<uint8_t>& uc = L->tmp_cig;
    uc.clear();
    int32_t end_pos = position0;
    for (int k = 0; k < n_cigar; ++k) {
        const char t = cigar_ops[k];
        uc.insert(uc.end(), cigar_lens[k], (uint8_t)t);
        if (t == 'D' || t == 'M' || t == 'X' || t == 'N' || t == '=') end_pos += (int32_t)cigar_lens[k];
    }
    const int32_t pos1 = position0 + 1;
    const std::string key(name);
    auto it = L->open.find(key);
    if (it == L->open.end()) {
        // new AlignmentRecord(alignment, readNames.size(), &readNames) (:171-172, ctor :186-209)
        L->recs.emplace_back();
        Rec& r = L->recs.back();
        r.name = (int32_t)L->names.size();
        L->names.push_back(key);
        const size_t at = L->seq.size();
        r.s[0] = L->append(pos1, end_pos, uc.data(), (int32_t)uc.size(), (const uint8_t*)bases, (const uint8_t*)quals, n_bases);
        for (size_t k = at; k < L->seq.size(); ++k) L->seq[k] = norm_base(L->seq[k]);
        L->open[key] = &r;
        return HC_OK;
    }
    // pairWith (:251-312)
    Rec& r = *it->second;
    if ((unsigned)pos1 > (unsigned)r.s[0].end) {                       // the mate lies behind: it becomes segment 2
        const size_t at = L->seq.size();
        r.s[1] = L->append(pos1, end_pos, uc.data(), (int32_t)uc.size(), (const uint8_t*)bases, (const uint8_t*)quals, n_bases);
        for (size_t k = at; k < L->seq.size(); ++k) L->seq[k] = norm_base(L->seq[k]);
        r.single = false;
    } else if ((unsigned)end_pos < (unsigned)r.s[0].start) {           // the mate lies in front: the record's read becomes segment 2
        r.s[1] = r.s[0];
        const size_t at = L->seq.size();
        r.s[0] = L->append(pos1, end_pos, uc.data(), (int32_t)uc.size(), (const uint8_t*)bases, (const uint8_t*)quals, n_bases);
        for (size_t k = at; k < L->seq.size(); ++k) L->seq[k] = norm_base(L->seq[k]);
        r.single = false;
    } else if (!merge_mates(L, r, pos1, end_pos, uc.data(), (int32_t)uc.size(), (const uint8_t*)bases, (const uint8_t*)quals, n_bases)) {
        L->err = "hc_loader_add: read " + key + ": the reference's mate merge is undefined for these two CIGARs";
        return HC_E_INVALID;
    }
    L->order.push_back(&r);                                           // reads->push_back (:168)
    L->open.erase(it);
    return HC_OK;
}

int hc_loader_finish(hc_loader* L, hc_reads* out, uint32_t* max_position) {
    if (!L || !out) { if (L) L->err = "hc_loader_finish: null argument"; return HC_E_INVALID; }
    if (!L->finished) {
        for (const auto& kv : L->open) L->order.push_back(kv.second);                         // leftovers in hash order (:178-180)
        if (L->order.empty()) { L->err = "No reads could be retrieved from the BamFile. "; return HC_E_INVALID; }   // (:183-186)
        auto comp = [](const Rec* a, const Rec* b) { return (unsigned)a->s[0].start < (unsigned)b->s[0].start; };   // getIntervalStart (:190)
        std::sort(L->order.begin(), L->order.end(), comp);
        const size_t n = L->order.size();
        L->o_single.resize(n); L->o_clique.resize(n); L->o_ncov.resize(n); L->o_first.resize(n); L->o_last.resize(n); L->o_name.resize(n);
        L->o_start.assign(2 * n, 0); L->o_end.assign(2 * n, 0); L->o_clen.assign(2 * n, 0); L->o_slen.assign(2 * n, 0);
        L->o_coff.assign(2 * n, 0); L->o_soff.assign(2 * n, 0);
        // the streams are rewritten in arrival order: a record's segments become contiguous, dead bytes of merged mates go
        L->o_cig.clear(); L->o_seq.clear(); L->o_qual.clear();
        L->o_cig.reserve(L->cig.size()); L->o_seq.reserve(L->seq.size()); L->o_qual.reserve(L->qual.size());
        for (size_t r = 0; r < n; ++r) {
            const Rec& rec = *L->order[r];
            L->o_single[r] = rec.single ? 1 : 0;
            L->o_name[r] = rec.name;
            L->o_clique[r] = L->names[(size_t)rec.name].find("Clique") != std::string::npos ? 1 : 0;
            int32_t ncov = 0, first = 0, last = 0;
            bool any = false;
            for (int s = 0; s < (rec.single ? 1 : 2); ++s) {
                const SegRef& sg = rec.s[s];
                const size_t k = 2 * r + s;
                L->o_start[k] = sg.start; L->o_end[k] = sg.end; L->o_clen[k] = sg.cig_len; L->o_slen[k] = sg.seq_len;
                L->o_coff[k] = (int64_t)L->o_cig.size(); L->o_soff[k] = (int64_t)L->o_seq.size();
                const uint8_t* c = L->cig.data() + sg.cig_off;
                L->o_cig.insert(L->o_cig.end(), c, c + sg.cig_len);
                L->o_seq.insert(L->o_seq.end(), L->seq.data() + sg.seq_off, L->seq.data() + sg.seq_off + sg.seq_len);
                L->o_qual.insert(L->o_qual.end(), L->qual.data() + sg.seq_off, L->qual.data() + sg.seq_off + sg.seq_len);
                // what coveredPositions (:315-378) would list: one entry per M op, positions counted over M and D
                int32_t ref = sg.start;
                for (int t = 0; t < sg.cig_len; ++t) {
                    if (c[t] == 'M') { if (!any) { first = ref; any = true; } last = ref; ++ncov; ++ref; }
                    else if (c[t] == 'D') ++ref;
                }
            }
            if (rec.single) { L->o_coff[2 * r + 1] = (int64_t)L->o_cig.size(); L->o_soff[2 * r + 1] = (int64_t)L->o_seq.size(); }
            L->o_ncov[r] = ncov; L->o_first[r] = first; L->o_last[r] = last;
        }
        L->finished = true;
    }
    out->n = (int32_t)L->order.size();
    out->single_end = L->o_single.data(); out->is_clique = L->o_clique.data(); out->ncov = L->o_ncov.data();
    out->cov_first = L->o_first.data(); out->cov_last = L->o_last.data(); out->seg_start = L->o_start.data(); out->seg_end = L->o_end.data();
    out->cig_off = L->o_coff.data(); out->cig_len = L->o_clen.data(); out->seq_off = L->o_soff.data(); out->seq_len = L->o_slen.data();
    out->cig = L->o_cig.data(); out->seq = L->o_seq.data(); out->qual = L->o_qual.data();
    if (max_position) {                                                   // max getIntervalEnd (:195-198)
        uint32_t mx = 0;
        for (const Rec* r : L->order) mx = std::max(mx, (uint32_t)(r->single ? r->s[0].end : r->s[1].end));
        *max_position = mx;
    }
    return HC_OK;
}

int32_t hc_loader_n_names(const hc_loader* L) { return L ? (int32_t)L->names.size() : 0; }
const char* hc_loader_name(const hc_loader* L, int32_t k) { return (L && k >= 0 && (size_t)k < L->names.size()) ? L->names[(size_t)k].c_str() : nullptr; }
int32_t hc_loader_record_name(const hc_loader* L, int32_t read) {
    return (L && L->finished && read >= 0 && (size_t)read < L->o_name.size()) ? L->o_name[(size_t)read] : -1;
}
int hc_loader_counts(const hc_loader* L, int64_t* alignments, int64_t* skipped, int64_t* merged_mates) {
    if (!L) return HC_E_INVALID;
    if (alignments) *alignments = L->alignments;
    if (skipped) *skipped = L->skipped;
    if (merged_mates) *merged_mates = L->merged_mates;
    return HC_OK;
}

}  // extern "C"
