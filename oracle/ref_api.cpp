// oracle/ref_api.cpp -- TEST INFRASTRUCTURE ONLY (never on the product path).
//
// A thin extern "C" face over the UNMODIFIED reference sources where they lie under
// /root/reference (nothing is copied into this repo).  Built by oracle/Makefile into
// oracle/_ref/libhcref.so.  Used to (1) pin oracle/edge_oracle.c against the real
// NewEdgeCalculator, (2) generate the golden fixtures under tests/golden/, and (3) time the
// reference's own CPU scorer for bench.py's reference arm / cpu_baseline.
//
// The reference CLI translation unit is included whole (the reference's own unit test does the
// same: test/unit/unit1.cpp:11-12), which gives us readBamFile (src/haploclique.cpp:143) and
// the CLI entry as HaploMain.
// QuasispeciesEdgeCalculator::computeOverlap is private; the checker wants its value, not only the boolean. The class
// body alone is read with `private` spelled `public` (its own includes come first, so nothing else is affected;
// access specifiers do not change the layout of this class, which has a single access section of data members).
#include <map>
#include <set>
#include <string>
#include "AlignmentRecord.h"
#include "EdgeCalculator.h"
#define private public
#include "QuasispeciesEdgeCalculator.h"
#undef private

#define main HaploMain
#include "haploclique.cpp"
#undef main

#include <chrono>
#ifdef HC_WITH_GPU_ADAPTER
#include "GpuBamLoader.h"
#include "GpuSuperReadBuilder.h"
#include "GpuEdgeCalculator.h"          // integration/: the product's reference-side binding, tested through the reference's finders
#endif

#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>

std::pair<char, char> computeEntry(const char& base1, const char& qual1, const char& base2, const char& qual2);   // src/AlignmentRecord.cpp:169

namespace {

struct RefCtx {
    std::vector<AlignmentRecord*> recs;   // position-sorted arrival order
    std::vector<std::string> names;       // readNameMap storage
    std::unordered_map<std::string, AlignmentRecord*> open_mates;
    NewEdgeCalculator* calc = nullptr;
    // parameters kept so worker threads can build their own calculator
    // (edgeBetween clears a mutable member: src/NewEdgeCalculator.h:43, .cpp:1265).
    double Q = 0.9, cut_cliques = 0.99, ov_cliques = 0.9, cut_single = 0.95, ov_single = 0.6, cut_mixed = 0.97;
    bool frameshift = false, no_prob0 = false;
    unsigned max_position = 0;
    size_t maxpos_done = 0;
    std::unordered_map<int, double> simpson;
    std::string err;
    BamTools::BamWriter* writer = nullptr;   // optional: every alignment fed to hcref_add_alignment is also written here
    ~RefCtx() {
        if (writer) { writer->Close(); delete writer; }
        for (auto* r : recs) delete r;
        for (auto& kv : open_mates) delete kv.second;
        delete calc;
    }
    NewEdgeCalculator* make_calc() {
        return new NewEdgeCalculator(Q, cut_cliques, ov_cliques, frameshift, simpson, cut_single, ov_single,
                                     cut_mixed, max_position, no_prob0);
    }
};

std::vector<BamTools::CigarOp> rle(const char* cig, int n) {
    std::vector<BamTools::CigarOp> out;
    for (int i = 0; i < n;) {
        int j = i;
        while (j < n && cig[j] == cig[i]) ++j;
        out.push_back(BamTools::CigarOp(cig[i], (uint32_t)(j - i)));
        i = j;
    }
    return out;
}

// The interval tests edgeBetween applies before the walk (src/NewEdgeCalculator.cpp:1252-1365),
// restated so the harness knows for which pairs calling the public checkGapsCigar is defined.
bool prefilter_pass(const AlignmentRecord& a, const AlignmentRecord& b) {
    const auto& c1 = a.getCovmap();
    const auto& c2 = b.getCovmap();
    if (c1.back().ref < c2[0].ref || c2.back().ref < c1[0].ref) return false;
    bool se1 = a.isSingleEnd(), se2 = b.isSingleEnd();
    unsigned e11 = a.getEnd1(), s11 = a.getStart1(), e21 = b.getEnd1(), s21 = b.getStart1();
    if (se1 && se2) {
        if (e11 < s21) return false;
        if (e21 < s11) return false;
    } else if (!se1 && !se2) {
        unsigned e12 = a.getEnd2(), s12 = a.getStart2(), e22 = b.getEnd2(), s22 = b.getStart2();
        if (e11 < s21 && e12 < s22) return false;
        if (e21 < s11 && e22 < s12) return false;
        if (e21 < s12 && e12 < s22) return false;
        if (e11 < s22 && e22 < s12) return false;
        if (e21 < s11 && e11 < s22) return false;
        if (e11 < s21 && e21 < s12) return false;
    } else if (se1 && !se2) {
        unsigned e22 = b.getEnd2(), s22 = b.getStart2();
        if (e11 < s21) return false;
        if (e21 < s11 && e11 < s22) return false;
        if (e22 < s11) return false;
    } else {
        unsigned e12 = a.getEnd2(), s12 = a.getStart2();
        if (e12 < s21) return false;
        if (e11 < s21 && e21 < s12) return false;
        if (e21 < s11) return false;
    }
    return true;
}

}  // namespace

extern "C" {

void* hcref_new() { return new RefCtx(); }
void hcref_free(void* p) { delete static_cast<RefCtx*>(p); }
const char* hcref_error(void* p) { return static_cast<RefCtx*>(p)->err.c_str(); }

// Runs the unmodified CLI (src/haploclique.cpp:203) with argv; cwd decides where the FASTA lands.
int hcref_cli(int argc, char** argv) { return HaploMain(argc, argv); }

// readBamFile (src/haploclique.cpp:143-201): mate pairing, leftover order and the sort included.
int hcref_load_bam(void* p, const char* path) {
    RefCtx* c = static_cast<RefCtx*>(p);
    try {
        BamTools::SamHeader header;
        BamTools::RefVector refs;
        unsigned maxpos = 0;
        std::deque<AlignmentRecord*>* reads = readBamFile(path, c->names, maxpos, header, refs);
        for (auto* r : *reads) c->recs.push_back(r);
        delete reads;
        c->max_position = std::max(c->max_position, maxpos);
    } catch (const std::exception& e) {
        c->err = e.what();
        return -1;
    }
    return (int)c->recs.size();
}

// One alignment in BAM order, handled as the loop body of readBamFile does (:160-175):
// first sight of a name -> new record, second sight -> pairWith + push.
int hcref_add_alignment(void* p, const char* name, int pos0, const char* cigar_unrolled, int ncig,
                        const char* bases, const char* quals, int nq) {
    RefCtx* c = static_cast<RefCtx*>(p);
    BamTools::BamAlignment al;
    al.Name = name;
    al.Position = pos0;
    al.CigarData = rle(cigar_unrolled, ncig);
    al.QueryBases.assign(bases, nq);
    al.Qualities.assign(quals, nq);
    al.Length = nq;
    if (c->writer) {
        al.RefID = 0; al.MateRefID = -1; al.MatePosition = -1; al.MapQuality = 60; al.AlignmentFlag = 0; al.InsertSize = 0; al.Bin = 0;
        c->writer->SaveAlignment(al);
    }
    bool valid = false;
    for (auto& op : al.CigarData) if (op.Type == 'M' && op.Length > 0) valid = true;
    if (!valid) return 0;
    auto it = c->open_mates.find(al.Name);
    if (it != c->open_mates.end()) {
        it->second->pairWith(al);
        c->recs.push_back(it->second);
        c->open_mates.erase(it);
    } else {
        c->open_mates[al.Name] = new AlignmentRecord(al, (int)c->names.size(), &c->names);
        c->names.push_back(al.Name);
    }
    return 1;
}

// Opens a BAM file (the reference's vendored BamWriter) that receives every alignment passed to hcref_add_alignment,
// so that synthetic read sets can be fed to the unmodified CLI. One reference sequence `ref_name` of length ref_len.
int hcref_bam_writer_open(void* p, const char* path, const char* ref_name, int ref_len) {
    RefCtx* c = static_cast<RefCtx*>(p);
    if (c->writer) { c->writer->Close(); delete c->writer; c->writer = nullptr; }
    BamTools::RefVector refs;
    refs.push_back(BamTools::RefData(ref_name, ref_len));
    const std::string header = std::string("@HD\tVN:1.0\tSO:unsorted\n@SQ\tSN:") + ref_name + "\tLN:" + std::to_string(ref_len) + "\n";
    c->writer = new BamTools::BamWriter();
    if (!c->writer->Open(path, header, refs)) { c->err = "BamWriter::Open failed"; delete c->writer; c->writer = nullptr; return -1; }
    return 0;
}
int hcref_bam_writer_close(void* p) {
    RefCtx* c = static_cast<RefCtx*>(p);
    if (c->writer) { c->writer->Close(); delete c->writer; c->writer = nullptr; }
    return 0;
}

// End of readBamFile: leftovers appended, optional sort by interval start, max_position (:178-198).
int hcref_finish_load(void* p, int do_sort) {
    RefCtx* c = static_cast<RefCtx*>(p);
    for (auto& kv : c->open_mates) c->recs.push_back(kv.second);
    c->open_mates.clear();
    if (do_sort) {
        std::sort(c->recs.begin(), c->recs.end(), [](AlignmentRecord* a, AlignmentRecord* b) {
            return a->getIntervalStart() < b->getIntervalStart();
        });
    }
    for (size_t k = c->maxpos_done; k < c->recs.size(); ++k)
        c->max_position = std::max(c->max_position, c->recs[k]->getIntervalEnd());
    c->maxpos_done = do_sort ? 0 : c->recs.size();
    return (int)c->recs.size();
}

// Unit-test fixture loader (src/AlignmentRecord.cpp:2346), appended in call order.
int hcref_restore_fixture(void* p, const char* path) {
    RefCtx* c = static_cast<RefCtx*>(p);
    AlignmentRecord* r = new AlignmentRecord();
    r->restoreCompleteAlignmentRecord(path);
    c->recs.push_back(r);
    return (int)c->recs.size() - 1;
}

// Super-read construction: AlignmentRecord(alignments, clique_id) (src/AlignmentRecord.cpp:211-249) over
// records[members[0..m)], appended to the record list like hcref_restore_fixture does. Returns the new record's index.
int hcref_merge_clique(void* p, const int* members, int m, unsigned clique_id) {
    RefCtx* c = static_cast<RefCtx*>(p);
    std::unique_ptr<std::vector<const AlignmentRecord*>> al(new std::vector<const AlignmentRecord*>());
    for (int k = 0; k < m; ++k) al->push_back(c->recs[members[k]]);
    c->recs.push_back(new AlignmentRecord(al, clique_id));
    return (int)c->recs.size() - 1;
}

// computeEntry (src/AlignmentRecord.cpp:169-184): consensus base and updated quality char of one merged column.
void hcref_compute_entry(char b1, char q1, char b2, char q2, char* out2) {
    std::pair<char, char> r = computeEntry(b1, q1, b2, q2);
    out2[0] = r.first;
    out2[1] = r.second;
}

int hcref_n(void* p) { return (int)static_cast<RefCtx*>(p)->recs.size(); }
unsigned hcref_max_position(void* p) { return static_cast<RefCtx*>(p)->max_position; }

// out[12]: single_end,start1,end1,start2,end2,ncig1,nq1,ncig2,nq2,ncov,cov_first,cov_last
int hcref_record_info(void* p, int i, int* out, char* name, int name_cap) {
    RefCtx* c = static_cast<RefCtx*>(p);
    const AlignmentRecord& r = *c->recs[i];
    bool se = r.isSingleEnd();
    out[0] = se;
    out[1] = r.getStart1();
    out[2] = r.getEnd1();
    out[3] = se ? 0 : r.getStart2();
    out[4] = se ? 0 : r.getEnd2();
    out[5] = (int)r.getCigar1Unrolled().size();
    out[6] = (int)r.getSequence1().size();
    out[7] = se ? 0 : (int)r.getCigar2Unrolled().size();
    out[8] = se ? 0 : (int)r.getSequence2().size();
    const auto& cov = r.getCovmap();
    out[9] = (int)cov.size();
    out[10] = cov.empty() ? 0 : cov.front().ref;
    out[11] = cov.empty() ? 0 : cov.back().ref;
    std::string nm = r.getName();
    std::snprintf(name, name_cap, "%s", nm.c_str());
    return 0;
}

int hcref_record_data(void* p, int i, char* cig1, char* seq1, char* qual1, char* cig2, char* seq2, char* qual2) {
    RefCtx* c = static_cast<RefCtx*>(p);
    const AlignmentRecord& r = *c->recs[i];
    const auto& c1 = r.getCigar1Unrolled();
    std::memcpy(cig1, c1.data(), c1.size());
    std::memcpy(seq1, r.getSequence1().toString().data(), r.getSequence1().size());
    std::memcpy(qual1, r.getSequence1().qualityString().data(), r.getSequence1().size());
    if (!r.isSingleEnd()) {
        const auto& c2 = r.getCigar2Unrolled();
        std::memcpy(cig2, c2.data(), c2.size());
        std::memcpy(seq2, r.getSequence2().toString().data(), r.getSequence2().size());
        std::memcpy(qual2, r.getSequence2().qualityString().data(), r.getSequence2().size());
    }
    return 0;
}

// NewEdgeCalculator construction (src/NewEdgeCalculator.h:58; call site src/haploclique.cpp:304).
// max_position < 0 keeps the value derived while loading.
int hcref_set_params(void* p, double Q, double cut_cliques, double ov_cliques, int frameshift,
                     const int* simpson_pos, const double* simpson_val, int n_simpson, double cut_single,
                     double ov_single, double cut_mixed, long max_position, int no_prob0) {
    RefCtx* c = static_cast<RefCtx*>(p);
    c->Q = Q; c->cut_cliques = cut_cliques; c->ov_cliques = ov_cliques; c->frameshift = frameshift != 0;
    c->cut_single = cut_single; c->ov_single = ov_single; c->cut_mixed = cut_mixed; c->no_prob0 = no_prob0 != 0;
    c->simpson.clear();
    for (int i = 0; i < n_simpson; ++i) c->simpson[simpson_pos[i]] = simpson_val[i];
    if (max_position >= 0) c->max_position = (unsigned)max_position;
    delete c->calc;
    c->calc = c->make_calc();
    return 0;
}

// edgeBetween(records[i], records[j]) (src/NewEdgeCalculator.cpp:1247).
int hcref_edge(void* p, int i, int j) {
    RefCtx* c = static_cast<RefCtx*>(p);
    return c->calc->edgeBetween(*c->recs[i], *c->recs[j]) ? 1 : 0;
}

int hcref_prefilter(void* p, int i, int j) {
    RefCtx* c = static_cast<RefCtx*>(p);
    return prefilter_pass(*c->recs[i], *c->recs[j]) ? 1 : 0;
}

// The walk alone through the public checkGapsCigar (src/NewEdgeCalculator.h:72), with the same
// initial state edgeBetween gives it (:1256-1259). Only defined when the prefilter passes.
int hcref_pair_detail(void* p, int i, int j, int* gap, int* cc, int* tc, double* probM, double* prob0) {
    RefCtx* c = static_cast<RefCtx*>(p);
    if (!prefilter_pass(*c->recs[i], *c->recs[j])) return 0;
    int tcv = 0, ccv = 0;
    double pm = 1.0, p0 = 0.0;
    bool g = c->calc->checkGapsCigar(*c->recs[i], *c->recs[j], pm, p0, ccv, tcv);
    *gap = g; *cc = ccv; *tc = tcv; *probM = pm; *prob0 = p0;
    return 1;
}

// QuasispeciesEdgeCalculator (src/QuasispeciesEdgeCalculator.cpp): q = computeOverlap(records[pi], records[pj], cutoff)
// with the cutoff edgeBetween picks from the two names (:59-66), and the boolean edgeBetween returns (:69).
// simpson: the calculator's map<int,double> (its constructor copies it, :42).
long hcref_quasi_pairs(void* p, const int* pi, const int* pj, long npairs, double cut_cliques, double ov_cliques,
                       int frameshift, const int* simpson_pos, const double* simpson_val, int n_simpson, double cut_single,
                       double ov_single, double cut_mixed, double* q_out, unsigned char* edge_out) {
    RefCtx* c = static_cast<RefCtx*>(p);
    std::map<int, double> simpson;
    for (int i = 0; i < n_simpson; ++i) simpson[simpson_pos[i]] = simpson_val[i];
    QuasispeciesEdgeCalculator calc(0.9, cut_cliques, ov_cliques, frameshift != 0, simpson, cut_single, ov_single, cut_mixed);
    long edges = 0;
    for (long k = 0; k < npairs; ++k) {
        const AlignmentRecord& a = *c->recs[pi[k]];
        const AlignmentRecord& b = *c->recs[pj[k]];
        const bool ca = a.getName().find("Clique") != std::string::npos, cb = b.getName().find("Clique") != std::string::npos;
        const double cutoff = (ca && cb) ? cut_cliques : ((ca || cb) ? cut_mixed : cut_single);
        if (q_out) q_out[k] = calc.computeOverlap(a, b, cutoff);
        const bool e = calc.edgeBetween(a, b);
        if (edge_out) edge_out[k] = e;
        edges += e;
    }
    return edges;
}

// All ordered pairs (i = later arrival, j < i), CLEVER's order (src/CLEVER.cpp:218-232).
// edge[k], and for prefilter-passing pairs the walk outputs; k enumerates (i,j) row-major over i>j.
// Arrays may be null. Returns number of edges.
long hcref_score_all(void* p, unsigned char* edge, unsigned char* pre, unsigned char* gap, int* cc, int* tc,
                     double* probM, double* prob0) {
    RefCtx* c = static_cast<RefCtx*>(p);
    long n = (long)c->recs.size(), k = 0, edges = 0;
    for (long i = 0; i < n; ++i) {
        for (long j = 0; j < i; ++j, ++k) {
            bool e = c->calc->edgeBetween(*c->recs[i], *c->recs[j]);
            edges += e;
            if (edge) edge[k] = e;
            bool pf = prefilter_pass(*c->recs[i], *c->recs[j]);
            if (pre) pre[k] = pf;
            if (pf && (gap || cc || tc || probM || prob0)) {
                int tcv = 0, ccv = 0; double pm = 1.0, p0 = 0.0;
                bool g = c->calc->checkGapsCigar(*c->recs[i], *c->recs[j], pm, p0, ccv, tcv);
                if (gap) gap[k] = g;
                if (cc) cc[k] = ccv;
                if (tc) tc[k] = tcv;
                if (probM) probM[k] = pm;
                if (prob0) prob0[k] = p0;
            }
        }
    }
    return edges;
}

// Times edgeBetween over an explicit pair list on `threads` host threads, one calculator per
// thread. Returns seconds; *edges_out = number of edges found; edge_out (optional) per pair.
double hcref_time_pairs(void* p, const int* pi, const int* pj, long npairs, int threads, long* edges_out,
                        unsigned char* edge_out) {
    RefCtx* c = static_cast<RefCtx*>(p);
    if (threads < 1) threads = 1;
    std::vector<NewEdgeCalculator*> calcs;
    for (int t = 0; t < threads; ++t) calcs.push_back(c->make_calc());
    std::atomic<long> total(0);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
        pool.emplace_back([&, t]() {
            long lo = npairs * t / threads, hi = npairs * (t + 1) / threads, e = 0;
            for (long k = lo; k < hi; ++k) {
                bool r = calcs[t]->edgeBetween(*c->recs[pi[k]], *c->recs[pj[k]]);
                e += r;
                if (edge_out) edge_out[k] = r;
            }
            total += e;
        });
    }
    for (auto& th : pool) th.join();
    auto t1 = std::chrono::steady_clock::now();
    for (auto* q : calcs) delete q;
    if (edges_out) *edges_out = total.load();
    return std::chrono::duration<double>(t1 - t0).count();
}

// One pass of the CLI's main loop body (src/haploclique.cpp:349-372) with CLEVER or BronKerbosch:
// consumes the current records, replaces them by the collector's position-sorted super-reads.
// state[0]=ct (in/out), state[1]=edges (out), state[2]=converged (out); stdev in/out.
// One pass of the CLI main loop (src/haploclique.cpp:349-372) with the given calculator. When `gpu` is set, the reads
// that survive the filter are first handed to GpuEdgeCalculator::scoreBatch (the driver hook of INTEGRATION.md).
static int iterate_impl(RefCtx* c, EdgeCalculator& calc, int use_bk, int filter_singletons, double significance, int* state,
                        double* stdev, const char* logfile, void* gpu) {
    int ct = state[0];
    std::vector<unsigned int> read_clique_counter(c->names.size());
    LogWriter* lw = (logfile && logfile[0]) ? new LogWriter(logfile, read_clique_counter) : nullptr;
    CliqueCollector collector(lw);
#ifdef HC_WITH_GPU_ADAPTER
    // with the GPU adapter the round's cliques are merged in one GPU batch (integration/CliqueCollector_batch.patch) unless
    // HC_GPU_NO_MERGE asks for the reference's own host fold
    std::unique_ptr<GpuSuperReadBuilder> builder;
    if (gpu && !std::getenv("HC_GPU_NO_MERGE")) {
        builder.reset(new GpuSuperReadBuilder(*static_cast<GpuEdgeCalculator*>(gpu)));
        collector.setSuperReadBuilder(builder.get());
    }
#endif
    CliqueFinder* finder;
    if (use_bk) finder = new BronKerbosch(calc, collector, lw);
    else finder = new CLEVER(calc, collector, lw, 0, 0, c->names.size(), filter_singletons != 0);
    finder->initialize();
    if (lw) lw->initialize();
    int size = (int)c->recs.size();
    int edgecounter = 0;
    double sd = *stdev;
    std::vector<AlignmentRecord*> kept;
    for (auto* r : c->recs) {
        bool drop = (ct == 1 && filter_singletons && r->getReadCount() <= 1) ||
                    (ct > 1 && significance != 0.0 && r->getProbability() < 1.0 / size - significance * sd);
        if (drop) delete r; else kept.push_back(r);
    }
    c->recs.clear();
#ifdef HC_WITH_GPU_ADAPTER
    if (gpu) {
        std::vector<const AlignmentRecord*> v(kept.begin(), kept.end());
        static_cast<GpuEdgeCalculator*>(gpu)->scoreBatch(v);
    }
#else
    (void)gpu;
#endif
    for (auto* r : kept) {
        std::unique_ptr<AlignmentRecord> al(r);
        finder->addAlignment(al, edgecounter);
    }
    finder->finish();
    std::deque<AlignmentRecord*>* next = collector.finish();
    if (lw) lw->finish();
    *stdev = setProbabilities(*next);
    for (auto* r : *next) c->recs.push_back(r);
    delete next;
    state[0] = ct + 1;
    state[1] = edgecounter;
    state[2] = finder->hasConverged() ? 1 : 0;
    delete finder;
    delete lw;
    return (int)c->recs.size();
}

int hcref_iterate(void* p, int use_bk, int filter_singletons, double significance, int* state, double* stdev,
                  const char* logfile) {
    RefCtx* c = static_cast<RefCtx*>(p);
    return iterate_impl(c, *c->calc, use_bk, filter_singletons, significance, state, stdev, logfile, nullptr);
}

// The same pass with the product's GpuEdgeCalculator plugged into the reference's finder. -1 = adapter not built
// into this library, -2 = the adapter threw (no device, ...): message in hcref_error.
int hcref_iterate_gpu(void* p, int use_bk, int filter_singletons, double significance, int* state, double* stdev,
                      const char* logfile, long* gpu_edges) {
#ifdef HC_WITH_GPU_ADAPTER
    RefCtx* c = static_cast<RefCtx*>(p);
    try {
        GpuEdgeCalculator calc(c->Q, c->cut_cliques, c->ov_cliques, c->frameshift, c->simpson, c->cut_single, c->ov_single,
                               c->cut_mixed, c->max_position, c->no_prob0);
        const int n = iterate_impl(c, calc, use_bk, filter_singletons, significance, state, stdev, logfile, &calc);
        if (gpu_edges) *gpu_edges = calc.edges();
        return n;
    } catch (const std::exception& e) {
        c->err = e.what();
        return -2;
    }
#else
    (void)p; (void)use_bk; (void)filter_singletons; (void)significance; (void)state; (void)stdev; (void)logfile; (void)gpu_edges;
    return -1;
#endif
}

// edgeBetween through the adapter for two stored records (the reference's unit-test style call). -1 on error.
int hcref_edge_gpu(void* p, int i, int j) {
#ifdef HC_WITH_GPU_ADAPTER
    RefCtx* c = static_cast<RefCtx*>(p);
    try {
        GpuEdgeCalculator calc(c->Q, c->cut_cliques, c->ov_cliques, c->frameshift, c->simpson, c->cut_single, c->ov_single,
                               c->cut_mixed, c->max_position, c->no_prob0);
        return calc.edgeBetween(*c->recs[i], *c->recs[j]) ? 1 : 0;
    } catch (const std::exception& e) {
        c->err = e.what();
        return -1;
    }
#else
    (void)p; (void)i; (void)j;
    return -1;
#endif
}

// QuasispeciesEdgeCalculator through the product's binding (integration/: GpuQuasispeciesEdgeCalculator): all stored
// records scored as one batch, then edgeBetween(records[pi], records[pj]) per pair. Returns the edge count, -1 / -2 as above.
long hcref_quasi_gpu_pairs(void* p, const int* pi, const int* pj, long npairs, double cut_cliques, double ov_cliques,
                           int frameshift, const int* simpson_pos, const double* simpson_val, int n_simpson, double cut_single,
                           double ov_single, double cut_mixed, unsigned char* edge_out) {
#ifdef HC_WITH_GPU_ADAPTER
    RefCtx* c = static_cast<RefCtx*>(p);
    try {
        std::unordered_map<int, double> simpson;
        for (int i = 0; i < n_simpson; ++i) simpson[simpson_pos[i]] = simpson_val[i];
        GpuQuasispeciesEdgeCalculator calc(0.9, cut_cliques, ov_cliques, frameshift != 0, simpson, cut_single, ov_single, cut_mixed,
                                           c->max_position);
        std::vector<const AlignmentRecord*> all(c->recs.begin(), c->recs.end());
        calc.scoreBatch(all);
        long edges = 0;
        for (long k = 0; k < npairs; ++k) {
            const bool e = calc.edgeBetween(*c->recs[pi[k]], *c->recs[pj[k]]);
            if (edge_out) edge_out[k] = e;
            edges += e;
        }
        return edges;
    } catch (const std::exception& e) {
        c->err = e.what();
        return -2;
    }
#else
    (void)p; (void)pi; (void)pj; (void)npairs; (void)cut_cliques; (void)ov_cliques; (void)frameshift; (void)simpson_pos;
    (void)simpson_val; (void)n_simpson; (void)cut_single; (void)ov_single; (void)cut_mixed; (void)edge_out;
    return -1;
#endif
}


// ---- the product's direct BAM -> SoA loader (integration/GpuBamLoader + hc_loader_*) for the N3 parity tests --------------
// Returns a handle (free with hcref_direct_free) or null; *n / *max_position / counts[3] = alignments, skipped, merged mates.
void* hcref_direct_load(const char* path, int* n, unsigned* max_position, long long* counts, double* seconds) {
#ifdef HC_WITH_GPU_ADAPTER
    try {
        const auto t0 = std::chrono::steady_clock::now();
        GpuBamLoader* L = new GpuBamLoader();
        L->load(path);
        if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (n) *n = L->reads().n;
        if (max_position) *max_position = L->maxPosition();
        if (counts) { int64_t a = 0, b = 0, c = 0; hc_loader_counts(L->handle(), &a, &b, &c); counts[0] = a; counts[1] = b; counts[2] = c; }
        return L;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "hcref_direct_load: %s\n", e.what());
        return nullptr;
    }
#else
    (void)path; (void)n; (void)max_position; (void)counts; (void)seconds;
    return nullptr;
#endif
}
// the hc_reads arrays of a loaded file (pointers into the handle) and the name of arrival r
const void* hcref_direct_reads(void* h) {
#ifdef HC_WITH_GPU_ADAPTER
    return &static_cast<GpuBamLoader*>(h)->reads();
#else
    (void)h; return nullptr;
#endif
}
int hcref_direct_name(void* h, int read, char* out, int cap) {
#ifdef HC_WITH_GPU_ADAPTER
    std::snprintf(out, cap, "%s", static_cast<GpuBamLoader*>(h)->recordName(read).c_str());
    return 0;
#else
    (void)h; (void)read; (void)out; (void)cap; return -1;
#endif
}
void hcref_direct_free(void* h) {
#ifdef HC_WITH_GPU_ADAPTER
    delete static_cast<GpuBamLoader*>(h);
#else
    (void)h;
#endif
}
// seconds the reference's own readBamFile needs for the same file (records freed afterwards)
double hcref_time_read_bam(const char* path, int* n) {
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::string> names;
    unsigned int mp = 0;
    BamTools::SamHeader header;
    BamTools::RefVector refs;
    std::deque<AlignmentRecord*>* reads = nullptr;
    {
        std::streambuf* old = std::cout.rdbuf(nullptr);        // "Read BamFile: done"
        try { reads = readBamFile(path, names, mp, header, refs); } catch (...) { std::cout.rdbuf(old); return -1.0; }
        std::cout.rdbuf(old);
    }
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (n) *n = (int)reads->size();
    for (auto* r : *reads) delete r;
    delete reads;
    return s;
}

}  // extern "C"
