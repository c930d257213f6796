// haploclique_gpu -- the reference CLI with its edge scoring on the GPU.
//
// The iteration driver of the reference (src/haploclique.cpp:203-419) restated around one extra call per iteration:
// after the filter that decides which reads reach the clique finder (:343-345, :357) and before the addAlignment
// loop (:353-359), the kept reads are handed to GpuEdgeCalculator::scoreBatch, which decides every pair of the
// iteration in one GPU pass. CLEVER / BronKerbosch, CliqueCollector, the super-read merge, setProbabilities and the
// writers are the reference's own, unchanged; their edgeBetween calls are answered from the adjacency words.
// Built inside the reference tree (oracle/Makefile target `gpu`); readBamFile, USAGE and the writers come from the
// reference's translation unit, included whole exactly as its own unit test does (test/unit/unit1.cpp:11-12).
// Not supported here: -I/-M indel calling (GaussianEdgeCalculator; "not supported yet" in the reference's USAGE).
#define main HaploMain
#include "haploclique.cpp"
#undef main

#include <sstream>

#include "GpuEdgeCalculator.h"
#include "GpuSuperReadBuilder.h"

int main(int argc, char* argv[]) {
    map<std::string, docopt::value> args = docopt::docopt(USAGE, {argv + 1, argv + argc}, true, HAPLOCLIQUE_VERSION);
    for (auto elem : args) cout << elem.first << " " << elem.second << "\n";       // the reference echoes its arguments (:210-213)
    const string bamfile = args["<bamfile>"].asString();
    const string outfile = args["<output>"] ? args["<output>"].asString() : string("quasispecies");
    const double cutoff_cliques = stod(args["--edge_quasi_cutoff_cliques"].asString());
    const double cutoff_single = stod(args["--edge_quasi_cutoff_single"].asString());
    const double cutoff_mixed = stod(args["--edge_quasi_cutoff_mixed"].asString());
    const double Q = stod(args["--random_overlap_quality"].asString());
    const double overlap_cliques = stod(args["--min_overlap_cliques"].asString());
    const double overlap_single = stod(args["--min_overlap_single"].asString());
    const bool frameshift_merge = args["--frame_shift_merge"].asBool();
    const int iterations = stoi(args["--iterations"].asString());
    const double filter = stod(args["--filter"].asString());
    const double significance = stod(args["--significance"].asString());
    const bool filter_singletons = args["--no_singletons"].asBool();
    const string logfile = args["--log"] ? args["--log"].asString() : string();
    const int doc_haplotypes = args["--doc_haplotypes"] ? stoi(args["--doc_haplotypes"].asString()) : 0;
    const bool no_prob0 = args["--no_prob0"].asBool();
    const int max_cliques = args["--max_cliques"] ? stoi(args["--max_cliques"].asString()) : 0;
    const int limit_clique_size = args["--limit_clique_size"] ? stoi(args["--limit_clique_size"].asString()) : 0;
    if (args["--call_indels"]) {
        cerr << "haploclique_gpu: -I/--call_indels is not supported by the GPU driver." << endl;
        return 1;
    }

    // allele frequencies -> log10 of the Simpson index per position (src/haploclique.cpp:259-279)
    std::unordered_map<int, double> simpson_map;
    unsigned int max_position2 = 0;
    if (args["--allel_frequencies"]) {
        ifstream in(args["--allel_frequencies"].asString().c_str());
        string line;
        while (getline(in, line)) {
            trim_right(line);
            std::vector<std::string> w;
            boost::split(w, line, boost::is_any_of("\t"), boost::token_compress_on);
            if (w.size() < 5 || w[0].find('.') != string::npos) continue;      // insertion rows ("pos.k") are skipped
            double s = 0.0;
            for (int k = 1; k <= 4; ++k) s += pow(atof(w[k].c_str()), 2);
            simpson_map[atoi(w[0].c_str())] = std::log10(s);
            max_position2 = atoi(w[0].c_str());
        }
    }

    const clock_t clock_start = clock();
    vector<string> original_read_names;
    unsigned int max_position1 = 0;
    BamTools::SamHeader header;
    BamTools::RefVector references;
    deque<AlignmentRecord*>* reads;
    try {
        reads = readBamFile(bamfile, original_read_names, max_position1, header, references);
    } catch (const runtime_error& error) {
        cerr << error.what() << endl;
        return 1;
    } catch (...) {
        cerr << "Unexpected error while reading BamFile." << endl;
        return 1;
    }
    max_position1 = std::max(max_position1, max_position2);

    // HC_GPU_DEVICES=0,1,...: the rows of every iteration's batch are cut across these GPUs by the library (default: GPU 0)
    std::vector<int> devices;
    if (const char* e = std::getenv("HC_GPU_DEVICES")) {
        std::string s(e), tok;
        std::stringstream ss(s);
        while (std::getline(ss, tok, ',')) if (!tok.empty()) devices.push_back(atoi(tok.c_str()));
    }
    if (devices.empty()) devices.push_back(0);
    GpuEdgeCalculator* calc;
    try {
        calc = new GpuEdgeCalculator(Q, cutoff_cliques, overlap_cliques, frameshift_merge, simpson_map, cutoff_single,
                                     overlap_single, cutoff_mixed, max_position1, no_prob0, devices);
    } catch (const runtime_error& error) {
        cerr << error.what() << endl;
        return 1;
    }
    // HC_GPU_NO_ROWS: the finders ask edgeBetween once per stored alignment, as the unpatched reference does (for timing
    // the row-consuming edge loops of integration/*_rows.patch against it)
    if (std::getenv("HC_GPU_NO_ROWS")) calc->setRowFeeding(false);
    std::vector<unsigned int> read_clique_counter(original_read_names.size());
    LogWriter* lw = logfile.empty() ? nullptr : new LogWriter(logfile, read_clique_counter);
    CliqueCollector collector(lw);
    // the cliques of a round become super-reads in one GPU batch when the collector finishes (integration/
    // CliqueCollector_batch.patch); HC_GPU_NO_MERGE keeps the reference's one-by-one host fold (for timing against it)
    GpuSuperReadBuilder* builder = nullptr;
    if (!std::getenv("HC_GPU_NO_MERGE")) {
        try {
            builder = new GpuSuperReadBuilder(*calc, devices[0]);
        } catch (const runtime_error& error) {
            cerr << error.what() << endl;
            return 1;
        }
        collector.setSuperReadBuilder(builder);
    }
    CliqueFinder* finder;
    if (args["bronkerbosch"].asBool()) finder = new BronKerbosch(*calc, collector, lw);
    else finder = new CLEVER(*calc, collector, lw, max_cliques, limit_clique_size, original_read_names.size(), filter_singletons);

    int ct = 0, edgecounter = 0;
    double stdev = 1.0, gpu_ms = 0.0;
    cout << "start: " << original_read_names.size();
    while (ct != iterations) {
        finder->initialize();
        if (lw) lw->initialize();
        const int size = (int)reads->size();
        // the reads this iteration will hand to the finder, in arrival order
        std::vector<const AlignmentRecord*> kept;
        std::vector<AlignmentRecord*> dropped;
        for (AlignmentRecord* r : *reads) {
            const bool drop = (ct == 1 && filter_singletons && r->getReadCount() <= 1) ||
                              (ct > 1 && significance != 0.0 && r->getProbability() < 1.0 / size - significance * stdev);
            if (drop) dropped.push_back(r); else kept.push_back(r);
        }
        calc->scoreBatch(kept);                            // <- the GPU pass: every edge of this iteration
        gpu_ms += calc->kernelMilliseconds();
        if (std::getenv("HC_GPU_CHECK_ROWS")) {            // self-check of the row-consuming hook against the per-pair answers
            long from_rows = 0, from_pairs = 0, mismatches = 0;
            std::vector<const AlignmentRecord*> nb;
            for (size_t i = 0; i < kept.size(); ++i) {
                calc->neighbours(*kept[i], nb);
                from_rows += (long)nb.size();
                size_t k = 0;
                for (size_t j = 0; j < i; ++j) {
                    if (!calc->edgeBetween(*kept[i], *kept[j])) continue;
                    from_pairs++;
                    if (k >= nb.size() || nb[k] != kept[j]) mismatches++;
                    k++;
                }
            }
            cout << "\trows " << (mismatches == 0 && from_rows == from_pairs ? "ok" : "MISMATCH") << ": " << from_rows << " neighbours";
        }
        for (AlignmentRecord* r : dropped) delete r;
        for (const AlignmentRecord* r : kept) {
            unique_ptr<AlignmentRecord> al(const_cast<AlignmentRecord*>(r));
            finder->addAlignment(al, edgecounter);
        }
        cout << "\tedges: " << edgecounter << endl;
        delete reads;
        finder->finish();
        try {
            reads = collector.finish();
        } catch (const runtime_error& error) {
            cerr << error.what() << endl;
            return 1;
        }
        if (lw) lw->finish();
        stdev = setProbabilities(*reads);
        if (finder->hasConverged()) break;
        cout << ct++ << ": " << reads->size();
        edgecounter = 0;
    }

    if (filter > 0.0) {
        auto low = [&](AlignmentRecord* al) { return al->getProbability() < filter; };
        auto new_end = std::remove_if(reads->begin(), reads->end(), low);
        for (auto it = new_end; it != reads->end(); ++it) delete *it;
        reads->erase(new_end, reads->end());
    }
    ofstream os(outfile + ".fasta", std::ofstream::out);
    setProbabilities(*reads);
    printReads(os, *reads, doc_haplotypes);
    if (args["--gff"].asBool()) { ofstream os1(outfile + ".gff", std::ofstream::out); printGFF(os1, *reads); }
    if (args["--bam"].asBool()) { ofstream os2(outfile + ".bam", std::ofstream::out); printBAM(os2, outfile + ".bam", *reads, header, references); }
    cout << "final: " << reads->size() << endl;
    for (auto&& r : *reads) delete r;
    delete reads;
    delete finder;
    delete lw;
    cout.precision(3);
    cout << std::fixed << "time: " << (double)(clock() - clock_start) / CLOCKS_PER_SEC << " (gpu scoring kernels " << gpu_ms << " ms";
    if (builder) cout << ", gpu merge kernels " << builder->kernelMilliseconds() << " ms for " << builder->cliquesMerged() << " cliques, folds in column space "
                      << builder->foldsInColumnSpace() << ", walked " << builder->foldsWalked();
    cout << ")" << endl;
    delete builder;
    delete calc;
    return 0;
}
