// GpuEdgeCalculator -- the reference-side binding of libhcgpu.so.
//
// A drop-in for the reference's NewEdgeCalculator behind its own plugin interface
// (class EdgeCalculator, src/EdgeCalculator.h:24-37). Compiled INSIDE the reference tree (it includes the
// reference's headers); nothing of the reference is copied here. Same constructor arguments as
// NewEdgeCalculator (src/NewEdgeCalculator.h:58) plus the CUDA device ordinal -- or a list of ordinals: the rows of a
// batch are then cut across the devices by the library (hc_multi_*: balanced by band pairs, halo from the covered
// intervals, every device copying its rows' words into this object's one adjacency buffer; no collective).
//
// Use (INTEGRATION.md shows the four-line change to src/haploclique.cpp):
//     GpuEdgeCalculator* calc = new GpuEdgeCalculator(Q, cutoff, overlap, fs_merge, simpson_map, cutoff_single,
//                                                      overlap_single, cutoff_mixed, max_position, no_prob0);
//     ... once per iteration, with the reads that will be handed to the clique finder, in that order:
//     calc->scoreBatch(reads_of_this_iteration);
//     ... CLEVER / BronKerbosch unchanged: their edgeBetween(new, old) calls are answered from the adjacency words.
//
// There is no CPU scoring path: every answer comes from the CUDA kernels; if the library cannot reach a device the
// constructor throws std::runtime_error.
#ifndef GPUEDGECALCULATOR_H_
#define GPUEDGECALCULATOR_H_

#include <deque>
#include <string>
#include <unordered_map>
#include <vector>

#include "AlignmentRecord.h"
#include "EdgeCalculator.h"

extern "C" {
#include "hc_edge.h"
}

class GpuEdgeCalculator : public EdgeCalculator {
public:
    GpuEdgeCalculator(double Q, double edge_quasi_cutoff, double overlap, bool frameshift_merge,
                      std::unordered_map<int, double>& simpson_map, double edge_quasi_cutoff_single, double overlap_single,
                      double edge_quasi_cutoff_mixed, unsigned int maxPosition, bool noProb0, int device = 0);
    GpuEdgeCalculator(double Q, double edge_quasi_cutoff, double overlap, bool frameshift_merge,
                      std::unordered_map<int, double>& simpson_map, double edge_quasi_cutoff_single, double overlap_single,
                      double edge_quasi_cutoff_mixed, unsigned int maxPosition, bool noProb0, const std::vector<int>& devices);
    virtual ~GpuEdgeCalculator();

    /** Packs the reads (arrival order = the order they will be added to the clique finder), uploads them and scores
     *  the whole band on the GPU. After this, edgeBetween(a, b) for two reads of the batch is a bit lookup. */
    void scoreBatch(const std::vector<const AlignmentRecord*>& reads);
    void scoreBatch(const std::deque<AlignmentRecord*>& reads);

    /** edgeBetween(new arrival, earlier arrival), as both clique finders call it (src/CLEVER.cpp:220,
     *  src/BronKerbosch.cpp:286). Any other two records (the reference's unit tests call it that way; two reads of the
     *  batch in the other order) are packed as a two-read batch and scored on a small second context, so the scored
     *  batch stays resident. */
    virtual bool edgeBetween(const AlignmentRecord& ap1, const AlignmentRecord& ap2) const;
    /** The earlier arrivals of the scored batch that are adjacent to `ap`, in arrival order, read straight from the set
     *  bits of its adjacency row: the row-consuming alternative to one edgeBetween call per active read in the finders'
     *  edge-feeding loops (src/CLEVER.cpp:218-232, src/BronKerbosch.cpp:286-292). Same answers as edgeBetween. */
    void neighbours(const AlignmentRecord& ap, std::vector<const AlignmentRecord*>& out) const;
    /** Row feeding for the finders' edge loops (integration/CLEVER_rows.patch, BronKerbosch_rows.patch): the arrival indices
     *  of the earlier reads adjacent to `ap`, ascending, provided `ap` is read number `id` of the scored batch (the finder's
     *  id == position in the batch when the batch is added in order after initialize()); false otherwise and the finder
     *  falls back to one edgeBetween call per stored alignment. No record is dereferenced: the finder may have freed
     *  earlier ones. */
    bool neighbourIndices(const AlignmentRecord& ap, int id, std::vector<int>& out) const;
    bool rowFeeding() const { return row_feeding_; }
    void setRowFeeding(bool on) { row_feeding_ = on; }
    /** Scratch the patched CLEVER keeps with the calculator (storage slot per id), so that class keeps its layout. */
    std::vector<size_t>& finderSlots() const { return finder_slots_; }
    /** The scored batch as the flat arrays hc_pack was given (valid until the next scoreBatch): the read set of this round
     *  for the GPU super-read merge (GpuSuperReadBuilder). */
    hc_reads batchView() const;
    /** Not implemented by NewEdgeCalculator either (src/NewEdgeCalculator.cpp:1378-1380). */
    virtual void getPartnerLengthRange(const AlignmentRecord& ap, unsigned int* min, unsigned int* max) const;
    virtual void setOverlapCliques(double d);

    long edges() const { return stats_.edges; }
    long scoredPairs() const { return stats_.pairs_scored; }
    /** Pairs whose score lay within 1e-12 (relative) of the threshold: the library re-decided them with std::log10 as
     *  similarityCriterion does (src/NewEdgeCalculator.cpp:1127-1140); this is the report of those pairs. */
    long nearThresholdPairs() const { return stats_.ambiguous; }
    const std::vector<hc_ambiguous>& nearThresholdReport() const { return near_; }
    /** Arrival index of a record in the scored batch (-1: not in it) and the record at an index. */
    int indexOf(const AlignmentRecord& ap) const;
    const AlignmentRecord* recordAt(int i) const { return reads_[(size_t)i]; }
    int batchSize() const { return (int)reads_.size(); }
    /** Row i of the banded adjacency: its words cover the earlier reads from firstColumn(i) on. */
    int firstColumn(int i) const { return col0_[(size_t)i]; }
    const uint32_t* rowWords(int i, int* n_words) const {
        *n_words = (int)(word_off_[(size_t)i + 1] - word_off_[(size_t)i]);
        return words_.data() + word_off_[(size_t)i];
    }
    double kernelMilliseconds() const { return stats_.kernel_ms; }

protected:
    /** The scoring calls behind scoreBatch / the out-of-batch edgeBetween; GpuQuasispeciesEdgeCalculator swaps them. */
    virtual int scoreResident(uint32_t* words, int64_t n_words, hc_band_stats* stats);
    virtual int scorePairOn(hc_ctx* ctx, int i, int j, bool* edge) const;
    hc_multi* handle() const { return multi_; }
    std::vector<hc_ambiguous> near_;

private:
    struct Flat;                                           // SoA staging of AlignmentRecord fields (hc_reads)
    static void flatten(const std::vector<const AlignmentRecord*>& reads, Flat& f);
    void fail(const char* what, int rc) const;

    void init(double edge_quasi_cutoff, double overlap, std::unordered_map<int, double>& simpson_map,
              double edge_quasi_cutoff_single, double overlap_single, double edge_quasi_cutoff_mixed, unsigned int maxPosition,
              bool noProb0, const std::vector<int>& devices);
    hc_params par_;
    std::vector<double> simpson_;
    std::vector<int> devices_;
    hc_multi* multi_;
    mutable hc_ctx* aux_;                                  // two-read batches of out-of-batch queries
    std::unordered_map<const AlignmentRecord*, int> index_;   // record -> arrival index in the scored batch
    std::vector<const AlignmentRecord*> reads_;              // arrival index -> record
    std::vector<int32_t> col0_;
    std::vector<int64_t> word_off_;
    std::vector<uint32_t> words_;
    hc_band_stats stats_;
    hc_packed* resident_;                                  // the scored batch (the devices hold slices of it)
    Flat* flat_;                                           // its flat arrays
    bool row_feeding_;
    mutable std::vector<size_t> finder_slots_;
};

/** The same binding for the reference's older scorer, QuasispeciesEdgeCalculator (src/QuasispeciesEdgeCalculator.h:57,
 *  .cpp:55-70): constructor arguments in the reference's order plus maxPosition (sizes the dense copy of simpson_map,
 *  whose values are frequencies here, not log10) and the device. scoreBatch / edgeBetween / neighbours as above, scored by
 *  hc_quasi_score_band / hc_quasi_score_pairs. The reference's CLI never constructs this class; a maintainer who wants
 *  it back plugs this one in the same way. Pairs on which the reference's walk would not terminate
 *  (src/QuasispeciesEdgeCalculator.cpp:258-267) are no edges. */
class GpuQuasispeciesEdgeCalculator : public GpuEdgeCalculator {
public:
    GpuQuasispeciesEdgeCalculator(double Q, double edge_quasi_cutoff, double overlap, bool frameshift_merge,
                                  std::unordered_map<int, double>& simpson_map, double edge_quasi_cutoff_single,
                                  double overlap_single, double edge_quasi_cutoff_mixed, unsigned int maxPosition, int device = 0);
    virtual bool edgeBetween(const AlignmentRecord& ap1, const AlignmentRecord& ap2) const;
    /** As GpuEdgeCalculator::neighbours, plus the same-name shortcut of edgeBetween (src/QuasispeciesEdgeCalculator.cpp:56)
     *  for batches that hold a name twice. */
    void neighbours(const AlignmentRecord& ap, std::vector<const AlignmentRecord*>& out) const;
    virtual void setOverlapCliques(double d) { qpar_.min_overlap_cliques = d; }

protected:
    virtual int scoreResident(uint32_t* words, int64_t n_words, hc_band_stats* stats);
    virtual int scorePairOn(hc_ctx* ctx, int i, int j, bool* edge) const;

private:
    hc_quasi_params qpar_;
};

#endif /* GPUEDGECALCULATOR_H_ */
