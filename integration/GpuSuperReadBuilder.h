// GpuSuperReadBuilder -- the reference-side binding of the GPU super-read merge (hc_merger_* / hc_merge_*).
//
// The reference turns every clique into a super-read inside CliqueCollector::add (src/CliqueCollector.h:59: new
// AlignmentRecord(alignments, id), src/AlignmentRecord.cpp:211-249), one at a time, folding the members on the host.
// With integration/CliqueCollector_batch.patch the collector keeps the cliques of a finder round and hands them to a
// SuperReadBuilder when finish() is called; this builder merges them in one GPU batch and wraps each result in an
// AlignmentRecord through the constructor the patch adds (everything the clique constructor sets, without the fold).
// The reads are the batch GpuEdgeCalculator::scoreBatch was given for this round: a clique member's id
// (AlignmentRecord::getID, assigned by the finder in arrival order) is its position in that batch.
// A clique on which the reference's own merge is undefined (it would not return) raises std::runtime_error.
#ifndef GPUSUPERREADBUILDER_H_
#define GPUSUPERREADBUILDER_H_

#include "CliqueCollector.h"
#include "GpuEdgeCalculator.h"

class GpuSuperReadBuilder : public SuperReadBuilder {
public:
    explicit GpuSuperReadBuilder(const GpuEdgeCalculator& calc, int device = 0);
    virtual ~GpuSuperReadBuilder();
    virtual void build(const std::vector<std::vector<int> >& members, const std::vector<unsigned int>& clique_ids,
                       const std::vector<std::set<int> >& read_names, std::vector<std::string>* read_name_map,
                       std::vector<AlignmentRecord*>& out);
    double kernelMilliseconds() const { return kernel_ms_; }
    long cliquesMerged() const { return cliques_; }
    long foldsInColumnSpace() const { return folds_plain_; }
    long foldsWalked() const { return folds_walked_; }

private:
    const GpuEdgeCalculator& calc_;
    hc_merger* merger_;
    double kernel_ms_;
    long cliques_, folds_plain_, folds_walked_;
};

#endif /* GPUSUPERREADBUILDER_H_ */
