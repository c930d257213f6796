// GpuBamLoader -- BAM file -> the position-sorted SoA batch, without AlignmentRecord objects.
//
// The reference-side half of hc_loader_* (include/hc_edge.h): the reference's vendored BamTools reader decodes the file,
// every alignment goes straight to the library, which pairs mates, merges overlapping mates and orders the records exactly
// as readBamFile (src/haploclique.cpp:143-201) with AlignmentRecord::pairWith (src/AlignmentRecord.cpp:251-312) does.
// The result is the hc_reads view hc_pack / GpuEdgeCalculator / hc_merger_set_reads take, plus what the driver needs from
// readBamFile's other outputs (read names, max_position, header, references).
#ifndef GPUBAMLOADER_H_
#define GPUBAMLOADER_H_

#include <string>
#include <vector>

#include <api/BamReader.h>

extern "C" {
#include "hc_edge.h"
}

class GpuBamLoader {
public:
    GpuBamLoader();
    ~GpuBamLoader();
    /** Reads the whole file. Throws std::runtime_error with the reference's messages ("Couldn't open Bamfile.",
     *  "No reads could be retrieved from the BamFile. "). */
    void load(const std::string& filename);
    const hc_reads& reads() const { return reads_; }
    unsigned int maxPosition() const { return max_position_; }
    /** readNames as readBamFile fills it; recordName(r) = the name of arrival r. */
    std::vector<std::string> readNames() const;
    std::string recordName(int read) const;
    hc_loader* handle() const { return loader_; }
    BamTools::SamHeader header;
    BamTools::RefVector references;

private:
    hc_loader* loader_;
    hc_reads reads_;
    unsigned int max_position_;
};

#endif /* GPUBAMLOADER_H_ */
