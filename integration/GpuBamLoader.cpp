// GpuBamLoader.cpp -- see GpuBamLoader.h. The decode loop of readBamFile (src/haploclique.cpp:159-176) around hc_loader_add.
#include "GpuBamLoader.h"

#include <cstring>
#include <stdexcept>

GpuBamLoader::GpuBamLoader() : loader_(nullptr), max_position_(0) {
    std::memset(&reads_, 0, sizeof reads_);
    if (hc_loader_create(&loader_) != HC_OK) throw std::runtime_error("GpuBamLoader: hc_loader_create failed");
}

GpuBamLoader::~GpuBamLoader() { hc_loader_destroy(loader_); }

void GpuBamLoader::load(const std::string& filename) {
    BamTools::BamReader bamreader;
    if (!bamreader.Open(filename)) throw std::runtime_error("Couldn't open Bamfile.");
    header = bamreader.GetHeader();
    references = bamreader.GetReferenceData();
    BamTools::BamAlignment alignment;
    std::vector<char> ops;
    std::vector<uint32_t> lens;
    while (bamreader.GetNextAlignment(alignment)) {
        ops.clear();
        lens.clear();
        for (const auto& op : alignment.CigarData) { ops.push_back(op.Type); lens.push_back(op.Length); }
        const int rc = hc_loader_add(loader_, alignment.Name.c_str(), alignment.Position, ops.data(), lens.data(), (int32_t)ops.size(),
                                     alignment.QueryBases.data(), alignment.Qualities.data(), (int32_t)alignment.QueryBases.size());
        if (rc != HC_OK) throw std::runtime_error(std::string("GpuBamLoader: ") + hc_loader_last_error(loader_));
    }
    bamreader.Close();
    const int rc = hc_loader_finish(loader_, &reads_, &max_position_);
    if (rc != HC_OK) throw std::runtime_error(hc_loader_last_error(loader_));
}

std::vector<std::string> GpuBamLoader::readNames() const {
    std::vector<std::string> out;
    const int n = hc_loader_n_names(loader_);
    out.reserve((size_t)n);
    for (int k = 0; k < n; ++k) out.push_back(hc_loader_name(loader_, k));
    return out;
}

std::string GpuBamLoader::recordName(int read) const { return hc_loader_name(loader_, hc_loader_record_name(loader_, read)); }
