// Host side of libhcgpu.so: AlignmentRecord fields -> packed SoA (headers + 16-bit op records).
// Reference for the fields: src/AlignmentRecord.h:54-77; clip-extended bounds: computeOffset /
// computeRevOffset (src/AlignmentRecord.cpp:126-156) as used at src/NewEdgeCalculator.cpp:413-421.
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <new>

#include <cuda_runtime_api.h>

#include <cstdlib>

#include "hc_internal.h"

namespace hc {

// Released page-locked buffers are kept (a few, bounded) and handed to the next alloc of a similar size: the driver
// packs one batch per iteration, and cudaHostAlloc of ~100 MB costs more than packing it.
namespace {
struct PoolSlot { void* p; size_t cap; };
std::mutex g_pool_mu;
std::vector<PoolSlot> g_pool;
size_t g_pool_bytes = 0;
constexpr size_t kPoolMaxSlots = 6;
constexpr size_t kPoolMaxBytes = size_t(2) << 30;
}  // namespace

bool HostBuf::alloc(size_t bytes) {
    release();
    if (bytes == 0) bytes = 64;
    static int have_gpu = -1;
    if (have_gpu < 0) {
        int n = 0;
        have_gpu = (cudaGetDeviceCount(&n) == cudaSuccess && n > 0) ? 1 : 0;
        if (!have_gpu) cudaGetLastError();
    }
    if (have_gpu) {
        {
            std::lock_guard<std::mutex> lk(g_pool_mu);
            int best = -1;
            for (int k = 0; k < (int)g_pool.size(); ++k)
                if (g_pool[k].cap >= bytes && g_pool[k].cap <= 2 * bytes + (size_t(1) << 20) && (best < 0 || g_pool[k].cap < g_pool[best].cap)) best = k;
            if (best >= 0) {
                p = g_pool[best].p; cap = g_pool[best].cap; pinned = true;
                g_pool_bytes -= cap;
                g_pool.erase(g_pool.begin() + best);
                return true;
            }
        }
        const size_t want = bytes + bytes / 8;                 // head room so that the next, slightly larger batch fits
        if (cudaHostAlloc(&p, want, cudaHostAllocDefault) == cudaSuccess) { pinned = true; cap = want; return true; }
        cudaGetLastError();
    }
    p = std::malloc(bytes);
    pinned = false;
    cap = bytes;
    return p != nullptr;
}

void HostBuf::release() {
    if (!p) return;
    if (pinned) {
        bool kept = false;
        {
            std::lock_guard<std::mutex> lk(g_pool_mu);
            if (g_pool.size() < kPoolMaxSlots && g_pool_bytes + cap <= kPoolMaxBytes) {
                g_pool.push_back({p, cap});
                g_pool_bytes += cap;
                kept = true;
            }
        }
        if (!kept) cudaFreeHost(p);
    } else {
        std::free(p);
    }
    p = nullptr;
    cap = 0;
}

static std::mutex g_err_mu;
static std::string g_err;

void set_global_error(const std::string& msg) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_err = msg;
}
const char* global_error() { return g_err.c_str(); }

void build_match_tables(double* eq, double* uneq) {
    std::memset(eq, 0, sizeof(double) * 127 * 127);
    std::memset(uneq, 0, sizeof(double) * 127 * 127);
    float e[127], e3[127];
    for (int c = 33; c < 127; ++c) {
        // error probability of quality char c: 10^(-(c-33)/10) evaluated in double, kept as float
        e[c] = static_cast<float>(std::pow(10.0, static_cast<double>(-(c - 33)) / 10.0));
        e3[c] = e[c] / 3;                        // float division, as the reference's e/3
    }
    for (int a = 33; a < 127; ++a) {
        const double ok_a = 1.0 - e[a];
        for (int b = 33; b < 127; ++b) {
            const double ok_b = 1.0 - e[b];
            const float both_wrong_same = e[a] * e3[b];          // float * float
            const float both_wrong_diff = 2 * e3[a] * e3[b];     // (2*float)*float in float
            eq[a * 127 + b] = ok_a * ok_b + both_wrong_same;
            uneq[a * 127 + b] = ok_a * e3[b] + ok_b * e3[a] + both_wrong_diff;
        }
    }
}

void build_phred_tables(uint8_t* tab) {
    constexpr int NQ = 94;
    float p[NQ];
    for (int q = 0; q < NQ; ++q) p[q] = (float)std::pow(10.0, (double)((float)(-q) / 10));
    for (int i = 0; i < NQ; ++i)
        for (int j = 0; j < NQ; ++j) {
            const float p1 = p[i], p2 = p[j];
            float agree = (p1 * p2 / 3) / (1 - p1 - p2 + 4 * p1 * p2 / 3);
            float dis = (p1 * (1 - p2 / 3)) / (p1 + p2 - 4 * p1 * p2 / 3);
            const int va = (int)std::round(-10 * std::log10(agree));
            const int vd = (int)std::round(-10 * std::log10(dis));
            tab[i * NQ + j] = (uint8_t)(char)std::min(va + 33, 126);
            tab[NQ * NQ + i * NQ + j] = (uint8_t)(char)std::min(vd + 33, 126);
        }
}

void build_prob0_ladder(std::vector<double>& F, int n_max) {
    F.assign(static_cast<size_t>(n_max) + 1, 0.0);
    const double step = std::log10(0.25);
    double acc = 0.0;
    for (int k = 1; k <= n_max; ++k) {
        acc += step;
        F[k] = acc;
    }
}

}  // namespace hc

namespace {

inline int op_code(uint8_t c) {
    switch (c) {
        case 'M': return HC_OP_M;
        case 'I': return HC_OP_I;
        case 'D': return HC_OP_D;
        case 'S': return HC_OP_S;
        case 'H': return HC_OP_H;
        default: return -1;
    }
}

inline int base_code(uint8_t c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return 4;                       // ShortDnaSequence maps everything else to 'N'
    }
}

}  // namespace

extern "C" {

int hc_abi_version(void) { return HC_ABI_VERSION; }

int hc_pack(const hc_reads* rd, hc_packed** out) {
    if (!rd || !out || rd->n < 0) {
        hc::set_global_error("hc_pack: null argument or negative read count");
        return HC_E_INVALID;
    }
    *out = nullptr;
    const int n = rd->n;
    hc_packed* p = nullptr;
    hc::PackedOwner* own = nullptr;
    try {
        p = new hc_packed();
        own = new hc::PackedOwner();
    } catch (const std::bad_alloc&) {
        delete p;
        hc::set_global_error("hc_pack: out of memory");
        return HC_E_NOMEM;
    }
    std::memset(p, 0, sizeof(*p));
    p->owner = own;

    // pass 1: which quality chars occur -> dense codes; where each read's op stream starts.
    // Every read's op stream starts on a 16-byte boundary (8 records) so that one TMA bulk copy moves it.
    // Both passes run on up to 16 host threads (reads are independent once the start offsets are known).
    std::vector<int64_t> start(static_cast<size_t>(n) + 1, 0);
    for (int r = 0; r < n; ++r) {
        int64_t ops = 0;
        for (int s = 0; s < (rd->single_end[r] ? 1 : 2); ++s) {
            const int k = 2 * r + s;
            if (rd->cig_len[k] < 0 || rd->cig_len[k] > 65534) {
                hc::set_global_error("hc_pack: a segment's unrolled CIGAR is longer than 65534 ops");
                hc_packed_free(p);
                return HC_E_INVALID;
            }
            ops += rd->cig_len[k];
        }
        start[r + 1] = (start[r] + ops + 7) & ~int64_t(7);
    }
    const int64_t total_ops = start[n];
    std::mutex mu;
    bool seen[256] = {false};
    hc::parallel_ranges(n, 2048, [&](int r0, int r1) {
        bool mine[256] = {false};
        for (int r = r0; r < r1; ++r)
            for (int s = 0; s < (rd->single_end[r] ? 1 : 2); ++s) {
                const int k = 2 * r + s;
                const uint8_t* q = rd->qual + rd->seq_off[k];
                for (int t = 0; t < rd->seq_len[k]; ++t) mine[q[t]] = true;
            }
        std::lock_guard<std::mutex> lk(mu);
        for (int c = 0; c < 256; ++c) seen[c] |= mine[c];
    });
    int code_of[256];
    int K = 0;
    for (int c = 0; c < 256; ++c) {
        code_of[c] = -1;
        if (!seen[c]) continue;
        if (c < 33 || c > 126) {
            hc::set_global_error("hc_pack: quality character outside 33..126 (the reference's tables are undefined there)");
            hc_packed_free(p);
            return HC_E_INVALID;
        }
        p->qual_chars[K] = static_cast<uint8_t>(c);
        code_of[c] = K++;
    }
    struct Lut { uint16_t op[256]; uint16_t base[256]; uint8_t cons[256]; uint8_t bad[256]; } lut;
    uint16_t qual_lut[256];
    for (int c = 0; c < 256; ++c) {
        const int oc = op_code(static_cast<uint8_t>(c));
        lut.bad[c] = oc < 0;
        lut.op[c] = static_cast<uint16_t>((oc < 0 ? 0 : oc) << HC_OP_SHIFT);
        lut.cons[c] = (oc == HC_OP_M || oc == HC_OP_I || oc == HC_OP_S) ? 1 : 0;
        lut.base[c] = static_cast<uint16_t>(base_code(static_cast<uint8_t>(c)) << HC_BASE_SHIFT);
        qual_lut[c] = static_cast<uint16_t>((code_of[c] < 0 ? 0 : code_of[c]) << HC_QUAL_SHIFT);
    }
    if (total_ops >= (int64_t(1) << 32) - 64) {
        hc::set_global_error("hc_pack: more than 2^32 CIGAR ops in one batch; split the batch");
        hc_packed_free(p);
        return HC_E_INVALID;
    }
    if (!own->hdr.alloc(sizeof(hc_read_hdr) * static_cast<size_t>(n)) ||
        !own->ops.alloc(sizeof(uint16_t) * (static_cast<size_t>(total_ops) + 64))) {   // padding: kernels may over-read
        hc_packed_free(p);
        hc::set_global_error("hc_pack: out of memory");
        return HC_E_NOMEM;
    }
    hc_read_hdr* hdr_out = static_cast<hc_read_hdr*>(own->hdr.p);
    uint16_t* ops_out = static_cast<uint16_t*>(own->ops.p);
    std::memset(ops_out + total_ops, 0, sizeof(uint16_t) * 64);

    // pass 2: headers and op records
    int max_ops = 0;
    std::string error;
    hc::parallel_ranges(n, 2048, [&](int r0, int r1) {
        int my_max = 0;
        std::string my_error;
        for (int r = r0; r < r1 && my_error.empty(); ++r) {
            int64_t w = start[r];
            hc_read_hdr& h = hdr_out[r];
            std::memset(&h, 0, sizeof(h));
            const bool single = rd->single_end[r] != 0;
            h.ncov_flags = (static_cast<uint32_t>(rd->ncov[r]) & 0x3fffffffu) | (single ? HC_FLAG_SINGLE_END : 0u) |
                           (rd->is_clique[r] ? HC_FLAG_IS_CLIQUE : 0u);
            h.cov_first = rd->cov_first[r];
            h.cov_last = rd->cov_last[r];
            int read_ops = 0;
            for (int s = 0; s < (single ? 1 : 2) && my_error.empty(); ++s) {
                const int k = 2 * r + s;
                const uint8_t* cg = rd->cig + rd->cig_off[k];
                const uint8_t* sq = rd->seq + rd->seq_off[k];
                const uint8_t* ql = rd->qual + rd->seq_off[k];
                const int nc = rd->cig_len[k], nq = rd->seq_len[k];
                int lead = 0, trail = 0;
                while (lead < nc && (cg[lead] == 'S' || cg[lead] == 'H')) ++lead;
                while (trail < nc && (cg[nc - 1 - trail] == 'S' || cg[nc - 1 - trail] == 'H')) ++trail;
                h.raw[2 * s] = rd->seg_start[k];
                h.raw[2 * s + 1] = rd->seg_end[k];
                h.ext[2 * s] = rd->seg_start[k] - lead;
                h.ext[2 * s + 1] = rd->seg_end[k] + trail;
                h.ops_off[s] = static_cast<uint32_t>(w);
                h.n_ops[s] = nc;
                // branch-free record build from look-up tables; the usual segment (clips, M only, clips) takes the
                // middle run without looking at the CIGAR again
                int ncons = 0;
                unsigned bad = 0, ins_acc = 0, nonm = 0;
                for (int t = 0; t < nc; ++t) { ncons += lut.cons[cg[t]]; bad |= lut.bad[cg[t]]; }
                if (bad) {
                    for (int t = 0; t < nc; ++t)
                        if (lut.bad[cg[t]]) {
                            my_error = std::string("hc_pack: CIGAR op '") + static_cast<char>(cg[t]) +
                                       "' is outside MIDSH (the reference's walk does not define it)";
                            break;
                        }
                    break;
                }
                if (ncons > nq) { my_error = "hc_pack: CIGAR consumes more query bases than the read has"; break; }
                uint16_t* dst = ops_out + static_cast<size_t>(w);
                int qi = 0;
                const int mid0 = std::min(lead, nc), mid1 = std::max(mid0, nc - trail);
                auto generic_run = [&](int t0, int t1, bool count_nonm) {
                    for (int t = t0; t < t1; ++t) {
                        const uint8_t c = cg[t];
                        const unsigned k = lut.cons[c];
                        const int qc = qi < nq ? qi : (nq > 0 ? nq - 1 : 0);
                        const uint16_t bq = nq > 0 ? static_cast<uint16_t>(lut.base[sq[qc]] | qual_lut[ql[qc]]) : uint16_t(0);
                        dst[t] = static_cast<uint16_t>(lut.op[c] | (bq & static_cast<uint16_t>(0u - k)));
                        qi += (int)k;
                        ins_acc |= (c == 'I');
                        if (count_nonm) nonm |= (c != 'M');
                    }
                };
                generic_run(0, mid0, false);
                bool all_m = true;
                for (int t = mid0; t < mid1; ++t) all_m &= (cg[t] == 'M');
                if (all_m) {
                    const uint16_t opm = static_cast<uint16_t>(HC_OP_M << HC_OP_SHIFT);
                    const int d = qi - mid0;
                    for (int t = mid0; t < mid1; ++t) dst[t] = static_cast<uint16_t>(opm | lut.base[sq[t + d]] | qual_lut[ql[t + d]]);
                    qi += mid1 - mid0;
                } else {
                    generic_run(mid0, mid1, true);
                }
                generic_run(mid1, nc, false);
                w += nc;
                const bool ins = ins_acc != 0;
                bool plain = nonm == 0;
                if (lead + trail > nc) plain = false;             // a segment made of clips only
                h.seg_flags[s] = static_cast<uint16_t>((ins ? HC_SEG_HAS_INS : 0) | (plain ? 0 : HC_SEG_NOT_PLAIN));
                read_ops += nc;
            }
            if (single) h.ops_off[1] = static_cast<uint32_t>(w);
            while (w < start[r + 1]) ops_out[static_cast<size_t>(w++)] = 0;      // the alignment padding behind the read
            if (read_ops > my_max) my_max = read_ops;
        }
        std::lock_guard<std::mutex> lk(mu);
        if (my_max > max_ops) max_ops = my_max;
        if (!my_error.empty() && error.empty()) error = my_error;
    });
    if (!error.empty()) {
        hc::set_global_error(error);
        hc_packed_free(p);
        return HC_E_INVALID;
    }
    p->n_reads = n;
    p->n_qual_codes = K;
    p->hdr = hdr_out;
    p->ops = ops_out;
    p->n_ops_total = total_ops;
    p->max_ops_per_read = max_ops;
    *out = p;
    return HC_OK;
}

void hc_packed_free(hc_packed* p) {
    if (!p) return;
    delete static_cast<hc::PackedOwner*>(p->owner);
    delete p;
}

// Rows -> parts for several GPUs: cuts at equal shares of the scoring work (scored pairs, weighted by the
// columns they share); a part's
// halo starts at the smallest band start (col0) of its rows. col0 exactly as compute_layout() in hc_gpu.cu defines it.
int hc_partition_rows(const int32_t* cov_first, const int32_t* cov_last, int32_t n, int32_t n_parts, int32_t* cuts, int32_t* first) {
    if (!cov_first || !cov_last || !cuts || !first || n < 0 || n_parts < 1) {
        hc::set_global_error("hc_partition_rows: bad argument");
        return HC_E_INVALID;
    }
    std::vector<int32_t> pmax((size_t)n), col0((size_t)n);
    int32_t run = INT32_MIN;
    for (int r = 0; r < n; ++r) { run = std::max(run, cov_last[r]); pmax[r] = run; }
    std::vector<int64_t> acc((size_t)n + 1, 0);
    hc::parallel_ranges(n, 4096, [&](int k0, int k1) {
        for (int i = k0; i < k1; ++i)
            col0[i] = (int32_t)(std::lower_bound(pmax.begin(), pmax.begin() + i, cov_first[i]) - pmax.begin()) & ~31;
    });
    // Weight of row i = what it costs to score: a per-pair part and a per-column part. Pairs: the earlier reads whose covered
    // interval reaches cov_first[i] (the band width i - col0[i] overstates them: one long earlier read widens the band of every
    // later row). Columns: the length of the intersection of the covered spans, min(last_i, last_j) - first_i + 1, summed over
    // those reads. Both come from two Fenwick trees (count, sum of cov_last) over the ranks of cov_last. The ratio -- one pair
    // costs as much as kColumnsPerPair shared columns -- is measured (C3 on B200: 5.3e-8 ms per pair, 2.5e-10 ms per column).
    {
        constexpr int64_t kColumnsPerPair = 215;
        std::vector<int32_t> lasts(cov_last, cov_last + n);
        std::sort(lasts.begin(), lasts.end());
        lasts.erase(std::unique(lasts.begin(), lasts.end()), lasts.end());
        const int m = (int)lasts.size();
        std::vector<int32_t> cnt((size_t)m + 1, 0);
        std::vector<int64_t> sum((size_t)m + 1, 0);
        auto prefix = [&](int r, int64_t& c, int64_t& sm) {          // over ranks 1..r
            c = 0; sm = 0;
            for (int k = r; k > 0; k -= k & -k) { c += cnt[k]; sm += sum[k]; }
        };
        for (int i = 0; i < n; ++i) {
            const int rf = (int)(std::lower_bound(lasts.begin(), lasts.end(), cov_first[i]) - lasts.begin());   // lasts[0..rf) < first_i
            const int rl = (int)(std::lower_bound(lasts.begin(), lasts.end(), cov_last[i]) - lasts.begin());    // lasts[0..rl) < last_i
            int64_t c_f, s_f, c_l, s_l;
            prefix(rf, c_f, s_f);
            prefix(std::max(rl, rf), c_l, s_l);
            const int64_t scored = i - c_f;                           // cov_last[j] >= first_i
            const int64_t inside = c_l - c_f, inside_sum = s_l - s_f; // first_i <= cov_last[j] < last_i
            const int64_t beyond = scored - inside;                   // cov_last[j] >= last_i
            const int64_t span = std::max<int64_t>(0, (int64_t)cov_last[i] - cov_first[i] + 1);
            const int64_t columns = inside_sum - inside * ((int64_t)cov_first[i] - 1) + beyond * span;
            acc[(size_t)i + 1] = acc[i] + kColumnsPerPair * scored + std::max<int64_t>(columns, 0);
            for (int k = rl + 1; k <= m; k += k & -k) { cnt[k]++; sum[k] += cov_last[i]; }
        }
    }
    cuts[0] = 0;
    for (int p = 1; p < n_parts; ++p) {
        const int64_t want = acc[n] / n_parts * p + acc[n] % n_parts * p / n_parts;
        int32_t c = (int32_t)(std::lower_bound(acc.begin(), acc.end(), want) - acc.begin());
        cuts[p] = std::min(std::max(c, cuts[p - 1]), n);
    }
    cuts[n_parts] = n;
    for (int p = 0; p < n_parts; ++p) {
        int32_t f = cuts[p];
        for (int i = cuts[p]; i < cuts[p + 1]; ++i) f = std::min(f, col0[i]);
        first[p] = f & ~31;
    }
    return HC_OK;
}

int hc_packed_slice(const hc_packed* p, int32_t first, int32_t end, hc_packed** out) {
    if (!p || !out || first < 0 || end < first || end > p->n_reads) {
        hc::set_global_error("hc_packed_slice: bad range (0 <= first <= end <= n_reads)");
        return HC_E_INVALID;
    }
    *out = nullptr;
    hc_packed* q = nullptr;
    hc::PackedOwner* own = nullptr;
    try {
        q = new hc_packed(*p);
        own = new hc::PackedOwner();
    } catch (const std::bad_alloc&) {
        delete q;
        hc::set_global_error("hc_packed_slice: out of memory");
        return HC_E_NOMEM;
    }
    q->owner = own;
    const int n = end - first;
    if (!own->hdr.alloc(sizeof(hc_read_hdr) * (size_t)std::max(n, 1))) {
        hc_packed_free(q);
        hc::set_global_error("hc_packed_slice: out of memory");
        return HC_E_NOMEM;
    }
    hc_read_hdr* h = static_cast<hc_read_hdr*>(own->hdr.p);
    // a read's op stream starts on a 16-byte boundary, so the slice's stream does too
    const uint32_t base = n > 0 ? p->hdr[first].ops_off[0] : 0u;
    const int64_t stop = end < p->n_reads ? (int64_t)p->hdr[end].ops_off[0] : p->n_ops_total;
    hc::parallel_ranges(n, 8192, [&](int k0, int k1) {
        for (int k = k0; k < k1; ++k) {
            h[k] = p->hdr[first + k];
            h[k].ops_off[0] -= base;
            h[k].ops_off[1] -= base;
        }
    });
    q->n_reads = n;
    q->hdr = h;
    q->ops = p->ops + base;                 // behind the slice: the next reads' records or the parent's padding (>= 64 records)
    q->n_ops_total = n > 0 ? stop - (int64_t)base : 0;
    *out = q;
    return HC_OK;
}

}  // extern "C"
