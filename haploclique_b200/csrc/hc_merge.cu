// Super-read construction for single-end cliques on the device (first slice).
//
// What it replaces: AlignmentRecord(alignments, clique_id) (src/AlignmentRecord.cpp:211-249) when every member is
// single-end -- a left fold of mergeAlignmentRecordsSingle(ar,1,1) (:581-734) over the members, each fold a lock-step
// walk of two unrolled CIGARs (noOverlapMerge :1539-1576, overlapMerge :1578-1674) whose shared M / I columns take
// the consensus base and the updated Phred char from computeEntry (:169-184, tables :40-90).
//
// First correct CUDA path: one lane per clique, the fold ping-pongs between the caller-visible output region and a
// scratch region of the same size in HBM, the two 94x94 quality tables sit in shared memory. Cliques are independent,
// so a batch of a finder round's cliques fills the grid; per-column work is byte loads and one table look-up
// (integer / byte work, HBM- and latency-bound). Paired / mixed merges are not built yet: a batch holding a
// paired-end member is refused (HC_E_INVALID), never routed elsewhere.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <string>
#include <vector>

#include "hc_internal.h"

namespace {

constexpr int NQ = 94;   // quality chars 33..126

// Phred update tables with the reference's arithmetic (src/AlignmentRecord.cpp:40-54, 68-86): error probabilities through
// a double pow rounded to float, the posterior in float, log10 and round in float; stored as the final quality char
// min(v + 33, 126) (:173, :177, :180). tab[0] = bases agree, tab[1] = bases differ (first index: the larger quality).
void build_phred_tables(uint8_t* tab) {
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

struct Seg {
    int start, end, n_ops, n_seq;
    const uint8_t *cig, *seq, *qual;
};

struct Out {
    uint8_t *cig, *seq, *qual;
    int n_ops, n_seq;
};

__device__ __forceinline__ bool clip(uint8_t c) { return c == 'S' || c == 'H'; }

__device__ __forceinline__ void emit(Out& o, uint8_t op, uint8_t base, uint8_t qual) {
    o.seq[o.n_seq] = base;
    o.qual[o.n_seq++] = qual;
    o.cig[o.n_ops++] = op;
}

// One column of a stretch only one read covers (noOverlapMerge).
__device__ __forceinline__ int lone_column(const Seg& s, Out& o, int& c, int& q, int& ref) {
    if (c >= s.n_ops) return HC_MERGE_OVERRUN;
    const uint8_t op = s.cig[c];
    if (op == 'M' || op == 'I') {
        emit(o, op, s.seq[q], s.qual[q]);
        ++q;
        ref += (op == 'M');
    } else if (op == 'D') {
        o.cig[o.n_ops++] = 'D';
        ++ref;
    } else if (op == 'S') {
        ++q;
        ++ref;
    } else if (op == 'H') {
        ++ref;
    } else {
        return HC_MERGE_UNDEFINED;
    }
    ++c;
    return HC_MERGE_OK;
}

// One column both reads cover (overlapMerge). Op pairs the reference does not list end the merge as UNDEFINED.
__device__ __forceinline__ int shared_column(const Seg& a, const Seg& b, Out& o, int& c1, int& c2, int& q1, int& q2, int& ref,
                                             const uint8_t* tab) {
    if (c1 >= a.n_ops || c2 >= b.n_ops) return HC_MERGE_OVERRUN;
    const uint8_t x = a.cig[c1], y = b.cig[c2];
    const bool xc = clip(x), yc = clip(y);
    if (x == y && (x == 'M' || x == 'I')) {
        const uint8_t b1 = a.seq[q1], b2 = b.seq[q2];
        const int k1 = a.qual[q1] - 33, k2 = b.qual[q2] - 33;
        uint8_t base, qual;
        const int i1 = max(k1, 0), i2 = max(k2, 0);
        if (b1 == b2) { base = b1; qual = tab[i1 * NQ + i2]; }
        else if (k1 >= k2) { base = b1; qual = tab[NQ * NQ + i1 * NQ + i2]; }
        else { base = b2; qual = tab[NQ * NQ + i2 * NQ + i1]; }
        // computeEntry itself can hand back a char below '!': agreement of Q0 with Q62, 64, 68, 70, 73 rounds to -1..-4
        // (29..32). When such a column is merged again the reference indexes table rows 0..32, which its builders never
        // fill (:68-86 start at 33); they sit in zero-initialised static storage, so the update reads 0 -> '!'.
        if ((k1 | k2) < 0) qual = 33;
        emit(o, x, base, qual);
        ref += (x == 'M');
        ++q1; ++q2; ++c1; ++c2;
    } else if ((x == 'D' || xc) && (y == 'D' || yc)) {
        // D/D, D/clip, clip/clip: a reference column without a merged base; it stays a deletion if either read has one
        if (x == 'D' || y == 'D') o.cig[o.n_ops++] = 'D';
        q1 += (x == 'S');
        q2 += (y == 'S');
        ++ref; ++c1; ++c2;
    } else if (x == 'M' && yc) {
        emit(o, 'M', a.seq[q1], a.qual[q1]);
        ++q1;
        q2 += (y == 'S');
        ++ref; ++c1; ++c2;
    } else if (xc && y == 'M') {
        emit(o, 'M', b.seq[q2], b.qual[q2]);
        ++q2;
        q1 += (x == 'S');
        ++ref; ++c1; ++c2;
    } else if (x == 'I' && yc) {
        emit(o, 'I', a.seq[q1], a.qual[q1]);
        ++c1; ++q1;
    } else if (xc && y == 'I') {
        emit(o, 'I', b.seq[q2], b.qual[q2]);
        ++c2; ++q2;
    } else {
        return HC_MERGE_UNDEFINED;
    }
    return HC_MERGE_OK;
}

__device__ int lead_clips(const Seg& s) { int k = 0; while (k < s.n_ops && clip(s.cig[k])) ++k; return k; }
__device__ int trail_clips(const Seg& s) { int k = 0; while (k < s.n_ops && clip(s.cig[s.n_ops - 1 - k])) ++k; return k; }

#define HC_STEP(expr) do { const int rc_ = (expr); if (rc_) return rc_; } while (0)

// mergeAlignmentRecordsSingle(ar,1,1): a = the running super-read, b = the next member.
__device__ int merge_pair(const Seg& a, const Seg& b, Out& o, const uint8_t* tab) {
    int s1 = a.start - lead_clips(a), e1 = a.end + trail_clips(a);
    int s2 = b.start - lead_clips(b), e2 = b.end + trail_clips(b);
    int q1 = 0, q2 = 0, c1 = 0, c2 = 0;
    o.n_ops = o.n_seq = 0;
    // The earlier start leads alone, then both reads up to the earlier end, then the later end trails alone. Ties go the
    // way the reference's four cases take them (:637, :649, :661, :673): equal starts -> a leads (no columns), equal
    // ends -> b trails (no columns).
    const bool a_leads = s1 <= s2, b_trails = e1 <= e2;
    const int both_from = a_leads ? s2 : s1, both_to = b_trails ? e1 : e2, all_to = b_trails ? e2 : e1;
    int ref = a_leads ? s1 : s2;
    if (a_leads) { while (ref < both_from) HC_STEP(lone_column(a, o, c1, q1, ref)); }
    else { while (ref < both_from) HC_STEP(lone_column(b, o, c2, q2, ref)); }
    while (ref <= both_to) HC_STEP(shared_column(a, b, o, c1, c2, q1, q2, ref, tab));
    if (b_trails) { while (ref <= all_to) HC_STEP(lone_column(b, o, c2, q2, ref)); }
    else { while (ref <= all_to) HC_STEP(lone_column(a, o, c1, q1, ref)); }
    return HC_MERGE_OK;
}

struct Batch {
    const int32_t *seg_start, *seg_end, *cig_len, *seq_len;
    const int64_t *cig_off, *seq_off;
    const uint8_t *cig, *seq, *qual;
    const int64_t *member_off, *out_off;
    const int32_t* members;
    uint8_t *out_cig, *out_seq, *out_qual;   // caller-visible result, out_off[c] .. out_off[c+1]
    uint8_t *tmp_cig, *tmp_seq, *tmp_qual;   // the other half of the ping-pong
    hc_merged_info* info;
    const uint8_t* tab;
    int32_t n_cliques;
};

__device__ __forceinline__ Seg load_seg(const Batch& B, int read) {
    const int r = 2 * read;
    return Seg{B.seg_start[r], B.seg_end[r], B.cig_len[r], B.seq_len[r], B.cig + B.cig_off[r], B.seq + B.seq_off[r], B.qual + B.seq_off[r]};
}

__global__ void __launch_bounds__(128) merge_cliques_kernel(Batch B) {
    __shared__ uint8_t tab[2 * NQ * NQ];
    for (int i = threadIdx.x; i < 2 * NQ * NQ; i += blockDim.x) tab[i] = B.tab[i];
    __syncthreads();
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < B.n_cliques; c += gridDim.x * blockDim.x) {
        const int64_t m0 = B.member_off[c];
        const int m = (int)(B.member_off[c + 1] - m0);
        const int64_t o0 = B.out_off[c];
        Out side[2] = {{B.out_cig + o0, B.out_seq + o0, B.out_qual + o0, 0, 0}, {B.tmp_cig + o0, B.tmp_seq + o0, B.tmp_qual + o0, 0, 0}};
        Seg acc = load_seg(B, B.members[m0]);
        int status = HC_MERGE_OK;
        // fold f = 1..m-1 writes side (m-1-f)&1, so the last fold lands in the caller-visible region
        for (int f = 1; f < m && status == HC_MERGE_OK; ++f) {
            const Seg b = load_seg(B, B.members[m0 + f]);
            Out& o = side[(m - 1 - f) & 1];
            status = merge_pair(acc, b, o, tab);
            acc.start = min(acc.start, b.start);   // :686-687
            acc.end = max(acc.end, b.end);
            acc.n_ops = o.n_ops;
            acc.n_seq = o.n_seq;
            acc.cig = o.cig; acc.seq = o.seq; acc.qual = o.qual;
        }
        if (m == 1) {   // the reference asserts size > 1 (:213); a lone member is copied through
            for (int k = 0; k < acc.n_ops; ++k) side[0].cig[k] = acc.cig[k];
            acc.n_seq = min(acc.n_seq, acc.n_ops);      // the slot is sized from the CIGAR length
            for (int k = 0; k < acc.n_seq; ++k) { side[0].seq[k] = acc.seq[k]; side[0].qual[k] = acc.qual[k]; }
        }
        hc_merged_info r;
        r.status = status;
        r.start = status ? 0 : acc.start;
        r.end = status ? 0 : acc.end;
        r.n_ops = status ? 0 : acc.n_ops;
        r.n_seq = status ? 0 : acc.n_seq;
        B.info[c] = r;
    }
}

struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t put(const void* src, size_t bytes, cudaStream_t s) {
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
        if (e != cudaSuccess) return e;
        return bytes ? cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, s) : cudaSuccess;
    }
    cudaError_t make(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};

int fail(int code, const std::string& msg) {
    hc::set_global_error(msg);
    return code;
}

int check_batch(const hc_reads* reads, int32_t n_cliques, const int64_t* member_off, const int32_t* members) {
    if (!reads || n_cliques < 0 || (n_cliques > 0 && (!member_off || !members))) return fail(HC_E_INVALID, "hc_merge: null argument");
    for (int c = 0; c < n_cliques; ++c) {
        if (member_off[c + 1] <= member_off[c]) return fail(HC_E_INVALID, "hc_merge: empty clique");
        for (int64_t k = member_off[c]; k < member_off[c + 1]; ++k) {
            const int r = members[k];
            if (r < 0 || r >= reads->n) return fail(HC_E_INVALID, "hc_merge: member index out of range");
            if (!reads->single_end[r]) return fail(HC_E_INVALID, "hc_merge_cliques_se: paired-end member (only single-end cliques are built)");
            const int nc = reads->cig_len[2 * r], nq = reads->seq_len[2 * r];
            if (nc < 0 || nq < 0) return fail(HC_E_INVALID, "hc_merge: negative CIGAR / sequence length");
            const uint8_t* cg = reads->cig + reads->cig_off[2 * r];
            int consumed = 0;
            for (int i = 0; i < nc; ++i) consumed += (cg[i] == 'M' || cg[i] == 'I' || cg[i] == 'S');
            if (consumed > nq) return fail(HC_E_INVALID, "hc_merge: CIGAR consumes more query bases than the read has");
            const uint8_t* q = reads->qual + reads->seq_off[2 * r];
            for (int i = 0; i < nq; ++i)
                if (q[i] < 33 || q[i] > 126) return fail(HC_E_INVALID, "hc_merge: quality char outside 33..126");
        }
    }
    return HC_OK;
}

}  // namespace

extern "C" {

int hc_merge_layout(const hc_reads* reads, int32_t n_cliques, const int64_t* member_off, const int32_t* members, int64_t* out_off) {
    int rc = check_batch(reads, n_cliques, member_off, members);
    if (rc != HC_OK) return rc;
    if (!out_off) return fail(HC_E_INVALID, "hc_merge_layout: null out_off");
    int64_t at = 0;
    for (int c = 0; c < n_cliques; ++c) {
        out_off[c] = at;
        int64_t cap = 0;
        for (int64_t k = member_off[c]; k < member_off[c + 1]; ++k) cap += reads->cig_len[2 * members[k]];
        at += (cap + 15) & ~(int64_t)15;
    }
    out_off[n_cliques] = at;
    return HC_OK;
}

int hc_merge_cliques_se(int32_t device, const hc_reads* reads, int32_t n_cliques, const int64_t* member_off, const int32_t* members,
                        const int64_t* out_off, uint8_t* out_cig, uint8_t* out_seq, uint8_t* out_qual, hc_merged_info* info,
                        float* kernel_ms) {
    int rc = check_batch(reads, n_cliques, member_off, members);
    if (rc != HC_OK) return rc;
    if (n_cliques == 0) return HC_OK;
    if (!out_off || !out_cig || !out_seq || !out_qual || !info) return fail(HC_E_INVALID, "hc_merge_cliques_se: null output");
    {   // the caller's out_off must leave every clique the room hc_merge_layout gives it
        std::vector<int64_t> want((size_t)n_cliques + 1);
        if ((rc = hc_merge_layout(reads, n_cliques, member_off, members, want.data())) != HC_OK) return rc;
        for (int c = 0; c < n_cliques; ++c)
            if (out_off[c] < 0 || out_off[c + 1] - out_off[c] < want[c + 1] - want[c])
                return fail(HC_E_INVALID, "hc_merge_cliques_se: out_off leaves a clique less room than hc_merge_layout");
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(HC_E_CUDA, "hc_merge_cliques_se: no such CUDA device (there is no CPU merging path)");
#define MG_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { if (stream) cudaStreamDestroy(stream); \
        return fail(HC_E_CUDA, std::string("hc_merge_cliques_se: ") + cudaGetErrorString(e_)); } } while (0)
    cudaStream_t stream = nullptr;
    MG_TRY(cudaSetDevice(device));
    MG_TRY(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    const int n = reads->n;
    // bytes of the three char streams actually referenced
    int64_t cig_bytes = 0, seq_bytes = 0;
    for (int r = 0; r < 2 * n; ++r) {
        cig_bytes = std::max<int64_t>(cig_bytes, reads->cig_off[r] + reads->cig_len[r]);
        seq_bytes = std::max<int64_t>(seq_bytes, reads->seq_off[r] + reads->seq_len[r]);
    }
    const int64_t n_members = member_off[n_cliques], out_bytes = out_off[n_cliques];
    std::vector<uint8_t> tab(2 * NQ * NQ);
    build_phred_tables(tab.data());
    DevBuf d_start, d_end, d_clen, d_slen, d_coff, d_soff, d_cig, d_seq, d_qual, d_moff, d_mem, d_ooff, d_tab, d_oc, d_os, d_oq, d_tc, d_ts, d_tq, d_info;
    MG_TRY(d_start.put(reads->seg_start, sizeof(int32_t) * 2 * n, stream));
    MG_TRY(d_end.put(reads->seg_end, sizeof(int32_t) * 2 * n, stream));
    MG_TRY(d_clen.put(reads->cig_len, sizeof(int32_t) * 2 * n, stream));
    MG_TRY(d_slen.put(reads->seq_len, sizeof(int32_t) * 2 * n, stream));
    MG_TRY(d_coff.put(reads->cig_off, sizeof(int64_t) * 2 * n, stream));
    MG_TRY(d_soff.put(reads->seq_off, sizeof(int64_t) * 2 * n, stream));
    MG_TRY(d_cig.put(reads->cig, (size_t)cig_bytes, stream));
    MG_TRY(d_seq.put(reads->seq, (size_t)seq_bytes, stream));
    MG_TRY(d_qual.put(reads->qual, (size_t)seq_bytes, stream));
    MG_TRY(d_moff.put(member_off, sizeof(int64_t) * ((size_t)n_cliques + 1), stream));
    MG_TRY(d_mem.put(members, sizeof(int32_t) * (size_t)n_members, stream));
    MG_TRY(d_ooff.put(out_off, sizeof(int64_t) * ((size_t)n_cliques + 1), stream));
    MG_TRY(d_tab.put(tab.data(), tab.size(), stream));
    MG_TRY(d_oc.make((size_t)out_bytes)); MG_TRY(d_os.make((size_t)out_bytes)); MG_TRY(d_oq.make((size_t)out_bytes));
    MG_TRY(d_tc.make((size_t)out_bytes)); MG_TRY(d_ts.make((size_t)out_bytes)); MG_TRY(d_tq.make((size_t)out_bytes));
    MG_TRY(d_info.make(sizeof(hc_merged_info) * (size_t)n_cliques));
    Batch B{};
    B.seg_start = (const int32_t*)d_start.p; B.seg_end = (const int32_t*)d_end.p;
    B.cig_len = (const int32_t*)d_clen.p; B.seq_len = (const int32_t*)d_slen.p;
    B.cig_off = (const int64_t*)d_coff.p; B.seq_off = (const int64_t*)d_soff.p;
    B.cig = (const uint8_t*)d_cig.p; B.seq = (const uint8_t*)d_seq.p; B.qual = (const uint8_t*)d_qual.p;
    B.member_off = (const int64_t*)d_moff.p; B.out_off = (const int64_t*)d_ooff.p; B.members = (const int32_t*)d_mem.p;
    B.out_cig = (uint8_t*)d_oc.p; B.out_seq = (uint8_t*)d_os.p; B.out_qual = (uint8_t*)d_oq.p;
    B.tmp_cig = (uint8_t*)d_tc.p; B.tmp_seq = (uint8_t*)d_ts.p; B.tmp_qual = (uint8_t*)d_tq.p;
    B.info = (hc_merged_info*)d_info.p; B.tab = (const uint8_t*)d_tab.p; B.n_cliques = n_cliques;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    MG_TRY(cudaEventCreate(&e0));
    MG_TRY(cudaEventCreate(&e1));
    // Resident CTAs per SM bound the number of cliques in flight, hence the ping-pong working set that has to live in L2
    // (HC_MERGE_BLOCKS_PER_SM overrides, 1..16; measurements in DESIGN.md 4.5).
    int sm_count = 148, per_sm = 8;
    MG_TRY(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    if (const char* e = std::getenv("HC_MERGE_BLOCKS_PER_SM")) per_sm = std::max(1, std::min(16, std::atoi(e)));
    const int blocks = std::max(1, std::min((n_cliques + 127) / 128, sm_count * per_sm));
    MG_TRY(cudaEventRecord(e0, stream));
    merge_cliques_kernel<<<blocks, 128, 0, stream>>>(B);
    MG_TRY(cudaGetLastError());
    MG_TRY(cudaEventRecord(e1, stream));
    MG_TRY(cudaMemcpyAsync(out_cig, d_oc.p, (size_t)out_bytes, cudaMemcpyDeviceToHost, stream));
    MG_TRY(cudaMemcpyAsync(out_seq, d_os.p, (size_t)out_bytes, cudaMemcpyDeviceToHost, stream));
    MG_TRY(cudaMemcpyAsync(out_qual, d_oq.p, (size_t)out_bytes, cudaMemcpyDeviceToHost, stream));
    MG_TRY(cudaMemcpyAsync(info, d_info.p, sizeof(hc_merged_info) * (size_t)n_cliques, cudaMemcpyDeviceToHost, stream));
    MG_TRY(cudaStreamSynchronize(stream));
    float ms = 0.f;
    MG_TRY(cudaEventElapsedTime(&ms, e0, e1));
    if (kernel_ms) *kernel_ms = ms;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaStreamDestroy(stream);
#undef MG_TRY
    return HC_OK;
}

}  // extern "C"
