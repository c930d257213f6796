/* hc_edge.h -- C ABI of libhcgpu.so: B200 (sm_100a) read-pair edge scoring.
 *
 * Drop-in boundary for ONE path of cbg-ethz/haploclique: the edge decision
 *     bool EdgeCalculator::edgeBetween(const AlignmentRecord&, const AlignmentRecord&) const
 * (reference: src/EdgeCalculator.h:29) as implemented by NewEdgeCalculator
 * (src/NewEdgeCalculator.cpp:1247-1376), evaluated for a whole batch of position-sorted reads at once.
 * Plain pointers and sizes only; no C++/torch types. All functions return HC_OK (0) or a negative
 * HC_E_* code; hc_last_error() gives the text. There is no CPU scoring path behind this API: without
 * a CUDA device every scoring call fails with HC_E_CUDA.
 *
 * Call sequence (INTEGRATION.md shows the reference-side binding):
 *     hc_create(&params, &ctx)            <- NewEdgeCalculator::NewEdgeCalculator (src/NewEdgeCalculator.cpp:77-90)
 *     hc_pack(&reads, &packed)            <- AlignmentRecord getters (src/AlignmentRecord.cpp:1703-1811)
 *     hc_upload(ctx, packed)
 *     hc_score_band(ctx, ...) / hc_score_pairs(ctx, ...)   <- the edgeBetween calls of
 *                                            CLEVER::addAlignment (src/CLEVER.cpp:218-232) and
 *                                            BronKerbosch::addAlignment (src/BronKerbosch.cpp:286)
 *     hc_packed_free(packed); hc_destroy(ctx)
 * Several GPUs behind one handle: hc_multi_*. Super-read construction for a round's cliques: hc_merger_* / hc_merge_*.
 */
#ifndef HC_EDGE_H_
#define HC_EDGE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HC_ABI_VERSION 2

enum {
    HC_OK = 0,
    HC_E_INVALID = -1,   /* bad argument / malformed read (unknown CIGAR op, quality outside 33..126, ...) */
    HC_E_CUDA = -2,      /* no device, or a CUDA call failed */
    HC_E_NOMEM = -3,
    HC_E_STATE = -4      /* call out of sequence (e.g. scoring before hc_upload) */
};

/* Constructor arguments of NewEdgeCalculator that the scorer reads (src/NewEdgeCalculator.h:58).
 * Q and frameshift_merge are stored but unused by the reference's scorer and are not taken here. */
typedef struct {
    double edge_quasi_cutoff_cliques;   /* EDGE_QUASI_CUTOFF        (default 0.99) */
    double edge_quasi_cutoff_mixed;     /* EDGE_QUASI_CUTOFF_MIXED  (default 0.97) */
    double edge_quasi_cutoff_single;    /* EDGE_QUASI_CUTOFF_SINGLE (default 0.95) */
    double min_overlap_cliques;         /* MIN_OVERLAP_CLIQUES      (default 0.9)  */
    double min_overlap_single;          /* MIN_OVERLAP_SINGLE       (default 0.6)  */
    int32_t no_prob0;                   /* NOPROB0 */
    int32_t simpson_len;                /* maxPosition+1; 0 = no allele-frequency file */
    const double* simpson;              /* dense SIMPSON_MAP, 0.0 = position absent (src/NewEdgeCalculator.cpp:85-88) */
    int32_t device;                     /* CUDA device ordinal */
    int32_t reserved;
    double ambiguity_rel_tol;           /* pairs with |score-cutoff| <= tol*|cutoff| are also listed (0 -> 1e-12) */
} hc_params;

/* A batch of reads in arrival (position-sorted) order, as the AlignmentRecord fields the scorer reads
 * (src/AlignmentRecord.h:54-77). Segment s (0/1) of read r is row 2*r+s of the per-segment arrays. */
typedef struct {
    int32_t n;
    const uint8_t* single_end;          /* [n]  AlignmentRecord::single_end */
    const uint8_t* is_clique;           /* [n]  name contains "Clique" (src/NewEdgeCalculator.cpp:1111) */
    const int32_t* ncov;                /* [n]  cov_pos.size() */
    const int32_t* cov_first;           /* [n]  cov_pos.front().ref */
    const int32_t* cov_last;            /* [n]  cov_pos.back().ref */
    const int32_t* seg_start;           /* [2n] start1/start2, 1-based inclusive */
    const int32_t* seg_end;             /* [2n] end1/end2 */
    const int64_t* cig_off;             /* [2n] offset of the segment's unrolled CIGAR in cig[] */
    const int32_t* cig_len;             /* [2n] 0 for the absent 2nd segment of a single-end read */
    const int64_t* seq_off;             /* [2n] offset of bases/qualities in seq[]/qual[] */
    const int32_t* seq_len;             /* [2n] */
    const uint8_t* cig;                 /* unrolled CIGAR chars, M I D S H only */
    const uint8_t* seq;                 /* bases; anything but ACGT counts as N (src/ShortDnaSequence.cpp:37-50) */
    const uint8_t* qual;                /* ASCII Phred+33, 33..126 */
} hc_reads;

/* Packed per-read header, 64 bytes, the unit the kernels load.
 * raw[]/ext[] follow the order {start1,end1,start2,end2}; ext = clip-extended interval
 * (start - leading S/H count, end + trailing S/H count: src/NewEdgeCalculator.cpp:413-421). */
typedef struct {
    int32_t raw[4];
    int32_t ext[4];
    uint32_t ops_off[2];                /* first op record of each segment in hc_packed.ops; a read's two segments are
                                           contiguous and ops_off[0] is a multiple of 8 (16-byte aligned stream) */
    int32_t n_ops[2];                   /* unrolled CIGAR length per segment */
    uint32_t ncov_flags;                /* bits 0..29 ncov, bit 30 single_end, bit 31 is_clique */
    int32_t cov_first;
    int32_t cov_last;
    uint16_t seg_flags[2];              /* HC_SEG_* per segment */
} hc_read_hdr;

#define HC_SEG_HAS_INS 1u               /* segment holds at least one I op */
#define HC_SEG_NOT_PLAIN 2u             /* some op between the leading and trailing clip runs is not M */

#define HC_FLAG_SINGLE_END 0x40000000u
#define HC_FLAG_IS_CLIQUE 0x80000000u

/* One 16-bit record per unrolled CIGAR op:
 *   bits 0..2  base code    A=0 C=1 G=2 T=3 N=4   (0 for D/H ops, which carry no base)
 *   bits 3..5  op           M=0 I=1 D=2 S=3 H=4
 *   bits 8..14 quality code (dense re-numbering of the quality chars present, see qual_chars[]): the high
 *              byte, so a byte extract yields the table index without masking */
#define HC_OP_SHIFT 3
#define HC_BASE_SHIFT 0
#define HC_QUAL_SHIFT 8
#define HC_REC_OP(r) (((r) >> HC_OP_SHIFT) & 7u)
#define HC_REC_BASE(r) ((r) & 7u)
#define HC_REC_QUAL(r) (((r) >> HC_QUAL_SHIFT) & 127u)
enum { HC_OP_M = 0, HC_OP_I = 1, HC_OP_D = 2, HC_OP_S = 3, HC_OP_H = 4 };

typedef struct {
    int32_t n_reads;
    int32_t n_qual_codes;               /* K <= 94 */
    uint8_t qual_chars[128];            /* code -> quality char */
    const hc_read_hdr* hdr;             /* [n_reads] */
    const uint16_t* ops;                /* [n_ops_total (+ 64 records of padding)] */
    int64_t n_ops_total;                /* records in ops[], including the per-read alignment padding */
    int32_t max_ops_per_read;           /* max over reads of n_ops[0]+n_ops[1] */
    int32_t reserved;
    void* owner;                        /* library-private */
} hc_packed;

/* Per-pair result of hc_score_pairs (the values edgeBetween computes on its way to the bool). */
enum { HC_ST_REJECT_COV = 0, HC_ST_REJECT_LAYOUT = 1, HC_ST_WALKED = 2, HC_ST_UNDEFINED = 3 };
typedef struct {
    int32_t status;                     /* HC_ST_*; UNDEFINED = the reference would read out of bounds here */
    int32_t edge;
    int32_t gap;                        /* checkGapsCigar returned true (src/NewEdgeCalculator.cpp:1367) */
    int32_t cc, tc;
    int32_t ambiguous;                  /* |score-cutoff| within ambiguity_rel_tol */
    double probM, prob0;
    double score;                       /* prob*potence (:1139); NaN when not reached */
} hc_pair_result;

/* A pair whose score lies within ambiguity_rel_tol of the threshold. The device's log10 is not glibc's, so such a pair
 * is re-decided on the host exactly as similarityCriterion does (src/NewEdgeCalculator.cpp:1127-1140: std::log10,
 * reciprocal, multiply, >=) from the bit-exact (probM, prob0, cc, tc); the adjacency bit the caller receives is the
 * host's decision. This list is the report of those pairs. */
typedef struct {
    int32_t i, j;                       /* ap1 = read i (later arrival), ap2 = read j */
    int32_t cc, tc;
    double probM, prob0;
    double score, cutoff;               /* score as re-computed on the host with std::log10; cutoff = log10(cutoff) */
    int32_t edge;                       /* the host's decision = the adjacency bit */
    int32_t edge_device;                /* what the device had decided */
} hc_ambiguous;

/* Totals of one hc_score_band call. */
typedef struct {
    int64_t pairs_total;                /* ordered pairs i>j looked at (inside the band) */
    int64_t pairs_scored;               /* passed the covered-interval reject (src/NewEdgeCalculator.cpp:1252) */
    int64_t pairs_walked;               /* also passed the layout rejects (:1268-1365) */
    int64_t edges;
    int64_t sum_cc, sum_tc;             /* column steps */
    int64_t undefined;
    int64_t ambiguous;                  /* near-threshold pairs found and re-decided on the host (may exceed the capacity
                                           of the caller's list; all of them are re-decided) */
    float kernel_ms;                    /* device time from the clearing of the adjacency words to the end of the last
                                           scoring kernel, CUDA events on the compute stream */
    int32_t launches;                   /* scoring kernels launched (one per row window) */
} hc_band_stats;

typedef struct hc_ctx hc_ctx;

int hc_abi_version(void);
int hc_device_count(void);
const char* hc_last_error(const hc_ctx* ctx);   /* ctx may be NULL: error of the last failed hc_create/hc_pack */

int hc_create(const hc_params* params, hc_ctx** out);
void hc_destroy(hc_ctx* ctx);

/* Host-side packing (C++, no GPU needed): validates, re-codes qualities, builds headers and op records.
 * The output buffers are page-locked when a CUDA device is present (faster hc_upload), plain otherwise. */
int hc_pack(const hc_reads* reads, hc_packed** out);
void hc_packed_free(hc_packed* p);

/* Copies a packed batch to the device (replaces the previous one). */
int hc_upload(hc_ctx* ctx, const hc_packed* p);

/* Band layout for rows [row_begin,row_end): row i holds bits for columns j in [col0[i], i), col0 a
 * multiple of 32, chosen so that every j < col0[i] fails the covered-interval reject against i.
 * word_off[k] (k = i-row_begin, row_end-row_begin+1 entries) = first 32-bit word of row i in the
 * adjacency buffer. Returns the total number of words through *n_words. */
int hc_band_layout(hc_ctx* ctx, int32_t row_begin, int32_t row_end, int32_t* col0, int64_t* word_off,
                   int64_t* n_words);

/* Scores every ordered pair (i, j<i) of rows [row_begin,row_end) that lies in the band and writes the
 * adjacency bits (bit (j-col0[i]) of row i = edgeBetween(read i, read j)) to host memory `adj`
 * (n_words words, see hc_band_layout; may be NULL to keep the result on the device only).
 * The rows are scored in windows (one kernel launch each); the words of a finished window travel to `adj` on a second
 * stream while the next window is scored. Pairs within ambiguity_rel_tol of the threshold are re-decided on the host with
 * std::log10 (see hc_ambiguous) and their bits patched, in `adj` and on the device; amb/amb_cap: optional report. */
int hc_score_band(hc_ctx* ctx, int32_t row_begin, int32_t row_end, uint32_t* adj, int64_t n_words,
                  hc_ambiguous* amb, int32_t amb_cap, hc_band_stats* stats);

/* Same pass, but additionally returns the per-pair values (status, gap, cc, tc, probM, prob0, score) of every
 * band slot: detail[w*32 + b] belongs to bit b of adjacency word w (slots with j >= i stay zeroed). Only the
 * tiled kernel produces this (parity testing of that kernel); HC_E_STATE otherwise. n_words as in hc_score_band. */
int hc_score_band_detail(hc_ctx* ctx, int32_t row_begin, int32_t row_end, uint32_t* adj, hc_pair_result* detail,
                         int64_t n_words, hc_band_stats* stats);

/* Which band kernel hc_score_band uses. AUTO takes the first eligible of
 *   TILED2   operands stored as table-address addends, 32 column reads staged in a shared-memory tile, rows fed by TMA;
 *            9-row table of 72*K*K bytes for K quality codes (K up to ~47 beside a tile of 2x250-column reads);
 *   TILED    16-bit op records, the same tile, 16*K*K bytes of tables (more quality codes than TILED2 can hold);
 *   STREAMED the work decomposition and lock-step loop of TILED with both op streams read in place from global
 *            memory, tables only in shared memory: reads too long or K too large for a tile (super-read rounds:
 *            thousands of columns, 60-94 quality codes); K <= 118, reads below 65536 ops;
 *   GENERIC  one lane per pair (what hc_score_pairs runs); as a band kernel only the cross-check of the others.
 * With an allele-frequency map every band kernel walks the tail columns instead of evaluating them in O(1).
 * GENERIC / TILED / TILED2 / STREAMED force one; a forced kernel fails with HC_E_STATE when the batch is not eligible.
 * All four produce identical results. hc_last_kernel: the kernel the last hc_score_band call launched (0 = none). */
enum { HC_KERNEL_AUTO = 0, HC_KERNEL_GENERIC = 1, HC_KERNEL_TILED = 2, HC_KERNEL_TILED2 = 3, HC_KERNEL_STREAMED = 4 };
int hc_set_kernel(hc_ctx* ctx, int32_t mode);
int hc_last_kernel(const hc_ctx* ctx);

/* Scores an explicit list of ordered pairs (ap1 = pi[k], ap2 = pj[k]), full detail per pair. */
int hc_score_pairs(hc_ctx* ctx, const int32_t* pi, const int32_t* pj, int64_t npairs, hc_pair_result* out);

/* hc_score_pairs for callers that must not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 * Same inputs, outputs and results as hc_score_pairs; the kernel's device-side scratch (the pair lists and the result
 * records) lives in a static device array of HC_STATIC_MAX_PAIRS * 56 bytes = 896 KiB that every GPU which loads the
 * library holds until the process exits. The call makes no cudaMalloc / cudaFree and calls no allocating library routine;
 * npairs > HC_STATIC_MAX_PAIRS fails with HC_E_INVALID and a message. All calls on a device share the one array (it keeps
 * the previous call's contents; the slots a call uses are cleared first), so calls that overlap on the same GPU -- from
 * two contexts or two host threads -- race: serialise them. The batch must already be resident (hc_upload allocates). */
#define HC_STATIC_MAX_PAIRS 16384
int hc_score_pairs_static(hc_ctx* ctx, const int32_t* pi, const int32_t* pj, int64_t npairs, hc_pair_result* out);

/* ---- QuasispeciesEdgeCalculator (src/QuasispeciesEdgeCalculator.cpp), the reference's older scorer: same
 * EdgeCalculator interface (src/EdgeCalculator.h:29), constructor src/QuasispeciesEdgeCalculator.h:57. It runs on
 * the batch hc_upload made resident and uses the context's allele-frequency map (its SIMPSON_MAP; entries equal to
 * 0.0 count as absent). q = computeOverlap(ap1, ap2, cutoff) (:72), edge = q >= cutoff (:69); the same-name
 * shortcut (:56) does not apply inside one batch. The per-column log terms are tabulated on the host, so the
 * running sum is bit-identical to the reference's; only the closing pow(exp(sum), 1/n) (:425) is device libm, and
 * pairs with |q - cutoff| <= ambiguity_rel_tol * |cutoff| are flagged. status = HC_ST_UNDEFINED where the reference
 * would not terminate (:258-267, :347-381) or would index out of range. */
typedef struct {
    double edge_quasi_cutoff_cliques, edge_quasi_cutoff_mixed, edge_quasi_cutoff_single;   /* :36-38 */
    double min_overlap_cliques, min_overlap_single;                                         /* :39-40 */
    int32_t frameshift_merge;                                                               /* :41 */
    int32_t reserved;
    double ambiguity_rel_tol;           /* 0 -> 1e-12 */
} hc_quasi_params;

typedef struct {
    int32_t status;                     /* HC_ST_WALKED or HC_ST_UNDEFINED */
    int32_t edge;
    int32_t n_single;                   /* singleOverlap calls computeOverlap made: 0 (rejected by the overlap gates), 1, 2 */
    int32_t flags;                      /* HC_QUASI_* */
    double q;                           /* computeOverlap's value */
    double cutoff;                      /* the cutoff edgeBetween picked (:59-66) */
} hc_quasi_result;

#define HC_QUASI_AMBIGUOUS 1             /* |q - cutoff| within ambiguity_rel_tol */
#define HC_QUASI_SUBNORMAL 2             /* a singleOverlap's log sum is below log(DBL_MIN): exp(sum) (:425) is subnormal or 0,
                                            q = pow(exp(sum), 1/n) has lost its precision in the reference itself (its exp
                                            result has few or no significant bits left); q < 0.06 there for n <= 250 */

/* Explicit list of ordered pairs (ap1 = pi[k], ap2 = pj[k]). */
int hc_quasi_score_pairs(hc_ctx* ctx, const hc_quasi_params* params, const int32_t* pi, const int32_t* pj,
                         int64_t npairs, hc_quasi_result* out);

/* Banded adjacency of rows [row_begin,row_end) in hc_band_layout's layout. Pairs outside the band have disjoint
 * covered intervals, for which computeOverlap returns 0: no edge as long as the cutoffs are > 0 (required here). */
int hc_quasi_score_band(hc_ctx* ctx, const hc_quasi_params* params, int32_t row_begin, int32_t row_end, uint32_t* adj,
                        int64_t n_words, hc_band_stats* stats);

/* ---- Several GPUs behind one handle ----------------------------------------------------------------
 * The band shards by rows with no exchange step: a pair depends on its two reads only. hc_partition_rows cuts the
 * position-sorted batch into n_parts consecutive row ranges of equal scoring work (scored pairs -- pairs inside the band whose
 * covered intervals intersect -- weighted by the columns they share), and gives for each part the first read it has to hold: its left halo, i.e. every earlier read whose
 * covered interval can reach one of its rows (from the running maximum of cov_last, not a constant). When everything
 * overlaps (deep amplicons, super-read rounds) the halo reaches read 0 and the parts become equal-area row blocks of the
 * replicated set. first[p] is a multiple of 32, so part p's local band layout is the global one shifted by first[p] and
 * its adjacency words are the global words [word_off[cuts[p]], word_off[cuts[p+1]]) unchanged.
 *   cov_first / cov_last: per read, as in hc_reads.   cuts: n_parts+1 entries, cuts[0] = 0, cuts[n_parts] = n_reads. */
int hc_partition_rows(const int32_t* cov_first, const int32_t* cov_last, int32_t n_reads, int32_t n_parts,
                      int32_t* cuts, int32_t* first);

/* Reads [first, end) of a packed batch as a batch of their own (headers copied and re-based, op records shared with
 * `p`, which must outlive the slice). Free with hc_packed_free. */
int hc_packed_slice(const hc_packed* p, int32_t first, int32_t end, hc_packed** out);

/* One host process driving n_devices GPUs (one context, two streams and one host thread per device; the same ordinal may
 * be listed more than once). hc_multi_upload partitions the batch with hc_partition_rows and uploads each part with its
 * halo; hc_multi_score_band scores all parts concurrently, every device copying its rows' words straight into the
 * caller's `adj` buffer (hc_multi_band_layout's layout = hc_band_layout's of the whole batch): the final gather is the
 * set of D2H copies, there is no collective. params->device is ignored. stats: sums over the devices, kernel_ms = the
 * slowest device's. per_device (optional, n_devices entries): each device's own stats. */
typedef struct hc_multi hc_multi;
int hc_multi_create(const hc_params* params, const int32_t* devices, int32_t n_devices, hc_multi** out);
void hc_multi_destroy(hc_multi* m);
const char* hc_multi_last_error(const hc_multi* m);
int hc_multi_n_devices(const hc_multi* m);
hc_ctx* hc_multi_context(hc_multi* m, int32_t k);            /* device k's context (hc_set_kernel, hc_l2_flush, ...) */
int hc_multi_set_params(hc_multi* m, const hc_params* params);
int hc_multi_upload(hc_multi* m, const hc_packed* p);
int hc_multi_partition(const hc_multi* m, int32_t* cuts, int32_t* first);     /* the partition of the resident batch */
int hc_multi_band_layout(hc_multi* m, int32_t* col0, int64_t* word_off, int64_t* n_words);
int hc_multi_score_band(hc_multi* m, uint32_t* adj, int64_t n_words, hc_ambiguous* amb, int32_t amb_cap,
                        hc_band_stats* stats, hc_band_stats* per_device);

/* Replaces the scalar scoring parameters (cutoffs, min overlaps, no_prob0, tolerance) of a live context;
 * the allele-frequency map and the device stay as given to hc_create. Mirrors
 * EdgeCalculator::setOverlapCliques (src/EdgeCalculator.h:35). */
int hc_set_params(hc_ctx* ctx, const hc_params* params);

/* Evicts the device L2 by overwriting a scratch buffer larger than it (benchmark hygiene between timed
 * iterations); asynchronous on the context's stream. */
int hc_l2_flush(hc_ctx* ctx);

/* Page-locked host memory for the adjacency / input buffers (plain cudaHostAlloc / cudaFreeHost). */
void* hc_host_alloc(int64_t bytes);
void hc_host_free(void* p);

/* On-chip peak probes for the roofline (bench.py): sustained shared-memory load bandwidth over all SMs (GB/s,
 * conflict-free 128-bit loads) and fp64 multiply rate (DMUL/s, independent chains), measured on `device` now. */
int hc_measure_peaks(int32_t device, double* lds_gbs, double* dmul_per_s);

/* The two 127x127 match-probability tables the context uploaded (row-major [q1][q2], doubles). */
int hc_get_tables(const hc_ctx* ctx, double* eq127x127, double* uneq127x127);

/* ---- Super-read construction -----------------------------------------------------------------------
 * Replaces AlignmentRecord::AlignmentRecord(unique_ptr<vector<const AlignmentRecord*>>& alignments, unsigned clique_id)
 * (src/AlignmentRecord.cpp:211-249; call site: a clique becomes a super-read in src/CliqueCollector.h:59) for whole batches
 * of cliques and any mix of single-end and paired-end members: the fold of mergeAlignmentRecordsSingle (:581-734),
 * mergeAlignmentRecordsPaired (:734-1293) and mergeAlignmentRecordsMixed (:1295-1537) over the members in the given order.
 * The caller keeps name ("Clique_<id>"), readNames and cov_pos, which are set unions / derived fields (:243-248).
 * There is no CPU merging path: without a CUDA device the calls fail with HC_E_CUDA.
 *     hc_merger_create(device, &m)                      once
 *     hc_merger_set_reads(m, &reads)                    per finder round: the reads the cliques refer to (same hc_reads as hc_pack)
 *     hc_merge_run(m, n, member_off, members, info, &total_ops, &total_seq, &ms)
 *     hc_merge_fetch(m, cig, seq, qual)                 the records' bytes, compacted on the device
 *     hc_merger_destroy(m) */
#define HC_MERGE_OK 0
#define HC_MERGE_UNDEFINED 1            /* an op pair overlapMerge does not list (:1671 assert(false): the Release build of
                                           the reference never leaves that loop) */
#define HC_MERGE_OVERRUN 2              /* a cursor ran past its CIGAR or its sequence (the layout's loops ask for columns
                                           the read does not have: the reference reads past the vector / throws from .at()) */
typedef struct {
    int32_t status;                     /* HC_MERGE_* */
    int32_t single_end;                 /* AlignmentRecord::single_end of the super-read */
    int32_t start[2], end[2];           /* start1/end1, start2/end2 */
    int32_t n_ops[2];                   /* lengths of cigar1_unrolled / cigar2_unrolled (M / I / D only: merges drop clips) */
    int32_t n_seq[2];                   /* lengths of sequence1 / sequence2 and their quality strings */
} hc_merged_record;

typedef struct hc_merger hc_merger;
int hc_merger_create(int32_t device, hc_merger** out);
void hc_merger_destroy(hc_merger* m);
const char* hc_merger_last_error(const hc_merger* m);
int hc_merger_set_reads(hc_merger* m, const hc_reads* reads);

/* Clique c holds reads members[member_off[c] .. member_off[c+1]) in fold order (the order of the reference's `alignments`
 * vector). info[c]: the super-read's header. The records' bytes stay on the device until hc_merge_fetch: unrolled CIGARs
 * back to back (clique-major, segment 1 then segment 2 of a paired result; total_ops bytes), and likewise bases and quality
 * chars (total_seq bytes each); a clique whose status is not HC_MERGE_OK contributes nothing. kernel_ms (optional): device
 * time of the merge kernel. */
int hc_merge_run(hc_merger* m, int32_t n_cliques, const int64_t* member_off, const int32_t* members, hc_merged_record* info,
                 int64_t* total_ops, int64_t* total_seq, float* kernel_ms);
int hc_merge_fetch(hc_merger* m, uint8_t* cig, uint8_t* seq, uint8_t* qual);
/* How the folds of the last hc_merge_run were evaluated: in reference-column space by a whole warp (segments that are
 * clips / M only / clips), or walked op by op by one lane (indels, cursor quirks). */
int hc_merge_fold_counts(const hc_merger* m, int64_t* column_space, int64_t* walked);

/* ---- Alignments in BAM order -> the position-sorted batch -----------------------------------------------
 * Replaces readBamFile (src/haploclique.cpp:143-201) and the record building it drives: AlignmentRecord(BamAlignment)
 * (src/AlignmentRecord.cpp:186-209), pairWith (:251-312) and getMergedDnaSequence for overlapping mates (:380-580), up to the
 * hc_reads arrays hc_pack / hc_merger_set_reads take -- without AlignmentRecord objects and without cov_pos (32 bytes per
 * covered column in the reference; only its size and its first / last position are needed and they are counted off the
 * CIGAR). The BAM decoder stays with the caller (integration/GpuBamLoader feeds the reference's vendored BamTools reader).
 * Host code only, no GPU needed.
 *     hc_loader_create(&l);  for every alignment in file order: hc_loader_add(l, ...);  hc_loader_finish(l, &reads, &max_position)
 * hc_loader_add: name, 0-based position, CIGAR as op chars + lengths (BamTools' CigarOp), query bases and qualities as
 * stored. Alignments without an M op of positive length are skipped as the reference skips them (:160-165). HC_E_INVALID
 * where the reference's mate merge is undefined for two overlapping mates (its loop would not return).
 * hc_loader_finish: the batch in the reference's arrival order (pairs in file order, leftovers in the reference's hash
 * order, std::sort by start); the arrays belong to the loader and live until hc_loader_destroy. max_position as :195-198.
 * hc_loader_record_name(read) indexes hc_loader_name(k), the reference's readNames vector. */
typedef struct hc_loader hc_loader;
int hc_loader_create(hc_loader** out);
void hc_loader_destroy(hc_loader* l);
const char* hc_loader_last_error(const hc_loader* l);
int hc_loader_add(hc_loader* l, const char* name, int32_t position0, const char* cigar_ops, const uint32_t* cigar_lens, int32_t n_cigar,
                  const char* bases, const char* quals, int32_t n_bases);
int hc_loader_finish(hc_loader* l, hc_reads* out, uint32_t* max_position);
int32_t hc_loader_n_names(const hc_loader* l);
const char* hc_loader_name(const hc_loader* l, int32_t k);
int32_t hc_loader_record_name(const hc_loader* l, int32_t read);
int hc_loader_counts(const hc_loader* l, int64_t* alignments, int64_t* skipped, int64_t* merged_mates);

#ifdef __cplusplus
}
#endif
#endif /* HC_EDGE_H_ */
