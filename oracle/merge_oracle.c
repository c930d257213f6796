/* oracle/merge_oracle.c -- TEST INFRASTRUCTURE ONLY (the checker; never on the product path).
 *
 * Plain-C restatement of the reference's super-read construction for SINGLE-END cliques (SURVEY.md 8(f) N2,
 * first slice): AlignmentRecord(alignments, clique_id) (src/AlignmentRecord.cpp:211-249) folding the members in
 * order through mergeAlignmentRecordsSingle(ar,1,1) (:581-734), built from noOverlapMerge (:1539-1576),
 * overlapMerge (:1578-1674) and computeEntry (:169-184) over the Phred update tables (:40-90).
 * Paired and mixed merges (mergeAlignmentRecordsPaired / Mixed, :734-1537) are NOT restated here.
 *
 * Parity: PINNED against the compiled reference (oracle/_ref/libhcref.so: hcref_merge_clique, hcref_compute_entry)
 * by tests/test_merge_oracle.py and the digests of tests/golden/merge_golden.json.
 *
 * Where the reference is undefined -- an op combination overlapMerge does not list (assert(false), compiled out in
 * the Release build: the loop then never advances) or a cursor running past its CIGAR -- the restatement stops and
 * reports ORC_MERGE_UNDEFINED / ORC_MERGE_OVERRUN instead.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_MERGE_OK 0
#define ORC_MERGE_UNDEFINED 1
#define ORC_MERGE_OVERRUN 2

/* agreement / disagreement (src/AlignmentRecord.cpp:40-54): float arithmetic, pow in double (std::pow(int,float)),
 * log10 and round on float. */
static int agreement(int q1, int q2) {
    float p1 = (float)pow(10.0, (double)((float)(-q1) / 10));
    float p2 = (float)pow(10.0, (double)((float)(-q2) / 10));
    float post = (p1 * p2 / 3) / (1 - p1 - p2 + 4 * p1 * p2 / 3);
    post = roundf(-10 * log10f(post));
    return (int)post;
}
static int disagreement(int q1, int q2) {
    float p1 = (float)pow(10.0, (double)((float)(-q1) / 10));
    float p2 = (float)pow(10.0, (double)((float)(-q2) / 10));
    float post = (p1 * (1 - p2 / 3)) / (p1 + p2 - 4 * p1 * p2 / 3);
    post = roundf(-10 * log10f(post));
    return (int)post;
}

/* Rows / columns 0..32 are never filled by the reference's builders (:68-86 start at 33) and sit in zero-initialised static
 * storage there as here. They ARE reached: agreement(Q62|64|68|70|73, Q0) rounds to -1..-4, so computeEntry can emit the
 * chars 29..32, and a later fold indexes the table with them, reading 0 -> '!'. */
static int tab_ready;
static int tab_agree[127][127], tab_dis[127][127];
static void tables(void) {
    if (tab_ready) return;
    for (int i = 33; i < 127; ++i)
        for (int j = 33; j < 127; ++j) {
            tab_agree[i][j] = agreement(i - 33, j - 33);    /* :68-76 */
            tab_dis[i][j] = disagreement(i - 33, j - 33);   /* :78-86 */
        }
    tab_ready = 1;
}

/* computeEntry (:169-184). Qualities are ASCII chars 33..126. */
void orc_compute_entry(uint8_t b1, uint8_t q1, uint8_t b2, uint8_t q2, uint8_t* out2) {
    tables();
    int v;
    if (b1 == b2) { out2[0] = b1; v = tab_agree[q1][q2] + 33; }
    else if (q1 >= q2) { out2[0] = b1; v = tab_dis[q1][q2] + 33; }
    else { out2[0] = b2; v = tab_dis[q2][q1] + 33; }
    out2[1] = (uint8_t)(v < 126 ? v : 126);
}

typedef struct {
    int start, end, n_ops, n_seq;
    const uint8_t *cig, *seq, *qual;
} mseg;

typedef struct {
    uint8_t *cig, *seq, *qual;
    int n_ops, n_seq;
} mout;

static int lead(const mseg* s) { int k = 0; while (k < s->n_ops && (s->cig[k] == 'S' || s->cig[k] == 'H')) ++k; return k; }
static int trail(const mseg* s) { int k = 0; while (k < s->n_ops && (s->cig[s->n_ops - 1 - k] == 'S' || s->cig[s->n_ops - 1 - k] == 'H')) ++k; return k; }

/* noOverlapMerge (:1539-1576) */
static int no_overlap(const mseg* s, mout* o, int* c, int* q, int* ref) {
    if (*c >= s->n_ops) return ORC_MERGE_OVERRUN;
    uint8_t op = s->cig[*c];
    if (op == 'H') { ++*ref; ++*c; }
    else if (op == 'I') { o->seq[o->n_seq] = s->seq[*q]; o->qual[o->n_seq++] = s->qual[*q]; o->cig[o->n_ops++] = 'I'; ++*q; ++*c; }
    else if (op == 'D') { o->cig[o->n_ops++] = 'D'; ++*ref; ++*c; }
    else if (op == 'S') { ++*ref; ++*q; ++*c; }
    else if (op == 'M') { o->seq[o->n_seq] = s->seq[*q]; o->qual[o->n_seq++] = s->qual[*q]; o->cig[o->n_ops++] = 'M'; ++*ref; ++*q; ++*c; }
    else return ORC_MERGE_UNDEFINED;
    return ORC_MERGE_OK;
}

static void take(const mseg* s, int q, mout* o, uint8_t op) {
    o->seq[o->n_seq] = s->seq[q];
    o->qual[o->n_seq++] = s->qual[q];
    o->cig[o->n_ops++] = op;
}

/* overlapMerge (:1578-1674) */
static int overlap(const mseg* a, const mseg* b, mout* o, int* c1p, int* c2p, int* q1p, int* q2p, int* ref) {
    if (*c1p >= a->n_ops || *c2p >= b->n_ops) return ORC_MERGE_OVERRUN;
    uint8_t c1 = a->cig[*c1p], c2 = b->cig[*c2p];
    if ((c1 == 'M' && c2 == 'M') || (c1 == 'S' && c2 == 'S') || (c1 == 'I' && c2 == 'I')) {
        if (c1 != 'S') {
            uint8_t e[2];
            orc_compute_entry(a->seq[*q1p], a->qual[*q1p], b->seq[*q2p], b->qual[*q2p], e);
            o->seq[o->n_seq] = e[0]; o->qual[o->n_seq++] = e[1]; o->cig[o->n_ops++] = c1;
        }
        if (c1 != 'I') ++*ref;
        ++*q1p; ++*q2p; ++*c1p; ++*c2p;
    } else if ((c1 == 'D' && c2 == 'D') || (c1 == 'H' && c2 == 'H') || (c1 == 'D' && c2 == 'H') || (c1 == 'H' && c2 == 'D') ||
               (c1 == 'D' && c2 == 'S') || (c1 == 'S' && c2 == 'D')) {
        ++*c1p; ++*c2p; ++*ref;
        if (c1 == 'D' || c2 == 'D') o->cig[o->n_ops++] = 'D';
        if (c1 == 'S') ++*q1p; else if (c2 == 'S') ++*q2p;
    } else if ((c1 == 'M' && (c2 == 'H' || c2 == 'S')) || ((c1 == 'H' || c1 == 'S') && c2 == 'M') || (c1 == 'S' && c2 == 'H') ||
               (c1 == 'H' && c2 == 'S')) {
        if (c1 == 'M') { take(a, *q1p, o, 'M'); ++*q1p; if (c2 == 'S') ++*q2p; }
        else if (c2 == 'M') { take(b, *q2p, o, 'M'); ++*q2p; if (c1 == 'S') ++*q1p; }
        else if (c1 == 'S') ++*q1p;
        else ++*q2p;
        ++*ref; ++*c1p; ++*c2p;
    } else if ((c1 == 'I' && (c2 == 'H' || c2 == 'S')) || ((c1 == 'H' || c1 == 'S') && c2 == 'I')) {
        if (c1 == 'I') { take(a, *q1p, o, 'I'); ++*c1p; ++*q1p; }
        else { take(b, *q2p, o, 'I'); ++*c2p; ++*q2p; }
    } else return ORC_MERGE_UNDEFINED;
    return ORC_MERGE_OK;
}

#define STEP(call) do { int rc_ = (call); if (rc_) return rc_; } while (0)

/* mergeAlignmentRecordsSingle(ar, 1, 1) (:581-734): a = this, b = ar. */
static int merge_single(const mseg* a, const mseg* b, mout* o, int* start, int* end) {
    int s1 = a->start - lead(a), e1 = a->end + trail(a);
    int s2 = b->start - lead(b), e2 = b->end + trail(b);
    int q1 = 0, q2 = 0, c1 = 0, c2 = 0;
    o->n_ops = o->n_seq = 0;
    if (s1 <= s2 && e1 <= e2) {                         /* :637 */
        while (s1 < s2) STEP(no_overlap(a, o, &c1, &q1, &s1));
        while (s1 <= e1) STEP(overlap(a, b, o, &c1, &c2, &q1, &q2, &s1));
        while (s1 <= e2) STEP(no_overlap(b, o, &c2, &q2, &s1));
    } else if (s1 <= s2 && e1 >= e2) {                  /* :649 */
        while (s1 < s2) STEP(no_overlap(a, o, &c1, &q1, &s1));
        while (s1 <= e2) STEP(overlap(a, b, o, &c1, &c2, &q1, &q2, &s1));
        while (s1 <= e1) STEP(no_overlap(a, o, &c1, &q1, &s1));
    } else if (s1 >= s2 && e1 <= e2) {                  /* :661 */
        while (s2 < s1) STEP(no_overlap(b, o, &c2, &q2, &s2));
        while (s2 <= e1) STEP(overlap(a, b, o, &c1, &c2, &q1, &q2, &s2));
        while (s2 <= e2) STEP(no_overlap(b, o, &c2, &q2, &s2));
    } else {                                            /* :673 */
        while (s2 < s1) STEP(no_overlap(b, o, &c2, &q2, &s2));
        while (s2 <= e2) STEP(overlap(a, b, o, &c1, &c2, &q1, &q2, &s2));
        while (s2 <= e1) STEP(no_overlap(a, o, &c1, &q1, &s2));
    }
    *start = a->start < b->start ? a->start : b->start;  /* :686-687 */
    *end = a->end > b->end ? a->end : b->end;
    return ORC_MERGE_OK;
}

/* AlignmentRecord(alignments, id) (:211-249) for single-end members: member 0 is copied, members 1.. are folded in.
 * Arrays follow hc_reads' per-segment rows (row 2*r of read r). Output buffers hold at least the sum of the members'
 * CIGAR lengths. info[5] = status, start1, end1, unrolled CIGAR length, sequence length. */
int orc_merge_clique_se(const int32_t* seg_start, const int32_t* seg_end, const int64_t* cig_off, const int32_t* cig_len,
                        const int64_t* seq_off, const int32_t* seq_len, const uint8_t* cig, const uint8_t* seq,
                        const uint8_t* qual, const int32_t* members, int32_t m, uint8_t* out_cig, uint8_t* out_seq,
                        uint8_t* out_qual, int32_t* info) {
    int64_t cap = 0;
    for (int k = 0; k < m; ++k) cap += cig_len[2 * members[k]];
    uint8_t* tmp = (uint8_t*)malloc((size_t)(6 * (cap + 1)));
    mout buf[2];
    for (int k = 0; k < 2; ++k) { buf[k].cig = tmp + (3 * k) * (cap + 1); buf[k].seq = tmp + (3 * k + 1) * (cap + 1); buf[k].qual = tmp + (3 * k + 2) * (cap + 1); }
    int r0 = 2 * members[0];
    mseg acc = { seg_start[r0], seg_end[r0], cig_len[r0], seq_len[r0], cig + cig_off[r0], seq + seq_off[r0], qual + seq_off[r0] };
    int status = ORC_MERGE_OK;
    for (int k = 1; k < m && !status; ++k) {
        int r = 2 * members[k];
        mseg b = { seg_start[r], seg_end[r], cig_len[r], seq_len[r], cig + cig_off[r], seq + seq_off[r], qual + seq_off[r] };
        mout* o = &buf[k & 1];
        int st = 0, en = 0;
        status = merge_single(&acc, &b, o, &st, &en);
        acc.start = st; acc.end = en; acc.n_ops = o->n_ops; acc.n_seq = o->n_seq;
        acc.cig = o->cig; acc.seq = o->seq; acc.qual = o->qual;
    }
    info[0] = status;
    if (!status) {
        info[1] = acc.start; info[2] = acc.end; info[3] = acc.n_ops; info[4] = acc.n_seq;
        memcpy(out_cig, acc.cig, (size_t)acc.n_ops);
        memcpy(out_seq, acc.seq, (size_t)acc.n_seq);
        memcpy(out_qual, acc.qual, (size_t)acc.n_seq);
    } else info[1] = info[2] = info[3] = info[4] = 0;
    free(tmp);
    return status;
}
