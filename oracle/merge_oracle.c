/* oracle/merge_oracle.c -- TEST INFRASTRUCTURE ONLY (the checker; never on the product path).
 *
 * Plain-C restatement of the reference's super-read construction for SINGLE-END cliques (SURVEY.md 8(f) N2,
 * first slice): AlignmentRecord(alignments, clique_id) (src/AlignmentRecord.cpp:211-249) folding the members in
 * order through mergeAlignmentRecordsSingle(ar,1,1) (:581-734), built from noOverlapMerge (:1539-1576),
 * overlapMerge (:1578-1674) and computeEntry (:169-184) over the Phred update tables (:40-90).
 * Paired and mixed merges (mergeAlignmentRecordsPaired / Mixed, :734-1537): second half of this file (orc_merge_clique).
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

/* ======================================================================================================================
 * General super-read construction: AlignmentRecord(alignments, clique_id) (src/AlignmentRecord.cpp:211-249) for cliques
 * with paired-end members -- mergeAlignmentRecordsSingle(ar,i,j) for any (i,j) (:581-734), mergeAlignmentRecordsPaired
 * (:734-1293) and mergeAlignmentRecordsMixed (:1295-1537).
 *
 * Shape of the restatement (ours): the reference spells each read layout out as hand-unrolled while-loops; here a layout
 * is a row of a phase-program table (tail / overlap / seat / finish steps with symbolic bounds) run by one interpreter,
 * plus a result rule saying where the merged string goes and how start/end are combined. Every row cites the branch it
 * restates. The reference's cursor slip in the sixth paired layout (:1000: the overlap of this.segment1 with
 * ar.segment2 is called with q_c2_pos1, the query cursor of this.segment2) is reproduced: the bases of those columns and of
 * the closing tail on this.segment1 come from where that leaves the cursors.
 * PINNED against hcref_merge_clique (oracle/_ref) by tests/test_merge_oracle.py. */
typedef struct { int start, end, n_ops, n_seq; const uint8_t *cig, *seq, *qual; } gseg;
typedef struct { int single_end; gseg s[2]; } grec;

#define GBID(ext, side, seg, is_end) ((ext) * 8 + (side) * 4 + (seg) * 2 + (is_end))
enum {
    gAs1 = GBID(0, 0, 0, 0), gAe1 = GBID(0, 0, 0, 1), gAs2 = GBID(0, 0, 1, 0), gAe2 = GBID(0, 0, 1, 1),
    gBs1 = GBID(0, 1, 0, 0), gBe1 = GBID(0, 1, 0, 1), gBs2 = GBID(0, 1, 1, 0), gBe2 = GBID(0, 1, 1, 1),
    gxAs1 = GBID(1, 0, 0, 0), gxAe1 = GBID(1, 0, 0, 1), gxAs2 = GBID(1, 0, 1, 0), gxAe2 = GBID(1, 0, 1, 1),
    gxBs1 = GBID(1, 1, 0, 0), gxBe1 = GBID(1, 1, 0, 1), gxBs2 = GBID(1, 1, 1, 0), gxBe2 = GBID(1, 1, 1, 1)
};
enum { G_END = 0, G_START, G_TLT, G_OVL, G_OVL2, G_OVLQX, G_SEAT, G_FIN };
typedef struct { uint8_t op, x, y, a, b; } gphase;
/* where the merged string goes: segment (0/1); start/end = min / max of the two named bounds; single_end afterwards;
 * copy: 0 nothing, 1 segment 2 := ar.segment2 afterwards, 2 segment 1 := ar.segment1 BEFORE start2 is formed (:1177-1186) */
typedef struct { uint8_t seg, smin_a, smin_b, emax_a, emax_b, single, copy; } gresult;
typedef struct { gphase ph[12]; gresult res; } gprog;

#define GP_START(a) {G_START, 0, 0, a, 0}
#define GP_TLT(sd, sg, a) {G_TLT, sd, sg, a, 0}
#define GP_OVL(sa, sb, a) {G_OVL, sa, sb, a, 0}
#define GP_OVL2(sa, sb, a, b) {G_OVL2, sa, sb, a, b}
#define GP_OVLQX(sa, sb, a, b) {G_OVLQX, sa, sb, a, b}
#define GP_SEAT(sd, sg) {G_SEAT, sd, sg, 0, 0}
#define GP_FIN(a, b) {G_FIN, 0, 0, a, b}
#define GP_END {G_END, 0, 0, 0, 0}

/* rows 0..11: the twelve walked layouts of mergeAlignmentRecordsPaired in source order (:773 :819 :868 :907 :946 :989
 * :1027 :1069 :1118 :1161 :1204 :1249); rows 12,13: this paired x ar single (:1335, :1368); rows 14,15: this single x ar
 * paired (:1461, :1496). A = this, B = ar. */
static const gprog g_prog[16] = {
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs1), GP_OVL(0, 0, gAe1), GP_TLT(1, 0, gAs2), GP_SEAT(0, 1), GP_OVL(1, 0, gBe1), GP_TLT(0, 1, gBs2),
      GP_SEAT(1, 1), GP_OVL2(1, 1, gxBe2, gxAe2), GP_FIN(gxBe2, gxAe2), GP_END}, {0, gAs1, gBs1, gBe2, gAe2, 1, 0}},
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs1), GP_OVL(0, 0, gAe1), GP_TLT(1, 0, gAs2), GP_SEAT(0, 1), GP_OVL2(1, 0, gxBe1, gxAe2),
      GP_FIN(gxBe1, gxAe2), GP_END}, {0, gAs1, gBs1, gBe1, gAe2, 0, 1}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs2), GP_OVL(1, 0, gBe1), GP_TLT(0, 1, gBs2), GP_SEAT(1, 1), GP_OVL2(1, 1, gxBe2, gxAe2),
      GP_FIN(gxBe2, gxAe2), GP_END}, {1, gAs2, gBs1, gBe2, gAe2, 0, 0}},
    {{GP_START(gxAs2), GP_TLT(0, 1, gxBs1), GP_OVL(1, 0, gBe1), GP_TLT(0, 1, gBs2), GP_SEAT(1, 1), GP_OVL2(1, 1, gxAe2, gxBe2),
      GP_FIN(gxAe2, gxBe2), GP_END}, {1, gAs2, gBs1, gBe2, gAe2, 0, 0}},
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs1), GP_OVL(0, 0, gBe1), GP_TLT(0, 0, gBs2), GP_SEAT(1, 1), GP_OVL(0, 1, gAe1), GP_TLT(1, 1, gAs2),
      GP_SEAT(0, 1), GP_OVL2(1, 1, gxBe2, gxAe2), GP_FIN(gxBe2, gxAe2), GP_END}, {0, gAs1, gBs1, gBe2, gAe2, 1, 0}},
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs1), GP_OVL(0, 0, gBe1), GP_TLT(0, 0, gBs2), GP_SEAT(1, 1), GP_OVLQX(0, 1, gxAe1, gxBe2),
      GP_FIN(gxAe1, gxBe2), GP_END}, {0, gAs1, gBs1, gBe2, gAe1, 0, 0}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs1), GP_OVL(0, 0, gAe1), GP_TLT(1, 0, gAs2), GP_SEAT(0, 1), GP_OVL(1, 0, gBe1), GP_TLT(0, 1, gBs2),
      GP_SEAT(1, 1), GP_OVL2(1, 1, gxAe2, gxBe2), GP_FIN(gxAe2, gxBe2), GP_END}, {0, gAs1, gBs1, gBe2, gAe2, 1, 0}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs1), GP_OVL(0, 0, gAe1), GP_TLT(1, 0, gAs2), GP_SEAT(0, 1), GP_OVL2(1, 0, gxBe1, gxAe2),
      GP_FIN(gxBe1, gxAe2), GP_END}, {0, gAs1, gBs1, gBe1, gAe2, 0, 1}},
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs2), GP_OVL(0, 1, gAe1), GP_TLT(1, 1, gAs2), GP_SEAT(0, 1), GP_OVL2(1, 1, gxAe2, gxBe2),
      GP_FIN(gxAe2, gxBe2), GP_END}, {1, gBs1, gBs2, gBe2, gAe2, 0, 2}},
    {{GP_START(gxBs2), GP_TLT(1, 1, gxAs1), GP_OVL(0, 1, gAe1), GP_TLT(1, 1, gAs2), GP_SEAT(0, 1), GP_OVL2(1, 1, gxAe2, gxBe2),
      GP_FIN(gxAe2, gxBe2), GP_END}, {1, gBs1, gBs2, gBe2, gAe2, 0, 2}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs1), GP_OVL(0, 0, gBe1), GP_TLT(0, 0, gBs2), GP_SEAT(1, 1), GP_OVL(0, 1, gAe1), GP_TLT(1, 1, gAs2),
      GP_SEAT(0, 1), GP_OVL2(1, 1, gxBe2, gxAe2), GP_FIN(gxBe2, gxAe2), GP_END}, {0, gAs1, gBs1, gBe2, gAe2, 1, 0}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs1), GP_OVL(0, 0, gBe1), GP_TLT(0, 0, gBs2), GP_SEAT(1, 1), GP_OVL2(0, 1, gxAe1, gxBe2),
      GP_FIN(gxAe1, gxBe2), GP_END}, {0, gAs1, gBs1, gBe2, gAe1, 0, 0}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs1), GP_OVL(0, 0, gAe1), GP_TLT(1, 0, gAs2), GP_SEAT(0, 1), GP_OVL2(1, 0, gxAe2, gxBe1),
      GP_FIN(gxAe2, gxBe1), GP_END}, {0, gAs1, gBs1, gBe1, gAe2, 1, 0}},
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs1), GP_OVL(0, 0, gAe1), GP_TLT(1, 0, gAs2), GP_SEAT(0, 1), GP_OVL2(1, 0, gxAe2, gxBe1),
      GP_FIN(gxAe2, gxBe1), GP_END}, {0, gAs1, gBs1, gBe1, gAe2, 1, 0}},
    {{GP_START(gxAs1), GP_TLT(0, 0, gxBs1), GP_OVL(0, 0, gBe1), GP_TLT(0, 0, gBs2), GP_SEAT(1, 1), GP_OVL2(0, 1, gxBe2, gxAe1),
      GP_FIN(gxBe2, gxAe1), GP_END}, {0, gAs1, gBs1, gBe2, gAe1, 1, 0}},
    {{GP_START(gxBs1), GP_TLT(1, 0, gxAs1), GP_OVL(0, 0, gBe1), GP_TLT(0, 0, gBs2), GP_SEAT(1, 1), GP_OVL2(0, 1, gxBe2, gxAe1),
      GP_FIN(gxBe2, gxAe1), GP_END}, {0, gAs1, gBs1, gBe2, gAe1, 1, 0}},
};

typedef struct {
    const gseg* seg[2][2];
    int c[2][2], q[2][2];
    int v[16];
    int ref;
    mout* o;
} gwalk;

static int g_lead(const gseg* s) { int k = 0; while (k < s->n_ops && (s->cig[k] == 'S' || s->cig[k] == 'H')) ++k; return k; }
static int g_trail(const gseg* s) { int k = 0; while (k < s->n_ops && (s->cig[s->n_ops - 1 - k] == 'S' || s->cig[s->n_ops - 1 - k] == 'H')) ++k; return k; }

/* noOverlapMerge (:1539-1576) with the reference's bounds-checked sequence access (.at(): past the end it throws) */
static int g_tail_step(gwalk* w, int side, int sg) {
    const gseg* s = w->seg[side][sg];
    int* c = &w->c[side][sg]; int* q = &w->q[side][sg];
    if (*c >= s->n_ops) return ORC_MERGE_OVERRUN;
    uint8_t op = s->cig[*c];
    if (op == 'M' || op == 'I') {
        if (*q >= s->n_seq) return ORC_MERGE_OVERRUN;
        w->o->seq[w->o->n_seq] = s->seq[*q]; w->o->qual[w->o->n_seq++] = s->qual[*q]; w->o->cig[w->o->n_ops++] = op;
        ++*q; ++*c;
        if (op == 'M') ++w->ref;
    } else if (op == 'D') { w->o->cig[w->o->n_ops++] = 'D'; ++w->ref; ++*c; }
    else if (op == 'S') { ++w->ref; ++*q; ++*c; }
    else if (op == 'H') { ++w->ref; ++*c; }
    else return ORC_MERGE_UNDEFINED;
    return ORC_MERGE_OK;
}

/* overlapMerge (:1578-1674): this.segment sa against ar.segment sb; qa = the query cursor the call was given for `this` */
static int g_ovl_step(gwalk* w, int sa, int sb, int* qa) {
    const gseg* a = w->seg[0][sa]; const gseg* b = w->seg[1][sb];
    int* c1 = &w->c[0][sa]; int* c2 = &w->c[1][sb]; int* q2 = &w->q[1][sb];
    mout* o = w->o;
    if (*c1 >= a->n_ops || *c2 >= b->n_ops) return ORC_MERGE_OVERRUN;
    uint8_t x = a->cig[*c1], y = b->cig[*c2];
#define G_NEED_A() do { if (*qa >= a->n_seq) return ORC_MERGE_OVERRUN; } while (0)
#define G_NEED_B() do { if (*q2 >= b->n_seq) return ORC_MERGE_OVERRUN; } while (0)
    if ((x == 'M' && y == 'M') || (x == 'S' && y == 'S') || (x == 'I' && y == 'I')) {
        if (x != 'S') {
            uint8_t e[2];
            G_NEED_A(); G_NEED_B();
            orc_compute_entry(a->seq[*qa], a->qual[*qa], b->seq[*q2], b->qual[*q2], e);
            o->seq[o->n_seq] = e[0]; o->qual[o->n_seq++] = e[1]; o->cig[o->n_ops++] = x;
        }
        if (x != 'I') ++w->ref;
        ++*qa; ++*q2; ++*c1; ++*c2;
    } else if ((x == 'D' && y == 'D') || (x == 'H' && y == 'H') || (x == 'D' && y == 'H') || (x == 'H' && y == 'D') ||
               (x == 'D' && y == 'S') || (x == 'S' && y == 'D')) {
        ++*c1; ++*c2; ++w->ref;
        if (x == 'D' || y == 'D') o->cig[o->n_ops++] = 'D';
        if (x == 'S') ++*qa; else if (y == 'S') ++*q2;
    } else if ((x == 'M' && (y == 'H' || y == 'S')) || ((x == 'H' || x == 'S') && y == 'M') || (x == 'S' && y == 'H') || (x == 'H' && y == 'S')) {
        if (x == 'M') { G_NEED_A(); o->seq[o->n_seq] = a->seq[*qa]; o->qual[o->n_seq++] = a->qual[*qa]; o->cig[o->n_ops++] = 'M'; ++*qa; if (y == 'S') ++*q2; }
        else if (y == 'M') { G_NEED_B(); o->seq[o->n_seq] = b->seq[*q2]; o->qual[o->n_seq++] = b->qual[*q2]; o->cig[o->n_ops++] = 'M'; ++*q2; if (x == 'S') ++*qa; }
        else if (x == 'S') ++*qa;
        else ++*q2;
        ++w->ref; ++*c1; ++*c2;
    } else if ((x == 'I' && (y == 'H' || y == 'S')) || ((x == 'H' || x == 'S') && y == 'I')) {
        if (x == 'I') { G_NEED_A(); o->seq[o->n_seq] = a->seq[*qa]; o->qual[o->n_seq++] = a->qual[*qa]; o->cig[o->n_ops++] = 'I'; ++*c1; ++*qa; }
        else { G_NEED_B(); o->seq[o->n_seq] = b->seq[*q2]; o->qual[o->n_seq++] = b->qual[*q2]; o->cig[o->n_ops++] = 'I'; ++*c2; ++*q2; }
    } else return ORC_MERGE_UNDEFINED;
#undef G_NEED_A
#undef G_NEED_B
    return ORC_MERGE_OK;
}

static void g_bounds(gwalk* w, const grec* A, const grec* B) {
    const grec* r[2] = {A, B};
    for (int side = 0; side < 2; ++side)
        for (int sg = 0; sg < 2; ++sg) {
            const gseg* s = &r[side]->s[sg];
            w->seg[side][sg] = s;
            w->c[side][sg] = w->q[side][sg] = 0;
            w->v[GBID(0, side, sg, 0)] = s->start; w->v[GBID(0, side, sg, 1)] = s->end;
            w->v[GBID(1, side, sg, 0)] = s->start - g_lead(s); w->v[GBID(1, side, sg, 1)] = s->end + g_trail(s);
        }
}

static int g_run(gwalk* w, const gphase* pg) {
    for (int k = 0; k < 12 && pg[k].op != G_END; ++k) {
        const gphase ph = pg[k];
        switch (ph.op) {
            case G_START: w->ref = w->v[ph.a]; break;
            case G_TLT: while (w->ref < w->v[ph.a]) STEP(g_tail_step(w, ph.x, ph.y)); break;
            case G_SEAT: {                                     /* computeSOffset (:136-145) */
                const gseg* s = w->seg[ph.x][ph.y];
                for (int t = 0; t < s->n_ops; ++t) {
                    if (s->cig[t] == 'S') { w->c[ph.x][ph.y]++; w->q[ph.x][ph.y]++; }
                    else if (s->cig[t] == 'H') w->c[ph.x][ph.y]++;
                    else break;
                }
                break;
            }
            case G_OVL: while (w->ref <= w->v[ph.a]) STEP(g_ovl_step(w, ph.x, ph.y, &w->q[0][ph.x])); break;
            case G_OVL2: while (w->ref <= w->v[ph.a] && w->ref <= w->v[ph.b]) STEP(g_ovl_step(w, ph.x, ph.y, &w->q[0][ph.x])); break;
            case G_OVLQX: while (w->ref <= w->v[ph.a] && w->ref <= w->v[ph.b]) STEP(g_ovl_step(w, ph.x, ph.y, &w->q[0][1 - ph.x])); break;
            case G_FIN: {
                int which = -1;
                if (w->ref - 1 == w->v[ph.a]) which = ph.b; else if (w->ref - 1 == w->v[ph.b]) which = ph.a;
                if (which >= 0) { const int sd = (which >> 2) & 1, sg = (which >> 1) & 1; while (w->ref <= w->v[which]) STEP(g_tail_step(w, sd, sg)); }
                break;
            }
        }
    }
    return ORC_MERGE_OK;
}

/* the running super-read: two segments with their own storage */
typedef struct { grec r; uint8_t* buf[2][3]; int64_t cap; } gacc;

static void g_store(gacc* a, int sg, const mout* o, int start, int end) {
    memcpy(a->buf[sg][0], o->cig, (size_t)o->n_ops); memcpy(a->buf[sg][1], o->seq, (size_t)o->n_seq); memcpy(a->buf[sg][2], o->qual, (size_t)o->n_seq);
    gseg* s = &a->r.s[sg];
    s->start = start; s->end = end; s->n_ops = o->n_ops; s->n_seq = o->n_seq;
    s->cig = a->buf[sg][0]; s->seq = a->buf[sg][1]; s->qual = a->buf[sg][2];
}
static void g_copy_seg(gacc* a, int sg, const gseg* src) {
    memcpy(a->buf[sg][0], src->cig, (size_t)src->n_ops); memcpy(a->buf[sg][1], src->seq, (size_t)src->n_seq); memcpy(a->buf[sg][2], src->qual, (size_t)src->n_seq);
    gseg* s = &a->r.s[sg];
    *s = *src;
    s->cig = a->buf[sg][0]; s->seq = a->buf[sg][1]; s->qual = a->buf[sg][2];
}
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* mergeAlignmentRecordsSingle(ar, i, j) (:581-734), i/j = 0/1 for segment 1/2 */
static int g_single(gacc* acc, const grec* ar, int i, int j, mout* o) {
    mseg a = { acc->r.s[i].start, acc->r.s[i].end, acc->r.s[i].n_ops, acc->r.s[i].n_seq, acc->r.s[i].cig, acc->r.s[i].seq, acc->r.s[i].qual };
    mseg b = { ar->s[j].start, ar->s[j].end, ar->s[j].n_ops, ar->s[j].n_seq, ar->s[j].cig, ar->s[j].seq, ar->s[j].qual };
    /* merge_single reads the sequences unchecked: make sure the CIGARs do not consume more than the reads hold */
    int ca = 0, cb = 0;
    for (int t = 0; t < a.n_ops; ++t) ca += (a.cig[t] == 'M' || a.cig[t] == 'I' || a.cig[t] == 'S');
    for (int t = 0; t < b.n_ops; ++t) cb += (b.cig[t] == 'M' || b.cig[t] == 'I' || b.cig[t] == 'S');
    if (ca > a.n_seq || cb > b.n_seq) return ORC_MERGE_OVERRUN;
    int st, en;
    STEP(merge_single(&a, &b, o, &st, &en));
    if (i == 0 && j == 0) { g_store(acc, 0, o, imin(a.start, b.start), imax(a.end, b.end)); acc->r.single_end = 1; }      /* :685-697 */
    else if (i == 0) g_store(acc, 1, o, imin(acc->r.s[0].start, b.start), imax(acc->r.s[0].end, b.end));                     /* :698-709: from start1 / end1 */
    else { g_store(acc, 1, o, imin(a.start, b.start), imax(a.end, b.end)); acc->r.single_end = 0; }                         /* :710-729 */
    return ORC_MERGE_OK;
}

/* how often each layout ran since the last orc_merge_layout_counts call (test coverage bookkeeping): 0..15 programs,
 * 16 two singles (:766), 17 no paired layout matched, 18..21 the mixed single-segment cases A B E F, 22 plain single/single */
static int64_t g_counts[24];
void orc_merge_layout_counts(int64_t* out24) { memcpy(out24, g_counts, sizeof g_counts); memset(g_counts, 0, sizeof g_counts); }

static int g_program(gacc* acc, const grec* ar, int k, mout* o) {
    gwalk w;
    g_counts[k]++;
    g_bounds(&w, &acc->r, ar);
    w.o = o; w.ref = 0;
    o->n_ops = o->n_seq = 0;
    STEP(g_run(&w, g_prog[k].ph));
    const gresult rs = g_prog[k].res;
    if (rs.copy == 2) g_copy_seg(acc, 0, &ar->s[0]);
    /* bounds as they are when the assignment runs: after copy 2 this->start1 is ar's (:1186 min(this->start1, ar.getStart2())) */
    int va = w.v[rs.smin_a], vb = w.v[rs.smin_b], ea = w.v[rs.emax_a], eb = w.v[rs.emax_b];
    g_store(acc, rs.seg, o, imin(va, vb), imax(ea, eb));
    acc->r.single_end = rs.single;
    if (rs.copy == 1) g_copy_seg(acc, 1, &ar->s[1]);
    return ORC_MERGE_OK;
}

static int g_paired(gacc* acc, const grec* ar, mout* o) {
    gwalk w;
    g_bounds(&w, &acc->r, ar);
    const int* v = w.v;
    if (v[gAe1] < v[gBs2] && v[gAs2] > v[gBe1]) {                                        /* :766 */
        g_counts[16]++;
        STEP(g_single(acc, ar, 0, 0, o));
        return g_single(acc, ar, 1, 1, o);
    }
    int k = -1;
    if      (v[gxAs1] <= v[gxBs1] && v[gAe1] >= v[gBs1] && v[gAs2] <= v[gBe1] && v[gAe2] >= v[gBs2]) k = 0;
    else if (v[gxAs1] <= v[gxBs1] && v[gAe1] >= v[gBs1] && v[gAs2] <= v[gBe1] && v[gAe2] <  v[gBs2]) k = 1;
    else if (v[gxBs1] <= v[gxAs2] && v[gAe1] <  v[gBs1] && v[gAs2] <= v[gBe1] && v[gAe2] >= v[gBs2]) k = 2;
    else if (v[gxAs2] <= v[gxBs1] && v[gAe1] <  v[gBs1] && v[gAs2] <= v[gBe1] && v[gAe2] >= v[gBs2]) k = 3;
    else if (v[gxAs1] <= v[gxBs1] && v[gAs1] <= v[gBe1] && v[gAe1] >= v[gBs2] && v[gAs2] <= v[gBe2]) k = 4;
    else if (v[gxAs1] <= v[gxBs1] && v[gAs2] >  v[gBe2] && v[gAe1] >= v[gBs2] && v[gAs1] <= v[gBe1]) k = 5;
    else if (v[gxBs1] <= v[gxAs1] && v[gAe1] >= v[gBs1] && v[gAs2] <= v[gBe1] && v[gAe2] >= v[gBs2]) k = 6;
    else if (v[gxBs1] <= v[gxAs1] && v[gAe1] >= v[gBs1] && v[gAs2] <= v[gBe1] && v[gAe2] <  v[gBs2]) k = 7;
    else if (v[gxAs1] <= v[gxBs2] && v[gAs1] >  v[gBe1] && v[gAe1] >= v[gBs2] && v[gAs2] <= v[gBe2]) k = 8;
    else if (v[gxBs2] <= v[gxAs1] && v[gBe1] <  v[gAs1] && v[gAe1] >= v[gBs2] && v[gAs2] <= v[gBe2]) k = 9;
    else if (v[gxBs1] <= v[gxAs1] && v[gAs1] <= v[gBe1] && v[gAe1] >= v[gBs2] && v[gAs2] <= v[gBe2]) k = 10;
    else if (v[gxBs1] <= v[gxAs1] && v[gAs2] >  v[gBe2] && v[gAs1] <= v[gBe1] && v[gAe1] >= v[gBs2]) k = 11;
    if (k < 0) { g_counts[17]++; return ORC_MERGE_OK; }   /* no layout matches: `this` stays as it is */
    return g_program(acc, ar, k, o);
}

static int g_mixed(gacc* acc, const grec* ar, mout* o) {
    gwalk w;
    g_bounds(&w, &acc->r, ar);
    const int* v = w.v;
    if (ar->single_end) {                                  /* this paired, ar single (:1296-1405) */
        if (v[gBe1] < v[gAs2]) { g_counts[18]++; STEP(g_single(acc, ar, 0, 0, o)); acc->r.single_end = 0; return ORC_MERGE_OK; }
        if (v[gBs1] > v[gAe1]) { g_counts[19]++; return g_single(acc, ar, 1, 0, o); }
        return g_program(acc, ar, v[gxBs1] <= v[gxAs1] ? 12 : 13, o);
    }
    /* this single, ar paired (:1406-1536) */
    if (v[gAe1] < v[gBs2]) { g_counts[20]++; STEP(g_single(acc, ar, 0, 0, o)); g_copy_seg(acc, 1, &ar->s[1]); acc->r.single_end = 0; return ORC_MERGE_OK; }
    if (v[gAs1] > v[gBe1]) { g_counts[21]++; STEP(g_single(acc, ar, 0, 1, o)); g_copy_seg(acc, 0, &ar->s[0]); acc->r.single_end = 0; return ORC_MERGE_OK; }
    return g_program(acc, ar, v[gxAs1] <= v[gxBs1] ? 14 : 15, o);
}

/* AlignmentRecord(alignments, id) for any mix of single-end and paired-end members. Arrays follow hc_reads (segment s of
 * read r = row 2r+s). out_*: two regions of `cap` bytes each (segment 1 at 0, segment 2 at cap), cap >= the summed CIGAR
 * lengths of all members' segments. info[10] = status, single_end, start1, end1, n_ops1, n_seq1, start2, end2, n_ops2, n_seq2. */
int orc_merge_clique(const uint8_t* single_end, const int32_t* seg_start, const int32_t* seg_end, const int64_t* cig_off,
                     const int32_t* cig_len, const int64_t* seq_off, const int32_t* seq_len, const uint8_t* cig, const uint8_t* seq,
                     const uint8_t* qual, const int32_t* members, int32_t m, int64_t cap, uint8_t* out_cig, uint8_t* out_seq,
                     uint8_t* out_qual, int32_t* info) {
    gacc acc;
    memset(&acc, 0, sizeof acc);
    acc.cap = cap + 16;
    uint8_t* pool = (uint8_t*)malloc((size_t)(9 * acc.cap));
    for (int sg = 0; sg < 2; ++sg) for (int t = 0; t < 3; ++t) acc.buf[sg][t] = pool + (size_t)(sg * 3 + t) * acc.cap;
    mout o = { pool + 6 * acc.cap, pool + 7 * acc.cap, pool + 8 * acc.cap, 0, 0 };
    int status = ORC_MERGE_OK;
    for (int k = 0; k < m && !status; ++k) {
        const int r = members[k];
        grec al;
        al.single_end = single_end[r] != 0;
        for (int sg = 0; sg < 2; ++sg) {
            const int row = 2 * r + sg;
            const int have = sg == 0 || !al.single_end;
            gseg s = { have ? seg_start[row] : 0, have ? seg_end[row] : 0, have ? cig_len[row] : 0, have ? seq_len[row] : 0,
                       cig + cig_off[row], seq + seq_off[row], qual + seq_off[row] };
            al.s[sg] = s;
        }
        if (k == 0) {                                          /* :216-232 */
            acc.r.single_end = al.single_end;
            g_copy_seg(&acc, 0, &al.s[0]);
            if (!al.single_end) g_copy_seg(&acc, 1, &al.s[1]);
            continue;
        }
        if (acc.r.single_end && al.single_end) { g_counts[22]++; status = g_single(&acc, &al, 0, 0, &o); }
        else if (!acc.r.single_end && !al.single_end) status = g_paired(&acc, &al, &o);
        else status = g_mixed(&acc, &al, &o);
    }
    memset(info, 0, 10 * sizeof(int32_t));
    info[0] = status;
    if (!status) {
        info[1] = acc.r.single_end;
        for (int sg = 0; sg < (acc.r.single_end ? 1 : 2); ++sg) {
            const gseg* s = &acc.r.s[sg];
            info[2 + 4 * sg] = s->start; info[3 + 4 * sg] = s->end; info[4 + 4 * sg] = s->n_ops; info[5 + 4 * sg] = s->n_seq;
            memcpy(out_cig + sg * cap, s->cig, (size_t)s->n_ops);
            memcpy(out_seq + sg * cap, s->seq, (size_t)s->n_seq);
            memcpy(out_qual + sg * cap, s->qual, (size_t)s->n_seq);
        }
    }
    free(pool);
    return status;
}
