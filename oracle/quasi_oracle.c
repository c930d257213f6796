/* oracle/quasi_oracle.c -- CPU ORACLE, TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's older scorer QuasispeciesEdgeCalculator::edgeBetween
 * (/root/reference/src/QuasispeciesEdgeCalculator.cpp:55-70, computeOverlap :72-131, overlapSize :133-179,
 * singleOverlap :181-426, computeOffsets :428-493), the checker of hc_quasi_score_pairs.
 * Only tests/ may load it; the product (haploclique_b200/) never does.
 *
 * Parity status: PINNED against oracle/_ref/libhcref.so (the unmodified reference compiled from /root/reference):
 * tests/test_quasi_oracle.py compares q = computeOverlap(...) and the boolean on every ordered pair of the bundled
 * read sets; digests under tests/golden/quasi_golden.json. The reference itself has no test for this class.
 *
 * Differences in shape (not in result): the reference keeps separate cursors into the CIGAR and into the
 * base/quality strings; so does this file. Where the reference would spin forever (the clip-skipping `for(;;)` at
 * :258-267 does not terminate once either cursor reaches its CIGAR's end) or index past a vector, the pair is
 * reported with status 3 (undefined) instead.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "orc_types.h"

typedef struct {
    double cutoff_cliques, cutoff_mixed, cutoff_single;   /* EDGE_QUASI_CUTOFF{,_MIXED,_SINGLE} (:36-38) */
    double overlap_cliques, overlap_single;               /* MIN_OVERLAP_{CLIQUES,SINGLE} (:39-40) */
    int32_t frameshift_merge;                             /* FRAMESHIFT_MERGE (:41) */
    int32_t simpson_len;
    const double* simpson;                                /* dense stand-in for map<int,double>: 0.0 = key absent */
} qorc_params;

typedef struct {
    int32_t status;          /* 0 = defined, 3 = the reference would hang or read out of bounds */
    int32_t edge;            /* q >= cutoff (:69) */
    int32_t n_single;        /* how many singleOverlap calls computeOverlap made (0, 1 or 2) */
    int32_t subnormal;       /* some singleOverlap had exp(sum) below DBL_MIN: q has lost its precision (:425) */
    double q;                /* computeOverlap's return value */
    double cutoff;
} qorc_result;

/* overlapSize (:133-179): an int difference returned as float */
static float overlap_size(int e1, int e2, int s1, int s2) {
    if (s1 == s2 && e1 == e2) return (float)(e1 - s1);
    if (s1 == s2) return (float)(e1 < e2 ? e1 - s1 : e2 - s2);
    if (e1 == e2) return (float)(s1 > s2 ? e1 - s1 : e2 - s2);
    if (e1 < e2) return (float)(s1 < s2 ? e1 - s2 : e1 - s1);
    return (float)(s2 < s1 ? e2 - s1 : e2 - s2);
}

/* computeOffsets (:428-493) */
static void compute_offsets(int s1, int s2, int e1, int e2, int* offset1, int* offset2, int* overlap) {
    const int x_l1 = e1 - s1, x_l2 = e2 - s2;
    if (s1 == s2 && e1 == e2) { *overlap = e1 - s1; *offset1 = 0; *offset2 = 0; }
    else if (s1 == s2) { *overlap = e1 < e2 ? e1 - s1 : e2 - s2; *offset1 = 0; *offset2 = 0; }
    else if (e1 == e2) {
        if (s1 > s2) { *overlap = e1 - s1; *offset1 = 0; *offset2 = x_l2 - *overlap; }
        else { *overlap = e2 - s2; *offset1 = x_l1 - *overlap; *offset2 = 0; }
    } else if (e1 < e2) {
        if (s1 < s2) { *overlap = e1 - s2; *offset1 = x_l1 - *overlap; *offset2 = 0; }
        else { *overlap = e1 - s1; *offset1 = 0; *offset2 = e2 - (e2 - e1) - *overlap - s2; }
    } else {
        if (s2 < s1) { *overlap = e2 - s1; *offset1 = 0; *offset2 = x_l2 - *overlap; }
        else { *overlap = e2 - s2; *offset1 = e1 - (e1 - e2) - *overlap - s1; *offset2 = 0; }
    }
}

/* ShortDnaSequence::qualityCorrect (src/ShortDnaSequence.cpp:117-119): 1 - pow(10, -(q-33)/10.0) */
static double quality_correct(char qual) { return 1 - pow(10, -(qual - 33) / 10.0); }

/* SIMPSON_MAP.find(pos) != end ? log(at(pos)) : nothing (:290-294, :388-392, :407-411) */
static void add_simpson(const qorc_params* P, int pos, double* acc) {
    if (P->simpson && pos >= 0 && pos < P->simpson_len && P->simpson[pos] != 0.0) *acc += log(P->simpson[pos]);
}

/* singleOverlap (:181-426). A/B: the chosen segments ("strains") of ap1/ap2. *undef is set where the reference would
 * not terminate or would index out of range. */
static double single_overlap(const orc_seg* A, const orc_seg* B, const qorc_params* P, double MIN_OVERLAP, double cutoff,
                             int* undef, int* subnormal) {
    const orc_seg *S1 = A, *S2 = B;
    if (S2->start < S1->start) { const orc_seg* t = S1; S1 = S2; S2 = t; }                 /* :215-220 */
    const int s1 = S1->start, e1 = S1->end, s2 = S2->start, e2 = S2->end;
    const char *cig1 = S1->cig, *cig2 = S2->cig;
    const unsigned long n1 = (unsigned long)S1->ncig, n2 = (unsigned long)S2->ncig;
    const unsigned long nq1 = (unsigned long)S1->nq, nq2 = (unsigned long)S2->nq;
    int offset1 = 0, offset2 = 0, overlap = 0;
    compute_offsets(s1, s2, e1, e2, &offset1, &offset2, &overlap);                          /* :227 */
    if (overlap < MIN_OVERLAP) return 0;                                                    /* :235 */
    double overlap_probability = 0.0, total_size = 0;
    if ((unsigned long)(long)offset1 >= nq1 || (unsigned long)(long)offset2 >= nq2) return 0;   /* :249, int -> size_t */
    const int perfect = cutoff == 1.0;                                                      /* :253 */
    unsigned long j_seq = 0, j_seq2 = 0, j_cigar = 0, j_cigar2 = 0;
    int j = 0, j2 = 0;
    for (;;) {                                                                              /* :258-267 */
        if (!(j_cigar < n1 && j_cigar2 < n2)) { *undef = 1; return 0; }                     /* the reference spins here */
        if (cig1[j_cigar] != 'S' && cig2[j_cigar2] != 'S' && cig1[j_cigar] != 'H' && cig2[j_cigar2] != 'H') break;
        if (cig1[j_cigar] == 'S') { j_cigar++; j_seq++; }
        if (cig2[j_cigar2] == 'S') { j_cigar2++; j_seq2++; }
        if (j_cigar >= n1 || j_cigar2 >= n2) { *undef = 1; return 0; }                      /* next two reads are out of range */
        if (cig1[j_cigar] == 'H') j_cigar++;
        if (cig2[j_cigar2] == 'H') j_cigar2++;
    }
    for (int j_prefix = 0;;) {                                                              /* :268-295 */
        if (j_prefix == offset1) break;
        if (j_cigar >= n1) { j_cigar++; j_prefix++; continue; }
        switch (cig1[j_cigar]) {
            case 'M':
                if (!perfect) add_simpson(P, s1 + j, &overlap_probability);
                j++;
                /* fall through, as the reference does */
            case 'I':
                j_seq++;
            default: break;
        }
        j_cigar++;
        j_prefix++;
    }
    while (j_cigar < n1 && j_cigar2 < n2 && cig1[j_cigar] != 'S' && cig2[j_cigar2] != 'S' && j_seq < nq1 && j_seq2 < nq2) {   /* :296 */
        const char c1 = cig1[j_cigar], c2 = cig2[j_cigar2];
        if (perfect) {                                                                      /* :298-306 */
            if (S1->seq[j_seq] != S2->seq[j_seq2]) return 0;
            j_cigar++; j_cigar2++; j_seq++; j_seq2++;
        } else if (c1 == c2 && (c1 == 'M' || c1 == 'I')) {                                  /* :307-343 */
            const double q_x1 = quality_correct(S1->qual[j_seq]), q_x2 = quality_correct(S2->qual[j_seq2]);
            const double anti_q_x1 = (1.0 - q_x1) / 3.0, anti_q_x2 = (1.0 - q_x2) / 3.0;
            static const char alphabet[4] = {'A', 'C', 'G', 'T'};
            double sum = 0.0;
            for (int k = 0; k < 4; ++k)
                sum += ((S1->seq[j_seq] == alphabet[k] ? q_x1 : anti_q_x1) * (S2->seq[j_seq2] == alphabet[k] ? q_x2 : anti_q_x2));
            overlap_probability += log(sum);
            total_size++;
            j_cigar++; j_cigar2++; j_seq++; j_seq2++;
            if (j_cigar < n1 && cig1[j_cigar] == 'M') { j++; j2++; }
        } else if (c1 == c2 && c1 == 'D') {                                                 /* :344-346 */
            j_cigar++; j_cigar2++;
        } else if (P->frameshift_merge) {                                                   /* :347-381 */
            if (c2 == 'M') {
                if (c1 == 'D') { j_cigar++; j_cigar2++; j_seq2++; j2++; }
                else if (c1 == 'I') { j_cigar++; j_seq++; }
                else { *undef = 1; return 0; }          /* `default: break`: nothing advances, the reference spins */
            } else if (c1 == 'M') {
                if (c2 == 'D') { j_cigar2++; j_cigar++; j_seq++; j++; }
                else if (c2 == 'I') { j_cigar2++; j_seq2++; }
                else { *undef = 1; return 0; }          /* no case matches: the reference spins */
            } else {
                return 0;
            }
        } else {
            return 0;
        }
    }
    if (!perfect) {                                                                         /* :384-420 */
        if (j_cigar < n1 && cig1[j_cigar] != 'S') {
            while (j_cigar < n1) {
                if (cig1[j_cigar] == 'S') break;
                if (cig1[j_cigar] == 'M') { add_simpson(P, s1 + j, &overlap_probability); j++; }
                j_cigar++;
            }
        } else if (j_cigar2 < n2 && cig2[j_cigar2] != 'S') {
            while (j_cigar2 < n2) {
                if (cig2[j_cigar2] == 'S') break;
                if (cig2[j_cigar2] == 'M') { add_simpson(P, s2 + j2, &overlap_probability); j2++; }
                j_cigar2++;
            }
        }
    }
    if (perfect) return 1;
    if (overlap_probability < -708.3964185322641) *subnormal = 1;                           /* log(DBL_MIN) */
    return pow(exp(overlap_probability), 1.0 / total_size);                                 /* :425 */
}

/* computeOverlap (:72-131) */
static double compute_overlap(const orc_read* a, const orc_read* b, const qorc_params* P, double cutoff, int* undef, int* ncalls,
                              int* subnormal) {
    const double MIN_OVERLAP = (a->is_clique && b->is_clique) ? P->overlap_cliques : P->overlap_single;   /* :74-79 */
    const orc_seg *a1 = &a->s[0], *a2 = &a->s[1], *b1 = &b->s[0], *b2 = &b->s[1];
#define MINI(x, y) ((x) < (y) ? (x) : (y))
    *ncalls = 0;
    if (a->single_end && b->single_end) {                                                   /* :81-86 */
        const int read_size1 = MINI(a1->end - a1->start, b1->end - b1->start);
        float overlap_size1 = overlap_size(a1->end, b1->end, a1->start, b1->start);
        if (MIN_OVERLAP <= 1) overlap_size1 /= (float)read_size1;
        if (overlap_size1 < MIN_OVERLAP) return 0;
        *ncalls = 1;
        return single_overlap(a1, b1, P, MIN_OVERLAP, cutoff, undef, subnormal);
    } else if (a->single_end || b->single_end) {                                            /* :87-117 */
        const orc_seg *x1, *y1, *x2, *y2;          /* the two singleOverlap calls: (x1,y1) and (x2,y2) */
        int read_size1, read_size2;
        float overlap_size1, overlap_size2;
        if (a->single_end) {
            read_size1 = MINI(a1->end - a1->start, b1->end - b1->start);
            read_size2 = MINI(a1->end - a1->start, b2->end - b2->start);
            overlap_size1 = overlap_size(a1->end, b1->end, a1->start, b1->start);
            overlap_size2 = overlap_size(a1->end, b2->end, a1->start, b2->start);
            x1 = a1; y1 = b1; x2 = a1; y2 = b2;
        } else {
            read_size1 = MINI(a1->end - a1->start, b1->end - b1->start);
            read_size2 = MINI(a2->end - a2->start, b1->end - b1->start);
            overlap_size1 = overlap_size(a1->end, b1->end, a1->start, b1->start);
            overlap_size2 = overlap_size(a2->end, b1->end, a2->start, b1->start);
            x1 = a1; y1 = b1; x2 = a2; y2 = b1;
        }
        if (MIN_OVERLAP <= 1) { overlap_size1 /= (float)read_size1; overlap_size2 /= (float)read_size2; }
        if ((overlap_size1 >= MIN_OVERLAP && overlap_size2 >= MIN_OVERLAP) ||
            (overlap_size1 + overlap_size2 >= MIN_OVERLAP && overlap_size1 > 0 && overlap_size2 > 0)) {
            *ncalls = 2;
            const double p = single_overlap(x1, y1, P, MIN_OVERLAP, cutoff, undef, subnormal);
            const double q = single_overlap(x2, y2, P, MIN_OVERLAP, cutoff, undef, subnormal);
            return p * q;
        }
        return 0;
    } else {                                                                                /* :118-129 */
        const int read_size1 = MINI(a1->end - a1->start, b1->end - b1->start);
        float overlap_size1 = overlap_size(a1->end, b1->end, a1->start, b1->start);
        if (MIN_OVERLAP <= 1) overlap_size1 /= (float)read_size1;
        if (overlap_size1 < MIN_OVERLAP) return 0;
        const int read_size2 = MINI(a2->end - a2->start, b2->end - b2->start);
        float overlap_size2 = overlap_size(a2->end, b2->end, a2->start, b2->start);
        if (MIN_OVERLAP <= 1) overlap_size2 /= (float)read_size2;
        if (overlap_size2 < MIN_OVERLAP) return 0;
        *ncalls = 2;
        const double p = single_overlap(a1, b1, P, MIN_OVERLAP, cutoff, undef, subnormal);
        const double q = single_overlap(a2, b2, P, MIN_OVERLAP, cutoff, undef, subnormal);
        return p * q;
    }
#undef MINI
}

/* edgeBetween (:55-70) without the same-name shortcut (:56): the batch holds one record per name */
void qorc_edge(const orc_read* a, const orc_read* b, const qorc_params* P, qorc_result* out) {
    double cutoff;
    if (a->is_clique && b->is_clique) cutoff = P->cutoff_cliques;
    else if (a->is_clique || b->is_clique) cutoff = P->cutoff_mixed;
    else cutoff = P->cutoff_single;
    int undef = 0, ncalls = 0, subnormal = 0;
    const double q = compute_overlap(a, b, P, cutoff, &undef, &ncalls, &subnormal);
    memset(out, 0, sizeof *out);
    out->status = undef ? 3 : 0;
    out->q = undef ? 0.0 : q;
    out->cutoff = cutoff;
    out->edge = !undef && q >= cutoff;
    out->n_single = ncalls;
    out->subnormal = !undef && subnormal;
}

void qorc_edge_pairs(const orc_read* reads, const int32_t* pi, const int32_t* pj, int64_t npairs, const qorc_params* P,
                     qorc_result* out) {
    for (int64_t k = 0; k < npairs; ++k) qorc_edge(&reads[pi[k]], &reads[pj[k]], P, &out[k]);
}

int32_t qorc_sizeof_result(void) { return (int32_t)sizeof(qorc_result); }
