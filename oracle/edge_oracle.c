/* oracle/edge_oracle.c -- CPU ORACLE, TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's read-pair edge decision
 * (NewEdgeCalculator::edgeBetween, /root/reference/src/NewEdgeCalculator.cpp:1247-1376) used as the
 * checker for the CUDA path. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load it; the product (haploclique_b200/) never does.
 *
 * Parity status: PINNED. tests/test_oracle_vs_reference.py compares every ordered pair of both bundled
 * BAMs (and the reference's unit-test fixtures) against oracle/_ref/libhcref.so, which is the unmodified
 * reference compiled from /root/reference; the resulting digests are committed under tests/golden/.
 *
 * Shape of the restatement (ours, not the reference's): the reference spells every read layout out as
 * hand-unrolled while-loops; here each layout is a row of a small PHASE PROGRAM table (tail / overlap /
 * seat / finish steps with symbolic bounds) run by one interpreter.  Each row cites the reference lines
 * it restates.  Cursor reads are bounds-checked: where the reference would read past a CIGAR or sequence
 * (undefined behaviour there) the pair is reported with HC_ST_UNDEFINED instead.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "orc_types.h"

typedef struct {
    double cutoff_cliques, cutoff_mixed, cutoff_single;   /* EDGE_QUASI_CUTOFF{,_MIXED,_SINGLE} */
    double overlap_cliques, overlap_single;               /* MIN_OVERLAP_{CLIQUES,SINGLE} */
    int32_t no_prob0;
    int32_t simpson_len;                                  /* SIMPSON_MAP.size() = maxPosition+1 (:85) */
    const double* simpson;                                /* dense, 0.0 = absent (:135-141) */
} orc_params;

enum { HC_ST_REJECT_COV = 0, HC_ST_REJECT_LAYOUT = 1, HC_ST_WALKED = 2, HC_ST_UNDEFINED = 3 };

typedef struct {
    int32_t status;          /* HC_ST_* */
    int32_t edge;
    int32_t gap;             /* checkGapsCigar returned true (:1367) */
    int32_t cc, tc;
    int32_t layout;          /* which phase program ran (diagnostic) */
    double probM, prob0;
    double score;            /* prob*potence (:1139), NaN if not reached */
} orc_result;

/* ---- probability tables (src/NewEdgeCalculator.cpp:41-74), same mixed float/double arithmetic ---- */
static double T_eq[127][127], T_uneq[127][127];
static int tables_ready = 0;

static void build_tables(void) {
    float e[127];
    int i, j;
    memset(T_eq, 0, sizeof T_eq);
    memset(T_uneq, 0, sizeof T_uneq);
    for (i = 33; i < 127; i++) {
        /* phredProb is called with the char (i-33): pow(10, -(q)/10.0) in double, narrowed to float (:41-48) */
        int q = i - 33;
        e[i] = (float)pow(10.0, (double)(-q) / 10.0);
    }
    for (i = 33; i < 127; i++) {
        for (j = 33; j < 127; j++) {
            float third_j = e[j] / 3;           /* float / int -> float */
            float third_i = e[i] / 3;
            float ff = e[i] * third_j;          /* float * float -> float (:56) */
            T_eq[i][j] = (1.0 - e[i]) * (1.0 - e[j]) + ff;
            /* (:67) (1.0-e_i)*(e_j/3) + (1.0-e_j)*(e_i/3) + 2*(e_i/3)*(e_j/3); last term: int*float*float in float */
            {
                float last = 2 * third_i * third_j;
                T_uneq[i][j] = (1.0 - e[i]) * third_j + (1.0 - e[j]) * third_i + last;
            }
        }
    }
    tables_ready = 1;
}

void orc_tables(double* eq_out, double* uneq_out) {
    if (!tables_ready) build_tables();
    memcpy(eq_out, T_eq, sizeof T_eq);
    memcpy(uneq_out, T_uneq, sizeof T_uneq);
}

/* ---- walk state ---- */
typedef struct { int c, q; } cursor;

typedef struct {
    const orc_read* rd[2];       /* rd[0] = ap1 (later arrival), rd[1] = ap2 */
    const orc_params* par;
    cursor cur[2][2];            /* [side][segment] */
    int ref;
    double probM, prob0;
    int cc, tc;
    int undefined;
} walk;

/* leading / trailing clip counts (computeOffset / computeRevOffset, src/AlignmentRecord.cpp:126-156) */
static int lead_clips(const orc_seg* s) {
    int k = 0;
    while (k < s->ncig && (s->cig[k] == 'S' || s->cig[k] == 'H')) k++;
    return k;
}
static int trail_clips(const orc_seg* s) {
    int k = 0;
    while (k < s->ncig && (s->cig[s->ncig - 1 - k] == 'S' || s->cig[s->ncig - 1 - k] == 'H')) k++;
    return k;
}

/* computeSOffset (src/AlignmentRecord.cpp:136-145): skip leading clips, S also consumes query */
static void seat(walk* w, int side, int seg) {
    const orc_seg* s = &w->rd[side]->s[seg];
    cursor* cu = &w->cur[side][seg];
    int k;
    for (k = 0; k < s->ncig; k++) {
        if (s->cig[k] == 'S') { cu->c++; cu->q++; }
        else if (s->cig[k] == 'H') { cu->c++; }
        else break;
    }
}

/* computeProb0 (src/NewEdgeCalculator.cpp:133-143) */
static void add_prob0(walk* w) {
    double d;
    if (w->ref < 0 || w->ref >= w->par->simpson_len) {
        /* reference reads SIMPSON_MAP out of bounds here; with no_prob0 the value is never used (:1132) */
        if (!w->par->no_prob0) w->undefined = 1;
        d = 0.0;
    } else {
        d = w->par->simpson[w->ref];
    }
    if (d != 0) w->prob0 += d; else w->prob0 += log10(0.25);
}

/* noOverlapCheckgap (:234-267): one step on a column only one read covers */
static void tail_step(walk* w, int side, int seg) {
    const orc_seg* s = &w->rd[side]->s[seg];
    cursor* cu = &w->cur[side][seg];
    char c;
    if (cu->c < 0 || cu->c >= s->ncig) { w->undefined = 1; return; }
    c = s->cig[cu->c];
    switch (c) {
    case 'H': w->ref++; cu->c++; break;
    case 'I': add_prob0(w); w->tc++; cu->q++; cu->c++; break;
    case 'D': w->ref++; cu->c++; break;
    case 'S': w->ref++; cu->q++; cu->c++; break;
    case 'M': add_prob0(w); w->tc++; w->ref++; cu->q++; cu->c++; break;
    default: w->undefined = 1; break;   /* reference: assert(0) / endless loop under NDEBUG (:262-266) */
    }
}

/* overlapCheckgap (:145-232): one lock-step on a column both reads cover. Returns 1 on gap conflict.
 * qa/qb are passed separately from the CIGAR cursors because one reference call site pairs a CIGAR
 * cursor with a different segment's query cursor (:629). */
static int ovl_step(walk* w, int sa, int sb, int* qa, int* qb) {
    const orc_seg* A = &w->rd[0]->s[sa];
    const orc_seg* B = &w->rd[1]->s[sb];
    int* ca = &w->cur[0][sa].c;
    int* cb = &w->cur[1][sb].c;
    char c1, c2;
    if (*ca < 0 || *ca >= A->ncig || *cb < 0 || *cb >= B->ncig) { w->undefined = 1; return 0; }
    c1 = A->cig[*ca];
    c2 = B->cig[*cb];
    if ((c1 == 'M' && c2 == 'M') || (c1 == 'S' && c2 == 'S') || (c1 == 'I' && c2 == 'I')) {
        if (c1 != 'S') {
            unsigned char q1, q2;
            if (*qa < 0 || *qa >= A->nq || *qb < 0 || *qb >= B->nq) { w->undefined = 1; return 0; }
            q1 = (unsigned char)A->qual[*qa];
            q2 = (unsigned char)B->qual[*qb];
            if (q1 < 33 || q1 > 126 || q2 < 33 || q2 > 126) { w->undefined = 1; return 0; } /* table rows <33 uninitialised (:54) */
            /* computeProbM (:122-131) */
            w->probM *= (A->seq[*qa] == B->seq[*qb]) ? T_eq[q1][q2] : T_uneq[q1][q2];
            w->cc++;
        }
        if (c1 != 'I') w->ref++;
        (*qa)++; (*qb)++; (*ca)++; (*cb)++;
        return 0;
    }
    {
        int d1 = c1 == 'D', d2 = c2 == 'D', h1 = c1 == 'H', h2 = c2 == 'H', s1 = c1 == 'S', s2 = c2 == 'S';
        int m1 = c1 == 'M', m2 = c2 == 'M', i1 = c1 == 'I', i2 = c2 == 'I';
        if ((d1 && d2) || (h1 && h2) || (d1 && h2) || (h1 && d2) || (d1 && s2) || (s1 && d2)) {   /* :183-194 */
            (*ca)++; (*cb)++; w->ref++;
            if (s1) (*qa)++; else if (s2) (*qb)++;
            return 0;
        }
        if ((m1 && (h2 || s2)) || ((h1 || s1) && m2) || (s1 && h2) || (h1 && s2)) {              /* :195-218 */
            w->ref++; (*ca)++; (*cb)++;
            if (m1) { (*qa)++; if (s2) (*qb)++; }
            else if (m2) { (*qb)++; if (s1) (*qa)++; }
            else if (s1) (*qa)++;
            else (*qb)++;
            return 0;
        }
        if ((i1 && (h2 || s2)) || ((h1 || s1) && i2)) {                                          /* :219-226 */
            if (i1) { (*ca)++; (*qa)++; } else { (*cb)++; (*qb)++; }
            return 0;
        }
    }
    return 1;                                                                                    /* :227-230 */
}

/* ---- symbolic bounds: id = ext*8 + side*4 + seg*2 + is_end ------------------------------------
 * raw  = AlignmentRecord::getStart/getEnd; ext = raw start - leading clips / raw end + trailing clips
 * (the ref_s_pos / ref_e_pos variables, src/NewEdgeCalculator.cpp:413-421). side 0 = ap1, 1 = ap2. */
#define BID(ext, side, seg, is_end) ((ext) * 8 + (side) * 4 + (seg) * 2 + (is_end))
enum {
    As1 = BID(0,0,0,0), Ae1 = BID(0,0,0,1), As2 = BID(0,0,1,0), Ae2 = BID(0,0,1,1),
    Bs1 = BID(0,1,0,0), Be1 = BID(0,1,0,1), Bs2 = BID(0,1,1,0), Be2 = BID(0,1,1,1),
    xAs1 = BID(1,0,0,0), xAe1 = BID(1,0,0,1), xAs2 = BID(1,0,1,0), xAe2 = BID(1,0,1,1),
    xBs1 = BID(1,1,0,0), xBe1 = BID(1,1,0,1), xBs2 = BID(1,1,1,0), xBe2 = BID(1,1,1,1)
};
#define B_SIDE(id) (((id) >> 2) & 1)
#define B_SEG(id) (((id) >> 1) & 1)

/* phase opcodes */
enum {
    P_END = 0,
    P_START,     /* ref = bound a                                   */
    P_TAIL_LT,   /* while (ref <  bound a) tail_step(side,seg)      */
    P_TAIL_LE,   /* while (ref <= bound a) tail_step(side,seg)      */
    P_OVL,       /* while (ref <= bound a) ovl_step(segA,segB)      */
    P_OVL2,      /* while (ref <= a && ref <= b) ovl_step(segA,segB)*/
    P_OVL2_QX,   /* as P_OVL2 but ap1's query cursor is the one of ap1's OTHER segment (:629) */
    P_SEAT,      /* computeSOffset on (side,seg)                    */
    P_FIN        /* if (ref-1 == a) tail owner(b) to b; else if (ref-1 == b) tail owner(a) to a */
};
typedef struct { uint8_t op, x, y, a, b; } phase;   /* x,y: (side,seg) for tails/seat, (segA,segB) for overlaps */

#define START(a)        {P_START, 0, 0, a, 0}
#define TAIL_LT(sd,sg,a) {P_TAIL_LT, sd, sg, a, 0}
#define TAIL_LE(sd,sg,a) {P_TAIL_LE, sd, sg, a, 0}
#define OVL(sa,sb,a)    {P_OVL, sa, sb, a, 0}
#define OVL2(sa,sb,a,b) {P_OVL2, sa, sb, a, b}
#define OVL2QX(sa,sb,a,b) {P_OVL2_QX, sa, sb, a, b}
#define SEAT(sd,sg)     {P_SEAT, sd, sg, 0, 0}
#define FIN(a,b)        {P_FIN, 0, 0, a, b}
#define END             {P_END, 0, 0, 0, 0}
#define A_ 0
#define B_ 1

#define MAXPH 12
/* Paired x paired programs, checkGapsPaired's 2nd..13th branches (src/NewEdgeCalculator.cpp:444-840). */
static const phase PP[12][MAXPH] = {
 /* 0  :444-479 */ { START(xAs1), TAIL_LT(A_,0,xBs1), OVL(0,0,Ae1), TAIL_LT(B_,0,As2), SEAT(A_,1), OVL(1,0,Be1),
                     TAIL_LT(A_,1,Bs2), SEAT(B_,1), OVL2(1,1,xBe2,xAe2), FIN(xBe2,xAe2), END },
 /* 1  :482-509 */ { START(xAs1), TAIL_LT(A_,0,xBs1), OVL(0,0,Ae1), TAIL_LT(B_,0,As2), SEAT(A_,1),
                     OVL2(1,0,xBe1,xAe2), FIN(xBe1,xAe2), END },
 /* 2  :513-540 */ { START(xBs1), TAIL_LT(B_,0,xAs2), OVL(1,0,Be1), TAIL_LT(A_,1,Bs2), SEAT(B_,1),
                     OVL2(1,1,xBe2,xAe2), FIN(xBe2,xAe2), END },
 /* 3  :543-570 */ { START(xAs2), TAIL_LT(A_,1,xBs1), OVL(1,0,Be1), TAIL_LT(A_,1,Bs2), SEAT(B_,1),
                     OVL2(1,1,xAe2,xBe2), FIN(xAe2,xBe2), END },
 /* 4  :574-610 */ { START(xAs1), TAIL_LT(A_,0,xBs1), OVL(0,0,Be1), TAIL_LT(A_,0,Bs2), SEAT(B_,1), OVL(0,1,Ae1),
                     TAIL_LT(B_,1,As2), SEAT(A_,1), OVL2(1,1,xBe2,xAe2), FIN(xBe2,xAe2), END },
 /* 5  :613-643 */ { START(xAs1), TAIL_LT(A_,0,xBs1), OVL(0,0,Be1), TAIL_LT(A_,0,Bs2), SEAT(B_,1),
                     OVL2QX(0,1,xAe1,xBe2), FIN(xAe1,xBe2), END },
 /* 6  :646-681 */ { START(xBs1), TAIL_LT(B_,0,xAs1), OVL(0,0,Ae1), TAIL_LT(B_,0,As2), SEAT(A_,1), OVL(1,0,Be1),
                     TAIL_LT(A_,1,Bs2), SEAT(B_,1), OVL2(1,1,xAe2,xBe2), FIN(xAe2,xBe2), END },
 /* 7  :684-710 */ { START(xBs1), TAIL_LT(B_,0,xAs1), OVL(0,0,Ae1), TAIL_LT(B_,0,As2), SEAT(A_,1),
                     OVL2(1,0,xBe1,xAe2), FIN(xBe1,xAe2), END },
 /* 8  :714-740 */ { START(xAs1), TAIL_LT(A_,0,xBs2), OVL(0,1,Ae1), TAIL_LT(B_,1,As2), SEAT(A_,1),
                     OVL2(1,1,xAe2,xBe2), FIN(xAe2,xBe2), END },
 /* 9  :744-770 */ { START(xBs2), TAIL_LT(B_,1,xAs1), OVL(0,1,Ae1), TAIL_LT(B_,1,As2), SEAT(A_,1),
                     OVL2(1,1,xAe2,xBe2), FIN(xAe2,xBe2), END },
 /* 10 :774-809 */ { START(xBs1), TAIL_LT(B_,0,xAs1), OVL(0,0,Be1), TAIL_LT(A_,0,Bs2), SEAT(B_,1), OVL(0,1,Ae1),
                     TAIL_LT(B_,1,As2), SEAT(A_,1), OVL2(1,1,xBe2,xAe2), FIN(xBe2,xAe2), END },
 /* 11 :813-839 */ { START(xBs1), TAIL_LT(B_,0,xAs1), OVL(0,0,Be1), TAIL_LT(A_,0,Bs2), SEAT(B_,1),
                     OVL2(0,1,xAe1,xBe2), FIN(xAe1,xBe2), END },
};
/* ap1 paired x ap2 single-end, checkGapsMixed first half, 3rd/4th branches (:893-949). */
static const phase PS[2][MAXPH] = {
 /* :893-919 */ { START(xBs1), TAIL_LT(B_,0,xAs1), OVL(0,0,Ae1), TAIL_LT(B_,0,As2), SEAT(A_,1),
                  OVL2(1,0,xAe2,xBe1), FIN(xAe2,xBe1), END },
 /* :922-948 */ { START(xAs1), TAIL_LT(A_,0,xBs1), OVL(0,0,Ae1), TAIL_LT(B_,0,As2), SEAT(A_,1),
                  OVL2(1,0,xAe2,xBe1), FIN(xAe2,xBe1), END },
};
/* ap1 single-end x ap2 paired, checkGapsMixed second half, 3rd/4th branches (:993-1052). */
static const phase SP[2][MAXPH] = {
 /* :993-1021  */ { START(xAs1), TAIL_LT(A_,0,xBs1), OVL(0,0,Be1), TAIL_LT(A_,0,Bs2), SEAT(B_,1),
                    OVL2(0,1,xBe2,xAe1), FIN(xBe2,xAe1), END },
 /* :1025-1051 */ { START(xBs1), TAIL_LT(B_,0,xAs1), OVL(0,0,Be1), TAIL_LT(A_,0,Bs2), SEAT(B_,1),
                    OVL2(0,1,xBe2,xAe1), FIN(xBe2,xAe1), END },
};

/* Runs one phase program. Returns 1 on gap conflict. */
static int run_program(walk* w, const phase* pg, const int* bnd) {
    int pc;
    for (pc = 0; pg[pc].op != P_END; pc++) {
        const phase* p = &pg[pc];
        switch (p->op) {
        case P_START: w->ref = bnd[p->a]; break;
        case P_TAIL_LT: while (w->ref < bnd[p->a] && !w->undefined) tail_step(w, p->x, p->y); break;
        case P_TAIL_LE: while (w->ref <= bnd[p->a] && !w->undefined) tail_step(w, p->x, p->y); break;
        case P_SEAT: seat(w, p->x, p->y); break;
        case P_OVL:
        case P_OVL2:
        case P_OVL2_QX: {
            int* qa = (p->op == P_OVL2_QX) ? &w->cur[0][1 - p->x].q : &w->cur[0][p->x].q;
            int* qb = &w->cur[1][p->y].q;
            while (w->ref <= bnd[p->a] && (p->op == P_OVL || w->ref <= bnd[p->b]) && !w->undefined) {
                if (ovl_step(w, p->x, p->y, qa, qb)) return 1;
            }
            break;
        }
        case P_FIN: {
            int first = p->a, second = p->b;
            if (w->ref - 1 == bnd[first]) {
                while (w->ref <= bnd[second] && !w->undefined) tail_step(w, B_SIDE(second), B_SEG(second));
            } else if (w->ref - 1 == bnd[second]) {
                while (w->ref <= bnd[first] && !w->undefined) tail_step(w, B_SIDE(first), B_SEG(first));
            }
            break;
        }
        default: break;
        }
        if (w->undefined) return 0;
    }
    return 0;
}

/* checkGapsSingle (:269-395): segment sa of ap1 against segment sb of ap2, fresh cursors. */
static int walk_single(walk* w, int sa, int sb, const int* bnd) {
    int xs1 = bnd[BID(1,0,sa,0)], xe1 = bnd[BID(1,0,sa,1)], xs2 = bnd[BID(1,1,sb,0)], xe2 = bnd[BID(1,1,sb,1)];
    phase pg[6];
    int n = 0;
    w->cur[0][sa].c = w->cur[0][sa].q = 0;
    w->cur[1][sb].c = w->cur[1][sb].q = 0;
#define PUSH(OP, X, Y, A) do { pg[n].op = OP; pg[n].x = X; pg[n].y = Y; pg[n].a = A; pg[n].b = 0; n++; } while (0)
    if (xs1 <= xs2 && xe1 <= xe2) {            /* :331-343 */
        PUSH(P_START, 0, 0, BID(1,0,sa,0)); PUSH(P_TAIL_LT, A_, sa, BID(1,1,sb,0));
        PUSH(P_OVL, sa, sb, BID(1,0,sa,1)); PUSH(P_TAIL_LE, B_, sb, BID(1,1,sb,1));
    } else if (xs1 <= xs2 && xe1 >= xe2) {     /* :346-358 */
        PUSH(P_START, 0, 0, BID(1,0,sa,0)); PUSH(P_TAIL_LT, A_, sa, BID(1,1,sb,0));
        PUSH(P_OVL, sa, sb, BID(1,1,sb,1)); PUSH(P_TAIL_LE, A_, sa, BID(1,0,sa,1));
    } else if (xs1 >= xs2 && xe1 <= xe2) {     /* :361-373 */
        PUSH(P_START, 0, 0, BID(1,1,sb,0)); PUSH(P_TAIL_LT, B_, sb, BID(1,0,sa,0));
        PUSH(P_OVL, sa, sb, BID(1,0,sa,1)); PUSH(P_TAIL_LE, B_, sb, BID(1,1,sb,1));
    } else {                                   /* :376-388 (the four conditions are exhaustive) */
        PUSH(P_START, 0, 0, BID(1,1,sb,0)); PUSH(P_TAIL_LT, B_, sb, BID(1,0,sa,0));
        PUSH(P_OVL, sa, sb, BID(1,1,sb,1)); PUSH(P_TAIL_LE, A_, sa, BID(1,0,sa,1));
    }
    PUSH(P_END, 0, 0, 0);
#undef PUSH
    return run_program(w, pg, bnd);
}

static void fill_bounds(const orc_read* a, const orc_read* b, int* bnd) {
    const orc_read* rd[2];
    int side, seg;
    rd[0] = a; rd[1] = b;
    for (side = 0; side < 2; side++) {
        for (seg = 0; seg < 2; seg++) {
            const orc_seg* s = &rd[side]->s[seg];
            int present = (seg == 0) || !rd[side]->single_end;
            bnd[BID(0, side, seg, 0)] = present ? s->start : 0;
            bnd[BID(0, side, seg, 1)] = present ? s->end : 0;
            bnd[BID(1, side, seg, 0)] = present ? s->start - lead_clips(s) : 0;
            bnd[BID(1, side, seg, 1)] = present ? s->end + trail_clips(s) : 0;
        }
    }
}

/* The interval rejects in front of the walk (src/NewEdgeCalculator.cpp:1268-1365). The reference compares
 * as unsigned int; starts/ends are positive so signed compares agree. Returns 1 = rejected. */
static int layout_reject(const orc_read* a, const orc_read* b, const int* v) {
    int se1 = a->single_end, se2 = b->single_end;
    if (se1 && se2) return v[Ae1] < v[Bs1] || v[Be1] < v[As1];
    if (!se1 && !se2) {
        return (v[Ae1] < v[Bs1] && v[Ae2] < v[Bs2]) || (v[Be1] < v[As1] && v[Be2] < v[As2]) ||
               (v[Be1] < v[As2] && v[Ae2] < v[Bs2]) || (v[Ae1] < v[Bs2] && v[Be2] < v[As2]) ||
               (v[Be1] < v[As1] && v[Ae1] < v[Bs2]) || (v[Ae1] < v[Bs1] && v[Be1] < v[As2]);
    }
    if (se1) return v[Ae1] < v[Bs1] || (v[Be1] < v[As1] && v[Ae1] < v[Bs2]) || v[Be2] < v[As1];
    return v[Ae2] < v[Bs1] || (v[Ae1] < v[Bs1] && v[Be1] < v[As2]) || v[Be1] < v[As1];
}

/* checkGapsCigar dispatch (:1063-1074) + layout selection of checkGapsPaired / checkGapsMixed. */
static int walk_pair(walk* w, const int* v, int* layout) {
    const orc_read* a = w->rd[0];
    const orc_read* b = w->rd[1];
    if (a->single_end && b->single_end) { *layout = 100; return walk_single(w, 0, 0, v); }
    if (!a->single_end && !b->single_end) {
        int k = -1;
        if (v[Ae1] < v[Bs2] && v[As2] > v[Be1]) {                     /* :436-440 */
            *layout = 200;
            if (walk_single(w, 0, 0, v)) return 1;
            if (w->undefined) return 0;
            return walk_single(w, 1, 1, v);
        }
        if      (v[xAs1] <= v[xBs1] && v[Ae1] >= v[Bs1] && v[As2] <= v[Be1] && v[Ae2] >= v[Bs2]) k = 0;
        else if (v[xAs1] <= v[xBs1] && v[Ae1] >= v[Bs1] && v[As2] <= v[Be1] && v[Ae2] <  v[Bs2]) k = 1;
        else if (v[xBs1] <= v[xAs2] && v[Ae1] <  v[Bs1] && v[As2] <= v[Be1] && v[Ae2] >= v[Bs2]) k = 2;
        else if (v[xAs2] <= v[xBs1] && v[Ae1] <  v[Bs1] && v[As2] <= v[Be1] && v[Ae2] >= v[Bs2]) k = 3;
        else if (v[xAs1] <= v[xBs1] && v[As1] <= v[Be1] && v[Ae1] >= v[Bs2] && v[As2] <= v[Be2]) k = 4;
        else if (v[xAs1] <= v[xBs1] && v[As2] >  v[Be2] && v[Ae1] >= v[Bs2] && v[As1] <= v[Be1]) k = 5;
        else if (v[xBs1] <= v[xAs1] && v[Ae1] >= v[Bs1] && v[As2] <= v[Be1] && v[Ae2] >= v[Bs2]) k = 6;
        else if (v[xBs1] <= v[xAs1] && v[Ae1] >= v[Bs1] && v[As2] <= v[Be1] && v[Ae2] <  v[Bs2]) k = 7;
        else if (v[xAs1] <= v[xBs2] && v[As1] >  v[Be1] && v[Ae1] >= v[Bs2] && v[As2] <= v[Be2]) k = 8;
        else if (v[xBs2] <= v[xAs1] && v[Be1] <  v[As1] && v[Ae1] >= v[Bs2] && v[As2] <= v[Be2]) k = 9;
        else if (v[xBs1] <= v[xAs1] && v[As1] <= v[Be1] && v[Ae1] >= v[Bs2] && v[As2] <= v[Be2]) k = 10;
        else if (v[xBs1] <= v[xAs1] && v[As2] >  v[Be2] && v[As1] <= v[Be1] && v[Ae1] >= v[Bs2]) k = 11;
        if (k < 0) { *layout = 299; return 0; }                       /* :841-844: assert compiled out, cc stays 0 */
        *layout = 201 + k;
        return run_program(w, PP[k], v);
    }
    if (b->single_end) {                                              /* ap1 paired, ap2 single (:850-954) */
        if (v[Be1] < v[As2]) { *layout = 300; return walk_single(w, 0, 0, v); }
        if (v[Bs1] > v[Ae1]) { *layout = 301; return walk_single(w, 1, 0, v); }
        if (v[xBs1] <= v[xAs1]) { *layout = 302; return run_program(w, PS[0], v); }
        *layout = 303;                                                /* xBs1 >= xAs1 is then always true */
        return run_program(w, PS[1], v);
    }
    /* ap1 single, ap2 paired (:955-1056) */
    if (v[Ae1] < v[Bs2]) { *layout = 400; return walk_single(w, 0, 0, v); }
    if (v[As1] > v[Be1]) { *layout = 401; return walk_single(w, 0, 1, v); }
    if (v[xAs1] <= v[xBs1]) { *layout = 402; return run_program(w, SP[0], v); }
    *layout = 403;
    return run_program(w, SP[1], v);
}

/* edgeBetween (:1247-1376) + similarityCriterion (:1107-1141). a = ap1 (the later arrival), b = ap2. */
void orc_edge(const orc_read* a, const orc_read* b, const orc_params* par, orc_result* out) {
    walk w;
    int v[16];
    int gap;
    if (!tables_ready) build_tables();
    memset(out, 0, sizeof *out);
    out->probM = 1.0;
    out->score = NAN;
    if (a->cov_last < b->cov_first || b->cov_last < a->cov_first) { out->status = HC_ST_REJECT_COV; return; }  /* :1252 */
    fill_bounds(a, b, v);
    if (layout_reject(a, b, v)) { out->status = HC_ST_REJECT_LAYOUT; return; }
    memset(&w, 0, sizeof w);
    w.rd[0] = a; w.rd[1] = b; w.par = par;
    w.probM = 1.0; w.prob0 = 0.0;
    gap = walk_pair(&w, v, &out->layout);
    out->gap = gap; out->cc = w.cc; out->tc = w.tc; out->probM = w.probM; out->prob0 = w.prob0;
    if (w.undefined) { out->status = HC_ST_UNDEFINED; return; }
    out->status = HC_ST_WALKED;
    if (gap) return;                                     /* :1367 */
    if (w.cc == 0) return;                               /* :1371 */
    {
        int both = a->is_clique && b->is_clique, any = a->is_clique || b->is_clique;
        double cutoff = log10(both ? par->cutoff_cliques : (any ? par->cutoff_mixed : par->cutoff_single));
        double min_overlap = both ? par->overlap_cliques : par->overlap_single;
        size_t smaller = (size_t)(a->ncov < b->ncov ? a->ncov : b->ncov);
        double prob, potence;
        if (w.cc <= min_overlap * smaller) return;       /* :1127 int vs double compare */
        if (par->no_prob0) { prob = log10(w.probM); potence = 1.0 / w.cc; }
        else { prob = log10(w.probM) + w.prob0; potence = 1.0 / (w.cc + w.tc); }
        out->score = prob * potence;
        out->edge = out->score >= cutoff;
    }
}

/* Batch helpers for the Python wrapper: reads come as a flat array of orc_read. */
void orc_edge_pairs(const orc_read* reads, const int32_t* pi, const int32_t* pj, int64_t npairs,
                    const orc_params* par, orc_result* out) {
    int64_t k;
    for (k = 0; k < npairs; k++) orc_edge(&reads[pi[k]], &reads[pj[k]], par, &out[k]);
}

/* All ordered pairs i > j (row-major over i), CLEVER's order (src/CLEVER.cpp:218-232). */
int64_t orc_edge_all(const orc_read* reads, int32_t n, const orc_params* par, orc_result* out, uint8_t* edge_only) {
    int64_t k = 0, edges = 0;
    int32_t i, j;
    orc_result tmp;
    for (i = 0; i < n; i++) {
        for (j = 0; j < i; j++, k++) {
            orc_result* r = out ? &out[k] : &tmp;
            orc_edge(&reads[i], &reads[j], par, r);
            edges += r->edge;
            if (edge_only) edge_only[k] = (uint8_t)r->edge;
        }
    }
    return edges;
}

int32_t orc_sizeof_read(void) { return (int32_t)sizeof(orc_read); }
int32_t orc_sizeof_result(void) { return (int32_t)sizeof(orc_result); }
