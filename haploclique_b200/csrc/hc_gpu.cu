// hc_gpu.cu -- sm_100a kernels and the CUDA half of the C ABI (include/hc_edge.h).
// This file: the phase programs, the one-lane-per-pair walker (pair lists; cross-check of the band kernels), work-item
// construction, and the host side of every entry point. The band kernels hc_score_band runs are in hc_tile2.cuh
// (address-addend formats, the default), hc_tile.cuh (16-bit records; tiled and streamed) -- the
// QuasispeciesEdgeCalculator kernels in hc_quasi.cuh.
//
// What one "pair" is: the decision NewEdgeCalculator::edgeBetween(ap1, ap2) makes
// (reference src/NewEdgeCalculator.cpp:1247-1376) for ap1 = the later arrival, ap2 = an earlier read.
// The reference walks two unrolled CIGARs in lock step; here every read is a run of 16-bit op records
// (op | base | quality code) so one load per side per column step feeds the classification, the base
// compare and the probability-table index. The fp64 match product is formed with one __dmul_rn per
// shared column in increasing reference order, exactly the reference's order, so probM is bit-identical
// (including its underflow to 0 on long weak overlaps).
//
// Layouts (which segment of which read is walked against which, and up to where) are rows of a small
// phase-program table in constant memory; one interpreter runs them. Row numbering and bound ids are
// defined below; each row cites the reference lines it stands for.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <mutex>
#include <thread>
#include <vector>

#include "hc_internal.h"

namespace {

// ---- bound ids: ext*8 + side*4 + seg*2 + is_end; side 0 = ap1 ("A"), 1 = ap2 ("B") -------------------
#define BID(ext, side, seg, is_end) ((ext) * 8 + (side) * 4 + (seg) * 2 + (is_end))
enum : uint8_t {
    As1 = BID(0, 0, 0, 0), Ae1 = BID(0, 0, 0, 1), As2 = BID(0, 0, 1, 0), Ae2 = BID(0, 0, 1, 1),
    Bs1 = BID(0, 1, 0, 0), Be1 = BID(0, 1, 0, 1), Bs2 = BID(0, 1, 1, 0), Be2 = BID(0, 1, 1, 1),
    xAs1 = BID(1, 0, 0, 0), xAe1 = BID(1, 0, 0, 1), xAs2 = BID(1, 0, 1, 0), xAe2 = BID(1, 0, 1, 1),
    xBs1 = BID(1, 1, 0, 0), xBe1 = BID(1, 1, 0, 1), xBs2 = BID(1, 1, 1, 0), xBe2 = BID(1, 1, 1, 1)
};

// ---- phase programs ---------------------------------------------------------------------------------
enum : uint8_t {
    PH_END = 0,
    PH_START,    // ref = v[a]
    PH_TAIL_LT,  // while (ref <  v[a]) tail step on (side x, seg y)
    PH_TAIL_LE,  // while (ref <= v[a]) tail step on (side x, seg y)
    PH_OVL,      // while (ref <= v[a]) lock step A.seg x / B.seg y
    PH_OVL2,     // while (ref <= v[a] && ref <= v[b]) lock step
    PH_OVL2_QX,  // same, but ap1's bases are indexed by the untouched query cursor of ap1's other segment (:629)
    PH_SEAT,     // cursor of (side x, seg y) skips the leading clips (computeSOffset)
    PH_FIN       // if (ref-1 == v[a]) tail owner(b) to v[b]; else if (ref-1 == v[b]) tail owner(a) to v[a]
};
struct Phase { uint8_t op, x, y, a, b; };
#define P_START(a) {PH_START, 0, 0, a, 0}
#define P_TLT(sd, sg, a) {PH_TAIL_LT, sd, sg, a, 0}
#define P_TLE(sd, sg, a) {PH_TAIL_LE, sd, sg, a, 0}
#define P_OVL(sa, sb, a) {PH_OVL, sa, sb, a, 0}
#define P_OVL2(sa, sb, a, b) {PH_OVL2, sa, sb, a, b}
#define P_OVLQX(sa, sb, a, b) {PH_OVL2_QX, sa, sb, a, b}
#define P_SEAT(sd, sg) {PH_SEAT, sd, sg, 0, 0}
#define P_FIN(a, b) {PH_FIN, 0, 0, a, b}
#define P_END {PH_END, 0, 0, 0, 0}
#define MAXPH 12
enum { PROG_PP0 = 0, PROG_PS0 = 12, PROG_SP0 = 14, N_PROGS = 16 };

// Linear phase programs. Order: checkGapsPaired branches 2..13 (:444 :482 :513 :543 :574 :613 :646 :684 :714 :744 :774 :813),
// then ap1 paired x ap2 single-end (checkGapsMixed :893, :922), then ap1 single-end x ap2 paired (:993, :1025).
#define HC_PROG_TABLE { \
    /* :444 */ {P_START(xAs1), P_TLT(0, 0, xBs1), P_OVL(0, 0, Ae1), P_TLT(1, 0, As2), P_SEAT(0, 1), P_OVL(1, 0, Be1), \
                P_TLT(0, 1, Bs2), P_SEAT(1, 1), P_OVL2(1, 1, xBe2, xAe2), P_FIN(xBe2, xAe2), P_END}, \
    /* :482 */ {P_START(xAs1), P_TLT(0, 0, xBs1), P_OVL(0, 0, Ae1), P_TLT(1, 0, As2), P_SEAT(0, 1), \
                P_OVL2(1, 0, xBe1, xAe2), P_FIN(xBe1, xAe2), P_END}, \
    /* :513 */ {P_START(xBs1), P_TLT(1, 0, xAs2), P_OVL(1, 0, Be1), P_TLT(0, 1, Bs2), P_SEAT(1, 1), \
                P_OVL2(1, 1, xBe2, xAe2), P_FIN(xBe2, xAe2), P_END}, \
    /* :543 */ {P_START(xAs2), P_TLT(0, 1, xBs1), P_OVL(1, 0, Be1), P_TLT(0, 1, Bs2), P_SEAT(1, 1), \
                P_OVL2(1, 1, xAe2, xBe2), P_FIN(xAe2, xBe2), P_END}, \
    /* :574 */ {P_START(xAs1), P_TLT(0, 0, xBs1), P_OVL(0, 0, Be1), P_TLT(0, 0, Bs2), P_SEAT(1, 1), P_OVL(0, 1, Ae1), \
                P_TLT(1, 1, As2), P_SEAT(0, 1), P_OVL2(1, 1, xBe2, xAe2), P_FIN(xBe2, xAe2), P_END}, \
    /* :613 */ {P_START(xAs1), P_TLT(0, 0, xBs1), P_OVL(0, 0, Be1), P_TLT(0, 0, Bs2), P_SEAT(1, 1), \
                P_OVLQX(0, 1, xAe1, xBe2), P_FIN(xAe1, xBe2), P_END}, \
    /* :646 */ {P_START(xBs1), P_TLT(1, 0, xAs1), P_OVL(0, 0, Ae1), P_TLT(1, 0, As2), P_SEAT(0, 1), P_OVL(1, 0, Be1), \
                P_TLT(0, 1, Bs2), P_SEAT(1, 1), P_OVL2(1, 1, xAe2, xBe2), P_FIN(xAe2, xBe2), P_END}, \
    /* :684 */ {P_START(xBs1), P_TLT(1, 0, xAs1), P_OVL(0, 0, Ae1), P_TLT(1, 0, As2), P_SEAT(0, 1), \
                P_OVL2(1, 0, xBe1, xAe2), P_FIN(xBe1, xAe2), P_END}, \
    /* :714 */ {P_START(xAs1), P_TLT(0, 0, xBs2), P_OVL(0, 1, Ae1), P_TLT(1, 1, As2), P_SEAT(0, 1), \
                P_OVL2(1, 1, xAe2, xBe2), P_FIN(xAe2, xBe2), P_END}, \
    /* :744 */ {P_START(xBs2), P_TLT(1, 1, xAs1), P_OVL(0, 1, Ae1), P_TLT(1, 1, As2), P_SEAT(0, 1), \
                P_OVL2(1, 1, xAe2, xBe2), P_FIN(xAe2, xBe2), P_END}, \
    /* :774 */ {P_START(xBs1), P_TLT(1, 0, xAs1), P_OVL(0, 0, Be1), P_TLT(0, 0, Bs2), P_SEAT(1, 1), P_OVL(0, 1, Ae1), \
                P_TLT(1, 1, As2), P_SEAT(0, 1), P_OVL2(1, 1, xBe2, xAe2), P_FIN(xBe2, xAe2), P_END}, \
    /* :813 */ {P_START(xBs1), P_TLT(1, 0, xAs1), P_OVL(0, 0, Be1), P_TLT(0, 0, Bs2), P_SEAT(1, 1), \
                P_OVL2(0, 1, xAe1, xBe2), P_FIN(xAe1, xBe2), P_END}, \
    {P_START(xBs1), P_TLT(1, 0, xAs1), P_OVL(0, 0, Ae1), P_TLT(1, 0, As2), P_SEAT(0, 1), P_OVL2(1, 0, xAe2, xBe1), \
     P_FIN(xAe2, xBe1), P_END}, \
    {P_START(xAs1), P_TLT(0, 0, xBs1), P_OVL(0, 0, Ae1), P_TLT(1, 0, As2), P_SEAT(0, 1), P_OVL2(1, 0, xAe2, xBe1), \
     P_FIN(xAe2, xBe1), P_END}, \
    {P_START(xAs1), P_TLT(0, 0, xBs1), P_OVL(0, 0, Be1), P_TLT(0, 0, Bs2), P_SEAT(1, 1), P_OVL2(0, 1, xBe2, xAe1), \
     P_FIN(xBe2, xAe1), P_END}, \
    {P_START(xBs1), P_TLT(1, 0, xAs1), P_OVL(0, 0, Be1), P_TLT(0, 0, Bs2), P_SEAT(1, 1), P_OVL2(0, 1, xBe2, xAe1), \
     P_FIN(xBe2, xAe1), P_END}, \
}
__constant__ Phase c_prog[N_PROGS][MAXPH] = HC_PROG_TABLE;
static const Phase h_prog[N_PROGS][MAXPH] = HC_PROG_TABLE;


// ---- scoring parameters on the device ----------------------------------------------------------------
struct DevParams {
    double log_cutoff[3];      // log10 of {cliques, mixed, single} cutoffs, host libm (:1112-1116)
    double min_overlap[2];     // {cliques, single}
    double log10_quarter;      // std::log10(0.25) from host libm (:140)
    double amb_tol;
    const double* table;       // [2][K][K]: equal-base block, then unequal-base block
    const double* simpson;     // dense map or nullptr
    const double* simpson_q;   // the addend of a tail column at each position: the map's entry, log10(0.25) where it has none
    const double* ladder;      // F[n] = n sequential additions of log10(0.25)
    int K, KK;
    int simpson_len;
    int ladder_len;
    int no_prob0;
};

struct PairOut {
    int status, edge, gap, cc, tc, ambiguous;
    double probM, prob0, score, cutoff;
};

// classification of an op pair in the lock step (overlapCheckgap, :173-231)
enum { ACT_SCORE = 0, ACT_SKIP = 1, ACT_ADV_A = 2, ACT_ADV_B = 3, ACT_GAP = 4 };
__device__ __forceinline__ int op_action(int oa, int ob) {
    // rows oa = M I D S H, cols ob = M I D S H, 3 bits each
    //  M: S G G K K | I: G S G A A | D: G G K K K | S: K B K K K | H: K B K K K   (S=score K=skip A/B=advance one G=gap)
    const unsigned long long lo =
        (0ull << 0) | (4ull << 3) | (4ull << 6) | (1ull << 9) | (1ull << 12) |          // M row
        (4ull << 15) | (0ull << 18) | (4ull << 21) | (2ull << 24) | (2ull << 27) |      // I row
        (4ull << 30) | (4ull << 33) | (1ull << 36) | (1ull << 39) | (1ull << 42) |      // D row
        (1ull << 45) | (3ull << 48) | (1ull << 51) | (1ull << 54) | (1ull << 57) |      // S row
        (1ull << 60);                                                                   // H,M
    const unsigned hi = (3u << 0) | (1u << 3) | (1u << 6) | (1u << 9);                  // H,I H,D H,S H,H
    const int idx = oa * 5 + ob;
    return idx < 21 ? (int)((lo >> (3 * idx)) & 7ull) : (int)((hi >> (3 * (idx - 21))) & 7u);
}

struct Seg {
    const uint16_t* ops;
    int n;
};

struct WalkState {
    int ref;
    int cc, tc;
    double probM, prob0;
    bool undefined;
};

__device__ __forceinline__ bool consumes_query(int op) { return op == HC_OP_M || op == HC_OP_I || op == HC_OP_S; }

// noOverlapCheckgap loop (:234-267): columns only one read covers. limit is exclusive.
__device__ __forceinline__ void tail_loop(const DevParams& P, WalkState& w, const Seg s, int& c, int limit) {
    while (w.ref < limit) {
        if (c >= s.n) { w.undefined = true; return; }
        const int op = (int)HC_REC_OP((unsigned)s.ops[c]);
        if (op == HC_OP_M || op == HC_OP_I) {
            w.tc++;
            if (P.simpson != nullptr) {                                        // computeProb0 (:133-143)
                double d = 0.0;
                if (w.ref < 0 || w.ref >= P.simpson_len) {
                    // the reference reads SIMPSON_MAP out of bounds here; harmless only when prob0 is ignored (:1132)
                    if (!P.no_prob0) { w.undefined = true; return; }
                } else {
                    d = __ldg(P.simpson + w.ref);
                }
                w.prob0 = __dadd_rn(w.prob0, d != 0.0 ? d : P.log10_quarter);
            }
        }
        if (op != HC_OP_I) w.ref++;
        c++;
    }
}

// overlapCheckgap loop (:145-232): columns both reads cover. hi is inclusive. Returns true on gap conflict.
template <bool QX>
__device__ __forceinline__ bool ovl_loop(const DevParams& P, WalkState& w, const Seg A, int& ca, const Seg B, int& cb,
                                         int hi) {
    int cq = 0;   // QX only: op index in A of the base the reference's stray query cursor points at
    if (QX) {
        while (cq < A.n && !consumes_query((int)HC_REC_OP((unsigned)A.ops[cq]))) cq++;
    }
    while (w.ref <= hi) {
        if (ca >= A.n || cb >= B.n) { w.undefined = true; return false; }
        const unsigned ra = A.ops[ca], rb = B.ops[cb];
        const int oa = (int)HC_REC_OP(ra), ob = (int)HC_REC_OP(rb);
        const int act = (oa | ob) == 0 ? ACT_SCORE : op_action(oa, ob);
        if (act == ACT_SCORE) {
            unsigned sa = ra;
            if (QX) {
                if (cq >= A.n) { w.undefined = true; return false; }
                sa = A.ops[cq];
            }
            const int idx = (int)HC_REC_QUAL(sa) * P.K + (int)HC_REC_QUAL(rb) + (HC_REC_BASE(sa ^ rb) ? P.KK : 0);
            w.probM = __dmul_rn(w.probM, __ldg(P.table + idx));                // computeProbM (:122-131)
            w.cc++;
            if (oa == HC_OP_M) w.ref++;
            ca++; cb++;
            if (QX) { cq++; while (cq < A.n && !consumes_query((int)HC_REC_OP((unsigned)A.ops[cq]))) cq++; }
        } else if (act == ACT_SKIP) {
            w.ref++; ca++; cb++;
            if (QX && consumes_query(oa)) { cq++; while (cq < A.n && !consumes_query((int)HC_REC_OP((unsigned)A.ops[cq]))) cq++; }
        } else if (act == ACT_ADV_A) {
            ca++;
            if (QX) { cq++; while (cq < A.n && !consumes_query((int)HC_REC_OP((unsigned)A.ops[cq]))) cq++; }
        } else if (act == ACT_ADV_B) {
            cb++;
        } else {
            return true;
        }
    }
    return false;
}

struct PairCtx {
    Seg seg[2][2];     // [side][segment]
    int cur[2][2];     // CIGAR cursors
    int lead[2][2];    // leading clip count (raw start - ext start)
    int v[16];         // bounds by id
};

// Runs one phase program. Returns true on gap conflict.
__device__ bool run_program(const DevParams& P, WalkState& w, PairCtx& pc, const Phase* __restrict__ pg) {
    for (int k = 0; k < MAXPH; ++k) {
        const Phase ph = pg[k];
        switch (ph.op) {
            case PH_END: return false;
            case PH_START: w.ref = pc.v[ph.a]; break;
            case PH_TAIL_LT: tail_loop(P, w, pc.seg[ph.x][ph.y], pc.cur[ph.x][ph.y], pc.v[ph.a]); break;
            case PH_TAIL_LE: tail_loop(P, w, pc.seg[ph.x][ph.y], pc.cur[ph.x][ph.y], pc.v[ph.a] + 1); break;
            case PH_SEAT: pc.cur[ph.x][ph.y] += pc.lead[ph.x][ph.y]; break;
            case PH_OVL:
                if (ovl_loop<false>(P, w, pc.seg[0][ph.x], pc.cur[0][ph.x], pc.seg[1][ph.y], pc.cur[1][ph.y], pc.v[ph.a]))
                    return true;
                break;
            case PH_OVL2:
                if (ovl_loop<false>(P, w, pc.seg[0][ph.x], pc.cur[0][ph.x], pc.seg[1][ph.y], pc.cur[1][ph.y],
                                    min(pc.v[ph.a], pc.v[ph.b])))
                    return true;
                break;
            case PH_OVL2_QX:
                if (ovl_loop<true>(P, w, pc.seg[0][ph.x], pc.cur[0][ph.x], pc.seg[1][ph.y], pc.cur[1][ph.y],
                                   min(pc.v[ph.a], pc.v[ph.b])))
                    return true;
                break;
            case PH_FIN: {
                int which = -1;                                   // bound id whose owner still has columns left
                if (w.ref - 1 == pc.v[ph.a]) which = ph.b;
                else if (w.ref - 1 == pc.v[ph.b]) which = ph.a;
                if (which >= 0) {
                    const int sd = (which >> 2) & 1, sg = (which >> 1) & 1;
                    tail_loop(P, w, pc.seg[sd][sg], pc.cur[sd][sg], pc.v[which] + 1);
                }
                break;
            }
        }
        if (w.undefined) return false;
    }
    return false;
}

// checkGapsSingle (:269-395): segment sa of ap1 against segment sb of ap2 with fresh cursors.
__device__ bool walk_single(const DevParams& P, WalkState& w, PairCtx& pc, int sa, int sb) {
    const int xs1 = pc.v[BID(1, 0, sa, 0)], xe1 = pc.v[BID(1, 0, sa, 1)];
    const int xs2 = pc.v[BID(1, 1, sb, 0)], xe2 = pc.v[BID(1, 1, sb, 1)];
    int ca = 0, cb = 0;
    const Seg A = pc.seg[0][sa], B = pc.seg[1][sb];
    // the reference's four shapes in order (:331, :346, :361, :376); together they are exhaustive
    int shape;
    if (xs1 <= xs2 && xe1 <= xe2) shape = 0;
    else if (xs1 <= xs2 && xe1 >= xe2) shape = 1;
    else if (xs1 >= xs2 && xe1 <= xe2) shape = 2;
    else shape = 3;
    if (shape < 2) { w.ref = xs1; tail_loop(P, w, A, ca, xs2); }
    else { w.ref = xs2; tail_loop(P, w, B, cb, xs1); }
    if (w.undefined) return false;
    const int ovl_hi = (shape == 0 || shape == 2) ? xe1 : xe2;
    if (ovl_loop<false>(P, w, A, ca, B, cb, ovl_hi)) return true;
    if (w.undefined) return false;
    if (shape == 0 || shape == 2) tail_loop(P, w, B, cb, xe2 + 1);
    else tail_loop(P, w, A, ca, xe1 + 1);
    return false;
}

// The interval rejects in front of the walk (:1268-1365). true = rejected.
__device__ __forceinline__ bool layout_reject(bool se1, bool se2, const int* v) {
    if (se1 && se2) return v[Ae1] < v[Bs1] || v[Be1] < v[As1];
    if (!se1 && !se2)
        return (v[Ae1] < v[Bs1] && v[Ae2] < v[Bs2]) || (v[Be1] < v[As1] && v[Be2] < v[As2]) ||
               (v[Be1] < v[As2] && v[Ae2] < v[Bs2]) || (v[Ae1] < v[Bs2] && v[Be2] < v[As2]) ||
               (v[Be1] < v[As1] && v[Ae1] < v[Bs2]) || (v[Ae1] < v[Bs1] && v[Be1] < v[As2]);
    if (se1) return v[Ae1] < v[Bs1] || (v[Be1] < v[As1] && v[Ae1] < v[Bs2]) || v[Be2] < v[As1];
    return v[Ae2] < v[Bs1] || (v[Ae1] < v[Bs1] && v[Be1] < v[As2]) || v[Be1] < v[As1];
}

// checkGapsCigar dispatch (:1063-1074) and the layout choice inside checkGapsPaired/Mixed.
__device__ bool walk_pair(const DevParams& P, WalkState& w, PairCtx& pc, bool se1, bool se2) {
    const int* v = pc.v;
    if (se1 && se2) return walk_single(P, w, pc, 0, 0);
    if (!se1 && !se2) {
        if (v[Ae1] < v[Bs2] && v[As2] > v[Be1]) {                                      // :436
            if (walk_single(P, w, pc, 0, 0)) return true;
            if (w.undefined) return false;
            return walk_single(P, w, pc, 1, 1);
        }
        int k = -1;
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
        if (k < 0) return false;                      // no layout matches: cc stays 0 (:841-844 under NDEBUG)
        return run_program(P, w, pc, c_prog[PROG_PP0 + k]);
    }
    if (se2) {                                        // ap1 paired, ap2 single-end (:850-954)
        if (v[Be1] < v[As2]) return walk_single(P, w, pc, 0, 0);
        if (v[Bs1] > v[Ae1]) return walk_single(P, w, pc, 1, 0);
        return run_program(P, w, pc, c_prog[PROG_PS0 + (v[xBs1] <= v[xAs1] ? 0 : 1)]);
    }
    if (v[Ae1] < v[Bs2]) return walk_single(P, w, pc, 0, 0);   // ap1 single-end, ap2 paired (:955-1056)
    if (v[As1] > v[Be1]) return walk_single(P, w, pc, 0, 1);
    return run_program(P, w, pc, c_prog[PROG_SP0 + (v[xAs1] <= v[xBs1] ? 0 : 1)]);
}

// edgeBetween (:1247-1376) + similarityCriterion (:1107-1141) for ap1 = hdr[i], ap2 = hdr[j].
__device__ void score_pair(const DevParams& P, const hc_read_hdr* __restrict__ hdr, const uint16_t* __restrict__ ops,
                           int i, int j, PairOut& o) {
    o.status = HC_ST_REJECT_COV; o.edge = 0; o.gap = 0; o.cc = 0; o.tc = 0; o.ambiguous = 0;
    o.probM = 1.0; o.prob0 = 0.0; o.score = __longlong_as_double(0x7ff8000000000000ll); o.cutoff = 0.0;
    // headers: 64 B each, four 16-B loads
    const int4* ha = reinterpret_cast<const int4*>(hdr + i);
    const int4* hb = reinterpret_cast<const int4*>(hdr + j);
    const int4 a3 = __ldg(ha + 3), b3 = __ldg(hb + 3);     // ncov_flags, cov_first, cov_last, has_ins
    if (a3.z < b3.y || b3.z < a3.y) return;                 // covered intervals disjoint (:1252)
    PairCtx pc;
    const int4 a0 = __ldg(ha), a1 = __ldg(ha + 1), a2 = __ldg(ha + 2);
    const int4 b0 = __ldg(hb), b1 = __ldg(hb + 1), b2 = __ldg(hb + 2);
    pc.v[0] = a0.x; pc.v[1] = a0.y; pc.v[2] = a0.z; pc.v[3] = a0.w;
    pc.v[4] = b0.x; pc.v[5] = b0.y; pc.v[6] = b0.z; pc.v[7] = b0.w;
    pc.v[8] = a1.x; pc.v[9] = a1.y; pc.v[10] = a1.z; pc.v[11] = a1.w;
    pc.v[12] = b1.x; pc.v[13] = b1.y; pc.v[14] = b1.z; pc.v[15] = b1.w;
    const unsigned fa = (unsigned)a3.x, fb = (unsigned)b3.x;
    const bool se1 = fa & HC_FLAG_SINGLE_END, se2 = fb & HC_FLAG_SINGLE_END;
    if (layout_reject(se1, se2, pc.v)) { o.status = HC_ST_REJECT_LAYOUT; return; }
    pc.seg[0][0] = Seg{ops + (unsigned)a2.x, a2.z};
    pc.seg[0][1] = Seg{ops + (unsigned)a2.y, a2.w};
    pc.seg[1][0] = Seg{ops + (unsigned)b2.x, b2.z};
    pc.seg[1][1] = Seg{ops + (unsigned)b2.y, b2.w};
    pc.lead[0][0] = a0.x - a1.x; pc.lead[0][1] = a0.z - a1.z;
    pc.lead[1][0] = b0.x - b1.x; pc.lead[1][1] = b0.z - b1.z;
    pc.cur[0][0] = pc.cur[0][1] = pc.cur[1][0] = pc.cur[1][1] = 0;
    WalkState w;
    w.ref = 0; w.cc = 0; w.tc = 0; w.probM = 1.0; w.prob0 = 0.0; w.undefined = false;
    const bool gap = walk_pair(P, w, pc, se1, se2);
    if (P.simpson == nullptr && !w.undefined) {
        if (w.tc >= P.ladder_len) w.undefined = true;
        else w.prob0 = __ldg(P.ladder + w.tc);        // tc sequential additions of log10(0.25)
    }
    o.gap = gap; o.cc = w.cc; o.tc = w.tc; o.probM = w.probM; o.prob0 = w.prob0;
    if (w.undefined) { o.status = HC_ST_UNDEFINED; return; }
    o.status = HC_ST_WALKED;
    if (gap || w.cc == 0) return;                     // :1367, :1371
    const bool cl1 = fa & HC_FLAG_IS_CLIQUE, cl2 = fb & HC_FLAG_IS_CLIQUE;
    const int sel = (cl1 && cl2) ? 0 : ((cl1 || cl2) ? 1 : 2);
    const double cutoff = P.log_cutoff[sel];
    const double min_overlap = P.min_overlap[(cl1 && cl2) ? 0 : 1];
    const unsigned smaller = min(fa & 0x3fffffffu, fb & 0x3fffffffu);
    o.cutoff = cutoff;
    if ((double)w.cc <= __dmul_rn(min_overlap, (double)smaller)) return;      // :1127
    double prob, potence;
    if (P.no_prob0) { prob = log10(w.probM); potence = __ddiv_rn(1.0, (double)w.cc); }
    else { prob = __dadd_rn(log10(w.probM), w.prob0); potence = __ddiv_rn(1.0, (double)(w.cc + w.tc)); }
    const double score = __dmul_rn(prob, potence);    // :1139
    o.score = score;
    o.edge = score >= cutoff;
    o.ambiguous = fabs(score - cutoff) <= P.amb_tol * fabs(cutoff);
}

// ---- kernels ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) score_pairs_kernel(DevParams P, const hc_read_hdr* __restrict__ hdr,
                                                         const uint16_t* __restrict__ ops, const int32_t* __restrict__ pi,
                                                         const int32_t* __restrict__ pj, long long npairs,
                                                         hc_pair_result* __restrict__ out) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < npairs; k += (long long)gridDim.x * blockDim.x) {
        PairOut o;
        score_pair(P, hdr, ops, pi[k], pj[k], o);
        hc_pair_result r;
        r.status = o.status; r.edge = o.edge; r.gap = o.gap; r.cc = o.cc; r.tc = o.tc; r.ambiguous = o.ambiguous;
        r.probM = o.probM; r.prob0 = o.prob0; r.score = o.score;
        out[k] = r;
    }
}

struct BandStatsDev {
    unsigned long long pairs_total, pairs_scored, pairs_walked, edges, sum_cc, sum_tc, undefined, ambiguous;
};

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long x) {
    for (int d = 16; d > 0; d >>= 1) x += __shfl_down_sync(0xffffffffu, x, d);
    return x;
}

// One warp per 32-bit adjacency word: the 32 lanes hold 32 consecutive earlier reads j against the same
// later read i, so ap1's header and op records are a broadcast and the edge bits come out of one ballot.
__global__ void __launch_bounds__(256) score_band_kernel(DevParams P, const hc_read_hdr* __restrict__ hdr,
                                                        const uint16_t* __restrict__ ops, int row_begin, int nrows,
                                                        const int32_t* __restrict__ col0,
                                                        const long long* __restrict__ word_off, long long total_words,
                                                        uint32_t* __restrict__ adj, BandStatsDev* __restrict__ stats,
                                                        hc_ambiguous* __restrict__ amb, int amb_cap) {
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    unsigned long long n_total = 0, n_scored = 0, n_walked = 0, n_edges = 0, s_cc = 0, s_tc = 0, n_undef = 0;
    for (long long wd = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; wd < total_words; wd += warps) {
        int lo = 0, hi = nrows;                       // last row with word_off[row] <= wd
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (word_off[mid] <= wd) lo = mid; else hi = mid;
        }
        const int i = row_begin + lo;
        const int j = col0[lo] + (int)(wd - word_off[lo]) * 32 + lane;
        bool edge = false;
        if (j < i) {
            PairOut o;
            score_pair(P, hdr, ops, i, j, o);
            edge = o.edge;
            n_total++;
            n_scored += o.status != HC_ST_REJECT_COV;
            if (o.status == HC_ST_WALKED) { n_walked++; s_cc += o.cc; s_tc += o.tc; }
            n_undef += o.status == HC_ST_UNDEFINED;
            if (o.ambiguous) {
                const unsigned long long slot = atomicAdd(&stats->ambiguous, 1ull);
                if (amb != nullptr && slot < (unsigned long long)amb_cap) {
                    hc_ambiguous a;
                    a.i = i; a.j = j; a.cc = o.cc; a.tc = o.tc; a.probM = o.probM; a.prob0 = o.prob0;
                    a.score = o.score; a.cutoff = o.cutoff; a.edge = o.edge; a.edge_device = o.edge;
                    amb[slot] = a;
                }
            }
        }
        const unsigned bits = __ballot_sync(0xffffffffu, edge);
        if (lane == 0) adj[wd] = bits;
        n_edges += edge;
    }
    n_total = warp_sum(n_total); n_scored = warp_sum(n_scored); n_walked = warp_sum(n_walked);
    n_edges = warp_sum(n_edges); s_cc = warp_sum(s_cc); s_tc = warp_sum(s_tc); n_undef = warp_sum(n_undef);
    if (lane == 0) {
        atomicAdd(&stats->pairs_total, n_total); atomicAdd(&stats->pairs_scored, n_scored);
        atomicAdd(&stats->pairs_walked, n_walked); atomicAdd(&stats->edges, n_edges);
        atomicAdd(&stats->sum_cc, s_cc); atomicAdd(&stats->sum_tc, s_tc); atomicAdd(&stats->undefined, n_undef);
    }
}

#include "hc_tile.cuh"
#include "hc_tile2.cuh"
#include "hc_quasi.cuh"

}  // namespace

// ======================================= host side of the C ABI =======================================
struct hc_ctx {
    hc_params par{};
    std::vector<double> simpson_host;
    std::vector<double> eq, uneq;          // 127x127 each
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // resident batch
    int n_reads = 0, K = 0, max_ops = 0;
    hc_read_hdr* d_hdr = nullptr;
    uint16_t* d_ops = nullptr;
    double* d_table = nullptr;
    double* d_simpson = nullptr; double* d_simpson_q = nullptr;
    double* d_ladder = nullptr;
    int ladder_len = 0;
    std::vector<int32_t> cov_first, cov_last_pmax;   // host copies for the band layout
    // band scratch (grown on demand)
    int32_t* d_col0 = nullptr; long long* d_word_off = nullptr; size_t col0_cap = 0, word_off_cap = 0;
    uint32_t* d_adj = nullptr; size_t adj_cap = 0;
    BandStatsDev* d_stats = nullptr;
    hc_ambiguous* d_amb = nullptr; size_t amb_cap = 0;
    // pair-list scratch
    // every buffer has its own capacity: hc_score_pairs and hc_quasi_score_pairs share d_pi / d_pj but not the results
    int32_t* d_pi = nullptr; int32_t* d_pj = nullptr; hc_pair_result* d_res = nullptr;
    size_t pi_cap = 0, pj_cap = 0, res_cap = 0;
    std::vector<int32_t> lay_col0; std::vector<long long> lay_off; int lay_begin = -1, lay_end = -1;
    void* d_flush = nullptr; size_t flush_bytes = 0;
    // tiled band kernel (hc_tile.cuh)
    uint32_t* d_aux = nullptr; uint16_t* d_col2op = nullptr; uint32_t* d_tprog = nullptr;
    // QuasispeciesEdgeCalculator path (hc_quasi.cuh): host-tabulated log terms per (base1, base2, code1, code2), ln of the map
    uint8_t qual_chars[128] = {0};
    double* d_logsum = nullptr; double* d_simpson_ln = nullptr; hc_quasi_result* d_qres = nullptr; size_t qres_cap = 0;
    int4* d_qseg = nullptr; size_t qseg_cap = 0; bool have_qseg = false;
    // second-generation tiled kernel (hc_tile2.cuh): row / column operand formats and the 9-row table
    uint32_t* d_opsA = nullptr; uint16_t* d_opsF = nullptr; uint32_t* d_tab9 = nullptr;
    bool have_formats = false;
    // capacities (elements) of the per-batch device buffers: they only grow, so a steady stream of hc_upload calls does
    // no cudaMalloc / cudaFree (both synchronise the device and would serialise the copies with the host-side layout work)
    size_t cap_hdr = 0, cap_ops = 0, cap_table = 0, cap_ladder = 0, cap_aux = 0, cap_col2op = 0, cap_opsA = 0, cap_opsF = 0, cap_tab9 = 0;
    int4* d_items = nullptr; size_t items_cap = 0; int* d_item_counter = nullptr;
    hc_pair_result* d_detail = nullptr; size_t detail_cap = 0;
    int kernel_mode = HC_KERNEL_AUTO;
    int last_kernel = 0;
    size_t smem_optin = 0;
    // scoring plan of the last range: row windows with their work items (build_plan)
    struct Win { int r0, r1; size_t item0, n_items; long long word0, n_words; };
    std::vector<int4> items; std::vector<int32_t> colidx; std::vector<Win> wins;
    int plan_begin = -1, plan_end = -1, plan_ctas = 0; bool plan_windows = false, plan_by_rows = false, plan_warp_items = false, plan_sig = false;
    cudaStream_t copy_stream = nullptr;                     // D2H of finished row windows, beside the compute streams
    cudaStream_t stream2 = nullptr;                         // odd row windows: their CTAs fill the SMs the previous window's tail frees
    cudaEvent_t ev_go = nullptr, ev_join = nullptr;
    cudaEvent_t win_ev[16] = {nullptr};                     // kernel of window k done
    std::vector<uint32_t> ncov_flags;                       // host copy of the headers' flags word (host re-decision)
    std::vector<hc_ambiguous> amb_host;                     // near-threshold pairs of the last pass, re-decided
    int32_t* d_colidx = nullptr; size_t colidx_cap = 0;
    std::vector<int32_t> cov_last, cov_first_smin; std::vector<uint8_t> is_single;
    std::vector<uint32_t> indel_sig;                        // per read: hash of its (position, I / D) events, 0 = none (build_items_range)
    std::string err;
};

namespace {

#define CU_TRY(ctx, call)                                                                      \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                  \
            return HC_E_CUDA;                                                                  \
        }                                                                                      \
    } while (0)

template <class T>
int grow(hc_ctx* c, T*& ptr, size_t& cap, size_t need) {
    if (need <= cap && ptr) return HC_OK;
    if (ptr) CU_TRY(c, cudaFree(ptr));
    ptr = nullptr;
    cap = 0;
    const size_t n = std::max<size_t>(need + need / 8, 1024);
    CU_TRY(c, cudaMalloc(reinterpret_cast<void**>(&ptr), n * sizeof(T)));
    cap = n;
    return HC_OK;
}

DevParams dev_params(const hc_ctx* c) {
    DevParams P{};
    P.log_cutoff[0] = std::log10(c->par.edge_quasi_cutoff_cliques);
    P.log_cutoff[1] = std::log10(c->par.edge_quasi_cutoff_mixed);
    P.log_cutoff[2] = std::log10(c->par.edge_quasi_cutoff_single);
    P.min_overlap[0] = c->par.min_overlap_cliques;
    P.min_overlap[1] = c->par.min_overlap_single;
    P.log10_quarter = std::log10(0.25);
    P.amb_tol = c->par.ambiguity_rel_tol > 0 ? c->par.ambiguity_rel_tol : 1e-12;
    P.table = c->d_table;
    P.simpson = c->d_simpson;
    P.simpson_q = c->d_simpson_q;
    P.ladder = c->d_ladder;
    P.K = c->K;
    P.KK = c->K * c->K;
    P.simpson_len = (int)c->simpson_host.size();
    P.ladder_len = c->ladder_len;
    P.no_prob0 = c->par.no_prob0;
    return P;
}

using hc::parallel_ranges;

int compute_layout(hc_ctx* c, int row_begin, int row_end) {
    if (c->lay_begin == row_begin && c->lay_end == row_end) return HC_OK;
    const int nrows = row_end - row_begin;
    c->lay_col0.assign(nrows, 0);
    c->lay_off.assign((size_t)nrows + 1, 0);
    parallel_ranges(nrows, 4096, [&](int k0, int k1) {
        for (int k = k0; k < k1; ++k) {
            const int i = row_begin + k;
            // first j whose running maximum of cov_last reaches cov_first[i]; everything before it is disjoint
            const auto it = std::lower_bound(c->cov_last_pmax.begin(), c->cov_last_pmax.begin() + i, c->cov_first[i]);
            c->lay_col0[k] = (int)(it - c->cov_last_pmax.begin()) & ~31;
        }
    });
    long long acc = 0;
    for (int k = 0; k < nrows; ++k) {
        c->lay_off[k] = acc;
        acc += (row_begin + k - c->lay_col0[k] + 31) / 32;
    }
    c->lay_off[nrows] = acc;
    c->lay_begin = row_begin;
    c->lay_end = row_end;
    return HC_OK;
}

// Slotted programs for the tiled kernel: the overlap phases of every program sit in slots 3, 7 and 11, the O(1)
// phases in front of them, the closing FIN / tail in slot 12 (hc_tile.cuh).
void build_tile_programs(std::vector<uint32_t>& out) {
    using namespace tile;
    out.assign((size_t)TP_COUNT * TP_WORDS, 0u);
    for (int k = 0; k < N_PROGS; ++k) {
        uint32_t* pg = &out[(size_t)k * TP_WORDS];
        int slot = 0, pre = 0;
        for (int t = 0; t < MAXPH && h_prog[k][t].op != PH_END; ++t) {
            const Phase ph = h_prog[k][t];
            switch (ph.op) {
                case PH_START: pg[slot * 4 + pre++] = TILE_MOP(M_START, 0, 0, ph.a, 0); break;
                case PH_TAIL_LT: pg[slot * 4 + pre++] = TILE_MOP(M_TLT, ph.x, ph.y, ph.a, 0); break;
                case PH_SEAT: pg[slot * 4 + pre++] = TILE_MOP(M_SEAT, ph.x, ph.y, 0, 0); break;
                case PH_OVL: pg[slot * 4 + 3] = TILE_MOP(M_OVL, ph.x, ph.y, ph.a, 0); slot++; pre = 0; break;
                case PH_OVL2: pg[slot * 4 + 3] = TILE_MOP(M_OVL2, ph.x, ph.y, ph.a, ph.b); slot++; pre = 0; break;
                case PH_OVL2_QX: pg[slot * 4 + 3] = TILE_MOP(M_OVLQX, ph.x, ph.y, ph.a, ph.b); slot++; pre = 0; break;
                case PH_FIN: pg[SLOT_FINAL] = TILE_MOP(M_FIN, 0, 0, ph.a, ph.b); break;
                default: break;
            }
            if (pre > 3 || slot > 3) abort();          // a program that does not fit the slot template
        }
    }
    const int single[3][2] = {{0, 0}, {1, 0}, {0, 1}};   // checkGapsSingle(segment of ap1, segment of ap2)
    for (int k = 0; k < 3; ++k) {
        uint32_t* pg = &out[(size_t)(TP_SINGLE00 + k) * TP_WORDS];
        pg[0] = TILE_MOP(M_SPRE, single[k][0], single[k][1], 0, 0);
        pg[SLOT_OVL0] = TILE_MOP(M_SOVL, single[k][0], single[k][1], 0, 0);
        pg[SLOT_FINAL] = TILE_MOP(M_SPOST, single[k][0], single[k][1], 0, 0);
    }
    uint32_t* two = &out[(size_t)TP_TWO_SINGLES * TP_WORDS];   // :436-440: segments 1/1 then segments 2/2
    two[0] = TILE_MOP(M_SPRE, 0, 0, 0, 0);
    two[SLOT_OVL0] = TILE_MOP(M_SOVL, 0, 0, 0, 0);
    two[4] = TILE_MOP(M_SPOST, 0, 0, 0, 0);
    two[5] = TILE_MOP(M_SPRE, 1, 1, 0, 0);
    two[SLOT_OVL1] = TILE_MOP(M_SOVL, 1, 1, 0, 0);
    two[SLOT_FINAL] = TILE_MOP(M_SPOST, 1, 1, 0, 0);
}

struct TileGeom { int maxw, abuf_bytes, nwarps; size_t smem; };

// First-generation tiled kernel: the warp count shrinks (24 -> 8) until tables + column tile + row buffers fit a CTA.
TileGeom tile_geometry(const hc_ctx* c) {
    TileGeom g;
    g.maxw = (c->max_ops + 1) / 2 + 4;
    g.abuf_bytes = ((c->max_ops * 2 + 15) & ~15) + 32;
    g.nwarps = 0;
    g.smem = ~size_t(0);
    for (int w = tile::kWarps; w >= 8; w -= 4) {
        size_t o = 0;
        o += ((size_t)2 * c->K * c->K * 8 + 15) & ~(size_t)15;
        o += (size_t)tile::TP_COUNT * tile::TP_WORDS * 4;
        o += 16 * 32 * 4;
        o += (size_t)g.maxw * 128;
        o += (size_t)w * 2 * (64 + g.abuf_bytes);
        o += (size_t)w * 2 * 8;
        o += 8 * 8;
        o += 16;
        if (o <= c->smem_optin) { g.nwarps = w; g.smem = o; break; }
    }
    return g;
}

// The instantiation of the first-generation kernel for (streamed, allele-frequency map, full 24-warp CTA).
typedef void (*Tile1Kernel)(const tile::Args);
Tile1Kernel tile1_kernel(bool stream, bool map, bool full) {
    const constexpr int W = tile::kWarps;
    if (stream) return map ? (full ? tile::score_tile_kernel<true, true, W> : tile::score_tile_kernel<true, true, 0>)
                           : (full ? tile::score_tile_kernel<true, false, W> : tile::score_tile_kernel<true, false, 0>);
    return map ? (full ? tile::score_tile_kernel<false, true, W> : tile::score_tile_kernel<false, true, 0>)
               : (full ? tile::score_tile_kernel<false, false, W> : tile::score_tile_kernel<false, false, 0>);
}

// The tiled kernel needs a column tile that fits (with an allele-frequency map its MAP variant walks the tails).
bool tile_eligible(const hc_ctx* c) {
    return tile_geometry(c).nwarps > 0;
}

// Streamed variant of the first-generation kernel: only the two probability tables, the programs and the headers live in
// shared memory (16*K*K + ~7 KB: any K up to 118), the op streams are read in place. Takes the batches neither tiled kernel
// can hold: super-read rounds (reads of several thousand columns, quality alphabets of 60-94 codes).
TileGeom stream_geometry(const hc_ctx* c) {
    TileGeom g;
    g.maxw = 0;
    g.abuf_bytes = 0;
    g.nwarps = tile::kWarps;
    size_t o = 0;
    o += ((size_t)2 * c->K * c->K * 8 + 15) & ~(size_t)15;
    o += (size_t)tile::TP_COUNT * tile::TP_WORDS * 4;
    o += (size_t)tile::kWarps * 16 * 32 * 4;                       // every warp's own copy of its column block's headers
    o += (size_t)tile::kWarps * 2 * 64;
    o += tile::kWarps * 2 * 8;
    o += 8 * 8;
    o += 16;
    g.smem = o;
    return g;
}
bool stream_eligible(const hc_ctx* c) {
    return c->max_ops < 65536 && stream_geometry(c).smem <= c->smem_optin;
}

// Second-generation tiled kernel: 9-row table (2304*K bytes split, 72*K*K bytes as doubles), 32-bit row records. The warp count shrinks
// (24 -> 8) until the CTA fits; below 8 warps the first-generation kernel or the generic walker takes over.
struct Tile2Geom { int maxw, abuf_bytes, nwarps; size_t smem; };

// Table layout of the second-generation tiled kernel: split low/high words for K <= split_max_k() codes, doubles above.
int split_max_k() {
    static const int v = [] { const char* e = std::getenv("HC_T2_SPLIT_MAXK"); return e ? std::atoi(e) : tile2::kSplitMaxK; }();
    return std::min(v, 32);
}
int row_pad() {      // extra doubles per table row (HC_T2_ROWPAD, experiments on the bank mapping of neighbouring rows)
    static const int v = [] { const char* e = std::getenv("HC_T2_ROWPAD"); return e ? std::max(0, std::atoi(e)) : 0; }();
    return v;
}
uint32_t tile2_row_bytes(int K) { return K <= split_max_k() ? 256u : 8u * (uint32_t)(K + row_pad()); }
uint32_t tile2_entry_bytes(int K) { return K <= split_max_k() ? 4u : 8u; }
size_t tile2_table_words(int K) { return (size_t)tile2::kRows * K * tile2_row_bytes(K) / 4; }

Tile2Geom tile2_geometry(const hc_ctx* c) {
    Tile2Geom g;
    g.maxw = (c->max_ops + 1) / 2 + 6;
    g.abuf_bytes = ((c->max_ops * 4 + 15) & ~15) + 32;
    g.nwarps = 0;
    g.smem = 0;
    for (int w = tile2::kMaxWarps; w >= 8; w -= 4) {
        size_t o = 0;
        o += ((size_t)tile2_table_words(c->K) * 4 + 15) & ~(size_t)15;
        o += (size_t)tile::TP_COUNT * tile::TP_WORDS * 4;
        o += 16 * 32 * 4;
        o += (size_t)g.maxw * 128;
        o += (size_t)w * 2 * (64 + g.abuf_bytes);
        o += (size_t)w * 2 * 8;
        o += 8 * 8;
        o += 16;
        if (o <= c->smem_optin) { g.nwarps = w; g.smem = o; break; }
    }
    return g;
}

bool tile2_eligible(const hc_ctx* c) {
    if (!c->have_formats) return false;        // an allele-frequency map is fine here: the MAP variant walks the tails
    return tile2_geometry(c).nwarps > 0;
}

// Column windows are (mean band width) / item_window_div() reads wide (HC_ITEM_WINDOW_DIV overrides, for experiments)
int item_window_div() {
    static const int v = [] { const char* e = std::getenv("HC_ITEM_WINDOW_DIV"); return e ? std::max(1, std::atoi(e)) : 16; }();
    return v;
}

int warp_item_min_rows() {
    static const int v = [] { const char* e = std::getenv("HC_WARP_ITEM_ROWS"); return e ? std::max(1, std::atoi(e)) : 8; }();
    return v;
}

// Column reads of a window are grouped by indel signature before the end sort (HC_ITEM_SIG=0 turns it off); a signature
// carried by fewer than item_sig_min() reads of the window (HC_ITEM_SIG_MIN) goes to the window's common "rare" group.
bool item_sig() { const char* e = std::getenv("HC_ITEM_SIG"); return !(e && e[0] == '0'); }     // read at every plan build
int item_sig_min() {
    static const int v = [] { const char* e = std::getenv("HC_ITEM_SIG_MIN"); return e ? std::max(1, std::atoi(e)) : 8; }();
    return v;
}

// The order in which the persistent CTAs take the work items: row-major unless HC_ITEM_ORDER=blocks (read at every plan build)
bool items_by_rows() { const char* e = std::getenv("HC_ITEM_ORDER"); return !(e && e[0] == 'b'); }

// Work items of the tiled kernels for rows [w0,w1) of the laid-out range (c->lay_begin .. c->lay_end).
// Columns (the earlier reads j) are taken in windows of W consecutive reads; inside a window they are sorted by
// (single/paired, indel signature, last covered position) and cut into blocks of 32. For a later row read i the length
// of the lock-step loop against j depends on min(end_i, end_j) but not on j's start (j starts first), so the 32 lanes of
// a block run loops of nearly the same length; and a block stops being scheduled at the first row that starts
// behind the block's furthest end, so lanes are not spent on the reference's first reject (:1252).
// The indel signature: a walk stops at the first column where one read shows an insertion or deletion the other lacks
// (overlapCheckgap's gap, :173-231) -- on C2 30 % of the walked pairs near a haplotype's indel -- and a warp runs as long as
// its longest lane. Column reads that carry the same indels stop against a given row read at the same column or not at
// all, so grouping them lets whole warps stop (CPU simulation over the oracle's per-pair column counts, C2 rows 50000-50300:
// lane-columns lost to stopped lanes 22.5 % -> 8.5-11 %). Which reads share a block changes no result. Measured on one B200:
// C4 (every read covers two of the ten haplotype indels) 19.61 -> 17.44 s per pass at full size, 839 -> 752 ms at 200k
// fragments; C2 (one read in eleven carries an indel) 56.6 -> 56.9 ms (profiles/r02_sig_sweep.json).
// An item = (block, chunk of rows that can reach it). Items and blocks are appended to `items` / `colidx`.
void build_items_range(hc_ctx* c, int w0, int w1, int n_ctas, bool warp_items, std::vector<int4>& items, std::vector<int32_t>& colidx) {
    const int nrows = w1 - w0;
    if (nrows <= 0 || w1 < 2) return;
    const int32_t* col0 = c->lay_col0.data() + (w0 - c->lay_begin);
    int32_t jlo = INT32_MAX;
    long long band = 0;
    for (int k = 0; k < nrows; ++k) { jlo = std::min(jlo, col0[k]); band += (w0 + k) - col0[k]; }
    const int jhi = w1 - 1;                                        // columns [jlo, jhi)
    int W = (int)(band / nrows / item_window_div());
    W = std::max(32, std::min(4096, (W + 16) / 32 * 32));
    struct Blk { int first_j; int maxend; };
    // windows are independent: block / colidx positions follow from the window sizes, so threads write in place
    const int nwin = jhi > jlo ? (jhi - jlo + W - 1) / W : 0;
    std::vector<size_t> blk0((size_t)nwin + 1, 0);
    for (int w = 0; w < nwin; ++w) {
        const int c0 = jlo + w * W, c1 = std::min(c0 + W, jhi);
        blk0[w + 1] = blk0[w] + (size_t)(c1 - c0 + 31) / 32;
    }
    std::vector<Blk> blks(blk0[nwin]);
    const size_t base_blk = colidx.size() / 32;
    colidx.resize((base_blk + blk0[nwin]) * 32, -1);
    int32_t* cix = colidx.data() + base_blk * 32;
    // not for the streamed kernel's per-warp items: on super-read rounds, where nearly every read carries indels of its own,
    // the grouping costs the end sort more than it gains (C5 6.72 -> 6.97 ms)
    const bool by_sig = item_sig() && !warp_items && c->indel_sig.size() == (size_t)c->n_reads;
    const size_t sig_min = (size_t)item_sig_min();
    parallel_ranges(nwin, 4, [&](int wa, int wb) {
        // key = single/paired (bit 63) | signature group (12 bits) | last covered position above the window's lowest (27 bits) |
        // read index inside the window (12 bits: W <= 4096)
        std::vector<uint64_t> order;
        std::vector<uint32_t> sigs;
        for (int w = wa; w < wb; ++w) {
            const int c0 = jlo + w * W, c1 = std::min(c0 + W, jhi);
            order.resize(c1 - c0);
            sigs.clear();
            int lowest = INT32_MAX;
            for (int j = c0; j < c1; ++j) lowest = std::min(lowest, c->cov_last[j]);
            if (by_sig) {                            // signatures carried by at least sig_min reads of this window, sorted
                for (int j = c0; j < c1; ++j) if (c->indel_sig[j]) sigs.push_back(c->indel_sig[j]);
                std::sort(sigs.begin(), sigs.end());
                size_t keep = 0;
                for (size_t a = 0; a < sigs.size();) {
                    size_t b = a;
                    while (b < sigs.size() && sigs[b] == sigs[a]) ++b;
                    if (b - a >= sig_min) sigs[keep++] = sigs[a];
                    a = b;
                }
                sigs.resize(keep);                   // at most W / sig_min <= 4096 groups
            }
            for (int j = c0; j < c1; ++j) {
                uint64_t group = 0;                  // no indel
                if (by_sig && c->indel_sig[j]) {
                    const auto it = std::lower_bound(sigs.begin(), sigs.end(), c->indel_sig[j]);
                    group = (it != sigs.end() && *it == c->indel_sig[j]) ? std::min<uint64_t>((uint64_t)(it - sigs.begin()) + 1, 4094) : 4095;
                }
                const uint64_t rel_end = std::min<uint64_t>((uint64_t)((long long)c->cov_last[j] - lowest), (1ull << 27) - 1);
                order[j - c0] = ((uint64_t)(c->is_single[j] ? 0 : 1) << 63) | (group << 39) | (rel_end << 12) | (uint64_t)(j - c0);
            }
            std::sort(order.begin(), order.end());
            for (size_t p = 0; p < order.size(); p += 32) {
                Blk b{INT32_MAX, INT32_MIN};
                const size_t bi = blk0[w] + p / 32;
                for (size_t q = p; q < p + 32 && q < order.size(); ++q) {
                    const int j = c0 + (int)(order[q] & 4095ull);
                    cix[bi * 32 + (q - p)] = j;
                    b.first_j = std::min(b.first_j, j);
                    b.maxend = std::max(b.maxend, c->cov_last[j]);
                }
                blks[bi] = b;
            }
        }
    });
    // rows that can reach a block: i > first_j and cov_first[i] <= maxend; cov_first_smin is the suffix minimum of
    // cov_first (monotone), so the admissible rows are a prefix
    std::vector<int> first(blks.size()), last(blks.size());
    long long total = 0;
    for (size_t b = 0; b < blks.size(); ++b) {
        first[b] = std::max(w0, blks[b].first_j + 1);
        last[b] = std::min(w1, (int)(std::upper_bound(c->cov_first_smin.begin(), c->cov_first_smin.end(), blks[b].maxend) -
                                     c->cov_first_smin.begin())) - 1;
        if (last[b] >= first[b]) total += last[b] - first[b] + 1;
    }
    long long chunk = (total + (long long)n_ctas * 16 - 1) / ((long long)n_ctas * 16);
    chunk = std::max<long long>(32, std::min<long long>(chunk, 4096));
    if (warp_items) {                            // streamed kernel: an item is taken by one warp, not by the 24 warps of a CTA
        chunk = (total + (long long)n_ctas * tile::kWarps * 8 - 1) / ((long long)n_ctas * tile::kWarps * 8);
        chunk = std::max<long long>(warp_item_min_rows(), std::min<long long>(chunk, 256));
    }
    const size_t item0 = items.size();
    for (size_t b = 0; b < blks.size(); ++b)
        for (long long r = first[b]; r <= last[b]; r += chunk)
            items.push_back(make_int4((int)(base_blk + b), (int)r, (int)std::min<long long>(r + chunk, (long long)last[b] + 1), 0));
    // Order of the items = order in which the persistent CTAs take them. Row-major: the CTAs running at the same time work on
    // the same few thousand rows against different column blocks, so a row's stream and its adjacency words come from HBM
    // once and from L2 afterwards. Block-major (HC_ITEM_ORDER=blocks, the order of generation) shares a column block instead
    // and re-fetches the rows: C3 on one B200, ncu: 784 GB read + 43 GB written per pass against 28 GB + 4 GB row-major (L2
    // hit rate 75 % -> 98 %), at the same kernel time (4933 -> 4927 ms) and 30 W less (profiles/r02_item_order_traffic.txt).
    if (items_by_rows())
        std::stable_sort(items.begin() + item0, items.end(), [chunk](const int4& a, const int4& b) { return a.y / chunk < b.y / chunk; });
}

// Adjacency words per row window (HC_WINDOW_WORDS overrides): the rows of a band are scored in windows so that the words
// of a finished window can travel to the host while the next window is scored.
long long window_words() {
    static const long long v = [] { const char* e = std::getenv("HC_WINDOW_WORDS"); return e ? std::max(1ll, std::atoll(e)) : (4ll << 20); }();
    return v;
}
constexpr int kMaxWindows = 16;

// The scoring plan of rows [row_begin,row_end) (needs compute_layout of the same range): row windows of about equal
// numbers of adjacency words, each with its own work items.
void build_plan(hc_ctx* c, int row_begin, int row_end, int n_ctas, bool windows, bool warp_items = false) {
    if (c->plan_begin == row_begin && c->plan_end == row_end && c->plan_ctas == n_ctas && c->plan_windows == windows &&
        c->plan_by_rows == items_by_rows() && c->plan_warp_items == warp_items && c->plan_sig == item_sig()) return;
    c->items.clear();
    c->colidx.clear();
    c->wins.clear();
    const int nrows = row_end - row_begin;
    const long long total = nrows > 0 ? c->lay_off[nrows] : 0;
    int nw = 1;
    if (windows) nw = (int)std::max<long long>(1, std::min<long long>(kMaxWindows, (total + window_words() - 1) / window_words()));
    int r = row_begin;
    for (int w = 0; w < nw && r < row_end; ++w) {
        int r1 = row_end;
        if (w + 1 < nw) {
            const long long want = total * (w + 1) / nw;
            r1 = row_begin + (int)(std::lower_bound(c->lay_off.begin(), c->lay_off.begin() + nrows + 1, want) - c->lay_off.begin());
            r1 = std::min(std::max(r1, r + 1), row_end);
        }
        hc_ctx::Win win;
        win.r0 = r; win.r1 = r1;
        win.item0 = c->items.size();
        build_items_range(c, r, r1, n_ctas, warp_items, c->items, c->colidx);
        win.n_items = c->items.size() - win.item0;
        win.word0 = c->lay_off[r - row_begin];
        win.n_words = c->lay_off[r1 - row_begin] - win.word0;
        c->wins.push_back(win);
        r = r1;
    }
    c->plan_begin = row_begin; c->plan_end = row_end; c->plan_ctas = n_ctas; c->plan_windows = windows; c->plan_by_rows = items_by_rows(); c->plan_warp_items = warp_items; c->plan_sig = item_sig();
}

// similarityCriterion's last lines (src/NewEdgeCalculator.cpp:1132-1140) on the host, with glibc's log10: the decision for
// a pair the device flagged as within ambiguity_rel_tol of the threshold. (probM, prob0, cc, tc) are bit-exact.
bool host_decide(const hc_ctx* c, double probM, double prob0, int cc, int tc, double log_cutoff, double* score_out) {
    double prob, potence;
    if (c->par.no_prob0) { prob = std::log10(probM); potence = 1.0 / cc; }
    else { prob = std::log10(probM) + prob0; potence = 1.0 / (cc + tc); }
    const double score = prob * potence;
    if (score_out) *score_out = score;
    return score >= log_cutoff;
}

}  // namespace

extern "C" {

int hc_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* hc_last_error(const hc_ctx* ctx) { return ctx ? ctx->err.c_str() : hc::global_error(); }

int hc_create(const hc_params* params, hc_ctx** out) {
    if (!params || !out) { hc::set_global_error("hc_create: null argument"); return HC_E_INVALID; }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        hc::set_global_error("hc_create: no CUDA device (this library has no CPU scoring path)");
        return HC_E_CUDA;
    }
    if (params->device < 0 || params->device >= ndev) { hc::set_global_error("hc_create: bad device ordinal"); return HC_E_INVALID; }
    hc_ctx* c = new hc_ctx();
    c->par = *params;
    c->device = params->device;
    if (params->simpson && params->simpson_len > 0) c->simpson_host.assign(params->simpson, params->simpson + params->simpson_len);
    bool any = false;
    for (double d : c->simpson_host) any |= (d != 0.0);
    if (!any) c->simpson_host.clear();           // no entry set: every tail column adds log10(0.25)
    c->par.simpson = nullptr;
    c->eq.resize(127 * 127);
    c->uneq.resize(127 * 127);
    hc::build_match_tables(c->eq.data(), c->uneq.data());
    auto fail = [&](const char* what, cudaError_t err) {
        hc::set_global_error(std::string("hc_create: ") + what + ": " + cudaGetErrorString(err));
        hc_destroy(c);
        return HC_E_CUDA;
    };
    if ((e = cudaSetDevice(c->device)) != cudaSuccess) return fail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, c->device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
    c->sm_count = prop.multiProcessorCount;
    c->smem_optin = prop.sharedMemPerBlockOptin;
    {
        std::vector<uint32_t> tp;
        build_tile_programs(tp);
        if ((e = cudaMalloc((void**)&c->d_tprog, tp.size() * sizeof(uint32_t))) != cudaSuccess) return fail("cudaMalloc", e);
        if ((e = cudaMemcpy(c->d_tprog, tp.data(), tp.size() * sizeof(uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy", e);
        if ((e = cudaMalloc((void**)&c->d_item_counter, kMaxWindows * sizeof(int))) != cudaSuccess) return fail("cudaMalloc", e);
    }
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&c->ev_go, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    for (int k = 0; k < kMaxWindows; ++k)
        if ((e = cudaEventCreateWithFlags(&c->win_ev[k], cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaMalloc((void**)&c->d_stats, 2 * sizeof(BandStatsDev))) != cudaSuccess) return fail("cudaMalloc", e);
    if (!c->simpson_host.empty()) {
        const size_t bytes = c->simpson_host.size() * sizeof(double);
        if ((e = cudaMalloc((void**)&c->d_simpson, bytes)) != cudaSuccess) return fail("cudaMalloc", e);
        if ((e = cudaMemcpy(c->d_simpson, c->simpson_host.data(), bytes, cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy", e);
        std::vector<double> addend(c->simpson_host);
        const double lq = std::log10(0.25);
        for (double& d : addend) if (d == 0.0) d = lq;
        if ((e = cudaMalloc((void**)&c->d_simpson_q, bytes)) != cudaSuccess) return fail("cudaMalloc", e);
        if ((e = cudaMemcpy(c->d_simpson_q, addend.data(), bytes, cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy", e);
    }
    *out = c;
    return HC_OK;
}

void hc_destroy(hc_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaFree(c->d_hdr); cudaFree(c->d_ops); cudaFree(c->d_table); cudaFree(c->d_simpson); cudaFree(c->d_simpson_q); cudaFree(c->d_ladder);
    cudaFree(c->d_col0); cudaFree(c->d_word_off); cudaFree(c->d_adj); cudaFree(c->d_stats); cudaFree(c->d_amb);
    cudaFree(c->d_pi); cudaFree(c->d_pj); cudaFree(c->d_res); cudaFree(c->d_flush);
    cudaFree(c->d_aux); cudaFree(c->d_col2op); cudaFree(c->d_tprog); cudaFree(c->d_items); cudaFree(c->d_item_counter);
    cudaFree(c->d_opsA); cudaFree(c->d_opsF); cudaFree(c->d_tab9);
    cudaFree(c->d_logsum); cudaFree(c->d_simpson_ln); cudaFree(c->d_qres); cudaFree(c->d_qseg);
    cudaFree(c->d_detail); cudaFree(c->d_colidx);
    for (int k = 0; k < kMaxWindows; ++k) if (c->win_ev[k]) cudaEventDestroy(c->win_ev[k]);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->ev_go) cudaEventDestroy(c->ev_go);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int hc_set_params(hc_ctx* c, const hc_params* p) {
    if (!c || !p) return HC_E_INVALID;
    const int dev = c->par.device;
    c->par = *p;
    c->par.device = dev;
    c->par.simpson = nullptr;
    return HC_OK;
}

int hc_l2_flush(hc_ctx* c) {
    if (!c) return HC_E_INVALID;
    CU_TRY(c, cudaSetDevice(c->device));
    if (!c->d_flush) {
        c->flush_bytes = (size_t)256 << 20;                      // 2x the 126 MB L2
        CU_TRY(c, cudaMalloc(&c->d_flush, c->flush_bytes));
    }
    CU_TRY(c, cudaMemsetAsync(c->d_flush, 0x5a, c->flush_bytes, c->stream));
    return HC_OK;
}

void* hc_host_alloc(int64_t bytes) {
    void* p = nullptr;
    if (bytes <= 0 || cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void hc_host_free(void* p) { if (p) cudaFreeHost(p); }

int hc_get_tables(const hc_ctx* c, double* eq, double* uneq) {
    if (!c || !eq || !uneq) return HC_E_INVALID;
    std::memcpy(eq, c->eq.data(), sizeof(double) * 127 * 127);
    std::memcpy(uneq, c->uneq.data(), sizeof(double) * 127 * 127);
    return HC_OK;
}

int hc_upload(hc_ctx* c, const hc_packed* p) {
    if (!c || !p) return HC_E_INVALID;
    CU_TRY(c, cudaSetDevice(c->device));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    if (c->d_logsum) { cudaFree(c->d_logsum); c->d_logsum = nullptr; }     // depends on the quality codes of the batch
    c->have_qseg = false;
    std::memcpy(c->qual_chars, p->qual_chars, sizeof(c->qual_chars));
    c->have_formats = false;
    int rc;
    c->plan_begin = c->plan_end = -1;
    c->n_reads = p->n_reads;
    c->K = std::max(p->n_qual_codes, 1);
    c->max_ops = p->max_ops_per_read;
    c->lay_begin = c->lay_end = -1;
    const size_t nh = std::max(p->n_reads, 1);
    if ((rc = grow(c, c->d_hdr, c->cap_hdr, nh)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_ops, c->cap_ops, (size_t)p->n_ops_total + 64)) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_hdr, p->hdr, (size_t)p->n_reads * sizeof(hc_read_hdr), cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_ops, p->ops, ((size_t)p->n_ops_total + 64) * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
    // probability tables re-indexed by quality code: [equal | unequal][code1][code2]
    const int K = c->K;
    std::vector<double> tab((size_t)2 * K * K, 0.0);
    for (int a = 0; a < p->n_qual_codes; ++a)
        for (int b = 0; b < p->n_qual_codes; ++b) {
            const int qa = p->qual_chars[a], qb = p->qual_chars[b];
            tab[(size_t)a * K + b] = c->eq[qa * 127 + qb];
            tab[(size_t)K * K + (size_t)a * K + b] = c->uneq[qa * 127 + qb];
        }
    if ((rc = grow(c, c->d_table, c->cap_table, tab.size())) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_table, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    std::vector<double> ladder;
    c->ladder_len = 2 * c->max_ops + 2;
    hc::build_prob0_ladder(ladder, c->ladder_len - 1);
    if ((rc = grow(c, c->d_ladder, c->cap_ladder, ladder.size())) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_ladder, ladder.data(), ladder.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    // prefix counts of the segments that are not "clips, plain M, clips" (the tiled kernel's O(1) tails)
    if ((rc = grow(c, c->d_aux, c->cap_aux, (size_t)p->n_ops_total + 2 * (size_t)p->n_reads + 64)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_col2op, c->cap_col2op, (size_t)p->n_ops_total + 64)) != HC_OK) return rc;
    if (p->n_reads > 0) {
        tile::build_aux_kernel<<<(2 * p->n_reads + 127) / 128, 128, 0, c->stream>>>(c->d_hdr, c->d_ops, p->n_reads, c->d_aux, c->d_col2op);
        CU_TRY(c, cudaGetLastError());
    }
    // operand formats of the second-generation tiled kernel (hc_tile2.cuh): A-words, B-records, and the table as
    // 9 rows per first quality (un x4, EQ, un x4), low words then high words. Skipped when the table alone
    // would not fit a CTA's shared memory (the kernel is then not eligible).
    std::vector<uint32_t> tab9;
    if (tile2_table_words(K) * 4 + 65536 <= c->smem_optin) {
        tab9.assign(tile2_table_words(K), 0u);
        // order of the quality codes inside a table row (HC_T2_PERM, experiments on bank conflicts): 0 identity,
        // 1 even codes first, 2 reversed, 3 stride 5
        tile2::ColumnOrder ord{};
        {
            static const int mode = [] { const char* e = std::getenv("HC_T2_PERM"); return e ? std::atoi(e) : 0; }();
            std::vector<int> seq(K);
            for (int q = 0; q < K; ++q) seq[q] = q;
            if (mode == 1) { int w = 0; for (int q = 0; q < K; q += 2) seq[w++] = q; for (int q = 1; q < K; q += 2) seq[w++] = q; }
            else if (mode == 2) std::reverse(seq.begin(), seq.end());
            else if (mode == 3) { std::vector<int> t; for (int s0 = 0; s0 < 5; ++s0) for (int q = s0; q < K; q += 5) t.push_back(q); seq = t; }
            for (int pos = 0; pos < K; ++pos) ord.slot[seq[pos]] = (uint8_t)pos;
        }
        const bool split = K <= split_max_k();
        for (int a = 0; a < K; ++a)
            for (int r = 0; r < tile2::kRows; ++r)
                for (int b = 0; b < K; ++b) {
                    const double v = tab[(r == 4 ? 0 : (size_t)K * K) + (size_t)a * K + b];
                    uint64_t bits;
                    std::memcpy(&bits, &v, 8);
                    const size_t row = (size_t)a * tile2::kRows + r;
                    if (split) {                       // 32 low words, then 32 high words
                        tab9[row * 64 + ord.slot[b]] = (uint32_t)bits;
                        tab9[row * 64 + 32 + ord.slot[b]] = (uint32_t)(bits >> 32);
                    } else {                           // K doubles
                        const size_t Kp = tile2_row_bytes(K) / 8;
                        tab9[(row * Kp + ord.slot[b]) * 2] = (uint32_t)bits;
                        tab9[(row * Kp + ord.slot[b]) * 2 + 1] = (uint32_t)(bits >> 32);
                    }
                }
        const long long nrec = (long long)p->n_ops_total + 64;
        if ((rc = grow(c, c->d_tab9, c->cap_tab9, tab9.size())) != HC_OK) return rc;
        CU_TRY(c, cudaMemcpyAsync(c->d_tab9, tab9.data(), tab9.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
        if ((rc = grow(c, c->d_opsA, c->cap_opsA, (size_t)nrec)) != HC_OK) return rc;
        if ((rc = grow(c, c->d_opsF, c->cap_opsF, (size_t)nrec)) != HC_OK) return rc;
        c->have_formats = true;
        tile2::build_formats_kernel<<<(unsigned)((nrec / 8 + 255) / 256), 256, 0, c->stream>>>(c->d_ops, nrec, tile2_row_bytes(K), tile2_entry_bytes(K), ord, c->d_opsA, c->d_opsF);
        CU_TRY(c, cudaGetLastError());
    }
    c->cov_first.resize(p->n_reads);
    c->cov_last_pmax.resize(p->n_reads);
    c->cov_last.resize(p->n_reads);
    c->is_single.resize(p->n_reads);
    c->cov_first_smin.resize(p->n_reads);
    c->ncov_flags.resize(p->n_reads);
    int run = INT32_MIN;
    for (int r = 0; r < p->n_reads; ++r) {
        c->ncov_flags[r] = p->hdr[r].ncov_flags;
        c->cov_first[r] = p->hdr[r].cov_first;
        c->cov_last[r] = p->hdr[r].cov_last;
        c->is_single[r] = (p->hdr[r].ncov_flags & HC_FLAG_SINGLE_END) ? 1 : 0;
        run = std::max(run, p->hdr[r].cov_last);
        c->cov_last_pmax[r] = run;
    }
    int smin = INT32_MAX;
    for (int r = p->n_reads - 1; r >= 0; --r) { smin = std::min(smin, c->cov_first[r]); c->cov_first_smin[r] = smin; }
    // indel signatures (build_items_range groups column reads by them): FNV-1a over the (reference position, op) of every I
    // and D op; only segments flagged "not plain" are read
    c->indel_sig.assign((size_t)p->n_reads, 0u);
    if (p->n_reads > 0 && (tile_eligible(c) || tile2_eligible(c)))      // the streamed kernel's per-warp items do not use them
        hc::parallel_ranges(p->n_reads, 1024, [&](int r0, int r1) {
            for (int r = r0; r < r1; ++r) {
                const hc_read_hdr& h = p->hdr[r];
                if (!((h.seg_flags[0] | h.seg_flags[1]) & HC_SEG_NOT_PLAIN)) continue;
                uint32_t f = 2166136261u;
                bool any = false;
                for (int sg = 0; sg < 2; ++sg) {
                    if (!(h.seg_flags[sg] & HC_SEG_NOT_PLAIN)) continue;
                    const uint16_t* o = p->ops + h.ops_off[sg];
                    int pos = h.ext[2 * sg];                      // clip-extended start: every op but I takes one position
                    for (int k = 0; k < h.n_ops[sg]; ++k) {
                        const uint32_t op = HC_REC_OP((uint32_t)o[k]);
                        if (op == HC_OP_I || op == HC_OP_D) {
                            f = (f ^ (uint32_t)pos) * 16777619u;
                            f = (f ^ op) * 16777619u;
                            any = true;
                        }
                        if (op != HC_OP_I) ++pos;
                    }
                }
                if (any) c->indel_sig[r] = f ? f : 1u;
            }
        });
    // the band layout and the tiled kernels' work items of the whole batch depend on the headers only: computed here, on
    // the host, while the copies and the device-side builds above are in flight (hc_score_band(0, n) finds them cached)
    if (p->n_reads > 1 && (tile_eligible(c) || tile2_eligible(c) || stream_eligible(c))) {
        compute_layout(c, 0, p->n_reads);
        build_plan(c, 0, p->n_reads, c->sm_count, true, !tile_eligible(c) && !tile2_eligible(c));
    }
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    return HC_OK;
}

int hc_band_layout(hc_ctx* c, int32_t row_begin, int32_t row_end, int32_t* col0, int64_t* word_off, int64_t* n_words) {
    if (!c) return HC_E_INVALID;
    if (!c->d_hdr) { c->err = "hc_band_layout: no batch uploaded"; return HC_E_STATE; }
    if (row_begin < 0 || row_end < row_begin || row_end > c->n_reads) { c->err = "hc_band_layout: bad row range"; return HC_E_INVALID; }
    compute_layout(c, row_begin, row_end);
    const int nrows = row_end - row_begin;
    if (col0) std::memcpy(col0, c->lay_col0.data(), sizeof(int32_t) * nrows);
    if (word_off) for (int k = 0; k <= nrows; ++k) word_off[k] = c->lay_off[k];
    if (n_words) *n_words = c->lay_off[nrows];
    return HC_OK;
}

// Host side of the near-threshold pairs: fetch the device's list, re-decide each pair with glibc's log10, patch the
// adjacency bit on the host (`adj`, may be null) and on the device, and fix the edge count. n = pairs in c->amb_host.
static int redecide_ambiguous(hc_ctx* c, int row_begin, uint32_t* adj, long long* edge_delta) {
    *edge_delta = 0;
    for (hc_ambiguous& a : c->amb_host) {
        double score = a.score;
        const bool host_edge = host_decide(c, a.probM, a.prob0, a.cc, a.tc, a.cutoff, &score);
        a.edge_device = a.edge;
        a.score = score;
        a.edge = host_edge ? 1 : 0;
        if (a.edge == a.edge_device) continue;
        *edge_delta += a.edge - a.edge_device;
        const int k = a.i - row_begin;
        const int rel = a.j - c->lay_col0[k];
        const long long w = c->lay_off[k] + (rel >> 5);
        const uint32_t bit = 1u << (rel & 31);
        uint32_t word = 0;
        CU_TRY(c, cudaMemcpy(&word, c->d_adj + w, sizeof(word), cudaMemcpyDeviceToHost));
        word = host_edge ? (word | bit) : (word & ~bit);
        CU_TRY(c, cudaMemcpy(c->d_adj + w, &word, sizeof(word), cudaMemcpyHostToDevice));
        if (adj) adj[w] = word;
    }
    return HC_OK;
}

static int score_band_impl(hc_ctx* c, int32_t row_begin, int32_t row_end, uint32_t* adj, int64_t n_words, hc_ambiguous* amb,
                           int32_t amb_cap, hc_band_stats* stats, hc_pair_result* detail) {
    if (!c) return HC_E_INVALID;
    if (!c->d_hdr) { c->err = "hc_score_band: no batch uploaded"; return HC_E_STATE; }
    if (row_begin < 0 || row_end < row_begin || row_end > c->n_reads) { c->err = "hc_score_band: bad row range"; return HC_E_INVALID; }
    if (amb_cap < 0 || (amb_cap > 0 && !amb)) { c->err = "hc_score_band: bad near-threshold list"; return HC_E_INVALID; }
    CU_TRY(c, cudaSetDevice(c->device));
    compute_layout(c, row_begin, row_end);
    const int nrows = row_end - row_begin;
    const long long total = c->lay_off[nrows];
    if ((adj || detail) && n_words < total) { c->err = "hc_score_band: output buffer smaller than hc_band_layout's n_words"; return HC_E_INVALID; }
    const bool can_tile1 = tile_eligible(c), can_tile2 = tile2_eligible(c), can_stream = stream_eligible(c);
    if ((c->kernel_mode == HC_KERNEL_TILED && !can_tile1) || (c->kernel_mode == HC_KERNEL_TILED2 && !can_tile2) ||
        (c->kernel_mode == HC_KERNEL_STREAMED && !can_stream)) {
        c->err = "hc_score_band: the tiled kernel was requested but this batch / parameter set is not eligible "
                 "(the tables and the column tile do not fit shared memory)";
        return HC_E_STATE;
    }
    const bool use_tile2 = can_tile2 && (c->kernel_mode == HC_KERNEL_AUTO || c->kernel_mode == HC_KERNEL_TILED2);
    const bool use_tile1 = !use_tile2 && can_tile1 && (c->kernel_mode == HC_KERNEL_AUTO || c->kernel_mode == HC_KERNEL_TILED);
    const bool use_stream = !use_tile2 && !use_tile1 && can_stream && (c->kernel_mode == HC_KERNEL_AUTO || c->kernel_mode == HC_KERNEL_STREAMED);
    const bool use_tile = use_tile1 || use_tile2 || use_stream;
    if (detail && !use_tile) { c->err = "hc_score_band_detail: only the tiled kernels write per-slot detail; use hc_score_pairs"; return HC_E_STATE; }
    int rc;
    if ((rc = grow(c, c->d_col0, c->col0_cap, (size_t)nrows + 1)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_word_off, c->word_off_cap, (size_t)nrows + 1)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_adj, c->adj_cap, (size_t)std::max<long long>(total, 1))) != HC_OK) return rc;
    if (detail && (rc = grow(c, c->d_detail, c->detail_cap, (size_t)std::max<long long>(total, 1) * 32)) != HC_OK) return rc;
    // the device always lists the near-threshold pairs: they are re-decided on the host whether or not the caller wants the report
    if ((rc = grow(c, c->d_amb, c->amb_cap, (size_t)std::max(amb_cap, 4096))) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_col0, c->lay_col0.data(), sizeof(int32_t) * nrows, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_word_off, c->lay_off.data(), sizeof(long long) * ((size_t)nrows + 1), cudaMemcpyHostToDevice, c->stream));
    BandStatsDev hs{};
    int launches = 0;
    for (int attempt = 0;; ++attempt) {
        CU_TRY(c, cudaMemsetAsync(c->d_stats, 0, 2 * sizeof(BandStatsDev), c->stream));    // [1]: HC_T2_PROFILE counters
        launches = 0;
        c->last_kernel = 0;
        if (use_tile && total > 0) {
            const TileGeom g = use_stream ? stream_geometry(c) : tile_geometry(c);
            const Tile2Geom g2 = tile2_geometry(c);
            const size_t smem_bytes = use_tile2 ? g2.smem : g.smem;
            const int threads = (use_tile2 ? g2.nwarps : g.nwarps) * 32;
            int per_sm = 0;
            const bool split = c->K <= split_max_k();
            const bool map = c->d_simpson != nullptr;
            if (use_tile2) {
                auto kern = split ? (map ? tile2::score_tile2_kernel<true, true> : tile2::score_tile2_kernel<true, false>)
                                  : (map ? tile2::score_tile2_kernel<false, true> : tile2::score_tile2_kernel<false, false>);
                CU_TRY(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
                CU_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem_bytes));
            } else {
                auto kern = tile1_kernel(use_stream, map, g.nwarps == tile::kWarps);
                CU_TRY(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
                CU_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem_bytes));
            }
            if (per_sm < 1) { c->err = "hc_score_band: tiled kernel does not fit on an SM"; return HC_E_CUDA; }
            const int n_ctas = c->sm_count * per_sm;
            build_plan(c, row_begin, row_end, n_ctas, detail == nullptr, use_stream);
            const size_t n_items = c->items.size();
            if ((rc = grow(c, c->d_items, c->items_cap, std::max<size_t>(n_items, 1))) != HC_OK) return rc;
            CU_TRY(c, cudaMemcpyAsync(c->d_items, c->items.data(), sizeof(int4) * n_items, cudaMemcpyHostToDevice, c->stream));
            if ((rc = grow(c, c->d_colidx, c->colidx_cap, std::max<size_t>(c->colidx.size(), 32))) != HC_OK) return rc;
            CU_TRY(c, cudaMemcpyAsync(c->d_colidx, c->colidx.data(), sizeof(int32_t) * c->colidx.size(), cudaMemcpyHostToDevice, c->stream));
            CU_TRY(c, cudaMemsetAsync(c->d_item_counter, 0, kMaxWindows * sizeof(int), c->stream));
            CU_TRY(c, cudaEventRecord(c->ev0, c->stream));
            // words of rows whose band holds no scored column block stay zero
            CU_TRY(c, cudaMemsetAsync(c->d_adj, 0, (size_t)total * sizeof(uint32_t), c->stream));
            if (detail) CU_TRY(c, cudaMemsetAsync(c->d_detail, 0, (size_t)total * 32 * sizeof(hc_pair_result), c->stream));
            tile2::Args2 a{};
            a.P = dev_params(c);
            a.opsA = c->d_opsA; a.opsF = c->d_opsF; a.tab9 = c->d_tab9; a.tab_words = (int)tile2_table_words(c->K);
            a.row_bytes = tile2_row_bytes(c->K); a.entry_bytes = tile2_entry_bytes(c->K);
            a.hdr = c->d_hdr; a.ops = c->d_ops; a.aux = c->d_aux; a.col2op = c->d_col2op; a.prog = c->d_tprog;
            a.n_reads = c->n_reads; a.row_begin = row_begin; a.col0 = c->d_col0; a.word_off = c->d_word_off;
            a.colidx = c->d_colidx;
            a.adj = c->d_adj; a.stats = c->d_stats; a.amb = c->d_amb; a.amb_cap = (int)std::min<size_t>(c->amb_cap, (size_t)INT32_MAX);
            a.detail = detail ? c->d_detail : nullptr;
            a.maxw = use_tile2 ? g2.maxw : g.maxw; a.abuf_bytes = use_tile2 ? g2.abuf_bytes : g.abuf_bytes;
            // windows alternate between two streams: a CTA of window k+1 starts on an SM as soon as window k's CTA there
            // has run out of work items, so the ragged end of one launch is filled by the next
            const bool two = c->wins.size() > 1;
            if (two) {
                CU_TRY(c, cudaEventRecord(c->ev_go, c->stream));
                CU_TRY(c, cudaStreamWaitEvent(c->stream2, c->ev_go, 0));
            }
            for (size_t k = 0; k < c->wins.size(); ++k) {
                const hc_ctx::Win& w = c->wins[k];
                cudaStream_t ks = (k & 1) ? c->stream2 : c->stream;
                if (w.n_items > 0) {
                    a.items = c->d_items + w.item0; a.n_items = (int)w.n_items; a.item_counter = c->d_item_counter + k;
                    const int blocks = (int)std::min<size_t>(use_stream ? (w.n_items + tile::kWarps - 1) / tile::kWarps : w.n_items, (size_t)n_ctas);
                    if (use_tile2 && split && map) tile2::score_tile2_kernel<true, true><<<blocks, threads, smem_bytes, ks>>>(a);
                    else if (use_tile2 && split) tile2::score_tile2_kernel<true, false><<<blocks, threads, smem_bytes, ks>>>(a);
                    else if (use_tile2 && map) tile2::score_tile2_kernel<false, true><<<blocks, threads, smem_bytes, ks>>>(a);
                    else if (use_tile2) tile2::score_tile2_kernel<false, false><<<blocks, threads, smem_bytes, ks>>>(a);
                    else tile1_kernel(use_stream, map, g.nwarps == tile::kWarps)<<<blocks, threads, smem_bytes, ks>>>(static_cast<const tile::Args&>(a));
                    CU_TRY(c, cudaGetLastError());
                    launches++;
                    c->last_kernel = use_tile2 ? HC_KERNEL_TILED2 : (use_stream ? HC_KERNEL_STREAMED : HC_KERNEL_TILED);
                }
                CU_TRY(c, cudaEventRecord(c->win_ev[k], ks));
            }
            if (two) {
                CU_TRY(c, cudaEventRecord(c->ev_join, c->stream2));
                CU_TRY(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
            }
            CU_TRY(c, cudaEventRecord(c->ev1, c->stream));
            // the words of a finished window go to the host while the later windows are scored
            if (adj)
                for (size_t k = 0; k < c->wins.size(); ++k) {
                    const hc_ctx::Win& w = c->wins[k];
                    if (w.n_words <= 0) continue;
                    CU_TRY(c, cudaStreamWaitEvent(c->copy_stream, c->win_ev[k], 0));
                    CU_TRY(c, cudaMemcpyAsync(adj + w.word0, c->d_adj + w.word0, (size_t)w.n_words * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->copy_stream));
                }
        } else {
            CU_TRY(c, cudaEventRecord(c->ev0, c->stream));
            if (total > 0) {
                const int threads = 256;
                const long long want = (total + (threads / 32) - 1) / (threads / 32);
                const int blocks = (int)std::min<long long>(want, (long long)c->sm_count * 8);
                score_band_kernel<<<blocks, threads, 0, c->stream>>>(dev_params(c), c->d_hdr, c->d_ops, row_begin, nrows, c->d_col0,
                                                                     c->d_word_off, total, c->d_adj, c->d_stats, c->d_amb,
                                                                     (int)std::min<size_t>(c->amb_cap, (size_t)INT32_MAX));
                CU_TRY(c, cudaGetLastError());
                launches = 1;
                c->last_kernel = HC_KERNEL_GENERIC;
            }
            CU_TRY(c, cudaEventRecord(c->ev1, c->stream));
            if (adj && total > 0) CU_TRY(c, cudaMemcpyAsync(adj, c->d_adj, (size_t)total * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
        }
        CU_TRY(c, cudaMemcpyAsync(&hs, c->d_stats, sizeof(hs), cudaMemcpyDeviceToHost, c->stream));
#ifdef HC_T2_PROFILE
        {
            BandStatsDev pr{};
            CU_TRY(c, cudaMemcpy(&pr, c->d_stats + 1, sizeof(pr), cudaMemcpyDeviceToHost));
            std::fprintf(stderr, "[HC_T2_PROFILE] 8-column iterations: first pass %llu (%.2f lanes), later passes %llu (%.2f lanes)\n",
                         pr.pairs_total, pr.pairs_total ? (double)pr.pairs_scored / pr.pairs_total : 0.0, pr.pairs_walked,
                         pr.pairs_walked ? (double)pr.edges / pr.pairs_walked : 0.0);
        }
#endif
        if (detail && total > 0) CU_TRY(c, cudaMemcpyAsync(detail, c->d_detail, (size_t)total * 32 * sizeof(hc_pair_result), cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(c, cudaStreamSynchronize(c->stream));
        CU_TRY(c, cudaStreamSynchronize(c->copy_stream));
        if (hs.ambiguous <= (unsigned long long)c->amb_cap || attempt > 0) break;
        // more near-threshold pairs than the list holds (a wide tolerance): a larger list, and the pass again
        if ((rc = grow(c, c->d_amb, c->amb_cap, (size_t)hs.ambiguous)) != HC_OK) return rc;
    }
    long long edge_delta = 0;
    c->amb_host.resize((size_t)std::min<unsigned long long>(hs.ambiguous, (unsigned long long)c->amb_cap));
    if (!c->amb_host.empty()) {
        CU_TRY(c, cudaMemcpy(c->amb_host.data(), c->d_amb, c->amb_host.size() * sizeof(hc_ambiguous), cudaMemcpyDeviceToHost));
        // the device lists them in completion order: sort by (i, j) so that the report does not depend on the schedule
        std::sort(c->amb_host.begin(), c->amb_host.end(), [](const hc_ambiguous& x, const hc_ambiguous& y) { return x.i != y.i ? x.i < y.i : x.j < y.j; });
        if ((rc = redecide_ambiguous(c, row_begin, adj, &edge_delta)) != HC_OK) return rc;
        if (amb) std::memcpy(amb, c->amb_host.data(), std::min<size_t>(c->amb_host.size(), (size_t)amb_cap) * sizeof(hc_ambiguous));
    }
    if (stats) {
        stats->pairs_total = (int64_t)hs.pairs_total; stats->pairs_scored = (int64_t)hs.pairs_scored;
        stats->pairs_walked = (int64_t)hs.pairs_walked; stats->edges = (int64_t)hs.edges + edge_delta;
        stats->sum_cc = (int64_t)hs.sum_cc; stats->sum_tc = (int64_t)hs.sum_tc; stats->undefined = (int64_t)hs.undefined;
        stats->ambiguous = (int64_t)hs.ambiguous;
        float ms = 0.f;
        CU_TRY(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        stats->kernel_ms = ms;
        stats->launches = launches;
    }
    return HC_OK;
}

int hc_score_band(hc_ctx* c, int32_t row_begin, int32_t row_end, uint32_t* adj, int64_t n_words, hc_ambiguous* amb,
                  int32_t amb_cap, hc_band_stats* stats) {
    return score_band_impl(c, row_begin, row_end, adj, n_words, amb, amb_cap, stats, nullptr);
}

int hc_score_band_detail(hc_ctx* c, int32_t row_begin, int32_t row_end, uint32_t* adj, hc_pair_result* detail, int64_t n_words,
                         hc_band_stats* stats) {
    if (!detail) { if (c) c->err = "hc_score_band_detail: null detail buffer"; return HC_E_INVALID; }
    return score_band_impl(c, row_begin, row_end, adj, n_words, nullptr, 0, stats, detail);
}

// ---- QuasispeciesEdgeCalculator path (hc_quasi.cuh) ----------------------------------------------------
static int quasi_prepare(hc_ctx* c, const hc_quasi_params* qp, quasi::QParams& P) {
    if (!c->d_hdr) { c->err = "hc_quasi_*: no batch uploaded"; return HC_E_STATE; }
    if (!(qp->edge_quasi_cutoff_cliques > 0 && qp->edge_quasi_cutoff_mixed > 0 && qp->edge_quasi_cutoff_single > 0)) {
        c->err = "hc_quasi_*: cutoffs must be > 0";
        return HC_E_INVALID;
    }
    CU_TRY(c, cudaSetDevice(c->device));
    const int K = c->K;
    if (!c->d_logsum) {
        // log(sum) of src/QuasispeciesEdgeCalculator.cpp:307-324 for every (base1, base2, quality1, quality2), the
        // reference's expression, summation order and libm
        std::vector<double> qc(K), anti(K);
        for (int a = 0; a < K; ++a) {
            const char qual = (char)c->qual_chars[a];
            qc[a] = 1 - std::pow(10, -(qual - 33) / 10.0);                 // ShortDnaSequence::qualityCorrect (:117-119)
            anti[a] = (1.0 - qc[a]) / 3.0;
        }
        std::vector<double> tab((size_t)25 * K * K);
        for (int b1 = 0; b1 < 5; ++b1)
            for (int b2 = 0; b2 < 5; ++b2)
                for (int a = 0; a < K; ++a)
                    for (int b = 0; b < K; ++b) {
                        double sum = 0.0;
                        for (int x = 0; x < 4; ++x) sum += ((b1 == x ? qc[a] : anti[a]) * (b2 == x ? qc[b] : anti[b]));
                        tab[(((size_t)b1 * 5 + b2) * K + a) * K + b] = std::log(sum);
                    }
        CU_TRY(c, cudaMalloc((void**)&c->d_logsum, tab.size() * sizeof(double)));
        CU_TRY(c, cudaMemcpy(c->d_logsum, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    if (!c->simpson_host.empty() && !c->d_simpson_ln) {
        std::vector<double> ln(c->simpson_host.size(), 0.0);
        for (size_t k = 0; k < ln.size(); ++k) if (c->simpson_host[k] != 0.0) ln[k] = std::log(c->simpson_host[k]);
        CU_TRY(c, cudaMalloc((void**)&c->d_simpson_ln, ln.size() * sizeof(double)));
        CU_TRY(c, cudaMemcpy(c->d_simpson_ln, ln.data(), ln.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    const bool no_fast = std::getenv("HC_QUASI_NO_FAST") != nullptr;             // tests: force the general walk
    if (!c->have_qseg && !no_fast) {
        int rc;
        if ((rc = grow(c, c->d_qseg, c->qseg_cap, (size_t)2 * c->n_reads)) != HC_OK) return rc;
        quasi::build_qseg_kernel<<<(2 * c->n_reads + 127) / 128, 128, 0, c->stream>>>(c->d_hdr, c->d_ops, c->n_reads, c->d_qseg);
        CU_TRY(c, cudaGetLastError());
        c->have_qseg = true;
    }
    P.qseg = no_fast ? nullptr : c->d_qseg;
    P.cutoff[0] = qp->edge_quasi_cutoff_cliques; P.cutoff[1] = qp->edge_quasi_cutoff_mixed; P.cutoff[2] = qp->edge_quasi_cutoff_single;
    P.min_overlap[0] = qp->min_overlap_cliques; P.min_overlap[1] = qp->min_overlap_single;
    P.amb_tol = qp->ambiguity_rel_tol > 0 ? qp->ambiguity_rel_tol : 1e-12;
    P.logsum = c->d_logsum;
    P.simpson_ln = c->d_simpson_ln;
    P.simpson_len = (int)c->simpson_host.size();
    P.K = K;
    P.frameshift = qp->frameshift_merge;
    return HC_OK;
}

int hc_quasi_score_pairs(hc_ctx* c, const hc_quasi_params* qp, const int32_t* pi, const int32_t* pj, int64_t npairs,
                         hc_quasi_result* out) {
    if (!c || !qp || npairs < 0 || (npairs > 0 && (!pi || !pj || !out))) return HC_E_INVALID;
    quasi::QParams P{};
    int rc = quasi_prepare(c, qp, P);
    if (rc != HC_OK) return rc;
    if (npairs == 0) return HC_OK;
    for (int64_t k = 0; k < npairs; ++k)
        if (pi[k] < 0 || pi[k] >= c->n_reads || pj[k] < 0 || pj[k] >= c->n_reads) { c->err = "hc_quasi_score_pairs: read index out of range"; return HC_E_INVALID; }
    if ((rc = grow(c, c->d_pi, c->pi_cap, (size_t)npairs)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_pj, c->pj_cap, (size_t)npairs)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_qres, c->qres_cap, (size_t)npairs)) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_pi, pi, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_pj, pj, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, c->stream));
    const int blocks = (int)std::min<int64_t>((npairs + 127) / 128, (int64_t)c->sm_count * 16);
    quasi::pairs_kernel<<<blocks, 128, 0, c->stream>>>(P, c->d_hdr, c->d_ops, c->d_pi, c->d_pj, npairs, c->d_qres);
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaMemcpyAsync(out, c->d_qres, sizeof(hc_quasi_result) * npairs, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    return HC_OK;
}

int hc_quasi_score_band(hc_ctx* c, const hc_quasi_params* qp, int32_t row_begin, int32_t row_end, uint32_t* adj, int64_t n_words,
                        hc_band_stats* stats) {
    if (!c || !qp) return HC_E_INVALID;
    quasi::QParams P{};
    int rc = quasi_prepare(c, qp, P);
    if (rc != HC_OK) return rc;
    if (row_begin < 0 || row_end < row_begin || row_end > c->n_reads) { c->err = "hc_quasi_score_band: bad row range"; return HC_E_INVALID; }
    compute_layout(c, row_begin, row_end);
    const int nrows = row_end - row_begin;
    const long long total = c->lay_off[nrows];
    if (adj && n_words < total) { c->err = "hc_quasi_score_band: output buffer smaller than hc_band_layout's n_words"; return HC_E_INVALID; }
    if ((rc = grow(c, c->d_col0, c->col0_cap, (size_t)nrows + 1)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_word_off, c->word_off_cap, (size_t)nrows + 1)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_adj, c->adj_cap, (size_t)std::max<long long>(total, 1))) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_col0, c->lay_col0.data(), sizeof(int32_t) * nrows, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_word_off, c->lay_off.data(), sizeof(long long) * ((size_t)nrows + 1), cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemsetAsync(c->d_stats, 0, sizeof(BandStatsDev), c->stream));
    CU_TRY(c, cudaEventRecord(c->ev0, c->stream));
    int launches = 0;
    if (total > 0) {
        const long long want = (total + 3) / 4;
        const int blocks = (int)std::min<long long>(want, (long long)c->sm_count * 16);
        quasi::band_kernel<<<blocks, 128, 0, c->stream>>>(P, c->d_hdr, c->d_ops, row_begin, nrows, c->d_col0, c->d_word_off, total,
                                                          c->d_adj, c->d_stats);
        CU_TRY(c, cudaGetLastError());
        launches = 1;
    }
    CU_TRY(c, cudaEventRecord(c->ev1, c->stream));
    BandStatsDev hs{};
    CU_TRY(c, cudaMemcpyAsync(&hs, c->d_stats, sizeof(hs), cudaMemcpyDeviceToHost, c->stream));
    if (adj && total > 0) CU_TRY(c, cudaMemcpyAsync(adj, c->d_adj, (size_t)total * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->pairs_total = (int64_t)hs.pairs_total; stats->pairs_scored = (int64_t)hs.pairs_scored;
        stats->pairs_walked = (int64_t)hs.pairs_walked; stats->edges = (int64_t)hs.edges;
        stats->undefined = (int64_t)hs.undefined; stats->ambiguous = (int64_t)hs.ambiguous;
        float ms = 0.f;
        CU_TRY(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        stats->kernel_ms = ms;
        stats->launches = launches;
    }
    return HC_OK;
}

int hc_set_kernel(hc_ctx* c, int32_t mode) {
    if (!c || mode < HC_KERNEL_AUTO || mode > HC_KERNEL_STREAMED) return HC_E_INVALID;
    c->kernel_mode = mode;
    return HC_OK;
}

int hc_last_kernel(const hc_ctx* c) { return c ? c->last_kernel : 0; }

int hc_score_pairs(hc_ctx* c, const int32_t* pi, const int32_t* pj, int64_t npairs, hc_pair_result* out) {
    if (!c || npairs < 0 || (npairs > 0 && (!pi || !pj || !out))) return HC_E_INVALID;
    if (!c->d_hdr) { c->err = "hc_score_pairs: no batch uploaded"; return HC_E_STATE; }
    for (int64_t k = 0; k < npairs; ++k)
        if (pi[k] < 0 || pi[k] >= c->n_reads || pj[k] < 0 || pj[k] >= c->n_reads) { c->err = "hc_score_pairs: read index out of range"; return HC_E_INVALID; }
    if (npairs == 0) return HC_OK;
    CU_TRY(c, cudaSetDevice(c->device));
    int rc;
    if ((rc = grow(c, c->d_pi, c->pi_cap, (size_t)npairs)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_pj, c->pj_cap, (size_t)npairs)) != HC_OK) return rc;
    if ((rc = grow(c, c->d_res, c->res_cap, (size_t)npairs)) != HC_OK) return rc;
    CU_TRY(c, cudaMemcpyAsync(c->d_pi, pi, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_pj, pj, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, c->stream));
    const int threads = 128;
    const int blocks = (int)std::min<long long>((npairs + threads - 1) / threads, (long long)c->sm_count * 16);
    score_pairs_kernel<<<blocks, threads, 0, c->stream>>>(dev_params(c), c->d_hdr, c->d_ops, c->d_pi, c->d_pj, npairs, c->d_res);
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaMemcpyAsync(out, c->d_res, sizeof(hc_pair_result) * npairs, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    // near-threshold pairs: the decision is the host's (glibc log10), as in hc_score_band
    const DevParams P = dev_params(c);
    for (int64_t k = 0; k < npairs; ++k) {
        if (!out[k].ambiguous) continue;
        const bool cl1 = c->ncov_flags[pi[k]] & HC_FLAG_IS_CLIQUE, cl2 = c->ncov_flags[pj[k]] & HC_FLAG_IS_CLIQUE;
        const double cutoff = P.log_cutoff[(cl1 && cl2) ? 0 : ((cl1 || cl2) ? 1 : 2)];
        out[k].edge = host_decide(c, out[k].probM, out[k].prob0, out[k].cc, out[k].tc, cutoff, &out[k].score) ? 1 : 0;
    }
    return HC_OK;
}

}  // extern "C"

// ---- hc_score_pairs without any allocation --------------------------------------------------------------------------
// The pair kernel's device-side scratch (the pair lists it reads and the result records it writes, which hc_score_pairs
// keeps in grow-only cudaMalloc buffers) in a file-scope array: HC_STATIC_MAX_PAIRS * (48 + 4 + 4) bytes = 896 KiB, held by
// every GPU that loads the library until the process ends. In host code `g_pairs_scratch` is only a stand-in for the
// symbol; the device address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_pairs_scratch[(size_t)HC_STATIC_MAX_PAIRS * (sizeof(hc_pair_result) + 2 * sizeof(int32_t))];

namespace {
constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_addr[kMaxStaticDevices] = {};

// Device address of g_pairs_scratch on c->device (the current device on entry).
int static_scratch(hc_ctx* c, unsigned char** out) {
    if (c->device < 0 || c->device >= kMaxStaticDevices) { c->err = "hc_score_pairs_static: device ordinal beyond 63"; return HC_E_INVALID; }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_addr[c->device]) {
        void* ptr = nullptr;
        CU_TRY(c, cudaGetSymbolAddress(&ptr, g_pairs_scratch));
        if (!ptr) { c->err = "hc_score_pairs_static: cudaGetSymbolAddress gave no address"; return HC_E_CUDA; }
        g_static_addr[c->device] = static_cast<unsigned char*>(ptr);
    }
    *out = g_static_addr[c->device];
    return HC_OK;
}
}  // namespace

extern "C" int hc_score_pairs_static(hc_ctx* c, const int32_t* pi, const int32_t* pj, int64_t npairs, hc_pair_result* out) {
    if (!c || npairs < 0 || (npairs > 0 && (!pi || !pj || !out))) return HC_E_INVALID;
    if (!c->d_hdr) { c->err = "hc_score_pairs_static: no batch uploaded"; return HC_E_STATE; }
    if (npairs > HC_STATIC_MAX_PAIRS) {
        c->err = "hc_score_pairs_static: " + std::to_string((long long)npairs) + " pairs in one call, the static scratch holds " +
                 std::to_string((int)HC_STATIC_MAX_PAIRS) + " (HC_STATIC_MAX_PAIRS); split the list or use hc_score_pairs";
        return HC_E_INVALID;
    }
    for (int64_t k = 0; k < npairs; ++k)
        if (pi[k] < 0 || pi[k] >= c->n_reads || pj[k] < 0 || pj[k] >= c->n_reads) { c->err = "hc_score_pairs_static: read index out of range"; return HC_E_INVALID; }
    if (npairs == 0) return HC_OK;
    CU_TRY(c, cudaSetDevice(c->device));
    unsigned char* base = nullptr;
    const int rc = static_scratch(c, &base);
    if (rc != HC_OK) return rc;
    hc_pair_result* d_res = reinterpret_cast<hc_pair_result*>(base);
    int32_t* d_pi = reinterpret_cast<int32_t*>(base + (size_t)HC_STATIC_MAX_PAIRS * sizeof(hc_pair_result));
    int32_t* d_pj = d_pi + HC_STATIC_MAX_PAIRS;
    // the array keeps the last call's records: the slots of this call start from zero
    CU_TRY(c, cudaMemsetAsync(d_res, 0, sizeof(hc_pair_result) * npairs, c->stream));
    CU_TRY(c, cudaMemcpyAsync(d_pi, pi, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(d_pj, pj, sizeof(int32_t) * npairs, cudaMemcpyHostToDevice, c->stream));
    const int threads = 128;
    const int blocks = (int)std::min<long long>((npairs + threads - 1) / threads, (long long)c->sm_count * 16);
    score_pairs_kernel<<<blocks, threads, 0, c->stream>>>(dev_params(c), c->d_hdr, c->d_ops, d_pi, d_pj, npairs, d_res);
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaMemcpyAsync(out, d_res, sizeof(hc_pair_result) * npairs, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    const DevParams P = dev_params(c);                   // near-threshold pairs: the host's decision, as in hc_score_pairs
    for (int64_t k = 0; k < npairs; ++k) {
        if (!out[k].ambiguous) continue;
        const bool cl1 = c->ncov_flags[pi[k]] & HC_FLAG_IS_CLIQUE, cl2 = c->ncov_flags[pj[k]] & HC_FLAG_IS_CLIQUE;
        const double cutoff = P.log_cutoff[(cl1 && cl2) ? 0 : ((cl1 || cl2) ? 1 : 2)];
        out[k].edge = host_decide(c, out[k].probM, out[k].prob0, out[k].cc, out[k].tc, cutoff, &out[k].score) ? 1 : 0;
    }
    return HC_OK;
}

// ---- several GPUs behind one handle ------------------------------------------------------------------------------
struct hc_multi {
    std::vector<hc_ctx*> ctx;
    std::vector<hc_packed*> part;              // slices of the resident batch (headers re-based), one per device
    std::vector<int32_t> cuts, first;
    std::vector<int32_t> col0; std::vector<int64_t> word_off;      // band layout of the whole batch
    int n_reads = 0;
    bool resident = false;
    std::string err;
};

namespace {
// f(k) for every device k on its own host thread (CUDA's current device is per thread); returns the first failure.
template <class F>
int for_each_device(hc_multi* m, F f) {
    const int n = (int)m->ctx.size();
    std::vector<int> rc((size_t)n, HC_OK);
    if (n == 1) rc[0] = f(0);
    else {
        std::vector<std::thread> pool;
        for (int k = 0; k < n; ++k) pool.emplace_back([&, k] { rc[k] = f(k); });
        for (auto& t : pool) t.join();
    }
    for (int k = 0; k < n; ++k)
        if (rc[k] != HC_OK) { m->err = "device " + std::to_string(k) + ": " + m->ctx[k]->err; return rc[k]; }
    return HC_OK;
}
void free_parts(hc_multi* m) {
    for (hc_packed* q : m->part) hc_packed_free(q);
    m->part.clear();
}
}  // namespace

extern "C" {

int hc_multi_create(const hc_params* params, const int32_t* devices, int32_t n_devices, hc_multi** out) {
    if (!params || !devices || !out || n_devices < 1 || n_devices > 64) { hc::set_global_error("hc_multi_create: bad argument"); return HC_E_INVALID; }
    *out = nullptr;
    hc_multi* m = new hc_multi();
    for (int k = 0; k < n_devices; ++k) {
        hc_params p = *params;
        p.device = devices[k];
        hc_ctx* c = nullptr;
        const int rc = hc_create(&p, &c);
        if (rc != HC_OK) { hc_multi_destroy(m); return rc; }
        m->ctx.push_back(c);
    }
    *out = m;
    return HC_OK;
}

void hc_multi_destroy(hc_multi* m) {
    if (!m) return;
    free_parts(m);
    for (hc_ctx* c : m->ctx) hc_destroy(c);
    delete m;
}

const char* hc_multi_last_error(const hc_multi* m) { return m ? m->err.c_str() : hc::global_error(); }
int hc_multi_n_devices(const hc_multi* m) { return m ? (int)m->ctx.size() : 0; }
hc_ctx* hc_multi_context(hc_multi* m, int32_t k) { return (m && k >= 0 && k < (int)m->ctx.size()) ? m->ctx[k] : nullptr; }

int hc_multi_set_params(hc_multi* m, const hc_params* params) {
    if (!m || !params) return HC_E_INVALID;
    for (hc_ctx* c : m->ctx) { const int rc = hc_set_params(c, params); if (rc != HC_OK) return rc; }
    return HC_OK;
}

int hc_multi_upload(hc_multi* m, const hc_packed* p) {
    if (!m || !p) return HC_E_INVALID;
    const int n = p->n_reads, D = (int)m->ctx.size();
    m->resident = false;
    free_parts(m);
    std::vector<int32_t> cf((size_t)std::max(n, 1)), cl((size_t)std::max(n, 1));
    for (int r = 0; r < n; ++r) { cf[r] = p->hdr[r].cov_first; cl[r] = p->hdr[r].cov_last; }
    m->cuts.assign((size_t)D + 1, 0);
    m->first.assign((size_t)D, 0);
    int rc = hc_partition_rows(cf.data(), cl.data(), n, D, m->cuts.data(), m->first.data());
    if (rc != HC_OK) { m->err = hc::global_error(); return rc; }
    m->part.assign((size_t)D, nullptr);
    for (int k = 0; k < D; ++k)
        if ((rc = hc_packed_slice(p, m->first[k], m->cuts[k + 1], &m->part[k])) != HC_OK) { m->err = hc::global_error(); return rc; }
    if ((rc = for_each_device(m, [&](int k) { return hc_upload(m->ctx[k], m->part[k]); })) != HC_OK) return rc;
    // the band layout of the whole batch = the parts' layouts shifted by first[k] (hc_partition_rows), concatenated
    m->col0.assign((size_t)std::max(n, 1), 0);
    m->word_off.assign((size_t)n + 1, 0);
    int64_t acc = 0;
    for (int k = 0; k < D; ++k) {
        hc_ctx* c = m->ctx[k];
        const int lo = m->cuts[k] - m->first[k], hi = m->cuts[k + 1] - m->first[k];
        compute_layout(c, lo, hi);
        for (int i = lo; i < hi; ++i) {
            m->col0[(size_t)m->first[k] + i] = c->lay_col0[i - lo] + m->first[k];
            m->word_off[(size_t)m->first[k] + i] = acc + c->lay_off[i - lo];
        }
        acc += c->lay_off[hi - lo];
    }
    m->word_off[n] = acc;
    m->n_reads = n;
    m->resident = true;
    return HC_OK;
}

int hc_multi_partition(const hc_multi* m, int32_t* cuts, int32_t* first) {
    if (!m || !m->resident) return HC_E_STATE;
    if (cuts) std::memcpy(cuts, m->cuts.data(), m->cuts.size() * sizeof(int32_t));
    if (first) std::memcpy(first, m->first.data(), m->first.size() * sizeof(int32_t));
    return HC_OK;
}

int hc_multi_band_layout(hc_multi* m, int32_t* col0, int64_t* word_off, int64_t* n_words) {
    if (!m) return HC_E_INVALID;
    if (!m->resident) { m->err = "hc_multi_band_layout: no batch uploaded"; return HC_E_STATE; }
    if (col0) std::memcpy(col0, m->col0.data(), sizeof(int32_t) * (size_t)m->n_reads);
    if (word_off) std::memcpy(word_off, m->word_off.data(), sizeof(int64_t) * ((size_t)m->n_reads + 1));
    if (n_words) *n_words = m->word_off[m->n_reads];
    return HC_OK;
}

int hc_multi_score_band(hc_multi* m, uint32_t* adj, int64_t n_words, hc_ambiguous* amb, int32_t amb_cap, hc_band_stats* stats,
                        hc_band_stats* per_device) {
    if (!m) return HC_E_INVALID;
    if (!m->resident) { m->err = "hc_multi_score_band: no batch uploaded"; return HC_E_STATE; }
    if (adj && n_words < m->word_off[m->n_reads]) { m->err = "hc_multi_score_band: output buffer smaller than hc_multi_band_layout's n_words"; return HC_E_INVALID; }
    if (amb_cap < 0 || (amb_cap > 0 && !amb)) { m->err = "hc_multi_score_band: bad near-threshold list"; return HC_E_INVALID; }
    const int D = (int)m->ctx.size();
    std::vector<hc_band_stats> st((size_t)D);
    int rc = for_each_device(m, [&](int k) {
        std::memset(&st[k], 0, sizeof(hc_band_stats));
        const int lo = m->cuts[k] - m->first[k], hi = m->cuts[k + 1] - m->first[k];
        const int64_t w0 = m->word_off[m->cuts[k]], w1 = m->word_off[m->cuts[k + 1]];
        return score_band_impl(m->ctx[k], lo, hi, adj ? adj + w0 : nullptr, w1 - w0, nullptr, 0, &st[k], nullptr);
    });
    if (rc != HC_OK) return rc;
    hc_band_stats tot;
    std::memset(&tot, 0, sizeof(tot));
    int32_t n_amb = 0;
    for (int k = 0; k < D; ++k) {
        tot.pairs_total += st[k].pairs_total; tot.pairs_scored += st[k].pairs_scored; tot.pairs_walked += st[k].pairs_walked;
        tot.edges += st[k].edges; tot.sum_cc += st[k].sum_cc; tot.sum_tc += st[k].sum_tc; tot.undefined += st[k].undefined;
        tot.ambiguous += st[k].ambiguous; tot.kernel_ms = std::max(tot.kernel_ms, st[k].kernel_ms); tot.launches += st[k].launches;
        for (const hc_ambiguous& a : m->ctx[k]->amb_host) {             // the parts' reports, in global read indices
            if (n_amb >= amb_cap) break;
            amb[n_amb] = a;
            amb[n_amb].i += m->first[k];
            amb[n_amb].j += m->first[k];
            n_amb++;
        }
    }
    if (stats) *stats = tot;
    if (per_device) std::memcpy(per_device, st.data(), sizeof(hc_band_stats) * (size_t)D);
    return HC_OK;
}

}  // extern "C"
