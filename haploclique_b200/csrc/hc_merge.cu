// Super-read construction on the device: AlignmentRecord(alignments, clique_id)
// (src/AlignmentRecord.cpp:211-249) for whole batches of cliques, any mix of single-end and paired-end members.
//
// What it replaces: the left fold of the clique's members through mergeAlignmentRecordsSingle(ar,i,j) (:581-734),
// mergeAlignmentRecordsPaired (:734-1293) and mergeAlignmentRecordsMixed (:1295-1537) -- lock-step walks of unrolled CIGARs
// (noOverlapMerge :1539-1576, overlapMerge :1578-1674) whose shared M / I columns take the consensus base and the updated
// Phred char from computeEntry (:169-184, tables :40-90).
//
// Formulation here. A warp owns a clique. The running super-read (two segments) lives in the warp's slice of shared memory
// across all folds (global scratch only for super-reads longer than the slice), members are read in place, and only the
// final record is written out -- cooperatively, 32 lanes wide. A fold is a row of a layout table: a phase program
// (tail / overlap / seat / finish steps with symbolic bounds: the reference's hand-unrolled while loops) plus a result
// rule (which segment receives the merged string, how start / end combine, what is taken over from the member).
//   * Plain folds -- every segment taking part is "clips, M only, clips" and the layout has no cursor quirk -- are
//     evaluated in reference-column space by all 32 lanes: a phase covers a column interval known in O(1), a column's
//     output slot follows from interval arithmetic (only M columns emit), so lanes work on 32 consecutive columns with
//     no scan. This is the common case (reads without indels).
//   * Every other fold (indels, the reference's cursor slip in the sixth paired layout, cursors that run off a CIGAR) is
//     walked by one lane exactly as the reference walks it.
// Results are compacted on the device; the host receives per clique the record header and the exact bytes.
// Undefined inputs (an op pair overlapMerge does not list; a cursor past its CIGAR or its sequence: the Release build of
// the reference never returns / throws there) end that clique with HC_MERGE_UNDEFINED / HC_MERGE_OVERRUN.
// There is no CPU merging path: without a CUDA device every call fails with HC_E_CUDA.
#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "hc_internal.h"

namespace {

constexpr int NQ = 94;            // quality chars 33..126
constexpr int kWarps = 24;        // warps (= cliques in flight) per CTA
constexpr int kSlice = 6 * 1024;  // shared memory per warp: 4 segment buffers (2 segments x 2 generations) of 3 byte streams
constexpr int kSegCap = kSlice / 12 / 16 * 16;   // ops a buffered segment may hold (512): longer running reads use global scratch
constexpr int kTabBytes = (2 * NQ * NQ + 255) / 256 * 256;

// ---- layout table: bound ids ext*8 + side*4 + seg*2 + is_end; side 0 = the running record ("this"), 1 = the member ("ar")
#define MB(ext, side, seg, is_end) ((ext) * 8 + (side) * 4 + (seg) * 2 + (is_end))
enum : uint8_t {
    As1 = MB(0, 0, 0, 0), Ae1 = MB(0, 0, 0, 1), As2 = MB(0, 0, 1, 0), Ae2 = MB(0, 0, 1, 1),
    Bs1 = MB(0, 1, 0, 0), Be1 = MB(0, 1, 0, 1), Bs2 = MB(0, 1, 1, 0), Be2 = MB(0, 1, 1, 1),
    xAs1 = MB(1, 0, 0, 0), xAe1 = MB(1, 0, 0, 1), xAs2 = MB(1, 0, 1, 0), xAe2 = MB(1, 0, 1, 1),
    xBs1 = MB(1, 1, 0, 0), xBe1 = MB(1, 1, 0, 1), xBs2 = MB(1, 1, 1, 0), xBe2 = MB(1, 1, 1, 1)
};
enum : uint8_t { P_END = 0, P_START, P_START_MIN, P_TLT, P_OVL, P_OVL2, P_OVLQX, P_SEAT, P_FIN };
struct Phase { uint8_t op, x, y, a, b; };
// where the merged string goes (segment 0/1), start = min(v[sa], v[sb]), end = max(v[ea], v[eb]), single_end afterwards
// (2 = leave the flag), take: 0 nothing, 1 segment 2 := ar.segment2 afterwards, 2 segment 1 := ar.segment1 beforehand
struct Result { uint8_t seg, sa, sb, ea, eb, single, take; };
struct Layout { Phase ph[11]; Result res; };
#define L_START(a) {P_START, 0, 0, a, 0}
#define L_START_MIN(a, b) {P_START_MIN, 0, 0, a, b}
#define L_TLT(sd, sg, a) {P_TLT, sd, sg, a, 0}
#define L_OVL(sa, sb, a) {P_OVL, sa, sb, a, 0}
#define L_OVL2(sa, sb, a, b) {P_OVL2, sa, sb, a, b}
#define L_OVLQX(sa, sb, a, b) {P_OVLQX, sa, sb, a, b}
#define L_SEAT(sd, sg) {P_SEAT, sd, sg, 0, 0}
#define L_FIN(a, b) {P_FIN, 0, 0, a, b}
#define L_END {P_END, 0, 0, 0, 0}
enum { LAY_PS0 = 12, LAY_SP0 = 14, LAY_SINGLE = 16, N_LAYOUTS = 20 };
// rows 0..11: mergeAlignmentRecordsPaired's walked layouts in source order (:773 :819 :868 :907 :946 :989 :1027 :1069 :1118
// :1161 :1204 :1249); 12, 13: this paired x ar single (:1335, :1368); 14, 15: this single x ar paired (:1461, :1496);
// 16..19: mergeAlignmentRecordsSingle(ar, i, j) for (1,1) (1,2) (2,1) (2,2): the four overlap shapes of :637-683 are one
// program -- the earlier extended start leads alone, both to the earlier extended end, the later end trails.
__constant__ Layout c_layout[N_LAYOUTS] = {
    {{L_START(xAs1), L_TLT(0, 0, xBs1), L_OVL(0, 0, Ae1), L_TLT(1, 0, As2), L_SEAT(0, 1), L_OVL(1, 0, Be1), L_TLT(0, 1, Bs2), L_SEAT(1, 1),
      L_OVL2(1, 1, xBe2, xAe2), L_FIN(xBe2, xAe2), L_END}, {0, As1, Bs1, Be2, Ae2, 1, 0}},
    {{L_START(xAs1), L_TLT(0, 0, xBs1), L_OVL(0, 0, Ae1), L_TLT(1, 0, As2), L_SEAT(0, 1), L_OVL2(1, 0, xBe1, xAe2), L_FIN(xBe1, xAe2), L_END},
     {0, As1, Bs1, Be1, Ae2, 0, 1}},
    {{L_START(xBs1), L_TLT(1, 0, xAs2), L_OVL(1, 0, Be1), L_TLT(0, 1, Bs2), L_SEAT(1, 1), L_OVL2(1, 1, xBe2, xAe2), L_FIN(xBe2, xAe2), L_END},
     {1, As2, Bs1, Be2, Ae2, 0, 0}},
    {{L_START(xAs2), L_TLT(0, 1, xBs1), L_OVL(1, 0, Be1), L_TLT(0, 1, Bs2), L_SEAT(1, 1), L_OVL2(1, 1, xAe2, xBe2), L_FIN(xAe2, xBe2), L_END},
     {1, As2, Bs1, Be2, Ae2, 0, 0}},
    {{L_START(xAs1), L_TLT(0, 0, xBs1), L_OVL(0, 0, Be1), L_TLT(0, 0, Bs2), L_SEAT(1, 1), L_OVL(0, 1, Ae1), L_TLT(1, 1, As2), L_SEAT(0, 1),
      L_OVL2(1, 1, xBe2, xAe2), L_FIN(xBe2, xAe2), L_END}, {0, As1, Bs1, Be2, Ae2, 1, 0}},
    {{L_START(xAs1), L_TLT(0, 0, xBs1), L_OVL(0, 0, Be1), L_TLT(0, 0, Bs2), L_SEAT(1, 1), L_OVLQX(0, 1, xAe1, xBe2), L_FIN(xAe1, xBe2), L_END},
     {0, As1, Bs1, Be2, Ae1, 0, 0}},
    {{L_START(xBs1), L_TLT(1, 0, xAs1), L_OVL(0, 0, Ae1), L_TLT(1, 0, As2), L_SEAT(0, 1), L_OVL(1, 0, Be1), L_TLT(0, 1, Bs2), L_SEAT(1, 1),
      L_OVL2(1, 1, xAe2, xBe2), L_FIN(xAe2, xBe2), L_END}, {0, As1, Bs1, Be2, Ae2, 1, 0}},
    {{L_START(xBs1), L_TLT(1, 0, xAs1), L_OVL(0, 0, Ae1), L_TLT(1, 0, As2), L_SEAT(0, 1), L_OVL2(1, 0, xBe1, xAe2), L_FIN(xBe1, xAe2), L_END},
     {0, As1, Bs1, Be1, Ae2, 0, 1}},
    {{L_START(xAs1), L_TLT(0, 0, xBs2), L_OVL(0, 1, Ae1), L_TLT(1, 1, As2), L_SEAT(0, 1), L_OVL2(1, 1, xAe2, xBe2), L_FIN(xAe2, xBe2), L_END},
     {1, Bs1, Bs2, Be2, Ae2, 0, 2}},
    {{L_START(xBs2), L_TLT(1, 1, xAs1), L_OVL(0, 1, Ae1), L_TLT(1, 1, As2), L_SEAT(0, 1), L_OVL2(1, 1, xAe2, xBe2), L_FIN(xAe2, xBe2), L_END},
     {1, Bs1, Bs2, Be2, Ae2, 0, 2}},
    {{L_START(xBs1), L_TLT(1, 0, xAs1), L_OVL(0, 0, Be1), L_TLT(0, 0, Bs2), L_SEAT(1, 1), L_OVL(0, 1, Ae1), L_TLT(1, 1, As2), L_SEAT(0, 1),
      L_OVL2(1, 1, xBe2, xAe2), L_FIN(xBe2, xAe2), L_END}, {0, As1, Bs1, Be2, Ae2, 1, 0}},
    {{L_START(xBs1), L_TLT(1, 0, xAs1), L_OVL(0, 0, Be1), L_TLT(0, 0, Bs2), L_SEAT(1, 1), L_OVL2(0, 1, xAe1, xBe2), L_FIN(xAe1, xBe2), L_END},
     {0, As1, Bs1, Be2, Ae1, 0, 0}},
    {{L_START(xBs1), L_TLT(1, 0, xAs1), L_OVL(0, 0, Ae1), L_TLT(1, 0, As2), L_SEAT(0, 1), L_OVL2(1, 0, xAe2, xBe1), L_FIN(xAe2, xBe1), L_END},
     {0, As1, Bs1, Be1, Ae2, 1, 0}},
    {{L_START(xAs1), L_TLT(0, 0, xBs1), L_OVL(0, 0, Ae1), L_TLT(1, 0, As2), L_SEAT(0, 1), L_OVL2(1, 0, xAe2, xBe1), L_FIN(xAe2, xBe1), L_END},
     {0, As1, Bs1, Be1, Ae2, 1, 0}},
    {{L_START(xAs1), L_TLT(0, 0, xBs1), L_OVL(0, 0, Be1), L_TLT(0, 0, Bs2), L_SEAT(1, 1), L_OVL2(0, 1, xBe2, xAe1), L_FIN(xBe2, xAe1), L_END},
     {0, As1, Bs1, Be2, Ae1, 1, 0}},
    {{L_START(xBs1), L_TLT(1, 0, xAs1), L_OVL(0, 0, Be1), L_TLT(0, 0, Bs2), L_SEAT(1, 1), L_OVL2(0, 1, xBe2, xAe1), L_FIN(xBe2, xAe1), L_END},
     {0, As1, Bs1, Be2, Ae1, 1, 0}},
    // mergeAlignmentRecordsSingle(ar, i, j). Result: (1,1) -> segment 1, single_end = true (:685-697); (1,2) -> segment 2
    // formed from start1 / end1, flag untouched (:698-709); (2,1), (2,2) -> segment 2, single_end = false (:710-729)
    {{L_START_MIN(xAs1, xBs1), L_TLT(0, 0, xBs1), L_TLT(1, 0, xAs1), L_OVL2(0, 0, xAe1, xBe1), L_FIN(xAe1, xBe1), L_END}, {0, As1, Bs1, Ae1, Be1, 1, 0}},
    {{L_START_MIN(xAs1, xBs2), L_TLT(0, 0, xBs2), L_TLT(1, 1, xAs1), L_OVL2(0, 1, xAe1, xBe2), L_FIN(xAe1, xBe2), L_END}, {1, As1, Bs2, Ae1, Be2, 2, 0}},
    {{L_START_MIN(xAs2, xBs1), L_TLT(0, 1, xBs1), L_TLT(1, 0, xAs2), L_OVL2(1, 0, xAe2, xBe1), L_FIN(xAe2, xBe1), L_END}, {1, As2, Bs1, Ae2, Be1, 0, 0}},
    {{L_START_MIN(xAs2, xBs2), L_TLT(0, 1, xBs2), L_TLT(1, 1, xAs2), L_OVL2(1, 1, xAe2, xBe2), L_FIN(xAe2, xBe2), L_END}, {1, As2, Bs2, Ae2, Be2, 0, 0}},
};

struct Seg {
    int start, end, n_ops, n_seq;
    int lead, trail;                  // leading / trailing clip counts (computeOffset / computeRevOffset, :126-156)
    int lead_h;                       // H ops among the leading clips (they consume no query)
    int plain;                        // clips, M only, clips -- and the CIGAR consumes no more query than there is
    const uint8_t *cig, *seq, *qual;
};

struct Rec { int single; Seg s[2]; };

struct OutBuf { uint8_t *cig, *seq, *qual; int n_ops, n_seq, cap; };

struct Job { int lay, after; };       // after: 1 single_end := false, 2 segment 2 := ar.segment2 + false, 3 segment 1 := ar.segment1 + false

__device__ __forceinline__ bool is_clip(uint8_t c) { return c == 'S' || c == 'H'; }

struct Walk {
    const Seg* seg[2][2];
    int c[2][2], q[2][2];
    int v[16];
    int ref;
    OutBuf* o;
};

// The walker's emits are bounded by the clique's capacity (columns spanned + insertions); an undefined walk that would emit
// more stops as an overrun.
__device__ __forceinline__ bool put(OutBuf& o, uint8_t op, uint8_t base, uint8_t qual) {
    if (o.n_ops >= o.cap) return false;
    o.seq[o.n_seq] = base;
    o.qual[o.n_seq++] = qual;
    o.cig[o.n_ops++] = op;
    return true;
}
__device__ __forceinline__ bool put_d(OutBuf& o) {
    if (o.n_ops >= o.cap) return false;
    o.cig[o.n_ops++] = 'D';
    return true;
}

// computeEntry (:169-184) from the byte tables. A char below '!' (the reference's own agreement of Q0 with Q62+ rounds to
// 29..32) indexes rows the reference's table builders never fill: they read as 0 there (zero-initialised storage in the g++
// build pinned under oracle/_ref; formally indeterminate), i.e. the update is '!'.
__device__ __forceinline__ void entry(const uint8_t* tab, uint8_t b1, uint8_t q1, uint8_t b2, uint8_t q2, uint8_t& base, uint8_t& qual) {
    const int k1 = (int)q1 - 33, k2 = (int)q2 - 33;
    const int i1 = max(k1, 0), i2 = max(k2, 0);
    if (b1 == b2) { base = b1; qual = tab[i1 * NQ + i2]; }
    else if (k1 >= k2) { base = b1; qual = tab[NQ * NQ + i1 * NQ + i2]; }
    else { base = b2; qual = tab[NQ * NQ + i2 * NQ + i1]; }
    if ((k1 | k2) < 0) qual = 33;
}

// noOverlapMerge (:1539-1576), one column of (side, sg)
__device__ int tail_step(Walk& w, int side, int sg) {
    const Seg& s = *w.seg[side][sg];
    int& c = w.c[side][sg]; int& q = w.q[side][sg];
    if (c >= s.n_ops) return HC_MERGE_OVERRUN;
    const uint8_t op = s.cig[c];
    if (op == 'M' || op == 'I') {
        if (q >= s.n_seq) return HC_MERGE_OVERRUN;
        if (!put(*w.o, op, s.seq[q], s.qual[q])) return HC_MERGE_OVERRUN;
        ++q; ++c;
        w.ref += (op == 'M');
    } else if (op == 'D') { if (!put_d(*w.o)) return HC_MERGE_OVERRUN; ++w.ref; ++c; }
    else if (op == 'S') { ++w.ref; ++q; ++c; }
    else if (op == 'H') { ++w.ref; ++c; }
    else return HC_MERGE_UNDEFINED;
    return HC_MERGE_OK;
}

// overlapMerge (:1578-1674), one step of this.segment sa against ar.segment sb; qa = the query cursor the call was given
__device__ int ovl_step(Walk& w, int sa, int sb, int& qa, const uint8_t* tab) {
    const Seg& a = *w.seg[0][sa]; const Seg& b = *w.seg[1][sb];
    int& c1 = w.c[0][sa]; int& c2 = w.c[1][sb]; int& q2 = w.q[1][sb];
    OutBuf& o = *w.o;
    if (c1 >= a.n_ops || c2 >= b.n_ops) return HC_MERGE_OVERRUN;
    const uint8_t x = a.cig[c1], y = b.cig[c2];
    const bool xc = is_clip(x), yc = is_clip(y);
    if (x == y && (x == 'M' || x == 'I')) {
        if (qa >= a.n_seq || q2 >= b.n_seq) return HC_MERGE_OVERRUN;
        uint8_t base, qual;
        entry(tab, a.seq[qa], a.qual[qa], b.seq[q2], b.qual[q2], base, qual);
        if (!put(o, x, base, qual)) return HC_MERGE_OVERRUN;
        w.ref += (x == 'M');
        ++qa; ++q2; ++c1; ++c2;
    } else if ((x == 'D' || xc) && (y == 'D' || yc)) {
        // D/D, D/clip, clip/clip: a reference column without a merged base; it stays a deletion if either read has one;
        // an S consumes its read's query base
        if ((x == 'D' || y == 'D') && !put_d(o)) return HC_MERGE_OVERRUN;
        qa += (x == 'S');
        q2 += (y == 'S');
        ++w.ref; ++c1; ++c2;
    } else if (x == 'M' && yc) {
        if (qa >= a.n_seq) return HC_MERGE_OVERRUN;
        if (!put(o, 'M', a.seq[qa], a.qual[qa])) return HC_MERGE_OVERRUN;
        ++qa; q2 += (y == 'S');
        ++w.ref; ++c1; ++c2;
    } else if (xc && y == 'M') {
        if (q2 >= b.n_seq) return HC_MERGE_OVERRUN;
        if (!put(o, 'M', b.seq[q2], b.qual[q2])) return HC_MERGE_OVERRUN;
        ++q2; qa += (x == 'S');
        ++w.ref; ++c1; ++c2;
    } else if (x == 'I' && yc) {
        if (qa >= a.n_seq) return HC_MERGE_OVERRUN;
        if (!put(o, 'I', a.seq[qa], a.qual[qa])) return HC_MERGE_OVERRUN;
        ++c1; ++qa;
    } else if (xc && y == 'I') {
        if (q2 >= b.n_seq) return HC_MERGE_OVERRUN;
        if (!put(o, 'I', b.seq[q2], b.qual[q2])) return HC_MERGE_OVERRUN;
        ++c2; ++q2;
    } else {
        return HC_MERGE_UNDEFINED;
    }
    return HC_MERGE_OK;
}

#define M_STEP(expr) do { const int rc_ = (expr); if (rc_) return rc_; } while (0)

__device__ void set_bounds(Walk& w, const Rec& A, const Rec& B) {
    const Rec* r[2] = {&A, &B};
#pragma unroll
    for (int side = 0; side < 2; ++side)
#pragma unroll
        for (int sg = 0; sg < 2; ++sg) {
            const Seg& s = r[side]->s[sg];
            w.seg[side][sg] = &s;
            w.c[side][sg] = w.q[side][sg] = 0;
            w.v[MB(0, side, sg, 0)] = s.start; w.v[MB(0, side, sg, 1)] = s.end;
            w.v[MB(1, side, sg, 0)] = s.start - s.lead; w.v[MB(1, side, sg, 1)] = s.end + s.trail;
        }
}

// One lane walks a layout exactly as the reference's loops do.
__device__ int walk_layout(Walk& w, const Layout& L, const uint8_t* tab) {
    for (int k = 0; k < 11; ++k) {
        const Phase ph = L.ph[k];
        switch (ph.op) {
            case P_END: return HC_MERGE_OK;
            case P_START: w.ref = w.v[ph.a]; break;
            case P_START_MIN: w.ref = min(w.v[ph.a], w.v[ph.b]); break;
            case P_TLT: while (w.ref < w.v[ph.a]) M_STEP(tail_step(w, ph.x, ph.y)); break;
            case P_SEAT: {                                     // computeSOffset (:136-145)
                const Seg& s = *w.seg[ph.x][ph.y];
                for (int t = 0; t < s.n_ops; ++t) {
                    if (s.cig[t] == 'S') { w.c[ph.x][ph.y]++; w.q[ph.x][ph.y]++; }
                    else if (s.cig[t] == 'H') w.c[ph.x][ph.y]++;
                    else break;
                }
                break;
            }
            case P_OVL: while (w.ref <= w.v[ph.a]) M_STEP(ovl_step(w, ph.x, ph.y, w.q[0][ph.x], tab)); break;
            case P_OVL2: while (w.ref <= w.v[ph.a] && w.ref <= w.v[ph.b]) M_STEP(ovl_step(w, ph.x, ph.y, w.q[0][ph.x], tab)); break;
            case P_OVLQX: while (w.ref <= w.v[ph.a] && w.ref <= w.v[ph.b]) M_STEP(ovl_step(w, ph.x, ph.y, w.q[0][1 - ph.x], tab)); break;
            case P_FIN: {
                int which = -1;
                if (w.ref - 1 == w.v[ph.a]) which = ph.b; else if (w.ref - 1 == w.v[ph.b]) which = ph.a;
                if (which >= 0) { const int sd = (which >> 2) & 1, sg = (which >> 1) & 1; while (w.ref <= w.v[which]) M_STEP(tail_step(w, sd, sg)); }
                break;
            }
        }
    }
    return HC_MERGE_OK;
}

// ---- plain folds in reference-column space, all 32 lanes ------------------------------------------------------------
// For a plain segment the cursors are functions of the column: op index c = p - xs (xs = extended start), the op is a
// clip outside [start, end] and M inside, and the query index is c - lead_h. A phase of the program is a column interval
// [lo, hi); the columns that emit are those where a participating read shows M.
struct PlainPhase { int lo, hi; int a, b; };     // a / b: reader (side*2 + seg), -1 = none

// Number of columns in [lo, x) on which segment s shows M.
__device__ __forceinline__ int m_before(const Seg& s, int lo, int x) { return max(0, min(x, s.end + 1) - max(lo, s.start)); }

// Lists the phases of a layout as column intervals when every cursor stays consistent with the columns (plain segments, no
// quirk). Returns the number of phases, or -1 when the layout has to be walked (a cursor would run off its CIGAR, the slip
// layout, a cursor that does not stand on the running column when its phase begins).
__device__ int plain_phases(const Walk& w, const Layout& L, PlainPhase* out) {
    int n = 0, ref = 0;
    int pos[2][2] = {{INT_MIN, INT_MIN}, {INT_MIN, INT_MIN}};     // column the cursor of (side, seg) stands on; MIN = untouched
    for (int k = 0; k < 11; ++k) {
        const Phase ph = L.ph[k];
        if (ph.op == P_END) break;
        if (ph.op == P_START) { ref = w.v[ph.a]; continue; }
        if (ph.op == P_START_MIN) { ref = min(w.v[ph.a], w.v[ph.b]); continue; }
        if (ph.op == P_SEAT) {
            if (pos[ph.x][ph.y] != INT_MIN) return -1;
            pos[ph.x][ph.y] = w.seg[ph.x][ph.y]->start;                  // past the leading clips
            continue;
        }
        if (ph.op == P_OVLQX) return -1;
        int hi, sa = -1, sb = -1;
        if (ph.op == P_TLT) { hi = w.v[ph.a]; sa = ph.x * 2 + ph.y; }
        else if (ph.op == P_OVL) { hi = w.v[ph.a] + 1; sa = ph.x; sb = 2 + ph.y; }
        else if (ph.op == P_OVL2) { hi = min(w.v[ph.a], w.v[ph.b]) + 1; sa = ph.x; sb = 2 + ph.y; }
        else {                                                           // P_FIN
            int which = -1;
            if (ref - 1 == w.v[ph.a]) which = ph.b; else if (ref - 1 == w.v[ph.b]) which = ph.a;
            if (which < 0) continue;
            hi = w.v[which] + 1;
            sa = ((which >> 2) & 1) * 2 + ((which >> 1) & 1);
        }
        if (hi <= ref) continue;                                         // the loop does not run
        for (int t = 0; t < 2; ++t) {
            const int sx = t ? sb : sa;
            if (sx < 0) continue;
            const int side = sx >> 1, sg = sx & 1;
            const Seg& s = *w.seg[side][sg];
            if (pos[side][sg] == INT_MIN) pos[side][sg] = s.start - s.lead;   // cursor 0 = the extended start
            if (pos[side][sg] != ref) return -1;
            if (hi - 1 > s.end + s.trail) return -1;                     // would read past the CIGAR
            pos[side][sg] = hi;
        }
        out[n].lo = ref; out[n].hi = hi; out[n].a = sa; out[n].b = sb;
        ++n;
        ref = hi;
    }
    return n;
}

// All lanes: the phases' columns, 32 at a time. Output slot of a column = emitting columns before it.
__device__ void plain_emit(const Walk& w, const PlainPhase* ph, int n_ph, OutBuf& o, const uint8_t* tab, int lane) {
    int base_out = 0;
    for (int k = 0; k < n_ph; ++k) {
        const PlainPhase P = ph[k];
        const Seg* a = P.a >= 0 ? w.seg[P.a >> 1][P.a & 1] : nullptr;
        const Seg* b = P.b >= 0 ? w.seg[P.b >> 1][P.b & 1] : nullptr;
        // emitting columns: M of a or M of b -- two intervals, their union counted by inclusion / exclusion
        auto emitted_before = [&](int x) -> int {
            int e = a ? m_before(*a, P.lo, x) : 0;
            if (b) {
                e += m_before(*b, P.lo, x);
                if (a) e -= max(0, min(min(x, a->end + 1), b->end + 1) - max(max(P.lo, a->start), b->start));
            }
            return e;
        };
        for (int p = P.lo + lane; p < P.hi; p += 32) {
            const bool ma = a && p >= a->start && p <= a->end, mb = b && p >= b->start && p <= b->end;
            if (!ma && !mb) continue;
            uint8_t base, qual;
            if (ma && mb) {
                const int qa = p - (a->start - a->lead) - a->lead_h, qb = p - (b->start - b->lead) - b->lead_h;
                entry(tab, a->seq[qa], a->qual[qa], b->seq[qb], b->qual[qb], base, qual);
            } else {
                const Seg* s = ma ? a : b;
                const int q = p - (s->start - s->lead) - s->lead_h;
                base = s->seq[q]; qual = s->qual[q];
            }
            const int slot = base_out + emitted_before(p);
            o.cig[slot] = 'M'; o.seq[slot] = base; o.qual[slot] = qual;
        }
        base_out += emitted_before(P.hi);
    }
    o.n_ops = o.n_seq = base_out;
}

// Per segment row (2*read + segment), host-built in hc_merger_set_reads: three 16-byte loads bring a segment in.
struct SegHdr {
    int32_t start, end, n_ops, n_seq;
    int32_t lead, trail, lead_h, plain;      // clip counts (computeOffset / computeRevOffset, :126-156), H among the leading clips, plain flag
    int64_t cig_off, seq_off;
};

// mergeAlignmentRecordsSingle(ar, i, j) on two plain segments that overlap or abut, all 32 lanes: the leader's tail, the
// overlap and the trailer's tail are one run of columns [min extended start, max extended end]; a column emits where either
// read shows M, and its slot is the number of emitting columns before it (interval arithmetic, no scan).
__device__ __forceinline__ void single_plain_fold(const Seg& a, const Seg& b, OutBuf& o, const uint8_t* tab, int lane) {
    const int as = a.start, ae = a.end, bs = b.start, be = b.end;               // the M ranges
    const int hi = max(ae + a.trail, be + b.trail) + 1;
    const int alen = ae - as + 1, blen = be - bs + 1;
    const int cs = max(as, bs), clen = max(0, min(ae, be) - cs + 1);            // both show M
    const uint8_t* aseq = a.seq + (a.lead - a.lead_h) - as; const uint8_t* aql = a.qual + (a.lead - a.lead_h) - as;   // index by column
    const uint8_t* bseq = b.seq + (b.lead - b.lead_h) - bs; const uint8_t* bql = b.qual + (b.lead - b.lead_h) - bs;
    // 128 columns per pass: every lane first issues the loads of its 4 columns (one memory latency per pass: the member's bytes
    // come from L2), then merges and stores them
    constexpr int U = 4;
    for (int p0 = min(as, bs) + lane; p0 < hi; p0 += 32 * U) {                  // columns left of the first M emit nothing
        uint8_t xa[U], qa[U], xb[U], qb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int p = p0 + 32 * u;
            const bool ma = p >= as && p <= ae, mb = p >= bs && p <= be;
            xa[u] = ma ? aseq[p] : 0; qa[u] = ma ? aql[p] : 0;
            xb[u] = mb ? bseq[p] : 0; qb[u] = mb ? bql[p] : 0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int p = p0 + 32 * u;
            const bool ma = p >= as && p <= ae, mb = p >= bs && p <= be;
            if (!ma && !mb) continue;
            uint8_t base, qual;
            if (ma && mb) entry(tab, xa[u], qa[u], xb[u], qb[u], base, qual);
            else if (ma) { base = xa[u]; qual = qa[u]; }
            else { base = xb[u]; qual = qb[u]; }
            const int slot = min(max(p - as, 0), alen) + min(max(p - bs, 0), blen) - min(max(p - cs, 0), clen);
            o.cig[slot] = 'M'; o.seq[slot] = base; o.qual[slot] = qual;
        }
    }
    o.n_ops = o.n_seq = alen + blen - clen;
}

struct Batch {
    const uint8_t* single_end;
    const SegHdr* seg;
    const uint8_t *cig, *seq, *qual;
    const int64_t* member_off;
    const int32_t* members;
    const int64_t* res_off;        // per clique: its capacity region in the result streams, 2 segments of cap = span / 2
    const int64_t* scratch_off;    // per clique: 4 segment buffers of cap in the scratch streams (only where cap > kSegCap)
    uint8_t *res_cig, *res_seq, *res_qual;
    uint8_t *scr_cig, *scr_seq, *scr_qual;
    hc_merged_record* info;
    const uint8_t* tab;
    int32_t n_cliques;
    unsigned long long* counters;  // [0] folds evaluated in column space, [1] folds walked
};

// Lanes 0 and 1 bring the two segments of a read into a record in shared memory (lane = segment): the loads are issued by
// fetch_rec (one fold ahead, so that their latency hides behind the current fold), commit_rec writes the record.
struct RecRegs { int4 a, f, o; int single; };
__device__ __forceinline__ RecRegs fetch_rec(const Batch& B, int read, int lane) {
    RecRegs r;
    r.a = r.f = r.o = make_int4(0, 0, 0, 0);
    r.single = 1;
    if (lane < 2 && read >= 0) {
        r.single = B.single_end[read] != 0;
        const int4* p = reinterpret_cast<const int4*>(B.seg + 2 * (size_t)read + lane);
        r.a = __ldg(p); r.f = __ldg(p + 1); r.o = __ldg(p + 2);
    }
    return r;
}
__device__ __forceinline__ void commit_rec(const Batch& B, const RecRegs& r, Rec& rec, int lane) {
    if (lane < 2) {
        const bool single = r.single != 0;
        const int4 a = r.a, f = r.f, o = r.o;
        const bool have = lane == 0 || !single;
        const long long cig_off = ((long long)(unsigned)o.y << 32) | (unsigned)o.x, seq_off = ((long long)(unsigned)o.w << 32) | (unsigned)o.z;
        Seg s;
        s.start = have ? a.x : 0; s.end = have ? a.y : 0; s.n_ops = have ? a.z : 0; s.n_seq = have ? a.w : 0;
        s.lead = have ? f.x : 0; s.trail = have ? f.y : 0; s.lead_h = have ? f.z : 0; s.plain = have ? f.w : 0;
        s.cig = B.cig + cig_off; s.seq = B.seq + seq_off; s.qual = B.qual + seq_off;
        rec.s[lane] = s;
        if (lane == 0) rec.single = single;
    }
}

// The layout(s) of one fold: AlignmentRecord's constructor dispatch (:233-242) and the branch heads of
// mergeAlignmentRecordsPaired / Mixed. Returns the number of jobs (0: no paired layout matches, the record stays as it is).
__device__ int pick_jobs(const Rec& acc, const Rec& mem, const int* v, Job* jobs) {
    if (acc.single && mem.single) { jobs[0] = Job{LAY_SINGLE, 0}; return 1; }
    if (!acc.single && !mem.single) {
        if (v[Ae1] < v[Bs2] && v[As2] > v[Be1]) { jobs[0] = Job{LAY_SINGLE, 0}; jobs[1] = Job{LAY_SINGLE + 3, 0}; return 2; }   // :766
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
        if (k < 0) return 0;
        jobs[0] = Job{k, 0};
        return 1;
    }
    if (mem.single) {                                                    // this paired, ar single (:1296-1405)
        if (v[Be1] < v[As2]) jobs[0] = Job{LAY_SINGLE, 1};
        else if (v[Bs1] > v[Ae1]) jobs[0] = Job{LAY_SINGLE + 2, 0};
        else jobs[0] = Job{LAY_PS0 + (v[xBs1] <= v[xAs1] ? 0 : 1), 0};
        return 1;
    }
    if (v[Ae1] < v[Bs2]) jobs[0] = Job{LAY_SINGLE, 2};                   // this single, ar paired (:1406-1536)
    else if (v[As1] > v[Be1]) jobs[0] = Job{LAY_SINGLE + 1, 3};
    else jobs[0] = Job{LAY_SP0 + (v[xAs1] <= v[xBs1] ? 0 : 1), 0};
    return 1;
}

// The merged string becomes a segment of the running record.
__device__ void set_merged(Seg& t, const OutBuf& o, int start, int end, bool known_plain_ops) {
    t.start = start; t.end = end;
    t.n_ops = o.n_ops; t.n_seq = o.n_seq;
    t.cig = o.cig; t.seq = o.seq; t.qual = o.qual;
    t.lead = t.trail = t.lead_h = 0;                     // merges drop the clips
    // A merged record takes min / max of the starts / ends but drops clip columns: where a clip column of one read met no M of
    // the other, its CIGAR is shorter than end - start + 1 (the reference walks such a record by cursor and runs off it); only
    // a record whose CIGAR is exactly its columns may be folded in column space
    int plain = t.n_ops == t.end - t.start + 1;
    if (!known_plain_ops) for (int k = 0; k < t.n_ops; ++k) plain &= t.cig[k] == 'M';
    t.plain = plain;
}

// Result rule of a paired / mixed layout (lane 0).
__device__ void apply_result(Rec& acc, const Rec& mem, const Result rs, const OutBuf& o, bool known_plain_ops) {
    int v[16];
    const Rec* r[2] = {&acc, &mem};
    for (int side = 0; side < 2; ++side)
        for (int sg = 0; sg < 2; ++sg) { v[MB(0, side, sg, 0)] = r[side]->s[sg].start; v[MB(0, side, sg, 1)] = r[side]->s[sg].end; }
    if (rs.take == 2) acc.s[0] = mem.s[0];
    set_merged(acc.s[rs.seg], o, min(v[rs.sa], v[rs.sb]), max(v[rs.ea], v[rs.eb]), known_plain_ops);
    if (rs.single != 2) acc.single = rs.single;
    if (rs.take == 1) acc.s[1] = mem.s[1];
}

// One job through the layout table: all lanes evaluate it in column space when every segment taking part is plain and the
// cursors stay on the columns; otherwise lane 0 walks it as the reference does. ctl[2] >= 0 afterwards: it ran in column space.
__device__ __noinline__ int generic_job(const Rec& acc, const Rec& mem, int lay, OutBuf& o, const uint8_t* tab, int lane, PlainPhase* ph_sm, int* ctl,
                                        unsigned long long& n_plain, unsigned long long& n_walk) {
    const Layout& L = c_layout[lay];
    Walk w;
    set_bounds(w, acc, mem);
    if (lane == 0) {
        bool all_plain = true;
        for (int k = 0; k < 11 && L.ph[k].op != P_END; ++k) {
            const Phase ph = L.ph[k];
            if (ph.op == P_TLT || ph.op == P_SEAT) all_plain &= w.seg[ph.x][ph.y]->plain != 0;
            else if (ph.op == P_OVL || ph.op == P_OVL2 || ph.op == P_OVLQX) all_plain &= w.seg[0][ph.x]->plain != 0 && w.seg[1][ph.y]->plain != 0;
        }
        ctl[2] = all_plain ? plain_phases(w, L, ph_sm) : -1;
    }
    __syncwarp();
    const int n_ph = ctl[2];
    if (n_ph >= 0) {
        plain_emit(w, ph_sm, n_ph, o, tab, lane);
        n_plain += lane == 0;
        return HC_MERGE_OK;
    }
    __syncwarp();
    if (lane == 0) {
        w.o = &o;
        ctl[3] = walk_layout(w, L, tab);
        ctl[0] = o.n_ops; ctl[1] = o.n_seq;
    }
    __syncwarp();
    o.n_ops = ctl[0]; o.n_seq = ctl[1];
    n_walk += lane == 0;
    return ctl[3];
}

__device__ __noinline__ int pick_jobs_general(const Rec& acc, const Rec& mem, Job* jobs) {
    Walk w0;
    set_bounds(w0, acc, mem);
    return pick_jobs(acc, mem, w0.v, jobs);
}

// Dynamic shared memory: the two Phred tables, then per warp 4 segment buffers (segment x generation) of 3 byte streams.
__global__ void __launch_bounds__(kWarps * 32, 1) merge_cliques_kernel(Batch B) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* tab = smem;
    for (int i = threadIdx.x; i < 2 * NQ * NQ; i += blockDim.x) tab[i] = B.tab[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* slice = smem + kTabBytes + (size_t)warp * kSlice;
    __shared__ Rec s_acc[kWarps], s_mem[kWarps];
    __shared__ PlainPhase s_ph[kWarps][11];
    __shared__ Job s_jobs[kWarps][2];
    __shared__ int s_ctl[kWarps][5];
    Rec& acc = s_acc[warp];
    Rec& mem = s_mem[warp];
    unsigned long long n_plain = 0, n_walk = 0;
    for (int c = blockIdx.x * kWarps + warp; c < B.n_cliques; c += gridDim.x * kWarps) {
        const int64_t m0 = B.member_off[c];
        const int m = (int)(B.member_off[c + 1] - m0);
        const int64_t ro = B.res_off[c], cap = (B.res_off[c + 1] - ro) / 2;      // capacity of one segment
        const bool in_smem = cap <= kSegCap;
        const int64_t so = B.scratch_off[c];
        auto buffer = [&](int sg, int gen) -> OutBuf {
            OutBuf o;
            if (in_smem) {
                uint8_t* p = slice + (size_t)(sg * 2 + gen) * 3 * kSegCap;
                o.cig = p; o.seq = p + kSegCap; o.qual = p + 2 * kSegCap;
            } else {
                const int64_t at = so + (int64_t)(sg * 2 + gen) * cap;
                o.cig = B.scr_cig + at; o.seq = B.scr_seq + at; o.qual = B.scr_qual + at;
            }
            o.n_ops = o.n_seq = 0;
            o.cap = (int)cap;
            return o;
        };
        int gen[2] = {0, 0};
        int status = HC_MERGE_OK;
        // the member indices of the clique, one per lane (cliques of more than 32 reads fetch the rest as they go)
        const int my_member = lane < m ? __ldg(B.members + m0 + lane) : -1;
        commit_rec(B, fetch_rec(B, __shfl_sync(0xffffffffu, my_member, 0), lane), acc, lane);
        RecRegs next = fetch_rec(B, m > 1 ? __shfl_sync(0xffffffffu, my_member, 1) : -1, lane);
        __syncwarp();
        for (int f = 1; f < m && status == HC_MERGE_OK; ++f) {
            commit_rec(B, next, mem, lane);
            {
                int nr = -1;
                if (f + 1 < m) nr = f + 1 < 32 ? __shfl_sync(0xffffffffu, my_member, (f + 1) & 31) : __ldg(B.members + m0 + f + 1);   // f, m: warp-uniform
                next = fetch_rec(B, nr, lane);
            }
            __syncwarp();
            if (lane == 0) {
                if (acc.single && mem.single) { s_jobs[warp][0] = Job{LAY_SINGLE, 0}; s_ctl[warp][4] = 1; }
                else s_ctl[warp][4] = pick_jobs_general(acc, mem, s_jobs[warp]);
            }
            __syncwarp();
            const int n_jobs = s_ctl[warp][4];
            for (int job = 0; job < n_jobs && status == HC_MERGE_OK; ++job) {
                const int lay = s_jobs[warp][job].lay, after = s_jobs[warp][job].after;
                int rc;
                if (lay >= LAY_SINGLE) {
                    // ---- mergeAlignmentRecordsSingle(ar, i, j): the bulk of all folds
                    const int i = (lay - LAY_SINGLE) >> 1, j = (lay - LAY_SINGLE) & 1;
                    const int tgt = (i | j) ? 1 : 0;
                    OutBuf o = buffer(tgt, gen[tgt] ^ 1);
                    const Seg sa = acc.s[i], sb = mem.s[j];
                    const bool fast = sa.plain && sb.plain && max(sa.start - sa.lead, sb.start - sb.lead) <= min(sa.end + sa.trail, sb.end + sb.trail) + 1;
                    if (fast) {
                        single_plain_fold(sa, sb, o, tab, lane);
                        rc = HC_MERGE_OK;
                        n_plain += lane == 0;
                    } else {
                        rc = generic_job(acc, mem, lay, o, tab, lane, s_ph[warp], s_ctl[warp], n_plain, n_walk);
                    }
                    __syncwarp();
                    if (rc != HC_MERGE_OK) { status = rc; break; }
                    if (lane == 0) {
                        // result rules of :685-729: (1,1) -> segment 1, single_end; (1,2) -> segment 2 formed from start1 / end1,
                        // flag untouched; (2,1), (2,2) -> segment 2, paired
                        Seg& t = acc.s[tgt];
                        const int st = min(sa.start, sb.start), en = max(sa.end, sb.end);
                        set_merged(t, o, st, en, fast);
                        if ((i | j) == 0) acc.single = 1; else if (i == 1) acc.single = 0;
                        if (after == 1) acc.single = 0;
                        if (after == 2) { acc.s[1] = mem.s[1]; acc.single = 0; }
                        if (after == 3) { acc.s[0] = mem.s[0]; acc.single = 0; }
                    }
                    gen[tgt] ^= 1;
                } else {
                    // ---- a paired / mixed layout: its program, walked or (plain segments) evaluated in column space
                    const int tgt = c_layout[lay].res.seg;
                    OutBuf o = buffer(tgt, gen[tgt] ^ 1);
                    rc = generic_job(acc, mem, lay, o, tab, lane, s_ph[warp], s_ctl[warp], n_plain, n_walk);
                    __syncwarp();
                    if (rc != HC_MERGE_OK) { status = rc; break; }
                    if (lane == 0) apply_result(acc, mem, c_layout[lay].res, o, s_ctl[warp][2] >= 0);
                    gen[tgt] ^= 1;
                }
                __syncwarp();
            }
        }
        // ---- the record goes out: header by lane 0, the streams by all lanes
        __syncwarp();
        if (lane == 0) {
            hc_merged_record r;
            memset(&r, 0, sizeof r);
            r.status = status;
            if (status == HC_MERGE_OK) {
                r.single_end = acc.single;
                for (int sg = 0; sg < (acc.single ? 1 : 2); ++sg) {
                    r.start[sg] = acc.s[sg].start; r.end[sg] = acc.s[sg].end;
                    r.n_ops[sg] = min(acc.s[sg].n_ops, (int)cap); r.n_seq[sg] = min(acc.s[sg].n_seq, (int)cap);
                }
            }
            B.info[c] = r;
        }
        if (status == HC_MERGE_OK) {
            for (int sg = 0; sg < (acc.single ? 1 : 2); ++sg) {
                const Seg& s = acc.s[sg];
                const int no = min(s.n_ops, (int)cap), ns = min(s.n_seq, (int)cap);
                uint8_t* dc = B.res_cig + ro + sg * cap; uint8_t* ds = B.res_seq + ro + sg * cap; uint8_t* dq = B.res_qual + ro + sg * cap;
                for (int k = lane; k < no; k += 32) dc[k] = s.cig[k];
                for (int k = lane; k < ns; k += 32) { ds[k] = s.seq[k]; dq[k] = s.qual[k]; }
            }
        }
        __syncwarp();
    }
    if (lane == 0 && B.counters) { atomicAdd(&B.counters[0], n_plain); atomicAdd(&B.counters[1], n_walk); }
}

// Compaction: record c's bytes move from its capacity region to their final offsets (a warp per record).
__global__ void compact_kernel(const hc_merged_record* __restrict__ info, const int64_t* __restrict__ res_off, const int64_t* __restrict__ ops_at,
                               const int64_t* __restrict__ seq_at, const uint8_t* __restrict__ res_cig, const uint8_t* __restrict__ res_seq,
                               const uint8_t* __restrict__ res_qual, uint8_t* __restrict__ out_cig, uint8_t* __restrict__ out_seq,
                               uint8_t* __restrict__ out_qual, int n_cliques) {
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < n_cliques; c += warps) {
        const hc_merged_record r = info[c];
        if (r.status != HC_MERGE_OK) continue;
        const int64_t ro = res_off[c], rcap = (res_off[c + 1] - ro) / 2;
        int64_t oa = ops_at[c], sa = seq_at[c];
        for (int sg = 0; sg < (r.single_end ? 1 : 2); ++sg) {
            for (int k = lane; k < r.n_ops[sg]; k += 32) out_cig[oa + k] = res_cig[ro + sg * rcap + k];
            for (int k = lane; k < r.n_seq[sg]; k += 32) { out_seq[sa + k] = res_seq[ro + sg * rcap + k]; out_qual[sa + k] = res_qual[ro + sg * rcap + k]; }
            oa += r.n_ops[sg]; sa += r.n_seq[sg];
        }
    }
}

int fail(int code, const std::string& msg) {
    hc::set_global_error(msg);
    return code;
}

}  // namespace

struct hc_merger {
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    // the read set (hc_merger_set_reads)
    int n_reads = 0;
    std::vector<uint8_t> single_end;
    std::vector<int32_t> cig_len, seg_xs, seg_xe, seg_ins;      // per segment row: CIGAR length, extended start / end, I ops
    uint8_t* d_single = nullptr;
    uint8_t *d_cig = nullptr, *d_seq = nullptr, *d_qual = nullptr, *d_tab = nullptr;
    SegHdr* d_seg = nullptr; size_t cap_seghdr = 0;
    size_t cap_single = 0, cap_cig = 0, cap_seq = 0, cap_qual = 0;
    // per batch of cliques (grow-only)
    int64_t *d_moff = nullptr, *d_resoff = nullptr, *d_scroff = nullptr, *d_opsat = nullptr, *d_seqat = nullptr; int32_t* d_members = nullptr;
    hc_merged_record* d_info = nullptr;
    uint8_t *d_scratch = nullptr, *d_res = nullptr, *d_out = nullptr;
    unsigned long long* d_counters = nullptr;
    size_t cap_moff = 0, cap_resoff = 0, cap_scroff = 0, cap_opsat = 0, cap_seqat = 0, cap_info = 0, cap_mem = 0, cap_scratch = 0, cap_res = 0, cap_out = 0;
    int64_t out_ops = 0, out_seq = 0;
    bool have_result = false;
    unsigned long long folds_plain = 0, folds_walked = 0;
    std::string err;
};

namespace {
#define MG_TRY(m, call)                                                                        \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            (m)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                    \
            hc::set_global_error((m)->err);                                                    \
            return HC_E_CUDA;                                                                  \
        }                                                                                      \
    } while (0)

template <class T>
int mgrow(hc_merger* m, T*& p, size_t& cap, size_t need) {
    if (need <= cap && p) return HC_OK;
    if (p) MG_TRY(m, cudaFree(p));
    p = nullptr; cap = 0;
    const size_t n = std::max<size_t>(need + need / 8, 256);
    MG_TRY(m, cudaMalloc(reinterpret_cast<void**>(&p), n * sizeof(T)));
    cap = n;
    return HC_OK;
}
int bad_arg(hc_merger* m, const char* msg) { m->err = msg; return fail(HC_E_INVALID, msg); }
}  // namespace

extern "C" {

int hc_merger_create(int32_t device, hc_merger** out) {
    if (!out) return fail(HC_E_INVALID, "hc_merger_create: null argument");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(HC_E_CUDA, "hc_merger_create: no CUDA device (there is no CPU merging path)");
    }
    if (device < 0 || device >= ndev) return fail(HC_E_INVALID, "hc_merger_create: bad device ordinal");
    hc_merger* m = new hc_merger();
    m->device = device;
    auto bad = [&](cudaError_t e) { hc::set_global_error(std::string("hc_merger_create: ") + cudaGetErrorString(e)); hc_merger_destroy(m); return HC_E_CUDA; };
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bad(e);
    if ((e = cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) return bad(e);
    if ((e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking)) != cudaSuccess) return bad(e);
    if ((e = cudaEventCreate(&m->e0)) != cudaSuccess) return bad(e);
    if ((e = cudaEventCreate(&m->e1)) != cudaSuccess) return bad(e);
    std::vector<uint8_t> tab(2 * NQ * NQ);
    hc::build_phred_tables(tab.data());
    if ((e = cudaMalloc((void**)&m->d_tab, tab.size())) != cudaSuccess) return bad(e);
    if ((e = cudaMemcpy(m->d_tab, tab.data(), tab.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return bad(e);
    if ((e = cudaMalloc((void**)&m->d_counters, 2 * sizeof(unsigned long long))) != cudaSuccess) return bad(e);
    *out = m;
    return HC_OK;
}

void hc_merger_destroy(hc_merger* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    cudaFree(m->d_single); cudaFree(m->d_seg); cudaFree(m->d_cig); cudaFree(m->d_seq); cudaFree(m->d_qual); cudaFree(m->d_tab); cudaFree(m->d_moff); cudaFree(m->d_resoff); cudaFree(m->d_scroff);
    cudaFree(m->d_opsat); cudaFree(m->d_seqat); cudaFree(m->d_members); cudaFree(m->d_info); cudaFree(m->d_scratch); cudaFree(m->d_res); cudaFree(m->d_out);
    cudaFree(m->d_counters);
    if (m->e0) cudaEventDestroy(m->e0);
    if (m->e1) cudaEventDestroy(m->e1);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

const char* hc_merger_last_error(const hc_merger* m) { return m ? m->err.c_str() : hc::global_error(); }

int hc_merger_set_reads(hc_merger* m, const hc_reads* rd) {
    if (!m) return fail(HC_E_INVALID, "hc_merger_set_reads: null merger");
    if (!rd || rd->n < 0) return bad_arg(m, "hc_merger_set_reads: bad argument");
    const int n = rd->n;
    int64_t cig_bytes = 0, seq_bytes = 0;
    for (int r = 0; r < n; ++r)
        for (int s = 0; s < (rd->single_end[r] ? 1 : 2); ++s) {
            const int k = 2 * r + s;
            const int nc = rd->cig_len[k], nq = rd->seq_len[k];
            if (nc < 0 || nq < 0 || rd->cig_off[k] < 0 || rd->seq_off[k] < 0) return bad_arg(m, "hc_merger_set_reads: negative length or offset");
            cig_bytes = std::max<int64_t>(cig_bytes, rd->cig_off[k] + nc);
            seq_bytes = std::max<int64_t>(seq_bytes, rd->seq_off[k] + nq);
        }
    // quality chars: the update tables cover 33..126. Chars below '!' are what computeEntry itself can emit (super-reads of
    // an earlier round) and are accepted; above 126 is refused. CIGAR ops outside MIDSH are refused.
    // In the same pass: the facts of every segment the folds need -- leading / trailing clip counts (computeOffset /
    // computeRevOffset, :126-156), H ops among the leading clips (they consume no query), and whether the segment is "plain":
    // clips, M only, clips, with a CIGAR that consumes no more query than there is and is exactly the segment's columns.
    int bad = 0;
    std::vector<int32_t> xs((size_t)std::max(2 * n, 2), 0), xe((size_t)std::max(2 * n, 2), 0), ins((size_t)std::max(2 * n, 2), 0);
    std::vector<SegHdr> hdr((size_t)std::max(2 * n, 2));
    std::memset(hdr.data(), 0, hdr.size() * sizeof(SegHdr));
    hc::parallel_ranges(n, 4096, [&](int r0, int r1) {
        int mine = 0;
        for (int r = r0; r < r1 && !mine; ++r)
            for (int s = 0; s < (rd->single_end[r] ? 1 : 2); ++s) {
                const int k = 2 * r + s;
                const uint8_t* q = rd->qual + rd->seq_off[k];
                for (int i = 0; i < rd->seq_len[k]; ++i) if (q[i] > 126) mine = 1;
                const uint8_t* cg = rd->cig + rd->cig_off[k];
                const int nc = rd->cig_len[k];
                int lead = 0, lead_h = 0, trail = 0, cons = 0;
                while (lead < nc && (cg[lead] == 'S' || cg[lead] == 'H')) { lead_h += cg[lead] == 'H'; ++lead; }
                while (trail < nc - lead && (cg[nc - 1 - trail] == 'S' || cg[nc - 1 - trail] == 'H')) ++trail;
                if (lead == nc) trail = nc;                      // clips only: computeRevOffset counts them all again
                int plain = lead + trail <= nc;
                for (int i = 0; i < nc; ++i) {
                    const uint8_t c = cg[i];
                    if (c != 'M' && c != 'I' && c != 'D' && c != 'S' && c != 'H') mine = 2;
                    cons += (c == 'M' || c == 'I' || c == 'S');
                    if (i >= lead && i < nc - trail && c != 'M') plain = 0;
                }
                if (cons > rd->seq_len[k]) plain = 0;
                if (nc != lead + trail + (rd->seg_end[k] - rd->seg_start[k] + 1)) plain = 0;
                int n_ins = 0;
                for (int i = 0; i < nc; ++i) n_ins += cg[i] == 'I';
                xs[(size_t)k] = rd->seg_start[k] - lead; xe[(size_t)k] = rd->seg_end[k] + trail; ins[(size_t)k] = n_ins;
                SegHdr& h = hdr[(size_t)k];
                h.start = rd->seg_start[k]; h.end = rd->seg_end[k]; h.n_ops = nc; h.n_seq = rd->seq_len[k];
                h.lead = lead; h.trail = trail; h.lead_h = lead_h; h.plain = plain;
                h.cig_off = rd->cig_off[k]; h.seq_off = rd->seq_off[k];
            }
        if (mine) bad = mine;
    });
    if (bad == 1) return bad_arg(m, "hc_merger_set_reads: quality char above 126");
    if (bad == 2) return bad_arg(m, "hc_merger_set_reads: CIGAR op outside MIDSH");
    MG_TRY(m, cudaSetDevice(m->device));
    MG_TRY(m, cudaStreamSynchronize(m->stream));
    m->have_result = false;
    int rc;
    const size_t n2 = (size_t)std::max(2 * n, 2);
    if ((rc = mgrow(m, m->d_single, m->cap_single, n2)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_seg, m->cap_seghdr, n2)) != HC_OK) return rc;
    MG_TRY(m, cudaMemcpyAsync(m->d_seg, hdr.data(), sizeof(SegHdr) * (size_t)(2 * n), cudaMemcpyHostToDevice, m->stream));
    if ((rc = mgrow(m, m->d_cig, m->cap_cig, (size_t)cig_bytes + 16)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_seq, m->cap_seq, (size_t)seq_bytes + 16)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_qual, m->cap_qual, (size_t)seq_bytes + 16)) != HC_OK) return rc;
    MG_TRY(m, cudaMemcpyAsync(m->d_single, rd->single_end, (size_t)n, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_cig, rd->cig, (size_t)cig_bytes, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_seq, rd->seq, (size_t)seq_bytes, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_qual, rd->qual, (size_t)seq_bytes, cudaMemcpyHostToDevice, m->stream));
    m->n_reads = n;
    m->single_end.assign(rd->single_end, rd->single_end + n);
    m->cig_len.assign(rd->cig_len, rd->cig_len + 2 * (size_t)n);
    m->seg_xs.swap(xs); m->seg_xe.swap(xe); m->seg_ins.swap(ins);
    MG_TRY(m, cudaStreamSynchronize(m->stream));
    return HC_OK;
}

int hc_merge_run(hc_merger* m, int32_t n_cliques, const int64_t* member_off, const int32_t* members, hc_merged_record* info,
                 int64_t* total_ops, int64_t* total_seq, float* kernel_ms) {
    if (!m) return fail(HC_E_INVALID, "hc_merge_run: null merger");
    if (n_cliques < 0 || (n_cliques > 0 && (!member_off || !members || !info))) return bad_arg(m, "hc_merge_run: null argument");
    if (!m->d_cig) { m->err = "hc_merge_run: no read set (hc_merger_set_reads)"; return fail(HC_E_STATE, m->err); }
    m->have_result = false;
    if (total_ops) *total_ops = 0;
    if (total_seq) *total_seq = 0;
    if (kernel_ms) *kernel_ms = 0.f;
    if (n_cliques == 0) { m->out_ops = m->out_seq = 0; m->have_result = true; return HC_OK; }
    // capacity of one segment of clique c: a fold emits one op per reference column it passes (M / D) or per insertion it
    // takes over, so a merged segment holds at most (columns the members span, clips included) + (their I ops) ops -- and never
    // more than the members' CIGAR lengths summed
    std::vector<int64_t> resoff((size_t)n_cliques + 1), scroff((size_t)n_cliques + 1);
    int64_t rat = 0, sat = 0;
    for (int c = 0; c < n_cliques; ++c) {
        if (member_off[c + 1] <= member_off[c]) return bad_arg(m, "hc_merge_run: empty clique");
        int64_t cap = 0, n_ins = 0;
        int32_t lo = INT_MAX, hi = INT_MIN;
        for (int64_t k = member_off[c]; k < member_off[c + 1]; ++k) {
            const int r = members[k];
            if (r < 0 || r >= m->n_reads) return bad_arg(m, "hc_merge_run: member index out of range");
            for (int sg = 0; sg < (m->single_end[r] ? 1 : 2); ++sg) {
                const size_t row = 2 * (size_t)r + sg;
                cap += m->cig_len[row]; n_ins += m->seg_ins[row];
                lo = std::min(lo, m->seg_xs[row]); hi = std::max(hi, m->seg_xe[row]);
            }
        }
        if (hi >= lo) cap = std::min<int64_t>(cap, (int64_t)hi - lo + 1 + n_ins);
        cap = std::max<int64_t>(16, (cap + 15) & ~(int64_t)15);
        resoff[c] = rat; scroff[c] = sat;
        rat += 2 * cap;
        sat += cap > kSegCap ? 4 * cap : 0;                // global scratch only where shared memory cannot hold the running record
    }
    resoff[n_cliques] = rat; scroff[n_cliques] = sat;
    MG_TRY(m, cudaSetDevice(m->device));
    int rc;
    const int64_t n_members = member_off[n_cliques];
    const size_t nc1 = (size_t)n_cliques + 1;
    if ((rc = mgrow(m, m->d_moff, m->cap_moff, nc1)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_resoff, m->cap_resoff, nc1)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_scroff, m->cap_scroff, nc1)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_opsat, m->cap_opsat, nc1)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_seqat, m->cap_seqat, nc1)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_info, m->cap_info, nc1)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_members, m->cap_mem, (size_t)n_members)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_scratch, m->cap_scratch, (size_t)sat * 3 + 64)) != HC_OK) return rc;
    if ((rc = mgrow(m, m->d_res, m->cap_res, (size_t)rat * 3 + 64)) != HC_OK) return rc;
    MG_TRY(m, cudaMemcpyAsync(m->d_moff, member_off, sizeof(int64_t) * nc1, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_members, members, sizeof(int32_t) * (size_t)n_members, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_resoff, resoff.data(), sizeof(int64_t) * nc1, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_scroff, scroff.data(), sizeof(int64_t) * nc1, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemsetAsync(m->d_counters, 0, 2 * sizeof(unsigned long long), m->stream));
    Batch B{};
    B.single_end = m->d_single; B.seg = m->d_seg; B.cig = m->d_cig; B.seq = m->d_seq; B.qual = m->d_qual;
    B.member_off = m->d_moff; B.members = m->d_members; B.res_off = m->d_resoff; B.scratch_off = m->d_scroff;
    B.res_cig = m->d_res; B.res_seq = m->d_res + rat; B.res_qual = m->d_res + 2 * rat;
    B.scr_cig = m->d_scratch; B.scr_seq = m->d_scratch + sat; B.scr_qual = m->d_scratch + 2 * sat;
    B.info = m->d_info; B.tab = m->d_tab; B.n_cliques = n_cliques; B.counters = m->d_counters;
    const size_t smem = (size_t)kTabBytes + (size_t)kWarps * kSlice;
    MG_TRY(m, cudaFuncSetAttribute(merge_cliques_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MG_TRY(m, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, merge_cliques_kernel, kWarps * 32, smem));
    const int blocks = std::max(1, std::min((n_cliques + kWarps - 1) / kWarps, m->sm_count * std::max(per_sm, 1)));
    MG_TRY(m, cudaEventRecord(m->e0, m->stream));
    merge_cliques_kernel<<<blocks, kWarps * 32, smem, m->stream>>>(B);
    MG_TRY(m, cudaGetLastError());
    MG_TRY(m, cudaEventRecord(m->e1, m->stream));
    MG_TRY(m, cudaMemcpyAsync(info, m->d_info, sizeof(hc_merged_record) * (size_t)n_cliques, cudaMemcpyDeviceToHost, m->stream));
    unsigned long long cnt[2] = {0, 0};
    MG_TRY(m, cudaMemcpyAsync(cnt, m->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost, m->stream));
    MG_TRY(m, cudaStreamSynchronize(m->stream));
    m->folds_plain = cnt[0]; m->folds_walked = cnt[1];
    // final offsets: clique-major, segment 1 then segment 2
    std::vector<int64_t> ops_at(nc1), seq_at(nc1);
    int64_t oa = 0, sa = 0;
    for (int c = 0; c < n_cliques; ++c) {
        ops_at[c] = oa; seq_at[c] = sa;
        if (info[c].status == HC_MERGE_OK)
            for (int sg = 0; sg < (info[c].single_end ? 1 : 2); ++sg) { oa += info[c].n_ops[sg]; sa += info[c].n_seq[sg]; }
    }
    ops_at[n_cliques] = oa; seq_at[n_cliques] = sa;
    m->out_ops = oa; m->out_seq = sa;
    if ((rc = mgrow(m, m->d_out, m->cap_out, (size_t)oa + 2 * (size_t)sa + 64)) != HC_OK) return rc;
    MG_TRY(m, cudaMemcpyAsync(m->d_opsat, ops_at.data(), sizeof(int64_t) * nc1, cudaMemcpyHostToDevice, m->stream));
    MG_TRY(m, cudaMemcpyAsync(m->d_seqat, seq_at.data(), sizeof(int64_t) * nc1, cudaMemcpyHostToDevice, m->stream));
    compact_kernel<<<std::max(1, std::min((n_cliques + 7) / 8, m->sm_count * 8)), 256, 0, m->stream>>>(
        m->d_info, m->d_resoff, m->d_opsat, m->d_seqat, m->d_res, m->d_res + rat, m->d_res + 2 * rat, m->d_out, m->d_out + oa, m->d_out + oa + sa, n_cliques);
    MG_TRY(m, cudaGetLastError());
    MG_TRY(m, cudaStreamSynchronize(m->stream));
    float ms = 0.f;
    MG_TRY(m, cudaEventElapsedTime(&ms, m->e0, m->e1));
    if (kernel_ms) *kernel_ms = ms;
    if (total_ops) *total_ops = oa;
    if (total_seq) *total_seq = sa;
    m->have_result = true;
    return HC_OK;
}

int hc_merge_fetch(hc_merger* m, uint8_t* cig, uint8_t* seq, uint8_t* qual) {
    if (!m) return fail(HC_E_INVALID, "hc_merge_fetch: null merger");
    if (!m->have_result) { m->err = "hc_merge_fetch: no result (hc_merge_run)"; return fail(HC_E_STATE, m->err); }
    if ((m->out_ops > 0 && !cig) || (m->out_seq > 0 && (!seq || !qual))) return bad_arg(m, "hc_merge_fetch: null buffer");
    MG_TRY(m, cudaSetDevice(m->device));
    if (m->out_ops > 0) MG_TRY(m, cudaMemcpyAsync(cig, m->d_out, (size_t)m->out_ops, cudaMemcpyDeviceToHost, m->stream));
    if (m->out_seq > 0) {
        MG_TRY(m, cudaMemcpyAsync(seq, m->d_out + m->out_ops, (size_t)m->out_seq, cudaMemcpyDeviceToHost, m->stream));
        MG_TRY(m, cudaMemcpyAsync(qual, m->d_out + m->out_ops + m->out_seq, (size_t)m->out_seq, cudaMemcpyDeviceToHost, m->stream));
    }
    MG_TRY(m, cudaStreamSynchronize(m->stream));
    return HC_OK;
}

int hc_merge_fold_counts(const hc_merger* m, int64_t* column_space, int64_t* walked) {
    if (!m) return HC_E_INVALID;
    if (column_space) *column_space = (int64_t)m->folds_plain;
    if (walked) *walked = (int64_t)m->folds_walked;
    return HC_OK;
}

}  // extern "C"
