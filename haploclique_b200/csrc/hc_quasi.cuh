// hc_quasi.cuh -- the reference's older scorer, QuasispeciesEdgeCalculator::edgeBetween
// (src/QuasispeciesEdgeCalculator.cpp:55-70; computeOverlap :72-131, overlapSize :133-179, singleOverlap :181-426,
// computeOffsets :428-493), SURVEY.md 8(a) row A13. Included by hc_gpu.cu; one lane per ordered pair.
//
// What is precomputed on the host (hc_gpu.cu: quasi_prepare), so that the device does no transcendental per column:
//   logsum[b1][b2][q1][q2] = log( sum over A,C,G,T of (b1==x ? qc1 : (1-qc1)/3) * (b2==x ? qc2 : (1-qc2)/3) )   (:307-324)
//     with qc = 1 - pow(10, -(Q)/10.0) (src/ShortDnaSequence.cpp:117-119), the reference's expression and summation
//     order, host libm: the running sum overlap_probability is then bit-identical to the reference's;
//   simpson_ln[pos] = log(SIMPSON_MAP[pos]) for the positions that have an entry, 0.0 elsewhere (:290-294).
// Only the closing pow(exp(sum), 1/n) (:425) uses device libm; pairs whose q lies within ambiguity_rel_tol of the
// cutoff are flagged, as in the NewEdgeCalculator path.
//
// The reference keeps separate cursors into the CIGAR and into the base/quality strings (they can drift apart in
// exact-match mode, :298-306). The op stream here holds base and quality on the query-consuming records (M, I, S),
// so the sequence cursor is the op index of the j_seq-th consuming record.
#pragma once

namespace quasi {

struct QParams {
    double cutoff[3];          // cliques, mixed, single (:59-66)
    double min_overlap[2];     // cliques, single (:74-79)
    double amb_tol;
    const double* logsum;      // [5][5][K][K]
    const double* simpson_ln;  // dense or nullptr
    int simpson_len;
    int K;
    int frameshift;
    const int4* qseg;          // per segment (2*read + s): {consuming ops, last consuming op, first non-clip op, first S at or after it}
};

struct SegV { const uint16_t* ops; int n, start, end; int4 q; };

// Per-segment facts the fast walk needs instead of scanning the op stream per pair (built once per upload).
__global__ void build_qseg_kernel(const hc_read_hdr* __restrict__ hdr, const uint16_t* __restrict__ ops, int n_reads, int4* __restrict__ qseg) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * n_reads) return;
    const hc_read_hdr& h = hdr[t >> 1];
    const int s = t & 1, n = h.n_ops[s];
    const uint16_t* o = ops + h.ops_off[s];
    int nq = 0, lastc = -1, first = n, mid_s = n;
    for (int c = 0; c < n; ++c) {
        const int op = (int)HC_REC_OP((uint32_t)o[c]);
        if (op == HC_OP_M || op == HC_OP_I || op == HC_OP_S) { nq++; lastc = c; }
        if (first == n && op != HC_OP_S && op != HC_OP_H) first = c;
        if (first != n && mid_s == n && op == HC_OP_S) mid_s = c;
    }
    qseg[t] = make_int4(nq, lastc, first, mid_s);
}

__device__ __forceinline__ int opc(const SegV& s, int c) { return (int)HC_REC_OP((uint32_t)__ldg(s.ops + c)); }
__device__ __forceinline__ bool consuming(int op) { return op == HC_OP_M || op == HC_OP_I || op == HC_OP_S; }
__device__ __forceinline__ int next_consuming(const SegV& s, int from) {
    while (from < s.n && !consuming(opc(s, from))) from++;
    return from;
}

// overlapSize (:133-179)
__device__ __forceinline__ float overlap_size(int e1, int e2, int s1, int s2) {
    if (s1 == s2 && e1 == e2) return (float)(e1 - s1);
    if (s1 == s2) return (float)(e1 < e2 ? e1 - s1 : e2 - s2);
    if (e1 == e2) return (float)(s1 > s2 ? e1 - s1 : e2 - s2);
    if (e1 < e2) return (float)(s1 < s2 ? e1 - s2 : e1 - s1);
    return (float)(s2 < s1 ? e2 - s1 : e2 - s2);
}

// computeOffsets (:428-493)
__device__ __forceinline__ void compute_offsets(int s1, int s2, int e1, int e2, int& offset1, int& offset2, int& overlap) {
    const int x_l1 = e1 - s1, x_l2 = e2 - s2;
    if (s1 == s2 && e1 == e2) { overlap = e1 - s1; offset1 = 0; offset2 = 0; }
    else if (s1 == s2) { overlap = e1 < e2 ? e1 - s1 : e2 - s2; offset1 = 0; offset2 = 0; }
    else if (e1 == e2) {
        if (s1 > s2) { overlap = e1 - s1; offset1 = 0; offset2 = x_l2 - overlap; }
        else { overlap = e2 - s2; offset1 = x_l1 - overlap; offset2 = 0; }
    } else if (e1 < e2) {
        if (s1 < s2) { overlap = e1 - s2; offset1 = x_l1 - overlap; offset2 = 0; }
        else { overlap = e1 - s1; offset1 = 0; offset2 = e2 - (e2 - e1) - overlap - s2; }
    } else {
        if (s2 < s1) { overlap = e2 - s1; offset1 = 0; offset2 = x_l2 - overlap; }
        else { overlap = e2 - s2; offset1 = e1 - (e1 - e2) - overlap - s1; offset2 = 0; }
    }
}

// exp() with gradual underflow. CUDA's exp(double) returns 0 once the result is subnormal; glibc (the reference) rounds
// to the subnormal grid, and pow(subnormal, 1/n) is an ordinary number (1e-310 ^ (1/700) = 0.36), so the difference
// decides edges at low cutoffs. Below log(DBL_MIN): exp(x + 128 ln 2) in the normal range, scaled by 2^-128 (one
// rounding to the subnormal grid). The pairs are flagged HC_QUASI_SUBNORMAL either way.
__device__ __forceinline__ double exp_gradual(double x) {
    if (x >= -708.3964185322641) return exp(x);
    if (x < -745.2) return 0.0;
    return scalbn(exp(x + 88.722839111672999), -128);
}

__device__ __forceinline__ void add_simpson(const QParams& P, int pos, double& acc) {
    if (P.simpson_ln != nullptr && pos >= 0 && pos < P.simpson_len) acc = __dadd_rn(acc, __ldg(P.simpson_ln + pos));
}

// singleOverlap (:181-426) without the per-pair scans of the general restatement below. The reference's sequence
// cursors (j_seq, j_seq2) are functions of its CIGAR cursors as long as every op a cursor passes advances both or
// neither in step with "consumes a base": true for the clip loop (:258-267), for M/I/D in the prefix loop (:268-295)
// and for every branch of the main loop (:307-381) -- not in exact-match mode (:298-306 advances all four cursors
// on any op) and not when the prefix loop runs over an S op (:268-295 passes it without touching j_seq). Under that
// invariant sequence[j_seq] is the record at the CIGAR cursor whenever that op is M or I, and "j_seq < size" is
// "a consuming op exists at or after the CIGAR cursor". Returns false (nothing decided) when the invariant cannot be
// guaranteed; the caller then runs the general walk. Same additions in the same order: bit-identical sums.
__device__ __forceinline__ bool single_overlap_fast(const SegV& A, const SegV& B, const QParams& P, double MIN_OVERLAP, double cutoff,
                                                    bool& undef, bool& subnormal, double& result) {
    if (cutoff == 1.0 || P.qseg == nullptr) return false;
    const bool swap = B.start < A.start;
    const SegV& S1 = swap ? B : A;
    const SegV& S2 = swap ? A : B;
    const int s1 = S1.start, e1 = S1.end, s2 = S2.start, e2 = S2.end;
    const int n1 = S1.n, n2 = S2.n;
    int offset1 = 0, offset2 = 0, overlap = 0;
    compute_offsets(s1, s2, e1, e2, offset1, offset2, overlap);
    result = 0.0;
    if ((double)overlap < MIN_OVERLAP) return true;
    if (offset1 < 0 || offset2 < 0 || offset1 >= S1.q.x || offset2 >= S2.q.x) return true;
    const int last1 = S1.q.y, last2 = S2.q.y;
    int jc1 = 0, jc2 = 0, j = 0, j2 = 0;
    for (;;) {
        if (!(jc1 < n1 && jc2 < n2)) { undef = true; return true; }
        const int c1 = opc(S1, jc1), c2 = opc(S2, jc2);
        if (c1 != HC_OP_S && c2 != HC_OP_S && c1 != HC_OP_H && c2 != HC_OP_H) break;
        if (c1 == HC_OP_S) jc1++;
        if (c2 == HC_OP_S) jc2++;
        if (jc1 >= n1 || jc2 >= n2) { undef = true; return true; }
        if (opc(S1, jc1) == HC_OP_H) jc1++;
        if (opc(S2, jc2) == HC_OP_H) jc2++;
    }
    if (jc1 < S1.q.z || min(jc1 + offset1, n1) > S1.q.w) { undef = false; return false; }     // the prefix would pass an S
    double acc = 0.0, total_size = 0.0;
    const bool smap = P.simpson_ln != nullptr;
    if (!smap) jc1 += offset1;
    else {
        for (int k = 0; k != offset1; ++k) {
            if (jc1 < n1 && opc(S1, jc1) == HC_OP_M) { add_simpson(P, s1 + j, acc); j++; }
            jc1++;
        }
    }
    while (jc1 < n1 && jc2 < n2 && jc1 <= last1 && jc2 <= last2) {
        const uint32_t r1 = __ldg(S1.ops + jc1), r2 = __ldg(S2.ops + jc2);
        const int c1 = (int)HC_REC_OP(r1), c2 = (int)HC_REC_OP(r2);
        if (c1 == HC_OP_S || c2 == HC_OP_S) break;
        if (c1 == c2 && (c1 == HC_OP_M || c1 == HC_OP_I)) {
            const size_t idx = (((size_t)HC_REC_BASE(r1) * 5 + HC_REC_BASE(r2)) * P.K + HC_REC_QUAL(r1)) * P.K + HC_REC_QUAL(r2);
            acc = __dadd_rn(acc, __ldg(P.logsum + idx));
            total_size += 1.0;
            jc1++; jc2++;
            if (smap && jc1 < n1 && opc(S1, jc1) == HC_OP_M) { j++; j2++; }
        } else if (c1 == c2 && c1 == HC_OP_D) {
            jc1++; jc2++;
        } else if (P.frameshift) {
            if (c2 == HC_OP_M) {
                if (c1 == HC_OP_D) { jc1++; jc2++; j2++; }
                else if (c1 == HC_OP_I) { jc1++; }
                else { undef = true; return true; }
            } else if (c1 == HC_OP_M) {
                if (c2 == HC_OP_D) { jc2++; jc1++; j++; }
                else if (c2 == HC_OP_I) { jc2++; }
                else { undef = true; return true; }
            } else {
                return true;
            }
        } else {
            return true;
        }
    }
    if (smap) {
        if (jc1 < n1 && opc(S1, jc1) != HC_OP_S) {
            while (jc1 < n1) {
                const int c = opc(S1, jc1);
                if (c == HC_OP_S) break;
                if (c == HC_OP_M) { add_simpson(P, s1 + j, acc); j++; }
                jc1++;
            }
        } else if (jc2 < n2 && opc(S2, jc2) != HC_OP_S) {
            while (jc2 < n2) {
                const int c = opc(S2, jc2);
                if (c == HC_OP_S) break;
                if (c == HC_OP_M) { add_simpson(P, s2 + j2, acc); j2++; }
                jc2++;
            }
        }
    }
    if (acc < -708.3964185322641) subnormal = true;
    result = pow(exp_gradual(acc), __ddiv_rn(1.0, total_size));
    return true;
}

// singleOverlap (:181-426), the general restatement. `undef`: the reference would spin forever or index out of range here.
__device__ double single_overlap(const SegV& A, const SegV& B, const QParams& P, double MIN_OVERLAP, double cutoff, bool& undef,
                                 bool& subnormal) {
    {
        double r = 0.0;
        if (single_overlap_fast(A, B, P, MIN_OVERLAP, cutoff, undef, subnormal, r)) return r;
    }
    const bool swap = B.start < A.start;                                                    // :215-220
    const SegV& S1 = swap ? B : A;
    const SegV& S2 = swap ? A : B;
    const int s1 = S1.start, e1 = S1.end, s2 = S2.start, e2 = S2.end;
    const int n1 = S1.n, n2 = S2.n;
    int offset1 = 0, offset2 = 0, overlap = 0;
    compute_offsets(s1, s2, e1, e2, offset1, offset2, overlap);                             // :227
    if ((double)overlap < MIN_OVERLAP) return 0.0;                                          // :235
    int nq1 = 0, nq2 = 0;                                                                   // sequence lengths
    for (int c = 0; c < n1; ++c) nq1 += consuming(opc(S1, c)) ? 1 : 0;
    for (int c = 0; c < n2; ++c) nq2 += consuming(opc(S2, c)) ? 1 : 0;
    if (offset1 < 0 || offset2 < 0 || offset1 >= nq1 || offset2 >= nq2) return 0.0;         // :249 (int -> size_t)
    const bool perfect = cutoff == 1.0;                                                     // :253
    double overlap_probability = 0.0, total_size = 0.0;
    int j_cigar = 0, j_cigar2 = 0, j = 0, j2 = 0;
    int sq1 = next_consuming(S1, 0), sq2 = next_consuming(S2, 0);      // op index of sequence1[j_seq] / sequence2[j_seq2]
#define Q_ADV1() (sq1 = sq1 < n1 ? next_consuming(S1, sq1 + 1) : n1)
#define Q_ADV2() (sq2 = sq2 < n2 ? next_consuming(S2, sq2 + 1) : n2)
    for (;;) {                                                                              // :258-267
        if (!(j_cigar < n1 && j_cigar2 < n2)) { undef = true; return 0.0; }
        const int c1 = opc(S1, j_cigar), c2 = opc(S2, j_cigar2);
        if (c1 != HC_OP_S && c2 != HC_OP_S && c1 != HC_OP_H && c2 != HC_OP_H) break;
        if (c1 == HC_OP_S) { j_cigar++; Q_ADV1(); }
        if (c2 == HC_OP_S) { j_cigar2++; Q_ADV2(); }
        if (j_cigar >= n1 || j_cigar2 >= n2) { undef = true; return 0.0; }
        if (opc(S1, j_cigar) == HC_OP_H) j_cigar++;
        if (opc(S2, j_cigar2) == HC_OP_H) j_cigar2++;
    }
    for (int j_prefix = 0; j_prefix != offset1; ++j_prefix) {                               // :268-295
        if (j_cigar < n1) {
            const int c = opc(S1, j_cigar);
            if (c == HC_OP_M) {
                if (!perfect) add_simpson(P, s1 + j, overlap_probability);
                j++;
                Q_ADV1();
            } else if (c == HC_OP_I) {
                Q_ADV1();
            }
        }
        j_cigar++;
    }
    while (j_cigar < n1 && j_cigar2 < n2 && sq1 < n1 && sq2 < n2) {                          // :296
        const int c1 = opc(S1, j_cigar), c2 = opc(S2, j_cigar2);
        if (c1 == HC_OP_S || c2 == HC_OP_S) break;
        if (perfect) {                                                                      // :298-306
            const uint32_t r1 = __ldg(S1.ops + sq1), r2 = __ldg(S2.ops + sq2);
            if (HC_REC_BASE(r1) != HC_REC_BASE(r2)) return 0.0;
            j_cigar++; j_cigar2++; Q_ADV1(); Q_ADV2();
        } else if (c1 == c2 && (c1 == HC_OP_M || c1 == HC_OP_I)) {                          // :307-343
            const uint32_t r1 = __ldg(S1.ops + sq1), r2 = __ldg(S2.ops + sq2);
            const size_t idx = (((size_t)HC_REC_BASE(r1) * 5 + HC_REC_BASE(r2)) * P.K + HC_REC_QUAL(r1)) * P.K + HC_REC_QUAL(r2);
            overlap_probability = __dadd_rn(overlap_probability, __ldg(P.logsum + idx));
            total_size += 1.0;
            j_cigar++; j_cigar2++; Q_ADV1(); Q_ADV2();
            if (j_cigar < n1 && opc(S1, j_cigar) == HC_OP_M) { j++; j2++; }
        } else if (c1 == c2 && c1 == HC_OP_D) {                                             // :344-346
            j_cigar++; j_cigar2++;
        } else if (P.frameshift) {                                                          // :347-381
            if (c2 == HC_OP_M) {
                if (c1 == HC_OP_D) { j_cigar++; j_cigar2++; Q_ADV2(); j2++; }
                else if (c1 == HC_OP_I) { j_cigar++; Q_ADV1(); }
                else { undef = true; return 0.0; }                 // `default: break`: the reference spins
            } else if (c1 == HC_OP_M) {
                if (c2 == HC_OP_D) { j_cigar2++; j_cigar++; Q_ADV1(); j++; }
                else if (c2 == HC_OP_I) { j_cigar2++; Q_ADV2(); }
                else { undef = true; return 0.0; }
            } else {
                return 0.0;
            }
        } else {
            return 0.0;
        }
    }
#undef Q_ADV1
#undef Q_ADV2
    if (!perfect) {                                                                         // :384-420
        if (j_cigar < n1 && opc(S1, j_cigar) != HC_OP_S) {
            while (j_cigar < n1) {
                const int c = opc(S1, j_cigar);
                if (c == HC_OP_S) break;
                if (c == HC_OP_M) { add_simpson(P, s1 + j, overlap_probability); j++; }
                j_cigar++;
            }
        } else if (j_cigar2 < n2 && opc(S2, j_cigar2) != HC_OP_S) {
            while (j_cigar2 < n2) {
                const int c = opc(S2, j_cigar2);
                if (c == HC_OP_S) break;
                if (c == HC_OP_M) { add_simpson(P, s2 + j2, overlap_probability); j2++; }
                j_cigar2++;
            }
        }
    }
    if (perfect) return 1.0;
    if (overlap_probability < -708.3964185322641) subnormal = true;                         // log(DBL_MIN)
    return pow(exp_gradual(overlap_probability), __ddiv_rn(1.0, total_size));                       // :425
}

__device__ __forceinline__ SegV seg_view(const hc_read_hdr& h, const uint16_t* ops, int s) {
    SegV v;
    v.ops = ops + h.ops_off[s];
    v.n = h.n_ops[s];
    v.start = h.raw[2 * s];
    v.end = h.raw[2 * s + 1];
    v.q = make_int4(0, 0, 0, 0);
    return v;
}

// edgeBetween (:55-70) + computeOverlap (:72-131) for ap1 = read i, ap2 = read j. The same-name shortcut (:56) does
// not apply: a batch holds one record per name and i != j.
__device__ void score_pair(const QParams& P, const hc_read_hdr* __restrict__ hdr, const uint16_t* __restrict__ ops, int i, int j,
                           hc_quasi_result& out) {
    const hc_read_hdr ha = hdr[i], hb = hdr[j];
    const bool cl1 = ha.ncov_flags & HC_FLAG_IS_CLIQUE, cl2 = hb.ncov_flags & HC_FLAG_IS_CLIQUE;
    const bool se1 = ha.ncov_flags & HC_FLAG_SINGLE_END, se2 = hb.ncov_flags & HC_FLAG_SINGLE_END;
    const double cutoff = P.cutoff[(cl1 && cl2) ? 0 : ((cl1 || cl2) ? 1 : 2)];
    const double MIN_OVERLAP = P.min_overlap[(cl1 && cl2) ? 0 : 1];
    SegV a1 = seg_view(ha, ops, 0), a2 = seg_view(ha, ops, 1), b1 = seg_view(hb, ops, 0), b2 = seg_view(hb, ops, 1);
    if (P.qseg != nullptr) {
        a1.q = __ldg(P.qseg + 2 * i); a2.q = __ldg(P.qseg + 2 * i + 1);
        b1.q = __ldg(P.qseg + 2 * j); b2.q = __ldg(P.qseg + 2 * j + 1);
    }
    bool undef = false, subnormal = false;
    int ncalls = 0;
    double q = 0.0;
    const bool ratio = MIN_OVERLAP <= 1.0;
    if (se1 && se2) {                                                                       // :81-86
        const int read_size1 = min(a1.end - a1.start, b1.end - b1.start);
        float o1 = overlap_size(a1.end, b1.end, a1.start, b1.start);
        if (ratio) o1 = __fdiv_rn(o1, (float)read_size1);
        if (!((double)o1 < MIN_OVERLAP)) { ncalls = 1; q = single_overlap(a1, b1, P, MIN_OVERLAP, cutoff, undef, subnormal); }
    } else if (se1 || se2) {                                                                // :87-117
        const SegV& x2 = se1 ? a1 : a2;
        const SegV& y2 = se1 ? b2 : b1;
        const int read_size1 = min(a1.end - a1.start, b1.end - b1.start);
        const int read_size2 = min(x2.end - x2.start, y2.end - y2.start);
        float o1 = overlap_size(a1.end, b1.end, a1.start, b1.start);
        float o2 = overlap_size(x2.end, y2.end, x2.start, y2.start);
        if (ratio) { o1 = __fdiv_rn(o1, (float)read_size1); o2 = __fdiv_rn(o2, (float)read_size2); }
        if (((double)o1 >= MIN_OVERLAP && (double)o2 >= MIN_OVERLAP) ||
            ((double)__fadd_rn(o1, o2) >= MIN_OVERLAP && o1 > 0.f && o2 > 0.f)) {
            ncalls = 2;
            const double p = single_overlap(a1, b1, P, MIN_OVERLAP, cutoff, undef, subnormal);
            const double r = single_overlap(x2, y2, P, MIN_OVERLAP, cutoff, undef, subnormal);
            q = __dmul_rn(p, r);
        }
    } else {                                                                                // :118-129
        const int read_size1 = min(a1.end - a1.start, b1.end - b1.start);
        float o1 = overlap_size(a1.end, b1.end, a1.start, b1.start);
        if (ratio) o1 = __fdiv_rn(o1, (float)read_size1);
        if (!((double)o1 < MIN_OVERLAP)) {
            const int read_size2 = min(a2.end - a2.start, b2.end - b2.start);
            float o2 = overlap_size(a2.end, b2.end, a2.start, b2.start);
            if (ratio) o2 = __fdiv_rn(o2, (float)read_size2);
            if (!((double)o2 < MIN_OVERLAP)) {
                ncalls = 2;
                const double p = single_overlap(a1, b1, P, MIN_OVERLAP, cutoff, undef, subnormal);
                const double r = single_overlap(a2, b2, P, MIN_OVERLAP, cutoff, undef, subnormal);
                q = __dmul_rn(p, r);
            }
        }
    }
    out.status = undef ? HC_ST_UNDEFINED : HC_ST_WALKED;
    out.n_single = ncalls;
    out.q = undef ? 0.0 : q;
    out.cutoff = cutoff;
    out.edge = (!undef && q >= cutoff) ? 1 : 0;
    out.flags = ((!undef && ncalls > 0 && fabs(q - cutoff) <= P.amb_tol * fabs(cutoff)) ? HC_QUASI_AMBIGUOUS : 0) |
                ((!undef && subnormal) ? HC_QUASI_SUBNORMAL : 0);
}

__global__ void __launch_bounds__(128) pairs_kernel(QParams P, const hc_read_hdr* __restrict__ hdr, const uint16_t* __restrict__ ops,
                                                    const int32_t* __restrict__ pi, const int32_t* __restrict__ pj, long long npairs,
                                                    hc_quasi_result* __restrict__ out) {
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npairs; k += (long long)gridDim.x * blockDim.x) {
        hc_quasi_result r;
        score_pair(P, hdr, ops, pi[k], pj[k], r);
        out[k] = r;
    }
}

// One warp per 32-bit adjacency word of the band layout (as score_band_kernel).
__global__ void __launch_bounds__(128) band_kernel(QParams P, const hc_read_hdr* __restrict__ hdr, const uint16_t* __restrict__ ops,
                                                   int row_begin, int nrows, const int32_t* __restrict__ col0,
                                                   const long long* __restrict__ word_off, long long total_words,
                                                   uint32_t* __restrict__ adj, BandStatsDev* __restrict__ stats) {
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    unsigned long long n_total = 0, n_walked = 0, n_edges = 0, n_undef = 0, n_amb = 0;
    for (long long wd = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; wd < total_words; wd += warps) {
        int lo = 0, hi = nrows;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (word_off[mid] <= wd) lo = mid; else hi = mid;
        }
        const int i = row_begin + lo;
        const int j = col0[lo] + (int)(wd - word_off[lo]) * 32 + lane;
        bool edge = false;
        if (j < i) {
            hc_quasi_result r;
            score_pair(P, hdr, ops, i, j, r);
            edge = r.edge != 0;
            n_total++;
            n_walked += r.n_single > 0;
            n_undef += r.status == HC_ST_UNDEFINED;
            n_amb += (r.flags & HC_QUASI_AMBIGUOUS) ? 1 : 0;
        }
        const unsigned bits = __ballot_sync(0xffffffffu, edge);
        if (lane == 0) adj[wd] = bits;
        n_edges += edge;
    }
    n_total = warp_sum(n_total); n_walked = warp_sum(n_walked); n_edges = warp_sum(n_edges);
    n_undef = warp_sum(n_undef); n_amb = warp_sum(n_amb);
    if (lane == 0) {
        atomicAdd(&stats->pairs_total, n_total); atomicAdd(&stats->pairs_scored, n_total); atomicAdd(&stats->pairs_walked, n_walked);
        atomicAdd(&stats->edges, n_edges); atomicAdd(&stats->undefined, n_undef); atomicAdd(&stats->ambiguous, n_amb);
    }
}

}  // namespace quasi
