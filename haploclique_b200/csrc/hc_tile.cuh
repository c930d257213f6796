// hc_tile.cuh -- the tiled band kernel (included by hc_gpu.cu after DevParams / op_action / BandStatsDev).
//
// Work shape. The adjacency band is cut into column blocks of 32 earlier reads ("j-block" = one 32-bit
// adjacency word per row). A CTA takes one work item = (j-block, chunk of rows). The 32 column reads' op
// streams are staged in shared memory once per item, transposed so that 32-bit word w of lane l sits at
// tile[w*32+l] (bank == lane for any cursor position: conflict-free per-lane streaming). Each warp then takes
// rows i from a shared counter; the row read's header and op stream arrive in the warp's own double buffer by
// TMA bulk copies (cp.async.bulk + mbarrier) issued one row ahead. Lane l of the warp decides the pair
// (ap1 = row read i, ap2 = column read 32*jb+l), so ap1's records are a shared-memory broadcast, the 32 edge
// bits come out of one ballot, and the two match tables are read from shared memory.
//
// Walk. The reference's layouts (src/NewEdgeCalculator.cpp:269-1061) are phase programs as in the v1 kernel, but
// re-slotted so that every program has its lock-step ("overlap") loops in the same three slots: the warp runs
// three overlap loops in which all 32 lanes iterate together; the phases between them are O(1) per lane:
//   * a tail (noOverlapCheckgap run, :234-267) only counts M/I ops and moves a cursor, so it is evaluated from
//     per-segment prefix counts (arithmetic for segments that are plain M between their clips, two small
//     lookups otherwise) instead of a loop; prob0 is then ladder[tc] (tc sequential additions of log10(0.25)),
//     which is why this kernel is used only when no allele-frequency map is set;
//   * the overlap loop multiplies probM by one table entry per shared column with __dmul_rn in increasing
//     column order -- bit-identical to the reference's running product -- two columns per iteration while both
//     reads show M (the common case), one generic step otherwise (clips, indels, the :629 cursor quirk).
#pragma once

namespace tile {

enum : uint32_t { M_NOP = 0, M_START, M_TLT, M_SEAT, M_OVL, M_OVL2, M_OVLQX, M_FIN, M_SPRE, M_SOVL, M_SPOST };
enum { TP_WORDS = 16, TP_SINGLE00 = 16, TP_SINGLE10 = 17, TP_SINGLE01 = 18, TP_TWO_SINGLES = 19, TP_NOP = 20, TP_COUNT = 21 };
enum { SLOT_OVL0 = 3, SLOT_OVL1 = 7, SLOT_OVL2 = 11, SLOT_FINAL = 12 };
#define TILE_MOP(op, x, y, a, b) ((uint32_t)(op) | ((uint32_t)(x) << 4) | ((uint32_t)(y) << 5) | ((uint32_t)(a) << 6) | ((uint32_t)(b) << 10))

#ifndef HC_TILE_WARPS
#define HC_TILE_WARPS 24
#endif
#ifndef HC_TILE_MINBLOCKS
#define HC_TILE_MINBLOCKS 1
#endif
constexpr int kWarps = HC_TILE_WARPS;
constexpr int kThreads = kWarps * 32;
constexpr unsigned FULL = 0xffffffffu;

struct Args {
    DevParams P;
    const hc_read_hdr* hdr;
    const uint16_t* ops;
    const uint32_t* aux;        // per segment, n_ops+1 entries at ops_off[s] + 2*read + s: (non-I ops before c) << 16 | (M/I ops before c)
    const uint16_t* col2op;     // per segment at ops_off[s]: op index of the k-th non-I op
    const uint32_t* prog;       // [TP_COUNT][TP_WORDS] slotted programs
    int n_reads, row_begin;
    const int32_t* col0;        // per row (row - row_begin)
    const long long* word_off;
    const int32_t* colidx;      // column blocks: 32 read indices per block (-1 = empty slot), see build_items()
    const int4* items;          // {column block, first row, end row, 0}
    int n_items;
    int* item_counter;
    uint32_t* adj;
    BandStatsDev* stats;
    hc_ambiguous* amb;
    int amb_cap;
    hc_pair_result* detail;     // optional: one record per band slot (word*32 + lane)
    int maxw;                   // 32-bit words per lane in the column tile (incl. read-ahead slack)
    int abuf_bytes;             // bytes of one row op buffer (multiple of 16, incl. slack)
};

struct RowInfo { unsigned ops_off; int n_ops; int col0; long long word_off; };

// ---- mbarrier / TMA bulk copy (1D) -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    unsigned long long t0 = 0;
    for (unsigned spins = 0;; ++spins) {
        uint32_t ok;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (ok) return;
        // a bulk copy that never lands must fail loudly, not hang the GPU: give up after two seconds of wall clock (the
        // nanosecond timer is read only every 4096 polls)
        if ((spins & 4095u) == 4095u) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 2000000000ull) __trap();
        }
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- header access: A = the row read (16 ints, linear), B = the lane's column read (field f at hB[f*32]) -----
// fields: 0-3 raw{s1,e1,s2,e2}  4-7 ext  8-9 ops_off  10-11 n_ops  12 ncov_flags  13 cov_first  14 cov_last  15 seg_flags
__device__ __forceinline__ int bound(const int* hA, const int* hB, int id) {
    const int f = ((id >> 3) << 2) | (id & 3);
    return (id & 4) ? hB[f * 32] : hA[f];
}

struct Lane {
    int ref, cc, tc;
    double probM;
    unsigned long long curs;   // four 16-bit CIGAR cursors, index side*2+seg
    bool gap, undef;
};

__device__ __forceinline__ int get_cur(const Lane& L, int side, int seg) { return (int)((L.curs >> (16 * (side * 2 + seg))) & 0xffffull); }
__device__ __forceinline__ void set_cur(Lane& L, int side, int seg, int c) {
    const int sh = 16 * (side * 2 + seg);
    L.curs = (L.curs & ~(0xffffull << sh)) | ((unsigned long long)(unsigned)c << sh);
}

// A run of noOverlapCheckgap steps on (side, seg) up to the exclusive column `limit`, in O(1).
__device__ __forceinline__ void tail(const Args& a, const int* hA, const int* hB, int rA, int rB, int side, int seg, int limit, Lane& L) {
    const int ncols = limit - L.ref;
    if (ncols <= 0) return;
    const int* h = side ? hB : hA;
    const int st = side ? 32 : 1;
    const int n = h[(10 + seg) * st];
    const int c0 = get_cur(L, side, seg);
    if (c0 >= n) { L.undef = true; return; }
    const unsigned fl = ((unsigned)h[15 * st] >> (16 * seg)) & 0xffffu;
    int c_end, dtc;
    if (fl == 0) {                                   // clips, plain M, clips
        c_end = c0 + ncols;
        if (c_end > n) { L.undef = true; return; }
        const int lead = h[(2 * seg) * st] - h[(4 + 2 * seg) * st];
        const int trail = h[(5 + 2 * seg) * st] - h[(1 + 2 * seg) * st];
        const int nm = n - lead - trail;
        dtc = min(max(c_end - lead, 0), nm) - min(max(c0 - lead, 0), nm);
    } else {
        const unsigned off = (unsigned)h[(8 + seg) * st];
        const uint32_t* ax = a.aux + (size_t)off + 2u * (unsigned)(side ? rB : rA) + seg;
        const uint32_t x0 = __ldg(ax + c0), xn = __ldg(ax + n);
        const int k0 = (int)(x0 >> 16), ktot = (int)(xn >> 16);
        if (k0 + ncols > ktot) { L.undef = true; return; }
        c_end = (int)__ldg(a.col2op + (size_t)off + k0 + ncols - 1) + 1;
        dtc = (int)(__ldg(ax + c_end) & 0xffffu) - (int)(x0 & 0xffffu);
    }
    L.tc += dtc;
    L.ref = limit;
    set_cur(L, side, seg, c_end);
}

// One of the cheap (non-loop) micro-ops. `shape` carries checkGapsSingle's case between SPRE/SOVL/SPOST.
// TF: how a run of noOverlapCheckgap steps is evaluated -- tf(side, seg, exclusive limit, L).
template <bool MERGE = true, class TF>
__device__ __forceinline__ void cheap_with(const int* hA, const int* hB, uint32_t m, Lane& L, int& shape, TF& tf) {
    const uint32_t op = m & 15u;
    if (op == M_NOP || L.gap || L.undef) return;
    const int x = (m >> 4) & 1, y = (m >> 5) & 1, ia = (m >> 6) & 15, ib = (m >> 10) & 15;
    switch (op) {
        case M_START: L.ref = bound(hA, hB, ia); break;
        case M_TLT: tf(x, y, bound(hA, hB, ia), L); break;
        case M_SEAT: {
            const int* h = x ? hB : hA;
            const int st = x ? 32 : 1;
            set_cur(L, x, y, get_cur(L, x, y) + h[(2 * y) * st] - h[(4 + 2 * y) * st]);
            break;
        }
        case M_FIN: {
            int which = -1;
            if (L.ref - 1 == bound(hA, hB, ia)) which = ib;
            else if (L.ref - 1 == bound(hA, hB, ib)) which = ia;
            if (which >= 0) tf((which >> 2) & 1, (which >> 1) & 1, bound(hA, hB, which) + 1, L);
            break;
        }
        case M_SPRE: {                               // checkGapsSingle (:269-395): shape, fresh cursors, leading tail
            const int xs1 = hA[4 + 2 * x], xe1 = hA[5 + 2 * x], xs2 = hB[(4 + 2 * y) * 32], xe2 = hB[(5 + 2 * y) * 32];
            if (xs1 <= xs2 && xe1 <= xe2) shape = 0;
            else if (xs1 <= xs2 && xe1 >= xe2) shape = 1;
            else if (xs1 >= xs2 && xe1 <= xe2) shape = 2;
            else shape = 3;
            set_cur(L, 0, x, 0);
            set_cur(L, 1, y, 0);
            if (MERGE) {   // one call site: lanes whose leading tail is on different sides walk them together
                const bool first = shape < 2;
                L.ref = first ? xs1 : xs2;
                tf(first ? 0 : 1, first ? x : y, first ? xs2 : xs1, L);
            } else if (shape < 2) { L.ref = xs1; tf(0, x, xs2, L); }
            else { L.ref = xs2; tf(1, y, xs1, L); }
            break;
        }
        case M_SPOST: {
            if (MERGE) {
                const bool onB = shape == 0 || shape == 2;
                const int eA = hA[5 + 2 * x], eB = hB[(5 + 2 * y) * 32];
                tf(onB ? 1 : 0, onB ? y : x, (onB ? eB : eA) + 1, L);
            } else if (shape == 0 || shape == 2) tf(1, y, hB[(5 + 2 * y) * 32] + 1, L);
            else tf(0, x, hA[5 + 2 * x] + 1, L);
            break;
        }
        default: break;
    }
}

struct TailO1 {            // the O(1) evaluation from prefix counts (no allele-frequency map)
    const Args& a; const int* hA; const int* hB; int rA, rB;
    __device__ __forceinline__ void operator()(int side, int seg, int limit, Lane& L) const { tail(a, hA, hB, rA, rB, side, seg, limit, L); }
};
__device__ __forceinline__ void cheap(const Args& a, const int* hA, const int* hB, int rA, int rB, uint32_t m, Lane& L, int& shape) {
    TailO1 tf{a, hA, hB, rA, rB};
    cheap_with(hA, hB, m, L, shape, tf);
}

// A run of noOverlapCheckgap steps (:234-267) when an allele-frequency map is set: prob0 is a running sum in column
// order (computeProb0, :133-143), so every tail column is visited (the O(1) evaluation above needs the ladder).
// The walk goes from one run of M ops to the next: inside a run a column is one addend (the map's entry at that position,
// or log10(0.25) where it has none -- tabulated in P.simpson_q) added in column order, with no record reads. Where the
// runs end is known from the header for "clips, plain M, clips" segments and from the prefix counts (aux) otherwise:
// ops [c, c+k) are all M iff both counters advance by k. Everything else (I, D, clips inside the range, positions
// outside the map) takes the single-column step.
struct TailWalk {
    const Args& a; const int* hA; const int* hB; const uint16_t* rawA; const uint16_t* rawB; double* prob0; int rA, rB;
    __device__ __forceinline__ void operator()(int side, int seg, int limit, Lane& L) const {
        if (limit - L.ref <= 0) return;
        const int* h = side ? hB : hA;
        const int st = side ? 32 : 1;
        const int n = h[(10 + seg) * st];
        int c = get_cur(L, side, seg);
        const unsigned fl = ((unsigned)h[15 * st] >> (16 * seg)) & 0xffffu;
        const int lead = h[(2 * seg) * st] - h[(4 + 2 * seg) * st];
        const int trail = h[(5 + 2 * seg) * st] - h[(1 + 2 * seg) * st];
        const uint16_t* raw = (side ? rawB : rawA) + (seg ? h[10 * st] : 0);
        const uint32_t* ax = a.aux + (size_t)(unsigned)h[(8 + seg) * st] + 2u * (unsigned)(side ? rB : rA) + seg;
        int ref = L.ref, tc = L.tc;
        double p0 = *prob0;
        while (ref < limit) {
            if (c >= n) { L.undef = true; break; }
            int k;                                               // length of the run of M ops at c, cut at the limit
            if (fl == 0) {
                if (c < lead || c >= n - trail) {                // a clip run: columns pass, nothing is added
                    const int s = min((c < lead ? lead : n) - c, limit - ref);
                    c += s; ref += s;
                    continue;
                }
                k = min(n - trail - c, limit - ref);
            } else {
                const int kmax = min(n - c, limit - ref);
                const uint32_t x0 = __ldg(ax + c);
                if (__ldg(ax + c + kmax) - x0 == (uint32_t)kmax * 0x10001u) k = kmax;
                else {
                    int lo = 0, hi = kmax;                       // ops [c, c+lo) are all M, [c, c+hi) are not
                    while (hi - lo > 1) {
                        const int mid = (lo + hi) >> 1;
                        if (__ldg(ax + c + mid) - x0 == (uint32_t)mid * 0x10001u) lo = mid; else hi = mid;
                    }
                    k = lo;
                }
            }
            if (k > 0 && ref >= 0 && ref + k <= a.P.simpson_len) {
                const double* mq = a.P.simpson_q + ref;
                int t = 0;
                if (ref & 1) { p0 = __dadd_rn(p0, __ldg(mq)); t = 1; }          // 16-byte loads from here on
                for (; t + 4 <= k; t += 4) {
                    const double2 u = __ldg(reinterpret_cast<const double2*>(mq + t));
                    const double2 v = __ldg(reinterpret_cast<const double2*>(mq + t + 2));
                    p0 = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(p0, u.x), u.y), v.x), v.y);
                }
                for (; t < k; ++t) p0 = __dadd_rn(p0, __ldg(mq + t));
                tc += k; c += k; ref += k;
                continue;
            }
            // one column, any op
            const int op = fl == 0 ? (int)HC_OP_M : (int)HC_REC_OP((uint32_t)__ldg(raw + c));
            if (op == HC_OP_M || op == HC_OP_I) {
                tc++;
                double d = 0.0;
                if (ref < 0 || ref >= a.P.simpson_len) {
                    if (!a.P.no_prob0) { L.undef = true; break; }       // the reference reads SIMPSON_MAP out of bounds here
                } else {
                    d = __ldg(a.P.simpson + ref);
                }
                p0 = __dadd_rn(p0, d != 0.0 ? d : a.P.log10_quarter);
            }
            if (op != HC_OP_I) ref++;
            c++;
        }
        L.ref = ref; L.tc = tc; *prob0 = p0;
        set_cur(L, side, seg, c);
    }
};

__device__ __forceinline__ bool consumes_q(uint32_t op) { return op == HC_OP_M || op == HC_OP_I || op == HC_OP_S; }

// Operand loads of the lock-step loop. G = false: the staged copies in shared memory (row buffer, transposed column
// tile: word w of the lane's read at opsB[w*32]). G = true ("streamed" kernel, reads too long or quality alphabets too
// wide for a tile): both op streams are read in place from global memory (L1/L2), word w at opsB[w].
template <bool G> __device__ __forceinline__ uint32_t ldw(const uint32_t* p) { return G ? __ldg(p) : *p; }
template <bool G> __device__ __forceinline__ uint32_t ldh(const uint16_t* p) { return G ? (uint32_t)__ldg(p) : (uint32_t)*p; }

// The lock-step loop of one slot, all 32 lanes together (lanes without an overlap phase in this slot idle).
template <bool G>
__device__ __forceinline__ void overlap(const Args& a, const int* hA, const int* hB, const uint16_t* opsA, const uint32_t* opsB,
                                        const unsigned char* Tb, uint32_t m, Lane& L, int shape) {
    const uint32_t op = m & 15u;
    bool more = (op == M_OVL || op == M_OVL2 || op == M_OVLQX || op == M_SOVL) && !L.gap && !L.undef;
    constexpr int BS = G ? 1 : 32;                  // stride of the lane's words in opsB
    const int sa = (m >> 4) & 1, sb = (m >> 5) & 1;
    const bool qx = op == M_OVLQX;
    int hi = 0;
    if (more) {
        if (op == M_SOVL) hi = (shape == 0 || shape == 2) ? hA[5 + 2 * sa] : hB[(5 + 2 * sb) * 32];
        else {
            hi = bound(hA, hB, (m >> 6) & 15);
            if (op != M_OVL) hi = min(hi, bound(hA, hB, (m >> 10) & 15));
        }
    }
    const int nA = hA[10 + sa], nB = hB[(10 + sb) * 32];
    const int aoff = sa ? hA[10] : 0, boff = sb ? hB[10 * 32] : 0;
    int ca = get_cur(L, 0, sa), cb = get_cur(L, 1, sb);
    int cq = 0;
    if (more && qx) { while (cq < nA && !consumes_q(HC_REC_OP(ldh<G>(opsA + aoff + cq)))) cq++; }
    const uint32_t K8 = (uint32_t)a.P.K * 8u;
    const uint32_t teq = smem_u32(Tb), tun = teq + (uint32_t)a.P.KK * 8u;   // shared-space addresses of the two tables
    const uint32_t* opsA32 = reinterpret_cast<const uint32_t*>(opsA);
    int ref = L.ref, cc = L.cc;
    double probM = L.probM;
    // probM *= T[neq][qa][qb]; the address is formed on the FMA pipe (two IMADs), the select is the only ALU op
#define TILE_MUL(QA, QB, NEQ)                                                             \
    do {                                                                                  \
        const uint32_t ad__ = (QB) * 8u + ((QA) * K8 + ((NEQ) ? tun : teq));              \
        double t__;                                                                       \
        asm("ld.shared.f64 %0, [%1];" : "=d"(t__) : "r"(ad__));                           \
        probM = __dmul_rn(probM, t__);                                                    \
    } while (0)
#define TILE_MUL2(AW, BW)                                                                 \
    do {                                                                                  \
        const uint32_t x__ = (AW) ^ (BW);                                                 \
        TILE_MUL(__byte_perm((AW), 0, 0x4441), __byte_perm((BW), 0, 0x4441), (x__ & 0x7u) != 0u);       \
        TILE_MUL((AW) >> 24, (BW) >> 24, (x__ & 0x70000u) != 0u);                         \
    } while (0)
    for (;;) {
        if (G) {
            // -- streamed operands: the lane's own stream comes in 16-byte loads (eight columns; a 4-byte load per lane costs the
            //    same 32 L1 wavefronts), the row's stream -- the same address in every lane -- in words. Single columns up to
            //    the lane's next 16-byte boundary, eight per iteration from there, single columns for the rest of the run.
            if (more && !qx) {
                const int run = min(min(hi - ref + 1, nA - ca), nB - cb);
                if (run >= 1) {
                    const int ia = aoff + ca, ib = boff + cb;
                    const uint16_t* qa16 = opsA + ia;
                    const uint16_t* qb16 = reinterpret_cast<const uint16_t*>(opsB) + ib;
                    int done = 0;
                    bool stop = false;
                    while (done < run && ((ib + done) & 7)) {
                        const uint32_t ra = ldh<true>(qa16 + done), rb = ldh<true>(qb16 + done);
                        if ((ra | rb) & 0x38u) { stop = true; break; }
                        TILE_MUL(ra >> 8, rb >> 8, ((ra ^ rb) & 0x7u) != 0u);
                        done++;
                    }
                    if (!stop && done + 8 <= run) {
                        const uint32_t* pa = opsA32 + ((ia + done) >> 1);
                        int sha = ((ia + done) & 1) * 16;
                        asm volatile("" : "+r"(sha));
                        const uint4* pb = reinterpret_cast<const uint4*>(qb16 + done);
                        uint32_t a0 = __ldg(pa);
                        uint4 v = __ldg(pb);
                        for (;;) {
                            const uint32_t a1 = __ldg(pa + 1), a2 = __ldg(pa + 2), a3 = __ldg(pa + 3), a4 = __ldg(pa + 4);
                            const bool again = done + 16 <= run;
                            uint4 vn = v;
                            if (again) vn = __ldg(pb + 1);
                            const uint32_t A01 = __funnelshift_r(a0, a1, sha), A23 = __funnelshift_r(a1, a2, sha);
                            const uint32_t A45 = __funnelshift_r(a2, a3, sha), A67 = __funnelshift_r(a3, a4, sha);
                            if ((A01 | A23 | A45 | A67 | v.x | v.y | v.z | v.w) & 0x00380038u) { stop = true; break; }
                            TILE_MUL2(A01, v.x);
                            TILE_MUL2(A23, v.y);
                            TILE_MUL2(A45, v.z);
                            TILE_MUL2(A67, v.w);
                            done += 8; pa += 4; pb += 1; a0 = a4; v = vn;
                            if (!again) break;
                        }
                    }
                    while (done < run) {
                        const uint32_t ra = ldh<true>(qa16 + done), rb = ldh<true>(qb16 + done);
                        if ((ra | rb) & 0x38u) break;
                        TILE_MUL(ra >> 8, rb >> 8, ((ra ^ rb) & 0x7u) != 0u);
                        done++;
                    }
                    ca += done; cb += done; ref += done; cc += done;
                }
            }
        } else
        // -- four (then two, then one) columns per iteration while both reads show M
        if (more && !qx) {
            const int run = min(min(hi - ref + 1, nA - ca), nB - cb);
            if (run >= 1) {
                const int ia = aoff + ca, ib = boff + cb;
                const uint32_t* pa = opsA32 + (ia >> 1);
                const uint32_t* pb = opsB + (ib >> 1) * BS;
                int sha = (ia & 1) * 16, shb = (ib & 1) * 16;
                asm volatile("" : "+r"(sha), "+r"(shb));          // keep the two shift amounts in registers
                uint32_t a0 = ldw<G>(pa), a1 = ldw<G>(pa + 1), a2 = ldw<G>(pa + 2);
                uint32_t b0 = ldw<G>(pb), b1 = ldw<G>(pb + BS), b2 = ldw<G>(pb + 2 * BS);
                int done = 0;
                while (done + 4 <= run) {
                    const uint32_t A01 = __funnelshift_r(a0, a1, sha), A23 = __funnelshift_r(a1, a2, sha);
                    const uint32_t B01 = __funnelshift_r(b0, b1, shb), B23 = __funnelshift_r(b1, b2, shb);
                    if ((A01 | B01 | A23 | B23) & 0x00380038u) break;
                    pa += 2; pb += 2 * BS;
                    a0 = a2; b0 = b2;
                    a1 = ldw<G>(pa + 1); a2 = ldw<G>(pa + 2); b1 = ldw<G>(pb + BS); b2 = ldw<G>(pb + 2 * BS);
                    TILE_MUL2(A01, B01);
                    TILE_MUL2(A23, B23);
                    done += 4;
                }
                uint32_t Aw = __funnelshift_r(a0, a1, sha), Bw = __funnelshift_r(b0, b1, shb);
                if (done + 2 <= run && !((Aw | Bw) & 0x00380038u)) {
                    TILE_MUL2(Aw, Bw);
                    done += 2;
                    Aw = __funnelshift_r(a1, a2, sha); Bw = __funnelshift_r(b1, b2, shb);
                }
                if (done + 1 <= run && !((Aw | Bw) & 0x38u)) {
                    TILE_MUL(__byte_perm(Aw, 0, 0x4441), __byte_perm(Bw, 0, 0x4441), ((Aw ^ Bw) & 0x7u) != 0u);
                    done += 1;
                }
                ca += done; cb += done; ref += done; cc += done;
            }
        }
        __syncwarp();
        // -- generic steps (overlapCheckgap, :145-232) until both reads show M again or the phase ends
        while (more) {
            if (ref > hi) { more = false; break; }
            if (ca >= nA || cb >= nB) { L.undef = true; more = false; break; }
            const uint32_t ra = ldh<G>(opsA + aoff + ca);
            const uint32_t wb = ldw<G>(opsB + ((boff + cb) >> 1) * BS);
            const uint32_t rb = ((boff + cb) & 1) ? (wb >> 16) : (wb & 0xffffu);
            const int oa = (int)HC_REC_OP(ra), ob = (int)HC_REC_OP(rb);
            const int act = (oa | ob) == 0 ? ACT_SCORE : op_action(oa, ob);
            if (act == ACT_SCORE) {
                uint32_t sq = ra;
                if (qx) {
                    if (cq >= nA) { L.undef = true; more = false; break; }
                    sq = ldh<G>(opsA + aoff + cq);
                }
                TILE_MUL(HC_REC_QUAL(sq), HC_REC_QUAL(rb), HC_REC_BASE(sq ^ rb) != 0u);
                cc++;
                if (oa == HC_OP_M) ref++;
                ca++; cb++;
                if (qx) { cq++; while (cq < nA && !consumes_q(HC_REC_OP(ldh<G>(opsA + aoff + cq)))) cq++; }
            } else if (act == ACT_SKIP) {
                ref++; ca++; cb++;
                if (qx && consumes_q(oa)) { cq++; while (cq < nA && !consumes_q(HC_REC_OP(ldh<G>(opsA + aoff + cq)))) cq++; }
            } else if (act == ACT_ADV_A) {
                ca++;
                if (qx) { cq++; while (cq < nA && !consumes_q(HC_REC_OP(ldh<G>(opsA + aoff + cq)))) cq++; }
            } else if (act == ACT_ADV_B) {
                cb++;
            } else {
                L.gap = true; more = false; break;
            }
            // hand back to the fast loop as soon as the next column is M against M again
            if (!qx && ref <= hi && ca < nA && cb < nB) {
                const uint32_t na = ldh<G>(opsA + aoff + ca);
                const uint32_t nw = ldw<G>(opsB + ((boff + cb) >> 1) * BS);
                const uint32_t nb = ((boff + cb) & 1) ? (nw >> 16) : (nw & 0xffffu);
                if (((na | nb) & 0x38u) == 0u) break;
            }
        }
        if (!__any_sync(FULL, more)) break;
    }
#undef TILE_MUL2
#undef TILE_MUL
    L.ref = ref; L.cc = cc; L.probM = probM;
    if (op == M_OVL || op == M_OVL2 || op == M_OVLQX || op == M_SOVL) { set_cur(L, 0, sa, ca); set_cur(L, 1, sb, cb); }
}

// checkGapsCigar dispatch (:1063-1074) + the layout choice of checkGapsPaired / checkGapsMixed -> program id.
__device__ __forceinline__ int pick_program(const int* hA, const int* hB, bool se1, bool se2) {
#define VA(f) hA[f]
#define VB(f) hB[(f) * 32]
    // raw: s1=0 e1=1 s2=2 e2=3; ext: +4
    if (se1 && se2) return TP_SINGLE00;
    if (!se1 && !se2) {
        const int As1 = VA(0), Ae1 = VA(1), As2 = VA(2), Ae2 = VA(3), xAs1 = VA(4), xAs2 = VA(6);
        const int Bs1 = VB(0), Be1 = VB(1), Bs2 = VB(2), Be2 = VB(3), xBs1 = VB(4), xBs2 = VB(6);
        if (Ae1 < Bs2 && As2 > Be1) return TP_TWO_SINGLES;                                  // :436
        if (xAs1 <= xBs1 && Ae1 >= Bs1 && As2 <= Be1 && Ae2 >= Bs2) return 0;
        if (xAs1 <= xBs1 && Ae1 >= Bs1 && As2 <= Be1 && Ae2 < Bs2) return 1;
        if (xBs1 <= xAs2 && Ae1 < Bs1 && As2 <= Be1 && Ae2 >= Bs2) return 2;
        if (xAs2 <= xBs1 && Ae1 < Bs1 && As2 <= Be1 && Ae2 >= Bs2) return 3;
        if (xAs1 <= xBs1 && As1 <= Be1 && Ae1 >= Bs2 && As2 <= Be2) return 4;
        if (xAs1 <= xBs1 && As2 > Be2 && Ae1 >= Bs2 && As1 <= Be1) return 5;
        if (xBs1 <= xAs1 && Ae1 >= Bs1 && As2 <= Be1 && Ae2 >= Bs2) return 6;
        if (xBs1 <= xAs1 && Ae1 >= Bs1 && As2 <= Be1 && Ae2 < Bs2) return 7;
        if (xAs1 <= xBs2 && As1 > Be1 && Ae1 >= Bs2 && As2 <= Be2) return 8;
        if (xBs2 <= xAs1 && Be1 < As1 && Ae1 >= Bs2 && As2 <= Be2) return 9;
        if (xBs1 <= xAs1 && As1 <= Be1 && Ae1 >= Bs2 && As2 <= Be2) return 10;
        if (xBs1 <= xAs1 && As2 > Be2 && As1 <= Be1 && Ae1 >= Bs2) return 11;
        return TP_NOP;                               // no layout matches: cc stays 0 (:841-844 under NDEBUG)
    }
    if (se2) {                                       // ap1 paired, ap2 single-end (:850-954)
        if (VB(1) < VA(2)) return TP_SINGLE00;
        if (VB(0) > VA(1)) return TP_SINGLE10;
        return PROG_PS0 + (VB(4) <= VA(4) ? 0 : 1);
    }
    if (VA(1) < VB(2)) return TP_SINGLE00;           // ap1 single-end, ap2 paired (:955-1056)
    if (VA(0) > VB(1)) return TP_SINGLE01;
    return PROG_SP0 + (VA(4) <= VB(4) ? 0 : 1);
#undef VA
#undef VB
}

__device__ __forceinline__ bool tile_layout_reject(const int* hA, const int* hB, bool se1, bool se2) {
    const int As1 = hA[0], Ae1 = hA[1], As2 = hA[2], Ae2 = hA[3];
    const int Bs1 = hB[0], Be1 = hB[32], Bs2 = hB[64], Be2 = hB[96];
    if (se1 && se2) return Ae1 < Bs1 || Be1 < As1;
    if (!se1 && !se2)
        return (Ae1 < Bs1 && Ae2 < Bs2) || (Be1 < As1 && Be2 < As2) || (Be1 < As2 && Ae2 < Bs2) || (Ae1 < Bs2 && Be2 < As2) ||
               (Be1 < As1 && Ae1 < Bs2) || (Ae1 < Bs1 && Be1 < As2);
    if (se1) return Ae1 < Bs1 || (Be1 < As1 && Ae1 < Bs2) || Be2 < As1;
    return Ae2 < Bs1 || (Ae1 < Bs1 && Be1 < As2) || Be1 < As1;
}

// G = true: the streamed variant -- no column tile and no row op buffers (a.maxw = 0, a.abuf_bytes = 0); the lock-step loop
// reads both op streams in place. Everything else (work items, headers, programs, tails, decision) is the same code.
// MAP = true: an allele-frequency map is set, the tails are walked (TailWalk) instead of evaluated in O(1).
// NW: warps per CTA as a compile-time constant (24, the usual launch) or 0 = read from the launch (a shrunken CTA).
template <bool G, bool MAP, int NW>
__global__ void __launch_bounds__(kThreads, HC_TILE_MINBLOCKS) score_tile_kernel(const Args a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwarps = NW ? NW : (int)(blockDim.x >> 5);          // the host shrinks the CTA (24 -> 8 warps) until it fits
    const int nthreads = nwarps * 32;
    const int K = a.P.K;
    // ---- shared memory carve-up (every piece a multiple of 16 bytes)
    size_t o = 0;
    double* Tsm = reinterpret_cast<double*>(smem + o);            o += ((size_t)2 * K * K * 8 + 15) & ~(size_t)15;
    uint32_t* progsm = reinterpret_cast<uint32_t*>(smem + o);     o += TP_COUNT * TP_WORDS * 4;
    int* hBsm = reinterpret_cast<int*>(smem + o);                 o += (size_t)(G ? nwarps : 1) * 16 * 32 * 4;
    uint32_t* tileB = reinterpret_cast<uint32_t*>(smem + o);      o += (size_t)a.maxw * 128;
    unsigned char* rowbuf = smem + o;                             o += (size_t)nwarps * 2 * (64 + a.abuf_bytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + o);       o += (size_t)nwarps * 2 * 8;
    unsigned long long* cstat = reinterpret_cast<unsigned long long*>(smem + o); o += 8 * 8;
    int* ctl = reinterpret_cast<int*>(smem + o);

    for (int k = tid; k < 2 * K * K; k += nthreads) Tsm[k] = a.P.table[k];
    for (int k = tid; k < TP_COUNT * TP_WORDS; k += nthreads) progsm[k] = a.prog[k];
    if (tid < 8) cstat[tid] = 0ull;
    unsigned a_total = 0, a_scored = 0, a_walked = 0, a_edges = 0, a_undef = 0;   // per-lane totals, reduced once at the end
    unsigned long long a_cc = 0, a_tc = 0;
    if (lane == 0) {
        mbar_init(&bars[warp * 2 + 0], 1);
        mbar_init(&bars[warp * 2 + 1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    const int bufstride = 64 + a.abuf_bytes;
    unsigned char* const mybuf0 = rowbuf + (size_t)(warp * 2) * bufstride;
    uint32_t parity = 0;       // bit k = phase parity of this warp's barrier k
    const unsigned char* Tb = reinterpret_cast<const unsigned char*>(Tsm);
    // G: nothing but the column headers is shared by the warps of a CTA, so every warp takes work items of its own (its
    // own copy of the headers, no CTA-wide barrier: rows of a super-read round differ in length by an order of magnitude)
    int* const hBmine = hBsm + (G ? warp * 16 * 32 : 0);
    const int* hB = hBmine + lane;

    auto load_info = [&](int row, int r1) {
        RowInfo ri;
        ri.ops_off = 0; ri.n_ops = 0; ri.col0 = 0; ri.word_off = 0;
        if (row < r1) {
            const int4 h2 = __ldg(reinterpret_cast<const int4*>(a.hdr + row) + 2);
            ri.ops_off = (unsigned)h2.x; ri.n_ops = h2.z + h2.w;
            ri.col0 = __ldg(a.col0 + (row - a.row_begin));
            ri.word_off = __ldg(a.word_off + (row - a.row_begin));
        }
        return ri;
    };
    auto issue = [&](int k, int row, const RowInfo& ri) {       // lane 0 only
        const uint32_t ob = G ? 0u : (((uint32_t)ri.n_ops * 2u + 15u) & ~15u);
        uint64_t* bar = &bars[warp * 2 + k];
        fence_proxy_async();
        mbar_expect_tx(bar, 64u + ob);
        bulk_g2s(mybuf0 + k * bufstride, a.hdr + row, 64u, bar);
        if (ob) bulk_g2s(mybuf0 + k * bufstride + 64, a.ops + ri.ops_off, ob, bar);
    };

    if (G) __syncthreads();                      // tables and programs are in place
    for (;;) {
        int item = 0;
        if (G) {
            if (lane == 0) item = atomicAdd(a.item_counter, 1);
            item = __shfl_sync(FULL, item, 0);
        } else {
            __syncthreads();
            if (tid == 0) ctl[0] = atomicAdd(a.item_counter, 1);
            __syncthreads();
            item = ctl[0];
        }
        if (item >= a.n_items) break;
        const int4 it = a.items[item];
        const int blk = it.x, r0 = it.y, r1 = it.z;
        const int jmine = __ldg(a.colidx + (size_t)blk * 32 + lane);       // this lane's column read (every warp agrees)
        // ---- stage the 32 column reads: headers transposed, op streams interleaved word-wise
        if (G) {
#pragma unroll
            for (int f4 = 0; f4 < 4; ++f4) {
                const int4 v = jmine >= 0 ? __ldg(reinterpret_cast<const int4*>(a.hdr + jmine) + f4) : make_int4(0, 0, 0, 0);
                hBmine[(4 * f4 + 0) * 32 + lane] = v.x; hBmine[(4 * f4 + 1) * 32 + lane] = v.y;
                hBmine[(4 * f4 + 2) * 32 + lane] = v.z; hBmine[(4 * f4 + 3) * 32 + lane] = v.w;
            }
        } else {
            for (int f = warp; f < 16; f += nwarps) hBsm[f * 32 + lane] = jmine >= 0 ? __ldg(reinterpret_cast<const int*>(a.hdr + jmine) + f) : 0;
        }
        const uint32_t* opsB = tileB + lane;
        const uint16_t* rawB = a.ops + (jmine >= 0 ? (unsigned)__ldg(reinterpret_cast<const int*>(a.hdr + jmine) + 8) : 0u);
        if (G) {
            opsB = reinterpret_cast<const uint32_t*>(rawB);
        } else {
            const int l = lane;
            const int j = jmine;
            if (j >= 0) {
                const int4 h2 = __ldg(reinterpret_cast<const int4*>(a.hdr + j) + 2);
                const int nchunk = (h2.z + h2.w + 7) >> 3;
                const uint4* src = reinterpret_cast<const uint4*>(a.ops + (unsigned)h2.x);
                for (int c = tid >> 5; c < nchunk; c += nwarps) {
                    const uint4 v = __ldg(src + c);
                    uint32_t* d = tileB + (size_t)(4 * c) * 32 + l;
                    d[0] = v.x; d[32] = v.y; d[64] = v.z; d[96] = v.w;
                }
            }
        }
        int next_row = r0;                           // G: the rows of the item are this warp's
        if (G) __syncwarp();
        else {
            if (tid == 0) ctl[1] = r0;
            __syncthreads();
        }

        // ---- rows: two grabbed ahead, the next row's bytes in flight while this one is walked
        auto grab = [&]() {
            if (G) return next_row++;
            int r = 0;
            if (lane == 0) r = atomicAdd(&ctl[1], 1);
            return __shfl_sync(FULL, r, 0);
        };
        int rowA = grab(), rowB = grab();
        RowInfo infA = load_info(rowA, r1), infB = load_info(rowB, r1);
        int kb = 0;
        if (rowA < r1 && lane == 0) issue(0, rowA, infA);
        while (rowA < r1) {
            if (rowB < r1 && lane == 0) issue(kb ^ 1, rowB, infB);
            const int rowC = grab();
            const RowInfo infC = load_info(rowC, r1);
            mbar_wait(&bars[warp * 2 + kb], (parity >> kb) & 1u);
            parity ^= 1u << kb;
            const int* hA = reinterpret_cast<const int*>(mybuf0 + kb * bufstride);
            const uint16_t* opsA = G ? a.ops + infA.ops_off : reinterpret_cast<const uint16_t*>(mybuf0 + kb * bufstride + 64);
            const int i = rowA;
            const int j = jmine;
            const bool active = j >= 0 && j < i;
            const bool covers = active && !(hA[14] < hB[13 * 32] || hB[14 * 32] < hA[13]);        // :1252
            if (__any_sync(FULL, covers)) {                         // some column read of the block reaches this row
                const int rel = j - infA.col0;                      // column inside the row's band (>= 0 whenever covers)
                int status = HC_ST_REJECT_COV;
                int pid = TP_NOP;
                const unsigned fa = (unsigned)hA[12], fb = (unsigned)hB[12 * 32];
                const bool se1 = fa & HC_FLAG_SINGLE_END, se2 = fb & HC_FLAG_SINGLE_END;
                if (covers) {
                    if (tile_layout_reject(hA, hB, se1, se2)) status = HC_ST_REJECT_LAYOUT;
                    else { status = HC_ST_WALKED; pid = pick_program(hA, hB, se1, se2); }
                }
                Lane L;
                L.ref = 0; L.cc = 0; L.tc = 0; L.probM = 1.0; L.curs = 0ull; L.gap = false; L.undef = false;
                int shape = 0;
                double prob0_run = 0.0;                             // MAP: the running sum of the tail columns
                TailWalk tail_walk{a, hA, hB, a.ops + infA.ops_off, rawB, &prob0_run, i, j};
                TailO1 tail_o1{a, hA, hB, i, j};
                auto cheap = [&](const Args&, const int*, const int*, int, int, uint32_t mo_, Lane& L_, int& shape_) {
                    if (MAP) cheap_with(hA, hB, mo_, L_, shape_, tail_walk);
                    else cheap_with<false>(hA, hB, mo_, L_, shape_, tail_o1);
                };
                if (__all_sync(FULL, pid == TP_SINGLE00 || pid == TP_NOP)) {
                    // the common case, checkGapsSingle on segments 1/1 for every lane: straight-line code
                    const bool mine = pid == TP_SINGLE00;
                    if (mine) cheap(a, hA, hB, i, j, TILE_MOP(M_SPRE, 0, 0, 0, 0), L, shape);
                    __syncwarp();
                    overlap<G>(a, hA, hB, opsA, opsB, Tb, mine ? TILE_MOP(M_SOVL, 0, 0, 0, 0) : 0u, L, shape);
                    if (mine) cheap(a, hA, hB, i, j, TILE_MOP(M_SPOST, 0, 0, 0, 0), L, shape);
                } else {
                    const uint32_t* pg = progsm + pid * TP_WORDS;
#pragma unroll 1
                    for (int s = 0; s < 3; ++s) {
                        cheap(a, hA, hB, i, j, pg[s * 4 + 0], L, shape);
                        cheap(a, hA, hB, i, j, pg[s * 4 + 1], L, shape);
                        cheap(a, hA, hB, i, j, pg[s * 4 + 2], L, shape);
                        __syncwarp();
                        const uint32_t mo = pg[s * 4 + 3];
                        if (__any_sync(FULL, (mo & 15u) != M_NOP)) overlap<G>(a, hA, hB, opsA, opsB, Tb, mo, L, shape);
                    }
                    cheap(a, hA, hB, i, j, pg[SLOT_FINAL], L, shape);
                }
                __syncwarp();
                // ---- similarityCriterion (:1107-1141)
                double prob0 = 0.0, score = __longlong_as_double(0x7ff8000000000000ll), cutoff = 0.0;
                bool edge = false, ambiguous = false;
                if (status == HC_ST_WALKED) {
                    if (!L.undef) {
                        if (MAP) prob0 = prob0_run;
                        else if (L.tc >= a.P.ladder_len) L.undef = true;
                        else prob0 = __ldg(a.P.ladder + L.tc);
                    }
                    if (L.undef) status = HC_ST_UNDEFINED;
                    else if (!L.gap && L.cc != 0) {
                        const bool cl1 = fa & HC_FLAG_IS_CLIQUE, cl2 = fb & HC_FLAG_IS_CLIQUE;
                        const int sel = (cl1 && cl2) ? 0 : ((cl1 || cl2) ? 1 : 2);
                        cutoff = a.P.log_cutoff[sel];
                        const double min_overlap = a.P.min_overlap[(cl1 && cl2) ? 0 : 1];
                        const unsigned smaller = min(fa & 0x3fffffffu, fb & 0x3fffffffu);
                        if (!((double)L.cc <= __dmul_rn(min_overlap, (double)smaller))) {
                            double prob, potence;
                            if (a.P.no_prob0) { prob = log10(L.probM); potence = __ddiv_rn(1.0, (double)L.cc); }
                            else { prob = __dadd_rn(log10(L.probM), prob0); potence = __ddiv_rn(1.0, (double)(L.cc + L.tc)); }
                            score = __dmul_rn(prob, potence);
                            edge = score >= cutoff;
                            ambiguous = fabs(score - cutoff) <= a.P.amb_tol * fabs(cutoff);
                        }
                    }
                }
                if (edge) atomicOr(&a.adj[infA.word_off + (rel >> 5)], 1u << (rel & 31));
                if (a.detail != nullptr && active && rel >= 0) {
                    hc_pair_result r;
                    r.status = status; r.edge = edge; r.gap = L.gap; r.cc = L.cc; r.tc = L.tc; r.ambiguous = ambiguous;
                    r.probM = L.probM; r.prob0 = prob0; r.score = score;
                    a.detail[infA.word_off * 32 + rel] = r;
                }
                if (ambiguous) {
                    const unsigned long long slot = atomicAdd(&a.stats->ambiguous, 1ull);
                    if (a.amb != nullptr && slot < (unsigned long long)a.amb_cap) {
                        hc_ambiguous am;
                        am.i = i; am.j = j; am.cc = L.cc; am.tc = L.tc; am.probM = L.probM; am.prob0 = prob0;
                        am.score = score; am.cutoff = cutoff; am.edge = edge; am.edge_device = edge;
                        a.amb[slot] = am;
                    }
                }
                const bool walked = status == HC_ST_WALKED;
                a_scored += (active && status != HC_ST_REJECT_COV) ? 1u : 0u;
                a_walked += walked ? 1u : 0u;
                a_edges += edge ? 1u : 0u;
                a_undef += status == HC_ST_UNDEFINED ? 1u : 0u;
                if (walked) { a_cc += (unsigned)L.cc; a_tc += (unsigned)L.tc; }
            }
            a_total += active ? 1u : 0u;
            __syncwarp();
            rowA = rowB; infA = infB; rowB = rowC; infB = infC; kb ^= 1;
        }
    }
    {
        const unsigned long long v[7] = {a_total, a_scored, a_walked, a_edges, a_cc, a_tc, a_undef};
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            unsigned long long x = v[k];
            for (int d = 16; d > 0; d >>= 1) x += __shfl_down_sync(FULL, x, d);
            if (lane == 0 && x) atomicAdd(&cstat[k], x);
        }
    }
    __syncthreads();
    if (tid < 7 && cstat[tid] != 0ull) {
        unsigned long long* dst = &a.stats->pairs_total;
        atomicAdd(dst + tid, cstat[tid]);
    }
}

// ---- per-segment prefix counts for segments that are not "clips, plain M, clips" ---------------------------
__global__ void build_aux_kernel(const hc_read_hdr* __restrict__ hdr, const uint16_t* __restrict__ ops, int n_reads,
                                 uint32_t* __restrict__ aux, uint16_t* __restrict__ col2op) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * n_reads) return;
    const int r = t >> 1, s = t & 1;
    const hc_read_hdr h = hdr[r];
    if (h.seg_flags[s] == 0 || h.n_ops[s] == 0) return;
    const unsigned off = h.ops_off[s];
    uint32_t* ax = aux + (size_t)off + 2u * (unsigned)r + s;
    uint32_t mi = 0, ni = 0;
    for (int c = 0; c < h.n_ops[s]; ++c) {
        ax[c] = (ni << 16) | mi;
        const int op = (int)HC_REC_OP((unsigned)ops[off + c]);
        if (op == HC_OP_M || op == HC_OP_I) mi++;
        if (op != HC_OP_I) { col2op[(size_t)off + ni] = (uint16_t)c; ni++; }
    }
    ax[h.n_ops[s]] = (ni << 16) | mi;
}

}  // namespace tile
