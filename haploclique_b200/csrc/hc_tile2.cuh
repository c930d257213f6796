// hc_tile2.cuh -- second-generation tiled band kernel (included by hc_gpu.cu after hc_tile.cuh, whose work
// decomposition, O(1) tails, phase programs and decision code it shares).
//
// What changes against hc_tile.cuh is the operand format of the lock-step loop. ncu on the first tiled kernel
// (profiles/r01_v4_*): 15 warp instructions per shared column, 86 % of the issue slots and 70 % of the shared-memory
// wavefront slots busy; 8.5 of the 15 instructions only turn two 16-bit op records into a table address
// (two field extracts, the base compare, the select of the equal/unequal table, two multiply-adds).
// Here both streams are stored as *addends of the table's shared-memory address*, so that per column the loop is
//      address = A-word + B-halfword          (one IDP.2A: halfword extract and add in one instruction)
//      load the factor, one DMUL
//   * probability table, 9 rows per row-read quality qa:  rows 0..8 = (un, un, un, un, EQ, un, un, un, un).
//     A-word   = ROW * (qa*9 + 4 - baseA)                 (32 bits per op, built once per upload in HBM)
//     B-record = table base + ROW * baseB + ESZ * qb      (16 bits per op; the base is added while staging the tile)
//     sum      = address of row (qa*9 + 4 + baseB - baseA), entry qb: row 4 (equal) iff baseB == baseA -- the base
//     compare and the table select are absorbed into the address (bases A,C,G,T,N = 0..4). Any op other than M
//     sets the record's escape bit (A: bit 31, B: bit 15) and the column goes through the generic step
//     (overlapCheckgap restated, reading the raw records from global memory).
//   * two table layouts. SPLIT (K <= 32 quality codes): a row is 32 low words then 32 high words of the factors
//     (ROW = 256, ESZ = 4): two LDS.32 at [addr] and [addr+128]; the lanes of a warp (same qa: the row read is shared;
//     almost always the same base) read K <= 32 consecutive words = distinct banks, two wavefronts per column.
//     PAIR (any K): a row is K doubles (ROW = 8K, ESZ = 8), one LDS.64; 16 bank pairs for K codes.
//   * the row read is one aligned LDS.128 per four columns (broadcast when the lanes agree on the cursor, which
//     they do unless a column read has an indel), the column reads stay 16 bits per op in the transposed tile.
// Eight columns per iteration; the escape bits of all sixteen records are tested with seven LOP3.
#pragma once

namespace tile2 {

using namespace tile;   // Lane, RowInfo, cheap(), tail(), pick_program(), mbarrier / bulk-copy helpers, M_* / TP_* / SLOT_*

constexpr uint32_t A_ESC = 0x80000000u;
constexpr uint32_t B_ESC = 0x8000u;
constexpr int kMaxWarps = 24;
constexpr int kSplitMaxK = 0;    // split-word layout off by default: measured 67.2 ms vs 57.4 ms for doubles on C2 (rows of different bases share banks); HC_T2_SPLIT_MAXK=32 turns it on

struct Args2 : tile::Args {
    const uint32_t* opsA;       // row format, one 32-bit word per op (same indexing as ops)
    const uint16_t* opsF;       // column format, one 16-bit record per op (same indexing as ops), table base not yet added
    const uint32_t* tab9;       // the 9-row table image as it sits in shared memory (32-bit words)
    int tab_words;              // its size in 32-bit words
    uint32_t row_bytes;         // ROW: 256 (split) or 8*K (pair)
    uint32_t entry_bytes;       // ESZ: 4 (split) or 8 (pair)
};
constexpr int kRows = 9;

// ---- row / column formats, built on the device once per upload ---------------------------------------
struct ColumnOrder { uint8_t slot[128]; };   // position of quality code q inside a table row (a permutation of 0..K-1)

// Eight records per thread (one 16-byte load, three 16-byte stores); the slot table sits in shared memory -- indexed
// per thread from the kernel parameters it would serialise in the constant cache. n is a multiple of 8 (streams are padded).
__global__ void __launch_bounds__(256) build_formats_kernel(const uint16_t* __restrict__ ops, long long n, uint32_t ROW, uint32_t ESZ, ColumnOrder ord,
                                                            uint32_t* __restrict__ opsA, uint16_t* __restrict__ opsF) {
    __shared__ uint32_t slot_sm[128];
    if (threadIdx.x < 128) slot_sm[threadIdx.x] = ESZ * ord.slot[threadIdx.x];
    __syncthreads();
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g * 8 >= n) return;
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(ops) + g);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t A[8], F[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t f2 = 0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t r = (w[k] >> (16 * h)) & 0xffffu;
            const uint32_t base = HC_REC_BASE(r), op = HC_REC_OP(r), q = HC_REC_QUAL(r);
            const bool esc = op != HC_OP_M || base > 4u;
            A[2 * k + h] = esc ? A_ESC : ROW * (q * 9u + 4u - base);
            f2 |= ((esc ? B_ESC : ROW * base + slot_sm[q]) & 0xffffu) << (16 * h);
        }
        F[k] = f2;
    }
    uint4* oa = reinterpret_cast<uint4*>(opsA) + 2 * g;
    oa[0] = make_uint4(A[0], A[1], A[2], A[3]);
    oa[1] = make_uint4(A[4], A[5], A[6], A[7]);
    reinterpret_cast<uint4*>(opsF)[g] = make_uint4(F[0], F[1], F[2], F[3]);
}

// The lock-step loop of one slot (see tile::overlap for the contract; same results, different operand format).
//   opsA  : the row read's A-words in this warp's shared-memory row buffer (segment 1 then segment 2)
//   opsB  : the lane's column of the transposed tile (32-bit word w at opsB[w*32], two B-records per word)
//   rawA/rawB : the raw 16-bit records of the two reads in global memory, for the generic steps only
// A-word + B-record is the shared-space address of the factor (the tile's records carry the table base).
#ifdef HC_T2_NOINLINE_OVERLAP
#define HC_T2_OVL_INLINE __noinline__
#else
#define HC_T2_OVL_INLINE __forceinline__
#endif
template <bool SPLIT>
__device__ HC_T2_OVL_INLINE void overlap2(const Args2& a, const int* hA, const int* hB, const uint32_t* opsA, const uint32_t* opsB,
                                         const uint16_t* rawA, const uint16_t* rawB, uint32_t m, Lane& L, int shape) {
    const uint32_t op = m & 15u;
    bool more = (op == M_OVL || op == M_OVL2 || op == M_OVLQX || op == M_SOVL) && !L.gap && !L.undef;
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
    if (more && qx) { while (cq < nA && !consumes_q(HC_REC_OP((uint32_t)__ldg(rawA + aoff + cq)))) cq++; }
    int ref = L.ref, cc = L.cc;
    double probM = L.probM;
    // probM *= factor at shared-space address AD (A-word + B-record)
#define T2_MUL(AD)                                                                         \
    do {                                                                                   \
        const uint32_t ad__ = (AD);                                                        \
        double t__;                                                                        \
        if (SPLIT) {                                                                       \
            uint32_t lo__, hi__;                                                           \
            asm("ld.shared.u32 %0, [%1];" : "=r"(lo__) : "r"(ad__));                       \
            asm("ld.shared.u32 %0, [%1+128];" : "=r"(hi__) : "r"(ad__));                   \
            t__ = __hiloint2double((int)hi__, (int)lo__);                                  \
        } else {                                                                           \
            asm("ld.shared.f64 %0, [%1];" : "=d"(t__) : "r"(ad__));                        \
        }                                                                                  \
        probM = __dmul_rn(probM, t__);                                                     \
    } while (0)
    // two columns: A-words AW0, AW1 against the two records of BW
#define T2_MUL2(AW0, AW1, BW)                                                              \
    do {                                                                                   \
        T2_MUL(__dp2a_lo((BW), 0x00000001u, (AW0)));                                       \
        T2_MUL(__dp2a_hi((BW), 0x01000000u, (AW1)));                                       \
    } while (0)
#ifdef HC_T2_PROFILE
    int pass__ = 0;
#endif
    for (;;) {
#ifdef HC_T2_PROFILE
        pass__++;
#endif
        if (more && !qx) {
            const int run = min(min(hi - ref + 1, nA - ca), nB - cb);
            if (run >= 1) {
                const int ia = aoff + ca, ib = boff + cb;
                int done = 0;
                bool stop = false;
                // one column: stops at an escape (non-M op on either side)
#define T2_SINGLE()                                                                        \
                {                                                                          \
                    const uint32_t wa__ = opsA[ia + done];                                 \
                    const uint32_t ww__ = opsB[((ib + done) >> 1) * 32];                   \
                    const uint32_t hb__ = ((ib + done) & 1) ? (ww__ >> 16) : (ww__ & 0xffffu); \
                    if ((wa__ & A_ESC) | (hb__ & B_ESC)) stop = true;                      \
                    else { T2_MUL(wa__ + hb__); done++; }                                  \
                }
                while (!stop && done < run && ((ia + done) & 3)) T2_SINGLE();              // until the A cursor is 16-byte aligned
                if (!stop && done + 4 <= run) {
                    const uint4* pa = reinterpret_cast<const uint4*>(opsA + ia + done);
                    const uint32_t* pb = opsB + ((ib + done) >> 1) * 32;
                    int shb = ((ib + done) & 1) * 16;
                    asm volatile("" : "+r"(shb));
                    uint32_t b0 = pb[0];
                    while (done + 8 <= run) {
#ifdef HC_T2_PROFILE
                        {   // [0] iterations, [1] active lanes, [2]/[3] the same for second and later passes of the slot
                            const unsigned am = __activemask();
                            if ((int)(threadIdx.x & 31) == __ffs(am) - 1) {
                                atomicAdd(&a.stats[1].pairs_total + (pass__ > 1 ? 2 : 0), 1ull);
                                atomicAdd(&a.stats[1].pairs_total + (pass__ > 1 ? 3 : 1), (unsigned long long)__popc(am));
                            }
                        }
#endif
                        const uint4 A0 = pa[0], A1 = pa[1];
                        const uint32_t b1 = pb[32], b2 = pb[64], b3 = pb[96], b4 = pb[128];
                        const uint32_t B01 = __funnelshift_r(b0, b1, shb), B23 = __funnelshift_r(b1, b2, shb);
                        const uint32_t B45 = __funnelshift_r(b2, b3, shb), B67 = __funnelshift_r(b3, b4, shb);
                        const uint32_t ea = A0.x | A0.y | A0.z | A0.w | A1.x | A1.y | A1.z | A1.w;
                        const uint32_t eb = B01 | B23 | B45 | B67;
                        if ((ea & A_ESC) | (eb & (B_ESC | (B_ESC << 16)))) break;
                        T2_MUL2(A0.x, A0.y, B01);
                        T2_MUL2(A0.z, A0.w, B23);
                        T2_MUL2(A1.x, A1.y, B45);
                        T2_MUL2(A1.z, A1.w, B67);
                        pa += 2; pb += 128; b0 = b4; done += 8;
                    }
                    if (done + 4 <= run) {
                        const uint4 A0 = pa[0];
                        const uint32_t b1 = pb[32], b2 = pb[64];
                        const uint32_t B01 = __funnelshift_r(b0, b1, shb), B23 = __funnelshift_r(b1, b2, shb);
                        if (!(((A0.x | A0.y | A0.z | A0.w) & A_ESC) | ((B01 | B23) & (B_ESC | (B_ESC << 16))))) {
                            T2_MUL2(A0.x, A0.y, B01);
                            T2_MUL2(A0.z, A0.w, B23);
                            done += 4;
                        }
                    }
                }
                while (!stop && done < run) T2_SINGLE();
#undef T2_SINGLE
                ca += done; cb += done; ref += done; cc += done;
            }
        }
        __syncwarp();
        // -- generic steps (overlapCheckgap, :145-232) on the raw records until both reads show a plain column again
        while (more) {
            if (ref > hi) { more = false; break; }
            if (ca >= nA || cb >= nB) { L.undef = true; more = false; break; }
            const uint32_t ra = __ldg(rawA + aoff + ca);
            const uint32_t rb = __ldg(rawB + boff + cb);
            const int oa = (int)HC_REC_OP(ra), ob = (int)HC_REC_OP(rb);
            const int act = (oa | ob) == 0 ? ACT_SCORE : op_action(oa, ob);
            if (act == ACT_SCORE) {
                uint32_t sq = ra;
                if (qx) {
                    if (cq >= nA) { L.undef = true; more = false; break; }
                    sq = __ldg(rawA + aoff + cq);
                }
                const int idx = (int)HC_REC_QUAL(sq) * a.P.K + (int)HC_REC_QUAL(rb) + (HC_REC_BASE(sq ^ rb) ? a.P.KK : 0);
                probM = __dmul_rn(probM, __ldg(a.P.table + idx));
                cc++;
                if (oa == HC_OP_M) ref++;
                ca++; cb++;
                if (qx) { cq++; while (cq < nA && !consumes_q(HC_REC_OP((uint32_t)__ldg(rawA + aoff + cq)))) cq++; }
            } else if (act == ACT_SKIP) {
                ref++; ca++; cb++;
                if (qx && consumes_q(oa)) { cq++; while (cq < nA && !consumes_q(HC_REC_OP((uint32_t)__ldg(rawA + aoff + cq)))) cq++; }
            } else if (act == ACT_ADV_A) {
                ca++;
                if (qx) { cq++; while (cq < nA && !consumes_q(HC_REC_OP((uint32_t)__ldg(rawA + aoff + cq)))) cq++; }
            } else if (act == ACT_ADV_B) {
                cb++;
            } else {
                L.gap = true; more = false; break;
            }
            // hand back to the fast loop as soon as the next column is escape-free on both sides
            if (!qx && ref <= hi && ca < nA && cb < nB) {
                const uint32_t na = opsA[aoff + ca];
                const uint32_t nw = opsB[((boff + cb) >> 1) * 32];
                const uint32_t nb = ((boff + cb) & 1) ? (nw >> 16) : (nw & 0xffffu);
                if (!((na & A_ESC) | (nb & B_ESC))) break;
            }
        }
        if (!__any_sync(FULL, more)) break;
    }
#undef T2_MUL2
#undef T2_MUL
    L.ref = ref; L.cc = cc; L.probM = probM;
    if (op == M_OVL || op == M_OVL2 || op == M_OVLQX || op == M_SOVL) { set_cur(L, 0, sa, ca); set_cur(L, 1, sb, cb); }
}

// Shared memory (bytes): 4*tab_words table | programs | 16*32*4 column headers | maxw*128 tile | nwarps*2*(64+abuf) row
// buffers | nwarps*2*8 barriers | 64 statistics | 16 control.  hc_gpu.cu: tile2_geometry() computes the same sum.
template <bool SPLIT, bool MAP>
__global__ void __launch_bounds__(kMaxWarps * 32, 1) score_tile2_kernel(const Args2 a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nthreads = blockDim.x, nwarps = nthreads >> 5;
    size_t o = 0;
    uint32_t* Tsm = reinterpret_cast<uint32_t*>(smem + o);        o += ((size_t)a.tab_words * 4 + 15) & ~(size_t)15;
    uint32_t* progsm = reinterpret_cast<uint32_t*>(smem + o);     o += TP_COUNT * TP_WORDS * 4;
    int* hBsm = reinterpret_cast<int*>(smem + o);                 o += 16 * 32 * 4;
    uint32_t* tileB = reinterpret_cast<uint32_t*>(smem + o);      o += (size_t)a.maxw * 128;
    unsigned char* rowbuf = smem + o;                             o += (size_t)nwarps * 2 * (64 + a.abuf_bytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + o);       o += (size_t)nwarps * 2 * 8;
    unsigned long long* cstat = reinterpret_cast<unsigned long long*>(smem + o); o += 8 * 8;
    int* ctl = reinterpret_cast<int*>(smem + o);

    for (int k = tid; k < a.tab_words; k += nthreads) Tsm[k] = a.tab9[k];
    for (int k = tid; k < TP_COUNT * TP_WORDS; k += nthreads) progsm[k] = a.prog[k];
    if (tid < 8) cstat[tid] = 0ull;
    unsigned a_total = 0, a_scored = 0, a_walked = 0, a_edges = 0, a_undef = 0;
    unsigned long long a_cc = 0, a_tc = 0;
    if (lane == 0) {
        mbar_init(&bars[warp * 2 + 0], 1);
        mbar_init(&bars[warp * 2 + 1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    const int bufstride = 64 + a.abuf_bytes;
    unsigned char* const mybuf0 = rowbuf + (size_t)(warp * 2) * bufstride;
    uint32_t parity = 0;
    // the table is the first thing in shared memory: its address (1 KB on sm_100) plus the largest B-record must stay
    // below the records' escape bit
    const uint32_t tbase = smem_u32(Tsm);
    if (tbase + 4u * a.row_bytes + a.entry_bytes * (uint32_t)a.P.K >= B_ESC) __trap();
    const uint32_t tbase2 = tbase | (tbase << 16);
    const int* hB = hBsm + lane;
    const uint32_t* opsB = tileB + lane;

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
        const uint32_t ob = ((uint32_t)ri.n_ops * 4u + 15u) & ~15u;
        uint64_t* bar = &bars[warp * 2 + k];
        fence_proxy_async();
        mbar_expect_tx(bar, 64u + ob);
        bulk_g2s(mybuf0 + k * bufstride, a.hdr + row, 64u, bar);
        if (ob) bulk_g2s(mybuf0 + k * bufstride + 64, a.opsA + ri.ops_off, ob, bar);
    };

    for (;;) {
        __syncthreads();
        if (tid == 0) ctl[0] = atomicAdd(a.item_counter, 1);
        __syncthreads();
        const int item = ctl[0];
        if (item >= a.n_items) break;
        const int4 it = a.items[item];
        const int blk = it.x, r0 = it.y, r1 = it.z;
        const int jmine = __ldg(a.colidx + (size_t)blk * 32 + lane);
        // ---- stage the 32 column reads: headers transposed, column-format records interleaved word-wise
        for (int f = warp; f < 16; f += nwarps) hBsm[f * 32 + lane] = jmine >= 0 ? __ldg(reinterpret_cast<const int*>(a.hdr + jmine) + f) : 0;
        unsigned rawB_off = 0;
        if (jmine >= 0) {
            const int4 h2 = __ldg(reinterpret_cast<const int4*>(a.hdr + jmine) + 2);
            rawB_off = (unsigned)h2.x;
            const int nchunk = (h2.z + h2.w + 7) >> 3;
            const uint4* src = reinterpret_cast<const uint4*>(a.opsF + rawB_off);
            for (int c = warp; c < nchunk; c += nwarps) {
                const uint4 v = __ldg(src + c);
                uint32_t* d = tileB + (size_t)(4 * c) * 32 + lane;
                d[0] = v.x + tbase2; d[32] = v.y + tbase2; d[64] = v.z + tbase2; d[96] = v.w + tbase2;
            }
        }
        const uint16_t* rawB = a.ops + rawB_off;
        if (tid == 0) ctl[1] = r0;
        __syncthreads();

        auto grab = [&]() {
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
            const uint32_t* opsA = reinterpret_cast<const uint32_t*>(mybuf0 + kb * bufstride + 64);
            const uint16_t* rawA = a.ops + infA.ops_off;
            const int i = rowA;
            const int j = jmine;
            const bool active = j >= 0 && j < i;
            const bool covers = active && !(hA[14] < hB[13 * 32] || hB[14 * 32] < hA[13]);        // :1252
            if (__any_sync(FULL, covers)) {
                const int rel = j - infA.col0;
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
                TailWalk tail_walk{a, hA, hB, rawA, rawB, &prob0_run, i, j};
                TailO1 tail_o1{a, hA, hB, i, j};
                auto cheap = [&](const Args2&, const int*, const int*, int, int, uint32_t mo_, Lane& L_, int& shape_) {
                    if (MAP) cheap_with(hA, hB, mo_, L_, shape_, tail_walk);
                    else cheap_with(hA, hB, mo_, L_, shape_, tail_o1);
                };
                if (__all_sync(FULL, pid == TP_SINGLE00 || pid == TP_NOP)) {
                    const bool mine = pid == TP_SINGLE00;
                    if (mine) cheap(a, hA, hB, i, j, TILE_MOP(M_SPRE, 0, 0, 0, 0), L, shape);
                    __syncwarp();
                    overlap2<SPLIT>(a, hA, hB, opsA, opsB, rawA, rawB, mine ? TILE_MOP(M_SOVL, 0, 0, 0, 0) : 0u, L, shape);
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
                        if (__any_sync(FULL, (mo & 15u) != M_NOP)) overlap2<SPLIT>(a, hA, hB, opsA, opsB, rawA, rawB, mo, L, shape);
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
                        else if (!a.P.no_prob0 || a.detail != nullptr) prob0 = __ldg(a.P.ladder + L.tc);   // with no_prob0 only the detail output shows it
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
#ifndef HC_T2_NO_ADJ_WRITE      /* experiment: how much of the pass the scattered atomicOr costs (profiles/r02_adjacency_write_cost.txt) */
                if (edge) atomicOr(&a.adj[infA.word_off + (rel >> 5)], 1u << (rel & 31));
#endif
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
                        am.i = i; am.j = j; am.cc = L.cc; am.tc = L.tc; am.probM = L.probM;
                        am.prob0 = (!MAP && a.P.no_prob0 && a.detail == nullptr) ? __ldg(a.P.ladder + L.tc) : prob0;
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

}  // namespace tile2
