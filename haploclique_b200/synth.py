"""Seeded synthetic quasispecies read sets of the shapes BASELINE.json names.

Produces a position-sorted :class:`ReadBatch` directly (the records the reference's ``readBamFile`` would
hand to the clique finder): mates that overlap become one single-end record, the rest stay paired
(``AlignmentRecord::pairWith``, AlignmentRecord.cpp:251-312). Everything is vectorised numpy so that
10^5..10^6 fragments are generated in seconds; no file or network access.
"""
from __future__ import annotations

import numpy as np

from .records import ReadBatch

_OP_M, _OP_I, _OP_D, _OP_S = ord("M"), ord("I"), ord("D"), ord("S")
_BASES = np.frombuffer(b"ACGTN", dtype=np.uint8)

CONFIGS = {
    # name: kwargs of quasispecies()
    "C2": dict(n_frag=100_000, genome_len=9719, read_len=250, frag_mean=450.0, frag_sd=60.0, frag_min=260, frag_max=800),
    "C3": dict(n_frag=1_000_000, genome_len=9719, read_len=250, frag_mean=450.0, frag_sd=60.0, frag_min=260, frag_max=800),
    "C4": dict(n_frag=1_000_000, genome_len=1500, read_len=150, frag_mean=320.0, frag_sd=40.0, frag_min=160, frag_max=600),
}
# BASELINE configs[4]: a super-read round (kwargs of super_reads())
C5 = dict(n_frag=4000, genome_len=9719, read_len=3500, frag_mean=3000.0, frag_sd=2500.0, frag_min=500, frag_max=9719)


def quasispecies(n_frag: int, genome_len: int = 9719, read_len: int = 250, frag_mean: float = 450.0,
                 frag_sd: float = 60.0, frag_min: int = 260, frag_max: int = 800,
                 snv_rates=(0.01, 0.02, 0.03, 0.04, 0.05), hap_freq=(0.40, 0.25, 0.15, 0.12, 0.08),
                 indel_err: float = 1e-4, clip_frac: float = 0.01, n_frac: float = 5e-4,
                 seed: int = 20261104, frag_seed=None) -> ReadBatch:
    """`seed` fixes the reference and the haplotypes; `frag_seed` (default: continue the same stream) the fragments,
    their qualities and errors."""
    rng = np.random.default_rng(seed)
    G = int(genome_len)
    H = len(hap_freq)
    # -- haplotypes in reference-column space ---------------------------------------------------------
    ref = rng.integers(0, 4, G, dtype=np.uint8)
    hap_base = np.tile(ref, (H, 1))
    hap_del = np.zeros((H, G), bool)
    hap_ins_col = np.zeros(H, np.int64)      # insertion sits before this 1-based column
    hap_ins_len = np.zeros(H, np.int64)
    for h in range(H):
        snv = rng.random(G) < snv_rates[h]
        hap_base[h, snv] = (hap_base[h, snv] + rng.integers(1, 4, int(snv.sum()), dtype=np.uint8)) % 4
        if h > 0:  # haplotype 0 is the reference itself
            d0 = int(rng.integers(G // 8, G - G // 8))
            hap_del[h, d0:d0 + int(rng.integers(1, 4))] = True
            hap_ins_col[h] = int(rng.integers(G // 8, G - G // 8))
            hap_ins_len[h] = int(rng.integers(1, 4))
    hap_ins_bases = rng.integers(0, 4, (H, 3), dtype=np.uint8)
    if frag_seed is not None:
        rng = np.random.default_rng(frag_seed)

    # -- fragments -> segments -------------------------------------------------------------------------
    L = np.clip(np.rint(rng.normal(frag_mean, frag_sd, n_frag)), frag_min, min(frag_max, G)).astype(np.int64)
    s = rng.integers(1, G - L + 2)                         # 1-based start column
    hap = rng.choice(H, n_frag, p=np.asarray(hap_freq) / np.sum(hap_freq))
    order = np.argsort(s, kind="stable")                   # readBamFile sorts by interval start
    L, s, hap = L[order], s[order], hap[order]
    merged = L < 2 * read_len                              # mates overlap -> one single-end record
    nseg = np.where(merged, 1, 2)
    n = n_frag
    seg_read = np.repeat(np.arange(n), nseg)
    seg_idx = np.arange(seg_read.size) - np.repeat(np.cumsum(nseg) - nseg, nseg)      # 0 or 1
    seg_s = np.where(seg_idx == 0, s[seg_read], s[seg_read] + L[seg_read] - read_len)
    seg_len = np.where(merged[seg_read], L[seg_read], read_len)
    S = seg_read.size

    # -- one entry per reference column of every segment ------------------------------------------------
    tot = int(seg_len.sum())
    col_seg = np.repeat(np.arange(S), seg_len)
    seg_first = np.cumsum(seg_len) - seg_len
    off = np.arange(tot) - seg_first[col_seg]
    col = seg_s[col_seg] + off                              # 1-based reference column
    hp = hap[seg_read[col_seg]]
    base = hap_base[hp, col - 1]
    seglen_c = seg_len[col_seg]
    inner = (off > 0) & (off < seglen_c - 1)
    is_del = (hap_del[hp, col - 1] | (rng.random(tot) < indel_err / 2)) & inner
    # Illumina-like quality: Q38 -> Q22 along the read (mate 2 is read from the far end), +-4, merged
    # records get the better of the two mates' profiles.
    frac_fwd = off / np.maximum(seglen_c - 1, 1)
    is_m = merged[seg_read[col_seg]]
    frac = np.where(is_m, np.minimum(frac_fwd, 1.0 - frac_fwd) * seglen_c / read_len,
                    np.where(seg_idx[col_seg] == 0, frac_fwd, 1.0 - frac_fwd))
    q = np.clip(np.rint(38.0 - 16.0 * np.minimum(frac, 1.0)) + rng.integers(-4, 5, tot), 2, 41).astype(np.int64)
    err = rng.random(tot) < np.power(10.0, -q / 10.0)
    base = np.where(err, (base + rng.integers(1, 4, tot, dtype=np.uint8)) % 4, base).astype(np.uint8)
    base[rng.random(tot) < n_frac] = 4
    op = np.where(is_del, _OP_D, _OP_M).astype(np.uint8)

    # -- insertions: the haplotype's own + sequencing errors, placed before an inner column -------------
    ins_hap = inner & (hap_ins_len[hp] > 0) & (col == hap_ins_col[hp]) & ~is_del
    ins_err = inner & (rng.random(tot) < indel_err / 2) & ~ins_hap
    pos_list, base_list, q_list = [], [], []
    idx_h = np.nonzero(ins_hap)[0]
    for k in range(3):                                      # k-th inserted base of the haplotype insertion
        sel = idx_h[hap_ins_len[hp[idx_h]] > k]
        pos_list.append(sel)
        base_list.append(hap_ins_bases[hp[sel], k])
        q_list.append(q[sel])
    idx_e = np.nonzero(ins_err)[0]
    pos_list.append(idx_e)
    base_list.append(rng.integers(0, 4, idx_e.size, dtype=np.uint8))
    q_list.append(q[idx_e])
    ipos = np.concatenate(pos_list)
    o = np.argsort(ipos, kind="stable")
    ipos = ipos[o]
    op = np.insert(op, ipos, _OP_I)
    base = np.insert(base, ipos, np.concatenate(base_list)[o])
    q = np.insert(q, ipos, np.concatenate(q_list)[o])
    col_seg = np.insert(col_seg, ipos, col_seg[ipos])
    cig_len = np.bincount(col_seg, minlength=S).astype(np.int64)
    cig_off = np.cumsum(cig_len) - cig_len

    # -- soft clips on a fraction of segments, only where the clipped ops are plain M -------------------
    raw_s = seg_s.copy()
    raw_e = seg_s + seg_len - 1
    clip = np.nonzero(rng.random(S) < clip_frac)[0]
    klen = rng.integers(5, 31, clip.size)
    at_end = rng.random(clip.size) < 0.5
    for sg, k, e in zip(clip.tolist(), klen.tolist(), at_end.tolist()):
        lo = cig_off[sg] + (cig_len[sg] - k if e else 0)
        if k * 2 >= seg_len[sg] or not (op[lo:lo + k] == _OP_M).all() or op[lo + k if not e else lo - 1] != _OP_M:
            continue
        op[lo:lo + k] = _OP_S
        if e:
            raw_e[sg] -= k
        else:
            raw_s[sg] += k

    # -- assemble the batch -----------------------------------------------------------------------------
    b = ReadBatch(n)
    has_q = op != _OP_D
    b.cig = op
    b.seq = _BASES[base[has_q]]
    b.qual = (q[has_q] + 33).astype(np.uint8)
    seq_len = np.bincount(col_seg[has_q], minlength=S).astype(np.int64)
    seq_off = np.cumsum(seq_len) - seq_len
    first_seg = np.cumsum(nseg) - nseg
    last_seg = first_seg + nseg - 1
    for arr, src in ((b.seg_start, raw_s), (b.seg_end, raw_e), (b.cig_off, cig_off), (b.cig_len, cig_len),
                     (b.seq_off, seq_off), (b.seq_len, seq_len)):
        arr[2 * seg_read + seg_idx] = src
    single = np.nonzero(merged)[0]
    b.cig_off[2 * single + 1] = cig_off[first_seg[single]] + cig_len[first_seg[single]]
    b.seq_off[2 * single + 1] = seq_off[first_seg[single]] + seq_len[first_seg[single]]
    b.single_end[:] = merged
    b.ncov[:] = np.bincount(seg_read[col_seg[op == _OP_M]], minlength=n)
    b.cov_first[:] = raw_s[first_seg]        # segments begin and end on an M op (after clips)
    b.cov_last[:] = raw_e[last_seg]
    # leading clips move start1; readBamFile's sort is on the clipped start (haploclique.cpp:190-192)
    order = np.argsort(b.seg_start[0::2], kind="stable")
    if (order != np.arange(n)).any():
        b = b.permute(order)
    return b


def quasispecies_sharded(n_frag: int, chunk: int = 50_000, start_window=None, seed: int = 20261104, **kw) -> ReadBatch:
    """The same kind of read set generated in chunks of `chunk` fragments (chunk k uses seed+1000*k for its
    fragments; the haplotypes come from `seed` and are shared), merged and sorted by start. `start_window` =
    (lo, hi) keeps only reads with lo <= start1 < hi: every rank of a sharded run generates the same global
    read set and keeps its own genome-position window plus the halo to its left, so memory stays bounded by
    one chunk plus the kept reads."""
    parts = []
    k = 0
    left = int(n_frag)
    while left > 0:
        n = min(chunk, left)
        b = quasispecies(n, seed=seed, frag_seed=seed + 1000 * (k + 1), **kw)
        if start_window is not None:
            st = b.seg_start[0::2]
            keep = np.nonzero((st >= start_window[0]) & (st < start_window[1]))[0]
            b = b.permute(keep)
        parts.append(b)
        left -= n
        k += 1
    return ReadBatch.concat_sorted(parts)


def super_reads(n_frag: int = 4000, clique_frac: float = 0.7, seed: int = 20261104, **kw) -> ReadBatch:
    """A batch shaped like a later iteration's input (BASELINE configs[4]): a few thousand records of 0.5-9.7 kb, most of
    them merged cliques (name Clique_*: cutoffs 0.99 / 0.97 and min overlap 0.9 apply, NewEdgeCalculator.cpp:1111-1125),
    the rest left-over reads. Columns of a super-read carry posterior qualities raised by every agreeing member
    (computeEntry, AlignmentRecord.cpp:169-184: up to char 126), so the quality alphabet is 60-90 codes wide instead of
    the ~25 of raw reads -- which is what decides the kernel (tables of 16*K*K bytes, reads of thousands of columns)."""
    args = dict(C5)
    args.update(kw)
    args.pop("n_frag", None)
    b = quasispecies(n_frag, indel_err=2e-4, clip_frac=0.02, seed=seed, **args)
    rng = np.random.default_rng(seed + 7)
    b.is_clique[:] = (rng.random(b.n) < clique_frac).astype(b.is_clique.dtype)
    depth = np.where(b.is_clique != 0, rng.integers(5, 56, b.n), 0)           # quality raised by the clique's depth
    seg_len = np.asarray(b.seq_len, np.int64)
    seg_boost = np.repeat(depth, 2)
    idx = np.repeat(np.asarray(b.seq_off, np.int64) - (np.cumsum(seg_len) - seg_len), seg_len) + np.arange(int(seg_len.sum()))
    boost = np.repeat(seg_boost, seg_len)
    q = b.qual.astype(np.int64)
    q[idx] = np.minimum(q[idx] + np.where(boost > 0, boost + rng.integers(-4, 5, boost.size), 0), 126)
    b.qual = q.astype(np.uint8)
    return b


def zoo(n_reads: int, genome_len: int = 400, seed: int = 1, max_seg: int = 120, clique_frac: float = 0.3,
        wide_qual=(2, 94), indel: float = 0.02) -> ReadBatch:
    """Small adversarial batch for parity tests: single/paired mix, many start ties, leading/trailing S and
    H clips, insertions and deletions anywhere inside, N bases, wide quality range (a fifth of the segments draw
    Phred values from `wide_qual`, the rest from 20..41), 'Clique' names."""
    from .records import AlignmentRecord
    rng = np.random.default_rng(seed)
    recs = []

    def segment(start_ext: int, ncol: int):
        lead = int(rng.integers(0, 8)) if rng.random() < 0.3 else 0
        trail = int(rng.integers(0, 8)) if rng.random() < 0.3 else 0
        lead, trail = min(lead, ncol // 3), min(trail, ncol // 3)
        lead_op = b"S" if rng.random() < 0.7 else b"H"
        trail_op = b"S" if rng.random() < 0.7 else b"H"
        mid = []
        ncore = ncol - lead - trail
        for c in range(ncore):
            inner = 0 < c < ncore - 1
            if inner and rng.random() < indel:
                mid.append(b"I" * int(rng.integers(1, 3)))
            mid.append(b"D" if inner and rng.random() < indel else b"M")
        cig = lead_op * lead + b"".join(mid) + trail_op * trail
        nq = sum(1 for ch in cig if ch in b"MIS")
        seq = _BASES[rng.choice(5, nq, p=[0.245, 0.245, 0.245, 0.245, 0.02])].tobytes()
        qual = (rng.integers(wide_qual[0], wide_qual[1], nq) + 33).astype(np.uint8).tobytes() if rng.random() < 0.2 else \
            (rng.integers(20, 42, nq) + 33).astype(np.uint8).tobytes()
        return start_ext + lead, start_ext + ncol - trail - 1, cig, seq, qual

    # a shared "true" sequence makes many pairs similar enough to pass the thresholds
    truth = _BASES[rng.integers(0, 4, genome_len + 4 * max_seg + 64)]
    for r in range(n_reads):
        name = ("Clique_%d" % r) if rng.random() < clique_frac else ("read_%d" % r)
        s1 = int(rng.integers(1, genome_len)) // 4 * 4 + 1        # coarse grid -> many ties
        n1 = int(rng.integers(12, max_seg))
        st1, en1, c1, q1s, q1q = segment(s1, n1)
        if rng.random() < 0.5:
            rec = AlignmentRecord(name, True, st1, en1, c1, q1s, q1q)
        else:
            gap = int(rng.integers(1, max_seg))
            n2 = int(rng.integers(12, max_seg))
            st2, en2, c2, q2s, q2q = segment(en1 + gap + int(rng.integers(0, 8)), n2)
            rec = AlignmentRecord(name, False, st1, en1, c1, q1s, q1q, st2, en2, c2, q2s, q2q)
        # overwrite most bases with the shared truth so that scores are high
        for segname, st, cig in (("sequence1", rec.start1, rec.cigar1_unrolled),) + \
                ((("sequence2", rec.start2, rec.cigar2_unrolled),) if not rec.single_end else ()):
            seq = bytearray(getattr(rec, segname))
            lead = len(cig) - len(cig.lstrip(b"SH"))
            refp, qi = st - lead, 0
            for ch in cig:
                if ch in b"MS":
                    if rng.random() < 0.97 and refp > 0:
                        seq[qi] = int(truth[refp])
                    qi += 1
                    refp += 1
                elif ch in b"I":
                    qi += 1
                else:
                    refp += 1
            setattr(rec, segname, bytes(seq))
        recs.append(rec)
    recs.sort(key=lambda x: x.start1)
    return ReadBatch.from_records(recs)


def hard_clipped(n_reads: int, genome_len: int = 300, seed: int = 1) -> ReadBatch:
    """Single-end reads whose CIGARs carry hard clips outside soft clips (H S core S H) and insertions / deletions right
    behind or in front of the clip runs -- legal BAM that the other generators do not produce. The core starts and ends
    on M. Sorted by start like readBamFile's output."""
    from .records import AlignmentRecord
    rng = np.random.default_rng(seed)
    recs = []
    for k in range(n_reads):
        lead = b"H" * int(rng.integers(0, 3)) + b"S" * int(rng.integers(0, 6))
        trail = b"S" * int(rng.integers(0, 6)) + b"H" * int(rng.integers(0, 3))
        inner = [rng.choice([b"M", b"M", b"M", b"M", b"M", b"M", b"I", b"D"]) for _ in range(int(rng.integers(8, 90)))]
        if rng.random() < 0.5:
            inner = [b"M"] * len(inner)
        core = b"M" + b"".join(inner) + b"M"
        cig = lead + core + trail
        nq = sum(cig.count(c) for c in (b"M", b"I", b"S"))
        seq = bytes(rng.choice(list(b"ACGT" * 6 + b"N"), nq).tolist())
        qual = bytes(rng.integers(35, 75, nq).astype(np.uint8).tolist())
        ref_len = cig.count(b"M") + cig.count(b"D")
        start = int(rng.integers(1, genome_len - 100))
        name = ("Clique_%d" if rng.random() < 0.3 else "read_%d") % k
        recs.append(AlignmentRecord(name, True, start, start + ref_len - 1, cig, seq, qual))
    recs.sort(key=lambda r: r.start1)
    return ReadBatch.from_records(recs)
