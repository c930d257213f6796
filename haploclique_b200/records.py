"""Host-side read model: the fields of the reference's ``AlignmentRecord`` that the edge-scoring
path reads (``/root/reference/src/AlignmentRecord.h:54-77``; getters ``AlignmentRecord.cpp:1703-1811``),
held as a struct-of-arrays batch in arrival (position-sorted) order.

Only what ``NewEdgeCalculator::edgeBetween`` touches is kept: per read ``single_end``, ``start/end`` x2,
whether the name contains ``"Clique"`` (``NewEdgeCalculator.cpp:1111``), ``cov_pos.size()`` and the first
and last covered reference position (``:1127``, ``:1252``); per segment the unrolled CIGAR, bases and
Phred+33 qualities.
"""
from __future__ import annotations

import dataclasses
from typing import Iterable, List, Optional, Sequence

import numpy as np

_VALID_OPS = frozenset(b"MIDSH")


@dataclasses.dataclass
class AlignmentRecord:
    """One read (or super-read). Mirrors the reference class of the same name for the scoring path."""

    name: str
    single_end: bool
    start1: int
    end1: int
    cigar1_unrolled: bytes
    sequence1: bytes
    qualities1: bytes
    start2: int = 0
    end2: int = 0
    cigar2_unrolled: bytes = b""
    sequence2: bytes = b""
    qualities2: bytes = b""

    # -- reference getters used by the scorer ------------------------------------------------------
    def isSingleEnd(self) -> bool:
        return self.single_end

    def isPairedEnd(self) -> bool:
        return not self.single_end

    def getName(self) -> str:
        return self.name

    def getIntervalStart(self) -> int:  # AlignmentRecord.cpp:1743
        return self.start1

    def getIntervalEnd(self) -> int:  # AlignmentRecord.cpp:1747
        return self.end1 if self.single_end else self.end2

    def covered(self):
        """(cov_pos.size(), front().ref, back().ref) as ``coveredPositions`` would give
        (AlignmentRecord.cpp:315-378): one entry per M op, segment 1 then segment 2."""
        n = 0
        first = last = None
        segs = [(self.start1, self.cigar1_unrolled)]
        if not self.single_end:
            segs.append((self.start2, self.cigar2_unrolled))
        for start, cig in segs:
            c = np.frombuffer(cig, dtype=np.uint8)
            is_m = c == ord("M")
            consumes = is_m | (c == ord("D"))
            if is_m.any():
                refpos = start + np.cumsum(consumes) - consumes
                mpos = refpos[is_m]
                if first is None:
                    first = int(mpos[0])
                last = int(mpos[-1])
                n += int(is_m.sum())
        if first is None:
            raise ValueError(f"read {self.name!r} covers no reference position (no M op)")
        return n, first, last

    @staticmethod
    def restoreCompleteAlignmentRecord(path: str) -> "AlignmentRecord":
        """Parses the reference's unit-test fixture layout written by ``saveCompleteAlignmentRecord``
        (AlignmentRecord.cpp:2183-2254), token by token like ``restoreCompleteAlignmentRecord`` (:2346-2427)."""
        with open(path, "rb") as fh:
            lines = fh.read().split(b"\n")
        pos = 0

        def tok() -> bytes:
            nonlocal pos
            while lines[pos].strip() == b"":
                pos += 1
            t = lines[pos].strip()
            pos += 1
            return t

        name = tok().decode()
        tok()  # phred_sum1
        start1, end1 = int(tok()), int(tok())
        n1 = int(tok())
        cig1 = tok()[:n1]
        tok(), tok()  # length_incl_deletions1 / longdeletions1
        seq1, qual1 = tok(), tok()
        single = int(tok()) != 0
        rec = AlignmentRecord(name, single, start1, end1, cig1, _norm_bases(seq1), qual1)
        if not single:
            tok()  # phred_sum2
            rec.start2, rec.end2 = int(tok()), int(tok())
            n2 = int(tok())
            rec.cigar2_unrolled = tok()[:n2]
            tok(), tok()
            rec.sequence2, rec.qualities2 = _norm_bases(tok()), tok()
        return rec


def _norm_bases(b: bytes) -> bytes:
    """ShortDnaSequence normalisation: upper-case, anything but ACGT becomes N (ShortDnaSequence.cpp:37-50)."""
    a = np.frombuffer(b.upper(), dtype=np.uint8).copy()
    ok = (a == 65) | (a == 67) | (a == 71) | (a == 84)
    a[~ok] = 78
    return a.tobytes()


class ReadBatch:
    """Struct-of-arrays batch of reads in arrival order (what the C ABI's ``hc_reads`` points at)."""

    def __init__(self, n: int):
        self.n = n
        self.single_end = np.zeros(n, np.uint8)
        self.is_clique = np.zeros(n, np.uint8)
        self.ncov = np.zeros(n, np.int32)
        self.cov_first = np.zeros(n, np.int32)
        self.cov_last = np.zeros(n, np.int32)
        # segment s of read r is row 2*r+s
        self.seg_start = np.zeros(2 * n, np.int32)
        self.seg_end = np.zeros(2 * n, np.int32)
        self.cig_off = np.zeros(2 * n, np.int64)
        self.cig_len = np.zeros(2 * n, np.int32)
        self.seq_off = np.zeros(2 * n, np.int64)
        self.seq_len = np.zeros(2 * n, np.int32)
        self.cig = np.zeros(0, np.uint8)
        self.seq = np.zeros(0, np.uint8)
        self.qual = np.zeros(0, np.uint8)
        self.names: Optional[List[str]] = None

    # -- construction --------------------------------------------------------------------------
    @staticmethod
    def from_records(recs: Sequence[AlignmentRecord]) -> "ReadBatch":
        b = ReadBatch(len(recs))
        cigs: List[bytes] = []
        seqs: List[bytes] = []
        quals: List[bytes] = []
        co = so = 0
        b.names = []
        for r, rec in enumerate(recs):
            b.single_end[r] = 1 if rec.single_end else 0
            b.is_clique[r] = 1 if "Clique" in rec.name else 0
            b.ncov[r], b.cov_first[r], b.cov_last[r] = rec.covered()
            b.names.append(rec.name)
            segs = [(rec.start1, rec.end1, rec.cigar1_unrolled, rec.sequence1, rec.qualities1)]
            if not rec.single_end:
                segs.append((rec.start2, rec.end2, rec.cigar2_unrolled, rec.sequence2, rec.qualities2))
            for s, (st, en, cg, sq, ql) in enumerate(segs):
                if len(sq) != len(ql):
                    raise ValueError(f"read {rec.name!r}: bases and qualities differ in length")
                k = 2 * r + s
                b.seg_start[k], b.seg_end[k] = st, en
                b.cig_off[k], b.cig_len[k] = co, len(cg)
                b.seq_off[k], b.seq_len[k] = so, len(sq)
                cigs.append(cg)
                seqs.append(sq)
                quals.append(ql)
                co += len(cg)
                so += len(sq)
            if rec.single_end:
                b.cig_off[2 * r + 1], b.seq_off[2 * r + 1] = co, so
        b.cig = np.frombuffer(b"".join(cigs), dtype=np.uint8).copy()
        b.seq = np.frombuffer(b"".join(seqs), dtype=np.uint8).copy()
        b.qual = np.frombuffer(b"".join(quals), dtype=np.uint8).copy()
        return b

    def record(self, r: int) -> AlignmentRecord:
        def seg(k):
            c0, cl = int(self.cig_off[k]), int(self.cig_len[k])
            s0, sl = int(self.seq_off[k]), int(self.seq_len[k])
            return (self.cig[c0:c0 + cl].tobytes(), self.seq[s0:s0 + sl].tobytes(), self.qual[s0:s0 + sl].tobytes())

        name = self.names[r] if self.names else ("Clique_%d" % r if self.is_clique[r] else "read_%d" % r)
        c1, s1, q1 = seg(2 * r)
        rec = AlignmentRecord(name, bool(self.single_end[r]), int(self.seg_start[2 * r]), int(self.seg_end[2 * r]), c1, s1, q1)
        if not rec.single_end:
            rec.start2, rec.end2 = int(self.seg_start[2 * r + 1]), int(self.seg_end[2 * r + 1])
            rec.cigar2_unrolled, rec.sequence2, rec.qualities2 = seg(2 * r + 1)
        return rec

    def records(self) -> List[AlignmentRecord]:
        return [self.record(r) for r in range(self.n)]

    def interval_end(self) -> np.ndarray:
        """getIntervalEnd per read (AlignmentRecord.cpp:1747)."""
        return np.where(self.single_end != 0, self.seg_end[0::2], self.seg_end[1::2])

    def max_position(self) -> int:
        """``max_position`` as readBamFile derives it (haploclique.cpp:195-198)."""
        return int(self.interval_end().max()) if self.n else 0

    def validate(self) -> None:
        bad = ~np.isin(self.cig, np.frombuffer(b"MIDSH", dtype=np.uint8))
        if bad.any():
            raise ValueError("unrolled CIGAR holds an op outside MIDSH: %r" % chr(int(self.cig[bad][0])))
        if self.qual.size and (self.qual.min() < 33 or self.qual.max() > 126):
            raise ValueError("quality character outside Phred+33 range 33..126")

    def select(self, idx: Iterable[int]) -> "ReadBatch":
        return ReadBatch.from_records([self.record(int(i)) for i in idx])

    def permute(self, order: np.ndarray) -> "ReadBatch":
        """Vectorised re-ordering / sub-setting of reads (new read k = old read order[k])."""
        order = np.asarray(order, np.int64)
        b = ReadBatch(order.shape[0])
        for f in ("single_end", "is_clique", "ncov", "cov_first", "cov_last"):
            setattr(b, f, getattr(self, f)[order].copy())
        segs = np.stack([2 * order, 2 * order + 1], axis=1).reshape(-1)
        b.seg_start, b.seg_end = self.seg_start[segs].copy(), self.seg_end[segs].copy()
        b.cig_len, b.seq_len = self.cig_len[segs].copy(), self.seq_len[segs].copy()
        b.cig_off = (np.cumsum(b.cig_len, dtype=np.int64) - b.cig_len).astype(np.int64)
        b.seq_off = (np.cumsum(b.seq_len, dtype=np.int64) - b.seq_len).astype(np.int64)

        def gather(src, old_off, lens, new_off):
            tot = int(lens.sum())
            if tot == 0:
                return np.zeros(0, src.dtype)
            idx = np.arange(tot, dtype=np.int64) + np.repeat(old_off - new_off, lens)
            return src[idx]

        b.cig = gather(self.cig, self.cig_off[segs], b.cig_len, b.cig_off)
        b.seq = gather(self.seq, self.seq_off[segs], b.seq_len, b.seq_off)
        b.qual = gather(self.qual, self.seq_off[segs], b.seq_len, b.seq_off)
        if self.names is not None:
            b.names = [self.names[int(i)] for i in order]
        return b

    @staticmethod
    def concat_sorted(parts: Sequence["ReadBatch"]) -> "ReadBatch":
        """Concatenates batches and re-sorts the reads by start1 (stable), the order readBamFile hands to the finder."""
        parts = [p for p in parts if p.n > 0]
        if not parts:
            return ReadBatch(0)
        b = ReadBatch(sum(p.n for p in parts))
        for f in ("single_end", "is_clique", "ncov", "cov_first", "cov_last", "seg_start", "seg_end", "cig_len", "seq_len",
                  "cig", "seq", "qual"):
            setattr(b, f, np.concatenate([getattr(p, f) for p in parts]))
        co = so = 0
        offs_c, offs_s = [], []
        for p in parts:
            offs_c.append(p.cig_off + co)
            offs_s.append(p.seq_off + so)
            co += int(p.cig.shape[0])
            so += int(p.seq.shape[0])
        b.cig_off, b.seq_off = np.concatenate(offs_c), np.concatenate(offs_s)
        if all(p.names is not None for p in parts):
            b.names = [nm for p in parts for nm in p.names]
        order = np.argsort(b.seg_start[0::2], kind="stable")
        return b.permute(order)

    # -- (de)serialisation for committed fixtures --------------------------------------------------
    _FIELDS = ("single_end", "is_clique", "ncov", "cov_first", "cov_last", "seg_start", "seg_end",
               "cig_off", "cig_len", "seq_off", "seq_len", "cig", "seq", "qual")

    def save(self, path: str) -> None:
        arrs = {f: getattr(self, f) for f in self._FIELDS}
        if self.names is not None:
            arrs["names"] = np.array(self.names)
        np.savez_compressed(path, **arrs)

    @staticmethod
    def load(path: str) -> "ReadBatch":
        z = np.load(path, allow_pickle=False)
        b = ReadBatch(int(z["single_end"].shape[0]))
        for f in ReadBatch._FIELDS:
            setattr(b, f, np.ascontiguousarray(z[f]))
        if "names" in z.files:
            b.names = [str(s) for s in z["names"]]
        return b
