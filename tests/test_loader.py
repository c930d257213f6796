"""CPU: the direct BAM -> SoA loader (hc_loader_* in libhcgpu.so, fed by integration/GpuBamLoader from the reference's
vendored BamTools reader) against the reference's own readBamFile + AlignmentRecord::pairWith
(oracle/_ref: hcref_load_bam) on the same files: same records, same fields, same arrival order (the unordered_map leftover
order and the unstable sort included), same max_position. Pins: the reference's unit1 test (test/unit/unit1.cpp:22-55: 1836
records on reads_HIV-1_50_01.bam; "No reads could be retrieved" on a BAM without M ops). No GPU is needed (host code)."""
import ctypes as C
import os

import numpy as np
import pytest

from haploclique_b200 import edge, synth
from haploclique_b200.records import ReadBatch
from oracle import pyref
from tests import common

pytestmark = pytest.mark.skipif(not pyref.gpu_adapter_available(), reason="oracle/_ref/libhcref_gpu.so not built (needs /root/reference)")


def direct_load(path):
    """(ReadBatch, max_position, counts, seconds) through GpuBamLoader / hc_loader_*."""
    L = pyref.lib_gpu()
    L.hcref_direct_load.restype = C.c_void_p
    L.hcref_direct_reads.restype = C.c_void_p
    L.hcref_direct_load.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.hcref_direct_reads.argtypes = [C.c_void_p]
    L.hcref_direct_name.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int]
    L.hcref_direct_free.argtypes = [C.c_void_p]
    n, mp, secs = C.c_int(0), C.c_uint(0), C.c_double(0)
    counts = (C.c_longlong * 3)()
    h = L.hcref_direct_load(path.encode(), C.byref(n), C.byref(mp), counts, C.byref(secs))
    if not h:
        return None
    rd = C.cast(L.hcref_direct_reads(h), C.POINTER(edge._Reads)).contents
    b = ReadBatch(rd.n)

    def arr(ptr, count, dtype):
        if count == 0:
            return np.zeros(0, dtype)
        return np.frombuffer((C.c_uint8 * (count * np.dtype(dtype).itemsize)).from_address(ptr), dtype=dtype).copy()

    nn = rd.n
    b.single_end, b.is_clique = arr(rd.single_end, nn, np.uint8), arr(rd.is_clique, nn, np.uint8)
    b.ncov, b.cov_first, b.cov_last = (arr(getattr(rd, k), nn, np.int32) for k in ("ncov", "cov_first", "cov_last"))
    b.seg_start, b.seg_end, b.cig_len, b.seq_len = (arr(getattr(rd, k), 2 * nn, np.int32) for k in ("seg_start", "seg_end", "cig_len", "seq_len"))
    b.cig_off, b.seq_off = arr(rd.cig_off, 2 * nn, np.int64), arr(rd.seq_off, 2 * nn, np.int64)
    tot_c = int(b.cig_off[-1] + b.cig_len[-1]) if nn else 0
    tot_s = int(b.seq_off[-1] + b.seq_len[-1]) if nn else 0
    b.cig, b.seq, b.qual = arr(rd.cig, tot_c, np.uint8), arr(rd.seq, tot_s, np.uint8), arr(rd.qual, tot_s, np.uint8)
    name = C.create_string_buffer(512)
    b.names = []
    for r in range(nn):
        L.hcref_direct_name(h, r, name, 512)
        b.names.append(name.value.decode())
    L.hcref_direct_free(h)
    return b, int(mp.value), [int(x) for x in counts], float(secs.value)


def reference_load(path):
    R = pyref.Reference()
    n = R.load_bam(path)
    return R.to_batch(), R.max_position(), n


def assert_same(got: ReadBatch, ref: ReadBatch):
    assert got.n == ref.n
    assert got.names == ref.names, "arrival order differs"
    for f in ("single_end", "is_clique", "ncov", "cov_first", "cov_last", "seg_start", "seg_end", "cig_len", "seq_len"):
        assert (getattr(got, f) == getattr(ref, f)).all(), f
    for r in range(ref.n):
        for s in range(1 if ref.single_end[r] else 2):
            k = 2 * r + s
            for stream, off, ln in (("cig", "cig_off", "cig_len"), ("seq", "seq_off", "seq_len"), ("qual", "seq_off", "seq_len")):
                a = getattr(got, stream)[getattr(got, off)[k]:getattr(got, off)[k] + getattr(got, ln)[k]]
                e = getattr(ref, stream)[getattr(ref, off)[k]:getattr(ref, off)[k] + getattr(ref, ln)[k]]
                assert (a == e).all(), "read %d (%s) segment %d: %s differs" % (r, ref.names[r], s, stream)


def test_bundled_hiv_reads_1836_records_like_the_references_unit1():
    path = os.path.join(common.GOLDEN, "reads_HIV-1_50_01.bam")
    got, mp, counts, _ = direct_load(path)
    ref, ref_mp, n = reference_load(path)
    assert got.n == n == 1836                                   # test/unit/unit1.cpp:30
    assert counts[2] > 1000 and got.single_end.all()           # every pair of mates overlaps and is merged into one record
    assert_same(got, ref)
    assert mp == ref_mp
    # and it is the batch the committed fixture holds (what the scorer tests run on)
    fix = common.load_batch("hiv50_records.npz")
    assert (got.cig == fix.cig).all() and (got.seq == fix.seq).all() and (got.qual == fix.qual).all()


def test_the_references_cram_input_paired_records_soft_clips_and_n_bases():
    path = os.path.join(common.GOLDEN, "arabis_large_500_01_reads.bam")
    got, mp, counts, _ = direct_load(path)
    ref, ref_mp, n = reference_load(path)
    assert got.n == n == 1104 and not got.single_end.any() and counts[2] == 0
    assert_same(got, ref)
    assert mp == ref_mp
    assert (got.seq == ord("N")).sum() == 211                    # 211 N bases in this file


@pytest.mark.parametrize("seed", [1, 2])
def test_synthetic_bams_mixed_pairs_leftovers_and_ties(tmp_path, seed):
    """Synthetic fragments written as a BAM by the reference's vendored BamWriter: overlapping mates (merged), separate
    mates (paired), single reads whose mate is missing (leftovers: unordered_map order), many equal starts (the unstable
    sort), clips and indels next to the overlap."""
    q = synth.quasispecies(1500, genome_len=1200, read_len=100, frag_mean=190.0, frag_sd=40.0, frag_min=110, frag_max=400, seed=seed,
                           indel_err=2e-3, clip_frac=0.2)
    # split every record into its mates again: single-end records of merged mates become one alignment; to exercise the mate
    # merge itself, cut long single-end records into two overlapping alignments of the same name
    from haploclique_b200.records import AlignmentRecord
    recs = []
    rng = np.random.default_rng(seed)
    for k, r in enumerate(q.records()):
        r.name = "frag_%d" % k
        if r.single_end and set(r.cigar1_unrolled) == {ord("M")} and len(r.cigar1_unrolled) > 140 and rng.random() < 0.7:
            n = len(r.cigar1_unrolled)
            a, b_ = int(n * 0.6), int(n * 0.35)
            first = AlignmentRecord(r.name, True, r.start1, r.start1 + a - 1, b"M" * a, r.sequence1[:a], r.qualities1[:a])
            second = AlignmentRecord(r.name, True, r.start1 + b_, r.end1, b"M" * (n - b_), r.sequence1[b_:], r.qualities1[b_:])
            lead = int(rng.integers(0, 6))
            if lead:                                                  # soft clips into the overlap
                second.cigar1_unrolled = b"S" * lead + second.cigar1_unrolled[lead:]
                second.start1 += lead
            recs += [first, second] if rng.random() < 0.8 else [second, first]
        elif not r.single_end and rng.random() < 0.15:
            recs.append(AlignmentRecord(r.name, True, r.start1, r.end1, r.cigar1_unrolled, r.sequence1, r.qualities1))   # mate lost
        else:
            recs.append(r)
    W = pyref.Reference()
    bam = str(tmp_path / "reads.bam")
    W.bam_writer_open(bam, "ref", 1500)
    # alignments in file order: mates of a pair need not be adjacent
    als = []
    for r in recs:
        als.append((r.name, r.start1, r.cigar1_unrolled, r.sequence1, r.qualities1))
        if not r.single_end:
            als.append((r.name, r.start2, r.cigar2_unrolled, r.sequence2, r.qualities2))
    for k in rng.permutation(len(als)):
        name, start, cig, seq, qual = als[int(k)]
        W.L.hcref_add_alignment(W.h, name.encode(), C.c_int(start - 1), cig, C.c_int(len(cig)), seq, qual, C.c_int(len(seq)))
    W.bam_writer_close()
    got = direct_load(bam)
    ref, ref_mp, n = reference_load(bam)
    assert got is not None
    assert_same(got[0], ref)
    assert got[1] == ref_mp and got[2][2] > 100 and (ref.single_end == 0).sum() > 100


def test_bam_without_m_ops_is_refused_like_the_reference(tmp_path):
    """test/unit/unit1.cpp:38-55: a BAM whose alignments have no M op -> "No reads could be retrieved from the BamFile. \""""
    from haploclique_b200.records import AlignmentRecord
    W = pyref.Reference()
    bam = str(tmp_path / "nom.bam")
    W.bam_writer_open(bam, "ref", 500)
    W.L.hcref_add_alignment(W.h, b"r1", C.c_int(10), b"SSSSS", C.c_int(5), b"ACGTA", b"IIIII", C.c_int(5))
    W.bam_writer_close()
    assert direct_load(bam) is None
    R = pyref.Reference()
    with pytest.raises(RuntimeError, match="No reads could be retrieved"):
        R.load_bam(bam)


def test_loader_is_faster_than_readBamFile_and_reports_per_read_time():
    path = os.path.join(common.GOLDEN, "reads_HIV-1_50_01.bam")
    L = pyref.lib_gpu()
    L.hcref_time_read_bam.restype = C.c_double
    n = C.c_int(0)
    ref_s = min(L.hcref_time_read_bam(path.encode(), C.byref(n)) for _ in range(3))
    ours = min(direct_load(path)[3] for _ in range(3))
    print("readBamFile %.1f ms, GpuBamLoader %.1f ms for %d records (%.2f / %.2f us per record)" % (ref_s * 1e3, ours * 1e3, n.value, ref_s / n.value * 1e6, ours / n.value * 1e6))
    assert n.value == 1836 and ours < ref_s
