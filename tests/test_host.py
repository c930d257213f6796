"""CPU: host-side logic -- record model, C++ packer behind the C ABI, exported symbols. No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from haploclique_b200 import edge, synth
from haploclique_b200.records import AlignmentRecord, ReadBatch
from tests import common

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "hc_edge.h")).read()
    body = hdr[hdr.index("typedef struct hc_ctx hc_ctx;"):]
    names = set(re.findall(r"\b(hc_[a-z_0-9]+)\s*\(", body))
    assert {"hc_create", "hc_destroy", "hc_pack", "hc_upload", "hc_score_band", "hc_score_pairs", "hc_band_layout",
            "hc_last_error"} <= names
    L = edge.load_library()
    for n in sorted(names):
        assert hasattr(L, n), "libhcgpu.so does not export " + n
    assert L.hc_abi_version() == 2


def test_struct_sizes_match_header():
    assert edge.HDR_DTYPE.itemsize == 64
    assert edge.PAIR_RESULT_DTYPE.itemsize == 48
    assert edge.AMBIGUOUS_DTYPE.itemsize == 56


def _python_pack(b: ReadBatch, qual_chars: bytes):
    code = {c: k for k, c in enumerate(qual_chars)}
    opc = {ord("M"): 0, ord("I"): 1, ord("D"): 2, ord("S"): 3, ord("H"): 4}
    bc = {ord(ch): v for v, pair in enumerate(("Aa", "Cc", "Gg", "Tt")) for ch in pair}   # toupper, else 'N' (src/ShortDnaSequence.cpp:38-49)
    out = []
    for r in range(b.n):
        for s in range(1 if b.single_end[r] else 2):
            k = 2 * r + s
            cig = b.cig[b.cig_off[k]:b.cig_off[k] + b.cig_len[k]]
            seq = b.seq[b.seq_off[k]:b.seq_off[k] + b.seq_len[k]]
            ql = b.qual[b.seq_off[k]:b.seq_off[k] + b.seq_len[k]]
            qi = 0
            for c in cig:
                rec = opc[int(c)] << 3
                if int(c) in (ord("M"), ord("I"), ord("S")):
                    rec |= bc.get(int(seq[qi]), 4) | (code[int(ql[qi])] << 8)
                    qi += 1
                out.append(rec)
    return np.array(out, np.uint16)


@pytest.mark.parametrize("fixture", ["unit_fixtures.npz", "superreads_arabis_it2.npz"])
def test_pack_matches_python_restatement(fixture):
    b = common.load_batch(fixture)
    pk = edge.pack_only(b)
    assert pk.n_reads == b.n and pk.n_ops_total >= int(b.cig_len.sum())
    qc = bytes(pk.c.qual_chars[:pk.c.n_qual_codes])
    assert qc == bytes(sorted(set(b.qual.tolist())))
    assert (pk.ops() == _python_pack(b, qc)).all()
    h = pk.headers()
    assert (h["raw"][:, 0] == b.seg_start[0::2]).all() and (h["raw"][:, 1] == b.seg_end[0::2]).all()
    assert ((h["ncov_flags"] & 0x3fffffff) == b.ncov).all()
    assert (((h["ncov_flags"] >> 30) & 1) == b.single_end).all()
    assert (((h["ncov_flags"] >> 31) & 1) == b.is_clique).all()
    for r in range(b.n):
        cig = b.cig[b.cig_off[2 * r]:b.cig_off[2 * r] + b.cig_len[2 * r]].tobytes()
        lead = len(cig) - len(cig.lstrip(b"SH"))
        trail = len(cig) - len(cig.rstrip(b"SH"))
        assert h["ext"][r, 0] == b.seg_start[2 * r] - lead and h["ext"][r, 1] == b.seg_end[2 * r] + trail
        assert (h["seg_flags"][r, 0] & 1) == (1 if b"I" in cig else 0)
        core = cig[lead:len(cig) - trail]
        assert ((h["seg_flags"][r, 0] >> 1) & 1) == (0 if core == b"M" * len(core) else 1)
        assert h["ops_off"][r, 0] % 8 == 0 and h["ops_off"][r, 1] == h["ops_off"][r, 0] + h["n_ops"][r, 0]


def test_pack_adversarial_cigars_match_python_restatement():
    """The packer's look-up-table record build against the plain restatement on CIGARs the synthetic sets do not hold:
    hard clips outside soft clips, insertions and deletions next to the clips, lower-case and ambiguous bases, every quality character 33..126."""
    rng = np.random.default_rng(11)
    recs = []
    for k in range(300):
        lead = b"H" * int(rng.integers(0, 3)) + b"S" * int(rng.integers(0, 5))
        trail = b"S" * int(rng.integers(0, 5)) + b"H" * int(rng.integers(0, 3))
        core = b"".join(rng.choice([b"M", b"M", b"M", b"M", b"I", b"D"]) for _ in range(int(rng.integers(0, 60))))
        if core.count(b"M") == 0:
            core += b"M"                                                   # a record without an M op covers nothing: not a read
        cig = lead + core + trail
        nq = sum(cig.count(c) for c in (b"M", b"I", b"S"))
        seq = bytes(rng.choice(list(b"ACGTNacgtRY"), nq).tolist())
        qual = bytes(rng.integers(33, 127, nq).astype(np.uint8).tolist())
        ref_len = cig.count(b"M") + cig.count(b"D")
        start = int(rng.integers(1, 500))
        recs.append(AlignmentRecord("Clique_%d" % k if k % 3 == 0 else "read_%d" % k, True, start, start + max(ref_len, 1) - 1, cig, seq, qual))
    b = ReadBatch.from_records(recs)
    pk = edge.pack_only(b)
    qc = bytes(pk.c.qual_chars[:pk.c.n_qual_codes])
    assert qc == bytes(sorted(set(b.qual.tolist())))
    assert (pk.ops() == _python_pack(b, qc)).all()
    h = pk.headers()
    assert pk.c.max_ops_per_read == int(b.cig_len.reshape(-1, 2).sum(1).max())
    for r in range(b.n):
        cig = recs[r].cigar1_unrolled
        lead = len(cig) - len(cig.lstrip(b"SH"))
        trail = len(cig) - len(cig.rstrip(b"SH"))
        assert h["ext"][r, 0] == b.seg_start[2 * r] - lead and h["ext"][r, 1] == b.seg_end[2 * r] + trail
        assert (h["seg_flags"][r, 0] & 1) == (1 if b"I" in cig else 0)
        core = cig[lead:len(cig) - trail] if lead + trail <= len(cig) else b"?"
        assert ((h["seg_flags"][r, 0] >> 1) & 1) == (0 if core == b"M" * len(core) else 1), (r, cig)
        assert h["n_ops"][r, 0] == len(cig) and h["ops_off"][r, 0] % 8 == 0


def test_pack_rejects_malformed_reads():
    good = AlignmentRecord("r", True, 1, 4, b"MMMM", b"ACGT", b"IIII")
    for bad in (AlignmentRecord("r", True, 1, 4, b"MMNM", b"ACGT", b"IIII"),       # op outside MIDSH
                AlignmentRecord("r", True, 1, 4, b"MMMM", b"ACGT", b"II I"),       # quality below '!'
                AlignmentRecord("r", True, 1, 4, b"MMMMM", b"ACGT", b"IIII")):     # CIGAR longer than the query
        with pytest.raises(edge.HcError):
            edge.pack_only(ReadBatch.from_records([good, bad]))
    assert edge.pack_only(ReadBatch.from_records([good])).n_reads == 1
    assert edge.pack_only(ReadBatch.from_records([])).n_reads == 0


def test_fixture_parser_matches_reference_dump():
    # unit_fixtures.npz was dumped from the reference's own restoreCompleteAlignmentRecord; our parser of the
    # same text format must agree when the files are reachable (build container only)
    path = "/root/reference/test/data/simulation/unit_data/alignment_sample%02d.txt"
    if not os.path.exists(path % 0):
        pytest.skip("reference fixtures not on this machine")
    b = common.load_batch("unit_fixtures.npz")
    mine = ReadBatch.from_records([AlignmentRecord.restoreCompleteAlignmentRecord(path % k) for k in range(10)])
    for f in ReadBatch._FIELDS:
        assert (getattr(b, f) == getattr(mine, f)).all(), f


def test_synthetic_generator_is_deterministic_and_valid():
    a = synth.quasispecies(300, genome_len=1500, seed=7, indel_err=2e-3, clip_frac=0.1)
    b = synth.quasispecies(300, genome_len=1500, seed=7, indel_err=2e-3, clip_frac=0.1)
    for f in ReadBatch._FIELDS:
        assert (getattr(a, f) == getattr(b, f)).all()
    a.validate()
    assert (np.diff(a.seg_start[0::2]) >= 0).all()                      # arrival order = sorted by start
    rb = ReadBatch.from_records(a.records())                           # cov fields agree with coveredPositions
    for f in ("ncov", "cov_first", "cov_last", "seg_start", "seg_end", "cig_len", "seq_len"):
        assert (getattr(a, f) == getattr(rb, f)).all(), f
    paired = a.single_end == 0
    assert paired.any() and (~paired).any()
    assert (a.seg_start[1::2][paired] > a.seg_end[0::2][paired]).all()  # pairWith's "no overlap" branch


def test_scoring_without_device_fails_loudly():
    if edge.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(edge.HcError, match="no CUDA device"):
        edge.NewEdgeCalculator()
