"""CPU: oracle/merge_oracle.c (single-end super-read merge) against the reference's golden digests
(tests/golden/merge_golden.json, made by tests/golden/make_merge_golden.py from the compiled reference) and, where
oracle/_ref/libhcref.so is present, against the reference directly."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import pyoracle, pyref
from tests import merge_cases
from tests.common import GOLDEN


def _golden():
    with open(os.path.join(GOLDEN, "merge_golden.json")) as fh:
        return json.load(fh)


def _digest(rows):
    h = hashlib.sha256()
    for start, end, cig, seq, qual in rows:
        h.update(np.array([start, end, len(cig), len(seq)], "<i4").tobytes())
        h.update(cig); h.update(seq); h.update(qual)
    return h.hexdigest()


def test_compute_entry_table_matches_reference_digest():
    tab = bytearray()
    for q1 in range(33, 127):
        for q2 in range(33, 127):
            for b2 in (65, 67):
                tab += bytes(pyoracle.compute_entry(65, q1, b2, q2))
    assert hashlib.sha256(bytes(tab)).hexdigest() == _golden()["compute_entry_sha256"]


def test_compute_entry_known_values():
    # equal bases raise the quality, unequal bases keep the better base and lower it; 126 is the cap (:173-180)
    assert pyoracle.compute_entry(ord("A"), 33 + 30, ord("A"), 33 + 30) == (ord("A"), 33 + 65)
    assert pyoracle.compute_entry(ord("A"), 33 + 30, ord("C"), 33 + 10)[0] == ord("A")
    assert pyoracle.compute_entry(ord("A"), 33 + 10, ord("C"), 33 + 30)[0] == ord("C")
    assert pyoracle.compute_entry(ord("G"), 126, ord("G"), 126) == (ord("G"), 126)
    # float rounding pushes agreement(Q70, Q0) to -4: a char below '!' (29), which later folds index the tables with
    assert pyoracle.compute_entry(ord("T"), 33 + 70, ord("T"), 33) == (ord("T"), 29)
    assert pyoracle.compute_entry(ord("T"), 29, ord("T"), 33 + 40) == (ord("T"), 33)


def unit3_expected():
    """(batch, members, expected fields): the reference's own test vector test/unit/unit3.cpp:33-60 -- sample00 merged with
    sample10 must equal sample11 (rows 0, 1, 2 of tests/golden/unit3_merge_fixtures.npz)."""
    from tests.common import load_batch
    b = load_batch("unit3_merge_fixtures.npz")
    k = 4
    exp = (int(b.seg_start[k]), int(b.seg_end[k]), b.cig[b.cig_off[k]:b.cig_off[k] + b.cig_len[k]].tobytes(),
           b.seq[b.seq_off[k]:b.seq_off[k] + b.seq_len[k]].tobytes(), b.qual[b.seq_off[k]:b.seq_off[k] + b.seq_len[k]].tobytes())
    return b, [0, 1], exp


def test_merge_oracle_reproduces_the_references_unit3_vector():
    b, members, exp = unit3_expected()
    assert b.single_end[:3].all() and b.names[2].startswith("Clique")
    st, *got = pyoracle.merge_clique_se(b, members)
    assert st == 0 and tuple(got) == exp


@pytest.mark.parametrize("case", ["hiv50", "hiv50_large_cliques", "clipped"])
def test_merge_oracle_matches_reference_digest(case):
    g = _golden()[case]
    name, batch, seed, n_cl, span, mm = next(c for c in merge_cases.CASES() if c[0] == case)
    rows, undefined = [], []
    for t, mem in enumerate(merge_cases.windows(batch.n, n_cl, seed, span, mm)):
        st, s, e, cig, seq, qual = pyoracle.merge_clique_se(batch, mem)
        if st:
            undefined.append(t)
        else:
            rows.append((s, e, cig, seq, qual))
    assert undefined == g["undefined"]
    assert _digest(rows) == g["sha256"]


@pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libhcref.so not built (needs /root/reference)")
def test_merge_oracle_matches_reference_directly():
    batch = merge_cases.clipped_batch(seed=9, n=250)
    R = pyref.Reference()
    R.add_batch(batch)
    n = 0
    for t, mem in enumerate(merge_cases.windows(batch.n, 600, 3, 8, 6)):
        o = pyoracle.merge_clique_se(batch, mem)
        if o[0]:
            continue
        r = R.merge_clique(mem, t)
        assert r[0] and r[1:] == o[1:], "clique %d %s" % (t, mem)
        n += 1
    assert n > 400


def test_merge_has_no_cpu_path():
    """Without a CUDA device the merger cannot be created (HC_E_CUDA): there is no CPU merging path in the product."""
    import torch
    from haploclique_b200 import edge
    if not torch.cuda.is_available():
        with pytest.raises(edge.HcError, match="no CPU merging path"):
            edge.Merger()


# ---- general merge: paired-end and mixed members (mergeAlignmentRecordsPaired / Mixed, src/AlignmentRecord.cpp:734-1537) ----
def _digest_general(rows):
    h = hashlib.sha256()
    for single, s1, s2 in rows:
        h.update(bytes([1 if single else 0]))
        for seg in (s1,) if single else (s1, s2):
            start, end, cig, seq, qual = seg
            h.update(np.array([start, end, len(cig), len(seq)], "<i4").tobytes())
            h.update(cig); h.update(seq); h.update(qual)
    return h.hexdigest()


def unit3_paired_expected():
    """The reference's second known-answer vector, test/unit/unit3.cpp:65-93: sample12 merged with sample13 (both paired-end)
    must equal sample14 (rows 3, 4, 5 of tests/golden/unit3_merge_fixtures.npz)."""
    from tests.common import load_batch
    b = load_batch("unit3_merge_fixtures.npz")
    r = b.record(5)
    exp = (False, (r.start1, r.end1, r.cigar1_unrolled, r.sequence1, r.qualities1),
           (r.start2, r.end2, r.cigar2_unrolled, r.sequence2, r.qualities2))
    return b, [3, 4], exp


def test_general_merge_oracle_reproduces_both_unit3_vectors():
    b, members, exp = unit3_paired_expected()
    assert not b.single_end[3:].any() and b.names[5].startswith("Clique")
    st, *got = pyoracle.merge_clique(b, members)
    assert st == 0 and tuple(got) == exp
    b, members, exp1 = unit3_expected()                       # the single-end vector through the general entry point
    st, single, s1, s2 = pyoracle.merge_clique(b, members)
    assert st == 0 and single and s2 is None and s1 == exp1


@pytest.mark.parametrize("case", ["arabis", "arabis_wide", "superreads_it1", "quasi_mix", "layout_zoo", "layout_zoo_indels"])
def test_general_merge_oracle_matches_reference_digest(case):
    g = _golden()["general_" + case]
    name, batch, seed, n_cl, span, mm = next(c for c in merge_cases.GENERAL_CASES() if c[0] == case)
    rows, undefined = [], []
    for t, mem in enumerate(merge_cases.windows(batch.n, n_cl, seed, span, mm)):
        st, *rec = pyoracle.merge_clique(batch, mem)
        if st:
            undefined.append(t)
        else:
            rows.append(tuple(rec))
    assert undefined == g["undefined"]
    assert _digest_general(rows) == g["sha256"]


def test_general_merge_agrees_with_the_single_end_restatement():
    name, batch, seed, n_cl, span, mm = next(c for c in merge_cases.CASES() if c[0] == "clipped")
    for mem in merge_cases.windows(batch.n, 400, seed, span, mm):
        a = pyoracle.merge_clique_se(batch, mem)
        st, single, s1, s2 = pyoracle.merge_clique(batch, mem)
        assert st == a[0] and (st != 0 or (single and s1 == a[1:]))


@pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libhcref.so not built (needs /root/reference)")
def test_general_merge_oracle_matches_reference_directly_and_reaches_every_layout():
    import ctypes as C
    from haploclique_b200 import synth

    def counts():
        out = np.zeros(24, np.int64)
        pyoracle.lib().orc_merge_layout_counts(out.ctypes.data_as(C.c_void_p))
        return out

    counts()
    n = 0
    for seed in (80, 81, 82, 83):
        batch = synth.zoo(400, seed=seed, genome_len=600, indel=0.0)
        R = pyref.Reference()
        R.add_batch(batch)
        for t, mem in enumerate(merge_cases.windows(batch.n, 3000, seed - 80, 16, 5)):
            st, *rec = pyoracle.merge_clique(batch, mem)
            if st:
                continue
            assert R.merge_clique_full(mem, t) == tuple(rec), "clique %d %s" % (t, mem)
            n += 1
    c = counts()
    assert n > 9000
    # every branch but the tenth paired layout (:1161, needs ar.segment2 to start left of this.segment1 with ar.segment1
    # ending before it) is reached by this generator; 17 = no paired layout matched (`this` stays as it is)
    assert (c[[0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22]] > 0).all(), c.tolist()
