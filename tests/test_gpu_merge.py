"""GPU: super-read construction (hc_merger_* / hc_merge_*) through the C ABI against the oracle, byte for
byte, and against the reference's golden digests: single-end cliques, paired-end and mixed cliques (every layout of
mergeAlignmentRecordsPaired / Mixed the generators reach), both of the reference's unit3 vectors."""
import hashlib
import json
import os

import numpy as np
import pytest

from haploclique_b200 import edge
from oracle import pyoracle
from tests import merge_cases
from tests.common import GOLDEN

pytestmark = pytest.mark.gpu


def _digest(rows):
    h = hashlib.sha256()
    for start, end, cig, seq, qual in rows:
        h.update(np.array([start, end, len(cig), len(seq)], "<i4").tobytes())
        h.update(cig); h.update(seq); h.update(qual)
    return h.hexdigest()


@pytest.mark.parametrize("case", ["hiv50", "hiv50_large_cliques", "clipped"])
def test_merge_matches_oracle_and_reference_digest(case):
    with open(os.path.join(GOLDEN, "merge_golden.json")) as fh:
        g = json.load(fh)[case]
    name, batch, seed, n_cl, span, mm = next(c for c in merge_cases.CASES() if c[0] == case)
    cliques = merge_cases.windows(batch.n, n_cl, seed, span, mm)
    dev = edge.merge_cliques(batch, cliques)
    assert len(dev) == n_cl
    rows = []
    for t, mem in enumerate(cliques):
        st, s, e, cig, seq, qual = pyoracle.merge_clique_se(batch, mem)
        assert int(dev.info[t]["status"]) == st, "clique %d: status" % t
        rec = dev.record(t)
        if st:
            assert rec is None
            continue
        assert (rec.start1, rec.end1) == (s, e), "clique %d: interval" % t
        assert rec.cigar1_unrolled == cig, "clique %d: CIGAR" % t
        assert rec.sequence1 == seq, "clique %d: bases" % t
        assert rec.qualities1 == qual, "clique %d: qualities" % t
        rows.append((rec.start1, rec.end1, rec.cigar1_unrolled, rec.sequence1, rec.qualities1))
    assert [t for t in range(n_cl) if dev.info[t]["status"] != 0] == g["undefined"]
    assert _digest(rows) == g["sha256"]


def test_merge_reproduces_the_references_unit3_vector():
    """test/unit/unit3.cpp:33-60: sample00 merged with sample10 equals sample11."""
    from tests.test_merge_oracle import unit3_expected
    b, members, exp = unit3_expected()
    rec = edge.merge_cliques(b, [members]).record(0)
    assert (rec.start1, rec.end1, rec.cigar1_unrolled, rec.sequence1, rec.qualities1) == exp


def test_merge_fold_order_and_single_member():
    from tests.common import load_batch
    b = load_batch("hiv50_records.npz")
    dev = edge.merge_cliques(b, [[5], [5, 6], [6, 5], [5, 6, 7, 8], [8, 7, 6, 5]])
    r = dev.record(0)
    k = 10
    assert r.cigar1_unrolled == b.cig[b.cig_off[k]:b.cig_off[k] + b.cig_len[k]].tobytes()
    for t, mem in enumerate([[5], [5, 6], [6, 5], [5, 6, 7, 8], [8, 7, 6, 5]]):
        if len(mem) < 2:
            continue
        st, s, e, cig, seq, qual = pyoracle.merge_clique_se(b, mem)
        rec = dev.record(t)
        assert st == 0 and (rec.start1, rec.end1, rec.cigar1_unrolled, rec.sequence1, rec.qualities1) == (s, e, cig, seq, qual)


def test_merged_super_reads_score_like_the_oracle():
    """The device-built super-reads go back into the scorer (the next round of the reference's loop,
    src/haploclique.cpp:353-397): edges among them agree with the oracle's on the same records."""
    from haploclique_b200.records import ReadBatch
    from oracle.pyoracle import Oracle
    from tests.common import load_batch, tri_pairs, assert_pairs_equal
    b = load_batch("hiv50_records.npz")
    cliques = [list(range(i, i + 4)) for i in range(0, 400, 3)]
    dev = edge.merge_cliques(b, cliques)
    recs = [dev.record(t, t) for t in range(len(cliques))]
    recs = [r for r in recs if r is not None]
    assert len(recs) > 100
    recs.sort(key=lambda r: r.start1)
    sb = ReadBatch.from_records(recs)
    kw = dict(noProb0=True, maxPosition=int(sb.seg_end.max()) + 1)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(sb)
    pi, pj = tri_pairs(sb.n)
    assert_pairs_equal(calc.score_pairs(pi, pj), Oracle(sb, **kw).pairs(pi, pj), "super-reads")


# ---- general merge: paired-end and mixed members -------------------------------------------------------------------
def _digest_general(rows):
    h = hashlib.sha256()
    for single, s1, s2 in rows:
        h.update(bytes([1 if single else 0]))
        for seg in (s1,) if single else (s1, s2):
            start, end, cig, seq, qual = seg
            h.update(np.array([start, end, len(cig), len(seq)], "<i4").tobytes())
            h.update(cig); h.update(seq); h.update(qual)
    return h.hexdigest()


@pytest.mark.parametrize("case", ["arabis", "arabis_wide", "superreads_it1", "quasi_mix", "layout_zoo", "layout_zoo_indels"])
def test_general_merge_matches_oracle_and_reference_digest(case):
    with open(os.path.join(GOLDEN, "merge_golden.json")) as fh:
        g = json.load(fh)["general_" + case]
    name, batch, seed, n_cl, span, mm = next(c for c in merge_cases.GENERAL_CASES() if c[0] == case)
    cliques = merge_cases.windows(batch.n, n_cl, seed, span, mm)
    dev = edge.merge_cliques(batch, cliques)
    rows, undefined = [], []
    for t, mem in enumerate(cliques):
        st, *rec = pyoracle.merge_clique(batch, mem)
        assert int(dev.info[t]["status"]) == st, "clique %d %s: status %d, oracle %d" % (t, list(mem), int(dev.info[t]["status"]), st)
        if st:
            undefined.append(t)
            assert dev.segments(t) is None
            continue
        assert dev.segments(t) == tuple(rec), "clique %d %s" % (t, list(mem))
        rows.append(tuple(rec))
    assert undefined == g["undefined"]
    assert _digest_general(rows) == g["sha256"]
    assert dev.folds["column_space"] + dev.folds["walked"] > 0
    if case in ("arabis", "layout_zoo"):
        assert dev.folds["column_space"] > dev.folds["walked"]       # reads without indels: folds run in column space


def test_general_merge_reproduces_the_references_second_unit3_vector():
    """test/unit/unit3.cpp:65-93: sample12 merged with sample13 (paired-end) equals sample14."""
    from tests.test_merge_oracle import unit3_paired_expected
    b, members, exp = unit3_paired_expected()
    assert edge.merge_cliques(b, [members]).segments(0) == exp


def test_long_super_reads_use_global_scratch():
    """Cliques of super-reads longer than a warp's shared-memory slice (2048 ops per segment): the running record lives in
    global scratch; same bytes as the oracle."""
    from haploclique_b200 import synth
    b = synth.super_reads(120, seed=3)
    cliques = [list(range(i, i + 3)) for i in range(0, b.n - 3, 2)]
    dev = edge.merge_cliques(b, cliques)
    n_ok = 0
    for t, mem in enumerate(cliques):
        st, *rec = pyoracle.merge_clique(b, np.asarray(mem, np.int32))
        assert int(dev.info[t]["status"]) == st
        if st == 0:
            assert dev.segments(t) == tuple(rec)
            n_ok += 1
            assert max(len(rec[1][2]), len(rec[2][2]) if rec[2] else 0) > 0
    assert n_ok > 10 and int(dev.info["n_ops"].max()) > 2048


def test_merger_is_reusable_and_validates():
    from tests.common import load_batch
    b = load_batch("arabis_records.npz")
    m = edge.Merger()
    with pytest.raises(edge.HcError, match="no read set"):
        m.merge([[0, 1]])
    m.set_reads(b)
    a1 = m.merge([[0, 1, 2], [5, 6]])
    a2 = m.merge([[5, 6]])
    assert a1.segments(1) == a2.segments(0)
    with pytest.raises(edge.HcError, match="out of range"):
        m.merge([[0, b.n]])
    with pytest.raises(edge.HcError, match="empty clique"):
        m.merge([[0, 1], []])
    h = load_batch("hiv50_records.npz")
    m.set_reads(h)                                              # a new round's reads on the same merger
    st, *rec = pyoracle.merge_clique(h, np.array([3, 4, 5], np.int32))
    assert st == 0 and m.merge([[3, 4, 5]]).segments(0) == tuple(rec)
