"""GPU: hc_merge_cliques_se (single-end super-read construction) through the C ABI against the
oracle, byte for byte, and against the reference's golden digests."""
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
    dev = edge.merge_cliques_se(batch, cliques)
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
    rec = edge.merge_cliques_se(b, [members]).record(0)
    assert (rec.start1, rec.end1, rec.cigar1_unrolled, rec.sequence1, rec.qualities1) == exp


def test_merge_fold_order_and_single_member():
    from tests.common import load_batch
    b = load_batch("hiv50_records.npz")
    dev = edge.merge_cliques_se(b, [[5], [5, 6], [6, 5], [5, 6, 7, 8], [8, 7, 6, 5]])
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
    dev = edge.merge_cliques_se(b, cliques)
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
