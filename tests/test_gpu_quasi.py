"""B200: the CUDA QuasispeciesEdgeCalculator path (hc_quasi_score_pairs / hc_quasi_score_band through the C ABI)
against oracle/quasi_oracle.c. The running log sum is bit-identical (host-tabulated terms); the closing
pow(exp(sum), 1/n) is device libm: q within 1e-12 relative, decision identical unless the device flags the pair."""
import numpy as np
import pytest

from haploclique_b200 import edge, synth
from oracle import pyoracle
from tests import common
from tests.golden.make_quasi_golden import FIXTURES, PARAM_SETS

pytestmark = pytest.mark.gpu
REL_TOL = 1e-12


@pytest.fixture(params=["fast", "general"], autouse=True)
def walk(request, monkeypatch):
    """Both device walks: the fast one (sequence cursors derived from the CIGAR cursors, per-segment facts precomputed)
    and the general restatement it falls back to (HC_QUASI_NO_FAST forces it for every pair)."""
    if request.param == "general":
        monkeypatch.setenv("HC_QUASI_NO_FAST", "1")
    else:
        monkeypatch.delenv("HC_QUASI_NO_FAST", raising=False)
    return request.param


def _compare(dev, orc, label):
    assert (dev["status"] == np.where(orc["status"] == 3, 3, 2)).all(), label + ": status"
    ok = orc["status"] != 3
    assert (dev["n_single"][ok] == orc["n_single"][ok]).all(), label + ": singleOverlap call count"
    assert (dev["cutoff"][ok] == orc["cutoff"][ok]).all()
    # exp(sum) subnormal: the reference's own q has lost its precision there (both sides flag the same pairs, from the
    # bit-identical log sum); q is compared on the others, the decision on all
    sub = orc["subnormal"][ok] != 0
    assert (((dev["flags"][ok] & edge.QUASI_SUBNORMAL) != 0) == sub).all(), label + ": subnormal flag"
    dq, oq = dev["q"][ok][~sub], orc["q"][ok][~sub]
    assert ((dq == 0) == (oq == 0)).all(), label + ": q == 0 pattern"
    rel = np.abs(dq - oq) / np.maximum(np.abs(oq), 1e-300)
    assert float(rel[oq != 0].max(initial=0.0)) <= REL_TOL, label + ": q differs by %g" % rel[oq != 0].max()
    differ = (dev["edge"][ok] != orc["edge"][ok])
    assert not (differ & ((dev["flags"][ok] & edge.QUASI_AMBIGUOUS) == 0)).any(), label + ": decision differs on unflagged pairs"
    return int(differ.sum())


@pytest.mark.parametrize("fixture", FIXTURES)
@pytest.mark.parametrize("pset", sorted(PARAM_SETS))
def test_quasi_pairs_match_oracle(fixture, pset):
    b = common.load_batch(fixture)
    kw = common.quasi_kwargs(PARAM_SETS[pset], b.max_position())
    pi, pj = common.quasi_pair_sample(b.n)
    calc = edge.QuasispeciesEdgeCalculator(**kw)
    calc.upload(b)
    dev = calc.score_pairs(pi, pj)
    orc = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
    _compare(dev, orc, "%s/%s" % (fixture, pset))


@pytest.mark.parametrize("seed", range(4))
def test_quasi_zoo_and_band(seed):
    b = synth.zoo(160, seed=40 + seed)
    kw = dict(overlap_single=0.1, overlap=0.1, edge_quasi_cutoff_single=0.5, edge_quasi_cutoff_mixed=0.5, edge_quasi_cutoff=0.5,
              frameshift_merge=bool(seed % 2), maxPosition=b.max_position())
    calc = edge.QuasispeciesEdgeCalculator(**kw)
    adj = calc.scoreBatch(b)
    pi, pj = common.tri_pairs(b.n)
    orc = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
    dev = calc.score_pairs(pi, pj)
    flips = _compare(dev, orc, "quasi-zoo%d" % seed)
    # the banded adjacency says the same as the pair list, and nothing outside the band is an edge
    ai, aj = adj.edges()
    if flips == 0:
        e = (orc["edge"] == 1) & (orc["status"] != 3)
        assert adj.count() == int(e.sum())
        assert (ai == pi[e]).all() and (aj == pj[e]).all()
    assert adj.stats["launches"] == 1 and adj.stats["pairs_total"] >= adj.count()


@pytest.mark.parametrize("frameshift", [False, True])
def test_quasi_super_read_round(frameshift):
    """BASELINE configs[4] names this scorer for the super-read rounds: bench.py's C5 generator at 140 records (reads of
    0.5-9.7 kb, ~75 quality codes, indels and clips), every ordered pair against the oracle, with and without an
    allele-frequency map."""
    b = synth.super_reads(140, seed=777)
    rng = np.random.default_rng(12)
    simpson = {int(p): float(rng.uniform(0.3, 1.0)) for p in rng.choice(np.arange(1, 9700), 4000, replace=False)}
    pi, pj = common.tri_pairs(b.n)
    for sm in (None, simpson):
        kw = dict(overlap_single=0.3, overlap=0.5, edge_quasi_cutoff_single=0.8, edge_quasi_cutoff_mixed=0.85, edge_quasi_cutoff=0.9,
                  frameshift_merge=frameshift, simpson_map=sm, maxPosition=max(b.max_position(), 9700) + 10)
        calc = edge.QuasispeciesEdgeCalculator(**kw)
        adj = calc.scoreBatch(b)
        dev = calc.score_pairs(pi, pj)
        orc = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
        flips = _compare(dev, orc, "quasi-C5")
        assert int((orc["n_single"] > 0).sum()) > 1000 and int(((orc["q"] > 0) & (orc["status"] != 3)).sum()) > 200
        if flips == 0:
            e = (orc["edge"] == 1) & (orc["status"] != 3)
            ai, aj = adj.edges()
            assert adj.count() == int(e.sum()) and (ai == pi[e]).all() and (aj == pj[e]).all()


@pytest.mark.parametrize("seed", [1, 2])
def test_quasi_hard_clipped_reads(seed):
    b = synth.hard_clipped(220, seed=seed)
    kw = dict(overlap_single=0.2, overlap=0.2, edge_quasi_cutoff_single=0.5, edge_quasi_cutoff_mixed=0.5, edge_quasi_cutoff=0.5,
              frameshift_merge=bool(seed % 2), maxPosition=b.max_position())
    pi, pj = common.tri_pairs(b.n)
    orc = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
    calc = edge.QuasispeciesEdgeCalculator(**kw)
    calc.upload(b)
    _compare(calc.score_pairs(pi, pj), orc, "quasi-hardclip%d" % seed)
    assert int((orc["q"] > 0).sum()) > 1000


def test_quasi_gradual_underflow_decides_edges():
    """exp(sum) in the subnormal range is not 0 for the reference (glibc rounds to the subnormal grid) and
    pow(subnormal, 1/n) is an ordinary number (1e-310 ^ (1/700) = 0.36): with low cutoffs such pairs ARE edges. The device
    has to underflow gradually too (CUDA's exp alone returns 0 there). Found by profiles/tools/fuzz_parity.py."""
    b = synth.super_reads(100, seed=5)
    kw = dict(overlap_single=0.3, overlap=0.3, edge_quasi_cutoff_single=0.2, edge_quasi_cutoff_mixed=0.2, edge_quasi_cutoff=0.2,
              maxPosition=b.max_position() + 10)
    pi, pj = common.tri_pairs(b.n)
    orc = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
    sub_edges = (orc["subnormal"] != 0) & (orc["status"] != 3) & (orc["edge"] == 1)
    assert int(sub_edges.sum()) >= 10
    calc = edge.QuasispeciesEdgeCalculator(**kw)
    calc.upload(b)
    dev = calc.score_pairs(pi, pj)
    assert _compare(dev, orc, "quasi-underflow") == 0
    assert (dev["edge"][sub_edges] == 1).all() and (dev["q"][sub_edges] > 0.2).all()


def test_quasi_reference_interface():
    b = synth.quasispecies(300, genome_len=1500, seed=9)
    calc = edge.QuasispeciesEdgeCalculator(overlap_single=0.3, edge_quasi_cutoff_single=0.8, maxPosition=b.max_position())
    adj = calc.scoreBatch(b)
    orc = pyoracle.QuasiOracle(b, overlap_single=0.3, edge_quasi_cutoff_single=0.8)
    pi, pj = common.tri_pairs(b.n)
    o = orc.pairs(pi, pj)
    assert adj.count() == int(o["edge"].sum()) > 0
    with pytest.raises(edge.HcError):
        edge.QuasispeciesEdgeCalculator(edge_quasi_cutoff_single=0.0).scoreBatch(b)     # cutoffs must be > 0
