"""CPU: the C oracle (oracle/edge_oracle.c) against the UNMODIFIED reference (oracle/_ref/libhcref.so, compiled from
/root/reference by oracle/Makefile) on inputs beyond the committed golden vectors. Skipped where the reference
library has not been built (it is built in the build container and travels to the GPU box)."""
import numpy as np
import pytest

from haploclique_b200 import synth
from oracle import pyoracle, pyref
from tests import common

pytestmark = pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libhcref.so not built on this machine")


def _compare(R, params, label):
    """Every ordered pair: reference edgeBetween + checkGapsCigar values vs the oracle on the reference's own records."""
    b = R.to_batch()
    R.set_params(**params)
    edges, ref = R.score_all()
    kw = dict(params)
    kw.setdefault("maxPosition", R.max_position())
    e2, orc = pyoracle.Oracle(b, **kw).all_pairs()
    assert (orc["edge"] == ref["edge"]).all(), label
    assert edges == e2
    walked = ref["pre"] == 1
    assert ((orc["status"] >= 2) == walked).all(), label + ": prefilter"
    assert (orc["gap"][walked] == ref["gap"][walked]).all(), label + ": gap"
    ok = walked & (ref["gap"] == 0) & (orc["status"] == 2)
    for f in ("cc", "tc"):
        assert (orc[f][ok] == ref[f][ok]).all(), label + ": " + f
    for f in ("probM", "prob0"):
        assert (orc[f][ok].view(np.uint64) == ref[f][ok].view(np.uint64)).all(), label + ": " + f
    return b, int(ok.sum())


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_zoo_through_the_reference_constructors(seed):
    # random single/paired reads with clips, indels, N, wide qualities, fed as BAM alignments through
    # AlignmentRecord(BamAlignment) / pairWith, then scored by the reference and by the oracle
    z = synth.zoo(120, seed=50 + seed)
    R = pyref.Reference()
    R.add_batch(z)
    params = dict(common.WEAK, noProb0=bool(seed % 2), maxPosition=z.max_position() + 64)
    b, n = _compare(R, params, "zoo%d" % seed)
    assert b.n == z.n and n > 200


def test_quasispecies_default_and_weak_params():
    q = synth.quasispecies(250, genome_len=1200, seed=3, indel_err=2e-3, clip_frac=0.1)
    R = pyref.Reference()
    R.add_batch(q)
    _compare(R, dict(maxPosition=q.max_position() + 64), "quasi-default")
    _compare(R, dict(common.WEAK, noProb0=True, maxPosition=q.max_position() + 64), "quasi-weak")


def test_allele_frequency_map():
    q = synth.quasispecies(150, genome_len=900, seed=4)
    rng = np.random.default_rng(1)
    simpson = {int(p): float(np.log10(rng.uniform(0.3, 1.0))) for p in rng.choice(np.arange(1, 900), 400, replace=False)}
    R = pyref.Reference()
    R.add_batch(q)
    _compare(R, dict(simpson_map=simpson, maxPosition=q.max_position() + 64), "simpson")


@pytest.mark.parametrize("seed", [1, 2])
def test_hard_clipped_reads(seed):
    """Hard clips outside soft clips and indels next to the clip runs (synth.hard_clipped): the restatements against the
    reference's NewEdgeCalculator on every pair, and against its QuasispeciesEdgeCalculator on the pairs it terminates on."""
    z = synth.hard_clipped(150, seed=seed)
    R = pyref.Reference()
    R.add_batch(z)
    for noprob0 in (False, True):
        b, n = _compare(R, dict(common.WEAK, noProb0=noprob0, maxPosition=z.max_position() + 64), "hardclip%d" % seed)
        assert n > 1000
    pi, pj = common.tri_pairs(b.n)
    kw = dict(overlap_single=0.2, overlap=0.2, edge_quasi_cutoff_single=0.5, edge_quasi_cutoff_mixed=0.5, edge_quasi_cutoff=0.5,
              frameshift_merge=bool(seed % 2))
    orc = pyoracle.QuasiOracle(b, maxPosition=b.max_position(), **kw).pairs(pi, pj)
    ok = orc["status"] != 3
    q, e = R.quasi_pairs(pi[ok], pj[ok], **kw)
    assert (q.view(np.uint64) == orc["q"][ok].view(np.uint64)).all() and (e == orc["edge"][ok]).all()
    assert int((q > 0).sum()) > 500
