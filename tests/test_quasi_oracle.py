"""CPU: oracle/quasi_oracle.c (restatement of QuasispeciesEdgeCalculator) against the digests
generated from the unmodified reference (tests/golden/make_quasi_golden.py), and directly against oracle/_ref when it
is present. The reference has no test of its own for this class."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import pyoracle, pyref
from tests import common
from tests.golden.make_quasi_golden import FIXTURES, PARAM_SETS

with open(os.path.join(common.GOLDEN, "quasi_golden.json")) as fh:
    QGOLD = json.load(fh)


@pytest.mark.parametrize("fixture", FIXTURES)
@pytest.mark.parametrize("pset", sorted(PARAM_SETS))
def test_quasi_oracle_matches_reference_digests(fixture, pset):
    g = QGOLD["%s/%s" % (fixture, pset)]
    b = common.load_batch(fixture)
    pi, pj = common.quasi_pair_sample(b.n)
    assert pi.size == g["pairs"]
    o = pyoracle.QuasiOracle(b, **common.quasi_kwargs(PARAM_SETS[pset], b.max_position())).pairs(pi, pj)
    assert int((o["status"] == 3).sum()) == 0
    assert int(o["edge"].sum()) == g["edges"] and int((o["q"] > 0).sum()) == g["q_positive"]
    assert hashlib.sha256(o["q"].astype("<f8").tobytes()).hexdigest() == g["sha256_q"]
    assert hashlib.sha256(np.packbits(o["edge"].astype(np.uint8)).tobytes()).hexdigest() == g["sha256_edges"]


@pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libhcref.so not built on this machine")
def test_quasi_oracle_matches_compiled_reference_on_the_zoo():
    from haploclique_b200 import synth
    kw = dict(overlap_single=0.1, overlap=0.1, edge_quasi_cutoff_single=0.5, edge_quasi_cutoff_mixed=0.5, edge_quasi_cutoff=0.5)
    for seed in (3, 4):
        b = synth.zoo(120, seed=seed)
        R = pyref.Reference()
        R.add_batch(b)
        pi, pj = common.tri_pairs(b.n)
        o = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
        ok = o["status"] != 3                   # the reference would hang on the others: not asked
        q, e = R.quasi_pairs(pi[ok], pj[ok], **kw)
        assert (q.view(np.uint64) == o["q"][ok].view(np.uint64)).all()
        assert (e == o["edge"][ok]).all()


def test_quasi_exact_match_mode_and_undefined_pairs():
    b = common.load_batch("arabis_records.npz")
    pi, pj = common.quasi_pair_sample(b.n, near=10, far=0)
    o = pyoracle.QuasiOracle(b, edge_quasi_cutoff_single=1.0, overlap_single=0.3).pairs(pi, pj)
    assert set(np.unique(o["q"])) <= {0.0, 1.0}            # cutoff == 1.0: exact-match mode (:253, :298-306, :424)
    assert int(o["edge"].sum()) == int((o["q"] == 1.0).sum()) > 0
