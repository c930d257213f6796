"""GPU: hc_score_pairs_static (include/hc_edge.h) -- the pair kernel run from a static device scratch array, for callers
that must not allocate device memory. Checked through the C ABI against the existing entry point hc_score_pairs: the same
records, bit for bit (the tolerance of every pair test here), on two calls in a row (the array keeps the first call's
contents: the second call is shorter and uses other pairs), plus the size limit. Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

from haploclique_b200 import edge, synth
from tests import common

pytestmark = pytest.mark.gpu


def _same(a, b, label):
    for f in edge.PAIR_RESULT_DTYPE.names:
        x, y = a[f], b[f]
        if x.dtype.kind == "f":
            assert (x.view(np.uint64) == y.view(np.uint64)).all(), "%s: %s differs" % (label, f)      # bit-exact, NaN included
        else:
            assert (x == y).all(), "%s: %s differs" % (label, f)


def test_static_scratch_entry_point_matches_hc_score_pairs_twice_in_a_row():
    b = synth.zoo(400, seed=91)
    kw = dict(common.WEAK, maxPosition=b.max_position() + 10, noProb0=True)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    rng = np.random.default_rng(5)
    n1, n2 = edge.STATIC_MAX_PAIRS, 5000
    i1 = rng.integers(1, b.n, n1).astype(np.int32)
    j1 = (rng.random(n1) * i1).astype(np.int32)
    i2 = rng.integers(1, b.n, n2).astype(np.int32)
    j2 = (rng.random(n2) * i2).astype(np.int32)
    want1, want2 = calc.score_pairs(i1, j1), calc.score_pairs(i2, j2)
    assert (want1["status"] == edge.ST_WALKED).sum() > 1000 and want1["edge"].sum() > 50
    got1 = calc.score_pairs_static(i1, j1)              # the full array
    got2 = calc.score_pairs_static(i2, j2)              # straight after, over the first call's leftovers
    _same(got1, want1, "first call")
    _same(got2, want2, "second call")
    # a second context on the same device shares the array (sequential calls are fine)
    other = edge.NewEdgeCalculator(**kw)
    other.upload(b)
    _same(other.score_pairs_static(i2, j2), want2, "second context")
    _same(calc.score_pairs_static(i1[:1], j1[:1]), want1[:1], "one pair")


def test_static_scratch_entry_point_rejects_more_pairs_than_it_holds():
    b = synth.zoo(60, seed=92)
    calc = edge.NewEdgeCalculator(**dict(common.WEAK, maxPosition=b.max_position() + 10))
    calc.upload(b)
    n = edge.STATIC_MAX_PAIRS + 1
    pi = np.ones(n, np.int32)
    pj = np.zeros(n, np.int32)
    out = np.zeros(n, edge.PAIR_RESULT_DTYPE)
    L = edge.load_library()
    rc = L.hc_score_pairs_static(calc._ctx, pi.ctypes.data_as(C.c_void_p), pj.ctypes.data_as(C.c_void_p), n, out.ctypes.data_as(C.c_void_p))
    assert rc == edge.HC_E_INVALID
    assert b"HC_STATIC_MAX_PAIRS" in L.hc_last_error(calc._ctx)
    assert L.hc_score_pairs_static(calc._ctx, None, None, 0, None) == edge.HC_OK
    with pytest.raises(edge.HcError):
        calc.score_pairs_static(pi, pj)
