"""CPU: the oracle (oracle/edge_oracle.c) against the golden vectors generated from the unmodified reference
(tests/golden/make_golden.py) and against the reference's own unit-test expectations (test/unit/unit2.cpp)."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import common


@pytest.mark.parametrize("label", sorted(common.DATASETS))
def test_oracle_matches_reference_digests(golden, label):
    g = golden[label]
    b = common.load_batch(common.DATASETS[label])
    assert b.n == g["n"]
    O = pyoracle.Oracle(b, **common.calc_kwargs(g["params"], g["max_position"]))
    edges, out = O.all_pairs()
    assert edges == g["edges"]
    assert int((out["status"] >= 2).sum()) == g["prefilter_pass"]
    assert int((out["status"] == 3).sum()) == 0
    ok = (out["status"] == 2) & (out["gap"] == 0)
    assert int(out["cc"][ok].sum()) == g["sum_cc"] and int(out["tc"][ok].sum()) == g["sum_tc"]
    h_edge, h_detail = common.pair_digests(out["status"], out["edge"], out["gap"], out["cc"], out["tc"],
                                           out["probM"], out["prob0"])
    assert h_edge == g["sha256_edges"]
    assert h_detail == g["sha256_detail"]


def test_headline_edge_counts(golden):
    # iteration-0 edge counts of the reference CLI on its bundled inputs (SURVEY.md 8(c) iv)
    assert golden["arabis_weak_noprob0"]["edges"] == 153685
    assert golden["hiv50_default"]["edges"] == 220
    assert golden["superreads_arabis_it1"]["edges"] == 48675
    assert golden["superreads_arabis_it2"]["edges"] == 421


def test_reference_unit_expectations(golden):
    b = common.load_batch("unit_fixtures.npz")
    O = pyoracle.Oracle(b, maxPosition=20000)
    for name, case in golden["unit2"].items():
        assert O.edge(case["ap1"], case["ap2"]) == case["expected"], name


def test_tables_shape_and_symmetry():
    eq, un = pyoracle.tables()
    assert (eq[:33] == 0).all() and (eq[:, :33] == 0).all()
    q = np.arange(33, 127)
    sub = eq[np.ix_(q, q)]
    assert (sub > 0).all() and (sub <= 1).all()
    # Q40 vs Q40 equal bases: (1-1e-4)^2 + 1e-4*(1e-4/3) within float rounding of the error probabilities
    assert abs(eq[73, 73] - ((1 - 1e-4) ** 2 + 1e-8 / 3)) < 1e-9
    assert abs(un[73, 73] - (2 * (1 - 1e-4) * 1e-4 / 3 + 2 * (1e-4 / 3) ** 2)) < 1e-9


def test_empty_and_singleton_batches():
    from haploclique_b200.records import ReadBatch
    b = common.load_batch("unit_fixtures.npz")
    one = b.select([0])
    edges, out = pyoracle.Oracle(one).all_pairs()
    assert edges == 0 and out.shape[0] == 0
    assert ReadBatch.from_records([]).n == 0
