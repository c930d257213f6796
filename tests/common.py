"""Shared helpers of the parity tests. The oracle (oracle/) is the checker; the CUDA path is what is checked."""
import hashlib
import json
import os

import numpy as np

from haploclique_b200.records import ReadBatch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

WEAK = dict(edge_quasi_cutoff=0.85, edge_quasi_cutoff_mixed=0.85, edge_quasi_cutoff_single=0.8, overlap=0.6,
            overlap_single=0.5)
# data set label in golden.json -> fixture file
DATASETS = {
    "hiv50_default": "hiv50_records.npz", "hiv50_weak": "hiv50_records.npz", "hiv50_weak_noprob0": "hiv50_records.npz",
    "arabis_weak_noprob0": "arabis_records.npz", "arabis_default": "arabis_records.npz",
    "superreads_arabis_it1": "superreads_arabis_it1.npz", "superreads_arabis_it1_default": "superreads_arabis_it1.npz",
    "superreads_arabis_it2": "superreads_arabis_it2.npz", "superreads_arabis_it2_default": "superreads_arabis_it2.npz",
}


def golden():
    with open(os.path.join(GOLDEN, "golden.json")) as fh:
        return json.load(fh)


def load_batch(name: str) -> ReadBatch:
    return ReadBatch.load(os.path.join(GOLDEN, name))


def pair_digests(status, edge, gap, cc, tc, probM, prob0):
    """Same digest as tests/golden/make_golden.py:digest, from per-pair arrays in all_pairs order.
    `status >= 2` (walked / undefined) plays the role of the reference harness' prefilter-pass flag."""
    w = np.asarray(status) >= 2
    h_edge = hashlib.sha256(np.packbits(np.asarray(edge).astype(np.uint8)).tobytes()).hexdigest()
    g = np.asarray(gap)[w].astype(np.uint8)
    ok = g == 0
    parts = [np.packbits(w).tobytes(), g.tobytes(), np.asarray(cc)[w][ok].astype("<i4").tobytes(),
             np.asarray(tc)[w][ok].astype("<i4").tobytes(), np.asarray(probM)[w][ok].astype("<f8").tobytes(),
             np.asarray(prob0)[w][ok].astype("<f8").tobytes()]
    return h_edge, hashlib.sha256(b"".join(parts)).hexdigest()


def tri_pairs(n: int):
    i, j = np.tril_indices(n, -1)
    return i.astype(np.int32), j.astype(np.int32)


def calc_kwargs(params: dict, max_position: int) -> dict:
    """golden.json params (reference constructor names) -> kwargs of NewEdgeCalculator / Oracle."""
    kw = dict(params)
    kw["maxPosition"] = max_position
    return kw


def assert_pairs_equal(dev, orc, label=""):
    """Bit-exact comparison of CUDA per-pair results with the oracle's, field by field."""
    assert (dev["status"] == orc["status"]).all(), label + ": status differs at %s" % np.nonzero(dev["status"] != orc["status"])[0][:5]
    walked = orc["status"] == 2
    assert (dev["gap"][walked] == orc["gap"][walked]).all(), label + ": gap flag differs"
    ok = walked & (orc["gap"] == 0)
    for f in ("cc", "tc"):
        bad = np.nonzero(dev[f][ok] != orc[f][ok])[0]
        assert bad.size == 0, "%s: %s differs at %s" % (label, f, np.nonzero(ok)[0][bad[:5]])
    for f in ("probM", "prob0"):
        a, b = dev[f][ok].view(np.uint64), orc[f][ok].view(np.uint64)
        bad = np.nonzero(a != b)[0]
        assert bad.size == 0, "%s: %s bits differ at %s" % (label, f, np.nonzero(ok)[0][bad[:5]])
    # decision: identical unless the device itself flagged the pair as within tolerance of the threshold
    differ = (dev["edge"] != orc["edge"]) & (orc["status"] != 3)
    unflagged = differ & (dev["ambiguous"] == 0)
    assert not unflagged.any(), label + ": edge differs on unflagged pairs %s" % np.nonzero(unflagged)[0][:5]
    return int(differ.sum())


# ---- QuasispeciesEdgeCalculator ------------------------------------------------------
def quasi_pair_sample(n: int, near: int = 40, far: int = 20000, seed: int = 7):
    """A fixed list of ordered pairs (i > j): every pair at most `near` arrivals apart plus `far` random ones."""
    i, j = tri_pairs(n)
    k = np.flatnonzero(i - j <= near)
    rng = np.random.default_rng(seed)
    if i.size > far:
        k = np.unique(np.concatenate([k, rng.choice(i.size, far, replace=False)]))
    else:
        k = np.arange(i.size)
    return i[k], j[k]


def quasi_kwargs(params: dict, max_position: int) -> dict:
    kw = dict(params)
    if kw.get("simpson_map") == "every7th":
        kw["simpson_map"] = {p: 0.05 + 0.9 * ((p * 37) % 101) / 101.0 for p in range(1, max_position + 1, 7)}
    kw["maxPosition"] = max_position
    return kw
