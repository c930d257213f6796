"""Generates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref/libhcref.so, compiled
from /root/reference). Run in the build container only:  python tests/golden/make_golden.py

Outputs (all under tests/golden/):
  hiv50_records.npz / arabis_records.npz   the reference's readBamFile() records of its two bundled BAMs
  unit_fixtures.npz                        test/data/simulation/unit_data/alignment_sample00..09 as restored by
                                           the reference's restoreCompleteAlignmentRecord
  superreads_arabis_it{1,2}.npz            the super-read deques the reference's CLEVER loop hands to iterations 1, 2
  golden.json                              per data set: edge counts, SHA-256 digests of the reference's per-pair
                                           edge bits and (gap, cc, tc, probM bits, prob0 bits), unit-test booleans
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/test/data/simulation"
WEAK = dict(edge_quasi_cutoff=0.85, edge_quasi_cutoff_mixed=0.85, edge_quasi_cutoff_single=0.8, overlap=0.6,
            overlap_single=0.5)


def digest(ref_out):
    """Digest definition shared with tests/common.py:pair_digests."""
    w = ref_out["pre"] == 1
    h_edge = hashlib.sha256(np.packbits(ref_out["edge"]).tobytes()).hexdigest()
    g = ref_out["gap"][w].astype(np.uint8)
    ok = g == 0
    parts = [np.packbits(w).tobytes(), g.tobytes(), ref_out["cc"][w][ok].astype("<i4").tobytes(),
             ref_out["tc"][w][ok].astype("<i4").tobytes(), ref_out["probM"][w][ok].astype("<f8").tobytes(),
             ref_out["prob0"][w][ok].astype("<f8").tobytes()]
    return h_edge, hashlib.sha256(b"".join(parts)).hexdigest()


def score(R, label, params, golden):
    R.set_params(**params)
    edges, out = R.score_all()
    h_edge, h_detail = digest(out)
    golden[label] = dict(params=params, n=R.n, max_position=R.max_position(), edges=edges,
                         prefilter_pass=int(out["pre"].sum()), gaps=int(out["gap"][out["pre"] == 1].sum()),
                         sum_cc=int(out["cc"][(out["pre"] == 1) & (out["gap"] == 0)].sum()),
                         sum_tc=int(out["tc"][(out["pre"] == 1) & (out["gap"] == 0)].sum()),
                         sha256_edges=h_edge, sha256_detail=h_detail)
    print(label, golden[label])


def main():
    golden = {}
    # --- bundled BAMs (config C1) ---
    R = pyref.Reference()
    R.load_bam(os.path.join(REF, "reads_HIV-1_50_01.bam"))
    R.to_batch().save(os.path.join(HERE, "hiv50_records.npz"))
    score(R, "hiv50_default", {}, golden)
    score(R, "hiv50_weak", dict(WEAK), golden)
    score(R, "hiv50_weak_noprob0", dict(WEAK, noProb0=True), golden)

    R = pyref.Reference()
    R.load_bam(os.path.join(REF, "arabis_large_500_01", "reads.bam"))
    R.to_batch().save(os.path.join(HERE, "arabis_records.npz"))
    score(R, "arabis_weak_noprob0", dict(WEAK, noProb0=True), golden)
    score(R, "arabis_default", {}, golden)
    # --- super-read rounds (config C5 shape): the reference's own CLEVER loop, cram-test parameters ---
    ct, sd = 0, 1.0
    for it in (1, 2):
        R.set_params(**dict(WEAK, noProb0=True))
        n, ct, edges, conv, sd = R.iterate(ct, sd, use_bk=False, filter_singletons=True, significance=4.0)
        print("iteration", it, "records", n, "edges", edges)
        b = R.to_batch()
        b.save(os.path.join(HERE, "superreads_arabis_it%d.npz" % it))
        # note: the CLI drops reads before scoring (filter_fn); the digest is over ALL records of the deque
        score(R, "superreads_arabis_it%d" % it, dict(WEAK, noProb0=True), golden)
        score(R, "superreads_arabis_it%d_default" % it, {}, golden)

    # --- the reference's unit-test fixtures and expectations (test/unit/unit2.cpp) ---
    R = pyref.Reference()
    for k in range(10):
        R.restore_fixture(os.path.join(REF, "unit_data", "alignment_sample%02d.txt" % k))
    # cov_pos of restored fixtures comes from the file; our records recompute it: checked inside to_batch()
    R.to_batch().save(os.path.join(HERE, "unit_fixtures.npz"))
    R.set_params(maxPosition=20000)      # unit2.cpp passes 0 (out-of-bounds SIMPSON_MAP reads); see SURVEY.md 4
    cases = {"edgeBetweenSameAlignments": (0, 0, True), "edgeBetweenSameAlignmentsSoftClip": (1, 1, True),
             "edgeBetweenIncompatibleAlignmentsStartDel": (2, 3, False),
             "edgeBetweenNonOverlappingAlignments": (4, 5, False)}
    src = open("/root/reference/test/unit/unit2.cpp").read()
    import re
    tests = re.findall(r"TEST\(edgeBetweenFunctionTest, (\w+)\)(.*?)EXPECT_EQ\(set_edge, (\w+)\)", src, re.S)
    unit = {}
    for name, body, expect in tests:
        files = re.findall(r"alignment_sample(\d+)\.txt", body)
        i, j = (int(files[0]), int(files[0])) if len(files) == 1 else (int(files[0]), int(files[1]))
        got = R.edge(i, j)
        unit[name] = dict(ap1=i, ap2=j, expected=(expect == "true"), reference_says=bool(got))
        assert got == (expect == "true"), name
    assert all(unit[k]["expected"] == v[2] and (unit[k]["ap1"], unit[k]["ap2"]) == v[:2] for k, v in cases.items())
    golden["unit2"] = unit
    print(unit)
    with open(os.path.join(HERE, "golden.json"), "w") as fh:
        json.dump(golden, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
