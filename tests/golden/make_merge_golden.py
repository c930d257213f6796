"""Generates tests/golden/merge_golden.json: digests of the UNMODIFIED reference's super-reads
(AlignmentRecord(alignments, id), src/AlignmentRecord.cpp:211-249, through oracle/_ref/libhcref.so) for the seeded
single-end clique batches of tests/merge_cases.py, plus the full computeEntry table (:169-184).
Run where /root/reference exists:  python tests/golden/make_merge_golden.py
Cliques on which the reference is undefined (the restatement says so first; the Release build of the reference would
not return) are listed by index and carry no digest."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle, pyref          # noqa: E402
from tests import merge_cases               # noqa: E402
from tests.common import load_batch         # noqa: E402


def digest(rows):
    h = hashlib.sha256()
    for start, end, cig, seq, qual in rows:
        h.update(np.array([start, end, len(cig), len(seq)], "<i4").tobytes())
        h.update(cig); h.update(seq); h.update(qual)
    return h.hexdigest()


def digest_general(rows):
    h = hashlib.sha256()
    for single, s1, s2 in rows:
        h.update(bytes([1 if single else 0]))
        for seg in (s1,) if single else (s1, s2):
            start, end, cig, seq, qual = seg
            h.update(np.array([start, end, len(cig), len(seq)], "<i4").tobytes())
            h.update(cig); h.update(seq); h.update(qual)
    return h.hexdigest()


def main():
    out = {}
    R = pyref.Reference()
    tab = bytearray()
    for q1 in range(33, 127):
        for q2 in range(33, 127):
            for b2 in (65, 67):
                tab += bytes(R.compute_entry(65, q1, b2, q2))
    out["compute_entry_sha256"] = hashlib.sha256(bytes(tab)).hexdigest()
    for name, batch, seed, n_cl, span, mm in merge_cases.CASES():
        R = pyref.Reference()
        R.add_batch(batch)
        rows, undefined = [], []
        for t, mem in enumerate(merge_cases.windows(batch.n, n_cl, seed, span, mm)):
            if pyoracle.merge_clique_se(batch, mem)[0] != 0:
                undefined.append(t)
                continue
            se, s, e, cig, seq, qual = R.merge_clique(mem, t)
            assert se
            rows.append((s, e, cig, seq, qual))
        out[name] = {"n_cliques": n_cl, "undefined": undefined, "sha256": digest(rows)}
        print(name, n_cl, "undefined", len(undefined))
    # general merge (paired / mixed members): digest over (single_end, both segments) of the reference's records
    for name, batch, seed, n_cl, span, mm in merge_cases.GENERAL_CASES():
        R = pyref.Reference()
        R.add_batch(batch)
        rows, undefined = [], []
        for t, mem in enumerate(merge_cases.windows(batch.n, n_cl, seed, span, mm)):
            if pyoracle.merge_clique(batch, mem)[0] != 0:
                undefined.append(t)
                continue
            rows.append(R.merge_clique_full(mem, t))
        out["general_" + name] = {"n_cliques": n_cl, "undefined": undefined, "sha256": digest_general(rows)}
        print("general", name, n_cl, "undefined", len(undefined))
    # The reference's own known-answer test for this path: test/unit/unit3.cpp:33-60 (mergeAlignmentRecordsMatch) merges
    # fixtures sample00 + sample10 and expects sample11 -- all three single-end. Their fields travel as a small npz.
    R = pyref.Reference()
    for k in (0, 10, 11, 12, 13, 14):      # unit3.cpp:65-93: sample12 + sample13 must equal sample14 (paired-end)
        R.restore_fixture(os.path.join(pyref.REFERENCE_ROOT, "test/data/simulation/unit_data/alignment_sample%02d.txt" % k))
    R.to_batch(check_cov=False).save(os.path.join(ROOT, "tests", "golden", "unit3_merge_fixtures.npz"))
    with open(os.path.join(ROOT, "tests", "golden", "merge_golden.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
