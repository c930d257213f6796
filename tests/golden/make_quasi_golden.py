#!/usr/bin/env python
"""Generates tests/golden/quasi_golden.json from the UNMODIFIED reference (oracle/_ref/libhcref.so, built by
`make -C oracle ref` where /root/reference exists): for each fixture and parameter set, QuasispeciesEdgeCalculator's
q = computeOverlap(...) and edgeBetween(...) on a fixed list of ordered pairs, as digests. Run from the repo root."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyref          # noqa: E402
from tests import common          # noqa: E402

PARAM_SETS = {
    "default": {},
    "low": dict(edge_quasi_cutoff=0.8, edge_quasi_cutoff_single=0.8, edge_quasi_cutoff_mixed=0.8, overlap=0.3, overlap_single=0.3),
    "frameshift": dict(frameshift_merge=True, overlap_single=0.2, overlap=0.2, edge_quasi_cutoff_single=0.9),
    "simpson": dict(overlap_single=0.3, overlap=0.3, edge_quasi_cutoff_single=0.7, edge_quasi_cutoff_mixed=0.7, edge_quasi_cutoff=0.7,
                    simpson_map="every7th"),
}
FIXTURES = ["arabis_records.npz", "hiv50_records.npz", "superreads_arabis_it1.npz", "superreads_arabis_it2.npz"]


def main():
    out = {}
    for fx in FIXTURES:
        b = common.load_batch(fx)
        R = pyref.Reference()
        R.add_batch(b)
        pi, pj = common.quasi_pair_sample(b.n)
        for name, kw in PARAM_SETS.items():
            kw = common.quasi_kwargs(kw, b.max_position())
            q, e = R.quasi_pairs(pi, pj, **{k: v for k, v in kw.items() if k != "maxPosition"})
            out["%s/%s" % (fx, name)] = {
                "pairs": int(pi.size), "edges": int(e.sum()), "q_positive": int((q > 0).sum()),
                "sha256_q": hashlib.sha256(q.astype("<f8").tobytes()).hexdigest(),
                "sha256_edges": hashlib.sha256(np.packbits(e).tobytes()).hexdigest()}
            print(fx, name, out["%s/%s" % (fx, name)])
    with open(os.path.join(common.GOLDEN, "quasi_golden.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
