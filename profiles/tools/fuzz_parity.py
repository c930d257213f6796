"""Randomised parity sweep (not part of the test suite; run on a B200 when kernels change): many seeds of the adversarial
zoo, small quasispecies sets with heavy indels/clips, super-read-shaped sets and hard-clipped reads, random parameters, with and without an
allele-frequency map, through every band kernel and both QuasispeciesEdgeCalculator walks, each against the oracle
(tests/test_gpu_parity.py:_tiled_detail_vs_oracle, tests/test_gpu_quasi.py:_compare). Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from haploclique_b200 import edge, synth  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests import common  # noqa: E402
from tests.test_gpu_parity import _tiled_detail_vs_oracle  # noqa: E402
from tests.test_gpu_quasi import _compare  # noqa: E402

budget_s = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
t0 = time.time()
done = {"band_cases": 0, "band_skipped_not_eligible": 0, "quasi_cases": 0, "pairs": 0}
seed = seed0
while time.time() - t0 < budget_s:
    rng = np.random.default_rng(seed)
    kind = seed % 4
    if kind == 0:
        lo = int(rng.integers(2, 40))
        b = synth.zoo(int(rng.integers(60, 260)), seed=seed, wide_qual=(lo, int(rng.integers(lo + 3, 95))),
                      clique_frac=float(rng.uniform(0, 1)))
        span = 700
    elif kind == 1:
        b = synth.quasispecies(int(rng.integers(200, 900)), genome_len=int(rng.integers(600, 2500)), seed=seed,
                               indel_err=float(rng.choice([1e-4, 2e-3, 1e-2])), clip_frac=float(rng.uniform(0, 0.3)),
                               n_frac=float(rng.choice([0, 5e-4, 5e-2])))
        span = 2600
    elif kind == 2:
        b = synth.super_reads(int(rng.integers(40, 120)), seed=seed, clique_frac=float(rng.uniform(0, 1)))
        span = 9800
    else:
        b = synth.hard_clipped(int(rng.integers(80, 300)), genome_len=int(rng.integers(150, 500)), seed=seed)
        span = 700
    kw = dict(noProb0=bool(rng.integers(0, 2)), edge_quasi_cutoff=float(rng.uniform(0.8, 0.995)),
              edge_quasi_cutoff_mixed=float(rng.uniform(0.8, 0.99)), edge_quasi_cutoff_single=float(rng.uniform(0.7, 0.97)),
              overlap=float(rng.uniform(0.1, 0.95)), overlap_single=float(rng.uniform(0.1, 0.8)))
    if rng.integers(0, 2):
        pos = rng.choice(np.arange(1, span), int(span * rng.uniform(0.05, 0.9)), replace=False)
        kw["simpson_map"] = {int(p): float(np.log10(rng.uniform(0.05, 1.0))) for p in pos}
    kw["maxPosition"] = max(b.max_position(), span) + int(rng.integers(0, 50))
    for kernel in (edge.KERNEL_TILED2, edge.KERNEL_TILED, edge.KERNEL_STREAMED):
        try:
            det, orc = _tiled_detail_vs_oracle(b, kw, "fuzz seed %d kernel %d" % (seed, kernel), kernel)
            done["band_cases"] += 1
            done["pairs"] += int(det.shape[0])
        except edge.HcError:
            done["band_skipped_not_eligible"] += 1
    if kind != 2 or b.n <= 80:
        qkw = dict(overlap_single=float(rng.uniform(0.05, 0.9)), overlap=float(rng.uniform(0.05, 0.9)),
                   edge_quasi_cutoff_single=float(rng.uniform(0.3, 0.99)), edge_quasi_cutoff_mixed=float(rng.uniform(0.3, 0.99)),
                   edge_quasi_cutoff=float(rng.uniform(0.3, 0.99)), frameshift_merge=bool(rng.integers(0, 2)),
                   maxPosition=kw["maxPosition"])
        if "simpson_map" in kw:
            qkw["simpson_map"] = {p: 10.0 ** v for p, v in kw["simpson_map"].items()}
        pi, pj = common.tri_pairs(b.n)
        orc = pyoracle.QuasiOracle(b, **qkw).pairs(pi, pj)
        for nofast in (False, True):
            if nofast:
                os.environ["HC_QUASI_NO_FAST"] = "1"
            else:
                os.environ.pop("HC_QUASI_NO_FAST", None)
            calc = edge.QuasispeciesEdgeCalculator(**qkw)
            calc.upload(b)
            _compare(calc.score_pairs(pi, pj), orc, "fuzz quasi seed %d nofast %d" % (seed, nofast))
            done["quasi_cases"] += 1
        os.environ.pop("HC_QUASI_NO_FAST", None)
    seed += 1
done.update(seeds=[seed0, seed], seconds=round(time.time() - t0, 1), result="all cases bit-identical to the oracle")
print(json.dumps(done))
