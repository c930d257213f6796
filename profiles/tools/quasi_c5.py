"""QuasispeciesEdgeCalculator band kernel on bench.py's C5 super-read round (BASELINE configs[4] names this scorer):
kernel ms / pass and scored pairs/s. (No CPU leg: the reference's class does not terminate on some clipped pairs,
src/QuasispeciesEdgeCalculator.cpp:258-267 -- the device reports those HC_ST_UNDEFINED.)"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from haploclique_b200 import edge, synth  # noqa: E402

b = synth.super_reads(int(sys.argv[1]) if len(sys.argv) > 1 else 4000)
kw = dict(maxPosition=b.max_position() + 10)
calc = edge.QuasispeciesEdgeCalculator(**kw)
calc.upload(b)
out = {}
for rep in range(3):
    t = time.perf_counter()
    adj = calc.score_band()
    wall = time.perf_counter() - t
    out = {"kernel_ms": adj.stats["kernel_ms"], "wall_ms": wall * 1e3, "pairs_scored": adj.stats["pairs_scored"],
           "edges": adj.stats["edges"], "undefined": adj.stats["undefined"],
           "pairs_per_s": adj.stats["pairs_scored"] / (adj.stats["kernel_ms"] * 1e-3)}
out["workload"] = "C5 synthetic super-read round, %d records" % b.n
print(json.dumps(out))
