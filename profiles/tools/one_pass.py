"""One upload and N scoring passes of one bench workload in a short process (for an ncu capture of its kernel).
  python profiles/tools/one_pass.py C5 [passes] [n_frag]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench                                     # noqa: E402
from haploclique_b200 import edge                # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C5"
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 3
n = int(sys.argv[3]) if len(sys.argv) > 3 else None
b, cfg = bench.generate(name, n, workers=16)
calc = edge.NewEdgeCalculator(**bench.calc_kwargs(b.max_position() + 10, "defaults + no_prob0"))
calc.upload(b)
for _ in range(passes):
    a = calc.score_band(amb_cap=0, fetch=False)
    print("%s kernel=%d launches=%d kernel_ms=%.3f edges=%d" % (name, calc.last_kernel(), a.stats["launches"], a.stats["kernel_ms"], a.stats["edges"]), flush=True)
