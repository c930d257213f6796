"""One pass of every bench workload except C3 (C2, C4 at 200k fragments, C5, C5r) in one process, for one ncu pass with
--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:score_ : the DRAM traffic per launch that
bench.py reports as roofline.traffic (profiles/traffic.json). Prints workload, kernel, launches in order.
  python profiles/tools/traffic_all.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench                                     # noqa: E402
from haploclique_b200 import edge                # noqa: E402

for name, n in (("C2", None), ("C4", 200_000), ("C5", None), ("C5r", None)):
    b, cfg = bench.generate(name, n, workers=16)
    calc = edge.NewEdgeCalculator(**bench.calc_kwargs(b.max_position() + 10, "defaults + no_prob0"))
    calc.upload(b)
    a = calc.score_band(amb_cap=0, fetch=False)
    print("%s kernel=%d launches=%d kernel_ms=%.2f edges=%d" % (name, calc.last_kernel(), a.stats["launches"], a.stats["kernel_ms"], a.stats["edges"]), flush=True)
    del calc
