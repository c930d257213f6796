"""C3 on one B200, both orders of the work items (HC_ITEM_ORDER=blocks, the default, and =rows) in one process, so that one
ncu pass (--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:score_tile2) lists the DRAM
traffic of every window launch under each order. Plain run: prints the kernel time of each order.
  python profiles/tools/item_order_traffic.py [n_frag]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench                                     # noqa: E402
from haploclique_b200 import edge                # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
b, cfg = bench.generate("C3", n, workers=16)
kw = bench.calc_kwargs(b.max_position() + 10, "defaults + no_prob0")
calc = edge.NewEdgeCalculator(**kw)
calc.upload(b)
for order in ("blocks", "rows", "blocks", "rows"):
    os.environ["HC_ITEM_ORDER"] = order
    t = time.time()
    a = calc.score_band(amb_cap=0, fetch=False)
    print("order=%s kernel_ms=%.1f wall_ms=%.1f edges=%d launches=%d" % (order, a.stats["kernel_ms"], 1e3 * (time.time() - t), a.stats["edges"], a.stats["launches"]), flush=True)
