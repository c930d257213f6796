#!/usr/bin/env python
"""How evenly hc_partition_rows splits C3 across N parts: every part scored on ONE GPU, one after the other (kernel ms, band
slots, scored pairs, shared columns per part). Usage: part_balance.py [n_parts] [n_frag]; writes gpurun_out/part_balance.json"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
from haploclique_b200 import edge

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n_frag = int(sys.argv[2]) if len(sys.argv) > 2 else None
batch, cfg = bench.generate("C3", n_frag, workers=16)
mp = batch.max_position() + 16
out = []
for rank in range(world):
    local, row_begin, first = bench.shard(batch, rank, world)
    calc = edge.NewEdgeCalculator(device=0, **bench.calc_kwargs(mp, "defaults + no_prob0"))
    calc.upload(local)
    ms = []
    for _ in range(4):
        st = calc.score_band(row_begin, local.n, fetch=False, amb_cap=0).stats
        ms.append(st["kernel_ms"])
    r = {"rank": rank, "rows": local.n - row_begin, "halo_reads": row_begin, "kernel_ms": float(np.mean(ms[1:])), "band_slots": st["pairs_total"],
         "scored": st["pairs_scored"], "sum_cc": st["sum_cc"], "launches": st["launches"]}
    print(json.dumps(r), flush=True)
    out.append(r)
    del calc
ms = np.array([r["kernel_ms"] for r in out])
print("max / mean kernel ms: %.3f" % (ms.max() / ms.mean()))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "part_balance.json"), "w"), indent=1)
