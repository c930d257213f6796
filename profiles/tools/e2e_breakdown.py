#!/usr/bin/env python
"""Where the end-to-end step of bench.py spends its time (C2): hc_upload vs hc_score_band (host layout + kernel + D2H)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from haploclique_b200 import edge, synth
import bench

b = synth.quasispecies(int(sys.argv[1]) if len(sys.argv) > 1 else 100_000, **{k: v for k, v in synth.CONFIGS["C2"].items() if k != "n_frag"})
calc = edge.NewEdgeCalculator(maxPosition=b.max_position() + 2000, device=0, **bench.PARAMS)
packed = edge.PackedBatch(b)
calc.upload(packed)
_, _, nw = calc.band_layout(0, b.n)
pinned = edge.PinnedBuffer(max(nw, 1), np.uint32)
for it in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); calc.upload(packed); t1 = time.perf_counter()
    adj = calc.score_band(0, b.n, out=pinned.array, amb_cap=1024); t2 = time.perf_counter()
    adj2 = calc.score_band(0, b.n, out=pinned.array, amb_cap=1024); t3 = time.perf_counter()
    print("upload %.2f ms | first score_band %.2f ms (kernel %.2f) | second score_band %.2f ms (kernel %.2f)" % (
        (t1 - t0) * 1e3, (t2 - t1) * 1e3, adj.stats["kernel_ms"], (t3 - t2) * 1e3, adj2.stats["kernel_ms"]))
