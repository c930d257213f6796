#!/usr/bin/env python
"""C2 with an allele-frequency map (every third position has an entry), no_prob0 off: the tiled MAP variant (tails walked
column by column) against the generic one-lane-per-pair walker -- the only band kernel that handled a map before."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from haploclique_b200 import edge, synth
b = synth.quasispecies(100_000, **{k: v for k, v in synth.CONFIGS["C2"].items() if k != "n_frag"})
rng = np.random.default_rng(1)
simpson = {int(p): float(np.log10(rng.uniform(0.3, 1.0))) for p in range(1, 9720, 3)}
calc = edge.NewEdgeCalculator(maxPosition=b.max_position() + 2000, device=0, simpson_map=simpson, noProb0=False)
calc.upload(b)
out = {}
for name, mode in (("tiled2_map", edge.KERNEL_TILED2), ("tiled_map", edge.KERNEL_TILED), ("streamed_map", edge.KERNEL_STREAMED),
                   ("generic", edge.KERNEL_GENERIC)):
    calc.set_kernel(mode)
    calc.score_band(fetch=False, amb_cap=0)
    adj = calc.score_band(amb_cap=0)
    out[name] = {"kernel_ms": adj.stats["kernel_ms"], "edges": adj.stats["edges"], "pairs_scored": adj.stats["pairs_scored"],
                 "sum_cc": adj.stats["sum_cc"], "sum_tc": adj.stats["sum_tc"], "words_sha": __import__("hashlib").sha256(adj.words.tobytes()).hexdigest()[:16]}
out["identical_words"] = len({out[k]["words_sha"] for k in ("tiled2_map", "tiled_map", "streamed_map", "generic")}) == 1
print(json.dumps(out))
