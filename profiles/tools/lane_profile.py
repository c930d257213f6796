#!/usr/bin/env python
"""Lane occupancy of the eight-column loop of the tile2 kernel on C2, from a library built with -DHC_T2_PROFILE
(profiles/tools/libhcgpu_profile.so): 8-column iterations and active lanes, first pass of a slot vs later passes
(a lane that left the loop for a generic step re-enters in a later pass, mostly alone)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from haploclique_b200 import edge, synth
import bench
edge.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libhcgpu_profile.so")
b = synth.quasispecies(100_000, **{k: v for k, v in synth.CONFIGS["C2"].items() if k != "n_frag"})
calc = edge.NewEdgeCalculator(maxPosition=b.max_position() + 2000, device=0, **bench.PARAMS)
calc.upload(b)
adj = calc.score_band(fetch=False, amb_cap=0)
print(adj.stats)
