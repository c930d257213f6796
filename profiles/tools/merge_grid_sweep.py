"""Kernel time of the super-read merge against the number of resident CTAs per SM (HC_MERGE_BLOCKS_PER_SM), same
workload as merge_bench.py. Writes gpurun_out/merge_grid_sweep.json."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from haploclique_b200 import edge                 # noqa: E402
from tests import merge_cases                     # noqa: E402
from tests.common import load_batch               # noqa: E402

b = load_batch("hiv50_records.npz")
cliques = merge_cases.windows(b.n, 200000, 123, 12, 7)
edge.merge_cliques_se(b, cliques[:1000])
res = {}
for per_sm in (16, 8, 4, 2, 1):
    os.environ["HC_MERGE_BLOCKS_PER_SM"] = str(per_sm)
    res[per_sm] = min(edge.merge_cliques_se(b, cliques).kernel_ms for _ in range(2))
with open(os.path.join(ROOT, "gpurun_out", "merge_grid_sweep.json"), "w") as fh:
    json.dump({"workload": "200000 cliques of 2..7 bundled HIV-1 reads", "kernel_ms_by_ctas_per_sm": res}, fh, indent=1)
print(json.dumps(res))
