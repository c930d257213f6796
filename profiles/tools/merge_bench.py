"""Timing of the single-end super-read merge (first slice) on one B200, with the reference's own
constructor timed beside it on a sample. Writes gpurun_out/merge_bench.json.
    python profiles/tools/merge_bench.py [n_cliques]
Input: the bundled HIV-1 reads (tests/golden/hiv50_records.npz), cliques = seeded windows of 2..7 consecutive reads
(tests/merge_cases.windows), i.e. the shape of a first CLEVER round on that data."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from haploclique_b200 import edge                 # noqa: E402
from tests import merge_cases                     # noqa: E402
from tests.common import load_batch               # noqa: E402


def main():
    n_cl = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    b = load_batch("hiv50_records.npz")
    cliques = merge_cases.windows(b.n, n_cl, 123, 12, 7)
    edge.merge_cliques_se(b, cliques[:1000])        # warm-up: context, module load
    runs = []
    for _ in range(3):
        t0 = time.perf_counter()
        dev = edge.merge_cliques_se(b, cliques)
        runs.append((dev.kernel_ms, (time.perf_counter() - t0) * 1e3))
    kernel_ms = min(r[0] for r in runs)
    wall_ms = min(r[1] for r in runs)
    ok = dev.info["status"] == 0
    member_ops = int(sum(int(b.cig_len[2 * np.asarray(c)].sum()) for c in cliques))
    out_ops = int(dev.info["n_ops"].sum())
    one_pass_bytes = 3 * (member_ops + out_ops)     # op, base, quality streams: every member read once, result written once
    res = {"workload": "bundled HIV-1 reads (1836 single-end records), %d cliques of 2..7 reads from 12-read windows" % n_cl,
           "cliques": n_cl, "defined": int(ok.sum()), "member_ops": member_ops, "merged_ops": out_ops,
           "kernel_ms": kernel_ms, "cliques_per_s_kernel": n_cl / (kernel_ms * 1e-3),
           "call_ms_host_buffers": wall_ms, "cliques_per_s_call": n_cl / (wall_ms * 1e-3),
           "one_pass_bytes": one_pass_bytes, "one_pass_GBps": one_pass_bytes / (kernel_ms * 1e-3) / 1e9,
           "note": "first correct path: one lane per clique, byte loads; the fold re-reads the running super-read, so real "
                   "traffic is above the one-pass figure"}
    try:
        from oracle import pyref, pyoracle
        if pyref.available():
            import ctypes as C
            R = pyref.Reference()
            R.add_batch(b)
            sample = [np.ascontiguousarray(c, np.int32) for c, k in zip(cliques[:20000], ok[:20000]) if k]
            t0 = time.perf_counter()
            for t, mem in enumerate(sample):
                R.L.hcref_merge_clique(R.h, mem.ctypes.data_as(C.c_void_p), C.c_int(mem.size), C.c_uint(t))
            dt = time.perf_counter() - t0
            res["cpu_baseline"] = {"value": len(sample) / dt, "unit": "cliques/s", "cores": 1, "kind": "reference",
                                   "sample": "the first %d defined cliques, AlignmentRecord(alignments,id) incl. coveredPositions" % len(sample)}
    except Exception as e:                           # the checker is optional here
        res["cpu_baseline"] = {"error": str(e)}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "merge_bench.json"), "w") as fh:
        json.dump(res, fh, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
