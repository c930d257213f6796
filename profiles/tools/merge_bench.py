"""Timing of the super-read merge on one B200, with the reference's own
constructor timed beside it on a sample. Writes gpurun_out/merge_bench.json.
    python profiles/tools/merge_bench.py [n_cliques [workload ...]]     workloads: hiv50 arabis C2
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


def workload(name, n_cl):
    if name == "hiv50":          # single-end reads with indels (bundled HIV-1 reads)
        b = load_batch("hiv50_records.npz")
        return b, merge_cases.windows(b.n, n_cl, 123, 12, 7), "bundled HIV-1 reads (1836 single-end records), %d cliques of 2..7 reads from 12-read windows" % n_cl
    if name == "arabis":         # the reference's cram input: 1104 paired-end records, plain CIGARs
        b = load_batch("arabis_records.npz")
        return b, merge_cases.windows(b.n, n_cl, 123, 12, 7), "the reference's cram input (1104 paired-end records), %d cliques of 2..7 reads from 12-read windows" % n_cl
    from haploclique_b200 import synth
    b = synth.quasispecies(20000, **{k: v for k, v in synth.CONFIGS["C2"].items() if k != "n_frag"})
    return b, merge_cases.windows(b.n, n_cl, 123, 40, 12), "C2-shaped reads (20000 fragments 2x250: single-end and paired-end records), %d cliques of 2..12 reads from 40-read windows" % n_cl


def main():
    n_cl = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    out = []
    for name in (sys.argv[2:] or ["hiv50", "arabis", "C2"]):
        b, cliques, text = workload(name, n_cl)
        m = edge.Merger()
        m.set_reads(b)
        m.merge(cliques[:1000])                         # warm-up: module load, buffers
        runs = []
        for _ in range(3):
            t0 = time.perf_counter()
            m.set_reads(b)
            dev = m.merge(cliques)
            runs.append((dev.kernel_ms, (time.perf_counter() - t0) * 1e3))
        kernel_ms = min(r[0] for r in runs)
        wall_ms = min(r[1] for r in runs)
        ok = dev.info["status"] == 0
        rows = lambda c: np.stack([2 * np.asarray(c), 2 * np.asarray(c) + 1], 1).reshape(-1)
        member_ops = int(sum(int(b.cig_len[rows(c)].sum()) for c in cliques))
        out_ops = int(dev.cig.shape[0])
        one_pass_bytes = 3 * (member_ops + out_ops)     # op, base, quality streams: every member read once, result written once
        peak = 6546.2
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            pass
        res = {"workload": text, "cliques": n_cl, "defined": int(ok.sum()), "member_ops": member_ops, "merged_ops": out_ops,
               "folds": dev.folds, "kernel_ms": kernel_ms, "cliques_per_s_kernel": n_cl / (kernel_ms * 1e-3),
               "call_ms_host_buffers": wall_ms, "cliques_per_s_call": n_cl / (wall_ms * 1e-3),
               "call_note": "hc_merger_set_reads + hc_merge_run + hc_merge_fetch through the Python mirror (includes building the member lists)",
               "one_pass_bytes": one_pass_bytes, "one_pass_GBps": one_pass_bytes / (kernel_ms * 1e-3) / 1e9,
               "roofline": {"bound": "hbm", "achieved": one_pass_bytes / (kernel_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": one_pass_bytes / (kernel_ms * 1e-3) / 1e9 / peak}}
        try:
            from oracle import pyref
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
        print(json.dumps(res), flush=True)
        out.append(res)
        del m
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "merge_bench.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
