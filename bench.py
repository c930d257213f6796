#!/usr/bin/env python
"""bench.py -- read-pair edges scored per second (BASELINE.json `metric`).

Workloads (synthetic, seeded; `config.workload` names the one used):
  C2  HIV-1 9.7 kb, 5 haplotypes, 100k 2x250 fragments                      -- N=1 (BASELINE configs[1])
  C3  the same mix at 1M fragments, sharded by genome-position windows      -- N>1 (BASELINE configs[2])
  C4  1.5 kb amplicon, 1M fragments = 2M 2x150 reads, dense tiles           -- --workload C4 (BASELINE configs[3])
  C5  a super-read round: 4000 records of 0.5-9.7 kb, 70 % merged cliques, 76 quality codes -- --workload C5
      (BASELINE configs[4]; N>1: every rank holds the set and scores an equal-area block of rows)
A step = one pass of the hot path over the whole read set: every ordered pair (later read i, earlier read j) inside
the sorted-start band is decided exactly as NewEdgeCalculator::edgeBetween would and the banded adjacency bitset is
produced. "Scored" pairs = pairs that pass the reference's first reject (covered intervals intersect,
src/NewEdgeCalculator.cpp:1252) -- the pairs the reference walks or layout-rejects.

  value : scored pairs/s with the packed reads resident in HBM (kernel timed with CUDA events on its stream,
          L2 flushed between steps), whole job: all ranks' pairs / max-over-ranks time
  e2e   : the same through the C ABI from HOST buffers per step: hc_upload (H2D of the packed reads + device-side
          prefix build) + hc_score_band + D2H of the adjacency bitset into pinned memory
  roofline : the kernel is bound on chip (shared-memory operand delivery), not by HBM and not by tensor cores
          (SURVEY.md 8(d)); achieved = algorithmic operand bytes (cc*(2*2 B records + 8 B table entry) + tc*2 B) / kernel
          time against the shared-memory load bandwidth measured now by the library's probe; the fp64 and HBM
          fractions are given beside it in `rooflines`
  --impl reference : the reference's own CPU scorer (oracle/_ref, the unmodified sources compiled in the build
          container) on all host threads over a bounded sample of the same workload

N>1 (torchrun): the genome is cut into N position windows; rank r keeps the reads that start in its window (its rows)
plus the reads that start up to one maximal read span to the left (the halo: every earlier read that can overlap a
row), scores only its rows, and returns its adjacency words to its host. No data-path collective; the only
communication is the max/sum of the timing and pair counters at the end.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "read_pair_edges_scored_per_sec"
UNIT = "pairs/s"
PARAMS = dict(edge_quasi_cutoff=0.99, overlap=0.9, edge_quasi_cutoff_single=0.95, overlap_single=0.6,
              edge_quasi_cutoff_mixed=0.97, noProb0=True)   # defaults + --no_prob0 (edges must exist)
HALO = 1000                                                 # >= longest fragment span (800) + clips
MAX_WORDS = 1 << 29                                         # adjacency words per row window (2 GiB)


def workload_name(args, world):
    return args.workload if args.workload != "auto" else ("C2" if world == 1 else "C3")


def make_workload(name, n_frag, rank=0, world=1):
    """Returns (local batch, first local row this rank owns, config dict). world == 1: the whole read set."""
    from haploclique_b200 import synth
    if name == "C5":
        kw = dict(synth.C5)
        n = n_frag or kw["n_frag"]
        b = synth.super_reads(n, **{k: v for k, v in kw.items() if k != "n_frag"})
        # every read overlaps most others: no position windows; rank r scores rows [cut_r, cut_r+1) of equal pair area
        cuts = [int(round(b.n * np.sqrt(r / world))) for r in range(world + 1)]
        if world > 1:
            b = b.permute(np.arange(0, cuts[rank + 1]))
        return b, cuts[rank], dict(kw, n_frag=n)
    kw = dict(synth.CONFIGS[name])
    if n_frag:
        kw["n_frag"] = n_frag
    n = kw.pop("n_frag")
    G = kw["genome_len"]
    if world == 1:
        # C2 is one draw; every other workload is the chunked draw, so that N=1 and N>1 see the same read set
        b = synth.quasispecies(n, **kw) if name == "C2" else synth.quasispecies_sharded(n, **kw)
        return b, 0, dict(kw, n_frag=n)
    lo = 1 + (G * rank) // world
    hi = 1 + (G * (rank + 1)) // world if rank + 1 < world else G + 2 * HALO
    b = synth.quasispecies_sharded(n, start_window=(lo - HALO, hi), **kw)
    row_begin = int(np.searchsorted(b.seg_start[0::2], lo, side="left"))
    return b, row_begin, dict(kw, n_frag=n)


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 7 for k in range(4) if r[3 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": reasons, "samples": len(sm)}


def cpu_sample(batch):
    """Bounded sample for the CPU arms: every ordered pair among 2000 consecutive reads from the middle of the set."""
    k = min(2000, batch.n)
    lo = max(0, batch.n // 2 - k // 2)
    sub = batch.permute(np.arange(lo, lo + k))
    i, j = np.tril_indices(k, -1)
    scored = int(((sub.cov_last[i] >= sub.cov_first[j]) & (sub.cov_last[j] >= sub.cov_first[i])).sum())
    return sub, i.astype(np.int32), j.astype(np.int32), scored, \
        "all %d ordered pairs among %d consecutive reads at the middle of the read set (%d scored)" % (i.size, k, scored)


def run_cpu(batch, steps, warmup, threads, max_position):
    """Times the reference's own edgeBetween (oracle/_ref) -- or, if that library is absent, the C oracle port."""
    sub, pi, pj, scored, desc = cpu_sample(batch)
    from oracle import pyref
    if pyref.available():
        R = pyref.Reference()
        R.add_batch(sub)
        R.set_params(maxPosition=max_position, **PARAMS)
        kind = "reference"

        def once():
            return R.time_pairs(pi, pj, threads)[0]
    else:
        from oracle import pyoracle
        O = pyoracle.Oracle(sub, maxPosition=max_position, **PARAMS)
        kind, threads = "port", 1

        def once():
            t = time.perf_counter()
            O.pairs(pi, pj)
            return time.perf_counter() - t
    for _ in range(warmup):
        once()
    secs = [once() for _ in range(steps)]
    return scored / float(np.mean(secs)), float(np.mean(secs)) * 1e3, kind, threads, desc


def workload_text(name, cfg):
    if name == "C5":
        return "C5: synthetic super-read round, %d records of %d-%d columns (70 %% Clique_*, posterior qualities up to char 126), genome %d, seed 20261104" % (
            cfg["n_frag"], cfg["frag_min"], cfg["frag_max"], cfg["genome_len"])
    return "%s: synthetic HIV-1 quasispecies, 5 haplotypes, %d fragments 2x%d bp, genome %d, seed 20261104" % (
        name, cfg["n_frag"], cfg["read_len"], cfg["genome_len"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "C2", "C3", "C4", "C5"])
    ap.add_argument("--n-frag", type=int, default=0, help="override the fragment count (debugging only)")
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "tiled", "tiled2", "streamed"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    wname = workload_name(args, max(world, args.gpus))

    if args.impl == "reference":
        if rank != 0:
            return 0
        from haploclique_b200 import synth
        cfg = dict(synth.C5 if wname == "C5" else synth.CONFIGS[wname])
        if args.n_frag:
            cfg["n_frag"] = args.n_frag
        # the CPU arm scores a bounded sample; a 100k-fragment draw of the same mix has the same local read density
        n_draw = min(cfg["n_frag"], 100_000)
        gen = synth.super_reads if wname == "C5" else synth.quasispecies
        batch = gen(n_draw, **{k: v for k, v in cfg.items() if k != "n_frag"})
        value, ms, kind, threads, desc = run_cpu(batch, max(args.steps, 1), min(args.warmup, 1), cores, batch.max_position())
        if n_draw != cfg["n_frag"]:
            desc += "; drawn at %d fragments: pairs/s of the CPU scorer does not depend on the set size" % n_draw
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_text(wname, cfg), "params": "defaults + no_prob0"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return 0

    import torch
    import torch.distributed as dist
    from haploclique_b200 import edge
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the scored path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # keep rank 0's stdout to the one JSON line: NCCL writes its version banner (and any debug output) to stdout
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            del os.environ["NCCL_DEBUG"]
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    t0 = time.perf_counter()
    local, row_begin, cfg = make_workload(wname, args.n_frag or None, rank, world)
    gen_s = time.perf_counter() - t0
    row_end = local.n
    max_position = cfg["genome_len"] + 2 * HALO

    calc = edge.NewEdgeCalculator(maxPosition=max_position, device=local_rank, **PARAMS)
    calc.set_kernel({"auto": edge.KERNEL_AUTO, "generic": edge.KERNEL_GENERIC, "tiled": edge.KERNEL_TILED, "tiled2": edge.KERNEL_TILED2,
                     "streamed": edge.KERNEL_STREAMED}[args.kernel])
    t0 = time.perf_counter()
    packed = edge.PackedBatch(local)
    pack_s = time.perf_counter() - t0
    calc.upload(packed)
    # the adjacency band is produced in row windows of at most MAX_WORDS 32-bit words (one window for C2/C3; the dense
    # amplicon C4 at full size has 1.6e10 words): each window is scored and, in the e2e leg, copied to the same pinned buffer
    _, off, nw = calc.band_layout(row_begin, row_end)
    cuts = [row_begin]
    while cuts[-1] < row_end:
        k0 = cuts[-1] - row_begin
        k1 = int(np.searchsorted(off, off[k0] + MAX_WORDS, side="right")) - 1
        cuts.append(row_begin + max(k1, k0 + 1))
    cuts[-1] = min(cuts[-1], row_end)
    windows = list(zip(cuts[:-1], cuts[1:]))
    win_words = max(int(off[b - row_begin] - off[a - row_begin]) for a, b in windows) if windows else 1
    pinned = edge.PinnedBuffer(max(win_words, 1), np.uint32)
    L, ctx = calc.L, calc._ctx

    def score_windows(**kw):
        tot = None
        for a, b in windows:
            st = calc.score_band(a, b, **kw).stats
            tot = dict(st) if tot is None else {k: tot[k] + st[k] for k in st}
        return tot

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step_resident():
        L.hc_l2_flush(ctx)
        return score_windows(fetch=False, amb_cap=0)

    def step_e2e():
        L.hc_l2_flush(ctx)
        torch.cuda.synchronize()
        t = time.perf_counter()
        calc.upload(packed)
        st = score_windows(out=pinned.array, amb_cap=1024)
        return time.perf_counter() - t, st

    def step_e2e_records():
        """As step_e2e, starting one call earlier: hc_pack of the AlignmentRecord fields (the SoA view the C++ adapter
        hands over) is inside the timed region too."""
        L.hc_l2_flush(ctx)
        torch.cuda.synchronize()
        t = time.perf_counter()
        pk = edge.PackedBatch(local)
        calc.upload(pk)
        score_windows(out=pinned.array, amb_cap=1024)
        s = time.perf_counter() - t
        del pk                                              # back to the library's pinned-buffer pool
        return s

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step_resident()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    wall0 = time.perf_counter()
    kms, launches, st = [], 0, None
    for _ in range(args.steps):
        st = step_resident()
        kms.append(st["kernel_ms"])
        launches += st["launches"]
    barrier()
    wall = time.perf_counter() - wall0
    e2e_s = []
    step_e2e()
    for _ in range(args.steps):
        barrier()
        s, adj = step_e2e()
        e2e_s.append(s)
    barrier()
    rec_s = []
    step_e2e_records()
    step_e2e_records()                                      # two untimed passes: the pinned-buffer pool is warm from the third on
    for _ in range(min(args.steps, 3)):
        barrier()
        rec_s.append(step_e2e_records())
    calc.upload(packed)
    barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    kernel_used = {0: "none", 1: "generic", 2: "tiled", 3: "tiled2", 4: "streamed"}[calc.last_kernel()]
    lds_gbs, dmul_s = edge.measure_peaks(local_rank) if rank == 0 else (0.0, 0.0)

    # whole-job numbers: units of all ranks / max-over-ranks time
    t_dev = torch.tensor([float(np.mean(kms)), float(np.mean(e2e_s)) * 1e3, wall / args.steps * 1e3, float(np.mean(rec_s)) * 1e3],
                         device="cuda", dtype=torch.float64)
    cnt = torch.tensor([st["pairs_scored"], st["pairs_total"], st["edges"], st["sum_cc"], st["sum_tc"], st["pairs_walked"],
                        packed.nbytes(), nw * 4, st["ambiguous"], st["undefined"], local.n - row_begin],
                       device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_kernel, ms_e2e, ms_wall, ms_rec = [float(x) for x in t_dev.tolist()]
    scored, total, edges, sum_cc, sum_tc, walked, h2d, d2h, amb, undef, rows = [int(x) for x in cnt.tolist()]

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(wname)
        except Exception:
            pass
        secs = ms_kernel * 1e-3
        # algorithmic figures of SURVEY.md 8(d), per launch (world launches run side by side, one per GPU)
        operand_bytes = sum_cc * 12 + sum_tc * 2
        hbm_bytes = h2d + d2h
        smem_ach = operand_bytes / secs / 1e9 / world
        fp64_ach = sum_cc / secs / world
        hbm_ach = hbm_bytes / secs / 1e9 / world
        out = {
            "metric": METRIC, "value": scored / secs, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_kernel, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_text(wname, cfg), "params": "defaults + no_prob0", "kernel": kernel_used,
                       "l2": "flushed between steps (256 MiB memset)", "row_windows": len(windows),
                       "sharding": "single GPU" if world == 1 else ("%d equal-area row blocks of the replicated set, no collective" % world) if wname == "C5"
                                   else ("%d genome-position windows + %d-column left halo, no collective" % (world, HALO))},
            "pairs": {"scored": scored, "examined": total, "walked": walked, "edges": edges, "sum_cc": sum_cc, "sum_tc": sum_tc,
                      "rows": rows, "all_pairs": rows * (rows - 1) // 2, "ambiguous_within_1e-12": amb, "undefined": undef},
            "all_pairs_per_s": rows * (rows - 1) / 2 / secs,
            "column_steps_per_s": (sum_cc + sum_tc) / secs,
            "wall_ms_per_step": ms_wall,
            "e2e": {"value": scored / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e, "host_pack_ms_once": pack_s * 1e3, "host_generate_s_once": gen_s,
                    "from_records": {"value": scored / (ms_rec * 1e-3), "unit": UNIT, "ms_per_step": ms_rec,
                                     "note": "hc_pack of the record fields + hc_upload + hc_score_band + D2H per step"}},
            "gpu_launches": launches * world,
            "roofline": {"bound": "smem", "achieved": smem_ach, "peak": lds_gbs, "unit": "GB/s", "frac": smem_ach / lds_gbs if lds_gbs else None,
                         "traffic": traffic, "peak_source": "hc_measure_peaks (ld.shared.v4 probe, this run, this GPU)",
                         "note": "per GPU; operand bytes = sum_cc*12 + sum_tc*2 (SURVEY.md 8(d)); HBM and tensor cores do not bound "
                                 "this kernel, see rooflines and DESIGN.md"},
            "rooflines": {
                "smem": {"achieved": smem_ach, "peak": lds_gbs, "unit": "GB/s", "frac": smem_ach / lds_gbs if lds_gbs else None},
                "fp64": {"achieved": fp64_ach, "peak": dmul_s, "unit": "DMUL/s", "frac": fp64_ach / dmul_s if dmul_s else None},
                "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}},
            "clocks": sampler.summary(),
        }
        if not args.no_cpu_baseline and world == 1:
            v, ms, kind, threads, desc = run_cpu(local, 1, 0, cores, max_position)
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": desc}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
