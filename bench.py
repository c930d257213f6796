#!/usr/bin/env python
"""bench.py -- read-pair edges scored per second (BASELINE.json `metric`).

Workloads (synthetic, seeded; `config.workload` names the one timed):
  C3  HIV-1 9.7 kb, 5 haplotypes, 1M 2x250 fragments (BASELINE configs[2])      -- the default at every N, so that the
      1 -> 8 GPU curve is one workload (strong scaling); the rows are cut by the library's partitioner
  C2  the same mix at 100k fragments (BASELINE configs[1])                      -- `also.C2` at N=1, or --workload C2
  C4  1.5 kb amplicon, 2x150 reads, dense tiles (BASELINE configs[3])           -- `also.C4` at 200k fragments; full size
      (1M fragments = 2M reads, 2.3e11 scored pairs, 20 s per pass) with --workload C4
  C5  a super-read round: 4000 records of 0.5-9.7 kb, 76 quality codes (BASELINE configs[4]) -- `also.C5` (synthetic shape up
      to full-length haplotypes) and `also.C5r` (a round captured from the reference's own loop: 3529 Clique_* records)
A step = one pass of the hot path over the whole read set: every ordered pair (later read i, earlier read j) inside
the sorted-start band is decided exactly as NewEdgeCalculator::edgeBetween would and the banded adjacency bitset is
produced. "Scored" pairs = pairs that pass the reference's first reject (covered intervals intersect,
src/NewEdgeCalculator.cpp:1252) -- the pairs the reference walks or layout-rejects.

  value : scored pairs/s with the packed reads resident in HBM (CUDA events on the compute stream from the clearing of
          the adjacency words to the end of the last scoring kernel; L2 flushed between steps), whole job: all ranks'
          pairs / max-over-ranks time
  e2e   : the same through the C ABI from HOST buffers per step: hc_upload (H2D of the packed reads + device-side
          builds) + hc_score_band + D2H of the adjacency bitset into pinned memory (row windows: the copy of a finished
          window overlaps the scoring of the next)
  roofline : the kernel is bound on chip (shared-memory operand delivery), not by HBM and not by tensor cores
          (SURVEY.md 8(d)); `frac` = algorithmic operand bytes (cc*(2*2 B records + 8 B table entry) + tc*2 B) / kernel
          time against the shared-memory load bandwidth measured now by the library's probe; `frac_fetched` counts
          only what this kernel must fetch (the tail columns are O(1) here): cc*12 B
  cpu_baseline / --impl reference : the reference's own edgeBetween (oracle/_ref, the unmodified sources compiled in
          the build container) on a seeded random sample of band pairs, on one thread and on all host threads

N>1 (torchrun, one process per GPU): every rank holds the same host read set, asks the library's partitioner
(hc_partition_rows: cuts balanced by pair count, halo from the covered intervals) for its part, packs / uploads / scores
that part and returns its adjacency words to its own pinned host buffer. No data-path collective; the only
communication is the max/sum of the timings and counters at the end. `--single-process` drives the N GPUs from one
process through hc_multi_* instead (what the C++ binding does).
"""
import argparse
import json
import multiprocessing
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
PARAM_SETS = {"defaults + no_prob0": {}, "defaults (prob0 on)": {"noProb0": False},
              "defaults + allele-frequency map (-A, every 3rd position), prob0 on": {"noProb0": False, "simpson": 3}}
MAX_WORDS = 3 << 30                                         # adjacency words per library call (12 GiB)
SEED = 20261104


# ---- workloads -------------------------------------------------------------------------------------------------------
def _chunk(args):
    from haploclique_b200 import synth
    n, k, kw = args
    return synth.quasispecies(n, seed=SEED, frag_seed=SEED + 1000 * (k + 1), **kw)


def generate(name, n_frag=None, workers=1):
    """The whole host read set of workload `name` (what readBamFile would hand to the clique finder, position-sorted).
    C3 / C4 are drawn in chunks of 50k fragments (the same reads whatever `workers` is)."""
    from haploclique_b200 import synth
    from haploclique_b200.records import ReadBatch
    if name == "C5r":
        # a super-read round captured from the unmodified reference's own loop (tests/golden/make_superread_rounds.py): the
        # input of round 1 on 1000 synthetic fragments, 3529 records (3515 Clique_*), 90 quality codes
        b = ReadBatch.load(os.path.join(ROOT, "tests", "golden", "superreads_synth_it1.npz"))
        return b, dict(n_frag=b.n, frag_min=int(b.cig_len.reshape(-1, 2).sum(1).min()), frag_max=int(b.cig_len.reshape(-1, 2).sum(1).max()),
                       genome_len=2500, captured=True)
    if name == "C5":
        kw = dict(synth.C5)
        n = n_frag or kw["n_frag"]
        return synth.super_reads(n, **{k: v for k, v in kw.items() if k != "n_frag"}), dict(kw, n_frag=n)
    kw = dict(synth.CONFIGS[name])
    n = n_frag or kw["n_frag"]
    kw.pop("n_frag")
    if name == "C2":
        return synth.quasispecies(n, **kw), dict(kw, n_frag=n)
    jobs = [(min(50_000, n - s), k, kw) for k, s in enumerate(range(0, n, 50_000))]
    if workers > 1 and len(jobs) > 1:
        with multiprocessing.get_context("fork").Pool(min(workers, len(jobs))) as pool:
            parts = pool.map(_chunk, jobs)
    else:
        parts = [_chunk(j) for j in jobs]
    return ReadBatch.concat_sorted(parts), dict(kw, n_frag=n)


def shard(batch, rank, world):
    """Part `rank` of `world` of one host read set, as the library cuts it (hc_partition_rows): the reads
    [first, cuts[rank+1]) and the index of the first row this part scores."""
    from haploclique_b200 import edge
    if world == 1:
        return batch, 0, 0
    cuts, first = edge.partition_rows(batch, world)
    lo, hi = int(first[rank]), int(cuts[rank + 1])
    return batch.permute(np.arange(lo, hi)), int(cuts[rank]) - lo, lo


def workload_text(name, cfg):
    if name == "C5r":
        return ("C5r: super-read round captured from the reference's own loop (input of round 1 on 1000 synthetic fragments, genome 2500, "
                "seed 11): %d records of %d-%d ops, Clique_* names, 90 quality codes" % (cfg["n_frag"], cfg["frag_min"], cfg["frag_max"]))
    if name == "C5":
        return "C5: synthetic super-read round, %d records of %d-%d columns (70 %% Clique_*, posterior qualities up to char 126), genome %d, seed %d" % (
            cfg["n_frag"], cfg["frag_min"], cfg["frag_max"], cfg["genome_len"], SEED)
    what = "HIV-1 quasispecies" if name in ("C2", "C3") else "deep amplicon"
    return "%s: synthetic %s, 5 haplotypes, %d fragments 2x%d bp, genome %d, seed %d" % (
        name, what, cfg["n_frag"], cfg["read_len"], cfg["genome_len"], SEED)


def simpson_map(max_position, every):
    return {p: float(np.log10(0.3 + 0.6 * ((p * 37) % 101) / 101.0)) for p in range(1, max_position + 1, every)}


def calc_kwargs(max_position, param_set):
    kw = dict(PARAMS, maxPosition=max_position)
    over = dict(PARAM_SETS[param_set])
    every = over.pop("simpson", 0)
    kw.update(over)
    if every:
        kw["simpson_map"] = simpson_map(max_position, every)
    return kw


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


# ---- the CPU arm: the reference's own edgeBetween on a bounded sample ------------------------------------------------
def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def band_pair_sample(batch, n_pairs, window=20_000, seed=SEED):
    """A seeded random sample of `n_pairs` scored band pairs (covered intervals intersect) among `window` consecutive reads
    from the middle of the read set; pair indices are relative to the window."""
    k = min(window, batch.n)
    lo = max(0, batch.n // 2 - k // 2)
    sub = batch.permute(np.arange(lo, lo + k))
    cf, cl = sub.cov_first.astype(np.int64), sub.cov_last.astype(np.int64)
    pmax = np.maximum.accumulate(cl)
    col0 = np.minimum(np.searchsorted(pmax, cf, side="left"), np.arange(k))      # first earlier read that can reach row i
    width = np.arange(k) - col0
    acc = np.concatenate([[0], np.cumsum(width)])
    rng = np.random.default_rng(seed)
    want = int(min(n_pairs, acc[-1]))
    pi = np.zeros(0, np.int64)
    pj = np.zeros(0, np.int64)
    while pi.size < want:
        slot = rng.integers(0, acc[-1], int(want * 1.2) + 16)
        i = np.searchsorted(acc, slot, side="right") - 1
        j = col0[i] + (slot - acc[i])
        ok = (cl[i] >= cf[j]) & (cl[j] >= cf[i])
        pi, pj = np.concatenate([pi, i[ok]]), np.concatenate([pj, j[ok]])
    pi, pj = pi[:want].astype(np.int32), pj[:want].astype(np.int32)
    desc = "seeded random sample of %d scored band pairs among %d consecutive reads (%d..%d of %d) of the same read set" % (
        want, k, lo, lo + k, batch.n)
    return sub, pi, pj, desc


def run_cpu(batch, n_pairs, threads, max_position, param_set, single_thread_pairs=150_000):
    """Times the reference's own edgeBetween (oracle/_ref) -- or, if that library is absent, the C oracle port -- on all
    `threads` host threads and on one thread (a prefix of the same sample)."""
    sub, pi, pj, desc = band_pair_sample(batch, n_pairs)
    kw = calc_kwargs(max_position, param_set)
    from oracle import pyref
    out = {"unit": UNIT, "sample": desc, "cpu_model": cpu_model(), "host_threads_available": os.cpu_count(), "params": param_set}
    if pyref.available():
        R = pyref.Reference()
        R.add_batch(sub)
        R.set_params(**kw)
        R.time_pairs(pi[:20_000], pj[:20_000], threads)                         # warm-up
        secs = R.time_pairs(pi, pj, threads)[0]
        k1 = min(single_thread_pairs, pi.size)
        secs1 = R.time_pairs(pi[:k1], pj[:k1], 1)[0]
        out.update(kind="reference", cores=threads, value=pi.size / secs, seconds=secs,
                   single_thread={"value": k1 / secs1, "pairs": int(k1), "seconds": secs1})
    else:
        from oracle import pyoracle
        O = pyoracle.Oracle(sub, **kw)
        k1 = min(single_thread_pairs, pi.size)
        t = time.perf_counter()
        O.pairs(pi[:k1], pj[:k1])
        secs1 = time.perf_counter() - t
        out.update(kind="port", cores=1, value=k1 / secs1, seconds=secs1,
                   single_thread={"value": k1 / secs1, "pairs": int(k1), "seconds": secs1})
    return out


# ---- one measured workload on this rank's GPU(s) ---------------------------------------------------------------------
class Measured:
    """Scores rows [row_begin, n) of `local` on `devices` (one context each): resident / e2e / from-records steps."""

    def __init__(self, local, row_begin, max_position, param_set, devices, kernel):
        from haploclique_b200 import edge
        self.edge, self.local, self.row_begin = edge, local, row_begin
        kw = calc_kwargs(max_position, param_set)
        self.multi = len(devices) > 1
        self.calc = edge.NewEdgeCalculator(devices=devices, **kw) if self.multi else edge.NewEdgeCalculator(device=devices[0], **kw)
        self.calc.set_kernel({"auto": edge.KERNEL_AUTO, "generic": edge.KERNEL_GENERIC, "tiled": edge.KERNEL_TILED,
                              "tiled2": edge.KERNEL_TILED2, "streamed": edge.KERNEL_STREAMED}[kernel])
        t0 = time.perf_counter()
        self.packed = edge.PackedBatch(local)
        self.pack_s = time.perf_counter() - t0
        self.calc.upload(self.packed)
        row_end = local.n
        # the band goes through the library in calls of at most MAX_WORDS adjacency words (one call except for full-size
        # C4); inside a call the library scores row windows and overlaps their D2H with the next window's kernel
        _, off, self.nw = self.calc.band_layout(row_begin, row_end)
        cuts = [row_begin]
        while cuts[-1] < row_end:
            k0 = cuts[-1] - row_begin
            k1 = int(np.searchsorted(off, off[k0] + MAX_WORDS, side="right")) - 1
            cuts.append(row_begin + max(k1, k0 + 1))
        cuts[-1] = min(cuts[-1], row_end)
        self.calls = list(zip(cuts[:-1], cuts[1:]))
        words = max([int(off[b - row_begin] - off[a - row_begin]) for a, b in self.calls] + [1])
        self.pinned = edge.PinnedBuffer(words, np.uint32)

    def _score(self, **kw):
        tot = None
        for a, b in self.calls:
            st = self.calc.score_band(a, b, **kw).stats
            tot = dict(st) if tot is None else {k: tot[k] + st[k] for k in st}
        return tot

    def flush(self):
        for ctx in self.calc.contexts():
            self.calc.L.hc_l2_flush(ctx)

    def step_resident(self):
        self.flush()
        return self._score(fetch=False, amb_cap=0)

    def step_e2e(self, sync):
        self.flush()
        sync()
        t = time.perf_counter()
        self.calc.upload(self.packed)
        st = self._score(out=self.pinned.array, amb_cap=1024)
        return time.perf_counter() - t, st

    def step_e2e_records(self, sync):
        """As step_e2e, starting one call earlier: hc_pack of the AlignmentRecord fields (the SoA view the C++ adapter
        hands over) is inside the timed region too."""
        self.flush()
        sync()
        t = time.perf_counter()
        pk = self.edge.PackedBatch(self.local)
        self.calc.upload(pk)
        self._score(out=self.pinned.array, amb_cap=1024)
        s = time.perf_counter() - t
        self.calc.upload(self.packed)
        del pk                                              # back to the library's pinned-buffer pool
        return s


def rooflines(st, ms_kernel, n_gpus, lds_gbs, dmul_s, hbm_peak, hbm_bytes, peaks_from_file, traffic):
    secs = ms_kernel * 1e-3
    operand = st["sum_cc"] * 12 + st["sum_tc"] * 2          # SURVEY.md 8(d), per launch set (n_gpus GPUs side by side)
    fetched = st["sum_cc"] * 12                             # what this kernel has to deliver: its tails are O(1)
    smem = operand / secs / 1e9 / n_gpus
    smem_f = fetched / secs / 1e9 / n_gpus
    fp64 = st["sum_cc"] / secs / n_gpus
    hbm = hbm_bytes / secs / 1e9 / n_gpus
    main = {"bound": "smem", "achieved": smem, "peak": lds_gbs, "unit": "GB/s", "frac": smem / lds_gbs if lds_gbs else None,
            "achieved_fetched": smem_f, "frac_fetched": smem_f / lds_gbs if lds_gbs else None, "traffic": traffic,
            "peak_source": "hc_measure_peaks (ld.shared.v4 probe, this run, this GPU)",
            "peak_measured": {"lds_gbs": lds_gbs, "dmul_per_s": dmul_s},
            "note": "per GPU; frac: operand bytes = sum_cc*12 + sum_tc*2 (SURVEY.md 8(d)); frac_fetched: sum_cc*12 only (the tail "
                    "columns are evaluated in O(1) and never fetched); HBM and tensor cores do not bound this kernel, see rooflines"}
    allr = {"smem": {"achieved": smem, "peak": lds_gbs, "unit": "GB/s", "frac": smem / lds_gbs if lds_gbs else None},
            "fp64": {"achieved": fp64, "peak": dmul_s, "unit": "DMUL/s", "frac": fp64 / dmul_s if dmul_s else None},
            "hbm": {"achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks_from_file else "fallback"}}
    return main, allr


def also_leg(name, n_frag, param_set, device, peaks, steps, cores, with_cpu, kernel="auto"):
    """A secondary workload on one GPU, reported under `also` in the same JSON line (resident + e2e + roofline + CPU arm)."""
    import torch
    batch, cfg = generate(name, n_frag, workers=min(cores, 8))
    max_position = batch.max_position() + 16
    m = Measured(batch, 0, max_position, param_set, [device], kernel)
    sync = torch.cuda.synchronize
    for _ in range(3):
        m.step_resident()
    sync()
    kms, st = [], None
    for _ in range(steps):
        st = m.step_resident()
        kms.append(st["kernel_ms"])
    m.step_e2e(sync)
    e2e = [m.step_e2e(sync)[0] for _ in range(steps)]
    ms, ms_e2e = float(np.mean(kms)), float(np.mean(e2e)) * 1e3
    h2d, d2h = m.packed.nbytes(), m.nw * 4
    main, allr = rooflines(st, ms, 1, peaks["lds"], peaks["dmul"], peaks["hbm"], h2d + d2h, peaks["from_file"], (peaks["traffic"] or {}).get(name))
    kern = {0: "none", 1: "generic", 2: "tiled", 3: "tiled2", 4: "streamed"}[m.calc.last_kernel()]
    out = {"workload": workload_text(name, cfg), "params": param_set, "kernel": kern, "steps": steps, "warmup": 3,
           "value": st["pairs_scored"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
           "e2e": {"value": st["pairs_scored"] / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                   "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
           "pairs": {"scored": st["pairs_scored"], "walked": st["pairs_walked"], "edges": st["edges"], "sum_cc": st["sum_cc"],
                     "sum_tc": st["sum_tc"], "near_threshold_redecided": st["ambiguous"], "undefined": st["undefined"]},
           "gpu_launches": st["launches"] * steps, "roofline": main, "rooflines": allr}
    if with_cpu:
        out["cpu_baseline"] = run_cpu(batch, 300_000, cores, max_position, param_set, single_thread_pairs=60_000)
    del m
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "C2", "C3", "C4", "C5", "C5r"])
    ap.add_argument("--params", default="defaults + no_prob0", choices=sorted(PARAM_SETS))
    ap.add_argument("--n-frag", type=int, default=0, help="override the fragment count (debugging only)")
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "tiled", "tiled2", "streamed"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the secondary workloads (also.C2 / C4 / C5 / parameter sets)")
    ap.add_argument("--single-process", action="store_true", help="drive --gpus N devices from this one process (hc_multi_*)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    wname = "C3" if args.workload == "auto" else args.workload

    if args.impl == "reference":
        if rank != 0:
            return 0
        # the CPU arm scores a bounded sample; a 100k-fragment draw of the same mix has the same local read density
        from haploclique_b200 import synth
        full = 3529 if wname == "C5r" else (synth.C5 if wname == "C5" else synth.CONFIGS[wname])["n_frag"]
        n_draw = min(args.n_frag or full, 100_000)
        batch, cfg = generate(wname, n_draw, workers=min(cores, 8))
        cfg["n_frag"] = args.n_frag or full
        max_position = batch.max_position() + 16
        reps = []
        for _ in range(max(min(args.warmup, 1), 0) + max(args.steps, 1)):
            reps.append(run_cpu(batch, 1_200_000, cores, max_position, args.params))
        reps = reps[-max(args.steps, 1):]
        cb = dict(reps[-1])
        cb["value"] = float(np.mean([r["value"] for r in reps]))
        cb["single_thread"] = dict(cb["single_thread"], value=float(np.mean([r["single_thread"]["value"] for r in reps])))
        if n_draw != cfg["n_frag"]:
            cb["sample"] += "; drawn at %d fragments: pairs/s of the CPU scorer does not depend on the set size" % n_draw
        ms = float(np.mean([r["seconds"] for r in reps])) * 1e3
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_text(wname, cfg), "params": args.params},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return 0

    # ---- the host read set (before CUDA is touched: the chunk generator forks workers) and this rank's part of it
    t0 = time.perf_counter()
    n_parts = args.gpus if args.single_process else world
    batch, cfg = generate(wname, args.n_frag or None, workers=max(1, min(cores // max(world, 1), 16)))
    gen_s = time.perf_counter() - t0
    max_position = batch.max_position() + 16

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
    if args.single_process:
        local, row_begin, first = batch, 0, 0
        devices = list(range(args.gpus))
    else:
        local, row_begin, first = shard(batch, rank, world)
        devices = [local_rank]
    shard_s = time.perf_counter() - t0
    if world > 1:
        del batch
    m = Measured(local, row_begin, max_position, args.params, devices, args.kernel)

    def barrier():
        torch.cuda.synchronize()
        for d in devices:
            torch.cuda.synchronize(d)
        if world > 1:
            dist.barrier()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        m.step_resident()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    wall0 = time.perf_counter()
    kms, launches, st = [], 0, None
    for _ in range(args.steps):
        st = m.step_resident()
        kms.append(st["kernel_ms"])
        launches += st["launches"]
    barrier()
    wall = time.perf_counter() - wall0
    n_e2e = min(args.steps, 5)
    m.step_e2e(barrier)
    e2e_s = [m.step_e2e(barrier)[0] for _ in range(n_e2e)]
    barrier()
    m.step_e2e_records(barrier)
    m.step_e2e_records(barrier)                             # two untimed passes: the pinned-buffer pool is warm from the third on
    rec_s = [m.step_e2e_records(barrier) for _ in range(min(args.steps, 3))]
    barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    kernel_used = {0: "none", 1: "generic", 2: "tiled", 3: "tiled2", 4: "streamed"}[m.calc.last_kernel()]
    lds_gbs, dmul_s = edge.measure_peaks(local_rank) if rank == 0 else (0.0, 0.0)

    # whole-job numbers: units of all ranks / max-over-ranks time
    t_dev = torch.tensor([float(np.mean(kms)), float(np.mean(e2e_s)) * 1e3, wall / args.steps * 1e3, float(np.mean(rec_s)) * 1e3,
                          m.pack_s * 1e3, shard_s * 1e3], device="cuda", dtype=torch.float64)
    cnt = torch.tensor([st["pairs_scored"], st["pairs_total"], st["edges"], st["sum_cc"], st["sum_tc"], st["pairs_walked"],
                        m.packed.nbytes(), m.nw * 4, st["ambiguous"], st["undefined"], local.n - row_begin],
                       device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_kernel, ms_e2e, ms_wall, ms_rec, ms_pack, ms_shard = [float(x) for x in t_dev.tolist()]
    scored, total, edges, sum_cc, sum_tc, walked, h2d, d2h, amb, undef, rows = [int(x) for x in cnt.tolist()]

    if rank == 0:
        n_gpus = max(world, len(devices))
        pk = {}
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            pass
        peaks = {"lds": lds_gbs, "dmul": dmul_s, "hbm": float(pk.get("hbm_gbs", 6650.0)), "from_file": bool(pk), "traffic": traffic}
        secs = ms_kernel * 1e-3
        main_roof, all_roof = rooflines({"sum_cc": sum_cc, "sum_tc": sum_tc}, ms_kernel, n_gpus, lds_gbs, dmul_s, peaks["hbm"],
                                        h2d + d2h, peaks["from_file"], (traffic or {}).get(wname))
        if n_gpus == 1:
            sharding = "single GPU"
        else:
            sharding = ("%d row ranges cut by hc_partition_rows (balanced by band pairs, halo from the covered intervals; row blocks of the "
                        "replicated set where everything overlaps), %s, no collective" %
                        (n_gpus, "one process, hc_multi_*" if args.single_process else "one process per GPU"))
        out = {
            "metric": METRIC, "value": scored / secs, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_kernel, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_text(wname, cfg), "params": args.params, "kernel": kernel_used,
                       "l2": "flushed between steps (256 MiB memset)", "library_calls_per_step": len(m.calls), "sharding": sharding},
            "pairs": {"scored": scored, "examined": total, "walked": walked, "edges": edges, "sum_cc": sum_cc, "sum_tc": sum_tc,
                      "rows": rows, "all_pairs": rows * (rows - 1) // 2, "near_threshold_redecided": amb, "undefined": undef},
            "all_pairs_per_s": rows * (rows - 1) / 2 / secs,
            "column_steps_per_s": (sum_cc + sum_tc) / secs,
            "wall_ms_per_step": ms_wall,
            "e2e": {"value": scored / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e, "steps": n_e2e, "host_pack_ms_once": ms_pack, "host_partition_ms_once": ms_shard,
                    "host_generate_s_once": gen_s,
                    "from_records": {"value": scored / (ms_rec * 1e-3), "unit": UNIT, "ms_per_step": ms_rec,
                                     "note": "hc_pack of the record fields + hc_upload + hc_score_band + D2H per step"}},
            "gpu_launches": launches * (world if not args.single_process else 1),
            "roofline": main_roof, "rooflines": all_roof,
            "clocks": sampler.summary(),
        }
        del m
        if n_gpus == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = run_cpu(local, 1_200_000, cores, max_position, args.params)
        if n_gpus == 1 and not args.no_also and args.workload == "auto":
            also = {}
            k = 5 if args.steps >= 5 else max(args.steps, 1)
            with_cpu = not args.no_cpu_baseline
            also["C2"] = also_leg("C2", None, "defaults + no_prob0", local_rank, peaks, k, cores, with_cpu)
            also["C2, prob0 on"] = also_leg("C2", None, "defaults (prob0 on)", local_rank, peaks, min(k, 3), cores, with_cpu)
            also["C2, allele-frequency map"] = also_leg("C2", None, "defaults + allele-frequency map (-A, every 3rd position), prob0 on",
                                                        local_rank, peaks, min(k, 3), cores, with_cpu)
            also["C4"] = also_leg("C4", 200_000, "defaults + no_prob0", local_rank, peaks, min(k, 3), cores, with_cpu)
            also["C5"] = also_leg("C5", None, "defaults + no_prob0", local_rank, peaks, k, cores, with_cpu)
            also["C5r"] = also_leg("C5r", None, "defaults + no_prob0", local_rank, peaks, k, cores, with_cpu)
            out["also"] = also
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
