#!/usr/bin/env python
"""End-to-end wall clock, BAM -> FASTA (BASELINE.json metric: "... e2e wall vs CPU"), three arms on the same BAM:
  reference_cpu   the unmodified reference CLI (oracle/_ref/haploclique_ref)
  gpu_pairs       the same CLI with the GPU calculator behind EdgeCalculator, the finders asking edgeBetween once per stored
                  alignment as the unpatched reference does (HC_GPU_NO_ROWS=1)
  gpu_rows        the GPU calculator with the finders' edge loops fed from adjacency rows (integration/*_rows.patch)
on synthetic quasispecies BAMs written with the reference's vendored BamWriter. Checks that all arms write the same FASTA.
The clique maintenance, super-read merging and I/O stay on the host in all arms. An arm that exceeds its time limit is
reported as such. Usage: cli_wall.py [fragments ...]   (default 1000 3000 10000); writes gpurun_out/cli_wall.json"""
import json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from haploclique_b200 import synth
from oracle import pyref

LIMIT = float(os.environ.get("HC_CLI_WALL_LIMIT", "240"))
sizes = [int(a) for a in sys.argv[1:]] or [1000, 3000, 10000]
out = []
for n_frag in sizes:
    genome = 2500 if n_frag <= 3000 else 9719
    q = synth.quasispecies(n_frag, genome_len=genome, seed=11)
    with tempfile.TemporaryDirectory() as tmp:
        bam = os.path.join(tmp, "reads.bam")
        W = pyref.Reference()
        W.bam_writer_open(bam, "ref", genome + 200)
        W.add_batch(q)
        W.bam_writer_close()
        res = {"fragments": n_frag, "records": q.n, "genome": genome, "limit_s": LIMIT}
        fasta = {}
        for tag, exe, env in (("gpu_rows", pyref.CLI_GPU, {}), ("gpu_pairs", pyref.CLI_GPU, {"HC_GPU_NO_ROWS": "1"}),
                              ("reference_cpu", pyref.CLI_REF, {})):
            d = os.path.join(tmp, tag)
            os.mkdir(d)
            t = time.perf_counter()
            try:
                r = subprocess.run([exe, "--no_prob0", "--no_singletons", "--significance=4", bam], cwd=d, capture_output=True,
                                   text=True, timeout=LIMIT, env=dict(os.environ, **env))
            except subprocess.TimeoutExpired:
                res[tag + "_wall_s"] = None
                res[tag + "_note"] = "exceeded %.0f s" % LIMIT
                continue
            dt = time.perf_counter() - t
            assert r.returncode == 0, r.stderr[-400:]
            res[tag + "_wall_s"] = round(dt, 3)
            fasta[tag] = open(os.path.join(d, "quasispecies.fasta"), "rb").read()
            res[tag + "_rounds"] = [ln.strip() for ln in r.stdout.splitlines() if "edges:" in ln]
            tl = [ln for ln in r.stdout.splitlines() if ln.startswith("time:")]
            res[tag + "_time_line"] = tl[-1] if tl else ""
        vals = list(fasta.values())
        res["fasta_identical"] = len(vals) > 0 and all(v == vals[0] for v in vals) and len(vals[0]) > 0
        res["arms_compared"] = sorted(fasta)
        for a, b in (("reference_cpu", "gpu_rows"), ("gpu_pairs", "gpu_rows")):
            if res.get(a + "_wall_s") and res.get(b + "_wall_s"):
                res["speedup_%s_over_%s" % (b, a)] = round(res[a + "_wall_s"] / res[b + "_wall_s"], 2)
        print(json.dumps(res), flush=True)
        out.append(res)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "cli_wall.json"), "w") as fh:
    json.dump(out, fh, indent=1)
