#!/usr/bin/env python
"""End-to-end wall clock, BAM -> FASTA: the unmodified reference CLI (oracle/_ref/haploclique_ref) against the same CLI with
the GPU calculator behind the EdgeCalculator interface (oracle/_ref/haploclique_gpu = integration/haploclique_gpu_main.cpp),
on synthetic quasispecies BAMs written with the reference's vendored BamWriter. Checks that both write the same FASTA.
BASELINE.json metric: "... e2e wall vs CPU". The clique finders, super-read merging and I/O stay on the host in both."""
import json, os, subprocess, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from haploclique_b200 import synth
from oracle import pyref

out = []
for n_frag, genome in ((500, 2500), (1000, 2500)):     # the reference CLI needs 30 s for 1000 fragments, minutes beyond
    q = synth.quasispecies(n_frag, genome_len=genome, seed=11)
    with tempfile.TemporaryDirectory() as tmp:
        bam = os.path.join(tmp, "reads.bam")
        W = pyref.Reference()
        W.bam_writer_open(bam, "ref", genome + 200)
        W.add_batch(q)
        W.bam_writer_close()
        res = {"fragments": n_frag, "records": q.n, "genome": genome}
        fasta = {}
        for tag, exe in (("reference_cpu", pyref.CLI_REF), ("gpu_calculator", pyref.CLI_GPU)):
            d = os.path.join(tmp, tag)
            os.mkdir(d)
            best = None
            for rep in range(1):
                t = time.perf_counter()
                r = subprocess.run([exe, "--no_prob0", "--no_singletons", "--significance=4", bam], cwd=d, capture_output=True, text=True, timeout=1500)
                dt = time.perf_counter() - t
                assert r.returncode == 0, r.stderr[-400:]
                best = dt if best is None else min(best, dt)
            res[tag + "_wall_s"] = round(best, 3)
            fasta[tag] = open(os.path.join(d, "quasispecies.fasta"), "rb").read()
            res[tag + "_edges_round0"] = next((ln for ln in r.stdout.splitlines() if "edges:" in ln), "")
        res["fasta_identical"] = fasta["reference_cpu"] == fasta["gpu_calculator"] and len(fasta["reference_cpu"]) > 0
        res["speedup"] = round(res["reference_cpu_wall_s"] / res["gpu_calculator_wall_s"], 2)
        print(json.dumps(res), flush=True)
        out.append(res)
