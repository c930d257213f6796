"""GPU: the reference-side binding (integration/GpuEdgeCalculator, compiled against the reference tree into
oracle/_ref/libhcref_gpu.so and oracle/_ref/haploclique_gpu) driving the reference's own clique finders: whole-program
parity -- same edges, same cliques, same super-reads, same FASTA as the unmodified reference on the same input."""
import os
import re
import subprocess

import numpy as np
import pytest

from haploclique_b200 import synth
from haploclique_b200.records import ReadBatch
from oracle import pyref
from tests import common

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not pyref.gpu_adapter_available(), reason="oracle/_ref/libhcref_gpu.so not built")]


def _same_batch(a: ReadBatch, b: ReadBatch):
    assert a.n == b.n
    for f in ReadBatch._FIELDS:
        assert (getattr(a, f) == getattr(b, f)).all(), f
    assert a.names == b.names


@pytest.mark.parametrize("use_bk", [False, True])
def test_clique_rounds_with_gpu_calculator_equal_the_reference(use_bk):
    # the reference's cram test input (arabis, 1104 paired records) and parameters, all rounds until convergence
    b = common.load_batch("arabis_records.npz")
    params = dict(common.WEAK, noProb0=True)
    ref, gpu = pyref.Reference(), pyref.Reference(gpu_adapter=True)     # unpatched finders / row-fed finders (integration/*_rows.patch)
    for R in (ref, gpu):
        R.add_batch(b)
        R.set_params(**params)
    ct_r = ct_g = 0
    sd_r = sd_g = 1.0
    rounds = []
    for it in range(6):
        n_r, ct_r, e_r, conv_r, sd_r = ref.iterate(ct_r, sd_r, use_bk=use_bk, filter_singletons=True, significance=4.0)
        n_g, ct_g, e_g, conv_g, sd_g, adj_edges = gpu.iterate_gpu(ct_g, sd_g, use_bk=use_bk, filter_singletons=True, significance=4.0)
        rounds.append((n_r, e_r))
        assert (n_g, e_g, conv_g) == (n_r, e_r, conv_r), "round %d" % it
        assert sd_g == sd_r
        if not use_bk:
            assert adj_edges == e_r                  # CLEVER asks about every earlier read: its count = the adjacency's
        _same_batch(ref.to_batch(), gpu.to_batch())  # the next round's super-reads
        if conv_r:
            break
    if not use_bk:
        assert rounds[0] == (447, 153685) and rounds[1][1] == 48675     # 1104 -> 447 -> 42 -> 2


def test_unit_expectations_through_the_adapter(golden):
    b = common.load_batch("unit_fixtures.npz")
    R = pyref.Reference(gpu_adapter=True)
    R.add_batch(b)
    R.set_params(maxPosition=20000)
    for name, case in golden["unit2"].items():
        assert R.edge_gpu(case["ap1"], case["ap2"]) == case["expected"] == R.edge(case["ap1"], case["ap2"]), name


def _run_cli(exe, d, finder, flags, bam, env=None):
    d.mkdir()
    cmd = [exe] + ([finder] if finder == "bronkerbosch" else []) + flags + ["--log=" + str(d / "edges.log"), bam]
    r = subprocess.run(cmd, cwd=str(d), capture_output=True, text=True, timeout=600, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stderr[-500:]
    return (open(d / "quasispecies.fasta", "rb").read(),
            [ln for ln in r.stdout.splitlines() if "edges:" in ln or ln.startswith("final")],
            open(d / "edges.log", "rb").read(), r.stdout)


def _clis_on_bam(tmp_path, bam, flags, finder):
    """The unmodified CLI and the GPU-backed CLI on one BAM: same per-round stdout lines (read and edge counts), byte-identical
    FASTA and byte-identical --log dumps (every iteration's adjacency lists and cliques, src/LogWriter.h:26-70) -- with the
    finders' edge loops fed from adjacency rows (integration/*_rows.patch, the default), with one edgeBetween call per stored
    alignment (HC_GPU_NO_ROWS), and with the rows self-checked against the per-pair answers in every round."""
    ref = _run_cli(pyref.CLI_REF, tmp_path / "ref", finder, flags, bam)
    assert len(ref[0]) > 0
    rows = _run_cli(pyref.CLI_GPU, tmp_path / "rows", finder, flags, bam)
    pairs = _run_cli(pyref.CLI_GPU, tmp_path / "pairs", finder, flags, bam, {"HC_GPU_NO_ROWS": "1"})
    check = _run_cli(pyref.CLI_GPU, tmp_path / "check", finder, flags, bam, {"HC_GPU_CHECK_ROWS": "1"})
    # `rows` also builds every round's super-reads in one GPU batch (integration/CliqueCollector_batch.patch); `hostfold` keeps
    # the reference's one-by-one constructor: same output either way
    hostfold = _run_cli(pyref.CLI_GPU, tmp_path / "hostfold", finder, flags, bam, {"HC_GPU_NO_MERGE": "1"})
    assert "gpu merge kernels" in rows[3] and "gpu merge kernels" not in hostfold[3]
    for tag, got in (("rows", rows), ("pairs", pairs), ("check", check), ("hostfold", hostfold)):
        assert got[0] == ref[0], tag + ": FASTA differs"
        lines = [re.sub(r"\trows ok: \d+ neighbours", "", ln) for ln in got[1]]      # the self-check's own note
        assert lines == ref[1], tag + ": per-round counts differ"
        assert got[2] == ref[2] and len(ref[2]) > 0, tag + ": --log adjacency / cliques differ"
    assert "rows ok" in check[3] and "MISMATCH" not in check[3]
    return ref


def _both_clis(tmp_path, batch, flags, finder, genome_len):
    """Writes `batch` as a BAM with the reference's vendored BamWriter and runs _clis_on_bam on it."""
    W = pyref.Reference()
    bam = str(tmp_path / "reads.bam")
    W.bam_writer_open(bam, "ref", genome_len)
    W.add_batch(batch)
    W.bam_writer_close()
    return _clis_on_bam(tmp_path, bam, flags, finder)[1]


@pytest.mark.parametrize("finder", ["clever", "bronkerbosch"])
def test_the_references_cram_vector_through_the_gpu_cli(tmp_path, finder):
    """The reference's own end-to-end test (test/cram/arabis_large_500_01.t): its reads.bam, its flags, its expected
    quasispecies.fasta.fasta -- both committed under tests/golden/ -- through haploclique_gpu (GPU edges, row-fed finders)."""
    if not (os.path.exists(pyref.CLI_REF) and os.path.exists(pyref.CLI_GPU)):
        pytest.skip("CLI binaries not built")
    bam = os.path.join(common.GOLDEN, "arabis_large_500_01_reads.bam")
    flags = ["--edge_quasi_cutoff_cliques=0.85", "--edge_quasi_cutoff_mixed=0.85", "--edge_quasi_cutoff_single=0.8",
             "--min_overlap_cliques=0.6", "--min_overlap_single=0.5", "--no_singletons", "--significance=4", "--no_prob0"]
    ref = _clis_on_bam(tmp_path, bam, flags, finder)
    assert ref[0] == open(os.path.join(common.GOLDEN, "arabis_large_500_01_quasispecies.fasta"), "rb").read()
    if finder == "clever":
        assert "edges: 153685" in ref[1][0]


@pytest.mark.parametrize("finder", ["clever", "bronkerbosch"])
def test_cli_fasta_identical(tmp_path, finder):
    if not (os.path.exists(pyref.CLI_REF) and os.path.exists(pyref.CLI_GPU)):
        pytest.skip("CLI binaries not built")
    q = synth.quasispecies(300, genome_len=1200, seed=9, indel_err=1e-3, clip_frac=0.05)
    # default cutoffs (the weak cram-test ones make CLEVER itself run for minutes on this input); 7 rounds, ~2 s on CPU
    _both_clis(tmp_path, q, ["--no_prob0", "--no_singletons", "--significance=4"], finder, 1300)


@pytest.mark.parametrize("no_prob0", [False, True])
def test_cli_fasta_identical_on_the_bundled_hiv_reads(tmp_path, no_prob0):
    """BASELINE configs[0]: the records of the reference's bundled test/data/simulation/reads_HIV-1_50_01.bam (1836 after
    mate merging, tests/golden/hiv50_records.npz), default NewEdgeCalculator parameters: BAM -> FASTA through both CLIs."""
    if not (os.path.exists(pyref.CLI_REF) and os.path.exists(pyref.CLI_GPU)):
        pytest.skip("CLI binaries not built")
    b = common.load_batch("hiv50_records.npz")
    flags = (["--no_prob0"] if no_prob0 else []) + ["--no_singletons", "--significance=4"]
    lines = _both_clis(tmp_path, b, flags, "clever", b.max_position() + 100)
    assert ("edges: 19233" if no_prob0 else "edges: 220") in lines[0]           # iteration 0 (220)


@pytest.mark.parametrize("fixture", ["hiv50_records.npz", "arabis_records.npz"])
def test_quasispecies_adapter_equals_the_reference_class(fixture):
    """integration/'s GpuQuasispeciesEdgeCalculator (scoreBatch + edgeBetween answered from the adjacency words) against
    the reference's own QuasispeciesEdgeCalculator::edgeBetween on the bundled reads. The reference does not terminate
    on some clipped pairs (src/QuasispeciesEdgeCalculator.cpp:258-267): it is only asked about pairs the oracle walks to
    the end; on the others the adapter must say "no edge"."""
    from oracle import pyoracle
    b = common.load_batch(fixture)
    kw = dict(overlap_single=0.3, overlap=0.5, edge_quasi_cutoff_single=0.8, edge_quasi_cutoff_mixed=0.85, edge_quasi_cutoff=0.9)
    pi, pj = common.quasi_pair_sample(b.n)
    later = pi > pj                                     # the finders' call order: (new arrival, earlier arrival)
    pi, pj = pi[later], pj[later]
    orc = pyoracle.QuasiOracle(b, maxPosition=b.max_position(), **kw).pairs(pi, pj)
    R = pyref.Reference(gpu_adapter=True)
    R.add_batch(b)
    R.set_params(maxPosition=b.max_position())
    dev = R.quasi_gpu_pairs(pi, pj, **kw)
    ok = orc["status"] != 3
    assert not dev[~ok].any()
    q, ref_edge = R.quasi_pairs(pi[ok], pj[ok], **kw)
    near = np.abs(q - orc["cutoff"][ok]) <= 1e-12 * orc["cutoff"][ok]
    assert ((dev[ok] == ref_edge) | near).all()
    assert int(ref_edge.sum()) > 100 and int(ok.sum()) > 10000
