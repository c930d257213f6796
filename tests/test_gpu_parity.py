"""GPU: the CUDA path (through the C ABI) against the oracle on the same inputs, and against the golden digests
of the unmodified reference. Bit-exact on status/gap/cc/tc/probM/prob0; the edge decision may differ only on
pairs the device itself reports as within 1e-12 (relative) of the threshold."""
import numpy as np
import pytest

from haploclique_b200 import edge, synth
from oracle import pyoracle
from tests import common

pytestmark = pytest.mark.gpu


def _calc(params, max_position):
    kw = common.calc_kwargs(params, max_position)
    return edge.NewEdgeCalculator(**kw)


@pytest.mark.parametrize("label", sorted(common.DATASETS))
def test_pairs_match_oracle_and_golden(golden, label):
    g = golden[label]
    b = common.load_batch(common.DATASETS[label])
    calc = _calc(g["params"], g["max_position"])
    calc.upload(b)
    pi, pj = common.tri_pairs(b.n)
    dev = calc.score_pairs(pi, pj)
    orc = pyoracle.Oracle(b, **common.calc_kwargs(g["params"], g["max_position"])).pairs(pi, pj)
    differing = common.assert_pairs_equal(dev, orc, label)
    h_edge, h_detail = common.pair_digests(dev["status"], dev["edge"], dev["gap"], dev["cc"], dev["tc"],
                                           dev["probM"], dev["prob0"])
    assert h_detail == g["sha256_detail"]
    if differing == 0:
        assert h_edge == g["sha256_edges"] and int(dev["edge"].sum()) == g["edges"]


@pytest.mark.parametrize("label", ["hiv50_default", "arabis_weak_noprob0", "superreads_arabis_it1"])
def test_band_adjacency_matches_oracle(golden, label):
    g = golden[label]
    b = common.load_batch(common.DATASETS[label])
    calc = _calc(g["params"], g["max_position"])
    adj = calc.scoreBatch(b)
    edges, e = pyoracle.Oracle(b, **common.calc_kwargs(g["params"], g["max_position"])).all_pairs(detail=False)
    pi, pj = common.tri_pairs(b.n)
    ai, aj = adj.edges()
    assert adj.stats["ambiguous"] == 0
    assert adj.count() == edges == g["edges"]
    assert (ai == pi[e == 1]).all() and (aj == pj[e == 1]).all()
    assert adj.stats["edges"] == edges
    assert adj.stats["pairs_scored"] == g["prefilter_pass"] + int(0) or adj.stats["pairs_walked"] == g["prefilter_pass"]
    assert adj.stats["pairs_walked"] == g["prefilter_pass"]
    assert adj.stats["sum_cc"] >= g["sum_cc"] and adj.stats["undefined"] == 0
    # spot-check the lookup the clique finders would do
    rng = np.random.default_rng(0)
    for k in rng.integers(0, pi.shape[0], 200):
        assert adj.edge(int(pi[k]), int(pj[k])) == bool(e[k])


def test_reference_unit_expectations_through_edgeBetween(golden):
    b = common.load_batch("unit_fixtures.npz")
    recs = b.records()
    calc = edge.NewEdgeCalculator(0.9, 0.99, 0.9, False, {}, 0.95, 0.6, 0.97, 20000, False)   # unit2.cpp:23-34 order
    for name, case in golden["unit2"].items():
        assert calc.edgeBetween(recs[case["ap1"]], recs[case["ap2"]]) == case["expected"], name


@pytest.mark.parametrize("seed", range(8))
def test_zoo_all_layouts_both_orders(seed):
    b = synth.zoo(140, seed=seed)
    kw = dict(noProb0=bool(seed % 2), edge_quasi_cutoff_single=0.8, edge_quasi_cutoff_mixed=0.85,
              edge_quasi_cutoff=0.9, overlap=0.3, overlap_single=0.2, maxPosition=b.max_position())
    pi, pj = common.tri_pairs(b.n)
    pi, pj = np.concatenate([pi, pj]), np.concatenate([pj, pi])        # also ap1 = the EARLIER read
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    dev = calc.score_pairs(pi, pj)
    orc = pyoracle.Oracle(b, **kw).pairs(pi, pj)
    common.assert_pairs_equal(dev, orc, "zoo%d" % seed)
    assert set(np.unique(orc["layout"])) >= {100, 200, 300, 400}


@pytest.mark.parametrize("noprob0", [False, True])
def test_synthetic_quasispecies_with_allele_frequencies(noprob0):
    b = synth.quasispecies(600, genome_len=1800, seed=11, indel_err=2e-3, clip_frac=0.1)
    rng = np.random.default_rng(5)
    simpson = {int(p): float(np.log10(rng.uniform(0.3, 1.0))) for p in rng.choice(np.arange(1, 1800), 900, replace=False)}
    kw = dict(noProb0=noprob0, simpson_map=simpson, maxPosition=max(b.max_position(), max(simpson)), edge_quasi_cutoff_single=0.9)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    pi, pj = common.tri_pairs(b.n)
    dev = calc.score_pairs(pi, pj)
    orc = pyoracle.Oracle(b, **kw).pairs(pi, pj)
    common.assert_pairs_equal(dev, orc, "simpson")
    adj = calc.score_band()
    assert adj.count() == int(orc["edge"].sum())


def test_long_overlap_underflow_is_reproduced():
    # 6000-column identical low-quality reads: the fp64 product underflows to 0 -> log10 = -inf -> no edge,
    # which a log-domain sum would get wrong (SURVEY.md 7 "hard parts")
    from haploclique_b200.records import AlignmentRecord, ReadBatch
    n = 6000
    rng = np.random.default_rng(3)
    seq = bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), n))
    a = AlignmentRecord("Clique_1", True, 1, n, b"M" * n, seq, bytes([33 + 3]) * n)
    c = AlignmentRecord("Clique_2", True, 1, n, b"M" * n, seq, bytes([33 + 3]) * n)
    b = ReadBatch.from_records([a, c])
    kw = dict(edge_quasi_cutoff=0.1, overlap=0.5, noProb0=True, maxPosition=n)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    dev = calc.score_pairs([1], [0])
    orc = pyoracle.Oracle(b, **kw).pairs([1], [0])
    common.assert_pairs_equal(dev, orc, "underflow")
    assert dev["probM"][0] == 0.0 and dev["edge"][0] == 0 and dev["cc"][0] == n


def test_empty_ragged_and_tiny_batches():
    from haploclique_b200.records import ReadBatch
    b = common.load_batch("unit_fixtures.npz")
    calc = edge.NewEdgeCalculator(maxPosition=20000)
    for sel in ([], [3], [0, 0], list(range(10))):
        sub = b.select(sel) if sel else ReadBatch.from_records([])
        adj = calc.scoreBatch(sub)
        edges, _ = pyoracle.Oracle(sub, maxPosition=20000).all_pairs(detail=False) if sub.n else (0, None)
        assert adj.count() == edges
    with pytest.raises(edge.HcError):
        calc.upload(b.select([0]))
        calc.score_pairs([5], [0])


def test_setOverlapCliques_changes_the_decision():
    b = common.load_batch("superreads_arabis_it2.npz")
    kw = dict(common.WEAK, noProb0=True, maxPosition=b.max_position())
    calc = edge.NewEdgeCalculator(**kw)
    e0 = calc.scoreBatch(b).count()
    calc.setOverlapCliques(0.99)
    e1 = calc.score_band().count()
    orc_kw = dict(kw, overlap=0.99)
    edges, _ = pyoracle.Oracle(b, **orc_kw).all_pairs(detail=False)
    assert e1 == edges and e1 <= e0


# ---- the tiled band kernel (shared-memory column tiles, TMA-fed rows): every band slot against the oracle ----
TILED_KERNELS = [pytest.param(edge.KERNEL_TILED, id="tiled"), pytest.param(edge.KERNEL_TILED2, id="tiled2"),
                 pytest.param(edge.KERNEL_STREAMED, id="streamed")]


def _tiled_detail_vs_oracle(b, kw, label, kernel=edge.KERNEL_TILED):
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    calc.set_kernel(kernel)
    adj, pi, pj, det = calc.score_band_detail()
    assert calc.last_kernel() == kernel or b.n < 2
    O = pyoracle.Oracle(b, **kw)
    orc = O.pairs(pi, pj)
    common.assert_pairs_equal(det, orc, label)
    # the adjacency words say the same as the per-slot detail, and nothing outside the band is an edge
    edges, e = O.all_pairs(detail=False)
    ai, aj = adj.edges()
    ti, tj = common.tri_pairs(b.n)
    differing = int(((det["edge"] != orc["edge"]) & (orc["status"] != 3)).sum())
    if differing == 0:
        assert adj.count() == edges
        assert (ai == ti[e == 1]).all() and (aj == tj[e == 1]).all()
    assert adj.stats["pairs_total"] >= adj.stats["pairs_scored"] == int((orc["status"] >= 1).sum())
    assert adj.stats["pairs_walked"] == int((orc["status"] == 2).sum())
    assert adj.stats["sum_cc"] == int(orc["cc"][orc["status"] == 2].sum())
    assert adj.stats["sum_tc"] == int(orc["tc"][orc["status"] == 2].sum())
    # same answer from the generic kernel
    calc.set_kernel(edge.KERNEL_GENERIC)
    adj2 = calc.score_band()
    assert calc.last_kernel() == edge.KERNEL_GENERIC or b.n < 2
    assert (adj2.words == adj.words).all()
    return det, orc


@pytest.mark.parametrize("kernel", TILED_KERNELS)
@pytest.mark.parametrize("label", sorted(common.DATASETS))
def test_tiled_kernel_matches_oracle_on_golden_sets(golden, label, kernel):
    g = golden[label]
    b = common.load_batch(common.DATASETS[label])
    kw = common.calc_kwargs(g["params"], g["max_position"])
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    calc.set_kernel(kernel)
    try:
        adj = calc.score_band()
    except edge.HcError:
        pytest.skip("batch not eligible for this tiled kernel (tables + longest read do not fit a CTA's shared memory)")
    assert calc.last_kernel() == kernel
    _tiled_detail_vs_oracle(b, kw, label, kernel)
    assert adj.count() == g["edges"]


@pytest.mark.parametrize("kernel", TILED_KERNELS)
@pytest.mark.parametrize("seed", [1, 2])
def test_hard_clipped_reads(seed, kernel):
    """Hard clips outside soft clips (H S core S H) and indels next to the clip runs: every band slot of every band
    kernel against the oracle (itself checked against the reference on the same generator, test_oracle_vs_reference)."""
    b = synth.hard_clipped(220, seed=seed)
    for noprob0 in (False, True):
        kw = dict(common.WEAK, noProb0=noprob0, maxPosition=b.max_position() + 64)
        det, orc = _tiled_detail_vs_oracle(b, kw, "hardclip%d" % seed, kernel)
        assert int((orc["status"] == 2).sum()) > 2000 and int(orc["gap"].sum()) > 100


def test_first_generation_tile_shrinks_its_cta_to_fit():
    """Early super-read rounds: reads of a few hundred columns with ~75 quality codes. Too many codes for the
    second-generation kernel's 9-row table; the first-generation kernel's tables + tile + 24 warps' row buffers exceed a
    CTA's shared memory by a little -- it runs with fewer warps instead of handing the batch to the streamed kernel."""
    b = synth.super_reads(500, genome_len=2500, read_len=450, frag_mean=700.0, frag_sd=120.0, frag_min=300, frag_max=880, seed=31)
    assert len(np.unique(b.qual)) > 70 and int(b.cig_len.reshape(-1, 2).sum(1).max()) > 870      # 243 KB at 24 warps, 228 KB at 20
    kw = dict(noProb0=True, maxPosition=b.max_position() + 10, edge_quasi_cutoff=0.9, edge_quasi_cutoff_mixed=0.9,
              edge_quasi_cutoff_single=0.9, overlap=0.5, overlap_single=0.5)
    calc = edge.NewEdgeCalculator(**kw)
    calc.scoreBatch(b)
    assert calc.last_kernel() == edge.KERNEL_TILED
    det, orc = _tiled_detail_vs_oracle(b, kw, "tile1-shrunk", edge.KERNEL_TILED)
    assert int(orc["edge"].sum()) > 100


def test_auto_prefers_the_second_generation_tiled_kernel():
    b = synth.quasispecies(300, genome_len=1500, seed=2)
    calc = edge.NewEdgeCalculator(noProb0=True, maxPosition=b.max_position())
    calc.scoreBatch(b)
    assert calc.last_kernel() == edge.KERNEL_TILED2


@pytest.mark.parametrize("kernel", TILED_KERNELS)
@pytest.mark.parametrize("seed", range(6))
def test_tiled_kernel_zoo(seed, kernel):
    # 92 quality codes for the first tiled kernel; the second one takes 22 codes (split-word table) on even seeds
    # and 44 codes (table of doubles) on odd seeds -- its 7-row table of 92 codes would not fit shared memory
    wide = (2, 94) if kernel != edge.KERNEL_TILED2 else ((20, 42) if seed % 2 == 0 else (2, 46))
    b = synth.zoo(200, seed=100 + seed, wide_qual=wide)
    kw = dict(noProb0=bool(seed % 2), edge_quasi_cutoff_single=0.8, edge_quasi_cutoff_mixed=0.85,
              edge_quasi_cutoff=0.9, overlap=0.3, overlap_single=0.2, maxPosition=b.max_position())
    det, orc = _tiled_detail_vs_oracle(b, kw, "tiled-zoo%d" % seed, kernel)
    assert set(np.unique(orc["layout"])) >= {100, 200, 300, 400}


@pytest.mark.parametrize("kernel", TILED_KERNELS)
def test_tiled_kernel_quasispecies_with_indels_and_clips(kernel):
    b = synth.quasispecies(1500, genome_len=2500, seed=21, indel_err=3e-3, clip_frac=0.15)
    kw = dict(noProb0=True, maxPosition=b.max_position(), edge_quasi_cutoff_single=0.9)
    det, orc = _tiled_detail_vs_oracle(b, kw, "tiled-quasi", kernel)
    assert int(orc["edge"].sum()) > 1000 and int(orc["gap"].sum()) > 100
    kw["noProb0"] = False
    _tiled_detail_vs_oracle(b, kw, "tiled-quasi-prob0", kernel)


@pytest.mark.parametrize("kernel", TILED_KERNELS)
def test_tiled_kernel_row_windows_concatenate(kernel):
    # sharding by row windows (what N>1 ranks do): the windows' words are the full run's words
    b = synth.quasispecies(900, genome_len=2000, seed=5)
    kw = dict(noProb0=True, maxPosition=b.max_position())
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    calc.set_kernel(kernel)
    full = calc.score_band()
    parts = [calc.score_band(r0, r1) for r0, r1 in ((0, 301), (301, 650), (650, 900))]
    assert sum(p.count() for p in parts) == full.count()
    fi, fj = full.edges()
    pi_ = np.concatenate([p.edges()[0] for p in parts])
    pj_ = np.concatenate([p.edges()[1] for p in parts])
    assert (fi == pi_).all() and (fj == pj_).all()


def test_full_size_C2_all_band_kernels_agree():
    """BASELINE configs[1] at full size (100k fragments, 4.7e8 scored pairs): the oracle cannot walk that many pairs in
    test time, so the check is a size-independent property -- the four band kernels (generic one-lane-per-pair walker,
    first- and second-generation tiled kernels, streamed kernel; the walker shares no walking code with the others) produce the same
    adjacency words and the same column totals, row windows concatenate to the full run, and a sample of 200k band
    pairs matches the oracle bit for bit."""
    import bench
    b = synth.quasispecies(100_000, **{k: v for k, v in synth.CONFIGS["C2"].items() if k != "n_frag"})
    kw = dict(bench.PARAMS, maxPosition=b.max_position() + 2000)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    res = {}
    for name, mode in (("tiled2", edge.KERNEL_TILED2), ("tiled", edge.KERNEL_TILED), ("streamed", edge.KERNEL_STREAMED),
                       ("generic", edge.KERNEL_GENERIC)):
        calc.set_kernel(mode)
        res[name] = calc.score_band(amb_cap=0)
        assert calc.last_kernel() == mode
    ref = res["generic"]
    for name in ("tiled2", "tiled", "streamed"):
        assert (res[name].words == ref.words).all(), name
        for k in ("pairs_scored", "pairs_walked", "edges", "sum_cc", "sum_tc", "undefined"):
            assert res[name].stats[k] == ref.stats[k], (name, k)
    assert ref.stats["pairs_scored"] > 4e8 and ref.stats["edges"] > 1e7 and ref.stats["undefined"] == 0
    # row windows (what N>1 ranks score) concatenate to the full run
    calc.set_kernel(edge.KERNEL_TILED2)
    cuts = [0, 20_000, 50_001, 77_777, b.n]
    parts = [calc.score_band(r0, r1, amb_cap=0) for r0, r1 in zip(cuts[:-1], cuts[1:])]
    assert (np.concatenate([p.words for p in parts]) == ref.words).all()
    # a sample of band pairs against the oracle, every field
    rng = np.random.default_rng(5)
    i = rng.integers(5000, b.n, 200_000).astype(np.int32)
    j = (i - rng.integers(1, 4000, i.size)).astype(np.int32)
    dev = calc.score_pairs(i, j)
    orc = pyoracle.Oracle(b, **kw).pairs(i, j)
    common.assert_pairs_equal(dev, orc, "C2-sample")
    assert (np.array([ref.edge(int(a), int(c)) for a, c in zip(i[:5000], j[:5000])]) == (orc["edge"][:5000] == 1)).all()


@pytest.mark.parametrize("seed", [1, 2])
def test_long_super_read_like_batch(seed):
    """BASELINE configs[4] shape: a few dozen records of 1-9.7 kb, most named Clique_* (cutoffs 0.99 / 0.97, min overlap
    0.9), long sequential products down to the underflow zone. Too long for a shared-memory tile: AUTO takes the streamed
    kernel for them; every pair against the oracle, every field, from the pair-list kernel and from the band kernel."""
    b = synth.quasispecies(70, genome_len=9719, read_len=3200, frag_mean=5200.0, frag_sd=1800.0, frag_min=900, frag_max=9719,
                           indel_err=2e-4, clip_frac=0.05, seed=900 + seed)
    rng = np.random.default_rng(seed)
    b.is_clique[:] = (rng.random(b.n) < 0.7).astype(b.is_clique.dtype)
    for no_prob0 in (True, False):
        kw = dict(noProb0=no_prob0, maxPosition=b.max_position(), overlap=0.5, overlap_single=0.3,
                  edge_quasi_cutoff=0.97, edge_quasi_cutoff_mixed=0.95, edge_quasi_cutoff_single=0.9)
        calc = edge.NewEdgeCalculator(**kw)
        adj = calc.scoreBatch(b)
        assert calc.last_kernel() == edge.KERNEL_STREAMED
        _tiled_detail_vs_oracle(b, kw, "long-reads-band", edge.KERNEL_STREAMED)
        pi, pj = common.tri_pairs(b.n)
        dev = calc.score_pairs(pi, pj)
        O = pyoracle.Oracle(b, **kw)
        orc = O.pairs(pi, pj)
        flips = common.assert_pairs_equal(dev, orc, "long-reads")
        assert int(orc["cc"].max()) > 3000
        if flips == 0:
            e = orc["edge"] == 1
            ai, aj = adj.edges()
            assert adj.count() == int(e.sum()) and (ai == pi[e]).all() and (aj == pj[e]).all()


@pytest.mark.parametrize("noprob0", [False, True])
def test_super_read_round_C5_shape(noprob0):
    """bench.py's C5 generator at 160 records (reads of 0.5-9.7 kb, 70 % Clique_*, ~75 quality codes): the tables of the
    second-generation kernel (72*K*K bytes) and the column tile of the first do not fit shared memory, AUTO streams the op
    records from global memory instead; every band slot against the oracle, and the same words from the generic walker."""
    b = synth.super_reads(160, seed=4242)
    assert len(np.unique(b.qual)) > 60 and int(b.cig_len.reshape(-1, 2).sum(1).max()) > 6000
    kw = dict(noProb0=noprob0, maxPosition=b.max_position() + 10)
    calc = edge.NewEdgeCalculator(**kw)
    calc.scoreBatch(b)
    assert calc.last_kernel() == edge.KERNEL_STREAMED
    for mode in (edge.KERNEL_TILED, edge.KERNEL_TILED2):
        calc.set_kernel(mode)
        with pytest.raises(edge.HcError):
            calc.score_band()
    det, orc = _tiled_detail_vs_oracle(b, kw, "C5-shape", edge.KERNEL_STREAMED)
    assert int(orc["cc"].max()) > 5000 and int((orc["status"] == 2).sum()) > 5000
    # looser parameters so that edges exist between the haplotypes' reads
    kw2 = dict(kw, edge_quasi_cutoff=0.9, edge_quasi_cutoff_mixed=0.9, edge_quasi_cutoff_single=0.85, overlap=0.5, overlap_single=0.3)
    det, orc = _tiled_detail_vs_oracle(b, kw2, "C5-shape-loose", edge.KERNEL_STREAMED)
    assert int(orc["edge"].sum()) > (50 if noprob0 else 0)      # with prob0 the tail penalty leaves a handful


@pytest.mark.parametrize("kernel", TILED_KERNELS)
@pytest.mark.parametrize("noprob0", [False, True])
def test_band_kernels_with_an_allele_frequency_map(noprob0, kernel):
    """With a SIMPSON_MAP prob0 is a running sum over the tail columns in column order (computeProb0, :133-143): the
    band kernels then walk the tails (MAP variants) instead of using the ladder; every band slot against the oracle,
    and the same words from the generic walker."""
    b = synth.quasispecies(1200, genome_len=1800, seed=33, indel_err=2e-3, clip_frac=0.1)
    rng = np.random.default_rng(3)
    simpson = {int(p): float(np.log10(rng.uniform(0.3, 1.0))) for p in rng.choice(np.arange(1, 1800), 900, replace=False)}
    kw = dict(noProb0=noprob0, simpson_map=simpson, maxPosition=max(b.max_position(), max(simpson)), edge_quasi_cutoff_single=0.9)
    det, orc = _tiled_detail_vs_oracle(b, kw, "band-simpson", kernel)
    assert int(orc["edge"].sum()) > 100 and int((orc["tc"] > 0).sum()) > 1000
    z = synth.zoo(200, seed=77, wide_qual=(20, 42))
    zs = {int(p): float(np.log10(rng.uniform(0.3, 1.0))) for p in rng.choice(np.arange(1, 700), 300, replace=False)}
    kwz = dict(noProb0=noprob0, simpson_map=zs, maxPosition=max(z.max_position(), max(zs)), edge_quasi_cutoff_single=0.8,
               edge_quasi_cutoff_mixed=0.85, edge_quasi_cutoff=0.9, overlap=0.3, overlap_single=0.2)
    _tiled_detail_vs_oracle(z, kwz, "band-simpson-zoo", kernel)


def test_super_read_round_with_an_allele_frequency_map():
    """Super-reads (too long and too many quality codes for a tile) plus a map: AUTO takes the streamed MAP variant."""
    b = synth.super_reads(120, seed=99)
    rng = np.random.default_rng(8)
    simpson = {int(p): float(np.log10(rng.uniform(0.3, 1.0))) for p in rng.choice(np.arange(1, 9700), 5000, replace=False)}
    kw = dict(noProb0=False, simpson_map=simpson, maxPosition=max(b.max_position(), 9700) + 10, edge_quasi_cutoff=0.9,
              edge_quasi_cutoff_mixed=0.9, edge_quasi_cutoff_single=0.85, overlap=0.5, overlap_single=0.3)
    calc = edge.NewEdgeCalculator(**kw)
    calc.scoreBatch(b)
    assert calc.last_kernel() == edge.KERNEL_STREAMED
    det, orc = _tiled_detail_vs_oracle(b, kw, "C5-simpson", edge.KERNEL_STREAMED)
    assert int((orc["tc"] > 1000).sum()) > 500


# ---- round 2: C4-shaped input, near-threshold pairs, shared context ------------------------------------------------
@pytest.mark.parametrize("kernel", TILED_KERNELS)
def test_C4_shaped_amplicon_every_band_slot(kernel):
    """BASELINE configs[3] at a size the oracle walks: 1.5 kb amplicon, 2x150 reads, nearly every pair overlaps (dense
    tiles, loops of ~100 columns). Every band slot of every band kernel against the oracle, with and without prob0."""
    cfg = {k: v for k, v in synth.CONFIGS["C4"].items() if k != "n_frag"}
    b = synth.quasispecies(1500, seed=44, **cfg)
    assert 0.2 < b.single_end.mean() < 0.8                        # merged mates and true pairs both present
    kw = dict(noProb0=True, maxPosition=b.max_position() + 10)
    det, orc = _tiled_detail_vs_oracle(b, kw, "C4-shape", kernel)
    n_pairs = b.n * (b.n - 1) // 2
    assert int((orc["status"] >= 1).sum()) > 0.4 * n_pairs        # dense: half of ALL pairs overlap (C2: one in eleven)
    assert int(orc["edge"].sum()) > 20_000 and set(np.unique(orc["layout"])) >= {100, 200, 300, 400}
    kw["noProb0"] = False
    _tiled_detail_vs_oracle(b, kw, "C4-shape-prob0", kernel)


@pytest.mark.parametrize("kernel", [pytest.param(edge.KERNEL_TILED, id="tiled"), pytest.param(edge.KERNEL_TILED2, id="tiled2")])
def test_grouping_column_reads_by_indel_signature_changes_no_word(kernel, monkeypatch):
    """The column reads of a window are grouped by indel signature before the end sort (build_items_range; HC_ITEM_SIG=0
    turns it off, read at every plan build). Which reads share a block of 32 lanes is scheduling only: the adjacency words and
    every counter of results are the same with and without it (pairs_total, the lanes that looked at a pair at all, is the one
    counter that depends on the blocks), on a C4-shaped set where one read in four carries an indel."""
    cfg = {k: v for k, v in synth.CONFIGS["C4"].items() if k != "n_frag"}
    b = synth.quasispecies(4000, seed=45, **cfg)
    has_indel = np.array([(b"I" in b.cig[int(b.cig_off[2 * r]):int(b.cig_off[2 * r + 1] + b.cig_len[2 * r + 1])].tobytes()) or
                          (b"D" in b.cig[int(b.cig_off[2 * r]):int(b.cig_off[2 * r + 1] + b.cig_len[2 * r + 1])].tobytes())
                          for r in range(b.n)])
    assert 0.2 < has_indel.mean() < 0.98
    calc = edge.NewEdgeCalculator(noProb0=True, maxPosition=b.max_position() + 10)
    calc.upload(b)
    calc.set_kernel(kernel)
    res = {}
    for sig in ("1", "0"):
        monkeypatch.setenv("HC_ITEM_SIG", sig)
        a = calc.score_band()
        assert calc.last_kernel() == kernel
        res[sig] = (a.words.copy(), {k: a.stats[k] for k in ("pairs_scored", "pairs_walked", "edges", "sum_cc", "sum_tc", "undefined")})
    assert (res["1"][0] == res["0"][0]).all() and res["1"][1] == res["0"][1]
    assert res["1"][1]["edges"] > 100_000
    calc.set_kernel(edge.KERNEL_GENERIC)
    assert (calc.score_band().words == res["1"][0]).all()


def test_near_threshold_pairs_are_decided_by_glibc_log10():
    """The device's log10 is not glibc's. Pairs whose score is within ambiguity_rel_tol of the threshold are listed by the
    kernel and re-decided on the host exactly as similarityCriterion does (src/NewEdgeCalculator.cpp:1127-1140). Here the
    cutoff is placed ON the score of known pairs (cutoff = 10**score), so those pairs sit inside the band of doubt; the
    adjacency must be the oracle's (host libm) for every one of them, through every band kernel and the pair list."""
    b = synth.quasispecies(500, genome_len=1200, seed=9)
    base = dict(noProb0=True, maxPosition=b.max_position() + 10)
    calc0 = edge.NewEdgeCalculator(**base)
    calc0.upload(b)
    pi, pj = common.tri_pairs(b.n)
    ref = calc0.score_pairs(pi, pj)
    ok = (ref["status"] == 2) & (ref["gap"] == 0) & np.isfinite(ref["score"]) & (ref["score"] < -1e-4)
    assert ok.sum() > 1000
    scores = np.unique(ref["score"][ok])
    rng = np.random.default_rng(4)
    listed = decided = 0
    for s in rng.choice(scores, 12, replace=False):
        kw = dict(base, edge_quasi_cutoff_single=float(10.0 ** s))
        O = pyoracle.Oracle(b, **kw)
        edges, e = O.all_pairs(detail=False)
        for mode in (edge.KERNEL_TILED2, edge.KERNEL_TILED, edge.KERNEL_STREAMED, edge.KERNEL_GENERIC):
            calc = edge.NewEdgeCalculator(**kw)
            calc.upload(b)
            calc.set_kernel(mode)
            adj = calc.score_band(amb_cap=64)
            assert adj.stats["ambiguous"] >= 1 and adj.ambiguous.shape[0] == adj.stats["ambiguous"]
            ai, aj = adj.edges()
            assert adj.count() == edges == adj.stats["edges"]
            assert (ai == pi[e == 1]).all() and (aj == pj[e == 1]).all()
            for a in adj.ambiguous:                               # the report carries the host's decision
                k = int(a["i"]) * (int(a["i"]) - 1) // 2 + int(a["j"])
                assert a["edge"] == e[k] and abs(a["score"] - a["cutoff"]) <= 1e-12 * abs(a["cutoff"])
                decided += int(a["edge"] != a["edge_device"])
            listed += adj.ambiguous.shape[0]
        dev = calc.score_pairs(pi, pj)                            # pair list: same rule
        assert (dev["edge"] == e).all() and int(dev["ambiguous"].sum()) >= 1
    assert listed >= 48
    print("near-threshold pairs listed: %d, decided differently by the host: %d" % (listed, decided))


def test_wide_tolerance_lists_more_pairs_than_the_default_capacity():
    """A wide tolerance (1e-2) puts thousands of pairs into the report: the device list grows and the pass is repeated;
    every listed pair is still decided by the host rule, so the adjacency is unchanged."""
    b = synth.quasispecies(1200, genome_len=1500, seed=10)
    kw = dict(noProb0=True, maxPosition=b.max_position() + 10, edge_quasi_cutoff_single=0.97)
    tight = edge.NewEdgeCalculator(**kw)
    wide = edge.NewEdgeCalculator(ambiguity_rel_tol=0.5, **kw)
    a0 = tight.scoreBatch(b)
    wide.upload(b)
    a1 = wide.score_band(amb_cap=100)
    assert a1.stats["ambiguous"] > 4096 and a1.ambiguous.shape[0] == 100
    assert (a1.words == a0.words).all() and a1.stats["edges"] == a0.stats["edges"]


def test_pair_list_and_quasi_pair_list_share_one_context():
    """hc_score_pairs and hc_quasi_score_pairs interleaved on ONE context with growing pair counts: every device buffer
    has its own capacity (round-1 advice: the pair-list result buffer used to trust a capacity the quasi call had grown)."""
    import ctypes as C
    b = synth.quasispecies(700, genome_len=1500, seed=12)
    kw = dict(noProb0=True, maxPosition=b.max_position() + 10)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    L, ctx = calc.L, calc._ctx
    qpar = edge._QuasiParams(0.99, 0.97, 0.95, 0.9, 0.6, 0, 0, 1e-12)
    pi, pj = common.tri_pairs(b.n)
    O = pyoracle.Oracle(b, **kw)

    def new_pairs(n):
        out = np.zeros(n, edge.PAIR_RESULT_DTYPE)
        assert L.hc_score_pairs(ctx, edge._ptr(pi[:n]), edge._ptr(pj[:n]), n, edge._ptr(out)) == 0
        common.assert_pairs_equal(out, O.pairs(pi[:n], pj[:n]), "shared-ctx %d" % n)

    def quasi_pairs(n):
        out = np.zeros(n, edge.QUASI_RESULT_DTYPE)
        assert L.hc_quasi_score_pairs(ctx, C.byref(qpar), edge._ptr(pi[:n]), edge._ptr(pj[:n]), n, edge._ptr(out)) == 0
        return out

    new_pairs(100)
    q1 = quasi_pairs(100_000)
    new_pairs(50_000)
    q2 = quasi_pairs(200_000)
    new_pairs(pi.shape[0])
    assert (q2[:100_000]["q"].view(np.uint64) == q1["q"].view(np.uint64)).all()


def test_super_read_rounds_captured_from_the_reference_loop():
    """BASELINE configs[4] on the real thing: the inputs of rounds 1-3 of the reference's own loop (unmodified CLEVER,
    CliqueCollector and merge; tests/golden/make_superread_rounds.py) on 1000 synthetic fragments. Round 1 (3529 Clique_*
    records, 90 quality codes): every band kernel that takes it produces the same words, the edge count is the one the
    reference's loop printed for that round (341035), and a sample of band pairs matches the oracle on every field. Rounds 2
    and 3 (391 / 215 records): every band slot against the oracle."""
    import bench
    b = common.load_batch("superreads_synth_it1.npz")
    assert b.n == 3529 and int(b.is_clique.sum()) == 3515
    kw = dict(bench.PARAMS, maxPosition=2700)
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    res = {}
    for name, mode in (("auto", edge.KERNEL_AUTO), ("streamed", edge.KERNEL_STREAMED), ("generic", edge.KERNEL_GENERIC)):
        calc.set_kernel(mode)
        res[name] = calc.score_band(amb_cap=0)
    assert (res["auto"].words == res["generic"].words).all() and (res["streamed"].words == res["generic"].words).all()
    assert res["generic"].count() == 341035                     # "1: 391  edges: 341035" of the reference CLI on this input
    rng = np.random.default_rng(2)
    i = rng.integers(200, b.n, 60_000).astype(np.int32)
    j = (i - rng.integers(1, 200, i.size)).astype(np.int32)
    common.assert_pairs_equal(calc.score_pairs(i, j), pyoracle.Oracle(b, **kw).pairs(i, j), "captured round 1")
    for it, n in ((2, 391), (3, 215)):
        bb = common.load_batch("superreads_synth_it%d.npz" % it)
        assert bb.n == n
        _tiled_detail_vs_oracle(bb, kw, "captured round %d" % it, edge.KERNEL_STREAMED)
    c2 = edge.NewEdgeCalculator(**kw)
    c2.upload(common.load_batch("superreads_synth_it2.npz"))
    assert c2.score_band(amb_cap=0).count() == 1153               # the reference loop's edge count on round 2's input
