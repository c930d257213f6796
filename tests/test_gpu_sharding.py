"""GPU: the row partition the library makes for several GPUs (hc_partition_rows / hc_multi_*) gives the
same adjacency WORDS as one unsharded pass -- what bench.py's ranks do (their own slice, packed and scored on their own
context) and what one process driving several devices does (hc_multi_*; the same ordinal listed several times on a
one-GPU box). The halo comes from the covered intervals, so a read longer than any constant halo cannot lose edges."""
import numpy as np
import pytest

import bench
from haploclique_b200 import edge, synth
from haploclique_b200.records import AlignmentRecord, ReadBatch

pytestmark = pytest.mark.gpu


def _unsharded(b, kw):
    calc = edge.NewEdgeCalculator(**kw)
    calc.upload(b)
    return calc.score_band()


def _as_ranks_do(b, kw, world):
    """Every rank: bench.shard() of the one host read set, its own pack / upload / score of rows >= row_begin."""
    words, stats = [], []
    for rank in range(world):
        local, row_begin, first = bench.shard(b, rank, world)
        calc = edge.NewEdgeCalculator(**kw)
        calc.upload(local)
        a = calc.score_band(row_begin, local.n)
        words.append(a.words)
        stats.append(a.stats)
    return np.concatenate(words), stats


@pytest.mark.parametrize("world", [2, 4])
def test_C3_shaped_position_windows_give_the_unsharded_words(world):
    cfg = {k: v for k, v in synth.CONFIGS["C3"].items() if k != "n_frag"}
    b = synth.quasispecies_sharded(20_000, chunk=7_000, **cfg)
    kw = dict(bench.PARAMS, maxPosition=b.max_position() + 10)
    full = _unsharded(b, kw)
    assert full.stats["edges"] > 100_000
    cuts, first = edge.partition_rows(b, world)
    assert (first % 32 == 0).all() and (first[1:] > 0).all() and (np.diff(cuts) > 0).all()     # real windows with a real halo
    words, stats = _as_ranks_do(b, kw, world)
    assert words.shape == full.words.shape and (words == full.words).all()
    for k in ("pairs_scored", "pairs_walked", "edges", "sum_cc", "sum_tc"):
        assert sum(s[k] for s in stats) == full.stats[k], k
    # balanced by scored pairs, not by reads
    share = np.array([s["pairs_scored"] for s in stats], float)
    assert share.max() / share.mean() < 1.1
    # one process, `world` contexts on this GPU
    multi = edge.NewEdgeCalculator(devices=[0] * world, **kw)
    multi.upload(b)
    m = multi.score_band()
    assert (m.col0 == full.col0).all() and (m.word_off == full.word_off).all()
    assert (m.words == full.words).all()
    for k in ("pairs_scored", "pairs_walked", "edges", "sum_cc", "sum_tc"):       # (pairs_total counts scheduled lanes: it depends on the cut)
        assert m.stats[k] == full.stats[k], k
    assert m.stats["launches"] >= world and len(multi.per_device_stats) == world


def test_C4_shaped_amplicon_parts_give_the_unsharded_words():
    """BASELINE configs[3] shape: half of all pairs overlap, the halo of a part is most of what lies to its left."""
    cfg = {k: v for k, v in synth.CONFIGS["C4"].items() if k != "n_frag"}
    b = synth.quasispecies(3_000, **cfg)
    kw = dict(bench.PARAMS, maxPosition=b.max_position() + 10)
    full = _unsharded(b, kw)
    assert full.stats["pairs_scored"] > 0.4 * b.n * (b.n - 1) / 2
    words, stats = _as_ranks_do(b, kw, 4)
    assert (words == full.words).all()
    assert sum(s["edges"] for s in stats) == full.stats["edges"] > 50_000
    multi = edge.NewEdgeCalculator(devices=[0, 0, 0], **kw)
    multi.upload(b)
    assert (multi.score_band().words == full.words).all()


def test_super_read_round_degenerates_to_row_blocks_of_the_replicated_set():
    """BASELINE configs[4] shape: everything overlaps, so every part's halo reaches read 0 and the cuts become equal-area
    row blocks of equal pair count (what SURVEY.md 8(e) prescribes when position windows degenerate)."""
    b = synth.super_reads(700, seed=5)
    kw = dict(bench.PARAMS, maxPosition=b.max_position() + 10)
    full = _unsharded(b, kw)
    cuts, first = edge.partition_rows(b, 4)
    assert (first == 0).all()
    words, stats = _as_ranks_do(b, kw, 4)
    assert (words == full.words).all()
    share = np.array([215 * s["pairs_scored"] + s["sum_cc"] for s in stats], float)
    # the cut estimates a pair's columns from the covered spans; super-read pairs often stop early at an indel conflict, so the
    # realised work is only roughly even
    assert share.max() / share.mean() < 2.0 and share.min() / share.mean() > 0.5
    multi = edge.NewEdgeCalculator(devices=[0, 0, 0, 0], **kw)
    multi.upload(b)
    assert (multi.score_band().words == full.words).all()


def test_a_read_longer_than_any_constant_halo_keeps_its_edges():
    """A 5 kb super-read in the middle of 2x250 reads: it overlaps rows thousands of positions to its right. The parts'
    halos follow the running maximum of cov_last, so every part that can reach it holds it."""
    cfg = {k: v for k, v in synth.CONFIGS["C3"].items() if k != "n_frag"}
    b = synth.quasispecies_sharded(6_000, chunk=3_000, **cfg)
    recs = b.records()
    rng = np.random.default_rng(3)
    donors = [r for r in recs if r.single_end and 2000 < r.start1 < 2300 and set(r.cigar1_unrolled) == {ord("M")}]
    assert donors
    d = donors[0]
    # stitch a long read from the haplotype-0 reads that follow the donor (plain M, contiguous coverage)
    seq, qual, pos = bytearray(d.sequence1), bytearray(d.qualities1), d.end1
    for r in recs:
        if r.single_end and set(r.cigar1_unrolled) == {ord("M")} and r.start1 <= pos + 1 < r.end1 and pos < 7500:
            k = pos + 1 - r.start1
            seq += r.sequence1[k:]
            qual += r.qualities1[k:]
            pos = r.end1
    assert pos - d.start1 > 4000
    long_read = AlignmentRecord("long_read", True, d.start1, pos, b"M" * len(seq), bytes(seq), bytes(qual))
    recs.append(long_read)
    recs.sort(key=lambda r: r.start1)
    b2 = ReadBatch.from_records(recs)
    at = [k for k, r in enumerate(recs) if r.name == "long_read"][0]
    kw = dict(bench.PARAMS, maxPosition=b2.max_position() + 10, edge_quasi_cutoff_single=0.8, overlap_single=0.05)
    full = _unsharded(b2, kw)
    rows_reaching_it = [i for i in range(at + 1, b2.n) if full.col0[i] <= at]
    assert len(rows_reaching_it) > 2500                           # far more rows than a 1000-column halo would serve
    cuts, first = edge.partition_rows(b2, 4)
    assert sum(1 for p in range(4) if first[p] <= at < cuts[p + 1]) >= 3
    words, _ = _as_ranks_do(b2, kw, 4)
    assert (words == full.words).all()
    multi = edge.NewEdgeCalculator(devices=[0, 0, 0, 0], **kw)
    multi.upload(b2)
    assert (multi.score_band().words == full.words).all()


def test_multi_device_handle_on_every_visible_gpu():
    """With more than one GPU on the box (gpurun --gpus N): one device per part."""
    n = edge.device_count()
    if n < 2:
        pytest.skip("one GPU visible")
    cfg = {k: v for k, v in synth.CONFIGS["C3"].items() if k != "n_frag"}
    b = synth.quasispecies_sharded(30_000, chunk=10_000, **cfg)
    kw = dict(bench.PARAMS, maxPosition=b.max_position() + 10)
    full = _unsharded(b, kw)
    multi = edge.NewEdgeCalculator(devices=list(range(n)), **kw)
    multi.upload(b)
    m = multi.score_band()
    assert (m.words == full.words).all() and m.stats["edges"] == full.stats["edges"]
