"""Clique batches for the super-read merge tests: the same seeded member lists for the golden
script, the oracle test and the GPU parity test."""
import numpy as np


def windows(n_reads: int, n_cliques: int, seed: int, span: int = 12, max_members: int = 7):
    """`n_cliques` member lists: 2..max_members reads out of a window of `span` consecutive arrivals; every third list
    in shuffled (non-arrival) fold order."""
    rng = np.random.default_rng(seed)
    out = []
    for t in range(n_cliques):
        i = int(rng.integers(0, n_reads - span))
        m = int(rng.integers(2, max_members + 1))
        mem = np.sort(rng.choice(np.arange(i, i + span), m, replace=False)).astype(np.int32)
        if t % 3 == 0:
            rng.shuffle(mem)
        out.append(mem)
    return out


def clipped_batch(seed: int = 5, n: int = 400, genome: int = 1500):
    """Single-end reads with soft / hard clips, insertions and deletions over a short genome: neighbours overlap heavily,
    so merges meet clip/M, clip/I, D/clip and clip/clip columns."""
    from haploclique_b200.records import AlignmentRecord, ReadBatch
    rng = np.random.default_rng(seed)
    ref = rng.choice(list(b"ACGT"), genome).astype(np.uint8)
    starts = np.sort(rng.integers(40, genome - 400, n))
    recs = []
    for k, st in enumerate(starts):
        core = int(rng.integers(120, 300))
        lead_h, lead_s = (int(rng.integers(0, 4)) if rng.random() < 0.3 else 0), (int(rng.integers(1, 12)) if rng.random() < 0.5 else 0)
        trail_s, trail_h = (int(rng.integers(1, 12)) if rng.random() < 0.5 else 0), (int(rng.integers(0, 4)) if rng.random() < 0.3 else 0)
        cig, seq = bytearray(b"H" * lead_h + b"S" * lead_s), bytearray(rng.choice(list(b"ACGT"), lead_s).astype(np.uint8).tobytes())
        pos, span = int(st), 0
        # indels sit at fixed genome positions (shared by all reads crossing them) plus a few private ones
        for _ in range(core):
            if pos % 97 == 0 and span > 0:
                cig += b"D"
            elif pos % 131 == 0 and span > 0 and cig[-1:] != b"I":
                cig += b"I"; seq.append(int(rng.choice(list(b"ACGT"))))
                cig += b"M"; seq.append(int(ref[pos]) if rng.random() > 0.02 else int(rng.choice(list(b"ACGT"))))
            else:
                cig += b"M"; seq.append(int(ref[pos]) if rng.random() > 0.02 else int(rng.choice(list(b"ACGT"))))
            pos += 1; span += 1
        while cig[-1:] in (b"D", b"I"):       # a read ends on M
            if cig[-1:] == b"I":
                seq.pop()
            else:
                pos -= 1
            cig = cig[:-1]
        end = pos - 1
        cig += b"S" * trail_s + b"H" * trail_h
        seq += rng.choice(list(b"ACGT"), trail_s).astype(np.uint8).tobytes()
        qual = rng.integers(33 + 2, 33 + 42, len(seq)).astype(np.uint8)
        if k % 17 == 0:
            qual[:] = rng.integers(33, 127, len(seq))      # the whole quality range, the 126 cap included
        recs.append(AlignmentRecord("read_%d" % k, True, int(st), int(end), bytes(cig), bytes(seq), qual.tobytes()))
    return ReadBatch.from_records(recs)


def CASES():
    """(label, batch, seed, n_cliques, window span, max members)"""
    from tests.common import load_batch
    yield "hiv50", load_batch("hiv50_records.npz"), 1, 3000, 12, 7
    yield "hiv50_large_cliques", load_batch("hiv50_records.npz"), 2, 300, 40, 30
    yield "clipped", clipped_batch(), 11, 1500, 8, 5


def GENERAL_CASES():
    """(label, batch, seed, n_cliques, window span, max members) for the general merge (paired and mixed members,
    SURVEY.md 8(f) N2): the reference's paired cram input, one of its super-read rounds, synthetic single/paired mixes
    with clips and haplotype indels, and adversarial indel-free layouts that reach every branch of
    mergeAlignmentRecordsPaired / Mixed."""
    from haploclique_b200 import synth
    from tests.common import load_batch
    yield "arabis", load_batch("arabis_records.npz"), 1, 1500, 12, 6
    yield "arabis_wide", load_batch("arabis_records.npz"), 2, 400, 60, 8
    yield "superreads_it1", load_batch("superreads_arabis_it1.npz"), 3, 500, 10, 5
    yield "quasi_mix", synth.quasispecies(700, genome_len=2000, read_len=120, frag_mean=300.0, frag_sd=50.0, frag_min=165, frag_max=540,
                                          seed=73, clip_frac=0.15, indel_err=3e-4), 3, 2500, 14, 6
    yield "layout_zoo", synth.zoo(400, seed=80, genome_len=600, indel=0.0), 0, 3000, 16, 5
    yield "layout_zoo_indels", synth.zoo(400, seed=84, genome_len=600, indel=0.002), 4, 3000, 16, 5
