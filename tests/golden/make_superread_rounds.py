"""Captures super-read rounds from the UNMODIFIED reference's own loop (oracle/_ref/libhcref.so: hcref_iterate = one pass of
src/haploclique.cpp:349-372 with the reference's NewEdgeCalculator, CLEVER and CliqueCollector) on a synthetic quasispecies set
small enough for the reference's host-side clique maintenance ("run the reference loop on a C2-style input
small enough for CLEVER, capture each round's super-read deque and score those"). Writes tests/golden/superreads_synth_it{1,2,3}.npz:
the inputs of rounds 1..3 (round r input = collector.finish() output of round r-1), names included (Clique_* / read names).
Run where /root/reference exists:  python tests/golden/make_superread_rounds.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from haploclique_b200 import synth          # noqa: E402
from oracle import pyref                    # noqa: E402

N_FRAG, GENOME, SEED = 1000, 2500, 11       # the set profiles/tools/cli_wall.py times (reference CLI: 35 s)
PARAMS = dict(noProb0=True)                 # defaults + --no_prob0, --no_singletons --significance=4 as the cram test


def main():
    q = synth.quasispecies(N_FRAG, genome_len=GENOME, seed=SEED)
    R = pyref.Reference()
    R.add_batch(q)
    R.set_params(maxPosition=GENOME + 200, **PARAMS)
    ct, sd = 0, 1.0
    for it in range(1, 4):
        n, ct, edges, conv, sd = R.iterate(ct, sd, use_bk=False, filter_singletons=True, significance=4.0)
        b = R.to_batch()
        path = os.path.join(ROOT, "tests", "golden", "superreads_synth_it%d.npz" % it)
        b.save(path)
        lens = b.cig_len.reshape(-1, 2).sum(1)
        print("round %d input: %d records (%d Clique_*), %d edges drawn in the round before, ops per record %d..%d, %d quality codes, %d bytes" % (
            it, b.n, int(b.is_clique.sum()), edges, lens.min(), lens.max(), len(set(b.qual.tolist())), os.path.getsize(path)))
        if conv:
            break


if __name__ == "__main__":
    main()
