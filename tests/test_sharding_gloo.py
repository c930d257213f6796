"""CPU, world_size 2 over gloo: the host-side sharding logic (hc_partition_rows through bench.shard, no GPU, no
scoring). Every rank takes its part of the same seeded host read set; the rows the ranks own partition the read set, and
the pairs that pass the reference's first reject (covered intervals intersect) add up to the unsharded count -- i.e. the
halo the library derives from the covered intervals is wide enough and no pair is scored twice."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_FRAG = 4000


def _band_pairs(batch, row_begin):
    """Scored pairs of rows >= row_begin against all earlier reads of the local batch."""
    total = 0
    for i in range(row_begin, batch.n):
        total += int(((batch.cov_last[:i] >= batch.cov_first[i]) & (batch.cov_last[i] >= batch.cov_first[:i])).sum())
    return total


def _worker(rank, world, port, out, workload="C3", n_frag=N_FRAG):
    sys.path.insert(0, ROOT)
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full, _ = bench.generate(workload, n_frag)
    local, row_begin, first = bench.shard(full, rank, world)
    t = torch.tensor([local.n - row_begin, _band_pairs(local, row_begin)], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    if rank == 0:
        out.put(t.tolist())
    dist.destroy_process_group()


def _run(world, workload, n_frag, port):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out, workload, n_frag)) for r in range(world)]
    for p in procs:
        p.start()
    got = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return got


@pytest.mark.timeout(300)
def test_position_windows_partition_rows_and_pairs():
    sys.path.insert(0, ROOT)
    import bench
    from haploclique_b200 import edge
    full, _ = bench.generate("C3", N_FRAG)
    want = [full.n, _band_pairs(full, 0)]
    cuts, first = edge.partition_rows(full, 2)
    assert cuts[0] == 0 and cuts[2] == full.n and first[0] == 0 and 0 < first[1] < cuts[1] and first[1] % 32 == 0
    # the halo reaches every earlier read that can overlap a row of the part, and not much more
    reach = min(int(np.flatnonzero(full.cov_last[:i] >= full.cov_first[i])[0]) for i in range(int(cuts[1]), min(int(cuts[1]) + 50, full.n)))
    assert first[1] <= reach
    assert got_close(first[1], reach, 64)
    assert _run(2, "C3", N_FRAG, 29600 + os.getpid() % 300) == want


def got_close(a, b, tol):
    return abs(int(a) - int(b)) <= tol


@pytest.mark.timeout(300)
def test_super_read_round_row_blocks_partition_rows_and_pairs():
    """C5 (every read overlaps most others): the halo reaches read 0, so each rank holds the reads up to the end of its
    block of rows; the blocks partition the rows, hold equal shares of the scoring work, and their scored pairs add up."""
    sys.path.insert(0, ROOT)
    import bench
    full, _ = bench.generate("C5", 600)
    assert full.n == 600
    want = [full.n, _band_pairs(full, 0)]
    l0, b0, f0 = bench.shard(full, 0, 2)
    l1, b1, f1 = bench.shard(full, 1, 2)
    assert b0 == 0 and f0 == 0 and f1 == 0 and l0.n == b1 and l1.n == full.n
    # equal shares of the scoring work: 215 per scored pair + the columns the covered spans share (super-reads: the columns count)
    def work(batch, row_begin):
        w = 0
        for i in range(row_begin, batch.n):
            hit = batch.cov_last[:i] >= batch.cov_first[i]
            w += 215 * int(hit.sum()) + int((np.minimum(batch.cov_last[:i][hit], batch.cov_last[i]) - batch.cov_first[i] + 1).clip(0).sum())
        return w
    w0, w1 = work(l0, 0), work(l1, b1)
    assert abs(w0 - w1) < 0.03 * (w0 + w1)
    assert _run(2, "C5", 600, 29300 + os.getpid() % 300) == want


def test_partition_is_balanced_by_band_pairs_and_slices_rebase():
    """hc_partition_rows / hc_packed_slice on the host: cuts at equal shares of the scoring work (scored pairs weighted by shared columns); a slice's
    headers are the parent's with the op offsets re-based and its records are the parent's."""
    import ctypes as C
    sys.path.insert(0, ROOT)
    import bench
    from haploclique_b200 import edge
    full, _ = bench.generate("C3", 6000)
    L = edge.load_library()
    for parts in (1, 3, 8):
        cuts, first = edge.partition_rows(full, parts)
        assert cuts[0] == 0 and cuts[-1] == full.n and (np.diff(cuts) >= 0).all() and (first % 32 == 0).all()
        pmax = np.maximum.accumulate(full.cov_last)
        col0 = np.minimum(np.searchsorted(pmax, full.cov_first, side="left"), np.arange(full.n)) & ~31
        # the weight of a row = the pairs it scores: earlier reads whose covered interval reaches its first covered position
        scored = np.array([int((full.cov_last[col0[i]:i] >= full.cov_first[i]).sum()) for i in range(full.n)])
        share = np.array([scored[cuts[p]:cuts[p + 1]].sum() for p in range(parts)], float)
        assert share.max() / share.mean() < 1.15             # the weight also counts the columns the pairs share
        for p in range(parts):
            assert first[p] == (col0[cuts[p]:cuts[p + 1]].min() if cuts[p + 1] > cuts[p] else cuts[p] & ~31)
    pk = edge.PackedBatch(full)
    cuts, first = edge.partition_rows(full, 3)
    h = pk.headers()
    ops_all = np.frombuffer((C.c_uint16 * pk.n_ops_total).from_address(pk.c.ops), dtype=np.uint16)
    for p in range(3):
        sl = C.c_void_p()
        assert L.hc_packed_slice(pk._h, int(first[p]), int(cuts[p + 1]), C.byref(sl)) == 0
        c = C.cast(sl, C.POINTER(edge._Packed)).contents
        n = int(cuts[p + 1] - first[p])
        assert c.n_reads == n and c.n_qual_codes == pk.c.n_qual_codes
        hs = np.frombuffer((C.c_uint8 * (64 * n)).from_address(c.hdr), dtype=edge.HDR_DTYPE)
        base = int(h["ops_off"][first[p], 0])
        assert (hs["ops_off"] == h["ops_off"][first[p]:cuts[p + 1]] - base).all()
        for f in ("raw", "ext", "n_ops", "ncov_flags", "cov_first", "cov_last", "seg_flags"):
            assert (hs[f] == h[f][first[p]:cuts[p + 1]]).all()
        ops = np.frombuffer((C.c_uint16 * c.n_ops_total).from_address(c.ops), dtype=np.uint16)
        assert (ops == ops_all[base:base + c.n_ops_total]).all()
        L.hc_packed_free(sl)
