"""CPU, world_size 2 over gloo: the host-side sharding logic of bench.py (genome-position windows + left halo).
Every rank builds its window of the same seeded read set; the rows it owns partition the read set and the pairs that
pass the reference's first reject (covered intervals intersect) add up to the unsharded count -- i.e. the halo is
wide enough and no pair is scored twice. No GPU, no scoring: only the partitioning."""
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
    local, row_begin, cfg = bench.make_workload(workload, n_frag, rank, world)
    t = torch.tensor([local.n - row_begin, _band_pairs(local, row_begin)], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    if rank == 0:
        out.put(t.tolist())
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_position_windows_partition_rows_and_pairs():
    sys.path.insert(0, ROOT)
    import bench
    full, rb, _ = bench.make_workload("C3", N_FRAG, 0, 1)
    assert rb == 0
    want = [full.n, _band_pairs(full, 0)]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    got = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == want


@pytest.mark.timeout(300)
def test_super_read_round_row_blocks_partition_rows_and_pairs():
    """C5 (every read overlaps most others): no position windows -- each rank holds the reads up to the end of its block
    of rows; the blocks partition the rows, have about equal pair area, and their scored pairs add up."""
    sys.path.insert(0, ROOT)
    import bench
    full, rb, _ = bench.make_workload("C5", 600, 0, 1)
    assert rb == 0 and full.n == 600
    want = [full.n, _band_pairs(full, 0)]
    l0, b0, _ = bench.make_workload("C5", 600, 0, 2)
    l1, b1, _ = bench.make_workload("C5", 600, 1, 2)
    assert b0 == 0 and l0.n == b1 and l1.n == full.n
    assert abs(l0.n * (l0.n - 1) / 2 - full.n * (full.n - 1) / 4) < 0.02 * full.n * full.n      # equal pair area
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29300 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out, "C5", 600)) for r in range(2)]
    for p in procs:
        p.start()
    got = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == want
