"""host-side multi-GPU logic on CPU: world_size 2 over gloo. Each rank takes its contiguous slice of a query batch
and of a pair batch, computes the slice's results with the checker (the oracle stands in for the device here: the
partition / rebase / concatenation logic is what is under test), rank 0 gathers and compares with the unsharded run."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from nvbio_gpl_b200 import sharding  # noqa: E402


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 8, 9, 1000, 1001):
        for world in (1, 2, 3, 4, 8):
            got = []
            for r in range(world):
                b, e = sharding.shard_bounds(n, world, r)
                assert 0 <= b <= e <= n
                got += list(range(b, e))
            assert got == list(range(n))


def _worker(rank, world, port, q):
    import pyoracle as po
    import util
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle = po.Oracle("oracle")
        rng = np.random.default_rng(5)                     # same data on every rank
        text = rng.integers(0, 4, 20000).astype(np.uint8)
        idx = oracle.build_index(text)                     # "replicated index"
        Q, qo = util.rand_queries(rng, text, 3001, len_range=(0, 25))
        P, Po, T, To = util.rand_pairs(rng, 1001, (1, 60), lambda M: M + 15, 4)
        oa = po.aligner(po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3))
        # this rank's slices
        qs, qso, (qb, qe) = sharding.shard_string_set(Q, qo, world, rank)
        ps, pso, (pb, pe) = sharding.shard_string_set(P, Po, world, rank)
        ts, tso, _ = sharding.shard_string_set(T, To, world, rank)
        assert len(qso) - 1 == qe - qb and len(pso) - 1 == pe - pb
        ranges = oracle.fm_match(idx, qs, qso)
        sinks = oracle.banded_score(oa, 15, ps, pso, ts, tso)
        all_ranges = sharding.gather_to_rank0(ranges)
        all_sinks = sharding.gather_to_rank0(sinks)
        if rank == 0:
            ok = (all_ranges == oracle.fm_match(idx, Q, qo)).all() and (all_sinks == oracle.banded_score(oa, 15, P, Po, T, To)).all()
            q.put(bool(ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_batches_concatenate_to_the_unsharded_result():
    import pyoracle as po
    if not po.Oracle.available("oracle"):
        pytest.skip("oracle not built")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True
