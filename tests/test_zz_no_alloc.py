"""-m gpu, collected last: nvb_batch_score_noalloc (full-matrix scoring with its scratch in a static device array of the
library; no counterpart in the reference) must return exactly what
nvb_batch_score returns -- integer results, so the tolerance is bit equality as everywhere else -- on two calls in a row that
share the static array, and must reject problems beyond its built-in limit."""
import ctypes as C

import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
import pyoracle as po
import util
import gpu_util as gu

pytestmark = pytest.mark.gpu

ALIGNERS = [(po.SW, po.LOCAL, (2, -1, -1, -1)), (po.GOTOH, po.SEMI_GLOBAL, (2, -2, -5, -3)), (po.ED, po.GLOBAL, (0, 0, 0, 0)),
            (po.GOTOH, po.LOCAL, (1, -1, -2, -1))]


def test_noalloc_score_equals_batch_score_twice_in_a_row(oracle):
    rng = np.random.default_rng(2025)
    ctx = nvb.Context.default(0)
    # two different batches back to back: the second call finds the first call's columns in the shared array
    batches = [util.rand_pairs(rng, 3000, (1, 150), lambda M: int(rng.integers(0, 400)), 4),
               util.rand_pairs(rng, 1777, (1, 90), lambda M: int(rng.integers(0, 250)), 2)]
    for kind, ty, scheme in ALIGNERS:
        a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
        for P, Po, T, To in batches:
            ps, ts = gu.sset(P, Po), gu.sset(T, To)
            want = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts, ctx=ctx))
            sinks = torch.full((ps.count, 3), 7, dtype=torch.int32, device=gu.dev())
            torch.cuda.synchronize()
            mem0 = torch.cuda.memory_allocated()
            got = gu.sinks_np(nvb.batch_alignment_score_noalloc(a, ps, ts, sinks, ctx=ctx))
            assert torch.cuda.memory_allocated() == mem0
            assert ctx.last_paths == 2
            assert (got == want).all(), (kind, ty, np.nonzero(got != want)[0][:5])
            assert (got == oracle.score(oa, P, Po, T, To)).all()


def test_noalloc_limit_is_enforced():
    L = nvb.lib()
    a = nvb.GotohAligner(nvb.LOCAL, 2, -2, -5, -3)
    assert L.nvb_batch_score_noalloc_bytes(a.c(), 500, 1000) == 2 * 500 * 1000 * 2
    n, N = 40000, 500                      # 2 * 500 * 40000 * 2 B = 80 MB > 32 MiB
    assert L.nvb_batch_score_noalloc_bytes(a.c(), N, n) > nvb.NOALLOC_SCRATCH_BYTES
    dev = gu.dev()
    off_p = (torch.arange(n + 1, device=dev, dtype=torch.int64) * 10).to(torch.int32)
    off_t = (torch.arange(n + 1, device=dev, dtype=torch.int64) * N).to(torch.int32)
    ps = nvb.StringSet(torch.zeros(n * 10, dtype=torch.uint8, device=dev), off_p, 8, 10)
    ts = nvb.StringSet(torch.zeros(n * N, dtype=torch.uint8, device=dev), off_t, 8, N)
    sinks = torch.zeros((n, 3), dtype=torch.int32, device=dev)
    ctx = nvb.Context.default(0)
    rc = L.nvb_batch_score_noalloc(ctx.h, a.c(), ps.c(), ts.c(), sinks.data_ptr(), 10, N, None)
    assert rc == -5 and b"limit" in L.nvb_status_string(rc)
    with pytest.raises(nvb.NvbError):
        nvb.batch_alignment_score_noalloc(a, ps, ts, sinks, ctx=ctx)
    # the same shape split into batches that fit goes through
    k = 8000
    ps2 = nvb.StringSet(ps.symbols[:k * 10], off_p[:k + 1].contiguous(), 8, 10)
    ts2 = nvb.StringSet(ts.symbols[:k * N], off_t[:k + 1].contiguous(), 8, N)
    nvb.batch_alignment_score_noalloc(a, ps2, ts2, sinks[:k], ctx=ctx)
    torch.cuda.synchronize()
