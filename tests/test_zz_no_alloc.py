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
            # driver-level free memory (cudaMemGetInfo): a cudaMalloc inside the library would show here, torch's own
            # allocator statistics would not see it
            free0, mem0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
            nvb.batch_alignment_score_noalloc(a, ps, ts, sinks, ctx=ctx)
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_allocated() == mem0
            got = gu.sinks_np(sinks)
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


def test_noalloc_runs_under_cuda_graph_capture(oracle):
    """the advertised use: captured in a CUDA graph (where cudaMalloc is illegal) and replayed on new inputs in the same buffers"""
    rng = np.random.default_rng(77)
    ctx = nvb.Context.default(0)           # created outside capture: resolves the static array's address for this device
    a, oa = nvb.GotohAligner(nvb.LOCAL, 2, -2, -5, -3), po.aligner(po.GOTOH, po.LOCAL, (2, -2, -5, -3))
    n, M, N = 2000, 60, 120
    batches = [util.rand_pairs(rng, n, (M, M + 1), lambda m: N, 4) for _ in range(2)]
    P, Po, T, To = batches[0]
    ps, ts = gu.sset(P, Po), gu.sset(T, To)
    sinks = torch.zeros((n, 3), dtype=torch.int32, device=gu.dev())
    nvb.batch_alignment_score_noalloc(a, ps, ts, sinks, ctx=ctx)          # warm: module load, lazy kernel load
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        nvb.batch_alignment_score_noalloc(a, ps, ts, sinks, max_pattern_len=M, max_text_len=N, ctx=ctx)
    for P, Po, T, To in batches[::-1]:     # replay on other symbols written into the captured buffers (same offsets: fixed lengths)
        ps.symbols.copy_(torch.from_numpy(P)); ts.symbols.copy_(torch.from_numpy(T))
        sinks.fill_(9)
        g.replay()
        torch.cuda.synchronize()
        assert (gu.sinks_np(sinks) == oracle.score(oa, P, Po, T, To)).all()


def test_noalloc_python_wrapper_validates_arguments():
    dev = gu.dev()
    a = nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1)
    off = (torch.arange(5, device=dev, dtype=torch.int64) * 4).to(torch.int32)
    ps = nvb.StringSet(torch.zeros(16, dtype=torch.uint8, device=dev), off, 8)        # no max_len given
    ts = nvb.StringSet(torch.zeros(16, dtype=torch.uint8, device=dev), off, 8, 4)
    good = torch.zeros((4, 3), dtype=torch.int32, device=dev)
    with pytest.raises(nvb.NvbError):      # would need a device reduction + .item() to learn max_len
        nvb.batch_alignment_score_noalloc(a, ps, ts, good)
    nvb.batch_alignment_score_noalloc(a, ps, ts, good, max_pattern_len=4)
    for bad in (torch.zeros((3, 3), dtype=torch.int32, device=dev), torch.zeros((4, 3), dtype=torch.int64, device=dev),
                torch.zeros((4, 3), dtype=torch.int32), torch.zeros((3, 4), dtype=torch.int32, device=dev).t()):
        with pytest.raises(nvb.NvbError):
            nvb.batch_alignment_score_noalloc(a, ps, ts, bad, max_pattern_len=4)
    # the byte count saturates instead of wrapping (2^31 x 2^31 x 2 x 2 B does not fit 64 bits)
    assert nvb.lib().nvb_batch_score_noalloc_bytes(nvb.GotohAligner(nvb.LOCAL, 2, -2, -5, -3).c(), 1 << 31, 1 << 31) > nvb.NOALLOC_SCRATCH_BYTES
    torch.cuda.synchronize()
