"""tools/sanitizer_workload.py -- a small pass through every kernel family of libnvbio_b200.so on reduced shapes, meant to be run
under compute-sanitizer (one tool per run):
    compute-sanitizer --tool memcheck  python tools/sanitizer_workload.py
    compute-sanitizer --tool racecheck python tools/sanitizer_workload.py
String sets are passed as views that start and end at odd addresses inside larger allocations, so that any read outside a
caller's array lands in memory the call does not own; results are still compared with the oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import nvbio_gpl_b200 as nvb  # noqa: E402
import pyoracle as po  # noqa: E402
import util  # noqa: E402

dev = torch.device("cuda", 0)
oracle = po.Oracle("oracle")
rng = np.random.default_rng(5)
ctx = nvb.Context.default(0)


def odd_view(arr, lead=3):
    """device copy of arr placed `lead` bytes into its own exactly sized allocation"""
    a = np.ascontiguousarray(arr, dtype=np.uint8)
    buf = torch.empty(len(a) + lead, dtype=torch.uint8, device=dev)
    v = buf[lead:lead + len(a)]
    v.copy_(torch.from_numpy(a))
    return v


def sset(sym, off):
    o = torch.from_numpy(np.ascontiguousarray(off, dtype=np.uint32).view(np.int32)).to(dev)
    ml = int(np.max(np.diff(off.astype(np.int64)))) if len(off) > 1 else 0
    return nvb.StringSet(odd_view(sym) if len(sym) else torch.zeros(16, dtype=torch.uint8, device=dev), o, 8, ml)


def sinks_np(t):
    a = t.cpu().numpy()
    out = np.zeros(a.shape[0], dtype=po.SINK_DT)
    out["score"], out["x"], out["y"] = a[:, 0], a[:, 1].view(np.uint32), a[:, 2].view(np.uint32)
    return out


# ---- FM-index: match (plain loop and per-batch table), match_reverse, rank, rank4, locate, filter
text = rng.integers(0, 4, 30000).astype(np.uint8)
fmi = nvb.build_index(torch.from_numpy(text).to(dev))
bwt_occ, ssa = fmi.host_arrays()
oidx = po.Index(fmi.length, fmi.primary, np.array(fmi.L2, dtype=np.uint32), bwt_occ, ssa)
Q, qo = util.rand_queries(rng, text, 3000, len_range=(0, 40))
qs = sset(Q, qo)
for share in (0xFFFFFFFF, 1):
    ctx.fm_batch_sharing(share)
    assert (nvb.match(fmi, qs.symbols, qs.offsets).cpu().numpy().view(np.uint32) == oracle.fm_match(oidx, Q, qo)).all()
    assert (nvb.match_reverse(fmi, qs.symbols, qs.offsets).cpu().numpy().view(np.uint32) == oracle.fm_match_reverse(oidx, Q, qo)).all()
ctx.fm_batch_sharing(1 << 16)
want = oracle.fm_match(oidx, Q, qo)
rows = want[want[:, 0] <= want[:, 1], 0].copy()
rd = torch.from_numpy(rows.view(np.int32)).to(dev)
assert (nvb.locate(fmi, rd).cpu().numpy().view(np.uint32) == oracle.fm_locate(oidx, rows)).all()
assert (nvb.locate(fmi, rd, two_phase=True).cpu().numpy().view(np.uint32) == oracle.fm_locate(oidx, rows)).all()
k = rng.integers(0, len(text) + 1, 500).astype(np.uint32)
assert (nvb.rank4(fmi, torch.from_numpy(k.view(np.int32)).to(dev)).cpu().numpy().view(np.uint32) == oracle.fm_rank4(oidx, k)).all()
flt = nvb.FMIndexFilter()
nh = flt.rank(fmi, qs.symbols, qs.offsets)
flt.locate(0, min(nh, 5000))
r_, c_ = nvb.match_approx(fmi, qs.symbols, qs.offsets, 10, True, cap=16)
print("fm ok")

# ---- alignment: every aligner x type, full and banded, scoring and traceback, all three kernel paths
for force in (0, 2, 1):
    ctx.force_generic(force)
    for kind in (po.ED, po.SW, po.GOTOH):
        scheme = util.SCHEMES[kind][-1]
        for ty in (0, 1, 2):
            a, oa = nvb.Aligner(kind, ty, *scheme), po.aligner(kind, ty, scheme)
            P, Po, T, To = util.rand_pairs(rng, 129, (1, 70), lambda M: int(rng.integers(0, 120)), 4)
            ps, ts = sset(P, Po), sset(T, To)
            assert (sinks_np(nvb.batch_alignment_score(a, ps, ts)) == oracle.score(oa, P, Po, T, To)).all(), (force, kind, ty)
            nvb.batch_alignment_traceback(a, 16, ps, ts)
            for band in (7, 15, 31):
                P, Po, T, To = util.rand_pairs(rng, 129, (1, 80), util.banded_text_len(rng, ty, band), 4)
                ps, ts = sset(P, Po), sset(T, To)
                assert (sinks_np(nvb.batch_banded_alignment_score(a, band, ps, ts)) == oracle.banded_score(oa, band, P, Po, T, To)).all(), (force, kind, ty, band)
                nvb.batch_banded_alignment_traceback(a, band, 16, ps, ts)
ctx.force_generic(0)
# wide shapes: the wavefront kernels over many lanes
for kind, ty, scheme in ((po.SW, 1, (2, -1, -1, -1)), (po.SW, 2, (2, -1, -1, -1)), (po.GOTOH, 1, (2, -2, -5, -3)), (po.GOTOH, 0, (2, -2, -5, -3))):
    a, oa = nvb.Aligner(kind, ty, *scheme), po.aligner(kind, ty, scheme)
    P, Po, T, To = util.rand_pairs(rng, 65, (140, 151), lambda M: 300, 4)
    ps, ts = sset(P, Po), sset(T, To)
    assert (sinks_np(nvb.batch_alignment_score(a, ps, ts)) == oracle.score(oa, P, Po, T, To)).all(), (kind, ty)
ma = nvb.MyersAligner(nvb.SEMI_GLOBAL, nvb.MyersTag(4))
P, Po, T, To = util.rand_pairs(rng, 200, (20, 60), lambda M: M + 16, 4)
nvb.batch_banded_myers_score(ma, 16, nvb.StringSet.from_numpy(P, Po, dev), nvb.StringSet.from_numpy(T, To, dev), min_score=-1000)   # needs 4-byte aligned symbol arrays
print("dp ok")

# ---- seed -> locate -> extend -> traceback
genome = torch.from_numpy(text).to(dev)
words = nvb.pack_genome(genome, ctx=ctx)
pos = rng.integers(0, len(text) - 160, 700)
reads = np.stack([text[p:p + 100] for p in pos]).reshape(-1).copy()
roff = (np.arange(701) * 100).astype(np.uint32)
mp = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), seed_len=20, seed_interval=10, max_hits=4, max_range=0, band=15,
                   both_strands=True, min_score=-60)
rs = sset(reads, roff)
hits, ops, st = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, mp, max_read_len=100, ctx=ctx)
torch.cuda.synchronize()
print("map ok", st)
