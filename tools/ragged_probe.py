"""Gotoh SEMI_GLOBAL scoring and traceback of pairs with MIXED pattern lengths (100..150 x 500): the packed sweeps need both pairs of a
couple to have the same pattern length, so the library couples the pairs in the order of a counting sort by length.
NVB_NO_ORDER=1 couples them as they come (almost every couple then goes to the int32 kernel)."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nvbio_gpl_b200 as nvb

dev = torch.device("cuda", 0)
ctx = nvb.Context.default(0)
n, N = 50000, 500
g = torch.Generator(device="cpu").manual_seed(3)
lens = torch.randint(100, 151, (n,), generator=g)
po = torch.zeros(n + 1, dtype=torch.int64); po[1:] = torch.cumsum(lens, 0)
P = torch.randint(0, 4, (int(po[-1]),), dtype=torch.uint8, generator=g).to(dev)
T = torch.randint(0, 4, (n * N,), dtype=torch.uint8, generator=g).to(dev)
Po = po.to(torch.int32).to(dev); To = (torch.arange(n + 1, dtype=torch.int64) * N).to(torch.int32).to(dev)
pt, tt = nvb.StringSet(P, Po, 8, 150), nvb.StringSet(T, To, 8, N)
a = nvb.GotohAligner(nvb.SEMI_GLOBAL, 2, -2, -5, -3)
for name, fn in (("score", lambda: nvb.batch_alignment_score(a, pt, tt, ctx=ctx)), ("traceback", lambda: nvb.batch_alignment_traceback(a, 64, pt, tt, ctx=ctx))):
    for _ in range(3): r = fn()
    torch.cuda.synchronize(); t = time.time()
    for _ in range(5): r = fn()
    torch.cuda.synchronize(); ms = (time.time() - t) / 5 * 1e3
    chk = int(r.sum()) if name == "score" else int(r[0].sum()) + int(r[2].sum())
    print("%s: %.3f ms per %d pairs (%.1f M pairs/s), checksum %d, NVB_NO_ORDER=%s" % (name, ms, n, n / ms / 1e3, chk, os.environ.get("NVB_NO_ORDER", "0")))
