"""times full-matrix scoring of every aligner kind/type at the configs[0] shape (profiling aid)"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nvbio_gpl_b200 as nvb
from bench import gen_pairs, time_steps

dev = torch.device("cuda", 0)
ctx = nvb.Context.default(0)
n, M, N = int(os.environ.get("PAIRS", 100000)), 150, 500
P, Po, T, To = gen_pairs(n, M, N, 1, dev)
ps, ts = nvb.StringSet(P, Po, 8, M), nvb.StringSet(T, To, 8, N)
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, mk in (("ed", lambda t: nvb.EditDistanceAligner(t)), ("sw", lambda t: nvb.SmithWatermanAligner(t, 2, -1, -1, -1)),
                 ("gotoh", lambda t: nvb.GotohAligner(t, 2, -2, -5, -3))):
    for tn, t in (("global", nvb.GLOBAL), ("local", nvb.LOCAL), ("semi", nvb.SEMI_GLOBAL)):
        a = mk(t)
        sinks = torch.empty((n, 3), dtype=torch.int32, device=dev)
        temp = torch.empty(int(nvb.lib().nvb_temp_bytes(0, a.c(), 0, 0, M, N, n)), dtype=torch.uint8, device=dev)
        fn = lambda: nvb.batch_alignment_score(a, ps, ts, sinks=sinks, temp=temp, ctx=ctx)
        tms = time_steps(fn, 3, 1, lambda: flush_buf.fill_(1))
        ms = sum(tms) / len(tms)
        print(json.dumps({"aligner": name, "type": tn, "gcups": round(n * M * N / ms / 1e6, 1), "ms": round(ms, 3), "paths": ctx.last_paths}))
