"""banded score / traceback rates over bands, aligner kinds and types (profiling aid: which calls leave the fast paths)"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nvbio_gpl_b200 as nvb
from bench import gen_pairs, time_steps
dev = torch.device("cuda", 0); ctx = nvb.Context.default(0)
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
n, M = 1_000_000, 150
for W in (3, 7, 15, 31):
    N = M + W
    P, Po, T, To = gen_pairs(n, M, N, 3, dev, sub=0.04, implant_frac=1.0)
    ps, ts = nvb.StringSet(P, Po, 8, M), nvb.StringSet(T, To, 8, N)
    for name, mk in (("ed", lambda t: nvb.EditDistanceAligner(t)), ("sw", lambda t: nvb.SmithWatermanAligner(t, 2, -1, -1, -1)), ("gotoh", lambda t: nvb.GotohAligner(t, 0, -5, -8, -3))):
        for tn, t in (("global", nvb.GLOBAL), ("local", nvb.LOCAL), ("semi", nvb.SEMI_GLOBAL)):
            a = mk(t)
            sinks = torch.empty((n, 3), dtype=torch.int32, device=dev)
            fn = lambda: nvb.batch_banded_alignment_score(a, W, ps, ts, sinks=sinks, ctx=ctx)
            tms = time_steps(fn, 2, 1, lambda: flush_buf.fill_(1)); ms = sum(tms) / len(tms); sp = ctx.last_paths
            nt = 200_000
            pt, tt = nvb.StringSet(P[:nt * M], Po[:nt + 1], 8, M), nvb.StringSet(T[:nt * N], To[:nt + 1], 8, N)
            ttemp = torch.empty(int(nvb.lib().nvb_temp_bytes(3, a.c(), W, 16, M, N, nt)), dtype=torch.uint8, device=dev)
            bufs = nvb.batch_banded_alignment_traceback(a, W, 16, pt, tt, temp=ttemp, ctx=ctx)
            fn2 = lambda: nvb.batch_banded_alignment_traceback(a, W, 16, pt, tt, temp=ttemp, ctx=ctx, bufs=bufs)
            t2 = time_steps(fn2, 2, 1, lambda: flush_buf.fill_(1)); ms2 = sum(t2) / len(t2)
            print(json.dumps({"band": W, "aligner": name, "type": tn, "score_gcups": round(n * M * W / ms / 1e6), "score_paths": sp,
                              "tb_Mpairs_s": round(nt / ms2 / 1e3, 1), "tb_paths": ctx.last_paths}))
