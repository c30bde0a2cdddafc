"""times banded / full scoring on 4-bit packed string sets vs 8-bit ones (profiling aid)"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import nvbio_gpl_b200 as nvb
from bench import gen_pairs, time_steps

dev = torch.device("cuda", 0); ctx = nvb.Context.default(0)
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def pack4(sym):      # uint8 device tensor -> 4-bit big-endian words (int32 tensor)
    s = sym.to(torch.int64) & 15
    pad = (-s.numel()) % 8
    if pad: s = torch.cat([s, torch.zeros(pad, dtype=torch.int64, device=dev)])
    sh = (28 - 4 * torch.arange(8, device=dev, dtype=torch.int64))
    return (s.view(-1, 8) << sh).sum(1).to(torch.int32)
n, M, W = 2_000_000, 250, 15; N = M + W
P, Po, T, To = gen_pairs(n, M, N, 3, dev, sub=0.04, implant_frac=1.0)
for bits in (8, 4):
    ps = nvb.StringSet(P if bits == 8 else pack4(P), Po, bits, M); ts = nvb.StringSet(T if bits == 8 else pack4(T), To, bits, N)
    for name, a in (("ed_semi", nvb.EditDistanceAligner(nvb.SEMI_GLOBAL)), ("gotoh_semi", nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3)),
                    ("sw_local", nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1))):
        sinks = torch.empty((n, 3), dtype=torch.int32, device=dev)
        fn = lambda: nvb.batch_banded_alignment_score(a, W, ps, ts, sinks=sinks, ctx=ctx)
        tms = time_steps(fn, 3, 1, lambda: flush_buf.fill_(1)); ms = sum(tms) / len(tms)
        if bits == 8: ref = {**globals().get("ref", {}), name: sinks.clone()}; globals()["ref"] = ref
        same = bool((sinks == globals()["ref"][name]).all().item())
        print(json.dumps({"banded": name, "bits": bits, "gcups": round(n * M * W / ms / 1e6, 1), "ms": round(ms, 3), "paths": ctx.last_paths, "same_as_8bit": same}))
n, M, N = 100000, 150, 500
P, Po, T, To = gen_pairs(n, M, N, 1, dev)
for bits in (8, 4):
    ps = nvb.StringSet(P if bits == 8 else pack4(P), Po, bits, M); ts = nvb.StringSet(T if bits == 8 else pack4(T), To, bits, N)
    a = nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1)
    sinks = torch.empty((n, 3), dtype=torch.int32, device=dev)
    temp = torch.empty(int(nvb.lib().nvb_temp_bytes(0, a.c(), 0, 0, M, N, n)), dtype=torch.uint8, device=dev)
    fn = lambda: nvb.batch_alignment_score(a, ps, ts, sinks=sinks, temp=temp, ctx=ctx)
    tms = time_steps(fn, 3, 1, lambda: flush_buf.fill_(1)); ms = sum(tms) / len(tms)
    print(json.dumps({"full": "sw_local", "bits": bits, "gcups": round(n * M * N / ms / 1e6, 1), "ms": round(ms, 3), "paths": ctx.last_paths}))
