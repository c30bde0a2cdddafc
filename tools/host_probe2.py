"""in-process A/B of nvb_fm_match_host_fixed (GPU box): plain loop / k-mer table per chunk / k-mer table per call (the
last needs the NVB_HOST_TABLE switch that was removed after this measurement: profiles/r2_host_probe.txt),
interleaved so that pinned-memory placement and bus state are shared by the variants"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import nvbio_gpl_b200 as nvb
import bench
dev = torch.device("cuda", 0)
genome = bench.gen_genome(100_000_000, 2, dev)
fmi = nvb.build_index(genome)
ctx = nvb.Context.default(0)
nq = 10_000_000
qsym, qoff = bench.gen_queries(genome, nq, 100, dev)
hs = torch.empty(qsym.numel(), dtype=torch.uint8).pin_memory(); hs.copy_(qsym)
hr = torch.empty((nq, 2), dtype=torch.int32).pin_memory()
hp = torch.from_numpy(nvb.pack_symbols(hs.numpy(), 4).view(np.int32)).pin_memory()
hr_n, hp_n = hr.numpy().view(np.uint32), hp.numpy().view(np.uint32)
def run(mode):
    os.environ.pop("NVB_FM_TABLE_K", None)
    if mode.startswith("per_call_k"):
        os.environ["NVB_FM_TABLE_K"] = mode[10:]; mode = "per_call"
    if mode == "plain":
        ctx.fm_batch_sharing(1 << 30)
    else:
        ctx.fm_batch_sharing(1 << 16)
        os.environ["NVB_HOST_TABLE"] = "1" if mode == "per_call" else "0"
    ts = []
    for _ in range(7):
        t0 = time.perf_counter(); nvb.match_host_fixed(fmi, hp_n, 20, nq, ranges_h=hr_n, bits=4, ctx=ctx); ts.append((time.perf_counter() - t0) * 1e3)
    return ts
MODES = ("plain", "per_chunk", "per_call", "per_call_k8", "per_call_k9", "per_call_k11")
for m in MODES: run(m)
for rnd in range(3):
    for m in MODES:
        ts = run(m)
        print("round %d %-12s min %.3f median %.3f max %.3f ms" % (rnd, m, min(ts), sorted(ts)[3], max(ts)))
if os.environ.get("TRACE"):
    os.environ["NVB_HOST_TRACE"] = "1"
    for m in ("plain", "per_chunk", "per_call"):
        os.environ.pop("NVB_HOST_TRACE", None); run(m); os.environ["NVB_HOST_TRACE"] = "1"
        print("== trace", m, flush=True)
        os.environ.pop("NVB_FM_TABLE_K", None)
        if m == "plain": ctx.fm_batch_sharing(1 << 30)
        else:
            ctx.fm_batch_sharing(1 << 16); os.environ["NVB_HOST_TABLE"] = "1" if m == "per_call" else "0"
        nvb.match_host_fixed(fmi, hp_n, 20, nq, ranges_h=hr_n, bits=4, ctx=ctx)
        sys.stderr.flush()
