"""nvb_fm_match_host_fixed2n (2-bit seeds + N list) end to end, 10 M 20-mers, 100 Mbp index: wall clock per call for the chunk size in
NVB_HOST_CHUNK (one process per setting), optional NVB_HOST_TRACE=1 stage timeline of one call"""
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
hs = qsym.cpu().numpy()
w2, nl = nvb.pack_symbols_2n(hs, 20)
h2 = torch.from_numpy(w2.view(np.int32)).pin_memory(); h2_n = h2.numpy().view(np.uint32)
hr = torch.empty((nq, 2), dtype=torch.int32).pin_memory(); hr_n = hr.numpy().view(np.uint32)
trace = os.environ.pop("NVB_HOST_TRACE", None)
for share in (1 << 16, 1 << 30):
    ctx.fm_batch_sharing(share)
    ts = []
    for _ in range(12):
        t0 = time.perf_counter(); nvb.match_host_fixed2n(fmi, h2_n, nl, 20, nq, ranges_h=hr_n, ctx=ctx); ts.append((time.perf_counter() - t0) * 1e3)
    print("chunk %s share_min %d: min %.3f median %.3f ms" % (os.environ.get("NVB_HOST_CHUNK"), share, min(ts[2:]), sorted(ts[2:])[5]), flush=True)
if trace:
    ctx.fm_batch_sharing(1 << 16)
    os.environ["NVB_HOST_TRACE"] = "1"
    nvb.match_host_fixed2n(fmi, h2_n, nl, 20, nq, ranges_h=hr_n, ctx=ctx)
