"""in-process A/B of nvb_fm_match_host_fixed (GPU box; the NVB_HOST_ZEROCOPY switch it toggles was removed after the measurement in
profiles/r2_host_probe.txt): ranges copied back by the copy engine (shipped) against ranges written by the
kernel straight into the caller's pinned array (NVB_HOST_ZEROCOPY=1), interleaved; 10 M 20-mers, 4-bit packed, 100 Mbp index"""
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
want = nvb.match(fmi, qsym, qoff, ctx=ctx).cpu().numpy().view(np.uint32)
hs = torch.empty(qsym.numel(), dtype=torch.uint8).pin_memory(); hs.copy_(qsym)
hr = torch.empty((nq, 2), dtype=torch.int32).pin_memory()
hp = torch.from_numpy(nvb.pack_symbols(hs.numpy(), 4).view(np.int32)).pin_memory()
hr_n, hp_n = hr.numpy().view(np.uint32), hp.numpy().view(np.uint32)
def run(mode, chunk=None):
    os.environ["NVB_HOST_ZEROCOPY"] = "1" if mode == "zerocopy" else "0"
    ts = []
    for _ in range(9):
        t0 = time.perf_counter(); nvb.match_host_fixed(fmi, hp_n, 20, nq, ranges_h=hr_n, bits=4, ctx=ctx); ts.append((time.perf_counter() - t0) * 1e3)
    return ts
for m in ("copy", "zerocopy"):
    hr_n[:] = 0; run(m); assert (hr_n == want).all(), m
for rnd in range(4):
    for m in ("copy", "zerocopy"):
        ts = run(m)
        print("round %d %-9s min %.3f median %.3f max %.3f ms" % (rnd, m, min(ts), sorted(ts)[4], max(ts)), flush=True)
