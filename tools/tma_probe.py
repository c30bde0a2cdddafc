"""fm_match with its query span staged by TMA bulk copy (default) vs per-thread 16-byte loads (NVB_FM_TMA=0): two contexts in
one process, interleaved, CUDA events, L2 flushed between calls"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import nvbio_gpl_b200 as nvb
import bench
dev = torch.device("cuda", 0)
os.environ["NVB_FM_TMA"] = "1"; c_tma = nvb.Context(0)
os.environ["NVB_FM_TMA"] = "0"; c_ldg = nvb.Context(0)
flushbuf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for G in (100_000_000, 1_000_000_000):
    genome = bench.gen_genome(G, 2, dev)
    fmi = nvb.build_index(genome)
    nq = 10_000_000
    qsym, qoff = bench.gen_queries(genome, nq, 100, dev)
    packed = torch.from_numpy(nvb.pack_symbols(qsym.cpu().numpy(), 4).view(np.int32)).to(dev)
    for share in (1 << 16, 1 << 30):
        res = {}
        for rnd in range(3):
            for name, c in (("tma", c_tma), ("ldg", c_ldg)):
                c.fm_batch_sharing(share)
                st = torch.cuda.ExternalStream(c.stream, device=dev)
                ts = []
                for _ in range(6):
                    flushbuf.fill_(1); torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    with torch.cuda.stream(st):
                        e0.record(st); r = nvb.match(fmi, qsym, qoff, ctx=c); e1.record(st)
                    e1.synchronize(); ts.append(e0.elapsed_time(e1))
                res.setdefault(name, []).append(min(ts[1:]))
                res.setdefault(name + "_r", r)
        assert bool((res["tma_r"] == res["ldg_r"]).all().item())
        print("genome %d share_min %d: tma %s ms  ldg %s ms" % (G, share, ["%.4f" % t for t in res["tma"]], ["%.4f" % t for t in res["ldg"]]), flush=True)
    del genome, fmi, qsym, qoff
