"""e2e probe of the host entry points (GPU box): chunk size / ramp sweep via NVB_HOST_CHUNK / NVB_HOST_RAMPS"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import nvbio_gpl_b200 as nvb
import bench
dev = torch.device("cuda", 0)
genome = bench.gen_genome(100_000_000, 2, dev)
fmi = nvb.build_index(genome)
nq = 10_000_000
qsym, qoff = bench.gen_queries(genome, nq, 100, dev)
hs = torch.empty(qsym.numel(), dtype=torch.uint8).pin_memory(); hs.copy_(qsym)
ho = torch.empty(qoff.numel(), dtype=torch.int32).pin_memory(); ho.copy_(qoff)
hr = torch.empty((nq, 2), dtype=torch.int32).pin_memory()
hp = torch.from_numpy(nvb.pack_symbols(hs.numpy(), 4).view(np.int32)).pin_memory()
hs_n, ho_n, hr_n, hp_n = hs.numpy(), ho.numpy().view(np.uint32), hr.numpy().view(np.uint32), hp.numpy().view(np.uint32)
def t(fn):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        t0 = time.time(); fn(); best = min(best, time.time() - t0)
    return best * 1e3
print("chunk %s ramps %s: 8bit %.2f ms  packed4 %.2f ms  fixed4 %.2f ms  fixed8 %.2f ms" % (
    os.environ.get("NVB_HOST_CHUNK"), os.environ.get("NVB_HOST_RAMPS"),
    t(lambda: nvb.match_host(fmi, hs_n, ho_n, hr_n)), t(lambda: nvb.match_host(fmi, hp_n, ho_n, hr_n, bits=4)),
    t(lambda: nvb.match_host_fixed(fmi, hp_n, 20, nq, ranges_h=hr_n, bits=4)), t(lambda: nvb.match_host_fixed(fmi, hs_n, 20, nq, ranges_h=hr_n, bits=8))))
