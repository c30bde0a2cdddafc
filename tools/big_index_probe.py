"""can the synthetic index builder (torch sorts) make a 3.1 Gbp index on this GPU? (profiling aid; prints memory and time)"""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nvbio_gpl_b200 as nvb
from bench import gen_genome, gen_queries
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_100_000_000
dev = torch.device("cuda", 0)
try:
    t = time.time(); genome = gen_genome(n, 2, dev); fmi = nvb.build_index(genome); torch.cuda.synchronize()
    print(json.dumps({"n": n, "build_s": round(time.time() - t, 2), "peak_gb": round(torch.cuda.max_memory_allocated() / 1e9, 1), "primary": fmi.primary}))
    q, qo = gen_queries(genome, 1_000_000, 5, dev)
    r = nvb.match(fmi, q, qo).cpu().numpy().view("uint32")
    found = float((r[:, 0] <= r[:, 1]).mean())
    pos = nvb.locate(fmi, torch.from_numpy(r[:1000, 0].copy().view("int32")).to(dev)).cpu().numpy().view("uint32")
    g = genome.cpu().numpy(); qh = q.cpu().numpy().reshape(-1, 20)
    ok = sum(1 for i in range(1000) if r[i, 0] <= r[i, 1] and (g[pos[i]:pos[i] + 20] == qh[i]).all())
    print(json.dumps({"found_frac": found, "located_ok_of_first_1000_found": ok}))
except Exception as e:
    print(json.dumps({"n": n, "error": str(e)[:300], "peak_gb": round(torch.cuda.max_memory_allocated() / 1e9, 1)}))
