"""timing probe of nvb_map_reads: batch size sweep (GPU box only)"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nvbio_gpl_b200 as nvb
import bench

dev = torch.device("cuda", 0)
G = int(os.environ.get("G", 100_000_000)); nr = int(os.environ.get("NR", 4_000_000))
genome = bench.gen_genome(G, 2, dev)
fmi = nvb.build_index(genome)
words = nvb.pack_genome(genome)
rsym, roff = bench.gen_reads(genome, nr, 150, 7, dev)
p = nvb.MapParams(min_score=-90)
ctx = nvb.Context.default(0)
for rpb in (1 << 19, 1 << 20, 1 << 21, 1 << 22):
    for want_ops in (True, False):
        for _ in range(2):
            nvb.map_reads(fmi, words, rsym, roff, p, max_read_len=150, reads_per_batch=rpb, want_ops=want_ops)
        torch.cuda.synchronize()
        t = time.time()
        for _ in range(3):
            h, o, st = nvb.map_reads(fmi, words, rsym, roff, p, max_read_len=150, reads_per_batch=rpb, want_ops=want_ops)
        torch.cuda.synchronize()
        dt = (time.time() - t) / 3
        print("rpb %8d ops %d: %.2f ms  %.1f M reads/s" % (rpb, want_ops, dt * 1e3, nr / dt / 1e6), st, flush=True)
