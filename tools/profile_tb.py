"""One full-matrix traceback call per aligner at configs[0]'s shape (150 x 500, 20 k pairs): the workload ncu captures the packed
traceback kernels on (dpf2_*_tb_forward_kernel, dpf2_*_tb_walk_kernel)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import nvbio_gpl_b200 as nvb

dev = torch.device("cuda", 0)
ctx = nvb.Context.default(0)
M, N, nt = 150, 500, 20000
P, Po, T, To = bench.gen_pairs(nt, M, N, 5, dev)
pt, tt = nvb.StringSet(P, Po, 8, M), nvb.StringSet(T, To, 8, N)
for a in (nvb.SmithWatermanAligner(nvb.SEMI_GLOBAL, 2, -1, -1, -1), nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1),
          nvb.GotohAligner(nvb.SEMI_GLOBAL, 2, -2, -5, -3), nvb.GotohAligner(nvb.LOCAL, 2, -2, -5, -3)):
    out, ops, n_ops = nvb.batch_alignment_traceback(a, 64, pt, tt, ctx=ctx)
    torch.cuda.synchronize()
    print(a.kind, a.type, int(n_ops.sum()), ctx.last_paths)
