"""tools/profile_workload.py -- one launch (after one warm-up) of every kernel family at its bench shape, for ncu:
    python tools/profile_workload.py && ncu --set full --clock-control none --import-source on \
        -k regex:'dpf2_|dpf_full|band2_|myers_|dpf_banded|fm_walk|fm_locate|fm_match_kernel|map_seed|map_walk' -c 60 -o gpurun_out/r2_kernels python tools/profile_workload.py
Shapes: configs[0] (150 x 500, reduced to 20 k pairs), configs[2] (250 x 265, band 15, 200 k pairs), band 31 (150 x 181),
FM match / locate on a 400 Mbp index (DRAM-resident, 200 MB) with 2 M queries, nvb_map_reads over 200 k reads of 150 bp."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nvbio_gpl_b200 as nvb  # noqa: E402
import bench  # noqa: E402

dev = torch.device("cuda", 0)
ctx = nvb.Context.default(0)
REP = 2   # warm-up + the launch worth reading


def sets(n, M, N, seed):
    P, Po, T, To = bench.gen_pairs(n, M, N, seed, dev)
    return nvb.StringSet(P, Po, 8, M), nvb.StringSet(T, To, 8, N)


# configs[0] shape: every packed full-matrix kernel
ps, ts = sets(20000, 150, 500, 1)
for a in (nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1), nvb.SmithWatermanAligner(nvb.SEMI_GLOBAL, 2, -1, -1, -1),
          nvb.EditDistanceAligner(nvb.GLOBAL), nvb.GotohAligner(nvb.LOCAL, 2, -2, -5, -3), nvb.GotohAligner(nvb.SEMI_GLOBAL, 2, -2, -5, -3)):
    for _ in range(REP):
        nvb.batch_alignment_score(a, ps, ts, ctx=ctx)
# full-matrix tracebacks (int32 kernels, 64 checkpoints)
ps2, ts2 = sets(5000, 150, 500, 2)
for a in (nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1), nvb.GotohAligner(nvb.SEMI_GLOBAL, 2, -2, -5, -3)):
    for _ in range(REP):
        nvb.batch_alignment_traceback(a, 64, ps2, ts2, ctx=ctx)
del ps, ts, ps2, ts2

# configs[2] shape: banded scoring, Myers, banded tracebacks
pb, tb = sets(200000, 250, 265, 3)
for a in (nvb.EditDistanceAligner(nvb.SEMI_GLOBAL), nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), nvb.SmithWatermanAligner(nvb.LOCAL, 2, -1, -1, -1)):
    for _ in range(REP):
        nvb.batch_banded_alignment_score(a, 15, pb, tb, ctx=ctx)
for _ in range(REP):
    nvb.batch_banded_myers_score(nvb.MyersAligner(nvb.SEMI_GLOBAL, nvb.MyersTag(4)), 16, pb, tb, min_score=-1000, ctx=ctx)
for a in (nvb.EditDistanceAligner(nvb.SEMI_GLOBAL), nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3)):
    for _ in range(REP):
        nvb.batch_banded_alignment_traceback(a, 15, 16, pb, tb, ctx=ctx)
del pb, tb
# band 31 (the widest band nvBowtie picks)
p31, t31 = sets(200000, 150, 181, 4)
for a in (nvb.EditDistanceAligner(nvb.SEMI_GLOBAL), nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3)):
    for _ in range(REP):
        nvb.batch_banded_alignment_score(a, 31, p31, t31, ctx=ctx)
del p31, t31

# FM-index: match (per-batch table), locate, and the read-mapping call on a DRAM-resident index
genome = bench.gen_genome(400_000_000, 2, dev)
fmi = nvb.build_index(genome)
nq = 2_000_000
qsym, qoff = bench.gen_queries(genome, nq, 100, dev)
for _ in range(REP):
    r = nvb.match(fmi, qsym, qoff, ctx=ctx)
rows = r[:, 0].contiguous()
rows = torch.where((r[:, 0].to(torch.int64) & 0xFFFFFFFF) <= (r[:, 1].to(torch.int64) & 0xFFFFFFFF), rows, torch.zeros_like(rows))
for _ in range(REP):
    nvb.locate(fmi, rows, ctx=ctx)
words = nvb.pack_genome(genome, ctx=ctx)
rsym, roff = bench.gen_reads(genome, 200000, 150, 7, dev)
mp = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), seed_len=20, seed_interval=10, max_hits=4, max_range=0, band=15,
                   both_strands=True, min_score=-90)
for _ in range(REP):
    hits, ops, st = nvb.map_reads(fmi, words, rsym, roff, mp, max_read_len=150, ctx=ctx)
torch.cuda.synchronize()
print("profile workload done: launches", ctx.launches, "aligned", st["aligned"])
