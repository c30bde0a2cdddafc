"""debug aid: map the first K bench reads on a big genome with the library, the C oracle and the compiled reference; print disagreements"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np, torch
import nvbio_gpl_b200 as nvb
import pyoracle as po
from bench import gen_genome, gen_reads
n = int(sys.argv[1]); K = int(sys.argv[2])
dev = torch.device("cuda", 0); ctx = nvb.Context.default(0)
genome = gen_genome(n, 2, dev); fmi = nvb.build_index(genome); words = nvb.pack_genome(genome, ctx=ctx); fmi.build_ktab(10, ctx=ctx)
rsym, roff = gen_reads(genome, 10_000_000, 150, 7, dev)
rsym, roff = rsym[:K * 150].contiguous(), roff[:K + 1].contiguous()
mp = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), seed_len=20, seed_interval=10, max_hits=4, max_range=0, band=15, both_strands=True, min_score=-90)
hits, ops, st = nvb.map_reads(fmi, words, rsym, roff, mp, max_read_len=150, ctx=ctx)
got = hits.cpu().numpy().view(np.uint32).reshape(-1, 6).copy().view(nvb.HIT_DT).reshape(-1)
bwt_occ, ssa = fmi.host_arrays()
oidx = po.Index(fmi.length, fmi.primary, np.array(fmi.L2, dtype=np.uint32), bwt_occ, ssa)
gh = genome.cpu().numpy(); R = rsym.cpu().numpy(); Ro = roff.cpu().numpy().view(np.uint32)
pc = po.MapParamsC(20, 10, 4, 0, 15, 1, -90, po.aligner(po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3)))
res = {}
for which in ("oracle", "ref"):
    o = po.Oracle(which)
    res[which], _, ncand = o.map_reads(oidx, gh, R, Ro, pc, ops_stride=nvb.ops_stride_for(150, 15))
    bad = np.nonzero((got.view(np.uint32).reshape(K, 6) != res[which].view(np.uint32).reshape(K, 6)).any(axis=1))[0]
    print(which, "mismatches", len(bad))
    for i in bad[:5]:
        print(" read", i, "gpu", got[i], which, res[which][i], "cands", ncand[i])
bad = np.nonzero((res["oracle"].view(np.uint32).reshape(K, 6) != res["ref"].view(np.uint32).reshape(K, 6)).any(axis=1))[0]
print("oracle vs ref mismatches", len(bad))
i = int(bad[0]) if len(bad) else None
which = "oracle"
bad = np.nonzero((got.view(np.uint32).reshape(K, 6) != res[which].view(np.uint32).reshape(K, 6)).any(axis=1))[0]
for i in bad[:2]:
    i = int(i)
    for lo, hi in ((i, i + 1), (max(0, i - 3), i + 4)):
        rs = rsym[lo * 150:hi * 150].contiguous(); ro = (roff[lo:hi + 1] - roff[lo]).contiguous()
        if hi == lo + 1: os.environ["NVB_MAP_DEBUG"] = "1"
        h2, _, st2 = nvb.map_reads(fmi, words, rs, ro, mp, max_read_len=150, ctx=ctx)
        os.environ.pop("NVB_MAP_DEBUG", None)
        g2 = h2.cpu().numpy().view(np.uint32).reshape(-1, 6).copy().view(nvb.HIT_DT).reshape(-1)
        print("alone", lo, hi, g2[i - lo], st2)
        if hi == lo + 1:
            for fg in (1, 2):
                ctx.force_generic(fg)
                h3, _, st3 = nvb.map_reads(fmi, words, rs, ro, mp, max_read_len=150, ctx=ctx)
                print("  force_generic", fg, h3.cpu().numpy().view(np.uint32).reshape(-1, 6).copy().view(nvb.HIT_DT).reshape(-1)[0], st3)
            ctx.force_generic(0)
            fmi.drop_ktab()
            h3, _, st3 = nvb.map_reads(fmi, words, rs, ro, mp, max_read_len=150, ctx=ctx)
            print("  no ktab", h3.cpu().numpy().view(np.uint32).reshape(-1, 6).copy().view(nvb.HIT_DT).reshape(-1)[0], st3)
            fmi.build_ktab(10, ctx=ctx)
    # every seed of the read on both strands: ranges + located rows through the plain entry points, with and without the table
    read = R[i * 150:(i + 1) * 150]
    rc = (3 - read)[::-1]
    seeds = np.concatenate([s[o:o + 20] for s in (read, rc) for o in range(0, 131, 10)]).astype(np.uint8)
    so = (np.arange(29) * 20).astype(np.uint32)
    want = po.Oracle("oracle").fm_match(oidx, seeds, so)
    sd, sod = torch.from_numpy(seeds).to(dev), torch.from_numpy(so.view(np.int32)).to(dev)
    g_k = nvb.match(fmi, sd, sod).cpu().numpy().view(np.uint32)
    fmi.drop_ktab()
    g_p = nvb.match(fmi, sd, sod).cpu().numpy().view(np.uint32)
    for k in range(28):
        if not ((want[k] == g_k[k]).all() and (want[k] == g_p[k]).all()) or want[k][0] <= want[k][1]:
            print(" seed", k, "oracle", want[k], "gpu ktab", g_k[k], "gpu plain", g_p[k])
    fmi.build_ktab(10, ctx=ctx)
    rows = np.array([want[k][0] for k in range(28) if want[k][0] <= want[k][1]], dtype=np.uint32)
    offs = np.array([(k % 14) * 10 for k in range(28) if want[k][0] <= want[k][1]], dtype=np.int64)
    lp = nvb.locate(fmi, torch.from_numpy(rows.view(np.int32)).to(dev)).cpu().numpy().view(np.uint32)
    lo_ = po.Oracle("oracle").fm_locate(oidx, rows)
    print(" rows", rows); print(" gpu locate", lp); print(" orc locate", lo_); print(" diag gpu", lp.astype(np.int64) - offs); print(" diag orc", lo_.astype(np.int64) - offs)
