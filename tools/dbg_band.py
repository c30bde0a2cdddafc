import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import nvbio_gpl_b200 as nvb
import pyoracle as po
import util
import gpu_util as gu

oracle = po.Oracle("oracle")
bad = 0
for kind, scheme in [(po.ED, (0, -1, -1, -1)), (po.SW, (2, -1, -1, -1)), (po.SW, (2, -3, -2, -4)), (po.GOTOH, (0, -5, -8, -3))]:
    for ty in (0, 2):
        a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
        for seed in range(6):
            rng = np.random.default_rng(seed)
            for band in (7, 15):
                P, Po, T, To = util.rand_pairs(rng, 500, (1, 80), util.banded_text_len(rng, ty, band), 4)
                ps, ts = gu.sset(P, Po), gu.sset(T, To)
                got = gu.sinks_np(nvb.batch_banded_alignment_score(a, band, ps, ts))
                want = oracle.banded_score(oa, band, P, Po, T, To)
                idx = np.nonzero(got != want)[0]
                for k in idx[:4]:
                    q = k ^ 1
                    print("MISMATCH kind", kind, scheme, "ty", ty, "band", band, "seed", seed, "pair", k, "M,N", Po[k+1]-Po[k], To[k+1]-To[k],
                          "partner M,N", Po[q+1]-Po[q] if q < 500 else None, To[q+1]-To[q] if q < 500 else None, "got", got[k], "want", want[k],
                          "Poff%16", Po[k] % 16, "Toff%16", To[k] % 16)
                bad += len(idx)
print("total mismatches", bad)
