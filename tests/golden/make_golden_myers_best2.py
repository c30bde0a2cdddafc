"""Generates tests/golden/myers_best2.npz from the REFERENCE ITSELF (oracle/_ref/libnvbio_ref.so, the reference's own
headers compiled from /root/reference): MyersTag banded edit distance and Best2Sink scoring on small seeded inputs.
Run in the build container:  python tests/golden/make_golden_myers_best2.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import pyoracle as po  # noqa: E402
import util  # noqa: E402

MIN = -(1 << 30)
MYERS = [(alphabet, band, ty, ms) for alphabet in (2, 4, 5) for band in (7, 16, 32) for ty in (0, 2) for ms in (MIN, -6, -1000)]
BEST2 = [(kind, par, ty, band, dd) for kind, par in ((po.ED, (0, 0, 0, 0)), (po.SW, (2, -1, -1, -1)), (po.GOTOH, (2, -2, -5, -3)))
         for ty in (0, 1, 2) for band in (0, 15) for dd in (0, 5)]


def main():
    ref = po.Oracle("ref")
    out = {}
    for i, (alphabet, band, ty, ms) in enumerate(MYERS):
        rng = np.random.default_rng(5000 + i)
        # texts stay in 0..3: MyersBitVectors<5> leaves vector 4 uninitialised in the reference (myers_banded_inl.h:139-145)
        P, Po, T, To = util.rand_pairs(rng, 150, (0, 80), lambda M: max(0, M + int(rng.integers(-1, band + 6))), alpha=min(alphabet, 4))
        out["m%d_P" % i] = P; out["m%d_Po" % i] = Po; out["m%d_T" % i] = T; out["m%d_To" % i] = To
        out["m%d" % i] = ref.banded_myers(ty, alphabet, band, ms, P, Po, T, To)
    for i, (kind, par, ty, band, dd) in enumerate(BEST2):
        rng = np.random.default_rng(7000 + i)
        n_of_m = util.banded_text_len(rng, ty, band) if band else (lambda M: M + int(rng.integers(0, 50)))
        P, Po, T, To = util.rand_pairs(rng, 120, (1, 60), n_of_m, alpha=2 if i % 3 == 0 else 4)
        out["b%d_P" % i] = P; out["b%d_Po" % i] = Po; out["b%d_T" % i] = T; out["b%d_To" % i] = To
        out["b%d" % i] = ref.score_best2(po.aligner(kind, ty, par), band, dd, P, Po, T, To)
    out["myers_cases"] = np.array(MYERS, dtype=np.int64)
    out["best2_cases"] = np.array([(k, ty, band, dd) + tuple(par) for k, par, ty, band, dd in BEST2], dtype=np.int64)
    path = os.path.join(HERE, "myers_best2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
