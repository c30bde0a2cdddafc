"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libnvbio_ref.so = the reference's own headers
compiled from /root/reference; see oracle/Makefile). Run in the build container:  python tests/golden/make_golden.py
The fixtures are small seeded inputs plus the reference's outputs; they travel with the repo so that the CPU oracle
tests and the GPU parity tests can check against reference outputs where /root/reference does not exist."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import pyoracle as po  # noqa: E402
import util  # noqa: E402


def main():
    ref = po.Oracle("ref")
    rng = np.random.default_rng(20260101)
    # ---- seeding: index image + queries -> SA ranges, located positions
    text = rng.integers(0, 4, 30011).astype(np.uint8)
    idx = ref.build_index(text)
    Q, qo = util.rand_queries(rng, text, 4000, len_range=(0, 30))
    ranges = ref.fm_match(idx, Q, qo)
    rows = ranges[ranges[:, 0] <= ranges[:, 1], 0][:1500].copy()
    pos = ref.fm_locate(idx, rows)
    np.savez_compressed(os.path.join(HERE, "fm.npz"), text=text, length=idx.length, primary=idx.primary, L2=idx.L2,
                        bwt_occ=idx.bwt_occ, ssa=idx.ssa, Q=Q, qo=qo, ranges=ranges, rows=rows, pos=pos)
    # ---- extension: full and banded, every aligner and type
    out = {}
    cases = []
    for kind in (po.ED, po.SW, po.GOTOH):
        for scheme in util.SCHEMES[kind][:2]:
            for ty in (0, 1, 2):
                cases.append((kind, scheme, ty))
    P, Po, T, To = util.rand_pairs(rng, 300, (1, 90), lambda M: int(rng.integers(0, 140)), 4)
    P2, Po2, T2, To2 = util.rand_pairs(rng, 300, (1, 60), lambda M: int(rng.integers(0, 90)), 2)
    out.update(full_P=P, full_Po=Po, full_T=T, full_To=To, full2_P=P2, full2_Po=Po2, full2_T=T2, full2_To=To2)
    for ci, (kind, scheme, ty) in enumerate(cases):
        a = po.aligner(kind, ty, scheme)
        out["full_%d" % ci] = ref.score(a, P, Po, T, To)
        out["full2_%d" % ci] = ref.score(a, P2, Po2, T2, To2)
        o, ops, n = ref.traceback(a, 16, P, Po, T, To)
        out["fulltb_%d_out" % ci] = o; out["fulltb_%d_ops" % ci] = ops; out["fulltb_%d_n" % ci] = n
        for band in (7, 15):
            r2 = np.random.default_rng(1000 * ci + band)
            BP, BPo, BT, BTo = util.rand_pairs(r2, 300, (1, 300), util.banded_text_len(r2, ty, band), 4 if ci % 2 == 0 else 2)
            key = "band%d_%d" % (band, ci)
            out[key + "_P"] = BP; out[key + "_Po"] = BPo; out[key + "_T"] = BT; out[key + "_To"] = BTo
            out[key] = ref.banded_score(a, band, BP, BPo, BT, BTo)
            o, ops, n = ref.banded_traceback(a, band, 16, BP, BPo, BT, BTo)
            out[key + "_tb_out"] = o; out[key + "_tb_ops"] = ops; out[key + "_tb_n"] = n
    out["cases"] = np.array([(k, t) + tuple(s) for k, s, t in cases], dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "aln.npz"), **out)
    print("wrote fm.npz, aln.npz:", sum(os.path.getsize(os.path.join(HERE, f)) for f in ("fm.npz", "aln.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
