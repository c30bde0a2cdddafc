"""CPU: the seed-and-extend composition (oracle/oracle_map.c) over the plain-C restatement equals the same
composition over the reference's own compiled headers, and recovers the sampled positions."""
import numpy as np
import pytest

import pyoracle as po
import map_util


@pytest.mark.parametrize("kind,scores,band", [(po.GOTOH, (0, -5, -8, -3), 15), (po.ED, (0, -1, -1, -1), 15), (po.SW, (1, -2, -3, -3), 7)])
def test_oracle_map_vs_ref(oracle, ref, kind, scores, band):
    rng = np.random.default_rng(11 + band + kind)
    genome = rng.integers(0, 4, 30000).astype(np.uint8)
    genome[5000:5400] = genome[9000:9400]                # a repeat: seeds with several hits
    idx = oracle.build_index(genome)
    sym, off = map_util.make_reads(rng, genome, 300)
    P = map_util.map_params(band=band, kind=kind, scores=scores, min_score=-40)
    h1, o1, c1 = oracle.map_reads(idx, genome, sym, off, P)
    h2, o2, c2 = ref.map_reads(idx, genome, sym, off, P)
    assert (h1 == h2).all() and (o1 == o2).all() and (c1 == c2).all()
    aligned = (h1["flags"] & 2) != 0
    assert aligned.sum() > 200
    assert (h1["end"][aligned] > h1["begin"][aligned]).all()
    assert (h1["n_ops"][~aligned] == 0).all() and (h1["begin"][~aligned] == 0xFFFFFFFF).all()
    assert (c1 > 1).any()


def test_oracle_map_strands_and_limits(oracle):
    rng = np.random.default_rng(5)
    genome = rng.integers(0, 4, 20000).astype(np.uint8)
    idx = oracle.build_index(genome)
    sym, off = map_util.make_reads(rng, genome, 120)
    both = oracle.map_reads(idx, genome, sym, off, map_util.map_params(both=1))[0]
    fwd = oracle.map_reads(idx, genome, sym, off, map_util.map_params(both=0))[0]
    assert ((fwd["flags"] & 1) == 0).all()
    assert ((both["flags"] & 2) != 0).sum() > ((fwd["flags"] & 2) != 0).sum()
    # forward-strand reads give the same result either way
    f = (both["flags"] & 3) == 2
    assert (both[f]["begin"] == fwd[f]["begin"]).all() and (both[f]["score"] == fwd[f]["score"]).all()
    none = oracle.map_reads(idx, genome, sym, off, map_util.map_params(min_score=1))[0]
    assert ((none["flags"] & 2) == 0).all()


def test_pair_hits_rules(oracle):
    """forward-reverse policy and fragment window of the pairing step on hand-made hits"""
    H = np.zeros(12, dtype=po.HIT_DT)
    def hit(i, begin, end, rev, score=-5, aligned=True):
        H[i] = (score, -(1 << 30), begin, end, (2 if aligned else 0) | (1 if rev else 0), 0)
    hit(0, 100, 200, False); hit(1, 300, 400, True)            # FR, fragment 300: concordant
    hit(2, 300, 400, True);  hit(3, 100, 200, False)           # mates swapped: still concordant
    hit(4, 100, 200, False); hit(5, 300, 400, False)           # same strand
    hit(6, 300, 400, False); hit(7, 100, 200, True)            # RF (reverse mate upstream)
    hit(8, 100, 200, False); hit(9, 0, 0, False, aligned=False)   # single
    hit(10, 100, 200, False); hit(11, 900, 1000, True)         # fragment 900 > window
    out = oracle.pair_hits(H, 0, 500)
    assert list(out["flags"]) == [7, 7, 3, 3, 1, 3]
    assert list(out["score"][:4]) == [-10, -10, -10, -10] and out["score"][4] == -5
    assert (out["frag_begin"][0], out["frag_end"][0]) == (100, 400)
    assert list(oracle.pair_hits(H, 350, 500)["flags"]) == [3, 3, 3, 3, 1, 3]
