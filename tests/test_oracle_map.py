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
