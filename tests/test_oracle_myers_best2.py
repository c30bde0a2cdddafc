"""Pins the C restatement of MyersTag banded edit distance (nvbio/alignment/myers/myers_banded_inl.h:183-296)
and of Best2Sink (nvbio/alignment/sink_inl.h:54-82) against the compiled reference (oracle/_ref), CPU only."""
import numpy as np
import pytest

import pyoracle as po
from util import rand_pairs, banded_text_len

HAVE_REF = po.Oracle.available("ref")
pytestmark = pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built (needs /root/reference)")

MIN = -(1 << 30)     # Field_traits<int32>::min() -- what the batch drivers pass (batched_inl.h:931)


@pytest.fixture(scope="module")
def libs():
    return po.Oracle("oracle"), po.Oracle("ref")


@pytest.mark.parametrize("alphabet", [2, 4, 5])
@pytest.mark.parametrize("band", [7, 15, 16, 31, 32])
def test_myers_vs_reference(libs, alphabet, band):
    o, r = libs
    rng = np.random.default_rng(100 * alphabet + band)
    for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
        # MyersBitVectors<5> leaves B[4] uninitialised (myers_banded_inl.h:139-145), so texts with an N are undefined
        # in the reference: symbols stay in 0..3 here; patterns with an N are fine (load only)
        S = rand_pairs(rng, 200, (0, 70), lambda M: max(0, M + int(rng.integers(-2, band + 6))), alpha=min(alphabet, 4))
        if alphabet == 5:
            P = S[0].copy(); P[rng.random(len(P)) < 0.02] = 4; S = (P,) + S[1:]
        for ms in (MIN, 0, -3, -10, -1000, -40000, 65536 - 4):   # the int16 truncation of min_score is part of the contract
            a = o.banded_myers(ty, alphabet, band, ms, *S)
            b = r.banded_myers(ty, alphabet, band, ms, *S)
            assert (a == b).all(), (ty, ms, int((a != b).sum()))


def test_myers_batch_min_score_quirk(libs):
    """through the batch drivers min_score = Field_traits<int32>::min() truncates to int16 0: only exact matches report"""
    o, _ = libs
    rng = np.random.default_rng(3)
    S = rand_pairs(rng, 300, (20, 60), lambda M: M + 15, related=1.0)
    a = o.banded_myers(po.SEMI_GLOBAL, 4, 16, MIN, *S)
    assert ((a["score"] == 0) | (a["score"] == MIN)).all() and (a["score"] == 0).any() and (a["score"] == MIN).any()


def test_myers_distance_equals_dp_in_band(libs):
    """with a permissive min_score the best Myers score equals the banded DP edit distance (the same band)"""
    o, _ = libs
    rng = np.random.default_rng(4)
    S = rand_pairs(rng, 500, (20, 60), lambda M: M + 15, related=1.0)
    a = o.banded_myers(po.SEMI_GLOBAL, 4, 16, -1000, *S)
    d = o.banded_score(po.aligner(po.ED, po.SEMI_GLOBAL, (0, 0, 0, 0)), 16, *S)
    assert (a["score"] >= d["score"]).mean() > 0.95


SCHEMES = [(po.ED, (0, 0, 0, 0)), (po.SW, (2, -1, -1, -1)), (po.SW, (1, -1, -1, -1)), (po.GOTOH, (2, -2, -5, -3)),
           (po.GOTOH, (1, -1, -1, -1))]


@pytest.mark.parametrize("kind,par", SCHEMES)
@pytest.mark.parametrize("band", [0, 3, 7, 15, 31])
def test_best2_vs_reference(libs, kind, par, band):
    o, r = libs
    rng = np.random.default_rng(7 + band + 10 * kind)
    for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
        al = po.aligner(kind, ty, par)
        for dd in (0, 1, 5, 30):
            n_of_m = banded_text_len(rng, ty, band) if band else (lambda M: M + int(rng.integers(0, 40)))
            S = rand_pairs(rng, 120, (1, 50), n_of_m, alpha=2 if dd in (1, 30) else 4)
            a = o.score_best2(al, band, dd, *S)
            b = r.score_best2(al, band, dd, *S)
            assert (a == b).all(), (ty, dd, int((a != b).sum()))
            if ty != po.GLOBAL:
                one = o.banded_score(al, band, *S) if band else o.score(al, *S)
                assert (one["score"] == a["score1"]).all() and (one["x"] == a["x1"]).all()
