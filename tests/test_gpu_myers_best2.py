"""-m gpu: MyersTag banded edit distance and Best2Sink scoring through the C ABI vs the oracle (which
tests/test_oracle_myers_best2.py pins against the compiled reference), bit for bit."""
import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
import pyoracle as po
import util
import gpu_util as gu

pytestmark = pytest.mark.gpu

MIN = nvb.MYERS_BATCH_MIN_SCORE


@pytest.mark.parametrize("alphabet", [2, 4, 5])
@pytest.mark.parametrize("band", [1, 7, 15, 16, 31, 32])
def test_myers_scores_and_sinks(oracle, alphabet, band):
    rng = np.random.default_rng(1000 * alphabet + band)
    for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
        # texts with symbols above 3 included: the oracle defines alphabet 5's vector 4 as empty at start
        P, Po, T, To = util.rand_pairs(rng, 1500, (0, 90), lambda M: max(0, M + int(rng.integers(-2, band + 8))), alpha=min(alphabet, 4))
        if alphabet != 2:
            P = P.copy(); T = T.copy()
            P[rng.random(len(P)) < 0.01] = 4; T[rng.random(len(T)) < 0.01] = 4
            P[rng.random(len(P)) < 0.002] = 9; T[rng.random(len(T)) < 0.002] = 200
        ps, ts = gu.sset(P, Po), gu.sset(T, To)
        a = nvb.MyersAligner(ty, nvb.MyersTag(alphabet))
        for ms in (MIN, 0, -2, -8, -1000, -40000, 65536 - 3):
            got = gu.sinks_np(nvb.batch_banded_myers_score(a, band, ps, ts, min_score=ms))
            want = oracle.banded_myers(ty, alphabet, band, ms, P, Po, T, To)
            assert (got == want).all(), (ty, ms, np.nonzero(got != want)[0][:5])


def test_myers_packed_sets_and_c3_shape(oracle):
    """4-bit packed string sets (the reference's read streams) and the configs[2] shape 250 x 265"""
    rng = np.random.default_rng(5)
    P, Po, T, To = util.rand_pairs(rng, 4000, (250, 251), lambda M: M + 15, related=1.0)
    ps, ts = gu.sset(P, Po), gu.sset(T, To)
    a = nvb.MyersAligner(po.SEMI_GLOBAL, nvb.MyersTag(4))
    got = gu.sinks_np(nvb.batch_banded_myers_score(a, 16, ps, ts, min_score=-1000))
    want = oracle.banded_myers(po.SEMI_GLOBAL, 4, 16, -1000, P, Po, T, To)
    assert (got == want).all()
    # the same scores as the DP edit distance over the 16-wide band whenever the optimum is in band
    dp = oracle.banded_score(po.aligner(po.ED, po.SEMI_GLOBAL, (0, 0, 0, 0)), 16, P, Po, T, To)
    assert (got["score"] >= dp["score"]).mean() > 0.95
    pk = torch.from_numpy(nvb.pack_symbols(P, 4).view(np.int32)).to(gu.dev())
    tk = torch.from_numpy(nvb.pack_symbols(T, 4).view(np.int32)).to(gu.dev())
    ps4 = nvb.StringSet(pk, ps.offsets, 4, 250)
    ts4 = nvb.StringSet(tk, ts.offsets, 4, 265)
    got4 = gu.sinks_np(nvb.batch_banded_myers_score(a, 16, ps4, ts4, min_score=-1000))
    assert (got4 == want).all()


def best2_np(t):
    a = t.cpu().numpy()
    out = np.zeros(a.shape[0], dtype=po.SINK2_DT)
    out["score1"], out["score2"] = a[:, 0], a[:, 1]
    for k, name in enumerate(("x1", "y1", "x2", "y2")):
        out[name] = a[:, 2 + k].view(np.uint32)
    return out


SCHEMES = [(po.ED, (0, 0, 0, 0)), (po.SW, (2, -1, -1, -1)), (po.SW, (1, -1, -1, -1)), (po.GOTOH, (2, -2, -5, -3)),
           (po.GOTOH, (1, -1, -1, -1))]


@pytest.mark.parametrize("kind,par", SCHEMES)
@pytest.mark.parametrize("band", [0, 3, 7, 15, 31, 12])
def test_best2_sinks(oracle, kind, par, band):
    rng = np.random.default_rng(17 + band + 10 * kind)
    for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
        a, oa = gu.aligner(kind, ty, par), po.aligner(kind, ty, par)
        for dd in (0, 1, 5, 30):
            n_of_m = util.banded_text_len(rng, ty, band) if band else (lambda M: M + int(rng.integers(0, 60)))
            P, Po, T, To = util.rand_pairs(rng, 400, (1, 70), n_of_m, alpha=2 if dd in (1, 30) else 4)
            ps, ts = gu.sset(P, Po), gu.sset(T, To)
            got = best2_np(nvb.batch_alignment_score_best2(a, band, dd, ps, ts))
            want = oracle.score_best2(oa, band, dd, P, Po, T, To)
            assert (got == want).all(), (ty, dd, np.nonzero(got != want)[0][:5])
