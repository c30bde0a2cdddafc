"""Pins the plain-C restatement (oracle/liboracle.so) bit-for-bit against the reference's own headers
compiled into oracle/_ref/libnvbio_ref.so, on seeded random inputs (incl. binary alphabets = many ties).
Skipped when oracle/_ref is not built."""
import numpy as np
import pytest
import pyoracle as po
import util


@pytest.mark.parametrize("n,alpha", [(1, 4), (2, 4), (63, 4), (64, 4), (65, 4), (1000, 2), (4096, 4), (100000, 4)])
def test_fm_index_build_and_queries(oracle, ref, n, alpha):
    rng = np.random.default_rng(n * 7 + alpha)
    text = rng.integers(0, alpha, n).astype(np.uint8)
    a, b = ref.build_index(text, True), oracle.build_index(text, True)
    assert a.primary == b.primary and (a.L2 == b.L2).all()
    assert (a.bwt_occ == b.bwt_occ).all() and (a.ssa == b.ssa).all() and (a.sa == b.sa).all()
    Q, qo = util.rand_queries(rng, text, 3000, alpha=alpha)
    ra, rb = ref.fm_match(a, Q, qo), oracle.fm_match(b, Q, qo)
    assert (ra == rb).all()
    assert (ref.fm_match_reverse(a, Q, qo) == oracle.fm_match_reverse(b, Q, qo)).all()      # fmindex_inl.h:252-278
    ok = ra[:, 0] <= ra[:, 1]
    rows = np.concatenate([ra[ok, 0], ra[ok, 1], rng.integers(0, n + 1, 500).astype(np.uint32)])
    pa, sa = ref.fm_locate(a, rows, True)
    pb, sb = oracle.fm_locate(b, rows, True)
    assert (pa == pb).all() and (sa == sb).all()
    k = rng.integers(0, n + 1, 1000).astype(np.uint32)
    k[:3] = [0xFFFFFFFF, n, 0]
    assert (ref.fm_rank4(a, k) == oracle.fm_rank4(b, k)).all()
    rg = np.stack([rng.integers(0, n + 1, 1000), rng.integers(0, n + 1, 1000)], 1).astype(np.uint32)
    rg.sort(1)
    rg[:50, 0] = 0xFFFFFFFF
    rg[50:60, 1] = n
    c = rng.integers(0, 4, 1000).astype(np.uint8)
    assert (ref.fm_rank(a, rg, c) == oracle.fm_rank(b, rg, c)).all()


CASES = [(k, sc, ty) for k in (po.ED, po.SW, po.GOTOH) for sc in util.SCHEMES[k] for ty in (0, 1, 2)]


@pytest.mark.parametrize("kind,scheme,ty", CASES)
def test_full_dp_score_and_traceback(oracle, ref, kind, scheme, ty):
    rng = np.random.default_rng(hash((kind, scheme, ty)) & 0xFFFF)
    a = po.aligner(kind, ty, scheme)
    for alpha in (4, 2):
        P, Po, T, To = util.rand_pairs(rng, 150, (1, 70), lambda M: int(rng.integers(0, 120)), alpha)
        assert (ref.score(a, P, Po, T, To) == oracle.score(a, P, Po, T, To)).all()
        for chk in (16, 32, 64):
            assert util.ops_equal(ref.traceback(a, chk, P, Po, T, To), oracle.traceback(a, chk, P, Po, T, To))


@pytest.mark.parametrize("kind,scheme,ty", CASES)
def test_banded_dp_score_and_traceback(oracle, ref, kind, scheme, ty):
    rng = np.random.default_rng(hash((kind, scheme, ty, 1)) & 0xFFFF)
    a = po.aligner(kind, ty, scheme)
    for alpha in (4, 2):
        for band in (3, 5, 7, 15, 31):
            P, Po, T, To = util.rand_pairs(rng, 120, (1, 80), util.banded_text_len(rng, ty, band), alpha)
            assert (ref.banded_score(a, band, P, Po, T, To) == oracle.banded_score(a, band, P, Po, T, To)).all()
            for chk in (16, 32):
                assert util.ops_equal(ref.banded_traceback(a, band, chk, P, Po, T, To),
                                      oracle.banded_traceback(a, band, chk, P, Po, T, To))


def test_config_shapes(oracle, ref):
    """the BASELINE.json shapes at small counts: C1 150x500 SW LOCAL, C3 250x265 band 15 ED/Gotoh SEMI_GLOBAL"""
    rng = np.random.default_rng(77)
    P, Po, T, To = util.rand_pairs(rng, 64, (150, 151), lambda M: 500)
    a = po.aligner(po.SW, po.LOCAL, (2, -1, -1, -1))
    assert (ref.score(a, P, Po, T, To) == oracle.score(a, P, Po, T, To)).all()
    P, Po, T, To = util.rand_pairs(rng, 64, (250, 251), lambda M: 265)
    for a in (po.aligner(po.ED, po.SEMI_GLOBAL), po.aligner(po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3))):
        assert (ref.banded_score(a, 15, P, Po, T, To) == oracle.banded_score(a, 15, P, Po, T, To)).all()
        assert util.ops_equal(ref.banded_traceback(a, 15, 16, P, Po, T, To), oracle.banded_traceback(a, 15, 16, P, Po, T, To))


@pytest.mark.parametrize("kind", [po.SW, po.GOTOH])
@pytest.mark.parametrize("ty", [0, 1, 2])
def test_quality_aware_schemes(oracle, ref, kind, ty):
    """quality-aware scoring schemes (alignment.h:411-436, 488-505): the restatement with match(q) / mismatch(a, b, q) tables and
    per-position pattern qualities against the reference's own quality overloads instantiated with table-driven schemes -- nvBowtie's
    (scoring.h:279-286: constant match, quality-scaled mismatch, distinct read/reference gap penalties) and an arbitrary one;
    full and banded, scoring and traceback"""
    rng = np.random.default_rng(1000 + 10 * kind + ty)
    bt_m, bt_x = po.QScheme.bowtie2(2 if ty == 1 else 0, 2, 6)
    schemes = [((0, 0, -8, -3) if kind == po.GOTOH else (0, 0, -3, -4), po.QScheme(bt_m, bt_x, -7, -2)),
               ((0, 0, -4, -2) if kind == po.GOTOH else (0, 0, -2, -2), po.QScheme(rng.integers(0, 5, 256), -rng.integers(0, 7, 256), -3, -3))]
    for params, qs in schemes:
        a = po.aligner(kind, ty, params)
        for alpha in (4, 2):
            P, Po, T, To = util.rand_pairs(rng, 120, (1, 60), lambda M: int(rng.integers(0, 100)), alpha)
            Q = rng.integers(0, 60, len(P)).astype(np.uint8)
            Q[rng.integers(0, max(1, len(P)), 40) % max(1, len(P))] = 255
            assert (ref.score(a, P, Po, T, To, qs=qs, quals=Q) == oracle.score(a, P, Po, T, To, qs=qs, quals=Q)).all()
            # the tables at quality 0 are what a NULL quality string reads
            assert (ref.score(a, P, Po, T, To, qs=qs) == oracle.score(a, P, Po, T, To, qs=qs)).all()
            assert util.ops_equal(ref.traceback(a, 16, P, Po, T, To, qs=qs, quals=Q), oracle.traceback(a, 16, P, Po, T, To, qs=qs, quals=Q))
            for band in (7, 15, 31):
                P, Po, T, To = util.rand_pairs(rng, 120, (1, 60), util.banded_text_len(rng, ty, band), alpha)
                Q = rng.integers(0, 60, len(P)).astype(np.uint8)
                assert (ref.banded_score(a, band, P, Po, T, To, qs=qs, quals=Q) == oracle.banded_score(a, band, P, Po, T, To, qs=qs, quals=Q)).all(), band
                assert util.ops_equal(ref.banded_traceback(a, band, 16, P, Po, T, To, qs=qs, quals=Q),
                                      oracle.banded_traceback(a, band, 16, P, Po, T, To, qs=qs, quals=Q)), band


@pytest.mark.parametrize("alpha", [4, 2])
def test_text_blocking_scoring(oracle, ref, alpha):
    """TextBlockingTag (alignment.h:266-279): the oracle's transposition identity against the reference's own text-blocking aligners
    (aln::transpose of every simple aligner, compiled in ref_wrap_aln.inl), incl. binary alphabets where LOCAL ties abound"""
    rng = np.random.default_rng(40 + alpha)
    pat, pat_off, txt, txt_off = util.rand_pairs(rng, 500, (0, 90), lambda M: max(0, M + int(rng.integers(-8, 40))), alpha=alpha)
    differs = 0
    for kind, sc in ((po.ED, (0, -1, -1, -1)), (po.SW, (2, -1, -2, -3)), (po.SW, (1, -1, -1, -1)), (po.GOTOH, (2, -2, -5, -3)), (po.GOTOH, (1, -1, -1, -1))):
        for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
            a = po.aligner(kind, ty, sc)
            want = ref.score_text_blocking(a, pat, pat_off, txt, txt_off)
            got = oracle.score_text_blocking(a, pat, pat_off, txt, txt_off)
            assert (got == want).all(), (kind, ty, sc)
            differs += int((want != ref.score(a, pat, pat_off, txt, txt_off)).sum())
    assert differs > 20          # the blocking does change LOCAL tie winners
