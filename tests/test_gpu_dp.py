"""-m gpu: DP kernels through the C ABI vs the oracle: scores, sinks (incl. tie-breaks on binary
alphabets), sources and op lists must be bit-identical. Golden vectors of the reference
(nvbio-test/alignment_test.cu:641-863) are checked on the GPU path too."""
import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
import pyoracle as po
import util
import gpu_util as gu

pytestmark = pytest.mark.gpu

CASES = [(k, sc, ty) for k in (po.ED, po.SW, po.GOTOH) for sc in util.SCHEMES[k] for ty in (0, 1, 2)]


@pytest.fixture(params=[0, 1, 2], ids=["auto", "generic", "v1"])
def ctx(request):
    """0 = the library's own choice, 1 = int32 kernels only, 2 = packed kernels without the sheared-coordinate ones"""
    c = nvb.Context.default(0)
    c.force_generic(request.param)
    yield c
    c.force_generic(0)


@pytest.mark.parametrize("kind,scheme,ty", CASES)
def test_full_score_and_traceback(oracle, ctx, kind, scheme, ty):
    rng = np.random.default_rng(hash((kind, scheme, ty)) & 0xFFFF)
    a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
    for alpha in (4, 2):
        P, Po, T, To = util.rand_pairs(rng, 700, (1, 70), lambda M: int(rng.integers(0, 120)), alpha)
        ps, ts = gu.sset(P, Po), gu.sset(T, To)
        got = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts))
        want = oracle.score(oa, P, Po, T, To)
        assert (got == want).all(), np.nonzero(got != want)[0][:5]
        for chk in (16, 64):
            got = gu.tb_np(nvb.batch_alignment_traceback(a, chk, ps, ts))
            assert gu.tb_equal(got, oracle.traceback(oa, chk, P, Po, T, To))


@pytest.mark.parametrize("kind,scheme,ty", CASES)
def test_full_score_wide_shapes(oracle, ctx, kind, scheme, ty):
    """full-matrix scoring at the shapes the wavefront kernels spread over many lanes (configs[0]: 150 x 500) and at ragged
    ones: couples of unequal pattern and text lengths, patterns of 1..16 stripes, texts shorter than patterns"""
    rng = np.random.default_rng(hash((kind, scheme, ty, 3)) & 0xFFFF)
    a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
    for alpha, m_range, n_of_m in ((4, (140, 151), lambda M: 500), (2, (1, 260), lambda M: int(rng.integers(1, 420))),
                                   (4, (150, 151), lambda M: int(rng.integers(100, 300)))):
        P, Po, T, To = util.rand_pairs(rng, 401, m_range, n_of_m, alpha)
        ps, ts = gu.sset(P, Po), gu.sset(T, To)
        got = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts))
        want = oracle.score(oa, P, Po, T, To)
        bad = np.nonzero(got != want)[0]
        assert len(bad) == 0, (bad[:5], got[bad[:5]], want[bad[:5]], np.diff(Po.astype(np.int64))[bad[:5]], np.diff(To.astype(np.int64))[bad[:5]])


@pytest.mark.parametrize("kind,scheme,ty", CASES)
def test_banded_score_and_traceback(oracle, ctx, kind, scheme, ty):
    rng = np.random.default_rng(hash((kind, scheme, ty, 1)) & 0xFFFF)
    a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
    for alpha in (4, 2):
        for band in (3, 5, 7, 15, 31, 12):
            P, Po, T, To = util.rand_pairs(rng, 500, (1, 80), util.banded_text_len(rng, ty, band), alpha)
            ps, ts = gu.sset(P, Po), gu.sset(T, To)
            got = gu.sinks_np(nvb.batch_banded_alignment_score(a, band, ps, ts))
            want = oracle.banded_score(oa, band, P, Po, T, To)
            assert (got == want).all(), (band, np.nonzero(got != want)[0][:5])
            for chk in (16, 32):
                got = gu.tb_np(nvb.batch_banded_alignment_traceback(a, band, chk, ps, ts))
                assert gu.tb_equal(got, oracle.banded_traceback(oa, band, chk, P, Po, T, To)), (band, chk)


def test_symbols_with_n_and_long_shapes(oracle, ctx):
    """N symbols (codes > 3) compare equal only to the same code; C1/C3-shaped pairs"""
    rng = np.random.default_rng(77)
    for kind, scheme, ty in [(po.SW, (2, -1, -1, -1), 1), (po.GOTOH, (2, -1, -2, -1), 1), (po.ED, (0, -1, -1, -1), 2)]:
        a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
        P, Po, T, To = util.rand_pairs(rng, 600, (140, 151), lambda M: 500, 4)
        P = P.copy(); T = T.copy()
        P[rng.integers(0, len(P), 300)] = 4
        T[rng.integers(0, len(T), 300)] = rng.integers(4, 7, 300)
        ps, ts = gu.sset(P, Po), gu.sset(T, To)
        got = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts))
        assert (got == oracle.score(oa, P, Po, T, To)).all()
        P, Po, T, To = util.rand_pairs(rng, 600, (240, 251), lambda M: M + 15, 4)
        P = P.copy(); P[rng.integers(0, len(P), 200)] = 4
        ps, ts = gu.sset(P, Po), gu.sset(T, To)
        got = gu.sinks_np(nvb.batch_banded_alignment_score(a, 15, ps, ts))
        assert (got == oracle.banded_score(oa, 15, P, Po, T, To)).all()
        got = gu.tb_np(nvb.batch_banded_alignment_traceback(a, 15, 16, ps, ts))
        assert gu.tb_equal(got, oracle.banded_traceback(oa, 15, 16, P, Po, T, To))


@pytest.mark.parametrize("kind,scheme", [(po.ED, (0, -1, -1, -1)), (po.SW, (2, -1, -1, -1)), (po.SW, (1, -3, -2, -2)),
                                         (po.SW, (3, -2, -4, -1)), (po.GOTOH, (0, -5, -8, -3)), (po.GOTOH, (2, -6, -5, -3)),
                                         (po.GOTOH, (1, -1, -1, -1)), (po.GOTOH, (2, -7, -4, -3))])
@pytest.mark.parametrize("ty", [0, 2])
def test_banded_long_ragged(oracle, ctx, kind, scheme, ty):
    """patterns spanning several staging chunks, couples of unequal lengths, texts ending inside the band, a few N"""
    rng = np.random.default_rng(hash((kind, scheme, ty, 7)) & 0xFFFF)
    a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
    for band in (15, 7):
        for alpha in (4, 2):
            P, Po, T, To = util.rand_pairs(rng, 1501, (1, 330), util.banded_text_len(rng, ty, band), alpha)
            if alpha == 4:
                P = P.copy(); T = T.copy()
                P[rng.integers(0, len(P), 20)] = 4
                T[rng.integers(0, len(T), 20)] = rng.integers(4, 256, 20)
            ps, ts = gu.sset(P, Po), gu.sset(T, To)
            got = gu.sinks_np(nvb.batch_banded_alignment_score(a, band, ps, ts))
            want = oracle.banded_score(oa, band, P, Po, T, To)
            assert (got == want).all(), (band, alpha, np.nonzero(got != want)[0][:5])


def test_empty_and_degenerate(oracle, ctx):
    a, oa = gu.aligner(po.SW, 1, (2, -1, -1, -1)), po.aligner(po.SW, 1, (2, -1, -1, -1))
    # empty batch
    z = nvb.StringSet.from_numpy(np.zeros(0, np.uint8), np.zeros(1, np.uint32), gu.dev())
    assert nvb.batch_alignment_score(a, z, z).shape[0] == 0
    # empty patterns / texts inside a batch
    P, Po = po.concat([np.zeros(0, np.uint8), po.dna("ACGT"), po.dna("A"), np.zeros(0, np.uint8)])
    T, To = po.concat([po.dna("ACGT"), np.zeros(0, np.uint8), po.dna("A"), np.zeros(0, np.uint8)])
    ps, ts = gu.sset(P, Po), gu.sset(T, To)
    for kind, scheme in [(po.SW, (2, -1, -1, -1)), (po.GOTOH, (2, -1, -2, -1)), (po.ED, (0, -1, -1, -1))]:
        a, oa = gu.aligner(kind, 1, scheme), po.aligner(kind, 1, scheme)
        assert (gu.sinks_np(nvb.batch_alignment_score(a, ps, ts)) == oracle.score(oa, P, Po, T, To)).all()
    # banded: text shorter than the pattern is refused exactly as the reference does
    a, oa = gu.aligner(po.ED, 2, (0, -1, -1, -1)), po.aligner(po.ED, 2, (0, -1, -1, -1))
    P, Po = po.concat([po.dna("ACGTACGT"), po.dna("ACGT")])
    T, To = po.concat([po.dna("ACGT"), po.dna("ACGTAC")])
    ps, ts = gu.sset(P, Po), gu.sset(T, To)
    assert (gu.sinks_np(nvb.batch_banded_alignment_score(a, 7, ps, ts)) == oracle.banded_score(oa, 7, P, Po, T, To)).all()


GOLD_P, GOLD_T = "ACAACTA", "AAACACCCTAACACACTAAA"


def cigar_of(ops, n):
    # the reference test's backtracer appends ops as pushed (end -> start): alignment_test.cu:96-135
    return po.rle(ops[:n])


@pytest.mark.parametrize("kind,ty,want", [
    (po.SW, 0, "1M2D3M1D3M10D"), (po.SW, 1, "4M1D3M"), (po.SW, 2, "4M1D3M"),
    (po.GOTOH, 0, "1M2D3M1D3M10D"), (po.GOTOH, 1, "4M1D3M"), (po.GOTOH, 2, "4M1D3M")])
def test_golden_cigars(ctx, kind, ty, want):
    """nvbio-test/alignment_test.cu:721-753"""
    a = gu.aligner(kind, ty, (2, -1, -1, -1))
    P, Po = po.concat([po.dna(GOLD_P)]); T, To = po.concat([po.dna(GOLD_T)])
    out, ops, n_ops = gu.tb_np(nvb.batch_alignment_traceback(a, 16, gu.sset(P, Po), gu.sset(T, To)))
    assert cigar_of(ops[0], n_ops[0]) == want
    if kind == po.GOTOH and ty != 0:
        out, ops, n_ops = gu.tb_np(nvb.batch_banded_alignment_traceback(a, 7, 16, gu.sset(P, Po), gu.sset(T, To)))
        assert cigar_of(ops[0], n_ops[0]) == want


def test_host_entry_points_and_cigar(oracle, ctx):
    rng = np.random.default_rng(9)
    a, oa = gu.aligner(po.GOTOH, 2, (0, -5, -8, -3)), po.aligner(po.GOTOH, 2, (0, -5, -8, -3))
    P, Po, T, To = util.rand_pairs(rng, 3000, (100, 151), lambda M: M + 15, 4)
    got = nvb.batch_score_host(a, 15, P, Po, T, To)
    want = oracle.banded_score(oa, 15, P, Po, T, To)
    assert (got["score"] == want["score"]).all() and (got["x"] == want["x"]).all() and (got["y"] == want["y"]).all()
    stride = 2 * 151 + 15 + 8
    got = nvb.batch_traceback_host(a, 15, 16, P, Po, T, To, stride)
    want = oracle.banded_traceback(oa, 15, 16, P, Po, T, To)
    assert gu.tb_equal(got, want)
    # full-matrix host forms
    a, oa = gu.aligner(po.SW, 1, (2, -1, -1, -1)), po.aligner(po.SW, 1, (2, -1, -1, -1))
    P, Po, T, To = util.rand_pairs(rng, 1000, (50, 151), lambda M: 300, 4)
    got = nvb.batch_score_host(a, 0, P, Po, T, To)
    want = oracle.score(oa, P, Po, T, To)
    assert (got["score"] == want["score"]).all() and (got["x"] == want["x"]).all() and (got["y"] == want["y"]).all()
    tb = nvb.batch_traceback_host(a, 0, 64, P, Po, T, To, 151 + 300 + 8)
    want = oracle.traceback(oa, 64, P, Po, T, To)
    assert gu.tb_equal(tb, want)
    # CIGAR packing (io::Cigar: op | len << 2, forward order)
    ops_d = torch.from_numpy(tb[1]).to(gu.dev()); nops_d = torch.from_numpy(tb[2].view(np.int32)).to(gu.dev())
    cig, ncig = nvb.ops_to_cigar(ops_d, nops_d)
    cig = cig.cpu().numpy().view(np.uint16); ncig = ncig.cpu().numpy()
    for i in range(50):
        s = "".join("%d%s" % (w >> 2, "MID"[w & 3]) for w in cig[i, :ncig[i]])
        assert s == po.rle(tb[1][i, :tb[2][i]][::-1])


@pytest.mark.parametrize("bits", [4, 2])
def test_banded_packed_string_sets(oracle, ctx, bits):
    """4- and 2-bit packed string sets (the reference's read / genome streams) through the banded entry points: expanded to
    bytes in context scratch for the fast kernels, read directly by the int32 ones; same sinks and op lists either way"""
    rng = np.random.default_rng(40 + bits)
    for kind, scheme in ((po.ED, (0, 0, 0, 0)), (po.GOTOH, (0, -5, -8, -3)), (po.SW, (2, -1, -1, -1))):
        for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
            for band in (15, 7, 12):
                a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
                P, Po, T, To = util.rand_pairs(rng, 600, (1, 120), util.banded_text_len(rng, ty, band), 4)
                if bits == 4:
                    P = P.copy(); T = T.copy()
                    P[rng.random(len(P)) < 0.003] = 4; T[rng.random(len(T)) < 0.003] = 9
                dev = gu.dev()
                pk = torch.from_numpy(nvb.pack_symbols(P, bits).view(np.int32)).to(dev)
                tk = torch.from_numpy(nvb.pack_symbols(T, bits).view(np.int32)).to(dev)
                po_d = torch.from_numpy(Po.view(np.int32)).to(dev); to_d = torch.from_numpy(To.view(np.int32)).to(dev)
                maxM = int(np.diff(Po.astype(np.int64)).max()); maxN = int(np.diff(To.astype(np.int64)).max())
                ps, ts = nvb.StringSet(pk, po_d, bits, maxM), nvb.StringSet(tk, to_d, bits, maxN)
                got = gu.sinks_np(nvb.batch_banded_alignment_score(a, band, ps, ts))
                want = oracle.banded_score(oa, band, P, Po, T, To)
                assert (got == want).all(), (kind, ty, band, np.nonzero(got != want)[0][:5])
                got = gu.tb_np(nvb.batch_banded_alignment_traceback(a, band, 16, ps, ts))
                assert gu.tb_equal(got, oracle.banded_traceback(oa, band, 16, P, Po, T, To)), (kind, ty, band)


@pytest.mark.parametrize("kind", [po.SW, po.GOTOH])
@pytest.mark.parametrize("ty", [0, 1, 2])
def test_quality_aware_schemes(oracle, kind, ty):
    """nvb_batch_*_q: quality-aware scoring schemes (alignment.h:411-436, 488-505) through the C ABI vs the oracle (itself pinned on
    the reference's quality overloads, tests/test_oracle_vs_ref.py): nvBowtie's scheme (scoring.h:279-286) and an arbitrary table,
    full and banded, scoring and traceback, packed pattern sets, and the NULL quality string"""
    rng = np.random.default_rng(2000 + 10 * kind + ty)
    bt_m, bt_x = po.QScheme.bowtie2(2 if ty == 1 else 0, 2, 6)
    schemes = [((0, 0, -8, -3) if kind == po.GOTOH else (0, 0, -3, -4), bt_m, bt_x, -7, -2),
               ((0, 0, -4, -2) if kind == po.GOTOH else (0, 0, -2, -2), rng.integers(0, 5, 256), -rng.integers(0, 7, 256), -3, -3)]
    for params, tm, tx, tgo, tge in schemes:
        a, oa = gu.aligner(kind, ty, params), po.aligner(kind, ty, params)
        qs, oqs = nvb.QualityScheme(tm, tx, tgo, tge), po.QScheme(tm, tx, tgo, tge)
        for alpha in (4, 2):
            P, Po, T, To = util.rand_pairs(rng, 300, (1, 70), lambda M: int(rng.integers(0, 120)), alpha)
            Q = rng.integers(0, 60, len(P)).astype(np.uint8)
            Q[rng.integers(0, len(P), 40)] = 255
            ps, ts = gu.sset(P, Po), gu.sset(T, To)
            qd = torch.from_numpy(Q).to(gu.dev())
            got = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts, scheme=qs, quals=qd))
            assert (got == oracle.score(oa, P, Po, T, To, qs=oqs, quals=Q)).all()
            got = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts, scheme=qs))
            assert (got == oracle.score(oa, P, Po, T, To, qs=oqs)).all()
            got = gu.tb_np(nvb.batch_alignment_traceback(a, 16, ps, ts, scheme=qs, quals=qd))
            assert gu.tb_equal(got, oracle.traceback(oa, 16, P, Po, T, To, qs=oqs, quals=Q))
            if alpha == 4:      # 4-bit packed patterns: the qualities stay one byte per symbol
                p4 = nvb.StringSet(torch.from_numpy(nvb.pack_symbols(P, 4).view(np.int32)).to(gu.dev()), ps.offsets, 4, ps.max_len)
                got = gu.sinks_np(nvb.batch_alignment_score(a, p4, ts, scheme=qs, quals=qd))
                assert (got == oracle.score(oa, P, Po, T, To, qs=oqs, quals=Q)).all()
            for band in (7, 15, 31, 12):
                P, Po, T, To = util.rand_pairs(rng, 300, (1, 80), util.banded_text_len(rng, ty, band), alpha)
                Q = rng.integers(0, 60, len(P)).astype(np.uint8)
                ps, ts = gu.sset(P, Po), gu.sset(T, To)
                qd = torch.from_numpy(Q).to(gu.dev())
                got = gu.sinks_np(nvb.batch_banded_alignment_score(a, band, ps, ts, scheme=qs, quals=qd))
                assert (got == oracle.banded_score(oa, band, P, Po, T, To, qs=oqs, quals=Q)).all(), band
                got = gu.tb_np(nvb.batch_banded_alignment_traceback(a, band, 16, ps, ts, scheme=qs, quals=qd))
                assert gu.tb_equal(got, oracle.banded_traceback(oa, band, 16, P, Po, T, To, qs=oqs, quals=Q)), band


@pytest.mark.parametrize("alpha", [4, 2])
def test_text_blocking_scoring(oracle, alpha):
    """nvb_batch_score_text_blocking (TextBlockingTag, alignment.h:266-279) vs the oracle (pinned on the reference's transposed
    aligners): every kind and type, binary alphabets with many LOCAL ties, the configs[0] shape through the packed kernels"""
    rng = np.random.default_rng(140 + alpha)
    pat, pat_off, txt, txt_off = util.rand_pairs(rng, 800, (0, 120), lambda M: max(0, M + int(rng.integers(-8, 60))), alpha=alpha)   # incl. empty strings
    ps, ts = gu.sset(pat, pat_off), gu.sset(txt, txt_off)
    ctx = nvb.Context.default(0)
    for kind, sc in ((po.ED, (0, -1, -1, -1)), (po.SW, (2, -1, -2, -3)), (po.SW, (1, -1, -1, -1)), (po.GOTOH, (2, -2, -5, -3))):
        for ty in (po.GLOBAL, po.LOCAL, po.SEMI_GLOBAL):
            want = oracle.score_text_blocking(po.aligner(kind, ty, sc), pat, pat_off, txt, txt_off)
            for force in (0, 1):
                ctx.force_generic(force)
                got = gu.sinks_np(nvb.batch_alignment_score(gu.aligner(kind, ty, sc), ps, ts, text_blocking=True))
                ctx.force_generic(0)
                assert (got == want).all(), (kind, ty, sc, force)
    # configs[0] shape: the transposed LOCAL batch (500-symbol "patterns") still takes a packed kernel
    P, Po, T, To = util.rand_pairs(rng, 300, (150, 151), lambda M: 500, alpha=alpha)
    a = po.aligner(po.SW, po.LOCAL, (2, -1, -1, -1))
    want = oracle.score_text_blocking(a, P, Po, T, To)
    got = gu.sinks_np(nvb.batch_alignment_score(gu.aligner(po.SW, po.LOCAL, (2, -1, -1, -1)), gu.sset(P, Po), gu.sset(T, To), text_blocking=True))
    assert (got == want).all()
    assert (want != oracle.score(a, P, Po, T, To)).any() or alpha == 4


@pytest.mark.parametrize("bits", [8, 4, 2])
def test_host_scoring_pipelined_and_packed(oracle, bits):
    """nvb_batch_score_host / _packed on batches large enough to be cut into pipelined chunks (ragged strings, chunk boundaries that
    do not fall on word boundaries of the packed streams), banded and full matrix, against the oracle"""
    rng = np.random.default_rng(300 + bits)
    n = 70001
    M = rng.integers(20, 61, n); N = M + rng.integers(15, 30, n)
    Po = np.concatenate([[0], np.cumsum(M)]).astype(np.uint32); To = np.concatenate([[0], np.cumsum(N)]).astype(np.uint32)
    P = rng.integers(0, 4, int(Po[-1])).astype(np.uint8); T = rng.integers(0, 4, int(To[-1])).astype(np.uint8)
    for i in range(0, n, 3):                                          # implant the pattern with a few substitutions
        s = int(rng.integers(0, N[i] - M[i] + 1)); seg = P[Po[i]:Po[i + 1]].copy()
        seg[rng.random(len(seg)) < 0.05] = rng.integers(0, 4)
        T[To[i] + s:To[i] + s + M[i]] = seg
    Ph, Th = (P, T) if bits == 8 else (nvb.pack_symbols(P, bits), nvb.pack_symbols(T, bits))
    for kind, ty, sc, band in ((po.GOTOH, 2, (0, -5, -8, -3), 15), (po.ED, 2, (0, -1, -1, -1), 15), (po.SW, 1, (2, -1, -1, -1), 0), (po.SW, 2, (2, -1, -2, -3), 0)):
        a, oa = gu.aligner(kind, ty, sc), po.aligner(kind, ty, sc)
        got = nvb.batch_score_host(a, band, Ph, Po, Th, To, bits=bits)
        want = oracle.banded_score(oa, band, P, Po, T, To) if band else oracle.score(oa, P, Po, T, To)
        assert (got["score"] == want["score"]).all() and (got["x"] == want["x"]).all() and (got["y"] == want["y"]).all(), (kind, ty, band)


@pytest.mark.parametrize("kind,scheme", [(po.ED, (0, -1, -1, -1)), (po.SW, (2, -1, -1, -1)), (po.SW, (1, -3, -2, -2)), (po.SW, (3, -2, -4, -1)),
                                         (po.SW, (0, -1, -1, -2)), (po.GOTOH, (2, -2, -5, -3)), (po.GOTOH, (0, -5, -8, -3)), (po.GOTOH, (1, -1, -1, -1)),
                                         (po.GOTOH, (2, -1, -2, -1)), (po.GOTOH, (2, -6, -5, -3))])
@pytest.mark.parametrize("ty", [0, 1, 2])
def test_full_traceback_packed(oracle, kind, scheme, ty):
    """full-matrix traceback of every aligner and type on the tagged packed forward passes + walks: configs[0]-shaped
    pairs, couples of equal pattern length (the ones the Gotoh kernel keeps), ragged couples over 1..17 stripes, a few N (those
    pairs go to the int32 kernel), a scratch that forces several launches, an unaligned op stride"""
    c = nvb.Context.default(0)
    c.force_generic(0)
    rng = np.random.default_rng(hash((kind, scheme, ty, 11)) & 0xFFFF)
    a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
    for alpha, m_range, n_of_m, n in ((4, (140, 151), lambda M: 500, 201), (4, (150, 151), lambda M: int(rng.integers(250, 500)), 301),
                                      (2, (100, 101), lambda M: int(rng.integers(1, 300)), 301), (2, (1, 270), lambda M: int(rng.integers(1, 420)), 401),
                                      (4, (1, 70), lambda M: int(rng.integers(1, 120)), 700)):
        P, Po, T, To = util.rand_pairs(rng, n, m_range, n_of_m, alpha)
        if alpha == 4:
            P = P.copy(); T = T.copy()
            P[rng.integers(0, len(P), 5)] = 4
            T[rng.integers(0, len(T), 5)] = rng.integers(4, 7, 5)
        ps, ts = gu.sset(P, Po), gu.sset(T, To)
        want = oracle.traceback(oa, 64, P, Po, T, To)
        got = gu.tb_np(nvb.batch_alignment_traceback(a, 64, ps, ts, ctx=c))
        assert c.last_paths & 4, "the packed traceback did not run"
        assert gu.tb_equal(got, want), (alpha, "default")
        small = torch.empty(int(nvb.lib().nvb_temp_bytes(2, a.c(), 0, 64, ps.max_len, ts.max_len, 128)), dtype=torch.uint8, device=gu.dev())
        got = gu.tb_np(nvb.batch_alignment_traceback(a, 64, ps, ts, temp=small, ops_stride=ps.max_len + ts.max_len + 3, ctx=c))
        assert gu.tb_equal(got, want), (alpha, "small scratch")


@pytest.mark.parametrize("kind,scheme,ty", [(po.SW, (2, -1, -1, -1), 1), (po.SW, (2, -1, -1, -1), 2), (po.ED, (0, -1, -1, -1), 0),
                                            (po.GOTOH, (2, -2, -5, -3), 2), (po.GOTOH, (2, -2, -5, -3), 1), (po.GOTOH, (0, -5, -8, -3), 0)])
def test_pattern_n_on_the_packed_kernels(oracle, kind, scheme, ty):
    """an N (code > 3) in the PATTERN only is a column that mismatches every text symbol: the packed full-matrix kernels keep such
    pairs (reads with an N against clean genome windows); a pair whose text holds a code > 3 as well still goes to the int32
    kernel, where equal codes match (every 16th pair here)"""
    c = nvb.Context.default(0)
    c.force_generic(0)
    rng = np.random.default_rng(hash((kind, scheme, ty, 23)) & 0xFFFF)
    a, oa = gu.aligner(kind, ty, scheme), po.aligner(kind, ty, scheme)
    n, M, N = 320, 150, 400
    P, Po, T, To = util.rand_pairs(rng, n, (M, M + 1), lambda m: N, 4)
    P = P.copy(); T = T.copy()
    for i in range(n):
        pos = rng.integers(0, M, 3)
        P[Po[i] + pos] = rng.integers(4, 6, 3)
        if i % 16 == 0:
            T[To[i] + rng.integers(0, N, 6)] = rng.integers(4, 6, 6)
    ps, ts = gu.sset(P, Po), gu.sset(T, To)
    got = gu.sinks_np(nvb.batch_alignment_score(a, ps, ts, ctx=c))
    want = oracle.score(oa, P, Po, T, To)
    assert (got == want).all(), np.nonzero(got != want)[0][:5]
    got = gu.tb_np(nvb.batch_alignment_traceback(a, 64, ps, ts, ctx=c))
    assert gu.tb_equal(got, oracle.traceback(oa, 64, P, Po, T, To))
