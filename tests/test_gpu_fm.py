"""-m gpu: FM-index kernels through the C ABI vs the oracle (bit-exact raw uint32 results).
Follows the reference's own method: nvbio-test/rank_test.cu:46-79 (rank vs counts) and
nvbio-test/fmindex_test.cu:576-707 (match/locate vs SA; host == device)."""
import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
from nvbio_gpl_b200 import fmindex as fmx
import pyoracle as po
import util
import gpu_util as gu

pytestmark = pytest.mark.gpu


def u32(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.fixture(autouse=True, params=["default", "share_all"])
def sharing_mode(request):
    """every test of this file runs twice: with the library's default choice between the plain one-query-per-thread kernel
    and the per-batch k-mer table (fm_batch_table in fm_kernels.cu), and with EVERY match() batch forced through the table"""
    ctx = nvb.Context.default(0)
    ctx.fm_batch_sharing(1 if request.param == "share_all" else 1 << 16)
    yield request.param
    ctx.fm_batch_sharing(1 << 16)


@pytest.mark.parametrize("n,alpha", [(1, 4), (2, 4), (63, 4), (64, 4), (65, 4), (1000, 2), (4096, 4), (300000, 4), (5000, 1)])
def test_match_rank_locate(oracle, n, alpha):
    rng = np.random.default_rng(n * 7 + alpha)
    text = rng.integers(0, alpha, n).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    ref_idx = oracle.build_index(text)
    assert fmi.primary == ref_idx.primary and (oidx.bwt_occ == ref_idx.bwt_occ).all() and (oidx.ssa == ref_idx.ssa).all()

    Q, qo = util.rand_queries(rng, text, 20000, len_range=(0, 40), alpha=alpha)
    qs = gu.sset(Q, qo)
    got = u32(nvb.match(fmi, qs.symbols, qs.offsets))
    want = oracle.fm_match(oidx, Q, qo)
    assert (got == want).all()

    # continuing from given ranges == matching the concatenation (fmindex_inl.h:212-239)
    start = torch.from_numpy(want.view(np.int32).copy()).to(gu.dev())
    got2 = u32(nvb.match(fmi, qs.symbols, qs.offsets, ranges=start))
    QQ = [np.concatenate([Q[qo[i]:qo[i + 1]], Q[qo[i]:qo[i + 1]]]) for i in range(2000)]
    Q2, q2o = po.concat(QQ)
    assert (got2[:2000] == oracle.fm_match(oidx, Q2, q2o)).all()

    ok = want[:, 0] <= want[:, 1]
    rows = np.concatenate([want[ok, 0], want[ok, 1], rng.integers(0, n + 1, 5000).astype(np.uint32),
                           np.array([0, n, fmi.primary], dtype=np.uint32)])
    rows_d = torch.from_numpy(rows.view(np.int32)).to(gu.dev())
    want_pos = oracle.fm_locate(oidx, rows)
    assert (u32(nvb.locate(fmi, rows_d)) == want_pos).all()
    assert (u32(nvb.locate(fmi, rows_d, two_phase=True)) == want_pos).all()

    k = rng.integers(0, n + 1, 5000).astype(np.uint32)
    k[:3] = [0xFFFFFFFF, n, 0]
    assert (u32(nvb.rank4(fmi, torch.from_numpy(k.view(np.int32)).to(gu.dev()))) == oracle.fm_rank4(oidx, k)).all()

    rg = np.stack([rng.integers(0, n + 1, 5000), rng.integers(0, n + 1, 5000)], 1).astype(np.uint32)
    rg.sort(1)
    rg[:50, 0] = 0xFFFFFFFF
    rg[50:60, 1] = n
    rg[60:80, 0] = rg[60:80, 1]
    c = rng.integers(0, 4, 5000).astype(np.uint8)
    got = u32(nvb.rank(fmi, torch.from_numpy(rg.view(np.int32)).to(gu.dev()), torch.from_numpy(c).to(gu.dev())))
    assert (got == oracle.fm_rank(oidx, rg, c)).all()


def test_located_positions_are_occurrences(oracle):
    """every locate() of a non-empty match points at an occurrence (fmindex_test.cu:604-656)"""
    rng = np.random.default_rng(5)
    n = 200000
    text = rng.integers(0, 4, n).astype(np.uint8)
    fmi, _ = gu.fm_index_pair(oracle, text)
    starts = rng.integers(0, n - 12, 4000)
    Q, qo = po.concat([text[s:s + 12] for s in starts])
    qs = gu.sset(Q, qo)
    r = u32(nvb.match(fmi, qs.symbols, qs.offsets))
    assert (r[:, 0] <= r[:, 1]).all()
    for col in (0, 1):
        pos = u32(nvb.locate(fmi, torch.from_numpy(r[:, col].copy().view(np.int32)).to(gu.dev())))
        for p, s in zip(pos[:500], starts[:500]):
            assert (text[p:p + 12] == text[s:s + 12]).all()


def test_host_entry_points_and_fuse(oracle):
    rng = np.random.default_rng(11)
    n = 100000
    text = rng.integers(0, 4, n).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    Q, qo = util.rand_queries(rng, text, 30000, len_range=(18, 23))
    want = oracle.fm_match(oidx, Q, qo)
    got = nvb.match_host(fmi, Q, qo)
    assert (got == want).all()
    rows = want[want[:, 0] <= want[:, 1], 0].copy()
    assert (nvb.locate_host(fmi, rows) == oracle.fm_locate(oidx, rows)).all()
    # fusing a plain BWT on the device reproduces the image (fmindex_impl.cu:283-318)
    words = oidx.bwt_occ.reshape(-1, 8)[:, :4].reshape(-1).copy()
    fused, L2 = fmx.fuse(torch.from_numpy(words.view(np.int32)).to(gu.dev()), n)
    assert (u32(fused) == oidx.bwt_occ).all() and L2 == fmi.L2


def test_packed_query_sets(oracle):
    """2-bit and 4-bit big-endian packed query sets (PackedStream<...,true>) give the same ranges"""
    rng = np.random.default_rng(3)
    n = 50000
    text = rng.integers(0, 4, n).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    Q, qo = util.rand_queries(rng, text, 5000, len_range=(1, 30), p_n=0.0)
    want = oracle.fm_match(oidx, Q, qo)
    off = torch.from_numpy(qo.view(np.int32)).to(gu.dev())
    for bits in (2, 4):
        per = 32 // bits
        nw = (len(Q) + per - 1) // per + 1
        sym = np.zeros(nw * per, dtype=np.uint64)
        sym[:len(Q)] = Q
        sh = (32 - bits - bits * np.arange(per)).astype(np.uint64)
        words = (sym.reshape(nw, per) << sh).sum(1).astype(np.uint32)
        wd = torch.from_numpy(words.view(np.int32)).to(gu.dev())
        got = u32(nvb.match(fmi, wd, off, bits=bits))
        assert (got == want).all()


def test_host_entry_packed_and_chunked(oracle):
    """host buffers: 8-bit and 4-bit packed streams, batch large enough to be pipelined in several chunks"""
    rng = np.random.default_rng(31)
    text = rng.integers(0, 4, 50000).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    Q, qo = util.rand_queries(rng, text, 1_200_000 // 8, len_range=(5, 28))
    Q = np.tile(Q, 9); qo = np.concatenate([[0], np.cumsum(np.tile(np.diff(qo.astype(np.int64)), 9))]).astype(np.uint32)
    want = oracle.fm_match(oidx, Q, qo)
    assert (nvb.match_host(fmi, Q, qo) == want).all()
    assert (nvb.match_host(fmi, nvb.pack_symbols(Q, 4), qo, bits=4) == want).all()
    Q2 = Q.copy(); Q2[Q2 > 3] = 0
    assert (nvb.match_host(fmi, nvb.pack_symbols(Q2, 2), qo, bits=2) == oracle.fm_match(oidx, Q2, qo)).all()


def test_host_entry_fixed_length(oracle):
    """queries of one length through nvb_fm_match_host_fixed (no offsets over PCIe), 8-bit and packed, batch long
    enough for the ramped chunk schedule and every buffer set of the three-stage pipeline"""
    rng = np.random.default_rng(32)
    text = rng.integers(0, 4, 80000).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    n, L = 2_400_007, 20
    pos = rng.integers(0, len(text) - L, n)
    Q = text[(pos[:, None] + np.arange(L)[None, :])].copy()
    junk = rng.random(n) < 0.2
    Q[junk] = rng.integers(0, 4, (int(junk.sum()), L))
    Q[rng.integers(0, n, 2000), rng.integers(0, L, 2000)] = 4
    Q = Q.reshape(-1)
    qo = (np.arange(n + 1, dtype=np.int64) * L).astype(np.uint32)
    want = oracle.fm_match(oidx, Q, qo)
    assert (nvb.match_host_fixed(fmi, Q, L, n) == want).all()
    assert (nvb.match_host_fixed(fmi, nvb.pack_symbols(Q, 4), L, n, bits=4) == want).all()
    assert (nvb.match_host(fmi, Q, qo) == want).all()
    Q2 = Q & 3
    assert (nvb.match_host_fixed(fmi, nvb.pack_symbols(Q2, 2), L, n, bits=2) == oracle.fm_match(oidx, Q2, qo)).all()
    assert (nvb.match_host_fixed(fmi, Q[:L * 5], L, 5) == want[:5]).all()


def test_fmindex_filter_rank_and_locate(oracle):
    """FMIndexFilter (filter_inl.h:75-110,261-393): scanned range sizes and (position, string id) hits, checked against the
    composition of the oracle's match / locate with the reference's slot arithmetic"""
    rng = np.random.default_rng(41)
    text = rng.integers(0, 4, 40000).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    Q, qo = util.rand_queries(rng, text, 5000, len_range=(0, 14))      # short queries: many multi-hit ranges, some empty
    qs = gu.sset(Q, qo)
    flt = nvb.FMIndexFilter()
    n_hits = flt.rank(fmi, qs.symbols, qs.offsets)
    want_r = oracle.fm_match(oidx, Q, qo)
    sizes = (1 + want_r[:, 1].astype(np.int64) - want_r[:, 0].astype(np.int64)) & 0xFFFFFFFF
    want_slots = np.cumsum(sizes)
    assert (flt.ranges.cpu().numpy().view(np.uint32) == want_r).all()
    assert (flt.slots.cpu().numpy() == want_slots).all() and n_hits == int(want_slots[-1])
    for b, e in [(0, min(n_hits, 200000)), (n_hits // 3, min(n_hits, n_hits // 3 + 50001)), (max(0, n_hits - 7), n_hits)]:
        hits = flt.locate(b, e).cpu().numpy().view(np.uint32)
        idx = np.arange(b, e, dtype=np.int64)
        slot = np.searchsorted(want_slots, idx, side="right")
        base = np.where(slot > 0, want_slots[np.maximum(slot, 1) - 1], 0)
        rows = (want_r[slot, 0].astype(np.int64) + (idx - base)).astype(np.uint32)
        assert (hits[:, 1] == slot).all()
        assert (hits[:, 0] == oracle.fm_locate(oidx, rows)).all()


def test_nvbwt_files_round_trip(oracle, tmp_path):
    """nvBWT on-disk format (nvBWT.cu:303-340 / fmindex_impl.cu:111-321): what the library writes is what an independent
    writer following the reference produces; both loaders (numpy + device fuse, and the C entry point) give back the index"""
    rng = np.random.default_rng(51)
    for n in (1, 63, 64, 1000, 40037):
        text = rng.integers(0, 4, n).astype(np.uint8)
        fmi, oidx = gu.fm_index_pair(oracle, text)
        prefix = str(tmp_path / ("idx%d" % n))
        nvb.save_index(fmi, prefix)
        # independent writer from the oracle's image: de-interleave the BWT words, header as nvBWT writes it
        img = oidx.bwt_occ.reshape(-1, 8)
        words = img[:, :4].reshape(-1)[: (n + 15) // 16]
        want_bwt = np.concatenate([np.array([oidx.primary] + [int(v) for v in oidx.L2[1:5]], dtype=np.uint32), words.astype(np.uint32)])
        got_bwt = np.fromfile(prefix + ".bwt", dtype=np.uint32)
        assert (got_bwt == want_bwt).all()
        got_sa = np.fromfile(prefix + ".sa", dtype=np.uint32)
        assert got_sa[0] == oidx.primary and got_sa[5] == 16 and got_sa[6] == n and (got_sa[7:] == oidx.ssa[1:]).all()
        back = nvb.load_index(prefix, gu.dev())
        b2, s2 = back.host_arrays()
        assert back.length == n and back.primary == oidx.primary and list(back.L2) == [int(v) for v in oidx.L2]
        assert (b2 == oidx.bwt_occ).all() and (s2 == oidx.ssa).all()
        c = nvb.CLoadedIndex(prefix)
        Q, qo = util.rand_queries(rng, text, 2000, len_range=(0, 12))
        qs = gu.sset(Q, qo)
        want = oracle.fm_match(oidx, Q, qo)
        assert (c.match(qs.symbols, qs.offsets).cpu().numpy().view(np.uint32) == want).all()
        rows = want[want[:, 0] <= want[:, 1], 0].copy()
        if len(rows):
            pos = c.locate(torch.from_numpy(rows.view(np.int32)).to(gu.dev())).cpu().numpy().view(np.uint32)
            assert (pos == oracle.fm_locate(oidx, rows)).all()
        c.close()


@pytest.mark.parametrize("k", [1, 3, 8, 10, 12])
def test_kmer_jump_table(oracle, k):
    """match() through the k-mer jump table returns the raw ranges of the plain loop (fmindex_inl.h:186-239), also for
    absent queries (the pair the loop stopped on), queries shorter than k and queries with an N in their last k symbols"""
    rng = np.random.default_rng(100 + k)
    for n, alpha in ((50000, 4), (3000, 2), (40, 4)):
        text = rng.integers(0, alpha, n).astype(np.uint8)
        fmi, oidx = gu.fm_index_pair(oracle, text)
        Q, qo = util.rand_queries(rng, text, 30000, len_range=(0, 30), alpha=4, p_from_text=0.6, p_n=0.1)
        qs = gu.sset(Q, qo)
        want = oracle.fm_match(oidx, Q, qo)
        plain = u32(nvb.match(fmi, qs.symbols, qs.offsets))
        fmi.build_ktab(k)
        tab = u32(fmi.ktab).reshape(-1, 2)
        assert tab.shape[0] == 4 ** k
        got = u32(nvb.match(fmi, qs.symbols, qs.offsets))
        assert (plain == want).all() and (got == want).all()
        # every table entry is match() of its k-mer
        ids = rng.integers(0, 4 ** k, min(4 ** k, 2000))
        kmers = ((ids[:, None] >> (2 * (k - 1 - np.arange(k)))[None, :]) & 3).astype(np.uint8)
        assert (tab[ids] == oracle.fm_match(oidx, kmers.reshape(-1), (np.arange(len(ids) + 1) * k).astype(np.uint32))).all()
        for bits in (4, 2):
            if bits == 2 and (Q > 3).any():
                Q2 = Q & 3
                w = torch.from_numpy(nvb.pack_symbols(Q2, 2).view(np.int32)).to(gu.dev())
                assert (u32(nvb.match(fmi, w, qs.offsets, bits=2)) == oracle.fm_match(oidx, Q2, qo)).all()
            else:
                w = torch.from_numpy(nvb.pack_symbols(Q, bits).view(np.int32)).to(gu.dev())
                assert (u32(nvb.match(fmi, w, qs.offsets, bits=bits)) == want).all()
        # the filter goes through the same kernel
        flt = nvb.FMIndexFilter()
        assert flt.rank(fmi, qs.symbols, qs.offsets) == int(((want[:, 1] + 1 - want[:, 0]).astype(np.uint32)).astype(np.uint64).sum())


@pytest.mark.parametrize("sa_int", [2, 4, 8])
def test_dense_ssa(oracle, sa_int):
    """locate through the optional denser sampled SA returns the positions of the plain walk (fmindex_inl.h:364-394): one-phase,
    two-phase and the FMIndexFilter form; every table entry is locate() of its row"""
    rng = np.random.default_rng(200 + sa_int)
    for n, alpha in ((60000, 4), (3000, 2), (41, 4), (16, 4)):
        text = rng.integers(0, alpha, n).astype(np.uint8)
        fmi, oidx = gu.fm_index_pair(oracle, text)
        rows = np.concatenate([rng.integers(0, n + 1, 20000), np.arange(min(n + 1, 200)), [n, fmi.primary]]).astype(np.uint32)
        want = oracle.fm_locate(oidx, rows)
        rd = torch.from_numpy(rows.view(np.int32)).to(gu.dev())
        assert (u32(nvb.locate(fmi, rd)) == want).all()
        fmi.build_dssa(sa_int)
        tab = u32(fmi.dssa)
        assert len(tab) == (n + sa_int) // sa_int
        assert (tab == oracle.fm_locate(oidx, (np.arange(len(tab)) * sa_int).astype(np.uint32))).all()
        assert (u32(nvb.locate(fmi, rd)) == want).all()
        assert (u32(nvb.locate(fmi, rd, two_phase=True)) == want).all()
        Q, qo = util.rand_queries(rng, text, 3000, len_range=(6, 20), alpha=alpha, p_from_text=0.9, p_n=0.0)
        qs = gu.sset(Q, qo)
        flt = nvb.FMIndexFilter()
        nh = flt.rank(fmi, qs.symbols, qs.offsets)
        hits_dense = u32(flt.locate(0, nh))
        fmi.drop_dssa()
        flt2 = nvb.FMIndexFilter()
        assert flt2.rank(fmi, qs.symbols, qs.offsets) == nh
        assert (hits_dense == u32(flt2.locate(0, nh))).all()


def test_match_approx(oracle):
    """nvb_fm_match_approx (nvBowtie's map<find_exact>, mapping_inl.h:114-184) vs the oracle composition: ranges in push order,
    counts, truncation at cap, N symbols, empty and short queries, packed sets"""
    rng = np.random.default_rng(31)
    for n, alpha in ((50000, 4), (3000, 2), (33, 4)):
        text = rng.integers(0, alpha, n).astype(np.uint8)
        fmi, oidx = gu.fm_index_pair(oracle, text)
        Q, qo = util.rand_queries(rng, text, 8000, len_range=(0, 30), alpha=alpha, p_from_text=0.8, p_n=0.05)
        qs = gu.sset(Q, qo)
        for len1, cap in ((0, 96), (10, 64), (14, 4), (40, 8)):
            for fe in (True, False):
                want_r, want_c = oracle.fm_match_approx(oidx, Q, qo, len1, fe, cap=cap)
                got_r, got_c = nvb.match_approx(fmi, qs.symbols, qs.offsets, len1, fe, cap=cap)
                assert (u32(got_c) == want_c).all(), (len1, fe)
                assert (u32(got_r).reshape(-1, cap, 2) == want_r).all(), (len1, fe)
        w = torch.from_numpy(nvb.pack_symbols(Q, 4).view(np.int32)).to(gu.dev())
        want_r, want_c = oracle.fm_match_approx(oidx, Q, qo, 10, True, cap=64)
        got_r, got_c = nvb.match_approx(fmi, w, qs.offsets, 10, True, cap=64, bits=4)
        assert (u32(got_c) == want_c).all() and (u32(got_r).reshape(-1, 64, 2) == want_r).all()


def test_match_reverse_and_reverse_index_files(oracle, tmp_path):
    """nvb_fm_match_reverse (fmindex_inl.h:252-278) vs the oracle (raw ranges, N symbols, empty patterns, packed sets, with
    and without the k-mer table), and nvBowtie's use of it: the reverse index written to / read from .rbwt + .rsa
    (nvb_fmindex_save/load with reverse = 1) finds as many occurrences as match() on the forward index"""
    rng = np.random.default_rng(61)
    for n, alpha in ((60000, 4), (3000, 2), (65, 4), (1, 4)):
        text = rng.integers(0, alpha, n).astype(np.uint8)
        fmi, oidx = gu.fm_index_pair(oracle, text)
        Q, qo = util.rand_queries(rng, text, 20000, len_range=(0, 36), alpha=alpha)
        qs = gu.sset(Q, qo)
        want = oracle.fm_match_reverse(oidx, Q, qo)
        assert (u32(nvb.match_reverse(fmi, qs.symbols, qs.offsets)) == want).all()
        w4 = torch.from_numpy(nvb.pack_symbols(Q, 4).view(np.int32)).to(gu.dev())
        assert (u32(nvb.match_reverse(fmi, w4, qs.offsets, bits=4)) == want).all()
        fmi.build_ktab(6)
        assert (u32(nvb.match_reverse(fmi, qs.symbols, qs.offsets)) == want).all()
        fmi.drop_ktab()
        # the reverse index through the nvBWT files
        rtext = text[::-1].copy()
        rfmi, ridx = gu.fm_index_pair(oracle, rtext)
        prefix = str(tmp_path / ("r%d" % n))
        nvb.save_index(rfmi, prefix, reverse=True)
        import os
        assert os.path.exists(prefix + ".rbwt") and os.path.exists(prefix + ".rsa") and not os.path.exists(prefix + ".bwt")
        back = nvb.load_index(prefix, gu.dev(), reverse=True)
        b2, s2 = back.host_arrays()
        assert back.primary == ridx.primary and (b2 == ridx.bwt_occ).all() and (s2 == ridx.ssa).all()
        c = nvb.CLoadedIndex(prefix, reverse=True)
        got_r = c.match(qs.symbols, qs.offsets, reverse=True).cpu().numpy().view(np.uint32)
        assert (got_r == oracle.fm_match_reverse(ridx, Q, qo)).all()
        fw = oracle.fm_match(oidx, Q, qo)
        cnt = lambda r: np.where(r[:, 0] <= r[:, 1], r[:, 1].astype(np.int64) - r[:, 0] + 1, 0)
        assert (cnt(got_r) == cnt(fw)).all()
        # a located row of the reverse index is an occurrence in the reversed text: position p <-> n - p - len in the forward text
        ok = np.nonzero((got_r[:, 0] <= got_r[:, 1]) & (np.diff(qo.astype(np.int64)) >= 8))[0][:300]
        if len(ok):
            pos = c.locate(torch.from_numpy(got_r[ok, 0].copy().view(np.int32)).to(gu.dev())).cpu().numpy().view(np.uint32)
            for p, i in zip(pos, ok):
                L = int(qo[i + 1] - qo[i])
                assert (rtext[p:p + L][::-1] == Q[qo[i]:qo[i + 1]]).all()
        c.close()


def test_batch_sharing_adversarial(oracle):
    """the per-batch k-mer table on batches large enough for every table depth the test can afford (K = 8, 9, 10), built to
    stress its entry rules: exact duplicates, queries that extend / truncate one another, absent k-mers repeated with different
    tails (dead table entries), N symbols at every position (inside and behind the first K consumed symbols), queries shorter
    than K, empty queries, a binary alphabet; match and match_reverse; 8-bit and packed sets"""
    rng = np.random.default_rng(909)
    ctx = nvb.Context.default(0)
    ctx.fm_batch_sharing(1)
    try:
        for n, alpha, lens, nq in ((200000, 4, (0, 34), 60000), (5000, 2, (0, 90), 400000), (70000, 4, (18, 23), 1500000)):
            text = rng.integers(0, alpha, n).astype(np.uint8)
            fmi, oidx = gu.fm_index_pair(oracle, text)
            base_n = min(nq, 60000)
            Q, qo = util.rand_queries(rng, text, base_n, len_range=lens, alpha=alpha, p_from_text=0.7, p_n=0.08)
            qs = [Q[qo[i]:qo[i + 1]] for i in range(len(qo) - 1)]
            extra = []
            for i in rng.integers(0, len(qs), 30000):
                q = qs[i]
                r = rng.random()
                if r < 0.3:
                    extra.append(q.copy())                                              # duplicate
                elif r < 0.5:
                    extra.append(np.concatenate([rng.integers(0, alpha, int(rng.integers(1, 40))).astype(np.uint8), q]))   # extends (match order)
                elif r < 0.7:
                    extra.append(np.concatenate([q, rng.integers(0, alpha, int(rng.integers(1, 40))).astype(np.uint8)]))   # extends (reverse order)
                elif r < 0.85:
                    extra.append(q[len(q) // 3:].copy())
                else:
                    extra.append(q[:len(q) // 2].copy())
            allq = qs + extra
            reps = max(1, nq // len(allq))                 # repeated to reach the batch size of the deeper tables
            perm = np.concatenate([rng.permutation(len(allq)) for _ in range(reps)])
            Q2, q2o = po.concat([allq[i] for i in perm])
            s2 = gu.sset(Q2, q2o)
            want_f, want_r = oracle.fm_match(oidx, Q2, q2o), oracle.fm_match_reverse(oidx, Q2, q2o)
            got = u32(nvb.match(fmi, s2.symbols, s2.offsets))
            bad = np.nonzero((got != want_f).any(axis=1))[0]
            assert len(bad) == 0, (n, bad[:5], got[bad[:5]], want_f[bad[:5]])
            assert (u32(nvb.match_reverse(fmi, s2.symbols, s2.offsets)) == want_r).all(), n
            w4 = torch.from_numpy(nvb.pack_symbols(Q2, 4).view(np.int32)).to(gu.dev())
            assert (u32(nvb.match(fmi, w4, s2.offsets, bits=4)) == want_f).all()
            Q3 = Q2 & 3
            w2 = torch.from_numpy(nvb.pack_symbols(Q3, 2).view(np.int32)).to(gu.dev())
            assert (u32(nvb.match_reverse(fmi, w2, s2.offsets, bits=2)) == oracle.fm_match_reverse(oidx, Q3, q2o)).all()
    finally:
        ctx.fm_batch_sharing(1 << 16)


@pytest.mark.parametrize("n,Lq,p_n", [(1, 20, 0.0), (70001, 20, 0.01), (300000, 20, 0.001), (5000, 7, 0.2), (1500000, 20, 0.002), (4097, 33, 0.05)])
def test_host_fixed_2bit_with_n_list(oracle, n, Lq, p_n):
    """nvb_fm_match_host_fixed2n: 2-bit seeds + (query, last N) exception list give the ranges of the 8-bit form, bit for bit --
    Ns at the last symbol, the first symbol, several per query, behind the point where the range is already empty"""
    rng = np.random.default_rng(n + Lq)
    G = 60000
    text = rng.integers(0, 4, G).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, text)
    pos = rng.integers(0, G - Lq, n)
    F = text[(pos[:, None] + np.arange(Lq)[None, :])].reshape(-1).copy()
    junk = rng.random(n) < 0.3
    F.reshape(n, Lq)[junk] = rng.integers(0, 4, (int(junk.sum()), Lq))
    F[rng.random(len(F)) < p_n] = 4
    if n > 10:
        F.reshape(n, Lq)[3, Lq - 1] = 4; F.reshape(n, Lq)[5, 0] = 4; F.reshape(n, Lq)[7, :] = 4
    fo = (np.arange(n + 1, dtype=np.int64) * Lq).astype(np.uint32)
    want = oracle.fm_match(oidx, F, fo)
    words, nlist = nvb.pack_symbols_2n(F, Lq)
    assert len(nlist) == int((F.reshape(n, Lq) > 3).any(axis=1).sum())
    got = nvb.match_host_fixed2n(fmi, words, nlist, Lq, n)
    assert (got == want).all()
    # argument checks: unsorted list, position past the query
    if len(nlist) > 1:
        with pytest.raises(nvb.NvbError):
            nvb.match_host_fixed2n(fmi, words, nlist[::-1], Lq, n)
    with pytest.raises(nvb.NvbError):
        nvb.match_host_fixed2n(fmi, words, np.array([[0, Lq]], np.uint32), Lq, n)
