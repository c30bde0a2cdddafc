"""-m gpu: nvb_map_reads (seed -> locate -> extend -> reduce -> traceback on the device) vs the same composition
of the oracle's pinned entry points (oracle/oracle_map.c): hits and op lists bit for bit."""
import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
import pyoracle as po
import gpu_util as gu
import map_util

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["default", "share_all"])
def sharing_mode(request):
    """every test of this file runs twice: with the library's default (the per-call k-mer table of nvb_map_reads only for
    calls of >= 65536 seeds) and with the table forced for every call"""
    ctx = nvb.Context.default(0)
    ctx.fm_batch_sharing(1 if request.param == "share_all" else 1 << 16)
    yield request.param
    ctx.fm_batch_sharing(1 << 16)


def hits_np(t):
    return t.cpu().numpy().view(np.uint32).reshape(-1, 6).copy().view(nvb.HIT_DT).reshape(-1)


def same(got_hits, got_ops, want_hits, want_ops):
    if not (got_hits.view(np.uint32) == want_hits.view(np.uint32)).all():
        bad = np.nonzero(got_hits != want_hits)[0][:5]
        print("hit mismatch", bad, got_hits[bad], want_hits[bad])
        return False
    for i in range(len(got_hits)):
        k = int(got_hits["n_ops"][i])
        if not (got_ops[i, :k] == want_ops[i, :k]).all():
            print("ops mismatch at", i)
            return False
    return True


CASES = [
    (po.GOTOH, (0, -5, -8, -3), 15, 4, 0, 1, -60),
    (po.ED, (0, -1, -1, -1), 15, 4, 0, 1, -20),
    (po.SW, (1, -2, -3, -3), 7, 2, 0, 1, -1000),
    (po.GOTOH, (2, -3, -6, -2), 15, 8, 6, 0, 0),
    (po.GOTOH, (0, -5, -8, -3), 12, 3, 0, 1, -60),      # a band only the int32 kernels take
    (po.ED, (0, -1, -1, -1), 15, 40, 0, 1, -20),         # more than 1024 possible hits per read: global sort path
    (po.ED, (0, -1, -1, -1), 15, 20, 0, 1, -20),         # up to 1024 possible hits per read: the 64 KB shared-memory sort
]


@pytest.mark.parametrize("kind,scores,band,max_hits,max_range,both,min_score", CASES)
def test_map_reads(oracle, kind, scores, band, max_hits, max_range, both, min_score):
    rng = np.random.default_rng(band * 31 + kind + max_hits)
    G = 60000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    genome[7000:7600] = genome[21000:21600]
    genome[30000:30300] = genome[21100:21400]
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 3000)
    pc = map_util.map_params(band, kind, scores, 20, 10, max_hits, max_range, both, min_score)
    want_hits, want_ops, want_nc = oracle.map_reads(oidx, genome, sym, off, pc)
    params = nvb.MapParams(nvb.Aligner(kind, nvb.SEMI_GLOBAL, *scores), 20, 10, max_hits, max_range, band, bool(both), min_score)
    rs = gu.sset(sym, off)
    for rpb in (0, 700, 1100):                            # one batch, several ragged batches; the last with the k-mer jump table
        if rpb == 1100:
            fmi.build_ktab(8)
        hits, ops, stats = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params, reads_per_batch=rpb)
        assert same(hits_np(hits), ops.cpu().numpy(), want_hits, want_ops)
        assert stats["candidates"] == int(want_nc.sum())
        assert stats["aligned"] == int(((want_hits["flags"] & 2) != 0).sum())
    # host entry point
    h_hits, h_ops, _ = nvb.map_reads_host(fmi, words, sym, off, params,
                                          ops_h=np.zeros((len(off) - 1, nvb.ops_stride_for(150, band)), np.uint8))
    assert same(h_hits, h_ops, want_hits, want_ops)
    # host entry point, several pipelined batches, CIGARs instead of op lists (io::Cigar: op | len << 2, forward order)
    cig = np.zeros((len(off) - 1, 24), np.uint16)
    c_hits, _, _ = nvb.map_reads_host(fmi, words, sym, off, params, cigar_h=cig, reads_per_batch=640)
    n_cig = c_hits["flags"] >> 8
    plain = c_hits.copy(); plain["flags"] &= 0xFF
    assert (plain.view(np.uint32) == want_hits.view(np.uint32)).all()
    for i in range(0, len(want_hits), 7):
        k = int(want_hits["n_ops"][i])
        fwd = want_ops[i, :k][::-1]
        got = "".join("%d%s" % (int(e) >> 2, "MID"[int(e) & 3]) for e in cig[i, :int(n_cig[i])])
        assert got == po.rle(fwd), (i, got, po.rle(fwd))
    assert (want_hits["flags"] & 2).astype(bool).sum() > 2000 or min_score >= 0


@pytest.mark.parametrize("seed_len,interval", [(6, 12), (9, 7), (12, 10), (31, 16)])
def test_map_reads_seed_lengths(oracle, seed_len, interval):
    """seeds shorter than, equal to and longer than the depth of the per-call k-mer table (no table below 8 symbols)"""
    rng = np.random.default_rng(seed_len)
    genome = rng.integers(0, 4, 20000).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 1500)
    pc = map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), seed_len, interval, 3, 0, 1, -60)
    want_hits, want_ops, want_nc = oracle.map_reads(oidx, genome, sym, off, pc)
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), seed_len, interval, 3, 0, 15, True, -60)
    rs = gu.sset(sym, off)
    hits, ops, stats = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params)
    assert same(hits_np(hits), ops.cpu().numpy(), want_hits, want_ops)
    assert stats["candidates"] == int(want_nc.sum())


MODES = [
    dict(seed_mismatches=1, max_seed_ranges=4),
    dict(seed_mismatches=1, max_seed_ranges=0),                                   # default: 8 ranges per seed
    dict(hit_priority=1),                                                         # no caps: same rows as the default selection
    dict(hit_priority=1, max_seed_hits=5, max_rows=9),
    dict(hit_priority=1, max_seed_hits=3),
    dict(hit_priority=1, max_rows=1),
    dict(seed_mismatches=1, max_seed_ranges=6, hit_priority=1, max_seed_hits=10, max_rows=14),
]


@pytest.mark.parametrize("mode", MODES)
def test_map_reads_modes(oracle, mode):
    """approximate seeds (nvBowtie's map<CHECK_EXACT> on the forward index) and hit selection by SA-range size inside the pipeline"""
    rng = np.random.default_rng(len(mode) * 13 + sum(mode.values()))
    G = 50000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    genome[7000:7600] = genome[21000:21600]
    genome[30000:30300] = genome[21100:21400]
    genome[40000:40200] = genome[21200:21400]
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 2500, sub=0.04)
    pc = map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), 20, 10, 4, 0, 1, -60, **mode)
    want_hits, want_ops, want_nc = oracle.map_reads(oidx, genome, sym, off, pc)
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), 20, 10, 4, 0, 15, True, -60, **mode)
    rs = gu.sset(sym, off)
    for rpb in (0, 900):
        hits, ops, stats = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params, reads_per_batch=rpb)
        assert same(hits_np(hits), ops.cpu().numpy(), want_hits, want_ops)
        assert stats["candidates"] == int(want_nc.sum())
    h_hits, h_ops, _ = nvb.map_reads_host(fmi, words, sym, off, params, ops_h=np.zeros((len(off) - 1, nvb.ops_stride_for(150, 15)), np.uint8), reads_per_batch=700)
    assert same(h_hits, h_ops, want_hits, want_ops)
    if mode == dict(hit_priority=1):
        plain, _, _ = oracle.map_reads(oidx, genome, sym, off, map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), 20, 10, 4, 0, 1, -60))
        assert (plain == want_hits).all()
    assert ((want_hits["flags"] & 2) != 0).sum() > 1500


@pytest.mark.parametrize("mode", [dict(max_seed_ranges=4), dict(max_seed_ranges=0), dict(max_seed_ranges=3, hit_priority=1, max_seed_hits=9, max_rows=11)])
@pytest.mark.parametrize("dense", [0, 2])
def test_map_reads_reverse_index(oracle, mode, dense):
    """both halves of nvBowtie's approximate seeds: forward index (substitution in the first half of a seed) and the index of the
    reversed text (substitution in the last half, mapping_inl.h:399-402), rows of the latter located on it and mapped back"""
    rng = np.random.default_rng(91 + len(mode) + dense)
    G = 50000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    genome[7000:7600] = genome[21000:21600]
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    rfmi, ridx = gu.fm_index_pair(oracle, genome[::-1].copy())
    if dense:
        fmi.build_dssa(dense); rfmi.build_dssa(dense)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 2500, sub=0.05)
    pc = map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), 20, 10, 4, 0, 1, -60, seed_mismatches=1, reverse_index=ridx, **mode)
    want_hits, want_ops, want_nc = oracle.map_reads(oidx, genome, sym, off, pc)
    fwd_hits, _, fwd_nc = oracle.map_reads(oidx, genome, sym, off, map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), 20, 10, 4, 0, 1, -60, seed_mismatches=1, **mode))
    assert want_nc.sum() > fwd_nc.sum()                                          # the reverse half contributes candidates
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), 20, 10, 4, 0, 15, True, -60, seed_mismatches=1, reverse_index=rfmi, **mode)
    rs = gu.sset(sym, off)
    for rpb in (0, 800):
        hits, ops, stats = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params, reads_per_batch=rpb)
        assert same(hits_np(hits), ops.cpu().numpy(), want_hits, want_ops)
        assert stats["candidates"] == int(want_nc.sum())
    h_hits, h_ops, _ = nvb.map_reads_host(fmi, words, sym, off, params, ops_h=np.zeros((len(off) - 1, nvb.ops_stride_for(150, 15)), np.uint8), reads_per_batch=700)
    assert same(h_hits, h_ops, want_hits, want_ops)


def test_map_reads_edges(oracle):
    rng = np.random.default_rng(3)
    genome = rng.integers(0, 4, 5000).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    params = nvb.MapParams()
    # no reads
    e = gu.sset(np.zeros(0, np.uint8), np.zeros(1, np.uint32))
    hits, ops, stats = nvb.map_reads(fmi, words, e.symbols, e.offsets, params, max_read_len=10)
    assert hits.shape[0] == 0 and stats["aligned"] == 0
    # reads that cannot seed (too short, all N) and junk
    reads = [genome[100:110], np.full(80, 4, np.uint8), rng.integers(0, 4, 100).astype(np.uint8), genome[4900:5000], genome[0:90]]
    sym, off = po.concat(reads)
    rs = gu.sset(sym, off)
    hits, ops, stats = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params)
    want_hits, want_ops, _ = oracle.map_reads(oidx, genome, sym, off, map_util.map_params(min_score=-(1 << 29)))
    assert same(hits_np(hits), ops.cpu().numpy(), want_hits, want_ops)
    assert (want_hits["flags"][:3] == 0).all() and (want_hits["flags"][3:] == 2).all()
    # out-of-contract arguments fail loudly
    with pytest.raises(nvb.NvbError):
        nvb.map_reads(fmi, words, rs.symbols, rs.offsets, nvb.MapParams(nvb.GotohAligner(nvb.LOCAL, 2, -3, -6, -2)))


def make_pairs(rng, genome, n, L=100, frag=(180, 420)):
    """mate 1 = forward prefix, mate 2 = reverse complement of the suffix of a fragment; every 8th pair swaps the strands,
    every 11th has an unrelated mate 2, every 13th a fragment longer than the window"""
    G = len(genome)
    reads = []
    for i in range(n):
        F = int(rng.integers(*frag)) if i % 13 else 900
        p = int(rng.integers(0, G - F))
        m1 = genome[p:p + L].copy()
        m2 = (3 - genome[p + F - L:p + F])[::-1].copy()
        if i % 11 == 10:
            m2 = rng.integers(0, 4, L).astype(np.uint8)
        for m in (m1, m2):
            k = rng.random(L) < 0.01
            m[k] = (m[k] + 1) % 4
        if i % 8 == 7:
            m1, m2 = m2, m1
        reads += [np.ascontiguousarray(m1, dtype=np.uint8), np.ascontiguousarray(m2, dtype=np.uint8)]
    return po.concat(reads)


def test_pair_hits(oracle):
    """nvb_pair_hits over independently mapped mates vs the oracle's pairing of the oracle's hits"""
    rng = np.random.default_rng(99)
    G = 80000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    R, Ro = make_pairs(rng, genome, 1500)
    mp = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), max_hits=4, band=15, min_score=-60)
    hits, _, _ = nvb.map_reads(fmi, words, torch.from_numpy(R).to(gu.dev()), torch.from_numpy(Ro.view(np.int32)).to(gu.dev()), mp)
    want_hits, _, _ = oracle.map_reads(oidx, genome, R, Ro, map_util.map_params(min_score=-60))
    assert (hits_np(hits).view(np.uint32) == want_hits.view(np.uint32)).all()
    for lo, hi in ((0, 500), (200, 400), (0, 1000)):
        got = nvb.pair_hits(hits, lo, hi).cpu().numpy().view(np.uint32).reshape(-1, 4).copy().view(nvb.PAIR_DT).reshape(-1)
        want = oracle.pair_hits(want_hits, lo, hi)
        assert (got.view(np.uint32) == want.view(np.uint32)).all()
    got = oracle.pair_hits(want_hits, 0, 500)
    conc = (got["flags"] & 4) != 0
    idx = np.arange(len(got))
    assert conc[(idx % 11 != 10) & (idx % 13 != 0)].mean() > 0.97       # sampled fragments pair up, whichever mate is forward
    assert not conc[idx % 13 == 0].any()                                  # 900-bp fragments fall outside the window


def test_rescue_mates(oracle):
    """nvb_rescue_mates (opposite-mate window search, score_inl.h:318-411) vs the oracle's composition of the pinned full-matrix
    score / traceback entry points, starting from the same mapped hits; then pairing again"""
    rng = np.random.default_rng(4242)
    G = 60000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    L, reads = 100, []
    n = 600
    for i in range(n):
        F = int(rng.integers(180, 420))
        p = int(rng.integers(0, G - F)) if i % 50 else (0 if i % 100 else G - F)      # some fragments hug the genome ends
        m1 = genome[p:p + L].copy()
        m2 = (3 - genome[p + F - L:p + F])[::-1].copy()
        victim = m2 if i % 2 else m1
        if i % 5 != 4:
            # a substitution every 11 symbols: no exact 20-mer seed survives, the alignment still scores 9 x -5 = -45
            for k in range(5, L, 11):
                victim[k] = (victim[k] + 1 + i % 3) % 4
            if i % 7 == 0:                                                          # and an indel
                victim = np.delete(victim, 50) if i % 14 else np.insert(victim, 50, 2)
        if i % 9 == 8:
            victim = rng.integers(0, 4, L).astype(np.uint8)                           # unrelated: must stay unaligned
        if i % 2:
            m2 = victim
        else:
            m1 = victim
        if i % 8 == 7:
            m1, m2 = m2, m1
        reads += [np.ascontiguousarray(m1, dtype=np.uint8), np.ascontiguousarray(m2, dtype=np.uint8)]
    sym, off = po.concat(reads)
    pc = map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), 20, 10, 4, 0, 1, -60)
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), 20, 10, 4, 0, 15, True, -60)
    want_hits, want_ops, _ = oracle.map_reads(oidx, genome, sym, off, pc)
    rs = gu.sset(sym, off)
    hits, ops, _ = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params)
    assert same(hits_np(hits), ops.cpu().numpy(), want_hits, want_ops)
    single = ((want_hits["flags"][0::2] & 2) != 0) != ((want_hits["flags"][1::2] & 2) != 0)
    assert single.sum() > 350
    stride = ops.shape[1]
    w_hits, w_ops = want_hits.copy(), np.zeros((2 * n, stride), np.uint8)
    w_ops[:, :want_ops.shape[1]] = want_ops[:, :stride]
    a = po.aligner(po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3))
    wc, wr = oracle.rescue_mates(genome, sym, off, a, -60, 500, w_hits, w_ops)
    gc, gr = nvb.rescue_mates(hits, ops, rs.symbols, rs.offsets, words, G, params.aligner, -60, 500)
    assert (gc, gr) == (wc, wr) and wr > 300 and wc > wr
    got = hits_np(hits)
    assert same(got, ops.cpu().numpy(), w_hits, w_ops)
    rescued = (got["flags"] & 4) != 0
    assert rescued.sum() == wr and ((got["flags"][rescued] & 2) != 0).all()
    # untouched mates keep their hits bit for bit
    assert (got[~rescued] == want_hits[~rescued]).all()
    # pairing after the rescue: most of the rescued pairs are concordant now
    pairs = nvb.pair_hits(hits, 0, 500).cpu().numpy().view(np.uint32).reshape(-1).view(nvb.PAIR_DT)
    want_pairs = oracle.pair_hits(w_hits, 0, 500)
    assert (pairs.view(np.uint32) == want_pairs.view(np.uint32)).all()
    was_single = np.nonzero(single)[0]
    assert ((pairs["flags"][was_single] & nvb.PAIR_CONCORDANT) != 0).sum() > 280
    # a second call finds nothing new to do for the rescued pairs
    gc2, gr2 = nvb.rescue_mates(hits, ops, rs.symbols, rs.offsets, words, G, params.aligner, -60, 500)
    assert gc2 == wc - wr and gr2 == 0


def test_multi_context_host_entries(oracle):
    """nvb_map_reads_host_multi / nvb_fm_match_host_multi (one host thread per context, contiguous slices of one host
    batch, disjoint ranges of one result array): three contexts -- on this box all on device 0, each with its own stream and
    scratch -- give the results of the single-context calls and of the oracle, for slice counts that do and do not divide the batch"""
    rng = np.random.default_rng(808)
    G = 50000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 2501)
    pc = map_util.map_params(15, po.GOTOH, (0, -5, -8, -3), 20, 10, 4, 0, 1, -60)
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), 20, 10, 4, 0, 15, True, -60)
    want_hits, want_ops, _ = oracle.map_reads(oidx, genome, sym, off, pc)
    L = int(np.max(np.diff(off.astype(np.int64))))
    stride = nvb.ops_stride_for(L, 15)
    ctxs = [nvb.Context(0) for _ in range(3)]
    try:
        for k in (1, 2, 3):
            ops_h = np.zeros((len(off) - 1, stride), dtype=np.uint8)
            hits_h, _, st = nvb.map_reads_host_multi(ctxs[:k], [fmi] * k, [words] * k, sym, off, params, ops_h=ops_h, reads_per_batch=500)
            assert same(hits_h, ops_h, want_hits, want_ops), k
            assert st["aligned"] == int((want_hits["flags"] & 2 != 0).sum())
        # seeding: ragged 8-bit queries, and fixed-length 4-bit packed seeds
        import util
        Q, qo = util.rand_queries(rng, genome, 70001, len_range=(0, 30))
        want = oracle.fm_match(oidx, Q, qo)
        for k in (1, 2, 3):
            assert (nvb.match_host_multi(ctxs[:k], [fmi] * k, Q, qo) == want).all(), k
        n, Lq = 100003, 20
        pos = rng.integers(0, G - Lq, n)
        F = genome[(pos[:, None] + np.arange(Lq)[None, :])].reshape(-1).copy()
        F[rng.integers(0, len(F), 500)] = 4
        fo = (np.arange(n + 1, dtype=np.int64) * Lq).astype(np.uint32)
        wantf = oracle.fm_match(oidx, F, fo)
        for k in (2, 3):
            got = nvb.match_host_multi(ctxs[:k], [fmi] * k, nvb.pack_symbols(F, 4), None, query_len=Lq, count=n, bits=4)
            assert (got == wantf).all(), k
    finally:
        for c in ctxs:
            c.close()
