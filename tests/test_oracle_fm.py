"""Property tests of the FM-index restatement, after the reference's own test method:
rank vs running counts (nvbio-test/rank_test.cu:46-79), SSA vs the true suffix array, match/locate vs
brute force (nvbio-test/fmindex_test.cu:576-656)."""
import numpy as np
import pytest
import pyoracle as po
import util


def brute_sa(text):
    n = len(text)
    b = bytes(text.tolist())
    return sorted(range(n + 1), key=lambda i: b[i:])


@pytest.mark.parametrize("n,alpha", [(1, 4), (17, 4), (64, 2), (200, 4), (777, 2), (3000, 4)])
def test_sa_bwt_layout(oracle, n, alpha):
    rng = np.random.default_rng(n)
    text = rng.integers(0, alpha, n).astype(np.uint8)
    idx = oracle.build_index(text, True)
    sa = brute_sa(text)
    assert idx.sa[0] == -1 and list(idx.sa[1:]) == sa[1:]
    assert idx.primary == sa.index(0)
    assert list(idx.ssa[1:]) == [sa[i] for i in range(16, n + 1, 16)] and idx.ssa[0] == 0xFFFFFFFF
    bwt = [int(text[s - 1]) for s in sa if s != 0]
    # unpack the fused layout: 4 big-endian 2-bit words then 4 counters per 64 symbols
    blocks = idx.bwt_occ.reshape(-1, 8)
    got = []
    for i in range(n):
        w = blocks[i >> 6, (i & 63) >> 4]
        got.append(int(w >> (30 - 2 * (i & 15))) & 3)
    assert got == bwt
    run = np.zeros(4, dtype=np.int64)
    for i in range(n):
        if i % 64 == 0:
            assert list(blocks[i >> 6, 4:]) == list(run)
        run[bwt[i]] += 1
    assert list(idx.L2) == [0] + list(np.cumsum(run))


def test_rank_matches_running_counts(oracle):
    rng = np.random.default_rng(3)
    n = 5000
    text = rng.integers(0, 4, n).astype(np.uint8)
    idx = oracle.build_index(text, True)
    sa = brute_sa(text)
    bwt_full = [int(text[s - 1]) if s else -1 for s in sa]          # n+1 rows, -1 at the primary row
    k = np.arange(n + 1, dtype=np.uint32)
    r4 = oracle.fm_rank4(idx, k)
    run = np.zeros(4, dtype=np.int64)
    for i in range(n + 1):
        if bwt_full[i] >= 0:
            run[bwt_full[i]] += 1
        if i < n:   # k == length returns the totals (fmindex_inl.h:41-42)
            assert list(r4[i]) == list(run), i
    assert list(r4[n]) == list(np.diff(idx.L2.astype(np.int64)))
    assert list(oracle.fm_rank4(idx, np.array([0xFFFFFFFF], dtype=np.uint32))[0]) == [0, 0, 0, 0]


@pytest.mark.parametrize("alpha", [4, 2])
def test_match_and_locate_vs_brute_force(oracle, alpha):
    rng = np.random.default_rng(alpha)
    n = 4000
    text = rng.integers(0, alpha, n).astype(np.uint8)
    idx = oracle.build_index(text, True)
    sa = brute_sa(text)
    b = bytes(text.tolist())
    Q, qo = util.rand_queries(rng, text, 400, (1, 14), alpha=alpha, p_n=0.03)
    ranges, blocks = oracle.fm_match(idx, Q, qo, want_blocks=True)
    for i in range(len(qo) - 1):
        q = bytes(Q[qo[i]:qo[i + 1]].tolist())
        rows = [r for r in range(n + 1) if b[sa[r]:sa[r] + len(q)] == q] if max(q) < 4 else []
        x, y = int(ranges[i, 0]), int(ranges[i, 1])
        if rows:
            assert (x, y) == (rows[0], rows[-1])
            pos = oracle.fm_locate(idx, np.arange(x, min(y, x + 20) + 1, dtype=np.uint32))
            assert [int(p) for p in pos] == [sa[r] for r in range(x, min(y, x + 20) + 1)]
        else:
            assert x > y
        if q[-1] >= 4:   # an N met while the range is still non-empty gives exactly (1,0) (fmindex_inl.h:226-228)
            assert (x, y) == (1, 0)
        assert blocks[i] <= 2 * len(q)
    # empty pattern returns the whole range (fmindex_inl.h:196)
    r = oracle.fm_match(idx, np.zeros(0, np.uint8), np.zeros(2, np.uint32))
    assert (int(r[0, 0]), int(r[0, 1])) == (0, n)


def test_locate_every_row(oracle):
    rng = np.random.default_rng(9)
    n = 2000
    text = rng.integers(0, 4, n).astype(np.uint8)
    idx = oracle.build_index(text, True)
    pos, steps = oracle.fm_locate(idx, np.arange(1, n + 1, dtype=np.uint32), want_steps=True)
    assert (pos.astype(np.int64) == idx.sa[1:]).all()
    assert steps.max() < 16 * 8 and (steps[15::16] == 0).all()


def test_match_approx_semantics(oracle):
    """nvBowtie's map<find_exact> (mapping_inl.h:114-184) restated over the pinned rank/rank4 entry points: every emitted range is
    the SA range of the query with exactly one substitution at a step >= len1 (or of the query itself), in push order, and the
    C-restatement and reference-header compositions agree"""
    rng = np.random.default_rng(12)
    text = rng.integers(0, 4, 3000).astype(np.uint8)
    idx = oracle.build_index(text)
    Q, qo = util.rand_queries(rng, text, 300, len_range=(4, 14), alpha=4, p_from_text=0.8, p_n=0.05)
    for len1 in (0, 3, 6, 20):
        for fe in (True, False):
            ranges, counts = oracle.fm_match_approx(idx, Q, qo, len1, fe, cap=64)
            for i in range(len(counts)):
                q = Q[qo[i]:qo[i + 1]]; L = len(q); want = []
                # exact prefix of the search (the LAST len1 symbols) must match somewhere, and the base range must stay non-empty up to step t
                for t in range(min(len1, L), L):
                    base = q[L - t:]                      # symbols consumed before step t
                    if (base > 3).any() or (len(base) and oracle.fm_match(idx, base, np.array([0, len(base)], np.uint32))[0][0] >
                                            oracle.fm_match(idx, base, np.array([0, len(base)], np.uint32))[0][1]):
                        break
                    for sub in range(4):
                        if sub == q[L - 1 - t]:
                            continue
                        m = q.copy(); m[L - 1 - t] = sub
                        r = oracle.fm_match(idx, m, np.array([0, L], np.uint32))[0]
                        if (m > 3).any():
                            continue
                        if r[0] <= r[1]:
                            want.append((int(r[0]), int(r[1])))
                    if q[L - 1 - t] > 3:
                        break
                if fe and not (q > 3).any():
                    r = oracle.fm_match(idx, q, np.array([0, L], np.uint32))[0]
                    if r[0] <= r[1]:
                        want.append((int(r[0]), int(r[1])))
                got = [(int(a), int(b)) for a, b in ranges[i, :counts[i]]]
                assert got == want, (len1, fe, i, got, want)


def test_match_approx_oracle_vs_reference_composition(oracle, ref):
    rng = np.random.default_rng(13)
    for n, alpha in ((20000, 4), (2000, 2)):
        text = rng.integers(0, alpha, n).astype(np.uint8)
        idx = oracle.build_index(text)
        Q, qo = util.rand_queries(rng, text, 2000, len_range=(0, 26), alpha=alpha, p_from_text=0.8, p_n=0.05)
        for len1 in (0, 10, 16):
            for fe in (True, False):
                a = oracle.fm_match_approx(idx, Q, qo, len1, fe, cap=80)
                b = ref.fm_match_approx(idx, Q, qo, len1, fe, cap=80)
                assert (a[1] == b[1]).all() and (a[0] == b[0]).all()


def test_match_reverse_on_the_reverse_index_counts_forward_occurrences(oracle):
    """how nvBowtie uses it (mapping_inl.h:71-86): match_reverse over the index of the REVERSED text finds the pattern's own
    occurrences -- the range sizes equal those of match() over the forward index, and match_reverse(q) == match(reversed q)"""
    rng = np.random.default_rng(77)
    for n, alpha in ((5000, 4), (700, 2)):
        text = rng.integers(0, alpha, n).astype(np.uint8)
        fwd, rev = oracle.build_index(text), oracle.build_index(text[::-1].copy())
        Q, qo = util.rand_queries(rng, text, 3000, len_range=(0, 16), alpha=alpha)
        a, b = oracle.fm_match(fwd, Q, qo), oracle.fm_match_reverse(rev, Q, qo)
        cnt = lambda r: np.where(r[:, 0] <= r[:, 1], r[:, 1].astype(np.int64) - r[:, 0] + 1, 0)
        assert (cnt(a) == cnt(b)).all()
        Qr, _ = po.concat([Q[qo[i]:qo[i + 1]][::-1] for i in range(len(qo) - 1)])
        assert (oracle.fm_match_reverse(fwd, Q, qo) == oracle.fm_match(fwd, Qr, qo)).all()
