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


def test_hit_priority_rule_is_the_references_deque(oracle, ref):
    """oracle_map.c's hit selection against the REFERENCE's own data structure (SeedHit + hit_compare in a priority_deque over a
    vector_view, pushes that drop the bottom of a full deque: compiled from nvBowtie/bowtie2/cuda/seed_hit.h and
    nvbio/basic/priority_deque.h in oracle/ref_wrap.cpp::ref_hit_order): identical visiting order for distinct sizes,
    identical size sequence when sizes tie (the reference leaves the order among equals to its interval heap)."""
    rng = np.random.default_rng(1)
    for it in range(1500):
        n = int(rng.integers(1, 60)); cap = int(rng.choice([0, 0, 1, 3, 8, 16]))
        sizes = rng.permutation(np.arange(1, 400))[:n].astype(np.uint32)
        sizes[rng.random(n) < 0.2] = 0
        take, order = oracle.hit_priority(sizes, 4, cap, 0)
        slots, rsizes = ref.ref_hit_order(sizes, cap)
        assert list(order) == list(slots)
        assert (take[order] == np.minimum(sizes[order], 4)).all() and take.sum() == np.minimum(sizes[order], 4).sum()
        tied = rng.integers(0, 6, n).astype(np.uint32)
        take, order = oracle.hit_priority(tied, 4, cap, 0)
        slots, rsizes = ref.ref_hit_order(tied, cap)
        assert list(tied[order]) == list(rsizes)
    # the row budget is spent in visiting order
    take, order = oracle.hit_priority(np.array([9, 2, 0, 5, 2], np.uint32), 4, 0, 7)
    assert list(order) == [1, 4, 3, 0] and list(take) == [0, 2, 0, 3, 2]
    # a full deque drops its largest range before every push, even for a larger newcomer (mapping_inl.h:158-160)
    take, order = oracle.hit_priority(np.array([1, 2, 3, 10], np.uint32), 4, 3, 0)
    assert list(order) == [0, 1, 3] and list(take) == [1, 2, 0, 4]


def test_oracle_map_modes(oracle, ref):
    """approximate seeds and hit priority in the composition: same results over the C restatement and over the reference's
    headers; approximate seeding finds reads whose every exact seed is broken by a substitution"""
    rng = np.random.default_rng(21)
    genome = rng.integers(0, 4, 30000).astype(np.uint8)
    genome[5000:5400] = genome[9000:9400]
    idx = oracle.build_index(genome)
    sym, off = map_util.make_reads(rng, genome, 200)
    for kw in (dict(seed_mismatches=1, max_seed_ranges=4), dict(hit_priority=1, max_seed_hits=6, max_rows=10),
               dict(seed_mismatches=1, max_seed_ranges=8, hit_priority=1, max_rows=12)):
        P = map_util.map_params(min_score=-40, **kw)
        h1, o1, c1 = oracle.map_reads(idx, genome, sym, off, P)
        h2, o2, c2 = ref.map_reads(idx, genome, sym, off, P)
        assert (h1 == h2).all() and (o1 == o2).all() and (c1 == c2).all(), kw
        assert ((h1["flags"] & 2) != 0).sum() > 120, kw
    # 40-bp reads with a substitution in the LAST half of each of their three 20-mer seeds (the half searched last is the seed's
    # first half; the exact part is its second half): exact seeding finds nothing, approximate seeding maps them
    reads = []
    for i in range(60):
        p = int(rng.integers(100, 29000)); r = genome[p:p + 40].copy()
        r[3] = (r[3] + 1) % 4; r[13] = (r[13] + 2) % 4; r[23] = (r[23] + 3) % 4
        reads.append(r)
    sym2, off2 = po.concat(reads)
    exact = map_util.map_params(min_score=-40, both=0)
    approx = map_util.map_params(min_score=-40, both=0, seed_mismatches=1, max_seed_ranges=8)
    he, _, _ = oracle.map_reads(idx, genome, sym2, off2, exact)
    ha, _, _ = oracle.map_reads(idx, genome, sym2, off2, approx)
    assert ((he["flags"] & 2) != 0).sum() == 0
    assert ((ha["flags"] & 2) != 0).sum() >= 58


def test_rescue_mates_composition(oracle, ref):
    """the opposite-mate window search over the C restatement and over the reference's headers agree; a mate no seed can find is
    recovered next to its anchor, on the strand and side the FR policy demands"""
    rng = np.random.default_rng(9)
    G = 20000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    idx = oracle.build_index(genome)
    L, reads, truth = 80, [], []
    for i in range(60):
        F = int(rng.integers(150, 400)); p = int(rng.integers(0, G - F))
        m1 = genome[p:p + L].copy(); m2 = (3 - genome[p + F - L:p + F])[::-1].copy()
        victim = m1 if i % 4 == 3 else m2                # i % 4 == 3: the forward mate is broken, the anchor lies on the reverse strand
        for k in range(4, L, 9):
            victim[k] = (victim[k] + 1) % 4               # no exact 20-mer seed survives
        reads += [m1, m2]; truth.append((p, F))
    sym, off = po.concat(reads)
    P = map_util.map_params(min_score=-60)
    h, o, _ = oracle.map_reads(idx, genome, sym, off, P)
    a = po.aligner(po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3))
    h1, o1 = h.copy(), o.copy(); h2, o2 = h.copy(), o.copy()
    r1 = oracle.rescue_mates(genome, sym, off, a, -60, 500, h1, o1)
    r2 = ref.rescue_mates(genome, sym, off, a, -60, 500, h2, o2)
    assert r1 == r2 and (h1 == h2).all() and (o1 == o2).all()
    assert r1[0] == 60 and r1[1] >= 58
    for i, (p, F) in enumerate(truth):
        broken = 2 * i + (0 if i % 4 == 3 else 1)
        if not h1["flags"][broken] & 4:
            continue
        if i % 4 == 3:                                    # left of the reverse-strand anchor, on the forward strand
            assert not (h1["flags"][broken] & 1) and abs(int(h1["begin"][broken]) - p) <= 3
        else:                                             # right of the forward-strand anchor, reverse-complemented
            assert h1["flags"][broken] & 1 and abs(int(h1["end"][broken]) - (p + F)) <= 3
        assert h1["score"][broken] >= -60
    pairs = oracle.pair_hits(h1, 0, 500)
    assert ((pairs["flags"] & 4) != 0).sum() >= 58


def test_mapq_v2_is_the_references_calculator(oracle, ref):
    """oracle_map.c::mapq_v2 (tables) against the compiled BowtieMapq2 (nvBowtie/bowtie2/cuda/mapq.h:145-316, instantiated in
    oracle/ref_wrap.cpp::ref_bowtie_mapq2 over a scheme that returns the given perfect / minimum scores): end-to-end and local
    scoring, with and without a runner-up, below the minimum score, pairs (summed scores and ranges, no runner-up)"""
    rng = np.random.default_rng(3)
    seen = set()
    for it in range(40000):
        L = int(rng.integers(30, 300)); mono = bool(rng.integers(0, 2))
        perfect = 0 if mono else 2 * L
        minsc = int(-0.6 - 0.6 * L) if mono else int(20 + 8 * np.log(L))
        best = int(rng.integers(minsc - 5, perfect + 1))
        hs = bool(rng.integers(0, 2)); second = int(rng.integers(minsc - 8, max(best, minsc - 7) + 1))
        a = oracle.mapq_v2(best, hs, second, perfect, minsc, mono)
        assert a == ref.ref_bowtie_mapq2(best, hs, second, perfect, minsc, mono), (best, hs, second, perfect, minsc, mono)
        seen.add(a)
    assert len(seen) >= 40                                    # the tables are exercised broadly
    for it in range(5000):
        best, ob = int(rng.integers(-60, 1)), int(rng.integers(-61, 1))
        assert oracle.mapq_v2(best + ob, False, 0, 0, -60 - 61, True) == ref.ref_bowtie_mapq2(best, False, 0, 0, -60, True, True, ob, 0, -61)
    # every boundary of the score range
    for best in range(-61, 1):
        for second in range(-61, best + 1):
            assert oracle.mapq_v2(best, True, second, 0, -60, True) == ref.ref_bowtie_mapq2(best, True, second, 0, -60, True)


def test_oracle_map_reverse_index_half(oracle, ref):
    """approximate seeds with the reverse index: reads whose every seed carries a substitution in its LAST half map only when the
    index of the reversed text takes part (the forward half covers substitutions in the first half of a seed)"""
    rng = np.random.default_rng(31)
    G = 30000
    genome = rng.integers(0, 4, G).astype(np.uint8)
    idx = oracle.build_index(genome)
    ridx = oracle.build_index(genome[::-1].copy())
    reads = []
    for i in range(80):
        p = int(rng.integers(100, G - 200)); r = genome[p:p + 40].copy()
        for k in (16, 26, 36):                                # seeds start at 0, 10, 20: positions 16 / 26 / 36 sit in their last halves
            r[k] = (r[k] + 1 + i % 3) % 4
        reads.append(r)
    sym, off = po.concat(reads)
    fwd_only = map_util.map_params(min_score=-40, both=0, seed_mismatches=1, max_seed_ranges=8)
    both = map_util.map_params(min_score=-40, both=0, seed_mismatches=1, max_seed_ranges=8, reverse_index=ridx)
    hf, _, _ = oracle.map_reads(idx, genome, sym, off, fwd_only)
    hb, ob, cb = oracle.map_reads(idx, genome, sym, off, both)
    assert ((hf["flags"] & 2) != 0).sum() == 0
    assert ((hb["flags"] & 2) != 0).sum() >= 78
    hr, orr, cr = ref.map_reads(idx, genome, sym, off, both)
    assert (hb == hr).all() and (ob == orr).all() and (cb == cr).all()
    # general reads: the union of both halves finds at least what the forward half finds
    sym2, off2 = map_util.make_reads(rng, genome, 300, sub=0.05)
    p1 = map_util.map_params(min_score=-60, seed_mismatches=1, max_seed_ranges=4)
    p2 = map_util.map_params(min_score=-60, seed_mismatches=1, max_seed_ranges=4, reverse_index=ridx)
    h1, _, c1 = oracle.map_reads(idx, genome, sym2, off2, p1)
    h2, _, c2 = oracle.map_reads(idx, genome, sym2, off2, p2)
    assert (c2 >= c1).all() and ((h2["flags"] & 2) != 0).sum() >= ((h1["flags"] & 2) != 0).sum()
    assert (h2["score"][(h1["flags"] & 2) != 0] >= h1["score"][(h1["flags"] & 2) != 0]).all()
