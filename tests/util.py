"""seeded input generators shared by the oracle tests, the GPU parity tests and the golden-fixture script"""
import numpy as np
import pyoracle as po

SCHEMES = {
    po.ED: [(0, -1, -1, -1)],
    po.SW: [(2, -1, -1, -1), (1, -1, -1, -1), (2, -3, -2, -4)],
    po.GOTOH: [(2, -1, -1, -1), (0, -5, -8, -3), (1, -1, -2, -1)],
}
KIND_NAME = {po.ED: "ed", po.SW: "sw", po.GOTOH: "gotoh"}
TYPE_NAME = {0: "global", 1: "local", 2: "semi"}


def mutate(rng, p, alpha, max_edits=4):
    pp = p.copy()
    for _ in range(int(rng.integers(0, max_edits))):
        if len(pp) > 2:
            j = int(rng.integers(0, len(pp)))
            r = rng.random()
            if r < 0.4:
                pp[j] = rng.integers(0, alpha)
            elif r < 0.7:
                pp = np.delete(pp, j)
            else:
                pp = np.insert(pp, j, rng.integers(0, alpha))
    return pp


def rand_pairs(rng, n, m_range, n_of_m, alpha=4, related=0.7):
    """n (pattern, text) pairs; text length = n_of_m(M); a mutated copy of the pattern is implanted
    in most texts so that alignments are not all noise; alpha=2 gives many score ties."""
    P, T = [], []
    for _ in range(n):
        M = int(rng.integers(*m_range))
        p = rng.integers(0, alpha, M).astype(np.uint8)
        N = int(n_of_m(M))
        t = rng.integers(0, alpha, N).astype(np.uint8)
        if N >= M and rng.random() < related:
            s = int(rng.integers(0, N - M + 1))
            pp = mutate(rng, p, alpha)
            e = min(N, s + len(pp))
            t[s:e] = pp[:e - s]
        P.append(p)
        T.append(t)
    return po.concat(P) + po.concat(T)


def banded_text_len(rng, ty, band):
    """text lengths for which the reference's banded code is well defined:
    SEMI_GLOBAL only needs N >= M (shorter = refused); GLOBAL/LOCAL need N >= M + band - 1."""
    if ty == po.SEMI_GLOBAL:
        return lambda M: M + int(rng.integers(0, band + 6)) - (1 if rng.random() < 0.05 else 0)
    return lambda M: M + band - 1 + int(rng.integers(0, 8))


def rand_queries(rng, text, n, len_range=(0, 25), alpha=4, p_from_text=0.7, p_n=0.05):
    qs = []
    N = len(text)
    for _ in range(n):
        L = int(rng.integers(*len_range))
        if N > L and rng.random() < p_from_text:
            s = int(rng.integers(0, N - L + 1))
            q = text[s:s + L].copy()
        else:
            q = rng.integers(0, alpha, L).astype(np.uint8)
        if L and rng.random() < p_n:
            q[rng.integers(0, L)] = 4
        qs.append(q)
    return po.concat(qs)


def ops_equal(a, b):
    (o1, ops1, n1), (o2, ops2, n2) = a, b
    if not ((o1 == o2).all() and (n1 == n2).all()):
        return False
    for i in range(len(n1)):
        if not (ops1[i, :n1[i]] == ops2[i, :n2[i]]).all():
            return False
    return True
