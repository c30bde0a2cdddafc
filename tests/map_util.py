"""seeded read sets for the seed-and-extend pipeline tests (SURVEY 8d C4 shape: reads sampled from the genome
with substitutions and indels, both strands; plus junk reads, reads with Ns, short reads, reads at the genome ends)"""
import numpy as np
import pyoracle as po


def make_reads(rng, genome, n, len_range=(60, 151), sub=0.02, indel=0.3, alpha=4):
    G = len(genome)
    reads = []
    for i in range(n):
        L = int(rng.integers(len_range[0], len_range[1]))
        kind = i % 16
        if kind == 15:                                   # unrelated read
            r = rng.integers(0, alpha, L).astype(np.uint8)
        else:
            if kind == 14:                               # hugging one end of the genome
                p = 0 if (i // 16) % 2 == 0 else G - L
            else:
                p = int(rng.integers(0, G - L))
            r = genome[p:p + L].copy()
            m = rng.random(L) < sub
            r[m] = (r[m] + rng.integers(1, 4, int(m.sum()))) % alpha
            if rng.random() < indel and L > 40:
                k = int(rng.integers(10, L - 10))
                if rng.random() < 0.5:
                    r = np.delete(r, k)
                else:
                    r = np.insert(r, k, rng.integers(0, alpha))
            if kind == 13:                               # an N in the read
                r[int(rng.integers(0, len(r)))] = 4
        if i % 2:
            r = np.where(r < 4, 3 - r, r)[::-1]
        reads.append(np.ascontiguousarray(r, dtype=np.uint8))
    if n > 3:
        reads[3] = reads[3][:12]                         # shorter than a seed
    return po.concat(reads)


def map_params(band=15, kind=po.GOTOH, scores=(0, -5, -8, -3), seed_len=20, interval=10, max_hits=4, max_range=0,
               both=1, min_score=-60, seed_mismatches=0, max_seed_ranges=0, hit_priority=0, max_seed_hits=0, max_rows=0,
               reverse_index=None):
    """reverse_index: a pyoracle.Index of the reversed genome (its descriptor is kept alive on the returned struct)"""
    import ctypes as C
    rc = reverse_index.c() if reverse_index is not None else None
    P = po.MapParamsC(seed_len, interval, max_hits, max_range, band, both, min_score, po.aligner(kind, po.SEMI_GLOBAL, scores),
                      seed_mismatches, max_seed_ranges, hit_priority, max_seed_hits, max_rows, C.addressof(rc) if rc is not None else None)
    P._rc = rc
    return P
