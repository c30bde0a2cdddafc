"""The reference's own golden vectors (nvbio-test/alignment_test.cu:637-864) against the C restatement
(and against the compiled reference when oracle/_ref is present)."""
import numpy as np
import pytest
import pyoracle as po

STR144 = "TAGGAGGTAACATGTATGGAGCATTTACCATAGGCCAAGCACTGTTCTAAGAACTTCGGACATGTTATCTCACTTGTATAAGTACTTAGGTGCCTACAACATAAGCAGCACCTGGTAAATTAAGTATTGAAAAAATGCAGATCG"
REF500 = ("CAGCACTGACCGGTGAGCATAAACCCTGGGGATGCCCAGAGCTGGTACAGCCAGGAGCTCCAGAAGCGTGGGATTCTCAGAGGGAAGTGGAGCTCACTGCTCTACAGGTCCTATTCAAGTTAGAAAGTAAGATACAATGCACACAAAGCCAAATTGTC"
          "ATCATTCAGCTCCTATTACAGGGGAACTAAGAGCTGCATTGAAAATTATTTGCAAAGCTTGTAAGTGGTTCTGCCACTTATTAGCCGTGTGAACCTTAGCAAATTACCTAGCGTCTCTGAGTTTCAACTTCCTCATCTACAAAATAGAAATGATAATAAT"
          "AACCGCATCGCAAGAGTTGTTGGAAAAATGAAAATGAGGTATCATAGGAGGTAACATGTATGGAGCATTTACCATAGGCCAAGCACTGTTCTAAGAACTTCGGACATGTTATCTCACTTGTATAAGTACTTAGGTGCCTACAACATAAACAGCACCTGGT"
          "AAATTAAGTATTGAAAAAATGC")
STR150 = "TTATGTAGGTGGTCTGGTTTTTGCCTTTTAAGCTTCTGCAAAAAACAACAACAAACTTGTGGTATTACACTGACTCTACAGATCAATTTGGGGACAACTTCCATGTGTTCCACCACCAATACTGAATCTTTCAATCGACTGACGTGGTAT"
REF181 = "ATCGGATTCTTTCTTACTTGTAGGTGGTCTGGTTTTTGCCTTTTAAGCTTCTGCAAAAAACAACAACAAACTTGTGGTATTACACTGACTCTACAGATCAATTTGGGGACAACTTCCATGTGTTCCACCACCAATACTGAATCTTTCAATCGACTGACGTGGTATCTCTCTCTCCATCTAT"

# (pattern, text, expected score) -- alignment_test.cu:641-707, banded_alignment_score<5>, ED SEMI_GLOBAL
BANDED_ED = [
    ("GGGTGCTCAA", "AAAAGGGTGCTCAA", 0), ("GGGTAAGCTC", "AAAAGGGTGCTCAA", -2), ("AAGGGTGCTC", "AAAAGGGTGCAATC", -2),
    ("AAAAGGGTGC", "AAAAGGGTGCTCAA", 0), ("AAAAGGGTG", "AAAAGGAAGTGCTC", -2), ("CACCGGGT", "AACAGGGTGCTC", -2),
]
# (kind, type, scheme, pattern, text, band(0=full), expected CIGAR in push order) -- :721-753,:770-785,:802-863
CIGARS = [
    (po.SW, po.GLOBAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 0, "1M2D3M1D3M10D"),
    (po.SW, po.LOCAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 0, "4M1D3M"),
    (po.SW, po.SEMI_GLOBAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 0, "4M1D3M"),
    (po.GOTOH, po.GLOBAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 0, "1M2D3M1D3M10D"),
    (po.GOTOH, po.LOCAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 0, "4M1D3M"),
    (po.GOTOH, po.SEMI_GLOBAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 0, "4M1D3M"),
    (po.GOTOH, po.SEMI_GLOBAL, (2, -1, -1, -1), "ACAACTA", "AAACACCCTAACACACTAAA", 7, "4M1D3M"),
    (po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3), STR150, REF181, 31, "147M2D3M"),
    (po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3), STR144, REF500, 0, "6I138M"),
    (po.ED, po.SEMI_GLOBAL, (0, -1, -1, -1), STR144, REF500, 0, "1I1M2I1M3I136M"),
]


def _impls(request):
    out = [request.getfixturevalue("oracle")]
    if po.Oracle.available("ref"):
        out.append(po.Oracle("ref"))
    return out


@pytest.mark.parametrize("case", BANDED_ED)
def test_banded_edit_distance_scores(request, case):
    p, t, expected = case
    P, Po = po.concat([po.dna(p)])
    T, To = po.concat([po.dna(t)])
    for impl in _impls(request):
        s = impl.banded_score(po.aligner(po.ED, po.SEMI_GLOBAL), 5, P, Po, T, To)
        assert int(s["score"][0]) == expected, impl.which


def rescore(kind, scheme, ops_fwd, p, t, src_x):
    """score of an op string, as TestBacktracker::score does (alignment_test_utils.h:636-781)"""
    if kind == po.ED:
        scheme = (0, -1, -1, -1)
    m, mm, a, b = scheme
    score, j, k, prev = 0, 0, src_x, None
    for op in ops_fwd:
        if op == 0:
            score += m if p[j] == t[k] else mm
            j += 1
            k += 1
        elif op == 1:  # insertion: consumes pattern
            score += (b if kind != po.GOTOH else (b if prev == 1 else a))
            j += 1
        else:          # deletion: consumes text
            score += (a if kind != po.GOTOH else (b if prev == 2 else a))
            k += 1
        prev = op
    return score


@pytest.mark.parametrize("case", CIGARS, ids=lambda c: "%d-%d-%d-%s" % (c[0], c[1], c[5], c[6]))
def test_golden_cigars(request, case):
    kind, ty, scheme, p, t, band, cigar = case
    pa, ta = po.dna(p), po.dna(t)
    P, Po = po.concat([pa])
    T, To = po.concat([ta])
    a = po.aligner(kind, ty, scheme)
    for impl in _impls(request):
        if band:
            sink = impl.banded_score(a, band, P, Po, T, To)[0]
            aln, ops, n = impl.banded_traceback(a, band, 32, P, Po, T, To)
        else:
            sink = impl.score(a, P, Po, T, To)[0]
            aln, ops, n = impl.traceback(a, 32, P, Po, T, To)
        assert po.rle(ops[0, :n[0]]) == cigar, impl.which
        assert aln["score"][0] == sink["score"] and aln["sink_x"][0] == sink["x"] and aln["sink_y"][0] == sink["y"]
        # re-scoring the alignment reproduces the score (the reference test's own check); in banded
        # mode insertion/deletion labels keep the same meaning for the op stream
        fwd = ops[0, :n[0]][::-1]
        got = rescore(kind, scheme, fwd, pa, ta, int(aln["src_x"][0]))
        assert got == int(aln["score"][0]), (impl.which, got, aln)


def test_checkpoint_interval_does_not_change_result(oracle):
    a = po.aligner(po.GOTOH, po.SEMI_GLOBAL, (0, -5, -8, -3))
    P, Po = po.concat([po.dna(STR144)])
    T, To = po.concat([po.dna(REF500)])
    base = oracle.traceback(a, 16, P, Po, T, To)
    for chk in (32, 64):
        o = oracle.traceback(a, chk, P, Po, T, To)
        assert (o[0] == base[0]).all() and (o[2] == base[2]).all()
        assert (o[1][0, :o[2][0]] == base[1][0, :base[2][0]]).all()
