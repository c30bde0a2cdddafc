"""CPU: the oracle's MDS byte-code (oracle/oracle_map.c::ops_to_mds, restating nvBowtie/bowtie2/cuda/traceback_inl.h:555-690)
against hand-worked vectors, against a direct definition (walk the alignment, classify every column), and through the
restatement of the reference's consumer (SamOutput::generate_md_string, nvbio/io/output/output_sam.cpp:216-289)."""
import numpy as np
import pytest

import pyoracle as po

M, I, D = 0, 1, 2


def push_order(forward_ops):
    return np.array(forward_ops[::-1], dtype=np.uint8)


def direct_mds(fwd, pattern, text, j0=0, k0=0):
    """definition of the byte-code, written from its description rather than from the reference's loop"""
    out, ed, j, k = [], 0, j0, k0
    i = 0
    while i < len(fwd):
        op = fwd[i]; l = 1
        while i + l < len(fwd) and fwd[i + l] == op:
            l += 1
        i += l
        if op == M:
            run = 0
            for _ in range(l):
                if pattern[j] == text[k]:
                    if run == 255:
                        out += [0, 255]; run = 0
                    run += 1
                else:
                    if run:
                        out += [0, run]; run = 0
                    out += [1, int(pattern[j])]; ed += 1
                j += 1; k += 1
            if run:
                out += [0, run]
        elif op == I:
            out += [2, l & 255] + [int(c) for c in pattern[j:j + l]]; j += l; ed += l
        else:
            out += [3, l & 255] + [int(c) for c in text[k:k + l]]; k += l; ed += l
    n = len(out) + 2
    return [n & 255, n >> 8] + out, ed


@pytest.fixture(scope="module")
def orc():
    return po.Oracle("oracle")


def test_hand_worked_vectors(orc):
    # read ACGTACGT vs text ACGAACGT: one mismatch at position 3 (read symbol T = 3)
    read, text = po.dna("ACGTACGT"), po.dna("ACGAACGT")
    code, ln, ed = orc.ops_to_mds(push_order([M] * 8), 8, read, text)
    assert list(code) == [8, 0, 0, 3, 1, 3, 0, 4] and ln == 8 and ed == 1
    assert po.mds_to_md_string(code)[0] == "3T4"
    # deletion of GG from the text, insertion of T in the read
    read, text = po.dna("ACTACA"), po.dna("ACGGACA")          # AC | GG deleted | T inserted | ACA
    fwd = [M, M, D, D, I, M, M, M]
    code, ln, ed = orc.ops_to_mds(push_order(fwd), len(fwd), read, text)
    assert list(code) == [13, 0, 0, 2, 3, 2, 2, 2, 2, 1, 3, 0, 3]
    want, wed = direct_mds(fwd, read, text)
    assert list(code) == want and ed == wed == 3
    md, mm, gapo, gape = po.mds_to_md_string(code)
    assert md == "2^GG03" and (mm, gapo, gape) == (0, 2, 1)
    # empty alignment
    code, ln, ed = orc.ops_to_mds(np.zeros(0, np.uint8), 0, read, text)
    assert list(code) == [2, 0] and ed == 0
    # alignment that starts inside both strings
    code, ln, ed = orc.ops_to_mds(push_order([M, M, M]), 3, po.dna("TTACG"), po.dna("GGGACG"), j0=2, k0=3)
    assert list(code) == [4, 0, 0, 3] and ed == 0


def test_random_alignments_against_definition(orc):
    rng = np.random.default_rng(11)
    for _ in range(300):
        n = int(rng.integers(1, 400))
        fwd = list(rng.choice([M, M, M, M, M, M, I, D], n))
        L = sum(1 for o in fwd if o != D); N = sum(1 for o in fwd if o != I)
        read = rng.integers(0, 4, L + 3).astype(np.uint8); text = rng.integers(0, 4, N + 5).astype(np.uint8)
        # make most columns match
        j = k = 0
        for o in fwd:
            if o == M:
                if rng.random() < 0.9:
                    text[k] = read[j]
                j += 1; k += 1
            elif o == I:
                j += 1
            else:
                k += 1
        code, ln, ed = orc.ops_to_mds(push_order(fwd), n, read, text, cap=4096)
        want, wed = direct_mds(fwd, read, text)
        assert list(code) == want and ed == wed and ln == len(want)


def test_long_match_runs_split_at_255(orc):
    read = np.zeros(700, np.uint8)
    code, ln, ed = orc.ops_to_mds(push_order([M] * 700), 700, read, read)
    assert list(code) == [8, 0, 0, 255, 0, 255, 0, 190] and ed == 0
    # truncated output keeps the full length in the header and in mds_len
    code, ln, ed = orc.ops_to_mds(push_order([M] * 700), 700, read, read, cap=5)
    assert ln == 8 and list(code) == [8, 0, 0, 255, 0]


@pytest.mark.skipif(not po.Oracle.available("ref"), reason="compiled reference absent")
def test_same_code_in_the_reference_build(orc):
    """oracle_map.c is compiled a second time into oracle/_ref (prefix ref_): the function does not touch the reference's
    headers, so this only checks that both builds export the same behaviour"""
    ref = po.Oracle("ref")
    rng = np.random.default_rng(5)
    fwd = list(rng.choice([M, M, M, I, D], 200))
    read = rng.integers(0, 4, 300).astype(np.uint8); text = rng.integers(0, 4, 300).astype(np.uint8)
    a = orc.ops_to_mds(push_order(fwd), 200, read, text); b = ref.ops_to_mds(push_order(fwd), 200, read, text)
    assert list(a[0]) == list(b[0]) and a[1:] == b[1:]
