"""-m gpu: nvb_ops_to_mds / nvb_map_hits_to_mds (nvBowtie's MDS byte-code and edit distance, traceback_inl.h:555-690)
against the oracle's literal restatement, alignment by alignment, byte for byte."""
import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
import pyoracle as po
import gpu_util as gu
import map_util
import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,ty,scores,band", [(po.GOTOH, nvb.SEMI_GLOBAL, (0, -5, -8, -3), 15), (po.ED, nvb.GLOBAL, (0, -1, -1, -1), 15),
                                                 (po.SW, nvb.LOCAL, (2, -1, -2, -2), 0)])
@pytest.mark.parametrize("bits", [8, 4])
def test_ops_to_mds(oracle, kind, ty, scores, band, bits):
    rng = np.random.default_rng(band + ty + bits)
    n = 400
    pat, pat_off, txt, txt_off = util.rand_pairs(rng, n, (20, 200), util.banded_text_len(rng, ty, band or 15), related=0.9)
    a = nvb.Aligner(kind, ty, *scores)
    ps, ts = gu.sset(pat, pat_off), gu.sset(txt, txt_off)
    if bits == 4:
        ps = nvb.StringSet(torch.from_numpy(nvb.pack_symbols(pat, 4).view(np.int32)).to(gu.dev()), ps.offsets, 4, ps.max_len)
        ts = nvb.StringSet(torch.from_numpy(nvb.pack_symbols(txt, 4).view(np.int32)).to(gu.dev()), ts.offsets, 4, ts.max_len)
    if band:
        out, ops, n_ops = nvb.batch_banded_alignment_traceback(a, band, 16, ps, ts)
    else:
        out, ops, n_ops = nvb.batch_alignment_traceback(a, 32, ps, ts)
    stride = 2 + 3 * ops.shape[1]
    mds, mds_len, ed = nvb.ops_to_mds(ops, n_ops, ps, ts, out, mds_stride=stride)
    oh, opsh, nh = gu.tb_np((out, ops, n_ops))
    mds, mds_len, ed = mds.cpu().numpy(), mds_len.cpu().numpy(), ed.cpu().numpy()
    some_indel = 0
    for i in range(n):
        P, T = pat[pat_off[i]:pat_off[i + 1]], txt[txt_off[i]:txt_off[i + 1]]
        want, wl, we = oracle.ops_to_mds(opsh[i], int(nh[i]), P, T, j0=int(oh["src_y"][i]), k0=int(oh["src_x"][i]), cap=stride)
        assert wl == mds_len[i] and we == ed[i], i
        assert (mds[i, :wl] == want).all(), i
        some_indel += int((opsh[i, :nh[i]] != 0).any())
        # the reference's consumer and the package's formatter agree on the text (runs below 256)
        assert nvb.mds_to_md_string(mds[i])[0] == po.mds_to_md_string(want)[0] or wl > 255
    assert some_indel > 20
    # a stride shorter than the code: truncated bytes, full length reported
    short, slen, _ = nvb.ops_to_mds(ops, n_ops, ps, ts, out, mds_stride=6)
    assert (slen.cpu().numpy() == mds_len).all() and (short.cpu().numpy()[:, 2:6] == mds[:, 2:6]).all()


def test_map_hits_to_mds(oracle):
    rng = np.random.default_rng(77)
    genome = rng.integers(0, 4, 40000).astype(np.uint8)
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 2000)
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, 0, -5, -8, -3), 20, 10, 4, 0, 15, True, -60)
    rs = gu.sset(sym, off)
    hits, ops, _ = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params)
    mds, mds_len, ed = nvb.map_hits_to_mds(hits, ops, rs.symbols, rs.offsets, words)
    h = hits.cpu().numpy().view(np.uint32).reshape(-1).view(nvb.HIT_DT)
    opsh, mds, mds_len, ed = ops.cpu().numpy(), mds.cpu().numpy(), mds_len.cpu().numpy(), ed.cpu().numpy()
    aligned = 0
    for r in range(len(h)):
        read = sym[off[r]:off[r + 1]]
        if not (h["flags"][r] & nvb.FLAG_ALIGNED):
            assert list(mds[r, :2]) == [2, 0] and mds_len[r] == 2 and ed[r] == 0
            continue
        aligned += 1
        if h["flags"][r] & nvb.FLAG_REVERSE:
            read = np.where(read < 4, 3 - read, read)[::-1]
        want, wl, we = oracle.ops_to_mds(opsh[r], int(h["n_ops"][r]), read, genome[int(h["begin"][r]):], cap=mds.shape[1])
        assert wl == mds_len[r] and we == ed[r], r
        assert (mds[r, :wl] == want).all(), r
        # edit distance consistent with the alignment's score under (0, -5, -8, -3): no gaps and k mismatches -> -5 k
        if (opsh[r, :h["n_ops"][r]] == 0).all():
            assert h["score"][r] == -5 * int(ed[r])
    assert aligned > 1500


@pytest.mark.parametrize("monotone", [True, False])
def test_map_hits_mapq(oracle, monotone):
    """nvb_map_hits_mapq (BowtieMapq2, mapq.h:145-316) vs the oracle restatement (itself pinned on the compiled calculator), on the
    hits of the pipeline, unpaired and with the pair records"""
    rng = np.random.default_rng(78 + monotone)
    genome = rng.integers(0, 4, 40000).astype(np.uint8)
    genome[5000:5600] = genome[20000:20600]                   # repeats: runners-up
    genome[9000:9300] = genome[20100:20400]
    fmi, oidx = gu.fm_index_pair(oracle, genome)
    words = nvb.pack_genome(torch.from_numpy(genome).to(gu.dev()))
    sym, off = map_util.make_reads(rng, genome, 3000, sub=0.03)
    scores = (0, -5, -8, -3) if monotone else (2, -5, -8, -3)
    params = nvb.MapParams(nvb.GotohAligner(nvb.SEMI_GLOBAL, *scores), 20, 10, 4, 0, 15, True, -60 if monotone else 30)
    rs = gu.sset(sym, off)
    hits, ops, _ = nvb.map_reads(fmi, words, rs.symbols, rs.offsets, params)
    max_len = 160
    lens = np.arange(max_len + 1)
    tab = (np.float32(-0.6) + np.float32(-0.6) * lens.astype(np.float32)).astype(np.int32) if monotone else \
          (np.float32(20) + np.float32(8) * np.log(np.maximum(lens, 1).astype(np.float32))).astype(np.int32)
    bonus = scores[0]
    h = hits.cpu().numpy().view(np.uint32).reshape(-1).view(nvb.HIT_DT)
    L = np.minimum(np.diff(off.astype(np.int64)), max_len)
    got = nvb.map_hits_mapq(hits, rs.offsets, tab, bonus, monotone).cpu().numpy()
    with_second = 0
    for r in range(len(h)):
        if not h["flags"][r] & 2:
            assert got[r] == 0
            continue
        hs = h["second_score"][r] != -(1 << 31)
        with_second += int(hs)
        assert got[r] == oracle.mapq_v2(h["score"][r], hs, h["second_score"][r], int(L[r]) * bonus, tab[L[r]], monotone), r
    assert with_second > 20 and len(set(got.tolist())) > 6
    # paired: reads 2 i, 2 i + 1 taken as mates (whatever their geometry, the rule only looks at the concordant flag)
    n2 = len(h) // 2 * 2
    pairs = nvb.pair_hits(hits[:n2], 0, 100000)
    pr = pairs.cpu().numpy().view(np.uint32).reshape(-1).view(nvb.PAIR_DT)
    gotp = nvb.map_hits_mapq(hits[:n2], rs.offsets[:n2 + 1], tab, bonus, monotone, pairs=pairs).cpu().numpy()
    conc = 0
    for r in range(n2):
        if not h["flags"][r] & 2:
            assert gotp[r] == 0
            continue
        if pr["flags"][r >> 1] & nvb.PAIR_CONCORDANT:
            o = r ^ 1; conc += 1
            want = oracle.mapq_v2(int(h["score"][r]) + int(h["score"][o]), False, 0, (int(L[r]) + int(L[o])) * bonus, int(tab[L[r]]) + int(tab[L[o]]), monotone)
        else:
            want = got[r]
        assert gotp[r] == want, r
    assert conc > 10
