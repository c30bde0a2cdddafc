"""Host-side mirror of the seed -> locate -> extend -> reduce -> traceback call (``nvb_map_reads``).

The steps are the ones nvBowtie's driver strings together (nvBowtie/bowtie2/cuda/mapping_inl.h:73-86 seed
matching, locate_inl.h:115-200, score_inl.h:85-130 window placement and banded scoring, reduce_inl.h,
traceback_inl.h:224-277), without its seed priorities, re-seeding, effort limits, MAPQ and mate pairing.
All compute runs in libnvbio_b200.so; there is no CPU implementation here.
"""
import ctypes as C

import numpy as np
import torch

from ._lib import lib, check, Context, MapParamsC, MapStatsC, StringSetC, current_stream_ptr
from .alignment import Aligner, GotohAligner, SEMI_GLOBAL

HIT_DT = np.dtype([("score", "<i4"), ("second_score", "<i4"), ("begin", "<u4"), ("end", "<u4"), ("flags", "<u4"), ("n_ops", "<u4")])
FLAG_REVERSE, FLAG_ALIGNED = 1, 2
PAIR_DT = np.dtype([("score", "<i4"), ("frag_begin", "<u4"), ("frag_end", "<u4"), ("flags", "<u4")])
PAIR_MATE1, PAIR_MATE2, PAIR_CONCORDANT = 1, 2, 4


class MapParams:
    """seed_len / seed_interval / max_hits follow nvBowtie's ``ParamsPOD`` (params.h:124-163); the window is
    ``[g - band/2, g - band/2 + band + read_len)`` (score_inl.h:103-106)"""

    def __init__(self, aligner=None, seed_len=20, seed_interval=10, max_hits=4, max_range=0, band=15,
                 both_strands=True, min_score=-(1 << 29), seed_mismatches=0, max_seed_ranges=0, hit_priority=False,
                 max_seed_hits=0, max_rows=0, reverse_index=None):
        """seed_mismatches = 1: nvBowtie's approximate seeds on the forward index (mapping_inl.h:114-184, :395-397), the first
        max_seed_ranges ranges per seed (0 = 8). hit_priority: ranges visited by size, smallest first, through a deque of
        max_seed_hits entries (0 = unbounded), max_rows rows per read (0 = no budget) (seed_hit.h:220-230, select_inl.h:93-113).
        reverse_index: the FMIndex of the reversed genome; with seed_mismatches = 1 it adds the other half of nvBowtie's approximate
        seeds (substitution in the last half of the seed, mapping_inl.h:399-402)"""
        self.aligner = aligner if aligner is not None else GotohAligner(SEMI_GLOBAL, 0, -5, -8, -3)
        assert isinstance(self.aligner, Aligner)
        self.seed_len, self.seed_interval, self.max_hits, self.max_range = seed_len, seed_interval, max_hits, max_range
        self.band, self.both_strands, self.min_score = band, both_strands, min_score
        self.seed_mismatches, self.max_seed_ranges = int(seed_mismatches), int(max_seed_ranges)
        self.hit_priority, self.max_seed_hits, self.max_rows = bool(hit_priority), int(max_seed_hits), int(max_rows)
        self.reverse_index = reverse_index

    def c(self):
        return MapParamsC(self.seed_len, self.seed_interval, self.max_hits, self.max_range, self.band,
                          1 if self.both_strands else 0, self.min_score, self.aligner.c(), self.seed_mismatches, self.max_seed_ranges,
                          1 if self.hit_priority else 0, self.max_seed_hits, self.max_rows, self._reverse_ptr())

    def _reverse_ptr(self):
        if self.reverse_index is None:
            return None
        self._rc = self.reverse_index.c()           # the descriptor must outlive the call
        return C.addressof(self._rc)


def ops_stride_for(max_read_len, band):
    return (2 * max_read_len + band + 16 + 15) // 16 * 16


def pack_genome(symbols, ctx=None):
    """uint8 device tensor of symbols 0..3 -> 2-bit big-endian words (int32 tensor, 16 symbols per word)"""
    if not symbols.is_cuda:
        raise RuntimeError("pack_genome needs a CUDA tensor; the hot path has no CPU fallback")
    ctx = ctx or Context.default(symbols.device.index)
    n = symbols.numel()
    words = torch.empty((n + 15) // 16, dtype=torch.int32, device=symbols.device)
    check(lib().nvb_pack_2bit(ctx.h, symbols.data_ptr(), n, words.data_ptr(), current_stream_ptr()), ctx.h)
    return words


def map_reads(fmi, genome_words, symbols, offsets, params, max_read_len=None, want_ops=True, reads_per_batch=0, ctx=None):
    """device tensors in, device tensors out: (hits int32[n, 6] viewable as HIT_DT, ops uint8[n, stride] or None, stats dict)"""
    if not symbols.is_cuda:
        raise RuntimeError("map_reads needs CUDA tensors; the hot path has no CPU fallback")
    dev = symbols.device
    ctx = ctx or Context.default(dev.index)
    n = offsets.numel() - 1
    if max_read_len is None:
        max_read_len = int((offsets[1:] - offsets[:-1]).max().item()) if n else 0
    stride = ops_stride_for(max_read_len, params.band)
    hits = torch.empty((n, 6), dtype=torch.int32, device=dev)
    ops = torch.zeros((n, stride), dtype=torch.uint8, device=dev) if want_ops else None
    stats = MapStatsC()
    fc, pc = fmi.c(), params.c()
    rs = StringSetC(symbols.data_ptr(), offsets.data_ptr(), n, 8)
    check(lib().nvb_map_reads(ctx.h, C.byref(fc), genome_words.data_ptr(), rs, max_read_len, C.byref(pc), reads_per_batch,
                              hits.data_ptr(), ops.data_ptr() if want_ops else None, stride, C.byref(stats),
                              current_stream_ptr()), ctx.h)
    return hits, ops, dict(seeds=stats.seeds, seed_hits=stats.seed_hits, candidates=stats.candidates, aligned=stats.aligned)


def map_reads_host(fmi, genome_words, symbols_h, offsets_h, params, hits_h=None, ops_h=None, cigar_h=None, reads_per_batch=0,
                   max_read_len=None, ctx=None):
    """numpy (ideally pinned) buffers in and out through ``nvb_map_reads_host``: batches pipelined over PCIe, synchronous.
    ``cigar_h`` (uint16 [n, stride]) receives io::Cigar entries (op | len << 2); their count is ``flags >> 8``."""
    ctx = ctx or Context.default(genome_words.device.index)
    n = len(offsets_h) - 1
    if max_read_len is None:
        max_read_len = int(np.max(np.diff(offsets_h.astype(np.int64)))) if n else 0
    stride = ops_h.shape[1] if ops_h is not None else ops_stride_for(max_read_len, params.band)
    if hits_h is None:
        hits_h = np.zeros(n, dtype=HIT_DT)
    stats = MapStatsC()
    fc, pc = fmi.c(), params.c()
    check(lib().nvb_map_reads_host(ctx.h, C.byref(fc), genome_words.data_ptr(), symbols_h.ctypes.data, offsets_h.ctypes.data, n,
                                   C.byref(pc), reads_per_batch, hits_h.ctypes.data,
                                   ops_h.ctypes.data if ops_h is not None else None, stride,
                                   cigar_h.ctypes.data if cigar_h is not None else None,
                                   cigar_h.shape[1] if cigar_h is not None else 0, C.byref(stats)), ctx.h)
    return hits_h, ops_h, dict(seeds=stats.seeds, seed_hits=stats.seed_hits, candidates=stats.candidates, aligned=stats.aligned)


def _ctx_array(ctxs):
    arr = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    return arr


def map_reads_host_multi(ctxs, fmis, genome_words, symbols_h, offsets_h, params, hits_h=None, ops_h=None, cigar_h=None,
                         reads_per_batch=0, max_read_len=None):
    """``nvb_map_reads_host_multi``: one host thread per context, contiguous slices of one host batch, every slice writing its own
    range of the result arrays. ctxs / fmis / genome_words: one per context (index replicas on the contexts' devices)."""
    from ._lib import FMIndexC
    n = len(offsets_h) - 1
    if max_read_len is None:
        max_read_len = int(np.max(np.diff(offsets_h.astype(np.int64)))) if n else 0
    stride = ops_h.shape[1] if ops_h is not None else ops_stride_for(max_read_len, params.band)
    if hits_h is None:
        hits_h = np.zeros(n, dtype=HIT_DT)
    k = len(ctxs)
    fcs = (FMIndexC * k)(*[f.c() for f in fmis])
    gws = (C.c_void_p * k)(*[g.data_ptr() for g in genome_words])
    stats = MapStatsC()
    pc = params.c()
    check(lib().nvb_map_reads_host_multi(_ctx_array(ctxs), fcs, gws, k, symbols_h.ctypes.data, offsets_h.ctypes.data, n, C.byref(pc),
                                         reads_per_batch, hits_h.ctypes.data, ops_h.ctypes.data if ops_h is not None else None, stride,
                                         cigar_h.ctypes.data if cigar_h is not None else None,
                                         cigar_h.shape[1] if cigar_h is not None else 0, C.byref(stats)), ctxs[0].h)
    return hits_h, ops_h, dict(seeds=stats.seeds, seed_hits=stats.seed_hits, candidates=stats.candidates, aligned=stats.aligned)


def match_host_multi(ctxs, fmis, symbols_h, offsets_h=None, query_len=0, count=None, ranges_h=None, bits=8):
    """``nvb_fm_match_host_multi``: the query batch cut into one contiguous slice per context (offsets_h = None: queries of
    query_len symbols each)"""
    from ._lib import FMIndexC
    k = len(ctxs)
    n = (len(offsets_h) - 1) if offsets_h is not None else int(count)
    if ranges_h is None:
        ranges_h = np.empty((n, 2), dtype=np.uint32)
    fcs = (FMIndexC * k)(*[f.c() for f in fmis])
    check(lib().nvb_fm_match_host_multi(_ctx_array(ctxs), fcs, k, symbols_h.ctypes.data, bits,
                                        offsets_h.ctypes.data if offsets_h is not None else None, int(query_len), n, ranges_h.ctypes.data), ctxs[0].h)
    return ranges_h


def rescue_mates(hits, ops, symbols, offsets, genome_words, genome_len, aligner, min_score, max_frag_len=500, max_read_len=None, ctx=None):
    """opposite-mate window search for pairs with exactly one aligned mate (``nvb_rescue_mates``; nvBowtie's
    BestOppositeScoreStream, score_inl.h:318-411). hits (and ops, may be None) of the 2 n mates are updated in place.
    Returns (n_candidates, n_rescued)."""
    if not hits.is_cuda:
        raise RuntimeError("rescue_mates needs CUDA tensors; the hot path has no CPU fallback")
    ctx = ctx or Context.default(hits.device.index)
    n = offsets.numel() - 1
    if max_read_len is None:
        o = offsets.to(torch.int64) & 0xFFFFFFFF
        max_read_len = int((o[1:] - o[:-1]).max().item()) if n else 0
    nc, nr = C.c_uint32(0), C.c_uint32(0)
    check(lib().nvb_rescue_mates(ctx.h, genome_words.data_ptr(), int(genome_len), StringSetC(symbols.data_ptr(), offsets.data_ptr(), n, 8),
                                 int(max_read_len), aligner.c(), int(min_score), int(max_frag_len), hits.data_ptr(),
                                 ops.data_ptr() if ops is not None else None, int(ops.shape[1]) if ops is not None else 0,
                                 C.byref(nc), C.byref(nr), current_stream_ptr()), ctx.h)
    return nc.value, nr.value


def map_hits_mapq(hits, offsets, min_score_by_len, match_bonus=0, monotone=True, pairs=None, ctx=None):
    """mapping quality of every hit (``nvb_map_hits_mapq``: nvBowtie's BowtieMapq2, mapq.h:145-316). min_score_by_len: host int32
    array, entry L = scheme.min_score(L) as the caller's scoring scheme computes it; match_bonus: scheme.match(0);
    monotone: end-to-end scoring. pairs: the tensor ``pair_hits`` returned for the same hits, or None. Returns uint8 [n]."""
    if not hits.is_cuda:
        raise RuntimeError("map_hits_mapq needs CUDA tensors; the hot path has no CPU fallback")
    ctx = ctx or Context.default(hits.device.index)
    n = offsets.numel() - 1
    tab = np.ascontiguousarray(min_score_by_len, dtype=np.int32)
    out = torch.empty(n, dtype=torch.uint8, device=hits.device)
    check(lib().nvb_map_hits_mapq(ctx.h, hits.data_ptr(), offsets.data_ptr(), n, pairs.data_ptr() if pairs is not None else None,
                                  tab.ctypes.data, len(tab) - 1, int(match_bonus), 1 if monotone else 0, out.data_ptr(),
                                  current_stream_ptr()), ctx.h)
    torch.cuda.current_stream().synchronize()      # the host table must outlive the staged copy
    return out


def map_hits_to_mds(hits, ops, symbols, offsets, genome_words, mds_stride=None, ctx=None):
    """MDS byte-code + edit distance of the hits of ``map_reads`` (``nvb_map_hits_to_mds``): read r against the genome from
    hits[r].begin on the hit's strand. Returns (mds uint8 [n, mds_stride], mds_len int32 [n], edit_distance int32 [n])."""
    if not hits.is_cuda:
        raise RuntimeError("map_hits_to_mds needs CUDA tensors; the hot path has no CPU fallback")
    ctx = ctx or Context.default(hits.device.index)
    n, stride = ops.shape
    if mds_stride is None:
        mds_stride = 2 + 3 * stride
    mds = torch.zeros((n, mds_stride), dtype=torch.uint8, device=hits.device)
    mds_len = torch.empty(n, dtype=torch.int32, device=hits.device)
    ed = torch.empty(n, dtype=torch.int32, device=hits.device)
    check(lib().nvb_map_hits_to_mds(ctx.h, hits.data_ptr(), ops.data_ptr(), stride, StringSetC(symbols.data_ptr(), offsets.data_ptr(), n, 8), genome_words.data_ptr(),
                                    mds.data_ptr(), mds_stride, mds_len.data_ptr(), ed.data_ptr(), current_stream_ptr()), ctx.h)
    return mds, mds_len, ed


def pair_hits(hits, min_frag_len=0, max_frag_len=500, ctx=None):
    """mate pairing (``nvb_pair_hits``) over the device hits of 2 n independently mapped mates laid out pair by pair; forward-reverse
    policy with nvBowtie's fragment window (params.h:147-152). Returns int32 [n, 4] viewable as PAIR_DT."""
    if not hits.is_cuda:
        raise RuntimeError("pair_hits needs CUDA tensors; the hot path has no CPU fallback")
    ctx = ctx or Context.default(hits.device.index)
    n = hits.shape[0] // 2
    out = torch.empty((n, 4), dtype=torch.int32, device=hits.device)
    check(lib().nvb_pair_hits(ctx.h, hits.data_ptr(), n, int(min_frag_len), int(max_frag_len), out.data_ptr(), current_stream_ptr()), ctx.h)
    return out
