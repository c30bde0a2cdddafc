"""Host-side mirror of the reference's FM-index interface for the hot path.

``FMIndex`` is the view ``nvbio::fm_index<rank_dictionary<2,64,...>, SSA_index_multiple_context<16>>``
(nvbio/fmindex/fmindex.h:320-361) over the fused nvBWT image of ``io::FMIndexData``
(nvbio/io/fmindex/fmindex.h:81-163); ``match / rank / rank4 / locate`` keep the reference's names and
argument meaning (nvbio/fmindex/fmindex_inl.h:31-460) but take whole batches, as
``FMIndexFilter::rank/locate`` does (nvbio/fmindex/filter_inl.h:193-393). All compute goes through
the C ABI (``include/nvbio_b200.h``); there is no CPU implementation here.

``build_index`` constructs a synthetic index (suffix sorting by prefix doubling with torch sorts, on
whatever device the text lives on). Index *construction* is outside the hot path (the reference
does it in nvBWT / sufsort); it exists so that benchmarks and tests have an index to query.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, Context, FMIndexC, StringSetC

OCC_INT, SA_INT = 64, 16   # FMIndexDataCore::OCC_INT / SA_INT (nvbio/io/fmindex/fmindex.h:85-86)


def _words_for(n):
    return (((n + 15) // 16) + 3) & ~3


def _u32(t):
    """torch int32 tensor -> numpy uint32 view (host copy)"""
    return t.detach().cpu().numpy().view(np.uint32)


class FMIndex:
    """fused nvBWT index image: 32 bytes per 64 BWT symbols = 4 big-endian 2-bit words + 4 counters"""

    def __init__(self, length, primary, L2, bwt_occ, ssa=None):
        self.length, self.primary = int(length), int(primary)
        self.L2 = [int(v) for v in L2]
        self.bwt_occ = bwt_occ      # torch.int32 [nvb_bwt_occ_words(length)]
        self.ssa = ssa              # torch.int32 [(length+16)//16] or None
        self.ktab, self.ktab_k = None, 0   # optional k-mer jump table (build_ktab)
        self.dssa, self.dssa_int = None, 0 # optional denser sampled SA (build_dssa)
        assert bwt_occ.dtype == torch.int32 and bwt_occ.numel() == _words_for(self.length) * 2

    @property
    def device(self):
        return self.bwt_occ.device

    def to(self, device):
        return FMIndex(self.length, self.primary, self.L2, self.bwt_occ.to(device),
                       None if self.ssa is None else self.ssa.to(device))

    def count(self, c):
        return self.L2[c + 1] - self.L2[c]

    def nbytes(self):
        return self.bwt_occ.numel() * 4 + (0 if self.ssa is None else self.ssa.numel() * 4)

    def c(self):
        f = FMIndexC()
        f.length, f.primary = self.length, self.primary
        for i in range(5):
            f.L2[i] = self.L2[i]
        f.occ_int, f.sa_int = OCC_INT, SA_INT
        f.bwt_occ = self.bwt_occ.data_ptr()
        f.ssa = self.ssa.data_ptr() if self.ssa is not None else None
        f.ktab = self.ktab.data_ptr() if self.ktab is not None else None
        f.ktab_k = self.ktab_k if self.ktab is not None else 0
        f.dssa = self.dssa.data_ptr() if self.dssa is not None else None
        f.dssa_int = self.dssa_int if self.dssa is not None else 0
        return f

    def build_ktab(self, k=10, ctx=None):
        """k-mer jump table (nvb_fmindex_build_ktab): match() then starts from the precomputed range of a query's
        last k symbols; identical results. 8 * 4**k bytes (k = 10: 8 MiB, L2-resident)."""
        ctx = ctx or Context.default(self.device.index)
        self.ktab, self.ktab_k = None, 0
        tab = torch.empty(int(lib().nvb_ktab_words(k)), dtype=torch.int32, device=self.device)
        f = self.c()
        check(lib().nvb_fmindex_build_ktab(ctx.h, C.byref(f), int(k), tab.data_ptr(), _lib.current_stream_ptr()), ctx.h)
        self.ktab, self.ktab_k = tab, int(k)
        return self

    def build_dssa(self, sa_int=4, ctx=None):
        """denser sampled SA (nvb_fmindex_build_dssa): entry i = locate(i * sa_int) through the 16-row SSA; every locate then walks
        to the nearest multiple of sa_int instead of 16. Identical positions; 4 * (length + sa_int) / sa_int bytes."""
        ctx = ctx or Context.default(self.device.index)
        self.dssa, self.dssa_int = None, 0
        tab = torch.empty(int(lib().nvb_dssa_words(self.length, sa_int)), dtype=torch.int32, device=self.device)
        f = self.c()
        check(lib().nvb_fmindex_build_dssa(ctx.h, C.byref(f), int(sa_int), tab.data_ptr(), _lib.current_stream_ptr()), ctx.h)
        self.dssa, self.dssa_int = tab, int(sa_int)
        return self

    def drop_dssa(self):
        self.dssa, self.dssa_int = None, 0
        return self

    def drop_ktab(self):
        self.ktab, self.ktab_k = None, 0
        return self

    def host_arrays(self):
        """(bwt_occ, ssa) as numpy uint32 -- what a host-side checker needs"""
        return _u32(self.bwt_occ), (None if self.ssa is None else _u32(self.ssa))

    def pin_l2(self, nbytes=None, ctx=None):
        """pin the hot BWT prefix in L2 for work on the ctx stream (nvb_fmindex_pin_l2)"""
        ctx = ctx or Context.default(self.device.index)
        f = self.c()
        check(lib().nvb_fmindex_pin_l2(ctx.h, C.byref(f), int(self.bwt_occ.numel() * 4 if nbytes is None else nbytes)), ctx.h)


# ------------------------------------------------------------------------------------------------
# synthetic index construction (not on the hot path)
# ------------------------------------------------------------------------------------------------
_SORT_MAX = (1 << 31) - 2      # torch.sort refuses dimensions above INT_MAX


def _sort_with_index(key):
    """(sorted keys, permutation) of an int64 key vector of any length: torch.sort where it applies, else the two halves are
    sorted separately and merged by rank (searchsorted); equal keys keep their original order"""
    m = key.numel()
    if m <= _SORT_MAX:
        return torch.sort(key, stable=True)
    h = m // 2
    ka, ia = torch.sort(key[:h], stable=True)
    kb, ib = torch.sort(key[h:], stable=True)
    ib += h
    del key
    pa = torch.searchsorted(kb, ka, right=False); pa += torch.arange(h, device=ka.device)
    ks = torch.empty(m, dtype=torch.int64, device=ka.device); idx = torch.empty(m, dtype=torch.int64, device=ka.device)
    ks[pa] = ka; idx[pa] = ia
    del pa, ia
    pb = torch.searchsorted(ka, kb, right=True); pb += torch.arange(m - h, device=kb.device)
    del ka
    ks[pb] = kb; idx[pb] = ib
    return ks, idx


def suffix_array(text):
    """suffix array of the n+1 suffixes of ``text`` (the empty suffix sorts first), prefix doubling"""
    dev = text.device
    n = text.numel()
    m = n + 1
    t = text.to(torch.int64) & 3
    # 21 symbols x 3 bits (0 = past the end, 1..4 = A..T) -> 63-bit keys
    K = 21
    pad = torch.zeros(K, dtype=torch.int64, device=dev)
    tp = torch.cat([t + 1, pad])
    del t
    key = torch.zeros(m, dtype=torch.int64, device=dev)
    for k in range(K):
        key = key * 8 + tp[k:k + m]
    del tp
    h = K
    while True:
        ks, idx = _sort_with_index(key)
        del key
        flag = torch.ones(m, dtype=torch.int64, device=dev)
        flag[1:] = (ks[1:] != ks[:-1]).to(torch.int64)
        del ks
        gid = torch.cumsum(flag, 0)          # 1-based dense group ids
        unique = int(gid[-1].item()) == m
        if unique or h > n:
            return idx
        rank = torch.empty(m, dtype=torch.int64, device=dev)
        rank[idx] = gid
        del gid, flag, idx
        nxt = torch.zeros(m, dtype=torch.int64, device=dev)
        if h < m:
            nxt[:m - h] = rank[h:]
        key = ((rank - (1 << 31)) << 32) + nxt      # (rank, next rank) as one signed key; both fit 32 bits for any uint32 text length
        del rank, nxt
        h *= 2


def build_index(text, with_ssa=True, return_sa=False):
    """text: uint8 tensor of symbols 0..3 (any device) -> FMIndex on that device.

    Layout follows nvbio/io/fmindex/fmindex_impl.cu:296-318 (fusing), rank_dictionary_inl.h:34-66
    (counters BEFORE each block), ssa_inl.h:255-300 (ssa[i] = SA[16 i], ssa[0] = -1)."""
    if isinstance(text, np.ndarray):
        text = torch.from_numpy(np.ascontiguousarray(text, dtype=np.uint8))
    dev = text.device
    n = text.numel()
    sa = suffix_array(text)                                   # int64 [n+1]
    primary = int(torch.nonzero(sa == 0)[0, 0].item())
    keep = torch.ones(n + 1, dtype=torch.bool, device=dev)
    keep[primary] = False
    prev = sa[keep] - 1                                       # n entries
    bwt = (text.to(torch.int64) & 3)[prev] if n else torch.zeros(0, dtype=torch.int64, device=dev)
    del prev, keep

    W = _words_for(n)
    nblk = W // 4
    sym = torch.zeros(W * 16, dtype=torch.int64, device=dev)
    sym[:n] = bwt
    shifts = (30 - 2 * torch.arange(16, device=dev, dtype=torch.int64))
    words = (sym.view(W, 16) << shifts).sum(1)               # big-endian 2-bit packing
    valid = torch.zeros(W * 16, dtype=torch.bool, device=dev)
    valid[:n] = True
    occ = torch.zeros(nblk, 4, dtype=torch.int64, device=dev)
    totals = []
    for c in range(4):
        per_blk = ((sym == c) & valid).view(nblk, 64).sum(1) if nblk else torch.zeros(0, dtype=torch.int64, device=dev)
        cs = torch.cumsum(per_blk, 0)
        occ[:, c] = cs - per_blk
        totals.append(int(cs[-1].item()) if nblk else 0)
    fused = torch.cat([words.view(nblk, 4), occ], 1).reshape(-1)
    fused = fused.to(torch.int32) if fused.numel() else torch.zeros(0, dtype=torch.int32, device=dev)
    L2 = [0]
    for c in range(4):
        L2.append(L2[-1] + totals[c])
    ssa = None
    if with_ssa:
        ssa = sa[::SA_INT].clone()
        ssa[0] = 0xFFFFFFFF
        ssa = ssa.to(torch.int32)
    fmi = FMIndex(n, primary, L2, fused.contiguous(), ssa)
    return (fmi, sa) if return_sa else fmi


# ------------------------------------------------------------------------------------------------
# queries (device tensors in, device tensors out; everything runs on torch's current stream)
# ------------------------------------------------------------------------------------------------
def _ctx(t, ctx):
    if not t.is_cuda:
        raise _lib.NvbError("nvbio_gpl_b200 has no CPU path: tensors must live on a CUDA device")
    return ctx or Context.default(t.device.index)


def _set_c(symbols, offsets, bits=8):
    s = StringSetC()
    s.symbols, s.offsets = symbols.data_ptr(), offsets.data_ptr()
    s.count, s.bits = offsets.numel() - 1, bits
    return s


def match(fmi, symbols, offsets, ranges=None, bits=8, ctx=None):
    """match(fmi, pattern, len) for every string (fmindex_inl.h:186-239). Returns int32 [n,2] holding the
    raw uint32 inclusive SA interval; pass ``ranges`` to continue from given intervals (match(.., range))."""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = offsets.numel() - 1
    cont = ranges is not None
    if ranges is None:
        ranges = torch.empty((n, 2), dtype=torch.int32, device=fmi.device)
    f = fmi.c()
    fn = lib().nvb_fm_match_from if cont else lib().nvb_fm_match
    check(fn(ctx.h, C.byref(f), _set_c(symbols, offsets, bits), ranges.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return ranges


def match_reverse(fmi, symbols, offsets, bits=8, ctx=None):
    """match_reverse(fmi, pattern, len) for every string (fmindex_inl.h:252-278): the pattern is consumed front to back, so on
    the index of the reversed text (nvBWT's .rbwt/.rsa) it finds the occurrences of the pattern itself. int32 [n,2] raw ranges."""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = offsets.numel() - 1
    ranges = torch.empty((n, 2), dtype=torch.int32, device=fmi.device)
    f = fmi.c()
    check(lib().nvb_fm_match_reverse(ctx.h, C.byref(f), _set_c(symbols, offsets, bits), ranges.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return ranges


def match_approx(fmi, symbols, offsets, len1, find_exact=True, cap=64, bits=8, ctx=None):
    """nvBowtie's map<find_exact> (mapping_inl.h:114-184): exact over the first len1 search steps, one substitution allowed in the
    rest. Returns (ranges int32 [n, cap, 2] holding raw uint32 SA ranges in the reference's push order, counts int32 [n])."""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = offsets.numel() - 1
    ranges = torch.zeros((n, cap, 2), dtype=torch.int32, device=fmi.device)
    counts = torch.empty(n, dtype=torch.int32, device=fmi.device)
    f = fmi.c()
    check(lib().nvb_fm_match_approx(ctx.h, C.byref(f), _set_c(symbols, offsets, bits), int(len1), 1 if find_exact else 0, int(cap),
                                    ranges.data_ptr(), counts.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return ranges, counts


def rank(fmi, ranges, c, ctx=None):
    """rank(fmi, range, c) (fmindex_inl.h:60-88): int32 [n,2]"""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = c.numel()
    out = torch.empty((n, 2), dtype=torch.int32, device=fmi.device)
    f = fmi.c()
    check(lib().nvb_fm_rank(ctx.h, C.byref(f), ranges.data_ptr(), c.data_ptr(), n, out.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return out


def rank4(fmi, k, ctx=None):
    """rank4(fmi, k) (fmindex_inl.h:100-125): int32 [n,4]"""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = k.numel()
    out = torch.empty((n, 4), dtype=torch.int32, device=fmi.device)
    f = fmi.c()
    check(lib().nvb_fm_rank4(ctx.h, C.byref(f), k.data_ptr(), n, out.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return out


def locate(fmi, rows, two_phase=False, ctx=None):
    """locate(fmi, row) (fmindex_inl.h:364-394); two_phase = locate_ssa_iterator + lookup_ssa_iterator (:408-460)"""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = rows.numel()
    pos = torch.empty(n, dtype=torch.int32, device=fmi.device)
    f = fmi.c()
    s = _lib.current_stream_ptr()
    if two_phase:
        walk = torch.empty((n, 2), dtype=torch.int32, device=fmi.device)
        check(lib().nvb_fm_locate_walk(ctx.h, C.byref(f), rows.data_ptr(), n, walk.data_ptr(), s), ctx.h)
        check(lib().nvb_fm_locate_lookup(ctx.h, C.byref(f), walk.data_ptr(), n, pos.data_ptr(), s), ctx.h)
    else:
        check(lib().nvb_fm_locate(ctx.h, C.byref(f), rows.data_ptr(), n, pos.data_ptr(), s), ctx.h)
    return pos


def fuse(bwt_words, length, ctx=None):
    """nvb_fmindex_fuse: plain BWT words (device int32) -> (fused image, L2)"""
    ctx = _ctx(bwt_words, ctx)
    out = torch.empty(lib().nvb_bwt_occ_words(length), dtype=torch.int32, device=bwt_words.device)
    L2 = (C.c_uint32 * 5)()
    check(lib().nvb_fmindex_fuse(ctx.h, bwt_words.data_ptr(), length, out.data_ptr(), C.cast(L2, C.c_void_p),
                                 _lib.current_stream_ptr()), ctx.h)
    return out, [int(v) for v in L2]


class FMIndexFilter:
    """FMIndexFilter<device_tag, fm_index_type> (nvbio/fmindex/filter.h:108-165): rank() matches every string of a set
    and scans the range sizes; locate(begin, end) enumerates hits [begin, end) as (text position, string id)."""

    def __init__(self, ctx=None):
        self.ctx = ctx
        self.index = None
        self.n_queries = 0
        self.n_occurrences = 0
        self.ranges = None     # int32 [n, 2]  (raw uint32 inclusive SA intervals)
        self.slots = None      # int64 [n]     (inclusive scan of the range sizes)

    def rank(self, fmi, symbols, offsets, bits=8):
        ctx = _ctx(fmi.bwt_occ, self.ctx)
        n = offsets.numel() - 1
        self.index, self.n_queries = fmi, n
        self.ranges = torch.empty((n, 2), dtype=torch.int32, device=fmi.device)
        self.slots = torch.empty(n, dtype=torch.int64, device=fmi.device)
        total = C.c_uint64(0)
        f = fmi.c()
        check(lib().nvb_fm_filter_rank(ctx.h, C.byref(f), _set_c(symbols, offsets, bits), self.ranges.data_ptr(), self.slots.data_ptr(),
                                       C.addressof(total), _lib.current_stream_ptr()), ctx.h)
        self.n_occurrences = int(total.value)
        return self.n_occurrences

    def n_hits(self):
        return self.n_occurrences

    def locate(self, begin, end, hits=None):
        fmi = self.index
        ctx = _ctx(fmi.bwt_occ, self.ctx)
        if hits is None:
            hits = torch.empty((max(0, end - begin), 2), dtype=torch.int32, device=fmi.device)
        f = fmi.c()
        check(lib().nvb_fm_filter_locate(ctx.h, C.byref(f), self.ranges.data_ptr(), self.slots.data_ptr(), self.n_queries,
                                         begin, end, hits.data_ptr(), _lib.current_stream_ptr()), ctx.h)
        return hits


# host-buffer forms: numpy (ideally pinned) in, numpy out; H2D + kernel + D2H inside the call
def match_host(fmi, symbols_h, offsets_h, ranges_h=None, ctx=None, bits=8):
    """bits = 8: one symbol per byte; bits = 4 | 2: symbols_h is the packed big-endian uint32 word stream (pack_symbols)"""
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = len(offsets_h) - 1
    if ranges_h is None:
        ranges_h = np.empty((n, 2), dtype=np.uint32)
    f = fmi.c()
    if bits == 8:
        check(lib().nvb_fm_match_host(ctx.h, C.byref(f), symbols_h.ctypes.data, offsets_h.ctypes.data, n, ranges_h.ctypes.data), ctx.h)
    else:
        check(lib().nvb_fm_match_host_packed(ctx.h, C.byref(f), symbols_h.ctypes.data, bits, offsets_h.ctypes.data, n, ranges_h.ctypes.data), ctx.h)
    return ranges_h


def match_host_fixed(fmi, symbols_h, query_len, count, ranges_h=None, ctx=None, bits=8):
    """queries of one length (seeds): query i = symbols [i * query_len, (i + 1) * query_len); no offsets cross PCIe"""
    ctx = _ctx(fmi.bwt_occ, ctx)
    if ranges_h is None:
        ranges_h = np.empty((count, 2), dtype=np.uint32)
    f = fmi.c()
    check(lib().nvb_fm_match_host_fixed(ctx.h, C.byref(f), symbols_h.ctypes.data, bits, query_len, count, ranges_h.ctypes.data), ctx.h)
    return ranges_h


def pack_symbols_2n(symbols, query_len):
    """fixed-length queries (uint8, 0..3 and > 3 for N) -> (2-bit big-endian words, N list): the wire format of
    ``nvb_fm_match_host_fixed2n``. N list: uint32 [k, 2] = (query index, position of its LAST N), ascending."""
    s = np.asarray(symbols, dtype=np.uint8)
    n = len(s) // query_len
    isn = (s[:n * query_len] > 3).reshape(n, query_len)
    has = isn.any(axis=1)
    qi = np.nonzero(has)[0].astype(np.uint32)
    last = (query_len - 1 - np.argmax(isn[qi][:, ::-1], axis=1)).astype(np.uint32) if len(qi) else np.zeros(0, np.uint32)
    return pack_symbols(s & 3, 2), np.ascontiguousarray(np.stack([qi, last], axis=1), dtype=np.uint32)


def match_host_fixed2n(fmi, words_h, n_list_h, query_len, count, ranges_h=None, ctx=None):
    """2-bit packed fixed-length queries + N exception list (``pack_symbols_2n``): the ranges of ``match_host_fixed``, bit for bit"""
    ctx = _ctx(fmi.bwt_occ, ctx)
    if ranges_h is None:
        ranges_h = np.empty((count, 2), dtype=np.uint32)
    f = fmi.c()
    nl = np.ascontiguousarray(n_list_h, dtype=np.uint32)
    check(lib().nvb_fm_match_host_fixed2n(ctx.h, C.byref(f), words_h.ctypes.data, query_len, count, nl.ctypes.data if len(nl) else None, len(nl),
                                          ranges_h.ctypes.data), ctx.h)
    return ranges_h


def pack_symbols(symbols, bits):
    """uint8 symbols -> big-endian packed uint32 words (PackedStream<uint32*, uint8, BITS, true>: symbol 0 in the top bits)"""
    spw = 32 // bits
    s = np.asarray(symbols, dtype=np.uint8)
    pad = (-len(s)) % spw
    if pad:
        s = np.concatenate([s, np.zeros(pad, dtype=np.uint8)])
    s = s.reshape(-1, spw).astype(np.uint32) & ((1 << bits) - 1)
    shifts = (32 - bits * (np.arange(spw, dtype=np.uint32) + 1)).astype(np.uint32)
    return (s << shifts[None, :]).sum(axis=1, dtype=np.uint64).astype(np.uint32)


def locate_host(fmi, rows_h, pos_h=None, ctx=None):
    ctx = _ctx(fmi.bwt_occ, ctx)
    n = len(rows_h)
    if pos_h is None:
        pos_h = np.empty(n, dtype=np.uint32)
    f = fmi.c()
    check(lib().nvb_fm_locate_host(ctx.h, C.byref(f), rows_h.ctypes.data, n, pos_h.ctypes.data), ctx.h)
    return pos_h
