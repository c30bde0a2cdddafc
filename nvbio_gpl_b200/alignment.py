"""Host-side mirror of the reference's batched alignment interface for the hot path.

Aligner tag objects follow nvbio/alignment/alignment.h:365-519 (``EditDistanceAligner<TYPE>``,
``SmithWatermanAligner<TYPE, SimpleSmithWatermanScheme>``, ``GotohAligner<TYPE, SimpleGotohScheme>``,
schemes at nvbio/alignment/utils.h:81-124); ``batch_alignment_score`` /
``batch_banded_alignment_score`` follow nvbio/alignment/batched.h:98-192 and the traceback calls the
``Batched[Banded]AlignmentTraceback::enact`` contexts (batched.h:335-438). Sinks are
``aln::BestSink<int32>`` records, alignments ``aln::Alignment<int32>`` records. All compute goes
through the C ABI; there is no CPU implementation here.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, Context, AlignerC, StringSetC

EDIT_DISTANCE, SMITH_WATERMAN, GOTOH = 0, 1, 2
GLOBAL, LOCAL, SEMI_GLOBAL = 0, 1, 2

SINK_DT = np.dtype([("score", "<i4"), ("x", "<u4"), ("y", "<u4")])
ALN_DT = np.dtype([("score", "<i4"), ("src_x", "<u4"), ("src_y", "<u4"), ("sink_x", "<u4"), ("sink_y", "<u4")])


class Aligner:
    def __init__(self, kind, type_, match=0, mismatch=-1, gap_a=-1, gap_b=-1):
        self.kind, self.type = int(kind), int(type_)
        self.match, self.mismatch, self.gap_a, self.gap_b = int(match), int(mismatch), int(gap_a), int(gap_b)

    def c(self):
        return AlignerC(self.kind, self.type, self.match, self.mismatch, self.gap_a, self.gap_b)


def EditDistanceAligner(type_):
    """aln::EditDistanceAligner<TYPE> (alignment.h:365-393); scheme fixed to (0,-1,-1,-1) (ed_utils.h:36-43)"""
    return Aligner(EDIT_DISTANCE, type_, 0, -1, -1, -1)


def SmithWatermanAligner(type_, match, mismatch, deletion, insertion):
    """aln::SmithWatermanAligner<TYPE, SimpleSmithWatermanScheme(match,mismatch,deletion,insertion)>"""
    return Aligner(SMITH_WATERMAN, type_, match, mismatch, deletion, insertion)


def GotohAligner(type_, match, mismatch, gap_open, gap_ext):
    """aln::GotohAligner<TYPE, SimpleGotohScheme(match,mismatch,gap_open,gap_ext)>"""
    return Aligner(GOTOH, type_, match, mismatch, gap_open, gap_ext)


class QualityScheme:
    """the quality-aware half of a SmithWatermanScoringScheme / GotohScoringScheme (alignment.h:411-436, 488-505): match(q) and
    mismatch(a, b, q) as 256-entry tables over the pattern quality, plus Gotoh's text gap penalties (the aligner's gap_a / gap_b are
    the pattern gap penalties). Pass it as ``scheme=`` together with ``quals=`` (uint8 device tensor, one per pattern symbol)."""

    def __init__(self, match, mismatch, text_gap_open=0, text_gap_ext=0):
        self.match = np.ascontiguousarray(match, dtype=np.int32)
        self.mismatch = np.ascontiguousarray(mismatch, dtype=np.int32)
        if len(self.match) != 256 or len(self.mismatch) != 256:
            raise _lib.NvbError("QualityScheme: the match and mismatch tables need 256 entries each")
        self.text_gap_open, self.text_gap_ext = int(text_gap_open), int(text_gap_ext)

    def c(self):
        return _lib.QualitySchemeC(self.match.ctypes.data, self.mismatch.ctypes.data, self.text_gap_open, self.text_gap_ext)


def _qptr(quals):
    return quals.data_ptr() if quals is not None else None


class StringSet:
    """ConcatenatedStringSet (nvbio/strings/string_set.h:405-492): symbols + count+1 offsets, on a CUDA device"""

    def __init__(self, symbols, offsets, bits=8, max_len=None):
        self.symbols, self.offsets, self.bits = symbols, offsets, bits
        self.count = offsets.numel() - 1
        self._max_len = max_len

    @classmethod
    def from_numpy(cls, symbols, offsets, device, bits=8):
        sym = torch.from_numpy(np.ascontiguousarray(symbols)).to(device)
        if sym.numel() == 0:
            sym = torch.zeros(16, dtype=torch.uint8, device=device)
        off = torch.from_numpy(np.ascontiguousarray(offsets, dtype=np.uint32).view(np.int32)).to(device)
        o64 = np.asarray(offsets, dtype=np.int64)
        ml = int(np.max(np.diff(o64))) if len(o64) > 1 else 0
        return cls(sym, off, bits, ml)

    @property
    def max_len(self):
        if self._max_len is None:
            o = self.offsets.to(torch.int64) & 0xFFFFFFFF
            self._max_len = int((o[1:] - o[:-1]).max().item()) if self.count else 0
        return self._max_len

    def c(self):
        s = StringSetC()
        s.symbols, s.offsets = self.symbols.data_ptr(), self.offsets.data_ptr()
        s.count, s.bits = self.count, self.bits
        return s


def _ctx(t, ctx):
    if not t.is_cuda:
        raise _lib.NvbError("nvbio_gpl_b200 has no CPU path: tensors must live on a CUDA device")
    return ctx or Context.default(t.device.index)


def _temp(op, aligner, band, chk, maxM, maxN, n, device, temp):
    need = int(lib().nvb_temp_bytes(op, aligner.c(), band, chk, maxM, maxN, n))
    if temp is None:
        temp = torch.empty(need, dtype=torch.uint8, device=device)
    return temp


def batch_alignment_score(aligner, patterns, texts, sinks=None, temp=None, ctx=None, scheme=None, quals=None, text_blocking=False):
    """aln::batch_alignment_score(aligner, patterns, [quals,] texts, sinks, scheduler, maxM, maxN) (batched.h:98-192).
    sinks: int32 [n,3] = (score, sink.x = text end, sink.y = pattern end). scheme / quals: a QualityScheme and the pattern
    qualities (the reference's quals overload, batched.h:126-148). text_blocking: the aligner's TextBlockingTag variant
    (``nvb_batch_score_text_blocking``; alignment.h:266-279: same scores, LOCAL ties in the text-blocking visiting order)"""
    ctx = _ctx(patterns.offsets, ctx)
    n = patterns.count
    dev = patterns.offsets.device
    if sinks is None:
        sinks = torch.empty((n, 3), dtype=torch.int32, device=dev)
    maxM, maxN = patterns.max_len, texts.max_len
    if text_blocking:
        if scheme is not None:
            raise _lib.NvbError("text_blocking is available for the simple schemes only")
        transposed = aligner.type != SEMI_GLOBAL
        temp = _temp(_lib.OP_SCORE, aligner, 0, 0, maxN if transposed else maxM, maxM if transposed else maxN, n, dev, temp)
        check(lib().nvb_batch_score_text_blocking(ctx.h, aligner.c(), patterns.c(), texts.c(), sinks.data_ptr(), maxM, maxN,
                                                  temp.data_ptr(), temp.numel(), _lib.current_stream_ptr()), ctx.h)
        return sinks
    temp = _temp(_lib.OP_SCORE, aligner, 0, 0, maxM, maxN, n, dev, temp)
    if scheme is not None:
        qs = scheme.c()
        check(lib().nvb_batch_score_q(ctx.h, aligner.c(), C.byref(qs), patterns.c(), _qptr(quals), texts.c(), sinks.data_ptr(), maxM, maxN,
                                      temp.data_ptr(), temp.numel(), _lib.current_stream_ptr()), ctx.h)
        return sinks
    check(lib().nvb_batch_score(ctx.h, aligner.c(), patterns.c(), texts.c(), sinks.data_ptr(), maxM, maxN,
                                temp.data_ptr(), temp.numel(), _lib.current_stream_ptr()), ctx.h)
    return sinks


NOALLOC_SCRATCH_BYTES = 32 << 20


def batch_alignment_score_noalloc(aligner, patterns, texts, sinks, max_pattern_len=None, max_text_len=None, ctx=None):
    """nvb_batch_score_noalloc: batch_alignment_score with its scratch in a static device array of the library (no allocation and
    no synchronisation in the call, so it can run under CUDA graph capture once a Context exists). `sinks` is the caller's int32
    [n,3] tensor on the sets' device. The maximum lengths must be known without touching the device: pass them, or build the
    StringSets with max_len (from_numpy does). Raises NvbError (status -5) beyond NOALLOC_SCRATCH_BYTES."""
    ctx = _ctx(patterns.offsets, ctx)
    maxM = patterns._max_len if max_pattern_len is None else int(max_pattern_len)
    maxN = texts._max_len if max_text_len is None else int(max_text_len)
    if maxM is None or maxN is None:
        raise _lib.NvbError("batch_alignment_score_noalloc: maximum string lengths unknown; computing them would allocate and "
                            "synchronise -- pass max_pattern_len / max_text_len or construct the StringSets with max_len")
    n = patterns.count
    if texts.count != n:
        raise _lib.NvbError("batch_alignment_score_noalloc: %d patterns but %d texts" % (n, texts.count))
    if not (isinstance(sinks, torch.Tensor) and sinks.is_cuda and sinks.device == patterns.offsets.device and sinks.dtype == torch.int32
            and sinks.is_contiguous() and tuple(sinks.shape) == (n, 3)):
        raise _lib.NvbError("batch_alignment_score_noalloc: sinks must be a contiguous int32 [%d, 3] tensor on %s" % (n, patterns.offsets.device))
    check(lib().nvb_batch_score_noalloc(ctx.h, aligner.c(), patterns.c(), texts.c(), sinks.data_ptr(), maxM, maxN,
                                        _lib.current_stream_ptr()), ctx.h)
    return sinks


def batch_banded_alignment_score(aligner, band, patterns, texts, sinks=None, ctx=None, scheme=None, quals=None):
    """aln::batch_banded_alignment_score<BAND_LEN>(...) (batched.h:142-192)"""
    ctx = _ctx(patterns.offsets, ctx)
    n = patterns.count
    dev = patterns.offsets.device
    if sinks is None:
        sinks = torch.empty((n, 3), dtype=torch.int32, device=dev)
    if scheme is not None:
        qs = scheme.c()
        check(lib().nvb_batch_banded_score_q(ctx.h, aligner.c(), C.byref(qs), band, patterns.c(), _qptr(quals), texts.c(), sinks.data_ptr(),
                                             patterns.max_len, texts.max_len, None, 0, _lib.current_stream_ptr()), ctx.h)
        return sinks
    check(lib().nvb_batch_banded_score(ctx.h, aligner.c(), band, patterns.c(), texts.c(), sinks.data_ptr(),
                                       patterns.max_len, texts.max_len, None, 0, _lib.current_stream_ptr()), ctx.h)
    return sinks


class MyersTag:
    """aln::MyersTag<ALPHABET_SIZE> (alignment.h:274-275): algorithm tag of the bit-vector edit distance"""

    def __init__(self, alphabet_size):
        self.alphabet_size = int(alphabet_size)


class MyersAligner:
    """aln::EditDistanceAligner<TYPE, MyersTag<ALPHABET>> (make_edit_distance_aligner<TYPE, MyersTag<A>>(), alignment.h:208)"""

    def __init__(self, type_, tag):
        self.type, self.alphabet = int(type_), int(tag.alphabet_size)


MYERS_BATCH_MIN_SCORE = -(1 << 30)   # Field_traits<int32>::min(): what the reference's batch drivers pass (-> int16 0)


def batch_banded_myers_score(aligner, band, patterns, texts, min_score=MYERS_BATCH_MIN_SCORE, sinks=None, ctx=None):
    """aln::batch_banded_alignment_score<BAND_LEN>(EditDistanceAligner<TYPE, MyersTag<A>>, ...) (batched.h:142-192 over
    myers/myers_banded_inl.h:183-296); min_score is truncated to int16 as the reference does"""
    ctx = _ctx(patterns.offsets, ctx)
    n = patterns.count
    if sinks is None:
        sinks = torch.empty((n, 3), dtype=torch.int32, device=patterns.offsets.device)
    check(lib().nvb_batch_banded_myers_score(ctx.h, aligner.type, aligner.alphabet, band, patterns.c(), texts.c(), int(min_score),
                                             sinks.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return sinks


def batch_alignment_score_best2(aligner, band, distinct_dist, patterns, texts, sinks=None, temp=None, ctx=None):
    """batch_alignment_score / batch_banded_alignment_score<band> with aln::Best2Sink<int32>(distinct_dist) sinks
    (sink.h:99-123). sinks: int32 [n,6] = (score1, score2, sink1.x, sink1.y, sink2.x, sink2.y); band 0 = full matrix"""
    ctx = _ctx(patterns.offsets, ctx)
    n = patterns.count
    dev = patterns.offsets.device
    if sinks is None:
        sinks = torch.empty((n, 6), dtype=torch.int32, device=dev)
    maxM, maxN = patterns.max_len, texts.max_len
    if band == 0:
        temp = _temp(_lib.OP_SCORE, aligner, 0, 0, maxM, maxN, n, dev, temp)
    tp, tn = (temp.data_ptr(), temp.numel()) if temp is not None else (None, 0)
    check(lib().nvb_batch_score_best2(ctx.h, aligner.c(), band, distinct_dist, patterns.c(), texts.c(), sinks.data_ptr(),
                                      maxM, maxN, tp, tn, _lib.current_stream_ptr()), ctx.h)
    return sinks


def batch_alignment_traceback(aligner, checkpoints, patterns, texts, ops_stride=None, temp=None, ctx=None, scheme=None, quals=None):
    """BatchedAlignmentTraceback<CHECKPOINTS,...>::enact (batched.h:335-380).
    Returns (alignments int32 [n,5], ops uint8 [n,ops_stride] in push order, n_ops int32 [n])"""
    ctx = _ctx(patterns.offsets, ctx)
    n = patterns.count
    dev = patterns.offsets.device
    maxM, maxN = patterns.max_len, texts.max_len
    if ops_stride is None:
        ops_stride = (maxM + maxN + 8 + 15) & ~15         # rows 16-byte aligned: the packed walk stores 16 ops at a time
    out = torch.empty((n, 5), dtype=torch.int32, device=dev)
    ops = torch.zeros((n, ops_stride), dtype=torch.uint8, device=dev)
    n_ops = torch.empty(n, dtype=torch.int32, device=dev)
    temp = _temp(_lib.OP_TRACEBACK, aligner, 0, checkpoints, maxM, maxN, n, dev, temp)
    if scheme is not None:
        qs = scheme.c()
        check(lib().nvb_batch_traceback_q(ctx.h, aligner.c(), C.byref(qs), checkpoints, patterns.c(), _qptr(quals), texts.c(), out.data_ptr(),
                                          ops.data_ptr(), ops_stride, n_ops.data_ptr(), maxM, maxN, temp.data_ptr(), temp.numel(),
                                          _lib.current_stream_ptr()), ctx.h)
        return out, ops, n_ops
    check(lib().nvb_batch_traceback(ctx.h, aligner.c(), checkpoints, patterns.c(), texts.c(), out.data_ptr(), ops.data_ptr(),
                                    ops_stride, n_ops.data_ptr(), maxM, maxN, temp.data_ptr(), temp.numel(),
                                    _lib.current_stream_ptr()), ctx.h)
    return out, ops, n_ops


def batch_banded_alignment_traceback(aligner, band, checkpoints, patterns, texts, ops_stride=None, temp=None, ctx=None, bufs=None, scheme=None, quals=None):
    """BatchedBandedAlignmentTraceback<BAND_LEN,CHECKPOINTS,...>::enact (batched.h:382-438)"""
    ctx = _ctx(patterns.offsets, ctx)
    n = patterns.count
    dev = patterns.offsets.device
    maxM, maxN = patterns.max_len, texts.max_len
    if ops_stride is None:
        ops_stride = (2 * maxM + band + 8 + 15) & ~15     # rows 16-byte aligned: the walk stores 16 ops at a time
    if bufs is not None:            # caller-owned (out, ops, n_ops) of a previous call with the same shapes
        out, ops, n_ops = bufs
        ops_stride = ops.shape[1]
    else:
        out = torch.empty((n, 5), dtype=torch.int32, device=dev)
        ops = torch.zeros((n, ops_stride), dtype=torch.uint8, device=dev)
        n_ops = torch.empty(n, dtype=torch.int32, device=dev)
    temp = _temp(_lib.OP_BANDED_TRACEBACK, aligner, band, checkpoints, maxM, maxN, n, dev, temp)
    if scheme is not None:
        qs = scheme.c()
        check(lib().nvb_batch_banded_traceback_q(ctx.h, aligner.c(), C.byref(qs), band, checkpoints, patterns.c(), _qptr(quals), texts.c(),
                                                 out.data_ptr(), ops.data_ptr(), ops_stride, n_ops.data_ptr(), maxM, maxN, temp.data_ptr(),
                                                 temp.numel(), _lib.current_stream_ptr()), ctx.h)
        return out, ops, n_ops
    check(lib().nvb_batch_banded_traceback(ctx.h, aligner.c(), band, checkpoints, patterns.c(), texts.c(), out.data_ptr(),
                                           ops.data_ptr(), ops_stride, n_ops.data_ptr(), maxM, maxN, temp.data_ptr(),
                                           temp.numel(), _lib.current_stream_ptr()), ctx.h)
    return out, ops, n_ops


def ops_to_cigar(ops, n_ops, cigar_stride=None, ctx=None):
    """push-order ops -> forward run-length io::Cigar words (nvbio/io/alignments.h:48-65)"""
    ctx = _ctx(ops, ctx)
    n, stride = ops.shape
    if cigar_stride is None:
        cigar_stride = stride
    cigar = torch.zeros((n, cigar_stride), dtype=torch.int16, device=ops.device)
    n_cigar = torch.empty(n, dtype=torch.int32, device=ops.device)
    check(lib().nvb_ops_to_cigar(ctx.h, ops.data_ptr(), stride, n_ops.data_ptr(), n, cigar.data_ptr(), cigar_stride,
                                 n_cigar.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return cigar, n_cigar


def ops_to_mds(ops, n_ops, patterns, texts, alignments, mds_stride=None, ctx=None):
    """nvBowtie's MDS byte-code and the edit distance of every alignment (``nvb_ops_to_mds``; finish_alignment_kernel,
    nvBowtie/bowtie2/cuda/traceback_inl.h:555-690). alignments: the int32 [n,5] records of a traceback call (their source
    gives the first aligned text / pattern symbol). Returns (mds uint8 [n, mds_stride], mds_len int32 [n], edit_distance int32 [n])."""
    ctx = _ctx(ops, ctx)
    n, stride = ops.shape
    if mds_stride is None:
        mds_stride = 2 + 3 * stride
    mds = torch.zeros((n, mds_stride), dtype=torch.uint8, device=ops.device)
    mds_len = torch.empty(n, dtype=torch.int32, device=ops.device)
    ed = torch.empty(n, dtype=torch.int32, device=ops.device)
    check(lib().nvb_ops_to_mds(ctx.h, ops.data_ptr(), stride, n_ops.data_ptr(), n, patterns.c(), texts.c(), alignments.data_ptr(),
                               mds.data_ptr(), mds_stride, mds_len.data_ptr(), ed.data_ptr(), _lib.current_stream_ptr()), ctx.h)
    return mds, mds_len, ed


def mds_to_md_string(mds):
    """SAM MD:Z text of one MDS byte-code, as SamOutput::generate_md_string writes it (nvbio/io/output/output_sam.cpp:216-289):
    match runs as numbers, a mismatch as the READ's symbol (what the byte-code stores), a deletion as ^symbols followed by 0,
    insertions skipped. Host-side formatting, not part of the device path. Returns (md, mismatches, gap_opens, gap_extensions)."""
    m = [int(v) for v in mds]
    n = m[0] | (m[1] << 8)
    out, mm, gapo, gape, i = [], 0, 0, 0, 2
    while i < n:
        op = m[i]; i += 1
        if op == 0:
            run = m[i]; i += 1
            while i + 1 < n and m[i] == 0:        # a run longer than 255 continues in further MATCH records
                run += m[i + 1]; i += 2
            out.append(str(run))
        elif op == 1:
            out.append("ACGTN"[min(m[i], 4)]); i += 1; mm += 1
        elif op == 2:
            l = m[i]; i += 1 + l; gapo += 1; gape += l - 1
        else:
            l = m[i]; i += 1
            out.append("^" + "".join("ACGTN"[min(c, 4)] for c in m[i:i + l]) + "0"); i += l; gapo += 1; gape += l - 1
    return "".join(out), mm, gapo, gape


# host-buffer forms --------------------------------------------------------------------------------
def batch_score_host(aligner, band, pat, pat_off, txt, txt_off, sinks_h=None, device=0, ctx=None, bits=8):
    """host string sets in, host sinks out (``nvb_batch_score_host[_packed]``); bits = 4 / 2: pat and txt are packed uint32 words
    (``pack_symbols``), offsets stay in symbols. Large batches are pipelined in chunks (upload under the kernels)."""
    ctx = ctx or Context.default(device)
    n = len(pat_off) - 1
    if sinks_h is None:
        sinks_h = np.empty(n, dtype=SINK_DT)
    if bits != 8:
        check(lib().nvb_batch_score_host_packed(ctx.h, aligner.c(), band, bits, pat.ctypes.data, pat_off.ctypes.data, txt.ctypes.data,
                                                txt_off.ctypes.data, n, sinks_h.ctypes.data), ctx.h)
        return sinks_h
    check(lib().nvb_batch_score_host(ctx.h, aligner.c(), band, pat.ctypes.data, pat_off.ctypes.data, txt.ctypes.data,
                                     txt_off.ctypes.data, n, sinks_h.ctypes.data), ctx.h)
    return sinks_h


def batch_traceback_host(aligner, band, checkpoints, pat, pat_off, txt, txt_off, ops_stride, device=0, ctx=None,
                         out=None, ops=None, n_ops=None):
    """host arrays in, host arrays out (``nvb_batch_traceback_host``). out / ops / n_ops: caller-owned result arrays (ALN_DT [n],
    uint8 [n, ops_stride], uint32 [n]); pinned ones (``nvb_host_alloc`` / ``torch.Tensor.pin_memory``) avoid the bounce copy"""
    ctx = ctx or Context.default(device)
    n = len(pat_off) - 1
    out = np.empty(n, dtype=ALN_DT) if out is None else out
    ops = np.zeros((n, ops_stride), dtype=np.uint8) if ops is None else ops
    n_ops = np.empty(n, dtype=np.uint32) if n_ops is None else n_ops
    if len(out) < n or ops.shape[0] < n or ops.shape[1] != ops_stride or len(n_ops) < n:
        raise ValueError("result arrays too small for %d pairs" % n)
    check(lib().nvb_batch_traceback_host(ctx.h, aligner.c(), band, checkpoints, pat.ctypes.data, pat_off.ctypes.data,
                                         txt.ctypes.data, txt_off.ctypes.data, n, out.ctypes.data, ops.ctypes.data,
                                         ops_stride, n_ops.ctypes.data), ctx.h)
    return out, ops, n_ops
