"""ctypes binding of oracle_abi.h -- TEST INFRASTRUCTURE ONLY.

Loads either oracle/liboracle.so (prefix ``orc_``: the plain-C restatement) or
oracle/_ref/libnvbio_ref.so (prefix ``ref_``: the reference's own headers, compiled by oracle/Makefile).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

ED, SW, GOTOH = 0, 1, 2
GLOBAL, LOCAL, SEMI_GLOBAL = 0, 1, 2

SINK_DT = np.dtype([("score", "<i4"), ("x", "<u4"), ("y", "<u4")])
PAIR_DT = np.dtype([("score", "<i4"), ("frag_begin", "<u4"), ("frag_end", "<u4"), ("flags", "<u4")])
SINK2_DT = np.dtype([("score1", "<i4"), ("score2", "<i4"), ("x1", "<u4"), ("y1", "<u4"), ("x2", "<u4"), ("y2", "<u4")])
ALN_DT = np.dtype([("score", "<i4"), ("src_x", "<u4"), ("src_y", "<u4"), ("sink_x", "<u4"), ("sink_y", "<u4")])


class FMIndexC(C.Structure):
    _fields_ = [("length", C.c_uint32), ("primary", C.c_uint32), ("L2", C.c_uint32 * 5),
                ("bwt_occ", C.c_void_p), ("ssa", C.c_void_p)]


class AlignerC(C.Structure):
    _fields_ = [("kind", C.c_int32), ("type", C.c_int32), ("p0", C.c_int32), ("p1", C.c_int32),
                ("p2", C.c_int32), ("p3", C.c_int32)]


class QSchemeC(C.Structure):
    _fields_ = [("match", C.c_void_p), ("mismatch", C.c_void_p), ("text_gap_open", C.c_int32), ("text_gap_ext", C.c_int32)]


class QScheme:
    """quality-aware scheme tables (alignment.h:411-436, 488-505): match(q), mismatch(a, b, q) for q = 0..255 and, for Gotoh, the
    text gap penalties; the aligner's p2/p3 are deletion/insertion (SW) or pattern_gap_open/extension (Gotoh)"""

    def __init__(self, match, mismatch, text_gap_open=0, text_gap_ext=0):
        self.match = np.ascontiguousarray(match, dtype=np.int32)
        self.mismatch = np.ascontiguousarray(mismatch, dtype=np.int32)
        assert len(self.match) == 256 and len(self.mismatch) == 256
        self.tgo, self.tge = int(text_gap_open), int(text_gap_ext)

    def c(self):
        return QSchemeC(self.match.ctypes.data, self.mismatch.ctypes.data, self.tgo, self.tge)

    @staticmethod
    def bowtie2(match=2, mm_min=2, mm_max=6):
        """nvBowtie's SmithWatermanScoringScheme<QualCost, ...> (nvBowtie/bowtie2/cuda/scoring.h:73-92,279-281): constant match bonus,
        mismatch = -(min + int(min(q, 40) / 40 * (max - min)))"""
        q = np.arange(256)
        frac = (np.minimum(q, 40) / np.float32(40.0)).astype(np.float32)
        mmp = mm_min + (frac * np.float32(mm_max - mm_min)).astype(np.int32)
        return np.full(256, match, np.int32), (-mmp).astype(np.int32)


class MapParamsC(C.Structure):
    _fields_ = [("seed_len", C.c_uint32), ("seed_interval", C.c_uint32), ("max_hits", C.c_uint32), ("max_range", C.c_uint32),
                ("band", C.c_uint32), ("both_strands", C.c_uint32), ("min_score", C.c_int32), ("aligner", AlignerC),
                ("seed_mismatches", C.c_uint32), ("max_seed_ranges", C.c_uint32), ("hit_priority", C.c_uint32),
                ("max_seed_hits", C.c_uint32), ("max_rows", C.c_uint32), ("reverse_index", C.c_void_p)]


HIT_DT = np.dtype([("score", "<i4"), ("second_score", "<i4"), ("begin", "<u4"), ("end", "<u4"), ("flags", "<u4"), ("n_ops", "<u4")])


def aligner(kind, type_, p=(0, -1, -1, -1)):
    return AlignerC(kind, type_, *[int(v) for v in p])


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def concat(strings):
    """list of uint8 arrays -> (symbols, offsets) ConcatenatedStringSet."""
    offs = np.zeros(len(strings) + 1, dtype=np.uint32)
    if len(strings):
        offs[1:] = np.cumsum([len(s) for s in strings])
    sym = np.concatenate([np.asarray(s, dtype=np.uint8) for s in strings]) if len(strings) and offs[-1] else np.zeros(0, np.uint8)
    return np.ascontiguousarray(sym, dtype=np.uint8), offs


def dna(s):
    """ASCII ACGT -> 0..3 (others -> 4)"""
    lut = np.full(256, 4, np.uint8)
    for i, ch in enumerate("ACGT"):
        lut[ord(ch)] = i
        lut[ord(ch.lower())] = i
    return lut[np.frombuffer(s.encode(), dtype=np.uint8)]


def rle(ops):
    """run-length encode an op array in the given order, e.g. [0,0,2] -> '2M1D'"""
    out, prev, cnt = [], None, 0
    for o in ops:
        if o == prev:
            cnt += 1
        else:
            if prev is not None:
                out.append("%d%s" % (cnt, "MID"[prev]))
            prev, cnt = int(o), 1
    if prev is not None:
        out.append("%d%s" % (cnt, "MID"[prev]))
    return "".join(out)


class Index:
    """host-side nvBWT index image (numpy arrays)"""

    def __init__(self, length, primary, L2, bwt_occ, ssa, sa=None):
        self.length, self.primary = int(length), int(primary)
        self.L2 = np.asarray(L2, dtype=np.uint32)
        self.bwt_occ, self.ssa, self.sa = bwt_occ, ssa, sa

    def c(self):
        f = FMIndexC()
        f.length, f.primary = self.length, self.primary
        for i in range(5):
            f.L2[i] = int(self.L2[i])
        f.bwt_occ = self.bwt_occ.ctypes.data
        f.ssa = self.ssa.ctypes.data if self.ssa is not None else None
        return f


class Oracle:
    def __init__(self, which="oracle"):
        if which == "oracle":
            path, self.prefix = os.path.join(HERE, "liboracle.so"), "orc_"
        elif which == "ref":
            path, self.prefix = os.path.join(HERE, "_ref", "libnvbio_ref.so"), "ref_"
        else:
            raise ValueError(which)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.which = which
        self.lib = C.CDLL(path)
        f = self._f
        f("bwt_occ_words").restype = C.c_size_t
        f("bwt_occ_words").argtypes = [C.c_uint32]
        f("ssa_words").restype = C.c_size_t
        f("ssa_words").argtypes = [C.c_uint32]

    @staticmethod
    def available(which):
        p = os.path.join(HERE, "liboracle.so") if which == "oracle" else os.path.join(HERE, "_ref", "libnvbio_ref.so")
        return os.path.exists(p)

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    # ---------------- FM-index ----------------
    def build_index(self, text, want_sa=False):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        n = len(text)
        bwt_occ = np.zeros(self._f("bwt_occ_words")(n), dtype=np.uint32)
        ssa = np.zeros(self._f("ssa_words")(n), dtype=np.uint32)
        sa = np.zeros(n + 1, dtype=np.int32) if want_sa else None
        primary = C.c_uint32(0)
        L2 = np.zeros(5, dtype=np.uint32)
        rc = self._f("build_index")(_ptr(text), C.c_uint32(n), C.byref(primary), _ptr(L2), _ptr(bwt_occ), _ptr(ssa), _ptr(sa))
        assert rc == 0, rc
        return Index(n, primary.value, L2, bwt_occ, ssa, sa)

    def fm_match(self, idx, symbols, offsets, want_blocks=False, threads=0):
        n = len(offsets) - 1
        symbols = np.ascontiguousarray(symbols, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        ranges = np.zeros((n, 2), dtype=np.uint32)
        blocks = np.zeros(n, dtype=np.uint32) if want_blocks else None
        f = idx.c()
        rc = self._f("fm_match")(C.byref(f), _ptr(symbols), _ptr(offsets), C.c_uint32(n), _ptr(ranges), _ptr(blocks), C.c_int(threads))
        assert rc == 0, rc
        return (ranges, blocks) if want_blocks else ranges

    def fm_match_reverse(self, idx, symbols, offsets, threads=0):
        n = len(offsets) - 1
        symbols = np.ascontiguousarray(symbols, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        ranges = np.zeros((n, 2), dtype=np.uint32)
        f = idx.c()
        rc = self._f("fm_match_reverse")(C.byref(f), _ptr(symbols), _ptr(offsets), C.c_uint32(n), _ptr(ranges), C.c_int(threads))
        assert rc == 0, rc
        return ranges

    def fm_rank(self, idx, ranges, c, threads=0):
        ranges = np.ascontiguousarray(ranges, dtype=np.uint32)
        c = np.ascontiguousarray(c, dtype=np.uint8)
        out = np.zeros((len(c), 2), dtype=np.uint32)
        f = idx.c()
        rc = self._f("fm_rank")(C.byref(f), _ptr(ranges), _ptr(c), C.c_uint32(len(c)), _ptr(out), C.c_int(threads))
        assert rc == 0, rc
        return out

    def fm_rank4(self, idx, k, threads=0):
        k = np.ascontiguousarray(k, dtype=np.uint32)
        out = np.zeros((len(k), 4), dtype=np.uint32)
        f = idx.c()
        rc = self._f("fm_rank4")(C.byref(f), _ptr(k), C.c_uint32(len(k)), _ptr(out), C.c_int(threads))
        assert rc == 0, rc
        return out

    def fm_locate(self, idx, rows, want_steps=False, threads=0):
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        pos = np.zeros(len(rows), dtype=np.uint32)
        steps = np.zeros(len(rows), dtype=np.uint32) if want_steps else None
        f = idx.c()
        rc = self._f("fm_locate")(C.byref(f), _ptr(rows), C.c_uint32(len(rows)), _ptr(pos), _ptr(steps), C.c_int(threads))
        assert rc == 0, rc
        return (pos, steps) if want_steps else pos

    # ---------------- alignment ----------------
    @staticmethod
    def _sets(pat, pat_off, txt, txt_off):
        return (np.ascontiguousarray(pat, dtype=np.uint8), np.ascontiguousarray(pat_off, dtype=np.uint32),
                np.ascontiguousarray(txt, dtype=np.uint8), np.ascontiguousarray(txt_off, dtype=np.uint32))

    @staticmethod
    def _q(qs, quals):
        """(pointer to the scheme struct or NULL, quality bytes or NULL); the struct is kept alive by the caller's frame"""
        qc = qs.c() if qs is not None else None
        qb = np.ascontiguousarray(quals, dtype=np.uint8) if quals is not None else None
        return qc, (C.byref(qc) if qc is not None else None), qb

    def score(self, a, pat, pat_off, txt, txt_off, threads=0, qs=None, quals=None):
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        sinks = np.zeros(n, dtype=SINK_DT)
        qc, qp, qb = self._q(qs, quals)
        rc = self._f("score_q")(a, qp, _ptr(qb), _ptr(pat), _ptr(pat_off), _ptr(txt), _ptr(txt_off), C.c_uint32(n), _ptr(sinks), C.c_int(threads))
        assert rc == 0, rc
        return sinks

    def score_text_blocking(self, a, pat, pat_off, txt, txt_off, threads=0):
        """scoring with the aligner's TextBlockingTag variant (alignment.h:266-279)"""
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        sinks = np.zeros(n, dtype=SINK_DT)
        rc = self._f("score_text_blocking")(a, _ptr(pat), _ptr(pat_off), _ptr(txt), _ptr(txt_off), C.c_uint32(n), _ptr(sinks), C.c_int(threads))
        assert rc == 0, rc
        return sinks

    def banded_score(self, a, band, pat, pat_off, txt, txt_off, threads=0, qs=None, quals=None):
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        sinks = np.zeros(n, dtype=SINK_DT)
        qc, qp, qb = self._q(qs, quals)
        rc = self._f("banded_score_q")(a, qp, _ptr(qb), C.c_uint32(band), _ptr(pat), _ptr(pat_off), _ptr(txt), _ptr(txt_off), C.c_uint32(n), _ptr(sinks), C.c_int(threads))
        assert rc == 0, rc
        return sinks

    def traceback(self, a, chk, pat, pat_off, txt, txt_off, threads=0, qs=None, quals=None):
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        M = int(np.max(np.diff(pat_off.astype(np.int64)))) if n else 0
        N = int(np.max(np.diff(txt_off.astype(np.int64)))) if n else 0
        stride = M + N + 8
        out = np.zeros(n, dtype=ALN_DT)
        ops = np.zeros((n, stride), dtype=np.uint8)
        n_ops = np.zeros(n, dtype=np.uint32)
        qc, qp, qb = self._q(qs, quals)
        rc = self._f("traceback_q")(a, qp, _ptr(qb), C.c_uint32(chk), _ptr(pat), _ptr(pat_off), _ptr(txt), _ptr(txt_off), C.c_uint32(n),
                                  _ptr(out), _ptr(ops), C.c_uint32(stride), _ptr(n_ops), C.c_int(threads))
        assert rc == 0, rc
        return out, ops, n_ops

    def banded_traceback(self, a, band, chk, pat, pat_off, txt, txt_off, threads=0, qs=None, quals=None):
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        M = int(np.max(np.diff(pat_off.astype(np.int64)))) if n else 0
        stride = 2 * M + band + 8
        out = np.zeros(n, dtype=ALN_DT)
        ops = np.zeros((n, stride), dtype=np.uint8)
        n_ops = np.zeros(n, dtype=np.uint32)
        qc, qp, qb = self._q(qs, quals)
        rc = self._f("banded_traceback_q")(a, qp, _ptr(qb), C.c_uint32(band), C.c_uint32(chk), _ptr(pat), _ptr(pat_off), _ptr(txt), _ptr(txt_off),
                                         C.c_uint32(n), _ptr(out), _ptr(ops), C.c_uint32(stride), _ptr(n_ops), C.c_int(threads))
        assert rc == 0, rc
        return out, ops, n_ops

    def score_best2(self, a, band, distinct, pat, pat_off, txt, txt_off, threads=0):
        """aln::Best2Sink<int32>(distinct) over the full (band = 0) or banded sweep"""
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        sinks = np.zeros(n, dtype=SINK2_DT)
        rc = self._f("score_best2")(a, C.c_uint32(band), C.c_uint32(distinct), _ptr(pat), _ptr(pat_off), _ptr(txt), _ptr(txt_off),
                                    C.c_uint32(n), _ptr(sinks), C.c_int(threads))
        assert rc == 0, rc
        return sinks

    def banded_myers(self, type_, alphabet, band, min_score, pat, pat_off, txt, txt_off, threads=0):
        """EditDistanceAligner<type, MyersTag<alphabet>> banded scoring into a BestSink"""
        pat, pat_off, txt, txt_off = self._sets(pat, pat_off, txt, txt_off)
        n = len(pat_off) - 1
        sinks = np.zeros(n, dtype=SINK_DT)
        rc = self._f("banded_myers")(C.c_int32(type_), C.c_uint32(alphabet), C.c_uint32(band), C.c_int32(min_score), _ptr(pat), _ptr(pat_off),
                                     _ptr(txt), _ptr(txt_off), C.c_uint32(n), _ptr(sinks), C.c_int(threads))
        assert rc == 0, rc
        return sinks

    # ---------------- seed-and-extend composition (oracle_map.c) ----------------
    def map_reads(self, idx, genome, reads, read_off, params, ops_stride=None, threads=0):
        """params: MapParamsC. Returns (hits[HIT_DT], ops[n, stride], n_candidates[n])."""
        genome = np.ascontiguousarray(genome, dtype=np.uint8)
        reads = np.ascontiguousarray(reads, dtype=np.uint8)
        read_off = np.ascontiguousarray(read_off, dtype=np.uint32)
        n = len(read_off) - 1
        L = int(np.max(np.diff(read_off.astype(np.int64)))) if n else 0
        stride = ops_stride if ops_stride is not None else (2 * L + params.band + 16 + 15) // 16 * 16
        hits = np.zeros(n, dtype=HIT_DT)
        ops = np.zeros((n, stride), dtype=np.uint8)
        ncand = np.zeros(n, dtype=np.uint32)
        f = idx.c()
        rc = self._f("map_reads")(C.byref(f), _ptr(genome), _ptr(reads), _ptr(read_off), C.c_uint32(n), C.byref(params),
                                  _ptr(hits), _ptr(ops), C.c_uint32(stride), _ptr(ncand), C.c_int(threads))
        assert rc == 0, rc
        return hits, ops, ncand

    def pair_hits(self, hits, min_frag, max_frag):
        """mate pairing over independently mapped mates (hits[2 i], hits[2 i + 1]) -> PAIR_DT records"""
        hits = np.ascontiguousarray(hits)
        n = len(hits) // 2
        out = np.zeros(n, dtype=PAIR_DT)
        rc = self._f("pair_hits")(_ptr(hits), C.c_uint32(n), C.c_uint32(min_frag), C.c_uint32(max_frag), _ptr(out))
        assert rc == 0, rc
        return out

    def rescue_mates(self, genome, reads, read_off, a, min_score, max_frag, hits, ops=None):
        """in-place opposite-mate window search over hits[2 n] (and ops[2 n, stride]); returns (n_candidates, n_rescued)"""
        genome = np.ascontiguousarray(genome, dtype=np.uint8); reads = np.ascontiguousarray(reads, dtype=np.uint8)
        read_off = np.ascontiguousarray(read_off, dtype=np.uint32)
        n_pairs = (len(read_off) - 1) // 2
        nc = C.c_uint32(0); nr = C.c_uint32(0)
        rc = self._f("rescue_mates")(_ptr(genome), C.c_uint32(len(genome)), _ptr(reads), _ptr(read_off), C.c_uint32(n_pairs), a, C.c_int32(min_score),
                                     C.c_uint32(max_frag), _ptr(hits), _ptr(ops) if ops is not None else None,
                                     C.c_uint32(ops.shape[1] if ops is not None else 0), C.byref(nc), C.byref(nr))
        assert rc == 0, rc
        return nc.value, nr.value

    def mapq_v2(self, best, has_second, second, perfect, min_score, monotone):
        """BowtieMapq2 on caller-supplied scores (mapq.h:145-316)"""
        return int(self._f("mapq_v2")(C.c_int32(int(best)), C.c_int32(1 if has_second else 0), C.c_int32(int(second)), C.c_int32(int(perfect)),
                                      C.c_int32(int(min_score)), C.c_int32(1 if monotone else 0)))

    def ref_bowtie_mapq2(self, best, has_second, second, perfect, min_score, monotone, paired=False, o_best=0, o_perfect=0, o_min_score=0):
        """the REFERENCE's compiled calculator (oracle/_ref only)"""
        return int(self.lib.ref_bowtie_mapq2(C.c_int32(int(best)), C.c_int32(1 if has_second else 0), C.c_int32(int(second)), C.c_int32(int(perfect)),
                                             C.c_int32(int(min_score)), C.c_int32(1 if monotone else 0), C.c_int32(1 if paired else 0),
                                             C.c_int32(int(o_best)), C.c_int32(int(o_perfect)), C.c_int32(int(o_min_score))))

    def hit_priority(self, sizes, max_hits, max_seed_hits=0, max_rows=0):
        """(take[n], visiting order) of oracle_map.c's hit selection by SA-range size"""
        sizes = np.ascontiguousarray(sizes, dtype=np.uint32); n = len(sizes)
        take = np.zeros(n, dtype=np.uint32); order = np.zeros(max(n, 1), dtype=np.uint32); no = C.c_uint32(0)
        fn = self._f("hit_priority"); fn.restype = C.c_size_t
        fn(_ptr(sizes), C.c_uint32(n), C.c_uint32(max_hits), C.c_uint32(max_seed_hits), C.c_uint32(max_rows), _ptr(take), _ptr(order), C.byref(no))
        return take, order[:no.value].copy()

    def ref_hit_order(self, sizes, cap=0):
        """pop order of the REFERENCE's own seed-hit deque (oracle/_ref only): (slots, sizes)"""
        sizes = np.ascontiguousarray(sizes, dtype=np.uint32); n = len(sizes)
        slots = np.zeros(max(n, 1), dtype=np.uint32); out = np.zeros(max(n, 1), dtype=np.uint32); no = C.c_uint32(0)
        rc = self.lib.ref_hit_order(_ptr(sizes), C.c_uint32(n), C.c_uint32(cap), _ptr(slots), _ptr(out), C.byref(no))
        assert rc == 0, rc
        return slots[:no.value].copy(), out[:no.value].copy()

    def ops_to_mds(self, ops, n_ops, pattern, text, j0=0, k0=0, cap=4096):
        """MDS byte-code (uint8 array of its full length) and edit distance of one alignment (traceback_inl.h:555-690)"""
        ops = np.ascontiguousarray(ops, dtype=np.uint8); pattern = np.ascontiguousarray(pattern, dtype=np.uint8)
        text = np.ascontiguousarray(text, dtype=np.uint8)
        out = np.zeros(cap, dtype=np.uint8); ln = C.c_uint32(0); ed = C.c_uint32(0)
        rc = self._f("ops_to_mds")(_ptr(ops), C.c_uint32(int(n_ops)), _ptr(pattern), _ptr(text), C.c_uint32(j0), C.c_uint32(k0),
                                   _ptr(out), C.c_uint32(cap), C.byref(ln), C.byref(ed))
        assert rc == 0, rc
        return out[:min(cap, ln.value)].copy(), ln.value, ed.value

    def fm_match_approx(self, idx, symbols, offsets, len1, find_exact=True, cap=64, threads=0):
        """nvBowtie's map<find_exact> (mapping_inl.h:114-184): (ranges[n, cap, 2], counts[n])"""
        symbols = np.ascontiguousarray(symbols, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        n = len(offsets) - 1
        ranges = np.zeros((n, cap, 2), dtype=np.uint32)
        counts = np.zeros(n, dtype=np.uint32)
        f = idx.c()
        rc = self._f("fm_match_approx")(C.byref(f), _ptr(symbols), _ptr(offsets), C.c_uint32(n), C.c_uint32(len1), C.c_int(1 if find_exact else 0),
                                        C.c_uint32(cap), _ptr(ranges), _ptr(counts), C.c_int(threads))
        assert rc == 0, rc
        return ranges, counts


def mds_to_md_string(mds):
    """SAM MD:Z string from an MDS byte-code: restatement of SamOutput::generate_md_string (nvbio/io/output/output_sam.cpp:216-289),
    including the '0' it appends after a deletion and its handling of a match run that continues in a second MATCH record
    (:239-240 adds the record's OP byte, 0, and then resumes parsing at its count byte -- runs longer than 255 come out
    garbled in the reference; reproduced literally, exercised only below that length by the pipeline's 150-bp reads).
    Returns (md, mismatches, gap_opens, gap_extensions)."""
    mds = [int(v) for v in mds]
    mds_len = mds[0] | (mds[1] << 8)
    out = []; mm = gapo = gape = 0
    if mds_len <= 2:                              # never produced by the reference (an aligned read has >= 1 op); its do-while would read past the code
        return "", 0, 0, 0
    i = 2
    while True:                                   # do { ... } while (i < mds_len)
        op = mds[i]; i += 1
        if op == 0:
            l = mds[i]; i += 1
            while i < mds_len and mds[i] == 0:    # :239-240, literally
                l = (l + mds[i]) & 0xFF; i += 1
            out.append(str(l))
        elif op == 1:
            out.append("ACGTN"[min(mds[i], 4)]); i += 1; mm += 1
        elif op == 2:
            l = mds[i]; i += 1 + l; gapo += 1; gape += l - 1
        elif op == 3:
            l = mds[i]; i += 1
            out.append("^" + "".join("ACGTN"[min(c, 4)] for c in mds[i:i + l]) + "0"); i += l; gapo += 1; gape += l - 1
        if not i < mds_len:
            break
    return "".join(out), mm, gapo, gape
