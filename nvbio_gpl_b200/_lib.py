"""ctypes loader of libnvbio_b200.so (the C ABI of include/nvbio_b200.h). Fails loudly."""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnvbio_b200.so")

NVB_OK = 0
OP_SCORE, OP_BANDED_SCORE, OP_TRACEBACK, OP_BANDED_TRACEBACK = 0, 1, 2, 3


class NvbError(RuntimeError):
    pass


class FMIndexC(C.Structure):
    _fields_ = [("length", C.c_uint32), ("primary", C.c_uint32), ("L2", C.c_uint32 * 5),
                ("occ_int", C.c_uint32), ("sa_int", C.c_uint32),
                ("bwt_occ", C.c_void_p), ("ssa", C.c_void_p), ("ktab", C.c_void_p), ("ktab_k", C.c_uint32),
                ("dssa", C.c_void_p), ("dssa_int", C.c_uint32)]


class StringSetC(C.Structure):
    _fields_ = [("symbols", C.c_void_p), ("offsets", C.c_void_p), ("count", C.c_uint32), ("bits", C.c_uint32)]


class AlignerC(C.Structure):
    _fields_ = [("kind", C.c_int32), ("type", C.c_int32), ("match", C.c_int32), ("mismatch", C.c_int32),
                ("gap_a", C.c_int32), ("gap_b", C.c_int32)]


class QualitySchemeC(C.Structure):
    _fields_ = [("match_h", C.c_void_p), ("mismatch_h", C.c_void_p), ("text_gap_open", C.c_int32), ("text_gap_ext", C.c_int32)]


class MapParamsC(C.Structure):
    _fields_ = [("seed_len", C.c_uint32), ("seed_interval", C.c_uint32), ("max_hits", C.c_uint32), ("max_range", C.c_uint32),
                ("band", C.c_uint32), ("both_strands", C.c_uint32), ("min_score", C.c_int32), ("aligner", AlignerC),
                ("seed_mismatches", C.c_uint32), ("max_seed_ranges", C.c_uint32), ("hit_priority", C.c_uint32),
                ("max_seed_hits", C.c_uint32), ("max_rows", C.c_uint32), ("reverse_index", C.c_void_p)]


class MapStatsC(C.Structure):
    _fields_ = [("seeds", C.c_uint64), ("seed_hits", C.c_uint64), ("candidates", C.c_uint64), ("aligned", C.c_uint64)]


def build(verbose=False):
    """compile libnvbio_b200.so for sm_100a (nvcc cross-compiles without a GPU)"""
    cmd = ["make", "-C", os.path.join(HERE, "csrc"), "-j8"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode:
        print(r.stdout)
    if r.returncode:
        raise NvbError("building libnvbio_b200.so failed")
    return LIB_PATH


_lib = None

_vp, _u32, _sz, _i = C.c_void_p, C.c_uint32, C.c_size_t, C.c_int
_PROTOS = {
    "nvb_version": (C.c_int, []),
    "nvb_status_string": (C.c_char_p, [_i]),
    "nvb_ctx_create": (_i, [_i, C.POINTER(_vp)]),
    "nvb_ctx_destroy": (None, [_vp]),
    "nvb_ctx_stream": (_vp, [_vp]),
    "nvb_ctx_synchronize": (_i, [_vp]),
    "nvb_last_cuda_error": (C.c_char_p, [_vp]),
    "nvb_ctx_launch_count": (C.c_uint64, [_vp]),
    "nvb_host_alloc": (_i, [_sz, C.POINTER(_vp)]),
    "nvb_host_free": (None, [_vp]),
    "nvb_fmindex_upload": (_i, [_vp, C.POINTER(FMIndexC), C.POINTER(FMIndexC)]),
    "nvb_fmindex_release": (_i, [_vp, C.POINTER(FMIndexC)]),
    "nvb_fmindex_pin_l2": (_i, [_vp, C.POINTER(FMIndexC), _sz]),
    "nvb_ktab_words": (_sz, [_u32]),
    "nvb_fmindex_build_ktab": (_i, [_vp, C.POINTER(FMIndexC), _u32, _vp, _vp]),
    "nvb_dssa_words": (_sz, [_u32, _u32]),
    "nvb_fmindex_build_dssa": (_i, [_vp, C.POINTER(FMIndexC), _u32, _vp, _vp]),
    "nvb_bwt_occ_words": (_sz, [_u32]),
    "nvb_ssa_words": (_sz, [_u32]),
    "nvb_fmindex_fuse": (_i, [_vp, _vp, _u32, _vp, _vp, _vp]),
    "nvb_fm_match": (_i, [_vp, C.POINTER(FMIndexC), StringSetC, _vp, _vp]),
    "nvb_fm_match_from": (_i, [_vp, C.POINTER(FMIndexC), StringSetC, _vp, _vp]),
    "nvb_fm_match_reverse": (_i, [_vp, C.POINTER(FMIndexC), StringSetC, _vp, _vp]),
    "nvb_fm_match_approx": (_i, [_vp, C.POINTER(FMIndexC), StringSetC, _u32, _i, _u32, _vp, _vp, _vp]),
    "nvb_fm_rank": (_i, [_vp, C.POINTER(FMIndexC), _vp, _vp, _u32, _vp, _vp]),
    "nvb_fm_rank4": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _vp, _vp]),
    "nvb_fm_locate": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _vp, _vp]),
    "nvb_fm_locate_walk": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _vp, _vp]),
    "nvb_fm_locate_lookup": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _vp, _vp]),
    "nvb_fmindex_load": (_i, [_vp, C.c_char_p, _i, C.POINTER(FMIndexC)]),
    "nvb_fmindex_save": (_i, [_vp, C.POINTER(FMIndexC), C.c_char_p, _i]),
    "nvb_fm_filter_rank": (_i, [_vp, C.POINTER(FMIndexC), StringSetC, _vp, _vp, _vp, _vp]),
    "nvb_fm_filter_locate": (_i, [_vp, C.POINTER(FMIndexC), _vp, _vp, _u32, C.c_uint64, C.c_uint64, _vp, _vp]),
    "nvb_fm_match_host": (_i, [_vp, C.POINTER(FMIndexC), _vp, _vp, _u32, _vp]),
    "nvb_fm_match_host_packed": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _vp, _u32, _vp]),
    "nvb_fm_match_host_fixed": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _u32, _u32, _vp]),
    "nvb_fm_match_host_fixed2n": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _u32, _vp, _u32, _vp]),
    "nvb_fm_locate_host": (_i, [_vp, C.POINTER(FMIndexC), _vp, _u32, _vp]),
    "nvb_temp_bytes": (_sz, [_i, AlignerC, _u32, _u32, _u32, _u32, _u32]),
    "nvb_batch_score": (_i, [_vp, AlignerC, StringSetC, StringSetC, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_score_text_blocking": (_i, [_vp, AlignerC, StringSetC, StringSetC, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_score_noalloc_bytes": (_sz, [AlignerC, _u32, _u32]),
    "nvb_batch_score_noalloc": (_i, [_vp, AlignerC, StringSetC, StringSetC, _vp, _u32, _u32, _vp]),
    "nvb_batch_score_q": (_i, [_vp, AlignerC, C.POINTER(QualitySchemeC), StringSetC, _vp, StringSetC, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_banded_score_q": (_i, [_vp, AlignerC, C.POINTER(QualitySchemeC), _u32, StringSetC, _vp, StringSetC, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_traceback_q": (_i, [_vp, AlignerC, C.POINTER(QualitySchemeC), _u32, StringSetC, _vp, StringSetC, _vp, _vp, _u32, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_banded_traceback_q": (_i, [_vp, AlignerC, C.POINTER(QualitySchemeC), _u32, _u32, StringSetC, _vp, StringSetC, _vp, _vp, _u32, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_banded_score": (_i, [_vp, AlignerC, _u32, StringSetC, StringSetC, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_score_best2": (_i, [_vp, AlignerC, _u32, _u32, StringSetC, StringSetC, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_banded_myers_score": (_i, [_vp, C.c_int32, _u32, _u32, StringSetC, StringSetC, C.c_int32, _vp, _vp]),
    "nvb_batch_traceback": (_i, [_vp, AlignerC, _u32, StringSetC, StringSetC, _vp, _vp, _u32, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_batch_banded_traceback": (_i, [_vp, AlignerC, _u32, _u32, StringSetC, StringSetC, _vp, _vp, _u32, _vp, _u32, _u32, _vp, _sz, _vp]),
    "nvb_ops_to_cigar": (_i, [_vp, _vp, _u32, _vp, _u32, _vp, _u32, _vp, _vp]),
    "nvb_ops_to_mds": (_i, [_vp, _vp, _u32, _vp, _u32, StringSetC, StringSetC, _vp, _vp, _u32, _vp, _vp, _vp]),
    "nvb_rescue_mates": (_i, [_vp, _vp, _u32, StringSetC, _u32, AlignerC, C.c_int32, _u32, _vp, _vp, _u32, _vp, _vp, _vp]),
    "nvb_map_hits_mapq": (_i, [_vp, _vp, _vp, _u32, _vp, _vp, _u32, C.c_int32, _u32, _vp, _vp]),
    "nvb_map_hits_to_mds": (_i, [_vp, _vp, _vp, _u32, StringSetC, _vp, _vp, _u32, _vp, _vp, _vp]),
    "nvb_sinks_to_reference_layout": (_i, [_vp, _vp, _u32, _vp, _vp]),
    "nvb_alignments_to_reference_layout": (_i, [_vp, _vp, _u32, _vp, _vp]),
    "nvb_batch_score_host": (_i, [_vp, AlignerC, _u32, _vp, _vp, _vp, _vp, _u32, _vp]),
    "nvb_batch_score_host_packed": (_i, [_vp, AlignerC, _u32, _u32, _vp, _vp, _vp, _vp, _u32, _vp]),
    "nvb_batch_traceback_host": (_i, [_vp, AlignerC, _u32, _u32, _vp, _vp, _vp, _vp, _u32, _vp, _vp, _u32, _vp]),
    "nvb_pack_2bit": (_i, [_vp, _vp, _u32, _vp, _vp]),
    "nvb_map_reads": (_i, [_vp, C.POINTER(FMIndexC), _vp, StringSetC, _u32, C.POINTER(MapParamsC), _u32, _vp, _vp, _u32,
                           C.POINTER(MapStatsC), _vp]),
    "nvb_map_reads_host": (_i, [_vp, C.POINTER(FMIndexC), _vp, _vp, _vp, _u32, C.POINTER(MapParamsC), _u32, _vp, _vp, _u32,
                                _vp, _u32, C.POINTER(MapStatsC)]),
    "nvb_fm_match_host_multi": (_i, [_vp, _vp, _u32, _vp, _u32, _vp, _u32, _u32, _vp]),
    "nvb_map_reads_host_multi": (_i, [_vp, _vp, _vp, _u32, _vp, _vp, _u32, C.POINTER(MapParamsC), _u32, _vp, _vp, _u32, _vp, _u32, C.POINTER(MapStatsC)]),
    "nvb_pair_hits": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _vp]),
    "nvb_last_paths": (_u32, [_vp]),
    "nvb_force_generic": (None, [_vp, _i]),
    "nvb_fm_batch_sharing": (None, [_vp, _u32]),
}
EXPORTS = sorted(_PROTOS)


def lib():
    """the loaded C-ABI library; raises (never falls back) when it has not been built"""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NvbError("libnvbio_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(rc, ctx=None):
    if rc != NVB_OK:
        msg = lib().nvb_status_string(rc).decode()
        if ctx is not None and rc == -2:
            msg += ": " + lib().nvb_last_cuda_error(ctx).decode()
        raise NvbError("libnvbio_b200: %s (status %d)" % (msg, rc))


class Context:
    """nvb_ctx: one per (device, host thread)"""
    _default = {}

    def __init__(self, device=0):
        self.h = C.c_void_p()
        check(lib().nvb_ctx_create(int(device), C.byref(self.h)))
        self.device = int(device)

    @classmethod
    def default(cls, device=None):
        import torch
        if device is None:
            device = torch.cuda.current_device()
        if device not in cls._default:
            cls._default[device] = cls(device)
        return cls._default[device]

    @property
    def launches(self):
        return int(lib().nvb_ctx_launch_count(self.h))

    @property
    def stream(self):
        return lib().nvb_ctx_stream(self.h)

    @property
    def last_paths(self):
        return int(lib().nvb_last_paths(self.h))

    def force_generic(self, on):
        lib().nvb_force_generic(self.h, int(on))

    def fm_batch_sharing(self, min_batch):
        """batches of at least min_batch queries share search prefixes in match() (0xFFFFFFFF = never)"""
        lib().nvb_fm_batch_sharing(self.h, int(min_batch))

    def synchronize(self):
        check(lib().nvb_ctx_synchronize(self.h), self.h)

    def close(self):
        if self.h:
            lib().nvb_ctx_destroy(self.h)
            self.h = C.c_void_p()


def current_stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
