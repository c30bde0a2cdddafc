"""nvBWT on-disk index files (the host-side mirror of io::FMIndexDataHost::load / the nvBWT writer):
<prefix>.bwt|.rbwt = uint32 primary, uint32 cumFreq[4], BWT words;  <prefix>.sa|.rsa = uint32 primary, 4 ignored words,
uint32 SA_INT, uint32 length, ssa[1..]  (nvBWT/nvBWT.cu:303-340, nvbio/io/fmindex/fmindex_impl.cu:111-321).
load_index() reads the files with numpy and fuses on the device; load_index_c()/save_index() go through the
C-ABI entry points nvb_fmindex_load / nvb_fmindex_save, which is what a C++ host would call."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, Context, FMIndexC
from .fmindex import FMIndex, fuse, SA_INT, _words_for


def load_index(prefix, device, reverse=False, ctx=None):
    ext_b, ext_s = (".rbwt", ".rsa") if reverse else (".bwt", ".sa")
    raw = np.fromfile(prefix + ext_b, dtype=np.uint32)
    primary, cum = int(raw[0]), [int(v) for v in raw[1:5]]
    length = cum[3]
    W = _words_for(length)
    words = np.zeros(W, dtype=np.uint32)
    n = (length + 15) // 16
    if len(raw) - 5 < n:
        raise ValueError("truncated bwt file %s" % (prefix + ext_b))
    words[:n] = raw[5:5 + n]
    fused, L2 = fuse(torch.from_numpy(words.view(np.int32)).to(device), length, ctx=ctx)
    if L2[1:] != cum:
        raise ValueError("cumulative frequencies of %s do not match its BWT" % (prefix + ext_b))
    ssa = None
    try:
        sraw = np.fromfile(prefix + ext_s, dtype=np.uint32)
        if int(sraw[0]) == primary and int(sraw[5]) == SA_INT and int(sraw[6]) == length:
            n_ssa = (length + SA_INT) // SA_INT
            s = np.full(n_ssa, 0xFFFFFFFF, dtype=np.uint32)
            s[1:] = sraw[7:7 + n_ssa - 1]
            ssa = torch.from_numpy(s.view(np.int32)).to(device)
    except (OSError, IndexError):
        ssa = None
    return FMIndex(length, primary, L2, fused, ssa)


def save_index(fmi, prefix, reverse=False, ctx=None):
    ctx = ctx or Context.default(fmi.device.index)
    f = fmi.c()
    check(lib().nvb_fmindex_save(ctx.h, C.byref(f), prefix.encode(), 1 if reverse else 0), ctx.h)


class CLoadedIndex:
    """an index loaded by nvb_fmindex_load: device memory owned by the library (released on close())"""

    def __init__(self, prefix, device_index=0, reverse=False, ctx=None):
        self.ctx = ctx or Context.default(device_index)
        self.f = FMIndexC()
        check(lib().nvb_fmindex_load(self.ctx.h, prefix.encode(), 1 if reverse else 0, C.byref(self.f)), self.ctx.h)

    def match(self, symbols, offsets, bits=8, reverse=False):
        from .fmindex import _set_c
        n = offsets.numel() - 1
        ranges = torch.empty((n, 2), dtype=torch.int32, device=offsets.device)
        fn = lib().nvb_fm_match_reverse if reverse else lib().nvb_fm_match
        check(fn(self.ctx.h, C.byref(self.f), _set_c(symbols, offsets, bits), ranges.data_ptr(), _lib.current_stream_ptr()), self.ctx.h)
        return ranges

    def locate(self, rows):
        pos = torch.empty(rows.numel(), dtype=torch.int32, device=rows.device)
        check(lib().nvb_fm_locate(self.ctx.h, C.byref(self.f), rows.data_ptr(), rows.numel(), pos.data_ptr(), _lib.current_stream_ptr()), self.ctx.h)
        return pos

    def close(self):
        if self.f.bwt_occ:
            torch.cuda.synchronize()
            check(lib().nvb_fmindex_release(self.ctx.h, C.byref(self.f)), self.ctx.h)
