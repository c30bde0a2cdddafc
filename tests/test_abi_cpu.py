"""CPU-side checks of the product: the C-ABI library loads and exports every symbol include/nvbio_b200.h
declares, compute entry points fail loudly without a GPU (no CPU fallback), and the synthetic index
builder (torch, device-agnostic, not on the hot path) produces the reference's nvBWT image."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import nvbio_gpl_b200 as nvb
from nvbio_gpl_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "nvbio_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nvb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = C.CDLL(nvb.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    for name in names:
        assert hasattr(L, name), name
    assert set(names) == set(_lib.EXPORTS), set(names) ^ set(_lib.EXPORTS)
    assert nvb.lib().nvb_version() == 100


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    h = C.c_void_p()
    assert nvb.lib().nvb_ctx_create(0, C.byref(h)) == -2      # NVB_ERR_CUDA
    text = torch.randint(0, 4, (1000,), dtype=torch.uint8)
    fmi = nvb.build_index(text)
    off = torch.tensor([0, 4], dtype=torch.int32)
    with pytest.raises(nvb.NvbError):
        nvb.match(fmi, text[:4].clone(), off)


def test_temp_bytes_and_words():
    L = nvb.lib()
    assert L.nvb_bwt_occ_words(64) == 8 and L.nvb_bwt_occ_words(65) == 16 and L.nvb_bwt_occ_words(0) == 0
    assert L.nvb_ssa_words(15) == 1 and L.nvb_ssa_words(16) == 2
    a = nvb.GotohAligner(nvb.LOCAL, 2, -1, -2, -1).c()
    assert L.nvb_temp_bytes(0, a, 0, 0, 150, 500, 128) == 500 * 4 * 128 + 256
    assert L.nvb_temp_bytes(1, a, 15, 0, 150, 500, 128) == 256


@pytest.mark.parametrize("n,alpha", [(1, 4), (2, 4), (63, 4), (64, 4), (65, 4), (1000, 2), (4096, 4), (50000, 4), (3000, 1)])
def test_build_index_matches_oracle(oracle, n, alpha):
    rng = np.random.default_rng(n + alpha)
    text = rng.integers(0, alpha, n).astype(np.uint8)
    ref = oracle.build_index(text, True)
    fmi, sa = nvb.build_index(torch.from_numpy(text), return_sa=True)
    bwt_occ, ssa = fmi.host_arrays()
    assert fmi.primary == ref.primary and fmi.L2 == [int(v) for v in ref.L2]
    assert (bwt_occ == ref.bwt_occ).all()
    assert (ssa == ref.ssa).all()
    assert (sa.numpy()[1:] == ref.sa[1:]).all()
