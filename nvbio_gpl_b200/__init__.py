"""nvbio_gpl_b200 -- B200-native (sm_100a) seed-and-extend hot path of NVBIO / nvBowtie.

The product is ``libnvbio_b200.so`` (hand-written CUDA behind the C ABI of ``include/nvbio_b200.h``).
This package is the thin host-side mirror of the reference's operator interface for that path
(``fm_index`` + ``match/rank/rank4/locate``; aligner tag objects + ``batch_alignment_score`` /
``batch_banded_alignment_score`` / batched traceback), used by the tests and ``bench.py``. PyTorch is
only plumbing here (device buffers, streams, torch.distributed). There is no CPU fallback: every
compute call raises if the CUDA library or a CUDA device is missing.
"""
from ._lib import lib, build, LIB_PATH, NvbError, Context  # noqa: F401
from .fmindex import FMIndex, FMIndexFilter, match, match_reverse, match_approx, match_host, match_host_fixed, match_host_fixed2n, pack_symbols, pack_symbols_2n, rank, rank4, locate, locate_host, build_index  # noqa: F401
from .io import load_index, save_index, CLoadedIndex  # noqa: F401
from .alignment import (  # noqa: F401
    EDIT_DISTANCE, SMITH_WATERMAN, GOTOH, GLOBAL, LOCAL, SEMI_GLOBAL,
    Aligner, EditDistanceAligner, SmithWatermanAligner, GotohAligner, QualityScheme, StringSet,
    batch_alignment_score, batch_banded_alignment_score, batch_alignment_score_noalloc, NOALLOC_SCRATCH_BYTES,
    MyersTag, MyersAligner, batch_banded_myers_score, batch_alignment_score_best2, MYERS_BATCH_MIN_SCORE,
    batch_alignment_traceback, batch_banded_alignment_traceback,
    batch_score_host, batch_traceback_host, ops_to_cigar, ops_to_mds, mds_to_md_string, SINK_DT, ALN_DT,
)
from .pipeline import MapParams, map_reads, map_reads_host, map_reads_host_multi, match_host_multi, pack_genome, ops_stride_for, HIT_DT, FLAG_REVERSE, FLAG_ALIGNED, pair_hits, PAIR_DT, PAIR_MATE1, PAIR_MATE2, PAIR_CONCORDANT, map_hits_to_mds, rescue_mates, map_hits_mapq  # noqa: F401
