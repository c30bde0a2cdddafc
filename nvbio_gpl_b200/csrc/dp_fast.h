// dp_fast.h -- the packed s16x2 DPX kernel family (dp_fast_full.cu, dp_fast_banded.cu).
// Each kernel scores two pairs per thread in the halves of 32-bit registers with
// VIADDMNMX.S16x2 / VIMNMX3.S16x2; pairs it cannot represent exactly (a symbol > 3, a length over
// the kernel's limits) are appended to worklist (worklist[0] = count, ids from worklist[1]) and left
// for the int32 kernels.
#pragma once
#include "dp_common.cuh"

#define NVB_MAX_BAND_LEN 64

bool dpf_full_score_supported(const nvb_aligner& a, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN);
int  dpf_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, nvb_best_sink* sinks, uint32_t maxM,
                    uint32_t maxN, uint32_t* worklist, void* temp, cudaStream_t s);
// full-matrix SW / ED GLOBAL + SEMI_GLOBAL and Gotoh, sheared coordinates (dp_fast_full2.cu)
bool dpf2_full_score_supported(const nvb_aligner& a, uint32_t maxM, uint32_t maxN);
int  dpf2_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, nvb_best_sink* sinks, uint32_t maxM,
                     uint32_t maxN, uint32_t* worklist, cudaStream_t s, const uint32_t* order = nullptr);
// the order in which the Gotoh GLOBAL / SEMI_GLOBAL sweeps couple the pairs of a batch: by pattern length (context-owned scratch)
bool dpf2_wants_order(const nvb_aligner& a);
uint32_t dpf2_order_entries(uint32_t n, uint32_t maxM);       // entries of the order array (runs padded to whole couples)
int  dpf2_order_by_length(nvb_ctx* ctx, const StrSet& P, uint32_t maxM, const uint32_t** order, cudaStream_t s);
// full-matrix traceback of SW / ED / Gotoh GLOBAL + SEMI_GLOBAL (dp_fast_full2.cu, part 3): one tagged forward pass writing two (Gotoh:
// four) flow bits per cell, then one walk thread per pair; pairs [first, first + n), first even
bool   dpf2_traceback_supported(const nvb_aligner& a, uint32_t maxM, uint32_t maxN);
size_t dpf2_tb_bytes_per_couple(const nvb_aligner& a, uint32_t maxM, uint32_t maxN);
int    dpf2_traceback(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, uint32_t first, uint32_t n, nvb_alignment* out, uint8_t* ops,
                      uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM, uint32_t maxN, uint8_t* flow, uint32_t* worklist, cudaStream_t s,
                      const uint32_t* order = nullptr);
bool dpf_banded_score_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN);
int  dpf_banded_score(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, nvb_best_sink* sinks,
                      uint32_t maxM, uint32_t maxN, uint32_t* worklist, cudaStream_t s);
// sheared-coordinate banded kernels (dp_band2.cu): GLOBAL / SEMI_GLOBAL, bands 7 and 15, 8-bit string sets
bool dpb2_score_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN);
int  dpb2_score(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, nvb_best_sink* sinks,
                uint32_t maxM, uint32_t maxN, uint32_t* worklist, cudaStream_t s);
// sheared-coordinate banded traceback (dp_band2_tb.cu): one tagged forward pass writing the flow of a launch to
// scratch ([row][couple]), then one walk thread per pair; pairs [first, first + n), first even
bool   dpb2_traceback_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM);
size_t dpb2_tb_bytes_per_couple(const nvb_aligner& a, uint32_t maxM);
int    dpb2_traceback(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, uint32_t first, uint32_t n,
                      nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM,
                      uint8_t* flow, uint32_t* worklist, cudaStream_t s);
