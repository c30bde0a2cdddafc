// dp_launch.h -- launchers of the DP kernel families (defined one family per translation unit).
#pragma once
#include "dp_common.cuh"

// int32 generic kernels (dp_generic_full.cu / dp_generic_banded.cu).
// list != NULL: score the pairs list[first .. first+n) (the packed kernels' dirty-pair worklist).
int dpg_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first,
                   uint32_t n, nvb_best_sink* sinks, int16_t* temp, uint32_t maxN, cudaStream_t s, Sink2 s2 = Sink2{nullptr, 0}, const QualDev* qd = nullptr);
int dpg_full_traceback(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n, uint32_t chk,
                       nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16,
                       uint8_t* flow, uint32_t maxM, uint32_t maxN, cudaStream_t s, const QualDev* qd = nullptr);
int dpg_banded_score(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, const uint32_t* list,
                     const uint32_t* d_count, uint32_t first, uint32_t n, nvb_best_sink* sinks, cudaStream_t s, Sink2 s2 = Sink2{nullptr, 0}, const QualDev* qd = nullptr);
int dpg_banded_traceback(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n,
                         uint32_t chk, nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops,
                         int16_t* temp16, uint8_t* flow, uint32_t maxM, cudaStream_t s, const QualDev* qd = nullptr);
