// dp_generic_banded.cu -- instantiations + launchers of the int32 banded kernels.
// Band lengths 3,5,7,15,31 (the ones nvBowtie picks, score_inl.h:463-507, plus 5) keep the band in
// registers; any other band in 2..64 runs the same code with the band in local memory.
#include "dp_generic_banded.cuh"
#include "dp_launch.h"

template <bool GOTOH, int TYPE, int WT>
static int score_w(nvb_ctx* ctx, Scheme sc, uint32_t band, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first,
                   uint32_t n, nvb_best_sink* sinks, cudaStream_t s, Sink2 s2)
{
    dpg_banded_score_kernel<GOTOH, TYPE, WT><<<(n + 127) / 128, 128, 0, s>>>(sc, band, P, T, list, d_count, first, n, sinks, s2);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
template <bool GOTOH, int TYPE>
static int score_t(nvb_ctx* ctx, Scheme sc, uint32_t band, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first,
                   uint32_t n, nvb_best_sink* sinks, cudaStream_t s, Sink2 s2)
{
    switch (band)
    {
    case 3:  return score_w<GOTOH, TYPE, 3>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2);
    case 5:  return score_w<GOTOH, TYPE, 5>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2);
    case 7:  return score_w<GOTOH, TYPE, 7>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2);
    case 15: return score_w<GOTOH, TYPE, 15>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2);
    case 31: return score_w<GOTOH, TYPE, 31>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2);
    default: return score_w<GOTOH, TYPE, 0>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2);
    }
}

int dpg_banded_score(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, const uint32_t* list,
                     const uint32_t* d_count, uint32_t first, uint32_t n, nvb_best_sink* sinks, cudaStream_t s, Sink2 s2, const QualDev* qd)
{
    if (n == 0) return NVB_OK;
    const Scheme sc = make_scheme_q(a, qd);
    const bool g = a.kind == NVB_GOTOH;
    #define GO(G, TY) return score_t<G, TY>(ctx, sc, band, P, T, list, d_count, first, n, sinks, s, s2)
    switch (a.type)
    {
    case NVB_GLOBAL:      if (g) GO(true, T_GLOBAL); else GO(false, T_GLOBAL);
    case NVB_LOCAL:       if (g) GO(true, T_LOCAL);  else GO(false, T_LOCAL);
    case NVB_SEMI_GLOBAL: if (g) GO(true, T_SEMI);   else GO(false, T_SEMI);
    }
    #undef GO
    return NVB_ERR_ARG;
}

template <bool GOTOH, int TYPE, int WT>
static int tb_w(nvb_ctx* ctx, Scheme sc, uint32_t band, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n, uint32_t chk,
                nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16, uint8_t* flow,
                uint32_t maxM, cudaStream_t s)
{
    dpg_banded_traceback_kernel<GOTOH, TYPE, WT><<<(n + 127) / 128, 128, 0, s>>>(
        sc, band, P, T, list, d_count, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
template <bool GOTOH, int TYPE>
static int tb_t(nvb_ctx* ctx, Scheme sc, uint32_t band, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n, uint32_t chk,
                nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16, uint8_t* flow,
                uint32_t maxM, cudaStream_t s)
{
    #define W_(WT) return tb_w<GOTOH, TYPE, WT>(ctx, sc, band, P, T, list, d_count, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, s)
    switch (band)
    {
    case 3:  W_(3);
    case 5:  W_(5);
    case 7:  W_(7);
    case 15: W_(15);
    case 31: W_(31);
    default: W_(0);
    }
    #undef W_
}

int dpg_banded_traceback(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n,
                         uint32_t chk, nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops,
                         int16_t* temp16, uint8_t* flow, uint32_t maxM, cudaStream_t s, const QualDev* qd)
{
    if (n == 0) return NVB_OK;
    const Scheme sc = make_scheme_q(a, qd);
    const bool g = a.kind == NVB_GOTOH;
    #define GO(G, TY) return tb_t<G, TY>(ctx, sc, band, P, T, list, d_count, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, s)
    switch (a.type)
    {
    case NVB_GLOBAL:      if (g) GO(true, T_GLOBAL); else GO(false, T_GLOBAL);
    case NVB_LOCAL:       if (g) GO(true, T_LOCAL);  else GO(false, T_LOCAL);
    case NVB_SEMI_GLOBAL: if (g) GO(true, T_SEMI);   else GO(false, T_SEMI);
    }
    #undef GO
    return NVB_ERR_ARG;
}
