// dp_generic_full.cu -- instantiations + launchers of the int32 full-matrix kernels.
#include "dp_generic_full.cuh"
#include "dp_launch.h"

template <bool GOTOH, int TYPE>
static int score_t(nvb_ctx* ctx, Scheme sc, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n,
                   nvb_best_sink* sinks, int16_t* temp, uint32_t maxN, cudaStream_t s, Sink2 s2)
{
    dpg_full_score_kernel<GOTOH, TYPE><<<(n + DPG_THREADS - 1) / DPG_THREADS, DPG_THREADS, 0, s>>>(
        sc, P, T, list, d_count, first, n, sinks, temp, maxN, s2);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

int dpg_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first,
                   uint32_t n, nvb_best_sink* sinks, int16_t* temp, uint32_t maxN, cudaStream_t s, Sink2 s2)
{
    if (n == 0) return NVB_OK;
    const Scheme sc = make_scheme(a);
    const bool g = a.kind == NVB_GOTOH;
    #define GO(G, TY) return score_t<G, TY>(ctx, sc, P, T, list, d_count, first, n, sinks, temp, maxN, s, s2)
    switch (a.type)
    {
    case NVB_GLOBAL:      if (g) GO(true, T_GLOBAL); else GO(false, T_GLOBAL);
    case NVB_LOCAL:       if (g) GO(true, T_LOCAL);  else GO(false, T_LOCAL);
    case NVB_SEMI_GLOBAL: if (g) GO(true, T_SEMI);   else GO(false, T_SEMI);
    }
    #undef GO
    return NVB_ERR_ARG;
}

template <bool GOTOH, int TYPE>
static int tb_t(nvb_ctx* ctx, Scheme sc, StrSet P, StrSet T, uint32_t first, uint32_t n, uint32_t chk,
                nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16,
                uint8_t* flow, uint32_t maxM, uint32_t maxN, cudaStream_t s)
{
    dpg_full_traceback_kernel<GOTOH, TYPE><<<(n + DPG_THREADS - 1) / DPG_THREADS, DPG_THREADS, 0, s>>>(
        sc, P, T, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, maxN);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

int dpg_full_traceback(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, uint32_t first, uint32_t n, uint32_t chk,
                       nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16,
                       uint8_t* flow, uint32_t maxM, uint32_t maxN, cudaStream_t s)
{
    if (n == 0) return NVB_OK;
    const Scheme sc = make_scheme(a);
    const bool g = a.kind == NVB_GOTOH;
    #define GO(G, TY) return tb_t<G, TY>(ctx, sc, P, T, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, maxN, s)
    switch (a.type)
    {
    case NVB_GLOBAL:      if (g) GO(true, T_GLOBAL); else GO(false, T_GLOBAL);
    case NVB_LOCAL:       if (g) GO(true, T_LOCAL);  else GO(false, T_LOCAL);
    case NVB_SEMI_GLOBAL: if (g) GO(true, T_SEMI);   else GO(false, T_SEMI);
    }
    #undef GO
    return NVB_ERR_ARG;
}
