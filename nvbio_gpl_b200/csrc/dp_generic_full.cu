// dp_generic_full.cu -- instantiations + launchers of the int32 full-matrix kernels.
#include "dp_generic_full.cuh"
#include "dp_launch.h"

template <bool GOTOH, int TYPE>
static int score_t(nvb_ctx* ctx, Scheme sc, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n,
                   nvb_best_sink* sinks, int16_t* temp, uint32_t maxN, cudaStream_t s, Sink2 s2)
{
    if (sc.qm) dpg_full_score_kernel<GOTOH, TYPE, true><<<(n + DPG_THREADS - 1) / DPG_THREADS, DPG_THREADS, 0, s>>>(
        sc, P, T, list, d_count, first, n, sinks, temp, maxN, s2);
    else dpg_full_score_kernel<GOTOH, TYPE, false><<<(n + DPG_THREADS - 1) / DPG_THREADS, DPG_THREADS, 0, s>>>(
        sc, P, T, list, d_count, first, n, sinks, temp, maxN, s2);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

int dpg_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first,
                   uint32_t n, nvb_best_sink* sinks, int16_t* temp, uint32_t maxN, cudaStream_t s, Sink2 s2, const QualDev* qd)
{
    if (n == 0) return NVB_OK;
    const Scheme sc = make_scheme_q(a, qd);
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
static int tb_t(nvb_ctx* ctx, Scheme sc, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n, uint32_t chk,
                nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16,
                uint8_t* flow, uint32_t maxM, uint32_t maxN, cudaStream_t s)
{
    if (sc.qm) dpg_full_traceback_kernel<GOTOH, TYPE, true><<<(n + DPG_THREADS - 1) / DPG_THREADS, DPG_THREADS, 0, s>>>(
        sc, P, T, list, d_count, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, maxN);
    else dpg_full_traceback_kernel<GOTOH, TYPE, false><<<(n + DPG_THREADS - 1) / DPG_THREADS, DPG_THREADS, 0, s>>>(
        sc, P, T, list, d_count, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, maxN);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

int dpg_full_traceback(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, const uint32_t* list, const uint32_t* d_count, uint32_t first, uint32_t n, uint32_t chk,
                       nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int16_t* temp16,
                       uint8_t* flow, uint32_t maxM, uint32_t maxN, cudaStream_t s, const QualDev* qd)
{
    if (n == 0) return NVB_OK;
    const Scheme sc = make_scheme_q(a, qd);
    const bool g = a.kind == NVB_GOTOH;
    #define GO(G, TY) return tb_t<G, TY>(ctx, sc, P, T, list, d_count, first, n, chk, out, ops, ops_stride, n_ops, temp16, flow, maxM, maxN, s)
    switch (a.type)
    {
    case NVB_GLOBAL:      if (g) GO(true, T_GLOBAL); else GO(false, T_GLOBAL);
    case NVB_LOCAL:       if (g) GO(true, T_LOCAL);  else GO(false, T_LOCAL);
    case NVB_SEMI_GLOBAL: if (g) GO(true, T_SEMI);   else GO(false, T_SEMI);
    }
    #undef GO
    return NVB_ERR_ARG;
}

// ---------------------------------------------------------------------------------------------------------------
// nvb_batch_score_noalloc -- full-matrix scoring that allocates NOTHING (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops,
// harnesses that forbid it). The score columns (one int16 column of max_text_len cells per pair, two for Gotoh) live in
// a file-scope device array of NVB_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit. It runs the
// int32 one-pair-per-thread kernels above (dpg_full_score_kernel), which initialise their columns themselves, so nothing
// in the array has to start at zero; every call shares the array, so calls that overlap on the GPU (different streams,
// different host threads) race on it.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[NVB_NOALLOC_SCRATCH_BYTES];

#include <mutex>
// In host code g_noalloc_scratch is only a stand-in for the symbol: its device address is asked for once per device, with
// THAT device current (the context's device, whatever the caller's current device is), and cached for the life of the
// process (cudaDeviceReset invalidates it, like every other device pointer the process holds). The first use per device
// must happen outside CUDA graph capture: nvb_ctx_create does it (nvb_noalloc_warmup), so a context created before
// capture begins is enough.
static int noalloc_scratch_ptr(nvb_ctx* ctx, void** out)
{
    enum { MAX_DEVICES = 64 };
    static void* ptrs[MAX_DEVICES] = { nullptr };
    static std::mutex mtx;
    const int dev = ctx->device;
    if (dev < 0 || dev >= MAX_DEVICES) return NVB_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(mtx);
    if (!ptrs[dev])
    {
        int cur = -1;
        NVB_CUDA(ctx, cudaGetDevice(&cur));
        if (cur != dev) NVB_CUDA(ctx, cudaSetDevice(dev));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (cur != dev && cur >= 0) cudaSetDevice(cur);
        if (e != cudaSuccess) { ctx->last_err = e; return NVB_ERR_CUDA; }
        if (!p) return NVB_ERR_CUDA;
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return NVB_OK;
}

// resolves the static array's address for ctx's device (called by nvb_ctx_create; harmless to call again)
int nvb_noalloc_warmup(nvb_ctx* ctx) { void* p = nullptr; return noalloc_scratch_ptr(ctx, &p); }

// saturating: a product that does not fit size_t reads as SIZE_MAX (> any limit)
extern "C" size_t nvb_batch_score_noalloc_bytes(nvb_aligner a, uint32_t max_text_len, uint32_t n)
{
    const unsigned __int128 b = (unsigned __int128)(a.kind == NVB_GOTOH ? 2u : 1u) * max_text_len * n * sizeof(int16_t);
    return b > (unsigned __int128)SIZE_MAX ? SIZE_MAX : (size_t)b;
}

extern "C" int nvb_batch_score_noalloc(nvb_ctx* ctx, nvb_aligner a, nvb_string_set patterns, nvb_string_set texts,
                                       nvb_best_sink* sinks, uint32_t max_pattern_len, uint32_t max_text_len, void* stream)
{
    (void)max_pattern_len;
    if (!ctx) return NVB_ERR_ARG;
    if (a.kind < NVB_EDIT_DISTANCE || a.kind > NVB_GOTOH || a.type < NVB_GLOBAL || a.type > NVB_SEMI_GLOBAL) return NVB_ERR_ARG;
    if (patterns.count != texts.count) return NVB_ERR_ARG;
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!sinks || !patterns.symbols || !patterns.offsets || !texts.symbols || !texts.offsets) return NVB_ERR_ARG;
    if (!(patterns.bits == 8 || patterns.bits == 4 || patterns.bits == 2) || !(texts.bits == 8 || texts.bits == 4 || texts.bits == 2)) return NVB_ERR_ARG;
    if (nvb_batch_score_noalloc_bytes(a, max_text_len, n) > (size_t)NVB_NOALLOC_SCRATCH_BYTES) return NVB_ERR_TOO_LARGE;
    void* scratch = nullptr;
    int rc = noalloc_scratch_ptr(ctx, &scratch);
    if (rc) return rc;
    ctx->last_paths = 2u;
    return dpg_full_score(ctx, a, make_set(patterns), make_set(texts), nullptr, nullptr, 0, n, sinks, (int16_t*)scratch, max_text_len,
                          (cudaStream_t)stream);
}
