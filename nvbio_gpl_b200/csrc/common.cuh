// common.cuh -- shared declarations of libnvbio_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include "../../include/nvbio_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libnvbio_b200 is written for sm_100a (B200) only"
#endif

#define NVB_SM_COUNT 148
#define NVB_STAGE_SLOTS 80
// stage slot map (nvb_ctx::d_stage): 0-7 *_host DP/locate entry points and FMIndexFilter | 8-32 the seed-and-extend pipeline
// (map_pipeline.cu S_*) | 33-44 nvb_fm_match_host* (4 buffer sets x 3) | 40-41 are NOT free: they belong to that range; the
// packed-set expansion of dp_api.cu uses 60-61 | 45-54 nvb_map_reads_host (2 sets x 5) | 55 quality-scheme tables (dp_api.cu) | 56 min-score table of nvb_map_hits_mapq (mds.cu) | 57 N list of nvb_fm_match_host_fixed2n | 64-75 pipelined nvb_batch_score_host* (2 sets x 5, temp) | 62-63 per-batch k-mer table (fm_kernels.cu) | 76 coupling order of the Gotoh sweeps (dp_fast_full2.cu)

struct nvb_ctx
{
    int          device;
    cudaStream_t stream;
    cudaError_t  last_err;
    uint64_t     launches;
    uint32_t     last_paths;
    int          force_generic;
    uint32_t     fm_share_min;  // nvb_fm_match batches of at least this many queries share search prefixes through a per-batch k-mer table (fm_kernels.cu)
    // grow-only device/pinned staging owned by the *_host entry points
    void*        d_stage[NVB_STAGE_SLOTS];      // 0..7: *_host entry points, 8..: the seed-and-extend pipeline
    size_t       d_stage_bytes[NVB_STAGE_SLOTS];
    // dirty-pair worklist of the packed DP kernels: [0] = count, [1..] = pair ids
    uint32_t*    d_worklist;
    size_t       d_worklist_cap;
    int          sm_count;
    size_t       l2_bytes;
    int          fm_tma;       // fm_match stages its queries by TMA bulk copy (NVB_FM_TMA, default 1)
    cudaEvent_t  host_ev[12];  // event chain of the pipelined *_host entry points (3 per buffer set)
    cudaStream_t aux[3];       // the *_host entry points pipeline chunks over these: copy-in, copy-out, second compute stream
    cudaEvent_t  tab_ev;       // the per-call k-mer table of the FM host entry points is ready
    uint32_t     dp_first[2];  // offsets[0] of the (pattern, text) sets of the next DP call when a host entry point passes a chunk with absolute offsets; else 0
};

// true the first time it is evaluated AT THIS PLACE for the context's device: cudaFuncSetAttribute is per device and, inside a
// template launcher, per instantiation (the static lives in the lambda of each instantiation)
#include <atomic>
#define NVB_FIRST_ON_DEVICE(ctx) ([&]() -> bool { static std::atomic<unsigned long long> m_{0}; const unsigned long long b_ = 1ull << ((ctx)->device & 63); \
                                                return !(m_.fetch_or(b_) & b_); }())

#define NVB_CUDA(ctx, call)                                                 \
    do { cudaError_t e_ = (call);                                           \
         if (e_ != cudaSuccess) { if (ctx) (ctx)->last_err = e_; return NVB_ERR_CUDA; } } while (0)

#define NVB_LAUNCHED(ctx)                                                   \
    do { (ctx)->launches++;                                                 \
         cudaError_t e_ = cudaPeekAtLastError();                            \
         if (e_ != cudaSuccess) { (ctx)->last_err = cudaGetLastError(); return NVB_ERR_CUDA; } } while (0)

int   nvb_stage(nvb_ctx* ctx, int slot, size_t bytes, void** out);
int   nvb_worklist(nvb_ctx* ctx, size_t entries, uint32_t** out);
int   nvb_noalloc_warmup(nvb_ctx* ctx);   // dp_generic_full.cu: resolves the static scratch array's device address for ctx->device

// ---------------------------------------------------------------------------------------------
// device-side view of the fused nvBWT index
// ---------------------------------------------------------------------------------------------
struct FMView
{
    uint32_t        length, primary;
    uint32_t        L2[5];
    const uint4*    blk;      // 2 uint4 per 64 symbols: words, counters
    const uint32_t* ssa;
    const uint2*    ktab;     // k-mer jump table (optional)
    uint32_t        ktab_k;
    const uint32_t* sa_tab;   // the sampled SA the walks end on: ssa (rows % 16 == 0) or the optional denser one
    uint32_t        sa_mask, sa_shift;
    uint32_t        hot_range; // backward-search steps whose range is wider than this load their blocks through L1
};

static inline FMView make_view(const nvb_fmindex* f)
{
    FMView v;
    v.length = f->length; v.primary = f->primary;
    for (int i = 0; i < 5; ++i) v.L2[i] = f->L2[i];
    v.blk = (const uint4*)f->bwt_occ; v.ssa = f->ssa;
    v.ktab = (const uint2*)f->ktab; v.ktab_k = f->ktab ? f->ktab_k : 0u;
    v.sa_tab = f->ssa; v.sa_mask = 15u; v.sa_shift = 4u;
    {
        // ranges wider than length / 4^k belong to the first k steps of every query: about 2.7 * 4^k distinct blocks
        static int shift = -1;
        if (shift < 0) { const char* e = getenv("NVB_FM_HOT_SHIFT"); shift = e ? atoi(e) : 10; }
        v.hot_range = shift >= 32 ? 0xFFFFFFFFu : (f->length >> shift);
        if (v.hot_range < 64u) v.hot_range = 64u;
    }
    if (f->dssa && (f->dssa_int == 2u || f->dssa_int == 4u || f->dssa_int == 8u))
    { v.sa_tab = f->dssa; v.sa_mask = f->dssa_int - 1u; v.sa_shift = f->dssa_int == 2u ? 1u : f->dssa_int == 4u ? 2u : 3u; }
    return v;
}

// per-batch k-mer table (fm_kernels.cu): replaces f.ktab by a table built in context scratch for a batch of n strings
// of at least max_k symbols each, when that is deeper than the caller's
int nvb_fm_batch_table(nvb_ctx* ctx, FMView& f, uint64_t n, uint32_t max_k, cudaStream_t s);

// ranges of the listed (query, last-N position) entries of a 2-bit query chunk, recomputed with the N rule (fm_kernels.cu)
int nvb_fm_fix_n(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, const uint32_t* list_d, uint32_t n_list, uint32_t first_query,
                 uint32_t* ranges, cudaStream_t s);

// the host entry points cut one batch into chunks: the table is built once for the whole batch and handed to the chunks
// as the k-mer table of a local copy of the caller's descriptor
static inline int nvb_with_batch_table(nvb_ctx* ctx, const nvb_fmindex* fmi, uint64_t n, uint32_t max_k, cudaStream_t s, nvb_fmindex* local)
{
    *local = *fmi;
    // (a descriptor the entry points would refuse is left to them)
    if (n < ctx->fm_share_min || !fmi->bwt_occ || fmi->occ_int != 64u || (((uintptr_t)fmi->bwt_occ) & 31u)) return NVB_OK;
    FMView v = make_view(fmi);
    const int rc = nvb_fm_batch_table(ctx, v, n, max_k, s);
    if (rc) return rc;
    local->ktab = (const uint32_t*)v.ktab; local->ktab_k = v.ktab_k;
    return NVB_OK;
}

// string-set view passed by value to kernels
struct StrSet
{
    const uint8_t*  sym;
    const uint32_t* off;
    uint32_t        count;
    uint32_t        bits;
};
static inline StrSet make_set(nvb_string_set s)
{
    StrSet r; r.sym = (const uint8_t*)s.symbols; r.off = s.offsets; r.count = s.count; r.bits = s.bits;
    return r;
}

#ifdef __CUDACC__
// symbol i of a string set's symbol stream (absolute symbol index)
template <int BITS>
__device__ __forceinline__ uint32_t sym_at(const uint8_t* base, uint32_t i)
{
    if (BITS == 8) return base[i];
    const uint32_t* w = (const uint32_t*)base;
    if (BITS == 4) return (__ldg(w + (i >> 3)) >> (28u - 4u * (i & 7u))) & 15u;
    return (__ldg(w + (i >> 4)) >> (30u - 2u * (i & 15u))) & 3u;
}
__device__ __forceinline__ uint32_t sym_at_rt(const uint8_t* base, uint32_t bits, uint32_t i)
{
    if (bits == 8) return base[i];
    if (bits == 4) return sym_at<4>(base, i);
    return sym_at<2>(base, i);
}
#endif
