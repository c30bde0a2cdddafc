// dp_api.cu -- C-ABI entry points of the alignment half: argument checks, scratch carving, choice
// between the packed DPX kernels (dp_fast_*.cu) and the int32 generic kernels, CIGAR packing.
#include "dp_launch.h"
#include "dp_fast.h"

static int check_aligner(const nvb_aligner& a)
{
    if (a.kind < NVB_EDIT_DISTANCE || a.kind > NVB_GOTOH) return NVB_ERR_ARG;
    if (a.type < NVB_GLOBAL || a.type > NVB_SEMI_GLOBAL) return NVB_ERR_ARG;
    return NVB_OK;
}
static int check_sets(const nvb_string_set& P, const nvb_string_set& T)
{
    if (P.count != T.count) return NVB_ERR_ARG;
    if (P.count == 0) return NVB_OK;
    if (!P.symbols || !P.offsets || !T.symbols || !T.offsets) return NVB_ERR_ARG;
    if (!(P.bits == 8 || P.bits == 4 || P.bits == 2) || !(T.bits == 8 || T.bits == 4 || T.bits == 2)) return NVB_ERR_ARG;
    return NVB_OK;
}

struct TempPlan { size_t i16_per_pair, flow_per_pair; };

static TempPlan plan(int op, const nvb_aligner& a, uint32_t band, uint32_t chk, uint32_t maxM, uint32_t maxN)
{
    TempPlan p = { 0, 0 };
    const bool g = a.kind == NVB_GOTOH;
    switch (op)
    {
    case NVB_OP_SCORE:           p.i16_per_pair = (size_t)maxN * (g ? 2 : 1); break;
    case NVB_OP_BANDED_SCORE:    break;
    case NVB_OP_TRACEBACK:
    {
        const size_t nchk = chk ? (maxM + chk - 1) / chk + 1 : 1;
        p.i16_per_pair = 2 * (size_t)maxN + 2 * nchk * maxN;
        p.flow_per_pair = (size_t)maxN * chk;
        break;
    }
    case NVB_OP_BANDED_TRACEBACK:
    {
        const size_t nrows = chk ? 2 + (maxM + chk - 1) / chk : 2;
        p.i16_per_pair = 2 * nrows * band;
        p.flow_per_pair = (size_t)band * chk;
        break;
    }
    }
    return p;
}

extern "C" size_t nvb_temp_bytes(int op, nvb_aligner a, uint32_t band, uint32_t chk, uint32_t maxM, uint32_t maxN, uint32_t n)
{
    const TempPlan p = plan(op, a, band, chk, maxM, maxN);
    const size_t n128 = ((size_t)n + 127) & ~(size_t)127;
    return (p.i16_per_pair * 2 + p.flow_per_pair) * n128 + 256;
}

// pairs per launch that fit the scratch (multiple of 128 unless everything fits)
static uint32_t batch_pairs(const TempPlan& p, size_t temp_bytes, uint32_t n)
{
    const size_t per = p.i16_per_pair * 2 + p.flow_per_pair;
    if (per == 0) return n;
    if (temp_bytes < 256) return 0;
    size_t fit = (temp_bytes - 256) / per;
    // the tracebacks place their flow region behind the int16 region of the batch rounded up to 128 pairs: "everything
    // fits" must hold for that rounded layout too, or a buffer sized per * n + 256 with n % 128 != 0 would be overrun
    if (fit >= n)
    {
        const size_t n128 = ((size_t)n + 127) & ~(size_t)127;
        if (p.flow_per_pair == 0 || p.i16_per_pair * 2 * n128 + p.flow_per_pair * (size_t)n + 256 <= temp_bytes) return n;
    }
    fit &= ~(size_t)127;
    if (fit > n) fit = n & ~(size_t)127;
    return (uint32_t)fit;
}

// ---------------------------------------------------------------------------------------------------------------
// Packed (4 / 2 bit) string sets -- the layout of the reference's read and genome streams -- are expanded to one byte per
// symbol into context-owned scratch so that the banded fast kernels (which stream 16-byte blocks of 8-bit symbols) take
// them too; the symbol index space, hence the offsets, is unchanged. max_len must be a true maximum (the reference sizes
// its own scratch from it as well): the expansion covers count * max_len symbols.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
dp_unpack_kernel(const uint32_t* __restrict__ words, uint32_t bits, const uint32_t* __restrict__ total_d, uint32_t first16, uint64_t cap, uint8_t* __restrict__ out)
{
    const uint64_t g = (uint64_t)blockIdx.x * 256u + threadIdx.x;        // one group of 16 symbols per thread
    const uint64_t total = min((uint64_t)*total_d, first16 + cap);
    const uint64_t i0 = first16 + g * 16u;                               // absolute symbol index of the group
    if (i0 >= total) return;
    const uint64_t gw = i0 >> 4;
    uint32_t b[4];
    if (bits == 4)
    {
        const uint32_t w0 = __ldg(words + 2 * gw), w1 = (i0 + 8u < total) ? __ldg(words + 2 * gw + 1) : 0u;
        #pragma unroll
        for (int k = 0; k < 4; ++k)
        {
            const uint32_t w = k < 2 ? w0 : w1; const int base = 28 - 16 * (k & 1);
            b[k] = ((w >> base) & 15u) | (((w >> (base - 4)) & 15u) << 8) | (((w >> (base - 8)) & 15u) << 16) | (((w >> (base - 12)) & 15u) << 24);
        }
    }
    else
    {
        const uint32_t w = __ldg(words + gw);
        #pragma unroll
        for (int k = 0; k < 4; ++k)
        {
            const int base = 30 - 8 * k;
            b[k] = ((w >> base) & 3u) | (((w >> (base - 2)) & 3u) << 8) | (((w >> (base - 4)) & 3u) << 16) | (((w >> (base - 6)) & 3u) << 24);
        }
    }
    *(uint4*)(out + g * 16u) = make_uint4(b[0], b[1], b[2], b[3]);       // the scratch is padded to whole groups
}

enum { DP_STAGE_UNPACK_P = 60, DP_STAGE_UNPACK_T = 61 };   // slot map in common.cuh (40/41 collided with nvb_fm_match_host's 4th buffer set)

// `first`: symbol index of the set's first symbol (offsets[0]); 0 for a caller's set, the chunk's first symbol when a host entry
// point hands over a chunk that keeps the absolute offsets of the caller's array (ctx->dp_first). The expansion covers the
// count * max_len symbols from there and the returned pointer is rebased so that absolute indices keep working.
static int unpack_set(nvb_ctx* ctx, StrSet& S, uint32_t max_len, int slot, cudaStream_t s, uint32_t first)
{
    if (S.bits == 8) return NVB_OK;
    const uint32_t first16 = first & ~15u;
    const uint64_t cap = (uint64_t)S.count * max_len + (first - first16);
    if (cap == 0 || cap > 0xFFFFFFF0ull) return NVB_ERR_UNSUPPORTED;
    void* out = nullptr;
    int rc = nvb_stage(ctx, slot, (size_t)((cap + 15u) & ~15ull) + 128, &out);
    if (rc) return rc;
    const uint64_t groups = (cap + 15u) / 16u;
    dp_unpack_kernel<<<(uint32_t)((groups + 255u) / 256u), 256, 0, s>>>((const uint32_t*)S.sym, S.bits, S.off + S.count, first16, cap, (uint8_t*)out + 64);
    NVB_LAUNCHED(ctx);
    S.sym = (const uint8_t*)out + 64 - first16; S.bits = 8;             // 64 spare bytes in front: aligned block fetches of the first string
    return NVB_OK;
}

// expands packed sets when that opens a fast path the 8-bit form would take
template <typename Supported>
static int unpack_for_fast_path(nvb_ctx* ctx, StrSet& P, StrSet& T, uint32_t maxM, uint32_t maxN, cudaStream_t s, Supported supported)
{
    if ((P.bits == 8 && T.bits == 8) || ctx->force_generic == 1) return NVB_OK;
    StrSet P8 = P, T8 = T; P8.bits = T8.bits = 8;
    if (!supported(P8, T8)) return NVB_OK;
    if ((uint64_t)P.count * maxM > 0xFFFFFFF0ull || (uint64_t)T.count * maxN > 0xFFFFFFF0ull) return NVB_OK;
    int rc;
    if ((rc = unpack_set(ctx, P, maxM, DP_STAGE_UNPACK_P, s, ctx->dp_first[0])) || (rc = unpack_set(ctx, T, maxN, DP_STAGE_UNPACK_T, s, ctx->dp_first[1]))) return rc;
    return NVB_OK;
}


// ---------------------------------------------------------------------------------------------------------------
// Quality-aware scoring schemes (nvbio/alignment/alignment.h:411-436,488-505; nvBowtie's is bowtie2/cuda/scoring.h:206-301): the
// host tables go to a small context-owned device buffer, the calls then run the int32 kernels, which read match(q) /
// mismatch(a, b, q) at the quality of each pattern position.
// ---------------------------------------------------------------------------------------------------------------
enum { DP_STAGE_QTABLES = 55 };   // slot map in common.cuh
static int quality_tables(nvb_ctx* ctx, const nvb_aligner& a, const nvb_quality_scheme* qs, const uint8_t* quals, cudaStream_t s, QualDev* qd, const QualDev** out)
{
    *out = nullptr;
    if (!qs) return quals ? NVB_ERR_ARG : NVB_OK;
    if (!qs->match_h || !qs->mismatch_h) return NVB_ERR_ARG;
    if (a.kind == NVB_EDIT_DISTANCE) return NVB_OK;                       // EditDistanceAligner has no scheme: qualities do not enter
    void* d = nullptr;
    int rc = nvb_stage(ctx, DP_STAGE_QTABLES, 512 * sizeof(int32_t), &d);
    if (rc) return rc;
    NVB_CUDA(ctx, cudaMemcpyAsync(d, qs->match_h, 256 * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    NVB_CUDA(ctx, cudaMemcpyAsync((int32_t*)d + 256, qs->mismatch_h, 256 * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    qd->qm = (const int32_t*)d; qd->qx = (const int32_t*)d + 256; qd->quals = quals;
    qd->tga = qs->text_gap_open; qd->tgb = qs->text_gap_ext;
    *out = qd;
    return NVB_OK;
}

static int batch_score_impl(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, const uint8_t* quals, nvb_string_set patterns, nvb_string_set texts,
                               nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc;
    if ((rc = check_aligner(a)) || (rc = check_sets(patterns, texts))) return rc;
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!sinks) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const StrSet P = make_set(patterns), T = make_set(texts);
    const TempPlan p = plan(NVB_OP_SCORE, a, 0, 0, maxM, maxN);
    ctx->last_paths = 0;
    QualDev qdev; const QualDev* qd = nullptr;
    if ((rc = quality_tables(ctx, a, qs, quals, s, &qdev, &qd))) return rc;

    const bool fast_local = dpf_full_score_supported(a, P, T, maxM, maxN);
    const bool fast_sheared = !fast_local && ctx->force_generic == 0 && dpf2_full_score_supported(a, maxM, maxN);
    const bool use_fast = !qd;                                                   // quality-aware schemes: int32 kernels   // force_generic 2: no sheared-coordinate kernels
    if (use_fast && ctx->force_generic != 1 && temp && batch_pairs(p, temp_bytes, n) >= n && (fast_local || fast_sheared))
    {
        // packed DPX kernel; pairs it cannot represent exactly are appended to the ctx worklist and
        // re-scored by the int32 kernel, which reads the count on the device (no host sync)
        uint32_t* wl = nullptr;
        if ((rc = nvb_worklist(ctx, n, &wl))) return rc;
        NVB_CUDA(ctx, cudaMemsetAsync(wl, 0, sizeof(uint32_t), s));
        if (fast_local) { if ((rc = dpf_full_score(ctx, a, P, T, sinks, maxM, maxN, wl, temp, s))) return rc; }
        else
        {
            const uint32_t* order = nullptr;
            if (dpf2_wants_order(a) && (rc = dpf2_order_by_length(ctx, P, maxM, &order, s))) return rc;
            if ((rc = dpf2_full_score(ctx, a, P, T, sinks, maxM, maxN, wl, s, order))) return rc;
            ctx->last_paths |= 4u;
        }
        ctx->last_paths |= 1u;
        if (!temp && p.i16_per_pair) return NVB_ERR_TEMP;
        uint32_t n_dirty = n;
        if (batch_pairs(p, temp_bytes, n) < n)
        {
            // scratch smaller than n columns: read the count back and batch
            NVB_CUDA(ctx, cudaMemcpyAsync(&n_dirty, wl, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
            NVB_CUDA(ctx, cudaStreamSynchronize(s));
            if (n_dirty == 0) return NVB_OK;
        }
        const uint32_t bp = batch_pairs(p, temp_bytes, n_dirty);
        if (bp == 0) return NVB_ERR_TEMP;
        for (uint32_t first = 0; first < n_dirty; first += bp)
        {
            const uint32_t cnt = n_dirty - first < bp ? n_dirty - first : bp;
            if ((rc = dpg_full_score(ctx, a, P, T, wl + 1, wl, first, cnt, sinks, (int16_t*)temp, maxN, s))) return rc;
        }
        return NVB_OK;
    }
    ctx->last_paths |= 2u;
    if (!temp && p.i16_per_pair) return NVB_ERR_TEMP;
    const uint32_t bp = batch_pairs(p, temp_bytes, n);
    if (bp == 0) return NVB_ERR_TEMP;
    for (uint32_t first = 0; first < n; first += bp)
    {
        const uint32_t cnt = n - first < bp ? n - first : bp;
        if ((rc = dpg_full_score(ctx, a, P, T, nullptr, nullptr, first, cnt, sinks, (int16_t*)temp, maxN, s, Sink2{nullptr, 0}, qd))) return rc;
    }
    return NVB_OK;
}

extern "C" int nvb_batch_score(nvb_ctx* ctx, nvb_aligner a, nvb_string_set patterns, nvb_string_set texts,
                               nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{ return batch_score_impl(ctx, a, nullptr, nullptr, patterns, texts, sinks, maxM, maxN, temp, temp_bytes, stream); }

// TextBlockingTag scoring (alignment.h:266-279; sw/sw_inl.h:872+, gotoh/gotoh_inl.h:848+; scoring only in the reference too): stripes of
// 8 text columns, pattern row by pattern row. Scores equal those of the default blocking; what changes is the order in which a LOCAL
// sink meets equal scores -- and that is the default order of the TRANSPOSED problem (aln::transpose, alignment.h:395-540: pattern
// and text exchanged, deletion and insertion exchanged). So LOCAL and GLOBAL run the transposed batch through the same kernels and swap
// the sink coordinates back; SEMI_GLOBAL (not symmetric) walks the last row in text order under both blockings.
__global__ void dp_swap_sinks_kernel(nvb_best_sink* sinks, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t x = sinks[i].sink_x; sinks[i].sink_x = sinks[i].sink_y; sinks[i].sink_y = x;
}
// SEMI_GLOBAL, empty pattern, non-empty text: the text-blocking sweep ends on the boundary cell of the last row, (N, 0) with score 0
__global__ void dp_tb_semi_empty_kernel(nvb_best_sink* sinks, const uint32_t* poff, const uint32_t* toff, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t N = toff[i + 1] - toff[i];
    if (poff[i + 1] == poff[i] && N > 0) { nvb_best_sink s; s.score = 0; s.sink_x = N; s.sink_y = 0; sinks[i] = s; }
}
extern "C" int nvb_batch_score_text_blocking(nvb_ctx* ctx, nvb_aligner a, nvb_string_set patterns, nvb_string_set texts,
                                             nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{
    const uint32_t n = patterns.count;
    if (a.type == NVB_SEMI_GLOBAL)
    {
        const int rc = batch_score_impl(ctx, a, nullptr, nullptr, patterns, texts, sinks, maxM, maxN, temp, temp_bytes, stream);
        if (rc || n == 0) return rc;
        dp_tb_semi_empty_kernel<<<(n + 255u) / 256u, 256, 0, (cudaStream_t)stream>>>(sinks, patterns.offsets, texts.offsets, n);
        NVB_LAUNCHED(ctx);
        return NVB_OK;
    }
    // GLOBAL and LOCAL are symmetric: the transposed batch (degenerate shapes included: an empty TEXT reports nothing here)
    nvb_aligner t = a;
    if (a.kind == NVB_SMITH_WATERMAN) { t.gap_a = a.gap_b; t.gap_b = a.gap_a; }
    const int rc = batch_score_impl(ctx, t, nullptr, nullptr, texts, patterns, sinks, maxN, maxM, temp, temp_bytes, stream);
    if (rc || n == 0) return rc;
    dp_swap_sinks_kernel<<<(n + 255u) / 256u, 256, 0, (cudaStream_t)stream>>>(sinks, n);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_batch_score_q(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, nvb_string_set patterns, const uint8_t* quals,
                                 nvb_string_set texts, nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{ if (!qs) return NVB_ERR_ARG; return batch_score_impl(ctx, a, qs, quals, patterns, texts, sinks, maxM, maxN, temp, temp_bytes, stream); }

static int batch_banded_score_impl(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, const uint8_t* quals, uint32_t band, nvb_string_set patterns,
                                      nvb_string_set texts, nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN,
                                      void* temp, size_t temp_bytes, void* stream)
{
    (void)temp; (void)temp_bytes;
    if (!ctx) return NVB_ERR_ARG;
    int rc;
    if ((rc = check_aligner(a)) || (rc = check_sets(patterns, texts))) return rc;
    if (band < 2 || band > NVB_MAX_BAND_LEN) return NVB_ERR_UNSUPPORTED;
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!sinks) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StrSet P = make_set(patterns), T = make_set(texts);
    QualDev qdev; const QualDev* qd = nullptr;
    if ((rc = quality_tables(ctx, a, qs, quals, s, &qdev, &qd))) return rc;
    if (qd) { ctx->last_paths = 2u; return dpg_banded_score(ctx, a, band, P, T, nullptr, nullptr, 0, n, sinks, s, Sink2{nullptr, 0}, qd); }
    if ((rc = unpack_for_fast_path(ctx, P, T, maxM, maxN, s, [&](const StrSet& p8, const StrSet& t8)
             { return (ctx->force_generic == 0 && dpb2_score_supported(a, band, p8, t8, maxM, maxN)) || dpf_banded_score_supported(a, band, p8, t8, maxM, maxN); }))) return rc;
    ctx->last_paths = 0;
    const bool use_b2 = ctx->force_generic == 0 && dpb2_score_supported(a, band, P, T, maxM, maxN);
    if (use_b2 || (ctx->force_generic != 1 && dpf_banded_score_supported(a, band, P, T, maxM, maxN)))   // force_generic: 1 = int32 kernels only, 2 = no sheared-coordinate kernels
    {
        uint32_t* wl = nullptr;
        if ((rc = nvb_worklist(ctx, n, &wl))) return rc;
        NVB_CUDA(ctx, cudaMemsetAsync(wl, 0, sizeof(uint32_t), s));
        if (use_b2) { if ((rc = dpb2_score(ctx, a, band, P, T, sinks, maxM, maxN, wl, s))) return rc; ctx->last_paths |= 4u; }
        else if ((rc = dpf_banded_score(ctx, a, band, P, T, sinks, maxM, maxN, wl, s))) return rc;
        ctx->last_paths |= 1u;
        return dpg_banded_score(ctx, a, band, P, T, wl + 1, wl, 0, n, sinks, s);
    }
    ctx->last_paths |= 2u;
    return dpg_banded_score(ctx, a, band, P, T, nullptr, nullptr, 0, n, sinks, s);
}

extern "C" int nvb_batch_banded_score(nvb_ctx* ctx, nvb_aligner a, uint32_t band, nvb_string_set patterns,
                                      nvb_string_set texts, nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN,
                                      void* temp, size_t temp_bytes, void* stream)
{ return batch_banded_score_impl(ctx, a, nullptr, nullptr, band, patterns, texts, sinks, maxM, maxN, temp, temp_bytes, stream); }

extern "C" int nvb_batch_banded_score_q(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, uint32_t band, nvb_string_set patterns,
                                        const uint8_t* quals, nvb_string_set texts, nvb_best_sink* sinks, uint32_t maxM, uint32_t maxN,
                                        void* temp, size_t temp_bytes, void* stream)
{ if (!qs) return NVB_ERR_ARG; return batch_banded_score_impl(ctx, a, qs, quals, band, patterns, texts, sinks, maxM, maxN, temp, temp_bytes, stream); }

// aln::Best2Sink<int32>(distinct_dist) in place of BestSink (sink.h:99-123, sink_inl.h:54-82); band 0 = full matrix.
// The sequential "else if" of Best2Sink::report is order-dependent, so this runs on the int32 one-pair-per-thread
// kernels, which visit cells in the reference's order.
extern "C" int nvb_batch_score_best2(nvb_ctx* ctx, nvb_aligner a, uint32_t band, uint32_t distinct_dist, nvb_string_set patterns,
                                     nvb_string_set texts, nvb_best2_sink* sinks, uint32_t maxM, uint32_t maxN,
                                     void* temp, size_t temp_bytes, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc;
    if ((rc = check_aligner(a)) || (rc = check_sets(patterns, texts))) return rc;
    if (band == 1 || band > NVB_MAX_BAND_LEN) return NVB_ERR_UNSUPPORTED;
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!sinks) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const StrSet P = make_set(patterns), T = make_set(texts);
    const Sink2 s2 = { sinks, distinct_dist };
    ctx->last_paths = 2u;
    if (band) return dpg_banded_score(ctx, a, band, P, T, nullptr, nullptr, 0, n, nullptr, s, s2);
    const TempPlan p = plan(NVB_OP_SCORE, a, 0, 0, maxM, maxN);
    if (!temp && p.i16_per_pair) return NVB_ERR_TEMP;
    const uint32_t bp = batch_pairs(p, temp_bytes, n);
    if (bp == 0) return NVB_ERR_TEMP;
    for (uint32_t first = 0; first < n; first += bp)
    {
        const uint32_t cnt = n - first < bp ? n - first : bp;
        if ((rc = dpg_full_score(ctx, a, P, T, nullptr, nullptr, first, cnt, nullptr, (int16_t*)temp, maxN, s, s2))) return rc;
    }
    return NVB_OK;
}

static int batch_traceback_impl(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, const uint8_t* quals, uint32_t chk, nvb_string_set patterns,
                                   nvb_string_set texts, nvb_alignment* out, uint8_t* ops, uint32_t ops_stride,
                                   uint32_t* n_ops, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc;
    if ((rc = check_aligner(a)) || (rc = check_sets(patterns, texts))) return rc;
    if (chk == 0 || (chk & 15u)) return NVB_ERR_UNSUPPORTED;   // stripes of 16/8 must tile a checkpoint window
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!out || !ops || !n_ops || !temp) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const StrSet P = make_set(patterns), T = make_set(texts);
    QualDev qdev; const QualDev* qd = nullptr;
    if ((rc = quality_tables(ctx, a, qs, quals, s, &qdev, &qd))) return rc;
    const TempPlan p = plan(NVB_OP_TRACEBACK, a, 0, chk, maxM, maxN);
    const uint32_t bp = batch_pairs(p, temp_bytes, n);
    if (bp == 0) return NVB_ERR_TEMP;
    ctx->last_paths = 0;
    const uint32_t bp128 = (bp + 127u) & ~127u;
    const uint32_t* list = nullptr; const uint32_t* d_count = nullptr;
    if (!qd && ctx->force_generic == 0 && dpf2_traceback_supported(a, maxM, maxN))
    {
        // SW / ED / Gotoh GLOBAL and SEMI_GLOBAL: one tagged packed forward pass + a walk per pair (the directions do not depend on the
        // checkpoints); the flow of a launch lives in temp. Pairs the packed kernel cannot represent are re-done below.
        const size_t per_couple = dpf2_tb_bytes_per_couple(a, maxM, maxN);
        uint8_t* flow = (uint8_t*)(((uintptr_t)temp + 15u) & ~(uintptr_t)15u);
        const size_t avail = temp_bytes > 32 ? temp_bytes - 32 : 0;
        uint64_t fit = 2ull * (avail / per_couple);
        if (fit >= 2)
        {
            uint32_t* wl = nullptr;
            if ((rc = nvb_worklist(ctx, n, &wl))) return rc;
            NVB_CUDA(ctx, cudaMemsetAsync(wl, 0, sizeof(uint32_t), s));
            const uint32_t* order = nullptr;
            if (dpf2_wants_order(a) && (rc = dpf2_order_by_length(ctx, P, maxM, &order, s))) return rc;
            const uint64_t n_sweep = order ? dpf2_order_entries(n, maxM) : n;   // the sweeps walk the (padded) coupling order
            if (fit > n_sweep) fit = (n_sweep + 1) & ~1ull;
            for (uint64_t first = 0; first < n_sweep; first += fit)
            {
                const uint32_t cnt = (uint32_t)(n_sweep - first < fit ? n_sweep - first : fit);
                if ((rc = dpf2_traceback(ctx, a, P, T, (uint32_t)first, cnt, out, ops, ops_stride, n_ops, maxM, maxN, flow, wl, s, order))) return rc;
            }
            ctx->last_paths |= 4u;
            list = wl + 1; d_count = wl;
        }
    }
    if (!list) ctx->last_paths |= 2u;
    for (uint32_t first = 0; first < n; first += bp)
    {
        const uint32_t cnt = n - first < bp ? n - first : bp;
        int16_t* t16 = (int16_t*)temp;
        uint8_t* flow = (uint8_t*)temp + p.i16_per_pair * 2 * (size_t)bp128;
        if ((rc = dpg_full_traceback(ctx, a, P, T, list, d_count, first, cnt, chk, out, ops, ops_stride, n_ops, t16, flow, maxM, maxN, s, qd))) return rc;
    }
    return NVB_OK;
}

extern "C" int nvb_batch_traceback(nvb_ctx* ctx, nvb_aligner a, uint32_t chk, nvb_string_set patterns,
                                   nvb_string_set texts, nvb_alignment* out, uint8_t* ops, uint32_t ops_stride,
                                   uint32_t* n_ops, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{ return batch_traceback_impl(ctx, a, nullptr, nullptr, chk, patterns, texts, out, ops, ops_stride, n_ops, maxM, maxN, temp, temp_bytes, stream); }

extern "C" int nvb_batch_traceback_q(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, uint32_t chk, nvb_string_set patterns,
                                     const uint8_t* quals, nvb_string_set texts, nvb_alignment* out, uint8_t* ops, uint32_t ops_stride,
                                     uint32_t* n_ops, uint32_t maxM, uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{ if (!qs) return NVB_ERR_ARG; return batch_traceback_impl(ctx, a, qs, quals, chk, patterns, texts, out, ops, ops_stride, n_ops, maxM, maxN, temp, temp_bytes, stream); }

static int batch_banded_traceback_impl(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, const uint8_t* quals, uint32_t band, uint32_t chk,
                                          nvb_string_set patterns, nvb_string_set texts, nvb_alignment* out,
                                          uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM,
                                          uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc;
    if ((rc = check_aligner(a)) || (rc = check_sets(patterns, texts))) return rc;
    if (band < 2 || band > NVB_MAX_BAND_LEN || chk == 0) return NVB_ERR_UNSUPPORTED;
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!out || !ops || !n_ops || !temp) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StrSet P = make_set(patterns), T = make_set(texts);
    QualDev qdev; const QualDev* qd = nullptr;
    if ((rc = quality_tables(ctx, a, qs, quals, s, &qdev, &qd))) return rc;
    if ((rc = unpack_for_fast_path(ctx, P, T, maxM, maxN, s, [&](const StrSet& p8, const StrSet& t8)
             { return !qd && ctx->force_generic == 0 && dpb2_traceback_supported(a, band, p8, t8, maxM); }))) return rc;
    const TempPlan p = plan(NVB_OP_BANDED_TRACEBACK, a, band, chk, maxM, maxN);
    const uint32_t bp = batch_pairs(p, temp_bytes, n);
    if (bp == 0) return NVB_ERR_TEMP;
    const uint32_t bp128 = (bp + 127u) & ~127u;
    ctx->last_paths = 0;
    const uint32_t* list = nullptr; const uint32_t* d_count = nullptr;
    if (!qd && ctx->force_generic == 0 && dpb2_traceback_supported(a, band, P, T, maxM))
    {
        // one tagged forward pass + a walk per pair; the flow of a launch lives in temp ([row][couple]).
        // Pairs the packed kernel cannot represent (a symbol > 3, ...) are re-done below by the int32 kernel.
        const size_t per_couple = dpb2_tb_bytes_per_couple(a, maxM);
        uint8_t* flow = (uint8_t*)(((uintptr_t)temp + 15u) & ~(uintptr_t)15u);
        const size_t avail = temp_bytes > 32 ? temp_bytes - 32 : 0;
        uint64_t fit = per_couple ? 2ull * (avail / per_couple) : 0;
        if (fit >= 2)
        {
            if (fit > n) fit = ((uint64_t)n + 1) & ~1ull;
            uint32_t* wl = nullptr;
            if ((rc = nvb_worklist(ctx, n, &wl))) return rc;
            NVB_CUDA(ctx, cudaMemsetAsync(wl, 0, sizeof(uint32_t), s));
            for (uint64_t first = 0; first < n; first += fit)
            {
                const uint32_t cnt = (uint32_t)(n - first < fit ? n - first : fit);
                if ((rc = dpb2_traceback(ctx, a, band, P, T, (uint32_t)first, cnt, out, ops, ops_stride, n_ops, maxM, flow, wl, s))) return rc;
            }
            ctx->last_paths |= 1u | 4u;
            list = wl + 1; d_count = wl;
        }
    }
    if (!list) ctx->last_paths |= 2u;
    for (uint32_t first = 0; first < n; first += bp)
    {
        const uint32_t cnt = n - first < bp ? n - first : bp;
        int16_t* t16 = (int16_t*)temp;
        uint8_t* flow = (uint8_t*)temp + p.i16_per_pair * 2 * (size_t)bp128;
        if ((rc = dpg_banded_traceback(ctx, a, band, P, T, list, d_count, first, cnt, chk, out, ops, ops_stride, n_ops, t16, flow, maxM, s, qd))) return rc;
    }
    return NVB_OK;
}

extern "C" int nvb_batch_banded_traceback(nvb_ctx* ctx, nvb_aligner a, uint32_t band, uint32_t chk,
                                          nvb_string_set patterns, nvb_string_set texts, nvb_alignment* out,
                                          uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM,
                                          uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{ return batch_banded_traceback_impl(ctx, a, nullptr, nullptr, band, chk, patterns, texts, out, ops, ops_stride, n_ops, maxM, maxN, temp, temp_bytes, stream); }

extern "C" int nvb_batch_banded_traceback_q(nvb_ctx* ctx, nvb_aligner a, const nvb_quality_scheme* qs, uint32_t band, uint32_t chk,
                                            nvb_string_set patterns, const uint8_t* quals, nvb_string_set texts, nvb_alignment* out,
                                            uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM,
                                            uint32_t maxN, void* temp, size_t temp_bytes, void* stream)
{ if (!qs) return NVB_ERR_ARG; return batch_banded_traceback_impl(ctx, a, qs, quals, band, chk, patterns, texts, out, ops, ops_stride, n_ops, maxM, maxN, temp, temp_bytes, stream); }

// ---------------------------------------------------------------------------------------------
// push-order ops -> forward run-length CIGAR (io::Cigar, nvbio/io/alignments.h:48-65)
// ---------------------------------------------------------------------------------------------
__global__ void ops_to_cigar_kernel(const uint8_t* __restrict__ ops, uint32_t ops_stride, const uint32_t* __restrict__ n_ops,
                                    uint32_t n, uint16_t* __restrict__ cigar, uint32_t cigar_stride, uint32_t* __restrict__ n_cigar)
{
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    const uint8_t* o = ops + (size_t)id * ops_stride;
    uint16_t* c = cigar + (size_t)id * cigar_stride;
    uint32_t len = n_ops[id]; if (len > ops_stride) len = ops_stride;
    uint32_t out = 0, run = 0, prev = 0xFFu;
    for (uint32_t k = len; k-- > 0;)
    {
        const uint32_t op = o[k] & 3u;
        if (op == prev && run < 16383u) { ++run; continue; }
        if (run) { if (out < cigar_stride) c[out] = (uint16_t)(prev | (run << 2)); ++out; }
        prev = op; run = 1;
    }
    if (run) { if (out < cigar_stride) c[out] = (uint16_t)(prev | (run << 2)); ++out; }
    n_cigar[id] = out;
}

extern "C" int nvb_ops_to_cigar(nvb_ctx* ctx, const uint8_t* ops, uint32_t ops_stride, const uint32_t* n_ops, uint32_t n,
                                uint16_t* cigar, uint32_t cigar_stride, uint32_t* n_cigar, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!ops || !n_ops || !cigar || !n_cigar) return NVB_ERR_ARG;
    ops_to_cigar_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(ops, ops_stride, n_ops, n, cigar, cigar_stride, n_cigar);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------
// The reference's in-memory records: CUDA's uint2 is 8-byte aligned, so aln::BestSink<int32> { int32 score; uint2 sink; } is
// 16 bytes (score, 4 bytes of padding, sink.x, sink.y) and aln::Alignment<int32> { int32 score; uint2 source, sink; } 24 --
// not the packed 12 / 20-byte records of this ABI. These two calls write an array of results in the reference's layout
// (what include/nvbio_b200_adaptor.h hands back to the reference's callers).
// ---------------------------------------------------------------------------------------------
__global__ void sinks_to_ref_kernel(const nvb_best_sink* __restrict__ in, uint32_t n, uint4* __restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const nvb_best_sink v = in[i]; out[i] = make_uint4((uint32_t)v.score, 0u, v.sink_x, v.sink_y); }
}
__global__ void alignments_to_ref_kernel(const nvb_alignment* __restrict__ in, uint32_t n, uint2* __restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
    {
        const nvb_alignment v = in[i];
        out[3 * (size_t)i]     = make_uint2((uint32_t)v.score, 0u);
        out[3 * (size_t)i + 1] = make_uint2(v.source_x, v.source_y);
        out[3 * (size_t)i + 2] = make_uint2(v.sink_x, v.sink_y);
    }
}
extern "C" int nvb_sinks_to_reference_layout(nvb_ctx* ctx, const nvb_best_sink* sinks, uint32_t n, void* best_sinks_out, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!sinks || !best_sinks_out || ((uintptr_t)best_sinks_out & 15u)) return NVB_ERR_ARG;
    sinks_to_ref_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(sinks, n, (uint4*)best_sinks_out);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
extern "C" int nvb_alignments_to_reference_layout(nvb_ctx* ctx, const nvb_alignment* alignments, uint32_t n, void* alignments_out, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!alignments || !alignments_out || ((uintptr_t)alignments_out & 7u)) return NVB_ERR_ARG;
    alignments_to_ref_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(alignments, n, (uint2*)alignments_out);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
