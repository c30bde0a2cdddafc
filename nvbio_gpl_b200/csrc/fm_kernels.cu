// fm_kernels.cu -- FM-index kernels and their C-ABI entry points.
//
//   fm_match_kernel    match(fmi, pattern, len[, range])  nvbio/fmindex/fmindex_inl.h:186-239
//   fm_rank_kernel     rank(fmi, range, c)                nvbio/fmindex/fmindex_inl.h:60-88
//   fm_rank4_kernel    rank4(fmi, k)                      nvbio/fmindex/fmindex_inl.h:100-125
//   fm_locate_*        locate / locate_ssa_iterator / lookup_ssa_iterator   fmindex_inl.h:364-460
//   fm_fuse_*          bwt + occ fusing                   nvbio/io/fmindex/fmindex_impl.cu:283-318
//
// The reference ships these only as device *functions* (its test kernel is
// nvbio-test/fmindex_test.cu:56-85, its batch form thrust::transform in filter_inl.h:261-291).
#include "fm_device.cuh"
#include <cub/device/device_scan.cuh>

#define FM_THREADS 256
#define FM_STAGE_BYTES 8192   // staged query symbols per CTA (queries up to ~31 symbols on average)

// ---------------------------------------------------------------------------------------------
// backward search, one query per thread.
// The query symbols of a CTA are one contiguous span of the concatenated set: it is staged into
// shared memory with coalesced 16-byte loads, so the dependent loop only touches the index.
// ---------------------------------------------------------------------------------------------
// REV: match_reverse (fmindex_inl.h:252-278) -- the same loop consuming the pattern front to back.
template <int BITS, bool FROM, bool KTAB, bool REV, bool TMA>
__global__ void __launch_bounds__(FM_THREADS, 8)
fm_match_kernel(const FMView f, const StrSet q, uint32_t* __restrict__ ranges)
{
    __shared__ __align__(16) uint8_t s_sym[FM_STAGE_BYTES];
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    __shared__ uint32_t s_L2[8];             // L2[c] by a shared-memory load: indexing the parameter struct would go through local memory
    if (threadIdx.x < 5) s_L2[threadIdx.x] = f.L2[threadIdx.x];
    fm_lut_init(s_lut);
    const uint32_t first = blockIdx.x * FM_THREADS;
    const uint32_t tid   = first + threadIdx.x;
    const uint32_t last  = min(first + FM_THREADS, q.count);

    // stage the CTA's symbol span: bytes for 8-bit sets, whole 32-bit words for packed sets
    constexpr uint32_t SPW = 32u / BITS;                               // symbols per word of a packed stream
    const uint32_t span_begin = BITS == 8 ? q.off[first] : (q.off[first] / SPW) * 4u;               // in bytes
    const uint32_t span_end   = BITS == 8 ? q.off[last]  : ((q.off[last] + SPW - 1u) / SPW) * 4u;
    const uint32_t span_len   = span_end - span_begin;
    const bool staged = span_len <= FM_STAGE_BYTES - 32u;
    uint32_t skew = 0;
    if (staged)
    {
        // align the global side to 16 bytes; the shared copy keeps the same skew
        const uintptr_t gaddr = (uintptr_t)(q.sym + span_begin);
        skew = (uint32_t)(gaddr & 15u);
        const uint4* g = (const uint4*)(gaddr - skew);
        const uint32_t n16 = (span_len + skew + 15u) >> 4;
        if (TMA && n16)
        {
            // one bulk copy (TMA, cp.async.bulk -> UBLKCP) of the CTA's span, issued by one thread and awaited on an mbarrier
            // by all: the span is contiguous, 16-byte aligned on both sides and a multiple of 16 bytes long
            __shared__ __align__(8) uint64_t s_bar;
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
            if (threadIdx.x == 0)
            {
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar));
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                const uint32_t bytes = n16 << 4;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"((uint32_t)__cvta_generic_to_shared(s_sym)), "l"(g), "r"(bytes), "r"(bar) : "memory");
            }
            __syncthreads();                                  // the barrier is initialised (and the LUT / L2 tables written)
            uint32_t done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(bar) : "memory");
        }
        else if (!TMA)
        {
            uint4* s = (uint4*)s_sym;
            for (uint32_t i = threadIdx.x; i < n16; i += FM_THREADS) s[i] = __ldg(g + i);
            __syncthreads();
        }
    }
    // symbol at absolute symbol index a of the set
    auto symbol = [&](uint32_t a) -> uint32_t
    {
        if (!staged) return sym_at<BITS>(q.sym, a);
        if (BITS == 8) return (uint32_t)s_sym[skew + (a - span_begin)];
        const uint32_t w = *(const uint32_t*)(s_sym + skew + ((a / SPW) * 4u - span_begin));
        return BITS == 4 ? (w >> (28u - 4u * (a & 7u))) & 15u : (w >> (30u - 2u * (a & 15u))) & 3u;
    };
    if (tid >= q.count) return;

    const uint32_t beg = q.off[tid];
    const uint32_t len = q.off[tid + 1] - beg;

    uint32_t x, y;
    if (FROM) { x = ranges[2 * tid]; y = ranges[2 * tid + 1]; }
    else      { x = 0; y = f.length; }

    // search step t consumes symbol len - 1 - t (match) or t (match_reverse)
    auto consumed = [&](uint32_t t) -> uint32_t { return symbol(beg + (REV ? t : len - 1u - t)); };
    uint32_t t = 0;
    if (KTAB && !FROM && len >= f.ktab_k)
    {
        // the first k consumed symbols select a precomputed (x, y): the pair the loop below reaches after its first k steps
        // (or the pair it stopped on); a symbol > 3 among them leaves the query to the loop
        uint32_t idx = 0, bad = 0;
        for (uint32_t j = 0; j < f.ktab_k; ++j)
        {
            const uint32_t c = consumed(j);
            bad |= c; idx |= (c & 3u) << (2u * j);
        }
        if (bad <= 3u) { const uint2 r = __ldg(f.ktab + idx); x = r.x; y = r.y; t = f.ktab_k; }
    }
    // first step: x - 1 may be -1 and y may be length (from [0, length] the reference's shortcuts touch no block at all)
    if (t < len && x <= y)
    {
        const uint32_t c = consumed(t);
        if (c > 3u) { x = 1; y = 0; t = len; }
        else
        {
            const uint2 r = fm_rank2(f, x - 1u, y, c, s_lut);
            x = s_L2[c] + r.x + 1u;
            y = s_L2[c] + r.y;
            ++t;
        }
    }
    // every later step has x >= 1: the lean form (an index whose primary row is 0 keeps the general one)
    if (f.primary != 0u)
    {
        for (; t < len && x <= y; ++t)
        {
            const uint32_t c = consumed(t);
            if (c > 3u) { x = 1; y = 0; break; }
            const uint2 r = fm_rank2_inner(f, x - 1u, y, c, s_lut);
            x = s_L2[c] + r.x + 1u;
            y = s_L2[c] + r.y;
        }
    }
    else
    {
        for (; t < len && x <= y; ++t)
        {
            const uint32_t c = consumed(t);
            if (c > 3u) { x = 1; y = 0; break; }
            const uint2 r = fm_rank2(f, x - 1u, y, c, s_lut);
            x = s_L2[c] + r.x + 1u;
            y = s_L2[c] + r.y;
        }
    }
    *(uint2*)(ranges + 2 * (size_t)tid) = make_uint2(x, y);
}

// ---------------------------------------------------------------------------------------------
// nvBowtie's approximate seed mapping, map<find_exact> (nvBowtie/bowtie2/cuda/mapping_inl.h:114-184): exact over the first
// len1 search steps, at most one substitution in the rest. One query per thread; a step of the second phase needs the counts
// of all four symbols at both ends of the range (rank4, :146): the same one or two 32-byte blocks as a plain step, four
// masked popcount pairs each. Behind a substitution the tail is searched like match_range (:74-86).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FM_THREADS)
fm_match_approx_kernel(const FMView f, const StrSet q, const uint32_t len1, const uint32_t find_exact, const uint32_t cap,
                       uint32_t* __restrict__ ranges, uint32_t* __restrict__ counts)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    __shared__ uint32_t s_L2[8];
    if (threadIdx.x < 5) s_L2[threadIdx.x] = f.L2[threadIdx.x];
    fm_lut_init(s_lut);
    const uint32_t tid = blockIdx.x * FM_THREADS + threadIdx.x;
    if (tid >= q.count) return;
    const uint32_t beg = q.off[tid], len = q.off[tid + 1] - beg;
    uint2* out = (uint2*)ranges + (size_t)tid * cap;
    uint32_t n = 0;
    fm_approx_search(f, s_lut, s_L2, len, len1, find_exact != 0u,
                     [&](uint32_t t) -> uint32_t { return sym_at_rt(q.sym, q.bits, beg + (len - 1u - t)); },   // step t consumes q[len-1-t]
                     [&](uint32_t x, uint32_t y) { if (n < cap) out[n] = make_uint2(x, y); ++n; });
    counts[tid] = n;
}

__global__ void __launch_bounds__(FM_THREADS)
fm_rank_kernel(const FMView f, const uint32_t* __restrict__ ranges, const uint8_t* __restrict__ c,
               uint32_t n, uint32_t* __restrict__ out)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    const uint32_t tid = blockIdx.x * FM_THREADS + threadIdx.x;
    if (tid >= n) return;
    const uint2 rg = *(const uint2*)(ranges + 2 * (size_t)tid);
    const uint32_t cc = c[tid] & 3u;
    uint2 r;
    if (rg.x == rg.y) { r.x = r.y = fm_rank1(f, rg.x, cc, s_lut); }   // fmindex_inl.h:66-70
    else r = fm_rank2(f, rg.x, rg.y, cc, s_lut);
    *(uint2*)(out + 2 * (size_t)tid) = r;
}

__global__ void __launch_bounds__(FM_THREADS)
fm_rank4_kernel(const FMView f, const uint32_t* __restrict__ k, uint32_t n, uint32_t* __restrict__ out)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    const uint32_t tid = blockIdx.x * FM_THREADS + threadIdx.x;
    if (tid >= n) return;
    uint32_t kk = k[tid];
    uint4 r = make_uint4(0, 0, 0, 0);
    if (kk == f.length)
        r = make_uint4(f.L2[1] - f.L2[0], f.L2[2] - f.L2[1], f.L2[3] - f.L2[2], f.L2[4] - f.L2[3]);
    else if (kk != 0xFFFFFFFFu)
    {
        if (kk >= f.primary) --kk;
        if (kk != 0xFFFFFFFFu)
        {
            const FMBlock b = fm_load_block(f.blk, kk >> 6);
            r = fm_block_rank4(b, kk & 63u, s_lut);
        }
    }
    *(uint4*)(out + 4 * (size_t)tid) = r;
}

template <bool DENSE>
__global__ void __launch_bounds__(FM_THREADS, 8)
fm_locate_kernel(const FMView f, const uint32_t* __restrict__ rows, uint32_t n, uint32_t* __restrict__ pos)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    fm_walk_balanced<DENSE>(f, n, blockIdx.x * FM_THREADS + threadIdx.x, gridDim.x * FM_THREADS, s_lut,
                     [&](uint32_t h) { return __ldg(rows + h); },
                     [&](uint32_t h, uint32_t j, uint32_t t) { pos[h] = __ldg(f.sa_tab + (j >> f.sa_shift)) + t; });
}

template <bool DENSE>
__global__ void __launch_bounds__(FM_THREADS)
fm_locate_walk_kernel(const FMView f, const uint32_t* __restrict__ rows, uint32_t n, uint32_t* __restrict__ walk)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    fm_walk_balanced<DENSE>(f, n, blockIdx.x * FM_THREADS + threadIdx.x, gridDim.x * FM_THREADS, s_lut,
                     [&](uint32_t h) { return __ldg(rows + h); },
                     [&](uint32_t h, uint32_t j, uint32_t t) { *(uint2*)(walk + 2 * (size_t)h) = make_uint2(j, t); });
}

__global__ void __launch_bounds__(FM_THREADS)
fm_locate_lookup_kernel(const FMView f, const uint32_t* __restrict__ walk, uint32_t n, uint32_t* __restrict__ pos)
{
    const uint32_t tid = blockIdx.x * FM_THREADS + threadIdx.x;
    if (tid >= n) return;
    const uint2 w = *(const uint2*)(walk + 2 * (size_t)tid);
    pos[tid] = __ldg(f.sa_tab + (w.x >> f.sa_shift)) + w.y;
}

// ---------------------------------------------------------------------------------------------
// fusing: per 64-symbol block, symbol counts -> exclusive scan -> interleaved image
// ---------------------------------------------------------------------------------------------
struct U4Add { __host__ __device__ uint4 operator()(const uint4& a, const uint4& b) const
{ return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); } };

__global__ void fm_fuse_count_kernel(const uint4* __restrict__ bwt, uint32_t n_blocks, uint32_t length,
                                     uint4* __restrict__ counts)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    const uint4 w = bwt[b];
    const uint32_t ws[4] = { w.x, w.y, w.z, w.w };
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    #pragma unroll
    for (int wi = 0; wi < 4; ++wi)
    {
        const int64_t remain = (int64_t)length - ((int64_t)b * 64 + wi * 16);
        const int32_t nv = remain < 0 ? 0 : remain > 16 ? 16 : (int32_t)remain;   // valid symbols
        const uint32_t m = (uint32_t)(0xFFFFFFFF00000000ull >> (2 * nv));
        #pragma unroll
        for (uint32_t c = 0; c < 4; ++c) cnt[c] += __popc(fm_eq_bits(ws[wi], c * 0x55555555u) & m);
    }
    counts[b] = make_uint4(cnt[0], cnt[1], cnt[2], cnt[3]);
}

__global__ void fm_fuse_write_kernel(const uint4* __restrict__ bwt, const uint4* __restrict__ occ,
                                     uint32_t n_blocks, uint4* __restrict__ fused)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    fused[2 * (size_t)b]     = bwt[b];
    fused[2 * (size_t)b + 1] = occ[b];
}

// ---------------------------------------------------------------------------------------------
// entry points
// ---------------------------------------------------------------------------------------------
static int check_fmi(const nvb_fmindex* f, bool need_ssa)
{
    if (!f || !f->bwt_occ) return NVB_ERR_ARG;
    if (f->occ_int != 64u || f->sa_int != 16u) return NVB_ERR_UNSUPPORTED;
    if (((uintptr_t)f->bwt_occ) & 31u) return NVB_ERR_ARG;
    if (need_ssa && !f->ssa) return NVB_ERR_ARG;
    return NVB_OK;
}
static inline uint32_t grid_for(uint32_t n) { return (n + FM_THREADS - 1) / FM_THREADS; }
// the load-balanced walks run on a grid that just fills the device (8 CTAs of 256 threads per SM)
static inline uint32_t walk_grid(const nvb_ctx* ctx, uint32_t n) { const uint32_t g = grid_for(n), p = (uint32_t)ctx->sm_count * 8u; return g < p ? g : p; }


template <bool FROM, bool REV>
static int fm_match_launch(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges,
                           void* stream)
{
    if (!ctx || !ranges) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    if (queries.count == 0) return NVB_OK;
    if (!queries.symbols || !queries.offsets) return NVB_ERR_ARG;
    FMView f = make_view(fmi);
    const StrSet q = make_set(queries);
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t grid = grid_for(q.count);
    if (f.ktab_k > 15u) return NVB_ERR_ARG;
    if (!(queries.bits == 8 || queries.bits == 4 || queries.bits == 2)) return NVB_ERR_ARG;
    // large batches share their search prefixes through a per-batch k-mer table (context-owned scratch, fm_batch_table);
    // continued ranges and small batches stay plain
    if (!FROM && q.count >= ctx->fm_share_min)
        if ((rc = nvb_fm_batch_table(ctx, f, q.count, 15u, s))) return rc;
    const bool kt = !FROM && f.ktab_k > 0;
    // the CTA's query span goes to shared memory by one TMA bulk copy (NVB_FM_TMA=0: by per-thread 16-byte loads)
    #define FM_LAUNCH(B, K, T) fm_match_kernel<B, FROM, K, REV, T><<<grid, FM_THREADS, 0, s>>>(f, q, ranges)
    #define FM_LAUNCH_B(B) do { if (kt) { if (ctx->fm_tma) FM_LAUNCH(B, true, true); else FM_LAUNCH(B, true, false); } \
                                else    { if (ctx->fm_tma) FM_LAUNCH(B, false, true); else FM_LAUNCH(B, false, false); } } while (0)
    switch (queries.bits)
    {
    case 8: FM_LAUNCH_B(8); break;
    case 4: FM_LAUNCH_B(4); break;
    case 2: FM_LAUNCH_B(2); break;
    default: return NVB_ERR_ARG;
    }
    #undef FM_LAUNCH_B
    #undef FM_LAUNCH
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// 2-bit query streams cannot hold an N: the host entry nvb_fm_match_host_fixed2n sends, beside the 2-bit words, a list of
// (query, position of its LAST N). match() consumes a query from its end and returns (1, 0) at the first N it reaches while the
// range is still non-empty, or the empty range it had already fallen into (fmindex_inl.h:186-239 as called with N symbols, the
// behaviour tests/test_gpu_fm.py pins): so the answer of such a query is the plain search over the symbols behind its last N,
// then that rule. One thread per listed query re-runs exactly that and overwrites the range the 2-bit search produced.
__global__ void __launch_bounds__(FM_THREADS)
fm_fix_n_kernel(const FMView f, const StrSet q, const uint2* __restrict__ list, uint32_t n_list, uint32_t first_query, uint32_t* __restrict__ ranges)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    const uint32_t e = blockIdx.x * FM_THREADS + threadIdx.x;
    if (e >= n_list) return;
    const uint2 it = list[e];
    const uint32_t qi = it.x - first_query;                    // query of this chunk
    const uint32_t beg = q.off[qi], len = q.off[qi + 1] - beg;
    if (it.y >= len) return;
    uint32_t x = 0, y = f.length;
    for (uint32_t p = len; p-- > it.y + 1u && x <= y;)
    {
        const uint32_t c = sym_at<2>(q.sym, beg + p);
        const uint2 r = fm_rank2(f, x - 1u, y, c, s_lut);
        x = f.L2[c] + r.x + 1u;
        y = f.L2[c] + r.y;
    }
    if (x <= y) { x = 1; y = 0; }
    *(uint2*)(ranges + 2 * (size_t)qi) = make_uint2(x, y);
}
int nvb_fm_fix_n(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, const uint32_t* list_d, uint32_t n_list, uint32_t first_query,
                 uint32_t* ranges, cudaStream_t s)
{
    if (n_list == 0) return NVB_OK;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    FMView f = make_view(fmi); f.ktab = nullptr; f.ktab_k = 0;
    fm_fix_n_kernel<<<grid_for(n_list), FM_THREADS, 0, s>>>(f, make_set(queries), (const uint2*)list_d, n_list, first_query, ranges);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// k-mer jump table: one thread per k-mer runs the plain backward search of its k symbols
__global__ void __launch_bounds__(FM_THREADS)
fm_ktab_kernel(const FMView f, uint32_t k, uint32_t n, uint2* __restrict__ tab)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    const uint32_t idx = blockIdx.x * FM_THREADS + threadIdx.x;
    if (idx >= n) return;
    uint32_t x = 0, y = f.length;
    for (uint32_t j = 0; j < k && x <= y; ++j)
    {
        const uint32_t c = (idx >> (2u * j)) & 3u;               // last symbol first
        const uint2 r = fm_rank2(f, x - 1u, y, c, s_lut);
        x = f.L2[c] + r.x + 1u;
        y = f.L2[c] + r.y;
    }
    tab[idx] = make_uint2(x, y);
}

// entries of a k_to-mer table from a k_from-mer table: the state after the first k_from consumed symbols is entry
// (idx mod 4^k_from); k_to - k_from more steps of the plain loop follow
__global__ void __launch_bounds__(FM_THREADS)
fm_ktab_extend_kernel(const FMView f, const uint2* __restrict__ from, uint32_t k_from, uint32_t k_to, uint32_t n, uint2* __restrict__ tab)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    const uint32_t idx = blockIdx.x * FM_THREADS + threadIdx.x;
    if (idx >= n) return;
    const uint2 st = __ldg(from + (idx & ((1u << (2u * k_from)) - 1u)));
    uint32_t x = st.x, y = st.y;
    for (uint32_t j = k_from; j < k_to && x <= y; ++j)
    {
        const uint32_t c = (idx >> (2u * j)) & 3u;
        const uint2 r = fm_rank2(f, x - 1u, y, c, s_lut);
        x = f.L2[c] + r.x + 1u;
        y = f.L2[c] + r.y;
    }
    tab[idx] = make_uint2(x, y);
}

// Per-batch k-mer table (batch sharing): among n queries the first K = log4(n / 2) consumed symbols take at most 4^K <= n / 2
// distinct values, so the states after K steps are computed once per value -- 4^8 eight-step searches, then 4^K entries
// extended from them -- into context-owned scratch, and every query starts from its entry exactly as it would from a
// caller-provided table (nvb_fmindex_build_ktab). It is rebuilt on EVERY call: per-batch work inside the call, not an index-side
// structure. Cost: 4^K (K - 8) + 8 * 4^8 rank steps (6 % of the 20 n steps of 10 M 20-mers); saves K steps per query.
enum { FM_SLOT_BT8 = 62, FM_SLOT_BTK = 63 };        // slot map in common.cuh
int nvb_fm_batch_table(nvb_ctx* ctx, FMView& f, uint64_t n, uint32_t max_k, cudaStream_t s)
{
    if (max_k < 8u) return NVB_OK;                                       // strings shorter than the smallest table
    uint32_t K = 8;
    while (K < 15u && K < max_k && (4ull << (2u * K)) <= n / 2u) ++K;    // largest K with 4^K <= n / 2
    // an index that would stay L2-resident must not be pushed out by its own table: index + table within half of L2
    // (each die keeps its own copy of what its SMs read). configs[1] (50 MB): K = 10 -> 12.4 G lookups/s, K = 11 -> 11.5
    const size_t idx_bytes = nvb_bwt_occ_words(f.length) * sizeof(uint32_t);
    if (idx_bytes <= ctx->l2_bytes * 3 / 4)
        while (K > 8u && idx_bytes + (sizeof(uint2) << (2u * K)) > ctx->l2_bytes / 2) --K;
    { const char* e = getenv("NVB_FM_TABLE_K"); if (e) { const uint32_t k = (uint32_t)atoi(e); if (k >= 8u && k <= 15u && k <= max_k) K = k; } }   // tuning aid (tools/fm_probe.py)
    if (f.ktab_k >= K) return NVB_OK;                                    // the caller's table is at least as deep
    int rc; void *t8 = nullptr, *tk = nullptr;
    if ((rc = nvb_stage(ctx, FM_SLOT_BT8, sizeof(uint2) << 16, &t8))) return rc;
    FMView plain = f; plain.ktab = nullptr; plain.ktab_k = 0;
    fm_ktab_kernel<<<(1u << 16) / FM_THREADS, FM_THREADS, 0, s>>>(plain, 8u, 1u << 16, (uint2*)t8);
    NVB_LAUNCHED(ctx);
    if (K > 8u)
    {
        const uint32_t m = 1u << (2u * K);
        if ((rc = nvb_stage(ctx, FM_SLOT_BTK, sizeof(uint2) * (size_t)m, &tk))) return rc;
        fm_ktab_extend_kernel<<<m / FM_THREADS, FM_THREADS, 0, s>>>(plain, (const uint2*)t8, 8u, K, m, (uint2*)tk);
        NVB_LAUNCHED(ctx);
    }
    f.ktab = (const uint2*)(K > 8u ? tk : t8); f.ktab_k = K;
    return NVB_OK;
}

extern "C" size_t nvb_ktab_words(uint32_t k) { return k && k <= 15u ? (size_t)2 << (2u * k) : 0; }

extern "C" int nvb_fmindex_build_ktab(nvb_ctx* ctx, const nvb_fmindex* fmi, uint32_t k, uint32_t* ktab, void* stream)
{
    if (!ctx || !ktab || k == 0 || k > 15u) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    FMView f = make_view(fmi); f.ktab = nullptr; f.ktab_k = 0;
    const uint32_t n = 1u << (2u * k);
    fm_ktab_kernel<<<grid_for(n), FM_THREADS, 0, (cudaStream_t)stream>>>(f, k, n, (uint2*)ktab);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// denser sampled SA: entry i = locate(fmi, i * sa_int) through the index's own SSA (rows are generated on the fly)
__global__ void __launch_bounds__(FM_THREADS)
fm_dssa_kernel(const FMView f, uint32_t shift, uint32_t n, uint32_t* __restrict__ out)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    fm_walk_balanced<false>(f, n, blockIdx.x * FM_THREADS + threadIdx.x, gridDim.x * FM_THREADS, s_lut,
                     [&](uint32_t h) { return h << shift; },
                     [&](uint32_t h, uint32_t j, uint32_t t) { out[h] = __ldg(f.sa_tab + (j >> f.sa_shift)) + t; });
}

extern "C" size_t nvb_dssa_words(uint32_t length, uint32_t sa_int)
{ return (sa_int == 2u || sa_int == 4u || sa_int == 8u) ? ((size_t)length + sa_int) / sa_int : 0; }

extern "C" int nvb_fmindex_build_dssa(nvb_ctx* ctx, const nvb_fmindex* fmi, uint32_t sa_int, uint32_t* dssa, void* stream)
{
    if (!ctx || !dssa || !(sa_int == 2u || sa_int == 4u || sa_int == 8u)) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, true); if (rc) return rc;
    nvb_fmindex plain = *fmi; plain.dssa = nullptr; plain.dssa_int = 0;      // always derived from the 16-row SSA
    const FMView f = make_view(&plain);
    const uint32_t n = (uint32_t)nvb_dssa_words(fmi->length, sa_int);
    fm_dssa_kernel<<<walk_grid(ctx, n), FM_THREADS, 0, (cudaStream_t)stream>>>(f, sa_int == 2u ? 1u : sa_int == 4u ? 2u : 3u, n, dssa);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_fm_match(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges, void* stream)
{ return fm_match_launch<false, false>(ctx, fmi, queries, ranges, stream); }

extern "C" int nvb_fm_match_reverse(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges, void* stream)
{ return fm_match_launch<false, true>(ctx, fmi, queries, ranges, stream); }

extern "C" int nvb_fm_match_from(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges, void* stream)
{ return fm_match_launch<true, false>(ctx, fmi, queries, ranges, stream); }

extern "C" int nvb_fm_match_approx(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t len1, int find_exact,
                                   uint32_t cap, uint32_t* ranges, uint32_t* counts, void* stream)
{
    if (!ctx || !counts || (cap && !ranges)) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    if (queries.count == 0) return NVB_OK;
    if (!queries.symbols || !queries.offsets || !(queries.bits == 8 || queries.bits == 4 || queries.bits == 2)) return NVB_ERR_ARG;
    fm_match_approx_kernel<<<grid_for(queries.count), FM_THREADS, 0, (cudaStream_t)stream>>>(make_view(fmi), make_set(queries), len1,
                                                                                             find_exact ? 1u : 0u, cap, ranges, counts);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_fm_rank(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* ranges, const uint8_t* c,
                           uint32_t n, uint32_t* out, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    if (n == 0) return NVB_OK;
    if (!ranges || !c || !out) return NVB_ERR_ARG;
    fm_rank_kernel<<<grid_for(n), FM_THREADS, 0, (cudaStream_t)stream>>>(make_view(fmi), ranges, c, n, out);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_fm_rank4(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* k, uint32_t n, uint32_t* out, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    if (n == 0) return NVB_OK;
    if (!k || !out) return NVB_ERR_ARG;
    fm_rank4_kernel<<<grid_for(n), FM_THREADS, 0, (cudaStream_t)stream>>>(make_view(fmi), k, n, out);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_fm_locate(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* rows, uint32_t n, uint32_t* pos, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, true); if (rc) return rc;
    if (n == 0) return NVB_OK;
    if (!rows || !pos) return NVB_ERR_ARG;
    const FMView f = make_view(fmi);
    if (f.sa_mask != 15u) fm_locate_kernel<true><<<walk_grid(ctx, n), FM_THREADS, 0, (cudaStream_t)stream>>>(f, rows, n, pos);
    else                  fm_locate_kernel<false><<<walk_grid(ctx, n), FM_THREADS, 0, (cudaStream_t)stream>>>(f, rows, n, pos);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_fm_locate_walk(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* rows, uint32_t n, uint32_t* walk, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, false); if (rc) return rc;
    if (n == 0) return NVB_OK;
    if (!rows || !walk) return NVB_ERR_ARG;
    const FMView f = make_view(fmi);
    if (f.sa_mask != 15u) fm_locate_walk_kernel<true><<<walk_grid(ctx, n), FM_THREADS, 0, (cudaStream_t)stream>>>(f, rows, n, walk);
    else                  fm_locate_walk_kernel<false><<<walk_grid(ctx, n), FM_THREADS, 0, (cudaStream_t)stream>>>(f, rows, n, walk);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_fm_locate_lookup(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* walk, uint32_t n, uint32_t* pos, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, true); if (rc) return rc;
    if (n == 0) return NVB_OK;
    if (!walk || !pos) return NVB_ERR_ARG;
    fm_locate_lookup_kernel<<<grid_for(n), FM_THREADS, 0, (cudaStream_t)stream>>>(make_view(fmi), walk, n, pos);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------
// FMIndexFilter (nvbio/fmindex/filter.h:40-200, filter_inl.h:193-393): rank = match every string + inclusive scan
// of the range sizes; locate = for output slot i the (text position, string id) of the i-th occurrence.
// The expansion is one thread per hit (upper_bound over the scanned sizes, filter_inl.h:75-110), so queries
// with many occurrences are spread over as many threads as they have hits.
// ---------------------------------------------------------------------------------------------
struct RangeSize
{
    const uint32_t* r;
    __host__ __device__ uint64_t operator()(uint32_t i) const { return (uint64_t)(uint32_t)(1u + r[2 * (size_t)i + 1] - r[2 * (size_t)i]); }   // filter_inl.h:30-33
};

__global__ void __launch_bounds__(FM_THREADS)
fm_range_size_kernel(const uint32_t* __restrict__ ranges, uint32_t n, uint64_t* __restrict__ sizes)
{
    const uint32_t tid = blockIdx.x * FM_THREADS + threadIdx.x;
    if (tid < n) sizes[tid] = (uint64_t)(uint32_t)(1u + ranges[2 * (size_t)tid + 1] - ranges[2 * (size_t)tid]);
}

__global__ void __launch_bounds__(FM_THREADS)
fm_filter_locate_kernel(const FMView f, const uint32_t* __restrict__ ranges, const uint64_t* __restrict__ slots, uint32_t n_queries,
                        uint64_t begin, uint64_t count, uint32_t* __restrict__ hits)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    const uint64_t tid = (uint64_t)blockIdx.x * FM_THREADS + threadIdx.x;
    if (tid >= count) return;
    const uint64_t out = begin + tid;
    // upper_bound(out, slots, n_queries): first slot with slots[slot] > out
    uint32_t lo = 0, hi = n_queries;
    while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if (__ldg(slots + mid) <= out) lo = mid + 1; else hi = mid; }
    const uint32_t slot = lo;
    uint32_t pos = 0xFFFFFFFFu;
    if (slot < n_queries)
    {
        const uint64_t base = slot ? __ldg(slots + slot - 1) : 0ull;
        const uint32_t row = __ldg(ranges + 2 * (size_t)slot) + (uint32_t)(out - base);
        const uint2 w = fm_locate_walk_one(f, row, s_lut);
        pos = __ldg(f.sa_tab + (w.x >> f.sa_shift)) + w.y;
    }
    *(uint2*)(hits + 2 * tid) = make_uint2(pos, slot);
}

// FMIndexFilter::rank: ranges[2n] and slots[n] (inclusive scan of the range sizes) stay on the device;
// *n_hits_h (optional) = slots[n-1], which costs one synchronisation of the stream
extern "C" int nvb_fm_filter_rank(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges,
                                  uint64_t* slots, uint64_t* n_hits_h, void* stream)
{
    if (!ctx || !ranges || !slots) return NVB_ERR_ARG;
    if (n_hits_h) *n_hits_h = 0;
    int rc = fm_match_launch<false, false>(ctx, fmi, queries, ranges, stream); if (rc) return rc;
    const uint32_t n = queries.count;
    if (n == 0) return NVB_OK;
    cudaStream_t s = (cudaStream_t)stream;
    fm_range_size_kernel<<<grid_for(n), FM_THREADS, 0, s>>>(ranges, n, slots);
    NVB_LAUNCHED(ctx);
    void* tmp = nullptr; size_t tmp_bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, slots, slots, (int)n, s);
    if ((rc = nvb_stage(ctx, 7, tmp_bytes, &tmp))) return rc;
    NVB_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, slots, slots, (int)n, s));
    ctx->launches++;
    if (n_hits_h)
    {
        NVB_CUDA(ctx, cudaMemcpyAsync(n_hits_h, slots + (n - 1), sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
        NVB_CUDA(ctx, cudaStreamSynchronize(s));
    }
    return NVB_OK;
}

// FMIndexFilter::locate(begin, end, hits): hits[2 i] = text position, hits[2 i + 1] = string id, i in [0, end - begin)
extern "C" int nvb_fm_filter_locate(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* ranges, const uint64_t* slots,
                                    uint32_t n_queries, uint64_t begin, uint64_t end, uint32_t* hits, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    int rc = check_fmi(fmi, true); if (rc) return rc;
    if (end <= begin || n_queries == 0) return NVB_OK;
    if (!ranges || !slots || !hits) return NVB_ERR_ARG;
    const uint64_t count = end - begin;
    if (count > 0x7FFFFFFFull * FM_THREADS) return NVB_ERR_ARG;
    const uint32_t grid = (uint32_t)((count + FM_THREADS - 1) / FM_THREADS);
    fm_filter_locate_kernel<<<grid, FM_THREADS, 0, (cudaStream_t)stream>>>(make_view(fmi), ranges, slots, n_queries, begin, count, hits);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" size_t nvb_bwt_occ_words(uint32_t length) { return (size_t)((((length + 15u) / 16u) + 3u) & ~3u) * 2u; }
extern "C" size_t nvb_ssa_words(uint32_t length) { return ((size_t)length + 16u) / 16u; }

extern "C" int nvb_fmindex_fuse(nvb_ctx* ctx, const uint32_t* bwt_words, uint32_t length, uint32_t* bwt_occ_out,
                                uint32_t* L2_out_h, void* stream)
{
    if (!ctx || !bwt_words || !bwt_occ_out || !L2_out_h) return NVB_ERR_ARG;
    if ((((uintptr_t)bwt_words) & 15u) || (((uintptr_t)bwt_occ_out) & 31u)) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t n_blocks = (uint32_t)(nvb_bwt_occ_words(length) / 8u);
    L2_out_h[0] = 0;
    if (n_blocks == 0) { for (int c = 0; c < 4; ++c) L2_out_h[c + 1] = 0; return NVB_OK; }
    uint4 *counts = nullptr, *occ = nullptr; void* tmp = nullptr; size_t tmp_bytes = 0;
    NVB_CUDA(ctx, cudaMallocAsync(&counts, sizeof(uint4) * ((size_t)n_blocks + 1), s));
    NVB_CUDA(ctx, cudaMallocAsync(&occ, sizeof(uint4) * ((size_t)n_blocks + 1), s));
    NVB_CUDA(ctx, cudaMemsetAsync(counts + n_blocks, 0, sizeof(uint4), s));
    const uint32_t g = (n_blocks + 255) / 256;
    fm_fuse_count_kernel<<<g, 256, 0, s>>>((const uint4*)bwt_words, n_blocks, length, counts);
    NVB_LAUNCHED(ctx);
    // exclusive scan over n_blocks+1 items: the last output is the total = symbol counts
    cub::DeviceScan::ExclusiveScan(nullptr, tmp_bytes, counts, occ, U4Add(), make_uint4(0, 0, 0, 0), (int)(n_blocks + 1), s);
    NVB_CUDA(ctx, cudaMallocAsync(&tmp, tmp_bytes, s));
    NVB_CUDA(ctx, cub::DeviceScan::ExclusiveScan(tmp, tmp_bytes, counts, occ, U4Add(), make_uint4(0, 0, 0, 0), (int)(n_blocks + 1), s));
    ctx->launches++;
    fm_fuse_write_kernel<<<g, 256, 0, s>>>((const uint4*)bwt_words, occ, n_blocks, (uint4*)bwt_occ_out);
    NVB_LAUNCHED(ctx);
    uint4 total;
    NVB_CUDA(ctx, cudaMemcpyAsync(&total, occ + n_blocks, sizeof(uint4), cudaMemcpyDeviceToHost, s));
    NVB_CUDA(ctx, cudaStreamSynchronize(s));
    NVB_CUDA(ctx, cudaFreeAsync(counts, s));
    NVB_CUDA(ctx, cudaFreeAsync(occ, s));
    NVB_CUDA(ctx, cudaFreeAsync(tmp, s));
    L2_out_h[1] = total.x; L2_out_h[2] = L2_out_h[1] + total.y;
    L2_out_h[3] = L2_out_h[2] + total.z; L2_out_h[4] = L2_out_h[3] + total.w;
    return NVB_OK;
}
