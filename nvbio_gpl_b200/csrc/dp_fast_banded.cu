// dp_fast_banded.cu -- packed s16x2 DPX kernels for banded scoring (edit distance, Smith-Waterman and
// Gotoh; GLOBAL / LOCAL / SEMI_GLOBAL), the extension step of nvBowtie
// (replaces batched_banded_alignment_score_kernel, nvbio/alignment/batched_banded_inl.h:73-81, and the
// per-pair banded::sw_/gotoh_alignment_score_dispatch::run, sw/sw_banded_inl.h:340-504,
// gotoh/gotoh_banded_inl.h:406-646).
//
// One thread scores TWO pairs (pair A in the low, pair B in the high half of every register). The band
// (W cells of H, plus F for Gotoh) stays in registers for the whole pattern. Substitution scores of
// both pairs come from one byte permute per packed cell:
//   prof*[p] = S(text[p], q) for q = A,C,G,T as 4 bytes (all mismatch past the end of the text), kept
//              for the W text positions under the band in a register window that rotates by renaming
//              (the row loop is unrolled W times);
//   sel      = per pattern row: picks byte q_A of profA and byte q_B of profB and replicates their signs.
// Both come out of small shared-memory tables indexed by a byte stream staged per thread, so the
// ALU pipe only sees PRMT + DPX:
//   SW/ED  : S = PRMT ; u = IMAD(top,1,G') ; v = VIADDMNMX(diag,S,u) ; l = IMAD(left,1,I') ; h = VIMNMX[3](l,v[,0])
//   Gotoh  : S = PRMT ; o = IMAD(Htop,1,open') ; F = VIADDMNMX(Ftop,ext,o) ; v = VIADDMNMX(diag,S,F) ;
//            h = VIMNMX[3](v,E[,0]) ; o' = IMAD(h,1,open') ; E = VIADDMNMX(E,ext,o')
// Scores are biased (LOCAL: by the largest gap penalty, else by 0x4000) so that the additions of the
// negative gap constants can run as plain IMADs on the FMA pipe without a borrow between the halves.
// Sinks follow the reference's visiting order: LOCAL reports every cell (rows ascending, band cells
// ascending, last maximum wins: sw_banded_inl.h:407-456), SEMI_GLOBAL the last row left to right
// (:489-501), GLOBAL the last band cell (:487-488). Pairs holding a symbol > 3 go to the worklist.
#include "dp_fast.h"

#define DPB_THREADS 128

struct BandFastParams
{
    StrSet         P, T;
    nvb_best_sink* sinks;
    uint32_t*      worklist;
    uint32_t       n_pairs, n_threads;
    uint32_t       rows_cap;      // staged rows per thread    (>= max pattern length)
    uint32_t       text_cap;      // staged text positions     (>= rows_cap + W - 1)
    int32_t        match, mismatch, ga, gb, bias;
    uint32_t       one, sixteen;
};

__device__ __forceinline__ uint32_t b_pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }
__device__ __forceinline__ uint32_t b_fix2(int32_t g) { return (((uint32_t)(g - 1) & 0xFFFFu) << 16) | ((uint32_t)g & 0xFFFFu); }
__device__ __forceinline__ uint32_t b_prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ uint32_t b_imad(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ int32_t b_lo(uint32_t x) { return (int32_t)(int16_t)(x & 0xFFFFu); }
__device__ __forceinline__ int32_t b_hi(uint32_t x) { return (int32_t)(int16_t)(x >> 16); }

// mask of the low n bytes of a word (n <= 0: none, n >= 4: all)
__device__ __forceinline__ uint32_t b_bytemask(int32_t n)
{ return n <= 0 ? 0u : n >= 4 ? 0xFFFFFFFFu : (1u << (8 * n)) - 1u; }
// 4 symbols starting at an arbitrarily aligned address, from one or two aligned words; `remain` = symbols
// left in the string from ptr on (nothing is read at or past the aligned word that holds the string's end)
__device__ __forceinline__ uint32_t b_load4(const uint8_t* ptr, uint32_t remain, uint32_t mask)
{
    if (mask == 0u) return 0u;
    const uintptr_t a = (uintptr_t)ptr;
    const uint32_t sh = (uint32_t)(a & 3u);
    const uint32_t* w = (const uint32_t*)(a - sh);
    const uint32_t lo = __ldg(w);
    const uint32_t hi = (sh != 0u && remain + sh > 4u) ? __ldg(w + 1) : 0u;
    return __funnelshift_r(lo, hi, 8u * sh);
}

#define DPB_INF2 0x83008300u   // packed -32000: "minus infinity" of the F band (never wins, never wraps)

template <int W, int TYPE, bool GOTOH>
__global__ void __launch_bounds__(DPB_THREADS)
dpf_banded_kernel(const BandFastParams p)
{
    extern __shared__ __align__(16) uint8_t s_raw[];
    uint2*    lut_prof = (uint2*)s_raw;                          // [64] text byte -> (profA, profB)
    uint32_t* lut_sel  = (uint32_t*)(s_raw + 512);               // [16] pattern byte -> PRMT selector
    uint8_t*  s_pat    = s_raw + 576;                            // [threads][rows_cap]  (strides = 4 mod 8: no bank conflicts)
    uint8_t*  s_txt    = s_pat + (size_t)p.rows_cap * DPB_THREADS;   // [threads][text_cap]
    const bool LOCAL = (TYPE == T_LOCAL);

    // ---- tables (one per CTA)
    {
        const uint32_t base  = ((uint32_t)p.mismatch & 0xFFu) * 0x01010101u;
        const uint32_t delta = ((uint32_t)(p.mismatch ^ p.match)) & 0xFFu;
        const uint32_t t = threadIdx.x;
        if (t < 64)
        {
            const uint32_t rA = t & 3u, peA = (t >> 2) & 1u, rB = (t >> 3) & 3u, peB = (t >> 5) & 1u;
            lut_prof[t] = make_uint2(peA ? base : base ^ (delta << (8u * rA)), peB ? base : base ^ (delta << (8u * rB)));
        }
        if (t < 16) lut_sel[t] = ((t & 3u) | ((t & 12u) << 6)) * 0x11u + 0xC480u;
    }
    __syncthreads();

    // ---- prologue: per-thread byte streams (pattern rows, text positions under the band), staged warp-
    // cooperatively: for each couple of the warp the 32 lanes read 32 consecutive symbols (one sector)
    const uint32_t tid = blockIdx.x * DPB_THREADS + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u;
    const bool live = tid < p.n_threads;
    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false;
    if (live)
    {
        idA = 2 * tid; hasB = idA + 1 < p.n_pairs; idB = hasB ? idA + 1 : idA;
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
    }
    const uint32_t Mmax = max(MA, MB);
    bool dirtyA = MA == 0 || MA > p.rows_cap, dirtyB = MB == 0 || MB > p.rows_cap;
    const uint32_t rows = min(Mmax, p.rows_cap);
    {
        // word-wise: a lane builds 4 consecutive stream bytes from one (funnel-shifted) 32-bit word of each
        // pair; ~25 instructions per 128 positions, 4 independent loads per lane and iteration
        const uint32_t wbase = threadIdx.x & ~31u;
        const uint8_t* Psym = p.P.sym; const uint8_t* Tsym = p.T.sym;
        uint32_t dmA = 0, dmB = 0;
        #pragma unroll 2
        for (uint32_t c = 0; c < 32; ++c)
        {
            const uint32_t c_rows = __shfl_sync(0xFFFFFFFFu, rows, c);
            const uint32_t c_pbA = __shfl_sync(0xFFFFFFFFu, pbA, c), c_pbB = __shfl_sync(0xFFFFFFFFu, pbB, c);
            const uint32_t c_tbA = __shfl_sync(0xFFFFFFFFu, tbA, c), c_tbB = __shfl_sync(0xFFFFFFFFu, tbB, c);
            const uint32_t c_MA = __shfl_sync(0xFFFFFFFFu, MA, c), c_MB = __shfl_sync(0xFFFFFFFFu, MB, c);
            const uint32_t c_NA = __shfl_sync(0xFFFFFFFFu, NA, c), c_NB = __shfl_sync(0xFFFFFFFFu, NB, c);
            uint32_t* c_pat = (uint32_t*)(s_pat + (size_t)(wbase + c) * p.rows_cap);
            uint32_t* c_txt = (uint32_t*)(s_txt + (size_t)(wbase + c) * p.text_cap);
            uint32_t bad = 0;
            for (uint32_t i = 4 * lane; i < c_rows; i += 128)
            {
                const uint32_t mA = b_bytemask((int32_t)(c_MA - i)), mB = b_bytemask((int32_t)(c_MB - i));
                const uint32_t a = b_load4(Psym + c_pbA + i, c_MA - i, mA) & mA;
                const uint32_t b = b_load4(Psym + c_pbB + i, c_MB - i, mB) & mB;
                bad |= ((a & 0xFCFCFCFCu) ? 1u : 0u) | ((b & 0xFCFCFCFCu) ? 2u : 0u);
                c_pat[i >> 2] = (a & 0x03030303u) | ((b & 0x03030303u) << 2);
            }
            const uint32_t c_text = c_rows ? c_rows + W - 1 : 0u;
            for (uint32_t q = 4 * lane; q < c_text; q += 128)
            {
                const uint32_t mA = b_bytemask((int32_t)(c_NA - q)), mB = b_bytemask((int32_t)(c_NB - q));
                const uint32_t a = b_load4(Tsym + c_tbA + q, c_NA - q, mA) & mA;
                const uint32_t b = b_load4(Tsym + c_tbB + q, c_NB - q, mB) & mB;
                bad |= ((a & 0xFCFCFCFCu) ? 1u : 0u) | ((b & 0xFCFCFCFCu) ? 2u : 0u);
                // past the end of a text: code 4 = mismatches everything
                const uint32_t wa = (a & 0x03030303u) | (0x04040404u & ~mA), wb = (b & 0x03030303u) | (0x04040404u & ~mB);
                c_txt[q >> 2] = wa | (wb << 3);
            }
            dmA |= (bad & 1u) << c; dmB |= ((bad >> 1) & 1u) << c;
        }
        dmA = __reduce_or_sync(0xFFFFFFFFu, dmA); dmB = __reduce_or_sync(0xFFFFFFFFu, dmB);
        dirtyA |= (dmA >> lane) & 1u; dirtyB |= (dmB >> lane) & 1u;
        __syncwarp();
    }
    if (!live) return;
    const uint8_t* my_pat = s_pat + (size_t)threadIdx.x * p.rows_cap;
    const uint8_t* my_txt = s_txt + (size_t)threadIdx.x * p.text_cap;

    const uint32_t B2 = b_pack2(p.bias);
    const uint32_t Af = b_fix2(p.ga);                 // SW: top + deletion ; Gotoh: H + gap_open
    const uint32_t Bf = b_fix2(p.gb);                 // SW: left + insertion
    const uint32_t ext2 = b_pack2(p.gb);              // Gotoh: gap extension (added inside VIADDMNMX)
    const uint32_t ins2 = ext2;                       // SW: insertion, likewise

    // rows are unrolled RU at a time: inside a group the window slots are static (slot = row-in-group + band
    // cell), after it the window is shifted down by RU (keeps the loop body within the instruction cache)
    constexpr int RU = W > 8 ? 5 : W;
    constexpr int WS = W + RU - 1;
    uint32_t X[W], F[W], TA[WS], TB[WS], fin[W];
    #pragma unroll
    for (int j = 0; j < W; ++j)
    {
        int32_t h0 = 0;                                                     // init_row_zero
        if (TYPE == T_GLOBAL) h0 = GOTOH ? (j > 0 ? p.ga + (j - 1) * p.gb : 0) : j * p.ga;
        X[j] = b_pack2(h0 + p.bias); F[j] = DPB_INF2; fin[j] = X[j];
    }
    #pragma unroll
    for (int j = 0; j < WS; ++j) { TA[j] = 0; TB[j] = 0; }
    #pragma unroll
    for (int j = 0; j < W - 1; ++j)
    {
        const uint2 pr = lut_prof[my_txt[j]];
        TA[j] = pr.x; TB[j] = pr.y;
    }
    constexpr uint32_t KB = W > 16 ? 5u : 4u;                    // bits of the band cell in the arg-max key (score << KB | cell)
    constexpr uint32_t KMASK = ~(((1u << KB) - 1u) * 0x00010001u);
    uint32_t bestkey = B2 << KB, bestrow = 0;

    for (uint32_t i0 = 0; i0 < rows; i0 += RU)
    {
        #pragma unroll
        for (int r = 0; r < RU; ++r)
        {
            const uint32_t i = i0 + r;
            if (i < rows)
            {
                // the text position entering the band on this row takes the slot of the one that left
                {
                    const uint2 pr = lut_prof[my_txt[i + W - 1]];
                    TA[r + W - 1] = pr.x; TB[r + W - 1] = pr.y;
                }
                const uint32_t sel = lut_sel[my_pat[i]];
                uint32_t left = 0, E = 0;
                uint32_t key[W];
                #pragma unroll
                for (int j = 0; j < W; ++j)
                {
                    const uint32_t S = b_prmt(TA[r + j], TB[r + j], sel);
                    uint32_t h;
                    if (!GOTOH)
                    {
                        if (j == 0)
                        {
                            const uint32_t u = b_imad(X[1], p.one, Af);
                            h = __viaddmax_s16x2(X[0], S, u);
                            if (LOCAL) h = __vmaxs2(h, B2);
                        }
                        else if (j < W - 1)
                        {
                            const uint32_t u = b_imad(X[j + 1 < W ? j + 1 : j], p.one, Af);
                            const uint32_t v = __viaddmax_s16x2(X[j], S, u);
                            if (LOCAL) h = __vimax3_s16x2(b_imad(left, p.one, Bf), v, B2);
                            else       h = __viaddmax_s16x2(left, ins2, v);                        // max(left + I, v)
                        }
                        else
                        {
                            const uint32_t l = b_imad(left, p.one, Bf);
                            h = __viaddmax_s16x2(X[j], S, l);
                            if (LOCAL) h = __vmaxs2(h, B2);
                        }
                    }
                    else
                    {
                        if (j < W - 1)
                        {
                            const uint32_t o = b_imad(X[j + 1 < W ? j + 1 : j], p.one, Af);
                            F[j] = __viaddmax_s16x2(F[j + 1 < W ? j + 1 : j], ext2, o);     // max(F + ext, H + open)
                            const uint32_t v = __viaddmax_s16x2(X[j], S, F[j]);
                            if (j == 0) h = LOCAL ? __vmaxs2(v, B2) : v;
                            else        h = LOCAL ? __vimax3_s16x2(v, E, B2) : __vmaxs2(v, E);
                            const uint32_t o2 = b_imad(h, p.one, Af);
                            E = (j == 0) ? o2 : __viaddmax_s16x2(E, ext2, o2);              // max(E + ext, H + open)
                        }
                        else
                        {
                            F[j] = DPB_INF2;
                            h = __viaddmax_s16x2(X[j], S, E);
                            if (LOCAL) h = __vmaxs2(h, B2);
                        }
                    }
                    X[j] = h; left = h;
                    if (LOCAL) key[j] = b_imad(h, p.sixteen, (uint32_t)j * 0x00010001u);
                }
                if (LOCAL)
                {
                    // later rows win ties: compare against the best with its column bits cleared, minus one
                    const uint32_t kb = __vadd2(bestkey & KMASK, 0xFFFFFFFFu);
                    uint32_t acc = kb;
                    #pragma unroll
                    for (int j = 0; j + 1 < W; j += 2) acc = __vimax3_s16x2(acc, key[j], key[j + 1]);
                    if (W & 1) acc = __vmaxs2(acc, key[W - 1]);
                    const uint32_t rmask = (i < MA ? 0x0000FFFFu : 0u) | (i < MB ? 0xFFFF0000u : 0u);
                    const uint32_t d = __vsub2(kb, acc);
                    const uint32_t m = b_prmt(d, d, 0xBB99u) & rmask;
                    bestkey = (acc & m) | (bestkey & ~m);
                    bestrow = ((i * 0x00010001u) & m) | (bestrow & ~m);
                }
                else if (i + 1 == MA || i + 1 == MB)
                {
                    const uint32_t m = (i + 1 == MA ? 0x0000FFFFu : 0u) | (i + 1 == MB ? 0xFFFF0000u : 0u);
                    #pragma unroll
                    for (int j = 0; j < W; ++j) fin[j] = (X[j] & m) | (fin[j] & ~m);
                }
            }
        }
        #pragma unroll
        for (int j = 0; j < W - 1; ++j) { TA[j] = TA[j + RU]; TB[j] = TB[j + RU]; }
    }

    // ---- epilogue
    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        if (h ? dirtyB : dirtyA)
        {
            const uint32_t slot = atomicAdd(p.worklist, 1u);
            p.worklist[1 + slot] = id;
            continue;
        }
        nvb_best_sink out;
        out.score = NVB_SINK_MIN; out.sink_x = out.sink_y = 0xFFFFFFFFu;
        if (N >= M)                                           // shorter texts are refused (sw_banded_inl.h:356-357)
        {
            if (LOCAL)
            {
                const uint32_t k = h ? (bestkey >> 16) : (bestkey & 0xFFFFu);
                const uint32_t row = h ? (bestrow >> 16) : (bestrow & 0xFFFFu);
                out.score = (int32_t)(k >> KB) - p.bias;
                out.sink_x = row + (k & ((1u << KB) - 1u)) + 1u; out.sink_y = row + 1u;
            }
            else if (TYPE == T_GLOBAL)
            {
                out.score = (h ? b_hi(fin[W - 1]) : b_lo(fin[W - 1])) - p.bias;
                out.sink_x = M + W - 1; out.sink_y = M;
            }
            else
            {
                const uint32_t lim = (M + W - 1 < N ? M + W - 1 : N) - (M - 1);
                #pragma unroll
                for (int j = 0; j < W; ++j)
                {
                    const int32_t s = (h ? b_hi(fin[j]) : b_lo(fin[j])) - p.bias;
                    if ((j == 0 || (uint32_t)j < lim) && out.score <= s) { out.score = s; out.sink_x = M + j; out.sink_y = M; }
                }
            }
        }
        p.sinks[id] = out;
    }
}

// ---------------------------------------------------------------------------------------------
static int32_t iabs32(int32_t v) { return v < 0 ? -v : v; }

static bool band_geometry(uint32_t band, uint32_t maxM, uint32_t* rows_cap, uint32_t* text_cap, size_t* smem)
{
    // per-thread strides of 4 mod 8 bytes: consecutive threads start in consecutive odd-spaced banks
    *rows_cap = ((maxM + 7u) & ~7u) + 4u; *text_cap = ((maxM + band - 1 + 7u) & ~7u) + 4u;
    *smem = 576 + ((size_t)*rows_cap + *text_cap) * DPB_THREADS;
    return *smem <= 200u * 1024u;
}

bool dpf_banded_score_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN)
{
    (void)maxN;
    if (band != 3 && band != 5 && band != 7 && band != 15 && band != 31) return false;
    if (P.bits != 8 || T.bits != 8) return false;       // packed sets are expanded by the caller (dp_api.cu)
    const Scheme s = make_scheme(a);
    if (s.match < -127 || s.match > 127 || s.mismatch < -127 || s.mismatch > 127) return false;
    if (s.ga >= 0 || s.gb >= 0 || s.ga < -512 || s.gb < -512) return false;
    if (maxM == 0) return false;
    const int32_t big = max(max(iabs32(s.match), iabs32(s.mismatch)), max(iabs32(s.ga), iabs32(s.gb)));
    if (a.type == NVB_LOCAL)
    {
        if (max(iabs32(s.ga), iabs32(s.gb)) > 64) return false;
        if ((uint64_t)max(s.match, 0) * maxM + 64u >= (band > 16 ? 1023u : 2047u)) return false;   // 11 (10 for band 31) score bits in the arg-max key
    }
    else if ((uint64_t)big * ((uint64_t)maxM + band + 2) >= 15000u) return false;  // bias 0x4000 keeps both halves positive
    uint32_t rc, tc; size_t smem;
    return band_geometry(band, maxM, &rc, &tc, &smem);
}

template <int W, int TYPE, bool GOTOH>
static int band_launch(nvb_ctx* ctx, const BandFastParams& p, size_t smem, cudaStream_t st)
{
    if (NVB_FIRST_ON_DEVICE(ctx))
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_banded_kernel<W, TYPE, GOTOH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    dpf_banded_kernel<W, TYPE, GOTOH><<<(p.n_threads + DPB_THREADS - 1) / DPB_THREADS, DPB_THREADS, smem, st>>>(p);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

template <int W>
static int band_dispatch(nvb_ctx* ctx, const nvb_aligner& a, const BandFastParams& p, size_t smem, cudaStream_t st)
{
    const bool g = a.kind == NVB_GOTOH;
    switch (a.type)
    {
    case NVB_GLOBAL:      return g ? band_launch<W, T_GLOBAL, true>(ctx, p, smem, st) : band_launch<W, T_GLOBAL, false>(ctx, p, smem, st);
    case NVB_LOCAL:       return g ? band_launch<W, T_LOCAL, true>(ctx, p, smem, st)  : band_launch<W, T_LOCAL, false>(ctx, p, smem, st);
    case NVB_SEMI_GLOBAL: return g ? band_launch<W, T_SEMI, true>(ctx, p, smem, st)   : band_launch<W, T_SEMI, false>(ctx, p, smem, st);
    }
    return NVB_ERR_ARG;
}

int dpf_banded_score(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, nvb_best_sink* sinks,
                     uint32_t maxM, uint32_t maxN, uint32_t* worklist, cudaStream_t st)
{
    (void)maxN;
    const Scheme s = make_scheme(a);
    BandFastParams p;
    size_t smem;
    if (!band_geometry(band, maxM, &p.rows_cap, &p.text_cap, &smem)) return NVB_ERR_UNSUPPORTED;
    p.P = P; p.T = T; p.sinks = sinks; p.worklist = worklist;
    p.n_pairs = P.count; p.n_threads = (P.count + 1) / 2;
    p.match = s.match; p.mismatch = s.mismatch; p.ga = s.ga; p.gb = s.gb;
    p.bias = a.type == NVB_LOCAL ? max(iabs32(s.ga), iabs32(s.gb)) : 0x4000;
    p.one = 1; p.sixteen = band > 16 ? 32 : 16;           // the key multiplier: 1 << (bits of the band cell)
    if (band == 3)  return band_dispatch<3>(ctx, a, p, smem, st);
    if (band == 5)  return band_dispatch<5>(ctx, a, p, smem, st);
    if (band == 7)  return band_dispatch<7>(ctx, a, p, smem, st);
    if (band == 15) return band_dispatch<15>(ctx, a, p, smem, st);
    if (band == 31) return band_dispatch<31>(ctx, a, p, smem, st);
    return NVB_ERR_UNSUPPORTED;
}
