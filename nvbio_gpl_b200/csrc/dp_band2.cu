// dp_band2.cu -- packed s16x2 banded scoring in SHEARED score coordinates: GLOBAL / SEMI_GLOBAL edit distance,
// Smith-Waterman and Gotoh, two pairs per thread (pair A in the low, pair B in the high half of every register).
// (replaces batched_banded_alignment_score_kernel, nvbio/alignment/batched_banded_inl.h:73-81, over
// banded::sw_/gotoh_alignment_score_dispatch::run, sw/sw_banded_inl.h:340-504, gotoh/gotoh_banded_inl.h:406-646.)
//
// The ALU pipe of an sm_100 SM issues 64 lanes/clk and carries every s16x2 min/max (VIMNMX, VIMNMX3,
// VIADDMNMX) as well as PRMT and LOP3; integer adds (IADD3 / IMAD) also run on the FMA pipe
// (tools/pipe_mix.cu, profiles/r1b_pipe_mix.jsonl). So the cell update is arranged to need as few
// ALU-pipe instructions as possible. With band state B_k (k pattern rows done), cell j <-> text position k+j:
//     H_{k+1}[j] = max( H_k[j] + S(text[k+j], pat[k]),  H_k[j+1] + ga,  H_{k+1}[j-1] + gb )          (SW / ED)
// Storing H'' = H - alpha*k - beta*j with beta = gb, alpha = ga + gb makes BOTH gap moves free:
//     H''_{k+1}[j] = max3( H''_k[j] + S', H''_k[j+1], H''_{k+1}[j-1] ),   S' = S - ga - gb  (>= 0)
// = PRMT (S' of both pairs from the row's byte profiles) + one integer add + ONE VIMNMX3 per packed cell:
// 2 ALU-pipe instructions per two cells instead of 3 (PRMT + 2 VIADDMNMX). Gotoh, with beta = ge,
// alpha = 2 ge, d = go - ge and the band kept as HO = H'' + d:
//     F''[j] = max(F''_k[j+1], HO_k[j+1])   h = max3(HO_k[j] + S'', F''[j], E'')   ho = h + d   E'' = max(E'', ho)
// with S'' = S - ge - go (>= 0): 4 ALU-pipe instructions instead of 5, none of the F/E additions left.
// Everything stored is non-negative (a small bias covers the GLOBAL first row), so the packed additions are
// plain 32-bit adds that never carry between the halves; the true score is recovered in the epilogue:
// H = stored - bias + alpha*M + beta*j.
//
// Staging: pattern rows and text positions are consumed front to back, so each thread copies its four strings
// R rows at a time with 16-byte loads into a private shared-memory slot (thread stride = 4 mod 8 bytes: byte
// loads of a warp fall in distinct banks) and the row loop reads them back with immediate-offset LDS.U8. The
// profile of a pattern row comes from a 16-entry shared table, the selector of a text position is two IMADs.
// A thread whose 16-byte blocks held a byte > 3 re-checks its own strings exactly and sends both of its pairs
// to the int32 kernel through the worklist (dp_fast.h).
#include "dp_fast.h"

#define B2_THREADS 128
#define B2_INF2    0x83008300u    // packed -32000: F'' before any gap was opened (only ever max'ed)

struct Band2Params
{
    StrSet         P, T;
    nvb_best_sink* sinks;
    uint32_t*      worklist;
    uint32_t       n_pairs, n_threads;
    int32_t        alpha, beta, bias;   // H = stored - bias + alpha*M + beta*j
    int32_t        init0, init_step;    // stored H''_0[j] = init0 (j == 0), init_step_base + ... see kernel
    int32_t        init_rest;
    uint32_t       s_match, s_mis;      // profile bytes S' (SW) / S'' (Gotoh), 0..127
    int32_t        d;                   // Gotoh: go - ge (<= 0)
    uint32_t       one;
};

__device__ __forceinline__ uint32_t c_pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }
__device__ __forceinline__ uint32_t c_fix2(int32_t g) { return (((uint32_t)(g - 1) & 0xFFFFu) << 16) | ((uint32_t)g & 0xFFFFu); }
__device__ __forceinline__ uint32_t c_prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ uint32_t c_imad(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ int32_t c_lo(uint32_t x) { return (int32_t)(int16_t)(x & 0xFFFFu); }
__device__ __forceinline__ int32_t c_hi(uint32_t x) { return (int32_t)(int16_t)(x >> 16); }

// copies string bytes [c, c+span) (clipped to the string) into `slot` in 16-byte blocks, zero-fills the rest of
// the span, ORs every loaded word into `seen`; returns the offset of byte c inside the slot
template <int MAXBLK>
__device__ __forceinline__ uint32_t c_stage(const uint8_t* sym, uint32_t beg, uint32_t L, uint32_t c, uint32_t span,
                                            uint8_t* slot, uint32_t& seen)
{
    const uint8_t* a = sym + (size_t)beg + c;
    const uint32_t lead = (uint32_t)((uintptr_t)a & 15u);
    const uint32_t avail = L > c ? min(L - c, span) : 0u;
    const uint4* g = (const uint4*)(a - lead);
    const uint32_t nblk = avail ? (lead + avail + 15u) >> 4 : 0u;
    #pragma unroll
    for (int m = 0; m < MAXBLK; ++m)
    {
        if ((uint32_t)m < nblk)
        {
            const uint4 v = __ldg(g + m);
            seen |= v.x | v.y | v.z | v.w;
            uint32_t* s = (uint32_t*)(slot + 16 * m);
            s[0] = v.x; s[1] = v.y; s[2] = v.z; s[3] = v.w;
        }
    }
    for (uint32_t x = avail; x < span; ++x) slot[lead + x] = 0;
    return lead;
}

__device__ __noinline__ bool c_has_high_symbol(const uint8_t* sym, uint32_t beg, uint32_t L)
{
    for (uint32_t x = 0; x < L; ++x) if (sym[(size_t)beg + x] > 3u) return true;
    return false;
}

template <int W, int TYPE, bool GOTOH, int RU, int R>
__global__ void __launch_bounds__(B2_THREADS, 4)
band2_kernel(const Band2Params p)
{
    constexpr int PS = ((15 + R + 15) / 16) * 16;               // pattern slot: R rows at any 16-byte phase
    constexpr int TS = ((15 + R + W - 1 + 15) / 16) * 16;       // text slot: R + W - 1 positions
    constexpr int STRIDE = 2 * PS + 2 * TS + 4;
    static_assert((STRIDE & 7) == 4, "thread stride must be 4 mod 8 bytes");
    static_assert(R % RU == 0, "chunk must hold whole row groups");
    constexpr int WS = W + RU - 1;

    extern __shared__ __align__(16) uint8_t s_raw[];
    uint2* lut = (uint2*)s_raw;                                  // [qA + 4 qB] -> (profile A, profile B); byte offset qA*8 + qB*32
    if (threadIdx.x < 16)
    {
        const uint32_t qa = threadIdx.x & 3u, qb = threadIdx.x >> 2;
        const uint32_t base = p.s_mis * 0x01010101u, delta = p.s_mis ^ p.s_match;
        lut[threadIdx.x] = make_uint2(base ^ (delta << (8u * qa)), base ^ (delta << (8u * qb)));
    }
    __syncthreads();

    const uint32_t tid = blockIdx.x * B2_THREADS + threadIdx.x;
    if (tid >= p.n_threads) return;
    const uint32_t idA = 2 * tid;
    const bool hasB = idA + 1 < p.n_pairs;
    const uint32_t idB = hasB ? idA + 1 : idA;
    const uint32_t pbA = p.P.off[idA], MA = p.P.off[idA + 1] - pbA, pbB = p.P.off[idB], MB = p.P.off[idB + 1] - pbB;
    const uint32_t tbA = p.T.off[idA], NA = p.T.off[idA + 1] - tbA, tbB = p.T.off[idB], NB = p.T.off[idB + 1] - tbB;
    const uint32_t Mmax = max(MA, MB), Mmin = min(MA, MB);
    const uint32_t rows = ((Mmax + RU - 1) / RU) * RU;

    uint8_t* slot = s_raw + 128 + (size_t)threadIdx.x * STRIDE;
    uint8_t* const sPA = slot, * const sPB = slot + PS, * const sTA = slot + 2 * PS, * const sTB = slot + 2 * PS + TS;
    const uint8_t* lutb = s_raw;

    const uint32_t dfix = p.d ? c_fix2(p.d) : 0u;        // c_fix2(0) would borrow from the high half
    uint32_t X[W], F[W], fin[W], win[WS];
    #pragma unroll
    for (int j = 0; j < W; ++j)
    {
        X[j] = c_pack2(j == 0 ? p.init0 : p.init_rest + j * p.init_step);
        F[j] = B2_INF2; fin[j] = X[j];
    }
    #pragma unroll
    for (int j = 0; j < WS; ++j) win[j] = 0;

    uint32_t seen = 0;
    uint32_t nextcap = Mmin;
    for (uint32_t c = 0; c < rows; c += R)
    {
        const uint8_t* pA = sPA + c_stage<PS / 16>(p.P.sym, pbA, MA, c, R, sPA, seen);
        const uint8_t* pB = sPB + c_stage<PS / 16>(p.P.sym, pbB, MB, c, R, sPB, seen);
        const uint8_t* tA = sTA + c_stage<TS / 16>(p.T.sym, tbA, NA, c, R + W - 1, sTA, seen);
        const uint8_t* tB = sTB + c_stage<TS / 16>(p.T.sym, tbB, NB, c, R + W - 1, sTB, seen);
        if (c == 0)
        {
            #pragma unroll
            for (int j = 0; j < W - 1; ++j) win[j] = c_imad(tB[j], 0x1100u, c_imad(tA[j], 0x11u, 0xC480u));
        }
        const uint32_t gend = min((uint32_t)R, rows - c);
        for (uint32_t g = 0; g < gend; g += RU, pA += RU, pB += RU, tA += RU, tB += RU)
        {
            #pragma unroll
            for (int r = 0; r < RU; ++r)
            {
                const uint32_t i = c + g + r;
                const uint2 pq = *(const uint2*)(lutb + c_imad(pB[r], 32u, (uint32_t)pA[r] * 8u));
                win[r + W - 1] = c_imad(tB[r + W - 1], 0x1100u, c_imad(tA[r + W - 1], 0x11u, 0xC480u));
                uint32_t left = 0, E = 0;
                #pragma unroll
                for (int j = 0; j < W; ++j)
                {
                    const uint32_t S = c_prmt(pq.x, pq.y, win[r + j]);
                    const uint32_t dg = X[j] + S;
                    uint32_t h;
                    if (!GOTOH)
                    {
                        if (j == 0)          h = __vmaxs2(dg, X[1 < W ? 1 : 0]);
                        else if (j < W - 1)  h = __vimax3_s16x2(dg, X[j + 1 < W ? j + 1 : j], left);
                        else                 h = __vmaxs2(dg, left);
                        X[j] = h; left = h;
                    }
                    else
                    {
                        if (j < W - 1) F[j] = __vmaxs2(F[j + 1 < W ? j + 1 : j], X[j + 1 < W ? j + 1 : j]);
                        else           F[j] = B2_INF2;
                        if (j == 0)          h = __vmaxs2(dg, F[0]);
                        else if (j < W - 1)  h = __vimax3_s16x2(dg, F[j], E);
                        else                 h = __vmaxs2(dg, E);
                        const uint32_t ho = c_imad(h, p.one, dfix);
                        E = (j == 0) ? ho : __vmaxs2(E, ho);
                        X[j] = ho;
                    }
                }
                if (i + 1 == nextcap)
                {
                    const uint32_t m = (i + 1 == MA ? 0x0000FFFFu : 0u) | (i + 1 == MB ? 0xFFFF0000u : 0u);
                    #pragma unroll
                    for (int j = 0; j < W; ++j) fin[j] = (X[j] & m) | (fin[j] & ~m);
                    nextcap = (nextcap == Mmax) ? 0xFFFFFFFFu : Mmax;
                }
            }
            #pragma unroll
            for (int j = 0; j < W - 1; ++j) win[j] = win[j + RU];
        }
    }

    // ---- epilogue
    bool dirty = MA == 0 || MB == 0;
    if (seen & 0xFCFCFCFCu)
        dirty = dirty || c_has_high_symbol(p.P.sym, pbA, MA) || c_has_high_symbol(p.P.sym, pbB, MB) ||
                c_has_high_symbol(p.T.sym, tbA, NA) || c_has_high_symbol(p.T.sym, tbB, NB);
    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        if (dirty)
        {
            const uint32_t at = atomicAdd(p.worklist, 1u);
            p.worklist[1 + at] = id;
            continue;
        }
        nvb_best_sink out;
        out.score = NVB_SINK_MIN; out.sink_x = out.sink_y = 0xFFFFFFFFu;
        if (N >= M)                                           // shorter texts are refused (sw_banded_inl.h:356-357)
        {
            const int32_t back = p.alpha * (int32_t)M - p.bias - (GOTOH ? p.d : 0);
            if (TYPE == T_GLOBAL)
            {
                out.score = (h ? c_hi(fin[W - 1]) : c_lo(fin[W - 1])) + back + p.beta * (W - 1);
                out.sink_x = M + W - 1; out.sink_y = M;
            }
            else
            {
                const uint32_t lim = (M + W - 1 < N ? M + W - 1 : N) - (M - 1);
                #pragma unroll
                for (int j = 0; j < W; ++j)
                {
                    const int32_t s = (h ? c_hi(fin[j]) : c_lo(fin[j])) + back + p.beta * j;
                    if ((j == 0 || (uint32_t)j < lim) && out.score <= s) { out.score = s; out.sink_x = M + j; out.sink_y = M; }
                }
            }
        }
        p.sinks[id] = out;
    }
}

// ---------------------------------------------------------------------------------------------
static int32_t iabs2(int32_t v) { return v < 0 ? -v : v; }

struct Band2Plan { Band2Params p; bool ok; };

static Band2Plan band2_plan(const nvb_aligner& a, uint32_t band, uint32_t maxM)
{
    Band2Plan pl; pl.ok = false;
    Band2Params& p = pl.p;
    const Scheme s = make_scheme(a);
    const bool g = a.kind == NVB_GOTOH;
    if (a.type == NVB_LOCAL) return pl;
    if (band != 7 && band != 15) return pl;
    if (maxM == 0) return pl;
    const int32_t W = (int32_t)band;
    int32_t smin, lo0, hi0;          // smallest diagonal increment; range of the true first row H''_0
    if (!g)
    {
        if (s.ga >= 0 || s.gb >= 0) return pl;
        p.beta = s.gb; p.alpha = s.ga + s.gb; p.d = 0;
        const int32_t sm = s.match - p.alpha, sx = s.mismatch - p.alpha;
        if (sm < 0 || sm > 127 || sx < 0 || sx > 127) return pl;
        p.s_match = (uint32_t)sm; p.s_mis = (uint32_t)sx; smin = 0;
        // H_0[j] = j*ga (GLOBAL) or 0 (sw_banded_inl.h:35-45)  ->  H''_0[j] = H_0[j] - beta*j
        const int32_t step = (a.type == NVB_GLOBAL ? s.ga : 0) - p.beta;
        p.init0 = 0; p.init_rest = 0; p.init_step = step;
        lo0 = step < 0 ? step * (W - 1) : 0; hi0 = step > 0 ? step * (W - 1) : 0;
    }
    else
    {
        if (s.ga > s.gb || s.gb >= 0) return pl;                       // open at least as costly as extend, both penalties
        p.beta = s.gb; p.alpha = 2 * s.gb; p.d = s.ga - s.gb;
        const int32_t sm = s.match - s.ga - s.gb, sx = s.mismatch - s.ga - s.gb;
        if (sm < 0 || sm > 127 || sx < 0 || sx > 127) return pl;
        p.s_match = (uint32_t)sm; p.s_mis = (uint32_t)sx;
        smin = min(0, min(s.match, s.mismatch) - p.alpha);              // H'' may sink along mismatching diagonals
        // H_0[0] = 0, H_0[j] = go + (j-1) ge (GLOBAL) or 0 (gotoh_banded_inl.h:37-68)
        if (a.type == NVB_GLOBAL) { p.init0 = 0; p.init_rest = s.ga - s.gb; p.init_step = 0; lo0 = p.d; hi0 = 0; }
        else                      { p.init0 = 0; p.init_rest = 0; p.init_step = -p.beta; lo0 = 0; hi0 = -p.beta * (W - 1); }
    }
    const int64_t low = (int64_t)lo0 + (int64_t)smin * maxM;            // lower bound of every true H''
    p.bias = (int32_t)(-low) + iabs2(p.d) + 1;
    const int64_t high = (int64_t)p.bias + hi0 + (int64_t)max(p.s_match, p.s_mis) * maxM + 256;
    if (p.bias > 30000 || high > 32000) return pl;
    // the band is kept biased; for Gotoh as HO = H'' + d
    p.init0 += p.bias + p.d; p.init_rest += p.bias + p.d;
    p.one = 1;
    pl.ok = true;
    return pl;
}

bool dpb2_score_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN)
{
    (void)maxN;
    if (P.bits != 8 || T.bits != 8) return false;
    return band2_plan(a, band, maxM).ok;
}

template <int W, int TYPE, bool GOTOH>
static int band2_launch(nvb_ctx* ctx, const Band2Params& p, cudaStream_t st)
{
    constexpr int RU = 8, R = 48;
    constexpr int PS = ((15 + R + 15) / 16) * 16, TS = ((15 + R + W - 1 + 15) / 16) * 16;
    constexpr size_t smem = 128 + (size_t)B2_THREADS * (2 * PS + 2 * TS + 4);
    static bool attr_set = false;
    if (!attr_set)
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(band2_kernel<W, TYPE, GOTOH, RU, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    band2_kernel<W, TYPE, GOTOH, RU, R><<<(p.n_threads + B2_THREADS - 1) / B2_THREADS, B2_THREADS, smem, st>>>(p);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

template <int W>
static int band2_dispatch(nvb_ctx* ctx, const nvb_aligner& a, const Band2Params& p, cudaStream_t st)
{
    const bool g = a.kind == NVB_GOTOH;
    if (a.type == NVB_GLOBAL) return g ? band2_launch<W, T_GLOBAL, true>(ctx, p, st) : band2_launch<W, T_GLOBAL, false>(ctx, p, st);
    return g ? band2_launch<W, T_SEMI, true>(ctx, p, st) : band2_launch<W, T_SEMI, false>(ctx, p, st);
}

int dpb2_score(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, nvb_best_sink* sinks,
               uint32_t maxM, uint32_t maxN, uint32_t* worklist, cudaStream_t st)
{
    (void)maxN;
    Band2Plan pl = band2_plan(a, band, maxM);
    if (!pl.ok) return NVB_ERR_UNSUPPORTED;
    Band2Params& p = pl.p;
    p.P = P; p.T = T; p.sinks = sinks; p.worklist = worklist;
    p.n_pairs = P.count; p.n_threads = (P.count + 1) / 2;
    if (band == 7)  return band2_dispatch<7>(ctx, a, p, st);
    if (band == 15) return band2_dispatch<15>(ctx, a, p, st);
    return NVB_ERR_UNSUPPORTED;
}
