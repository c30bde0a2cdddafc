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
// Staging: pattern rows and text positions are consumed front to back, one byte per row, so each of a thread's
// four strings streams through a private 64-byte shared-memory ring (struct Stream below): one 16-byte global load
// per string and row group, issued a whole group ahead of its use, and immediate-offset LDS.U8 in the row loop.
// The profile of a pattern row comes from a 16-entry shared table, the selector of a text position is two IMADs.
// A thread whose 16-byte blocks held a byte > 3 re-checks its own strings exactly and sends both of its pairs
// to the int32 kernel through the worklist (dp_fast.h).
#include "dp_fast.h"
#include <cstdlib>

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
    uint32_t       one, k8, k32;        // runtime 1, 8, 32
};

__device__ __forceinline__ uint32_t c_pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }
__device__ __forceinline__ uint32_t c_fix2(int32_t g) { return (((uint32_t)(g - 1) & 0xFFFFu) << 16) | ((uint32_t)g & 0xFFFFu); }
__device__ __forceinline__ uint32_t c_prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ uint32_t c_imad(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ int32_t c_lo(uint32_t x) { return (int32_t)(int16_t)(x & 0xFFFFu); }
__device__ __forceinline__ int32_t c_hi(uint32_t x) { return (int32_t)(int16_t)(x >> 16); }

__device__ __noinline__ bool c_has_high_symbol(const uint8_t* sym, uint32_t beg, uint32_t L)
{
    for (uint32_t x = 0; x < L; ++x) if (sym[(size_t)beg + x] > 3u) return true;
    return false;
}

__device__ __forceinline__ uint2 c_lds64(uint32_t addr)
{ uint2 v; asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr)); return v; }

// ---- byte streams. Each of a thread's four strings (pattern A/B; text A/B from position W-1 on) is consumed one
// byte per row. It is kept as a ring of three aligned 16-byte blocks of the symbol array plus a copy of the first
// slot after the last, so that the 16 bytes of a row group are contiguous wherever they start. While a group is
// computed, the block after the two resident ones sits in registers (one LDG.128 per stream and group, issued
// before the group, stored after it): the load latency is covered by the thread's own arithmetic.
struct Stream
{
    const uint8_t* base;   // symbol array rounded down to 16 bytes
    uint32_t       nb;     // offset (from base) of the next block to fetch = block of the current byte + 32
    uint32_t       end;    // offset of the end of the string
    uint8_t*       ring;   // 64 bytes of shared memory
    const uint8_t* ptr;    // shared-memory address of the byte the next row reads: ring <= ptr < ring + 48
};

// the aligned block at offset bstart, or zeros when it holds no byte of the string (nothing past the string's last
// block is ever read). volatile + "memory": the load is issued where it is written, a whole row group ahead of c_put
__device__ __forceinline__ uint4 c_load(const Stream& s, uint32_t bstart)
{
    uint4 v;
    const uint32_t ok = bstart < s.end ? 1u : 0u;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %5, 0;\n\tmov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\tmov.u32 %2, 0;\n\tmov.u32 %3, 0;\n\t"
                 "@p ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "=&r"(v.x), "=&r"(v.y), "=&r"(v.z), "=&r"(v.w) : "l"(s.base + bstart), "r"(ok) : "memory");
    return v;
}
// zeroes the bytes of block bstart that lie past the end of the string
__device__ __forceinline__ uint4 c_clip(const Stream& s, uint32_t bstart, uint4 v)
{
    if (bstart < s.end && s.end - bstart < 16u)
    {
        const uint32_t n = s.end - bstart;
        uint32_t w[4] = { v.x, v.y, v.z, v.w };
        #pragma unroll
        for (int k = 0; k < 4; ++k)
        {
            const uint32_t keep = n > 4u * k ? min(n - 4u * k, 4u) : 0u;
            w[k] = keep == 4u ? w[k] : (keep == 0u ? 0u : (w[k] & ((1u << (8u * keep)) - 1u)));
        }
        v = make_uint4(w[0], w[1], w[2], w[3]);
    }
    return v;
}
// stores a block in ring slot slot_off (slot 0 is mirrored after slot 2)
__device__ __forceinline__ void c_put(uint8_t* ring, uint32_t slot_off, const uint4 v)
{
    uint32_t* d = (uint32_t*)(ring + slot_off);
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    if (slot_off == 0u) { d[12] = v.x; d[13] = v.y; d[14] = v.z; d[15] = v.w; }
}
// skip = bytes of the string that are not streamed (the text positions the window starts with)
__device__ __forceinline__ void c_open(Stream& s, const uint8_t* sym, uint32_t beg, uint32_t L, uint32_t skip, uint8_t* ring, uint32_t& seen)
{
    const uint32_t delta = (uint32_t)((uintptr_t)sym & 15u);
    s.base = sym - delta;
    const uint32_t pos = delta + beg + skip, b0 = pos & ~15u;
    s.end = delta + beg + L; s.nb = b0 + 32u;
    s.ring = ring; s.ptr = ring + (pos & 15u);
    const uint4 v0 = c_load(s, b0), v1 = c_load(s, b0 + 16u);
    seen |= v0.x | v0.y | v0.z | v0.w | v1.x | v1.y | v1.z | v1.w;      // neighbours' bytes included: a filter, confirmed exactly later
    c_put(ring, 0u, c_clip(s, b0, v0));
    c_put(ring, 16u, c_clip(s, b0 + 16u, v1));
}
// one byte on (single rows): fetches synchronously when the stream enters its next block
__device__ __forceinline__ void c_step(Stream& s, uint32_t& seen)
{
    const uint32_t at = (uint32_t)(s.ptr - s.ring);
    if ((at & 15u) == 15u)
    {
        const uint4 v = c_load(s, s.nb);
        seen |= v.x | v.y | v.z | v.w;
        const uint32_t o = at & 0x30u;                                  // slot of the current block: 0, 16, 32
        c_put(s.ring, o == 0u ? 32u : o - 16u, c_clip(s, s.nb, v));     // slot (o + 32) mod 48
        s.nb += 16u;
        if (o == 32u) s.ptr -= 48;
    }
    ++s.ptr;
}

// one pattern row: cell j reads window slot (R0 + j) % WS
template <int W, bool GOTOH, int WS, int R0>
__device__ __forceinline__ void c_row(uint32_t (&X)[W], uint32_t (&F)[W], const uint32_t (&win)[WS], const uint2 pq,
                                      const uint32_t dfix, const uint32_t one)
{
    uint32_t left = 0, E = 0;
    #pragma unroll
    for (int j = 0; j < W; ++j)
    {
        const uint32_t S = c_prmt(pq.x, pq.y, win[(R0 + j) % WS]);
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
            const uint32_t ho = c_imad(h, one, dfix);
            E = (j == 0) ? ho : __vmaxs2(E, ho);
            X[j] = ho;
        }
    }
}

// rows R0 .. RU-1 of a group
template <int W, int RU, int R0>
struct RowUnroll
{
    template <bool GOTOH, int WS>
    static __device__ __forceinline__ void run(uint32_t (&X)[W], uint32_t (&F)[W], uint32_t (&win)[WS], const uint8_t* pA, const uint8_t* pB,
                                               const uint8_t* tA, const uint8_t* tB, uint32_t lut_addr, uint32_t dfix, uint32_t one,
                                               uint32_t k8, uint32_t k32)
    {
        const uint2 pq = c_lds64(c_imad(pA[R0], k8, c_imad(pB[R0], k32, lut_addr)));     // runtime 8 and 32: IMADs stay IMADs (FMA pipe)
        win[(R0 + W - 1) % WS] = c_imad(tB[R0], 0x1100u, c_imad(tA[R0], 0x11u, 0xC480u));
        c_row<W, GOTOH, WS, R0>(X, F, win, pq, dfix, one);
        RowUnroll<W, RU, R0 + 1>::template run<GOTOH, WS>(X, F, win, pA, pB, tA, tB, lut_addr, dfix, one, k8, k32);
    }
};
template <int W, int RU>
struct RowUnroll<W, RU, RU>
{
    template <bool GOTOH, int WS>
    static __device__ __forceinline__ void run(uint32_t (&)[W], uint32_t (&)[W], uint32_t (&)[WS], const uint8_t*, const uint8_t*,
                                               const uint8_t*, const uint8_t*, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {}
};

template <int W, int TYPE, bool GOTOH, int MINB>
__global__ void __launch_bounds__(B2_THREADS, MINB)
band2_kernel(const Band2Params p)
{
    constexpr int WS = W + 1;                                    // circular window of text selectors; one group = WS rows
    constexpr int RU = WS;
    static_assert(RU <= 16, "a row group may not outrun a 16-byte block");
    constexpr int STRIDE = 4 * 64 + 4;                           // 4 mod 8 bytes: the slots of a warp start in distinct banks

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

    uint8_t* slot = s_raw + 128 + (size_t)threadIdx.x * STRIDE;
    const uint32_t lut_addr = (uint32_t)__cvta_generic_to_shared(s_raw);
    const uint32_t dfix = p.d ? c_fix2(p.d) : 0u;        // c_fix2(0) would borrow from the high half

    uint32_t seen = 0;
    Stream sPA, sPB, sTA, sTB;
    c_open(sPA, p.P.sym, pbA, MA, 0, slot, seen);
    c_open(sPB, p.P.sym, pbB, MB, 0, slot + 64, seen);
    c_open(sTA, p.T.sym, tbA, NA, W - 1, slot + 128, seen);
    c_open(sTB, p.T.sym, tbB, NB, W - 1, slot + 192, seen);

    uint32_t X[W], F[W], win[WS];
    #pragma unroll
    for (int j = 0; j < W; ++j) { X[j] = c_pack2(j == 0 ? p.init0 : p.init_rest + j * p.init_step); F[j] = B2_INF2; }
    win[WS - 1] = win[WS - 2] = 0;
    #pragma unroll
    for (int j = 0; j < W - 1; ++j)                       // text positions 0 .. W-2 (their bytes also pass the filter: first blocks)
    {
        const uint32_t ra = (uint32_t)j < NA ? __ldg(p.T.sym + (size_t)tbA + j) : 0u;
        const uint32_t rb = (uint32_t)j < NB ? __ldg(p.T.sym + (size_t)tbB + j) : 0u;
        seen |= ra | rb;
        win[j] = c_imad(rb, 0x1100u, c_imad(ra, 0x11u, 0xC480u));
    }

    int32_t scoreA = NVB_SINK_MIN, scoreB = NVB_SINK_MIN;
    uint32_t sxA = 0xFFFFFFFFu, sxB = 0xFFFFFFFFu;
    uint32_t i = 0;

    #pragma unroll 1
    for (int seg = 0; seg < 2; ++seg)
    {
        const uint32_t target = seg ? Mmax : Mmin;
        // groups of RU rows, all four streams in step (only before the first single row, i.e. in the first segment):
        // the window rotates in place and is linear again (slot s <-> text position i + s) after each group
        if (seg == 0)
        {
            uint32_t o = 0;                                    // ring slot of the streams' current blocks: 0, 16, 32
            #pragma unroll 1
            for (uint32_t g = target / RU; g > 0; --g)
            {
                const uint4 vPA = c_load(sPA, sPA.nb), vPB = c_load(sPB, sPB.nb), vTA = c_load(sTA, sTA.nb), vTB = c_load(sTB, sTB.nb);
                RowUnroll<W, RU, 0>::template run<GOTOH, WS>(X, F, win, sPA.ptr, sPB.ptr, sTA.ptr, sTB.ptr, lut_addr, dfix, p.one, p.k8, p.k32);
                seen |= vPA.x | vPA.y | vPA.z | vPA.w | vPB.x | vPB.y | vPB.z | vPB.w;
                seen |= vTA.x | vTA.y | vTA.z | vTA.w | vTB.x | vTB.y | vTB.z | vTB.w;
                // bytes past the end of a string are read: text right of the band's last cells, and the shorter pattern of
                // a couple while the longer one goes on; they must be symbols (the profile table is indexed by both patterns)
                const uint4 cPA = c_clip(sPA, sPA.nb, vPA), cPB = c_clip(sPB, sPB.nb, vPB);
                const uint4 cTA = c_clip(sTA, sTA.nb, vTA), cTB = c_clip(sTB, sTB.nb, vTB);
                if (RU == 16)
                {
                    const uint32_t so = o == 0u ? 32u : o - 16u;   // (o + 32) mod 48
                    c_put(sPA.ring, so, cPA); c_put(sPB.ring, so, cPB); c_put(sTA.ring, so, cTA); c_put(sTB.ring, so, cTB);
                    const int32_t step = o == 32u ? -32 : 16;
                    o = o == 32u ? 0u : o + 16u;
                    sPA.ptr += step; sPB.ptr += step; sTA.ptr += step; sTB.ptr += step;
                    sPA.nb += 16u; sPB.nb += 16u; sTA.nb += 16u; sTB.nb += 16u;
                }
                else
                {
                    // shorter groups: a stream enters its next block only every other group or so
                    Stream* st[4] = { &sPA, &sPB, &sTA, &sTB };
                    const uint4 vv[4] = { cPA, cPB, cTA, cTB };
                    #pragma unroll
                    for (int k = 0; k < 4; ++k)
                    {
                        Stream& t = *st[k];
                        const uint32_t at = (uint32_t)(t.ptr - t.ring), ob = at & 0x30u;
                        c_put(t.ring, ob == 0u ? 32u : ob - 16u, vv[k]);
                        if ((at & 15u) + RU >= 16u) { t.nb += 16u; if (at + RU >= 48u) t.ptr -= 48; }
                        t.ptr += RU;
                    }
                }
                i += RU;
            }
        }
        // single rows (the window is shifted down after each and stays linear)
        #pragma unroll 1
        while (i < target)
        {
            RowUnroll<W, 1, 0>::template run<GOTOH, WS>(X, F, win, sPA.ptr, sPB.ptr, sTA.ptr, sTB.ptr, lut_addr, dfix, p.one, p.k8, p.k32);
            #pragma unroll
            for (int j = 0; j < W - 1; ++j) win[j] = win[j + 1];
            c_step(sPA, seen); c_step(sPB, seen); c_step(sTA, seen); c_step(sTB, seen);
            ++i;
        }
        // the pairs that end here: read their result off the band (H = stored - bias + alpha*M + beta*j)
        #pragma unroll
        for (int h = 0; h < 2; ++h)
        {
            const uint32_t M = h ? MB : MA, N = h ? NB : NA;
            if (M != target || (seg == 1 && Mmin == Mmax) || M == 0 || N < M) continue;   // shorter texts are refused (sw_banded_inl.h:356-357)
            const int32_t back = p.alpha * (int32_t)M - p.bias - (GOTOH ? p.d : 0);
            int32_t best = NVB_SINK_MIN; uint32_t bx = 0xFFFFFFFFu;
            if (TYPE == T_GLOBAL) { best = (h ? c_hi(X[W - 1]) : c_lo(X[W - 1])) + back + p.beta * (W - 1); bx = M + W - 1; }
            else
            {
                const uint32_t lim = (M + W - 1 < N ? M + W - 1 : N) - (M - 1);
                #pragma unroll
                for (int j = 0; j < W; ++j)
                {
                    const int32_t s = (h ? c_hi(X[j]) : c_lo(X[j])) + back + p.beta * j;
                    if ((j == 0 || (uint32_t)j < lim) && best <= s) { best = s; bx = M + j; }
                }
            }
            if (h) { scoreB = best; sxB = bx; } else { scoreA = best; sxA = bx; }
        }
    }

    // ---- epilogue
    bool dirty = MA == 0 || MB == 0;
    // GLOBAL reports the band's last cell; when that lies right of the text the reference's score depends on the 255s it
    // reads there (sw_banded_inl.h:444), which this kernel's text stream does not reproduce
    if (TYPE == T_GLOBAL) dirty = dirty || NA < MA + W - 1 || NB < MB + W - 1;
    if (seen & 0xFCFCFCFCu)
        dirty = dirty || c_has_high_symbol(p.P.sym, pbA, MA) || c_has_high_symbol(p.P.sym, pbB, MB) ||
                c_has_high_symbol(p.T.sym, tbA, NA) || c_has_high_symbol(p.T.sym, tbB, NB);
    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        if (dirty)
        {
            const uint32_t at = atomicAdd(p.worklist, 1u);
            p.worklist[1 + at] = id;
            continue;
        }
        nvb_best_sink out;
        out.score = h ? scoreB : scoreA; out.sink_x = h ? sxB : sxA;
        out.sink_y = out.sink_x == 0xFFFFFFFFu ? 0xFFFFFFFFu : (h ? MB : MA);
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
    p.one = 1; p.k8 = 8; p.k32 = 32;
    pl.ok = true;
    return pl;
}

bool dpb2_score_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN)
{
    (void)maxN;
    if (P.bits != 8 || T.bits != 8) return false;
    return band2_plan(a, band, maxM).ok;
}

template <int W, int TYPE, bool GOTOH, int MINB>
static int band2_launch_as(nvb_ctx* ctx, const Band2Params& p, cudaStream_t st)
{
    constexpr size_t smem = 128 + (size_t)B2_THREADS * (4 * 64 + 4);
    static bool attr_set = false;
    if (!attr_set)
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(band2_kernel<W, TYPE, GOTOH, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    band2_kernel<W, TYPE, GOTOH, MINB><<<(p.n_threads + B2_THREADS - 1) / B2_THREADS, B2_THREADS, smem, st>>>(p);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
template <int W, int TYPE, bool GOTOH>
static int band2_launch(nvb_ctx* ctx, const Band2Params& p, cudaStream_t st)
{
    static int minb = -1;
    if (minb < 0) { const char* e = getenv("NVB_B2_MINB"); minb = e ? atoi(e) : 4; }
    if (minb == 6) return band2_launch_as<W, TYPE, GOTOH, 6>(ctx, p, st);
    if (minb == 4) return band2_launch_as<W, TYPE, GOTOH, 4>(ctx, p, st);
    return band2_launch_as<W, TYPE, GOTOH, 5>(ctx, p, st);
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
