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
#include "dp_band2.cuh"

// per-launch constants of the row loop, and the running arg-max of LOCAL alignments
struct RowConst
{
    uint32_t lut_addr, dfix, one, k8, k32;
    uint32_t Af, Bf, ext2, B2, sixteen;      // LOCAL (unsheared, biased): top/open addend, left addend, packed extension, packed zero, runtime 16
};
struct LocalBest { uint32_t key, row, rmask; };   // packed (biased score << 4 | band cell), packed pattern rows, halves still running

// one pattern row: cell j reads window slot (R0 + j) % WS
template <int W, int TYPE, bool GOTOH, int WS, int R0>
__device__ __forceinline__ void c_row(uint32_t (&X)[W], uint32_t (&F)[W], const uint32_t (&win)[WS], const uint2 pq,
                                      const RowConst& rc, LocalBest& lb, const uint32_t i)
{
    uint32_t left = 0, E = 0;
    if (TYPE != T_LOCAL)
    {
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
                const uint32_t ho = c_imad(h, rc.one, rc.dfix);
                E = (j == 0) ? ho : __vmaxs2(E, ho);
                X[j] = ho;
            }
        }
    }
    else
    {
        // LOCAL: unsheared scores biased by the largest gap penalty (zero = B2), so that the gap additions are plain IMADs;
        // every cell is reported: rows ascending, band cells ascending, the last maximum wins (sw_banded_inl.h:407-456)
        uint32_t key[W];
        #pragma unroll
        for (int j = 0; j < W; ++j)
        {
            const uint32_t S = c_prmt(pq.x, pq.y, win[(R0 + j) % WS]);
            uint32_t h;
            if (!GOTOH)
            {
                if (j == 0)
                {
                    const uint32_t u = c_imad(X[1 < W ? 1 : 0], rc.one, rc.Af);
                    h = __vmaxs2(__viaddmax_s16x2(X[0], S, u), rc.B2);
                }
                else if (j < W - 1)
                {
                    const uint32_t u = c_imad(X[j + 1 < W ? j + 1 : j], rc.one, rc.Af);
                    const uint32_t v = __viaddmax_s16x2(X[j], S, u);
                    h = __vimax3_s16x2(c_imad(left, rc.one, rc.Bf), v, rc.B2);
                }
                else h = __vmaxs2(__viaddmax_s16x2(X[j], S, c_imad(left, rc.one, rc.Bf)), rc.B2);
            }
            else
            {
                if (j < W - 1)
                {
                    const uint32_t o = c_imad(X[j + 1 < W ? j + 1 : j], rc.one, rc.Af);
                    F[j] = __viaddmax_s16x2(F[j + 1 < W ? j + 1 : j], rc.ext2, o);        // max(F + ext, H + open)
                    const uint32_t v = __viaddmax_s16x2(X[j], S, F[j]);
                    h = (j == 0) ? __vmaxs2(v, rc.B2) : __vimax3_s16x2(v, E, rc.B2);
                    const uint32_t o2 = c_imad(h, rc.one, rc.Af);
                    E = (j == 0) ? o2 : __viaddmax_s16x2(E, rc.ext2, o2);                  // max(E + ext, H + open)
                }
                else
                {
                    F[j] = B2_INF2;
                    h = __vmaxs2(__viaddmax_s16x2(X[j], S, E), rc.B2);
                }
            }
            X[j] = h; left = h;
            key[j] = c_imad(h, rc.sixteen, (uint32_t)j * 0x00010001u);
        }
        // later rows win ties: compare against the best with its cell bits cleared, minus one
        const uint32_t kb = __vadd2(lb.key & 0xFFF0FFF0u, 0xFFFFFFFFu);
        uint32_t acc = kb;
        #pragma unroll
        for (int j = 0; j + 1 < W; j += 2) acc = __vimax3_s16x2(acc, key[j], key[j + 1]);
        if (W & 1) acc = __vmaxs2(acc, key[W - 1]);
        const uint32_t d = __vsub2(kb, acc);                              // < 0 in a half that improved
        const uint32_t m = c_prmt(d, d, 0xBB99u) & lb.rmask;             // 0xFFFF per improved, running half
        lb.key = (acc & m) | (lb.key & ~m);
        lb.row = ((i * 0x00010001u) & m) | (lb.row & ~m);
    }
}

// rows R0 .. RU-1 of a group starting at pattern row i0
template <int W, int RU, int R0>
struct RowUnroll
{
    template <int TYPE, bool GOTOH, int WS>
    static __device__ __forceinline__ void run(uint32_t (&X)[W], uint32_t (&F)[W], uint32_t (&win)[WS], const uint8_t* pA, const uint8_t* pB,
                                               const uint8_t* tA, const uint8_t* tB, const RowConst& rc, LocalBest& lb, uint32_t i0)
    {
        const uint2 pq = c_lds64(c_imad(pA[R0], rc.k8, c_imad(pB[R0], rc.k32, rc.lut_addr)));     // runtime 8 and 32: IMADs stay IMADs (FMA pipe)
        win[(R0 + W - 1) % WS] = c_imad(tB[R0], 0x1100u, c_imad(tA[R0], 0x11u, 0xC480u));
        c_row<W, TYPE, GOTOH, WS, R0>(X, F, win, pq, rc, lb, i0 + R0);
        RowUnroll<W, RU, R0 + 1>::template run<TYPE, GOTOH, WS>(X, F, win, pA, pB, tA, tB, rc, lb, i0);
    }
};
template <int W, int RU>
struct RowUnroll<W, RU, RU>
{
    template <int TYPE, bool GOTOH, int WS>
    static __device__ __forceinline__ void run(uint32_t (&)[W], uint32_t (&)[W], uint32_t (&)[WS], const uint8_t*, const uint8_t*,
                                               const uint8_t*, const uint8_t*, const RowConst&, LocalBest&, uint32_t) {}
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
    RowConst rc;
    rc.lut_addr = (uint32_t)__cvta_generic_to_shared(s_raw);
    rc.dfix = p.d ? c_fix2(p.d) : 0u;                    // c_fix2(0) would borrow from the high half
    rc.one = p.one; rc.k8 = p.k8; rc.k32 = p.k32;
    rc.Af = c_fix2(p.ga); rc.Bf = c_fix2(p.gb); rc.ext2 = c_pack2(p.gb); rc.B2 = c_pack2(p.bias); rc.sixteen = p.k8 * 2u;
    LocalBest lb; lb.key = rc.B2 << 4; lb.row = 0; lb.rmask = 0xFFFFFFFFu;

    uint32_t seen = 0;
    Stream sPA, sPB, sTA, sTB;
    const uint32_t Ptot = __ldg(p.P.off + p.P.count), Ttot = __ldg(p.T.off + p.T.count);
    bool edge = c_open(sPA, p.P.sym, Ptot, pbA, MA, 0, slot, seen);
    edge |= c_open(sPB, p.P.sym, Ptot, pbB, MB, 0, slot + 64, seen);
    edge |= c_open(sTA, p.T.sym, Ttot, tbA, NA, W - 1, slot + 128, seen);
    edge |= c_open(sTB, p.T.sym, Ttot, tbB, NB, W - 1, slot + 192, seen);

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
    uint32_t sxA = 0xFFFFFFFFu, sxB = 0xFFFFFFFFu, syA = 0xFFFFFFFFu, syB = 0xFFFFFFFFu;
    uint32_t i = 0;

    #pragma unroll 1
    for (int seg = 0; seg < 2; ++seg)
    {
        const uint32_t target = seg ? Mmax : Mmin;
        if (seg) lb.rmask = MA > MB ? 0x0000FFFFu : 0xFFFF0000u;     // only the couple's longer pattern goes on
        // groups of RU rows, all four streams in step (only before the first single row, i.e. in the first segment):
        // the window rotates in place and is linear again (slot s <-> text position i + s) after each group
        if (seg == 0)
        {
            uint32_t o = 0;                                    // ring slot of the streams' current blocks: 0, 16, 32
            #pragma unroll 1
            for (uint32_t g = target / RU; g > 0; --g)
            {
                const uint4 vPA = c_load(sPA, sPA.nb), vPB = c_load(sPB, sPB.nb), vTA = c_load(sTA, sTA.nb), vTB = c_load(sTB, sTB.nb);
                RowUnroll<W, RU, 0>::template run<TYPE, GOTOH, WS>(X, F, win, sPA.ptr, sPB.ptr, sTA.ptr, sTB.ptr, rc, lb, i);
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
            RowUnroll<W, 1, 0>::template run<TYPE, GOTOH, WS>(X, F, win, sPA.ptr, sPB.ptr, sTA.ptr, sTB.ptr, rc, lb, i);
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
            if (TYPE == T_LOCAL)
            {
                const uint32_t k = h ? (lb.key >> 16) : (lb.key & 0xFFFFu), row = h ? (lb.row >> 16) : (lb.row & 0xFFFFu);
                const int32_t best = (int32_t)(k >> 4) - p.bias;
                const uint32_t bx = row + (k & 15u) + 1u;
                if (h) { scoreB = best; sxB = bx; syB = row + 1u; } else { scoreA = best; sxA = bx; syA = row + 1u; }
                continue;
            }
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
            if (h) { scoreB = best; sxB = bx; syB = M; } else { scoreA = best; sxA = bx; syA = M; }
        }
    }

    // ---- epilogue
    bool dirty = MA == 0 || MB == 0 || edge;     // edge: a string at the end of an unaligned symbol array (c_open)
    // GLOBAL reports the band's last cell; when that lies right of the text the reference's score depends on the 255s it
    // reads there (sw_banded_inl.h:444), which this kernel's text stream does not reproduce
    // LOCAL reports every band cell, the ones right of the text included (where the reference reads 255s as well)
    if (TYPE != T_SEMI) dirty = dirty || NA < MA + W - 1 || NB < MB + W - 1;
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
        out.sink_y = out.sink_x == 0xFFFFFFFFu ? 0xFFFFFFFFu : (h ? syB : syA);
        p.sinks[id] = out;
    }
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
    if (NVB_FIRST_ON_DEVICE(ctx))
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(band2_kernel<W, TYPE, GOTOH, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
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
    if (a.type == NVB_LOCAL)  return g ? band2_launch<W, T_LOCAL, true>(ctx, p, st) : band2_launch<W, T_LOCAL, false>(ctx, p, st);
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
