// dp_band2_tb.cu -- banded traceback (GLOBAL / SEMI_GLOBAL edit distance, Smith-Waterman, Gotoh; bands 7 and 15) on
// the sheared-coordinate packed kernel of dp_band2.cu.
// (replaces BatchedBandedAlignmentTraceback::enact, nvbio/alignment/batched_banded_inl.h:243-321,331-453, over
// banded_alignment_traceback, banded_inl.h:354-430, and the walkers sw_banded_inl.h:740-800, gotoh_banded_inl.h:883-950.)
//
// The reference scores the band once saving a checkpoint every CHK rows, then re-scores CHK-row windows backwards
// storing 2-bit (SW/ED) or 4-bit (Gotoh) flow cells and walks them. The directions only depend on comparisons between
// the three (SW) or two (E, F) candidates of a cell, so ONE forward pass can emit them -- provided the comparisons of
// both packed pairs come out of the s16x2 max instructions themselves. They do, by tagging: every score is kept
// times four and each candidate of a max carries its priority in the two low bits:
//     SW/ED   h' = max3(4(diag + S') + 2, 4 left + 1, 4 top + 0)          tag 2 = SUBSTITUTION, 1 = DELETION, 0 = INSERTION
// which is exactly `top > left ? (top > diag ? INS : SUB) : (left > diag ? DEL : SUB)` (sw_banded_inl.h:438-440:
// the diagonal wins ties, then the move along the text). The tag is cut off (LOP3 + subtract) before the value is
// used again and accumulated, 2 bits per cell and pair, into the row's flow words by IMADs. Gotoh tags three maxima
// per cell the same way: F'' = max(F0_top, HO1_top) (low bit 1 = gap opened), h' = max3(diag|2, E|1, F|0),
// E'' = max(E1, ho3) (bit 1 set = opened); one nibble per cell and pair.
// The flow of a launch goes to global scratch, [row][thread] so that a warp's stores are contiguous; the walk runs
// in a second kernel, one thread per pair, and pushes the ops end -> start as the reference's Backtracer receives
// them. Checkpoint distance is irrelevant here (no windows); the reference's int16 clamp of checkpoints
// (gotoh_banded_inl.h:225-230) only touches the -inf lanes of E/F, which never win a comparison against a real score.
#include "dp_band2.cuh"
#include "dp_launch.h"

struct Band2TbParams
{
    Band2Params    b;            // scores scaled by 4 (see band2_tb_plan)
    nvb_alignment* out;
    uint32_t*      n_ops;
    uint8_t*       flow;         // [row][thread] of uint2 (SW/ED) or uint4 (Gotoh)
    uint32_t       first;        // first pair of this launch (even)
    uint32_t       n_launch;     // pairs in this launch
    uint32_t       threads;      // couples in this launch = row stride of flow
    uint32_t       maxM;
    uint8_t*       ops;
    uint32_t       ops_stride;
    uint32_t       dfix1;        // Gotoh: packed 4d + 1 with the borrow fix
};

// one pattern row with direction tags; fl = this thread's flow cell of the row
template <int W, bool GOTOH, int WS, int R0>
__device__ __forceinline__ void t_row(uint32_t (&X)[W], uint32_t (&F)[W], const uint32_t (&win)[WS], const uint2 pq,
                                      const uint32_t dfix1, const uint32_t one, uint8_t* fl)
{
    if (!GOTOH)
    {
        uint32_t acc0 = 0, acc1 = 0, left = 0;
        #pragma unroll
        for (int j = 0; j < W; ++j)
        {
            const uint32_t S = c_prmt(pq.x, pq.y, win[(R0 + j) % WS]);        // 4 S' + 2
            const uint32_t dg = X[j] + S;
            uint32_t h;
            if (j == 0)          h = __vmaxs2(dg, X[1 < W ? 1 : 0]);
            else if (j < W - 1)  h = __vimax3_s16x2(dg, X[j + 1 < W ? j + 1 : j], left + 0x00010001u);
            else                 h = __vmaxs2(dg, left + 0x00010001u);
            const uint32_t t = h & 0x00030003u;
            h -= t;
            X[j] = h; left = h;
            if (j < 8) acc0 = c_imad(t, 1u << (2 * (j & 7)), acc0);
            else       acc1 = c_imad(t, 1u << (2 * (j & 7)), acc1);
        }
        *(uint2*)fl = make_uint2(acc0, acc1);
    }
    else
    {
        uint32_t acc[4] = { 0, 0, 0, 0 };
        uint32_t E = 0, xEprev = 0x00020002u;                                 // cell 0 takes no E: "not an extension"
        #pragma unroll
        for (int j = 0; j < W; ++j)
        {
            const uint32_t S = c_prmt(pq.x, pq.y, win[(R0 + j) % WS]);        // 4 S'' + 1
            const uint32_t dg = X[j] + S;                                     // (4 HO + 1) + 4 S'' + 1: tag 2
            uint32_t xF = 0x00010001u;                                        // last cell takes no F
            if (j < W - 1)
            {
                const uint32_t ft = __vmaxs2(F[j + 1 < W ? j + 1 : j], X[j + 1 < W ? j + 1 : j]);   // F0 | HO1: extension only if strictly larger
                xF = ft & 0x00010001u;
                F[j] = ft - xF;
            }
            else F[j] = B2_INF2;
            uint32_t h;
            if (j == 0)          h = __vmaxs2(dg, F[0]);
            else if (j < W - 1)  h = __vimax3_s16x2(dg, F[j], E);
            else                 h = __vmaxs2(dg, E);
            const uint32_t th = h & 0x00030003u;
            h -= th;
            const uint32_t ho1 = c_imad(h, one, dfix1);                       // 4 (H'' + d) + 1
            X[j] = ho1;
            const uint32_t m = 1u << (4 * (j & 3));
            acc[j >> 2] = c_imad(th, m, c_imad(xEprev, 2u * m, c_imad(xF, 8u * m, acc[j >> 2])));
            if (j == 0) { E = ho1; xEprev = 0x00020002u; }                    // E = H + open (gotoh_banded_inl.h:497)
            else
            {
                const uint32_t et = __vmaxs2(E, ho1 + 0x00020002u);           // E1 | ho3: extension only if strictly larger
                xEprev = et & 0x00020002u;
                E = et - xEprev;
            }
        }
        *(uint4*)fl = make_uint4(acc[0], acc[1], acc[2], acc[3]);
    }
}

template <int W, int RU, int R0>
struct TbRowUnroll
{
    template <bool GOTOH, int WS>
    static __device__ __forceinline__ void run(uint32_t (&X)[W], uint32_t (&F)[W], uint32_t (&win)[WS], const uint8_t* pA, const uint8_t* pB,
                                               const uint8_t* tA, const uint8_t* tB, uint32_t lut_addr, uint32_t dfix1, uint32_t one,
                                               uint32_t k8, uint32_t k32, uint8_t* fl, size_t row_bytes)
    {
        const uint2 pq = c_lds64(c_imad(pA[R0], k8, c_imad(pB[R0], k32, lut_addr)));
        win[(R0 + W - 1) % WS] = c_imad(tB[R0], 0x1100u, c_imad(tA[R0], 0x11u, 0xC480u));
        t_row<W, GOTOH, WS, R0>(X, F, win, pq, dfix1, one, fl);
        TbRowUnroll<W, RU, R0 + 1>::template run<GOTOH, WS>(X, F, win, pA, pB, tA, tB, lut_addr, dfix1, one, k8, k32, fl + row_bytes, row_bytes);
    }
};
template <int W, int RU>
struct TbRowUnroll<W, RU, RU>
{
    template <bool GOTOH, int WS>
    static __device__ __forceinline__ void run(uint32_t (&)[W], uint32_t (&)[W], uint32_t (&)[WS], const uint8_t*, const uint8_t*,
                                               const uint8_t*, const uint8_t*, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint8_t*, size_t) {}
};

template <int W, int TYPE, bool GOTOH>
__global__ void __launch_bounds__(B2_THREADS, 4)
band2_tb_forward_kernel(const Band2TbParams q)
{
    const Band2Params& p = q.b;
    constexpr int WS = W + 1;
    constexpr int RU = WS;
    constexpr int STRIDE = 4 * 64 + 4;
    constexpr int CELL = GOTOH ? 16 : 8;

    extern __shared__ __align__(16) uint8_t s_raw[];
    uint2* lut = (uint2*)s_raw;
    if (threadIdx.x < 16)
    {
        const uint32_t qa = threadIdx.x & 3u, qb = threadIdx.x >> 2;
        const uint32_t base = p.s_mis * 0x01010101u, delta = p.s_mis ^ p.s_match;
        lut[threadIdx.x] = make_uint2(base ^ (delta << (8u * qa)), base ^ (delta << (8u * qb)));
    }
    __syncthreads();

    const uint32_t tid = blockIdx.x * B2_THREADS + threadIdx.x;
    if (tid >= q.threads) return;
    const uint32_t idA = q.first + 2 * tid;
    const bool hasB = 2 * tid + 1 < q.n_launch;
    const uint32_t idB = hasB ? idA + 1 : idA;
    const uint32_t pbA = p.P.off[idA], MA = p.P.off[idA + 1] - pbA, pbB = p.P.off[idB], MB = p.P.off[idB + 1] - pbB;
    const uint32_t tbA = p.T.off[idA], NA = p.T.off[idA + 1] - tbA, tbB = p.T.off[idB], NB = p.T.off[idB + 1] - tbB;
    const uint32_t Mmax = max(MA, MB), Mmin = min(MA, MB);

    uint8_t* slot = s_raw + 128 + (size_t)threadIdx.x * STRIDE;
    const uint32_t lut_addr = (uint32_t)__cvta_generic_to_shared(s_raw);
    const size_t row_bytes = (size_t)q.threads * CELL;
    uint8_t* fl = q.flow + (size_t)tid * CELL;

    // rows beyond maxM have no flow row: such pairs are invalid for the reference too (batched_banded_inl.h traceback contexts)
    const bool too_long = Mmax > q.maxM;

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
    for (int j = 0; j < W - 1; ++j)
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
    for (int seg = 0; seg < 2 && !too_long; ++seg)
    {
        const uint32_t target = seg ? Mmax : Mmin;
        if (seg == 0)
        {
            uint32_t o = 0;
            #pragma unroll 1
            for (uint32_t g = target / RU; g > 0; --g)
            {
                const uint4 vPA = c_load(sPA, sPA.nb), vPB = c_load(sPB, sPB.nb), vTA = c_load(sTA, sTA.nb), vTB = c_load(sTB, sTB.nb);
                TbRowUnroll<W, RU, 0>::template run<GOTOH, WS>(X, F, win, sPA.ptr, sPB.ptr, sTA.ptr, sTB.ptr, lut_addr, q.dfix1, p.one, p.k8, p.k32, fl, row_bytes);
                fl += row_bytes * RU;
                seen |= vPA.x | vPA.y | vPA.z | vPA.w | vPB.x | vPB.y | vPB.z | vPB.w;
                seen |= vTA.x | vTA.y | vTA.z | vTA.w | vTB.x | vTB.y | vTB.z | vTB.w;
                const uint4 cPA = c_clip(sPA, sPA.nb, vPA), cPB = c_clip(sPB, sPB.nb, vPB);
                const uint4 cTA = c_clip(sTA, sTA.nb, vTA), cTB = c_clip(sTB, sTB.nb, vTB);
                if (RU == 16)
                {
                    const uint32_t so = o == 0u ? 32u : o - 16u;
                    c_put(sPA.ring, so, cPA); c_put(sPB.ring, so, cPB); c_put(sTA.ring, so, cTA); c_put(sTB.ring, so, cTB);
                    const int32_t step = o == 32u ? -32 : 16;
                    o = o == 32u ? 0u : o + 16u;
                    sPA.ptr += step; sPB.ptr += step; sTA.ptr += step; sTB.ptr += step;
                    sPA.nb += 16u; sPB.nb += 16u; sTA.nb += 16u; sTB.nb += 16u;
                }
                else
                {
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
        #pragma unroll 1
        while (i < target)
        {
            TbRowUnroll<W, 1, 0>::template run<GOTOH, WS>(X, F, win, sPA.ptr, sPB.ptr, sTA.ptr, sTB.ptr, lut_addr, q.dfix1, p.one, p.k8, p.k32, fl, row_bytes);
            fl += row_bytes;
            #pragma unroll
            for (int j = 0; j < W - 1; ++j) win[j] = win[j + 1];
            c_step(sPA, seen); c_step(sPB, seen); c_step(sTA, seen); c_step(sTB, seen);
            ++i;
        }
        // the pairs that end here (scores are kept times four; Gotoh as 4 HO + 1)
        #pragma unroll
        for (int h = 0; h < 2; ++h)
        {
            const uint32_t M = h ? MB : MA, N = h ? NB : NA;
            if (M != target || (seg == 1 && Mmin == Mmax) || M == 0 || N < M) continue;
            const int32_t back = p.alpha * (int32_t)M - p.bias - (GOTOH ? p.d : 0);
            int32_t best = NVB_SINK_MIN; uint32_t bx = 0xFFFFFFFFu;
            if (TYPE == T_GLOBAL) { best = ((h ? c_hi(X[W - 1]) : c_lo(X[W - 1])) >> 2) + back + p.beta * (W - 1); bx = M + W - 1; }
            else
            {
                const uint32_t lim = (M + W - 1 < N ? M + W - 1 : N) - (M - 1);
                #pragma unroll
                for (int j = 0; j < W; ++j)
                {
                    const int32_t s = ((h ? c_hi(X[j]) : c_lo(X[j])) >> 2) + back + p.beta * j;
                    if ((j == 0 || (uint32_t)j < lim) && best <= s) { best = s; bx = M + j; }
                }
            }
            if (h) { scoreB = best; sxB = bx; } else { scoreA = best; sxA = bx; }
        }
    }

    bool dirty = MA == 0 || MB == 0 || edge;     // edge: a string at the end of an unaligned symbol array (c_open)
    if (TYPE == T_GLOBAL) dirty = dirty || NA < MA + W - 1 || NB < MB + W - 1;
    if (seen & 0xFCFCFCFCu)
        dirty = dirty || c_has_high_symbol(p.P.sym, pbA, MA) || c_has_high_symbol(p.P.sym, pbB, MB) ||
                c_has_high_symbol(p.T.sym, tbA, NA) || c_has_high_symbol(p.T.sym, tbB, NB);
    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        nvb_alignment a;
        a.score = NVB_SINK_MIN; a.source_x = a.source_y = a.sink_x = a.sink_y = 0xFFFFFFFFu;
        if (dirty || too_long)
        {
            // left to the int32 kernel (which also reports patterns longer than maxM as invalid)
            const uint32_t at = atomicAdd(p.worklist, 1u);
            p.worklist[1 + at] = id;
        }
        else if ((h ? sxB : sxA) != 0xFFFFFFFFu)
        {
            a.score = h ? scoreB : scoreA; a.sink_x = h ? sxB : sxA; a.sink_y = h ? MB : MA;
        }
        q.out[id] = a;            // the walk fills in the source
        q.n_ops[id] = 0;
    }
}

// ---------------------------------------------------------------------------------------------
// the walk: one thread per pair (sw_banded_inl.h:740-800, gotoh_banded_inl.h:883-950 without the windows)
// ---------------------------------------------------------------------------------------------
template <int W, bool GOTOH>
__global__ void __launch_bounds__(256)
band2_tb_walk_kernel(const Band2TbParams q)
{
    const uint32_t lp = blockIdx.x * 256 + threadIdx.x;
    if (lp >= q.n_launch) return;
    const uint32_t id = q.first + lp;
    const nvb_alignment a = q.out[id];
    if (a.sink_x == 0xFFFFFFFFu) return;                       // invalid, or left to the int32 kernel
    const uint32_t half = lp & 1u, t = lp >> 1;
    OpSink sink; sink.init(q.ops + (size_t)id * q.ops_stride, q.ops_stride);
    int32_t entry = (int32_t)(a.sink_x - a.sink_y);
    int32_t row = (int32_t)a.sink_y - 1;
    uint32_t state = ST_H;
    int32_t have = -1;
    uint32_t bits0 = 0, bits1 = 0;                              // this pair's flow of row `have`
    while (row >= 0)
    {
        if (have != row)
        {
            if (!GOTOH)
            {
                const uint2 w = *(const uint2*)(q.flow + ((size_t)row * q.threads + t) * 8);
                bits0 = half ? ((w.x >> 16) | (w.y & 0xFFFF0000u)) : ((w.x & 0xFFFFu) | (w.y << 16));
            }
            else
            {
                const uint4 w = *(const uint4*)(q.flow + ((size_t)row * q.threads + t) * 16);
                bits0 = half ? ((w.x >> 16) | (w.y & 0xFFFF0000u)) : ((w.x & 0xFFFFu) | (w.y << 16));   // cells 0..7
                bits1 = half ? ((w.z >> 16) | (w.w & 0xFFFF0000u)) : ((w.z & 0xFFFFu) | (w.w << 16));   // cells 8..14
            }
            have = row;
        }
        if (!GOTOH)
        {
            const uint32_t tag = (bits0 >> (2 * entry)) & 3u;   // 2 = SUBSTITUTION, 1 = DELETION (text), 0 = INSERTION
            if (tag == 1u)      { --entry; sink.push(OP_DEL); }
            else if (tag == 0u) { ++entry; --row; sink.push(OP_INS); }
            else                { --row; sink.push(OP_SUB); }
        }
        else
        {
            const uint32_t nib = ((entry < 8 ? bits0 : bits1) >> (4 * (entry & 7))) & 15u;
            const uint32_t tag = nib & 3u;
            const bool e_ext = (nib & 4u) == 0u, f_ext = (nib & 8u) == 0u;
            if (state == ST_E)      { if (!e_ext) state = ST_H; --entry; sink.push(OP_DEL); }
            else if (state == ST_F) { if (!f_ext) state = ST_H; ++entry; --row; sink.push(OP_INS); }
            else
            {
                if (tag == 1u)      state = ST_E;
                else if (tag == 0u) state = ST_F;
                else                { --row; sink.push(OP_SUB); }
            }
        }
        if (entry < 0 || entry >= W) break;
    }
    sink.finish();
    const uint32_t sy = row < 0 ? 0u : (uint32_t)row + 1u;
    q.out[id].source_x = (uint32_t)entry + sy;
    q.out[id].source_y = sy;
    q.n_ops[id] = sink.n;
}

// ---------------------------------------------------------------------------------------------
struct Band2TbPlan { Band2TbParams q; bool ok; };

static Band2TbPlan band2_tb_plan(const nvb_aligner& a, uint32_t band, uint32_t maxM)
{
    Band2TbPlan t; t.ok = false;
    Band2Plan pl = band2_plan(a, band, maxM);
    if (!pl.ok) return t;
    Band2Params& p = pl.p;
    const bool g = a.kind == NVB_GOTOH;
    // everything times four; profile bytes 4 S' + 2 (SW/ED: the diagonal's tag) or 4 S'' + 1 (Gotoh: on top of HO1's 1)
    const uint32_t sm = 4 * p.s_match + (g ? 1 : 2), sx = 4 * p.s_mis + (g ? 1 : 2);
    if (sm > 127 || sx > 127) return t;
    const int64_t high = 4 * ((int64_t)p.bias + (int64_t)max(p.s_match, p.s_mis) * maxM + 16 * 64 + 256) + 8;
    if (high > 32000) return t;
    p.s_match = sm; p.s_mis = sx;
    p.init0 = 4 * p.init0 + (g ? 1 : 0); p.init_rest = 4 * p.init_rest + (g ? 1 : 0); p.init_step = 4 * p.init_step;
    const int32_t d1 = 4 * p.d + 1;                                                   // h (clean) -> HO1
    t.q.dfix1 = d1 < 0 ? ((((uint32_t)(d1 - 1) & 0xFFFFu) << 16) | ((uint32_t)d1 & 0xFFFFu)) : (((uint32_t)d1 & 0xFFFFu) * 0x00010001u);
    t.q.b = p;
    t.ok = true;
    return t;
}

size_t dpb2_tb_bytes_per_couple(const nvb_aligner& a, uint32_t maxM) { return (size_t)maxM * (a.kind == NVB_GOTOH ? 16 : 8); }

bool dpb2_traceback_supported(const nvb_aligner& a, uint32_t band, const StrSet& P, const StrSet& T, uint32_t maxM)
{
    if (P.bits != 8 || T.bits != 8) return false;
    return band2_tb_plan(a, band, maxM).ok;
}

template <int W, int TYPE, bool GOTOH>
static int band2_tb_launch(nvb_ctx* ctx, const Band2TbParams& q, cudaStream_t st)
{
    constexpr size_t smem = 128 + (size_t)B2_THREADS * (4 * 64 + 4);
    if (NVB_FIRST_ON_DEVICE(ctx))
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(band2_tb_forward_kernel<W, TYPE, GOTOH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    band2_tb_forward_kernel<W, TYPE, GOTOH><<<(q.threads + B2_THREADS - 1) / B2_THREADS, B2_THREADS, smem, st>>>(q);
    NVB_LAUNCHED(ctx);
    band2_tb_walk_kernel<W, GOTOH><<<(q.n_launch + 255) / 256, 256, 0, st>>>(q);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// pairs [first, first + n) (first even); flow scratch of at least ceil(n / 2) * dpb2_tb_bytes_per_couple bytes
int dpb2_traceback(nvb_ctx* ctx, const nvb_aligner& a, uint32_t band, StrSet P, StrSet T, uint32_t first, uint32_t n,
                   nvb_alignment* out, uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM,
                   uint8_t* flow, uint32_t* worklist, cudaStream_t st)
{
    Band2TbPlan t = band2_tb_plan(a, band, maxM);
    if (!t.ok) return NVB_ERR_UNSUPPORTED;
    Band2TbParams& q = t.q;
    q.b.P = P; q.b.T = T; q.b.sinks = nullptr; q.b.worklist = worklist;
    q.b.n_pairs = P.count; q.b.n_threads = (n + 1) / 2;
    q.out = out; q.n_ops = n_ops; q.flow = flow; q.first = first; q.n_launch = n; q.threads = (n + 1) / 2; q.maxM = maxM;
    q.ops = ops; q.ops_stride = ops_stride;
    const bool g = a.kind == NVB_GOTOH;
    if (band == 7)
    {
        if (a.type == NVB_GLOBAL) return g ? band2_tb_launch<7, T_GLOBAL, true>(ctx, q, st) : band2_tb_launch<7, T_GLOBAL, false>(ctx, q, st);
        return g ? band2_tb_launch<7, T_SEMI, true>(ctx, q, st) : band2_tb_launch<7, T_SEMI, false>(ctx, q, st);
    }
    if (band == 15)
    {
        if (a.type == NVB_GLOBAL) return g ? band2_tb_launch<15, T_GLOBAL, true>(ctx, q, st) : band2_tb_launch<15, T_GLOBAL, false>(ctx, q, st);
        return g ? band2_tb_launch<15, T_SEMI, true>(ctx, q, st) : band2_tb_launch<15, T_SEMI, false>(ctx, q, st);
    }
    return NVB_ERR_UNSUPPORTED;
}
