// dp_generic_banded.cuh -- int32, one-pair-per-thread banded DP (scoring + checkpointed traceback).
//
// Restates banded::sw_alignment_score_dispatch / gotoh_alignment_score_dispatch
// (nvbio/alignment/sw/sw_banded_inl.h:340-504, gotoh/gotoh_banded_inl.h:406-646) and
// banded_alignment_traceback (banded_inl.h:354-430): W cells per pattern row, cell j <-> text
// position row+j, text past the end reads 255, checkpoints clamped to INT16_MIN+32 when stored
// (sw_banded_inl.h:168-215, gotoh_banded_inl.h:217-249). Note the banded code labels the move from
// the previous pattern row INSERTION and the move along the text DELETION (sw_banded_inl.h:438-440),
// the opposite of the full-matrix code.
#pragma once
#include "dp_common.cuh"

#define NVB_MAX_BAND 64

template <bool GOTOH, int TYPE, int WT>
struct BandDP
{
    static const int WA = WT ? WT : NVB_MAX_BAND;
    static const int UN = WT ? WT : 1;
    Scheme   s;
    PairView v;
    uint32_t Wrt;
    size_t   cs;
    int16_t* chkH;   // [row_id*W + j]
    int16_t* chkF;
    uint8_t* flow;   // [(i - id*CHK)*W + j]
    uint32_t CHK;
    Best*    best;

    __device__ __forceinline__ uint32_t W() const { return WT ? (uint32_t)WT : Wrt; }
    __device__ __forceinline__ static int16_t clamp16(int32_t x) { return (int16_t)(x > -32768 + 32 ? x : -32768 + 32); }
    __device__ __forceinline__ uint32_t text_at(uint32_t p) const { return p < v.N ? v.t(p) : 255u; }

    // rows [wb, we); false = the reference refuses the problem (text shorter than pattern)
    __device__ bool sweep(uint32_t wb, uint32_t we, bool from_chk, bool save_chk, uint32_t chk_id)
    {
        const uint32_t Wd = W(), M = v.M, N = v.N;
        const bool local = (TYPE == T_LOCAL);
        if (N < M) return false;                                              // sw_banded_inl.h:356-357

        int32_t H[WA], F[WA];
        const bool packed_cache = !(Wd == 3 || Wd == 5 || Wd == 7 || Wd == 15);   // alignment_base_inl.h:66-90
        const int32_t G = s.ga, I = s.gb;
        int32_t gmax = s.ga > s.gb ? s.ga : s.gb;
        if (GOTOH) gmax = max(gmax, max(s.tga, s.tgb));                          // the text gaps count too: gotoh_banded_inl.h:437-439
        const int32_t inf = GOTOH ? -32768 - gmax : -32768 + gmax;

        if (from_chk)
        {
            #pragma unroll UN
            for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j < Wd)
            {
                H[j] = chkH[((size_t)chk_id * Wd + j) * cs];
                F[j] = GOTOH ? (int32_t)chkF[((size_t)chk_id * Wd + j) * cs] : 0;
            }
        }
        else
        {
            #pragma unroll UN
            for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j < Wd)
            {
                if (!GOTOH) H[j] = (TYPE == T_GLOBAL) ? (int32_t)j * s.ga : 0;                         // sw_banded_inl.h:35-45
                else H[j] = (TYPE == T_GLOBAL && j > 0) ? s.tga + (int32_t)(j - 1) * s.tgb : 0;        // text gaps: gotoh_banded_inl.h:37-68
                F[j] = inf;
            }
        }

        for (uint32_t i = wb; i < we; ++i)
        {
            if (save_chk && (i % CHK) == 0)
            {
                #pragma unroll UN
                for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j < Wd)
                {
                    chkH[((size_t)(i / CHK) * Wd + j) * cs] = clamp16(H[j]);
                    if (GOTOH) chkF[((size_t)(i / CHK) * Wd + j) * cs] = clamp16(F[j]);
                }
            }
            const uint32_t q = v.p(i);
            // match(qq) / mismatch(g, q, qq) at this pattern row's quality (sw_banded_inl.h:391, gotoh_banded_inl.h:461)
            const int32_t s_match = s.qm ? __ldg(s.qm + v.pq(i)) : s.match, s_mis = s.qx ? __ldg(s.qx + v.pq(i)) : s.mismatch;
            int32_t E = 0; uint32_t edir = 0;
            #pragma unroll UN
            for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j < Wd)
            {
                uint32_t g = text_at(i + j);
                if (packed_cache && j < Wd - 1) g &= 3u;
                const int32_t S = (g == q) ? s_match : s_mis;
                const int32_t dg = H[j] + S;
                int32_t hi; uint32_t dir, cell;
                if (!GOTOH)
                {
                    if (j == 0)           { const int32_t top = H[j + 1 < (uint32_t)WA ? j + 1 : j] + G; hi = max(top, dg); dir = top > dg ? OP_INS : OP_SUB; }
                    else if (j == Wd - 1) { const int32_t left = H[j - 1] + I; hi = max(left, dg); dir = left > dg ? OP_DEL : OP_SUB; }
                    else
                    {
                        const int32_t top = H[j + 1 < (uint32_t)WA ? j + 1 : j] + G, left = H[j - 1] + I;
                        hi = max(max(top, left), dg);
                        dir = top > left ? (top > dg ? OP_INS : OP_SUB) : (left > dg ? OP_DEL : OP_SUB);
                    }
                    if (local) hi = max(hi, 0);
                    cell = dir;   // banded SW/ED never stores SINK (sw_banded_inl.h:259-270)
                }
                else
                {
                    uint32_t fdir = 0;
                    if (j < Wd - 1)
                    {
                        const uint32_t j1 = j + 1 < (uint32_t)WA ? j + 1 : j;
                        const int32_t ftop = F[j1] + s.gb, htop = H[j1] + s.ga;
                        F[j] = max(ftop, htop);
                        fdir = ftop > htop ? OP_DEL_EXT : 0;
                    }
                    else F[j] = inf;
                    if (j == 0)           { const int32_t top = F[0]; hi = max(top, dg); dir = top > dg ? OP_INS : OP_SUB; }
                    else if (j == Wd - 1) { const int32_t left = E; hi = max(left, dg); dir = left > dg ? OP_DEL : OP_SUB; }
                    else
                    {
                        const int32_t top = F[j], left = E;
                        hi = max(max(top, left), dg);
                        dir = top > left ? (top > dg ? OP_INS : OP_SUB) : (left > dg ? OP_DEL : OP_SUB);
                    }
                    if (local) { hi = max(hi, 0); if (hi == 0) dir = OP_SINK; }                     // gotoh_banded_inl.h:325-326
                    cell = dir | (j == 0 ? 0u : edir) | fdir;
                    if (j == 0) { E = hi + s.ga; edir = 0; }                                        // gotoh_banded_inl.h:497
                    else
                    {
                        const int32_t eleft = E + s.gb, ediag = hi + s.ga;                         // :560-564
                        edir = eleft > ediag ? OP_INS_EXT : 0;
                        E = max(ediag, eleft);
                    }
                }
                H[j] = hi;
                if (local && best) best->report(hi, i + j + 1, i + 1);
                if (flow) flow[((size_t)(i - chk_id * CHK) * Wd + j) * cs] = (uint8_t)cell;
            }
        }
        if (save_chk)   // last_row: sw_banded_inl.h:196-215
        {
            const uint32_t row = (we + CHK - 1) / CHK;
            #pragma unroll UN
            for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j < Wd)
            {
                chkH[((size_t)row * Wd + j) * cs] = clamp16(H[j]);
                if (GOTOH) chkF[((size_t)row * Wd + j) * cs] = clamp16(F[j]);
            }
        }
        if (we == M && best)
        {
            if (TYPE == T_GLOBAL)
            {
                int32_t h = 0;
                #pragma unroll UN
                for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j == Wd - 1) h = H[j];
                best->report(h, M + Wd - 1, M);                                                   // sw_banded_inl.h:487-488
            }
            else if (TYPE == T_SEMI)
            {
                const uint32_t lim = (M + Wd - 1 < N ? M + Wd - 1 : N) - (M - 1);                  // :489-501
                #pragma unroll UN
                for (uint32_t j = 0; j < (uint32_t)WA; ++j) if (j < Wd)
                    if (j == 0 || j < lim) best->report(H[j], M + j, M);
            }
        }
        return true;
    }

    // sw_banded_inl.h:740-800, gotoh_banded_inl.h:883-950
    __device__ bool walk(uint32_t chk_id, uint32_t& state, uint32_t& sx, uint32_t& sy,
                         uint8_t* ops, uint32_t ops_cap, uint32_t& n_ops) const
    {
        const bool local = (TYPE == T_LOCAL);
        const uint32_t Wd = W();
        int32_t entry = (int32_t)(sx - sy);
        int32_t row   = (int32_t)(sy - chk_id * CHK - 1u);
        #define PUSH(o) do { if (n_ops < ops_cap) ops[n_ops] = (uint8_t)(o); ++n_ops; } while (0)
        while (row >= 0)
        {
            const uint32_t op = flow[((size_t)row * Wd + (uint32_t)entry) * cs];
            if (!GOTOH)
            {
                if (local && op == OP_SINK) { sy = (uint32_t)row + chk_id * CHK + 1u; sx = (uint32_t)entry + sy; return true; }
                if (op == OP_DEL)      { --entry; PUSH(OP_DEL); }
                else if (op == OP_INS) { ++entry; --row; PUSH(OP_INS); }
                else                   { --row; PUSH(OP_SUB); }
            }
            else
            {
                const uint32_t h_op = op & 3u;
                if (local && state == ST_H && h_op == OP_SINK)
                { sy = (uint32_t)row + chk_id * CHK + 1u; sx = (uint32_t)entry + sy; return true; }
                if (state == ST_E)      { if ((op & OP_INS_EXT) == 0) state = ST_H; --entry; PUSH(OP_DEL); }
                else if (state == ST_F) { if ((op & OP_DEL_EXT) == 0) state = ST_H; ++entry; --row; PUSH(OP_INS); }
                else
                {
                    if (h_op == OP_DEL) state = ST_E;
                    else if (h_op == OP_INS) state = ST_F;
                    else { --row; PUSH(OP_SUB); }
                }
            }
            if (entry < 0 || entry >= (int32_t)Wd) break;   // the reference only asserts here
        }
        #undef PUSH
        sy = chk_id * CHK;
        sx = (uint32_t)entry + sy;
        return false;
    }

    __device__ void traceback(nvb_alignment* res, uint8_t* ops, uint32_t ops_cap, uint32_t* n_ops_out)
    {
        const uint32_t M = v.M;
        Best b; b.init();
        best = &b; uint8_t* fl = flow; flow = nullptr;
        sweep(0, M, false, true, 0);
        best = nullptr; flow = fl;
        res->score = b.score; res->sink_x = b.x; res->sink_y = b.y;
        uint32_t n_ops = 0;
        if (b.x == 0xFFFFFFFFu || b.y == 0xFFFFFFFFu)
        {
            res->source_x = res->source_y = res->sink_x = res->sink_y = 0xFFFFFFFFu;
            *n_ops_out = 0;
            return;
        }
        uint32_t sx = b.x, sy = b.y;
        int32_t id = (int32_t)(1u + (M + CHK - 1) / CHK) - 2;                      // banded_inl.h:378
        if (TYPE == T_LOCAL) for (; id >= 0; --id) if ((uint32_t)id * CHK < sy) break;
        uint32_t state = ST_H;
        for (; id >= 0; --id)
        {
            const uint32_t wb = (uint32_t)id * CHK;
            const uint32_t we = wb + CHK < M ? wb + CHK : M;
            sweep(wb, we, true, false, (uint32_t)id);
            if (walk((uint32_t)id, state, sx, sy, ops, ops_cap, n_ops)) break;
        }
        res->source_x = sx; res->source_y = sy;
        *n_ops_out = n_ops;
    }
};

template <bool GOTOH, int TYPE, int WT>
__global__ void __launch_bounds__(128)
dpg_banded_score_kernel(Scheme s, uint32_t W, StrSet P, StrSet T, const uint32_t* __restrict__ list, const uint32_t* __restrict__ d_count, uint32_t first,
                        uint32_t n, nvb_best_sink* __restrict__ sinks, Sink2 s2)
{
    const uint32_t lp = blockIdx.x * 128 + threadIdx.x;
    if (lp >= n) return;
    if (d_count && first + lp >= *d_count) return;
    const uint32_t id = list ? list[first + lp] : first + lp;
    BandDP<GOTOH, TYPE, WT> dp;
    dp.s = s; dp.v = make_pair_view(P, T, id, s.quals); dp.Wrt = W; dp.cs = 1;
    dp.chkH = dp.chkF = nullptr; dp.flow = nullptr; dp.CHK = 1u << 30;
    Best b; b.init(); dp.best = &b;
    if (s2.out) { b.two = true; b.dd = s2.distinct; }
    dp.sweep(0, dp.v.M, false, false, 0);
    if (s2.out) { nvb_best2_sink o; o.score1 = b.score; o.score2 = b.score2; o.sink1_x = b.x; o.sink1_y = b.y; o.sink2_x = b.x2; o.sink2_y = b.y2; s2.out[id] = o; }
    else { sinks[id].score = b.score; sinks[id].sink_x = b.x; sinks[id].sink_y = b.y; }
}

template <bool GOTOH, int TYPE, int WT>
__global__ void __launch_bounds__(128)
dpg_banded_traceback_kernel(Scheme s, uint32_t W, StrSet P, StrSet T, const uint32_t* __restrict__ list, const uint32_t* __restrict__ d_count,
                            uint32_t first, uint32_t n, uint32_t CHK,
                            nvb_alignment* __restrict__ out, uint8_t* __restrict__ ops, uint32_t ops_stride,
                            uint32_t* __restrict__ n_ops, int16_t* __restrict__ temp16, uint8_t* __restrict__ flow,
                            uint32_t maxM)
{
    const uint32_t lp = blockIdx.x * 128 + threadIdx.x;
    if (lp >= n) return;
    if (d_count && first + lp >= *d_count) return;
    const uint32_t id = list ? list[first + lp] : first + lp;
    const uint32_t nrows = 2u + (maxM + CHK - 1) / CHK;
    BandDP<GOTOH, TYPE, WT> dp;
    dp.s = s; dp.v = make_pair_view(P, T, id, s.quals); dp.Wrt = W; dp.cs = n;
    dp.chkH = temp16 + lp;
    dp.chkF = temp16 + (size_t)nrows * W * n + lp;
    dp.flow = flow + lp; dp.CHK = CHK; dp.best = nullptr;
    if (dp.v.M > maxM)
    {
        out[id].score = NVB_SINK_MIN;
        out[id].source_x = out[id].source_y = out[id].sink_x = out[id].sink_y = 0xFFFFFFFFu;
        n_ops[id] = 0;
        return;
    }
    dp.traceback(out + id, ops + (size_t)id * ops_stride, ops_stride, n_ops + id);
}
