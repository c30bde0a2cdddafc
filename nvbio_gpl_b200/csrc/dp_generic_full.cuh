// dp_generic_full.cuh -- int32, one-pair-per-thread full-matrix DP (scoring + checkpointed traceback).
//
// This is the parity anchor of the DP half: it visits cells in exactly the reference's order
//   pattern stripes of BAND_LEN columns (16 SW/ED: sw_inl.h:1321-1325, 8 Gotoh: gotoh_inl.h:1259-1262)
//   -> text rows -> columns of the stripe            (sw_inl.h:610-749, gotoh_inl.h:611-840)
// so BestSink's "last maximal cell wins" (sink_inl.h:44) falls out of the loop order. The stripe
// boundary column and all checkpoints round-trip through int16 as in the reference (utils.h:51-53).
// It serves every combination the packed DPX kernels do not (symbols > 3, wide scores, GLOBAL /
// SEMI_GLOBAL full matrices, Gotoh full matrices) and re-does the pairs they flag.
#pragma once
#include "dp_common.cuh"

template <bool GOTOH, int TYPE, bool QUAL = false>
struct FullDP
{
    static const int BL = GOTOH ? 8 : 16;
    Scheme   s;
    PairView v;
    size_t   cs;            // element stride between consecutive cells of this pair's scratch
    int16_t* colH;          // [i]          boundary column H
    int16_t* colE;          // [i]          boundary column E (Gotoh)
    int16_t* chkH;          // [id*N + i]   checkpoints
    int16_t* chkE;
    uint8_t* flow;          // [i*CHK + c]  submatrix flow cells (NULL when only scoring)
    uint32_t CHK;
    Best*    best;          // NULL = null sink

    __device__ __forceinline__ int32_t infimum() const
    { return -32768 - (s.ga < s.gb ? s.ga : s.gb); }   // gotoh_inl.h:644

    // one stripe [block, block+ncols); last = the stripe holding column M
    __device__ void stripe(uint32_t block, uint32_t ncols, bool last, bool save_chk, uint32_t chk_id)
    {
        const uint32_t N = v.N;
        const bool local = (TYPE == T_LOCAL);
        const int32_t inf = infimum();
        int32_t H[BL + 1], F[BL + 1];
        uint32_t q[BL + 1];
        int32_t sm[QUAL ? BL + 1 : 1], sx[QUAL ? BL + 1 : 1];       // match(qq) / mismatch(r, q, qq) of the stripe's columns (sw_inl.h:470-472)

        if (save_chk && (block % CHK) == 0)
        {
            const uint32_t id = block / CHK;
            for (uint32_t i = 0; i < N; ++i)
            {
                chkH[((size_t)id * N + i) * cs] = colH[i * cs];
                if (GOTOH) chkE[((size_t)id * N + i) * cs] = colE[i * cs];
            }
        }
        #pragma unroll
        for (int j = 0; j <= BL; ++j)
        {
            if (local) H[j] = 0;
            else if (!GOTOH) H[j] = s.gb * (int32_t)(block + j);                                   // sw_inl.h:655
            else H[j] = (block + j > 0) ? s.ga + s.gb * (int32_t)(block + j - 1u) : 0;             // gotoh_inl.h:675
            F[j] = inf;
            q[j] = (j >= 1 && (uint32_t)j <= ncols) ? v.p(block + j - 1) : 0x100u;
            if (QUAL)
            {
                const uint32_t qq = (j >= 1 && (uint32_t)j <= ncols) ? v.pq(block + j - 1) : 0u;
                sm[j] = __ldg(s.qm + qq); sx[j] = __ldg(s.qx + qq);
            }
        }
        int32_t carry = H[0];
        int32_t hlast = 0;
        #pragma unroll
        for (int j = 0; j <= BL; ++j) if ((uint32_t)j == ncols) hlast = H[j];

        const uint32_t fbase = block - chk_id * CHK;
        // BestSink over a LOCAL stripe row: cells are visited left to right and the last maximum wins, i.e. the maximum of
        // (score << 5 | column) -- one IMNMX per cell instead of a compare and three selects; the row then reports once.
        // (Best2Sink needs every cell in order, so it keeps the per-cell report.)
        const bool keyed = local && best && !best->two;
        for (uint32_t i = 0; i < N; ++i)
        {
            const uint32_t r = v.t(i);
            int32_t rk = -1;
            int32_t diag = carry;
            H[0] = carry = colH[i * cs];
            int32_t E = GOTOH ? (int32_t)colE[i * cs] : 0;
            int32_t Eout = E;
            #pragma unroll
            for (int j = 1; j <= BL; ++j)
            {
                if ((uint32_t)j <= ncols)
                {
                    const int32_t S = QUAL ? ((r == q[j]) ? sm[j] : sx[j]) : ((r == q[j]) ? s.match : s.mismatch);
                    int32_t hi; uint32_t dir;
                    if (!GOTOH)
                    {
                        const int32_t top = H[j] + s.ga, left = H[j - 1] + s.gb, dg = diag + S;
                        hi = max(max(top, left), dg);
                        if (local) hi = max(hi, 0);
                        dir = top > left ? (top > dg ? OP_DEL : OP_SUB) : (left > dg ? OP_INS : OP_SUB);   // sw_inl.h:488-490
                        if (flow) flow[((size_t)i * CHK + fbase + j - 1) * cs] = (local && hi == 0) ? OP_SINK : dir;
                    }
                    else
                    {
                        const int32_t ftop = F[j] + s.gb, htop = H[j] + s.ga;
                        F[j] = max(ftop, htop);
                        const uint32_t fdir = ftop > htop ? OP_DEL_EXT : 0;
                        const int32_t eleft = E + s.gb, hleft = H[j - 1] + s.ga;
                        E = max(eleft, hleft);
                        const uint32_t edir = eleft > hleft ? OP_INS_EXT : 0;
                        const int32_t dg = diag + S, top = F[j], left = E;
                        hi = max(max(left, top), dg);
                        if (local) hi = max(hi, 0);
                        dir = top > left ? (top > dg ? OP_DEL : OP_SUB) : (left > dg ? OP_INS : OP_SUB);
                        if (flow) flow[((size_t)i * CHK + fbase + j - 1) * cs] =
                            (uint8_t)(((local && hi == 0) ? OP_SINK : dir) | edir | fdir);
                        Eout = E;
                    }
                    diag = H[j];
                    H[j] = hi;
                    if (keyed) rk = max(rk, (hi << 5) | j);
                    else if (local && best) best->report(hi, i + 1, block + j);
                    if ((uint32_t)j == ncols) hlast = hi;
                }
            }
            if (keyed && rk >= 0) best->report(rk >> 5, i + 1, block + (uint32_t)(rk & 31));
            if (ncols == 0) hlast = H[0];
            colH[i * cs] = (int16_t)hlast;
            if (GOTOH) colE[i * cs] = (int16_t)Eout;
            if (best && last && TYPE == T_SEMI) best->report(hlast, i + 1, v.M);
        }
        if (last && TYPE == T_GLOBAL && best) best->report(hlast, N, v.M);   // sw_inl.h:745-746
    }

    // sweep the pattern window [wb, we) in stripes (sw_inl.h:640-747)
    __device__ void sweep(uint32_t wb, uint32_t we, bool from_chk, bool save_chk, uint32_t chk_id)
    {
        const uint32_t N = v.N, M = v.M;
        const int32_t inf = infimum();
        if (from_chk)
        {
            for (uint32_t i = 0; i < N; ++i)
            {
                colH[i * cs] = chkH[((size_t)chk_id * N + i) * cs];
                if (GOTOH) colE[i * cs] = chkE[((size_t)chk_id * N + i) * cs];
            }
        }
        else if (wb == 0)
        {
            for (uint32_t i = 0; i < N; ++i)
            {
                if (!GOTOH) colH[i * cs] = (int16_t)(TYPE == T_GLOBAL ? s.ga * (int32_t)(i + 1) : 0);   // sw_inl.h:68-70
                else
                {
                    colH[i * cs] = (int16_t)(TYPE == T_GLOBAL ? s.tga + s.tgb * (int32_t)i : 0);      // text gaps: gotoh_inl.h:68-71
                    colE[i * cs] = (int16_t)(TYPE == T_LOCAL ? 0 : inf);
                }
            }
        }
        uint32_t end_block;
        if (we == M) { end_block = BL * ((M + BL - 1) / BL); if (end_block < (uint32_t)BL) end_block = BL; }
        else end_block = we + BL;

        uint32_t block = wb;
        for (; block + BL < end_block; block += BL) stripe(block, BL, false, save_chk, chk_id);
        if (we == M)
        {
            block = end_block - BL;
            const uint32_t ncols = (M - block < (uint32_t)BL) ? M - block : BL;
            if (M > 0 || TYPE == T_LOCAL) stripe(block, ncols, true, save_chk, chk_id);
        }
    }

    // walk one submatrix (sw_inl.h:1644-1693, gotoh_inl.h:1573-1639); true = source found
    __device__ bool walk(uint32_t chk_id, uint32_t& state, uint32_t& sx, uint32_t& sy,
                         uint8_t* ops, uint32_t ops_cap, uint32_t& n_ops) const
    {
        const bool local = (TYPE == T_LOCAL);
        int32_t row = (int32_t)sx;
        int32_t col = (int32_t)(sy - chk_id * CHK - 1u);
        #define PUSH(o) do { if (n_ops < ops_cap) ops[n_ops] = (uint8_t)(o); ++n_ops; } while (0)
        while (row > 0 && col >= 0)
        {
            const uint32_t op = flow[((size_t)(row - 1) * CHK + (uint32_t)col) * cs];
            if (!GOTOH)
            {
                if (local && op == OP_SINK) { sx = (uint32_t)row; sy = (uint32_t)col + chk_id * CHK + 1u; return true; }
                if (op != OP_DEL) --col;
                if (op != OP_INS) --row;
                PUSH(op);
            }
            else
            {
                const uint32_t h_op = op & 3u;
                if (local && state == ST_H && h_op == OP_SINK)
                { sx = (uint32_t)row; sy = (uint32_t)col + chk_id * CHK + 1u; return true; }
                if (state == ST_E)      { if ((op & OP_INS_EXT) == 0) state = ST_H; --col; PUSH(OP_INS); }
                else if (state == ST_F) { if ((op & OP_DEL_EXT) == 0) state = ST_H; --row; PUSH(OP_DEL); }
                else
                {
                    if (h_op == OP_INS) state = ST_E;
                    else if (h_op == OP_DEL) state = ST_F;
                    else { --col; --row; PUSH(OP_SUB); }
                }
            }
        }
        sx = (uint32_t)row;
        sy = (uint32_t)(col + (int32_t)(chk_id * CHK) + 1);
        return row == 0;
    }

    // aln::alignment_traceback (alignment_inl.h:355-455)
    __device__ void traceback(nvb_alignment* res, uint8_t* ops, uint32_t ops_cap, uint32_t* n_ops_out)
    {
        const uint32_t M = v.M;
        Best b; b.init();
        best = &b; uint8_t* fl = flow; flow = nullptr;
        sweep(0, M, false, true, 0);
        best = nullptr; flow = fl;

        uint32_t n_ops = 0;
        res->score = b.score; res->sink_x = b.x; res->sink_y = b.y;
        if (b.x == 0xFFFFFFFFu || b.y == 0xFFFFFFFFu)
        {
            res->source_x = res->source_y = res->sink_x = res->sink_y = 0xFFFFFFFFu;
            *n_ops_out = 0;
            return;
        }
        uint32_t sx = b.x, sy = b.y;
        int32_t id = (int32_t)((M + CHK - 1) / CHK) - 1;
        if (TYPE == T_LOCAL) for (; id >= 0; --id) if ((uint32_t)id * CHK < sy) break;
        uint32_t state = ST_H;
        for (; id >= 0; --id)
        {
            const uint32_t wb = (uint32_t)id * CHK;
            const uint32_t we = wb + CHK < M ? wb + CHK : M;
            sweep(wb, we, true, false, (uint32_t)id);
            if (walk((uint32_t)id, state, sx, sy, ops, ops_cap, n_ops)) break;
        }
        if (TYPE == T_SEMI || TYPE == T_GLOBAL)
            if (sx == 0) for (; sy > 0; --sy) PUSH(OP_INS);
        if (TYPE == T_GLOBAL)
            if (sy == 0) for (; sx > 0; --sx) PUSH(OP_DEL);
        #undef PUSH
        res->source_x = sx; res->source_y = sy;
        *n_ops_out = n_ops;
    }
};

#define DPG_THREADS 128

// scoring: scratch = colH[,colE], strided by the pairs of the launch
template <bool GOTOH, int TYPE, bool QUAL>
__global__ void __launch_bounds__(DPG_THREADS)
dpg_full_score_kernel(Scheme s, StrSet P, StrSet T, const uint32_t* __restrict__ list, const uint32_t* __restrict__ d_count, uint32_t first,
                      uint32_t n, nvb_best_sink* __restrict__ sinks, int16_t* __restrict__ temp, uint32_t maxN, Sink2 s2)
{
    const uint32_t lp = blockIdx.x * DPG_THREADS + threadIdx.x;
    if (lp >= n) return;
    if (d_count && first + lp >= *d_count) return;
    const uint32_t id = list ? list[first + lp] : first + lp;
    FullDP<GOTOH, TYPE, QUAL> dp;
    dp.s = s; dp.v = make_pair_view(P, T, id, s.quals);
    dp.cs = n;
    dp.colH = temp + lp;
    dp.colE = temp + (size_t)maxN * n + lp;
    dp.chkH = dp.chkE = nullptr; dp.flow = nullptr; dp.CHK = 1u << 30;
    Best b; b.init(); dp.best = &b;
    if (s2.out) { b.two = true; b.dd = s2.distinct; }
    if (dp.v.N <= maxN) dp.sweep(0, dp.v.M, false, false, 0);
    if (s2.out) { nvb_best2_sink o; o.score1 = b.score; o.score2 = b.score2; o.sink1_x = b.x; o.sink1_y = b.y; o.sink2_x = b.x2; o.sink2_y = b.y2; s2.out[id] = o; }
    else { sinks[id].score = b.score; sinks[id].sink_x = b.x; sinks[id].sink_y = b.y; }
}

// traceback: scratch per pair = 2 columns + 2*nchk*N checkpoints (int16) + N*CHK flow bytes
template <bool GOTOH, int TYPE, bool QUAL>
__global__ void __launch_bounds__(DPG_THREADS)
dpg_full_traceback_kernel(Scheme s, StrSet P, StrSet T, const uint32_t* __restrict__ list, const uint32_t* __restrict__ d_count,
                          uint32_t first, uint32_t n, uint32_t CHK,
                          nvb_alignment* __restrict__ out, uint8_t* __restrict__ ops, uint32_t ops_stride,
                          uint32_t* __restrict__ n_ops, int16_t* __restrict__ temp16, uint8_t* __restrict__ flow,
                          uint32_t maxM, uint32_t maxN)
{
    const uint32_t lp = blockIdx.x * DPG_THREADS + threadIdx.x;
    if (lp >= n) return;
    if (d_count && first + lp >= *d_count) return;
    const uint32_t id = list ? list[first + lp] : first + lp;     // list: the pairs a packed kernel left behind
    const uint32_t nchk = (maxM + CHK - 1) / CHK + 1;
    FullDP<GOTOH, TYPE, QUAL> dp;
    dp.s = s; dp.v = make_pair_view(P, T, id, s.quals);
    dp.cs = n;
    dp.colH = temp16 + lp;
    dp.colE = temp16 + (size_t)maxN * n + lp;
    dp.chkH = temp16 + (size_t)2 * maxN * n + lp;
    dp.chkE = dp.chkH + (size_t)nchk * maxN * n;
    dp.flow = flow + lp; dp.CHK = CHK; dp.best = nullptr;
    if (dp.v.N > maxN || dp.v.M > maxM)
    {
        out[id].score = NVB_SINK_MIN;
        out[id].source_x = out[id].source_y = out[id].sink_x = out[id].sink_y = 0xFFFFFFFFu;
        n_ops[id] = 0;
        return;
    }
    dp.traceback(out + id, ops + (size_t)id * ops_stride, ops_stride, n_ops + id);
}
