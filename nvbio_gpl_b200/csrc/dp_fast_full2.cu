// dp_fast_full2.cu -- packed s16x2 DPX wavefront kernels for the full-matrix aligners the LOCAL Smith-Waterman kernel
// (dp_fast_full.cu) does not cover: Smith-Waterman / edit distance GLOBAL and SEMI_GLOBAL (this file, part 1) and Gotoh
// (part 2: LOCAL, and GLOBAL / SEMI_GLOBAL). They replace the reference's one-thread-per-pair sweeps (nvbio/alignment/batched_inl.h:79-128 over
// sw/sw_inl.h:419-749 and gotoh/gotoh_inl.h:458-840) for pairs of 2-bit symbols whose scores fit 15 bits; everything else
// goes to the int32 kernels through the worklist, bit for bit as before.
//
// Layout as in dp_fast_full.cu: two pairs per 32-bit register (a "couple"), the pattern cut into stripes of 16 columns
// spread over adjacent lanes, lane k one text row behind lane k-1 (anti-diagonal wavefront, boundary cells by warp
// shuffle), the couple's two texts as a nibble stream in shared memory.
//
// SW / ED, GLOBAL and SEMI_GLOBAL: SHEARED in both directions. With r text rows and j pattern columns consumed,
//   h[r][j] = H[r][j] - G r - I j          (G = deletion, I = insertion, both <= 0)
// turns the recurrence of sw_inl.h:470-486 into
//   h[r][j] = max3( h[r-1][j-1] + S', h[r-1][j], h[r][j-1] ),   S' = S - G - I >= 0
// so both gap moves are free, every stored value is >= 0 (the packed addition is a plain 32-bit IMAD on the FMA pipe:
// no carry between the halves) and a cell costs PRMT + VIMNMX3 on the ALU pipe. The boundaries become constants:
// row 0 (I j, sw_inl.h:655) -> 0; column 0 of GLOBAL (G r) -> 0, of SEMI_GLOBAL (0) -> -G r.
// Pattern columns past M get S' = 0: there the free move from the left makes h[r][j] = h[r][M] exactly (the diagonal and
// top candidates are older values of the same non-decreasing column), so the LAST lane of a couple always ends its row on
// h[r][M] of both pairs, whatever their lengths. That value, un-sheared along the rows, goes to a per-couple array in
// shared memory (one IMAD + one STS per row, nothing on the ALU pipe); the sinks are taken from it after the sweep:
// GLOBAL = row N (sw_inl.h:745-746), SEMI_GLOBAL = the maximum over the rows, the last one among equals (sw_inl.h:518-524
// with BestSink's <=, sink_inl.h:44).
#include "dp_fast.h"

#define DF2_WARPS   4
#define F2_NO_PAIR  0xFFFFFFFFu     // entry of a coupling order: no pair here (dpf2_order_by_length)
#define DF2_THREADS (DF2_WARPS * 32)

struct Full2Params
{
    StrSet         P, T;
    nvb_best_sink* sinks;
    uint32_t*      worklist;   // [0] = count, [1..] = pair ids
    uint32_t       n_pairs, n_tasks;
    uint32_t       lanes, tpw;     // stripes per couple, couples per warp
    uint32_t       tstride;        // text words per couple = ceil(maxN / 8)
    uint32_t       cstride;        // candidate words per couple = maxN (+ lanes in the selector-stream kernels)
    uint32_t       sstride;        // selector-stream words per couple = ceil((maxN + 2 lanes) / 2)
    int32_t        match, mismatch, ga, gb;
    int32_t        semi;           // 1 = SEMI_GLOBAL, 0 = GLOBAL
    uint32_t       one;
    const uint32_t* order;         // pair ids in the order they are coupled (NULL: 0, 1, 2, ...); see dpf2_order_by_length
};

__host__ __device__ __forceinline__ uint32_t f2_pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }
__device__ __forceinline__ uint32_t f2_prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ uint32_t f2_imad(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t f2_squeeze4(uint32_t x)
{
    x &= 0x03030303u; x = (x | (x >> 4)) & 0x00FF00FFu; x = (x | (x >> 8)) & 0x0000FFFFu;
    return x;
}
__device__ __forceinline__ void f2_mark_dirty(uint32_t* wl, uint32_t id)
{
    const uint32_t slot = atomicAdd(wl, 1u);
    wl[1 + slot] = id;
}

// the couple's two texts as a nibble stream (8 rows per word, nibble = rA | rB << 2), written by the couple's lanes
__device__ __forceinline__ void f2_pack_text(const StrSet& T, uint32_t* my_text, uint32_t k, uint32_t lanes, uint32_t tbA, uint32_t NA,
                                             uint32_t tbB, uint32_t NB, uint32_t Nmax, bool& dirtyA, bool& dirtyB)
{
    for (uint32_t w = k; w * 8 < Nmax; w += lanes)
    {
        auto rows8 = [&](const uint32_t tb, const uint32_t N, bool& dirty) -> uint32_t
        {
            const uint32_t i0 = w * 8;
            const uint8_t* src = T.sym + tb + i0;
            if (T.bits == 8 && i0 + 8 <= N && ((uintptr_t)src & 3u) == 0)
            {
                const uint32_t x0 = __ldg((const uint32_t*)src), x1 = __ldg((const uint32_t*)src + 1);
                dirty |= ((x0 | x1) & 0xFCFCFCFCu) != 0;
                return f2_squeeze4(x0) | (f2_squeeze4(x1) << 16);
            }
            uint32_t word = 0;
            #pragma unroll
            for (uint32_t e = 0; e < 8; ++e)
            {
                const uint32_t r = i0 + e < N ? sym_at_rt(T.sym, T.bits, tb + i0 + e) : 0u;
                dirty |= r > 3u;
                word |= (r & 3u) << (4 * e);
            }
            return word;
        };
        my_text[w] = rows8(tbA, NA, dirtyA) | (rows8(tbB, NB, dirtyB) << 2);
    }
}

// the couple's two texts as a stream of ready-made PRMT selectors, one 16-bit entry per row: entry pad + r selects byte rA of
// the first profile word into the low half and byte rB of the second into the high half, sign-extended
// (((rA | rB << 8) * 0x11) + 0xC480). `pad` entries before row 0 and `pad + 1` after the last row hold a valid selector, so the
// ramp-up and drain steps of the wavefront read the stream without a bounds check. One LDS.U16 per row replaces the nibble
// extraction and selector arithmetic of the nibble stream (5 ALU-pipe instructions per row).
__device__ __forceinline__ void f2_pack_selectors(const StrSet& T, uint16_t* my_sel, uint32_t k, uint32_t lanes, uint32_t pad, uint32_t tbA, uint32_t NA,
                                                  uint32_t tbB, uint32_t NB, uint32_t Nmax, bool& dirtyA, bool& dirtyB)
{
    for (uint32_t e = k; e < pad; e += lanes) my_sel[e] = 0xC480u;
    for (uint32_t e = k; e < pad + 1u; e += lanes) my_sel[pad + Nmax + e] = 0xC480u;
    for (uint32_t w = k; w * 8 < Nmax; w += lanes)
    {
        const uint32_t i0 = w * 8;
        auto rows8 = [&](const uint32_t tb, const uint32_t N, bool& dirty, uint32_t& lo, uint32_t& hi)
        {
            const uint8_t* src = T.sym + tb + i0;
            if (T.bits == 8 && i0 + 8 <= N && ((uintptr_t)src & 3u) == 0)
            {
                lo = __ldg((const uint32_t*)src); hi = __ldg((const uint32_t*)src + 1);
                dirty |= ((lo | hi) & 0xFCFCFCFCu) != 0;
                lo &= 0x03030303u; hi &= 0x03030303u;
                return;
            }
            lo = hi = 0;
            #pragma unroll
            for (uint32_t e = 0; e < 8; ++e)
            {
                const uint32_t r = i0 + e < N ? sym_at_rt(T.sym, T.bits, tb + i0 + e) : 0u;
                dirty |= r > 3u;
                if (e < 4) lo |= (r & 3u) << (8 * e); else hi |= (r & 3u) << (8 * (e - 4));
            }
        };
        uint32_t al, ah, bl, bh;
        rows8(tbA, NA, dirtyA, al, ah); rows8(tbB, NB, dirtyB, bl, bh);
        #pragma unroll
        for (uint32_t e = 0; e < 8; ++e)
        {
            if (i0 + e >= Nmax) break;
            const uint32_t a = ((e < 4 ? al : ah) >> (8 * (e & 3))) & 3u, b = ((e < 4 ? bl : bh) >> (8 * (e & 3))) & 3u;
            my_sel[pad + i0 + e] = (uint16_t)((a | (b << 8)) * 0x11u + 0xC480u);
        }
    }
}

// the stripe's 16 pattern symbols of one pair as 4 words of bytes (0xFF past the pattern)
__device__ __forceinline__ void f2_stripe16(const StrSet& P, uint32_t pb, uint32_t M, uint32_t block, uint32_t* q4)
{
    const uint8_t* src = P.sym + pb + block;
    if (P.bits == 8 && block + 16u <= M && ((uintptr_t)src & 3u) == 0)
    {
        #pragma unroll
        for (int w = 0; w < 4; ++w) q4[w] = __ldg((const uint32_t*)src + w);
    }
    else
    {
        #pragma unroll
        for (int w = 0; w < 4; ++w)
        {
            uint32_t x = 0;
            #pragma unroll
            for (int e = 0; e < 4; ++e)
            {
                const uint32_t c = block + 4 * w + e;
                x |= (c < M ? sym_at_rt(P.sym, P.bits, pb + c) : 0xFFu) << (8 * e);
            }
            q4[w] = x;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// part 1: Smith-Waterman / edit distance, GLOBAL and SEMI_GLOBAL
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DF2_THREADS, 5)
dpf2_sw_kernel(const Full2Params p)
{
    extern __shared__ uint32_t s_mem[];    // per warp and couple slot: tstride text words, then cstride candidate words
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;       // couple slot, stripe
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks;
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false;
    if (live)
    {
        idA = 2 * task; hasB = idA + 1 < p.n_pairs; idB = hasB ? idA + 1 : idA;
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;

    uint32_t* my_words = s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * (p.sstride + p.cstride);
    uint16_t* my_sel   = (uint16_t*)my_words;                // selector of row r at entry pad + r
    uint32_t* my_cand  = my_words + p.sstride;               // h[r][M] of both pairs as the last lane sends it, at entry r + pad
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();

    // substitution profiles in sheared scores: S' = S - G - I (0 past the pattern)
    const int32_t sm = p.match - p.ga - p.gb, sx = p.mismatch - p.ga - p.gb;
    const uint32_t base  = ((uint32_t)sx & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)(sx ^ sm)) & 0xFFu;
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], H[16];
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, b = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0u, wb = 0u;
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = b > 3u ? base : base ^ (delta << (8u * b));
            TA[j] = wa; TB[j] = wb; H[j] = 0u;
        }
    }
    uint32_t r_one; asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    const uint32_t g2 = f2_pack2(-p.ga);                                   // -G in both halves
    // column 0 of the matrix, the left input of lane 0: GLOBAL 0, SEMI_GLOBAL -G r; other lanes take their neighbour's last cell
    const uint32_t cin_mul = k == 0 ? 0u : r_one, col_mul = (k == 0 && p.semi) ? g2 : 0u;
    const bool last_lane = live && k == p.lanes - 1u;
    uint32_t carry = 0u, send = 0u, colacc = 0u;
    const uint16_t* sel_at = my_sel + pad - k;               // step t reads row t - k
    uint32_t* cand_at = my_cand + pad - k;                   // row r of the storing (last) lane at entry pad + r

    // step t: this lane's text row t - k (a dummy row while t < k, swept away by the reset at t == k)
    auto row_step = [&](const uint32_t t, const bool first)
    {
        colacc = f2_imad(col_mul, r_one, colacc);            // column 0 of lane 0: (row + 1) * col_mul
        const uint32_t cin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send, 1), cin_mul, colacc);
        if (first && t == k)
        {
            #pragma unroll
            for (int j = 0; j < 16; ++j) H[j] = 0u;
            carry = 0u;
        }
        const uint32_t sel = sel_at[t];
        // phase A: diagonal candidates from the previous row (FMA pipe)
        uint32_t d[16];
        {
            uint32_t diag = carry;
            #pragma unroll
            for (int j = 0; j < 16; ++j)
            {
                d[j] = f2_imad(diag, r_one, f2_prmt(TA[j], TB[j], sel));
                diag = H[j];
            }
        }
        carry = cin;
        // phase B: one VIMNMX3 per packed cell along the row
        uint32_t left = cin;
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            left = __vimax3_s16x2(d[j], H[j], left);
            H[j] = left;
        }
        send = left;
        // h[r][M] of both pairs (still sheared along the rows) for the epilogue
        if (last_lane) cand_at[t] = send;
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t t_lo  = min(p.lanes, steps);
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true);
    for (; t < steps; ++t) row_step(t, false);
    __syncwarp();

    // ---- sinks from the candidate column: key = value << 16 | row, so the last row wins among equal values
    uint32_t bA = 0u, bB = 0u;
    if (live)
    {
        if (p.semi)
        {
            // un-sheared along the rows: + G r + const (const = -G Nmax keeps it >= 0)
            for (uint32_t r = k; r < Nmax; r += p.lanes)
            {
                const uint32_t c = f2_imad(Nmax - 1u - r, g2, my_cand[r + pad]);
                if (r < NA) bA = max(bA, ((c & 0xFFFFu) << 16) | r);
                if (r < NB) bB = max(bB, (c & 0xFFFF0000u) | r);
            }
        }
        else if (k == 0)
        {
            if (NA) bA = ((f2_imad(Nmax - NA, g2, my_cand[NA - 1u + pad]) & 0xFFFFu) << 16) | (NA - 1u);
            if (NB) bB = (f2_imad(Nmax - NB, g2, my_cand[NB - 1u + pad]) & 0xFFFF0000u) | (NB - 1u);
        }
    }
    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wA = 0u, wB = 0u;
    for (uint32_t dd = 0; dd < p.lanes; ++dd)
    {
        const uint32_t src = min(gbase + dd, 31u);
        wA = max(wA, __shfl_sync(0xFFFFFFFFu, bA, src));
        wB = max(wB, __shfl_sync(0xFFFFFFFFu, bB, src));
    }
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0;
        if (M == 0 || N == 0) dirty = true;                                 // degenerate shapes: the int32 kernel knows the reference's answers
        // 15-bit range of everything stored: the other pair's rows and the padded columns are swept too
        const uint32_t pos = p.match > 0 ? (uint32_t)p.match : 0u;
        if ((uint64_t)pos * min(max(MA, MB), Nmax) + 2ull * (uint32_t)(-p.ga) * Nmax + (uint64_t)(uint32_t)(-p.gb) * (16u * p.lanes) >= 32000ull) dirty = true;
        if (dirty) { f2_mark_dirty(p.worklist, id); continue; }
        const uint32_t w = h ? wB : wA;
        nvb_best_sink out;
        out.score  = (int32_t)(w >> 16) + p.ga * (int32_t)Nmax + p.gb * (int32_t)M;
        out.sink_x = (w & 0xFFFFu) + 1u;
        out.sink_y = M;
        p.sinks[id] = out;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// part 2: Gotoh. Reference recurrence (gotoh_inl.h:458-575), go = gap open, ge = gap extension, d = go - ge <= 0:
//   F[r][c] = max(F[r-1][c] + ge, H[r-1][c] + go)      E[r][c] = max(E[r][c-1] + ge, H[r][c-1] + go)
//   H[r][c] = max(E[r][c], F[r][c], H[r-1][c-1] + S)   (LOCAL: and 0)
// Both kernels keep HO = H + d (H with the opening surcharge already applied) so that extension and opening become one
// maximum each, and shear the coordinates by ge so that extending costs nothing.
// ---------------------------------------------------------------------------------------------------------------
// addend that adds g <= 0 to both halves of a register whose low half is >= -g (the low half's carry completes the high half's g - 1)
__host__ __device__ __forceinline__ uint32_t f2_fix2(int32_t g) { return g == 0 ? 0u : ((((uint32_t)(g - 1) & 0xFFFFu) << 16) | ((uint32_t)g & 0xFFFFu)); }

struct Gotoh2Params
{
    Full2Params f;
    uint32_t    sixteen;
    uint32_t    unshear;        // LOCAL: fix2(-16 b), the shear step between neighbouring lanes
    uint32_t    FL[16];         // LOCAL floor of local column j: pack2(B + b j)
    uint32_t    CJ[16];         // LOCAL key addend: (h' << 4) + CJ[j] = (H + B) << 4 | j
    int32_t     B;              // bias
};

#define DF2_BIAS 64

// ---- Gotoh LOCAL: sheared along the columns of a stripe only (the floor of local column j is the constant B + b j, b = -ge):
//   x'[j] = x + B + b j  for x = H, E, F;   HO' = H' + d
//   F'  = max(F', HO'top) + ge          E' = max(E'left, HO'left)
//   h'  = max3( max(HO'diag + (S - go), floor_j), F', E' )              HO' = h' + d
// per packed cell: PRMT + VIMNMX (F) + VIMNMX (E) + VIADDMNMX (diagonal with the floor folded in) + VIMNMX3 + 1/2 VIMNMX3 (key)
// on the ALU pipe; the three additions are IMADs. BestSink's order: the reference sweeps Gotoh in stripes of EIGHT columns
// (gotoh_inl.h:1259-1262), so inside a lane's 16 columns the upper half belongs to a later stripe: among equal scores the
// running best is replaced when (score, column >> 3) is >= -- keys compared with their low 3 bits cleared.
template <int BT>
__global__ void __launch_bounds__(DF2_THREADS, 4)
dpf2_gotoh_local_kernel(const Gotoh2Params q)
{
    const Full2Params& p = q.f;
    auto FLc = [&](const int j) -> uint32_t { return BT ? (uint32_t)(DF2_BIAS + BT * j) * 0x00010001u : q.FL[j]; };
    auto CJc = [&](const int j) -> uint32_t
    {
        if (!BT) return q.CJ[j];
        const int32_t cj = j - 16 * BT * j;
        return cj < 0 ? ((((uint32_t)(cj - 1) & 0xFFFFu) << 16) | ((uint32_t)cj & 0xFFFFu)) : 0u;
    };
    extern __shared__ uint32_t s_mem[];
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks;
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false;
    if (live)
    {
        idA = 2 * task; hasB = idA + 1 < p.n_pairs; idB = hasB ? idA + 1 : idA;
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;
    uint16_t* my_sel = (uint16_t*)(s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * p.sstride);   // selector of row r at entry pad + r
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();
    const uint16_t* sel_at = my_sel + pad - k;               // step t reads row t - k

    const int32_t go = p.ga, ge = p.gb, b = -ge, dd = go - ge, B = q.B;
    // profile bytes: S - go (the diagonal reads HO' = H' + d of the previous column: -d + b = -go)
    const uint32_t base  = ((uint32_t)(p.mismatch - go) & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)((p.mismatch - go) ^ (p.match - go))) & 0xFFu;
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], HO[16], F[16];
    const uint32_t d2 = f2_fix2(dd), ge2 = f2_fix2(ge);
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, bb = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0x80808080u, wb = 0x80808080u;       // columns past the pattern: S = -128
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = bb > 3u ? base : base ^ (delta << (8u * bb));
            TA[j] = wa; TB[j] = wb;
        }
    }
    const uint32_t smask = (block < MA ? 0x0000FFFFu : 0u) | (block < MB ? 0xFFFF0000u : 0u);
    uint32_t r_one, r_sixteen, r_m1;
    asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_sixteen) : "r"(q.sixteen));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_m1) : "r"(0u - p.one));
    // column 0 of the matrix in the shear of local column -1: H = 0 -> HO' = B - b + d; E = 0 (gotoh_inl.h:68-71) -> B - b
    const uint32_t ho_zero = f2_pack2(B - b + dd), e_zero = f2_pack2(B - b);
    // lane 0 has no left neighbour; other lanes take (HO', E') of their neighbour's column 15 out of its shear (-16 b)
    const uint32_t cin_mul = k == 0 ? 0u : r_one;
    const uint32_t ho_add = k == 0 ? ho_zero : q.unshear, e_add = k == 0 ? e_zero : q.unshear;
    uint32_t bestkey = f2_pack2(B) << 4, bestrow = 0u;
    uint32_t carry = ho_zero, send_h = ho_zero, send_e = e_zero;

    auto reset = [&]()
    {
        #pragma unroll
        for (int j = 0; j < 16; ++j) { HO[j] = f2_imad(FLc(j), r_one, d2); F[j] = 0u; }     // row 0: H = 0, F = -inf
        carry = ho_zero;
    };
    reset();

    auto row_step = [&](const uint32_t t, const bool masked, const bool first)
    {
        const uint32_t hin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_h, 1), cin_mul, ho_add);
        const uint32_t ein = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_e, 1), cin_mul, e_add);
        const uint32_t i = t - k;
        if (first && t == k) reset();
        const uint32_t sel = sel_at[t];

        const uint32_t kb = f2_imad(bestkey & 0xFFF8FFF8u, r_one, 0xFFFEFFFFu);
        uint32_t acc = kb, left = hin, e = ein, diag = carry;
        carry = hin;
        #pragma unroll
        for (int j = 0; j < 16; j += 2)
        {
            uint32_t hh[2];
            #pragma unroll
            for (int u = 0; u < 2; ++u)
            {
                const int jj = j + u;
                const uint32_t S = f2_prmt(TA[jj], TB[jj], sel);
                const uint32_t top = HO[jj];
                const uint32_t fn = f2_imad(__vmaxs2(F[jj], top), r_one, ge2);     // max(F, HO top) + ge
                const uint32_t dg = __viaddmax_s16x2(diag, S, FLc(jj));                   // max(diagonal, floor)
                e = __vmaxs2(e, left);                                               // E' of this column
                const uint32_t h = __vimax3_s16x2(dg, fn, e);
                F[jj] = fn; diag = top;
                left = f2_imad(h, r_one, d2);                                             // HO' = h' + d
                HO[jj] = left;
                hh[u] = h;
            }
            acc = __vimax3_s16x2(acc, f2_imad(hh[0], r_sixteen, CJc(j)), f2_imad(hh[1], r_sixteen, CJc(j + 1)));
        }
        send_h = left; send_e = e;

        const uint32_t df = f2_imad(kb, r_m1, acc);
        const uint32_t sgn = f2_imad(df, r_one, 0x7FFF7FFFu);
        uint32_t m = f2_prmt(sgn, sgn, 0xBB99u) & smask;
        if (masked) m &= (i < NA ? 0x0000FFFFu : 0u) | (i < NB ? 0xFFFF0000u : 0u);
        bestkey = (acc & m) | (bestkey & ~m);
        bestrow = (f2_imad(i, 0x00010001u, 0u) & m) | (bestrow & ~m);
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t Nmin  = __reduce_min_sync(0xFFFFFFFFu, live ? min(NA, hasB ? NB : NA) : 0xFFFFFFFFu);
    const uint32_t t_lo  = min(p.lanes, steps);
    const uint32_t t_hi  = max(t_lo, min(Nmin, steps));
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true, true);
    for (; t < t_hi; ++t)  row_step(t, false, false);
    for (; t < steps; ++t) row_step(t, true, false);

    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wkA = 0, wkB = 0, wlaneA = 0, wlaneB = 0;
    for (uint32_t l = 0; l < p.lanes; ++l)
    {
        const uint32_t src = min(gbase + l, 31u);
        const uint32_t ok = __shfl_sync(0xFFFFFFFFu, bestkey, src);
        const uint32_t sA = (ok & 0xFFFFu) >> 4, sB = ok >> 20;
        if (sA >= (wkA >> 4)) { wkA = ok & 0xFFFFu; wlaneA = l; }
        if (sB >= (wkB >> 4)) { wkB = ok >> 16; wlaneB = l; }
    }
    const uint32_t rowsA = __shfl_sync(0xFFFFFFFFu, bestrow, min(gbase + wlaneA, 31u));
    const uint32_t rowsB = __shfl_sync(0xFFFFFFFFu, bestrow, min(gbase + wlaneB, 31u));
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0;
        if (p.match > 0 && (uint64_t)p.match * min(max(MA, MB), Nmax) + (uint32_t)B + 15u * (uint32_t)b >= 2047u) dirty = true;
        if (dirty) { f2_mark_dirty(p.worklist, id); continue; }
        const uint32_t wk = h ? wkB : wkA;
        const uint32_t stripe = h ? wlaneB : wlaneA;
        const uint32_t row = h ? (rowsB >> 16) : (rowsA & 0xFFFFu);
        const int32_t score = (int32_t)(wk >> 4) - B;
        nvb_best_sink out;
        if (M == 0 || N == 0) { out.score = NVB_SINK_MIN; out.sink_x = out.sink_y = 0xFFFFFFFFu; }
        else if (score <= 0) { out.score = 0; out.sink_x = N; out.sink_y = M; }   // all cells tie at 0: the last one visited
        else { out.score = score; out.sink_x = row + 1u; out.sink_y = stripe * 16u + (wk & 15u) + 1u; }
        p.sinks[id] = out;
    }
}

// ---- Gotoh GLOBAL / SEMI_GLOBAL: sheared by ge in BOTH directions, x''[r][c] = x[r][c] - ge (r + c) + bias:
//   F''  = max(F'', HO''top)          E'' = max(E''left, HO''left)
//   h''  = max3( HO''diag + (S - ge - go), F'', E'' )                   HO'' = h'' + d
// per packed cell PRMT + 2 VIMNMX + VIMNMX3 on the ALU pipe, two IMADs. Boundaries: row 0 (go + ge (c - 1),
// gotoh_inl.h:675) and column 0 of GLOBAL become the constant d, column 0 of SEMI_GLOBAL (0) becomes -ge r; E of column 0
// and F of row 0 are -infinity = 0 below every biased value. The opening surcharge is paid on every move to the right, so
// columns past M do NOT carry h[r][M] along (as they do in part 1): the lane holding column M picks it out of its 16
// registers with a one-hot multiply-add chain on the FMA pipe, which needs both pairs of a couple to have the same M:
// the pairs of a batch are coupled in the order of a counting sort by pattern length (dpf2_order_by_length below).
__global__ void __launch_bounds__(DF2_THREADS, 4)
dpf2_gotoh_kernel(const Gotoh2Params q)
{
    const Full2Params& p = q.f;
    extern __shared__ uint32_t s_mem[];
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks && !(p.order && p.order[2 * min(task, p.n_tasks - 1u)] == F2_NO_PAIR);   // pads past the last run
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false;
    if (live)
    {
        hasB = 2 * task + 1 < p.n_pairs;
        idA = p.order ? p.order[2 * task] : 2 * task; idB = hasB ? (p.order ? p.order[2 * task + 1] : idA + 1) : idA;
        if (idB == F2_NO_PAIR) { hasB = false; idB = idA; }                  // a run of odd count: the pair sweeps alone
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;
    uint32_t* my_words = s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * (p.sstride + p.cstride);
    uint16_t* my_sel   = (uint16_t*)my_words;                // selector of row r at entry pad + r
    uint32_t* my_cand  = my_words + p.sstride;               // captured column M of row r at entry pad + r
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();
    const uint16_t* sel_at = my_sel + pad - k;               // step t reads row t - k
    uint32_t* cand_at = my_cand + pad - k;

    const int32_t go = p.ga, ge = p.gb, b = -ge, dd = go - ge, B = q.B;
    const int32_t sm = p.match - ge - go, sx = p.mismatch - ge - go;       // >= 0
    const uint32_t base  = ((uint32_t)sx & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)(sx ^ sm)) & 0xFFu;
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], HO[16], F[16], OH[16];
    const uint32_t d2 = f2_fix2(dd), b2 = f2_pack2(b);
    const bool capture = live && MA > 0 && k == (MA - 1u) / 16u;          // this lane holds column M (MA == MB, else the couple is dirty)
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, bb = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0u, wb = 0u;
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = bb > 3u ? base : base ^ (delta << (8u * bb));
            TA[j] = wa; TB[j] = wb;
            OH[j] = (capture && (uint32_t)j == ((MA - 1u) & 15u)) ? 1u : 0u;
        }
    }
    uint32_t r_one; asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    // row 0 and column 0 in sheared, biased form
    const uint32_t ho_row0 = f2_pack2(2 * dd + B), ho_origin = f2_pack2(dd + B);
    const uint32_t cin_mul = k == 0 ? 0u : r_one;
    const uint32_t col_mul = (k == 0 && p.semi) ? b2 : 0u;                 // SEMI_GLOBAL: H[r][0] = 0 -> -ge r
    const uint32_t col_add = k == 0 ? (p.semi ? ho_origin : ho_row0) : 0u; // GLOBAL: H[r][0] = go + ge (r - 1) -> d, like row 0
    uint32_t carry = 0u, send_h = 0u, send_e = 0u, colacc = col_add;

    auto reset = [&]()
    {
        #pragma unroll
        for (int j = 0; j < 16; ++j) { HO[j] = ho_row0; F[j] = 0u; }
        carry = k == 0 ? ho_origin : ho_row0;                               // H[0][0] = 0, H[0][c > 0] = go + ge (c - 1)
    };
    reset();

    auto row_step = [&](const uint32_t t, const bool first)
    {
        colacc = f2_imad(col_mul, r_one, colacc);            // column 0 of lane 0: col_add + (row + 1) * col_mul
        const uint32_t hin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_h, 1), cin_mul, colacc);
        const uint32_t ein = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_e, 1), cin_mul, 0u);        // column 0: E = -infinity
        if (first && t == k) reset();
        const uint32_t sel = sel_at[t];
        uint32_t left = hin, e = ein, diag = carry, cap = 0u;
        carry = hin;
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t top = HO[j];
            const uint32_t fn = __vmaxs2(F[j], top);
            const uint32_t dg = f2_imad(diag, r_one, f2_prmt(TA[j], TB[j], sel));
            e = __vmaxs2(e, left);
            const uint32_t h = __vimax3_s16x2(dg, fn, e);
            F[j] = fn; diag = top;
            left = f2_imad(h, r_one, d2);
            HO[j] = left;
            cap = f2_imad(left, OH[j], cap);
        }
        send_h = left; send_e = e;
        // HO''[r][M], still sheared along the rows, for the epilogue
        if (capture) cand_at[t] = cap;
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t t_lo  = min(p.lanes, steps);
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true);
    for (; t < steps; ++t) row_step(t, false);
    __syncwarp();

    uint32_t bA = 0u, bB = 0u;
    if (live)
    {
        if (p.semi)
        {
            // un-sheared along the rows: + ge r, shifted by -ge Nmax to stay >= 0
            for (uint32_t r = k; r < Nmax; r += p.lanes)
            {
                const uint32_t c = f2_imad(Nmax - 1u - r, b2, my_cand[r + pad]);
                if (r < NA) bA = max(bA, ((c & 0xFFFFu) << 16) | r);
                if (r < NB) bB = max(bB, (c & 0xFFFF0000u) | r);
            }
        }
        else if (k == 0)
        {
            if (NA) bA = ((f2_imad(Nmax - NA, b2, my_cand[NA - 1u + pad]) & 0xFFFFu) << 16) | (NA - 1u);
            if (NB) bB = (f2_imad(Nmax - NB, b2, my_cand[NB - 1u + pad]) & 0xFFFF0000u) | (NB - 1u);
        }
    }
    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wA = 0u, wB = 0u;
    for (uint32_t l = 0; l < p.lanes; ++l)
    {
        const uint32_t src = min(gbase + l, 31u);
        wA = max(wA, __shfl_sync(0xFFFFFFFFu, bA, src));
        wB = max(wB, __shfl_sync(0xFFFFFFFFu, bB, src));
    }
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0;
        if (M == 0 || N == 0 || (hasB && MA != MB)) dirty = true;
        const uint32_t pos = p.match > 0 ? (uint32_t)p.match : 0u;
        if ((uint64_t)pos * min(M, Nmax) + (uint64_t)(uint32_t)b * (2ull * Nmax + 16u * p.lanes) + (uint32_t)B >= 32000ull) dirty = true;
        if (dirty) { f2_mark_dirty(p.worklist, id); continue; }
        const uint32_t w = h ? wB : wA;
        nvb_best_sink out;
        // value = H''[r][M] + d + B - ge (Nmax - r)  ->  H = value - d - B + ge (Nmax + M)
        out.score  = (int32_t)(w >> 16) - dd - B + ge * (int32_t)(Nmax + M);
        out.sink_x = (w & 0xFFFFu) + 1u;
        out.sink_y = M;
        p.sinks[id] = out;
    }
}

// the largest shared-memory footprint among the kernels: the selector-stream kernels keep (maxN + 2 lanes) 16-bit selectors and
// (maxN + lanes) candidate words per couple, the nibble-stream kernels maxN / 8 text words and maxN candidate words
static bool full2_geometry(uint32_t maxM, uint32_t maxN, uint32_t* lanes, uint32_t* tpw, uint32_t* tstride, uint32_t* cstride, size_t* smem,
                           uint32_t* sstride = nullptr, size_t* smem_sel = nullptr)
{
    const uint32_t stripes = (maxM + 15u) / 16u;
    if (stripes == 0 || stripes > 32u) return false;
    *lanes = stripes; *tpw = 32u / stripes; *tstride = (maxN + 7u) / 8u; *cstride = maxN;
    *smem = (size_t)DF2_WARPS * *tpw * (*tstride + *cstride) * sizeof(uint32_t);
    const uint32_t ss = (maxN + 2u * stripes + 1u) / 2u;
    const size_t sel_bytes = (size_t)DF2_WARPS * *tpw * (ss + maxN + 2u * stripes) * sizeof(uint32_t);
    if (sstride) *sstride = ss;
    if (smem_sel) *smem_sel = sel_bytes;
    return *smem <= 160u * 1024u && sel_bytes <= 160u * 1024u;
}

static bool sw_supported(const nvb_aligner& a)
{
    if (a.type != NVB_GLOBAL && a.type != NVB_SEMI_GLOBAL) return false;
    if (a.kind != NVB_SMITH_WATERMAN && a.kind != NVB_EDIT_DISTANCE) return false;
    const Scheme s = make_scheme(a);
    if (s.ga > 0 || s.gb > 0 || s.ga < -255 || s.gb < -255) return false;
    const int32_t sm = s.match - s.ga - s.gb, sx = s.mismatch - s.ga - s.gb;     // sheared substitution scores: bytes 0..127
    return !(sm < 0 || sm > 127 || sx < 0 || sx > 127);
}
static bool gotoh_supported(const nvb_aligner& a)
{
    if (a.kind != NVB_GOTOH) return false;
    const Scheme s = make_scheme(a);
    const int32_t go = s.ga, ge = s.gb;
    if (ge > 0 || go > ge || go < -64) return false;            // d = go - ge <= 0: opening never cheaper than extending
    if (a.type == NVB_LOCAL)
        return s.match - go <= 127 && s.match - go >= -127 && s.mismatch - go <= 127 && s.mismatch - go >= -127;
    const int32_t sm = s.match - ge - go, sx = s.mismatch - ge - go;
    return !(sm < 0 || sm > 127 || sx < 0 || sx > 127);
}

// The Gotoh GLOBAL / SEMI_GLOBAL sweeps capture column M of a couple in ONE lane and register, so both pairs of a couple must have
// the same pattern length. Reads of one length are the common case; for a batch of mixed lengths the pairs are coupled in the order of
// a counting sort by pattern length (histogram, scan, scatter: three small launches, warp-aggregated atomics). Every length's run is
// padded to an even count with the entry 0xFFFFFFFF ("no second pair"), so no couple ever mixes two lengths -- a single pair left to
// the int32 kernel would cost that kernel's full one-thread latency (milliseconds at 150 x 500). The order array therefore has
// dpf2_order_entries(n, maxM) >= n entries; entries past the last run are 0xFFFFFFFF too. Which pairs share a register does not
// change any pair's result.
__global__ void __launch_bounds__(256)
f2_len_hist_kernel(const uint32_t* __restrict__ off, uint32_t n, uint32_t cap, uint32_t* __restrict__ bins)
{
    const uint32_t i = blockIdx.x * 256 + threadIdx.x;
    const bool on = i < n;
    const uint32_t len = on ? min(off[i + 1] - off[i], cap) : 0xFFFFFFFFu;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, len);
    if (on && (threadIdx.x & 31u) == (uint32_t)(__ffs(peers) - 1)) atomicAdd(bins + len, (uint32_t)__popc(peers));
}
// counts -> start of every length's run, each run padded to an even number of entries (one block)
__global__ void __launch_bounds__(1024)
f2_len_scan_kernel(uint32_t* __restrict__ bins, uint32_t nb)
{
    __shared__ uint32_t s_warp[32];
    uint32_t run = 0;
    for (uint32_t b0 = 0; b0 < nb; b0 += 1024u)
    {
        const uint32_t b = b0 + threadIdx.x;
        const uint32_t c = b < nb ? ((bins[b] + 1u) & ~1u) : 0u;
        uint32_t x = c;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, d); if ((threadIdx.x & 31u) >= (uint32_t)d) x += y; }
        if ((threadIdx.x & 31u) == 31u) s_warp[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32u)
        {
            uint32_t w = s_warp[threadIdx.x];
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, w, d); if (threadIdx.x >= (uint32_t)d) w += y; }
            s_warp[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t before = (threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1u] : 0u;
        if (b < nb) bins[b] = run + before + x - c;
        run += s_warp[31];
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256)
f2_len_scatter_kernel(const uint32_t* __restrict__ off, uint32_t n, uint32_t cap, uint32_t* __restrict__ bins, uint32_t* __restrict__ order)
{
    const uint32_t i = blockIdx.x * 256 + threadIdx.x;
    const bool on = i < n;
    const uint32_t len = on ? min(off[i + 1] - off[i], cap) : 0xFFFFFFFFu;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, len);
    const uint32_t lane = threadIdx.x & 31u, leader = (uint32_t)(__ffs(peers) - 1);
    uint32_t base = 0;
    if (on && lane == leader) base = atomicAdd(bins + len, (uint32_t)__popc(peers));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (on) order[base + __popc(peers & ((1u << lane) - 1u))] = i;
}

enum { F2_SLOT_ORDER = 76 };        // slot map in common.cuh
bool dpf2_wants_order(const nvb_aligner& a)
{
    static int off = -1;                                   // NVB_NO_ORDER=1: couple the pairs as they come (tools/ragged_probe.py)
    if (off < 0) { const char* e = getenv("NVB_NO_ORDER"); off = e ? atoi(e) : 0; }
    return !off && a.kind == NVB_GOTOH && a.type != NVB_LOCAL;
}
// entries of the order array: n pairs + at most one pad per length bin, rounded up to whole couples
uint32_t dpf2_order_entries(uint32_t n, uint32_t maxM) { return (n + maxM + 2u + 1u) & ~1u; }
int dpf2_order_by_length(nvb_ctx* ctx, const StrSet& P, uint32_t maxM, const uint32_t** order, cudaStream_t st)
{
    const uint32_t n = P.count, nb = maxM + 2u;            // lengths 0 .. maxM, and "longer"
    const uint32_t entries = dpf2_order_entries(n, maxM);
    void* buf = nullptr;
    int rc = nvb_stage(ctx, F2_SLOT_ORDER, sizeof(uint32_t) * ((size_t)entries + nb), &buf);
    if (rc) return rc;
    uint32_t* bins = (uint32_t*)buf; uint32_t* ord = bins + nb;
    NVB_CUDA(ctx, cudaMemsetAsync(bins, 0, sizeof(uint32_t) * nb, st));
    NVB_CUDA(ctx, cudaMemsetAsync(ord, 0xFF, sizeof(uint32_t) * (size_t)entries, st));
    f2_len_hist_kernel<<<(n + 255) / 256, 256, 0, st>>>(P.off, n, maxM + 1u, bins);
    NVB_LAUNCHED(ctx);
    f2_len_scan_kernel<<<1, 1024, 0, st>>>(bins, nb);
    NVB_LAUNCHED(ctx);
    f2_len_scatter_kernel<<<(n + 255) / 256, 256, 0, st>>>(P.off, n, maxM + 1u, bins, ord);
    NVB_LAUNCHED(ctx);
    *order = ord;
    return NVB_OK;
}

bool dpf2_full_score_supported(const nvb_aligner& a, uint32_t maxM, uint32_t maxN)
{
    if (!sw_supported(a) && !gotoh_supported(a)) return false;
    if (maxN == 0 || maxM == 0 || maxN > 65535u) return false;
    uint32_t lanes, tpw, tstride, cstride; size_t smem;
    return full2_geometry(maxM, maxN, &lanes, &tpw, &tstride, &cstride, &smem);
}

int dpf2_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, nvb_best_sink* sinks, uint32_t maxM,
                    uint32_t maxN, uint32_t* worklist, cudaStream_t st, const uint32_t* order)
{
    const Scheme s = make_scheme(a);
    Gotoh2Params q;
    Full2Params& p = q.f;
    size_t smem, smem_sel;
    if (!full2_geometry(maxM, maxN, &p.lanes, &p.tpw, &p.tstride, &p.cstride, &smem, &p.sstride, &smem_sel)) return NVB_ERR_UNSUPPORTED;
    p.P = P; p.T = T; p.sinks = sinks; p.worklist = worklist; p.order = dpf2_wants_order(a) ? order : nullptr;
    p.n_pairs = p.order ? dpf2_order_entries(P.count, maxM) : P.count; p.n_tasks = (p.n_pairs + 1) / 2;
    p.match = s.match; p.mismatch = s.mismatch; p.ga = s.ga; p.gb = s.gb;
    p.semi = a.type == NVB_SEMI_GLOBAL ? 1 : 0;
    p.one = 1;
    if (NVB_FIRST_ON_DEVICE(ctx))
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_local_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_local_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_local_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_local_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    }
    p.cstride = maxN + 2u * p.lanes;                         // candidate rows are stored unconditionally: pad on both sides
    const uint32_t n_warps = (p.n_tasks + p.tpw - 1) / p.tpw;
    const uint32_t grid = (n_warps + DF2_WARPS - 1) / DF2_WARPS;
    if (a.kind != NVB_GOTOH)
        dpf2_sw_kernel<<<grid, DF2_THREADS, smem_sel, st>>>(p);
    else if (a.type == NVB_LOCAL)
    {
        const int32_t b = -s.gb;
        const int bt = (b >= 1 && b <= 3) ? b : 0;
        q.B = DF2_BIAS;
        q.sixteen = 16;
        q.unshear = (((uint32_t)(-16 * b - 1) & 0xFFFFu) << 16) | ((uint32_t)(-16 * b) & 0xFFFFu);
        if (b == 0) q.unshear = 0u;
        for (int j = 0; j < 16; ++j)
        {
            q.FL[j] = ((uint32_t)(q.B + b * j) & 0xFFFFu) * 0x00010001u;
            const int32_t cj = j - 16 * b * j;
            q.CJ[j] = cj < 0 ? ((((uint32_t)(cj - 1) & 0xFFFFu) << 16) | ((uint32_t)cj & 0xFFFFu)) : (uint32_t)cj * 0x00010001u;
        }
        const size_t smem_l = (size_t)DF2_WARPS * p.tpw * p.sstride * sizeof(uint32_t);
        switch (bt)
        {
        case 1:  dpf2_gotoh_local_kernel<1><<<grid, DF2_THREADS, smem_l, st>>>(q); break;
        case 2:  dpf2_gotoh_local_kernel<2><<<grid, DF2_THREADS, smem_l, st>>>(q); break;
        case 3:  dpf2_gotoh_local_kernel<3><<<grid, DF2_THREADS, smem_l, st>>>(q); break;
        default: dpf2_gotoh_local_kernel<0><<<grid, DF2_THREADS, smem_l, st>>>(q); break;
        }
    }
    else
    {
        const int32_t dd = s.ga - s.gb;
        q.B = 3 * (-dd) + 4;
        q.sixteen = 16; q.unshear = 0;
        dpf2_gotoh_kernel<<<grid, DF2_THREADS, smem_sel, st>>>(q);
    }
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// part 3: full-matrix traceback, Smith-Waterman / edit distance GLOBAL and SEMI_GLOBAL (aln::alignment_traceback,
// alignment_inl.h:355-455 over sw_inl.h:1321-1693). The reference sweeps the matrix once for the sink while saving a
// checkpoint column every CHECKPOINTS pattern columns, then recomputes one window of columns at a time, keeping a flow byte per
// cell, and walks it; the int32 kernel (dp_generic_full.cuh) does exactly that. Here ONE packed forward pass -- the sweep of
// dpf2_sw_kernel -- writes the direction of every cell, and one thread per pair walks them: the directions of a cell do not depend
// on where the checkpoints are.
//
// The direction comes out of the maximum itself (the trick of dp_band2_tb.cu): every value is kept times four and each candidate
// carries its move in the two low bits -- diagonal 2, left (insertion) 1, top (deletion) 0 -- so VIMNMX3 picks, among equal scores,
// exactly what sw_inl.h:488-490 picks: the diagonal unless a gap is strictly better, the insertion unless the deletion is strictly
// better. A cell costs PRMT + VIMNMX3 + LOP3 (re-tag as `left`) on the ALU pipe and four IMADs on the FMA pipe (the tag is h - clean).
// Flow: 2 bits per cell and pair, one 64-bit word per stripe and row (columns 0..7 of both pairs in .x, 8..15 in .y, column j at
// bits 2 (7 - j mod 8), pair B 16 bits up); for the layout see f2_flow_base.
// ---------------------------------------------------------------------------------------------------------------
struct Full2TbParams
{
    Full2Params    f;              // sinks unused
    nvb_alignment* out;
    uint32_t*      n_ops;
    uint8_t*       ops;
    uint32_t       ops_stride;
    uint2*         flow;
    uint32_t       first, n_launch;        // pairs [first, first + n_launch)
    uint32_t       maxM, maxN;
    uint32_t       by_step;                // flow layout: 1 = [couple][step][stripe], 0 = [couple][stripe][row]
    uint32_t       ahead;                  // rows the walk prefetches ahead of itself
};

// where stripe k stores the flow word of step t (row t - k); the first form makes the lanes of a couple write one contiguous run per
// step (the SW / ED sweeps were bound by their scattered stores), the second keeps a stripe's rows together for the walk (the Gotoh
// sweeps are bound by the ALU pipe, their walks by the 16-byte words they fetch). Rows that do not exist are stored too (the scratch
// holds maxN + lanes entries per stripe either way).
__device__ __forceinline__ size_t f2_flow_base(const Full2TbParams& q, uint32_t task, uint32_t k)
{ return (size_t)task * q.f.lanes * (q.maxN + q.f.lanes) + (q.by_step ? (size_t)k : (size_t)k * (q.maxN + q.f.lanes)); }
__device__ __forceinline__ size_t f2_flow_step(const Full2TbParams& q) { return q.by_step ? (size_t)q.f.lanes : (size_t)1; }
__device__ __forceinline__ size_t f2_flow_at(const Full2TbParams& q, uint32_t couple, uint32_t stripe, uint32_t row)
{ return f2_flow_base(q, couple, stripe) + (size_t)(row + stripe) * f2_flow_step(q); }

__global__ void __launch_bounds__(DF2_THREADS, 4)
dpf2_sw_tb_forward_kernel(const Full2TbParams q)
{
    const Full2Params& p = q.f;
    extern __shared__ uint32_t s_mem[];
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks;
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false, longA = false, longB = false;
    if (live)
    {
        idA = q.first + 2 * task; hasB = idA + 1 < q.first + q.n_launch; idB = hasB ? idA + 1 : idA;
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
        // a pair longer than the launch was sized for is left to the int32 kernel (which reports it as invalid) and swept as empty
        longA = MA > q.maxM || NA > q.maxN; longB = MB > q.maxM || NB > q.maxN;
        if (longA) MA = NA = 0;
        if (longB) MB = NB = 0;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;

    uint32_t* my_words = s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * (p.sstride + p.cstride);
    uint16_t* my_sel   = (uint16_t*)my_words;
    uint32_t* my_cand  = my_words + p.sstride;
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();

    // profiles: 4 S' + 2 (the diagonal's tag) inside the pattern, 0 past it
    const int32_t sm = 4 * (p.match - p.ga - p.gb) + 2, sx = 4 * (p.mismatch - p.ga - p.gb) + 2;
    const uint32_t base  = ((uint32_t)sx & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)(sx ^ sm)) & 0xFFu;
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], H[16];          // H: the row above, clean (tag 0 = the deletion's)
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, b = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0u, wb = 0u;
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = b > 3u ? base : base ^ (delta << (8u * b));
            TA[j] = wa; TB[j] = wb; H[j] = 0u;
        }
    }
    uint32_t r_one, r_m1;
    asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_m1) : "r"(0u - p.one));
    const uint32_t g2 = f2_pack2(-4 * p.ga);                               // -4 G in both halves
    const uint32_t minus1 = 0u - 0x00010001u;                              // takes the `left` tag off both halves (each half >= 1)
    const uint32_t cin_mul = k == 0 ? 0u : r_one, col_mul = (k == 0 && p.semi) ? g2 : 0u;
    const bool last_lane = live && k == p.lanes - 1u;
    // column 0 of the matrix as lane 0's left input, tagged as a left candidate: 4 h + 1
    uint32_t carry = 0u, send = 0u, colacc = k == 0 ? 0x00010001u : 0u;
    const uint16_t* sel_at = my_sel + pad - k;
    uint32_t* cand_at = my_cand + pad - k;
    uint2* flow_at = q.flow + f2_flow_base(q, task, k);
    const size_t flow_step = f2_flow_step(q);

    auto row_step = [&](const uint32_t t, const bool first)
    {
        colacc = f2_imad(col_mul, r_one, colacc);
        const uint32_t cin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send, 1), cin_mul, colacc);
        if (first && t == k)
        {
            #pragma unroll
            for (int j = 0; j < 16; ++j) H[j] = 0u;
            carry = 0u;
        }
        const uint32_t sel = sel_at[t];
        uint32_t d[16];
        {
            uint32_t diag = carry;
            #pragma unroll
            for (int j = 0; j < 16; ++j)
            {
                d[j] = f2_imad(diag, r_one, f2_prmt(TA[j], TB[j], sel));
                diag = H[j];
            }
        }
        carry = f2_imad(r_one, minus1, cin);                 // this row's column-0 neighbour, clean: the next row's first diagonal
        uint32_t left = cin, acc0 = 0u, acc1 = 0u;
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t hn = __vimax3_s16x2(d[j], H[j], left);
            left = (hn & 0xFFFCFFFCu) | 0x00010001u;
            H[j] = f2_imad(r_one, minus1, left);
            // acc * 4 + tag with tag = hn - H[j] (each half of hn is >= that of H[j]): two IMADs, nothing on the ALU pipe
            if (j < 8) acc0 = f2_imad(H[j], r_m1, f2_imad(acc0, 4u, hn)); else acc1 = f2_imad(H[j], r_m1, f2_imad(acc1, 4u, hn));
        }
        send = left;
        if (last_lane) cand_at[t] = send;
        if (live) flow_at[t * flow_step] = make_uint2(acc0, acc1);
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t t_lo  = min(p.lanes, steps);
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true);
    for (; t < steps; ++t) row_step(t, false);
    __syncwarp();

    // ---- sinks from the candidate column (4 h + 1, sheared along the rows), as in dpf2_sw_kernel
    uint32_t bA = 0u, bB = 0u;
    if (live)
    {
        if (p.semi)
        {
            for (uint32_t r = k; r < Nmax; r += p.lanes)
            {
                const uint32_t c = f2_imad(Nmax - 1u - r, g2, my_cand[r + pad]);
                if (r < NA) bA = max(bA, ((c & 0xFFFFu) << 16) | r);
                if (r < NB) bB = max(bB, (c & 0xFFFF0000u) | r);
            }
        }
        else if (k == 0)
        {
            if (NA) bA = ((f2_imad(Nmax - NA, g2, my_cand[NA - 1u + pad]) & 0xFFFFu) << 16) | (NA - 1u);
            if (NB) bB = (f2_imad(Nmax - NB, g2, my_cand[NB - 1u + pad]) & 0xFFFF0000u) | (NB - 1u);
        }
    }
    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wA = 0u, wB = 0u;
    for (uint32_t dd = 0; dd < p.lanes; ++dd)
    {
        const uint32_t src = min(gbase + dd, 31u);
        wA = max(wA, __shfl_sync(0xFFFFFFFFu, bA, src));
        wB = max(wB, __shfl_sync(0xFFFFFFFFu, bB, src));
    }
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0 || (h ? longB : longA);
        if (M == 0 || N == 0) dirty = true;
        const uint32_t pos = p.match > 0 ? (uint32_t)p.match : 0u;
        if (4ull * ((uint64_t)pos * min(max(MA, MB), Nmax) + 2ull * (uint32_t)(-p.ga) * Nmax + (uint64_t)(uint32_t)(-p.gb) * (16u * p.lanes)) + 8ull >= 32000ull) dirty = true;
        nvb_alignment a;
        a.score = NVB_SINK_MIN; a.source_x = a.source_y = a.sink_x = a.sink_y = 0xFFFFFFFFu;
        if (dirty) f2_mark_dirty(p.worklist, id);
        else
        {
            const uint32_t w = h ? wB : wA;
            a.score  = (int32_t)(w >> 18) + p.ga * (int32_t)Nmax + p.gb * (int32_t)M;
            a.sink_x = (w & 0xFFFFu) + 1u;
            a.sink_y = M;
        }
        q.out[id] = a;                // the walk fills in the source
        q.n_ops[id] = 0;
    }
}

// the walk (sw_inl.h:1644-1693 without the windows, then alignment_inl.h:437-450): one thread per pair
__global__ void __launch_bounds__(256)
dpf2_sw_tb_walk_kernel(const Full2TbParams q)
{
    const uint32_t lp = blockIdx.x * 256 + threadIdx.x;
    if (lp >= q.n_launch) return;
    const uint32_t id = q.first + lp;
    const nvb_alignment a = q.out[id];
    if (a.sink_x == 0xFFFFFFFFu) return;                       // left to the int32 kernel
    const uint32_t up = (lp & 1u) * 16u;
    const uint2* flow = q.flow;
    OpSink sink; sink.init(q.ops + (size_t)id * q.ops_stride, q.ops_stride);
    uint32_t row = a.sink_x, col = a.sink_y;                   // cells still to explain: text rows, pattern columns
    while (row > 0 && col > 0)
    {
        const uint32_t c = col - 1u;
        const uint2 w = flow[f2_flow_at(q, lp >> 1, c >> 4, row - 1u)];
        // the path climbs a row at most every step: ask for the word it will need `ahead` rows up now (a dependent DRAM round trip per row otherwise)
        if (row > q.ahead) asm volatile("prefetch.global.L1 [%0];" :: "l"(flow + f2_flow_at(q, lp >> 1, c >> 4, row - 1u - q.ahead)));
        const uint32_t tag = (((c & 8u) ? w.y : w.x) >> (2u * (7u - (c & 7u)) + up)) & 3u;
        if (tag == 3u)      break;                             // LOCAL: a cell of score 0 is the source
        if (tag == 2u)      { --row; --col; sink.push(OP_SUB); }
        else if (tag == 1u) { --col; sink.push(OP_INS); }
        else                { --row; sink.push(OP_DEL); }
    }
    if (q.f.semi != 2)                                         // GLOBAL / SEMI_GLOBAL: the rest of the pattern, GLOBAL: and of the text
    {
        if (row == 0) for (; col > 0; --col) sink.push(OP_INS);
        if (!q.f.semi && col == 0) for (; row > 0; --row) sink.push(OP_DEL);
    }
    sink.finish();
    q.out[id].source_x = row;
    q.out[id].source_y = col;
    q.n_ops[id] = sink.n;
}

// ---- Gotoh GLOBAL / SEMI_GLOBAL traceback (gotoh_inl.h:1259-1639): the sweep of dpf2_gotoh_kernel with three tagged maxima per
// cell, as in dp_band2_tb.cu. Everything is kept times four; HO travels as 4 HO + 1 ("HO1"), E as 4 E + 1, F clean:
//   F  = max(F, HO1 top)              low bit set = the gap was opened (the extension only wins when strictly larger, gotoh_inl.h:506)
//   E  = max(E1, HO1 left + 2)        bit 1 set   = opened
//   h  = max3(HO1 diag + (4 S'' + 1), F, E1)      tags 2 = substitution, 1 = insertion (E), 0 = deletion (F)
// One nibble per cell and pair: tag | opened-E << 2 | opened-F << 3; one 128-bit word per stripe and row (columns 4 q .. 4 q + 3 in
// word q, column j at bits 4 (j mod 4), pair B 16 bits up).
struct Gotoh2TbParams
{
    Full2TbParams t;
    int32_t       B;              // bias
};

__global__ void __launch_bounds__(DF2_THREADS, 3)
dpf2_gotoh_tb_forward_kernel(const Gotoh2TbParams gq)
{
    const Full2TbParams& q = gq.t;
    const Full2Params& p = q.f;
    extern __shared__ uint32_t s_mem[];
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks && !(p.order && p.order[q.first + 2 * min(task, p.n_tasks - 1u)] == F2_NO_PAIR);
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false, longA = false, longB = false;
    if (live)
    {
        hasB = 2 * task + 1 < q.n_launch;
        idA = p.order ? p.order[q.first + 2 * task] : q.first + 2 * task; idB = hasB ? (p.order ? p.order[q.first + 2 * task + 1] : idA + 1) : idA;
        if (idB == F2_NO_PAIR) { hasB = false; idB = idA; }
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
        longA = MA > q.maxM || NA > q.maxN; longB = MB > q.maxM || NB > q.maxN;
        if (longA) MA = NA = 0;
        if (longB) MB = NB = 0;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;
    uint32_t* my_words = s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * (p.sstride + p.cstride);
    uint16_t* my_sel   = (uint16_t*)my_words;
    uint32_t* my_cand  = my_words + p.sstride;
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();
    const uint16_t* sel_at = my_sel + pad - k;
    uint32_t* cand_at = my_cand + pad - k;
    uint4* flow_at = (uint4*)q.flow + f2_flow_base(q, task, k);
    const size_t flow_step = f2_flow_step(q);

    const int32_t go = p.ga, ge = p.gb, b = -ge, dd = go - ge, B = gq.B;
    const int32_t sm = 4 * (p.match - ge - go) + 1, sx = 4 * (p.mismatch - ge - go) + 1;
    const uint32_t base  = ((uint32_t)sx & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)(sx ^ sm)) & 0xFFu;
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], HO[16], F[16], OH[16];
    const int32_t d1 = 4 * dd + 1;                                          // clean h -> HO1
    const uint32_t dfix1 = d1 < 0 ? f2_fix2(d1) : f2_pack2(d1);
    const uint32_t b2 = f2_pack2(4 * b);
    const bool capture = live && MA > 0 && k == (MA - 1u) / 16u;
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, bb = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0x01010101u, wb = 0x01010101u;                    // past the pattern: S'' = 0, still the diagonal's tag
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = bb > 3u ? base : base ^ (delta << (8u * bb));
            TA[j] = wa; TB[j] = wb;
            OH[j] = (capture && (uint32_t)j == ((MA - 1u) & 15u)) ? 1u : 0u;
        }
    }
    uint32_t r_one; asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    const uint32_t ho_row0 = f2_pack2(4 * (2 * dd + B) + 1), ho_origin = f2_pack2(4 * (dd + B) + 1);
    const uint32_t cin_mul = k == 0 ? 0u : r_one;
    const uint32_t col_mul = (k == 0 && p.semi) ? b2 : 0u;
    const uint32_t col_add = k == 0 ? (p.semi ? ho_origin : ho_row0) : 0u;
    uint32_t carry = 0u, send_h = 0u, send_e = 0u, colacc = col_add;

    auto reset = [&]()
    {
        #pragma unroll
        for (int j = 0; j < 16; ++j) { HO[j] = ho_row0; F[j] = 0u; }
        carry = k == 0 ? ho_origin : ho_row0;
    };
    reset();

    auto row_step = [&](const uint32_t t, const bool first)
    {
        colacc = f2_imad(col_mul, r_one, colacc);
        const uint32_t hin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_h, 1), cin_mul, colacc);
        const uint32_t ein = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_e, 1), cin_mul, 0u);
        if (first && t == k) reset();
        const uint32_t sel = sel_at[t];
        uint32_t left = hin, e = ein, diag = carry, cap = 0u;
        uint32_t acc[4] = { 0u, 0u, 0u, 0u };
        carry = hin;
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t top = HO[j];
            const uint32_t ft = __vmaxs2(F[j], top);
            const uint32_t xF = ft & 0x00010001u;
            const uint32_t fn = ft - xF;
            const uint32_t dg = f2_imad(diag, r_one, f2_prmt(TA[j], TB[j], sel));
            const uint32_t et = __vmaxs2(e, left + 0x00020002u);
            const uint32_t xE = et & 0x00020002u;
            e = et - xE;
            const uint32_t h = __vimax3_s16x2(dg, fn, e);
            const uint32_t th = h & 0x00030003u;
            F[j] = fn; diag = top;
            left = f2_imad(h - th, r_one, dfix1);
            HO[j] = left;
            cap = f2_imad(left, OH[j], cap);
            const uint32_t m = 1u << (4 * (j & 3));
            acc[j >> 2] = f2_imad(th, m, f2_imad(xE, 2u * m, f2_imad(xF, 8u * m, acc[j >> 2])));
        }
        send_h = left; send_e = e;
        if (capture) cand_at[t] = cap;
        if (live) flow_at[t * flow_step] = make_uint4(acc[0], acc[1], acc[2], acc[3]);
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t t_lo  = min(p.lanes, steps);
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true);
    for (; t < steps; ++t) row_step(t, false);
    __syncwarp();

    uint32_t bA = 0u, bB = 0u;
    if (live)
    {
        if (p.semi)
        {
            for (uint32_t r = k; r < Nmax; r += p.lanes)
            {
                const uint32_t c = f2_imad(Nmax - 1u - r, b2, my_cand[r + pad]);
                if (r < NA) bA = max(bA, ((c & 0xFFFFu) << 16) | r);
                if (r < NB) bB = max(bB, (c & 0xFFFF0000u) | r);
            }
        }
        else if (k == 0)
        {
            if (NA) bA = ((f2_imad(Nmax - NA, b2, my_cand[NA - 1u + pad]) & 0xFFFFu) << 16) | (NA - 1u);
            if (NB) bB = (f2_imad(Nmax - NB, b2, my_cand[NB - 1u + pad]) & 0xFFFF0000u) | (NB - 1u);
        }
    }
    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wA = 0u, wB = 0u;
    for (uint32_t l = 0; l < p.lanes; ++l)
    {
        const uint32_t src = min(gbase + l, 31u);
        wA = max(wA, __shfl_sync(0xFFFFFFFFu, bA, src));
        wB = max(wB, __shfl_sync(0xFFFFFFFFu, bB, src));
    }
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0 || longA || longB;
        if (M == 0 || N == 0 || (hasB && MA != MB)) dirty = true;
        const uint32_t pos = p.match > 0 ? (uint32_t)p.match : 0u;
        if (4ull * ((uint64_t)pos * min(M, Nmax) + (uint64_t)(uint32_t)b * (2ull * Nmax + 16u * p.lanes) + (uint32_t)B) + 8ull >= 32000ull) dirty = true;
        nvb_alignment a;
        a.score = NVB_SINK_MIN; a.source_x = a.source_y = a.sink_x = a.sink_y = 0xFFFFFFFFu;
        if (dirty) f2_mark_dirty(p.worklist, id);
        else
        {
            const uint32_t w = h ? wB : wA;
            a.score  = (int32_t)(w >> 18) - dd - B + ge * (int32_t)(Nmax + M);
            a.sink_x = (w & 0xFFFFu) + 1u;
            a.sink_y = M;
        }
        q.out[id] = a;
        q.n_ops[id] = 0;
    }
}

// the Gotoh walk (gotoh_inl.h:1573-1639 without the windows, then alignment_inl.h:437-450)
__global__ void __launch_bounds__(256)
dpf2_gotoh_tb_walk_kernel(const Full2TbParams q)
{
    const uint32_t lp = blockIdx.x * 256 + threadIdx.x;
    if (lp >= q.n_launch) return;
    const uint32_t id = q.f.order ? q.f.order[q.first + lp] : q.first + lp;
    if (id == F2_NO_PAIR) return;
    const nvb_alignment a = q.out[id];
    if (a.sink_x == 0xFFFFFFFFu) return;
    const uint32_t up = (lp & 1u) * 16u;
    const uint4* flow = (const uint4*)q.flow;
    OpSink sink; sink.init(q.ops + (size_t)id * q.ops_stride, q.ops_stride);
    uint32_t row = a.sink_x, col = a.sink_y, state = ST_H;
    while (row > 0 && col > 0)
    {
        const uint32_t c = col - 1u;
        const uint4 w = flow[f2_flow_at(q, lp >> 1, c >> 4, row - 1u)];
        if (row > q.ahead) asm volatile("prefetch.global.L1 [%0];" :: "l"(flow + f2_flow_at(q, lp >> 1, c >> 4, row - 1u - q.ahead)));
        const uint32_t word = (c & 8u) ? ((c & 4u) ? w.w : w.z) : ((c & 4u) ? w.y : w.x);
        const uint32_t nib = (word >> (4u * (c & 3u) + up)) & 15u;
        if (state == ST_E)      { if (nib & 4u) state = ST_H; --col; sink.push(OP_INS); }
        else if (state == ST_F) { if (nib & 8u) state = ST_H; --row; sink.push(OP_DEL); }
        else
        {
            const uint32_t tag = nib & 3u;
            if (tag == 3u)      break;                         // LOCAL: a cell of score 0 is the source (gotoh_inl.h:1597-1604)
            if (tag == 1u)      state = ST_E;
            else if (tag == 0u) state = ST_F;
            else                { --row; --col; sink.push(OP_SUB); }
        }
    }
    if (q.f.semi != 2)
    {
        if (row == 0) for (; col > 0; --col) sink.push(OP_INS);
        if (!q.f.semi && col == 0) for (; row > 0; --row) sink.push(OP_DEL);
    }
    sink.finish();
    q.out[id].source_x = row;
    q.out[id].source_y = col;
    q.n_ops[id] = sink.n;
}

// ---- Smith-Waterman / edit distance LOCAL traceback: the sweep of dpf_full_sw_local_kernel (dp_fast_full.cu: sheared along the
// columns of a stripe, biased by B, arg-max keys score << 4 | column) on values kept times four, with the move in the two low bits:
//   h = max3( left | 1, max(diag + (4 (S + b) + 2), top + 4 G), floor_j | 3 )
// The floor carries the highest tag: a cell whose score is 0 is a SINK whatever else ties with it (sw_inl.h:1376). The clean value
// (4 h') times four is the scoring kernel's key, so the sink comes out of the same pass.
struct Local2TbParams
{
    Full2TbParams t;
    uint32_t      FL[16];         // floor of stripe column j, tagged: pack2(4 (B + b j) + 3)
    uint32_t      CJ[16];         // key addend, as in dp_fast_full.cu
    uint32_t      unshear;        // fix2(-64 b)
    int32_t       B;
};

__global__ void __launch_bounds__(DF2_THREADS, 4)
dpf2_sw_local_tb_forward_kernel(const Local2TbParams lq)
{
    const Full2TbParams& q = lq.t;
    const Full2Params& p = q.f;
    extern __shared__ uint32_t s_mem[];
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks;
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false, longA = false, longB = false;
    if (live)
    {
        idA = q.first + 2 * task; hasB = idA + 1 < q.first + q.n_launch; idB = hasB ? idA + 1 : idA;
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
        longA = MA > q.maxM || NA > q.maxN; longB = MB > q.maxM || NB > q.maxN;
        if (longA) MA = NA = 0;
        if (longB) MB = NB = 0;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;
    uint16_t* my_sel = (uint16_t*)(s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * p.sstride);
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();
    const uint16_t* sel_at = my_sel + pad - k;
    uint2* flow_at = q.flow + f2_flow_base(q, task, k);
    const size_t flow_step = f2_flow_step(q);

    const int32_t G = p.ga, I = p.gb, B = lq.B;
    const int32_t sm = 4 * (p.match - I) + 2, sx = 4 * (p.mismatch - I) + 2;
    const uint32_t base  = ((uint32_t)sx & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)(sx ^ sm)) & 0xFFu;
    const uint32_t block = k * 16u;
    const uint32_t Gf = f2_fix2(4 * G);
    uint32_t TA[16], TB[16], H[16];          // H: the row above, clean
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, bb = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0x80808080u, wb = 0x80808080u;       // columns past the pattern: -128
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = bb > 3u ? base : base ^ (delta << (8u * bb));
            TA[j] = wa; TB[j] = wb; H[j] = lq.FL[j] - 0x00030003u;
        }
    }
    const uint32_t smask = (block < MA ? 0x0000FFFFu : 0u) | (block < MB ? 0xFFFF0000u : 0u);
    uint32_t bestkey = f2_pack2(B) << 4, bestrow = 0u;
    uint32_t r_one, r_four, r_m1;
    asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_four) : "r"(4u * p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_m1) : "r"(0u - p.one));
    const uint32_t minus1 = 0u - 0x00010001u;
    const uint32_t zero_in = f2_pack2(4 * (B + I) + 1);         // H = 0 of the column left of the stripe, tagged as a left candidate
    uint32_t carry = zero_in - 0x00010001u, send = zero_in;
    const uint32_t cin_mul = k == 0 ? 0u : r_one, cin_add = k == 0 ? zero_in : lq.unshear;

    auto row_step = [&](const uint32_t t, const bool masked, const bool first)
    {
        const uint32_t cin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send, 1), cin_mul, cin_add);
        const uint32_t i = t - k;
        if (first && t == k)
        {
            #pragma unroll
            for (int j = 0; j < 16; ++j) H[j] = lq.FL[j] - 0x00030003u;
            carry = zero_in - 0x00010001u;
        }
        const uint32_t sel = sel_at[t];
        uint32_t v[16];
        {
            uint32_t diag = carry;
            #pragma unroll
            for (int j = 0; j < 16; ++j)
            {
                const uint32_t S = f2_prmt(TA[j], TB[j], sel);
                const uint32_t u = f2_imad(H[j], r_one, Gf);
                v[j] = __viaddmax_s16x2(diag, S, u);
                diag = H[j];
            }
        }
        carry = f2_imad(r_one, minus1, cin);
        const uint32_t kb = f2_imad(bestkey & 0xFFF0FFF0u, r_one, 0xFFFEFFFFu);
        uint32_t acc = kb, left = cin, a0 = 0u, a1 = 0u;
        #pragma unroll
        for (int j = 0; j < 16; j += 2)
        {
            const uint32_t h0 = __vimax3_s16x2(left, v[j], lq.FL[j]);
            const uint32_t t0 = h0 & 0x00030003u;
            const uint32_t c0 = h0 - t0;
            const uint32_t h1 = __vimax3_s16x2((h0 & 0xFFFCFFFCu) | 0x00010001u, v[j + 1], lq.FL[j + 1]);
            const uint32_t t1 = h1 & 0x00030003u;
            const uint32_t c1 = h1 - t1;
            H[j] = c0; H[j + 1] = c1; left = (h1 & 0xFFFCFFFCu) | 0x00010001u;
            acc = __vimax3_s16x2(acc, f2_imad(c0, r_four, lq.CJ[j]), f2_imad(c1, r_four, lq.CJ[j + 1]));
            if (j < 8) a0 = f2_imad(f2_imad(a0, 4u, t0), 4u, t1); else a1 = f2_imad(f2_imad(a1, 4u, t0), 4u, t1);
        }
        send = left;
        const uint32_t d = f2_imad(kb, r_m1, acc);
        const uint32_t sgn = f2_imad(d, r_one, 0x7FFF7FFFu);
        uint32_t m = f2_prmt(sgn, sgn, 0xBB99u) & smask;
        if (masked) m &= (i < NA ? 0x0000FFFFu : 0u) | (i < NB ? 0xFFFF0000u : 0u);
        bestkey = (acc & m) | (bestkey & ~m);
        bestrow = (f2_imad(i, 0x00010001u, 0u) & m) | (bestrow & ~m);
        if (live) flow_at[t * flow_step] = make_uint2(a0, a1);
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t Nmin  = __reduce_min_sync(0xFFFFFFFFu, live ? min(NA, hasB ? NB : NA) : 0xFFFFFFFFu);
    const uint32_t t_lo  = min(p.lanes, steps);
    const uint32_t t_hi  = max(t_lo, min(Nmin, steps));
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true, true);
    for (; t < t_hi; ++t)  row_step(t, false, false);
    for (; t < steps; ++t) row_step(t, true, false);

    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wkA = 0, wkB = 0, wlaneA = 0, wlaneB = 0;
    for (uint32_t l = 0; l < p.lanes; ++l)
    {
        const uint32_t src = min(gbase + l, 31u);
        const uint32_t ok = __shfl_sync(0xFFFFFFFFu, bestkey, src);
        const uint32_t sA = (ok & 0xFFFFu) >> 4, sB = ok >> 20;
        if (sA >= (wkA >> 4)) { wkA = ok & 0xFFFFu; wlaneA = l; }
        if (sB >= (wkB >> 4)) { wkB = ok >> 16; wlaneB = l; }
    }
    const uint32_t rowsA = __shfl_sync(0xFFFFFFFFu, bestrow, min(gbase + wlaneA, 31u));
    const uint32_t rowsB = __shfl_sync(0xFFFFFFFFu, bestrow, min(gbase + wlaneB, 31u));
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0 || (h ? longB : longA) || M == 0 || N == 0;
        if (p.match > 0 && (uint64_t)p.match * min(M, Nmax) + (uint32_t)B + 15u * (uint32_t)(-I) >= 2047u) dirty = true;
        nvb_alignment a;
        a.score = NVB_SINK_MIN; a.source_x = a.source_y = a.sink_x = a.sink_y = 0xFFFFFFFFu;
        if (dirty) f2_mark_dirty(p.worklist, id);
        else
        {
            const uint32_t wk = h ? wkB : wkA;
            const uint32_t stripe = h ? wlaneB : wlaneA;
            const uint32_t row = h ? (rowsB >> 16) : (rowsA & 0xFFFFu);
            const int32_t score = (int32_t)(wk >> 4) - B;
            if (score <= 0) { a.score = 0; a.sink_x = N; a.sink_y = M; }          // all cells tie at 0: the last one visited
            else { a.score = score; a.sink_x = row + 1u; a.sink_y = stripe * 16u + (wk & 15u) + 1u; }
        }
        q.out[id] = a;
        q.n_ops[id] = 0;
    }
}

// ---- Gotoh LOCAL traceback: the sweep of dpf2_gotoh_local_kernel (sheared along the columns of a stripe, biased, keys compared per
// 8-column stripe of the reference) with the three tagged maxima of dpf2_gotoh_tb_forward_kernel and the floor tagged 3 (SINK).
struct GotohLocal2TbParams
{
    Full2TbParams t;
    uint32_t      FL[16];         // pack2(4 (B + b j) + 3)
    uint32_t      CJ[16];
    uint32_t      unshear;        // fix2(-64 b)
    int32_t       B;
};

__global__ void __launch_bounds__(DF2_THREADS, 3)
dpf2_gotoh_local_tb_forward_kernel(const GotohLocal2TbParams lq)
{
    const Full2TbParams& q = lq.t;
    const Full2Params& p = q.f;
    extern __shared__ uint32_t s_mem[];
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DF2_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks;
    const uint32_t gbase = g * p.lanes;

    uint32_t idA = 0, idB = 0, pbA = 0, pbB = 0, tbA = 0, tbB = 0, MA = 0, MB = 0, NA = 0, NB = 0;
    bool hasB = false, longA = false, longB = false;
    if (live)
    {
        idA = q.first + 2 * task; hasB = idA + 1 < q.first + q.n_launch; idB = hasB ? idA + 1 : idA;
        pbA = p.P.off[idA]; MA = p.P.off[idA + 1] - pbA; pbB = p.P.off[idB]; MB = p.P.off[idB + 1] - pbB;
        tbA = p.T.off[idA]; NA = p.T.off[idA + 1] - tbA; tbB = p.T.off[idB]; NB = p.T.off[idB + 1] - tbB;
        longA = MA > q.maxM || NA > q.maxN; longB = MB > q.maxM || NB > q.maxN;
        if (longA) MA = NA = 0;
        if (longB) MB = NB = 0;
    }
    const uint32_t Nmax = max(NA, NB);
    bool dirtyA = false, dirtyB = false;
    uint16_t* my_sel = (uint16_t*)(s_mem + ((size_t)wib * p.tpw + (live ? g : 0)) * p.sstride);
    const uint32_t pad = p.lanes - 1u;
    if (live) f2_pack_selectors(p.T, my_sel, k, p.lanes, pad, tbA, NA, tbB, NB, Nmax, dirtyA, dirtyB);
    __syncwarp();
    const uint16_t* sel_at = my_sel + pad - k;
    uint4* flow_at = (uint4*)q.flow + f2_flow_base(q, task, k);
    const size_t flow_step = f2_flow_step(q);

    const int32_t go = p.ga, ge = p.gb, b = -ge, dd = go - ge, B = lq.B;
    const int32_t sm = 4 * (p.match - go) + 1, sx = 4 * (p.mismatch - go) + 1;
    const uint32_t base  = ((uint32_t)sx & 0xFFu) * 0x01010101u;
    const uint32_t delta = ((uint32_t)(sx ^ sm)) & 0xFFu;
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], HO[16], F[16];
    const int32_t d1 = 4 * dd + 1;
    const uint32_t dfix1 = d1 < 0 ? f2_fix2(d1) : f2_pack2(d1);
    const uint32_t ge4 = f2_fix2(4 * ge);
    {
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { f2_stripe16(p.P, pbA, MA, block, qa); f2_stripe16(p.P, pbB, MB, block, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, bb = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0x80808080u, wb = 0x80808080u;
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = bb > 3u ? base : base ^ (delta << (8u * bb));
            TA[j] = wa; TB[j] = wb;
        }
    }
    const uint32_t smask = (block < MA ? 0x0000FFFFu : 0u) | (block < MB ? 0xFFFF0000u : 0u);
    uint32_t r_one, r_four, r_m1;
    asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_four) : "r"(4u * p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_m1) : "r"(0u - p.one));
    // column 0 of the matrix in the shear of local column -1: H = 0 -> HO1 = 4 (B - b + d) + 1; E = 0 -> 4 (B - b) + 1
    const uint32_t ho_zero = f2_pack2(4 * (B - b + dd) + 1), e_zero = f2_pack2(4 * (B - b) + 1);
    const uint32_t cin_mul = k == 0 ? 0u : r_one;
    const uint32_t ho_add = k == 0 ? ho_zero : lq.unshear, e_add = k == 0 ? e_zero : lq.unshear;
    uint32_t bestkey = f2_pack2(B) << 4, bestrow = 0u;
    uint32_t carry = ho_zero, send_h = ho_zero, send_e = e_zero;

    auto reset = [&]()
    {
        #pragma unroll
        for (int j = 0; j < 16; ++j) { HO[j] = f2_imad(lq.FL[j] - 0x00030003u, r_one, dfix1); F[j] = 0u; }     // row 0: H = 0, F = -inf
        carry = ho_zero;
    };
    reset();

    auto row_step = [&](const uint32_t t, const bool masked, const bool first)
    {
        const uint32_t hin = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_h, 1), cin_mul, ho_add);
        const uint32_t ein = f2_imad(__shfl_up_sync(0xFFFFFFFFu, send_e, 1), cin_mul, e_add);
        const uint32_t i = t - k;
        if (first && t == k) reset();
        const uint32_t sel = sel_at[t];
        const uint32_t kb = f2_imad(bestkey & 0xFFF8FFF8u, r_one, 0xFFFEFFFFu);
        uint32_t acc = kb, left = hin, e = ein, diag = carry;
        uint32_t fl[4] = { 0u, 0u, 0u, 0u };
        carry = hin;
        #pragma unroll
        for (int j = 0; j < 16; j += 2)
        {
            uint32_t cc[2];
            #pragma unroll
            for (int u = 0; u < 2; ++u)
            {
                const int jj = j + u;
                const uint32_t S = f2_prmt(TA[jj], TB[jj], sel);
                const uint32_t top = HO[jj];
                const uint32_t ft = __vmaxs2(F[jj], top);
                const uint32_t xF = ft & 0x00010001u;
                const uint32_t fn = f2_imad(ft - xF, r_one, ge4);                       // max(F, HO top) + ge, clean
                const uint32_t dg = __viaddmax_s16x2(diag, S, lq.FL[jj]);               // max(diagonal | 2, floor | 3)
                const uint32_t et = __vmaxs2(e, left + 0x00020002u);
                const uint32_t xE = et & 0x00020002u;
                e = et - xE;
                const uint32_t h = __vimax3_s16x2(dg, fn, e);
                const uint32_t th = h & 0x00030003u;
                const uint32_t clean = h - th;
                F[jj] = fn; diag = top;
                left = f2_imad(clean, r_one, dfix1);
                HO[jj] = left;
                cc[u] = clean;
                const uint32_t m = 1u << (4 * (jj & 3));
                fl[jj >> 2] = f2_imad(th, m, f2_imad(xE, 2u * m, f2_imad(xF, 8u * m, fl[jj >> 2])));
            }
            acc = __vimax3_s16x2(acc, f2_imad(cc[0], r_four, lq.CJ[j]), f2_imad(cc[1], r_four, lq.CJ[j + 1]));
        }
        send_h = left; send_e = e;

        const uint32_t df = f2_imad(kb, r_m1, acc);
        const uint32_t sgn = f2_imad(df, r_one, 0x7FFF7FFFu);
        uint32_t m = f2_prmt(sgn, sgn, 0xBB99u) & smask;
        if (masked) m &= (i < NA ? 0x0000FFFFu : 0u) | (i < NB ? 0xFFFF0000u : 0u);
        bestkey = (acc & m) | (bestkey & ~m);
        bestrow = (f2_imad(i, 0x00010001u, 0u) & m) | (bestrow & ~m);
        if (live) flow_at[t * flow_step] = make_uint4(fl[0], fl[1], fl[2], fl[3]);
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    const uint32_t Nmin  = __reduce_min_sync(0xFFFFFFFFu, live ? min(NA, hasB ? NB : NA) : 0xFFFFFFFFu);
    const uint32_t t_lo  = min(p.lanes, steps);
    const uint32_t t_hi  = max(t_lo, min(Nmin, steps));
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true, true);
    for (; t < t_hi; ++t)  row_step(t, false, false);
    for (; t < steps; ++t) row_step(t, true, false);

    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wkA = 0, wkB = 0, wlaneA = 0, wlaneB = 0;
    for (uint32_t l = 0; l < p.lanes; ++l)
    {
        const uint32_t src = min(gbase + l, 31u);
        const uint32_t ok = __shfl_sync(0xFFFFFFFFu, bestkey, src);
        const uint32_t sA = (ok & 0xFFFFu) >> 4, sB = ok >> 20;
        if (sA >= (wkA >> 4)) { wkA = ok & 0xFFFFu; wlaneA = l; }
        if (sB >= (wkB >> 4)) { wkB = ok >> 16; wlaneB = l; }
    }
    const uint32_t rowsA = __shfl_sync(0xFFFFFFFFu, bestrow, min(gbase + wlaneA, 31u));
    const uint32_t rowsB = __shfl_sync(0xFFFFFFFFu, bestrow, min(gbase + wlaneB, 31u));
    if (!live || k != 0) return;

    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        bool dirty = ((h ? dmaskB : dmaskA) & gmask) != 0 || (h ? longB : longA) || M == 0 || N == 0;
        if (p.match > 0 && (uint64_t)p.match * min(max(MA, MB), Nmax) + (uint32_t)B + 15u * (uint32_t)b >= 2047u) dirty = true;
        nvb_alignment a;
        a.score = NVB_SINK_MIN; a.source_x = a.source_y = a.sink_x = a.sink_y = 0xFFFFFFFFu;
        if (dirty) f2_mark_dirty(p.worklist, id);
        else
        {
            const uint32_t wk = h ? wkB : wkA;
            const uint32_t stripe = h ? wlaneB : wlaneA;
            const uint32_t row = h ? (rowsB >> 16) : (rowsA & 0xFFFFu);
            const int32_t score = (int32_t)(wk >> 4) - B;
            if (score <= 0) { a.score = 0; a.sink_x = N; a.sink_y = M; }
            else { a.score = score; a.sink_x = row + 1u; a.sink_y = stripe * 16u + (wk & 15u) + 1u; }
        }
        q.out[id] = a;
        q.n_ops[id] = 0;
    }
}

bool dpf2_traceback_supported(const nvb_aligner& a, uint32_t maxM, uint32_t maxN)
{
    if (maxN == 0 || maxM == 0 || maxN > 65535u) return false;
    const Scheme s = make_scheme(a);
    if (sw_supported(a))
    {
        if (4 * (s.match - s.ga - s.gb) + 2 > 127 || 4 * (s.mismatch - s.ga - s.gb) + 2 > 127) return false;   // profile bytes
    }
    else if (a.type == NVB_LOCAL && (a.kind == NVB_SMITH_WATERMAN || a.kind == NVB_EDIT_DISTANCE))
    {
        if (s.ga >= 0 || s.gb >= 0 || s.ga < -64 || s.gb < -64) return false;
        if (s.match - s.gb < -32 || s.match - s.gb > 31 || s.mismatch - s.gb < -32 || s.mismatch - s.gb > 31) return false;   // bytes 4 (S + b) + 2
    }
    else if (gotoh_supported(a) && a.type != NVB_LOCAL)
    {
        if (4 * (s.match - s.ga - s.gb) + 1 > 127 || 4 * (s.mismatch - s.ga - s.gb) + 1 > 127) return false;
    }
    else if (gotoh_supported(a))
    {
        if (s.match - s.ga > 31 || s.match - s.ga < -32 || s.mismatch - s.ga > 31 || s.mismatch - s.ga < -32) return false;      // bytes 4 (S - go) + 1
    }
    else return false;
    uint32_t lanes, tpw, tstride, cstride; size_t smem;
    return full2_geometry(maxM, maxN, &lanes, &tpw, &tstride, &cstride, &smem);
}

size_t dpf2_tb_bytes_per_couple(const nvb_aligner& a, uint32_t maxM, uint32_t maxN)
{ const size_t lanes = (maxM + 15u) / 16u; return lanes * (maxN + lanes) * (a.kind == NVB_GOTOH ? sizeof(uint4) : sizeof(uint2)); }

// pairs [first, first + n) (first even); flow scratch of at least ceil(n / 2) * dpf2_tb_bytes_per_couple bytes, 16-byte aligned
int dpf2_traceback(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, uint32_t first, uint32_t n, nvb_alignment* out, uint8_t* ops,
                   uint32_t ops_stride, uint32_t* n_ops, uint32_t maxM, uint32_t maxN, uint8_t* flow, uint32_t* worklist, cudaStream_t st,
                   const uint32_t* order)
{
    const Scheme s = make_scheme(a);
    Gotoh2TbParams gq;
    Full2TbParams& q = gq.t;
    Full2Params& p = q.f;
    size_t smem, smem_sel;
    if (!full2_geometry(maxM, maxN, &p.lanes, &p.tpw, &p.tstride, &p.cstride, &smem, &p.sstride, &smem_sel)) return NVB_ERR_UNSUPPORTED;
    p.P = P; p.T = T; p.sinks = nullptr; p.worklist = worklist; p.order = dpf2_wants_order(a) ? order : nullptr;
    p.n_pairs = P.count; p.n_tasks = (n + 1) / 2;
    p.match = s.match; p.mismatch = s.mismatch; p.ga = s.ga; p.gb = s.gb;
    p.semi = a.type == NVB_SEMI_GLOBAL ? 1 : 0;
    p.one = 1;
    p.cstride = maxN + 2u * p.lanes;
    q.out = out; q.n_ops = n_ops; q.ops = ops; q.ops_stride = ops_stride; q.flow = (uint2*)flow;
    q.first = first; q.n_launch = n; q.maxM = maxM; q.maxN = maxN;
    {
        // bit 0 SW / ED GLOBAL + SEMI_GLOBAL, 1 SW / ED LOCAL, 2 Gotoh GLOBAL + SEMI_GLOBAL, 3 Gotoh LOCAL (NVB_TB_LAYOUT: tuning aid)
        static int mask = -1;
        if (mask < 0) { const char* e = getenv("NVB_TB_LAYOUT"); mask = e ? atoi(e) : 1; }
        const int bit = (a.kind == NVB_GOTOH ? 2 : 0) + (a.type == NVB_LOCAL ? 1 : 0);
        q.by_step = (mask >> bit) & 1;
        static int ahead = -1;
        if (ahead < 0) { const char* e = getenv("NVB_TB_AHEAD"); ahead = e ? atoi(e) : 4; if (ahead < 1) ahead = 0x7FFFFFFF; }
        q.ahead = (uint32_t)ahead;
    }
    if (NVB_FIRST_ON_DEVICE(ctx))
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_sw_tb_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_tb_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    }
    const uint32_t n_warps = (p.n_tasks + p.tpw - 1) / p.tpw;
    const uint32_t grid = (n_warps + DF2_WARPS - 1) / DF2_WARPS;
    if (a.type == NVB_LOCAL && a.kind == NVB_GOTOH)
    {
        GotohLocal2TbParams lq;
        lq.t = q; lq.t.f.semi = 2;
        const int32_t b = -s.gb;
        lq.B = DF2_BIAS;
        lq.unshear = f2_fix2(-64 * b);
        for (int j = 0; j < 16; ++j)
        {
            lq.FL[j] = f2_pack2(4 * (lq.B + b * j) + 3);
            const int32_t cj = j - 16 * b * j;
            lq.CJ[j] = cj < 0 ? f2_fix2(cj) : (uint32_t)cj * 0x00010001u;
        }
        if (NVB_FIRST_ON_DEVICE(ctx))
            NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_gotoh_local_tb_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        dpf2_gotoh_local_tb_forward_kernel<<<grid, DF2_THREADS, (size_t)DF2_WARPS * p.tpw * p.sstride * sizeof(uint32_t), st>>>(lq);
        NVB_LAUNCHED(ctx);
        dpf2_gotoh_tb_walk_kernel<<<(n + 255) / 256, 256, 0, st>>>(lq.t);
    }
    else if (a.type == NVB_LOCAL)
    {
        Local2TbParams lq;
        lq.t = q; lq.t.f.semi = 2;                                          // the walk's "LOCAL" mark
        const int32_t b = -s.gb;
        lq.B = 64;
        lq.unshear = f2_fix2(-64 * b);
        for (int j = 0; j < 16; ++j)
        {
            lq.FL[j] = f2_pack2(4 * (lq.B + b * j) + 3);
            const int32_t cj = j - 16 * b * j;
            lq.CJ[j] = cj < 0 ? f2_fix2(cj) : 0u;
        }
        if (NVB_FIRST_ON_DEVICE(ctx))
            NVB_CUDA(ctx, cudaFuncSetAttribute(dpf2_sw_local_tb_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        dpf2_sw_local_tb_forward_kernel<<<grid, DF2_THREADS, (size_t)DF2_WARPS * p.tpw * p.sstride * sizeof(uint32_t), st>>>(lq);
        NVB_LAUNCHED(ctx);
        dpf2_sw_tb_walk_kernel<<<(n + 255) / 256, 256, 0, st>>>(lq.t);
    }
    else if (a.kind == NVB_GOTOH)
    {
        gq.B = 3 * (-(s.ga - s.gb)) + 4;
        dpf2_gotoh_tb_forward_kernel<<<grid, DF2_THREADS, smem_sel, st>>>(gq);
        NVB_LAUNCHED(ctx);
        dpf2_gotoh_tb_walk_kernel<<<(n + 255) / 256, 256, 0, st>>>(q);
    }
    else
    {
        dpf2_sw_tb_forward_kernel<<<grid, DF2_THREADS, smem_sel, st>>>(q);
        NVB_LAUNCHED(ctx);
        dpf2_sw_tb_walk_kernel<<<(n + 255) / 256, 256, 0, st>>>(q);
    }
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
