// dp_fast_full.cu -- packed s16x2 DPX wavefront kernel for full-matrix LOCAL Smith-Waterman / edit
// distance scoring (replaces lmem_batched_alignment_score_kernel + sw_alignment_score_dispatch and the
// warp-per-alignment kernel of the reference for its speed-test shape:
// nvbio/alignment/batched_inl.h:79-91,183-205, sw/sw_inl.h:419-526,610-749, sw/sw_warp_inl.h).
//
// Parallelism, both kinds at once:
//   inter-sequence : every 32-bit register holds TWO pairs (pair A in the low, pair B in the high half);
//   intra-sequence : the pattern stripes of 16 columns (the reference's BAND_LEN) of one such couple are
//                    spread over adjacent lanes; lane k sweeps stripe k down the text one row per step,
//                    one row behind lane k-1, and receives the stripe boundary cell by warp shuffle
//                    (anti-diagonal wavefront). A warp carries floor(32 / stripes) couples.
// Nothing but the packed text (4 bits per row and couple, shared memory) is read inside the sweep.
//
// Per packed cell (= 2 DP cells), scores biased by B = max(|G|,|I|) so that the two gap additions
// can be plain 32-bit IMADs on the FMA pipe (the low half never borrows):
//   S   = PRMT(profA[j], profB[j], sel_row)   both substitution scores by one byte permute: prof*[j] holds
//                                             S(r, q_j) for r = A,C,G,T as 4 bytes; the row selector picks
//                                             byte r and replicates its sign into the upper byte
//   u   = IMAD(top, 1, G')                    top + G
//   v   = VIADDMNMX.S16x2(diag, S, u)         max(diag + S, top + G)
//   l   = IMAD(left, 1, I')                   left + I
//   h   = VIMNMX3.S16x2(l, v, B)              max(left + I, v, 0)          (LOCAL floor)
//   key = IMAD(h, 16, j)                      arg-max key
// BestSink's "last maximal cell wins" (sink_inl.h:44) with the reference's visiting order
// (stripe, text row, column) is reproduced by keys: inside a stripe row (score << 4 | column) is
// maximised with VIMNMX3.S16x2; a later row replaces the running best whenever its score is >= ; across
// stripes the higher stripe wins ties (final shuffle reduction). Pairs with a symbol > 3, or whose
// score could reach 2048 - B, go to the worklist for the int32 kernel.
#include "dp_fast.h"

#define DPF_WARPS   4
#define DPF_THREADS (DPF_WARPS * 32)

struct FullFastParams
{
    StrSet         P, T;
    nvb_best_sink* sinks;
    uint32_t*      worklist;   // [0] = count, [1..] = pair ids
    uint32_t       n_pairs;
    uint32_t       n_tasks;    // couples = ceil(n_pairs / 2)
    uint32_t       lanes;      // lanes (= stripes) per couple
    uint32_t       tpw;        // couples per warp = 32 / lanes
    uint32_t       tstride;    // shared-memory words per couple = ceil((maxN + 2 lanes) / 2): one 16-bit selector per row + padding
    int32_t        match, mismatch, G, I, B;
    uint32_t       one, sixteen;   // runtime 1 and 16: keep the IMADs IMADs
    uint32_t       unshear;        // fix2(-16 b): the value handed to the next lane leaves the stripe's shear
    uint32_t       FL[16];         // LOCAL floor of stripe column j in sheared coordinates: pack2(B + b j)
    uint32_t       CJ[16];         // key addend: (h' << 4) + CJ[j] = (H + B) << 4 | j
};

__device__ __forceinline__ uint32_t pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }
// addend that adds the negative g to both halves of a register whose low half never drops below -g
__device__ __forceinline__ uint32_t fix2(int32_t g) { return (((uint32_t)(g - 1) & 0xFFFFu) << 16) | ((uint32_t)g & 0xFFFFu); }
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }

// 4 bytes holding 2-bit symbols -> 4 nibbles in the low half
__device__ __forceinline__ uint32_t squeeze4(uint32_t x)
{
    x &= 0x03030303u; x = (x | (x >> 4)) & 0x00FF00FFu; x = (x | (x >> 8)) & 0x0000FFFFu;
    return x;
}

__device__ __forceinline__ void mark_dirty(uint32_t* wl, uint32_t id)
{
    const uint32_t slot = atomicAdd(wl, 1u);
    wl[1 + slot] = id;
}

// BT > 0: b = -I = BT and the bias B = DPF_BIAS are compile-time, so the per-column floors FL[j] and key addends CJ[j]
// are immediates; BT = 0 reads them from the parameter block (any I in -64..-1).
#define DPF_BIAS 64
template <int BT>
__global__ void __launch_bounds__(DPF_THREADS, 5)
dpf_full_sw_local_kernel(const FullFastParams p)
{
    auto FLc = [&](const int j) -> uint32_t { return BT ? (uint32_t)(DPF_BIAS + BT * j) * 0x00010001u : p.FL[j]; };
    auto CJc = [&](const int j) -> uint32_t
    {
        if (!BT) return p.CJ[j];
        const int32_t cj = j - 16 * BT * j;
        return cj < 0 ? ((((uint32_t)(cj - 1) & 0xFFFFu) << 16) | ((uint32_t)cj & 0xFFFFu)) : 0u;
    };
    extern __shared__ uint32_t s_text[];   // [warp][couple slot][tstride]: the couple's selector stream
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * DPF_WARPS + wib;
    const uint32_t g = lane / p.lanes, k = lane - g * p.lanes;       // couple slot, stripe
    const uint32_t task = warp * p.tpw + g;
    const bool live = g < p.tpw && task < p.n_tasks;
    const uint32_t gbase = g * p.lanes;                               // first lane of this couple

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

    // ---- prologue: the couple's lanes turn its two texts into a stream of ready-made PRMT selectors, one 16-bit entry per row:
    // entry pad + r selects byte rA of the first profile word into the low half and byte rB of the second into the high half,
    // sign-extended (((rA | rB << 8) * 0x11) + 0xC480). pad entries before row 0 and pad + 1 after the last row hold a valid selector,
    // so the ramp-up and drain steps of the wavefront read the stream unconditionally: one LDS.U16 per row instead of the nibble
    // extraction and selector arithmetic of a 4-bit stream (5 ALU-pipe instructions per row)
    uint16_t* my_sel = (uint16_t*)(s_text + ((size_t)wib * p.tpw + (live ? g : 0)) * p.tstride);
    const uint32_t pad = p.lanes - 1u;
    if (live)
    {
        for (uint32_t e = k; e < pad; e += p.lanes) my_sel[e] = 0xC480u;
        for (uint32_t e = k; e < pad + 1u; e += p.lanes) my_sel[pad + Nmax + e] = 0xC480u;
        for (uint32_t w = k; w * 8 < Nmax; w += p.lanes)
        {
            const uint32_t i0 = w * 8;
            // 8 rows of one text as 8 bytes; whole aligned words of an 8-bit set are fetched with two 32-bit loads
            auto rows8 = [&](const uint32_t tb, const uint32_t N, bool& dirty, uint32_t& lo, uint32_t& hi)
            {
                const uint8_t* src = p.T.sym + tb + i0;
                if (p.T.bits == 8 && i0 + 8 <= N && ((uintptr_t)src & 3u) == 0)
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
                    const uint32_t r = i0 + e < N ? sym_at_rt(p.T.sym, p.T.bits, tb + i0 + e) : 0u;
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
    const uint16_t* sel_at = my_sel + pad - k;               // step t reads row t - k
    __syncwarp();

    // ---- this lane's stripe: substitution profiles of 16 pattern columns of both pairs
    const uint32_t B2 = pack2(p.B), Gf = fix2(p.G), If = fix2(p.I);
    const uint32_t base  = ((uint32_t)(p.mismatch - p.I) & 0xFFu) * 0x01010101u;            // 4 x (mismatch + b)
    const uint32_t delta = ((uint32_t)((p.mismatch - p.I) ^ (p.match - p.I))) & 0xFFu;       // flips one byte to match + b
    const uint32_t block = k * 16u;
    uint32_t TA[16], TB[16], H[16];
    {
        // the stripe's 16 pattern symbols of one pair: four aligned 32-bit loads when the stripe is whole, else bytes
        auto stripe16 = [&](const uint32_t pb, const uint32_t M, bool& dirty, uint32_t* q4)
        {
            const uint8_t* src = p.P.sym + pb + block;
            if (p.P.bits == 8 && block + 16u <= M && ((uintptr_t)src & 3u) == 0)
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
                        x |= (c < M ? sym_at_rt(p.P.sym, p.P.bits, pb + c) : 0xFFu) << (8 * e);   // 0xFF = past the pattern
                    }
                    q4[w] = x;
                }
            }
        };
        uint32_t qa[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu }, qb[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu };
        if (live) { stripe16(pbA, MA, dirtyA, qa); stripe16(pbB, MB, dirtyB, qb); }
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            const uint32_t a = (qa[j >> 2] >> (8 * (j & 3))) & 255u, b = (qb[j >> 2] >> (8 * (j & 3))) & 255u;
            uint32_t wa = 0x80808080u, wb = 0x80808080u;       // columns past the pattern: S = -128
            if (live && c < MA) wa = a > 3u ? base : base ^ (delta << (8u * a));                 // a pattern N matches no text symbol 0..3: mismatch in every row
            if (live && c < MB) wb = b > 3u ? base : base ^ (delta << (8u * b));
            TA[j] = wa; TB[j] = wb; H[j] = FLc(j);
        }
    }
    // halves that own cells in this stripe
    const uint32_t smask = (block < MA ? 0x0000FFFFu : 0u) | (block < MB ? 0xFFFF0000u : 0u);

    uint32_t bestkey = B2 << 4;          // packed (biased score << 4 | column in stripe)
    uint32_t bestrow = 0;                // packed text rows of the best
    // Rows are kept SHEARED along the stripe: stored h'[j] = H + B + b*j (b = -I), so the move from the left costs
    // nothing (h'[j-1] already equals H[j-1] + I + B + b*j), the diagonal's +b sits in the profile bytes and the LOCAL
    // floor of column j is the constant FL[j] = B + b*j. Only the value handed to the next lane is un-sheared (-16 b).
    uint32_t r_one, r_sixteen;          // in registers: an instruction takes one constant-bank operand, FL[] / CJ[] / Gf use it
    asm volatile("mov.u32 %0, %1;" : "=r"(r_one) : "r"(p.one));
    asm volatile("mov.u32 %0, %1;" : "=r"(r_sixteen) : "r"(p.sixteen));
    const uint32_t zero_in = pack2(p.B + p.I);                  // H = 0 of the column left of the stripe (local j = -1)
    uint32_t carry = zero_in, send = FLc(15);

    // one text row of the stripe. MASKED = false: every lane of the warp is on a row that exists in both of its pairs.
    // lane 0 of a couple has no left neighbour: column 0 of a LOCAL matrix is all zeros. Folded into the un-shear IMAD.
    const uint32_t cin_mul = k == 0 ? 0u : r_one, cin_add = k == 0 ? zero_in : p.unshear;
    uint32_t r_m1; asm volatile("mov.u32 %0, %1;" : "=r"(r_m1) : "r"(0u - p.one));

    // one text row of the stripe. masked = false: every lane of the warp is on a row that exists in both of its pairs;
    // first = true: some lane may still be on the dummy rows above its first row (they leave garbage that is reset here)
    auto row_step = [&](const uint32_t t, const bool masked, const bool first)
    {
        const uint32_t cin = imad(__shfl_up_sync(0xFFFFFFFFu, send, 1), cin_mul, cin_add);
        const uint32_t i = t - k;                               // this lane's text row (wraps while t < k: a dummy row)
        if (first && t == k)
        {
            #pragma unroll
            for (int j = 0; j < 16; ++j) H[j] = FLc(j);
            carry = zero_in;
        }
        // selector: byte rA of profA -> byte 0, its sign -> byte 1; byte rB of profB -> byte 2, sign -> byte 3
        const uint32_t sel = sel_at[t];

        // phase A: everything that only needs the previous row
        uint32_t v[16];
        {
            uint32_t diag = carry;
            #pragma unroll
            for (int j = 0; j < 16; ++j)
            {
                const uint32_t S = prmt(TA[j], TB[j], sel);
                const uint32_t u = imad(H[j], r_one, Gf);
                v[j] = __viaddmax_s16x2(diag, S, u);
                diag = H[j];
            }
        }
        carry = cin;
        // phase B: the serial chain along the row, keys reduced pairwise on the fly.
        // kb = running best with its column bits cleared, minus one: a cell of equal score compares greater. Both halves
        // of every operand below are non-negative and the differences too, so plain 32-bit IMADs (FMA pipe) do the
        // packed arithmetic without borrows.
        const uint32_t kb = imad(bestkey & 0xFFF0FFF0u, r_one, 0xFFFEFFFFu);          // - 0x00010001
        uint32_t acc = kb, left = cin;
        #pragma unroll
        for (int j = 0; j < 16; j += 2)
        {
            const uint32_t h0 = __vimax3_s16x2(left, v[j], FLc(j));
            const uint32_t h1 = __vimax3_s16x2(h0, v[j + 1], FLc(j + 1));
            H[j] = h0; H[j + 1] = h1; left = h1;
            acc = __vimax3_s16x2(acc, imad(h0, r_sixteen, CJc(j)), imad(h1, r_sixteen, CJc(j + 1)));
        }
        send = left;

        // halves whose row holds a score >= the best so far take the row (later rows win ties): acc - kb > 0 there
        const uint32_t d = imad(kb, r_m1, acc);                        // acc - kb, >= 0 in both halves
        const uint32_t sgn = imad(d, r_one, 0x7FFF7FFFu);              // bit 15 of a half set iff its difference is > 0
        uint32_t m = prmt(sgn, sgn, 0xBB99u) & smask;                  // 0xFFFF per improved half that owns cells here
        if (masked) m &= (i < NA ? 0x0000FFFFu : 0u) | (i < NB ? 0xFFFF0000u : 0u);
        bestkey = (acc & m) | (bestkey & ~m);
        bestrow = (imad(i, 0x00010001u, 0u) & m) | (bestrow & ~m);
    };

    const uint32_t steps = __reduce_max_sync(0xFFFFFFFFu, live ? Nmax + p.lanes - 1 : 0u);
    // rows [t_lo, t_hi): all lanes on existing rows of both pairs (dead lanes never report: their smask is 0)
    const uint32_t Nmin  = __reduce_min_sync(0xFFFFFFFFu, live ? min(NA, hasB ? NB : NA) : 0xFFFFFFFFu);
    const uint32_t t_lo  = min(p.lanes, steps);
    const uint32_t t_hi  = max(t_lo, min(Nmin, steps));
    uint32_t t = 0;
    for (; t < t_lo; ++t)  row_step(t, true, true);
    for (; t < t_hi; ++t)  row_step(t, false, false);
    for (; t < steps; ++t) row_step(t, true, false);

    // ---- epilogue: the highest stripe among those with the best score wins (stripes are visited in order)
    const uint32_t dmaskA = __ballot_sync(0xFFFFFFFFu, dirtyA), dmaskB = __ballot_sync(0xFFFFFFFFu, dirtyB);
    const uint32_t gmask = (p.lanes >= 32 ? 0xFFFFFFFFu : ((1u << p.lanes) - 1u)) << gbase;
    uint32_t wkA = 0, wkB = 0, wlaneA = 0, wlaneB = 0;
    for (uint32_t d = 0; d < p.lanes; ++d)
    {
        const uint32_t src = min(gbase + d, 31u);
        const uint32_t ok = __shfl_sync(0xFFFFFFFFu, bestkey, src);
        const uint32_t sA = (ok & 0xFFFFu) >> 4, sB = ok >> 20;
        if (sA >= (wkA >> 4)) { wkA = ok & 0xFFFFu; wlaneA = d; }
        if (sB >= (wkB >> 4)) { wkB = ok >> 16; wlaneB = d; }
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
        // key overflow; the other pair's rows are swept too, hence the couple's longer text in the bound
        if (p.match > 0 && (uint64_t)p.match * min(M, Nmax) + (uint32_t)p.B + 15u * (uint32_t)(-p.I) >= 2047u) dirty = true;
        if (dirty) { mark_dirty(p.worklist, id); continue; }
        const uint32_t wk = h ? wkB : wkA;
        const uint32_t stripe = h ? wlaneB : wlaneA;
        const uint32_t row = h ? (rowsB >> 16) : (rowsA & 0xFFFFu);
        const int32_t score = (int32_t)(wk >> 4) - p.B;
        nvb_best_sink out;
        if (M == 0 || N == 0) { out.score = NVB_SINK_MIN; out.sink_x = out.sink_y = 0xFFFFFFFFu; }
        else if (score <= 0) { out.score = 0; out.sink_x = N; out.sink_y = M; }   // all cells tie at 0: the last one visited
        else { out.score = score; out.sink_x = row + 1u; out.sink_y = stripe * 16u + (wk & 15u) + 1u; }
        p.sinks[id] = out;
    }
}

static bool full_geometry(uint32_t maxM, uint32_t maxN, uint32_t* lanes, uint32_t* tpw, uint32_t* tstride, size_t* smem)
{
    const uint32_t stripes = (maxM + 15u) / 16u;
    if (stripes == 0 || stripes > 32u) return false;
    *lanes = stripes; *tpw = 32u / stripes; *tstride = (maxN + 2u * stripes + 1u) / 2u;
    *smem = (size_t)DPF_WARPS * *tpw * *tstride * sizeof(uint32_t);
    return *smem <= 160u * 1024u;
}

bool dpf_full_score_supported(const nvb_aligner& a, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN)
{
    (void)P; (void)T;
    if (a.type != NVB_LOCAL) return false;
    if (a.kind != NVB_SMITH_WATERMAN && a.kind != NVB_EDIT_DISTANCE) return false;
    const Scheme s = make_scheme(a);
    if (s.match - s.gb < -127 || s.match - s.gb > 127 || s.mismatch - s.gb < -127 || s.mismatch - s.gb > 127) return false;   // profile bytes hold S + b
    if (s.ga >= 0 || s.gb >= 0 || s.ga < -64 || s.gb < -64) return false;
    if (maxN == 0 || maxM == 0 || maxN > 65535u) return false;
    uint32_t lanes, tpw, tstride; size_t smem;
    return full_geometry(maxM, maxN, &lanes, &tpw, &tstride, &smem);
}

int dpf_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, nvb_best_sink* sinks, uint32_t maxM,
                   uint32_t maxN, uint32_t* worklist, void* temp, cudaStream_t st)
{
    (void)temp;
    const Scheme s = make_scheme(a);
    FullFastParams p;
    size_t smem;
    if (!full_geometry(maxM, maxN, &p.lanes, &p.tpw, &p.tstride, &smem)) return NVB_ERR_UNSUPPORTED;
    p.P = P; p.T = T; p.sinks = sinks; p.worklist = worklist;
    p.n_pairs = P.count; p.n_tasks = (P.count + 1) / 2;
    p.match = s.match; p.mismatch = s.mismatch; p.G = s.ga; p.I = s.gb;
    const int bt = (-s.gb >= 1 && -s.gb <= 4) ? -s.gb : 0;      // instantiated immediates
    p.B = bt ? DPF_BIAS : -(s.ga < s.gb ? s.ga : s.gb);
    p.one = 1; p.sixteen = 16;
    {
        const int32_t b = -s.gb;
        p.unshear = (((uint32_t)(-16 * b - 1) & 0xFFFFu) << 16) | ((uint32_t)(-16 * b) & 0xFFFFu);
        for (int j = 0; j < 16; ++j)
        {
            p.FL[j] = ((uint32_t)(p.B + b * j) & 0xFFFFu) * 0x00010001u;
            const int32_t cj = j - 16 * b * j;                 // <= 0; a negative addend borrows one from the upper half
            p.CJ[j] = cj < 0 ? ((((uint32_t)(cj - 1) & 0xFFFFu) << 16) | ((uint32_t)cj & 0xFFFFu)) : 0u;
        }
    }
    if (NVB_FIRST_ON_DEVICE(ctx))
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_full_sw_local_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_full_sw_local_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_full_sw_local_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_full_sw_local_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_full_sw_local_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    }
    const uint32_t n_warps = (p.n_tasks + p.tpw - 1) / p.tpw;
    const uint32_t grid = (n_warps + DPF_WARPS - 1) / DPF_WARPS;
    switch (bt)
    {
    case 1:  dpf_full_sw_local_kernel<1><<<grid, DPF_THREADS, smem, st>>>(p); break;
    case 2:  dpf_full_sw_local_kernel<2><<<grid, DPF_THREADS, smem, st>>>(p); break;
    case 3:  dpf_full_sw_local_kernel<3><<<grid, DPF_THREADS, smem, st>>>(p); break;
    case 4:  dpf_full_sw_local_kernel<4><<<grid, DPF_THREADS, smem, st>>>(p); break;
    default: dpf_full_sw_local_kernel<0><<<grid, DPF_THREADS, smem, st>>>(p); break;
    }
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
