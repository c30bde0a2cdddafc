// dp_fast_full.cu -- packed s16x2 DPX kernel for full-matrix LOCAL Smith-Waterman / edit distance
// scoring (replaces lmem_batched_alignment_score_kernel / sw_alignment_score_dispatch for the
// reference's speed-test shape: nvbio/alignment/batched_inl.h:79-91, sw/sw_inl.h:419-526,610-749).
//
// Inter-sequence parallelism: one thread scores TWO pairs, pair A in the low and pair B in the high
// half of every 32-bit register. Cells are visited exactly like the reference (pattern stripes of 16
// columns -> text rows -> columns), the stripe lives in 16 registers, the boundary column goes through
// global scratch ([row][thread], coalesced, software-prefetched one row ahead).
//
// Per packed cell (= 2 DP cells):
//   S  = PRMT(profA[j], profB[j], sel_row)     substitution scores of both pairs in one byte permute:
//                                              prof*[j] holds S(r, q_j) for r = A,C,G,T as 4 bytes, the
//                                              selector picks byte r (+ its sign replication) per half
//   u  = VIADD.16x2(top, G)
//   v  = VIADDMNMX.S16x2(diag, S, u)           max(diag + S, top + G)
//   h  = VIADDMNMX.S16x2.RELU(left, I, v)      max(left + I, v, 0)
//   k  = IMAD(h, 16, j)                        arg-max key, issued on the FMA pipe
// BestSink's "last maximal cell wins" (sink_inl.h:44) is reproduced with keys: within a stripe row the
// key (score << 4 | column) is maximised with VIMNMX3.S16x2; across rows and stripes a later row
// replaces the best whenever its score is >= (the column bits of the running best are cleared before
// the row is folded in). Scores must stay below 2048 (match * min(M,N)); pairs beyond that, or holding
// a symbol > 3, are appended to the worklist for the int32 kernel.
#include "dp_fast.h"

#define DPF_THREADS 128

struct FullFastParams
{
    StrSet         P, T;
    nvb_best_sink* sinks;
    uint32_t*      col;        // [maxN][threads] packed boundary column
    uint32_t*      worklist;   // [0] = count, [1..] = pair ids
    uint32_t       n_pairs;
    uint32_t       n_threads;  // total threads = ceil(n_pairs / 2)
    uint32_t       tstride;    // per-thread text words in shared memory = ceil(maxN / 8)
    int32_t        match, mismatch, G, I;
    uint32_t       sixteen;    // = 16, a runtime value so that the key is an IMAD
};

__device__ __forceinline__ uint32_t pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }

__device__ __forceinline__ void mark_dirty(uint32_t* wl, uint32_t id)
{
    const uint32_t slot = atomicAdd(wl, 1u);
    wl[1 + slot] = id;
}

__global__ void __launch_bounds__(DPF_THREADS)
dpf_full_sw_local_kernel(const FullFastParams p)
{
    extern __shared__ uint32_t s_text[];   // [tstride][DPF_THREADS]: 8 rows per word, nibble = rA | rB << 2
    const uint32_t tid = blockIdx.x * DPF_THREADS + threadIdx.x;
    if (tid >= p.n_threads) return;        // no block-wide synchronisation below
    const uint32_t idA = 2 * tid;
    const bool     hasB = idA + 1 < p.n_pairs;
    const uint32_t idB = hasB ? idA + 1 : idA;

    const uint32_t pbA = p.P.off[idA], MA = p.P.off[idA + 1] - pbA;
    const uint32_t pbB = p.P.off[idB], MB = p.P.off[idB + 1] - pbB;
    const uint32_t tbA = p.T.off[idA], NA = p.T.off[idA + 1] - tbA;
    const uint32_t tbB = p.T.off[idB], NB = p.T.off[idB + 1] - tbB;
    const uint32_t Nmax = max(NA, NB), Mmax = max(MA, MB);
    bool dirtyA = false, dirtyB = false;

    // scores must fit the 11 score bits of the arg-max key
    if (p.match > 0)
    {
        if ((uint64_t)p.match * min(MA, NA) >= 2048u) dirtyA = true;
        if ((uint64_t)p.match * min(MB, NB) >= 2048u) dirtyB = true;
    }

    // ---- prologue: text -> nibble stream in shared memory
    uint32_t* my_text = s_text + threadIdx.x;
    for (uint32_t w = 0; w * 8 < Nmax; ++w)
    {
        uint32_t word = 0;
        #pragma unroll
        for (uint32_t k = 0; k < 8; ++k)
        {
            const uint32_t i = w * 8 + k;
            const uint32_t rA = i < NA ? sym_at_rt(p.T.sym, p.T.bits, tbA + i) : 0u;
            const uint32_t rB = i < NB ? sym_at_rt(p.T.sym, p.T.bits, tbB + i) : 0u;
            dirtyA |= rA > 3u; dirtyB |= rB > 3u;
            word |= ((rA & 3u) | ((rB & 3u) << 2)) << (4 * k);
        }
        my_text[w * DPF_THREADS] = word;
    }

    const uint32_t G2 = pack2(p.G), I2 = pack2(p.I);
    const uint32_t base  = ((uint32_t)p.mismatch & 0xFFu) * 0x01010101u;     // 4 x mismatch
    const uint32_t delta = ((uint32_t)(p.mismatch ^ p.match)) & 0xFFu;       // flips one byte to match
    const uint32_t nstripes = (Mmax + 15u) / 16u;
    uint32_t* colp = p.col + tid;
    const size_t  cstride = p.n_threads;

    uint32_t bestkey = 0;               // packed (score << 4 | column-in-stripe)
    uint32_t posA = 0, posB = 0;        // stripe << 16 | row of the best

    for (uint32_t s = 0; s < nstripes; ++s)
    {
        const uint32_t block = s * 16u;
        uint32_t TA[16], TB[16], H[16];
        #pragma unroll
        for (int j = 0; j < 16; ++j)
        {
            const uint32_t c = block + j;
            uint32_t wa = 0x80808080u, wb = 0x80808080u;       // columns past the pattern: S = -128
            if (c < MA) { const uint32_t q = sym_at_rt(p.P.sym, p.P.bits, pbA + c); dirtyA |= q > 3u; wa = base ^ (delta << (8u * (q & 3u))); }
            if (c < MB) { const uint32_t q = sym_at_rt(p.P.sym, p.P.bits, pbB + c); dirtyB |= q > 3u; wb = base ^ (delta << (8u * (q & 3u))); }
            TA[j] = wa; TB[j] = wb; H[j] = 0;
        }
        const bool svA = block < MA, svB = block < MB;
        const bool rd = s > 0, wr = s + 1 < nstripes;
        uint32_t carry = 0;
        uint32_t cin_next = rd && Nmax ? colp[0] : 0u;
        uint32_t tw = 0;

        for (uint32_t i = 0; i < Nmax; ++i)
        {
            const uint32_t cin = cin_next;
            if (rd && i + 1 < Nmax) cin_next = colp[(size_t)(i + 1) * cstride];
            if ((i & 7u) == 0) tw = my_text[(i >> 3) * DPF_THREADS];
            const uint32_t nib = tw & 15u; tw >>= 4;
            // selector: byte rA of profA -> byte 0 (+ sign replica in byte 1), byte rB of profB -> bytes 2,3
            const uint32_t sel = ((nib & 3u) | ((nib & 12u) << 6)) * 0x11u + 0xC480u;

            uint32_t diag = carry, left = cin;
            carry = cin;
            const uint32_t kb = bestkey & 0xFFF0FFF0u;
            uint32_t acc = kb;
            uint32_t key[16];
            #pragma unroll
            for (int j = 0; j < 16; ++j)
            {
                const uint32_t S = __byte_perm(TA[j], TB[j], sel);
                const uint32_t u = __vadd2(H[j], G2);
                const uint32_t v = __viaddmax_s16x2(diag, S, u);
                const uint32_t h = __viaddmax_s16x2_relu(left, I2, v);
                diag = H[j]; H[j] = h; left = h;
                asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(key[j]) : "r"(h), "r"(p.sixteen), "r"((uint32_t)j * 0x00010001u));
            }
            #pragma unroll
            for (int j = 0; j < 16; j += 2) acc = __vimax3_s16x2(acc, key[j], key[j + 1]);
            if (wr) colp[(size_t)i * cstride] = left;

            // a later row wins whenever its score is >= the best so far
            bool keep_hi, keep_lo;
            (void)__vibmax_s16x2(kb, acc, &keep_hi, &keep_lo);      // keep = kb >= acc  (nothing >= in this row)
            const bool updA = !keep_lo && svA && i < NA;
            const bool updB = !keep_hi && svB && i < NB;
            const uint32_t rowpos = (s << 16) | i;
            if (updA) { bestkey = (bestkey & 0xFFFF0000u) | (acc & 0x0000FFFFu); posA = rowpos; }
            if (updB) { bestkey = (bestkey & 0x0000FFFFu) | (acc & 0xFFFF0000u); posB = rowpos; }
        }
    }

    // ---- epilogue
    #pragma unroll
    for (int h = 0; h < 2; ++h)
    {
        if (h == 1 && !hasB) break;
        const uint32_t id = h ? idB : idA;
        const bool dirty = h ? dirtyB : dirtyA;
        if (dirty) { mark_dirty(p.worklist, id); continue; }
        const uint32_t M = h ? MB : MA, N = h ? NB : NA;
        const uint32_t k = h ? (bestkey >> 16) : (bestkey & 0xFFFFu);
        const uint32_t pos = h ? posB : posA;
        nvb_best_sink out;
        if (M == 0 || N == 0) { out.score = NVB_SINK_MIN; out.sink_x = out.sink_y = 0xFFFFFFFFu; }
        else if ((k >> 4) == 0) { out.score = 0; out.sink_x = N; out.sink_y = M; }   // all cells tie at 0: the last one visited
        else { out.score = (int32_t)(k >> 4); out.sink_x = (pos & 0xFFFFu) + 1u; out.sink_y = (pos >> 16) * 16u + (k & 15u) + 1u; }
        p.sinks[id] = out;
    }
}

bool dpf_full_score_supported(const nvb_aligner& a, const StrSet& P, const StrSet& T, uint32_t maxM, uint32_t maxN)
{
    (void)P; (void)T;
    if (a.type != NVB_LOCAL) return false;
    if (a.kind != NVB_SMITH_WATERMAN && a.kind != NVB_EDIT_DISTANCE) return false;
    const Scheme s = make_scheme(a);
    if (s.match < -127 || s.match > 127 || s.mismatch < -127 || s.mismatch > 127) return false;
    if (s.ga >= 0 || s.gb >= 0 || s.ga < -1024 || s.gb < -1024) return false;
    if (maxN == 0 || maxM == 0 || maxN > 2048 || maxM > 65535) return false;
    return true;
}

int dpf_full_score(nvb_ctx* ctx, const nvb_aligner& a, StrSet P, StrSet T, nvb_best_sink* sinks, uint32_t maxM,
                   uint32_t maxN, uint32_t* worklist, void* temp, cudaStream_t st)
{
    (void)maxM;
    const Scheme s = make_scheme(a);
    FullFastParams p;
    p.P = P; p.T = T; p.sinks = sinks; p.col = (uint32_t*)temp; p.worklist = worklist;
    p.n_pairs = P.count; p.n_threads = (P.count + 1) / 2;
    p.tstride = (maxN + 7) / 8;
    p.match = s.match; p.mismatch = s.mismatch; p.G = s.ga; p.I = s.gb; p.sixteen = 16;
    const size_t smem = (size_t)p.tstride * DPF_THREADS * sizeof(uint32_t);
    static bool attr_set = false;
    if (!attr_set)
    {
        NVB_CUDA(ctx, cudaFuncSetAttribute(dpf_full_sw_local_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    const uint32_t grid = (p.n_threads + DPF_THREADS - 1) / DPF_THREADS;
    dpf_full_sw_local_kernel<<<grid, DPF_THREADS, smem, st>>>(p);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
