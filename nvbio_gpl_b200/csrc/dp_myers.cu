// dp_myers.cu -- EditDistanceAligner<TYPE, MyersTag<ALPHABET>>: banded bit-vector edit distance
// (replaces priv::banded_myers<BAND_WIDTH,0,TYPE,ALPHABET>, nvbio/alignment/myers/myers_banded_inl.h:183-262,
// reached from batch_banded_alignment_score<BAND_LEN> as examples/fmmap/fmmap.cu:347-358 does).
//
// One pair per thread, the whole band (<= 32 cells) in ONE register pair of vertical deltas (VP, VN). The
// reference keeps one match vector per symbol and shifts all of them every column; here the pattern window
// is kept as bit PLANES of the symbol codes (2 planes for DNA, +1 "is N" and +1 "loads at all" plane for the
// 5-letter alphabet), so a column costs 2-4 funnel shifts and two LOP3 to form the match vector instead of
// ALPHABET shifts and a select chain:
//     eq = valid & ~((lo ^ TLO) | (hi ^ THI))        TLO/THI = the text symbol's code bits smeared to 32 bits
// which is bit-for-bit MyersBitVectors<A>::get(text[i]) (:96-182): alphabets 2 and 4 mask the codes (c & 1,
// c & 3), alphabet 5 loads nothing for pattern codes above 4 and reads vector 4 for every text code above 3.
// Symbols are fetched as aligned 32-bit words (one LDG per 4 columns and string).
//
// Phase 1 follows the main diagonal while pattern symbols enter at bit BAND-1, phase 2 slides along the last
// pattern row reading the horizontal delta at bit s (:211-234). The running value is the NEGATED distance;
// SEMI_GLOBAL reports every phase-2 column with dist >= min_score to a BestSink ('<=': the last best column
// wins), GLOBAL reports the final column, LOCAL reports nothing. min_score is truncated to int16 exactly like
// the reference's parameter (:189) -- so the Field_traits<int32>::min() the batch drivers pass becomes 0.
#include "dp_common.cuh"

namespace {

// symbol i of an 8-bit or packed (4 / 2 bit, big-endian words) stream through a cached aligned word
struct SymReader
{
    const uint8_t* base; uint32_t bits; uint32_t beg; uint32_t word; uint32_t have;   // have = index of cached word + 1
    __device__ __forceinline__ void init(const uint8_t* b, uint32_t bt, uint32_t bg) { base = b; bits = bt; beg = bg; have = 0; word = 0; }
    __device__ __forceinline__ uint32_t get(uint32_t i)
    {
        const uint32_t a = beg + i;
        const uint32_t lg = bits == 8 ? 2u : bits == 4 ? 3u : 4u;          // log2(symbols per 32-bit word)
        const uint32_t w = a >> lg;
        if (have != w + 1u) { word = __ldg((const uint32_t*)base + w); have = w + 1u; }
        if (bits == 8) return (word >> (8u * (a & 3u))) & 255u;
        if (bits == 4) return (word >> (28u - 4u * (a & 7u))) & 15u;        // big-endian packed streams (packedstream_inl.h:30-52)
        return (word >> (30u - 2u * (a & 15u))) & 3u;
    }
};

template <int ALPHABET>
struct Planes
{
    uint32_t lo, hi, isn, valid;
    __device__ __forceinline__ void init() { lo = hi = isn = 0u; valid = 0u; }
    __device__ __forceinline__ void shift() { lo >>= 1; if (ALPHABET > 2) hi >>= 1; if (ALPHABET == 5) { isn >>= 1; } valid >>= 1; }
    __device__ __forceinline__ void load(uint32_t c, uint32_t bit)
    {
        if (ALPHABET == 5)
        {
            if (c <= 4u) { valid |= bit; if (c == 4u) isn |= bit; else { if (c & 1u) lo |= bit; if (c & 2u) hi |= bit; } }
        }
        else
        {
            valid |= bit;
            if (c & 1u) lo |= bit;
            if (ALPHABET > 2) { if (c & 2u) hi |= bit; }
        }
    }
    __device__ __forceinline__ uint32_t eq(uint32_t c) const
    {
        if (ALPHABET == 5)
        {
            if (c > 3u) return isn;                                  // get(c > 3) = B[4]
            const uint32_t tl = 0u - (c & 1u), th = 0u - ((c >> 1) & 1u);
            return valid & ~isn & ~((lo ^ tl) | (hi ^ th));
        }
        const uint32_t tl = 0u - (c & 1u);
        if (ALPHABET == 2) return valid & ~(lo ^ tl);
        const uint32_t th = 0u - ((c >> 1) & 1u);
        return valid & ~((lo ^ tl) | (hi ^ th));
    }
};

template <int ALPHABET>
__global__ void __launch_bounds__(128)
myers_banded_kernel(const StrSet P, const StrSet T, const int32_t type, const uint32_t band, const int32_t min_score,
                    nvb_best_sink* __restrict__ sinks)
{
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= P.count) return;
    const uint32_t pb = P.off[id], M = P.off[id + 1] - pb;
    const uint32_t tb = T.off[id], N = T.off[id + 1] - tb;
    Best best; best.init();
    if (N >= M)                                                            // :195-196 (else: refused, sink untouched)
    {
        SymReader pr, tr; pr.init(P.sym, P.bits, pb); tr.init(T.sym, T.bits, tb);
        Planes<ALPHABET> B; B.init();
        uint32_t VP = 0xFFFFFFFFu, VN = 0u;
        int32_t dist = 0;
        const uint32_t top = 1u << (band - 1u);
        const uint32_t last = min(N - 1u, M);                              // uint32 arithmetic as the reference (:211)
        for (uint32_t i = 0; i < last; ++i)
        {
            B.shift();
            B.load(pr.get(i), top);
            uint32_t X = B.eq(tr.get(i)) | VN;
            const uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
            const uint32_t HN = VP & D0;
            const uint32_t HP = VN | ~(VP | D0);
            X = D0 >> 1;
            VN = X & HP;
            VP = HN | ~(X | HP);
            dist -= 1 - (int32_t)((D0 >> (band - 1u)) & 1u);
        }
        int32_t s = (int32_t)(band - 1u + M - last);
        for (uint32_t i = last; i < N && s >= 0; ++i, --s)
        {
            B.shift();
            uint32_t X = B.eq(tr.get(i)) | VN;
            const uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
            const uint32_t HN = VP & D0;
            const uint32_t HP = VN | ~(VP | D0);
            X = D0 >> 1;
            VN = X & HP;
            VP = HN | ~(X | HP);
            // s == 32 happens for BAND = 32 and N == M: undefined in the reference's C++ (:179); its host build shifts
            // by s mod 32 (x86), which is the behaviour the parity contract pins
            dist -= (int32_t)((HP >> (s & 31)) & 1u) - (int32_t)((HN >> (s & 31)) & 1u);
            if (type == T_SEMI && dist >= min_score) best.report(dist, i + 1u, M);
        }
        if (type == T_GLOBAL && dist >= min_score) best.report(dist, N, M);
    }
    sinks[id].score = best.score; sinks[id].sink_x = best.x; sinks[id].sink_y = best.y;
}

// four consecutive symbols of an 8-bit stream per call, from aligned 32-bit words joined by a funnel shift (one LDG per call);
// words past the end of the symbol array are not touched
struct Sym4Reader
{
    const uint32_t* w; uint32_t sh, w0, w1, left;      // left = aligned words still readable after w1
    __device__ __forceinline__ void init(const uint8_t* base, uint32_t beg, uint32_t total)
    {
        const uint32_t words = (total + 3u) >> 2, wi = beg >> 2;
        w = (const uint32_t*)base + wi; sh = (beg & 3u) * 8u;
        w0 = wi < words ? __ldg(w) : 0u;
        w1 = wi + 1u < words ? __ldg(w + 1) : 0u;
        left = wi + 2u < words ? words - (wi + 2u) : 0u;
    }
    __device__ __forceinline__ uint32_t next4()
    {
        const uint32_t r = __funnelshift_r(w0, w1, sh);
        w0 = w1; ++w;
        if (left) { w1 = __ldg(w + 1); --left; } else w1 = 0u;
        return r;
    }
};

// the same algorithm over 8-bit string sets, four text columns per loop iteration: symbols arrive as whole words and both phases
// share one column loop (column i is on the diagonal while i < last, then slides along the last pattern row until the band's
// bottom leaves it: i < min(N, M + band), i.e. s = band - 1 + M - i >= 0)
template <int ALPHABET>
__global__ void __launch_bounds__(128)
myers_banded_kernel8(const StrSet P, const StrSet T, const int32_t type, const uint32_t band, const int32_t min_score,
                     nvb_best_sink* __restrict__ sinks)
{
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= P.count) return;
    const uint32_t pb = P.off[id], M = P.off[id + 1] - pb;
    const uint32_t tb = T.off[id], N = T.off[id + 1] - tb;
    Best best; best.init();
    if (N >= M)
    {
        Sym4Reader pr, tr; pr.init(P.sym, pb, P.off[P.count]); tr.init(T.sym, tb, T.off[T.count]);
        Planes<ALPHABET> B; B.init();
        uint32_t VP = 0xFFFFFFFFu, VN = 0u;
        int32_t dist = 0;
        const uint32_t top = 1u << (band - 1u);
        const uint32_t last = min(N - 1u, M);
        const uint32_t ncols = (uint32_t)min((uint64_t)N, (uint64_t)M + band);
        for (uint32_t i0 = 0; i0 < ncols; i0 += 4u)
        {
            const uint32_t p4 = i0 < last ? pr.next4() : 0u;
            const uint32_t t4 = tr.next4();
            #pragma unroll
            for (uint32_t k = 0; k < 4u; ++k)
            {
                const uint32_t i = i0 + k;
                if (i >= ncols) break;
                B.shift();
                if (i < last) B.load((p4 >> (8u * k)) & 255u, top);
                uint32_t X = B.eq((t4 >> (8u * k)) & 255u) | VN;
                const uint32_t D0 = ((VP + (X & VP)) ^ VP) | X;
                const uint32_t HN = VP & D0;
                const uint32_t HP = VN | ~(VP | D0);
                X = D0 >> 1;
                VN = X & HP;
                VP = HN | ~(X | HP);
                if (i < last) dist -= 1 - (int32_t)((D0 >> (band - 1u)) & 1u);
                else
                {
                    const uint32_t s = (band - 1u + M - i) & 31u;          // (s == 32 behaves as on the reference's x86 host build)
                    dist -= (int32_t)((HP >> s) & 1u) - (int32_t)((HN >> s) & 1u);
                    if (type == T_SEMI && dist >= min_score) best.report(dist, i + 1u, M);
                }
            }
        }
        if (type == T_GLOBAL && dist >= min_score) best.report(dist, N, M);
    }
    sinks[id].score = best.score; sinks[id].sink_x = best.x; sinks[id].sink_y = best.y;
}

} // namespace

extern "C" int nvb_batch_banded_myers_score(nvb_ctx* ctx, int32_t type, uint32_t alphabet, uint32_t band, nvb_string_set patterns,
                                            nvb_string_set texts, int32_t min_score, nvb_best_sink* sinks, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (type < NVB_GLOBAL || type > NVB_SEMI_GLOBAL) return NVB_ERR_ARG;
    if (band < 1u || band > 32u) return NVB_ERR_ARG;
    if (alphabet != 2u && alphabet != 4u && alphabet != 5u) return NVB_ERR_UNSUPPORTED;
    if (patterns.count != texts.count) return NVB_ERR_ARG;
    const uint32_t n = patterns.count;
    if (n == 0) return NVB_OK;
    if (!sinks || !patterns.symbols || !patterns.offsets || !texts.symbols || !texts.offsets) return NVB_ERR_ARG;
    if (!(patterns.bits == 8 || patterns.bits == 4 || patterns.bits == 2) || !(texts.bits == 8 || texts.bits == 4 || texts.bits == 2)) return NVB_ERR_ARG;
    if ((patterns.bits == 8 && ((uintptr_t)patterns.symbols & 3u)) || (texts.bits == 8 && ((uintptr_t)texts.symbols & 3u))) return NVB_ERR_ARG;
    const StrSet P = make_set(patterns), T = make_set(texts);
    const int32_t ms = (int32_t)(int16_t)min_score;                        // myers_banded_inl.h:189
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t grid = (n + 127u) / 128u;
    const bool fast8 = patterns.bits == 8 && texts.bits == 8;
    #define NVB_MYERS_GO(A) do { if (fast8) myers_banded_kernel8<A><<<grid, 128, 0, s>>>(P, T, type, band, ms, sinks); \
                                 else myers_banded_kernel<A><<<grid, 128, 0, s>>>(P, T, type, band, ms, sinks); } while (0)
    switch (alphabet)
    {
    case 2:  NVB_MYERS_GO(2); break;
    case 4:  NVB_MYERS_GO(4); break;
    default: NVB_MYERS_GO(5); break;
    }
    #undef NVB_MYERS_GO
    ctx->last_paths = 1;
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
