// dp_common.cuh -- shared pieces of the DP (alignment) kernels.
#pragma once
#include "common.cuh"

enum { T_GLOBAL = NVB_GLOBAL, T_LOCAL = NVB_LOCAL, T_SEMI = NVB_SEMI_GLOBAL };
enum { OP_SUB = 0, OP_INS = 1, OP_DEL = 2, OP_SINK = 3, OP_INS_EXT = 4, OP_DEL_EXT = 8 };   // alignment.h:326-337
enum { ST_H = 0, ST_E = 1, ST_F = 2 };                                                      // alignment.h:341-346

// SW/ED: ga = deletion (consumes text), gb = insertion (consumes pattern); Gotoh: ga = open, gb = ext
// Quality-aware schemes (alignment.h:411-436, 488-505): qm/qx = match(q) / mismatch(a, b, q) tables in device memory (NULL: the
// constants), quals = one byte per pattern symbol addressed by the patterns' offsets (NULL: all 0); tga/tgb = Gotoh's
// text_gap_open/extension, which the reference uses only for GLOBAL first columns / band rows and the banded infimum.
struct Scheme { int32_t match, mismatch, ga, gb; int32_t tga, tgb; const int32_t* qm; const int32_t* qx; const uint8_t* quals; };
struct QualDev { const int32_t* qm; const int32_t* qx; const uint8_t* quals; int32_t tga, tgb; };

static inline Scheme make_scheme(const nvb_aligner& a);
static inline Scheme make_scheme_q(const nvb_aligner& a, const QualDev* q)
{
    Scheme s = make_scheme(a);
    if (q && a.kind != NVB_EDIT_DISTANCE) { s.qm = q->qm; s.qx = q->qx; s.quals = q->quals; if (a.kind == NVB_GOTOH) { s.tga = q->tga; s.tgb = q->tgb; } }
    return s;
}
static inline Scheme make_scheme(const nvb_aligner& a)
{
    Scheme s;
    if (a.kind == NVB_EDIT_DISTANCE) { s.match = 0; s.mismatch = -1; s.ga = -1; s.gb = -1; }   // ed_utils.h:36-43
    else { s.match = a.match; s.mismatch = a.mismatch; s.ga = a.gap_a; s.gb = a.gap_b; }
    s.tga = s.ga; s.tgb = s.gb; s.qm = s.qx = nullptr; s.quals = nullptr;
    return s;
}

// optional Best2Sink output of the int32 scoring kernels (out == NULL: plain BestSink into `sinks`)
struct Sink2 { nvb_best2_sink* out; uint32_t distinct; };

#define NVB_SINK_MIN (-(1 << 30))   // Field_traits<int32>::min(), nvbio/basic/numbers.h:738-741

#ifdef __CUDACC__
// BestSink, and with `two` set Best2Sink (sink_inl.h:54-82): the runner-up must end more than `dd` text positions
// away from the current best (uint32 arithmetic, as the reference)
struct Best
{
    int32_t score; uint32_t x, y;
    int32_t score2; uint32_t x2, y2, dd; bool two;
    __device__ __forceinline__ void init() { score = score2 = NVB_SINK_MIN; x = y = x2 = y2 = 0xFFFFFFFFu; dd = 0; two = false; }
    // BestSink::report uses '<=': the last visited maximal cell wins (sink_inl.h:40-48)
    __device__ __forceinline__ void report(int32_t s, uint32_t xx, uint32_t yy)
    {
        if (score <= s) { score = s; x = xx; y = yy; }
        else if (two && score2 <= s && (xx + dd < x || xx > x + dd)) { score2 = s; x2 = xx; y2 = yy; }
    }
};

struct PairView
{
    const uint8_t* pat; uint32_t pbits; uint32_t pbeg; uint32_t M;
    const uint8_t* txt; uint32_t tbits; uint32_t tbeg; uint32_t N;
    const uint8_t* qual;                                   // pattern qualities of this pair (NULL: all 0)
    __device__ __forceinline__ uint32_t pq(uint32_t j) const { return qual ? (uint32_t)__ldg(qual + j) : 0u; }
    __device__ __forceinline__ uint32_t p(uint32_t j) const { return sym_at_rt(pat, pbits, pbeg + j); }
    __device__ __forceinline__ uint32_t t(uint32_t i) const { return sym_at_rt(txt, tbits, tbeg + i); }
};

__device__ __forceinline__ PairView make_pair_view(const StrSet& P, const StrSet& T, uint32_t id, const uint8_t* quals = nullptr)
{
    PairView v;
    v.pat = P.sym; v.pbits = P.bits; v.pbeg = P.off[id]; v.M = P.off[id + 1] - v.pbeg;
    v.txt = T.sym; v.tbits = T.bits; v.tbeg = T.off[id]; v.N = T.off[id + 1] - v.tbeg;
    v.qual = quals ? quals + v.pbeg : nullptr;
    return v;
}

// ops of one pair, pushed one at a time; written 16 at a time when the pair's op row is 16-byte aligned (a byte
// store per op costs a 32-byte sector write each: the walk would be bound by them)
struct OpSink
{
    uint8_t* ops; uint32_t cap, n; bool wide;
    uint32_t cur, w0, w1, w2;          // the block of 16 ops being filled: three finished words + the current one
    __device__ __forceinline__ void init(uint8_t* o, uint32_t c)
    { ops = o; cap = c; n = 0; cur = w0 = w1 = w2 = 0; wide = (((uintptr_t)o) & 15u) == 0u; }
    // bytes [0, cnt) of the current block go to ops[base ..], clipped to cap; `done` words of the block are finished
    __device__ __forceinline__ void spill(uint32_t base, uint32_t cnt, uint32_t done)
    {
        for (uint32_t k = 0; k < cnt; ++k)
        {
            const uint32_t wi = k >> 2;
            const uint32_t word = wi >= done ? cur : (wi == 0u ? w0 : wi == 1u ? w1 : w2);
            if (base + k < cap) ops[base + k] = (uint8_t)(word >> (8u * (k & 3u)));
        }
    }
    __device__ __forceinline__ void push(uint32_t op)
    {
        if (!wide) { if (n < cap) ops[n] = (uint8_t)op; ++n; return; }
        cur |= op << (8u * (n & 3u));
        if ((n & 3u) == 3u)
        {
            const uint32_t k = (n >> 2) & 3u;
            if (k == 0u) w0 = cur; else if (k == 1u) w1 = cur; else if (k == 2u) w2 = cur;
            else if (n < cap) *(uint4*)(ops + (n & ~15u)) = make_uint4(w0, w1, w2, cur);
            else spill(n & ~15u, 16u, 3u);
            cur = 0;
        }
        ++n;
    }
    __device__ __forceinline__ void finish()
    {
        const uint32_t pend = n & 15u;
        if (wide && pend) spill(n - pend, pend, pend >> 2);
    }
};
#endif
