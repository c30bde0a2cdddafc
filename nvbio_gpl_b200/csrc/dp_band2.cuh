// dp_band2.cuh -- pieces shared by the sheared-coordinate banded kernels (dp_band2.cu: scoring, dp_band2_tb.cu:
// traceback): launch parameters, the per-thread byte streams, the host-side plan that maps a scoring scheme to the
// sheared constants. See dp_band2.cu for the derivation.
#pragma once
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
    int32_t        ga, gb;              // LOCAL (unsheared): SW deletion / insertion, Gotoh open / extension
};

__device__ __forceinline__ uint32_t c_pack2(int32_t v) { return ((uint32_t)v & 0xFFFFu) * 0x00010001u; }
__device__ __forceinline__ uint32_t c_fix2(int32_t g) { return (((uint32_t)(g - 1) & 0xFFFFu) << 16) | ((uint32_t)g & 0xFFFFu); }
__device__ __forceinline__ uint32_t c_prmt(uint32_t a, uint32_t b, uint32_t s)
{ uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ uint32_t c_imad(uint32_t a, uint32_t b, uint32_t c)
{ uint32_t d; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ int32_t c_lo(uint32_t x) { return (int32_t)(int16_t)(x & 0xFFFFu); }
__device__ __forceinline__ int32_t c_hi(uint32_t x) { return (int32_t)(int16_t)(x >> 16); }

static __device__ __noinline__ bool c_has_high_symbol(const uint8_t* sym, uint32_t beg, uint32_t L)
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
// skip = bytes of the string that are not streamed (the text positions the window starts with); total = symbols in the whole
// set. The streams fetch ALIGNED 16-byte blocks: a string whose first or last block sticks out of the symbol array
// [sym, sym + total) -- only the strings at the two ends of a set whose array is not 16-byte aligned there -- is not streamed
// at all (every load predicated off); the function then returns true and the caller hands the pair to the int32 kernel. No byte
// outside the caller's array is ever read.
__device__ __forceinline__ bool c_open(Stream& s, const uint8_t* sym, uint32_t total, uint32_t beg, uint32_t L, uint32_t skip, uint8_t* ring, uint32_t& seen)
{
    const uint32_t delta = (uint32_t)((uintptr_t)sym & 15u);
    s.base = sym - delta;
    const uint32_t pos = delta + beg + skip, b0 = pos & ~15u;
    s.end = delta + beg + L; s.nb = b0 + 32u;
    s.ring = ring; s.ptr = ring + (pos & 15u);
    const bool edge = L > skip && (b0 < delta || ((s.end + 15u) & ~15u) > delta + total);
    if (edge) s.end = 0u;
    const uint4 v0 = c_load(s, b0), v1 = c_load(s, b0 + 16u);
    seen |= v0.x | v0.y | v0.z | v0.w | v1.x | v1.y | v1.z | v1.w;      // neighbours' bytes included: a filter, confirmed exactly later
    c_put(ring, 0u, c_clip(s, b0, v0));
    c_put(ring, 16u, c_clip(s, b0 + 16u, v1));
    return edge;
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

// ---------------------------------------------------------------------------------------------
static inline int32_t iabs2(int32_t v) { return v < 0 ? -v : v; }

struct Band2Plan { Band2Params p; bool ok; };

static inline Band2Plan band2_plan(const nvb_aligner& a, uint32_t band, uint32_t maxM)
{
    Band2Plan pl; pl.ok = false;
    Band2Params& p = pl.p;
    const Scheme s = make_scheme(a);
    const bool g = a.kind == NVB_GOTOH;
    if (band != 7 && band != 15) return pl;
    if (maxM == 0) return pl;
    p.ga = s.ga; p.gb = s.gb;
    if (a.type == NVB_LOCAL)
    {
        // unsheared, biased by the largest gap penalty; raw signed profile bytes; 11 score bits in the arg-max key
        if (s.match < -127 || s.match > 127 || s.mismatch < -127 || s.mismatch > 127) return pl;
        if (s.ga >= 0 || s.gb >= 0 || s.ga < -64 || s.gb < -64) return pl;
        if (g && s.ga > s.gb) return pl;
        if ((uint64_t)max(s.match, 0) * maxM + 64u >= 2047u) return pl;
        p.alpha = p.beta = 0; p.d = 0;
        p.bias = max(iabs2(s.ga), iabs2(s.gb));
        p.s_match = (uint32_t)s.match & 0xFFu; p.s_mis = (uint32_t)s.mismatch & 0xFFu;
        p.init0 = p.init_rest = p.bias; p.init_step = 0;
        p.one = 1; p.k8 = 8; p.k32 = 32;
        pl.ok = true;
        return pl;
    }
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

