// fm_device.cuh -- device functions of the FM-index half (rank over the fused nvBWT blocks).
//
// Replaces, for the one layout nvBowtie uses (2-bit symbols, OCC_INT = 64, uint4 words | uint4
// counters interleaved), the template stack
//   popc_2bit            nvbio/basic/popcount_inl.h:230-246,318-353
//   dispatch_rank<2,64>  nvbio/fmindex/rank_dictionary_inl.h:351-448
//   rank(fmi, ...)       nvbio/fmindex/fmindex_inl.h:31-88
// One rank touches exactly one 32-byte block (= one DRAM/L2 sector): it is fetched with a single
// 256-bit load (LDG.E.ENL2.256 on sm_100a) so that the L1 sees one wavefront per block.
#pragma once
#include "common.cuh"

struct FMBlock { uint32_t w0, w1, w2, w3; uint32_t o0, o1, o2, o3; };

// the few thousand blocks under the top of the search tree (range still wider than FMView::hot_range rows: every query
// of a batch walks through them) are kept in L1; everything below is a one-off and bypasses it
__device__ __forceinline__ FMBlock fm_load_block_hot(const uint4* blk, uint32_t b)
{
    FMBlock r;
    const void* p = (const void*)(blk + 2ull * b);
    asm volatile("ld.global.nc.L1::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w0), "=r"(r.w1), "=r"(r.w2), "=r"(r.w3), "=r"(r.o0), "=r"(r.o1), "=r"(r.o2), "=r"(r.o3)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ FMBlock fm_load_block(const uint4* blk, uint32_t b)
{
    FMBlock r;
    const void* p = (const void*)(blk + 2ull * b);
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w0), "=r"(r.w1), "=r"(r.w2), "=r"(r.w3), "=r"(r.o0), "=r"(r.o1), "=r"(r.o2), "=r"(r.o3)
                 : "l"(p));
    return r;
}

// Position masks of a block, one entry per pos = 0..63, staged in shared memory by every kernel:
// x = mask over the flag word of words 0|1, y = over words 2|3. Flags of an even word sit on the even
// bit of each 2-bit slot, flags of an odd word on the odd bit, so two words share one popcount.
#define FM_LUT_ENTRIES 64
__device__ __forceinline__ uint32_t fm_top_bits(int32_t k) { return k > 0 ? 0xFFFFFFFFu << (32 - 2 * k) : 0u; }
__device__ __forceinline__ void fm_lut_init(uint2* lut)
{
    for (uint32_t pos = threadIdx.x; pos < FM_LUT_ENTRIES; pos += blockDim.x)
    {
        const int32_t n = (int32_t)pos + 1;   // big-endian packing: symbol 0 sits in bits 31:30
        uint2 m;
        m.x = (fm_top_bits(min(n, 16)) & 0x55555555u) | (fm_top_bits(min(max(n - 16, 0), 16)) & 0xAAAAAAAAu);
        m.y = (fm_top_bits(min(max(n - 32, 0), 16)) & 0x55555555u) | (fm_top_bits(min(max(n - 48, 0), 16)) & 0xAAAAAAAAu);
        lut[pos] = m;
    }
    __syncthreads();
}

// one flag per symbol equal to c: on the even bit of the slot for w_even, on the odd bit for w_odd
__device__ __forceinline__ uint32_t fm_eq_pair(uint32_t w_even, uint32_t w_odd, uint32_t cpat)
{
    const uint32_t x0 = w_even ^ cpat, x1 = w_odd ^ cpat;   // 00 where the symbol equals c
    const uint32_t a = x0 | (x0 >> 1), b = x1 | (x1 << 1);
    return ~((a & 0x55555555u) | (b & 0xAAAAAAAAu));
}

__device__ __forceinline__ uint32_t fm_eq_bits(uint32_t w, uint32_t cpat)
{
    const uint32_t x = w ^ cpat;
    return ~(x | (x >> 1)) & 0x55555555u;
}

// occurrences of c among symbols 0..pos (inclusive, pos in 0..63) of the block, plus the counter
__device__ __forceinline__ uint32_t fm_block_rank(const FMBlock& b, uint32_t pos, uint32_t c, const uint2* lut)
{
    const uint32_t cpat = c * 0x55555555u;
    const uint2 m = lut[pos];
    const uint32_t f01 = fm_eq_pair(b.w0, b.w1, cpat) & m.x;
    const uint32_t f23 = fm_eq_pair(b.w2, b.w3, cpat) & m.y;
    const uint32_t base = (c & 2u) ? ((c & 1u) ? b.o3 : b.o2) : ((c & 1u) ? b.o1 : b.o0);
    return base + __popc(f01) + __popc(f23);
}

// all four symbol counts (rank4: nvbio/fmindex/rank_dictionary_inl.h:450-478)
__device__ __forceinline__ uint4 fm_block_rank4(const FMBlock& b, uint32_t pos, const uint2* lut)
{
    return make_uint4(fm_block_rank(b, pos, 0, lut), fm_block_rank(b, pos, 1, lut),
                      fm_block_rank(b, pos, 2, lut), fm_block_rank(b, pos, 3, lut));
}

// A rank request after the FM-index special cases (fmindex_inl.h:31-47):
//   k == -1 -> 0 ; k == length -> count(c) ; k >= primary -> k-1 (the '$' row is not stored)
struct FMReq { uint32_t k; uint32_t value; bool load; };

__device__ __forceinline__ FMReq fm_make_req(const FMView& f, uint32_t k, uint32_t c)
{
    FMReq r; r.load = false; r.value = 0; r.k = 0;
    if (k == 0xFFFFFFFFu) return r;
    if (k == f.length) { r.value = f.L2[c + 1] - f.L2[c]; return r; }
    if (k >= f.primary) --k;
    if (k == 0xFFFFFFFFu) return r;   // primary == 0 and k == 0
    r.k = k; r.load = true;
    return r;
}

__device__ __forceinline__ uint32_t fm_rank1(const FMView& f, uint32_t k, uint32_t c, const uint2* lut)
{
    const FMReq q = fm_make_req(f, k, c);
    if (!q.load) return q.value;
    const FMBlock b = fm_load_block(f.blk, q.k >> 6);
    return fm_block_rank(b, q.k & 63u, c, lut);
}

// rank(fmi, (x,y), c): both blocks are requested before either is consumed (2 loads in flight);
// when both indices fall in one block it is fetched once (rank_dictionary_inl.h:425-448)
__device__ __forceinline__ uint2 fm_rank2(const FMView& f, uint32_t x, uint32_t y, uint32_t c, const uint2* lut)
{
    const FMReq qx = fm_make_req(f, x, c);
    const FMReq qy = fm_make_req(f, y, c);
    const bool same = qx.load && qy.load && ((qx.k >> 6) == (qy.k >> 6));
    FMBlock by = { 0, 0, 0, 0, 0, 0, 0, 0 }, bx = by;
    if (qy.load) by = fm_load_block(f.blk, qy.k >> 6);
    if (qx.load && !same) bx = fm_load_block(f.blk, qx.k >> 6);
    if (same) bx = by;
    uint2 r;
    r.y = qy.load ? fm_block_rank(by, qy.k & 63u, c, lut) : qy.value;
    r.x = qx.load ? fm_block_rank(bx, qx.k & 63u, c, lut) : qx.value;
    return r;
}

// rank(fmi, (x, y), c) inside the backward-search loop, after its first step: there x = previous x - 1 >= 0 and
// y <= length, so of the FM special cases (fmindex_inl.h:31-47) only "k >= primary -> k - 1" is left (k == length falls
// under it: BWT[0..length-1] holds every occurrence), and with primary > 0 the adjusted index cannot underflow. When both
// indices fall into one block (the common case once the range is narrow) the block is fetched once and its match flags
// are computed once for both ranks (rank_dictionary_inl.h:425-448 fetches it once, too).
__device__ __forceinline__ uint2 fm_rank2_inner(const FMView& f, uint32_t x, uint32_t y, uint32_t c, const uint2* lut)
{
    const uint32_t kx = x - (x >= f.primary ? 1u : 0u), ky = y - (y >= f.primary ? 1u : 0u);
    const uint32_t bxi = kx >> 6, byi = ky >> 6;
    const uint32_t cpat = c * 0x55555555u;
    const bool hot = y - x > f.hot_range;
    const FMBlock by = hot ? fm_load_block_hot(f.blk, byi) : fm_load_block(f.blk, byi);
    uint32_t f01 = fm_eq_pair(by.w0, by.w1, cpat), f23 = fm_eq_pair(by.w2, by.w3, cpat);
    uint32_t base = (c & 2u) ? ((c & 1u) ? by.o3 : by.o2) : ((c & 1u) ? by.o1 : by.o0);
    const uint2 my = lut[ky & 63u], mx = lut[kx & 63u];
    uint2 r;
    r.y = base + __popc(f01 & my.x) + __popc(f23 & my.y);
    if (bxi != byi)
    {
        const FMBlock bx = hot ? fm_load_block_hot(f.blk, bxi) : fm_load_block(f.blk, bxi);
        f01 = fm_eq_pair(bx.w0, bx.w1, cpat); f23 = fm_eq_pair(bx.w2, bx.w3, cpat);
        base = (c & 2u) ? ((c & 1u) ? bx.o3 : bx.o2) : ((c & 1u) ? bx.o1 : bx.o0);
    }
    r.x = base + __popc(f01 & mx.x) + __popc(f23 & mx.y);
    return r;
}

__device__ __forceinline__ uint4 fm_rank4_at(const FMView& f, uint32_t k, const uint2* lut)
{
    // rank4(fmi, k) with the FM special cases (fmindex_inl.h:100-125)
    if (k == 0xFFFFFFFFu) return make_uint4(0, 0, 0, 0);
    if (k == f.length) return make_uint4(f.L2[1] - f.L2[0], f.L2[2] - f.L2[1], f.L2[3] - f.L2[2], f.L2[4] - f.L2[3]);
    if (k >= f.primary) --k;
    if (k == 0xFFFFFFFFu) return make_uint4(0, 0, 0, 0);
    const FMBlock b = fm_load_block(f.blk, k >> 6);
    return fm_block_rank4(b, k & 63u, lut);
}
__device__ __forceinline__ uint32_t fm_comp(const uint4& v, uint32_t c) { return (c & 2u) ? ((c & 1u) ? v.w : v.z) : ((c & 1u) ? v.y : v.x); }


// map<find_exact> for one string: symbol(t) = the symbol search step t consumes, emit(x, y) receives every non-empty range in the
// reference's push order (substitutions by position and symbol, then the exact match)
template <typename Symbol, typename Emit>
__device__ __forceinline__ void fm_approx_search(const FMView& f, const uint2* s_lut, const uint32_t* s_L2, uint32_t len, uint32_t len1,
                                                 bool find_exact, Symbol symbol, Emit emit)
{
    const uint32_t l1 = min(len1, len);
    // match_range over steps [t0, t_end)
    auto tail = [&](uint32_t t0, uint32_t& x, uint32_t& y, uint32_t t_end)
    {
        for (uint32_t t = t0; t < t_end && x <= y; ++t)
        {
            const uint32_t c = symbol(t);
            if (c > 3u) { x = 1; y = 0; return; }
            const uint2 r = (x != 0u && f.primary != 0u) ? fm_rank2_inner(f, x - 1u, y, c, s_lut) : fm_rank2(f, x - 1u, y, c, s_lut);
            x = s_L2[c] + r.x + 1u; y = s_L2[c] + r.y;
        }
    };
    uint32_t bx = 0, by = f.length;
    tail(0, bx, by, l1);
    for (uint32_t t = l1; t < len && bx <= by; ++t)
    {
        const uint32_t c = symbol(t);
        const uint4 lo = fm_rank4_at(f, bx - 1u, s_lut), hi = fm_rank4_at(f, by, s_lut);
        for (uint32_t sub = 0; sub < 4u; ++sub)
        {
            const uint32_t ol = fm_comp(lo, sub), oh = fm_comp(hi, sub);
            if (sub != c && oh > ol)
            {
                uint32_t x = s_L2[sub] + ol + 1u, y = s_L2[sub] + oh;
                tail(t + 1u, x, y, len);
                if (x <= y) emit(x, y);
            }
        }
        if (c < 4u) { bx = s_L2[c] + fm_comp(lo, c) + 1u; by = s_L2[c] + fm_comp(hi, c); }
        else { bx = 1; by = 0; break; }
    }
    if (find_exact && bx <= by) emit(bx, by);
}

__device__ __forceinline__ uint32_t fm_bwt_symbol(const FMBlock& b, uint32_t pos)
{
    const uint32_t wi = pos >> 4;
    const uint32_t w = (wi & 2u) ? ((wi & 1u) ? b.w3 : b.w2) : ((wi & 1u) ? b.w1 : b.w0);
    return (w >> (30u - 2u * (pos & 15u))) & 3u;
}

// LF-walk from SA row j to the nearest sampled row (j % 16 == 0). One block load per step: the
// BWT symbol and its rank come from the same 32-byte block.
__device__ __forceinline__ uint2 fm_locate_walk_one(const FMView& f, uint32_t j, const uint2* lut)
{
    uint32_t t = 0;
    while (j & f.sa_mask)
    {
        if (j != f.primary)
        {
            const uint32_t jj = j < f.primary ? j : j - 1u;        // fmindex_inl.h:384
            const FMBlock b = fm_load_block(f.blk, jj >> 6);
            const uint32_t c = fm_bwt_symbol(b, jj & 63u);
            // rank(fmi, j, c) with the FM special cases (fmindex_inl.h:31-47)
            uint32_t r;
            if (j == f.length) r = f.L2[c + 1] - f.L2[c];
            else r = fm_block_rank(b, jj & 63u, c, lut);
            j = f.L2[c] + r;
        }
        else j = 0;
        ++t;
    }
    return make_uint2(j, t);
}

// Many LF-walks per thread, load-balanced: walk lengths vary (geometric, 15 steps on average), so a thread owns the strided list of walks
// first, first + stride, ... and starts its next one as soon as one ends; the warp vote keeps the loop converged.
//   fetch(h) -> SA row of walk h ;  emit(h, sampled_row, steps)
// DENSE = false: the 16-row SSA of the nvBWT layout (mask known at compile time); true: the optional denser table
template <bool DENSE, typename Fetch, typename Emit>
__device__ __forceinline__ void fm_walk_balanced(const FMView& f, uint32_t n, uint32_t first, uint32_t stride, const uint2* lut,
                                                 Fetch fetch, Emit emit)
{
    const uint32_t sa_mask = DENSE ? f.sa_mask : 15u;
    uint32_t h = first;
    bool alive = h < n;
    uint32_t j = alive ? fetch(h) : 0u, t = 0;
    uint32_t hn = h + stride;
    uint32_t jn = (alive && hn < n) ? fetch(hn) : 0u;          // the next walk's row is fetched one walk ahead
    while (__any_sync(0xFFFFFFFFu, alive))
    {
        if (alive && (j & sa_mask) == 0u)
        {
            emit(h, j, t);
            h = hn; j = jn; t = 0;
            alive = h < n;
            hn = h + stride;
            jn = (alive && hn < n) ? fetch(hn) : 0u;
        }
        if (alive && (j & sa_mask) != 0u)
        {
            uint32_t nj = 0;                                    // j == primary -> 0 (fmindex_inl.h:380-383)
            if (j != f.primary)
            {
                const uint32_t jj = j < f.primary ? j : j - 1u;             // fmindex_inl.h:384
                const FMBlock b = fm_load_block(f.blk, jj >> 6);
                const uint32_t c = fm_bwt_symbol(b, jj & 63u);
                const uint32_t r = (j == f.length) ? f.L2[c + 1] - f.L2[c] : fm_block_rank(b, jj & 63u, c, lut);
                nj = f.L2[c] + r;
            }
            j = nj; ++t;
        }
    }
}
