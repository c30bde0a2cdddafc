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

struct __align__(32) FMBlock { uint32_t w[4]; uint32_t occ[4]; };

__device__ __forceinline__ FMBlock fm_load_block(const uint4* blk, uint32_t b)
{
    FMBlock r;
    const void* p = (const void*)(blk + 2ull * b);
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]),
                   "=r"(r.occ[0]), "=r"(r.occ[1]), "=r"(r.occ[2]), "=r"(r.occ[3])
                 : "l"(p));
    return r;
}

// one bit (at the even bit of its 2-bit slot) per symbol of w equal to c
__device__ __forceinline__ uint32_t fm_eq_bits(uint32_t w, uint32_t cpat)
{
    const uint32_t x = w ^ cpat;              // 00 where the symbol equals c
    return ~(x | (x >> 1)) & 0x55555555u;
}

// occurrences of c among symbols 0..pos (inclusive, pos in 0..63) of the block, plus the counter
__device__ __forceinline__ uint32_t fm_block_rank(const FMBlock& b, uint32_t pos, uint32_t c)
{
    const uint32_t cpat = c * 0x55555555u;
    const int32_t  n    = (int32_t)pos + 1;   // symbols to count
    uint32_t acc = 0;
    #pragma unroll
    for (int wi = 0; wi < 4; wi += 2)
    {
        const int32_t n0 = min(max(n - 16 * wi, 0), 16);
        const int32_t n1 = min(max(n - 16 * (wi + 1), 0), 16);
        // big-endian packing: symbol 0 sits in bits 31:30 -> keep the top 2*n bits
        const uint32_t m0 = (uint32_t)(0xFFFFFFFF00000000ull >> (2 * n0));
        const uint32_t m1 = (uint32_t)(0xFFFFFFFF00000000ull >> (2 * n1));
        const uint32_t t0 = fm_eq_bits(b.w[wi], cpat) & m0;
        const uint32_t t1 = fm_eq_bits(b.w[wi + 1], cpat) & m1;
        acc += __popc(t0 | (t1 << 1));
    }
    const uint32_t base = c == 0 ? b.occ[0] : c == 1 ? b.occ[1] : c == 2 ? b.occ[2] : b.occ[3];
    return base + acc;
}

// all four symbol counts (rank4: nvbio/fmindex/rank_dictionary_inl.h:450-478)
__device__ __forceinline__ uint4 fm_block_rank4(const FMBlock& b, uint32_t pos)
{
    return make_uint4(fm_block_rank(b, pos, 0), fm_block_rank(b, pos, 1),
                      fm_block_rank(b, pos, 2), fm_block_rank(b, pos, 3));
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

__device__ __forceinline__ uint32_t fm_rank1(const FMView& f, uint32_t k, uint32_t c)
{
    const FMReq q = fm_make_req(f, k, c);
    if (!q.load) return q.value;
    const FMBlock b = fm_load_block(f.blk, q.k >> 6);
    return fm_block_rank(b, q.k & 63u, c);
}

// rank(fmi, (x,y), c): both blocks are requested before either is consumed (2 loads in flight);
// when both indices fall in one block it is fetched once (rank_dictionary_inl.h:425-448)
__device__ __forceinline__ uint2 fm_rank2(const FMView& f, uint32_t x, uint32_t y, uint32_t c)
{
    const FMReq qx = fm_make_req(f, x, c);
    const FMReq qy = fm_make_req(f, y, c);
    FMBlock bx, by;
    const bool same = qx.load && qy.load && ((qx.k >> 6) == (qy.k >> 6));
    if (qy.load) by = fm_load_block(f.blk, qy.k >> 6);
    if (qx.load && !same) bx = fm_load_block(f.blk, qx.k >> 6);
    uint2 r;
    r.y = qy.load ? fm_block_rank(by, qy.k & 63u, c) : qy.value;
    r.x = qx.load ? fm_block_rank(same ? by : bx, qx.k & 63u, c) : qx.value;
    return r;
}

__device__ __forceinline__ uint32_t fm_bwt_symbol(const FMBlock& b, uint32_t pos)
{
    const uint32_t wi = pos >> 4;
    const uint32_t w = wi == 0 ? b.w[0] : wi == 1 ? b.w[1] : wi == 2 ? b.w[2] : b.w[3];
    return (w >> (30u - 2u * (pos & 15u))) & 3u;
}
