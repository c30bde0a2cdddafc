// mds.cu -- nvBowtie's MDS byte-code (the device-side precursor of SAM's MD:Z tag) and the edit distance of an alignment,
// from our push-order op lists.
//
//   reference: finish_alignment_kernel, nvBowtie/bowtie2/cuda/traceback_inl.h:555-690 (the loop over the CIGAR that builds
//   mds_local[] and counts `ed`), op codes nvbio/io/alignments.h:37-42, consumer nvbio/io/output/output_sam.cpp:216-289.
//
// Byte-code: [len lo, len hi] then a sequence of
//   MDS_MATCH(0), run (1..255)            a run of equal symbols; a longer run continues in another MATCH record
//   MDS_MISMATCH(1), read symbol          (the READ's symbol, as the reference stores it)
//   MDS_INSERTION(2) | MDS_DELETION(3), l (low 8 bits of the run length), then l symbols (read symbols / reference symbols)
// The reference patches the last MATCH record in place; here a pending run is flushed when anything else is emitted.
#include "common.cuh"

#define MDS_THREADS 128
enum { MDS_MATCH = 0, MDS_MISMATCH = 1, MDS_INSERTION = 2, MDS_DELETION = 3 };
enum { OP_SUB = 0, OP_INS = 1, OP_DEL = 2, OP_CLP = 3 };      // io::Cigar::{SUBSTITUTION, INSERTION, DELETION, SOFT_CLIPPING}

struct MdsOut
{
    uint8_t* p; uint32_t cap, len, run;
    __device__ __forceinline__ void put(uint32_t b) { if (len < cap) p[len] = (uint8_t)b; ++len; }
    __device__ __forceinline__ void flush() { if (run) { put(MDS_MATCH); put(run); run = 0; } }
};

// o[0 .. n_ops): ops in push order (last op of the alignment first); read(j) / ref(k): symbol j of the aligned strand read /
// symbol k of the text, 0-based; (j0, k0) = first aligned read / text symbol. Returns the byte-code length (may exceed cap:
// truncated output) and the edit distance.
template <typename ReadAt, typename RefAt>
__device__ __forceinline__ uint32_t mds_encode(const uint8_t* __restrict__ o, uint32_t n_ops, ReadAt read, RefAt ref,
                                               uint32_t j0, uint32_t k0, uint8_t* out, uint32_t cap, uint32_t& ed_out)
{
    MdsOut m = { out, cap, 2u, 0u };
    uint32_t ed = 0, j = j0, k = k0;
    for (uint32_t i = n_ops; i > 0;)
    {
        const uint32_t op = o[i - 1] & 3u;
        uint32_t l = 1;
        while (l < i && (o[i - 1 - l] & 3u) == op && l < 16383u) ++l;      // one io::Cigar entry (14-bit length)
        i -= l;
        if (op == OP_SUB)
        {
            for (uint32_t n = 0; n < l; ++n, ++j, ++k)
            {
                const uint32_t rc = read(j), gc = ref(k);
                if (rc == gc) { if (m.run == 255u) m.flush(); ++m.run; }
                else          { m.flush(); m.put(MDS_MISMATCH); m.put(rc); ++ed; }
            }
        }
        else
        {
            m.flush();
            m.put(op == OP_DEL ? MDS_DELETION : MDS_INSERTION); m.put(l & 0xFFu);
            for (uint32_t n = 0; n < l; ++n)
            {
                if (op == OP_DEL) m.put(ref(k++)); else m.put(read(j++));
                if (op != OP_CLP) ++ed;
            }
        }
    }
    m.flush();
    if (cap > 0) out[0] = (uint8_t)(m.len & 0xFFu);
    if (cap > 1) out[1] = (uint8_t)(m.len >> 8);
    ed_out = ed;
    return m.len;
}

__global__ void __launch_bounds__(MDS_THREADS)
mds_sets_kernel(const uint8_t* __restrict__ ops, uint32_t ops_stride, const uint32_t* __restrict__ n_ops, uint32_t n,
                const StrSet P, const StrSet T, const nvb_alignment* __restrict__ aln,
                uint8_t* __restrict__ mds, uint32_t mds_stride, uint32_t* __restrict__ mds_len, uint32_t* __restrict__ ed)
{
    const uint32_t i = blockIdx.x * MDS_THREADS + threadIdx.x;
    if (i >= n) return;
    const uint32_t pb = P.off[i], tb = T.off[i];
    uint32_t len = n_ops[i]; if (len > ops_stride) len = ops_stride;
    uint32_t e = 0;
    const uint32_t l = mds_encode(ops + (size_t)i * ops_stride, len,
                                  [&](uint32_t j) { return sym_at_rt(P.sym, P.bits, pb + j); }, [&](uint32_t k) { return sym_at_rt(T.sym, T.bits, tb + k); },
                                  aln[i].source_y, aln[i].source_x, mds + (size_t)i * mds_stride, mds_stride, e);
    if (mds_len) mds_len[i] = l;
    if (ed) ed[i] = e;
}

__global__ void __launch_bounds__(MDS_THREADS)
mds_hits_kernel(const nvb_read_hit* __restrict__ hits, const uint8_t* __restrict__ ops, uint32_t ops_stride, uint32_t n,
                const StrSet reads, const uint32_t* __restrict__ genome,
                uint8_t* __restrict__ mds, uint32_t mds_stride, uint32_t* __restrict__ mds_len, uint32_t* __restrict__ ed)
{
    const uint32_t r = blockIdx.x * MDS_THREADS + threadIdx.x;
    if (r >= n) return;
    const nvb_read_hit h = hits[r];
    uint8_t* out = mds + (size_t)r * mds_stride;
    if (!(h.flags & 2u))
    {
        // unaligned read: an empty byte-code
        if (mds_stride > 0) out[0] = 2; if (mds_stride > 1) out[1] = 0;
        if (mds_len) mds_len[r] = 2; if (ed) ed[r] = 0;
        return;
    }
    const uint32_t rb = reads.off[r], L = reads.off[r + 1] - rb, strand = h.flags & 1u;
    uint32_t len = h.n_ops; if (len > ops_stride) len = ops_stride;
    uint32_t e = 0;
    const uint32_t l = mds_encode(ops + (size_t)r * ops_stride, len,
                                  [&](uint32_t j) -> uint32_t { if (!strand) return sym_at_rt(reads.sym, reads.bits, rb + j); const uint32_t c = sym_at_rt(reads.sym, reads.bits, rb + L - 1u - j); return c < 4u ? 3u - c : c; },
                                  [&](uint32_t k) -> uint32_t { const uint32_t p = h.begin + k; return (__ldg(genome + (p >> 4)) >> (30u - 2u * (p & 15u))) & 3u; },
                                  0u, 0u, out, mds_stride, e);
    if (mds_len) mds_len[r] = l;
    if (ed) ed[r] = e;
}

extern "C" int nvb_ops_to_mds(nvb_ctx* ctx, const uint8_t* ops, uint32_t ops_stride, const uint32_t* n_ops, uint32_t n,
                              nvb_string_set patterns, nvb_string_set texts, const nvb_alignment* alignments,
                              uint8_t* mds, uint32_t mds_stride, uint32_t* mds_len, uint32_t* edit_distance, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!ops || !n_ops || !alignments || !mds || mds_stride < 2u || patterns.count < n || texts.count < n) return NVB_ERR_ARG;
    if (!patterns.symbols || !patterns.offsets || !texts.symbols || !texts.offsets) return NVB_ERR_ARG;
    if (!(patterns.bits == 8 || patterns.bits == 4 || patterns.bits == 2) || !(texts.bits == 8 || texts.bits == 4 || texts.bits == 2)) return NVB_ERR_ARG;
    mds_sets_kernel<<<(n + MDS_THREADS - 1) / MDS_THREADS, MDS_THREADS, 0, (cudaStream_t)stream>>>(
        ops, ops_stride, n_ops, n, make_set(patterns), make_set(texts), alignments, mds, mds_stride, mds_len, edit_distance);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_map_hits_to_mds(nvb_ctx* ctx, const nvb_read_hit* hits, const uint8_t* ops, uint32_t ops_stride,
                                   nvb_string_set reads, const uint32_t* genome_words, uint8_t* mds, uint32_t mds_stride,
                                   uint32_t* mds_len, uint32_t* edit_distance, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    const uint32_t n = reads.count;
    if (n == 0) return NVB_OK;
    if (!hits || !ops || !genome_words || !mds || mds_stride < 2u || !reads.symbols || !reads.offsets) return NVB_ERR_ARG;
    if (!(reads.bits == 8 || reads.bits == 4 || reads.bits == 2)) return NVB_ERR_ARG;
    mds_hits_kernel<<<(n + MDS_THREADS - 1) / MDS_THREADS, MDS_THREADS, 0, (cudaStream_t)stream>>>(
        hits, ops, ops_stride, n, make_set(reads), genome_words, mds, mds_stride, mds_len, edit_distance);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
