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

// ---------------------------------------------------------------------------------------------
// Mapping quality of the hits of nvb_map_reads: nvBowtie's BowtieMapq2 (nvBowtie/bowtie2/cuda/mapq.h:145-316, Bowtie 2's "V2"
// calculator), restated as tables. The scheme's min_score(len) is a float expression of the caller's scoring scheme
// (nvBowtie/bowtie2/cuda/params.h:95-100), so the caller tabulates it by read length -- the same boundary as the quality tables --
// and perfect_score(len) = len * match_bonus (scoring.h:254). All comparisons are single float operations in the reference's order.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mapq_v2(int32_t best_i, bool has_second, int32_t second_i, int32_t perfect, int32_t min_score_i, bool monotone)
{
    // no runner-up: fraction of the score range the best score must reach -> value; last entry = otherwise     [local, end-to-end]
    const float uniq_t[6] = { 0.8f, 0.7f, 0.6f, 0.5f, 0.4f, 0.3f };
    const uint8_t uniq_v[2][7] = { { 44, 42, 41, 36, 28, 24, 22 }, { 42, 40, 24, 23, 8, 3, 0 } };
    // runner-up: bands of | |best| - |second| | over the range, then tiers of the best score inside a band
    const float top_t[4] = { 0.9f, 0.8f, 0.7f, 0.6f };
    const uint8_t top_v[2][4][2] = { { {40,40}, {39,39}, {38,38}, {37,37} }, { {39,33}, {38,27}, {37,26}, {36,22} } };
    const float band_t[5] = { 0.5f, 0.4f, 0.3f, 0.2f, 0.1f };
    const float tier_t[2][5][2] = { { {.5f,.5f}, {.5f,.5f}, {.5f,.5f}, {.5f,.5f}, {.5f,.5f} },
                                    { {.84f,.68f}, {.84f,.68f}, {.88f,.67f}, {.88f,.67f}, {.88f,.67f} } };
    const uint8_t tier_v[2][5][4] = { { {35,25,25,20}, {34,21,21,19}, {33,18,18,16}, {32,17,17,12}, {31,14,14,9} },
                                      { {35,25,16,5},  {34,21,14,4},  {32,18,15,3},  {31,17,11,0},  {30,12,7,0} } };
    const float low_t[2] = { 0.5f, 0.67f };
    const uint8_t low_v[2][2][2] = { { {11, 2}, {1, 0} }, { {6, 2}, {1, 0} } };
    const int m = monotone ? 1 : 0;
    const float min_score = (float)min_score_i, diff = (float)perfect - min_score, best = (float)best_i;
    if (best < min_score) return 0u;
    const float best_over = best - min_score;
    if (!has_second)
    {
        #pragma unroll
        for (int i = 0; i < 6; ++i) if (best_over >= __fmul_rn(diff, uniq_t[i])) return uniq_v[m][i];
        return uniq_v[m][6];
    }
    const float best_diff = fabsf(fabsf(best) - fabsf((float)second_i));
    #pragma unroll
    for (int i = 0; i < 4; ++i) if (best_diff >= __fmul_rn(diff, top_t[i])) return top_v[m][i][best_over == diff ? 0 : 1];
    #pragma unroll
    for (int i = 0; i < 5; ++i)
        if (best_diff >= __fmul_rn(diff, band_t[i]))
        {
            if (best_over == diff) return tier_v[m][i][0];
            if (best_over >= __fmul_rn(diff, tier_t[m][i][0])) return tier_v[m][i][1];
            if (best_over >= __fmul_rn(diff, tier_t[m][i][1])) return tier_v[m][i][2];
            return tier_v[m][i][3];
        }
    return low_v[m][best_diff > 0.f ? 0 : 1][best_over >= __fmul_rn(diff, low_t[m]) ? 0 : 1];
}

__global__ void __launch_bounds__(MDS_THREADS)
mapq_hits_kernel(const nvb_read_hit* __restrict__ hits, const uint32_t* __restrict__ off, uint32_t n, const nvb_pair_hit* __restrict__ pairs,
                 const int32_t* __restrict__ min_by_len, uint32_t max_len, int32_t match_bonus, uint32_t monotone, uint8_t* __restrict__ mapq)
{
    const uint32_t r = blockIdx.x * MDS_THREADS + threadIdx.x;
    if (r >= n) return;
    const nvb_read_hit h = hits[r];
    if (!(h.flags & 2u)) { mapq[r] = 0; return; }
    auto len_of = [&](uint32_t i) { const uint32_t l = off[i + 1] - off[i]; return l < max_len ? l : max_len; };
    const uint32_t L = len_of(r);
    int32_t best = h.score, perfect = (int32_t)L * match_bonus, minsc = min_by_len[L];
    bool has_second = h.second_score != INT32_MIN;
    if (pairs && (pairs[r >> 1].flags & 4u))
    {
        // a concordant pair: scores, perfect and minimum scores of both mates summed (mapq.h:153-156, io/alignments.h:225); our
        // reduction keeps per-mate runners-up only, so a pair has no paired second best (has_second_paired() false)
        const uint32_t o = r ^ 1u, Lo = len_of(o);
        best += hits[o].score; perfect += (int32_t)Lo * match_bonus; minsc += min_by_len[Lo];
        has_second = false;
    }
    mapq[r] = (uint8_t)mapq_v2(best, has_second, h.second_score, perfect, minsc, monotone != 0u);
}

extern "C" int nvb_map_hits_mapq(nvb_ctx* ctx, const nvb_read_hit* hits, const uint32_t* read_offsets, uint32_t n, const nvb_pair_hit* pairs,
                                 const int32_t* min_score_by_len_h, uint32_t max_len, int32_t match_bonus, uint32_t monotone,
                                 uint8_t* mapq, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!hits || !read_offsets || !min_score_by_len_h || !mapq || (pairs && (n & 1u))) return NVB_ERR_ARG;
    void* d_tab; int rc;
    if ((rc = nvb_stage(ctx, 56, sizeof(int32_t) * ((size_t)max_len + 1), &d_tab))) return rc;
    NVB_CUDA(ctx, cudaMemcpyAsync(d_tab, min_score_by_len_h, sizeof(int32_t) * ((size_t)max_len + 1), cudaMemcpyHostToDevice, (cudaStream_t)stream));
    mapq_hits_kernel<<<(n + MDS_THREADS - 1) / MDS_THREADS, MDS_THREADS, 0, (cudaStream_t)stream>>>(
        hits, read_offsets, n, pairs, (const int32_t*)d_tab, max_len, match_bonus, monotone, mapq);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}
