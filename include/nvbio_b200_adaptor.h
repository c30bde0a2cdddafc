/*
 * nvbio_b200_adaptor.h -- the ONE header a maintainer of the reference (nvbio-gpl) adds next to its own to run the seed-and-extend
 * hot path on libnvbio_b200.so with the reference's call shapes unchanged. It is compiled against the reference's own headers
 * (tests/adaptor/adaptor_test.cu does that in this repository and checks the results against the reference's host code).
 *
 *   nvbio::B200Context c(0);
 *   nvb_fmindex f = nvbio::b200_view( fmi_data );                       // io::FMIndexDataDevice (any FMIndexDataCore)
 *   nvbio::match( c, f, queries, ranges_d );                             // match(fmi, pattern, len) for a ConcatenatedStringSet
 *   aln::batch_alignment_score( aligner, patterns, texts, sinks_d, aln::DeviceB200Scheduler( c ), maxM, maxN );
 *   aln::batch_banded_alignment_score<15>( aligner, patterns, texts, sinks_d, aln::DeviceB200Scheduler( c ), maxM, maxN );
 *
 * Host code, C++11; include after the reference headers it adapts are on the include path.
 */
#pragma once
#include <nvbio_b200.h>
#include <nvbio/basic/types.h>
#include <nvbio/basic/exceptions.h>
#include <nvbio/strings/string_set.h>
#include <nvbio/alignment/alignment.h>
#include <nvbio/alignment/utils.h>
#include <nvbio/alignment/sink.h>
#include <nvbio/alignment/batched.h>
#include <nvbio/io/fmindex/fmindex.h>
#include <cuda_runtime.h>
#include <stdexcept>
#include <string>

namespace nvbio {

/// one per (device, host thread), like a Batched* context object (nvbio/alignment/batched.h:279-289)
struct B200Context
{
    nvb_ctx* ctx;
    void*    temp;          ///< grow-only scratch handed to the scoring / traceback calls (the reference's temp_storage protocol)
    size_t   temp_bytes;
    void*    packed;        ///< grow-only staging for the library's packed result records (repacked into the reference's padded structs)
    size_t   packed_bytes;
    explicit B200Context(const int device = 0) : ctx(NULL), temp(NULL), temp_bytes(0), packed(NULL), packed_bytes(0)
    { if (nvb_ctx_create( device, &ctx ) != NVB_OK) throw std::runtime_error( "nvb_ctx_create failed: no sm_100a device" ); }
    ~B200Context() { if (temp) cudaFree( temp ); if (packed) cudaFree( packed ); nvb_ctx_destroy( ctx ); }
    void* staging(const size_t bytes)
    {
        if (bytes > packed_bytes)
        {
            if (packed) cudaFree( packed );
            if (cudaMalloc( &packed, bytes ) != cudaSuccess) { packed = NULL; packed_bytes = 0; throw std::runtime_error( "B200Context: out of device memory" ); }
            packed_bytes = bytes;
        }
        return packed;
    }
    void* scratch(const size_t bytes)
    {
        if (bytes > temp_bytes)
        {
            if (temp) cudaFree( temp );
            if (cudaMalloc( &temp, bytes ) != cudaSuccess) { temp = NULL; temp_bytes = 0; throw std::runtime_error( "B200Context: out of device memory" ); }
            temp_bytes = bytes;
        }
        return temp;
    }
    void check(const int rc) const
    { if (rc != NVB_OK) throw std::runtime_error( std::string( "libnvbio_b200: " ) + nvb_status_string( rc ) + " / " + nvb_last_cuda_error( ctx ) ); }
private:
    B200Context(const B200Context&); B200Context& operator=(const B200Context&);
};

// ---- seeding: io::FMIndexDataCore (nvbio/io/fmindex/fmindex.h:74-139; FMIndexDataDevice derives from it) -> nvb_fmindex ------------
/// the L2 table of a device index lives in device memory (fmindex_impl.cu:756-760): it is fetched once here
inline nvb_fmindex b200_view(const io::FMIndexDataCore& d, const bool reverse = false, const bool device_L2 = true)
{
    nvb_fmindex f;
    f.length  = d.m_seq_length;
    f.primary = reverse ? d.m_rprimary : d.m_primary;
    if (device_L2) { if (cudaMemcpy( f.L2, d.m_L2, 5 * sizeof(uint32), cudaMemcpyDeviceToHost ) != cudaSuccess) throw std::runtime_error( "b200_view: L2" ); }
    else for (int i = 0; i < 5; ++i) f.L2[i] = d.m_L2[i];
    f.occ_int = io::FMIndexDataCore::OCC_INT;
    f.sa_int  = io::FMIndexDataCore::SA_INT;
    f.bwt_occ = reverse ? d.m_rbwt_occ : d.m_bwt_occ;                   // the fused bwt|occ image, as the reference lays it out
    f.ssa     = (reverse ? d.m_rssa : d.m_ssa).m_ssa;                   // SSA_index_multiple_context<16>::m_ssa (nvbio/fmindex/ssa.h)
    f.ktab = NULL; f.ktab_k = 0; f.dssa = NULL; f.dssa_int = 0;
    return f;
}

template <typename string_set_type>
inline nvb_string_set b200_set(const string_set_type& s)               // ConcatenatedStringSet<const uint8*, const uint32*> in device memory
{ nvb_string_set r = { (const void*)s.base_string(), (const uint32_t*)s.offsets(), s.size(), 8u }; return r; }

/// match(fmi, pattern, len) for every string of a set (nvbio/fmindex/fmindex_inl.h:186-239; batch form filter_inl.h:261-291)
template <typename string_set_type>
inline void match(B200Context& c, const nvb_fmindex& f, const string_set_type& queries, uint2* ranges, cudaStream_t s = 0)
{ c.check( nvb_fm_match( c.ctx, &f, b200_set( queries ), (uint32_t*)ranges, s ) ); }
/// match_reverse (fmindex_inl.h:252-278)
template <typename string_set_type>
inline void match_reverse(B200Context& c, const nvb_fmindex& f, const string_set_type& queries, uint2* ranges, cudaStream_t s = 0)
{ c.check( nvb_fm_match_reverse( c.ctx, &f, b200_set( queries ), (uint32_t*)ranges, s ) ); }
/// locate(fmi, row) (fmindex_inl.h:364-394)
inline void locate(B200Context& c, const nvb_fmindex& f, const uint32* sa_rows, const uint32 n, uint32* pos, cudaStream_t s = 0)
{ c.check( nvb_fm_locate( c.ctx, &f, sa_rows, n, pos, s ) ); }

namespace aln {

/// the scheduler tag that selects the B200 path: batch_alignment_score( aligner, patterns, texts, sinks, DeviceB200Scheduler( c ), M, N )
struct DeviceB200Scheduler
{
    B200Context* c; cudaStream_t stream;
    explicit DeviceB200Scheduler(B200Context& _c, cudaStream_t _s = 0) : c( &_c ), stream( _s ) {}
};

// ---- aligner tag objects (alignment.h:365-519) with the simple schemes (utils.h:81-124, ed/ed_utils.h) -> nvb_aligner ---------------
template <AlignmentType T, typename A>
inline nvb_aligner b200_aligner(const EditDistanceAligner<T, A>&)
{ nvb_aligner a = { NVB_EDIT_DISTANCE, (int32_t)T, 0, -1, -1, -1 }; return a; }
template <AlignmentType T, typename A>
inline nvb_aligner b200_aligner(const SmithWatermanAligner<T, SimpleSmithWatermanScheme, A>& al)
{ nvb_aligner a = { NVB_SMITH_WATERMAN, (int32_t)T, al.scheme.m_match, al.scheme.m_mismatch, al.scheme.m_deletion, al.scheme.m_insertion }; return a; }
template <AlignmentType T, typename A>
inline nvb_aligner b200_aligner(const GotohAligner<T, SimpleGotohScheme, A>& al)
{ nvb_aligner a = { NVB_GOTOH, (int32_t)T, al.scheme.m_match, al.scheme.m_mismatch, al.scheme.m_gap_open, al.scheme.m_gap_ext }; return a; }

// BestSink<int32> { int32 score; uint2 sink; } (sink.h:60-75) is 16 bytes (uint2 is 8-byte aligned) and Alignment<int32> 24: the
// library's packed 12 / 20-byte records are staged and repacked by nvb_sinks_to_reference_layout / nvb_alignments_to_reference_layout
static_assert( sizeof(BestSink<int32>) == 16 && sizeof(Alignment<int32>) == 24, "the reference's padded result records" );
inline void b200_store_sinks(const DeviceB200Scheduler& sch, const nvb_best_sink* packed, const uint32 n, BestSink<int32>* sinks)
{ sch.c->check( nvb_sinks_to_reference_layout( sch.c->ctx, packed, n, sinks, sch.stream ) ); }

/// aln::batch_alignment_score with the B200 scheduler (same arguments as nvbio/alignment/batched.h:98-140)
template <typename aligner_type, typename pattern_set_type, typename text_set_type>
void batch_alignment_score(const aligner_type aligner, const pattern_set_type patterns, const text_set_type texts,
                           BestSink<int32>* sinks, const DeviceB200Scheduler sch, const uint32 max_pattern_length, const uint32 max_text_length)
{
    const nvb_aligner a = b200_aligner( aligner );
    // aligners carrying TextBlockingTag (aln::transpose( aligner ), alignment.h:266-279, :395-540) keep their own tie order
    const bool text_blocking = equal<typename aligner_type::algorithm_tag, TextBlockingTag>();
    const uint32 longest = max_pattern_length > max_text_length ? max_pattern_length : max_text_length;
    const size_t bytes = text_blocking ? nvb_temp_bytes( NVB_OP_SCORE, a, 0, 0, longest, longest, patterns.size() )
                                       : nvb_temp_bytes( NVB_OP_SCORE, a, 0, 0, max_pattern_length, max_text_length, patterns.size() );
    nvb_best_sink* packed = (nvb_best_sink*)sch.c->staging( sizeof(nvb_best_sink) * patterns.size() );
    if (text_blocking)
        sch.c->check( nvb_batch_score_text_blocking( sch.c->ctx, a, b200_set( patterns ), b200_set( texts ), packed,
                                                     max_pattern_length, max_text_length, sch.c->scratch( bytes ), bytes, sch.stream ) );
    else
        sch.c->check( nvb_batch_score( sch.c->ctx, a, b200_set( patterns ), b200_set( texts ), packed,
                                       max_pattern_length, max_text_length, sch.c->scratch( bytes ), bytes, sch.stream ) );
    b200_store_sinks( sch, packed, patterns.size(), sinks );
}
/// aln::batch_banded_alignment_score<BAND_LEN> (batched.h:142-192)
template <uint32 BAND_LEN, typename aligner_type, typename pattern_set_type, typename text_set_type>
void batch_banded_alignment_score(const aligner_type aligner, const pattern_set_type patterns, const text_set_type texts,
                                  BestSink<int32>* sinks, const DeviceB200Scheduler sch, const uint32 max_pattern_length, const uint32 max_text_length)
{
    nvb_best_sink* packed = (nvb_best_sink*)sch.c->staging( sizeof(nvb_best_sink) * patterns.size() );
    sch.c->check( nvb_batch_banded_score( sch.c->ctx, b200_aligner( aligner ), BAND_LEN, b200_set( patterns ), b200_set( texts ), packed,
                                          max_pattern_length, max_text_length, NULL, 0, sch.stream ) );
    b200_store_sinks( sch, packed, patterns.size(), sinks );
}
/// EditDistanceAligner<TYPE, MyersTag<ALPHABET>> (alignment.h:274-275; examples/fmmap/fmmap.cu:347-358) goes to the bit-vector kernel
template <uint32 BAND_LEN, AlignmentType T, uint32 ALPHABET, typename pattern_set_type, typename text_set_type>
void batch_banded_alignment_score(const EditDistanceAligner<T, MyersTag<ALPHABET> >, const pattern_set_type patterns, const text_set_type texts,
                                  BestSink<int32>* sinks, const DeviceB200Scheduler sch, const uint32, const uint32)
{
    nvb_best_sink* packed = (nvb_best_sink*)sch.c->staging( sizeof(nvb_best_sink) * patterns.size() );
    sch.c->check( nvb_batch_banded_myers_score( sch.c->ctx, (int32_t)T, ALPHABET, BAND_LEN, b200_set( patterns ), b200_set( texts ),
                                                NVB_MYERS_BATCH_MIN_SCORE, packed, sch.stream ) );
    b200_store_sinks( sch, packed, patterns.size(), sinks );
}
/// BatchedAlignmentTraceback<CHECKPOINTS,...>::enact (batched.h:335-380): alignments + the op list each Backtracer would have been
/// pushed (end -> start), ops_stride bytes per pair
template <uint32 CHECKPOINTS, typename aligner_type, typename pattern_set_type, typename text_set_type>
void batch_alignment_traceback(const aligner_type aligner, const pattern_set_type patterns, const text_set_type texts,
                               Alignment<int32>* alignments, uint8* ops, const uint32 ops_stride, uint32* n_ops,
                               const DeviceB200Scheduler sch, const uint32 max_pattern_length, const uint32 max_text_length)
{
    const nvb_aligner a = b200_aligner( aligner );
    const size_t bytes = nvb_temp_bytes( NVB_OP_TRACEBACK, a, 0, CHECKPOINTS, max_pattern_length, max_text_length, patterns.size() );
    nvb_alignment* packed = (nvb_alignment*)sch.c->staging( sizeof(nvb_alignment) * patterns.size() );
    sch.c->check( nvb_batch_traceback( sch.c->ctx, a, CHECKPOINTS, b200_set( patterns ), b200_set( texts ), packed, ops, ops_stride, n_ops,
                                       max_pattern_length, max_text_length, sch.c->scratch( bytes ), bytes, sch.stream ) );
    sch.c->check( nvb_alignments_to_reference_layout( sch.c->ctx, packed, patterns.size(), alignments, sch.stream ) );
}
/// BatchedBandedAlignmentTraceback<BAND_LEN,CHECKPOINTS,...>::enact (batched.h:382-438)
template <uint32 BAND_LEN, uint32 CHECKPOINTS, typename aligner_type, typename pattern_set_type, typename text_set_type>
void batch_banded_alignment_traceback(const aligner_type aligner, const pattern_set_type patterns, const text_set_type texts,
                                      Alignment<int32>* alignments, uint8* ops, const uint32 ops_stride, uint32* n_ops,
                                      const DeviceB200Scheduler sch, const uint32 max_pattern_length, const uint32 max_text_length)
{
    const nvb_aligner a = b200_aligner( aligner );
    const size_t bytes = nvb_temp_bytes( NVB_OP_BANDED_TRACEBACK, a, BAND_LEN, CHECKPOINTS, max_pattern_length, max_text_length, patterns.size() );
    nvb_alignment* packed = (nvb_alignment*)sch.c->staging( sizeof(nvb_alignment) * patterns.size() );
    sch.c->check( nvb_batch_banded_traceback( sch.c->ctx, a, BAND_LEN, CHECKPOINTS, b200_set( patterns ), b200_set( texts ), packed, ops, ops_stride,
                                              n_ops, max_pattern_length, max_text_length, sch.c->scratch( bytes ), bytes, sch.stream ) );
    sch.c->check( nvb_alignments_to_reference_layout( sch.c->ctx, packed, patterns.size(), alignments, sch.stream ) );
}

} // namespace aln
} // namespace nvbio
