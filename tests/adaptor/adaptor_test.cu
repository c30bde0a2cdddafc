/*
 * adaptor_test.cu -- TEST INFRASTRUCTURE: include/nvbio_b200_adaptor.h compiled against the reference's OWN headers
 * (/root/reference, never copied), driven with the reference's own types -- io::FMIndexDataCore over a device image,
 * ConcatenatedStringSet<const uint8*, const uint32*>, aln::make_*_aligner objects, BestSink<int32> / Alignment<int32> -- and
 * compared in the same binary with the reference's HOST results (oracle/_ref/libnvbio_ref.so: nvbio's header-only host code).
 * Built by oracle/Makefile (target adaptor) into oracle/_ref/adaptor_test; tests/test_gpu_adaptor.py runs it on the GPU box.
 */
#include <nvbio_b200_adaptor.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <random>

#define ORC_PREFIX ref_
#include "oracle_abi.h"

using namespace nvbio;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 2; } } while (0)

typedef ConcatenatedStringSet<const uint8*, const uint32*> string_set_type;

template <typename T> static T* to_device(const std::vector<T>& v)
{
    T* d = NULL;
    cudaMalloc( &d, (v.size() + 16) * sizeof(T) );
    cudaMemcpy( d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice );
    return d;
}

static void make_pairs(uint32 n, uint32 Mmax, uint32 Nfix, std::mt19937& rng, std::vector<uint8>& P, std::vector<uint32>& Po,
                       std::vector<uint8>& T, std::vector<uint32>& To, bool banded, uint32 band)
{
    P.clear(); T.clear(); Po.assign( 1, 0u ); To.assign( 1, 0u );
    for (uint32 i = 0; i < n; ++i)
    {
        const uint32 M = 1u + rng() % Mmax;
        const uint32 N = banded ? M + band - 1u + rng() % 6u : (Nfix ? Nfix : rng() % 200u);
        std::vector<uint8> t( N ), p( M );
        for (uint32 j = 0; j < N; ++j) t[j] = rng() & 3u;
        const uint32 s = N > M ? rng() % (N - M + 1u) : 0u;
        for (uint32 j = 0; j < M; ++j) p[j] = (s + j < N && (rng() % 100u) >= 8u) ? t[s + j] : (uint8)(rng() & 3u);
        P.insert( P.end(), p.begin(), p.end() ); T.insert( T.end(), t.begin(), t.end() );
        Po.push_back( (uint32)P.size() ); To.push_back( (uint32)T.size() );
    }
}

static int n_fail = 0;
static void report(const char* what, size_t bad, size_t n)
{
    printf( "%-58s %s (%zu of %zu differ)\n", what, bad ? "FAIL" : "ok", bad, n );
    if (bad) ++n_fail;
}

template <typename aligner_type>
static int check_scoring(B200Context& c, const char* name, const aligner_type aligner, const orc_aligner oa, std::mt19937& rng)
{
    std::vector<uint8> P, T; std::vector<uint32> Po, To;
    // full matrix, ragged
    make_pairs( 2001, 150, 0, rng, P, Po, T, To, false, 0 );
    const uint32 n = (uint32)Po.size() - 1u;
    uint8 *dP = to_device( P ), *dT = to_device( T ); uint32 *dPo = to_device( Po ), *dTo = to_device( To );
    aln::BestSink<int32>* d_sinks; CK( cudaMalloc( &d_sinks, n * sizeof(aln::BestSink<int32>) ) );
    const string_set_type patterns( n, dP, dPo ), texts( n, dT, dTo );
    aln::batch_alignment_score( aligner, patterns, texts, d_sinks, aln::DeviceB200Scheduler( c ), 150u, 200u );
    std::vector< aln::BestSink<int32> > got( n );
    CK( cudaMemcpy( got.data(), d_sinks, n * sizeof(aln::BestSink<int32>), cudaMemcpyDeviceToHost ) );
    std::vector<orc_sink> want( n );
    ref_score( oa, P.data(), Po.data(), T.data(), To.data(), n, want.data(), 0 );
    size_t bad = 0;
    for (uint32 i = 0; i < n; ++i) bad += (got[i].score != want[i].score || got[i].sink.x != want[i].x || got[i].sink.y != want[i].y);
    char label[128]; snprintf( label, sizeof(label), "batch_alignment_score             %s", name ); report( label, bad, n );

    // the same aligner with TextBlockingTag (aln::transpose, alignment.h:395-540): its own tie order, same call shape
    aln::batch_alignment_score( aln::transpose( aligner ), patterns, texts, d_sinks, aln::DeviceB200Scheduler( c ), 150u, 200u );
    CK( cudaMemcpy( got.data(), d_sinks, n * sizeof(aln::BestSink<int32>), cudaMemcpyDeviceToHost ) );
    ref_score_text_blocking( oa, P.data(), Po.data(), T.data(), To.data(), n, want.data(), 0 );
    bad = 0;
    for (uint32 i = 0; i < n; ++i) bad += (got[i].score != want[i].score || got[i].sink.x != want[i].x || got[i].sink.y != want[i].y);
    snprintf( label, sizeof(label), "batch_alignment_score TextBlocking %s", name ); report( label, bad, n );

    // traceback<64>
    aln::Alignment<int32>* d_aln; uint8* d_ops; uint32* d_nops; const uint32 stride = 150u + 200u + 8u;
    CK( cudaMalloc( &d_aln, n * sizeof(aln::Alignment<int32>) ) ); CK( cudaMalloc( &d_ops, (size_t)n * stride ) ); CK( cudaMalloc( &d_nops, n * 4 ) );
    aln::batch_alignment_traceback<64u>( aligner, patterns, texts, d_aln, d_ops, stride, d_nops, aln::DeviceB200Scheduler( c ), 150u, 200u );
    std::vector< aln::Alignment<int32> > galn( n ); std::vector<uint8> gops( (size_t)n * stride ); std::vector<uint32> gn( n );
    CK( cudaMemcpy( galn.data(), d_aln, n * sizeof(aln::Alignment<int32>), cudaMemcpyDeviceToHost ) );
    CK( cudaMemcpy( gops.data(), d_ops, gops.size(), cudaMemcpyDeviceToHost ) ); CK( cudaMemcpy( gn.data(), d_nops, n * 4, cudaMemcpyDeviceToHost ) );
    std::vector<orc_alignment> waln( n ); std::vector<uint8> wops( (size_t)n * stride ); std::vector<uint32> wn( n );
    ref_traceback( oa, 64u, P.data(), Po.data(), T.data(), To.data(), n, waln.data(), wops.data(), stride, wn.data(), 0 );
    bad = 0;
    for (uint32 i = 0; i < n; ++i)
    {
        bool b = galn[i].score != waln[i].score || galn[i].source.x != waln[i].src_x || galn[i].source.y != waln[i].src_y ||
                 galn[i].sink.x != waln[i].sink_x || galn[i].sink.y != waln[i].sink_y || gn[i] != wn[i];
        for (uint32 k = 0; !b && k < wn[i] && k < stride; ++k) b = gops[(size_t)i * stride + k] != wops[(size_t)i * stride + k];
        bad += b;
    }
    snprintf( label, sizeof(label), "batch_alignment_traceback<64>     %s", name ); report( label, bad, n );
    cudaFree( dP ); cudaFree( dT ); cudaFree( dPo ); cudaFree( dTo ); cudaFree( d_ops );

    // banded <15>
    make_pairs( 2001, 150, 0, rng, P, Po, T, To, true, 15 );
    dP = to_device( P ); dT = to_device( T ); dPo = to_device( Po ); dTo = to_device( To );
    const string_set_type bp( n, dP, dPo ), bt( n, dT, dTo );
    aln::batch_banded_alignment_score<15u>( aligner, bp, bt, d_sinks, aln::DeviceB200Scheduler( c ), 150u, 170u );
    CK( cudaMemcpy( got.data(), d_sinks, n * sizeof(aln::BestSink<int32>), cudaMemcpyDeviceToHost ) );
    ref_banded_score( oa, 15u, P.data(), Po.data(), T.data(), To.data(), n, want.data(), 0 );
    bad = 0;
    for (uint32 i = 0; i < n; ++i) bad += (got[i].score != want[i].score || got[i].sink.x != want[i].x || got[i].sink.y != want[i].y);
    snprintf( label, sizeof(label), "batch_banded_alignment_score<15>  %s", name ); report( label, bad, n );
    const uint32 bstride = 2u * 150u + 15u + 8u + 13u;      // a multiple of 16
    CK( cudaMalloc( &d_ops, (size_t)n * bstride ) );
    aln::batch_banded_alignment_traceback<15u, 16u>( aligner, bp, bt, d_aln, d_ops, bstride, d_nops, aln::DeviceB200Scheduler( c ), 150u, 170u );
    gops.assign( (size_t)n * bstride, 0 ); wops.assign( (size_t)n * bstride, 0 );
    CK( cudaMemcpy( galn.data(), d_aln, n * sizeof(aln::Alignment<int32>), cudaMemcpyDeviceToHost ) );
    CK( cudaMemcpy( gops.data(), d_ops, gops.size(), cudaMemcpyDeviceToHost ) ); CK( cudaMemcpy( gn.data(), d_nops, n * 4, cudaMemcpyDeviceToHost ) );
    ref_banded_traceback( oa, 15u, 16u, P.data(), Po.data(), T.data(), To.data(), n, waln.data(), wops.data(), bstride, wn.data(), 0 );
    bad = 0;
    for (uint32 i = 0; i < n; ++i)
    {
        bool b = galn[i].score != waln[i].score || galn[i].source.x != waln[i].src_x || galn[i].source.y != waln[i].src_y ||
                 galn[i].sink.x != waln[i].sink_x || galn[i].sink.y != waln[i].sink_y || gn[i] != wn[i];
        for (uint32 k = 0; !b && k < wn[i] && k < bstride; ++k) b = gops[(size_t)i * bstride + k] != wops[(size_t)i * bstride + k];
        bad += b;
    }
    snprintf( label, sizeof(label), "batch_banded_alignment_traceback  %s", name ); report( label, bad, n );
    cudaFree( dP ); cudaFree( dT ); cudaFree( dPo ); cudaFree( dTo ); cudaFree( d_ops ); cudaFree( d_aln ); cudaFree( d_nops ); cudaFree( d_sinks );
    return 0;
}

int main()
{
    std::mt19937 rng( 12345 );
    B200Context c( 0 );

    // ---- seeding: an index built by the reference's host code, held the way io::FMIndexDataDevice holds it
    const uint32 n = 200000;
    std::vector<uint8> text( n );
    for (uint32 i = 0; i < n; ++i) text[i] = rng() & 3u;
    std::vector<uint32> bwt_occ( ref_bwt_occ_words( n ) ), ssa( ref_ssa_words( n ) ), L2( 5 );
    uint32 primary = 0;
    ref_build_index( text.data(), n, &primary, L2.data(), bwt_occ.data(), ssa.data(), NULL );
    io::FMIndexDataCore d;                                               // the fields FMIndexDataDevice fills (fmindex_impl.cu:740-826)
    d.m_seq_length = n; d.m_primary = primary; d.m_rprimary = 0;
    d.m_L2 = to_device( L2 ); d.m_bwt_occ = to_device( bwt_occ ); d.m_rbwt_occ = NULL;
    d.m_bwt_occ_words = (uint32)bwt_occ.size(); d.m_sa_words = (uint32)ssa.size();
    d.m_ssa = io::FMIndexDataCore::ssa_type( to_device( ssa ) );
    const nvb_fmindex f = b200_view( d );

    const uint32 nq = 300000;                                            // large enough for the per-batch k-mer table path
    std::vector<uint8> Q; std::vector<uint32> qo( 1, 0u );
    for (uint32 i = 0; i < nq; ++i)
    {
        const uint32 len = rng() % 30u, s = rng() % (n - 32u);
        for (uint32 j = 0; j < len; ++j) Q.push_back( (rng() % 10u) ? text[s + j] : (uint8)(rng() % 5u) );
        qo.push_back( (uint32)Q.size() );
    }
    uint8* dQ = to_device( Q ); uint32* dqo = to_device( qo );
    uint2* d_ranges; CK( cudaMalloc( &d_ranges, nq * sizeof(uint2) ) );
    const string_set_type queries( nq, dQ, dqo );
    match( c, f, queries, d_ranges );
    std::vector<uint2> got( nq ); CK( cudaMemcpy( got.data(), d_ranges, nq * sizeof(uint2), cudaMemcpyDeviceToHost ) );
    orc_fmindex of = { n, primary, { L2[0], L2[1], L2[2], L2[3], L2[4] }, bwt_occ.data(), ssa.data() };
    std::vector<uint32> want( 2 * (size_t)nq );
    ref_fm_match( &of, Q.data(), qo.data(), nq, want.data(), NULL, 0 );
    size_t bad = 0;
    for (uint32 i = 0; i < nq; ++i) bad += (got[i].x != want[2 * i] || got[i].y != want[2 * i + 1]);
    report( "match( fm_index, ConcatenatedStringSet )", bad, nq );
    match_reverse( c, f, queries, d_ranges );
    CK( cudaMemcpy( got.data(), d_ranges, nq * sizeof(uint2), cudaMemcpyDeviceToHost ) );
    ref_fm_match_reverse( &of, Q.data(), qo.data(), nq, want.data(), 0 );
    bad = 0;
    for (uint32 i = 0; i < nq; ++i) bad += (got[i].x != want[2 * i] || got[i].y != want[2 * i + 1]);
    report( "match_reverse( fm_index, ConcatenatedStringSet )", bad, nq );
    std::vector<uint32> rows;
    ref_fm_match( &of, Q.data(), qo.data(), nq, want.data(), NULL, 0 );
    for (uint32 i = 0; i < nq && rows.size() < 100000; ++i) if (want[2 * i] <= want[2 * i + 1]) rows.push_back( want[2 * i] );
    uint32* d_rows = to_device( rows ); uint32* d_pos; CK( cudaMalloc( &d_pos, rows.size() * 4 ) );
    locate( c, f, d_rows, (uint32)rows.size(), d_pos );
    std::vector<uint32> gpos( rows.size() ), wpos( rows.size() );
    CK( cudaMemcpy( gpos.data(), d_pos, rows.size() * 4, cudaMemcpyDeviceToHost ) );
    ref_fm_locate( &of, rows.data(), (uint32)rows.size(), wpos.data(), NULL, 0 );
    bad = 0;
    for (size_t i = 0; i < rows.size(); ++i) bad += gpos[i] != wpos[i];
    report( "locate( fm_index, row )", bad, rows.size() );

    // ---- extension: the reference's aligner objects through the reference's call shapes
    {
        const orc_aligner oa = { 1, 1, 2, -1, -1, -1 };
        if (check_scoring( c, "SmithWatermanAligner<LOCAL>", aln::make_smith_waterman_aligner<aln::LOCAL>( aln::SimpleSmithWatermanScheme( 2, -1, -1, -1 ) ), oa, rng )) return 2;
    }
    {
        const orc_aligner oa = { 1, 2, 2, -1, -1, -1 };
        if (check_scoring( c, "SmithWatermanAligner<SEMI_GLOBAL>", aln::make_smith_waterman_aligner<aln::SEMI_GLOBAL>( aln::SimpleSmithWatermanScheme( 2, -1, -1, -1 ) ), oa, rng )) return 2;
    }
    {
        const orc_aligner oa = { 2, 2, 0, -5, -8, -3 };
        if (check_scoring( c, "GotohAligner<SEMI_GLOBAL>", aln::make_gotoh_aligner<aln::SEMI_GLOBAL>( aln::SimpleGotohScheme( 0, -5, -8, -3 ) ), oa, rng )) return 2;
    }
    {
        const orc_aligner oa = { 2, 1, 2, -2, -5, -3 };
        if (check_scoring( c, "GotohAligner<LOCAL>", aln::make_gotoh_aligner<aln::LOCAL>( aln::SimpleGotohScheme( 2, -2, -5, -3 ) ), oa, rng )) return 2;
    }
    {
        const orc_aligner oa = { 0, 0, 0, -1, -1, -1 };
        if (check_scoring( c, "EditDistanceAligner<GLOBAL>", aln::make_edit_distance_aligner<aln::GLOBAL>(), oa, rng )) return 2;
    }
    CK( cudaDeviceSynchronize() );
    printf( "%s\n", n_fail ? "ADAPTOR TEST FAILED" : "adaptor test ok: every result equals the reference's host code" );
    return n_fail ? 1 : 0;
}
