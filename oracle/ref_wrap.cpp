/*
 * ref_wrap.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * Thin extern "C" wrapper that instantiates the reference's OWN header-only host code
 * (from /root/reference, never copied into this repo) behind oracle_abi.h with prefix ref_.
 * Built by oracle/Makefile into oracle/_ref/libnvbio_ref.so. It is used
 *   (1) to pin oracle/oracle.c (the plain-C restatement) on randomised inputs,
 *   (2) to generate tests/golden fixtures (tests/golden/make_golden.py),
 *   (3) as the timed CPU baseline (bench.py cpu_baseline / --impl reference, kind "reference").
 *
 * Per-item reference calls are made from OUR OWN omp loop with private scratch per thread:
 * the reference's HostThreadScheduler is racy with >1 thread
 * (nvbio/alignment/batched_inl.h:274-276,302 strides columns by the pattern length).
 */
#include <nvbio/basic/types.h>
#include <nvbio/basic/packedstream.h>
#include <nvbio/basic/deinterleaved_iterator.h>
#include <nvbio/basic/vector_view.h>
#include <nvbio/fmindex/bwt.h>
#include <nvbio/fmindex/rank_dictionary.h>
#include <nvbio/fmindex/ssa.h>
#include <nvbio/fmindex/fmindex.h>
#include <nvbio/alignment/alignment.h>
#include <nvbio/alignment/sink.h>
#include <vector>
#include <cstring>
#include <omp.h>

#define ORC_PREFIX ref_
#include "oracle_abi.h"

using namespace nvbio;

namespace {

// the exact typedefs of io::FMIndexData (nvbio/io/fmindex/fmindex.h:150-163)
typedef const uint4*                                   bwt_occ_type;
typedef deinterleaved_iterator<2,0,bwt_occ_type>       bwt_type;
typedef deinterleaved_iterator<2,1,bwt_occ_type>       occ_type;
typedef PackedStream<bwt_type,uint8,2u,true>           bwt_stream_type;
typedef rank_dictionary<2u,64u,bwt_stream_type,occ_type,const uint32*> rank_dict_type;
typedef SSA_index_multiple_context<16u,const uint32*>  ssa_type;
typedef fm_index<rank_dict_type,ssa_type>              fm_index_type;

static uint32 g_count_table[256];
static bool   g_count_table_init = false;

fm_index_type make_fmi(const orc_fmindex* f)
{
    if (!g_count_table_init)
    {
        #pragma omp critical
        { gen_bwt_count_table( g_count_table ); g_count_table_init = true; }
    }
    const bwt_occ_type base = (bwt_occ_type)f->bwt_occ;
    return fm_index_type(
        f->length, f->primary, f->L2,
        rank_dict_type( bwt_stream_type( bwt_type( base ) ), occ_type( base ), g_count_table ),
        ssa_type( f->ssa ) );
}

inline uint32 words_for(const uint32 n) { return ((n + 15u) / 16u + 3u) & ~3u; }

} // anonymous namespace

extern "C" {

size_t ref_bwt_occ_words(uint32_t n) { return size_t( words_for(n) ) * 2u; }
size_t ref_ssa_words(uint32_t n)     { return (size_t(n) + 16u) / 16u; }

// follows nvbio-test/fmindex_test.cu:438-539 and nvbio/io/fmindex/fmindex_impl.cu:296-318
int ref_build_index(const uint8_t* text_sym, uint32_t n, uint32_t* primary, uint32_t* L2,
                    uint32_t* bwt_occ, uint32_t* ssa, int32_t* sa_out)
{
    const uint32 W = words_for( n );
    std::vector<uint32> text_words( W + 4u, 0u ), bwt_words( W + 4u, 0u ), occ( W, 0u ); // +4: gen_bwt_from_sa writes slot n
    typedef PackedStream<uint32*,uint8,2,true> stream_type;
    stream_type text( &text_words[0] );
    for (uint32 i = 0; i < n; ++i) text[i] = text_sym[i] & 3u;

    std::vector<int32> sa_vec;
    int32* sa = sa_out;
    if (!sa) { sa_vec.resize( size_t(n) + 1u ); sa = &sa_vec[0]; }

    gen_sa( n, text, sa );
    stream_type bwt( &bwt_words[0] );
    *primary = gen_bwt_from_sa( n, text, sa, bwt );
    // gen_bwt_from_sa leaves a stale symbol at slot n when n is not the last slot: clear padding
    for (uint32 i = n; i < (W + 4u)*16u; ++i) bwt[i] = 0;

    uint32 cnt[4];
    build_occurrence_table<64u>( bwt, bwt + n, &occ[0], cnt );
    L2[0] = 0;
    for (uint32 c = 0; c < 4; ++c) L2[c+1] = L2[c] + cnt[c];

    for (uint32 w = 0; w < W; w += 4)
    {
        for (uint32 k = 0; k < 4; ++k)
        {
            bwt_occ[ w*2+k   ] = bwt_words[ w+k ];
            bwt_occ[ w*2+4+k ] = occ[ w+k ];
        }
    }
    // SSA_index_multiple<K>(n, sa) : ssa[i] = sa[i*K]  (nvbio/fmindex/ssa_inl.h:255-268), ssa[0] = -1
    const uint32 n_items = (n + 16u) / 16u;
    for (uint32 i = 0; i < n_items; ++i) ssa[i] = uint32( sa[i*16u] );
    ssa[0] = uint32(-1);
    if (sa_out) sa_out[0] = -1;
    return 0;
}

int ref_fm_match(const orc_fmindex* f, const uint8_t* symbols, const uint32_t* offsets,
                 uint32_t count, uint32_t* ranges, uint32_t* blocks, int threads)
{
    const fm_index_type fmi = make_fmi( f );
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64 q = 0; q < int64(count); ++q)
    {
        const uint8* p   = symbols + offsets[q];
        const uint32 len = offsets[q+1] - offsets[q];
        const uint2 r = match( fmi, p, len );
        ranges[2*q] = r.x; ranges[2*q+1] = r.y;
        if (blocks) blocks[q] = 0; // only the C restatement counts blocks
    }
    return 0;
}

int ref_fm_match_reverse(const orc_fmindex* f, const uint8_t* symbols, const uint32_t* offsets,
                         uint32_t count, uint32_t* ranges, int threads)
{
    const fm_index_type fmi = make_fmi( f );
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64 q = 0; q < int64(count); ++q)
    {
        const uint2 r = match_reverse( fmi, symbols + offsets[q], offsets[q+1] - offsets[q] );
        ranges[2*q] = r.x; ranges[2*q+1] = r.y;
    }
    return 0;
}

int ref_fm_rank(const orc_fmindex* f, const uint32_t* ranges, const uint8_t* c, uint32_t count,
                uint32_t* out, int threads)
{
    const fm_index_type fmi = make_fmi( f );
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64 q = 0; q < int64(count); ++q)
    {
        const uint2 r = rank( fmi, make_uint2( ranges[2*q], ranges[2*q+1] ), c[q] );
        out[2*q] = r.x; out[2*q+1] = r.y;
    }
    return 0;
}

int ref_fm_rank4(const orc_fmindex* f, const uint32_t* k, uint32_t count, uint32_t* out, int threads)
{
    const fm_index_type fmi = make_fmi( f );
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64 q = 0; q < int64(count); ++q)
    {
        const uint4 r = rank4( fmi, k[q] );
        out[4*q] = r.x; out[4*q+1] = r.y; out[4*q+2] = r.z; out[4*q+3] = r.w;
    }
    return 0;
}

int ref_fm_locate(const orc_fmindex* f, const uint32_t* rows, uint32_t count, uint32_t* pos,
                  uint32_t* steps, int threads)
{
    const fm_index_type fmi = make_fmi( f );
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64 q = 0; q < int64(count); ++q)
    {
        pos[q] = locate( fmi, rows[q] );
        if (steps) steps[q] = inv_psi( fmi, rows[q] ).y;
    }
    return 0;
}

} // extern "C"

#include "ref_wrap_aln.inl"

// ---------------------------------------------------------------------------------------------
// nvBowtie's seed-hit deque, as its mapping kernels use it: SeedHit + hit_compare (nvBowtie/bowtie2/cuda/seed_hit.h:41-230) in
// a priority_deque over a vector_view (seed_hit_deque_array.h:144-145; built like mapping_inl.h:441-444), pushes that drop the
// bottom of a full deque (mapping_inl.h:158-160), then pop_top until empty (select_inl.h:93-113 visits hits in this order).
// The hit's 10-bit pos field carries the caller's slot id. Pins oracle_map.c's hit-priority rule wherever sizes are distinct.
// ---------------------------------------------------------------------------------------------
#include <nvBowtie/bowtie2/cuda/seed_hit.h>
#include <nvbio/basic/priority_deque.h>
extern "C" int ref_hit_order(const uint32_t* sizes, uint32_t n, uint32_t cap, uint32_t* out_slots, uint32_t* out_sizes, uint32_t* n_out)
{
    using namespace nvbio::bowtie2::cuda;
    if (n >= 1024u) return -1;
    std::vector<SeedHit> storage(n + 1u);
    typedef nvbio::vector_view<SeedHit*> hit_vector_type;
    typedef nvbio::priority_deque<SeedHit, hit_vector_type, hit_compare> hit_deque_type;
    hit_deque_type heap(hit_vector_type(0, &storage[0]), true);
    for (uint32_t i = 0; i < n; ++i)
    {
        if (sizes[i] == 0 || sizes[i] >= (1u << 20)) continue;
        if (cap && heap.size() == cap) heap.pop_bottom();
        heap.push(SeedHit(SeedHit::build_flags(STANDARD, FORWARD, i), make_uint2(1000u, 1000u + sizes[i])));   // exclusive range: delta = size
    }
    uint32_t k = 0;
    while (heap.size())
    {
        const SeedHit h = heap.top();
        out_slots[k] = h.get_posinread(); out_sizes[k] = h.get_range_size(); ++k;
        heap.pop_top();
    }
    *n_out = k;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// nvBowtie's mapping quality: BowtieMapq2 (nvBowtie/bowtie2/cuda/mapq.h:132-316) instantiated over a scheme that returns the
// caller's perfect / minimum scores (the calculator only asks its scheme for perfect_score(len), min_score(len) and m_monotone),
// fed an io::BestPairedAlignments built the way the driver builds it (io/alignments.h:68-230).
// ---------------------------------------------------------------------------------------------
#include <nvBowtie/bowtie2/cuda/mapq.h>
namespace {
struct PinnedScheme
{
    int32 perfect, o_perfect, minsc, o_minsc; uint32 len; bool m_monotone;
    int32 perfect_score(const uint32 l) const { return l == len ? perfect : o_perfect; }
    int32 min_score(const uint32 l)     const { return l == len ? minsc : o_minsc; }
};
}
extern "C" int ref_bowtie_mapq2(int32_t best, int32_t has_second, int32_t second, int32_t perfect, int32_t min_score, int32_t monotone,
                           int32_t paired, int32_t o_best, int32_t o_perfect, int32_t o_min_score)
{
    using namespace nvbio;
    PinnedScheme sc; sc.perfect = perfect; sc.o_perfect = o_perfect; sc.minsc = min_score; sc.o_minsc = o_min_score; sc.len = 100; sc.m_monotone = monotone != 0;
    const io::Alignment a1(1000u, 0u, best, 0u, 0u, paired != 0);
    const io::Alignment a2 = has_second ? io::Alignment(2000u, 0u, second, 0u, 0u, false) : io::Alignment::invalid();
    const io::Alignment o1 = paired ? io::Alignment(1200u, 0u, o_best, 1u, 1u, true) : io::Alignment::invalid();
    const io::BestPairedAlignments bp(io::BestAlignments(a1, a2), io::BestAlignments(o1, io::Alignment::invalid()));
    nvbio::bowtie2::cuda::BowtieMapq2<PinnedScheme> mq(sc);
    return mq(bp, 100u, 101u);
}
