// ref_wrap_aln.inl -- alignment half of ref_wrap.cpp (TEST INFRASTRUCTURE ONLY).
// Calls the reference's per-pair host functions:
//   aln::alignment_score                 nvbio/alignment/alignment_inl.h:55-73
//   aln::banded_alignment_score<BAND>    nvbio/alignment/banded_inl.h
//   aln::alignment_traceback<CHK>        nvbio/alignment/alignment_inl.h:355-455
//   aln::banded_alignment_traceback<BAND,CHK>  nvbio/alignment/banded_inl.h:354-430
// with min_score = INT_MIN exactly as priv::AlignmentStream::init_context does
// (nvbio/alignment/batched_inl.h:924-933).

namespace {

typedef vector_view<const uint8*> sym_string;

struct RecBacktracer
{
    uint8* ops; uint32 n; uint32 cap;
    void clip(const uint32) {}
    void push(const uint8 op) { if (n < cap) ops[n] = op; ++n; }
};

const int32 MIN_SCORE = Field_traits<int32>::min();

template <typename aligner_type>
void score_batch(const aligner_type aligner, const uint8_t* pat, const uint32_t* pat_off,
                 const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    typedef typename aln::column_storage_type<aligner_type>::type cell_type;
    #pragma omp parallel num_threads(threads)
    {
        std::vector<cell_type> column;
        #pragma omp for schedule(dynamic,64)
        for (int64 i = 0; i < int64(count); ++i)
        {
            const uint32 M = pat_off[i+1] - pat_off[i];
            const uint32 N = txt_off[i+1] - txt_off[i];
            if (column.size() < N + 1u) column.resize( N + 1u );
            aln::BestSink<int32> sink;
            aln::alignment_score( aligner, sym_string( M, pat + pat_off[i] ), aln::trivial_quality_string(),
                                  sym_string( N, txt + txt_off[i] ), MIN_SCORE, sink, &column[0] );
            sinks[i].score = sink.score; sinks[i].x = sink.sink.x; sinks[i].y = sink.sink.y;
        }
    }
}

template <uint32 BAND, typename aligner_type>
void banded_score_batch(const aligner_type aligner, const uint8_t* pat, const uint32_t* pat_off,
                        const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    #pragma omp parallel for schedule(dynamic,64) num_threads(threads)
    for (int64 i = 0; i < int64(count); ++i)
    {
        const uint32 M = pat_off[i+1] - pat_off[i];
        const uint32 N = txt_off[i+1] - txt_off[i];
        aln::BestSink<int32> sink;
        aln::banded_alignment_score<BAND>( aligner, sym_string( M, pat + pat_off[i] ), aln::trivial_quality_string(),
                                           sym_string( N, txt + txt_off[i] ), MIN_SCORE, sink );
        sinks[i].score = sink.score; sinks[i].x = sink.sink.x; sinks[i].y = sink.sink.y;
    }
}

inline void store_aln(orc_alignment* out, const aln::Alignment<int32>& a)
{
    out->score = a.score;
    out->src_x = a.source.x; out->src_y = a.source.y;
    out->sink_x = a.sink.x;  out->sink_y = a.sink.y;
}

template <uint32 CHK, typename aligner_type>
void traceback_batch(const aligner_type aligner, const uint8_t* pat, const uint32_t* pat_off,
                     const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                     uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{
    typedef typename aln::column_storage_type<aligner_type>::type     cell_type;
    typedef typename aln::checkpoint_storage_type<aligner_type>::type chk_type;
    const uint32 BITS = aln::direction_vector_traits<aligner_type>::BITS;
    #pragma omp parallel num_threads(threads)
    {
        std::vector<cell_type> column;
        std::vector<chk_type>  checkpoints;
        std::vector<uint32>    submatrix_words;
        #pragma omp for schedule(dynamic,16)
        for (int64 i = 0; i < int64(count); ++i)
        {
            const uint32 M = pat_off[i+1] - pat_off[i];
            const uint32 N = txt_off[i+1] - txt_off[i];
            column.assign( N + 1u, cell_type() );
            checkpoints.assign( size_t(N + 1u) * ((M + CHK - 1u) / CHK + 1u), chk_type() );
            submatrix_words.assign( (size_t(N) * CHK * BITS + 31u) / 32u + 1u, 0u );
            PackedStream<uint32*,uint8,BITS,false> submatrix( &submatrix_words[0] );
            RecBacktracer bt = { ops + size_t(i) * ops_stride, 0u, ops_stride };
            const aln::Alignment<int32> a = aln::alignment_traceback<CHK>(
                aligner, sym_string( M, pat + pat_off[i] ), aln::trivial_quality_string(),
                sym_string( N, txt + txt_off[i] ), MIN_SCORE, bt, &checkpoints[0], submatrix, &column[0] );
            store_aln( out + i, a );
            n_ops[i] = bt.n;
        }
    }
}

template <uint32 BAND, uint32 CHK, typename aligner_type>
void banded_traceback_batch(const aligner_type aligner, const uint8_t* pat, const uint32_t* pat_off,
                            const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                            uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{
    typedef typename aln::checkpoint_storage_type<aligner_type>::type chk_type;
    const uint32 BITS = aln::direction_vector_traits<aligner_type>::BITS;
    #pragma omp parallel num_threads(threads)
    {
        std::vector<chk_type>  checkpoints;
        std::vector<uint32>    submatrix_words;
        #pragma omp for schedule(dynamic,16)
        for (int64 i = 0; i < int64(count); ++i)
        {
            const uint32 M = pat_off[i+1] - pat_off[i];
            const uint32 N = txt_off[i+1] - txt_off[i];
            checkpoints.assign( size_t(BAND) * (2u + (M + CHK - 1u) / CHK), chk_type() );
            submatrix_words.assign( (size_t(BAND) * CHK * BITS + 31u) / 32u + 1u, 0u );
            PackedStream<uint32*,uint8,BITS,false> submatrix( &submatrix_words[0] );
            RecBacktracer bt = { ops + size_t(i) * ops_stride, 0u, ops_stride };
            const aln::Alignment<int32> a = aln::banded_alignment_traceback<BAND,CHK>(
                aligner, sym_string( M, pat + pat_off[i] ), aln::trivial_quality_string(),
                sym_string( N, txt + txt_off[i] ), MIN_SCORE, bt, &checkpoints[0], submatrix );
            store_aln( out + i, a );
            n_ops[i] = bt.n;
        }
    }
}

// ---- runtime -> compile-time dispatch ---------------------------------------------------------

template <typename F>
int dispatch_aligner(const orc_aligner a, F f)
{
    using namespace aln;
    #define NVB_TYPES(MAKE) \
        switch (a.type) { \
        case 0: f( MAKE(GLOBAL) );      return 0; \
        case 1: f( MAKE(LOCAL) );       return 0; \
        case 2: f( MAKE(SEMI_GLOBAL) ); return 0; \
        default: return -1; }
    if (a.kind == 0)
    {
        #define NVB_MK_ED(T) make_edit_distance_aligner<T>()
        NVB_TYPES(NVB_MK_ED)
    }
    else if (a.kind == 1)
    {
        const SimpleSmithWatermanScheme s( a.p0, a.p1, a.p2, a.p3 );
        #define NVB_MK_SW(T) make_smith_waterman_aligner<T>( s )
        NVB_TYPES(NVB_MK_SW)
    }
    else if (a.kind == 2)
    {
        const SimpleGotohScheme s( a.p0, a.p1, a.p2, a.p3 );
        #define NVB_MK_GT(T) make_gotoh_aligner<T>( s )
        NVB_TYPES(NVB_MK_GT)
    }
    return -1;
}

struct Args
{
    const uint8_t* pat; const uint32_t* pat_off; const uint8_t* txt; const uint32_t* txt_off;
    uint32_t count; orc_sink* sinks; orc_alignment* out; uint8_t* ops; uint32_t ops_stride; uint32_t* n_ops;
    int threads;
};

struct ScoreF { Args g;
    template <typename A> void operator()(const A a) const
    { score_batch( a, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); } };

template <uint32 BAND> struct BandedScoreF { Args g;
    template <typename A> void operator()(const A a) const
    { banded_score_batch<BAND>( a, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); } };

template <uint32 CHK> struct TracebackF { Args g;
    template <typename A> void operator()(const A a) const
    { traceback_batch<CHK>( a, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.out, g.ops, g.ops_stride, g.n_ops, g.threads ); } };

template <uint32 BAND, uint32 CHK> struct BandedTracebackF { Args g;
    template <typename A> void operator()(const A a) const
    { banded_traceback_batch<BAND,CHK>( a, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.out, g.ops, g.ops_stride, g.n_ops, g.threads ); } };

template <uint32 BAND>
int banded_traceback_chk(const orc_aligner a, const uint32 chk, const Args& g)
{
    switch (chk) {
    case 16: { BandedTracebackF<BAND,16> f = { g }; return dispatch_aligner( a, f ); }
    case 32: { BandedTracebackF<BAND,32> f = { g }; return dispatch_aligner( a, f ); }
    case 64: { BandedTracebackF<BAND,64> f = { g }; return dispatch_aligner( a, f ); }
    default: return -2; }
}

} // anonymous namespace

extern "C" {

int ref_score(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
              const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, sinks, 0, 0, 0, 0, threads };
    ScoreF f = { g };
    return dispatch_aligner( a, f );
}

int ref_banded_score(orc_aligner a, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                     const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, sinks, 0, 0, 0, 0, threads };
    switch (band) {
    case 3:  { BandedScoreF<3>  f = { g }; return dispatch_aligner( a, f ); }
    case 5:  { BandedScoreF<5>  f = { g }; return dispatch_aligner( a, f ); }
    case 7:  { BandedScoreF<7>  f = { g }; return dispatch_aligner( a, f ); }
    case 15: { BandedScoreF<15> f = { g }; return dispatch_aligner( a, f ); }
    case 31: { BandedScoreF<31> f = { g }; return dispatch_aligner( a, f ); }
    default: return -2; }
}

int ref_traceback(orc_aligner a, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                  const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                  uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, 0, out, ops, ops_stride, n_ops, threads };
    switch (checkpoints) {
    case 16: { TracebackF<16> f = { g }; return dispatch_aligner( a, f ); }
    case 32: { TracebackF<32> f = { g }; return dispatch_aligner( a, f ); }
    case 64: { TracebackF<64> f = { g }; return dispatch_aligner( a, f ); }
    default: return -2; }
}

int ref_banded_traceback(orc_aligner a, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                         const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                         uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                         uint32_t* n_ops, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, 0, out, ops, ops_stride, n_ops, threads };
    switch (band) {
    case 3:  return banded_traceback_chk<3>( a, checkpoints, g );
    case 5:  return banded_traceback_chk<5>( a, checkpoints, g );
    case 7:  return banded_traceback_chk<7>( a, checkpoints, g );
    case 15: return banded_traceback_chk<15>( a, checkpoints, g );
    case 31: return banded_traceback_chk<31>( a, checkpoints, g );
    default: return -2; }
}

} // extern "C"
