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

// pattern qualities of pair i: trivial_quality_string (all 0) or a view of the caller's bytes
struct NoQuals   { typedef aln::trivial_quality_string type; type get(int64, uint32) const { return type(); } };
struct ByteQuals { typedef sym_string type; const uint8_t* q; const uint32_t* off; type get(int64 i, uint32 M) const { return sym_string( M, q + off[i] ); } };

// table-driven models of the reference's quality-aware scheme interfaces (alignment.h:411-436, 488-505): what nvBowtie's
// SmithWatermanScoringScheme (nvBowtie/bowtie2/cuda/scoring.h:279-286) evaluates, with the functors tabulated over q
struct TableSWScheme
{
    const int32* m; const int32* x; int32 del, ins;
    int32 match(const uint8 q = 0) const { return m[q]; }
    int32 mismatch(const uint8 q = 0) const { return x[q]; }
    int32 mismatch(const uint8 a, const uint8 b, const uint8 q = 0) const { return x[q]; }
    int32 deletion()  const { return del; }
    int32 insertion() const { return ins; }
};
struct TableGotohScheme
{
    const int32* m; const int32* x; int32 pgo, pge, tgo, tge;
    int32 match(const uint8 q = 0) const { return m[q]; }
    int32 mismatch(const uint8 q = 0) const { return x[q]; }
    int32 mismatch(const uint8 a, const uint8 b, const uint8 q = 0) const { return x[q]; }
    int32 pattern_gap_open()      const { return pgo; }
    int32 pattern_gap_extension() const { return pge; }
    int32 text_gap_open()         const { return tgo; }
    int32 text_gap_extension()    const { return tge; }
};

template <typename aligner_type, typename quals_type>
void score_batch(const aligner_type aligner, const quals_type quals, const uint8_t* pat, const uint32_t* pat_off,
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
            if (column.size() < (M > N ? M : N) + 1u) column.resize( (M > N ? M : N) + 1u );   // TextBlockingTag keeps a column over the pattern
            aln::BestSink<int32> sink;
            aln::alignment_score( aligner, sym_string( M, pat + pat_off[i] ), quals.get( i, M ),
                                  sym_string( N, txt + txt_off[i] ), MIN_SCORE, sink, &column[0] );
            sinks[i].score = sink.score; sinks[i].x = sink.sink.x; sinks[i].y = sink.sink.y;
        }
    }
}

template <uint32 BAND, typename aligner_type, typename quals_type>
void banded_score_batch(const aligner_type aligner, const quals_type quals, const uint8_t* pat, const uint32_t* pat_off,
                        const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    #pragma omp parallel for schedule(dynamic,64) num_threads(threads)
    for (int64 i = 0; i < int64(count); ++i)
    {
        const uint32 M = pat_off[i+1] - pat_off[i];
        const uint32 N = txt_off[i+1] - txt_off[i];
        aln::BestSink<int32> sink;
        aln::banded_alignment_score<BAND>( aligner, sym_string( M, pat + pat_off[i] ), quals.get( i, M ),
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

template <uint32 CHK, typename aligner_type, typename quals_type>
void traceback_batch(const aligner_type aligner, const quals_type quals, const uint8_t* pat, const uint32_t* pat_off,
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
                aligner, sym_string( M, pat + pat_off[i] ), quals.get( i, M ),
                sym_string( N, txt + txt_off[i] ), MIN_SCORE, bt, &checkpoints[0], submatrix, &column[0] );
            store_aln( out + i, a );
            n_ops[i] = bt.n;
        }
    }
}

template <uint32 BAND, uint32 CHK, typename aligner_type, typename quals_type>
void banded_traceback_batch(const aligner_type aligner, const quals_type quals, const uint8_t* pat, const uint32_t* pat_off,
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
                aligner, sym_string( M, pat + pat_off[i] ), quals.get( i, M ),
                sym_string( N, txt + txt_off[i] ), MIN_SCORE, bt, &checkpoints[0], submatrix );
            store_aln( out + i, a );
            n_ops[i] = bt.n;
        }
    }
}

// ---- runtime -> compile-time dispatch ---------------------------------------------------------

template <typename F>
int dispatch_aligner(const orc_aligner a, F f, const orc_qscheme* qs = 0)
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
    else if (a.kind == 1 && qs)
    {
        const TableSWScheme s = { qs->match, qs->mismatch, a.p2, a.p3 };
        #define NVB_MK_SWQ(T) make_smith_waterman_aligner<T>( s )
        NVB_TYPES(NVB_MK_SWQ)
    }
    else if (a.kind == 2 && qs)
    {
        const TableGotohScheme s = { qs->match, qs->mismatch, a.p2, a.p3, qs->text_gap_open, qs->text_gap_ext };
        #define NVB_MK_GTQ(T) make_gotoh_aligner<T>( s )
        NVB_TYPES(NVB_MK_GTQ)
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
    const uint8_t* quals;   // NULL: trivial_quality_string
};

struct ScoreF { Args g;
    template <typename A> void operator()(const A a) const
    { if (g.quals) { const ByteQuals q = { g.quals, g.pat_off }; score_batch( a, q, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); }
      else score_batch( a, NoQuals(), g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); } };

// the same through the aligner's TextBlockingTag variant (aln::transpose, alignment.h:395-540; sw_inl.h:872+, gotoh_inl.h:848+)
struct ScoreTextBlockingF { Args g;
    template <typename A> void operator()(const A a) const
    { score_batch( aln::transpose( a ), NoQuals(), g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); } };

template <uint32 BAND> struct BandedScoreF { Args g;
    template <typename A> void operator()(const A a) const
    { if (g.quals) { const ByteQuals q = { g.quals, g.pat_off }; banded_score_batch<BAND>( a, q, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); }
      else banded_score_batch<BAND>( a, NoQuals(), g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.sinks, g.threads ); } };

template <uint32 CHK> struct TracebackF { Args g;
    template <typename A> void operator()(const A a) const
    { if (g.quals) { const ByteQuals q = { g.quals, g.pat_off }; traceback_batch<CHK>( a, q, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.out, g.ops, g.ops_stride, g.n_ops, g.threads ); }
      else traceback_batch<CHK>( a, NoQuals(), g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.out, g.ops, g.ops_stride, g.n_ops, g.threads ); } };

template <uint32 BAND, uint32 CHK> struct BandedTracebackF { Args g;
    template <typename A> void operator()(const A a) const
    { if (g.quals) { const ByteQuals q = { g.quals, g.pat_off }; banded_traceback_batch<BAND,CHK>( a, q, g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.out, g.ops, g.ops_stride, g.n_ops, g.threads ); }
      else banded_traceback_batch<BAND,CHK>( a, NoQuals(), g.pat, g.pat_off, g.txt, g.txt_off, g.count, g.out, g.ops, g.ops_stride, g.n_ops, g.threads ); } };

template <uint32 BAND>
int banded_traceback_chk(const orc_aligner a, const uint32 chk, const Args& g, const orc_qscheme* qs)
{
    switch (chk) {
    case 16: { BandedTracebackF<BAND,16> f = { g }; return dispatch_aligner( a, f, qs ); }
    case 32: { BandedTracebackF<BAND,32> f = { g }; return dispatch_aligner( a, f, qs ); }
    case 64: { BandedTracebackF<BAND,64> f = { g }; return dispatch_aligner( a, f, qs ); }
    default: return -2; }
}

} // anonymous namespace

extern "C" {

int ref_score_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, sinks, 0, 0, 0, 0, threads, quals };
    ScoreF f = { g };
    return dispatch_aligner( a, f, qs );
}
int ref_score(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
              const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{ return ref_score_q( a, 0, 0, pat, pat_off, txt, txt_off, count, sinks, threads ); }

int ref_score_text_blocking(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                            const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, sinks, 0, 0, 0, 0, threads, 0 };
    ScoreTextBlockingF f = { g };
    return dispatch_aligner( a, f, 0 );
}

int ref_banded_score_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                       const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, sinks, 0, 0, 0, 0, threads, quals };
    switch (band) {
    case 3:  { BandedScoreF<3>  f = { g }; return dispatch_aligner( a, f, qs ); }
    case 5:  { BandedScoreF<5>  f = { g }; return dispatch_aligner( a, f, qs ); }
    case 7:  { BandedScoreF<7>  f = { g }; return dispatch_aligner( a, f, qs ); }
    case 15: { BandedScoreF<15> f = { g }; return dispatch_aligner( a, f, qs ); }
    case 31: { BandedScoreF<31> f = { g }; return dispatch_aligner( a, f, qs ); }
    default: return -2; }
}
int ref_banded_score(orc_aligner a, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                     const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{ return ref_banded_score_q( a, 0, 0, band, pat, pat_off, txt, txt_off, count, sinks, threads ); }

int ref_traceback_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                    const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                    uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, 0, out, ops, ops_stride, n_ops, threads, quals };
    switch (checkpoints) {
    case 16: { TracebackF<16> f = { g }; return dispatch_aligner( a, f, qs ); }
    case 32: { TracebackF<32> f = { g }; return dispatch_aligner( a, f, qs ); }
    case 64: { TracebackF<64> f = { g }; return dispatch_aligner( a, f, qs ); }
    default: return -2; }
}
int ref_traceback(orc_aligner a, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                  const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                  uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{ return ref_traceback_q( a, 0, 0, checkpoints, pat, pat_off, txt, txt_off, count, out, ops, ops_stride, n_ops, threads ); }

int ref_banded_traceback_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                           const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                           uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                           uint32_t* n_ops, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args g = { pat, pat_off, txt, txt_off, count, 0, out, ops, ops_stride, n_ops, threads, quals };
    switch (band) {
    case 3:  return banded_traceback_chk<3>( a, checkpoints, g, qs );
    case 5:  return banded_traceback_chk<5>( a, checkpoints, g, qs );
    case 7:  return banded_traceback_chk<7>( a, checkpoints, g, qs );
    case 15: return banded_traceback_chk<15>( a, checkpoints, g, qs );
    case 31: return banded_traceback_chk<31>( a, checkpoints, g, qs );
    default: return -2; }
}
int ref_banded_traceback(orc_aligner a, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                         const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                         uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                         uint32_t* n_ops, int threads)
{ return ref_banded_traceback_q( a, 0, 0, band, checkpoints, pat, pat_off, txt, txt_off, count, out, ops, ops_stride, n_ops, threads ); }

} // extern "C"

// ---- Best2Sink (nvbio/alignment/sink.h:99-123) and MyersTag (myers/myers_banded_inl.h) ----------------------
namespace {

inline void store_sink2(orc_sink2* out, const aln::Best2Sink<int32>& s)
{
    out->score1 = s.score1; out->score2 = s.score2;
    out->x1 = s.sink1.x; out->y1 = s.sink1.y; out->x2 = s.sink2.x; out->y2 = s.sink2.y;
}

struct Args2
{
    const uint8_t* pat; const uint32_t* pat_off; const uint8_t* txt; const uint32_t* txt_off;
    uint32_t count; uint32_t distinct; orc_sink2* sinks; int threads;
};

struct Score2F { Args2 g;
    template <typename aligner_type> void operator()(const aligner_type aligner) const
    {
        typedef typename aln::column_storage_type<aligner_type>::type cell_type;
        #pragma omp parallel num_threads(g.threads)
        {
            std::vector<cell_type> column;
            #pragma omp for schedule(dynamic,64)
            for (int64 i = 0; i < int64(g.count); ++i)
            {
                const uint32 M = g.pat_off[i+1] - g.pat_off[i];
                const uint32 N = g.txt_off[i+1] - g.txt_off[i];
                if (column.size() < (M > N ? M : N) + 1u) column.resize( (M > N ? M : N) + 1u );   // TextBlockingTag keeps a column over the pattern
                aln::Best2Sink<int32> sink( g.distinct );
                aln::alignment_score( aligner, sym_string( M, g.pat + g.pat_off[i] ), aln::trivial_quality_string(),
                                      sym_string( N, g.txt + g.txt_off[i] ), MIN_SCORE, sink, &column[0] );
                store_sink2( g.sinks + i, sink );
            }
        }
    } };

template <uint32 BAND> struct BandedScore2F { Args2 g;
    template <typename aligner_type> void operator()(const aligner_type aligner) const
    {
        #pragma omp parallel for schedule(dynamic,64) num_threads(g.threads)
        for (int64 i = 0; i < int64(g.count); ++i)
        {
            const uint32 M = g.pat_off[i+1] - g.pat_off[i];
            const uint32 N = g.txt_off[i+1] - g.txt_off[i];
            aln::Best2Sink<int32> sink( g.distinct );
            aln::banded_alignment_score<BAND>( aligner, sym_string( M, g.pat + g.pat_off[i] ), aln::trivial_quality_string(),
                                               sym_string( N, g.txt + g.txt_off[i] ), MIN_SCORE, sink );
            store_sink2( g.sinks + i, sink );
        }
    } };

template <uint32 BAND, aln::AlignmentType TYPE, uint32 ALPHABET>
void myers_batch(const int32 min_score, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                 const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    const aln::EditDistanceAligner<TYPE, aln::MyersTag<ALPHABET> > aligner =
        aln::make_edit_distance_aligner<TYPE, aln::MyersTag<ALPHABET> >();
    #pragma omp parallel for schedule(dynamic,64) num_threads(threads)
    for (int64 i = 0; i < int64(count); ++i)
    {
        const uint32 M = pat_off[i+1] - pat_off[i];
        const uint32 N = txt_off[i+1] - txt_off[i];
        aln::BestSink<int32> sink;
        aln::banded_alignment_score<BAND>( aligner, sym_string( M, pat + pat_off[i] ), aln::trivial_quality_string(),
                                           sym_string( N, txt + txt_off[i] ), min_score, sink );
        sinks[i].score = sink.score; sinks[i].x = sink.sink.x; sinks[i].y = sink.sink.y;
    }
}

template <uint32 BAND, uint32 ALPHABET>
int myers_type(const int32 type, const int32 min_score, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
               const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    switch (type) {
    case 0: myers_batch<BAND,aln::GLOBAL,ALPHABET>( min_score, pat, pat_off, txt, txt_off, count, sinks, threads ); return 0;
    case 1: myers_batch<BAND,aln::LOCAL,ALPHABET>( min_score, pat, pat_off, txt, txt_off, count, sinks, threads ); return 0;
    case 2: myers_batch<BAND,aln::SEMI_GLOBAL,ALPHABET>( min_score, pat, pat_off, txt, txt_off, count, sinks, threads ); return 0;
    default: return -1; }
}

template <uint32 BAND>
int myers_alphabet(const uint32 alphabet, const int32 type, const int32 min_score, const uint8_t* pat, const uint32_t* pat_off,
                   const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    switch (alphabet) {
    case 2: return myers_type<BAND,2>( type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    case 4: return myers_type<BAND,4>( type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    case 5: return myers_type<BAND,5>( type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    default: return -2; }
}

} // anonymous namespace

extern "C" {

int ref_score_best2(orc_aligner a, uint32_t band, uint32_t distinct_dist, const uint8_t* pat, const uint32_t* pat_off,
                    const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink2* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    Args2 g = { pat, pat_off, txt, txt_off, count, distinct_dist, sinks, threads };
    switch (band) {
    case 0:  { Score2F f = { g }; return dispatch_aligner( a, f ); }
    case 3:  { BandedScore2F<3>  f = { g }; return dispatch_aligner( a, f ); }
    case 5:  { BandedScore2F<5>  f = { g }; return dispatch_aligner( a, f ); }
    case 7:  { BandedScore2F<7>  f = { g }; return dispatch_aligner( a, f ); }
    case 15: { BandedScore2F<15> f = { g }; return dispatch_aligner( a, f ); }
    case 31: { BandedScore2F<31> f = { g }; return dispatch_aligner( a, f ); }
    default: return -2; }
}

int ref_banded_myers(int32_t type, uint32_t alphabet, uint32_t band, int32_t min_score, const uint8_t* pat,
                     const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off, uint32_t count,
                     orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    switch (band) {
    case 7:  return myers_alphabet<7>(  alphabet, type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    case 15: return myers_alphabet<15>( alphabet, type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    case 16: return myers_alphabet<16>( alphabet, type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    case 31: return myers_alphabet<31>( alphabet, type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    case 32: return myers_alphabet<32>( alphabet, type, min_score, pat, pat_off, txt, txt_off, count, sinks, threads );
    default: return -2; }
}

} // extern "C"
