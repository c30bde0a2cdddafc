/*
 * oracle_abi.h -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * One C ABI implemented twice:
 *   - oracle/oracle.c        (prefix orc_)  plain-C restatement of the reference algorithms
 *   - oracle/ref_wrap.cpp    (prefix ref_)  thin wrapper that compiles the reference's OWN
 *                                           headers from /root/reference (built into oracle/_ref/)
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load either library.
 *
 * Conventions (all follow the reference, cited in oracle.c):
 *   symbols   : one uint8 per symbol, 0..3 = A,C,G,T, >3 = N
 *   string set: concatenated symbols + offsets[count+1]  (ConcatenatedStringSet)
 *   aligner   : kind 0=EDIT_DISTANCE 1=SMITH_WATERMAN 2=GOTOH ; type 0=GLOBAL 1=LOCAL 2=SEMI_GLOBAL
 *               SW   : p0=match p1=mismatch p2=deletion p3=insertion
 *               GOTOH: p0=match p1=mismatch p2=gap_open p3=gap_ext
 *               ED   : params ignored (0,-1,-1,-1)
 *   sink      : {score, text_end(x), pattern_end(y)}  == aln::BestSink<int32>
 *   alignment : {score, source.x, source.y, sink.x, sink.y} == aln::Alignment<int32>
 *   ops       : uint8 per op in PUSH ORDER (end -> start): 0=M(substitution) 1=I 2=D
 */
#ifndef NVB_ORACLE_ABI_H
#define NVB_ORACLE_ABI_H

#include <stdint.h>
#include <stddef.h>

#ifndef ORC_PREFIX
#define ORC_PREFIX orc_
#endif
#define ORC_CAT2(a, b) a##b
#define ORC_CAT(a, b) ORC_CAT2(a, b)
#define ORC(name) ORC_CAT(ORC_PREFIX, name)

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    uint32_t        length;     /* number of text symbols                       */
    uint32_t        primary;    /* row of the full text in the BWT matrix       */
    uint32_t        L2[5];      /* cumulative symbol counts                     */
    const uint32_t* bwt_occ;    /* fused nvBWT array: per 64 symbols 4 bwt words + 4 occ counters */
    const uint32_t* ssa;        /* (length+16)/16 entries, ssa[0] = 0xFFFFFFFF  */
} orc_fmindex;

typedef struct { int32_t kind, type, p0, p1, p2, p3; } orc_aligner;
typedef struct { int32_t score; uint32_t x, y; } orc_sink;
typedef struct { int32_t score; uint32_t src_x, src_y, sink_x, sink_y; } orc_alignment;

/* number of uint32 words of the fused bwt_occ array for a text of n symbols */
size_t ORC(bwt_occ_words)(uint32_t n);
/* number of uint32 entries of the sampled suffix array */
size_t ORC(ssa_words)(uint32_t n);

/* build SA -> BWT -> occ -> fused layout -> SSA(16). sa_out (n+1 int32) may be NULL. */
int ORC(build_index)(const uint8_t* text, uint32_t n, uint32_t* primary, uint32_t* L2,
                     uint32_t* bwt_occ, uint32_t* ssa, int32_t* sa_out);

/* exact backward search; ranges = 2 uint32 per query (inclusive SA interval, raw).
 * blocks (optional, may be NULL) receives per query the number of distinct fused 32-byte
 * blocks the reference algorithm touches (the roofline's algorithmic bytes / 32). */
int ORC(fm_match)(const orc_fmindex* fmi, const uint8_t* symbols, const uint32_t* offsets,
                  uint32_t count, uint32_t* ranges, uint32_t* blocks, int threads);

/* match_reverse (fmindex_inl.h:252-278): the pattern consumed front to back */
int ORC(fm_match_reverse)(const orc_fmindex* fmi, const uint8_t* symbols, const uint32_t* offsets,
                          uint32_t count, uint32_t* ranges, int threads);

/* rank of symbol c[i] over the range pairs (x,y); out = 2 uint32 per query */
int ORC(fm_rank)(const orc_fmindex* fmi, const uint32_t* ranges, const uint8_t* c,
                 uint32_t count, uint32_t* out, int threads);

/* all four symbol counts in [0,k]; out = 4 uint32 per query */
int ORC(fm_rank4)(const orc_fmindex* fmi, const uint32_t* k, uint32_t count, uint32_t* out, int threads);

/* SA row -> text coordinate; steps (optional) = LF steps walked */
int ORC(fm_locate)(const orc_fmindex* fmi, const uint32_t* rows, uint32_t count, uint32_t* pos,
                   uint32_t* steps, int threads);

/* full-matrix scoring, one sink per pair */
int ORC(score)(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
               const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads);

/* the same with the aligner's TextBlockingTag variant (alignment.h:266-279); see oracle_aln.c */
int ORC(score_text_blocking)(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                             const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads);

/* banded scoring */
int ORC(banded_score)(orc_aligner a, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                      const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks,
                      int threads);

/* full-matrix checkpointed traceback. ops: ops_stride bytes per pair, n_ops[i] valid entries. */
int ORC(traceback)(orc_aligner a, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                   const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                   uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads);

/* banded checkpointed traceback */
int ORC(banded_traceback)(orc_aligner a, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                          const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                          uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                          uint32_t* n_ops, int threads);

/* Quality-aware scoring schemes (alignment.h:411-436 GotohScoringScheme, :488-505 SmithWatermanScoringScheme; nvBowtie's is
 * nvBowtie/bowtie2/cuda/scoring.h:206-301): match(q) and mismatch(a, b, q) as 256-entry tables indexed by the quality of the
 * PATTERN position; the aligner's p2/p3 stay deletion/insertion (SW) or pattern_gap_open/extension (Gotoh); text_gap_* are
 * what the reference uses to initialise GLOBAL first columns / band rows and the banded infimum. quals = one uint8 per
 * pattern symbol, addressed by the patterns' offsets (NULL = trivial_quality_string: all 0). */
typedef struct { const int32_t* match; const int32_t* mismatch; int32_t text_gap_open, text_gap_ext; } orc_qscheme;
int ORC(score_q)(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                 const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads);
int ORC(banded_score_q)(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                        const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads);
int ORC(traceback_q)(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                     const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                     uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads);
int ORC(banded_traceback_q)(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                            const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                            uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                            uint32_t* n_ops, int threads);

/* aln::Best2Sink<int32> (sink.h:99-123): best and runner-up whose text end is more than distinct_dist away */
typedef struct { int32_t score1, score2; uint32_t x1, y1, x2, y2; } orc_sink2;
/* scoring into a Best2Sink; band 0 = full matrix */
int ORC(score_best2)(orc_aligner a, uint32_t band, uint32_t distinct_dist, const uint8_t* pat, const uint32_t* pat_off,
                     const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink2* sinks, int threads);
/* EditDistanceAligner<type, MyersTag<alphabet>> banded scoring (myers/myers_banded_inl.h:183-296); min_score is
 * truncated to int16 as the reference does; type 0 = GLOBAL, 2 = SEMI_GLOBAL (LOCAL reports nothing) */
int ORC(banded_myers)(int32_t type, uint32_t alphabet, uint32_t band, int32_t min_score, const uint8_t* pat,
                      const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off, uint32_t count,
                      orc_sink* sinks, int threads);

/* seed-and-extend composition (oracle_map.c); flags: bit 0 = reverse-complement strand, bit 1 = aligned.
 * begin/end = genome coordinates of the first aligned symbol / one past the last; ops in push order (end -> start). */
typedef struct { uint32_t seed_len, seed_interval, max_hits, max_range, band, both_strands; int32_t min_score;
                 orc_aligner aligner;
                 uint32_t seed_mismatches, max_seed_ranges;           /* nvBowtie's approximate seeds, forward index (mapping_inl.h:114-184) */
                 uint32_t hit_priority, max_seed_hits, max_rows;      /* hit selection by SA-range size (seed_hit.h:220-230, select_inl.h:93-113) */
                 const orc_fmindex* reverse_index;                    /* index of the reversed text: the other half of the approximate seeds (mapping_inl.h:399-402) */
               } orc_map_params;
typedef struct { int32_t score, second_score; uint32_t begin, end, flags, n_ops; } orc_read_hit;
int ORC(map_reads)(const orc_fmindex* fmi, const uint8_t* genome, const uint8_t* reads, const uint32_t* read_off,
                   uint32_t n_reads, const orc_map_params* p, orc_read_hit* hits, uint8_t* ops, uint32_t ops_stride,
                   uint32_t* n_candidates, int threads);

/* nvBowtie's approximate seed mapping map<find_exact> (mapping_inl.h:114-184); see oracle_map.c */
int ORC(fm_match_approx)(const orc_fmindex* fmi, const uint8_t* symbols, const uint32_t* offsets, uint32_t count, uint32_t len1,
                         int find_exact, uint32_t cap, uint32_t* ranges, uint32_t* counts, int threads);
typedef struct { int32_t score; uint32_t frag_begin, frag_end, flags; } orc_pair_hit;
int ORC(pair_hits)(const orc_read_hit* hits, uint32_t n_pairs, uint32_t min_frag, uint32_t max_frag, orc_pair_hit* out);
/* hit selection by SA-range size (seed_hit.h:220-230, mapping_inl.h:158-160, select_inl.h:93-113); see oracle_map.c */
size_t ORC(hit_priority)(const uint32_t* sizes, uint32_t n, uint32_t max_hits, uint32_t max_seed_hits, uint32_t max_rows,
                         uint32_t* take, uint32_t* order, uint32_t* n_order);
/* opposite-mate window search for pairs with one aligned mate (score_inl.h:318-411); see oracle_map.c */
int ORC(rescue_mates)(const uint8_t* genome, uint32_t G, const uint8_t* reads, const uint32_t* read_off, uint32_t n_pairs,
                      orc_aligner aligner, int32_t min_score, uint32_t max_frag, orc_read_hit* hits, uint8_t* ops, uint32_t ops_stride,
                      uint32_t* n_candidates, uint32_t* n_rescued);
/* BowtieMapq2 (nvBowtie/bowtie2/cuda/mapq.h:145-316) on scores the caller supplies; see oracle_map.c */
int ORC(mapq_v2)(int32_t best, int32_t has_second, int32_t second, int32_t perfect, int32_t min_score, int32_t monotone);
/* MDS byte-code + edit distance of one alignment (nvBowtie/bowtie2/cuda/traceback_inl.h:555-690) */
int ORC(ops_to_mds)(const uint8_t* ops, uint32_t n_ops, const uint8_t* pattern, const uint8_t* text, uint32_t j0, uint32_t k0,
                    uint8_t* mds, uint32_t cap, uint32_t* mds_len_out, uint32_t* ed_out);

#ifdef __cplusplus
}
#endif
#endif
