/*
 * oracle_aln.c -- TEST INFRASTRUCTURE ONLY: plain-C restatement of the reference's DP alignment path
 * (Smith-Waterman / Gotoh / edit distance; full-matrix and banded; scoring and checkpointed traceback).
 *
 * PARITY PINNED: (a) the reference's golden vectors (nvbio-test/alignment_test.cu:641-707 banded ED
 * scores, :721-753 ACAACTA CIGARs, :770-785 147M2D3M, :802-863 6I138M / 1I1M2I1M3I136M) are checked
 * by tests/test_oracle_golden.py; (b) randomised bit-for-bit comparison against the compiled
 * reference (oracle/_ref) in tests/test_oracle_vs_ref.py (scores, sinks incl. ties, sources, ops).
 *
 * Never linked into the product library.
 *
 * What is restated (not "fixed"):
 *  - full DP runs in pattern stripes of 16 (SW/ED) or 8 (Gotoh) columns, text rows inside a stripe,
 *    columns inside a row; sinks use '<=' so the LAST visited maximal cell wins
 *    (sw_inl.h:610-749, gotoh_inl.h:611-840, sink_inl.h:40-48);
 *  - the stripe boundary column and all checkpoints round-trip through int16 (utils.h:51-53);
 *  - traceback = score pass saving a column every CHK pattern columns, then per checkpoint a
 *    re-computation storing flow cells and a walk (alignment_inl.h:355-455);
 *  - banded DP keeps W cells per pattern row, cell j <-> text position row+j, text past the end
 *    reads as 255 (sw_banded_inl.h:340-504, gotoh_banded_inl.h:406-646); banded checkpoints are
 *    clamped to INT16_MIN+32 when stored (sw_banded_inl.h:168-215, gotoh_banded_inl.h:217-249).
 *  - deviation, documented: the reference pre-loads text[0..W-2] without a bounds check (undefined
 *    when the text is shorter than W-1); here every out-of-range text position reads as 255.
 */
#include "oracle_abi.h"
#include <stdlib.h>
#include <string.h>
#include <limits.h>
#include <omp.h>

enum { T_GLOBAL = 0, T_LOCAL = 1, T_SEMI = 2 };
enum { K_ED = 0, K_SW = 1, K_GOTOH = 2 };
enum { OP_SUB = 0, OP_INS = 1, OP_DEL = 2, OP_SINK = 3, OP_INS_EXT = 4, OP_DEL_EXT = 8 };
enum { ST_H = 0, ST_E = 1, ST_F = 2 };

typedef struct { int gotoh, type, match, mismatch, ga, gb;
                 int tga, tgb;                  /* Gotoh: text_gap_open / text_gap_extension, used ONLY to initialise the first text column (full) /
                                                   band row (banded) of GLOBAL alignments and in the banded infimum (gotoh_inl.h:69-70, gotoh_banded_inl.h:48,437-439) */
                 const int32_t* qm; const int32_t* qx; /* quality-aware schemes: match(q) / mismatch(a,b,q) tables (alignment.h:411-436,488-505), NULL = constants */
               } scheme_t;
/* SW/ED: ga = deletion (vertical move, consumes text), gb = insertion (consumes pattern)
 * Gotoh: ga = gap open, gb = gap extension */

static scheme_t make_scheme(orc_aligner a)
{
    scheme_t s;
    s.gotoh = (a.kind == K_GOTOH);
    s.type = a.type;
    if (a.kind == K_ED) { s.match = 0; s.mismatch = -1; s.ga = -1; s.gb = -1; } /* ed_utils.h:36-43 */
    else { s.match = a.p0; s.mismatch = a.p1; s.ga = a.p2; s.gb = a.p3; }
    s.tga = s.ga; s.tgb = s.gb; s.qm = s.qx = NULL;
    return s;
}

static scheme_t make_scheme_q(orc_aligner a, const orc_qscheme* q)
{
    scheme_t s = make_scheme(a);
    if (q && a.kind != K_ED)
    {
        s.qm = q->match; s.qx = q->mismatch;
        if (s.gotoh) { s.tga = q->text_gap_open; s.tgb = q->text_gap_ext; }
    }
    return s;
}
/* substitution score at pattern quality qq: scoring.match(qq) / scoring.mismatch(r, q, qq) (sw_inl.h:470-472, gotoh_inl.h:506-508) */
static inline int32_t subst(const scheme_t* s, int equal, uint8_t qq)
{ return equal ? (s->qm ? s->qm[qq] : s->match) : (s->qx ? s->qx[qq] : s->mismatch); }

/* BestSink and, when `two` is set, Best2Sink (sink_inl.h:54-82): the runner-up must end more than `dd` text
 * positions away from the current best */
typedef struct { int32_t score; uint32_t x, y; int two; uint32_t dd; int32_t score2; uint32_t x2, y2; } best_t;
/* Field_traits<int32>::min() is -(1<<30) in the reference (nvbio/basic/numbers.h:738-741) */
static inline void best_init(best_t* b)
{ b->score = b->score2 = -(1 << 30); b->x = b->y = b->x2 = b->y2 = 0xFFFFFFFFu; b->two = 0; b->dd = 0; }
static inline void best_report(best_t* b, int32_t s, uint32_t x, uint32_t y)
{
    if (b->score <= s) { b->score = s; b->x = x; b->y = y; } /* sink_inl.h:44: last one wins */
    else if (b->two && b->score2 <= s && (x + b->dd < b->x || x > b->x + b->dd)) /* sink_inl.h:77-81 (uint32 arithmetic) */
    { b->score2 = s; b->x2 = x; b->y2 = y; }
}
static inline int32_t imax(int32_t a, int32_t b) { return a > b ? a : b; }

typedef struct { uint8_t* ops; uint32_t n, cap; } tracer_t;
static inline void push(tracer_t* t, uint8_t op) { if (t->n < t->cap) t->ops[t->n] = op; t->n++; }

/* ========================================================================================== */
/*  FULL MATRIX                                                                               */
/* ========================================================================================== */
typedef struct {
    scheme_t       s;
    const uint8_t* pat; uint32_t M;
    const uint8_t* qual;   /* pattern qualities (NULL = all 0: trivial_quality_string) */
    const uint8_t* txt; uint32_t N;
    int16_t*       colH;   /* stripe boundary column (H) : N cells                       */
    int16_t*       colE;   /* Gotoh only: E of the boundary column                        */
    best_t*        best;   /* NULL = null sink                                            */
    int16_t*       chkH;   /* checkpoints [id*N + i] (may be NULL)                         */
    int16_t*       chkE;
    uint32_t       CHK;
    uint8_t*       flow;   /* submatrix flow cells [i*CHK + (col - id*CHK)] (may be NULL)  */
} full_t;

enum { INIT_ZERO_COLUMN = 0, INIT_FROM_CHECKPOINT = 1 };

/* one stripe [block, block+ncols) of the matrix; last = this is the stripe holding column M */
static void full_stripe(full_t* f, uint32_t block, uint32_t ncols, int last, int save_chk,
                        uint32_t chk_id, int32_t infimum)
{
    const scheme_t* s = &f->s;
    const uint32_t N = f->N, M = f->M;
    const int local = (s->type == T_LOCAL);
    int32_t H[17], F[17];

    if (save_chk && (block % f->CHK) == 0)
    {
        const uint32_t id = block / f->CHK;
        for (uint32_t i = 0; i < N; ++i)
        {
            f->chkH[(size_t)id * N + i] = f->colH[i];
            if (s->gotoh) f->chkE[(size_t)id * N + i] = f->colE[i];
        }
    }
    /* row 0 of the stripe: sw_inl.h:655 / gotoh_inl.h:675 */
    for (uint32_t j = 0; j <= ncols; ++j)
    {
        if (local) H[j] = 0;
        else if (!s->gotoh) H[j] = s->gb * (int32_t)(block + j);
        else H[j] = (block + j > 0) ? s->ga + s->gb * (int32_t)(block + j - 1u) : 0;
        F[j] = infimum;
    }
    int32_t carry = H[0]; /* value of the boundary column on the previous row (diagonal term) */

    for (uint32_t i = 0; i < N; ++i)
    {
        const uint8_t r = f->txt[i];
        int32_t diag = carry;
        H[0] = carry = f->colH[i];
        int32_t E = s->gotoh ? f->colE[i] : 0;

        for (uint32_t j = 1; j <= ncols; ++j)
        {
            const uint8_t q = f->pat[block + j - 1];
            const int32_t S = subst(s, r == q, f->qual ? f->qual[block + j - 1] : 0);
            int32_t hi; uint8_t dir;
            if (!s->gotoh)
            {
                const int32_t top = H[j] + s->ga, left = H[j - 1] + s->gb, dg = diag + S;
                hi = imax(imax(top, left), dg);
                if (local) hi = imax(hi, 0);
                dir = top > left ? (top > dg ? OP_DEL : OP_SUB) : (left > dg ? OP_INS : OP_SUB);
                if (f->flow) /* sw_inl.h:382-387 */
                    f->flow[(size_t)i * f->CHK + (block + j - 1 - chk_id * f->CHK)] =
                        (local && hi == 0) ? OP_SINK : dir;
            }
            else
            {
                const int32_t ftop = F[j] + s->gb, htop = H[j] + s->ga;
                F[j] = imax(ftop, htop);
                const uint8_t fdir = ftop > htop ? OP_DEL_EXT : 0;
                const int32_t eleft = E + s->gb, hleft = H[j - 1] + s->ga;
                E = imax(eleft, hleft);
                const uint8_t edir = eleft > hleft ? OP_INS_EXT : 0;
                const int32_t dg = diag + S, top = F[j], left = E;
                hi = imax(imax(left, top), dg);
                if (local) hi = imax(hi, 0);
                dir = top > left ? (top > dg ? OP_DEL : OP_SUB) : (left > dg ? OP_INS : OP_SUB);
                if (f->flow) /* gotoh_inl.h:424-429 */
                    f->flow[(size_t)i * f->CHK + (block + j - 1 - chk_id * f->CHK)] =
                        (uint8_t)(((local && hi == 0) ? OP_SINK : dir) | edir | fdir);
            }
            diag = H[j];
            H[j] = hi;
        }
        /* right boundary of the stripe goes back to the column, through int16 */
        f->colH[i] = (int16_t)H[ncols];
        if (s->gotoh) f->colE[i] = (int16_t)E;

        if (f->best)
        {
            if (local)
                for (uint32_t j = 1; j <= ncols; ++j) best_report(f->best, H[j], i + 1, block + j);
            else if (last && s->type == T_SEMI)
                best_report(f->best, H[M - block], i + 1, M);
        }
    }
    if (last && s->type == T_GLOBAL && f->best) /* save_Mth with i = N-1: sw_inl.h:745-746 */
        best_report(f->best, H[M - block], N, M);
}

/* sweep the pattern window [wb, we) in stripes: sw_inl.h:640-747 */
static void full_sweep(full_t* f, uint32_t wb, uint32_t we, int init, int save_chk, uint32_t chk_id)
{
    const scheme_t* s = &f->s;
    const uint32_t BL = s->gotoh ? 8u : 16u; /* sw_inl.h:1321-1325, gotoh_inl.h:1259-1262 */
    const uint32_t N = f->N, M = f->M;
    const int32_t infimum = INT16_MIN - (s->ga < s->gb ? s->ga : s->gb); /* gotoh_inl.h:644 */

    if (init == INIT_FROM_CHECKPOINT)
    {
        for (uint32_t i = 0; i < N; ++i)
        {
            f->colH[i] = f->chkH[(size_t)chk_id * N + i];
            if (s->gotoh) f->colE[i] = f->chkE[(size_t)chk_id * N + i];
        }
    }
    else if (wb == 0)
    {
        for (uint32_t i = 0; i < N; ++i)
        {
            if (!s->gotoh) f->colH[i] = (int16_t)(s->type == T_GLOBAL ? s->ga * (int32_t)(i + 1) : 0);
            else
            {
                f->colH[i] = (int16_t)(s->type == T_GLOBAL ? s->tga + s->tgb * (int32_t)i : 0);
                f->colE[i] = (int16_t)(s->type == T_LOCAL ? 0 : infimum);
            }
        }
    }
    uint32_t end_block;
    if (we == M) { end_block = BL * ((M + BL - 1) / BL); if (end_block < BL) end_block = BL; }
    else end_block = we + BL;

    uint32_t block = wb;
    for (; block + BL < end_block; block += BL)
        full_stripe(f, block, BL, 0, save_chk, chk_id, infimum);
    if (we == M)
    {
        block = end_block - BL;
        const uint32_t ncols = (M - block < BL) ? M - block : BL;
        if (M > 0 || s->type == T_LOCAL) /* M == 0 is undefined in the reference for non-LOCAL */
            full_stripe(f, block, ncols, 1, save_chk, chk_id, infimum);
    }
}

/* walk one submatrix: sw_inl.h:1644-1693 (SW/ED), gotoh_inl.h:1573-1639 (Gotoh).
 * returns 1 when the alignment source has been found */
static int full_walk(const full_t* f, uint32_t chk_id, uint8_t* state, uint32_t* sx, uint32_t* sy,
                     tracer_t* out)
{
    const int local = (f->s.type == T_LOCAL);
    int32_t row = (int32_t)*sx;
    int32_t col = (int32_t)(*sy - chk_id * f->CHK - 1u);
    while (row > 0 && col >= 0)
    {
        const uint8_t op = f->flow[(size_t)(row - 1) * f->CHK + (uint32_t)col];
        if (!f->s.gotoh)
        {
            if (local && op == OP_SINK) { *sx = (uint32_t)row; *sy = (uint32_t)col + chk_id * f->CHK + 1u; return 1; }
            if (op != OP_DEL) --col;
            if (op != OP_INS) --row;
            push(out, op);
        }
        else
        {
            const uint8_t h_op = op & 3u;
            if (local && *state == ST_H && h_op == OP_SINK)
            { *sx = (uint32_t)row; *sy = (uint32_t)col + chk_id * f->CHK + 1u; return 1; }
            if (*state == ST_E)
            {
                if ((op & OP_INS_EXT) == 0) *state = ST_H;
                --col; push(out, OP_INS);
            }
            else if (*state == ST_F)
            {
                if ((op & OP_DEL_EXT) == 0) *state = ST_H;
                --row; push(out, OP_DEL);
            }
            else
            {
                if (h_op == OP_INS) *state = ST_E;
                else if (h_op == OP_DEL) *state = ST_F;
                else { --col; --row; push(out, OP_SUB); }
            }
        }
    }
    *sx = (uint32_t)row;
    *sy = (uint32_t)(col + (int32_t)(chk_id * f->CHK) + 1);
    return row ? 0 : 1;
}

/* alignment_inl.h:355-455 */
static void full_traceback(full_t* f, orc_alignment* res, tracer_t* out)
{
    const uint32_t M = f->M, CHK = f->CHK;
    best_t best; best_init(&best);
    f->best = &best; uint8_t* flow = f->flow; f->flow = NULL;
    full_sweep(f, 0, M, INIT_ZERO_COLUMN, 1, 0);
    f->best = NULL; f->flow = flow;

    res->score = best.score; res->sink_x = best.x; res->sink_y = best.y;
    if (best.x == 0xFFFFFFFFu || best.y == 0xFFFFFFFFu)
    { res->src_x = res->src_y = res->sink_x = res->sink_y = 0xFFFFFFFFu; return; }

    uint32_t sx = best.x, sy = best.y;
    int32_t id = (int32_t)((M + CHK - 1) / CHK) - 1;
    if (f->s.type == T_LOCAL)
        for (; id >= 0; --id) if ((uint32_t)id * CHK < sy) break;

    uint8_t state = ST_H;
    for (; id >= 0; --id)
    {
        const uint32_t wb = (uint32_t)id * CHK;
        const uint32_t we = wb + CHK < M ? wb + CHK : M;
        full_sweep(f, wb, we, INIT_FROM_CHECKPOINT, 0, (uint32_t)id);
        if (full_walk(f, (uint32_t)id, &state, &sx, &sy, out)) break;
    }
    /* borders that are not stored explicitly */
    if (f->s.type == T_SEMI || f->s.type == T_GLOBAL)
        if (sx == 0) for (; sy > 0; --sy) push(out, OP_INS);
    if (f->s.type == T_GLOBAL)
        if (sy == 0) for (; sx > 0; --sx) push(out, OP_DEL);
    res->src_x = sx; res->src_y = sy;
}

int orc_score_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel num_threads(threads)
    {
        int16_t* col = NULL; size_t cap = 0;
        #pragma omp for schedule(dynamic,64)
        for (int64_t p = 0; p < (int64_t)count; ++p)
        {
            full_t f; memset(&f, 0, sizeof(f));
            f.s = make_scheme_q(a, qs);
            f.qual = quals ? quals + pat_off[p] : NULL;
            f.pat = pat + pat_off[p]; f.M = pat_off[p + 1] - pat_off[p];
            f.txt = txt + txt_off[p]; f.N = txt_off[p + 1] - txt_off[p];
            if (cap < (size_t)f.N + 1) { cap = (size_t)f.N + 1; free(col); col = (int16_t*)malloc(cap * 2 * sizeof(int16_t)); }
            f.colH = col; f.colE = col + cap; f.CHK = 1u << 30;
            best_t best; best_init(&best); f.best = &best;
            full_sweep(&f, 0, f.M, INIT_ZERO_COLUMN, 0, 0);
            sinks[p].score = best.score; sinks[p].x = best.x; sinks[p].y = best.y;
        }
        free(col);
    }
    return 0;
}

int orc_score(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
              const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{ return orc_score_q(a, NULL, NULL, pat, pat_off, txt, txt_off, count, sinks, threads); }

/* TextBlockingTag scoring (alignment.h:266-279; sw/sw_inl.h:872+, gotoh/gotoh_inl.h:848+): the matrix is swept in stripes of 8
 * TEXT columns, pattern row by pattern row. Scores are those of the default blocking; only the order in which a LOCAL sink sees
 * equal scores changes. That order is the pattern-blocking order of the TRANSPOSED problem (pattern and text exchanged, deletion
 * and insertion exchanged: aln::transpose, alignment.h:395-540), so LOCAL and GLOBAL = transposed call with the sink coordinates
 * swapped back (this also reproduces the degenerate shapes: an empty TEXT reports nothing, an empty pattern the boundary cell);
 * SEMI_GLOBAL walks the last row in text order under both blockings. Pinned on the reference's
 * TextBlockingTag aligners (oracle/ref_wrap_aln.inl::ref_score_text_blocking). */
int orc_score_text_blocking(orc_aligner a, const uint8_t* pat, const uint32_t* pat_off, const uint8_t* txt,
                            const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (a.type == 2)
    {
        /* SEMI_GLOBAL is not symmetric: same sweep results as the default blocking; an EMPTY pattern, which the pattern-blocking
         * sweep reports nothing for, here ends on the last row's boundary: score 0 at (N, 0) */
        const int rc = orc_score(a, pat, pat_off, txt, txt_off, count, sinks, threads);
        for (uint32_t i = 0; i < count; ++i)
            if (pat_off[i + 1] == pat_off[i] && txt_off[i + 1] > txt_off[i]) { sinks[i].score = 0; sinks[i].x = txt_off[i + 1] - txt_off[i]; sinks[i].y = 0; }
        return rc;
    }
    orc_aligner t = a;
    if (a.kind == 1) { t.p2 = a.p3; t.p3 = a.p2; }
    const int rc = orc_score(t, txt, txt_off, pat, pat_off, count, sinks, threads);
    for (uint32_t i = 0; i < count; ++i) { const uint32_t x = sinks[i].x; sinks[i].x = sinks[i].y; sinks[i].y = x; }
    return rc;
}

int orc_traceback_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                    const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                    uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{
    if (checkpoints == 0 || (checkpoints & 15u)) return -2;
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,16) num_threads(threads)
    for (int64_t p = 0; p < (int64_t)count; ++p)
    {
        full_t f; memset(&f, 0, sizeof(f));
        f.s = make_scheme_q(a, qs);
        f.qual = quals ? quals + pat_off[p] : NULL;
        f.pat = pat + pat_off[p]; f.M = pat_off[p + 1] - pat_off[p];
        f.txt = txt + txt_off[p]; f.N = txt_off[p + 1] - txt_off[p];
        f.CHK = checkpoints;
        const size_t N1 = (size_t)f.N + 1, nchk = (f.M + checkpoints - 1) / checkpoints + 1;
        int16_t* mem = (int16_t*)calloc(2 * N1 + 2 * N1 * nchk, sizeof(int16_t));
        f.colH = mem; f.colE = mem + N1; f.chkH = mem + 2 * N1; f.chkE = f.chkH + N1 * nchk;
        f.flow = (uint8_t*)calloc(N1 * checkpoints, 1);
        tracer_t tr = { ops + (size_t)p * ops_stride, 0, ops_stride };
        full_traceback(&f, out + p, &tr);
        n_ops[p] = tr.n;
        free(mem); free(f.flow);
    }
    return 0;
}

int orc_traceback(orc_aligner a, uint32_t checkpoints, const uint8_t* pat, const uint32_t* pat_off,
                  const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_alignment* out,
                  uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, int threads)
{ return orc_traceback_q(a, NULL, NULL, checkpoints, pat, pat_off, txt, txt_off, count, out, ops, ops_stride, n_ops, threads); }

/* ========================================================================================== */
/*  BANDED                                                                                    */
/* ========================================================================================== */
#define MAX_BAND 64
typedef struct {
    scheme_t       s;
    uint32_t       W;      /* band length */
    const uint8_t* pat; uint32_t M;
    const uint8_t* qual;   /* pattern qualities (NULL = all 0) */
    const uint8_t* txt; uint32_t N;
    best_t*        best;   /* NULL = null sink */
    int16_t*       chkH;   /* [W*row_id + j], clamped to INT16_MIN+32 */
    int16_t*       chkF;   /* Gotoh only */
    uint32_t       CHK;
    uint8_t*       flow;   /* [(i - id*CHK)*W + j] */
} band_t;

static inline int16_t clamp16(int32_t v) { return (int16_t)(v > INT16_MIN + 32 ? v : INT16_MIN + 32); }
static inline uint8_t text_at(const band_t* b, uint32_t p) { return p < b->N ? b->txt[p] : 255u; }

/* rows [wb, we) of the band: sw_banded_inl.h:340-504, gotoh_banded_inl.h:406-646.
 * returns 0 when the reference refuses the problem (text shorter than pattern) */
static int banded_sweep(band_t* b, uint32_t wb, uint32_t we, int from_chk, int save_chk, uint32_t chk_id)
{
    const scheme_t* s = &b->s;
    const uint32_t W = b->W, M = b->M, N = b->N;
    const int local = (s->type == T_LOCAL);
    if (N < M) return 0;

    int32_t H[MAX_BAND], F[MAX_BAND];
    const int packed_cache = !(W == 3 || W == 5 || W == 7 || W == 15);
    /* SW: G = deletion() is charged on the move from the previous pattern row, I = insertion() on
     * the move from the previous text position (sw_banded_inl.h:365-366,399,417-418) */
    const int32_t G = s->ga, I = s->gb;
    int32_t gmax = s->ga > s->gb ? s->ga : s->gb;
    if (s->gotoh) { const int32_t t = s->tga > s->tgb ? s->tga : s->tgb; if (t > gmax) gmax = t; }   /* gotoh_banded_inl.h:437-439 */
    const int32_t infimum = s->gotoh ? INT16_MIN - gmax : INT16_MIN + gmax;

    if (from_chk)
        for (uint32_t j = 0; j < W; ++j)
        {
            H[j] = b->chkH[(size_t)chk_id * W + j];
            F[j] = s->gotoh ? b->chkF[(size_t)chk_id * W + j] : 0;
        }
    else /* init_row_zero: sw_banded_inl.h:35-45, gotoh_banded_inl.h:37-68 */
        for (uint32_t j = 0; j < W; ++j)
        {
            if (!s->gotoh) H[j] = (s->type == T_GLOBAL) ? (int32_t)j * s->ga : 0;
            else H[j] = (s->type == T_GLOBAL && j > 0) ? s->tga + (int32_t)(j - 1) * s->tgb : 0;   /* text gaps: gotoh_banded_inl.h:48 */
            F[j] = infimum;
        }

    for (uint32_t i = wb; i < we; ++i)
    {
        if (save_chk && (i % b->CHK) == 0)
            for (uint32_t j = 0; j < W; ++j)
            {
                b->chkH[(size_t)(i / b->CHK) * W + j] = clamp16(H[j]);
                if (s->gotoh) b->chkF[(size_t)(i / b->CHK) * W + j] = clamp16(F[j]);
            }
        const uint8_t q = b->pat[i];
        int32_t E = 0; uint8_t edir = 0;
        for (uint32_t j = 0; j < W; ++j)
        {
            /* every band cell but the last reads the text through the row cache, which for band
             * lengths other than 3,5,7,15 is a 2-bit packed stream (alignment_base_inl.h:66-90):
             * symbols are truncated to 2 bits there (N -> A, the 255 past-the-end marker -> T) */
            uint8_t g = text_at(b, i + j);
            if (packed_cache && j < W - 1) g &= 3u;
            const int32_t S = subst(s, g == q, b->qual ? b->qual[i] : 0);
            const int32_t dg = H[j] + S;
            int32_t hi; uint8_t dir, cell;
            if (!s->gotoh)
            {
                if (j == 0)          { const int32_t top = H[1] + G; hi = imax(top, dg); dir = top > dg ? OP_INS : OP_SUB; }
                else if (j == W - 1) { const int32_t left = H[j - 1] + I; hi = imax(left, dg); dir = left > dg ? OP_DEL : OP_SUB; }
                else
                {
                    const int32_t top = H[j + 1] + G, left = H[j - 1] + I;
                    hi = imax(imax(top, left), dg);
                    dir = top > left ? (top > dg ? OP_INS : OP_SUB) : (left > dg ? OP_DEL : OP_SUB);
                }
                if (local) hi = imax(hi, 0);
                cell = dir; /* banded SW/ED never stores SINK: sw_banded_inl.h:259-270 */
            }
            else
            {
                uint8_t fdir = 0;
                if (j < W - 1)
                {
                    const int32_t ftop = F[j + 1] + s->gb, htop = H[j + 1] + s->ga;
                    F[j] = imax(ftop, htop);
                    fdir = ftop > htop ? OP_DEL_EXT : 0;
                }
                else F[j] = infimum;
                if (j == 0)          { const int32_t top = F[0]; hi = imax(top, dg); dir = top > dg ? OP_INS : OP_SUB; }
                else if (j == W - 1) { const int32_t left = E; hi = imax(left, dg); dir = left > dg ? OP_DEL : OP_SUB; }
                else
                {
                    const int32_t top = F[j], left = E;
                    hi = imax(imax(top, left), dg);
                    dir = top > left ? (top > dg ? OP_INS : OP_SUB) : (left > dg ? OP_DEL : OP_SUB);
                }
                if (local) { hi = imax(hi, 0); if (hi == 0) dir = OP_SINK; }
                cell = (uint8_t)(dir | (j == 0 ? 0 : edir) | fdir);
                /* E for the next cell of this row: gotoh_banded_inl.h:497,560-564 */
                if (j == 0) { E = hi + s->ga; edir = 0; }
                else
                {
                    const int32_t eleft = E + s->gb, ediag = hi + s->ga;
                    edir = eleft > ediag ? OP_INS_EXT : 0;
                    E = imax(ediag, eleft);
                }
            }
            H[j] = hi;
            if (local && b->best) best_report(b->best, hi, i + j + 1, i + 1);
            if (b->flow) b->flow[(size_t)(i - chk_id * b->CHK) * W + j] = cell;
        }
    }
    if (save_chk) /* last_row: sw_banded_inl.h:196-215 */
    {
        const uint32_t row = (we + b->CHK - 1) / b->CHK;
        for (uint32_t j = 0; j < W; ++j)
        {
            b->chkH[(size_t)row * W + j] = clamp16(H[j]);
            if (s->gotoh) b->chkF[(size_t)row * W + j] = clamp16(F[j]);
        }
    }
    if (we == M && b->best)
    {
        if (s->type == T_GLOBAL) best_report(b->best, H[W - 1], M + W - 1, M);
        else if (s->type == T_SEMI)
        {
            const uint32_t lim = (M + W - 1 < N ? M + W - 1 : N) - (M - 1);
            best_report(b->best, H[0], M, M);
            for (uint32_t j = 1; j < W; ++j) if (j < lim) best_report(b->best, H[j], M + j, M);
        }
    }
    return 1;
}

/* sw_banded_inl.h:740-800 (SW/ED), gotoh_banded_inl.h:883-950 (Gotoh) */
static int banded_walk(const band_t* b, uint32_t chk_id, uint8_t* state, uint32_t* sx, uint32_t* sy,
                       tracer_t* out)
{
    const int local = (b->s.type == T_LOCAL);
    const uint32_t W = b->W;
    int32_t entry = (int32_t)(*sx - *sy);
    int32_t row   = (int32_t)(*sy - chk_id * b->CHK - 1u);
    while (row >= 0)
    {
        const uint8_t op = b->flow[(size_t)row * W + (uint32_t)entry];
        if (!b->s.gotoh)
        {
            if (local && op == OP_SINK)
            { *sy = (uint32_t)row + chk_id * b->CHK + 1u; *sx = (uint32_t)entry + *sy; return 1; }
            if (op == OP_DEL)      { --entry; push(out, OP_DEL); }
            else if (op == OP_INS) { ++entry; --row; push(out, OP_INS); }
            else                   { --row; push(out, OP_SUB); }
        }
        else
        {
            const uint8_t h_op = op & 3u;
            if (local && *state == ST_H && h_op == OP_SINK)
            { *sy = (uint32_t)row + chk_id * b->CHK + 1u; *sx = (uint32_t)entry + *sy; return 1; }
            if (*state == ST_E)
            {
                if ((op & OP_INS_EXT) == 0) *state = ST_H;
                --entry; push(out, OP_DEL);
            }
            else if (*state == ST_F)
            {
                if ((op & OP_DEL_EXT) == 0) *state = ST_H;
                ++entry; --row; push(out, OP_INS);
            }
            else
            {
                if (h_op == OP_DEL) *state = ST_E;
                else if (h_op == OP_INS) *state = ST_F;
                else { --row; push(out, OP_SUB); }
            }
        }
        if (entry < 0 || entry >= (int32_t)W) break; /* the reference only asserts here */
    }
    *sy = chk_id * b->CHK;
    *sx = (uint32_t)entry + *sy;
    return 0;
}

/* banded_inl.h:354-430 */
static void banded_traceback(band_t* b, orc_alignment* res, tracer_t* out)
{
    const uint32_t M = b->M, CHK = b->CHK;
    best_t best; best_init(&best);
    b->best = &best; uint8_t* flow = b->flow; b->flow = NULL;
    banded_sweep(b, 0, M, 0, 1, 0);
    b->best = NULL; b->flow = flow;

    res->score = best.score; res->sink_x = best.x; res->sink_y = best.y;
    if (best.x == 0xFFFFFFFFu || best.y == 0xFFFFFFFFu)
    { res->src_x = res->src_y = res->sink_x = res->sink_y = 0xFFFFFFFFu; return; }

    uint32_t sx = best.x, sy = best.y;
    int32_t id = (int32_t)(1u + (M + CHK - 1) / CHK) - 2;
    if (b->s.type == T_LOCAL)
        for (; id >= 0; --id) if ((uint32_t)id * CHK < sy) break;
    uint8_t state = ST_H;
    for (; id >= 0; --id)
    {
        const uint32_t wb = (uint32_t)id * CHK;
        const uint32_t we = wb + CHK < M ? wb + CHK : M;
        banded_sweep(b, wb, we, 1, 0, (uint32_t)id);
        if (banded_walk(b, (uint32_t)id, &state, &sx, &sy, out)) break;
    }
    res->src_x = sx; res->src_y = sy;
}

int orc_banded_score_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                       const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{
    if (band < 2 || band > MAX_BAND) return -2;
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,64) num_threads(threads)
    for (int64_t p = 0; p < (int64_t)count; ++p)
    {
        band_t b; memset(&b, 0, sizeof(b));
        b.s = make_scheme_q(a, qs); b.W = band; b.CHK = 1u << 30;
        b.qual = quals ? quals + pat_off[p] : NULL;
        b.pat = pat + pat_off[p]; b.M = pat_off[p + 1] - pat_off[p];
        b.txt = txt + txt_off[p]; b.N = txt_off[p + 1] - txt_off[p];
        best_t best; best_init(&best); b.best = &best;
        banded_sweep(&b, 0, b.M, 0, 0, 0);
        sinks[p].score = best.score; sinks[p].x = best.x; sinks[p].y = best.y;
    }
    return 0;
}

int orc_banded_score(orc_aligner a, uint32_t band, const uint8_t* pat, const uint32_t* pat_off,
                     const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink* sinks, int threads)
{ return orc_banded_score_q(a, NULL, NULL, band, pat, pat_off, txt, txt_off, count, sinks, threads); }

int orc_banded_traceback_q(orc_aligner a, const orc_qscheme* qs, const uint8_t* quals, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                           const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                           uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                           uint32_t* n_ops, int threads)
{
    if (band < 2 || band > MAX_BAND || checkpoints == 0) return -2;
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,16) num_threads(threads)
    for (int64_t p = 0; p < (int64_t)count; ++p)
    {
        band_t b; memset(&b, 0, sizeof(b));
        b.s = make_scheme_q(a, qs); b.W = band; b.CHK = checkpoints;
        b.qual = quals ? quals + pat_off[p] : NULL;
        b.pat = pat + pat_off[p]; b.M = pat_off[p + 1] - pat_off[p];
        b.txt = txt + txt_off[p]; b.N = txt_off[p + 1] - txt_off[p];
        const size_t nchk = (size_t)band * (2u + (b.M + checkpoints - 1) / checkpoints);
        int16_t* mem = (int16_t*)calloc(2 * nchk, sizeof(int16_t));
        b.chkH = mem; b.chkF = mem + nchk;
        b.flow = (uint8_t*)calloc((size_t)band * checkpoints, 1);
        tracer_t tr = { ops + (size_t)p * ops_stride, 0, ops_stride };
        banded_traceback(&b, out + p, &tr);
        n_ops[p] = tr.n;
        free(mem); free(b.flow);
    }
    return 0;
}

int orc_banded_traceback(orc_aligner a, uint32_t band, uint32_t checkpoints, const uint8_t* pat,
                         const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off,
                         uint32_t count, orc_alignment* out, uint8_t* ops, uint32_t ops_stride,
                         uint32_t* n_ops, int threads)
{ return orc_banded_traceback_q(a, NULL, NULL, band, checkpoints, pat, pat_off, txt, txt_off, count, out, ops, ops_stride, n_ops, threads); }

/* ========================================================================================== */
/*  Best2Sink scoring (sink.h:99-123, sink_inl.h:54-82) over the same sweeps; band 0 = full    */
/* ========================================================================================== */
int orc_score_best2(orc_aligner a, uint32_t band, uint32_t distinct_dist, const uint8_t* pat, const uint32_t* pat_off,
                    const uint8_t* txt, const uint32_t* txt_off, uint32_t count, orc_sink2* sinks, int threads)
{
    if (band == 1 || band > MAX_BAND) return -2;
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel num_threads(threads)
    {
        int16_t* col = NULL; size_t cap = 0;
        #pragma omp for schedule(dynamic,64)
        for (int64_t p = 0; p < (int64_t)count; ++p)
        {
            best_t best; best_init(&best); best.two = 1; best.dd = distinct_dist;
            const uint32_t M = pat_off[p + 1] - pat_off[p], N = txt_off[p + 1] - txt_off[p];
            if (band == 0)
            {
                full_t f; memset(&f, 0, sizeof(f));
                f.s = make_scheme(a);
                f.pat = pat + pat_off[p]; f.M = M; f.txt = txt + txt_off[p]; f.N = N;
                if (cap < (size_t)N + 1) { cap = (size_t)N + 1; free(col); col = (int16_t*)malloc(cap * 2 * sizeof(int16_t)); }
                f.colH = col; f.colE = col + cap; f.CHK = 1u << 30; f.best = &best;
                full_sweep(&f, 0, f.M, INIT_ZERO_COLUMN, 0, 0);
            }
            else
            {
                band_t b; memset(&b, 0, sizeof(b));
                b.s = make_scheme(a); b.W = band; b.CHK = 1u << 30;
                b.pat = pat + pat_off[p]; b.M = M; b.txt = txt + txt_off[p]; b.N = N; b.best = &best;
                banded_sweep(&b, 0, b.M, 0, 0, 0);
            }
            sinks[p].score1 = best.score; sinks[p].score2 = best.score2;
            sinks[p].x1 = best.x; sinks[p].y1 = best.y; sinks[p].x2 = best.x2; sinks[p].y2 = best.y2;
        }
        free(col);
    }
    return 0;
}

/* ========================================================================================== */
/*  MyersTag<ALPHABET>: banded bit-vector edit distance (myers/myers_banded_inl.h:183-262)     */
/* ========================================================================================== */
/* One 32-bit column of vertical deltas slides down the diagonal: bit BAND-1 is the main diagonal's
 * cell. `Peq[c]` holds, for the pattern symbols currently inside the window, the positions equal to c
 * (loaded at bit BAND-1, shifted right once per text column). Phase 1 follows the diagonal while new
 * pattern symbols enter, phase 2 walks right along the last pattern row; the running distance is the
 * NEGATED edit distance. min_score is truncated to int16 exactly as banded_myers' parameter list does
 * (:189), so the INT_MIN of the batch drivers becomes 0. */
static inline uint32_t myers_peq(const uint32_t* B, uint32_t alphabet, uint8_t c)
{
    if (alphabet == 2) return B[c & 1u];       /* MyersBitVectors<2>::get (:85-90) */
    if (alphabet == 4) return B[c & 3u];       /* MyersBitVectors<4>::get (:121-128) */
    return B[c <= 4 ? c : 4];                  /* MyersBitVectors<5>::get (:160-166): anything above 3 reads B[4] */
}
static inline void myers_load(uint32_t* B, uint32_t alphabet, uint8_t c, uint32_t v)
{
    if (alphabet == 2) B[c & 1u] |= v;
    else if (alphabet == 4) B[c & 3u] |= v;
    else if (c <= 4) B[c] |= v;                /* MyersBitVectors<5>::load (:150-158): symbols above 4 load nothing */
}
static inline uint32_t myers_column(uint32_t eq, uint32_t* VP, uint32_t* VN, uint32_t* HPo, uint32_t* HNo)
{
    uint32_t X = eq | *VN;
    const uint32_t D0 = ((*VP + (X & *VP)) ^ *VP) | X;
    const uint32_t HN = *VP & D0;
    const uint32_t HP = *VN | ~(*VP | D0);
    X = D0 >> 1;
    *VN = X & HP;
    *VP = HN | ~(X | HP);
    *HPo = HP; *HNo = HN;
    return D0;
}

int orc_banded_myers(int32_t type, uint32_t alphabet, uint32_t band, int32_t min_score_in, const uint8_t* pat,
                     const uint32_t* pat_off, const uint8_t* txt, const uint32_t* txt_off, uint32_t count,
                     orc_sink* sinks, int threads)
{
    if (band < 1 || band > 32 || (alphabet != 2 && alphabet != 4 && alphabet != 5)) return -2;
    if (threads <= 0) threads = omp_get_max_threads();
    const int16_t min_score = (int16_t)min_score_in;                       /* :189 */
    #pragma omp parallel for schedule(dynamic,64) num_threads(threads)
    for (int64_t p = 0; p < (int64_t)count; ++p)
    {
        const uint8_t* P = pat + pat_off[p]; const uint32_t M = pat_off[p + 1] - pat_off[p];
        const uint8_t* T = txt + txt_off[p]; const uint32_t N = txt_off[p + 1] - txt_off[p];
        best_t best; best_init(&best);
        if (N >= M)                                                        /* :195-196 */
        {
            uint32_t B[5] = { 0, 0, 0, 0, 0 }, VP = 0xFFFFFFFFu, VN = 0, HP, HN;
            int dist = 0;                                                  /* -C with C = 0 (:203, :288) */
            const uint32_t last = (N - 1u < M) ? N - 1u : M;               /* min(text_len-1, pattern_len-C), uint32 (:211) */
            for (uint32_t i = 0; i < last; ++i)                            /* phase 1 (:212-219) */
            {
                for (uint32_t c = 0; c < 5; ++c) B[c] >>= 1;
                myers_load(B, alphabet, P[i], 1u << (band - 1));
                const uint32_t D0 = myers_column(myers_peq(B, alphabet, T[i]), &VP, &VN, &HP, &HN);
                dist -= 1 - (int)((D0 >> (band - 1)) & 1u);
            }
            int s = (int)(band - 1u + M - last);                           /* :222 */
            for (uint32_t i = last; i < N && s >= 0; ++i)                  /* phase 2 (:223-234) */
            {
                for (uint32_t c = 0; c < 5; ++c) B[c] >>= 1;
                myers_column(myers_peq(B, alphabet, T[i]), &VP, &VN, &HP, &HN);
                /* s == 32 (band 32, N == M) is an undefined shift in the reference (:179); its x86 host build shifts by s mod 32 */
                dist -= (int)((HP >> (s & 31)) & 1u) - (int)((HN >> (s & 31)) & 1u);
                if (type == T_SEMI && dist >= min_score) best_report(&best, dist, i + 1, M);
                --s;
            }
            if (type == T_GLOBAL && dist >= min_score) best_report(&best, dist, N, M);
        }
        sinks[p].score = best.score; sinks[p].x = best.x; sinks[p].y = best.y;
    }
    return 0;
}
