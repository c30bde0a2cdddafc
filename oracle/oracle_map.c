/*
 * oracle_map.c -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * CPU composition of the seed-and-extend pipeline (SURVEY 8f row 4, 8d "C4/C5 inputs") out of the
 * oracle's own entry points (oracle_abi.h): it is compiled twice, once over the plain-C restatement
 * (prefix orc_) and once over the reference's own headers (prefix ref_, oracle/_ref), so the
 * match / locate / banded score / banded traceback calls it makes are the pinned ones.
 *
 * The composition itself is OURS ("parity unpinned" for the glue): nvBowtie's driver
 * (nvBowtie/bowtie2/cuda/mapping_inl.h, locate_inl.h, score_inl.h, reduce_inl.h, traceback_inl.h) adds
 * seed-priority heuristics, re-seeding, effort limits and MAPQ that are out of scope.
 * What is kept from it, with the reference lines each step follows:
 *   - seeds: seed_len symbols every seed_interval symbols of the read and of its reverse complement,
 *     exact backward search of each (mapping_inl.h:73-86 -> match(fmi, seed, len))
 *   - hits: the first max_hits rows of every SA range (ranges wider than max_range are skipped),
 *     located by locate(fmi, row); the hit's diagonal is g_pos = pos - pos_in_read
 *     (locate_inl.h:130-136; clamped at 0 here where the reference lets the unsigned value wrap)
 *   - extension window: genome_begin = g_pos > band/2 ? g_pos - band/2 : 0,
 *     genome_end = min(genome_begin + band + read_len, genome_length)  (score_inl.h:103-106)
 *   - banded SEMI_GLOBAL scoring of the (strand) read against the window (score_inl.h:482-583
 *     -> batch_banded_alignment_score), candidates de-duplicated on (strand, genome_begin)
 *   - best / second-best reduction per read (reduce_inl.h), best = first maximum in
 *     (strand, genome_begin) order; reads whose best is below min_score stay unaligned
 *   - banded traceback of the best with 16-row checkpoints (traceback_inl.h:224-277, defs.h:95)
 */
#include "oracle_abi.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAP_INVALID (-(1 << 30))

static int cmp_u64(const void* a, const void* b)
{
    const uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
    return x < y ? -1 : x > y ? 1 : 0;
}

/* Hit selection by SA-range size. nvBowtie keeps a read's seed hits in a priority deque ordered by range size (hit_compare,
 * nvBowtie/bowtie2/cuda/seed_hit.h:220-230; seed_hit_deque_array.h:144-145); a push into a deque that already holds max_seed_hits
 * entries first drops the LARGEST range (mapping_inl.h:158-160,176-178); extension then takes rows from the smallest range first,
 * front to back (select_inl.h:93-113). sizes[n] in push (slot) order, 0 = no range. take[i] = rows range i contributes:
 * min(size, max_hits, what is left of max_rows) in visiting order; order[] (optional) = the visiting order, *n_order its length.
 * Ties between equal sizes: dropped = the latest slot, visited = the earliest slot first (the reference's interval heap leaves
 * them unspecified; pinned on the reference's own deque, ref_hit_order, for distinct sizes). Returns the total rows. */
size_t ORC(hit_priority)(const uint32_t* sizes, uint32_t n, uint32_t max_hits, uint32_t max_seed_hits, uint32_t max_rows,
                         uint32_t* take, uint32_t* order, uint32_t* n_order)
{
    uint8_t* in = (uint8_t*)calloc(n ? n : 1, 1);
    uint32_t held = 0; size_t total = 0;
    for (uint32_t t = 0; t < n; ++t)
    {
        take[t] = 0;
        if (!sizes[t]) continue;
        if (max_seed_hits && held == max_seed_hits)
        {
            uint32_t worst = 0, wsize = 0;
            for (uint32_t u = 0; u < t; ++u) if (in[u] && sizes[u] >= wsize) { wsize = sizes[u]; worst = u; }
            in[worst] = 0; --held;
        }
        in[t] = 1; ++held;
    }
    uint32_t budget = max_rows ? max_rows : 0xFFFFFFFFu;
    for (uint32_t done = 0; done < held; ++done)
    {
        uint32_t pick = 0, psize = 0xFFFFFFFFu;
        for (uint32_t u = 0; u < n; ++u) if (in[u] == 1 && sizes[u] < psize) { psize = sizes[u]; pick = u; }
        uint32_t c = psize < max_hits ? psize : max_hits;
        if (c > budget) c = budget;
        budget -= c; take[pick] = c; in[pick] = 2; total += c;
        if (order) order[done] = pick;
    }
    if (n_order) *n_order = held;
    free(in);
    return total;
}

int ORC(map_reads)(const orc_fmindex* fmi, const uint8_t* genome, const uint8_t* reads, const uint32_t* read_off,
                   uint32_t n_reads, const orc_map_params* p, orc_read_hit* hits, uint8_t* ops, uint32_t ops_stride,
                   uint32_t* n_candidates, int threads)
{
    if (!fmi || !genome || !read_off || !p || !hits) return -1;
    if (p->seed_len == 0 || p->seed_interval == 0 || p->max_hits == 0 || p->band < 2) return -1;
    const uint32_t G = fmi->length;
    int failed = 0;
    (void)threads;
#ifdef _OPENMP
    #pragma omp parallel num_threads(threads > 0 ? threads : omp_get_max_threads())
#endif
    {
        uint8_t*  strand_read = NULL; size_t cap_read = 0;
        uint64_t* keys = NULL;        size_t cap_keys = 0;
        uint8_t*  tb_ops = NULL;      size_t cap_ops = 0;
        uint32_t *sl_x = NULL, *sl_size = NULL, *sl_take = NULL, *sl_o = NULL, *ap = NULL; uint32_t* order = NULL; uint8_t* rseed = NULL; size_t cap_slots = 0, cap_order = 0;
#ifdef _OPENMP
        #pragma omp for schedule(dynamic, 64)
#endif
        for (uint32_t r = 0; r < n_reads; ++r)
        {
            const uint32_t L = read_off[r + 1] - read_off[r];
            const uint8_t* fwd = reads + read_off[r];
            orc_read_hit h; h.score = h.second_score = INT32_MIN; h.begin = h.end = 0xFFFFFFFFu; h.flags = 0; h.n_ops = 0;
            if (n_candidates) n_candidates[r] = 0;
            hits[r] = h;
            if (L == 0) continue;
            if (2 * (size_t)L > cap_read) { cap_read = 2 * (size_t)L; strand_read = (uint8_t*)realloc(strand_read, cap_read); }
            memcpy(strand_read, fwd, L);
            for (uint32_t i = 0; i < L; ++i) { const uint8_t c = fwd[L - 1 - i]; strand_read[L + i] = c < 4 ? (uint8_t)(3 - c) : c; }
            const uint32_t n_strands = p->both_strands ? 2u : 1u;
            const uint32_t n_seeds = L >= p->seed_len ? (L - p->seed_len) / p->seed_interval + 1 : 0;
            /* ---- the SA ranges of the read, in slot order (strand, seed, range): exact seeds give one, approximate seeds
             * (mapping_inl.h:114-184 through fm_match_approx, first R ranges in its push order) up to R */
            const uint32_t Rf = p->seed_mismatches ? (p->max_seed_ranges ? p->max_seed_ranges : 8u) : 1u;
            const orc_fmindex* rfmi = p->seed_mismatches ? p->reverse_index : NULL;
            const uint32_t R = rfmi ? 2 * Rf : Rf;              /* slots [0, Rf): forward-index ranges, [Rf, R): reverse-index ranges */
            const size_t n_slots = (size_t)n_strands * n_seeds * R;
            if (n_slots > cap_slots)
            {
                cap_slots = n_slots;
                sl_x = (uint32_t*)realloc(sl_x, 4 * cap_slots); sl_size = (uint32_t*)realloc(sl_size, 4 * cap_slots);
                sl_take = (uint32_t*)realloc(sl_take, 4 * cap_slots); sl_o = (uint32_t*)realloc(sl_o, 4 * cap_slots);
                rseed = (uint8_t*)realloc(rseed, p->seed_len + 1);
                ap = (uint32_t*)realloc(ap, 8 * (size_t)R + 8);
            }
            size_t ns = 0;
            for (uint32_t s = 0; s < n_strands; ++s)
                for (uint32_t k = 0; k < n_seeds; ++k)
                {
                    const uint32_t o = k * p->seed_interval;
                    const uint32_t off[2] = { 0, p->seed_len };
                    uint32_t n_rng = 0;
                    if (!p->seed_mismatches)
                    {
                        if (ORC(fm_match)(fmi, strand_read + s * L + o, off, 1, ap, NULL, 1)) { failed = 1; continue; }
                        n_rng = ap[0] <= ap[1] ? 1u : 0u;
                    }
                    else
                    {
                        if (ORC(fm_match_approx)(fmi, strand_read + s * L + o, off, 1, p->seed_len / 2u, 1, Rf, ap, &n_rng, 1)) { failed = 1; continue; }
                        if (n_rng > Rf) n_rng = Rf;
                    }
                    for (uint32_t j = 0; j < Rf; ++j, ++ns)
                    {
                        sl_o[ns] = o; sl_x[ns] = 0; sl_size[ns] = 0;
                        if (j >= n_rng) continue;
                        const uint32_t size = ap[2 * j + 1] - ap[2 * j] + 1;
                        if (p->max_range && size > p->max_range) continue;
                        sl_x[ns] = ap[2 * j]; sl_size[ns] = size;
                    }
                    if (rfmi)
                    {
                        /* the other half (mapping_inl.h:399-402, map<IGNORE_EXACT> on rfmi): the REVERSED seed on the index of the
                         * reversed text is consumed from the seed's start; exact over (seed_len + 1) / 2 steps, no exact match */
                        for (uint32_t e = 0; e < p->seed_len; ++e) rseed[e] = strand_read[s * L + o + p->seed_len - 1 - e];
                        uint32_t m_rng = 0;
                        if (ORC(fm_match_approx)(rfmi, rseed, off, 1, (p->seed_len + 1u) / 2u, 0, Rf, ap, &m_rng, 1)) { failed = 1; m_rng = 0; }
                        if (m_rng > Rf) m_rng = Rf;
                        for (uint32_t j = 0; j < Rf; ++j, ++ns)
                        {
                            sl_o[ns] = o; sl_x[ns] = 0; sl_size[ns] = 0;
                            if (j >= m_rng) continue;
                            const uint32_t size = ap[2 * j + 1] - ap[2 * j] + 1;
                            if (p->max_range && size > p->max_range) continue;
                            sl_x[ns] = ap[2 * j]; sl_size[ns] = size;
                        }
                    }
                }
            /* ---- rows each range contributes */
            size_t need = 0;
            if (!p->hit_priority)
                for (size_t t = 0; t < ns; ++t) { sl_take[t] = sl_size[t] < p->max_hits ? sl_size[t] : p->max_hits; need += sl_take[t]; }
            else
            {
                if (ns > cap_order) { cap_order = ns; order = (uint32_t*)realloc(order, 4 * cap_order); }
                need = ORC(hit_priority)(sl_size, (uint32_t)ns, p->max_hits, p->max_seed_hits, p->max_rows, sl_take, order, NULL);
            }
            if (need > cap_keys) { cap_keys = need; keys = (uint64_t*)realloc(keys, sizeof(uint64_t) * cap_keys); }
            size_t nk = 0;
            for (size_t t = 0; t < ns; ++t)
            {
                const uint32_t s = (uint32_t)(t / ((size_t)n_seeds * R)), o = sl_o[t];
                const int rev = (t % R) >= Rf;
                for (uint32_t q = 0; q < sl_take[t]; ++q)
                {
                    const uint32_t row = sl_x[t] + q;
                    uint32_t pos;
                    if (ORC(fm_locate)(rev ? rfmi : fmi, &row, 1, &pos, NULL, 1)) { failed = 1; continue; }
                    if (rev) pos = G - pos - p->seed_len;           /* the reversed seed at q in the reversed text = the seed at G - q - seed_len */
                    const uint32_t g_pos = pos > o ? pos - o : 0;
                    const uint32_t gb = g_pos > p->band / 2 ? g_pos - p->band / 2 : 0;
                    keys[nk++] = ((uint64_t)s << 32) | gb;
                }
            }
            if (nk == 0) continue;
            qsort(keys, nk, sizeof(uint64_t), cmp_u64);
            size_t nu = 1;
            for (size_t i = 1; i < nk; ++i) if (keys[i] != keys[nu - 1]) keys[nu++] = keys[i];
            if (n_candidates) n_candidates[r] = (uint32_t)nu;
            int32_t best = INT32_MIN, second = INT32_MIN; size_t best_i = 0;
            for (size_t i = 0; i < nu; ++i)
            {
                const uint32_t s = (uint32_t)(keys[i] >> 32), gb = (uint32_t)keys[i];
                const uint64_t ge64 = (uint64_t)gb + p->band + L;
                const uint32_t ge = ge64 < G ? (uint32_t)ge64 : G;
                const uint32_t poff[2] = { 0, L }, toff[2] = { 0, ge - gb };
                orc_sink sink;
                if (ORC(banded_score)(p->aligner, p->band, strand_read + s * L, poff, genome + gb, toff, 1, &sink, 1)) { failed = 1; continue; }
                if (sink.score <= MAP_INVALID) continue;
                if (sink.score > best) { if (best != INT32_MIN) second = best > second ? best : second; best = sink.score; best_i = i; }
                else if (sink.score > second) second = sink.score;
            }
            if (best == INT32_MIN || best < p->min_score) continue;
            {
                const uint32_t s = (uint32_t)(keys[best_i] >> 32), gb = (uint32_t)keys[best_i];
                const uint64_t ge64 = (uint64_t)gb + p->band + L;
                const uint32_t ge = ge64 < G ? (uint32_t)ge64 : G;
                const uint32_t poff[2] = { 0, L }, toff[2] = { 0, ge - gb };
                const uint32_t stride = 2 * L + p->band + 16;
                if (stride > cap_ops) { cap_ops = stride; tb_ops = (uint8_t*)realloc(tb_ops, cap_ops); }
                orc_alignment a; uint32_t n_ops = 0;
                if (ORC(banded_traceback)(p->aligner, p->band, 16, strand_read + s * L, poff, genome + gb, toff, 1, &a, tb_ops, stride, &n_ops, 1)) { failed = 1; continue; }
                h.score = a.score; h.second_score = second;
                h.begin = gb + a.src_x; h.end = gb + a.sink_x;
                h.flags = 2u | s; h.n_ops = n_ops;
                hits[r] = h;
                if (ops && ops_stride) memcpy(ops + (size_t)r * ops_stride, tb_ops, n_ops < ops_stride ? n_ops : ops_stride);
            }
        }
        free(strand_read); free(keys); free(tb_ops); free(sl_x); free(sl_size); free(sl_take); free(sl_o); free(order); free(ap); free(rseed);
    }
    return failed ? -2 : 0;
}

/* mate pairing over independently mapped mates (read 2 i / 2 i + 1): forward-reverse policy with a fragment-length window
 * (nvBowtie/bowtie2/cuda/params.h:147-152); the composition is ours, see the header */
int ORC(pair_hits)(const orc_read_hit* hits, uint32_t n_pairs, uint32_t min_frag, uint32_t max_frag, orc_pair_hit* out)
{
    for (uint32_t i = 0; i < n_pairs; ++i)
    {
        const orc_read_hit a = hits[2 * (size_t)i], b = hits[2 * (size_t)i + 1];
        const int al1 = (a.flags & 2u) != 0, al2 = (b.flags & 2u) != 0;
        orc_pair_hit o; o.score = INT32_MIN; o.frag_begin = o.frag_end = 0xFFFFFFFFu; o.flags = (al1 ? 1u : 0u) | (al2 ? 2u : 0u);
        if (al1 && al2)
        {
            const int r1 = (a.flags & 1u) != 0, r2 = (b.flags & 1u) != 0;
            const orc_read_hit* f = r1 ? &b : &a; const orc_read_hit* r = r1 ? &a : &b;
            o.score = a.score + b.score;
            o.frag_begin = a.begin < b.begin ? a.begin : b.begin;
            o.frag_end = a.end > b.end ? a.end : b.end;
            const uint32_t frag = r->end >= f->begin ? r->end - f->begin : 0u;
            if (r1 != r2 && f->begin <= r->begin && f->end <= r->end && frag >= min_frag && frag <= max_frag) o.flags |= 4u;
        }
        else if (al1) { o.score = a.score; o.frag_begin = a.begin; o.frag_end = a.end; }
        else if (al2) { o.score = b.score; o.frag_begin = b.begin; o.frag_end = b.end; }
        out[i] = o;
    }
    return 0;
}

/* nvBowtie's approximate seed mapping, `map<find_exact>` (nvBowtie/bowtie2/cuda/mapping_inl.h:114-184): the seed is matched
 * exactly over its first len1 search steps and with at most one substitution in the remaining ones, using rank4 for the four
 * candidate symbols of a step (:146) and match_range (:74-86) for the tail behind a substitution. Search order here is the order
 * of nvb_fm_match / match(): step t consumes symbol q[len - 1 - t] (the reference iterates a reversed seed stream). Every
 * non-empty SA range is emitted in the reference's push order (substitutions by step, then by symbol; the exact range last when
 * find_exact); counts[q] = ranges found, of which the first min(counts, cap) are stored. Composed of the pinned entry points
 * ORC(fm_rank4) / ORC(fm_rank), compiled over the C restatement and over the reference's own headers. */
static void ORC(approx_tail)(const orc_fmindex* fmi, const uint8_t* q, uint32_t len, uint32_t t0, uint32_t* x, uint32_t* y)
{
    for (uint32_t t = t0; t < len && *x <= *y; ++t)                                    /* match_range, :74-86 */
    {
        const uint8_t c = q[len - 1 - t];
        if (c > 3) { *x = 1; *y = 0; return; }
        const uint32_t rg[2] = { *x - 1u, *y }; uint32_t r[2];
        ORC(fm_rank)(fmi, rg, &c, 1, r, 1);
        *x = fmi->L2[c] + r[0] + 1u; *y = fmi->L2[c] + r[1];
    }
}

int ORC(fm_match_approx)(const orc_fmindex* fmi, const uint8_t* symbols, const uint32_t* offsets, uint32_t count, uint32_t len1,
                         int find_exact, uint32_t cap, uint32_t* ranges, uint32_t* counts, int threads)
{
    if (!fmi || !offsets || !counts || (cap && !ranges)) return -1;
    (void)threads;
#ifdef _OPENMP
    #pragma omp parallel for schedule(dynamic, 64) num_threads(threads > 0 ? threads : omp_get_max_threads())
#endif
    for (int64_t qi = 0; qi < (int64_t)count; ++qi)
    {
        const uint8_t* q = symbols + offsets[qi]; const uint32_t len = offsets[qi + 1] - offsets[qi];
        const uint32_t l1 = len1 < len ? len1 : len;
        uint32_t* out = ranges + 2 * (size_t)qi * cap; uint32_t n = 0;
        uint32_t bx = 0, by = fmi->length;
        ORC(approx_tail)(fmi, q + (len - l1), l1, 0, &bx, &by);                         /* exact over the first len1 steps, :136-137 */
        for (uint32_t t = l1; t < len && bx <= by; ++t)                                /* :140-173 */
        {
            const uint8_t c = q[len - 1 - t];
            const uint32_t k[2] = { bx - 1u, by }; uint32_t occ[8];
            ORC(fm_rank4)(fmi, k, 2, occ, 1);                                          /* :146 */
            for (uint32_t sub = 0; sub < 4; ++sub)
                if (sub != c && occ[4 + sub] > occ[sub])
                {
                    uint32_t x = fmi->L2[sub] + occ[sub] + 1u, y = fmi->L2[sub] + occ[4 + sub];
                    ORC(approx_tail)(fmi, q, len, t + 1, &x, &y);
                    if (x <= y) { if (n < cap) { out[2 * n] = x; out[2 * n + 1] = y; } ++n; }
                }
            if (c < 4) { bx = fmi->L2[c] + occ[c] + 1u; by = fmi->L2[c] + occ[4 + c]; }
            else { bx = 1; by = 0; break; }
        }
        if (find_exact && bx <= by) { if (n < cap) { out[2 * n] = bx; out[2 * n + 1] = by; } ++n; }
        counts[qi] = n;
    }
    return 0;
}

/*
 * MDS byte-code + edit distance of one alignment: a literal restatement of the CIGAR loop of finish_alignment_kernel
 * (nvBowtie/bowtie2/cuda/traceback_inl.h:555-690; the reference function is a __global__ kernel over its pipeline state, not
 * host-callable: "parity unpinned", checked against hand-worked vectors and against the reference's own consumer,
 * nvbio/io/output/output_sam.cpp:216-289, restated in tests). ops: push order (the last op of the alignment first), as the
 * reference's cigar_vector is indexed back to front (:574-575); runs longer than a 14-bit io::Cigar length are split.
 * (j0, k0): first aligned pattern / text symbol (the reference starts at j = 0, k = cigar_offset, :572).
 */
int ORC(ops_to_mds)(const uint8_t* ops, uint32_t n_ops, const uint8_t* pattern, const uint8_t* text, uint32_t j0, uint32_t k0,
                    uint8_t* mds, uint32_t cap, uint32_t* mds_len_out, uint32_t* ed_out)
{
    enum { SUB = 0, INS = 1, DEL = 2, CLP = 3, MDS_MATCH = 0, MDS_MISMATCH = 1, MDS_INSERTION = 2, MDS_DELETION = 3, MDS_INVALID = 4 };
    uint8_t* local = (uint8_t*)malloc(2 + 3 * (size_t)n_ops + 8);
    if (!local) return -1;
    uint32_t mds_len = 2, ed = 0; uint8_t mds_op = MDS_INVALID;                              /* :566-570 */
    uint32_t j = j0, k = k0, i = n_ops;
    while (i > 0)
    {
        const uint32_t t = ops[i - 1] & 3u; uint32_t l = 1;
        while (l < i && (ops[i - 1 - l] & 3u) == t && l < 16383u) ++l;
        i -= l;
        if (t == INS || t == DEL || t == CLP)                                                 /* :582-590 */
        {
            mds_op = t == DEL ? MDS_DELETION : MDS_INSERTION;
            local[mds_len++] = mds_op;
            local[mds_len++] = (uint8_t)l;
        }
        for (uint32_t n = 0; n < l; ++n)                                                      /* :592-641 */
        {
            j += (t == SUB || t == INS || t == CLP) ? 1u : 0u;
            k += (t == SUB || t == DEL) ? 1u : 0u;
            const uint8_t readc = j > 0 ? pattern[j - 1] : 255u;
            const uint8_t refc  = k > 0 ? text[k - 1] : 255u;
            if (t == SUB)
            {
                if (readc == refc)
                {
                    if (mds_op == MDS_MATCH && local[mds_len - 1] < 255) local[mds_len - 1]++;
                    else { local[mds_len++] = mds_op = MDS_MATCH; local[mds_len++] = 1; }
                }
                else { local[mds_len++] = mds_op = MDS_MISMATCH; local[mds_len++] = readc; ++ed; }
            }
            else
            {
                local[mds_len++] = t == DEL ? refc : readc;
                if (t != CLP) ++ed;
            }
        }
    }
    local[0] = (uint8_t)(mds_len & 0xFF); local[1] = (uint8_t)(mds_len >> 8);               /* :684-685 */
    memcpy(mds, local, mds_len < cap ? mds_len : cap);
    free(local);
    if (mds_len_out) *mds_len_out = mds_len;
    if (ed_out) *ed_out = ed;
    return 0;
}

/*
 * Opposite-mate window search for pairs with exactly one aligned mate: composition of the pinned full-matrix score / traceback
 * entry points over the window nvBowtie frames for the opposite mate (BestOppositeScoreStream::init_context,
 * nvBowtie/bowtie2/cuda/score_inl.h:318-411, PE_POLICY_FR, pe_overlap): anchor forward -> [anchor.begin, anchor.begin + max_frag),
 * mate reverse-complemented (:398-404); anchor reverse -> [anchor.end - max_frag, anchor.end), mate forward (:391-397); clipped to
 * the genome (:405). The pair-score targets and the max_text_gaps trimming (:341-392) are left out -- the composition is ours.
 * hits[2 n_pairs] is updated in place (flags |= 4 on rescued mates), ops likewise (truncated to ops_stride).
 */
int ORC(rescue_mates)(const uint8_t* genome, uint32_t G, const uint8_t* reads, const uint32_t* read_off, uint32_t n_pairs,
                      orc_aligner aligner, int32_t min_score, uint32_t max_frag, orc_read_hit* hits, uint8_t* ops, uint32_t ops_stride,
                      uint32_t* n_candidates, uint32_t* n_rescued)
{
    uint32_t nc = 0, nr = 0; int failed = 0;
    for (uint32_t i = 0; i < n_pairs; ++i)
    {
        const orc_read_hit a = hits[2 * (size_t)i], b = hits[2 * (size_t)i + 1];
        const int al1 = (a.flags & 2u) != 0, al2 = (b.flags & 2u) != 0;
        if (al1 == al2) continue;
        const orc_read_hit an = al1 ? a : b;
        const uint32_t read = 2 * i + (al1 ? 1u : 0u);
        uint32_t strand, gb, ge;
        if (!(an.flags & 1u)) { strand = 1; gb = an.begin; const uint64_t e = (uint64_t)an.begin + max_frag; ge = (uint32_t)(e < G ? e : G); }
        else                  { strand = 0; ge = an.end < G ? an.end : G; gb = ge > max_frag ? ge - max_frag : 0; }
        const uint32_t L = read_off[read + 1] - read_off[read];
        if (ge <= gb || L == 0 || L > ge - gb) continue;
        ++nc;
        uint8_t* pat = (uint8_t*)malloc(L);
        const uint8_t* fwd = reads + read_off[read];
        for (uint32_t k = 0; k < L; ++k) { const uint8_t c = strand ? fwd[L - 1 - k] : fwd[k]; pat[k] = (strand && c < 4) ? (uint8_t)(3 - c) : c; }
        const uint32_t poff[2] = { 0, L }, toff[2] = { 0, ge - gb };
        orc_sink sink;
        if (ORC(score)(aligner, pat, poff, genome + gb, toff, 1, &sink, 1)) { failed = 1; free(pat); continue; }
        if (sink.score > MAP_INVALID && sink.score >= min_score)
        {
            const uint32_t stride = L + (ge - gb) + 16;
            uint8_t* tb = (uint8_t*)malloc(stride);
            orc_alignment al; uint32_t n_ops = 0;
            if (ORC(traceback)(aligner, 32, pat, poff, genome + gb, toff, 1, &al, tb, stride, &n_ops, 1)) failed = 1;
            else
            {
                orc_read_hit h; h.score = al.score; h.second_score = INT32_MIN; h.begin = gb + al.src_x; h.end = gb + al.sink_x;
                h.flags = 2u | 4u | strand; h.n_ops = n_ops;
                hits[read] = h;
                if (ops && ops_stride)
                {
                    memset(ops + (size_t)read * ops_stride, 0, ops_stride);
                    memcpy(ops + (size_t)read * ops_stride, tb, n_ops < ops_stride ? n_ops : ops_stride);
                }
                ++nr;
            }
            free(tb);
        }
        free(pat);
    }
    if (n_candidates) *n_candidates = nc;
    if (n_rescued) *n_rescued = nr;
    return failed ? -2 : 0;
}

/*
 * Mapping quality: BowtieMapq2::operator() (nvBowtie/bowtie2/cuda/mapq.h:145-316; Bowtie 2's "V2" calculator), restated as tables.
 * best / second: alignment scores (a pair: the sum of both mates, has_second = 0 there: our reduction keeps per-mate runners-up only);
 * perfect / min_score: the scheme's perfect_score() and min_score() for the read (pair: sums), which the CALLER evaluates
 * (min_score(len) is a float expression of the scheme, nvBowtie/bowtie2/cuda/params.h:95-100). All comparisons are made on floats in
 * the reference's order. Pinned on the compiled calculator (oracle/ref_wrap.cpp::ref_bowtie_mapq2).
 */
int ORC(mapq_v2)(int32_t best_i, int32_t has_second, int32_t second_i, int32_t perfect, int32_t min_score_i, int32_t monotone)
{
    /* fraction of the score range the best score must reach, value returned (no runner-up); last entry = otherwise */
    static const float uniq_t[6] = { 0.8f, 0.7f, 0.6f, 0.5f, 0.4f, 0.3f };
    static const int   uniq_v[2][7] = { { 44, 42, 41, 36, 28, 24, 22 }, { 42, 40, 24, 23, 8, 3, 0 } };           /* [local, end-to-end] */
    /* runner-up present: bands of |best| - |second| over the range, then tiers of the best score inside a band */
    static const float top_t[4] = { 0.9f, 0.8f, 0.7f, 0.6f };
    static const int   top_v[2][4][2] = { { {40,40}, {39,39}, {38,38}, {37,37} }, { {39,33}, {38,27}, {37,26}, {36,22} } };   /* {best == range, else} */
    static const float band_t[5] = { 0.5f, 0.4f, 0.3f, 0.2f, 0.1f };
    static const float tier_t[2][5][2] = { { {.5f,.5f}, {.5f,.5f}, {.5f,.5f}, {.5f,.5f}, {.5f,.5f} },
                                           { {.84f,.68f}, {.84f,.68f}, {.88f,.67f}, {.88f,.67f}, {.88f,.67f} } };
    static const int   tier_v[2][5][4] = { { {35,25,25,20}, {34,21,21,19}, {33,18,18,16}, {32,17,17,12}, {31,14,14,9} },
                                           { {35,25,16,5},  {34,21,14,4},  {32,18,15,3},  {31,17,11,0},  {30,12,7,0} } };
    static const float low_t[2] = { 0.5f, 0.67f };
    static const int   low_v[2][2][2] = { { {11, 2}, {1, 0} }, { {6, 2}, {1, 0} } };                           /* [diff > 0, diff == 0][reached, not] */
    const int m = monotone ? 1 : 0;
    const float max_score = (float)perfect, min_score = (float)min_score_i;
    const float diff = max_score - min_score, best = (float)best_i;
    if (best < min_score) return 0;                                                                           /* :163-164 */
    const float best_over = best - min_score;
    if (!has_second)
    {
        for (int i = 0; i < 6; ++i) if (best_over >= diff * uniq_t[i]) return uniq_v[m][i];
        return uniq_v[m][6];
    }
    const float second = (float)second_i;
    float a = best < 0 ? -best : best, b = second < 0 ? -second : second;
    float best_diff = a - b; if (best_diff < 0) best_diff = -best_diff;                                       /* fabsf(fabsf(best) - fabsf(second)) */
    for (int i = 0; i < 4; ++i) if (best_diff >= diff * top_t[i]) return top_v[m][i][best_over == diff ? 0 : 1];
    for (int i = 0; i < 5; ++i)
        if (best_diff >= diff * band_t[i])
        {
            if (best_over == diff) return tier_v[m][i][0];
            if (best_over >= diff * tier_t[m][i][0]) return tier_v[m][i][1];
            if (best_over >= diff * tier_t[m][i][1]) return tier_v[m][i][2];
            return tier_v[m][i][3];
        }
    return low_v[m][best_diff > 0 ? 0 : 1][best_over >= diff * low_t[m] ? 0 : 1];
}
