/*
 * oracle_fm.c -- TEST INFRASTRUCTURE ONLY: plain-C restatement of the reference's FM-index path.
 *
 * PARITY PINNED: checked bit-for-bit against oracle/_ref/libnvbio_ref.so (the reference's own
 * headers compiled from /root/reference) by tests/test_oracle_vs_ref.py, and against the
 * reference's property tests (rank vs running counts, match/locate vs the true suffix array:
 * nvbio-test/rank_test.cu:46-79, nvbio-test/fmindex_test.cu:576-656) by tests/test_oracle_fm.py.
 *
 * Never linked into the product library; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg load it.
 *
 * Layout restated (nvBWT, nvbio/io/fmindex/fmindex.h:81-86, fmindex_impl.cu:296-318):
 *   2-bit symbols, 16 per uint32 word, BIG-endian inside the word (symbol 0 in bits 31:30);
 *   per 64 symbols one 32-byte block = 4 BWT words followed by 4 uint32 counters (A,C,G,T)
 *   holding the counts BEFORE the block (rank_dictionary_inl.h:34-66);
 *   L2[c+1] = L2[c] + count(c); SSA keeps SA[i] for i % 16 == 0 with ssa[0] = 0xFFFFFFFF.
 */
#include "oracle_abi.h"
#include <stdlib.h>
#include <string.h>
#include <omp.h>

static uint32_t words_for(uint32_t n) { return ((n + 15u) / 16u + 3u) & ~3u; }

size_t orc_bwt_occ_words(uint32_t n) { return (size_t)words_for(n) * 2u; }
size_t orc_ssa_words(uint32_t n)     { return ((size_t)n + 16u) / 16u; }

/* ------------------------------------------------------------------------------------------ */
/* suffix array by prefix doubling with an LSD radix sort (any correct SA gives the same BWT)  */
/* ------------------------------------------------------------------------------------------ */
typedef struct { uint64_t key; uint32_t idx; } kv_t;

static void radix_sort_kv(kv_t* a, kv_t* tmp, size_t n, int key_bits)
{
    for (int shift = 0; shift < key_bits; shift += 11)
    {
        size_t hist[2049];
        memset(hist, 0, sizeof(hist));
        for (size_t i = 0; i < n; ++i) hist[((a[i].key >> shift) & 2047u) + 1]++;
        for (int b = 0; b < 2048; ++b) hist[b + 1] += hist[b];
        for (size_t i = 0; i < n; ++i) tmp[hist[(a[i].key >> shift) & 2047u]++] = a[i];
        kv_t* t = a; a = tmp; tmp = t;
    }
    /* the caller passes an even/odd-agnostic pair: copy back if the result sits in tmp */
    int passes = (key_bits + 10) / 11;
    if (passes & 1) memcpy(tmp, a, n * sizeof(kv_t)); /* after swap: 'tmp' is the original a */
}

static int bits_for(uint64_t v) { int b = 1; while ((v >> b) != 0 && b < 64) ++b; return b; }

/* sa[0..n] over the n+1 suffixes (the empty suffix sorts first) */
static void build_sa(const uint8_t* t, uint32_t n, uint32_t* sa)
{
    const size_t m = (size_t)n + 1u;
    kv_t* a   = (kv_t*)malloc(m * sizeof(kv_t));
    kv_t* tmp = (kv_t*)malloc(m * sizeof(kv_t));
    uint32_t* rank = (uint32_t*)malloc(m * sizeof(uint32_t));

    /* initial key: first 16 symbols (zero padded) then the remaining length capped at 16 */
    for (size_t i = 0; i < m; ++i)
    {
        uint64_t k = 0;
        const uint32_t rem = n - (uint32_t)i;
        for (uint32_t j = 0; j < 16; ++j)
            k = (k << 2) | (j < rem ? (uint64_t)(t[i + j] & 3u) : 0u);
        a[i].key = (k << 5) | (rem < 16u ? rem : 16u);
        a[i].idx = (uint32_t)i;
    }
    radix_sort_kv(a, tmp, m, 37);

    uint32_t h = 16;
    for (;;)
    {
        /* dense ranks = index of the first member of each equal-key group */
        int all_unique = 1;
        uint32_t r = 0;
        for (size_t i = 0; i < m; ++i)
        {
            if (i > 0 && a[i].key != a[i - 1].key) r = (uint32_t)i;
            else if (i > 0) all_unique = 0;
            rank[a[i].idx] = r;
        }
        if (all_unique || h > n) break;
        const int rb = bits_for((uint64_t)m + 1u);
        for (size_t i = 0; i < m; ++i)
        {
            const uint32_t s = a[i].idx;
            const uint64_t r2 = ((uint64_t)s + h <= n) ? (uint64_t)rank[s + h] + 1u : 0u;
            a[i].key = ((uint64_t)rank[s] << rb) | r2;
        }
        radix_sort_kv(a, tmp, m, 2 * rb);
        h *= 2;
    }
    for (size_t i = 0; i < m; ++i) sa[i] = a[i].idx;
    free(a); free(tmp); free(rank);
}

/* ------------------------------------------------------------------------------------------ */
/* index construction: nvbio/fmindex/bwt.h:44-55 (BWT from SA), rank_dictionary_inl.h:34-66     */
/* (occurrence table), io/fmindex/fmindex_impl.cu:296-318 (fusing + L2), ssa_inl.h:255-268     */
/* ------------------------------------------------------------------------------------------ */
int orc_build_index(const uint8_t* text, uint32_t n, uint32_t* primary, uint32_t* L2,
                    uint32_t* bwt_occ, uint32_t* ssa, int32_t* sa_out)
{
    const uint32_t W = words_for(n);
    uint32_t* sa = (uint32_t*)malloc(((size_t)n + 1u) * sizeof(uint32_t));
    uint8_t* bwt = (uint8_t*)malloc((size_t)n + 1u);
    build_sa(text, n, sa);

    /* the row whose suffix is the whole text has no preceding symbol: it is dropped */
    uint32_t prim = 0, o = 0;
    for (uint32_t i = 0; i <= n; ++i)
    {
        if (sa[i] == 0) prim = i;
        else bwt[o++] = text[sa[i] - 1] & 3u;
    }
    *primary = prim;

    memset(bwt_occ, 0, (size_t)W * 2u * sizeof(uint32_t));
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    for (uint32_t i = 0; i < n; ++i)
    {
        uint32_t* blk = bwt_occ + (size_t)(i >> 6) * 8u;
        if ((i & 63u) == 0)
            for (int c = 0; c < 4; ++c) blk[4 + c] = cnt[c];
        blk[(i & 63u) >> 4] |= (uint32_t)bwt[i] << (30u - 2u * (i & 15u));
        cnt[bwt[i]]++;
    }
    L2[0] = 0;
    for (int c = 0; c < 4; ++c) L2[c + 1] = L2[c] + cnt[c];

    const uint32_t n_ssa = (uint32_t)orc_ssa_words(n);
    for (uint32_t i = 0; i < n_ssa; ++i) ssa[i] = sa[(size_t)i * 16u];
    ssa[0] = 0xFFFFFFFFu;
    if (sa_out)
    {
        for (uint32_t i = 0; i <= n; ++i) sa_out[i] = (int32_t)sa[i];
        sa_out[0] = -1;
    }
    free(sa); free(bwt);
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* rank: nvbio/basic/popcount_inl.h:230-236,318-341 ; rank_dictionary_inl.h:351-448            */
/* ------------------------------------------------------------------------------------------ */
static inline uint32_t popc_sym(uint32_t w, uint32_t c)
{
    const uint32_t odd  = ((c & 2u) ? w : ~w) >> 1;
    const uint32_t even = ((c & 1u) ? w : ~w);
    return (uint32_t)__builtin_popcount(odd & even & 0x55555555u);
}
/* occurrences of c among the symbols of w except the LAST `skip` ones (low bits) */
static inline uint32_t popc_sym_head(uint32_t w, uint32_t c, uint32_t skip)
{
    const uint32_t r = popc_sym(w & ~((1u << (skip << 1)) - 1u), c);
    return c == 0 ? r - skip : r; /* the cleared symbols read as 'A' */
}

/* occurrences of c in bwt[0..k] (inclusive); k == -1 -> 0 */
static inline uint32_t dict_rank(const uint32_t* bwt_occ, uint32_t k, uint32_t c)
{
    if (k == 0xFFFFFFFFu) return 0;
    const uint32_t* blk = bwt_occ + (size_t)(k >> 6) * 8u;
    const uint32_t m = (k & 63u) >> 4;
    uint32_t r = blk[4 + c];
    for (uint32_t w = 0; w < m; ++w) r += popc_sym(blk[w], c);
    return r + popc_sym_head(blk[m], c, ~k & 15u);
}

/* FM-index rank with the primary shift: nvbio/fmindex/fmindex_inl.h:31-47 */
static inline uint32_t fm_rank1(const orc_fmindex* f, uint32_t k, uint32_t c, uint32_t* nblk)
{
    if (k == 0xFFFFFFFFu) return 0;
    if (k == f->length) return f->L2[c + 1] - f->L2[c];
    if (k >= f->primary) --k;
    if (nblk) ++*nblk;
    return dict_rank(f->bwt_occ, k, c);
}

/* range version: nvbio/fmindex/fmindex_inl.h:60-88 + rank_dictionary_inl.h:425-448 */
static inline void fm_rank2(const orc_fmindex* f, uint32_t x, uint32_t y, uint32_t c,
                            uint32_t* rx, uint32_t* ry, uint32_t* nblk)
{
    if (x == y) { *rx = *ry = fm_rank1(f, x, c, nblk); return; }
    if (x == 0xFFFFFFFFu) { *rx = 0; *ry = fm_rank1(f, y, c, nblk); return; }
    if (y == f->length) { *rx = fm_rank1(f, x, c, nblk); *ry = f->L2[c + 1] - f->L2[c]; return; }
    if (x >= f->primary) --x;
    if (y >= f->primary) --y;
    if (x == y) /* the dictionary's own single-index shortcut */
    {
        if (nblk) ++*nblk;
        *rx = *ry = dict_rank(f->bwt_occ, y, c);
        return;
    }
    if (nblk) *nblk += (x == 0xFFFFFFFFu || (x >> 6) == (y >> 6)) ? 1u : 2u; /* x == -1 only if primary == 0 */
    *rx = dict_rank(f->bwt_occ, x, c);
    *ry = dict_rank(f->bwt_occ, y, c);
}

/* backward search: nvbio/fmindex/fmindex_inl.h:186-239 */
int orc_fm_match(const orc_fmindex* f, const uint8_t* symbols, const uint32_t* offsets,
                 uint32_t count, uint32_t* ranges, uint32_t* blocks, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64_t q = 0; q < (int64_t)count; ++q)
    {
        const uint8_t* p = symbols + offsets[q];
        const uint32_t len = offsets[q + 1] - offsets[q];
        uint32_t x = 0, y = f->length, nb = 0;
        for (int32_t i = (int32_t)len - 1; i >= 0 && x <= y; --i)
        {
            const uint32_t c = p[i];
            if (c > 3) { x = 1; y = 0; break; }
            uint32_t rx, ry;
            fm_rank2(f, x - 1u, y, c, &rx, &ry, &nb);
            x = f->L2[c] + rx + 1u;
            y = f->L2[c] + ry;
        }
        ranges[2 * q] = x; ranges[2 * q + 1] = y;
        if (blocks) blocks[q] = nb;
    }
    return 0;
}

/* forward search over the pattern = backward search of the reversed pattern: nvbio/fmindex/fmindex_inl.h:252-278 */
int orc_fm_match_reverse(const orc_fmindex* f, const uint8_t* symbols, const uint32_t* offsets,
                         uint32_t count, uint32_t* ranges, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64_t q = 0; q < (int64_t)count; ++q)
    {
        const uint8_t* p = symbols + offsets[q];
        const uint32_t len = offsets[q + 1] - offsets[q];
        uint32_t x = 0, y = f->length;
        for (uint32_t i = 0; i < len && x <= y; ++i)
        {
            const uint32_t c = p[i];
            if (c > 3) { x = 1; y = 0; break; }
            uint32_t rx, ry;
            fm_rank2(f, x - 1u, y, c, &rx, &ry, NULL);
            x = f->L2[c] + rx + 1u;
            y = f->L2[c] + ry;
        }
        ranges[2 * q] = x; ranges[2 * q + 1] = y;
    }
    return 0;
}

int orc_fm_rank(const orc_fmindex* f, const uint32_t* ranges, const uint8_t* c, uint32_t count,
                uint32_t* out, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64_t q = 0; q < (int64_t)count; ++q)
        fm_rank2(f, ranges[2 * q], ranges[2 * q + 1], c[q], &out[2 * q], &out[2 * q + 1], NULL);
    return 0;
}

/* rank4: nvbio/fmindex/fmindex_inl.h:100-125 ; all four symbols at once */
int orc_fm_rank4(const orc_fmindex* f, const uint32_t* k, uint32_t count, uint32_t* out, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64_t q = 0; q < (int64_t)count; ++q)
        for (uint32_t c = 0; c < 4; ++c) out[4 * q + c] = fm_rank1(f, k[q], c, NULL);
    return 0;
}

/* symbol i of the packed BWT */
static inline uint32_t bwt_at(const uint32_t* bwt_occ, uint32_t i)
{
    return (bwt_occ[(size_t)(i >> 6) * 8u + ((i & 63u) >> 4)] >> (30u - 2u * (i & 15u))) & 3u;
}

/* locate: nvbio/fmindex/fmindex_inl.h:364-394, SSA fetch ssa_inl.h:478-487 */
int orc_fm_locate(const orc_fmindex* f, const uint32_t* rows, uint32_t count, uint32_t* pos,
                  uint32_t* steps, int threads)
{
    if (threads <= 0) threads = omp_get_max_threads();
    #pragma omp parallel for schedule(dynamic,256) num_threads(threads)
    for (int64_t q = 0; q < (int64_t)count; ++q)
    {
        uint32_t j = rows[q], t = 0;
        while ((j & 15u) != 0)
        {
            if (j != f->primary)
            {
                const uint32_t c = bwt_at(f->bwt_occ, j < f->primary ? j : j - 1u);
                j = f->L2[c] + fm_rank1(f, j, c, NULL);
            }
            else j = 0;
            ++t;
        }
        pos[q] = f->ssa[j >> 4] + t;
        if (steps) steps[q] = t;
    }
    return 0;
}
