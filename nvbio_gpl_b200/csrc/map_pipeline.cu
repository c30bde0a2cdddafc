// map_pipeline.cu -- seed -> locate -> extend -> reduce -> traceback over a batch of reads.
//
// The glue between the two halves of the hot path, as nvBowtie's driver arranges them:
//   map_seed_kernel     exact backward search of every seed of every read, both strands
//                       (nvBowtie/bowtie2/cuda/mapping_inl.h:73-86 -> match(fmi, seed, len), fmindex_inl.h:186-239)
//   map_locate_kernel   one thread per selected SA row: locate (fmindex_inl.h:290-304), diagonal of the hit
//                       g_pos = pos - pos_in_read (locate_inl.h:130-136), window begin g_pos - band/2 (score_inl.h:103-104)
//   sort + unique       candidates = distinct (read, strand, window begin)
//   map_extract_kernel  strand read and 2-bit genome window -> 8-bit string sets for the banded DP kernels
//   nvb_batch_banded_score / nvb_batch_banded_traceback   the extension itself (dp_api.cu)
//   map_reduce_kernel   best / second best window per read (reduce_inl.h), first maximum in (strand, begin) order
//   map_finish_kernel   genome coordinates + op list of the best window
// Everything a read needs stays on the device; the host only learns three counters per batch (seed hits,
// candidates, aligned reads) to size the next launch.
#include "fm_device.cuh"
#include "dp_fast.h"
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <cstdio>
#include <cstdlib>
#include <climits>
#include <chrono>
#include <vector>

#define MAP_THREADS 256
#define MAP_NONE    0xFFFFFFFFu
#define MAP_INVALID (-(1 << 30))

enum { S_SEED = 8, S_COUNT, S_SCAN, S_KEYS_A, S_KEYS_B, S_CUB, S_NUM, S_POFF, S_TOFF, S_PSYM, S_TSYM, S_SINKS,
       S_BEST, S_SECOND, S_ALIST, S_ALN, S_TBOPS, S_TBNOPS, S_TBTEMP, S_HREADS, S_HOFF, S_HHITS, S_HOPS, S_ROWS, S_STEPS };

struct MapGeom
{
    uint32_t r0, n_reads;            // the batch: reads [r0, r0 + n_reads)
    uint32_t S, n_strands;           // seeds per strand (for the longest read), strands
    uint32_t R;                      // range slots per seed: 1 (exact seeding) or Rf + Rr (approximate seeding)
    uint32_t Rf;                     // slots [0, Rf) of a seed hold forward-index ranges, [Rf, R) reverse-index ranges
    uint32_t seed_len, interval, max_hits, max_range, band;
    uint32_t len1;                   // approximate seeding: search steps that must match exactly
    uint32_t priority, deque_cap, max_rows;   // range-size hit priority (map_priority_kernel)
};
// slot t of a batch = ((read * n_strands + strand) * S + seed) * R + range
struct MapSlot { uint32_t r, strand, k, j; };
__device__ __forceinline__ MapSlot map_slot(const MapGeom& g, uint32_t t)
{
    MapSlot s;
    s.j = t % g.R; t /= g.R;
    s.k = t % g.S; t /= g.S;
    s.strand = t % g.n_strands; s.r = t / g.n_strands;
    return s;
}

__device__ __forceinline__ uint32_t map_strand_symbol(const uint8_t* read, uint32_t L, uint32_t strand, uint32_t i)
{
    if (!strand) return read[i];
    const uint32_t c = read[L - 1u - i];
    return c < 4u ? 3u - c : c;
}

// rows a range of `size` SA rows contributes: none when wider than max_range, else its first max_hits; under hit priority the
// full size is kept here and map_priority_kernel allots the rows
__device__ __forceinline__ uint32_t map_rows_of(const MapGeom& g, uint32_t size)
{
    if (g.max_range && size > g.max_range) return 0u;
    if (g.priority) return size < 0x7FFFFFFFu ? size : 0x7FFFFFFFu;
    return size < g.max_hits ? size : g.max_hits;
}

__global__ void __launch_bounds__(MAP_THREADS)
map_seed_kernel(const FMView f, const StrSet reads, const MapGeom g, uint2* __restrict__ seeds, uint32_t* __restrict__ counts)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    __shared__ uint32_t s_L2[8];
    if (threadIdx.x < 5) s_L2[threadIdx.x] = f.L2[threadIdx.x];
    fm_lut_init(s_lut);
    // a grid that just fills the device walks the slots with a stride: the LUT set-up of a CTA serves many seeds (a seed lives for
    // 3 - 5 dependent DRAM accesses only)
    const uint32_t n_slots = g.n_reads * g.n_strands * g.S;
    for (uint32_t t = blockIdx.x * MAP_THREADS + threadIdx.x; t < n_slots; t += gridDim.x * MAP_THREADS)
    {
    const MapSlot sl = map_slot(g, t);                       // R == 1 on this path
    const uint32_t r = sl.r, strand = sl.strand, k = sl.k;
    const uint32_t beg = reads.off[g.r0 + r], L = reads.off[g.r0 + r + 1] - beg;
    const uint8_t* read = reads.sym + beg;
    const uint32_t o = k * g.interval;
    uint32_t x = 1, y = 0, cnt = 0;
    if (o + g.seed_len <= L)
    {
        x = 0; y = f.length;
        int32_t i = (int32_t)g.seed_len - 1;
        if (f.ktab_k && g.seed_len >= f.ktab_k)
        {
            uint32_t idx = 0, bad = 0;
            for (uint32_t j = g.seed_len - f.ktab_k; j < g.seed_len; ++j)
            {
                const uint32_t c = map_strand_symbol(read, L, strand, o + j);
                bad |= c; idx = idx * 4u + (c & 3u);
            }
            if (bad <= 3u) { const uint2 r = __ldg(f.ktab + idx); x = r.x; y = r.y; i -= (int32_t)f.ktab_k; }
        }
        for (; i >= 0 && x <= y; --i)
        {
            const uint32_t c = map_strand_symbol(read, L, strand, o + (uint32_t)i);
            if (c > 3u) { x = 1; y = 0; break; }
            // x >= 1 after any step (and from the table): the lean form without the FM special cases applies
            const uint2 rk = (x != 0u && f.primary != 0u) ? fm_rank2_inner(f, x - 1u, y, c, s_lut) : fm_rank2(f, x - 1u, y, c, s_lut);
            x = s_L2[c] + rk.x + 1u;
            y = s_L2[c] + rk.y;
        }
        if (x <= y) cnt = map_rows_of(g, y - x + 1u);
    }
    seeds[t] = make_uint2(x, cnt);
    counts[t] = cnt;
    }
}

// approximate seeding (nvb_map_params.seed_mismatches = 1): nvBowtie's map<CHECK_EXACT> on the forward index
// (mapping_inl.h:114-184 as called at :395-397: exact over the first seed_len / 2 search steps, one substitution allowed in
// the rest, plus the exact match); the first R ranges in the reference's push order fill the seed's R slots
__global__ void __launch_bounds__(MAP_THREADS)
map_seed_approx_kernel(const FMView f, const FMView rf, const StrSet reads, const MapGeom g, uint2* __restrict__ seeds, uint32_t* __restrict__ counts)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    __shared__ uint32_t s_L2[16];
    if (threadIdx.x < 5) { s_L2[threadIdx.x] = f.L2[threadIdx.x]; s_L2[8 + threadIdx.x] = rf.L2[threadIdx.x]; }
    fm_lut_init(s_lut);
    const uint32_t t = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (t >= g.n_reads * g.n_strands * g.S) return;
    const MapSlot sl = map_slot(g, t * g.R);
    const uint32_t beg = reads.off[g.r0 + sl.r], L = reads.off[g.r0 + sl.r + 1] - beg;
    const uint8_t* read = reads.sym + beg;
    const uint32_t o = sl.k * g.interval;
    uint2* out = seeds + (size_t)t * g.R; uint32_t* cnt = counts + (size_t)t * g.R;
    const bool seeded = o + g.seed_len <= L;
    // forward index: the search consumes the seed from its end; exact over its last half, one substitution in its first half
    uint32_t n = 0;
    if (seeded)
        fm_approx_search(f, s_lut, s_L2, g.seed_len, g.len1, true,
                         [&](uint32_t st) -> uint32_t { return map_strand_symbol(read, L, sl.strand, o + g.seed_len - 1u - st); },
                         [&](uint32_t x, uint32_t y) { if (n < g.Rf) { const uint32_t c = map_rows_of(g, y - x + 1u); out[n] = make_uint2(x, c); cnt[n] = c; } ++n; });
    for (uint32_t j = n < g.Rf ? n : g.Rf; j < g.Rf; ++j) { out[j] = make_uint2(1u, 0u); cnt[j] = 0u; }
    // reverse index (the FM-index of the reversed text, nvBWT's .rbwt/.rsa): the search consumes the seed from its START, so
    // the first half is exact and the substitution falls into the last half -- map<IGNORE_EXACT>(r_reader, (seed_len + 1) / 2,
    // seed_len, rfmi, ...) of mapping_inl.h:399-402; the exact match was reported by the forward search already
    if (g.R > g.Rf)
    {
        uint32_t m = 0;
        const uint32_t Rr = g.R - g.Rf;
        if (seeded)
            fm_approx_search(rf, s_lut, s_L2 + 8, g.seed_len, (g.seed_len + 1u) / 2u, false,
                             [&](uint32_t st) -> uint32_t { return map_strand_symbol(read, L, sl.strand, o + st); },
                             [&](uint32_t x, uint32_t y) { if (m < Rr) { const uint32_t c = map_rows_of(g, y - x + 1u); out[g.Rf + m] = make_uint2(x, c); cnt[g.Rf + m] = c; } ++m; });
        for (uint32_t j = m < Rr ? m : Rr; j < Rr; ++j) { out[g.Rf + j] = make_uint2(1u, 0u); cnt[g.Rf + j] = 0u; }
    }
}

// Range-size hit priority (nvb_map_params.hit_priority = 1), per read: nvBowtie keeps a read's seed hits in a priority deque
// ordered by SA-range size (seed_hit.h:220-230 hit_compare, seed_hit_deque_array.h), pushes every new hit after dropping the
// LARGEST one when the deque already holds max_hits entries (mapping_inl.h:158-160,176-178), and extends rows from the smallest
// range first, front to back (select_inl.h:93-113). Here: the ranges of a read are pushed in slot order into a deque of
// deque_cap entries (0 = unbounded) under that rule, the survivors are visited by (size, slot) ascending and each takes
// min(size, max_hits, what is left of max_rows) rows from its front. Ties between equal sizes go by slot order (the reference's
// interval heap leaves them unspecified). One thread per read; a read has a few dozen slots, so the quadratic scans stay in L1.
__global__ void __launch_bounds__(MAP_THREADS)
map_priority_kernel(const MapGeom g, uint2* __restrict__ seeds, uint32_t* __restrict__ counts)
{
    const uint32_t r = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (r >= g.n_reads) return;
    const uint32_t per_read = g.n_strands * g.S * g.R;
    uint2* sd = seeds + (size_t)r * per_read; uint32_t* cn = counts + (size_t)r * per_read;
    // on entry sd[t].y = full size of the range (0 = empty or wider than max_range); cn[t] is scratch: 1 = in the deque
    uint32_t held = 0;
    for (uint32_t t = 0; t < per_read; ++t)
    {
        cn[t] = 0;
        if (sd[t].y == 0) continue;
        if (g.deque_cap && held == g.deque_cap)
        {
            uint32_t worst = 0, wsize = 0;
            for (uint32_t u = 0; u < t; ++u) if (cn[u] && sd[u].y >= wsize) { wsize = sd[u].y; worst = u; }   // the largest, latest slot on ties
            cn[worst] = 0; --held;
        }
        cn[t] = 1; ++held;
    }
    uint32_t budget = g.max_rows ? g.max_rows : 0xFFFFFFFFu;
    for (uint32_t done = 0; done < held; ++done)
    {
        uint32_t pick = 0, psize = 0xFFFFFFFFu;
        for (uint32_t u = 0; u < per_read; ++u) if (cn[u] == 1 && sd[u].y < psize) { psize = sd[u].y; pick = u; }   // the smallest, earliest slot on ties
        uint32_t c = psize < g.max_hits ? psize : g.max_hits;
        if (c > budget) c = budget;
        budget -= c;
        sd[pick].y = c | 0x80000000u;                        // allotted (the flag keeps it out of the next scans)
        cn[pick] = 2;
    }
    for (uint32_t t = 0; t < per_read; ++t)
    {
        const uint32_t c = cn[t] == 2 ? sd[t].y & 0x7FFFFFFFu : 0u;
        sd[t].y = c; cn[t] = c;
    }
}

// one thread per seed slot: its selected SA rows go to their places in hit order (row, slot)
__global__ void __launch_bounds__(MAP_THREADS)
map_expand_kernel(const uint2* __restrict__ seeds, const uint32_t* __restrict__ scan, uint32_t n_slots, uint2* __restrict__ rows)
{
    const uint32_t sl = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (sl >= n_slots) return;
    const uint2 sd = seeds[sl];
    const uint32_t base = scan[sl];
    for (uint32_t t = 0; t < sd.y; ++t) rows[base + t] = make_uint2(sd.x + t, sl);
}

// LF-walk of every hit to its sampled SA row (locate_ssa_iterator, fmindex_inl.h:364-400). The sampled rows are every
// 16th ROW of the suffix array, so a walk is a geometric trial: 15 steps on average, unbounded in principle (one in
// 10^7 hits walks more than 255 steps -- seen on the 3.1 Gbp genome), hence 32-bit step counts:
// each thread owns a strided list of hits and starts its next walk as soon as one ends, so the lanes of a warp do
// not wait for the longest walk among them. rows[h].x becomes the sampled row, steps[h] the distance walked.
// REV selects the rows of one index when a reverse index takes part (slots [Rf, R) of a seed): rows of the other index pass
// through as already-sampled dummies and are left untouched
template <bool DENSE>
__global__ void __launch_bounds__(MAP_THREADS, 8)
map_walk_kernel(const FMView f, uint32_t n_hits, uint2* __restrict__ rows, uint32_t* __restrict__ steps, uint32_t R, uint32_t Rf, uint32_t rev)
{
    __shared__ uint2 s_lut[FM_LUT_ENTRIES];
    fm_lut_init(s_lut);
    auto mine = [&](uint32_t slot) { return R == Rf || ((slot % R) >= Rf) == (rev != 0u); };
    fm_walk_balanced<DENSE>(f, n_hits, blockIdx.x * MAP_THREADS + threadIdx.x, gridDim.x * MAP_THREADS, s_lut,
                     [&](uint32_t h) { const uint2 r = rows[h]; return mine(r.y) ? r.x : 0u; },
                     [&](uint32_t h, uint32_t j, uint32_t t) { if (mine(rows[h].y)) { rows[h].x = j; steps[h] = t; } });
}

// sampled row + steps -> text position -> candidate key ((read * 2 + strand) << 32 | window begin)
__global__ void __launch_bounds__(MAP_THREADS)
map_keys_kernel(const FMView f, const FMView rf, const MapGeom g, const uint2* __restrict__ rows, const uint32_t* __restrict__ steps,
                uint32_t n_hits, uint64_t* __restrict__ keys)
{
    const uint32_t h = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (h >= n_hits) return;
    const uint2 rec = rows[h];
    const MapSlot sl = map_slot(g, rec.y);
    uint32_t pos;
    if (sl.j < g.Rf) pos = __ldg(f.sa_tab + (rec.x >> f.sa_shift)) + steps[h];
    else
    {
        // an occurrence of the reversed seed at q in the reversed text is an occurrence of the seed at length - q - seed_len
        const uint32_t q = __ldg(rf.sa_tab + (rec.x >> rf.sa_shift)) + steps[h];
        pos = f.length - q - g.seed_len;
    }
    const uint32_t r = sl.r, strand = sl.strand;
    const uint32_t o = sl.k * g.interval;
    const uint32_t g_pos = pos > o ? pos - o : 0u;
    const uint32_t gb = g_pos > g.band / 2u ? g_pos - g.band / 2u : 0u;
    keys[h] = ((uint64_t)(r * 2u + strand) << 32) | gb;
}

// The hits of a read are contiguous (slots are in read order): one warp sorts them and keeps the distinct keys,
// in place; the freed places are filled with MAP_DUP and dropped by the stream compaction that follows.
#define MAP_DUP 0xFFFFFFFFFFFFFFFFull
#define MAP_DEDUPE_WARPS 8
__device__ __forceinline__ uint64_t map_shfl_xor64(uint64_t v, int m)
{
    return ((uint64_t)__shfl_xor_sync(0xFFFFFFFFu, (uint32_t)(v >> 32), m) << 32) | __shfl_xor_sync(0xFFFFFFFFu, (uint32_t)v, m);
}
__global__ void __launch_bounds__(MAP_DEDUPE_WARPS * 32)
map_dedupe_kernel(const uint32_t* __restrict__ scan, uint32_t per_read, uint32_t n_reads, uint32_t cap, uint64_t* __restrict__ keys)
{
    extern __shared__ uint64_t s_keys[];                    // MAP_DEDUPE_WARPS * cap (only reads with more than 32 hits use it)
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t r = blockIdx.x * MAP_DEDUPE_WARPS + warp;
    if (r >= n_reads) return;
    const uint32_t lo = scan[r * per_read], cnt = scan[(r + 1u) * per_read] - lo;
    if (cnt == 0) return;
    if (cnt <= 32u)
    {
        uint64_t v = lane < cnt ? keys[lo + lane] : MAP_DUP;
        #pragma unroll
        for (int k = 2; k <= 32; k <<= 1)
            #pragma unroll
            for (int j = k >> 1; j > 0; j >>= 1)
            {
                const uint64_t o = map_shfl_xor64(v, j);
                const bool keep_min = ((lane & k) == 0) == ((lane & j) == 0);
                v = keep_min ? (v < o ? v : o) : (v > o ? v : o);
            }
        const uint32_t pl = __shfl_up_sync(0xFFFFFFFFu, (uint32_t)v, 1), ph = __shfl_up_sync(0xFFFFFFFFu, (uint32_t)(v >> 32), 1);
        const bool keep = v != MAP_DUP && (lane == 0 || v != (((uint64_t)ph << 32) | pl));
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
        const uint32_t kept = __popc(m);
        if (keep) keys[lo + __popc(m & ((1u << lane) - 1u))] = v;
        if (lane >= kept && lane < cnt) keys[lo + lane] = MAP_DUP;
        return;
    }
    uint64_t* sk = s_keys + (size_t)warp * cap;
    uint32_t N = 64; while (N < cnt) N <<= 1;                // cnt <= cap by construction
    for (uint32_t i = lane; i < N; i += 32u) sk[i] = i < cnt ? keys[lo + i] : MAP_DUP;
    __syncwarp();
    for (uint32_t k = 2; k <= N; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1)
        {
            for (uint32_t i = lane; i < N; i += 32u)
            {
                const uint32_t p = i ^ j;
                if (p > i)
                {
                    const uint64_t a = sk[i], b = sk[p];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { sk[i] = b; sk[p] = a; }
                }
            }
            __syncwarp();
        }
    uint32_t kept = 0;
    for (uint32_t base = 0; base < N; base += 32u)
    {
        const uint32_t i = base + lane;
        const uint64_t v = sk[i];
        const bool keep = v != MAP_DUP && (i == 0 || v != sk[i - 1]);
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
        if (keep) keys[lo + kept + __popc(m & ((1u << lane) - 1u))] = v;
        kept += __popc(m);
    }
    for (uint32_t i = kept + lane; i < cnt; i += 32u) keys[lo + i] = MAP_DUP;
}
struct MapNotDup { __host__ __device__ bool operator()(uint64_t v) const { return v != MAP_DUP; } };

// candidate j of a launch: a plain index into the sorted keys, or the best candidate of an aligned read
__device__ __forceinline__ uint64_t map_key(const uint64_t* keys, const uint32_t* alist, const uint32_t* best, uint32_t j)
{
    return keys[alist ? best[alist[j]] : j];
}

__global__ void __launch_bounds__(MAP_THREADS)
map_lengths_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ alist, const uint32_t* __restrict__ best,
                   uint32_t c0, uint32_t cnt, const StrSet reads, uint32_t r0, uint32_t band, uint32_t G,
                   uint32_t* __restrict__ plen, uint32_t* __restrict__ tlen)
{
    const uint32_t j = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (j > cnt) return;
    if (j == cnt) { plen[j] = 0; tlen[j] = 0; return; }
    const uint64_t key = map_key(keys, alist, best, c0 + j);
    const uint32_t r = (uint32_t)(key >> 33), gb = (uint32_t)key;
    const uint32_t L = reads.off[r0 + r + 1] - reads.off[r0 + r];
    const uint64_t ge = (uint64_t)gb + band + L;
    plen[j] = L;
    tlen[j] = (uint32_t)(ge < G ? ge : G) - gb;             // score_inl.h:105-106
}

// one warp per candidate
__global__ void __launch_bounds__(MAP_THREADS)
map_extract_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ alist, const uint32_t* __restrict__ best,
                   uint32_t c0, uint32_t cnt, const StrSet reads, uint32_t r0, const uint32_t* __restrict__ genome,
                   const uint32_t* __restrict__ poff, const uint32_t* __restrict__ toff,
                   uint8_t* __restrict__ psym, uint8_t* __restrict__ tsym)
{
    const uint32_t w = (blockIdx.x * MAP_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= cnt) return;
    const uint64_t key = map_key(keys, alist, best, c0 + w);
    const uint32_t r = (uint32_t)(key >> 33), strand = (uint32_t)(key >> 32) & 1u, gb = (uint32_t)key;
    const uint32_t p0 = poff[w], L = poff[w + 1] - p0, t0 = toff[w], N = toff[w + 1] - t0;
    const uint8_t* read = reads.sym + reads.off[r0 + r];
    for (uint32_t i = lane; i < L; i += 32u) psym[p0 + i] = (uint8_t)map_strand_symbol(read, L, strand, i);
    for (uint32_t i = lane; i < N; i += 32u)
    {
        const uint32_t p = gb + i;
        tsym[t0 + i] = (uint8_t)((__ldg(genome + (p >> 4)) >> (30u - 2u * (p & 15u))) & 3u);
    }
}

__global__ void __launch_bounds__(MAP_THREADS)
map_reduce_kernel(const uint64_t* __restrict__ keys, uint32_t n_cand, const nvb_best_sink* __restrict__ sinks, uint32_t n_reads,
                  int32_t min_score, uint32_t* __restrict__ best, int32_t* __restrict__ second, nvb_read_hit* __restrict__ hits)
{
    const uint32_t r = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (r >= n_reads) return;
    uint32_t range[2];
    #pragma unroll
    for (int e = 0; e < 2; ++e)
    {
        const uint64_t v = (uint64_t)(2u * r + 2u * e) << 32;   // lower_bound
        uint32_t lo = 0, hi = n_cand;
        while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if (keys[mid] < v) lo = mid + 1u; else hi = mid; }
        range[e] = lo;
    }
    int32_t b = INT32_MIN, s2 = INT32_MIN; uint32_t bi = MAP_NONE;
    for (uint32_t i = range[0]; i < range[1]; ++i)
    {
        const int32_t sc = sinks[i].score;
        if (sc <= MAP_INVALID) continue;
        if (sc > b) { if (b != INT32_MIN && b > s2) s2 = b; b = sc; bi = i; }
        else if (sc > s2) s2 = sc;
    }
    if (bi != MAP_NONE && b < min_score) bi = MAP_NONE;
    best[r] = bi; second[r] = s2;
    nvb_read_hit h; h.score = h.second_score = INT32_MIN; h.begin = h.end = 0xFFFFFFFFu; h.flags = 0; h.n_ops = 0;
    hits[r] = h;
}

struct MapAligned
{
    const uint32_t* best;
    __host__ __device__ bool operator()(uint32_t r) const { return best[r] != MAP_NONE; }
};

// one warp per aligned read
__global__ void __launch_bounds__(MAP_THREADS)
map_finish_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ alist, const uint32_t* __restrict__ best,
                  const int32_t* __restrict__ second, uint32_t c0, uint32_t cnt, const nvb_alignment* __restrict__ aln,
                  const uint8_t* __restrict__ tb_ops, uint32_t tb_stride, const uint32_t* __restrict__ tb_n_ops,
                  nvb_read_hit* __restrict__ hits, uint8_t* __restrict__ ops, uint32_t ops_stride)
{
    const uint32_t w = (blockIdx.x * MAP_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= cnt) return;
    const uint32_t r = alist[c0 + w];
    const uint64_t key = keys[best[r]];
    const uint32_t gb = (uint32_t)key;
    const uint32_t n = tb_n_ops[w];
    if (lane == 0)
    {
        const nvb_alignment a = aln[w];
        nvb_read_hit h;
        h.score = a.score; h.second_score = second[r];
        h.begin = gb + a.source_x; h.end = gb + a.sink_x;
        h.flags = 2u | ((uint32_t)(key >> 32) & 1u); h.n_ops = n;
        hits[r] = h;
    }
    if (ops)
    {
        const uint32_t bytes = n < ops_stride ? n : ops_stride;
        const uint4* src = (const uint4*)(tb_ops + (size_t)w * tb_stride);
        uint4* dst = (uint4*)(ops + (size_t)r * ops_stride);
        for (uint32_t i = lane; i * 16u < bytes; i += 32u) dst[i] = src[i];
    }
}

__global__ void __launch_bounds__(MAP_THREADS)
pack_2bit_kernel(const uint8_t* __restrict__ sym, uint32_t n, uint32_t n_words, uint32_t* __restrict__ words)
{
    const uint32_t w = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (w >= n_words) return;
    uint32_t v = 0;
    const uint32_t base = w * 16u;
    if (base + 16u <= n && (((uintptr_t)sym) & 15u) == 0)
    {
        const uint4 q = *(const uint4*)(sym + base);
        const uint32_t qs[4] = { q.x, q.y, q.z, q.w };
        #pragma unroll
        for (int k = 0; k < 4; ++k)
            #pragma unroll
            for (int b = 0; b < 4; ++b) v |= ((qs[k] >> (8 * b)) & 3u) << (30 - 2 * (4 * k + b));
    }
    else
        for (uint32_t i = 0; i < 16u && base + i < n; ++i) v |= (uint32_t)(sym[base + i] & 3u) << (30u - 2u * i);
    words[w] = v;
}

extern "C" int nvb_pack_2bit(nvb_ctx* ctx, const uint8_t* symbols_d, uint32_t n, uint32_t* words_d, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!symbols_d || !words_d) return NVB_ERR_ARG;
    const uint32_t n_words = (uint32_t)(((uint64_t)n + 15u) / 16u);
    pack_2bit_kernel<<<(n_words + MAP_THREADS - 1) / MAP_THREADS, MAP_THREADS, 0, (cudaStream_t)stream>>>(symbols_d, n, n_words, words_d);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

// NVB_MAP_TRACE=1: per-phase device time of a call (CUDA events on the call's stream), printed to stderr
struct MapTrace
{
    enum { SEED, LOCATE, SORT, EXTRACT, SCORE, REDUCE, TB_EXTRACT, TRACEBACK, FINISH, N_PHASES };
    bool on; cudaStream_t s; cudaEvent_t ev[2]; float ms[N_PHASES]; int cur; std::chrono::steady_clock::time_point t0;
    MapTrace(cudaStream_t st) : on(false), s(st), cur(0)
    {
        t0 = std::chrono::steady_clock::now();
        const char* e = getenv("NVB_MAP_TRACE");
        on = e && e[0] == '1';
        for (int i = 0; i < N_PHASES; ++i) ms[i] = 0.f;
        if (on) { cudaEventCreate(&ev[0]); cudaEventCreate(&ev[1]); cudaEventRecord(ev[0], s); }
    }
    void mark(int phase)
    {
        if (!on) return;
        cudaEventRecord(ev[cur ^ 1], s); cudaEventSynchronize(ev[cur ^ 1]);
        float t = 0.f; cudaEventElapsedTime(&t, ev[cur], ev[cur ^ 1]); ms[phase] += t; cur ^= 1;
    }
    ~MapTrace()
    {
        if (!on) return;
        static const char* names[N_PHASES] = { "seed+scan", "locate", "sort+unique", "extract", "score", "reduce+select", "tb extract", "traceback", "finish" };
        float tot = 0.f; for (int i = 0; i < N_PHASES; ++i) tot += ms[i];
        cudaStreamSynchronize(s);
        const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        fprintf(stderr, "[nvb_map_reads] wall %.3f ms, phases %.3f ms:", wall, tot);
        for (int i = 0; i < N_PHASES; ++i) fprintf(stderr, " %s %.3f |", names[i], ms[i]);
        fprintf(stderr, "\n");
        cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
    }
};

static inline uint32_t map_grid(uint64_t threads) { return (uint32_t)((threads + MAP_THREADS - 1) / MAP_THREADS); }
static inline uint32_t bits_for(uint64_t v) { uint32_t b = 0; while (b < 64 && (v >> b)) ++b; return b; }

#define MAP_STAGE(slot, bytes, ptr) do { void* p_ = nullptr; if ((rc = nvb_stage(ctx, slot, (bytes), &p_))) return rc; ptr = (decltype(ptr))p_; } while (0)

// string sets of `cnt` candidates starting at c0 (or of the best candidates of aligned reads alist[c0..]) into the slots
static int map_build_sets(nvb_ctx* ctx, const uint64_t* keys, const uint32_t* alist, const uint32_t* best, uint32_t c0, uint32_t cnt,
                          const StrSet& reads, uint32_t r0, const uint32_t* genome, uint32_t band, uint32_t G,
                          uint32_t maxL, nvb_string_set* P, nvb_string_set* T, cudaStream_t s)
{
    int rc;
    uint32_t *poff, *toff; uint8_t *psym, *tsym; void* tmp;
    MAP_STAGE(S_POFF, sizeof(uint32_t) * ((size_t)cnt + 1), poff);
    MAP_STAGE(S_TOFF, sizeof(uint32_t) * ((size_t)cnt + 1), toff);
    MAP_STAGE(S_PSYM, (size_t)cnt * maxL + 128, psym);
    MAP_STAGE(S_TSYM, (size_t)cnt * ((size_t)maxL + band) + 128, tsym);
    psym += 64; tsym += 64;        // the banded kernels read whole 16-byte blocks around a string
    map_lengths_kernel<<<map_grid((uint64_t)cnt + 1), MAP_THREADS, 0, s>>>(keys, alist, best, c0, cnt, reads, r0, band, G, poff, toff);
    NVB_LAUNCHED(ctx);
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, poff, poff, (int)(cnt + 1), s);
    MAP_STAGE(S_CUB, tmp_bytes, tmp);
    NVB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, poff, poff, (int)(cnt + 1), s));
    NVB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, toff, toff, (int)(cnt + 1), s));
    ctx->launches += 2;
    map_extract_kernel<<<map_grid((uint64_t)cnt * 32u), MAP_THREADS, 0, s>>>(keys, alist, best, c0, cnt, reads, r0, genome, poff, toff, psym, tsym);
    NVB_LAUNCHED(ctx);
    P->symbols = psym; P->offsets = poff; P->count = cnt; P->bits = 8;
    T->symbols = tsym; T->offsets = toff; T->count = cnt; T->bits = 8;
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// mate pairing over the hits of 2 n reads (read 2 i = mate 1, read 2 i + 1 = mate 2 of pair i): classifies each pair
// as concordant / discordant / single / unaligned under the forward-reverse policy with a fragment-length window
// (nvBowtie's ParamsPOD::pe_policy = PE_POLICY_FR, min_frag_len / max_frag_len: params.h:147-152; its anchor + opposite-
// mate window search, score_inl.h:385-411, is replaced by the independent mapping of both mates).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
map_pair_kernel(const nvb_read_hit* __restrict__ hits, uint32_t n_pairs, uint32_t min_frag, uint32_t max_frag,
                nvb_pair_hit* __restrict__ out)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n_pairs) return;
    // 24-byte records: 8-byte aligned loads
    const uint2* h2 = (const uint2*)(hits + 2 * (size_t)i);
    const uint2 pa = __ldg(h2), qa = __ldg(h2 + 1), a1 = __ldg(h2 + 2);      // (score, second), (begin, end), (flags, n_ops) of mate 1
    const uint2 pb = __ldg(h2 + 3), qb = __ldg(h2 + 4), b1 = __ldg(h2 + 5);
    const uint4 a0 = make_uint4(pa.x, pa.y, qa.x, qa.y), b0 = make_uint4(pb.x, pb.y, qb.x, qb.y);
    const bool al1 = (a1.x & 2u) != 0, al2 = (b1.x & 2u) != 0;
    nvb_pair_hit o; o.score = INT32_MIN; o.frag_begin = o.frag_end = 0xFFFFFFFFu; o.flags = (al1 ? 1u : 0u) | (al2 ? 2u : 0u);
    if (al1 && al2)
    {
        o.score = (int32_t)a0.x + (int32_t)b0.x;
        const bool r1 = (a1.x & 1u) != 0, r2 = (b1.x & 1u) != 0;
        // the forward mate starts the fragment, the reverse-complemented mate ends it
        const uint32_t fb = r1 ? b0.z : a0.z, fe = r1 ? b0.w : a0.w, rb = r1 ? a0.z : b0.z, re = r1 ? a0.w : b0.w;
        o.frag_begin = min(a0.z, b0.z); o.frag_end = max(a0.w, b0.w);
        const uint32_t frag = re >= fb ? re - fb : 0u;
        if (r1 != r2 && fb <= rb && fe <= re && frag >= min_frag && frag <= max_frag) o.flags |= 4u;   // concordant
    }
    else if (al1) { o.score = (int32_t)a0.x; o.frag_begin = a0.z; o.frag_end = a0.w; }
    else if (al2) { o.score = (int32_t)b0.x; o.frag_begin = b0.z; o.frag_end = b0.w; }
    out[i] = o;
}

extern "C" int nvb_pair_hits(nvb_ctx* ctx, const nvb_read_hit* hits_d, uint32_t n_pairs, uint32_t min_frag_len, uint32_t max_frag_len,
                             nvb_pair_hit* pairs_d, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n_pairs == 0) return NVB_OK;
    if (!hits_d || !pairs_d || min_frag_len > max_frag_len) return NVB_ERR_ARG;
    map_pair_kernel<<<(n_pairs + 255u) / 256u, 256, 0, (cudaStream_t)stream>>>(hits_d, n_pairs, min_frag_len, max_frag_len, pairs_d);
    NVB_LAUNCHED(ctx);
    return NVB_OK;
}

extern "C" int nvb_map_reads(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* genome, nvb_string_set reads_set,
                             uint32_t maxL, const nvb_map_params* mp, uint32_t reads_per_batch,
                             nvb_read_hit* hits, uint8_t* ops, uint32_t ops_stride, nvb_map_stats* stats, void* stream)
{
    if (!ctx || !fmi || !mp) return NVB_ERR_ARG;
    if (stats) { stats->seeds = stats->seed_hits = stats->candidates = stats->aligned = 0; }
    const uint32_t n_total = reads_set.count;
    if (n_total == 0) return NVB_OK;
    if (!genome || !reads_set.symbols || !reads_set.offsets || !hits || !fmi->bwt_occ || !fmi->ssa) return NVB_ERR_ARG;
    if (reads_set.bits != 8 || fmi->occ_int != 64 || fmi->sa_int != 16) return NVB_ERR_UNSUPPORTED;
    if (mp->aligner.type != NVB_SEMI_GLOBAL || mp->seed_len == 0 || mp->seed_interval == 0 || mp->max_hits == 0) return NVB_ERR_UNSUPPORTED;
    if (mp->band < 2 || mp->band > NVB_MAX_BAND_LEN || maxL == 0) return NVB_ERR_UNSUPPORTED;
    if (mp->seed_mismatches > 1u || mp->max_seed_ranges > 64u || mp->hit_priority > 1u) return NVB_ERR_UNSUPPORTED;
    const uint32_t tb_stride = ops ? ops_stride : ((2u * maxL + mp->band + 16u + 15u) & ~15u);
    if ((tb_stride & 15u) || tb_stride < 2u * maxL + mp->band || (ops && (((uintptr_t)ops) & 15u))) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    FMView f = make_view(fmi);
    const StrSet reads = make_set(reads_set);
    const uint32_t G = fmi->length;
    int rc;

    MapGeom g;
    g.n_strands = mp->both_strands ? 2u : 1u;
    g.S = maxL >= mp->seed_len ? (maxL - mp->seed_len) / mp->seed_interval + 1u : 0u;
    g.seed_len = mp->seed_len; g.interval = mp->seed_interval; g.max_hits = mp->max_hits; g.max_range = mp->max_range; g.band = mp->band;
    // approximate seeds: Rf forward-index range slots per seed, and as many reverse-index slots when a reverse index is given
    const nvb_fmindex* rfmi = mp->seed_mismatches ? mp->reverse_index : nullptr;
    if (rfmi && (!rfmi->bwt_occ || !rfmi->ssa || rfmi->occ_int != 64 || rfmi->sa_int != 16 || rfmi->length != fmi->length || (((uintptr_t)rfmi->bwt_occ) & 31u))) return NVB_ERR_ARG;
    const FMView rf = rfmi ? make_view(rfmi) : f;
    g.Rf = mp->seed_mismatches ? (mp->max_seed_ranges ? mp->max_seed_ranges : 8u) : 1u;
    g.R = rfmi ? 2u * g.Rf : g.Rf;
    g.len1 = mp->seed_len / 2u;                                       // mapping_inl.h:397
    g.priority = mp->hit_priority; g.deque_cap = mp->max_seed_hits; g.max_rows = mp->max_rows;
    const uint32_t per_read = g.S * g.n_strands * g.R;               // range slots per read
    // worst-case rows of one read
    uint64_t rows_per_read = (uint64_t)per_read * mp->max_hits;
    if (g.priority && g.max_rows && g.max_rows < rows_per_read) rows_per_read = g.max_rows;
    // batch: bounded so that slots and worst-case hits stay below 2^31
    uint64_t rb = reads_per_batch ? reads_per_batch : (1u << 20);
    if (per_read) { const uint64_t cap = 0x7FFFFFFFull / ((uint64_t)per_read * mp->max_hits); if (rb > cap) rb = cap; }
    if (rb == 0) return NVB_ERR_UNSUPPORTED;
    const uint32_t cand_chunk = 4u << 20, tb_chunk = 2u << 20;

    uint32_t* d_num; MAP_STAGE(S_NUM, 64, d_num);
    // the seeds of the whole call share their first steps through one per-call k-mer table (as nvb_fm_match does per batch)
    {
        const uint64_t seeds = (uint64_t)n_total * ((uint64_t)(maxL >= mp->seed_len ? (maxL - mp->seed_len) / mp->seed_interval + 1u : 0u) * (mp->both_strands ? 2u : 1u));
        if (!mp->seed_mismatches && seeds >= ctx->fm_share_min && (rc = nvb_fm_batch_table(ctx, f, seeds, mp->seed_len, s))) return rc;
    }
    MapTrace trace(s);
    for (uint32_t r0 = 0; r0 < n_total; r0 += (uint32_t)rb)
    {
        const uint32_t nr = (uint32_t)(n_total - r0 < rb ? n_total - r0 : rb);
        g.r0 = r0; g.n_reads = nr;
        nvb_read_hit* bhits = hits + r0;
        uint32_t* best; int32_t* second;
        MAP_STAGE(S_BEST, sizeof(uint32_t) * (size_t)nr, best);
        MAP_STAGE(S_SECOND, sizeof(int32_t) * (size_t)nr, second);
        const uint32_t n_slots = nr * per_read;
        uint32_t n_hits = 0, n_cand = 0, n_aligned = 0;
        uint64_t *keys_a = nullptr, *keys_b = nullptr;
        void* tmp; size_t tmp_bytes;
        if (n_slots)
        {
            // ---- seeding
            uint2* seeds; uint32_t *counts, *scan;
            MAP_STAGE(S_SEED, sizeof(uint2) * (size_t)n_slots, seeds);
            MAP_STAGE(S_COUNT, sizeof(uint32_t) * ((size_t)n_slots + 1), counts);
            MAP_STAGE(S_SCAN, sizeof(uint32_t) * ((size_t)n_slots + 1), scan);
            NVB_CUDA(ctx, cudaMemsetAsync(counts + n_slots, 0, sizeof(uint32_t), s));
            if (g.R == 1u)
            {
                static int seed_waves = -1;
                if (seed_waves < 0) { const char* e = getenv("NVB_MAP_SEED_WAVES"); seed_waves = e ? atoi(e) : 16; }
                const uint32_t want = map_grid(n_slots), fill = (uint32_t)ctx->sm_count * 6u * (uint32_t)(seed_waves > 0 ? seed_waves : 1);
                map_seed_kernel<<<(seed_waves > 0 && fill < want) ? fill : want, MAP_THREADS, 0, s>>>(f, reads, g, seeds, counts);
            }
            else           map_seed_approx_kernel<<<map_grid(n_slots / g.R), MAP_THREADS, 0, s>>>(f, rf, reads, g, seeds, counts);
            NVB_LAUNCHED(ctx);
            if (g.priority)
            {
                map_priority_kernel<<<map_grid(nr), MAP_THREADS, 0, s>>>(g, seeds, counts);
                NVB_LAUNCHED(ctx);
            }
            tmp_bytes = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, scan, (int)(n_slots + 1), s);
            MAP_STAGE(S_CUB, tmp_bytes, tmp);
            NVB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, scan, (int)(n_slots + 1), s));
            ctx->launches++;
            NVB_CUDA(ctx, cudaMemcpyAsync(&n_hits, scan + n_slots, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
            NVB_CUDA(ctx, cudaStreamSynchronize(s));
            trace.mark(MapTrace::SEED);
            if (n_hits)
            {
                // ---- locate + candidate keys, de-duplicated per read
                uint2* rows; uint32_t* steps;
                MAP_STAGE(S_KEYS_A, sizeof(uint64_t) * (size_t)n_hits, keys_a);
                MAP_STAGE(S_KEYS_B, sizeof(uint64_t) * (size_t)n_hits, keys_b);
                MAP_STAGE(S_ROWS, sizeof(uint2) * (size_t)n_hits, rows);
                MAP_STAGE(S_STEPS, sizeof(uint32_t) * (size_t)n_hits, steps);
                map_expand_kernel<<<map_grid(n_slots), MAP_THREADS, 0, s>>>(seeds, scan, n_slots, rows);
                NVB_LAUNCHED(ctx);
                {
                    const uint32_t want = map_grid(n_hits), persistent = (uint32_t)ctx->sm_count * 8u;
                    const uint32_t wg = want < persistent ? want : persistent;
                    if (f.sa_mask != 15u) map_walk_kernel<true><<<wg, MAP_THREADS, 0, s>>>(f, n_hits, rows, steps, g.R, g.Rf, 0u);
                    else                  map_walk_kernel<false><<<wg, MAP_THREADS, 0, s>>>(f, n_hits, rows, steps, g.R, g.Rf, 0u);
                    NVB_LAUNCHED(ctx);
                    if (g.R > g.Rf)
                    {
                        if (rf.sa_mask != 15u) map_walk_kernel<true><<<wg, MAP_THREADS, 0, s>>>(rf, n_hits, rows, steps, g.R, g.Rf, 1u);
                        else                   map_walk_kernel<false><<<wg, MAP_THREADS, 0, s>>>(rf, n_hits, rows, steps, g.R, g.Rf, 1u);
                        NVB_LAUNCHED(ctx);
                    }
                }
                map_keys_kernel<<<map_grid(n_hits), MAP_THREADS, 0, s>>>(f, rf, g, rows, steps, n_hits, keys_b);
                NVB_LAUNCHED(ctx);
                trace.mark(MapTrace::LOCATE);
                uint32_t cap = 64; while (cap < rows_per_read && cap < (1u << 30)) cap <<= 1;
                size_t sel_bytes = 0;
                if (cap <= 1024u)
                {
                    if (NVB_FIRST_ON_DEVICE(ctx))              // up to 64 KB of dynamic shared memory (8 warps x 1024 keys)
                        NVB_CUDA(ctx, cudaFuncSetAttribute(map_dedupe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MAP_DEDUPE_WARPS * 1024 * (int)sizeof(uint64_t)));
                    map_dedupe_kernel<<<(nr + MAP_DEDUPE_WARPS - 1) / MAP_DEDUPE_WARPS, MAP_DEDUPE_WARPS * 32, MAP_DEDUPE_WARPS * cap * sizeof(uint64_t), s>>>(
                        scan, per_read, nr, cap, keys_b);
                    NVB_LAUNCHED(ctx);
                    cub::DeviceSelect::If(nullptr, sel_bytes, keys_b, keys_a, d_num, (int)n_hits, MapNotDup(), s);
                    MAP_STAGE(S_CUB, sel_bytes, tmp);
                    NVB_CUDA(ctx, cub::DeviceSelect::If(tmp, sel_bytes, keys_b, keys_a, d_num, (int)n_hits, MapNotDup(), s));
                    ctx->launches++;
                }
                else
                {
                    // very many hits per read: global radix sort + unique
                    const int end_bit = (int)(32u + bits_for(2ull * nr));
                    size_t sort_bytes = 0, uniq_bytes = 0;
                    uint64_t* keys_c = (uint64_t*)rows;          // rows are consumed
                    cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, keys_b, keys_c, (int)n_hits, 0, end_bit, s);
                    cub::DeviceSelect::Unique(nullptr, uniq_bytes, keys_c, keys_a, d_num, (int)n_hits, s);
                    tmp_bytes = sort_bytes > uniq_bytes ? sort_bytes : uniq_bytes;
                    MAP_STAGE(S_CUB, tmp_bytes, tmp);
                    NVB_CUDA(ctx, cub::DeviceRadixSort::SortKeys(tmp, sort_bytes, keys_b, keys_c, (int)n_hits, 0, end_bit, s));
                    NVB_CUDA(ctx, cub::DeviceSelect::Unique(tmp, uniq_bytes, keys_c, keys_a, d_num, (int)n_hits, s));
                    ctx->launches += 2;
                }
                NVB_CUDA(ctx, cudaMemcpyAsync(&n_cand, d_num, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
                NVB_CUDA(ctx, cudaStreamSynchronize(s));
                trace.mark(MapTrace::SORT);
            }
        }
        const uint64_t* keys = keys_a;
        // ---- extension: banded scoring of every candidate
        nvb_best_sink* sinks = nullptr;
        if (n_cand)
        {
            MAP_STAGE(S_SINKS, sizeof(nvb_best_sink) * (size_t)n_cand, sinks);
            for (uint32_t c0 = 0; c0 < n_cand; c0 += cand_chunk)
            {
                const uint32_t cnt = n_cand - c0 < cand_chunk ? n_cand - c0 : cand_chunk;
                nvb_string_set P, T;
                if ((rc = map_build_sets(ctx, keys, nullptr, nullptr, c0, cnt, reads, r0, genome, mp->band, G, maxL, &P, &T, s))) return rc;
                trace.mark(MapTrace::EXTRACT);
                if ((rc = nvb_batch_banded_score(ctx, mp->aligner, mp->band, P, T, sinks + c0, maxL, maxL + mp->band, nullptr, 0, s))) return rc;
                trace.mark(MapTrace::SCORE);
            }
        }
        // ---- best / second best per read, list of aligned reads
        map_reduce_kernel<<<map_grid(nr), MAP_THREADS, 0, s>>>(keys, n_cand, sinks, nr, mp->min_score, best, second, bhits);
        NVB_LAUNCHED(ctx);
        if (n_cand)
        {
            uint32_t* alist;
            MAP_STAGE(S_ALIST, sizeof(uint32_t) * (size_t)nr, alist);
            MapAligned pred; pred.best = best;
            thrust::counting_iterator<uint32_t> iota(0u);
            tmp_bytes = 0;
            cub::DeviceSelect::If(nullptr, tmp_bytes, iota, alist, d_num, (int)nr, pred, s);
            MAP_STAGE(S_CUB, tmp_bytes, tmp);
            NVB_CUDA(ctx, cub::DeviceSelect::If(tmp, tmp_bytes, iota, alist, d_num, (int)nr, pred, s));
            ctx->launches++;
            NVB_CUDA(ctx, cudaMemcpyAsync(&n_aligned, d_num, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
            NVB_CUDA(ctx, cudaStreamSynchronize(s));
            trace.mark(MapTrace::REDUCE);
            // ---- traceback of the best window of every aligned read
            for (uint32_t c0 = 0; c0 < n_aligned; c0 += tb_chunk)
            {
                const uint32_t cnt = n_aligned - c0 < tb_chunk ? n_aligned - c0 : tb_chunk;
                nvb_string_set P, T;
                if ((rc = map_build_sets(ctx, keys, alist + c0, best, 0, cnt, reads, r0, genome, mp->band, G, maxL, &P, &T, s))) return rc;
                trace.mark(MapTrace::TB_EXTRACT);
                nvb_alignment* aln; uint8_t* tb_ops; uint32_t* tb_n; void* tb_temp;
                MAP_STAGE(S_ALN, sizeof(nvb_alignment) * (size_t)cnt, aln);
                MAP_STAGE(S_TBOPS, (size_t)cnt * tb_stride + 16, tb_ops);
                MAP_STAGE(S_TBNOPS, sizeof(uint32_t) * (size_t)cnt, tb_n);
                size_t want = nvb_temp_bytes(NVB_OP_BANDED_TRACEBACK, mp->aligner, mp->band, 16, maxL, maxL + mp->band, cnt);
                const size_t fast = (size_t)((cnt + 1u) / 2u) * 16u * maxL + 4096;   // flow words of the packed forward pass
                if (want < fast) want = fast;
                if (want > ((size_t)3 << 30)) want = (size_t)3 << 30;
                MAP_STAGE(S_TBTEMP, want, tb_temp);
                if ((rc = nvb_batch_banded_traceback(ctx, mp->aligner, mp->band, 16, P, T, aln, tb_ops, tb_stride, tb_n, maxL, maxL + mp->band,
                                                     tb_temp, want, s))) return rc;
                trace.mark(MapTrace::TRACEBACK);
                map_finish_kernel<<<map_grid((uint64_t)cnt * 32u), MAP_THREADS, 0, s>>>(keys, alist + c0, best, second, 0, cnt, aln, tb_ops, tb_stride, tb_n,
                                                                                       bhits, ops ? ops + (size_t)r0 * ops_stride : nullptr, ops_stride);
                NVB_LAUNCHED(ctx);
                trace.mark(MapTrace::FINISH);
            }
        }
        if (stats) { stats->seeds += n_slots / g.R; stats->seed_hits += n_hits; stats->candidates += n_cand; stats->aligned += n_aligned; }
    }
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Opposite-mate window search (mate rescue) for pairs with exactly one aligned mate. nvBowtie scores the opposite mate of an
// anchor in the genome window where it would complete a concordant pair (BestOppositeScoreStream::init_context,
// nvBowtie/bowtie2/cuda/score_inl.h:318-411; frame_opposite_mate under PE_POLICY_FR) with the full, unbanded aligner
// (max_text_length = max_frag_len, :312). Kept from it: the window -- anchor on the forward strand: the mate lies to the right
// on the reverse strand, [anchor.begin, anchor.begin + max_frag) (pe_overlap: the mates may overlap, :400-402); anchor on the
// reverse strand: the mate lies to the left on the forward strand, [anchor.end - max_frag, anchor.end) (:393-397) -- clipped to
// the genome, and full-matrix SEMI_GLOBAL scoring followed by the full-matrix traceback of the mates that reach min_score.
// Not kept: the pair-score targets (compute_target_score) and max_text_gaps trimming of the window (:381-392).
// ---------------------------------------------------------------------------------------------------------------
struct RescueTask { uint32_t read, strand, gb, ge; };     // read index (the mate to rescue), its strand, genome window [gb, ge)

__device__ __forceinline__ bool map_rescue_task(const nvb_read_hit* hits, uint32_t pair, uint32_t max_frag, uint32_t G, RescueTask& t)
{
    const nvb_read_hit a = hits[2 * (size_t)pair], b = hits[2 * (size_t)pair + 1];
    const bool al1 = (a.flags & 2u) != 0, al2 = (b.flags & 2u) != 0;
    if (al1 == al2) return false;
    const nvb_read_hit& an = al1 ? a : b;
    t.read = 2u * pair + (al1 ? 1u : 0u);
    if (!(an.flags & 1u)) { t.strand = 1u; t.gb = an.begin; const uint64_t e = (uint64_t)an.begin + max_frag; t.ge = (uint32_t)(e < G ? e : G); }
    else                  { t.strand = 0u; t.ge = an.end < G ? an.end : G; t.gb = t.ge > max_frag ? t.ge - max_frag : 0u; }
    return t.ge > t.gb;
}
__global__ void __launch_bounds__(MAP_THREADS)
map_rescue_flag_kernel(const nvb_read_hit* __restrict__ hits, uint32_t n_pairs, uint32_t max_frag, uint32_t G, const StrSet reads, uint8_t* __restrict__ flag)
{
    const uint32_t i = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (i >= n_pairs) return;
    RescueTask t;
    bool ok = map_rescue_task(hits, i, max_frag, G, t);
    if (ok) { const uint32_t L = reads.off[t.read + 1] - reads.off[t.read]; ok = L > 0 && L <= t.ge - t.gb; }
    flag[i] = ok ? 1 : 0;
}
// list[j] = pair index; lengths of task j (one extra zero entry for the scans)
__global__ void __launch_bounds__(MAP_THREADS)
map_rescue_lengths_kernel(const nvb_read_hit* __restrict__ hits, const uint32_t* __restrict__ list, uint32_t cnt, uint32_t max_frag, uint32_t G,
                          const StrSet reads, uint32_t* __restrict__ plen, uint32_t* __restrict__ tlen)
{
    const uint32_t j = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (j > cnt) return;
    if (j == cnt) { plen[j] = 0; tlen[j] = 0; return; }
    RescueTask t; map_rescue_task(hits, list[j], max_frag, G, t);
    plen[j] = reads.off[t.read + 1] - reads.off[t.read];
    tlen[j] = t.ge - t.gb;
}
// one warp per task: strand read and genome window -> 8-bit string sets
__global__ void __launch_bounds__(MAP_THREADS)
map_rescue_extract_kernel(const nvb_read_hit* __restrict__ hits, const uint32_t* __restrict__ list, uint32_t cnt, uint32_t max_frag, uint32_t G,
                          const StrSet reads, const uint32_t* __restrict__ genome, const uint32_t* __restrict__ poff, const uint32_t* __restrict__ toff,
                          uint8_t* __restrict__ psym, uint8_t* __restrict__ tsym)
{
    const uint32_t w = (blockIdx.x * MAP_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= cnt) return;
    RescueTask t; map_rescue_task(hits, list[w], max_frag, G, t);
    const uint32_t p0 = poff[w], L = poff[w + 1] - p0, t0 = toff[w], N = toff[w + 1] - t0;
    const uint8_t* read = reads.sym + reads.off[t.read];
    for (uint32_t i = lane; i < L; i += 32u) psym[p0 + i] = (uint8_t)map_strand_symbol(read, L, t.strand, i);
    for (uint32_t i = lane; i < N; i += 32u)
    {
        const uint32_t p = t.gb + i;
        tsym[t0 + i] = (uint8_t)((__ldg(genome + (p >> 4)) >> (30u - 2u * (p & 15u))) & 3u);
    }
}
__global__ void __launch_bounds__(MAP_THREADS)
map_rescue_pass_kernel(const nvb_best_sink* __restrict__ sinks, uint32_t cnt, int32_t min_score, uint8_t* __restrict__ flag)
{
    const uint32_t j = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (j < cnt) flag[j] = (sinks[j].score > MAP_INVALID && sinks[j].score >= min_score) ? 1 : 0;
}
// the rescued mate's hit: genome coordinates from the window, strand, aligned + rescued (bit 2) flags; its op list
__global__ void __launch_bounds__(MAP_THREADS)
map_rescue_finish_kernel(nvb_read_hit* __restrict__ hits, const uint32_t* __restrict__ list, uint32_t cnt, uint32_t max_frag, uint32_t G,
                         const nvb_alignment* __restrict__ aln, const uint32_t* __restrict__ n_ops, const uint8_t* __restrict__ tb_ops, uint32_t tb_stride,
                         uint8_t* __restrict__ ops, uint32_t ops_stride)
{
    const uint32_t w = (blockIdx.x * MAP_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= cnt) return;
    RescueTask t; map_rescue_task(hits, list[w], max_frag, G, t);
    __syncwarp();                                            // every lane has read the pair before lane 0 rewrites it
    const nvb_alignment a = aln[w];
    const uint32_t n = n_ops[w];
    if (lane == 0)
    {
        nvb_read_hit h; h.score = a.score; h.second_score = INT32_MIN;
        h.begin = t.gb + a.source_x; h.end = t.gb + a.sink_x; h.flags = 2u | 4u | t.strand; h.n_ops = n;
        hits[t.read] = h;
    }
    if (ops)
    {
        uint8_t* o = ops + (size_t)t.read * ops_stride; const uint8_t* src = tb_ops + (size_t)w * tb_stride;
        const uint32_t m = n < ops_stride ? n : ops_stride;
        for (uint32_t i = lane; i < ops_stride; i += 32u) o[i] = i < m ? src[i] : 0;
    }
}

static int map_rescue_sets(nvb_ctx* ctx, const nvb_read_hit* hits, const uint32_t* list, uint32_t cnt, uint32_t max_frag, uint32_t G,
                           const StrSet& reads, const uint32_t* genome, uint32_t maxL, nvb_string_set* P, nvb_string_set* T, cudaStream_t s)
{
    int rc;
    uint32_t *poff, *toff; uint8_t *psym, *tsym; void* tmp;
    MAP_STAGE(S_POFF, sizeof(uint32_t) * ((size_t)cnt + 1), poff);
    MAP_STAGE(S_TOFF, sizeof(uint32_t) * ((size_t)cnt + 1), toff);
    MAP_STAGE(S_PSYM, (size_t)cnt * maxL + 128, psym);
    MAP_STAGE(S_TSYM, (size_t)cnt * max_frag + 128, tsym);
    psym += 64; tsym += 64;
    map_rescue_lengths_kernel<<<map_grid((uint64_t)cnt + 1), MAP_THREADS, 0, s>>>(hits, list, cnt, max_frag, G, reads, poff, toff);
    NVB_LAUNCHED(ctx);
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, poff, poff, (int)(cnt + 1), s);
    MAP_STAGE(S_CUB, tmp_bytes, tmp);
    NVB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, poff, poff, (int)(cnt + 1), s));
    NVB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, toff, toff, (int)(cnt + 1), s));
    ctx->launches += 2;
    map_rescue_extract_kernel<<<map_grid((uint64_t)cnt * 32u), MAP_THREADS, 0, s>>>(hits, list, cnt, max_frag, G, reads, genome, poff, toff, psym, tsym);
    NVB_LAUNCHED(ctx);
    P->symbols = psym; P->offsets = poff; P->count = cnt; P->bits = 8;
    T->symbols = tsym; T->offsets = toff; T->count = cnt; T->bits = 8;
    return NVB_OK;
}

extern "C" int nvb_rescue_mates(nvb_ctx* ctx, const uint32_t* genome, uint32_t genome_len, nvb_string_set mates, uint32_t maxL,
                                nvb_aligner aligner, int32_t min_score, uint32_t max_frag_len, nvb_read_hit* hits,
                                uint8_t* ops, uint32_t ops_stride, uint32_t* n_candidates_h, uint32_t* n_rescued_h, void* stream)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n_candidates_h) *n_candidates_h = 0;
    if (n_rescued_h) *n_rescued_h = 0;
    const uint32_t n_pairs = mates.count / 2u;
    if (n_pairs == 0) return NVB_OK;
    if (!genome || !mates.symbols || !mates.offsets || !hits || maxL == 0 || max_frag_len == 0) return NVB_ERR_ARG;
    if (mates.bits != 8 || aligner.type != NVB_SEMI_GLOBAL || max_frag_len > 4096u || maxL > max_frag_len) return NVB_ERR_UNSUPPORTED;
    if (ops && (ops_stride == 0 || (ops_stride & 15u) || (((uintptr_t)ops) & 15u))) return NVB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const StrSet reads = make_set(mates);
    int rc; void* tmp; size_t tmp_bytes = 0;
    uint32_t* d_num; MAP_STAGE(S_NUM, 64, d_num);
    // ---- pairs with exactly one aligned mate and a window the other mate fits into
    uint8_t* flag; uint32_t* list;
    MAP_STAGE(S_COUNT, (size_t)n_pairs, flag);
    MAP_STAGE(S_SCAN, sizeof(uint32_t) * (size_t)n_pairs, list);
    map_rescue_flag_kernel<<<map_grid(n_pairs), MAP_THREADS, 0, s>>>(hits, n_pairs, max_frag_len, genome_len, reads, flag);
    NVB_LAUNCHED(ctx);
    cub::DeviceSelect::Flagged(nullptr, tmp_bytes, thrust::counting_iterator<uint32_t>(0), flag, list, d_num, (int)n_pairs, s);
    MAP_STAGE(S_CUB, tmp_bytes, tmp);
    NVB_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, tmp_bytes, thrust::counting_iterator<uint32_t>(0), flag, list, d_num, (int)n_pairs, s));
    ctx->launches++;
    uint32_t n_cand = 0;
    NVB_CUDA(ctx, cudaMemcpyAsync(&n_cand, d_num, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    NVB_CUDA(ctx, cudaStreamSynchronize(s));
    if (n_candidates_h) *n_candidates_h = n_cand;
    if (n_cand == 0) return NVB_OK;
    // ---- full-matrix scoring of every window, in chunks
    const uint32_t chunk = 1u << 20;
    nvb_best_sink* sinks; uint8_t* pass; uint32_t* surv;
    MAP_STAGE(S_SINKS, sizeof(nvb_best_sink) * (size_t)n_cand, sinks);
    MAP_STAGE(S_KEYS_A, (size_t)n_cand, pass);
    MAP_STAGE(S_KEYS_B, sizeof(uint32_t) * (size_t)n_cand, surv);
    for (uint32_t c0 = 0; c0 < n_cand; c0 += chunk)
    {
        const uint32_t cnt = n_cand - c0 < chunk ? n_cand - c0 : chunk;
        nvb_string_set P, T;
        if ((rc = map_rescue_sets(ctx, hits, list + c0, cnt, max_frag_len, genome_len, reads, genome, maxL, &P, &T, s))) return rc;
        const size_t tb = nvb_temp_bytes(NVB_OP_SCORE, aligner, 0, 0, maxL, max_frag_len, cnt);
        void* temp; MAP_STAGE(S_TBTEMP, tb, temp);
        if ((rc = nvb_batch_score(ctx, aligner, P, T, sinks + c0, maxL, max_frag_len, temp, tb, s))) return rc;
    }
    map_rescue_pass_kernel<<<map_grid(n_cand), MAP_THREADS, 0, s>>>(sinks, n_cand, min_score, pass);
    NVB_LAUNCHED(ctx);
    tmp_bytes = 0;
    cub::DeviceSelect::Flagged(nullptr, tmp_bytes, list, pass, surv, d_num, (int)n_cand, s);
    MAP_STAGE(S_CUB, tmp_bytes, tmp);
    NVB_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, tmp_bytes, list, pass, surv, d_num, (int)n_cand, s));
    ctx->launches++;
    uint32_t n_surv = 0;
    NVB_CUDA(ctx, cudaMemcpyAsync(&n_surv, d_num, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    NVB_CUDA(ctx, cudaStreamSynchronize(s));
    if (n_rescued_h) *n_rescued_h = n_surv;
    // ---- full-matrix traceback of the mates that passed
    const uint32_t tb_stride = (maxL + max_frag_len + 15u) & ~15u, tb_chunk = 1u << 18;
    for (uint32_t c0 = 0; c0 < n_surv; c0 += tb_chunk)
    {
        const uint32_t cnt = n_surv - c0 < tb_chunk ? n_surv - c0 : tb_chunk;
        nvb_string_set P, T;
        if ((rc = map_rescue_sets(ctx, hits, surv + c0, cnt, max_frag_len, genome_len, reads, genome, maxL, &P, &T, s))) return rc;
        nvb_alignment* aln; uint8_t* tb_ops; uint32_t* tb_nops; void* temp;
        const size_t tb = nvb_temp_bytes(NVB_OP_TRACEBACK, aligner, 0, 32, maxL, max_frag_len, cnt);
        MAP_STAGE(S_ALN, sizeof(nvb_alignment) * (size_t)cnt, aln);
        MAP_STAGE(S_TBOPS, (size_t)cnt * tb_stride, tb_ops);
        MAP_STAGE(S_TBNOPS, sizeof(uint32_t) * (size_t)cnt, tb_nops);
        MAP_STAGE(S_TBTEMP, tb, temp);
        if ((rc = nvb_batch_traceback(ctx, aligner, 32, P, T, aln, tb_ops, tb_stride, tb_nops, maxL, max_frag_len, temp, tb, s))) return rc;
        map_rescue_finish_kernel<<<map_grid((uint64_t)cnt * 32u), MAP_THREADS, 0, s>>>(hits, surv + c0, cnt, max_frag_len, genome_len, aln, tb_nops, tb_ops, tb_stride, ops, ops_stride);
        NVB_LAUNCHED(ctx);
    }
    return NVB_OK;
}

// op lists (push order) -> forward run-length CIGARs (io::Cigar: op | len << 2, nvbio/io/alignments.h:48-65);
// the number of entries goes to bits 8.. of the hit's flags
__global__ void __launch_bounds__(MAP_THREADS)
map_cigar_kernel(nvb_read_hit* __restrict__ hits, const uint8_t* __restrict__ ops, uint32_t ops_stride, uint32_t n,
                 uint16_t* __restrict__ cigar, uint32_t cigar_stride)
{
    const uint32_t r = blockIdx.x * MAP_THREADS + threadIdx.x;
    if (r >= n) return;
    const uint8_t* o = ops + (size_t)r * ops_stride;
    uint16_t* c = cigar + (size_t)r * cigar_stride;
    uint32_t len = hits[r].n_ops; if (len > ops_stride) len = ops_stride;
    uint32_t out = 0, run = 0, prev = 0xFFu;
    for (uint32_t k = len; k-- > 0;)
    {
        const uint32_t op = o[k] & 3u;
        if (op == prev && run < 16383u) { ++run; continue; }
        if (run) { if (out < cigar_stride) c[out] = (uint16_t)(prev | (run << 2)); ++out; }
        prev = op; run = 1;
    }
    if (run) { if (out < cigar_stride) c[out] = (uint16_t)(prev | (run << 2)); ++out; }
    hits[r].flags |= out << 8;
}

// Host buffers in and out, pipelined over batches of reads: the reads of batch k+1 cross PCIe while batch k is mapped,
// and the results of batch k-1 leave (copy-in / compute / copy-out streams chained by events, two buffer sets).
// cigar_h (optional): cigar_stride uint16 entries per read, io::Cigar encoding, entry count in flags >> 8; asking for
// CIGARs instead of raw op lists (ops_h) cuts the D2H traffic from ~2 L to 2 * cigar_stride bytes per read.
#define MAP_HOST_SLOT0 45
extern "C" int nvb_map_reads_host(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* genome, const uint8_t* reads_h,
                                  const uint32_t* off_h, uint32_t n, const nvb_map_params* mp, uint32_t reads_per_batch,
                                  nvb_read_hit* hits_h, uint8_t* ops_h, uint32_t ops_stride,
                                  uint16_t* cigar_h, uint32_t cigar_stride, nvb_map_stats* stats)
{
    if (!ctx || !mp || !fmi) return NVB_ERR_ARG;
    if (stats) { stats->seeds = stats->seed_hits = stats->candidates = stats->aligned = 0; }
    if (n == 0) return NVB_OK;
    if (!reads_h || !off_h || !hits_h || (cigar_h && cigar_stride == 0)) return NVB_ERR_ARG;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc;
    const bool want_ops = ops_h || cigar_h;
    const uint32_t rb = reads_per_batch ? reads_per_batch : (1u << 20);
    const uint32_t n_batches = (n + rb - 1) / rb;
    const int n_sets = n_batches > 1 ? 2 : 1;
    size_t max_bytes = 0;
    for (uint32_t k = 0; k < n_batches; ++k)
    {
        const uint32_t r0 = k * rb, r1 = r0 + rb < n ? r0 + rb : n;
        const size_t nb = (size_t)off_h[r1] - off_h[r0];
        if (nb > max_bytes) max_bytes = nb;
    }
    const uint32_t cap = n < rb ? n : rb;
    uint8_t *d_sym[2], *d_ops[2] = { nullptr, nullptr }; uint32_t* d_off[2]; nvb_read_hit* d_hits[2]; uint16_t* d_cig[2] = { nullptr, nullptr };
    for (int b = 0; b < n_sets; ++b)
    {
        MAP_STAGE(MAP_HOST_SLOT0 + 5 * b + 0, max_bytes + 64, d_sym[b]);
        MAP_STAGE(MAP_HOST_SLOT0 + 5 * b + 1, sizeof(uint32_t) * ((size_t)cap + 1), d_off[b]);
        MAP_STAGE(MAP_HOST_SLOT0 + 5 * b + 2, sizeof(nvb_read_hit) * (size_t)cap, d_hits[b]);
        if (want_ops) MAP_STAGE(MAP_HOST_SLOT0 + 5 * b + 3, (size_t)cap * ops_stride + 16, d_ops[b]);
        if (cigar_h)  MAP_STAGE(MAP_HOST_SLOT0 + 5 * b + 4, sizeof(uint16_t) * (size_t)cap * cigar_stride, d_cig[b]);
    }
    if (!ctx->host_ev[0])
        for (int i = 0; i < 12; ++i) NVB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->host_ev[i], cudaEventDisableTiming));
    cudaStream_t s_in = ctx->aux[0], s_k = ctx->stream, s_out = ctx->aux[1];
    auto copy_in = [&](uint32_t k) -> int
    {
        const int b = (int)(k & 1u) % n_sets;
        const uint32_t r0 = k * rb, r1 = r0 + rb < n ? r0 + rb : n;
        if (k >= (uint32_t)n_sets) NVB_CUDA(ctx, cudaStreamWaitEvent(s_in, ctx->host_ev[3 * b + 1], 0));   // batch k-2 has been mapped
        NVB_CUDA(ctx, cudaMemcpyAsync(d_sym[b], reads_h + off_h[r0], (size_t)off_h[r1] - off_h[r0], cudaMemcpyHostToDevice, s_in));
        NVB_CUDA(ctx, cudaMemcpyAsync(d_off[b], off_h + r0, sizeof(uint32_t) * ((size_t)(r1 - r0) + 1), cudaMemcpyHostToDevice, s_in));
        NVB_CUDA(ctx, cudaEventRecord(ctx->host_ev[3 * b], s_in));
        return NVB_OK;
    };
    if ((rc = copy_in(0))) return rc;
    // one k-mer table for the seeds of the whole call, sized on the longest read of the first batch
    nvb_fmindex fmi_call;
    {
        uint32_t maxL0 = 0;
        for (uint32_t i = 0; i < cap; ++i) { const uint32_t l = off_h[i + 1] - off_h[i]; if (l > maxL0) maxL0 = l; }
        const uint64_t per_read = (mp->seed_len && mp->seed_interval && maxL0 >= mp->seed_len ? (maxL0 - mp->seed_len) / mp->seed_interval + 1u : 0u) * (mp->both_strands ? 2u : 1u);
        fmi_call = *fmi;
        if (!mp->seed_mismatches && (rc = nvb_with_batch_table(ctx, fmi, (uint64_t)n * per_read, mp->seed_len, s_k, &fmi_call))) return rc;
        fmi = &fmi_call;
    }
    for (uint32_t k = 0; k < n_batches; ++k)
    {
        const int b = (int)(k & 1u) % n_sets;
        const uint32_t r0 = k * rb, r1 = r0 + rb < n ? r0 + rb : n, nr = r1 - r0;
        if (k + 1 < n_batches && (rc = copy_in(k + 1))) return rc;          // queued before this batch's compute blocks the host
        uint32_t maxL = 0;
        for (uint32_t i = r0; i < r1; ++i) { const uint32_t l = off_h[i + 1] - off_h[i]; if (l > maxL) maxL = l; }
        NVB_CUDA(ctx, cudaStreamWaitEvent(s_k, ctx->host_ev[3 * b], 0));
        if (k >= (uint32_t)n_sets) NVB_CUDA(ctx, cudaStreamWaitEvent(s_k, ctx->host_ev[3 * b + 2], 0));     // results of batch k-2 have left
        if (want_ops) NVB_CUDA(ctx, cudaMemsetAsync(d_ops[b], 0, (size_t)nr * ops_stride, s_k));
        nvb_map_stats st;
        if (maxL == 0)
        {
            // nothing to map: every read of the batch is unaligned
            std::vector<nvb_read_hit> none(nr);
            for (uint32_t i = 0; i < nr; ++i) { none[i].score = none[i].second_score = INT32_MIN; none[i].begin = none[i].end = 0xFFFFFFFFu; none[i].flags = 0; none[i].n_ops = 0; }
            NVB_CUDA(ctx, cudaMemcpyAsync(d_hits[b], none.data(), sizeof(nvb_read_hit) * nr, cudaMemcpyHostToDevice, s_k));
            NVB_CUDA(ctx, cudaStreamSynchronize(s_k));
            st.seeds = st.seed_hits = st.candidates = st.aligned = 0;
        }
        else
        {
            nvb_string_set q; q.symbols = d_sym[b] - off_h[r0]; q.offsets = d_off[b]; q.count = nr; q.bits = 8;
            if ((rc = nvb_map_reads(ctx, fmi, genome, q, maxL, mp, nr, d_hits[b], d_ops[b], ops_stride, &st, s_k))) return rc;
        }
        if (cigar_h)
        {
            map_cigar_kernel<<<map_grid(nr), MAP_THREADS, 0, s_k>>>(d_hits[b], d_ops[b], ops_stride, nr, d_cig[b], cigar_stride);
            NVB_LAUNCHED(ctx);
        }
        NVB_CUDA(ctx, cudaEventRecord(ctx->host_ev[3 * b + 1], s_k));
        NVB_CUDA(ctx, cudaStreamWaitEvent(s_out, ctx->host_ev[3 * b + 1], 0));
        NVB_CUDA(ctx, cudaMemcpyAsync(hits_h + r0, d_hits[b], sizeof(nvb_read_hit) * (size_t)nr, cudaMemcpyDeviceToHost, s_out));
        if (ops_h)   NVB_CUDA(ctx, cudaMemcpyAsync(ops_h + (size_t)r0 * ops_stride, d_ops[b], (size_t)nr * ops_stride, cudaMemcpyDeviceToHost, s_out));
        if (cigar_h) NVB_CUDA(ctx, cudaMemcpyAsync(cigar_h + (size_t)r0 * cigar_stride, d_cig[b], sizeof(uint16_t) * (size_t)nr * cigar_stride, cudaMemcpyDeviceToHost, s_out));
        NVB_CUDA(ctx, cudaEventRecord(ctx->host_ev[3 * b + 2], s_out));
        if (stats) { stats->seeds += st.seeds; stats->seed_hits += st.seed_hits; stats->candidates += st.candidates; stats->aligned += st.aligned; }
    }
    NVB_CUDA(ctx, cudaStreamSynchronize(s_out));
    NVB_CUDA(ctx, cudaStreamSynchronize(s_k));
    NVB_CUDA(ctx, cudaStreamSynchronize(s_in));
    return NVB_OK;
}
