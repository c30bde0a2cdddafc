// multi.cu -- one-process, many-context forms of the host entry points ("one host thread + stream per device, contiguous
// shards of one host batch, results written into disjoint ranges of one pinned array"). The path shards by independent units
// (queries, reads): there is no collective; every context maps its slice against its own index replica (contexts on different
// devices) or a shared one (several contexts on one device). The reference has no counterpart: nvBowtie drives one device per
// process thread the same way (nvBowtie/bowtie2/cuda/driver.cpp: one thread per GPU over a shared input stream).
#include "common.cuh"
#include <thread>
#include <vector>

// slice i of n units over k workers, boundaries on multiples of `grain`
static inline uint32_t shard_begin(uint32_t n, uint32_t k, uint32_t i, uint32_t grain)
{
    if (i >= k) return n;
    const uint64_t per = ((uint64_t)n + k - 1) / k;
    uint64_t b = per * i;
    b = (b / grain) * grain;
    return b > n ? n : (uint32_t)b;
}

extern "C" int nvb_fm_match_host_multi(nvb_ctx* const* ctxs, const nvb_fmindex* fmis_d, uint32_t n_ctx, const void* symbols_h, uint32_t bits,
                                       const uint32_t* offsets_h, uint32_t fixed_len, uint32_t count, uint32_t* ranges_h)
{
    if (!ctxs || !fmis_d || n_ctx == 0) return NVB_ERR_ARG;
    if (count == 0) return NVB_OK;
    if (!symbols_h || !ranges_h || (!offsets_h && fixed_len == 0)) return NVB_ERR_ARG;
    if (bits != 8 && bits != 4 && bits != 2) return NVB_ERR_ARG;
    for (uint32_t i = 0; i < n_ctx; ++i) if (!ctxs[i]) return NVB_ERR_ARG;
    std::vector<int> rcs(n_ctx, NVB_OK);
    std::vector<std::thread> workers;
    for (uint32_t i = 0; i < n_ctx; ++i)
    {
        // fixed-length packed streams: slices start on whole 32-bit words (32 queries of any length do)
        const uint32_t b = shard_begin(count, n_ctx, i, 32u), e = shard_begin(count, n_ctx, i + 1, 32u);
        if (e <= b) continue;
        workers.emplace_back([=, &rcs]()
        {
            if (offsets_h)
                rcs[i] = bits == 8 ? nvb_fm_match_host(ctxs[i], fmis_d + i, (const uint8_t*)symbols_h, offsets_h + b, e - b, ranges_h + 2 * (size_t)b)
                                   : nvb_fm_match_host_packed(ctxs[i], fmis_d + i, symbols_h, bits, offsets_h + b, e - b, ranges_h + 2 * (size_t)b);
            else
                rcs[i] = nvb_fm_match_host_fixed(ctxs[i], fmis_d + i, (const uint8_t*)symbols_h + (size_t)b * fixed_len * bits / 8u, bits, fixed_len, e - b,
                                                 ranges_h + 2 * (size_t)b);
        });
    }
    for (auto& w : workers) w.join();
    for (uint32_t i = 0; i < n_ctx; ++i) if (rcs[i]) return rcs[i];
    return NVB_OK;
}

extern "C" int nvb_map_reads_host_multi(nvb_ctx* const* ctxs, const nvb_fmindex* fmis_d, const uint32_t* const* genome_words_d, uint32_t n_ctx,
                                        const uint8_t* reads_h, const uint32_t* read_off_h, uint32_t n_reads, const nvb_map_params* params,
                                        uint32_t reads_per_batch, nvb_read_hit* hits_h, uint8_t* ops_h, uint32_t ops_stride,
                                        uint16_t* cigar_h, uint32_t cigar_stride, nvb_map_stats* stats_h)
{
    if (!ctxs || !fmis_d || !genome_words_d || n_ctx == 0 || !params) return NVB_ERR_ARG;
    if (stats_h) { stats_h->seeds = stats_h->seed_hits = stats_h->candidates = stats_h->aligned = 0; }
    if (n_reads == 0) return NVB_OK;
    if (!reads_h || !read_off_h || !hits_h) return NVB_ERR_ARG;
    for (uint32_t i = 0; i < n_ctx; ++i) if (!ctxs[i] || !genome_words_d[i]) return NVB_ERR_ARG;
    std::vector<int> rcs(n_ctx, NVB_OK);
    std::vector<nvb_map_stats> sts(n_ctx);
    std::vector<std::thread> workers;
    for (uint32_t i = 0; i < n_ctx; ++i)
    {
        sts[i].seeds = sts[i].seed_hits = sts[i].candidates = sts[i].aligned = 0;
        // even boundaries keep the mates of a pair (reads 2 p, 2 p + 1) in one slice
        const uint32_t b = shard_begin(n_reads, n_ctx, i, 2u), e = shard_begin(n_reads, n_ctx, i + 1, 2u);
        if (e <= b) continue;
        workers.emplace_back([=, &rcs, &sts]()
        {
            rcs[i] = nvb_map_reads_host(ctxs[i], fmis_d + i, genome_words_d[i], reads_h, read_off_h + b, e - b, params, reads_per_batch, hits_h + b,
                                        ops_h ? ops_h + (size_t)b * ops_stride : nullptr, ops_stride,
                                        cigar_h ? cigar_h + (size_t)b * cigar_stride : nullptr, cigar_stride, &sts[i]);
        });
    }
    for (auto& w : workers) w.join();
    for (uint32_t i = 0; i < n_ctx; ++i)
    {
        if (rcs[i]) return rcs[i];
        if (stats_h) { stats_h->seeds += sts[i].seeds; stats_h->seed_hits += sts[i].seed_hits; stats_h->candidates += sts[i].candidates; stats_h->aligned += sts[i].aligned; }
    }
    return NVB_OK;
}
