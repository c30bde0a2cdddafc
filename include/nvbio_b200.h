/*
 * nvbio_b200.h -- C ABI of libnvbio_b200.so: the B200-native (sm_100a) seed-and-extend hot path.
 *
 * This is the drop-in boundary. Every entry point names the reference interface it replaces
 * (file:line into the nvbio-gpl tree). The reference API is header-only C++ templates over
 * "storage-free" views with no error channel; here the same views are plain structs of pointers
 * and sizes, every call returns an nvb_status, and nothing but CUDA is required to link.
 *
 * Conventions
 *   - pointers are DEVICE pointers unless the parameter name ends in _h or the function name
 *     ends in _host; `stream` is a cudaStream_t passed as void* (NULL = the default stream);
 *   - device entry points never synchronise and never allocate (scratch is passed in as `temp`,
 *     sized by nvb_temp_bytes); the *_host entry points stage through buffers owned by the
 *     nvb_ctx, run on the ctx stream and return after the results are in the host buffers;
 *   - one nvb_ctx per (device, host thread): it holds the device ordinal, a stream, the staging
 *     buffers of the *_host calls and the small "dirty pair" worklists of the DP kernels;
 *   - symbols: one uint8 per symbol (bits = 8), 0..3 = A,C,G,T, anything else = N (compares equal
 *     only to the same code, as `r == q` does in nvbio/alignment/sw/sw_inl.h:452);
 *     bits = 2 / 4: big-endian packed 32-bit words, the layout of
 *     PackedStream<const uint32*,uint8,BITS,true> (nvbio/basic/packedstream_inl.h:30-52);
 *   - there is NO CPU fallback: without a CUDA device every compute entry point returns
 *     NVB_ERR_CUDA.
 */
#ifndef NVBIO_B200_H
#define NVBIO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library itself is built with -fvisibility=hidden */
#endif

#define NVB_VERSION 100

typedef enum {
    NVB_OK            = 0,
    NVB_ERR_ARG       = -1,  /* NULL / out-of-contract argument                         */
    NVB_ERR_CUDA      = -2,  /* a CUDA runtime call failed (see nvb_last_cuda_error)    */
    NVB_ERR_TEMP      = -3,  /* temp buffer smaller than nvb_temp_bytes(...)            */
    NVB_ERR_UNSUPPORTED = -4,/* combination outside the contract        */
    NVB_ERR_TOO_LARGE = -5   /* problem larger than a fixed built-in limit (nvb_batch_score_noalloc) */
} nvb_status;

typedef struct nvb_ctx nvb_ctx;

/* ---------------------------------------------------------------------------------------------
 * FM-index view == nvbio::fm_index< rank_dictionary<2,64,...>, SSA_index_multiple_context<16> >
 * (nvbio/fmindex/fmindex.h:320-361) over the fused nvBWT image of io::FMIndexData
 * (nvbio/io/fmindex/fmindex.h:81-86,126-163; fusing loop nvbio/io/fmindex/fmindex_impl.cu:296-318):
 * per 64 BWT symbols one 32-byte block = 4 big-endian 2-bit words + 4 uint32 counters (A,C,G,T
 * occurrences BEFORE the block). bwt_occ must be 32-byte aligned.
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t        length;     /* fm_index::m_length  : text symbols                           */
    uint32_t        primary;    /* fm_index::m_primary : BWT row of the full text               */
    uint32_t        L2[5];      /* fm_index::m_L2      : cumulative symbol counts               */
    uint32_t        occ_int;    /* must be 64  (FMIndexDataCore::OCC_INT)                       */
    uint32_t        sa_int;     /* must be 16  (FMIndexDataCore::SA_INT)                        */
    const void*     bwt_occ;    /* fused array, 32 B per 64 symbols                             */
    const uint32_t* ssa;        /* (length+16)/16 entries, ssa[0] = 0xFFFFFFFF; may be NULL     */
    const uint32_t* ktab;       /* optional k-mer jump table (nvb_fmindex_build_ktab), else NULL */
    uint32_t        ktab_k;     /* symbols per table entry (0 = no table)                       */
    const uint32_t* dssa;       /* optional denser sampled SA (nvb_fmindex_build_dssa), else NULL */
    uint32_t        dssa_int;   /* its row interval: 2, 4 or 8 (0 = none)                       */
} nvb_fmindex;

/* ConcatenatedStringSet<symbols, offsets> (nvbio/strings/string_set.h:405-492) */
typedef struct {
    const void*     symbols;    /* bits==8: uint8 per symbol; bits==2|4: packed uint32 words    */
    const uint32_t* offsets;    /* count+1 symbol offsets                                       */
    uint32_t        count;
    uint32_t        bits;       /* 8, 4 or 2                                                    */
} nvb_string_set;

/* aligner tag objects (nvbio/alignment/alignment.h:365-519) with the "simple" scoring schemes
 * (nvbio/alignment/utils.h:81-124, ed/ed_utils.h:36-43) */
enum { NVB_EDIT_DISTANCE = 0, NVB_SMITH_WATERMAN = 1, NVB_GOTOH = 2 };
enum { NVB_GLOBAL = 0, NVB_LOCAL = 1, NVB_SEMI_GLOBAL = 2 };          /* alignment.h:242 */
typedef struct {
    int32_t kind;      /* NVB_EDIT_DISTANCE | NVB_SMITH_WATERMAN | NVB_GOTOH                    */
    int32_t type;      /* NVB_GLOBAL | NVB_LOCAL | NVB_SEMI_GLOBAL                              */
    int32_t match;     /* SimpleSmithWatermanScheme / SimpleGotohScheme ::m_match               */
    int32_t mismatch;  /* ::m_mismatch                                                          */
    int32_t gap_a;     /* SW: m_deletion   ; Gotoh: m_gap_open                                  */
    int32_t gap_b;     /* SW: m_insertion  ; Gotoh: m_gap_ext                                   */
} nvb_aligner;         /* kind == ED ignores the four scores: (0,-1,-1,-1)                      */

/* the fields of aln::BestSink<int32> (nvbio/alignment/sink.h, sink_inl.h:31-48), PACKED (12 bytes; the reference's struct is 16:
 * see nvb_sinks_to_reference_layout): sink = (text end, pattern end), 1-based; "no cell reported" = score -(1<<30), sink
 * (0xFFFFFFFF,0xFFFFFFFF) */
typedef struct { int32_t score; uint32_t sink_x, sink_y; } nvb_best_sink;

/* aln::Best2Sink<int32> (nvbio/alignment/sink.h:99-123, sink_inl.h:54-82): the best cell and the best runner-up whose
 * text end lies more than distinct_dist away from it; same "no cell reported" values as nvb_best_sink */
typedef struct { int32_t score1, score2; uint32_t sink1_x, sink1_y, sink2_x, sink2_y; } nvb_best2_sink;

/* the fields of aln::Alignment<int32> (nvbio/alignment/alignment.h:310-322), packed (20 bytes; the reference's struct is 24) */
typedef struct { int32_t score; uint32_t source_x, source_y, sink_x, sink_y; } nvb_alignment;

/* traceback ops, one uint8 per op in PUSH order (end -> start), as a Backtracer receives them
 * (nvbio/alignment/alignment.h:126-143, DirectionVector :326-337) */
enum { NVB_OP_SUBSTITUTION = 0, NVB_OP_INSERTION = 1, NVB_OP_DELETION = 2 };

/* ----------------------------------------------------------------------------------------- */
/* context                                                                                     */
/* ----------------------------------------------------------------------------------------- */
int         nvb_version(void);
const char* nvb_status_string(int status);
int         nvb_ctx_create(int device, nvb_ctx** ctx);
void        nvb_ctx_destroy(nvb_ctx* ctx);
void*       nvb_ctx_stream(nvb_ctx* ctx);              /* the ctx's own cudaStream_t */
int         nvb_ctx_synchronize(nvb_ctx* ctx);
const char* nvb_last_cuda_error(nvb_ctx* ctx);
/* number of kernels this library has launched through ctx since creation */
uint64_t    nvb_ctx_launch_count(nvb_ctx* ctx);

/* pinned host memory for the *_host entry points (plain cudaHostAlloc / cudaFreeHost) */
int         nvb_host_alloc(size_t bytes, void** ptr_h);
void        nvb_host_free(void* ptr_h);

/* ----------------------------------------------------------------------------------------- */
/* FM-index: index image                                                                       */
/* ----------------------------------------------------------------------------------------- */
/* io::FMIndexDataDevice(host_data) (nvbio/io/fmindex/fmindex_impl.cu:740-826): upload a host image
 * (fmi_h: bwt_occ / ssa are HOST pointers) and fill *fmi_d with device pointers owned by ctx. */
int nvb_fmindex_upload(nvb_ctx* ctx, const nvb_fmindex* fmi_h, nvb_fmindex* fmi_d);
int nvb_fmindex_release(nvb_ctx* ctx, nvb_fmindex* fmi_d);
/* io::FMIndexDataHost::load + FMIndexDataDevice (nvbio/io/fmindex/fmindex_impl.cu:111-321,333-479): reads the files the
 * reference's nvBWT tool writes (nvBWT/nvBWT.cu:303-340) -- <prefix>.bwt + .sa, or .rbwt + .rsa when reverse != 0 --
 * and rebuilds the fused occurrence image on the device. A missing .sa leaves fmi_d->ssa NULL (match/rank only). */
int nvb_fmindex_load(nvb_ctx* ctx, const char* prefix, int reverse, nvb_fmindex* fmi_d);
/* the inverse: writes a device index in nvBWT's on-disk format */
int nvb_fmindex_save(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const char* prefix, int reverse);
/* pin the first `bytes` of the fused array (the hot BWT prefix) in L2 for work on the ctx stream
 * (cudaAccessPolicyWindow); bytes = 0 removes the window. Replaces the reference's texture/__ldg
 * cache hint (nvbio/fmindex/fmindex.h:564-583). */
/* Optional k-mer jump table: entry i = the raw (x, y) pair match() returns for the k-mer whose symbols are the
 * base-4 digits of i (first symbol most significant), 2 uint32 per entry, 4^k entries (k <= 15: 8 GiB). With fmi->ktab set,
 * nvb_fm_match / nvb_fm_filter_rank / nvb_map_reads start a query of >= k symbols (none > 3 among its last k) from the
 * entry of its last k symbols instead of walking the first k steps of fmindex_inl.h:186-239; results are identical,
 * bit for bit (the loop would have stopped on the same pair). The table is the caller's memory, like the index. */
size_t nvb_ktab_words(uint32_t k);
int nvb_fmindex_build_ktab(nvb_ctx* ctx, const nvb_fmindex* fmi_d, uint32_t k, uint32_t* ktab_d, void* stream);
int nvb_fmindex_pin_l2(nvb_ctx* ctx, const nvb_fmindex* fmi_d, size_t bytes);
/* Optional denser sampled suffix array: entry i = locate(fmi, i * sa_int) computed once through the index's own SSA
 * (sa_int = 2, 4 or 8; (length + sa_int) / sa_int entries). With fmi->dssa set, every locate (nvb_fm_locate[_walk/_lookup],
 * nvb_fm_filter_locate, nvb_map_reads) stops its LF walk at the first row that is a multiple of sa_int instead of 16 and
 * reads the position there: the same positions, bit for bit, after a quarter (sa_int = 4) of the walk. Like the k-mer
 * table it is an index-side structure in the caller's memory; the nvBWT files and the reference layout are unchanged. */
size_t nvb_dssa_words(uint32_t length, uint32_t sa_int);
int nvb_fmindex_build_dssa(nvb_ctx* ctx, const nvb_fmindex* fmi_d, uint32_t sa_int, uint32_t* dssa_d, void* stream);
/* fuse a plain BWT (2-bit big-endian words, as in a .bwt file: nvBWT/nvBWT.cu:303-321) into the
 * interleaved bwt|occ image and compute L2; all pointers DEVICE; bwt_occ_out has
 * nvb_bwt_occ_words(length) words. Replaces nvbio/io/fmindex/fmindex_impl.cu:283-318 +
 * build_occurrence_table<64> (nvbio/fmindex/rank_dictionary_inl.h:34-66). */
size_t nvb_bwt_occ_words(uint32_t length);
size_t nvb_ssa_words(uint32_t length);
int nvb_fmindex_fuse(nvb_ctx* ctx, const uint32_t* bwt_words, uint32_t length, uint32_t* bwt_occ_out,
                     uint32_t* L2_out_h, void* stream);

/* ----------------------------------------------------------------------------------------- */
/* FM-index: queries                                                                           */
/* ----------------------------------------------------------------------------------------- */
/* match(fmi, pattern, len) for every string of the set (nvbio/fmindex/fmindex_inl.h:186-239; batch
 * form FMIndexFilter::rank, nvbio/fmindex/filter_inl.h:36-67,261-291).
 * ranges: 2 uint32 per query = the raw inclusive SA interval (x,y); empty iff x > y. */
int nvb_fm_match(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges,
                 void* stream);
/* Batches of at least min_batch queries run nvb_fm_match / nvb_fm_match_reverse / nvb_fm_filter_rank with their search
 * prefixes SHARED across the batch: the state after k steps depends only on the first k symbols consumed (fmindex_inl.h:186-239),
 * and among n queries the first K = log4(n / 2) of them take at most 4^K <= n / 2 distinct values. Each call first computes the
 * states of all 4^K K-mers (4^8 eight-step searches, then 4^K entries extended from those) into context-owned scratch and
 * starts every query from its entry -- the mechanism of nvb_fmindex_build_ktab, but rebuilt per call, sized by the batch
 * (8 * 4^K <= 4 bytes per query) and charged to the call. The ranges are identical bit for bit. Calls through one ctx must be
 * stream-ordered, and the first call at a larger batch size allocates (grow-only). Default 65536 (NVB_FM_SHARE_MIN in the
 * environment); 0xFFFFFFFF = always the plain loop. Continued ranges (nvb_fm_match_from) and smaller batches stay plain; a
 * caller's own table (fmi->ktab) is used instead when it is at least as deep. */
void nvb_fm_batch_sharing(nvb_ctx* ctx, uint32_t min_batch);
/* match_reverse(fmi, pattern, len) (nvbio/fmindex/fmindex_inl.h:252-278): the same search consuming the pattern front to
 * back, i.e. match() of the reversed pattern -- what nvBowtie runs against the index of the REVERSED text (.rbwt/.rsa,
 * nvb_fmindex_load(reverse = 1); nvBowtie/bowtie2/cuda/mapping_inl.h:71-86) */
int nvb_fm_match_reverse(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges,
                         void* stream);
/* match continued from given ranges: match(fmi, pattern, len, range) (fmindex_inl.h:212-239);
 * ranges is in/out */
int nvb_fm_match_from(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges,
                      void* stream);
/* nvBowtie's approximate seed mapping, map<find_exact>(query, len1, len2 = length, ...) (nvBowtie/bowtie2/cuda/mapping_inl.h:114-184,
 * driven by seed_mapper<APPROX_MAPPING>): the query is matched exactly over its first len1 search steps and with at most one
 * substitution over the remaining ones (rank4 at every such step, :146). Search order is match()'s: step t consumes symbol
 * length - 1 - t (the reference iterates a reversed seed stream). Every non-empty raw SA range is emitted in the reference's
 * push order -- substitutions by step then by symbol, the exact range last when find_exact != 0: counts[q] ranges were found,
 * the first min(counts[q], cap) of them are at ranges[2 * (q * cap + k)]. The priority deque the reference pushes them into
 * (seed_hit_deque_array.h) is the caller's. */
int nvb_fm_match_approx(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t len1, int find_exact,
                        uint32_t cap, uint32_t* ranges, uint32_t* counts, void* stream);
/* rank(fmi, range, c) (fmindex_inl.h:60-88): out[2i],out[2i+1] = occurrences of c[i] in
 * BWT[0..ranges[2i]] and BWT[0..ranges[2i+1]] (0xFFFFFFFF = -1 -> 0) */
int nvb_fm_rank(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* ranges, const uint8_t* c,
                uint32_t n, uint32_t* out, void* stream);
/* rank4(fmi, k) (fmindex_inl.h:100-125): out[4i+c] for c = A,C,G,T */
int nvb_fm_rank4(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* k, uint32_t n, uint32_t* out,
                 void* stream);
/* locate(fmi, row) (fmindex_inl.h:364-394): SA row -> text coordinate through the sampled SA */
int nvb_fm_locate(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* rows, uint32_t n,
                  uint32_t* pos, void* stream);
/* the two-phase form nvBowtie uses: locate_ssa_iterator / lookup_ssa_iterator
 * (fmindex_inl.h:408-460): walk[2i] = sampled row, walk[2i+1] = steps; then pos = ssa[row/16]+steps */
int nvb_fm_locate_walk(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* rows, uint32_t n,
                       uint32_t* walk, void* stream);
int nvb_fm_locate_lookup(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* walk, uint32_t n,
                         uint32_t* pos, void* stream);

/* host-buffer forms: H2D + kernel + D2H on the ctx stream, synchronous */
int nvb_fm_match_host(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint8_t* symbols_h,
                      const uint32_t* offsets_h, uint32_t count, uint32_t* ranges_h);
/* FMIndexFilter<device_tag, fm_index>::rank(index, string_set) (nvbio/fmindex/filter.h:108-165, filter_inl.h:261-300):
 * ranges[2 n] = match() of every string, slots[n] = inclusive scan of the range sizes (1 + y - x); both stay on the
 * device. *n_hits_h (optional) receives slots[n-1] = the total number of hits and synchronises the stream. */
int nvb_fm_filter_rank(nvb_ctx* ctx, const nvb_fmindex* fmi, nvb_string_set queries, uint32_t* ranges,
                       uint64_t* slots, uint64_t* n_hits_h, void* stream);
/* FMIndexFilter::locate(begin, end, hits) (filter_inl.h:75-110,302-393): output slot begin + i is the
 * (slot - slots[q-1])-th suffix-array row of query q = upper_bound(slot, slots); hits[2 i] = its text position,
 * hits[2 i + 1] = q. One thread per hit. */
int nvb_fm_filter_locate(nvb_ctx* ctx, const nvb_fmindex* fmi, const uint32_t* ranges, const uint64_t* slots,
                         uint32_t n_queries, uint64_t begin, uint64_t end, uint32_t* hits, void* stream);

/* the same over a packed host stream (bits = 4 or 2; big-endian 32-bit words as in the reference's read streams,
 * PackedStream<uint32*,uint8,BITS,true>: nvbio/io/sequence/sequence.h, nvbio/basic/packedstream.h) */
int nvb_fm_match_host_packed(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const void* words_h, uint32_t bits,
                             const uint32_t* offsets_h, uint32_t count, uint32_t* ranges_h);
/* queries of one length (seeds): query i = symbols [i * query_len, (i + 1) * query_len) of the 8-bit or packed
 * (4 / 2 bits, big-endian words) host stream; the offsets never cross PCIe */
int nvb_fm_match_host_fixed(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const void* symbols_h, uint32_t bits,
                            uint32_t query_len, uint32_t count, uint32_t* ranges_h);
/* the same over a 2-bit stream (the reference's genome-stream layout, PackedStream<uint32,uint8,2,true>), which cannot hold an N:
 * n_list_h = n_count pairs (query index, position of the LAST N of that query), ascending by query index; the 2-bit code stored at
 * an N position is ignored. Results are those of the 4/8-bit forms bit for bit (a query is answered by the search over the symbols
 * behind its last N and match()'s N rule); 5 bytes per 20-mer cross the bus instead of 10. */
int nvb_fm_match_host_fixed2n(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const void* words_h, uint32_t query_len, uint32_t count,
                              const uint32_t* n_list_h, uint32_t n_count, uint32_t* ranges_h);
int nvb_fm_locate_host(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint32_t* rows_h, uint32_t n,
                       uint32_t* pos_h);

/* ----------------------------------------------------------------------------------------- */
/* alignment                                                                                   */
/* ----------------------------------------------------------------------------------------- */
enum { NVB_OP_SCORE = 0, NVB_OP_BANDED_SCORE = 1, NVB_OP_TRACEBACK = 2, NVB_OP_BANDED_TRACEBACK = 3 };

/* scratch bytes for n pairs: BatchedAlignmentScore::min/max_temp_storage
 * (nvbio/alignment/batched.h:279-289, batched_inl.h:340-363,699-745) */
size_t nvb_temp_bytes(int op, nvb_aligner aligner, uint32_t band, uint32_t checkpoints,
                      uint32_t max_pattern_len, uint32_t max_text_len, uint32_t n);

/* aln::batch_alignment_score(aligner, patterns, texts, sinks, DeviceThreadScheduler(), maxM, maxN)
 * (nvbio/alignment/batched.h:98-140, batched_inl.h:980-1007; kernels batched_inl.h:79-128) */
int nvb_batch_score(nvb_ctx* ctx, nvb_aligner aligner, nvb_string_set patterns, nvb_string_set texts,
                    nvb_best_sink* sinks, uint32_t max_pattern_len, uint32_t max_text_len,
                    void* temp, size_t temp_bytes, void* stream);

/* the same for the aligner's TextBlockingTag variant (nvbio/alignment/alignment.h:266-279, aln::transpose :395-540; kernels
 * sw/sw_inl.h:872+, gotoh/gotoh_inl.h:848+; scoring only, as in the reference): same scores, LOCAL ties resolved in the
 * text-blocking visiting order (stripes of 8 text columns, pattern row by pattern row). temp: size it with
 * nvb_temp_bytes(NVB_OP_SCORE, aligner, 0, 0, max_text_len, max_pattern_len, n) for GLOBAL and LOCAL (the batch runs transposed). */
int nvb_batch_score_text_blocking(nvb_ctx* ctx, nvb_aligner aligner, nvb_string_set patterns, nvb_string_set texts,
                                  nvb_best_sink* sinks, uint32_t max_pattern_len, uint32_t max_text_len,
                                  void* temp, size_t temp_bytes, void* stream);

/* nvb_batch_score without any allocation. NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results as nvb_batch_score; the scratch (one int16 score column of max_text_len cells per pair,
 * two for Gotoh) is a static device array of NVB_NOALLOC_SCRATCH_BYTES inside the library, held by every GPU that loads it
 * until exit. Problems needing more -- nvb_batch_score_noalloc_bytes(aligner, max_text_len, n) > NVB_NOALLOC_SCRATCH_BYTES,
 * i.e. more than 16 Mi cells (8 Mi for Gotoh) -- are rejected with NVB_ERR_TOO_LARGE: split the batch. The call makes no
 * cudaMalloc / cudaFree and touches no context-owned buffer; it runs the int32 one-pair-per-thread kernels. All calls share
 * the array: calls that overlap on the GPU (other streams, other host threads, other contexts on the same device) race. */
#define NVB_NOALLOC_SCRATCH_BYTES (32u << 20)
size_t nvb_batch_score_noalloc_bytes(nvb_aligner aligner, uint32_t max_text_len, uint32_t n);
int nvb_batch_score_noalloc(nvb_ctx* ctx, nvb_aligner aligner, nvb_string_set patterns, nvb_string_set texts,
                            nvb_best_sink* sinks, uint32_t max_pattern_len, uint32_t max_text_len, void* stream);

/* aln::batch_banded_alignment_score<BAND_LEN>(...) (nvbio/alignment/batched.h:142-192,
 * batched_banded_inl.h:34-81,152-157). band in 2..64; pairs with text shorter than the pattern
 * are refused exactly as the reference does (sink left at "no cell reported"). */
int nvb_batch_banded_score(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band, nvb_string_set patterns,
                           nvb_string_set texts, nvb_best_sink* sinks, uint32_t max_pattern_len,
                           uint32_t max_text_len, void* temp, size_t temp_bytes, void* stream);

/* the same two calls with an aln::Best2Sink<int32>(distinct_dist) per pair instead of a BestSink (the sink nvBowtie's
 * scoring uses to get a second-best score, nvBowtie/bowtie2/cuda/score_inl.h); band = 0: full matrix, else banded */
int nvb_batch_score_best2(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band, uint32_t distinct_dist,
                          nvb_string_set patterns, nvb_string_set texts, nvb_best2_sink* sinks,
                          uint32_t max_pattern_len, uint32_t max_text_len, void* temp, size_t temp_bytes, void* stream);

/* batch_banded_alignment_score<BAND_LEN>(EditDistanceAligner<TYPE, MyersTag<ALPHABET>>, ...): Myers' bit-vector banded
 * edit distance (nvbio/alignment/alignment.h:274-275, myers/myers_banded_inl.h:183-296; used by examples/fmmap/fmmap.cu:347-358).
 * alphabet = 2, 4 or 5 (MyersTag's ALPHABET_SIZE), band = 1..32. Scores are negated distances; SEMI_GLOBAL reports every
 * column of the last pattern row that lies in the band and scores >= min_score, GLOBAL the last one, LOCAL nothing.
 * min_score is truncated to int16 as the reference's banded_myers does (:189) -- pass NVB_MYERS_BATCH_MIN_SCORE to get
 * the value the reference's batch drivers end up with (Field_traits<int32>::min() -> 0: exact matches only).
 * 8-bit symbol arrays must be 4-byte aligned (the kernel reads them as aligned words; NVB_ERR_ARG otherwise). */
#define NVB_MYERS_BATCH_MIN_SCORE (-(1 << 30))
int nvb_batch_banded_myers_score(nvb_ctx* ctx, int32_t type, uint32_t alphabet, uint32_t band, nvb_string_set patterns,
                                 nvb_string_set texts, int32_t min_score, nvb_best_sink* sinks, void* stream);

/* BatchedAlignmentTraceback<CHECKPOINTS,...>::enact (nvbio/alignment/batched.h:335-380,
 * batched_inl.h:599-680,760-830; per pair aln::alignment_traceback, alignment_inl.h:355-455).
 * ops: ops_stride bytes per pair, n_ops[i] of them valid (if n_ops[i] > ops_stride the list was
 * truncated); checkpoints must be a positive multiple of 16 (64 in nvBowtie, defs.h:96). The result does not depend on
 * it: pairs of 2-bit symbols whose scores fit 15 bits are traced by one packed forward pass that keeps the direction of
 * every cell plus a walk, everything else by the reference's checkpointed scheme. ops rows that are 16-byte aligned
 * (ops and ops_stride multiples of 16) are written 16 ops at a time. */
int nvb_batch_traceback(nvb_ctx* ctx, nvb_aligner aligner, uint32_t checkpoints,
                        nvb_string_set patterns, nvb_string_set texts, nvb_alignment* out,
                        uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops,
                        uint32_t max_pattern_len, uint32_t max_text_len,
                        void* temp, size_t temp_bytes, void* stream);

/* BatchedBandedAlignmentTraceback<BAND_LEN,CHECKPOINTS,...>::enact
 * (nvbio/alignment/batched.h:382-438, batched_banded_inl.h:243-321,331-453; per pair
 * aln::banded_alignment_traceback, banded_inl.h:354-430) */
int nvb_batch_banded_traceback(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band, uint32_t checkpoints,
                               nvb_string_set patterns, nvb_string_set texts, nvb_alignment* out,
                               uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops,
                               uint32_t max_pattern_len, uint32_t max_text_len,
                               void* temp, size_t temp_bytes, void* stream);

/* push-order ops -> run-length CIGAR in forward (start -> end) order, io::Cigar encoding
 * (nvbio/io/alignments.h:48-65: 2-bit op in the low bits | 14-bit length << 2); cigar_stride
 * uint16 per pair, n_cigar[i] entries written */
int nvb_ops_to_cigar(nvb_ctx* ctx, const uint8_t* ops, uint32_t ops_stride, const uint32_t* n_ops,
                     uint32_t n, uint16_t* cigar, uint32_t cigar_stride, uint32_t* n_cigar,
                     void* stream);

/* nvBowtie's MDS byte-code (the device-side precursor of SAM's MD:Z tag; finish_alignment_kernel,
 * nvBowtie/bowtie2/cuda/traceback_inl.h:555-690, op codes nvbio/io/alignments.h:37-42, turned into the MD string by
 * nvbio/io/output/output_sam.cpp:216-289) and the edit distance, from push-order op lists. Per alignment mds_stride
 * bytes: [len lo, len hi], then (0, run <= 255) for equal symbols | (1, read symbol) for a mismatch |
 * (2 insertion | 3 deletion, run & 255, the run's read / text symbols). alignments[i].source = (first text symbol,
 * first pattern symbol) of the alignment; mds_len (may be NULL) receives the full length, which may exceed mds_stride
 * (the stored code is then truncated); edit_distance may be NULL. 8/4/2-bit string sets. */
int nvb_ops_to_mds(nvb_ctx* ctx, const uint8_t* ops, uint32_t ops_stride, const uint32_t* n_ops, uint32_t n,
                   nvb_string_set patterns, nvb_string_set texts, const nvb_alignment* alignments,
                   uint8_t* mds, uint32_t mds_stride, uint32_t* mds_len, uint32_t* edit_distance, void* stream);

/* Quality-aware scoring schemes: the scheme interfaces of nvbio/alignment/alignment.h:411-436 (GotohScoringScheme) and :488-505
 * (SmithWatermanScoringScheme) evaluate match(q) and mismatch(a, b, q) at the quality q of each PATTERN position
 * (sw_inl.h:470-472, gotoh_inl.h:666,729, sw_banded_inl.h:391, gotoh_banded_inl.h:461); nvBowtie's scheme is one
 * (nvBowtie/bowtie2/cuda/scoring.h:206-301: constant match bonus, quality-scaled mismatch penalty, separate read and reference gap
 * penalties). Across the C ABI the two functors are 256-entry HOST tables indexed by q (schemes whose mismatch also depends on
 * the two symbols stay out of contract); aligner.match / .mismatch are ignored, aligner.gap_a / gap_b stay deletion / insertion
 * (SW) or pattern_gap_open / pattern_gap_extension (Gotoh: the reference runs the whole recurrence on the PATTERN gap
 * penalties, gotoh_inl.h:641-642); text_gap_open / text_gap_ext are Gotoh's text_gap_open() / text_gap_extension(), which the
 * reference reads only to initialise GLOBAL first columns / band rows and in the banded infimum (gotoh_inl.h:69-70,
 * gotoh_banded_inl.h:48,437-439). quals (DEVICE) holds one byte per pattern symbol, addressed by the patterns' offsets; NULL =
 * aln::trivial_quality_string (q = 0 everywhere). Edit distance has no scheme: qualities do not enter. These calls run the
 * int32 one-pair-per-thread kernels and copy the tables (2 KB) to a context-owned buffer first; otherwise they behave exactly
 * like the entry points without _q (same scratch, results, refusal rules). */
typedef struct {
    const int32_t* match_h;        /* 256 entries: scheme.match(q)                                  */
    const int32_t* mismatch_h;     /* 256 entries: scheme.mismatch(a, b, q), a != b                 */
    int32_t        text_gap_open;  /* Gotoh: scheme.text_gap_open()                                 */
    int32_t        text_gap_ext;   /* Gotoh: scheme.text_gap_extension()                            */
} nvb_quality_scheme;
int nvb_batch_score_q(nvb_ctx* ctx, nvb_aligner aligner, const nvb_quality_scheme* scheme, nvb_string_set patterns, const uint8_t* quals,
                      nvb_string_set texts, nvb_best_sink* sinks, uint32_t max_pattern_len, uint32_t max_text_len,
                      void* temp, size_t temp_bytes, void* stream);
int nvb_batch_banded_score_q(nvb_ctx* ctx, nvb_aligner aligner, const nvb_quality_scheme* scheme, uint32_t band, nvb_string_set patterns,
                             const uint8_t* quals, nvb_string_set texts, nvb_best_sink* sinks, uint32_t max_pattern_len,
                             uint32_t max_text_len, void* temp, size_t temp_bytes, void* stream);
int nvb_batch_traceback_q(nvb_ctx* ctx, nvb_aligner aligner, const nvb_quality_scheme* scheme, uint32_t checkpoints,
                          nvb_string_set patterns, const uint8_t* quals, nvb_string_set texts, nvb_alignment* out,
                          uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t max_pattern_len, uint32_t max_text_len,
                          void* temp, size_t temp_bytes, void* stream);
int nvb_batch_banded_traceback_q(nvb_ctx* ctx, nvb_aligner aligner, const nvb_quality_scheme* scheme, uint32_t band, uint32_t checkpoints,
                                 nvb_string_set patterns, const uint8_t* quals, nvb_string_set texts, nvb_alignment* out,
                                 uint8_t* ops, uint32_t ops_stride, uint32_t* n_ops, uint32_t max_pattern_len, uint32_t max_text_len,
                                 void* temp, size_t temp_bytes, void* stream);

/* The reference's own records are padded: CUDA aligns uint2 to 8 bytes, so aln::BestSink<int32> { int32 score; uint2 sink; }
 * (nvbio/alignment/sink.h:60-75) occupies 16 bytes (score, 4 bytes of padding, sink.x, sink.y) and aln::Alignment<int32>
 * { int32 score; uint2 source, sink; } (alignment.h:310-322) 24 bytes -- nvb_best_sink / nvb_alignment are the packed 12 / 20-byte
 * forms. These calls write n results in the reference's layout (best_sinks_out: 16-byte aligned, 16 n bytes; alignments_out:
 * 8-byte aligned, 24 n bytes); include/nvbio_b200_adaptor.h uses them to fill the caller's BestSink<int32>* / Alignment<int32>*. */
int nvb_sinks_to_reference_layout(nvb_ctx* ctx, const nvb_best_sink* sinks, uint32_t n, void* best_sinks_out, void* stream);
int nvb_alignments_to_reference_layout(nvb_ctx* ctx, const nvb_alignment* alignments, uint32_t n, void* alignments_out, void* stream);

/* host-buffer forms (H2D of both string sets, kernel(s), D2H of the results; synchronous) */
int nvb_batch_score_host(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band /* 0 = full matrix */,
                         const uint8_t* pat_h, const uint32_t* pat_off_h, const uint8_t* txt_h,
                         const uint32_t* txt_off_h, uint32_t n, nvb_best_sink* sinks_h);
/* the same over packed host string sets (bits = 4 or 2: PackedStream<uint32,uint8,BITS,true> words, symbol 0 of the set in the top
 * bits of word 0 -- the layout the reference's own alignment benchmark feeds): 2x / 4x fewer bytes cross the bus. Both forms cut
 * batches of >= 65536 pairs into chunks and overlap the upload of one chunk with the kernels of another. */
int nvb_batch_score_host_packed(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band /* 0 = full matrix */, uint32_t bits,
                                const void* pat_words_h, const uint32_t* pat_off_h, const void* txt_words_h,
                                const uint32_t* txt_off_h, uint32_t n, nvb_best_sink* sinks_h);
int nvb_batch_traceback_host(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band /* 0 = full matrix */,
                             uint32_t checkpoints, const uint8_t* pat_h, const uint32_t* pat_off_h,
                             const uint8_t* txt_h, const uint32_t* txt_off_h, uint32_t n,
                             nvb_alignment* out_h, uint8_t* ops_h, uint32_t ops_stride,
                             uint32_t* n_ops_h);

/* which kernel family served the last alignment call on ctx: bit 0 = packed DPX fast path,
 * bit 1 = int32 generic path (for tests and profiling; both may be set) */
uint32_t nvb_last_paths(nvb_ctx* ctx);
/* force the int32 generic kernels (parity anchor) : 0 = auto, 1 = generic only */
void nvb_force_generic(nvb_ctx* ctx, int on);

/* ----------------------------------------------------------------------------------------- */
/* seed -> locate -> extend -> reduce -> traceback over a read batch.         */
/* The nvBowtie steps it replaces, as one call: seed matching (nvBowtie/bowtie2/cuda/          */
/* mapping_inl.h:73-86), SA-row selection and locate (locate_inl.h:115-200), banded window     */
/* scoring (score_inl.h:85-130,482-583), best/second-best reduction (reduce_inl.h) and the     */
/* banded traceback of the best hit (traceback_inl.h:224-277). Re-seeding and effort limits    */
/* beyond a row budget are NOT part of it; approximate seeds and hit selection by range size   */
/* are modes of nvb_map_params, pairing / mate rescue / MDS / MAPQ separate calls below.        */
/* ----------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t    seed_len;       /* symbols per seed                              (params.h:131 seed_len)  */
    uint32_t    seed_interval;  /* a seed every seed_interval symbols, both strands                        */
    uint32_t    max_hits;       /* SA rows located per seed: the first max_hits rows of its range          */
    uint32_t    max_range;      /* seeds whose SA range is wider are skipped; 0 = no limit                 */
    uint32_t    band;           /* band of the extension; window = [g - band/2, g - band/2 + band + len)   */
    uint32_t    both_strands;   /* 0 = forward read only, 1 = also its reverse complement                  */
    int32_t     min_score;      /* reads whose best window scores less stay unaligned                      */
    nvb_aligner aligner;        /* type must be NVB_SEMI_GLOBAL                                            */
    /* nvBowtie's approximate seed mapping (mapping_inl.h:114-184, as called at :395-397 on the forward index): exact over the
     * first seed_len / 2 search steps, one substitution allowed in the rest, plus the exact match. 0 = exact seeds (default). The
     * first max_seed_ranges SA ranges of a seed in the reference's push order are kept (0 = 8, at most 64). The reverse-index
     * half of nvBowtie's scheme (:399-402, mismatches in the other half of the seed) needs reverse_index below. */
    uint32_t    seed_mismatches, max_seed_ranges;
    /* nvBowtie's hit selection by SA-range size (seed_hit.h:220-230, select_inl.h:93-113): 0 = every range contributes its first
     * max_hits rows (default); 1 = the ranges of a read go through a deque of max_seed_hits entries that drops its largest
     * range when full (0 = unbounded; mapping_inl.h:158-160), are visited smallest first (ties: seed order) and take at most
     * max_hits rows each until max_rows rows of the read are spent (0 = no budget) */
    uint32_t    hit_priority, max_seed_hits, max_rows;
    /* the FM-index of the REVERSED text (nvBWT's .rbwt / .rsa, loaded with nvb_fmindex_load(..., reverse = 1)), or NULL. With
     * seed_mismatches = 1 it adds the other half of nvBowtie's approximate seeds: exact over the first (seed_len + 1) / 2 symbols
     * of the seed, one substitution in the rest, the exact match excluded (map<IGNORE_EXACT> on rfmi, mapping_inl.h:399-402);
     * max_seed_ranges more range slots per seed; rows are located on this index and mapped back (length - q - seed_len) */
    const nvb_fmindex* reverse_index;
} nvb_map_params;

/* flags: bit 0 = reverse-complement strand, bit 1 = aligned. begin/end = genome coordinates of the first
 * aligned symbol / one past the last. Unaligned: score = second_score = INT32_MIN, begin = end = 0xFFFFFFFF. */
typedef struct { int32_t score, second_score; uint32_t begin, end, flags, n_ops; } nvb_read_hit;
typedef struct { uint64_t seeds, seed_hits, candidates, aligned; } nvb_map_stats;

/* 8-bit symbols (0..3) -> 2-bit big-endian words, 16 per uint32 (the layout of the reference's packed genome,
 * nvbio/io/sequence: PackedStream<uint32,uint8,2,true>); words_d holds (n + 15) / 16 words */
int nvb_pack_2bit(nvb_ctx* ctx, const uint8_t* symbols_d, uint32_t n, uint32_t* words_d, void* stream);

/* reads: 8-bit device string set; genome_words_d: the indexed text, 2-bit packed (nvb_pack_2bit);
 * hits_d[reads.count]; ops_d (may be NULL): ops_stride bytes per read (a multiple of 16, >= 2 * max_read_len + band),
 * push order like nvb_batch_banded_traceback; stats_h (may be NULL). Scratch is owned by the context (grow-only);
 * the call synchronises the stream a few times per batch of reads_per_batch reads (0 = default 2^20). */
int nvb_map_reads(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint32_t* genome_words_d, nvb_string_set reads,
                  uint32_t max_read_len, const nvb_map_params* params, uint32_t reads_per_batch,
                  nvb_read_hit* hits_d, uint8_t* ops_d, uint32_t ops_stride, nvb_map_stats* stats_h, void* stream);
/* Opposite-mate window search for pairs with exactly one aligned mate (nvBowtie: BestOppositeScoreStream::init_context,
 * nvBowtie/bowtie2/cuda/score_inl.h:318-411, frame_opposite_mate under PE_POLICY_FR; the opposite mate is scored with the full,
 * unbanded aligner over at most max_frag_len text symbols, :312). mates: the 2 n reads of n pairs laid out pair by pair (the set
 * nvb_map_reads mapped), hits_d[2 n] its hits, updated IN PLACE: the unaligned mate of such a pair is aligned on the strand the
 * policy demands against the window in which it completes a concordant pair -- anchor forward: [anchor.begin, anchor.begin +
 * max_frag_len), mate reverse-complemented; anchor reverse: [anchor.end - max_frag_len, anchor.end), mate forward; mates may
 * overlap -- by full-matrix SEMI_GLOBAL scoring, and, where the score reaches min_score, the full-matrix traceback. A rescued
 * mate gets flags = strand | 2 (aligned) | 4 (rescued), second_score = INT32_MIN, and its op list in ops_d (may be NULL; longer
 * lists are truncated to ops_stride, n_ops keeps the full length). The pair-score targets and the max_text_gaps trimming of
 * the reference's window (:341-392) are not part of it. Run nvb_pair_hits afterwards. Synchronises the stream twice. */
int nvb_rescue_mates(nvb_ctx* ctx, const uint32_t* genome_words_d, uint32_t genome_len, nvb_string_set mates,
                     uint32_t max_read_len, nvb_aligner aligner, int32_t min_score, uint32_t max_frag_len,
                     nvb_read_hit* hits_d, uint8_t* ops_d, uint32_t ops_stride, uint32_t* n_candidates_h, uint32_t* n_rescued_h,
                     void* stream);

/* the same byte-code for the hits of nvb_map_reads: read r against the genome from hits_d[r].begin on the hit's strand
 * (unaligned reads get the empty code [2, 0]); reads = the set that was mapped, ops_d / ops_stride = its op lists */
int nvb_map_hits_to_mds(nvb_ctx* ctx, const nvb_read_hit* hits_d, const uint8_t* ops_d, uint32_t ops_stride,
                        nvb_string_set reads, const uint32_t* genome_words_d, uint8_t* mds_d, uint32_t mds_stride,
                        uint32_t* mds_len_d, uint32_t* edit_distance_d, void* stream);
/* host buffers in / out, pipelined over read batches (reads of batch k+1 in, batch k mapped, results of batch k-1
 * out); synchronous. ops_h and cigar_h are both optional: cigar_h receives cigar_stride uint16 entries per read in
 * io::Cigar encoding (op | len << 2, forward order), the entry count in bits 8.. of the hit's flags; ops_stride is
 * needed (scratch layout) whenever either is asked for. */
int nvb_map_reads_host(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint32_t* genome_words_d,
                       const uint8_t* reads_h, const uint32_t* read_off_h, uint32_t n_reads,
                       const nvb_map_params* params, uint32_t reads_per_batch,
                       nvb_read_hit* hits_h, uint8_t* ops_h, uint32_t ops_stride,
                       uint16_t* cigar_h, uint32_t cigar_stride, nvb_map_stats* stats_h);

/* Mate pairing over the hits of 2 n reads laid out pair by pair (read 2 i = mate 1, 2 i + 1 = mate 2), both mapped
 * independently by nvb_map_reads: forward-reverse policy with a fragment-length window (nvBowtie's pe_policy = PE_POLICY_FR,
 * min_frag_len / max_frag_len, nvBowtie/bowtie2/cuda/params.h:147-152). flags: bit 0 = mate 1 aligned, bit 1 = mate 2
 * aligned, bit 2 = concordant (opposite strands, the forward mate starts and the reverse mate ends a fragment of
 * min_frag_len..max_frag_len). score = sum of the aligned mates' scores (INT32_MIN if none); frag_begin/frag_end = the hull
 * of the aligned mates. nvBowtie's anchor + opposite-mate window search (score_inl.h:385-411) is NOT part of it. */
typedef struct { int32_t score; uint32_t frag_begin, frag_end, flags; } nvb_pair_hit;
int nvb_pair_hits(nvb_ctx* ctx, const nvb_read_hit* hits_d, uint32_t n_pairs, uint32_t min_frag_len, uint32_t max_frag_len,
                  nvb_pair_hit* pairs_d, void* stream);

/* Mapping quality of every hit: nvBowtie's BowtieMapq2 (nvBowtie/bowtie2/cuda/mapq.h:145-316, Bowtie 2's V2 calculator) over
 * (score, second_score) of the hit. The scheme enters through two things only, as in the reference (it asks its scheme for
 * perfect_score(len), min_score(len) and m_monotone): match_bonus (perfect_score = len * match(0), scoring.h:254) and
 * min_score_by_len_h[0 .. max_len], a HOST table of scheme.min_score(len) that the caller evaluates (a float expression,
 * params.h:95-100; longer reads use entry max_len). monotone: 1 = end-to-end scoring, 0 = local. pairs_d (may be NULL): the
 * records of nvb_pair_hits over the same hits (read 2 i, 2 i + 1 = mates of pair i); mates of a concordant pair are rated on
 * the summed scores and ranges without a runner-up (io/alignments.h:220-228). Unaligned reads get 0. */
int nvb_map_hits_mapq(nvb_ctx* ctx, const nvb_read_hit* hits_d, const uint32_t* read_offsets_d, uint32_t n_reads,
                      const nvb_pair_hit* pairs_d, const int32_t* min_score_by_len_h, uint32_t max_len, int32_t match_bonus,
                      uint32_t monotone, uint8_t* mapq_d, void* stream);

/* ----------------------------------------------------------------------------------------- */
/* One process, several contexts: the batch is cut into n_ctx contiguous slices, one host thread per context runs the   */
/* single-context host entry point on its slice against ITS index (fmis_d[i], genome_words_d[i]: a replica per device, or the same */
/* pointers for contexts that share a device), and every slice writes its own range of the caller's result arrays. No collective,  */
/* no counterpart in the reference's library (nvBowtie runs one thread per GPU the same way). Slices start on multiples of 32       */
/* queries / 2 reads (packed fixed-length streams stay word-aligned, mates stay together). stats_h receives the sums.               */
/* ----------------------------------------------------------------------------------------- */
int nvb_fm_match_host_multi(nvb_ctx* const* ctxs, const nvb_fmindex* fmis_d, uint32_t n_ctx, const void* symbols_h, uint32_t bits,
                            const uint32_t* offsets_h /* NULL: queries of fixed_len symbols */, uint32_t fixed_len, uint32_t count,
                            uint32_t* ranges_h);
int nvb_map_reads_host_multi(nvb_ctx* const* ctxs, const nvb_fmindex* fmis_d, const uint32_t* const* genome_words_d, uint32_t n_ctx,
                             const uint8_t* reads_h, const uint32_t* read_off_h, uint32_t n_reads, const nvb_map_params* params,
                             uint32_t reads_per_batch, nvb_read_hit* hits_h, uint8_t* ops_h, uint32_t ops_stride,
                             uint16_t* cigar_h, uint32_t cigar_stride, nvb_map_stats* stats_h);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* NVBIO_B200_H */
