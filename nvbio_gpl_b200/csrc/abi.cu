// abi.cu -- context, index image management and the host-buffer (*_host) entry points.
#include "common.cuh"
#include <string>
#include <stdlib.h>
#include <string.h>
#include <new>

extern "C" int nvb_version(void) { return NVB_VERSION; }

extern "C" const char* nvb_status_string(int status)
{
    switch (status)
    {
    case NVB_OK:              return "ok";
    case NVB_ERR_ARG:         return "invalid argument";
    case NVB_ERR_CUDA:        return "CUDA error (no CPU fallback exists)";
    case NVB_ERR_TEMP:        return "temp buffer too small";
    case NVB_ERR_UNSUPPORTED: return "combination outside the contract";
    case NVB_ERR_TOO_LARGE:   return "problem larger than the built-in limit of a no-allocation entry point (split the batch)";
    default:                  return "unknown status";
    }
}

extern "C" int nvb_ctx_create(int device, nvb_ctx** out)
{
    if (!out) return NVB_ERR_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) { cudaGetLastError(); return NVB_ERR_CUDA; }
    if (cudaSetDevice(device) != cudaSuccess) return NVB_ERR_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return NVB_ERR_CUDA;
    if (prop.major < 10) return NVB_ERR_UNSUPPORTED;   // sm_100a cubins only
    nvb_ctx* c = new (std::nothrow) nvb_ctx();
    if (!c) return NVB_ERR_ARG;
    memset(c, 0, sizeof(*c));
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->l2_bytes = (size_t)prop.l2CacheSize;
    { const char* e = getenv("NVB_FM_TMA"); c->fm_tma = e ? atoi(e) != 0 : 1; }
    { const char* e = getenv("NVB_FM_SHARE_MIN"); c->fm_share_min = e ? (uint32_t)strtoul(e, nullptr, 10) : (1u << 16); }
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return NVB_ERR_CUDA; }
    for (int i = 0; i < 3; ++i)
        if (cudaStreamCreateWithFlags(&c->aux[i], cudaStreamNonBlocking) != cudaSuccess) { delete c; return NVB_ERR_CUDA; }
    // nvb_batch_score_noalloc: the address of its static scratch array is resolved here, once per device, so that the entry
    // point itself can be the first thing a caller runs under CUDA graph capture (a failure only surfaces there)
    (void)nvb_noalloc_warmup(c);
    c->last_err = cudaSuccess; cudaGetLastError();
    *out = c;
    return NVB_OK;
}

extern "C" void nvb_ctx_destroy(nvb_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (int i = 0; i < NVB_STAGE_SLOTS; ++i) if (c->d_stage[i]) cudaFree(c->d_stage[i]);
    if (c->d_worklist) cudaFree(c->d_worklist);
    for (int i = 0; i < 3; ++i) if (c->aux[i]) { cudaStreamSynchronize(c->aux[i]); cudaStreamDestroy(c->aux[i]); }
    if (c->tab_ev) cudaEventDestroy(c->tab_ev);
    for (int i = 0; i < 12; ++i) if (c->host_ev[i]) cudaEventDestroy(c->host_ev[i]);
    cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" void nvb_fm_batch_sharing(nvb_ctx* c, uint32_t min_batch) { if (c) c->fm_share_min = min_batch; }
extern "C" void* nvb_ctx_stream(nvb_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int nvb_ctx_synchronize(nvb_ctx* c)
{
    if (!c) return NVB_ERR_ARG;
    NVB_CUDA(c, cudaStreamSynchronize(c->stream));
    return NVB_OK;
}
extern "C" const char* nvb_last_cuda_error(nvb_ctx* c) { return c ? cudaGetErrorString(c->last_err) : "no context"; }
extern "C" uint64_t nvb_ctx_launch_count(nvb_ctx* c) { return c ? c->launches : 0; }
extern "C" uint32_t nvb_last_paths(nvb_ctx* c) { return c ? c->last_paths : 0; }
extern "C" void nvb_force_generic(nvb_ctx* c, int on) { if (c) c->force_generic = on; }

extern "C" int nvb_host_alloc(size_t bytes, void** p)
{
    if (!p) return NVB_ERR_ARG;
    return cudaHostAlloc(p, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? NVB_OK : NVB_ERR_CUDA;
}
extern "C" void nvb_host_free(void* p) { if (p) cudaFreeHost(p); }

int nvb_stage(nvb_ctx* ctx, int slot, size_t bytes, void** out)
{
    if (bytes == 0) bytes = 256;
    if (ctx->d_stage_bytes[slot] < bytes)
    {
        if (ctx->d_stage[slot]) { NVB_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); NVB_CUDA(ctx, cudaFree(ctx->d_stage[slot])); }
        ctx->d_stage[slot] = nullptr; ctx->d_stage_bytes[slot] = 0;
        const size_t cap = bytes + bytes / 4 + 4096;
        NVB_CUDA(ctx, cudaMalloc(&ctx->d_stage[slot], cap));
        ctx->d_stage_bytes[slot] = cap;
    }
    *out = ctx->d_stage[slot];
    return NVB_OK;
}

int nvb_worklist(nvb_ctx* ctx, size_t entries, uint32_t** out)
{
    if (ctx->d_worklist_cap < entries + 4)
    {
        if (ctx->d_worklist) { NVB_CUDA(ctx, cudaDeviceSynchronize()); NVB_CUDA(ctx, cudaFree(ctx->d_worklist)); }
        ctx->d_worklist = nullptr; ctx->d_worklist_cap = 0;
        NVB_CUDA(ctx, cudaMalloc(&ctx->d_worklist, sizeof(uint32_t) * (entries + 4)));
        ctx->d_worklist_cap = entries + 4;
    }
    *out = ctx->d_worklist;
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------
// index image
// ---------------------------------------------------------------------------------------------
extern "C" int nvb_fmindex_upload(nvb_ctx* ctx, const nvb_fmindex* h, nvb_fmindex* d)
{
    if (!ctx || !h || !d || !h->bwt_occ) return NVB_ERR_ARG;
    if (h->occ_int != 64u || h->sa_int != 16u) return NVB_ERR_UNSUPPORTED;
    *d = *h; d->bwt_occ = nullptr; d->ssa = nullptr; d->ktab = nullptr; d->ktab_k = 0; d->dssa = nullptr; d->dssa_int = 0;
    const size_t bw = nvb_bwt_occ_words(h->length) * sizeof(uint32_t);
    void* p = nullptr;
    NVB_CUDA(ctx, cudaMalloc(&p, bw ? bw : 32));
    NVB_CUDA(ctx, cudaMemcpyAsync(p, h->bwt_occ, bw, cudaMemcpyHostToDevice, ctx->stream));
    d->bwt_occ = p;
    if (h->ssa)
    {
        const size_t sw = nvb_ssa_words(h->length) * sizeof(uint32_t);
        void* s = nullptr;
        NVB_CUDA(ctx, cudaMalloc(&s, sw));
        NVB_CUDA(ctx, cudaMemcpyAsync(s, h->ssa, sw, cudaMemcpyHostToDevice, ctx->stream));
        d->ssa = (const uint32_t*)s;
    }
    NVB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------
// nvBWT on-disk index (written by the reference's nvBWT tool, nvBWT/nvBWT.cu:303-340; read by
// io::FMIndexDataHost::load, nvbio/io/fmindex/fmindex_impl.cu:111-321,333-479):
//   <prefix>.bwt / .rbwt : uint32 primary, uint32 cumFreq[4] (cumFreq[3] = text length), BWT words
//                          (2-bit symbols, 16 per big-endian uint32; padded here to a multiple of 4 words with zeros)
//   <prefix>.sa  / .rsa  : uint32 primary, 4 x uint32 (ignored: nvBWT.cu:337 writes pointer bytes there), uint32 SA_INT,
//                          uint32 text length, then ssa[1 ..] (ssa[0] = 0xFFFFFFFF is implied)
// The occurrence counters are not stored: the fused image is rebuilt on the device (nvb_fmindex_fuse).
// ---------------------------------------------------------------------------------------------
#include <cstdio>
#include <vector>
extern "C" int nvb_fmindex_load(nvb_ctx* ctx, const char* prefix, int reverse, nvb_fmindex* d)
{
    if (!ctx || !prefix || !d) return NVB_ERR_ARG;
    memset(d, 0, sizeof(*d));
    d->occ_int = 64; d->sa_int = 16;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    std::string name = std::string(prefix) + (reverse ? ".rbwt" : ".bwt");
    FILE* f = fopen(name.c_str(), "rb");
    if (!f) return NVB_ERR_ARG;
    uint32_t hdr[5];
    if (fread(hdr, sizeof(uint32_t), 5, f) != 5) { fclose(f); return NVB_ERR_ARG; }
    const uint32_t primary = hdr[0], length = hdr[4];
    const size_t words = nvb_bwt_occ_words(length) / 2;                  // ceil(length / 16) rounded up to 4
    std::vector<uint32_t> bwt(words ? words : 4, 0u);
    const size_t got = fread(bwt.data(), sizeof(uint32_t), words, f);
    fclose(f);
    if (((got + 3) & ~(size_t)3) != words) return NVB_ERR_ARG;           // fmindex_impl.cu:153-158
    void *d_bwt = nullptr, *d_img = nullptr;
    NVB_CUDA(ctx, cudaMalloc(&d_bwt, (words ? words : 4) * sizeof(uint32_t)));
    NVB_CUDA(ctx, cudaMalloc(&d_img, (words ? words : 4) * 2 * sizeof(uint32_t)));
    NVB_CUDA(ctx, cudaMemcpyAsync(d_bwt, bwt.data(), words * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    uint32_t L2[5];
    int rc = nvb_fmindex_fuse(ctx, (const uint32_t*)d_bwt, length, (uint32_t*)d_img, L2, ctx->stream);
    cudaFree(d_bwt);
    if (rc) { cudaFree(d_img); return rc; }
    d->length = length; d->primary = primary;
    for (int i = 0; i < 5; ++i) d->L2[i] = L2[i];
    d->bwt_occ = d_img;
    // the counts recomputed from the words must agree with the file's cumulative frequencies
    if (L2[1] != hdr[1] || L2[2] != hdr[2] || L2[3] != hdr[3] || L2[4] != hdr[4]) { nvb_fmindex_release(ctx, d); return NVB_ERR_ARG; }
    // sampled suffix array (optional, like the reference: a missing or mismatching file leaves ssa = NULL)
    name = std::string(prefix) + (reverse ? ".rsa" : ".sa");
    f = fopen(name.c_str(), "rb");
    if (f)
    {
        uint32_t sh[7];
        const size_t n_ssa = nvb_ssa_words(length);
        if (fread(sh, sizeof(uint32_t), 7, f) == 7 && sh[0] == primary && sh[5] == 16u && sh[6] == length)
        {
            std::vector<uint32_t> ssa(n_ssa, 0xFFFFFFFFu);
            if (n_ssa <= 1 || fread(ssa.data() + 1, sizeof(uint32_t), n_ssa - 1, f) == n_ssa - 1)
            {
                void* d_ssa = nullptr;
                if (cudaMalloc(&d_ssa, n_ssa * sizeof(uint32_t)) == cudaSuccess &&
                    cudaMemcpy(d_ssa, ssa.data(), n_ssa * sizeof(uint32_t), cudaMemcpyHostToDevice) == cudaSuccess)
                    d->ssa = (const uint32_t*)d_ssa;
            }
        }
        fclose(f);
    }
    return NVB_OK;
}

// writes <prefix>.bwt/.sa (or .rbwt/.rsa) from a device index: the inverse of nvb_fmindex_load, in nvBWT's format
extern "C" int nvb_fmindex_save(nvb_ctx* ctx, const nvb_fmindex* d, const char* prefix, int reverse)
{
    if (!ctx || !d || !prefix || !d->bwt_occ) return NVB_ERR_ARG;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t img_words = nvb_bwt_occ_words(d->length);
    std::vector<uint32_t> img(img_words ? img_words : 8);
    NVB_CUDA(ctx, cudaMemcpy(img.data(), d->bwt_occ, img_words * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    std::string name = std::string(prefix) + (reverse ? ".rbwt" : ".bwt");
    FILE* f = fopen(name.c_str(), "wb");
    if (!f) return NVB_ERR_ARG;
    const uint32_t hdr[5] = { d->primary, d->L2[1], d->L2[2], d->L2[3], d->L2[4] };
    fwrite(hdr, sizeof(uint32_t), 5, f);
    const size_t n_words = ((size_t)d->length + 15) / 16;                // nvBWT writes exactly the used words
    for (size_t w = 0; w < n_words; ++w) fwrite(&img[(w / 4) * 8 + (w & 3)], sizeof(uint32_t), 1, f);
    fclose(f);
    if (d->ssa)
    {
        const size_t n_ssa = nvb_ssa_words(d->length);
        std::vector<uint32_t> ssa(n_ssa);
        NVB_CUDA(ctx, cudaMemcpy(ssa.data(), d->ssa, n_ssa * sizeof(uint32_t), cudaMemcpyDeviceToHost));
        name = std::string(prefix) + (reverse ? ".rsa" : ".sa");
        f = fopen(name.c_str(), "wb");
        if (!f) return NVB_ERR_ARG;
        const uint32_t sh[7] = { d->primary, 0, 0, 0, 0, 16u, d->length };
        fwrite(sh, sizeof(uint32_t), 7, f);
        if (n_ssa > 1) fwrite(ssa.data() + 1, sizeof(uint32_t), n_ssa - 1, f);
        fclose(f);
    }
    return NVB_OK;
}

extern "C" int nvb_fmindex_release(nvb_ctx* ctx, nvb_fmindex* d)
{
    if (!ctx || !d) return NVB_ERR_ARG;
    if (d->bwt_occ) NVB_CUDA(ctx, cudaFree((void*)d->bwt_occ));
    if (d->ssa) NVB_CUDA(ctx, cudaFree((void*)d->ssa));
    d->bwt_occ = nullptr; d->ssa = nullptr;
    return NVB_OK;
}

extern "C" int nvb_fmindex_pin_l2(nvb_ctx* ctx, const nvb_fmindex* d, size_t bytes)
{
    if (!ctx || !d || !d->bwt_occ) return NVB_ERR_ARG;
    cudaDeviceProp prop;
    NVB_CUDA(ctx, cudaGetDeviceProperties(&prop, ctx->device));
    cudaStreamAttrValue attr;
    memset(&attr, 0, sizeof(attr));
    if (bytes)
    {
        size_t win = bytes;
        if ((size_t)prop.accessPolicyMaxWindowSize < win) win = (size_t)prop.accessPolicyMaxWindowSize;
        size_t carve = win;
        if ((size_t)prop.persistingL2CacheMaxSize < carve) carve = (size_t)prop.persistingL2CacheMaxSize;
        NVB_CUDA(ctx, cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve));
        attr.accessPolicyWindow.base_ptr  = (void*)d->bwt_occ;
        attr.accessPolicyWindow.num_bytes = win;
        attr.accessPolicyWindow.hitRatio  = win ? (float)((double)carve / (double)win) : 0.f;
        attr.accessPolicyWindow.hitProp   = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp  = cudaAccessPropertyStreaming;
    }
    else
    {
        attr.accessPolicyWindow.num_bytes = 0;
        attr.accessPolicyWindow.hitProp   = cudaAccessPropertyNormal;
        attr.accessPolicyWindow.missProp  = cudaAccessPropertyNormal;
    }
    NVB_CUDA(ctx, cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
    if (!bytes) NVB_CUDA(ctx, cudaCtxResetPersistingL2Cache());
    return NVB_OK;
}

// ---------------------------------------------------------------------------------------------
// host-buffer forms
// ---------------------------------------------------------------------------------------------
#define NVB_HOST_CHUNK (1u << 19)   // queries per pipelined chunk of the host entry points
#define NVB_HOST_SETS  4            // buffer sets in flight
#define NVB_HOST_SLOT0 33           // stage slots 33 .. 33 + 3 * NVB_HOST_SETS - 1

__global__ void fixed_offsets_kernel(uint32_t* off, uint32_t first, uint32_t n, uint32_t len)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) off[i] = (first + i) * len;
}

// Host buffers in, host buffers out, as a three-stage pipeline over chunks of the batch: copy-in stream (H2D of the
// symbols and offsets of chunk k), compute stream (the kernel), copy-out stream (D2H of the ranges), chained by
// events, NVB_HOST_SETS buffer sets in flight. PCIe is full duplex, so the call costs about max(H2D, D2H) instead
// of their sum (NVB_HOST_CHUNK / NVB_HOST_RAMPS in the environment tune the chunking). Pinned host
// memory (nvb_host_alloc) is needed for the overlap. offsets_h == NULL means `fixed_len` symbols per query
// (offsets are then generated on the device and never cross PCIe).
static int fm_match_host_impl(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint8_t* symbols_h, uint32_t bits,
                              const uint32_t* offsets_h, uint32_t fixed_len, uint32_t count, uint32_t* ranges_h,
                              const uint32_t* nlist_h = nullptr, uint32_t n_nlist = 0)
{
    if (!ctx || !fmi_d) return NVB_ERR_ARG;
    if (count == 0) return NVB_OK;
    if (!symbols_h || !ranges_h || (!offsets_h && fixed_len == 0)) return NVB_ERR_ARG;
    if (bits != 8 && bits != 4 && bits != 2) return NVB_ERR_ARG;
    if (!offsets_h && (uint64_t)count * fixed_len > 0xFFFFFFFFull) return NVB_ERR_ARG;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc;
    const uint32_t spw = 32u / bits;                                  // symbols per 32-bit word of a packed stream
    // bytes [b0, b1) of the host symbol array that hold symbols [s0, s1) (whole words for packed streams)
    #define SPAN_BEGIN(s0) (bits == 8 ? (size_t)(s0) : (size_t)((s0) / spw) * 4u)
    #define SPAN_END(s1)   (bits == 8 ? (size_t)(s1) : (size_t)(((s1) + spw - 1) / spw) * 4u)
    #define OFFSET_OF(i)   (offsets_h ? offsets_h[i] : (uint32_t)(i) * fixed_len)
    // chunk schedule: 1/4, 1/2, 1, 1, ..., 1, 1/2, 1/4 of NVB_HOST_CHUNK
    std::vector<uint32_t> cut; cut.push_back(0);
    {
        static uint32_t env_chunk = 0, env_ramps = 0;   // ramped first/last chunks measured no gain on B200 + PCIe 5 (tools/host_probe.py)
        if (!env_chunk)
        {
            const char* e = getenv("NVB_HOST_CHUNK"); env_chunk = e ? (uint32_t)atoi(e) : NVB_HOST_CHUNK;
            if (env_chunk < 4096u) env_chunk = NVB_HOST_CHUNK;
            const char* r = getenv("NVB_HOST_RAMPS"); if (r) env_ramps = (uint32_t)atoi(r);
        }
        const uint32_t C = env_chunk;
        const bool ramps = env_ramps && count >= 4u * C;
        uint32_t done = 0;
        if (ramps) { done += C / 4; cut.push_back(done); done += C / 2; cut.push_back(done); }
        const uint32_t mid = count - done - (ramps ? C / 2 + C / 4 : 0u);
        const uint32_t n_mid = (mid + C - 1) / C, each = (mid + n_mid - 1) / n_mid;
        for (uint32_t i = 0, left = mid; i < n_mid; ++i) { const uint32_t c = left < each ? left : each; done += c; left -= c; cut.push_back(done); }
        if (ramps) { done += C / 2; cut.push_back(done); done += C / 4; cut.push_back(done); }
    }
    const uint32_t n_chunks = (uint32_t)cut.size() - 1;
    size_t max_bytes = 0; uint32_t max_cnt = 0;
    for (uint32_t k = 0; k < n_chunks; ++k)
    {
        const size_t nb = SPAN_END(OFFSET_OF(cut[k + 1])) - SPAN_BEGIN(OFFSET_OF(cut[k]));
        if (nb > max_bytes) max_bytes = nb;
        if (cut[k + 1] - cut[k] > max_cnt) max_cnt = cut[k + 1] - cut[k];
    }
    const int n_sets = n_chunks < NVB_HOST_SETS ? (int)n_chunks : NVB_HOST_SETS;
    void *d_sym[NVB_HOST_SETS], *d_off[NVB_HOST_SETS], *d_rng[NVB_HOST_SETS];
    for (int b = 0; b < n_sets; ++b)
    {
        if ((rc = nvb_stage(ctx, NVB_HOST_SLOT0 + 3 * b + 0, max_bytes + 32, &d_sym[b]))) return rc;
        if ((rc = nvb_stage(ctx, NVB_HOST_SLOT0 + 3 * b + 1, sizeof(uint32_t) * ((size_t)max_cnt + 1), &d_off[b]))) return rc;
        if ((rc = nvb_stage(ctx, NVB_HOST_SLOT0 + 3 * b + 2, sizeof(uint32_t) * 2 * (size_t)max_cnt, &d_rng[b]))) return rc;
    }
    if (!ctx->host_ev[0])
        for (int i = 0; i < 3 * NVB_HOST_SETS; ++i) NVB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->host_ev[i], cudaEventDisableTiming));
    cudaStream_t s_in = ctx->aux[0], s_k = ctx->stream, s_out = ctx->aux[1];
    // the N exceptions of a 2-bit stream: (query, position of its last N), ascending by query; on the device before the first kernel
    void* d_nlist = nullptr; uint32_t nl_next = 0;
    if (n_nlist)
    {
        if ((rc = nvb_stage(ctx, 57, sizeof(uint32_t) * 2 * (size_t)n_nlist, &d_nlist))) return rc;
        NVB_CUDA(ctx, cudaMemcpyAsync(d_nlist, nlist_h, sizeof(uint32_t) * 2 * (size_t)n_nlist, cudaMemcpyHostToDevice, s_k));
    }
    // The kernels of a chunk are short and latency-bound (offsets, table, match, N fix-up: ~125 us in a row for 2^19 queries,
    // against 42 us of match work). When the upload is lighter than the download (2-bit seeds: 5 B in, 8 B out per query) they are
    // the critical path, so (1) the k-mer table is built ONCE for the call, before the first chunk, and handed to the chunks as the
    // table of a local descriptor, and (2) chunks alternate between two compute streams, so the tail kernels of one overlap the match
    // kernel of the next: 2.45 -> 2.05 ms for 10 M 20-mers. With a heavier upload (4-bit and 8-bit streams) the call is bound by the
    // two copy directions sharing the bus and the same change measured 8 % SLOWER (3.82 -> 3.50 G lookups/s), so those keep one table
    // per chunk on one stream. NVB_HOST_TABLE / NVB_HOST_STREAMS force either form (tools/host_probe4.py, profiles/r2_host_probe.txt).
    static int env_tab = -2, env_str = -2;
    if (env_tab == -2) { const char* e = getenv("NVB_HOST_TABLE"); env_tab = e ? (atoi(e) != 0) : -1; const char* t = getenv("NVB_HOST_STREAMS"); env_str = t ? atoi(t) : -1; }
    const bool light_upload = (SPAN_END(OFFSET_OF(count)) + (offsets_h ? 4ull * count : 0ull)) < 8ull * count;
    const int use_tab = env_tab >= 0 ? env_tab : (light_upload ? 1 : 0), use_str = env_str >= 0 ? env_str : (light_upload ? 2 : 1);
    nvb_fmindex fmi_call = *fmi_d;
    if (use_tab) { if ((rc = nvb_with_batch_table(ctx, fmi_d, count, 15u, s_k, &fmi_call))) return rc; fmi_d = &fmi_call; }
    const bool two = use_str >= 2 && use_tab && n_chunks > 1;       // a second stream needs the shared, read-only table
    if (two)
    {
        if (!ctx->tab_ev) NVB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->tab_ev, cudaEventDisableTiming));
        NVB_CUDA(ctx, cudaEventRecord(ctx->tab_ev, s_k));
        NVB_CUDA(ctx, cudaStreamWaitEvent(ctx->aux[2], ctx->tab_ev, 0));
    }
    // NVB_HOST_TRACE=1: per-chunk completion times of the three stages on stderr (tuning aid, tools/host_probe2.py)
    const bool trace = getenv("NVB_HOST_TRACE") != nullptr;
    std::vector<cudaEvent_t> tev;
    if (trace)
    {
        tev.resize(1 + 3 * (size_t)n_chunks);
        for (auto& e : tev) cudaEventCreate(&e);
        cudaEventRecord(tev[0], s_in);
    }
    for (uint32_t k = 0; k < n_chunks; ++k)
    {
        const int b = (int)(k % (uint32_t)n_sets);
        cudaEvent_t ev_in = ctx->host_ev[3 * b], ev_k = ctx->host_ev[3 * b + 1], ev_out = ctx->host_ev[3 * b + 2];
        const uint32_t c0 = cut[k], c1 = cut[k + 1], cnt = c1 - c0;
        const size_t b0 = SPAN_BEGIN(OFFSET_OF(c0)), nb = SPAN_END(OFFSET_OF(c1)) - b0;
        cudaStream_t s_k = (two && (k & 1u)) ? ctx->aux[2] : ctx->stream;                       // this chunk's compute stream
        if (k >= (uint32_t)n_sets) NVB_CUDA(ctx, cudaStreamWaitEvent(s_in, ev_out, 0));       // the set's previous ranges have left
        NVB_CUDA(ctx, cudaMemcpyAsync(d_sym[b], symbols_h + b0, nb, cudaMemcpyHostToDevice, s_in));
        if (offsets_h) NVB_CUDA(ctx, cudaMemcpyAsync(d_off[b], offsets_h + c0, sizeof(uint32_t) * ((size_t)cnt + 1), cudaMemcpyHostToDevice, s_in));
        NVB_CUDA(ctx, cudaEventRecord(ev_in, s_in));
        NVB_CUDA(ctx, cudaStreamWaitEvent(s_k, ev_in, 0));
        if (!offsets_h)
        {
            fixed_offsets_kernel<<<(cnt + 1 + 255) / 256, 256, 0, s_k>>>((uint32_t*)d_off[b], c0, cnt + 1, fixed_len);
            NVB_LAUNCHED(ctx);
        }
        // the chunk keeps its absolute offsets: rebase the symbol pointer instead
        nvb_string_set q; q.symbols = (const uint8_t*)d_sym[b] - b0; q.offsets = (const uint32_t*)d_off[b]; q.count = cnt; q.bits = bits;
        if ((rc = nvb_fm_match(ctx, fmi_d, q, (uint32_t*)d_rng[b], s_k))) return rc;
        if (n_nlist)
        {
            uint32_t nl_end = nl_next;
            while (nl_end < n_nlist && nlist_h[2 * (size_t)nl_end] < c1) ++nl_end;
            if ((rc = nvb_fm_fix_n(ctx, fmi_d, q, (const uint32_t*)d_nlist + 2 * (size_t)nl_next, nl_end - nl_next, c0, (uint32_t*)d_rng[b], s_k))) return rc;
            nl_next = nl_end;
        }
        NVB_CUDA(ctx, cudaEventRecord(ev_k, s_k));
        NVB_CUDA(ctx, cudaStreamWaitEvent(s_out, ev_k, 0));
        NVB_CUDA(ctx, cudaMemcpyAsync(ranges_h + 2 * (size_t)c0, d_rng[b], sizeof(uint32_t) * 2 * (size_t)cnt, cudaMemcpyDeviceToHost, s_out));
        NVB_CUDA(ctx, cudaEventRecord(ev_out, s_out));
        if (trace) { cudaEventRecord(tev[1 + 3 * k], s_in); cudaEventRecord(tev[2 + 3 * k], s_k); cudaEventRecord(tev[3 + 3 * k], s_out); }
    }
    #undef SPAN_BEGIN
    #undef SPAN_END
    #undef OFFSET_OF
    NVB_CUDA(ctx, cudaStreamSynchronize(s_out));
    NVB_CUDA(ctx, cudaStreamSynchronize(s_in));
    NVB_CUDA(ctx, cudaStreamSynchronize(s_k));
    NVB_CUDA(ctx, cudaStreamSynchronize(ctx->aux[2]));
    if (trace)
    {
        for (uint32_t k = 0; k < n_chunks; ++k)
        {
            float a = 0, b = 0, c = 0;
            cudaEventElapsedTime(&a, tev[0], tev[1 + 3 * k]); cudaEventElapsedTime(&b, tev[0], tev[2 + 3 * k]); cudaEventElapsedTime(&c, tev[0], tev[3 + 3 * k]);
            fprintf(stderr, "[nvb host trace] chunk %2u: in %.3f  kernel %.3f  out %.3f ms\n", k, a, b, c);
        }
        for (auto& e : tev) cudaEventDestroy(e);
    }
    return NVB_OK;
}

// Host buffers in, host buffers out. Large batches are cut into chunks that alternate between two streams, so that
// the H2D copy of one chunk, the kernel of another and the D2H copy of a third overlap (PCIe is full duplex):
// the call costs about max(H2D, D2H) instead of their sum. Pinned host memory (nvb_host_alloc) is needed for that.
extern "C" int nvb_fm_match_host(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint8_t* symbols_h,
                                 const uint32_t* offsets_h, uint32_t count, uint32_t* ranges_h)
{ return fm_match_host_impl(ctx, fmi_d, symbols_h, 8, offsets_h, 0, count, ranges_h); }

// the same over a packed host stream (bits = 4 or 2, big-endian 32-bit words: the layout of the reference's
// io::SequenceData read streams, nvbio/io/sequence/sequence.h), which moves 2x / 4x fewer bytes over PCIe
extern "C" int nvb_fm_match_host_packed(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const void* words_h, uint32_t bits,
                                        const uint32_t* offsets_h, uint32_t count, uint32_t* ranges_h)
{ return fm_match_host_impl(ctx, fmi_d, (const uint8_t*)words_h, bits, offsets_h, 0, count, ranges_h); }

// queries of one length (seeds): no offsets cross PCIe. bits = 8, 4 or 2; query i = symbols [i * len, (i + 1) * len)
extern "C" int nvb_fm_match_host_fixed(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const void* symbols_h, uint32_t bits,
                                       uint32_t query_len, uint32_t count, uint32_t* ranges_h)
{
    if (query_len == 0) return NVB_ERR_ARG;
    return fm_match_host_impl(ctx, fmi_d, (const uint8_t*)symbols_h, bits, nullptr, query_len, count, ranges_h);
}

// 2-bit queries of one length plus the list of queries that hold an N: n_list_h = n_count pairs (query index, position of the LAST N
// in that query), ascending by query index; the 2-bit code at an N position is ignored. 5 bytes per 20-mer cross the bus instead of 10.
extern "C" int nvb_fm_match_host_fixed2n(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const void* words_h, uint32_t query_len, uint32_t count,
                                         const uint32_t* n_list_h, uint32_t n_count, uint32_t* ranges_h)
{
    if (query_len == 0 || (n_count && !n_list_h)) return NVB_ERR_ARG;
    for (uint32_t i = 0; i < n_count; ++i)
        if (n_list_h[2 * (size_t)i] >= count || n_list_h[2 * (size_t)i + 1] >= query_len || (i && n_list_h[2 * (size_t)i] <= n_list_h[2 * (size_t)i - 2])) return NVB_ERR_ARG;
    return fm_match_host_impl(ctx, fmi_d, (const uint8_t*)words_h, 2, nullptr, query_len, count, ranges_h, n_list_h, n_count);
}

extern "C" int nvb_fm_locate_host(nvb_ctx* ctx, const nvb_fmindex* fmi_d, const uint32_t* rows_h, uint32_t n, uint32_t* pos_h)
{
    if (!ctx || !fmi_d) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!rows_h || !pos_h) return NVB_ERR_ARG;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    void *d_rows, *d_pos; int rc;
    if ((rc = nvb_stage(ctx, 0, sizeof(uint32_t) * (size_t)n, &d_rows))) return rc;
    if ((rc = nvb_stage(ctx, 1, sizeof(uint32_t) * (size_t)n, &d_pos))) return rc;
    NVB_CUDA(ctx, cudaMemcpyAsync(d_rows, rows_h, sizeof(uint32_t) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = nvb_fm_locate(ctx, fmi_d, (const uint32_t*)d_rows, n, (uint32_t*)d_pos, ctx->stream))) return rc;
    NVB_CUDA(ctx, cudaMemcpyAsync(pos_h, d_pos, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    NVB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NVB_OK;
}

static uint32_t max_len(const uint32_t* off, uint32_t n)
{
    uint32_t m = 0;
    for (uint32_t i = 0; i < n; ++i) { const uint32_t l = off[i + 1] - off[i]; if (l > m) m = l; }
    return m;
}

// Host string sets in, host sinks out. Batches of >= 65536 pairs are cut into chunks that run as a three-stage pipeline (H2D of the
// symbols and offsets of chunk k+1 on the copy-in stream, the kernels of chunk k, D2H of the sinks of chunk k-1; two buffer sets),
// so a call costs about max(copies, kernels) instead of their sum. bits = 8, 4 or 2: packed streams (the reference's
// PackedStream<uint32,uint8,BITS,true> layout, what its own sw-benchmark feeds) put 2x / 4x fewer bytes on the bus; a chunk keeps
// its absolute offsets and the device symbol pointer is rebased to the chunk's first (word-aligned) byte.
static int batch_score_host_impl(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band, uint32_t bits,
                                 const uint8_t* pat_h, const uint32_t* pat_off_h, const uint8_t* txt_h,
                                 const uint32_t* txt_off_h, uint32_t n, nvb_best_sink* sinks_h)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!pat_h || !pat_off_h || !txt_h || !txt_off_h || !sinks_h) return NVB_ERR_ARG;
    if (bits != 8 && bits != 4 && bits != 2) return NVB_ERR_ARG;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint32_t spw = 32u / bits;
    auto span_begin = [&](uint32_t s0) -> size_t { return bits == 8 ? (size_t)s0 : (size_t)(s0 / spw) * 4u; };
    auto span_end   = [&](uint32_t s1) -> size_t { return bits == 8 ? (size_t)s1 : (size_t)((s1 + spw - 1u) / spw) * 4u; };
    uint32_t chunk = n;
    if (n >= 65536u) { chunk = ((n + 7u) / 8u + 1023u) & ~1023u; if (chunk < 16384u) chunk = 16384u; }
    const uint32_t n_chunks = (n + chunk - 1u) / chunk;
    const int n_sets = n_chunks > 1 ? 2 : 1;
    // per-chunk sizes
    size_t max_pb = 0, max_tb = 0; uint32_t maxM = 0, maxN = 0;
    for (uint32_t k = 0; k < n_chunks; ++k)
    {
        const uint32_t c0 = k * chunk, c1 = c0 + chunk < n ? c0 + chunk : n;
        const size_t pb = span_end(pat_off_h[c1]) - span_begin(pat_off_h[c0]), tb = span_end(txt_off_h[c1]) - span_begin(txt_off_h[c0]);
        if (pb > max_pb) max_pb = pb;
        if (tb > max_tb) max_tb = tb;
    }
    maxM = max_len(pat_off_h, n); maxN = max_len(txt_off_h, n);
    const uint32_t cap = n < chunk ? n : chunk;
    const size_t temp_bytes = nvb_temp_bytes(band ? NVB_OP_BANDED_SCORE : NVB_OP_SCORE, aligner, band, 0, maxM, maxN, cap);
    const size_t ob = sizeof(uint32_t) * ((size_t)cap + 1);
    void *d_p[2], *d_po[2], *d_t[2], *d_to[2], *d_s[2], *d_tmp; int rc;
    for (int b = 0; b < n_sets; ++b)
    {
        // 64 spare bytes on both sides: a chunk's first / last string sits at the edge of ITS copy, and the banded kernels fetch whole
        // aligned 16-byte blocks (they guard the ends of the set they are given, which here spans the caller's whole array)
        if ((rc = nvb_stage(ctx, 64 + 5 * b + 0, max_pb + 128, &d_p[b]))) return rc;
        if ((rc = nvb_stage(ctx, 64 + 5 * b + 1, ob, &d_po[b]))) return rc;
        if ((rc = nvb_stage(ctx, 64 + 5 * b + 2, max_tb + 128, &d_t[b]))) return rc;
        if ((rc = nvb_stage(ctx, 64 + 5 * b + 3, ob, &d_to[b]))) return rc;
        if ((rc = nvb_stage(ctx, 64 + 5 * b + 4, sizeof(nvb_best_sink) * (size_t)cap, &d_s[b]))) return rc;
    }
    if ((rc = nvb_stage(ctx, 74, temp_bytes, &d_tmp))) return rc;
    if (!ctx->host_ev[0])
        for (int i = 0; i < 12; ++i) NVB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->host_ev[i], cudaEventDisableTiming));
    cudaStream_t s_in = n_sets > 1 ? ctx->aux[0] : ctx->stream, s_k = ctx->stream, s_out = n_sets > 1 ? ctx->aux[1] : ctx->stream;
    for (uint32_t k = 0; k < n_chunks; ++k)
    {
        const int b = (int)(k & 1u) % n_sets;
        cudaEvent_t ev_in = ctx->host_ev[3 * b], ev_k = ctx->host_ev[3 * b + 1], ev_out = ctx->host_ev[3 * b + 2];
        const uint32_t c0 = k * chunk, c1 = c0 + chunk < n ? c0 + chunk : n, cnt = c1 - c0;
        const size_t pb0 = span_begin(pat_off_h[c0]), pbn = span_end(pat_off_h[c1]) - pb0;
        const size_t tb0 = span_begin(txt_off_h[c0]), tbn = span_end(txt_off_h[c1]) - tb0;
        if (k >= (uint32_t)n_sets) NVB_CUDA(ctx, cudaStreamWaitEvent(s_in, ev_out, 0));          // the set's previous sinks have left
        NVB_CUDA(ctx, cudaMemcpyAsync((uint8_t*)d_p[b] + 64, pat_h + pb0, pbn, cudaMemcpyHostToDevice, s_in));
        NVB_CUDA(ctx, cudaMemcpyAsync(d_po[b], pat_off_h + c0, sizeof(uint32_t) * ((size_t)cnt + 1), cudaMemcpyHostToDevice, s_in));
        NVB_CUDA(ctx, cudaMemcpyAsync((uint8_t*)d_t[b] + 64, txt_h + tb0, tbn, cudaMemcpyHostToDevice, s_in));
        NVB_CUDA(ctx, cudaMemcpyAsync(d_to[b], txt_off_h + c0, sizeof(uint32_t) * ((size_t)cnt + 1), cudaMemcpyHostToDevice, s_in));
        if (n_sets > 1)
        {
            NVB_CUDA(ctx, cudaEventRecord(ev_in, s_in));
            NVB_CUDA(ctx, cudaStreamWaitEvent(s_k, ev_in, 0));
        }
        nvb_string_set P = { (const uint8_t*)d_p[b] + 64 - pb0, (const uint32_t*)d_po[b], cnt, bits }, T = { (const uint8_t*)d_t[b] + 64 - tb0, (const uint32_t*)d_to[b], cnt, bits };
        ctx->dp_first[0] = pat_off_h[c0]; ctx->dp_first[1] = txt_off_h[c0];
        if (band) rc = nvb_batch_banded_score(ctx, aligner, band, P, T, (nvb_best_sink*)d_s[b], maxM, maxN, d_tmp, temp_bytes, s_k);
        else      rc = nvb_batch_score(ctx, aligner, P, T, (nvb_best_sink*)d_s[b], maxM, maxN, d_tmp, temp_bytes, s_k);
        ctx->dp_first[0] = ctx->dp_first[1] = 0;
        if (rc) return rc;
        if (n_sets > 1)
        {
            NVB_CUDA(ctx, cudaEventRecord(ev_k, s_k));
            NVB_CUDA(ctx, cudaStreamWaitEvent(s_out, ev_k, 0));
        }
        NVB_CUDA(ctx, cudaMemcpyAsync(sinks_h + c0, d_s[b], sizeof(nvb_best_sink) * (size_t)cnt, cudaMemcpyDeviceToHost, s_out));
        if (n_sets > 1) NVB_CUDA(ctx, cudaEventRecord(ev_out, s_out));
    }
    NVB_CUDA(ctx, cudaStreamSynchronize(s_out));
    NVB_CUDA(ctx, cudaStreamSynchronize(s_in));
    NVB_CUDA(ctx, cudaStreamSynchronize(s_k));
    return NVB_OK;
}

extern "C" int nvb_batch_score_host(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band,
                                    const uint8_t* pat_h, const uint32_t* pat_off_h, const uint8_t* txt_h,
                                    const uint32_t* txt_off_h, uint32_t n, nvb_best_sink* sinks_h)
{ return batch_score_host_impl(ctx, aligner, band, 8, pat_h, pat_off_h, txt_h, txt_off_h, n, sinks_h); }

// packed host string sets: bits = 4 or 2, big-endian 32-bit words
extern "C" int nvb_batch_score_host_packed(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band, uint32_t bits,
                                           const void* pat_words_h, const uint32_t* pat_off_h, const void* txt_words_h,
                                           const uint32_t* txt_off_h, uint32_t n, nvb_best_sink* sinks_h)
{
    if (bits != 4 && bits != 2) return NVB_ERR_ARG;
    return batch_score_host_impl(ctx, aligner, band, bits, (const uint8_t*)pat_words_h, pat_off_h, (const uint8_t*)txt_words_h, txt_off_h, n, sinks_h);
}

extern "C" int nvb_batch_traceback_host(nvb_ctx* ctx, nvb_aligner aligner, uint32_t band, uint32_t checkpoints,
                                        const uint8_t* pat_h, const uint32_t* pat_off_h, const uint8_t* txt_h,
                                        const uint32_t* txt_off_h, uint32_t n, nvb_alignment* out_h,
                                        uint8_t* ops_h, uint32_t ops_stride, uint32_t* n_ops_h)
{
    if (!ctx) return NVB_ERR_ARG;
    if (n == 0) return NVB_OK;
    if (!pat_h || !pat_off_h || !txt_h || !txt_off_h || !out_h || !ops_h || !n_ops_h) return NVB_ERR_ARG;
    NVB_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t np = pat_off_h[n], nt = txt_off_h[n];
    const uint32_t maxM = max_len(pat_off_h, n), maxN = max_len(txt_off_h, n);
    const size_t tb = nvb_temp_bytes(band ? NVB_OP_BANDED_TRACEBACK : NVB_OP_TRACEBACK, aligner, band, checkpoints, maxM, maxN, n);
    void *d_p, *d_po, *d_t, *d_to, *d_o, *d_tmp, *d_ops, *d_nops; int rc;
    const size_t ob = sizeof(uint32_t) * ((size_t)n + 1);
    if ((rc = nvb_stage(ctx, 0, np + 32, &d_p))) return rc;
    if ((rc = nvb_stage(ctx, 1, ob, &d_po))) return rc;
    if ((rc = nvb_stage(ctx, 2, nt + 32, &d_t))) return rc;
    if ((rc = nvb_stage(ctx, 3, ob, &d_to))) return rc;
    if ((rc = nvb_stage(ctx, 4, sizeof(nvb_alignment) * (size_t)n, &d_o))) return rc;
    if ((rc = nvb_stage(ctx, 5, tb, &d_tmp))) return rc;
    if ((rc = nvb_stage(ctx, 6, (size_t)ops_stride * n, &d_ops))) return rc;
    if ((rc = nvb_stage(ctx, 7, sizeof(uint32_t) * (size_t)n, &d_nops))) return rc;
    NVB_CUDA(ctx, cudaMemcpyAsync(d_p, pat_h, np, cudaMemcpyHostToDevice, ctx->stream));
    NVB_CUDA(ctx, cudaMemcpyAsync(d_po, pat_off_h, ob, cudaMemcpyHostToDevice, ctx->stream));
    NVB_CUDA(ctx, cudaMemcpyAsync(d_t, txt_h, nt, cudaMemcpyHostToDevice, ctx->stream));
    NVB_CUDA(ctx, cudaMemcpyAsync(d_to, txt_off_h, ob, cudaMemcpyHostToDevice, ctx->stream));
    nvb_string_set P = { d_p, (const uint32_t*)d_po, n, 8 }, T = { d_t, (const uint32_t*)d_to, n, 8 };
    if (band) rc = nvb_batch_banded_traceback(ctx, aligner, band, checkpoints, P, T, (nvb_alignment*)d_o, (uint8_t*)d_ops,
                                              ops_stride, (uint32_t*)d_nops, maxM, maxN, d_tmp, tb, ctx->stream);
    else      rc = nvb_batch_traceback(ctx, aligner, checkpoints, P, T, (nvb_alignment*)d_o, (uint8_t*)d_ops,
                                       ops_stride, (uint32_t*)d_nops, maxM, maxN, d_tmp, tb, ctx->stream);
    if (rc) return rc;
    NVB_CUDA(ctx, cudaMemcpyAsync(out_h, d_o, sizeof(nvb_alignment) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    NVB_CUDA(ctx, cudaMemcpyAsync(ops_h, d_ops, (size_t)ops_stride * n, cudaMemcpyDeviceToHost, ctx->stream));
    NVB_CUDA(ctx, cudaMemcpyAsync(n_ops_h, d_nops, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    NVB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NVB_OK;
}
