/*
 * ref_gpu.cu -- TEST / MEASUREMENT INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * "The reference's GPU path on this box" (informational): the reference's OWN CUDA kernels -- included from
 * /root/reference, never copied -- compiled for sm_100a and timed on the same B200 next to ours:
 *   aln::batch_alignment_score(aligner, patterns, texts, sinks, DeviceThreadScheduler(), M, N)     nvbio/alignment/batched.h:98-140
 *   aln::batch_banded_alignment_score<15>(...)                                                     nvbio/alignment/batched.h:142-192
 *   match(fmi, pattern, len) from one thread per query, as nvbio-test/fmindex_test.cu:56-85 does
 * Built by oracle/Makefile (target refgpu) into oracle/_ref/ref_gpu; prints one JSON line per measurement.
 *   usage: ref_gpu [sw_pairs] [banded_pairs] [genome_symbols] [queries]
 */
#include <nvbio/basic/types.h>
#include <nvbio/basic/packedstream.h>
#include <nvbio/basic/deinterleaved_iterator.h>
#include <nvbio/basic/vector_view.h>
#include <nvbio/strings/string_set.h>
#include <nvbio/fmindex/bwt.h>
#include <nvbio/fmindex/rank_dictionary.h>
#include <nvbio/fmindex/ssa.h>
#include <nvbio/fmindex/fmindex.h>
#include <nvbio/alignment/alignment.h>
#include <nvbio/alignment/batched.h>
#include <nvbio/alignment/sink.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <random>

#define ORC_PREFIX ref_
#include "oracle_abi.h"

using namespace nvbio;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("{\"error\": \"%s at %s:%d\"}\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

typedef ConcatenatedStringSet<const uint8*, const uint32*> string_set_type;

static void make_pairs(uint32 n, uint32 M, uint32 N, uint32 seed, std::vector<uint8>& P, std::vector<uint32>& Po,
                       std::vector<uint8>& T, std::vector<uint32>& To, bool implant_all)
{
    std::mt19937 rng(seed);
    P.resize(size_t(n) * M); T.resize(size_t(n) * N); Po.resize(n + 1); To.resize(n + 1);
    for (uint32 i = 0; i <= n; ++i) { Po[i] = i * M; To[i] = i * N; }
    for (size_t i = 0; i < T.size(); ++i) T[i] = rng() & 3u;
    for (uint32 i = 0; i < n; ++i)
    {
        if (implant_all || (i & 1u))
        {
            const uint32 s = N > M ? rng() % (N - M + 1) : 0u;
            for (uint32 j = 0; j < M; ++j) P[size_t(i) * M + j] = ((rng() % 100) < 5) ? (rng() & 3u) : T[size_t(i) * N + s + j];
        }
        else for (uint32 j = 0; j < M; ++j) P[size_t(i) * M + j] = rng() & 3u;
    }
}

template <typename F>
static float time_ms(F f, int reps)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r)
    {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}

// ---- FM-index: the exact typedefs of io::FMIndexData (nvbio/io/fmindex/fmindex.h:150-163), over device memory
typedef const uint4*                                   bwt_occ_type;
typedef deinterleaved_iterator<2,0,bwt_occ_type>       bwt_type;
typedef deinterleaved_iterator<2,1,bwt_occ_type>       occ_type;
typedef PackedStream<bwt_type,uint8,2u,true>           bwt_stream_type;
typedef rank_dictionary<2u,64u,bwt_stream_type,occ_type,const uint32*> rank_dict_type;
typedef SSA_index_multiple_context<16u,const uint32*>  ssa_type;
typedef fm_index<rank_dict_type,ssa_type>              fm_index_type;

__global__ void ref_match_kernel(const fm_index_type fmi, const uint8* q, const uint32 qlen, const uint32 n, uint2* ranges)
{
    const uint32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ranges[i] = match( fmi, q + size_t(i) * qlen, qlen );
}

int main(int argc, char** argv)
{
    const uint32 n_sw   = argc > 1 ? atoi(argv[1]) : 100000;
    const uint32 n_band = argc > 2 ? atoi(argv[2]) : 1000000;
    const uint32 G      = argc > 3 ? atoi(argv[3]) : 100000000;
    const uint32 n_q    = argc > 4 ? atoi(argv[4]) : 10000000;

    // ---------------- configs[0]: SW local, 150 x 500
    {
        const uint32 M = 150, N = 500;
        std::vector<uint8> P, T; std::vector<uint32> Po, To;
        make_pairs(n_sw, M, N, 1, P, Po, T, To, false);
        uint8 *dP, *dT; uint32 *dPo, *dTo; aln::BestSink<int32>* dS;
        CK(cudaMalloc(&dP, P.size())); CK(cudaMalloc(&dT, T.size())); CK(cudaMalloc(&dPo, Po.size() * 4)); CK(cudaMalloc(&dTo, To.size() * 4));
        CK(cudaMalloc(&dS, sizeof(aln::BestSink<int32>) * n_sw));
        CK(cudaMemcpy(dP, P.data(), P.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dT, T.data(), T.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dPo, Po.data(), Po.size() * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dTo, To.data(), To.size() * 4, cudaMemcpyHostToDevice));
        const string_set_type patterns(n_sw, dP, dPo), texts(n_sw, dT, dTo);
        const aln::SimpleSmithWatermanScheme scheme(2, -1, -1, -1);
        const float ms = time_ms([&]() {
            aln::batch_alignment_score(aln::make_smith_waterman_aligner<aln::LOCAL>(scheme), patterns, texts, dS, aln::DeviceThreadScheduler(), M, N);
        }, 3);
        CK(cudaGetLastError());
        std::vector<aln::BestSink<int32> > hs(n_sw); CK(cudaMemcpy(hs.data(), dS, sizeof(hs[0]) * n_sw, cudaMemcpyDeviceToHost));
        long long sum = 0; for (uint32 i = 0; i < n_sw; ++i) sum += hs[i].score;
        printf("{\"what\": \"reference CUDA batch_alignment_score, SW LOCAL (2,-1,-1,-1), DeviceThreadScheduler\", \"pairs\": %u, \"M\": %u, \"N\": %u, \"ms\": %.3f, \"gcups\": %.1f, \"score_sum\": %lld}\n",
               n_sw, M, N, ms, double(n_sw) * M * N / ms / 1e6, sum);
        cudaFree(dP); cudaFree(dT); cudaFree(dPo); cudaFree(dTo); cudaFree(dS);
    }
    // ---------------- configs[2]: banded (15), 250 x 265, edit distance and Gotoh, SEMI_GLOBAL
    {
        const uint32 M = 250, N = 265;
        std::vector<uint8> P, T; std::vector<uint32> Po, To;
        make_pairs(n_band, M, N, 3, P, Po, T, To, true);
        uint8 *dP, *dT; uint32 *dPo, *dTo; aln::BestSink<int32>* dS;
        CK(cudaMalloc(&dP, P.size())); CK(cudaMalloc(&dT, T.size())); CK(cudaMalloc(&dPo, Po.size() * 4)); CK(cudaMalloc(&dTo, To.size() * 4));
        CK(cudaMalloc(&dS, sizeof(aln::BestSink<int32>) * n_band));
        CK(cudaMemcpy(dP, P.data(), P.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dT, T.data(), T.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dPo, Po.data(), Po.size() * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dTo, To.data(), To.size() * 4, cudaMemcpyHostToDevice));
        const string_set_type patterns(n_band, dP, dPo), texts(n_band, dT, dTo);
        float ms = time_ms([&]() {
            aln::batch_banded_alignment_score<15>(aln::make_edit_distance_aligner<aln::SEMI_GLOBAL>(), patterns, texts, dS, aln::DeviceThreadScheduler(), M, N);
        }, 3);
        CK(cudaGetLastError());
        printf("{\"what\": \"reference CUDA batch_banded_alignment_score<15>, edit distance SEMI_GLOBAL\", \"pairs\": %u, \"ms\": %.3f, \"gcups\": %.1f}\n",
               n_band, ms, double(n_band) * M * 15 / ms / 1e6);
        const aln::SimpleGotohScheme gscheme(0, -5, -8, -3);
        ms = time_ms([&]() {
            aln::batch_banded_alignment_score<15>(aln::make_gotoh_aligner<aln::SEMI_GLOBAL>(gscheme), patterns, texts, dS, aln::DeviceThreadScheduler(), M, N);
        }, 3);
        CK(cudaGetLastError());
        printf("{\"what\": \"reference CUDA batch_banded_alignment_score<15>, Gotoh (0,-5,-8,-3) SEMI_GLOBAL\", \"pairs\": %u, \"ms\": %.3f, \"gcups\": %.1f}\n",
               n_band, ms, double(n_band) * M * 15 / ms / 1e6);
        cudaFree(dP); cudaFree(dT); cudaFree(dPo); cudaFree(dTo); cudaFree(dS);
    }
    // ---------------- configs[1]: exact 20-mer backward search, one thread per query
    {
        std::vector<uint8> text(G);
        std::mt19937_64 rng(2);
        for (uint32 i = 0; i < G; ++i) text[i] = rng() & 3u;
        std::vector<uint32> bwt_occ(ref_bwt_occ_words(G)), ssa(ref_ssa_words(G));
        uint32 primary = 0, L2[5];
        if (ref_build_index(text.data(), G, &primary, L2, bwt_occ.data(), ssa.data(), NULL)) { printf("{\"error\": \"index build\"}\n"); return 1; }
        const uint32 QL = 20;
        std::vector<uint8> q(size_t(n_q) * QL);
        for (uint32 i = 0; i < n_q; ++i)
        {
            if (rng() % 10) { const uint32 p = rng() % (G - QL); for (uint32 j = 0; j < QL; ++j) q[size_t(i) * QL + j] = text[p + j]; }
            else for (uint32 j = 0; j < QL; ++j) q[size_t(i) * QL + j] = rng() & 3u;
        }
        uint32 *d_img, *d_ssa, *d_L2, *d_ct; uint8* d_q; uint2* d_r;
        CK(cudaMalloc(&d_img, bwt_occ.size() * 4)); CK(cudaMalloc(&d_ssa, ssa.size() * 4)); CK(cudaMalloc(&d_L2, 20)); CK(cudaMalloc(&d_ct, 1024));
        CK(cudaMalloc(&d_q, q.size())); CK(cudaMalloc(&d_r, sizeof(uint2) * n_q));
        uint32 ct[256]; gen_bwt_count_table(ct);
        CK(cudaMemcpy(d_img, bwt_occ.data(), bwt_occ.size() * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_ssa, ssa.data(), ssa.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_L2, L2, 20, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_ct, ct, 1024, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_q, q.data(), q.size(), cudaMemcpyHostToDevice));
        const bwt_occ_type base = (bwt_occ_type)d_img;
        const fm_index_type fmi(G, primary, d_L2, rank_dict_type(bwt_stream_type(bwt_type(base)), occ_type(base), d_ct), ssa_type(d_ssa));
        const float ms = time_ms([&]() { ref_match_kernel<<<(n_q + 127) / 128, 128>>>(fmi, d_q, QL, n_q, d_r); }, 5);
        CK(cudaGetLastError());
        std::vector<uint2> hr(n_q); CK(cudaMemcpy(hr.data(), d_r, sizeof(uint2) * n_q, cudaMemcpyDeviceToHost));
        uint32 found = 0; for (uint32 i = 0; i < n_q; ++i) found += hr[i].x <= hr[i].y;
        printf("{\"what\": \"reference CUDA match(fmi, q, 20), one thread per query (fmindex_test.cu:56-85)\", \"queries\": %u, \"genome\": %u, \"ms\": %.3f, \"lookups_per_s\": %.4g, \"found\": %u}\n",
               n_q, G, ms, double(n_q) / (ms * 1e-3), found);
    }
    return 0;
}
