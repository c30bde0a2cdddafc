// gather_peak.cu -- the memory system's ceiling for what the FM-index kernels do on an index larger than L2:
// random, independent 32-byte block reads (one 256-bit load each, the same instruction fm_load_block issues).
// Reported per table size: blocks/s and GB/s of useful bytes, with 1, 2 and 4 independent loads in flight per
// thread, and a DEPENDENT chain (next index derived from the loaded block) like a backward-search step.
// This is the denominator to read fm_match's DRAM-regime throughput against.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_peak tools/gather_peak.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

struct Blk { uint32_t w[8]; };

__device__ __forceinline__ Blk load_blk(const uint4* base, uint64_t b)
{
    Blk r; const void* p = (const void*)(base + 2ull * b);
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7]) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// ILP independent streams per thread, STEPS loads per stream; DEP: the next index also depends on the loaded data
template <int ILP, bool DEP>
__global__ void __launch_bounds__(256) gather(const uint4* base, uint32_t n_blocks, uint32_t steps, uint32_t* out)
{
    uint32_t s[ILP], acc = 0;
    #pragma unroll
    for (int i = 0; i < ILP; ++i) s[i] = mix((blockIdx.x * 256u + threadIdx.x) * ILP + i + 1u);
    for (uint32_t t = 0; t < steps; ++t)
    {
        Blk b[ILP];
        #pragma unroll
        for (int i = 0; i < ILP; ++i) b[i] = load_blk(base, (uint64_t)(((uint64_t)s[i] * n_blocks) >> 32));
        #pragma unroll
        for (int i = 0; i < ILP; ++i)
        {
            const uint32_t v = b[i].w[0] ^ b[i].w[3] ^ b[i].w[4] ^ b[i].w[7];
            acc += v;
            s[i] = mix(s[i] + (DEP ? v : 0u) + 0x9e3779b9u);
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int ILP, bool DEP>
static void run(const uint4* d, uint32_t n_blocks, uint32_t* out, size_t bytes)
{
    const uint32_t threads = 148 * 2048 * 4, steps = 64 / ILP;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        cudaEventRecord(e0);
        gather<ILP, DEP><<<threads / 256, 256>>>(d, n_blocks, steps, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double loads = (double)threads * steps * ILP;
    printf("{\"table_mb\": %.0f, \"ilp\": %d, \"dependent\": %s, \"blocks_per_s\": %.4g, \"useful_gbs\": %.1f, \"ms\": %.3f}\n",
           bytes / 1e6, ILP, DEP ? "true" : "false", loads / (best * 1e-3), loads * 32 / (best * 1e-3) / 1e9, best);
}

int main()
{
    const size_t sizes[] = { 50ull << 20, 500ull * 1000 * 1000, 1550ull * 1000 * 1000, 8ull << 30 };
    uint32_t* out; cudaMalloc(&out, 64);
    for (size_t bytes : sizes)
    {
        uint4* d; if (cudaMalloc(&d, bytes) != cudaSuccess) { printf("{\"error\": \"alloc %zu\"}\n", bytes); return 1; }
        cudaMemset(d, 1, bytes);
        const uint32_t n_blocks = (uint32_t)(bytes / 32);
        run<1, false>(d, n_blocks, out, bytes);
        run<2, false>(d, n_blocks, out, bytes);
        run<4, false>(d, n_blocks, out, bytes);
        run<1, true>(d, n_blocks, out, bytes);
        run<2, true>(d, n_blocks, out, bytes);
        cudaFree(d);
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
