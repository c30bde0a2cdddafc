// gather_gran.cu -- what does one random 32-byte block read cost in DRAM traffic on B200? Random independent block reads over
// a 1.55 GB table (the 3.1 Gbp index) with different load instructions and cudaLimitMaxL2FetchGranularity settings; the
// blocks/s reached x 32 / 64 / 128 B against the 6.55 TB/s copy peak tells the fetch granularity.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_gran tools/gather_gran.cu ; tools/gather_gran [limit]
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE>
__device__ __forceinline__ uint32_t load_blk(const uint4* base, uint64_t b)
{
    const void* p = (const void*)(base + 2ull * b);
    uint32_t w[8];
    if (MODE == 0)
        asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
    else if (MODE == 1)
    {
        asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
        asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4+16];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
    }
    else if (MODE == 2)
    {
        const uint4 a = __ldg((const uint4*)p), c = __ldg((const uint4*)p + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = c.x; w[5] = c.y; w[6] = c.z; w[7] = c.w;
    }
    else if (MODE == 3)
        asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
    else if (MODE == 4)
    {
        asm volatile("ld.global.cg.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
        asm volatile("ld.global.cg.v4.b32 {%0,%1,%2,%3}, [%4+16];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
    }
    else if (MODE == 5)
    {
        asm volatile("ld.global.cv.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
        asm volatile("ld.global.cv.v4.b32 {%0,%1,%2,%3}, [%4+16];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
    }
    else if (MODE == 6)
    {
        // only 16 bytes of the block (half a sector): does DRAM traffic follow the request size?
        asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
        w[4] = w[5] = w[6] = w[7] = 0;
    }
    else
    {
        uint64_t pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.b32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p), "l"(pol));
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.b32 {%0,%1,%2,%3}, [%4+16], %5;" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p), "l"(pol));
    }
    return w[0] ^ w[3] ^ w[4] ^ w[7];
}

template <int MODE>
__global__ void __launch_bounds__(256) gather(const uint4* base, uint32_t n_blocks, uint32_t steps, uint32_t* out)
{
    uint32_t s = mix(blockIdx.x * 256u + threadIdx.x + 1u), acc = 0;
    for (uint32_t t = 0; t < steps; ++t)
    {
        acc += load_blk<MODE>(base, (uint64_t)(((uint64_t)s * n_blocks) >> 32));
        s = mix(s + 0x9e3779b9u);
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int MODE>
static void run(const char* name, const uint4* d, uint32_t n_blocks, uint32_t* out, size_t limit)
{
    const uint32_t threads = 148 * 2048 * 4, steps = 32;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        cudaEventRecord(e0);
        gather<MODE><<<threads / 256, 256>>>(d, n_blocks, steps, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double loads = (double)threads * steps;
    printf("{\"limit\": %zu, \"mode\": \"%s\", \"blocks_per_s\": %.4g, \"ms\": %.3f, \"tb_s_if_32B\": %.2f, \"tb_s_if_64B\": %.2f, \"tb_s_if_128B\": %.2f}\n",
           limit, name, loads / (best * 1e-3), best, loads * 32 / (best * 1e-3) / 1e12, loads * 64 / (best * 1e-3) / 1e12, loads * 128 / (best * 1e-3) / 1e12);
}

int main(int argc, char** argv)
{
    if (argc > 1)
    {
        const size_t want = (size_t)atoi(argv[1]);
        cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, want);
        if (e != cudaSuccess) printf("{\"set_limit_error\": \"%s\"}\n", cudaGetErrorString(e));
    }
    size_t limit = 0; cudaDeviceGetLimit(&limit, cudaLimitMaxL2FetchGranularity);
    const size_t bytes = 1550ull * 1000 * 1000;
    uint32_t* out; cudaMalloc(&out, 64);
    uint4* d; if (cudaMalloc(&d, bytes) != cudaSuccess) return 1;
    cudaMemset(d, 1, bytes);
    const uint32_t n_blocks = (uint32_t)(bytes / 32);
    run<0>("nc.v8 (256-bit)", d, n_blocks, out, limit);
    run<1>("nc.v4 x2", d, n_blocks, out, limit);
    run<2>("__ldg uint4 x2 (L1 allocating)", d, n_blocks, out, limit);
    run<3>("v8 without .nc", d, n_blocks, out, limit);
    run<4>("cg.v4 x2", d, n_blocks, out, limit);
    run<5>("cv.v4 x2", d, n_blocks, out, limit);
    run<6>("nc.v4 x1 (16 B only)", d, n_blocks, out, limit);
    run<7>("nc.v4 x2 L2 evict_first policy", d, n_blocks, out, limit);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
