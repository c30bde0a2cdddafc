// pipe_mix.cu -- second microbenchmark: which of the instructions a packed DP cell can be built from share an
// issue pipe on sm_100a, measured with NON-idempotent dependency rings so that neither nvcc nor ptxas can fold
// repeated applications (the first tool's max/min chains were idempotent). Every mode runs NACC independent
// rings per thread; ops_per_iter counts every instruction of the body.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_mix tools/pipe_mix.cu
//   cuobjdump -sass tools/pipe_mix | grep -c VIADDMNMX   (check the mnemonics before trusting a line)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
#define NACC 8

__device__ __forceinline__ uint32_t op_dpx(uint32_t a, uint32_t b, uint32_t c) { return __viaddmax_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t op_max3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t op_max2(uint32_t a, uint32_t b) { return __vmaxs2(a, b); }
__device__ __forceinline__ uint32_t op_add2(uint32_t a, uint32_t b) { return __vadd2(a, b); }
__device__ __forceinline__ uint32_t op_hmax(uint32_t a, uint32_t b) { uint32_t d; asm volatile("max.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t op_hmax3(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; d = op_hmax(op_hmax(a, b), c); return d; }
__device__ __forceinline__ uint32_t op_hadd(uint32_t a, uint32_t b) { uint32_t d; asm volatile("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t op_hfma(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t op_hset(uint32_t a, uint32_t b) { uint32_t d; asm volatile("set.eq.f16x2.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t op_imad(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t op_iadd(uint32_t a, uint32_t b) { uint32_t d; asm volatile("add.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t op_imax(uint32_t a, uint32_t b) { uint32_t d; asm volatile("max.s32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t op_prmt(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t op_lop(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t op_ffma(uint32_t a, uint32_t b, uint32_t c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b)), "f"(__uint_as_float(c))); return __float_as_uint(d); }
__device__ __forceinline__ uint32_t op_fmax(uint32_t a, uint32_t b) { float d; asm volatile("max.f32 %0, %1, %2;" : "=f"(d) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); return __float_as_uint(d); }
__device__ __forceinline__ uint32_t op_dpx32(uint32_t a, uint32_t b, uint32_t c) { return (uint32_t)__viaddmax_s32((int)a, (int)b, (int)c); }
__device__ __forceinline__ uint32_t op_max3_32(uint32_t a, uint32_t b, uint32_t c) { return (uint32_t)__vimax3_s32((int)a, (int)b, (int)c); }

template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t a0, uint32_t b0, uint32_t one)
{
    uint32_t a[NACC], x[NACC];
    #pragma unroll
    for (int i = 0; i < NACC; ++i) { a[i] = a0 + threadIdx.x * (i + 1); x[i] = b0 ^ (i * 0x01010101u) ^ threadIdx.x; }
    for (int it = 0; it < ITERS; ++it)
    {
        #pragma unroll
        for (int i = 0; i < NACC; ++i)
        {
            const int n = (i + 1) % NACC;
            // ---- single instructions (ring over a[])
            if (MODE == 0) a[i] = op_dpx(a[i], b0, a[n]);
            if (MODE == 1) a[i] = op_max3(a[i], b0, a[n]);
            if (MODE == 2) a[i] = op_max2(a[i], a[n]);
            if (MODE == 3) a[i] = op_add2(a[i], a[n]);
            if (MODE == 4) a[i] = op_hmax(a[i], a[n]);
            if (MODE == 5) a[i] = op_hadd(a[i], a[n]);
            if (MODE == 6) a[i] = op_hfma(a[i], b0, a[n]);
            if (MODE == 7) a[i] = op_hset(a[i], a[n]);
            if (MODE == 8) a[i] = op_imad(a[i], one, a[n]);
            if (MODE == 9) a[i] = op_iadd(a[i], a[n]);
            if (MODE == 10) a[i] = op_imax(a[i], a[n]);
            if (MODE == 11) a[i] = op_prmt(a[i], a[n], b0);
            if (MODE == 12) a[i] = op_lop(a[i], a[n], b0);
            if (MODE == 13) a[i] = op_ffma(a[i], b0, a[n]);
            if (MODE == 14) a[i] = op_fmax(a[i], a[n]);
            if (MODE == 15) a[i] = op_dpx32(a[i], b0, a[n]);
            if (MODE == 16) a[i] = op_max3_32(a[i], b0, a[n]);
            if (MODE == 17) a[i] = op_hmax3(a[i], b0, a[n]);
            // ---- pairs: one DPX ring (a) + one other ring (x)
            if (MODE >= 20 && MODE < 40)
            {
                a[i] = op_dpx(a[i], b0, a[n]);
                if (MODE == 20) x[i] = op_hmax(x[i], x[n]);
                if (MODE == 21) x[i] = op_hadd(x[i], x[n]);
                if (MODE == 22) x[i] = op_hfma(x[i], b0, x[n]);
                if (MODE == 23) x[i] = op_hset(x[i], x[n]);
                if (MODE == 24) x[i] = op_imad(x[i], one, x[n]);
                if (MODE == 25) x[i] = op_iadd(x[i], x[n]);
                if (MODE == 26) x[i] = op_imax(x[i], x[n]);
                if (MODE == 27) x[i] = op_prmt(x[i], x[n], b0);
                if (MODE == 28) x[i] = op_lop(x[i], x[n], b0);
                if (MODE == 29) x[i] = op_ffma(x[i], b0, x[n]);
                if (MODE == 30) x[i] = op_fmax(x[i], x[n]);
                if (MODE == 31) x[i] = op_max2(x[i], x[n]);
                if (MODE == 32) x[i] = op_add2(x[i], x[n]);
                if (MODE == 33) x[i] = op_dpx32(x[i], b0, x[n]);
            }
            // ---- pairs not involving DPX
            if (MODE == 40) { a[i] = op_hmax(a[i], a[n]); x[i] = op_hadd(x[i], x[n]); }
            if (MODE == 41) { a[i] = op_hmax(a[i], a[n]); x[i] = op_imad(x[i], one, x[n]); }
            if (MODE == 42) { a[i] = op_hmax(a[i], a[n]); x[i] = op_prmt(x[i], x[n], b0); }
            if (MODE == 43) { a[i] = op_hadd(a[i], a[n]); x[i] = op_imad(x[i], one, x[n]); }
            if (MODE == 44) { a[i] = op_iadd(a[i], a[n]); x[i] = op_imax(x[i], x[n]); }
            if (MODE == 45) { a[i] = op_iadd(a[i], a[n]); x[i] = op_prmt(x[i], x[n], b0); }
            if (MODE == 46) { a[i] = op_hfma(a[i], b0, a[n]); x[i] = op_imad(x[i], one, x[n]); }
            if (MODE == 47) { a[i] = op_hset(a[i], a[n]); x[i] = op_imad(x[i], one, x[n]); }
            if (MODE == 48) { a[i] = op_hset(a[i], a[n]); x[i] = op_hmax(x[i], x[n]); }
            if (MODE == 49) { a[i] = op_ffma(a[i], b0, a[n]); x[i] = op_imad(x[i], one, x[n]); }
            // ---- candidate packed-cell bodies (ops_per_iter = instruction count given to run<>)
            if (MODE == 60)      // today's SW cell: PRMT + IMAD + 2 DPX
            {
                const uint32_t s = op_prmt(x[i], b0, a[n]);
                const uint32_t u = op_imad(a[i], one, b0);
                const uint32_t v = op_dpx(x[i], s, u);
                a[i] = op_dpx(a[n], b0, v);
            }
            if (MODE == 61)      // fp16 SW cell: HSET2 + HFMA2 + 2 HADD2 + 2 HMNMX2
            {
                const uint32_t e = op_hset(x[i], a[n]);
                const uint32_t d = op_hfma(e, b0, x[n]);
                const uint32_t u = op_hadd(a[i], b0);
                const uint32_t l = op_hadd(a[n], one);
                a[i] = op_hmax(op_hmax(u, l), d);
            }
            if (MODE == 62)      // mixed: PRMT + HADD2(top+G) + DPX(diag+S, u) + HADD2(left+I)... (int/half mix is not value-compatible; rate only)
            {
                const uint32_t s = op_prmt(x[i], b0, a[n]);
                const uint32_t u = op_hadd(a[i], b0);
                const uint32_t v = op_dpx(x[i], s, u);
                const uint32_t l = op_hadd(a[n], one);
                a[i] = op_hmax(l, v);
            }
            if (MODE == 63)      // 32-bit cell: IADD + IADD + VIADDMNMX.S32 + VIMNMX.S32 (one pair per register)
            {
                const uint32_t u = op_iadd(a[i], b0);
                const uint32_t v = op_dpx32(x[i], a[n], u);
                const uint32_t l = op_iadd(a[n], one);
                a[i] = op_imax(l, v);
            }
            if (MODE == 64)      // PRMT + IMAD + DPX + IMAD + HMNMX2
            {
                const uint32_t s = op_prmt(x[i], b0, a[n]);
                const uint32_t u = op_imad(a[i], one, b0);
                const uint32_t v = op_dpx(x[i], s, u);
                const uint32_t l = op_imad(a[n], one, b0);
                a[i] = op_hmax(l, v);
            }
        }
    }
    uint32_t r = 0;
    #pragma unroll
    for (int i = 0; i < NACC; ++i) r ^= a[i] ^ x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
static void run(const char* name, double instr_per_iter, int sms)
{
    uint32_t* out; cudaMalloc(&out, sizeof(uint32_t) * sms * 8 * 256);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<sms * 8, 256>>>(out, 3, 0x00010001u, 1);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep)
    {
        cudaEventRecord(e0); k<MODE><<<sms * 8, 256>>>(out, 3, 0x00010001u, 1); cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double bodies = (double)sms * 8 * 256 * ITERS * NACC;
    const double rate = bodies / (best * 1e-3);
    printf("{\"mix\": \"%s\", \"instr_per_body\": %.0f, \"ms\": %.4f, \"bodies_per_s_per_sm\": %.4e, \"lane_instr_per_clk_per_sm_at_1965\": %.1f}\n",
           name, instr_per_iter, best, rate / sms, rate / sms * instr_per_iter / 1.965e9);
    cudaFree(out);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("{\"device\": \"%s\", \"sms\": %d}\n", p.name, sms);
    run<0>("VIADDMNMX.S16x2", 1, sms);   run<1>("VIMNMX3.S16x2", 1, sms);   run<2>("VIMNMX.S16x2", 1, sms);
    run<3>("VIADD.16x2", 1, sms);        run<4>("HMNMX2", 1, sms);          run<5>("HADD2", 1, sms);
    run<6>("HFMA2", 1, sms);             run<7>("HSET2", 1, sms);           run<8>("IMAD", 1, sms);
    run<9>("IADD3", 1, sms);             run<10>("VIMNMX.S32", 1, sms);     run<11>("PRMT", 1, sms);
    run<12>("LOP3", 1, sms);             run<13>("FFMA", 1, sms);           run<14>("FMNMX", 1, sms);
    run<15>("VIADDMNMX.S32", 1, sms);    run<16>("VIMNMX3.S32", 1, sms);    run<17>("HMNMX2.3in", 1, sms);
    run<20>("DPX+HMNMX2", 2, sms);       run<21>("DPX+HADD2", 2, sms);      run<22>("DPX+HFMA2", 2, sms);
    run<23>("DPX+HSET2", 2, sms);        run<24>("DPX+IMAD", 2, sms);       run<25>("DPX+IADD3", 2, sms);
    run<26>("DPX+VIMNMX.S32", 2, sms);   run<27>("DPX+PRMT", 2, sms);       run<28>("DPX+LOP3", 2, sms);
    run<29>("DPX+FFMA", 2, sms);         run<30>("DPX+FMNMX", 2, sms);      run<31>("DPX+VIMNMX.S16x2", 2, sms);
    run<32>("DPX+VIADD.16x2", 2, sms);   run<33>("DPX+VIADDMNMX.S32", 2, sms);
    run<40>("HMNMX2+HADD2", 2, sms);     run<41>("HMNMX2+IMAD", 2, sms);    run<42>("HMNMX2+PRMT", 2, sms);
    run<43>("HADD2+IMAD", 2, sms);       run<44>("IADD3+VIMNMX.S32", 2, sms); run<45>("IADD3+PRMT", 2, sms);
    run<46>("HFMA2+IMAD", 2, sms);       run<47>("HSET2+IMAD", 2, sms);     run<48>("HSET2+HMNMX2", 2, sms);
    run<49>("FFMA+IMAD", 2, sms);
    run<60>("cell PRMT+IMAD+2DPX", 4, sms);
    run<61>("cell fp16 HSET2+HFMA2+2HADD2+2HMNMX2", 6, sms);
    run<62>("cell PRMT+2HADD2+DPX+HMNMX2", 5, sms);
    run<63>("cell s32 2IADD+VIADDMNMX.S32+VIMNMX.S32", 4, sms);
    run<64>("cell PRMT+2IMAD+DPX+HMNMX2", 5, sms);
    return 0;
}
