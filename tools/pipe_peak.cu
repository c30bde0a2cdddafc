// pipe_peak.cu -- measures the per-SM issue rate of the integer instructions the DP kernels are built
// from (DPX VIADDMNMX.S16x2 / VIMNMX3.S16x2, PRMT, VIADD.16x2, IMAD, LOP3) on all SMs, alone and mixed.
// The measured DPX rate is the denominator of the DP roofline: peak GCUPS =
// DPX lane-ops/s x 2 cells per s16x2 op / DPX ops per packed cell.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_peak tools/pipe_peak.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

#define ITERS 4096
#define NACC 8

template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t a0, uint32_t b0, uint32_t one)
{
    uint32_t acc[NACC], x[NACC];
    #pragma unroll
    for (int i = 0; i < NACC; ++i) { acc[i] = a0 + threadIdx.x * (i + 1); x[i] = b0 ^ (i * 0x01010101u); }
    for (int it = 0; it < ITERS; ++it)
    {
        #pragma unroll
        for (int i = 0; i < NACC; ++i)
        {
            if (MODE == 0) acc[i] = __viaddmax_s16x2(acc[i], b0, x[i]);                 // VIADDMNMX.S16x2
            if (MODE == 1) acc[i] = __viaddmax_s16x2_relu(acc[i], b0, x[i]);            // VIADDMNMX.S16x2.RELU
            if (MODE == 2) acc[i] = __vimax3_s16x2(acc[i], b0, x[i]);                   // VIMNMX3.S16x2
            if (MODE == 3) acc[i] = __byte_perm(acc[i], x[i], b0);                      // PRMT
            if (MODE == 4) acc[i] = __vadd2(acc[i], x[i]);                              // VIADD.16x2
            if (MODE == 5) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(acc[i]) : "r"(one), "r"(x[i]));   // IMAD
            if (MODE == 6) acc[i] = (acc[i] & x[i]) ^ b0;                               // LOP3
            if (MODE == 7)                                                              // DP cell mix: PRMT + IMAD + 2 DPX
            {
                const uint32_t s = __byte_perm(x[i], b0, acc[i]);
                uint32_t u; asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(u) : "r"(acc[i]), "r"(one), "r"(b0));
                const uint32_t v = __viaddmax_s16x2(x[i], s, u);
                acc[i] = __viaddmax_s16x2(acc[i], b0, v);
            }
            if (MODE == 8)                                                              // 2 DPX + PRMT + VIADD (all ALU)
            {
                const uint32_t s = __byte_perm(x[i], b0, acc[i]);
                const uint32_t u = __vadd2(acc[i], b0);
                const uint32_t v = __viaddmax_s16x2(x[i], s, u);
                acc[i] = __viaddmax_s16x2_relu(acc[i], b0, v);
            }
            if (MODE == 9) acc[i] = __viaddmax_s32(acc[i], b0, x[i]);                   // VIADDMNMX (s32)
            if (MODE == 10) acc[i] = __popc(acc[i] ^ x[i]) + b0;                        // POPC (+IADD)
            if (MODE == 11) asm volatile("add.u32 %0, %0, %1;" : "+r"(acc[i]) : "r"(x[i]));                               // IADD3
            if (MODE == 12) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(acc[i]) : "r"(x[i]), "r"(b0));            // LOP3
            if (MODE == 13) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(acc[i]) : "r"(x[i]), "r"(b0));            // SHF
            if (MODE == 14) { float f = __uint_as_float(acc[i]); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__uint_as_float(x[i])), "f"(__uint_as_float(b0))); acc[i] = __float_as_uint(f); }   // FFMA
            if (MODE == 15) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(acc[i]) : "r"(x[i]), "r"(b0));              // HFMA2
            if (MODE == 16) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(acc[i]) : "r"(x[i]));                              // HMNMX2
            if (MODE == 17) asm volatile("add.f16x2 %0, %0, %1;" : "+r"(acc[i]) : "r"(x[i]));                              // HADD2
            if (MODE == 18) asm volatile("max.s32 %0, %0, %1;" : "+r"(acc[i]) : "r"(x[i]));                                // VIMNMX.S32
            if (MODE == 19) asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %1, %2, p;}" : "+r"(acc[i]) : "r"(x[i]), "r"(b0));   // ISETP+SEL
            // pair mixes: one DPX + one other op per iteration
            if (MODE >= 20 && MODE <= 27)
            {
                acc[i] = __viaddmax_s16x2(acc[i], b0, x[i]);
                if (MODE == 20) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(one));
                if (MODE == 21) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(one), "r"(b0));
                if (MODE == 22) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(one), "r"(b0));
                if (MODE == 23) { float f = __uint_as_float(x[i]); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__uint_as_float(one)), "f"(__uint_as_float(b0))); x[i] = __float_as_uint(f); }
                if (MODE == 24) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(x[i]) : "r"(one));
                if (MODE == 25) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(one), "r"(b0));
                if (MODE == 26) x[i] = __byte_perm(x[i], one, b0);
                if (MODE == 27) asm volatile("add.f16x2 %0, %0, %1;" : "+r"(x[i]) : "r"(one));
            }
            if (MODE == 28) { asm volatile("add.u32 %0, %0, %1;" : "+r"(acc[i]) : "r"(one)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(one), "r"(b0)); }   // IADD3 + LOP3
            if (MODE == 29) { asm volatile("max.f16x2 %0, %0, %1;" : "+r"(acc[i]) : "r"(one)); asm volatile("add.f16x2 %0, %0, %1;" : "+r"(x[i]) : "r"(one)); }              // HMNMX2 + HADD2
            if (MODE == 30) { asm volatile("max.f16x2 %0, %0, %1;" : "+r"(acc[i]) : "r"(one)); x[i] = __byte_perm(x[i], one, b0); }                                          // HMNMX2 + PRMT
        }
    }
    uint32_t r = 0;
    #pragma unroll
    for (int i = 0; i < NACC; ++i) r ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
static void run(const char* name, double ops_per_iter, int sms, double* dpx_rate)
{
    uint32_t* out; cudaMalloc(&out, sizeof(uint32_t) * sms * 8 * 256);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<sms * 8, 256>>>(out, 3, 0x00010001u, 1);   // warm-up
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep)
    {
        cudaEventRecord(e0); k<MODE><<<sms * 8, 256>>>(out, 3, 0x00010001u, 1); cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double lane_ops = (double)sms * 8 * 256 * ITERS * NACC * ops_per_iter;
    const double rate = lane_ops / (best * 1e-3);
    printf("{\"pipe\": \"%s\", \"ms\": %.4f, \"lane_ops_per_s\": %.4e, \"lane_ops_per_s_per_sm\": %.4e}\n", name, best, rate, rate / sms);
    if (dpx_rate) *dpx_rate = rate;
    cudaFree(out);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d}\n", p.name, p.multiProcessorCount, clk);
    const int sms = p.multiProcessorCount;
    double dpx = 0;
    run<0>("VIADDMNMX.S16x2", 1, sms, &dpx);
    run<1>("VIADDMNMX.S16x2.RELU", 1, sms, nullptr);
    run<2>("VIMNMX3.S16x2", 1, sms, nullptr);
    run<9>("VIADDMNMX.S32", 1, sms, nullptr);
    run<3>("PRMT", 1, sms, nullptr);
    run<4>("VIADD.16x2", 1, sms, nullptr);
    run<5>("IMAD", 1, sms, nullptr);
    run<6>("LOP3", 1, sms, nullptr);
    run<10>("POPC+IADD", 1, sms, nullptr);
    run<11>("IADD3", 1, sms, nullptr);
    run<12>("LOP3(asm)", 1, sms, nullptr);
    run<13>("SHF", 1, sms, nullptr);
    run<14>("FFMA", 1, sms, nullptr);
    run<15>("HFMA2", 1, sms, nullptr);
    run<16>("HMNMX2", 1, sms, nullptr);
    run<17>("HADD2", 1, sms, nullptr);
    run<18>("VIMNMX.S32", 1, sms, nullptr);
    run<19>("ISETP+SEL (pairs)", 1, sms, nullptr);
    run<20>("pair DPX+IADD3", 1, sms, nullptr);
    run<21>("pair DPX+LOP3", 1, sms, nullptr);
    run<22>("pair DPX+SHF", 1, sms, nullptr);
    run<23>("pair DPX+FFMA", 1, sms, nullptr);
    run<24>("pair DPX+HMNMX2", 1, sms, nullptr);
    run<25>("pair DPX+IMAD", 1, sms, nullptr);
    run<26>("pair DPX+PRMT", 1, sms, nullptr);
    run<27>("pair DPX+HADD2", 1, sms, nullptr);
    run<28>("pair IADD3+LOP3", 1, sms, nullptr);
    run<29>("pair HMNMX2+HADD2", 1, sms, nullptr);
    run<30>("pair HMNMX2+PRMT", 1, sms, nullptr);
    run<7>("cell_mix(PRMT+IMAD+2DPX) per cell", 1, sms, nullptr);
    run<8>("cell_mix(PRMT+VIADD+2DPX) per cell", 1, sms, nullptr);
    printf("{\"dpx_peak_lane_ops_per_s\": %.4e, \"note\": \"GCUPS peak = this x 2 cells / DPX ops per packed cell / 1e9\"}\n", dpx);
    return 0;
}
