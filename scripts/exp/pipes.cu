// Throughput of the integer instructions the butterflies are made of, alone and mixed, in cycles per warp
// instruction per SM sub-partition (experiment; not part of the product).
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int KIND, int CH>
__global__ void __launch_bounds__(256) k(uint64_t *out, int iters, uint64_t a, uint64_t b)
{
    uint64_t t[CH], w[CH];
#pragma unroll
    for (int i = 0; i < CH; i++)
    {
        t[i] = threadIdx.x * 3 + i;
        w[i] = threadIdx.x * 5 + i;
    }
    for (int it = 0; it < iters; it++)
    {
#pragma unroll
        for (int i = 0; i < CH; i++)
        {
            if (KIND == 0) // 64-bit three-input add: IADD3 + IADD3.X
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(w[i]));
            else if (KIND == 1) // 64-bit two-input add
                asm("add.u64 %0, %0, %1;" : "+l"(t[i]) : "l"(w[i]));
            else if (KIND == 2) // IMAD.WIDE alone
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(static_cast<uint32_t>(a)));
            else if (KIND == 3) // IMAD.WIDE + independent three-input 64-bit add
            {
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(static_cast<uint32_t>(a)));
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(b));
            }
            else if (KIND == 4) // IMAD.WIDE + dependent three-input add (butterfly-like chaining)
            {
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(t[i])), "r"(static_cast<uint32_t>(a)));
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(w[i]));
            }
            else if (KIND == 5) // IMAD (narrow) + independent three-input add
            {
                uint32_t lo = static_cast<uint32_t>(w[i]);
                asm("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)), "r"(static_cast<uint32_t>(b)));
                w[i] = lo;
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(b));
            }
            else if (KIND == 6) // LOP3
            {
                uint32_t lo = static_cast<uint32_t>(t[i]);
                asm("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)), "r"(static_cast<uint32_t>(w[i])));
                t[i] = lo;
            }
            else if (KIND == 8) // IMAD.HI + independent three-input add
            {
                uint32_t lo = static_cast<uint32_t>(w[i]);
                asm("mul.hi.u32 %0, %0, %1;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)));
                w[i] = lo | 0x80000000u;
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(b));
            }
            else if (KIND == 9) // IMAD.WIDE with 64-bit addend + independent three-input add
            {
                asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i])), "r"(static_cast<uint32_t>(a)));
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(b));
            }
            else if (KIND == 10) // IMAD.WIDE with 64-bit addend alone
                asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i])), "r"(static_cast<uint32_t>(a)));
            else if (KIND == 11) // IMAD.HI alone
            {
                uint32_t lo = static_cast<uint32_t>(w[i]);
                asm("mul.hi.u32 %0, %0, %1;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)));
                w[i] = lo | 0x80000000u;
            }
            else if (KIND == 12) // IMAD.WIDE + IMAD, no adds
            {
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(static_cast<uint32_t>(a)));
                uint32_t lo = static_cast<uint32_t>(t[i]);
                asm("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)), "r"(static_cast<uint32_t>(b)));
                t[i] = lo;
            }
            else if (KIND == 13) // IMAD.WIDE + 2 IMAD (FMA 8, tax 2)
            {
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(static_cast<uint32_t>(a)));
                uint32_t lo = static_cast<uint32_t>(t[i]), hi = static_cast<uint32_t>(t[i] >> 32);
                asm("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)), "r"(static_cast<uint32_t>(b)));
                asm("mad.lo.u32 %0, %0, %1, %2;" : "+r"(hi) : "r"(static_cast<uint32_t>(b)), "r"(static_cast<uint32_t>(a)));
                t[i] = (static_cast<uint64_t>(hi) << 32) | lo;
            }
            else if (KIND == 14) // IMAD + one two-input add (IADD3) : FMA 2, ALU 2
            {
                uint32_t lo = static_cast<uint32_t>(w[i]);
                asm("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(static_cast<uint32_t>(a)), "r"(static_cast<uint32_t>(b)));
                w[i] = lo;
                uint32_t tl = static_cast<uint32_t>(t[i]);
                asm("add.u32 %0, %0, %1;" : "+r"(tl) : "r"(static_cast<uint32_t>(it)));
                t[i] = tl;
            }
            else if (KIND == 7) // 2 IMAD.WIDE + 1 three-input add (FMA 8 : ALU 4)
            {
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(static_cast<uint32_t>(a)));
                asm("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(static_cast<uint32_t>(b)));
                asm("add.u64 %0, %0, %1;\n\tadd.u64 %0, %0, %2;" : "+l"(t[i]) : "l"(a), "l"(b));
            }
        }
    }
    uint64_t acc = 0;
#pragma unroll
    for (int i = 0; i < CH; i++)
        acc ^= t[i] ^ w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int KIND, int CH>
void run(const char *what, int sms, double mhz, uint64_t *buf)
{
    const int iters = 4096;
    for (int ctas = 1; ctas <= 4; ctas *= 2)
    {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        k<KIND, CH><<<sms * ctas, 256>>>(buf, iters, 0x9e3779b97f4a7c15ull, 12345);
        cudaEventRecord(e0);
        for (int r = 0; r < 5; r++)
            k<KIND, CH><<<sms * ctas, 256>>>(buf, iters, 0x9e3779b97f4a7c15ull, 12345);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        ms /= 5;
        // warps per SMSP = ctas * 8 / 4; cycles per (chain step) per warp-slot
        const double cyc = ms * 1e-3 * mhz * 1e6;
        const double steps_per_smsp = static_cast<double>(iters) * CH * ctas * 2;
        printf("%-52s CH=%2d ctas/SM=%d  %.2f cycles per chain step per SMSP\n", what, CH, ctas, cyc / steps_per_smsp);
    }
}

int main()
{
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = 1965.0; // observed under load (nvidia-smi), attribute says khz/1000
    printf("SMs %d, attribute clock %d kHz, using %.0f MHz\n", sms, khz, mhz);
    uint64_t *buf;
    cudaMalloc(&buf, sizeof(uint64_t) * sms * 8 * 256);
    run<0, 8>("add64 three-input (IADD3 + IADD3.X)", sms, mhz, buf);
    run<1, 8>("add64 two-input", sms, mhz, buf);
    run<2, 8>("IMAD.WIDE", sms, mhz, buf);
    run<3, 8>("IMAD.WIDE + independent add64x3", sms, mhz, buf);
    run<4, 8>("IMAD.WIDE + dependent add64x3", sms, mhz, buf);
    run<5, 8>("IMAD + independent add64x3", sms, mhz, buf);
    run<6, 8>("LOP3", sms, mhz, buf);
    run<7, 8>("2 IMAD.WIDE + independent add64x3", sms, mhz, buf);
    run<8, 8>("IMAD.HI + independent add64x3", sms, mhz, buf);
    run<9, 8>("IMAD.WIDE(acc) + independent add64x3", sms, mhz, buf);
    run<10, 8>("IMAD.WIDE(acc)", sms, mhz, buf);
    run<11, 8>("IMAD.HI", sms, mhz, buf);
    run<12, 8>("IMAD.WIDE + IMAD", sms, mhz, buf);
    run<13, 8>("IMAD.WIDE + 2 IMAD", sms, mhz, buf);
    run<14, 8>("IMAD + IADD3", sms, mhz, buf);
    return cudaDeviceSynchronize() != cudaSuccess;
}
