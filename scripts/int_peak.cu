// MEASUREMENT AID (bench side, not part of the product library or its C ABI).
// Register-resident integer-pipe microbenchmarks that define the INT roofline bench.py and DESIGN.md
// quote.  Built by __graft_entry__.build() into scripts/_C/libgsb_intpeak.so and loaded by bench.py /
// scripts/int_peak.py only.
//
//   imad       : CH independent chains of 32-bit IMAD        (mad.lo.u32)   -> multiply-adds / s
//   imad_wide  : CH independent chains of IMAD.WIDE.U32      (mad.wide.u32) -> 32x32+64 multiply-adds / s
//   imad_hi    : CH independent chains of IMAD.HI.U32        (mad.hi.u32)   -> multiply-adds / s
//   mul_wide   : CH independent chains of IMAD.WIDE.U32 with no addend (mul.wide.u32; ptxas splits the
//                64-bit addend of mad.wide.u32 off into an IADD3 pair on the ALU pipe, so this is the
//                multiplier pipe alone)                                     -> multiplies / s
//   butterfly  : the forward lazy Harvey butterfly of csrc/ntt_core.cuh (radix-8 round on 8 register
//                words, twiddles in registers, no memory traffic), literal and wide-range forms
//                                                                           -> butterflies / s
// Each is swept over chains per thread and CTAs per SM by the caller; the roofline uses the maximum.
// Every timing is one warm-up launch followed by `reps` timed launches bracketed by CUDA events.
#include "../gemini_seal_b200/csrc/ntt_core.cuh"
#include <cstdio>
#include <vector>

namespace
{
    using namespace gsb;

    // KIND 0 = IMAD, 1 = IMAD.WIDE.U32 with 64-bit addend, 2 = IMAD.HI.U32, 3 = IMAD.WIDE.U32 without addend
    template <int KIND, int CH>
    __global__ void __launch_bounds__(256) chain_kernel(uint32_t *out, int iters, uint32_t m)
    {
        uint32_t a[CH];
        uint64_t w[CH];
#pragma unroll
        for (int i = 0; i < CH; i++)
        {
            a[i] = threadIdx.x * 2654435761u + i;
            w[i] = threadIdx.x * 3 + i;
        }
        for (int it = 0; it < iters; it++)
        {
#pragma unroll
            for (int i = 0; i < CH; i++)
            {
                if (KIND == 1)
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i])), "r"(m));
                else if (KIND == 3)
                    asm volatile("mul.wide.u32 %0, %1, %2;" : "+l"(w[i]) : "r"(static_cast<uint32_t>(w[i] >> 32)), "r"(m));
                else if (KIND == 2)
                    asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(static_cast<uint32_t>(it)));
                else
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(static_cast<uint32_t>(it)));
            }
        }
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < CH; i++)
            acc += (KIND == 1 || KIND == 3) ? static_cast<uint32_t>(w[i] >> 7) : a[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    }

    struct TwRegs
    {
        Twiddle w[7];
        GSB_DEV Twiddle fwd(int g, uint32_t, int grp) const { return w[(g == 2 ? 0 : g == 1 ? 1 : 3) + grp]; }
    };

    // FAST = the wide-range butterfly of the library-internal transforms (fwd_round).  Values are not
    // kept in range here: timing does not depend on them and nothing reads the result.
    template <bool FAST>
    __global__ void __launch_bounds__(256) butterfly_kernel(uint64_t *out, LimbC c, Twiddle t0, int iters)
    {
        TwRegs tw;
#pragma unroll
        for (int i = 0; i < 7; i++)
            tw.w[i] = make_ulonglong2(t0.x + i + (threadIdx.x & 7), t0.y + 3 * i + (threadIdx.x & 7));
        uint64_t x[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
            x[i] = threadIdx.x * 8 + i;
        for (int it = 0; it < iters; it++)
            fwd_round<2, 3, 8, false, (FAST ? 7 : 0)>(x, threadIdx.x, tw, c); // 12 butterflies; 7 = the tile kernels' form (GSB_TILE_FWD_FORM)
        uint64_t acc = 0;
#pragma unroll
        for (int i = 0; i < 8; i++)
            acc ^= x[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    }

    template <class L>
    double time_ms(L &&launch, int reps)
    {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        launch();
        cudaEventRecord(e0, nullptr);
        for (int r = 0; r < reps; r++)
            launch();
        cudaEventRecord(e1, nullptr);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        return static_cast<double>(ms) / reps;
    }

    template <int KIND, int CH>
    double chain_rate(int grid, uint32_t *buf, int iters, int reps)
    {
        const double ms = time_ms([&] { chain_kernel<KIND, CH><<<grid, 256>>>(buf, iters, 2654435761u); }, reps);
        return static_cast<double>(grid) * 256 * iters * CH / (ms * 1e-3);
    }

    template <int KIND>
    double chain_rate_ch(int ch, int grid, uint32_t *buf, int iters, int reps)
    {
        switch (ch)
        {
        case 4: return chain_rate<KIND, 4>(grid, buf, iters, reps);
        case 8: return chain_rate<KIND, 8>(grid, buf, iters, reps);
        case 12: return chain_rate<KIND, 12>(grid, buf, iters, reps);
        case 16: return chain_rate<KIND, 16>(grid, buf, iters, reps);
        default: return -1.0;
        }
    }
} // namespace

extern "C"
{
    // kind: 0 IMAD, 1 IMAD.WIDE.U32 + 64-bit addend, 2 IMAD.HI.U32, 3 IMAD.WIDE.U32 alone; chains in {4, 8, 12, 16}; returns multiply-adds / s
    // (negative on a bad argument or a CUDA error).
    double gsb_bench_chain_rate(int device, int kind, int chains, int ctas_per_sm)
    {
        if (cudaSetDevice(device) != cudaSuccess || ctas_per_sm < 1 || ctas_per_sm > 8)
            return -1.0;
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const int grid = sms * ctas_per_sm, iters = 4096, reps = 5;
        uint32_t *buf = nullptr;
        if (cudaMalloc(&buf, static_cast<size_t>(grid) * 256 * sizeof(uint32_t)) != cudaSuccess)
            return -1.0;
        double r = kind == 0   ? chain_rate_ch<0>(chains, grid, buf, iters, reps)
                   : kind == 1 ? chain_rate_ch<1>(chains, grid, buf, iters, reps)
                   : kind == 2 ? chain_rate_ch<2>(chains, grid, buf, iters, reps)
                   : kind == 3 ? chain_rate_ch<3>(chains, grid, buf, iters, reps)
                               : -1.0;
        if (cudaDeviceSynchronize() != cudaSuccess)
            r = -1.0;
        cudaFree(buf);
        return r;
    }

    // wide: 0 = literal butterfly (ForwardLazy, 6 IMAD.WIDE + 4 IMAD), 1 = wide-range (5 + 4);
    // q = a modulus of <= 56 bits; returns butterflies / s.
    double gsb_bench_butterfly_rate(int device, int wide, int ctas_per_sm, uint64_t q)
    {
        if (cudaSetDevice(device) != cudaSuccess || ctas_per_sm < 1 || ctas_per_sm > 8 || q < 3)
            return -1.0;
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const int grid = sms * ctas_per_sm, iters = 4096, reps = 5;
        uint64_t *buf = nullptr;
        if (cudaMalloc(&buf, static_cast<size_t>(grid) * 256 * sizeof(uint64_t)) != cudaSuccess)
            return -1.0;
        LimbC c{};
        c.q = q;
        c.two_q = q << 1;
        c.rdp = static_cast<uint64_t>((static_cast<unsigned __int128>(1) << 64) / q);
        c.pad = 0;
        const Twiddle t0 = make_ulonglong2(q / 3, static_cast<uint64_t>((static_cast<unsigned __int128>(q / 3) << 64) / q));
        double ms;
        if (wide)
            ms = time_ms([&] { butterfly_kernel<true><<<grid, 256>>>(buf, c, t0, iters); }, reps);
        else
            ms = time_ms([&] { butterfly_kernel<false><<<grid, 256>>>(buf, c, t0, iters); }, reps);
        double r = static_cast<double>(grid) * 256 * iters * 12 / (ms * 1e-3);
        if (cudaDeviceSynchronize() != cudaSuccess)
            r = -1.0;
        cudaFree(buf);
        return r;
    }
}
