// Register-resident integer-pipe microbenchmarks: the INT roofline bench.py and DESIGN.md quote.
//   imad      : independent chains of 32-bit IMAD (mad.lo.u32)            -> 32-bit multiply-adds / s
//   imad_wide : independent chains of IMAD.WIDE.U32 (mad.wide.u32)        -> 32x32+64 multiply-adds / s
//   butterfly : the forward lazy Harvey butterfly of ntt_core.cuh (radix-8 rounds on 8 register words,
//               twiddles in registers, no memory traffic), in its literal form (the reference's words)
//               and in the wide-range form of the library-internal transforms -> butterflies / s
// Included at the end of gsb_api.cu.
#pragma once
#include "ntt_core.cuh"

namespace gsb
{
    template <int WIDE>
    __global__ void __launch_bounds__(256) imad_peak_kernel(uint32_t *out, int iters, uint32_t m)
    {
        uint32_t a[8];
        uint64_t w[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
        {
            a[i] = threadIdx.x + i;
            w[i] = threadIdx.x * 3 + i;
        }
        for (int it = 0; it < iters; it++)
        {
#pragma unroll
            for (int i = 0; i < 8; i++)
            {
                if (WIDE)
                    w[i] = static_cast<uint64_t>(static_cast<uint32_t>(w[i])) * m + w[i];
                else
                    a[i] = a[i] * m + static_cast<uint32_t>(it);
            }
        }
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < 8; i++)
            acc += WIDE ? static_cast<uint32_t>(w[i] >> 7) : a[i];
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
    __global__ void __launch_bounds__(256) butterfly_peak_kernel(uint64_t *out, const LimbC *limb, const Twiddle *WW, int iters)
    {
        const LimbC c = limb[0];
        TwRegs tw;
#pragma unroll
        for (int i = 0; i < 7; i++)
            tw.w[i] = WW[1 + i + (threadIdx.x & 7)];
        uint64_t x[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
            x[i] = threadIdx.x * 8 + i;
        for (int it = 0; it < iters; it++)
            fwd_round<2, 3, 8, false, FAST>(x, threadIdx.x, tw, c); // 12 butterflies
        uint64_t acc = 0;
#pragma unroll
        for (int i = 0; i < 8; i++)
            acc ^= x[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    }
} // namespace gsb

static int int_peak_impl(gsb_ctx *ctx, int ctas_per_sm, double *imad_per_s, double *imad_wide_per_s, double *butterflies_per_s, double *wide_butterflies_per_s)
{
    using namespace gsb;
    return guarded([&] {
        require(ctx != nullptr && ctas_per_sm >= 1 && ctas_per_sm <= 8, GSB_E_INVALID_ARG, "bad argument");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        int sms = 0;
        GSB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        const int grid = sms * ctas_per_sm, iters = 4096;
        void *buf = nullptr;
        GSB_CUDA_CHECK(cudaMalloc(&buf, static_cast<size_t>(grid) * 256 * sizeof(uint64_t)));
        cudaEvent_t e0, e1;
        GSB_CUDA_CHECK(cudaEventCreate(&e0));
        GSB_CUDA_CHECK(cudaEventCreate(&e1));
        auto time_ms = [&](auto &&launch) {
            launch();
            GSB_CUDA_CHECK(cudaEventRecord(e0, nullptr));
            launch();
            GSB_CUDA_CHECK(cudaEventRecord(e1, nullptr));
            GSB_CUDA_CHECK(cudaEventSynchronize(e1));
            float ms = 0.f;
            GSB_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            return static_cast<double>(ms);
        };
        const double threads = static_cast<double>(grid) * 256;
        if (imad_per_s)
        {
            const double ms = time_ms([&] { GSB_LAUNCH("int_peak", "imad", imad_peak_kernel<0>, dim3(grid), dim3(256), 0, nullptr, static_cast<uint32_t *>(buf), iters, 2654435761u); });
            *imad_per_s = threads * iters * 8 / (ms * 1e-3);
        }
        if (imad_wide_per_s)
        {
            const double ms = time_ms([&] { GSB_LAUNCH("int_peak", "imad_wide", imad_peak_kernel<1>, dim3(grid), dim3(256), 0, nullptr, static_cast<uint32_t *>(buf), iters, 2654435761u); });
            *imad_wide_per_s = threads * iters * 8 / (ms * 1e-3);
        }
        if (butterflies_per_s)
        {
            const double ms = time_ms([&] { GSB_LAUNCH("int_peak", "butterfly", butterfly_peak_kernel<false>, dim3(grid), dim3(256), 0, nullptr, static_cast<uint64_t *>(buf), ctx->d_limb, ctx->d_WW, iters); });
            *butterflies_per_s = threads * iters * 12 / (ms * 1e-3);
        }
        if (wide_butterflies_per_s)
        {
            const double ms = time_ms([&] { GSB_LAUNCH("int_peak", "butterfly_wide", butterfly_peak_kernel<true>, dim3(grid), dim3(256), 0, nullptr, static_cast<uint64_t *>(buf), ctx->d_limb, ctx->d_WW, iters); });
            *wide_butterflies_per_s = threads * iters * 12 / (ms * 1e-3);
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(buf);
    });
}

extern "C" int gsb_int_peak(gsb_ctx *ctx, int ctas_per_sm, double *imad_per_s, double *imad_wide_per_s, double *butterflies_per_s)
{
    return int_peak_impl(ctx, ctas_per_sm, imad_per_s, imad_wide_per_s, butterflies_per_s, nullptr);
}

extern "C" int gsb_int_peak_wide(gsb_ctx *ctx, int ctas_per_sm, double *wide_butterflies_per_s)
{
    return int_peak_impl(ctx, ctas_per_sm, nullptr, nullptr, nullptr, wide_butterflies_per_s);
}
