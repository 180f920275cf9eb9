// Allocation-free variant of gsb_rescale_dev (Evaluator::rescale_to_next, evaluator.cpp:1090-1126,829-892;
// util/rns.cpp:777-851).  No counterpart in the reference: added for
// callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  Included at the end of gsb_api.cu, i.e. compiled in the same translation unit as everything
// that names the array below.
//
// gsb_rescale_dev needs two pieces of device memory besides its operands: the rounded last limb `t`
// (n_polys * N words, from the caller's workspace, grown with cudaMalloc on demand) and the per-level
// constants q_last^-1 mod q_i / q_i - (q_last/2 mod q_i) (gsb_ctx::consts_for, cudaMalloc on first use of a
// level).  Here both live in ONE file-scope device array of fixed size, which every GPU that loads the
// library holds until the process exits:
//     words [0, 128)    q_last^-1 mod q_i with its Shoup scaling, [L-1][2]
//     words [128, 256)  q_i - (q_last/2 mod q_i),                 [L-1]
//     words [256, ...)  t, up to GSB_NOALLOC_MAX_WORDS words
// The constants are computed on the host (plain arithmetic) and written by a one-block kernel from its launch
// parameters, so the call issues no allocation, no free and no memcpy.  Nothing has to start at zero: every
// word the kernels read is written earlier in the same call, in stream order.  All calls share the array:
// calls that overlap on the GPU (different streams, or different host threads without a common stream) race.
#pragma once
#include "gsb_evaluator.cuh"

#ifndef GSB_CUSIM

#define GSB_NOALLOC_MAX_WORDS (size_t(1) << 19) /* 4 MiB of t: batch * size * N <= 524288 */
#define GSB_NOALLOC_HDR_WORDS 256

// In HOST code this name is a host-side stand-in, not a device address: host code must go through
// cudaGetSymbolAddress (noalloc_scratch_ptr below).  Device code may name it directly (noalloc_consts_kernel).
static __device__ __align__(256) uint64_t g_noalloc_scratch[GSB_NOALLOC_HDR_WORDS + GSB_NOALLOC_MAX_WORDS];

namespace gsb
{
    struct NoallocConsts
    {
        uint64_t inv_qlast[128]; // [L-1][2]
        uint64_t neg_half[64];   // [L-1]
        uint32_t n;              // L-1
    };

    __global__ void __launch_bounds__(128) noalloc_consts_kernel(NoallocConsts c)
    {
        const uint32_t i = threadIdx.x;
        if (i < 2 * c.n)
            g_noalloc_scratch[i] = c.inv_qlast[i];
        if (i < c.n)
            g_noalloc_scratch[128 + i] = c.neg_half[i];
    }

    // Device address of the array on `device`, resolved once per device with that device current.
    static uint64_t *noalloc_scratch_ptr(int device)
    {
        static std::mutex mu;
        static uint64_t *ptrs[64] = {};
        require(device >= 0 && device < 64, GSB_E_INVALID_ARG, "device index out of range");
        std::lock_guard<std::mutex> lock(mu);
        if (!ptrs[device])
        {
            GSB_CUDA_CHECK(cudaSetDevice(device));
            void *p = nullptr;
            GSB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            require(p != nullptr, GSB_E_CUDA, "cudaGetSymbolAddress returned a null address");
            ptrs[device] = static_cast<uint64_t *>(p);
        }
        return ptrs[device];
    }

    // rescale_core (gsb_evaluator.cuh) with `t` and the level constants at caller-given device addresses.
    template <int LOGN>
    static void rescale_noalloc_core(
        gsb_ctx *ctx, uint64_t *t, const uint64_t *d_inv_qlast, const uint64_t *d_neg_half, const uint64_t *in, uint64_t *out,
        size_t n_polys, size_t L, cudaStream_t st)
    {
        constexpr bool two = Split<LOGN>::two_pass;
        const size_t N = ctx->N;
        const NttTables tab = ctx->ntt_tables();
        const int nz = static_cast<int>(n_polys);
        const ModMap mlast{ 1 << 30, 0, static_cast<int>(L - 1) };
        // stage 1 (util/rns.cpp:798-806): t = barrett63(strict_iNTT(last limb) + half)
        if constexpr (two)
        {
            launch_inv_contig<LOGN>(IoPlain<OUT_RAW>{ in + (L - 1) * N, t, L * N, N }, tab, mlast, 1, nz, st, "rescale_last");
            IoPair<PlainLoad, RsLastStore> io{ PlainArgs{ t, N }, PlainStoreArgs{ t, N } };
            launch_inv_cols<LOGN>(io, tab, mlast, 1, nz, st, "rescale_last");
        }
        else
        {
            IoPair<PlainLoad, RsLastStore> io{ PlainArgs{ in + (L - 1) * N, L * N }, PlainStoreArgs{ t, N } };
            launch_inv_contig<LOGN>(io, tab, mlast, 1, nz, st, "rescale_last");
        }
        // stage 2 (:811-848), literal butterflies (valid for every modulus the context accepts)
        const RsLoadArgs la{ t, d_neg_half, ctx->tables[L - 1].mod.q };
        const RsStoreArgs sa{ in, out, d_inv_qlast, static_cast<uint32_t>(L) };
        const ModMap mm{ 1 << 30, 0, 0 };
        const int ny = static_cast<int>(L - 1);
        if constexpr (two)
        {
            IoPair<RsLoad, PlainStore<OUT_RAW>> io1{ la, PlainStoreArgs{ out, (L - 1) * N } };
            IoPair<PlainLoad, RsStore> io2{ PlainArgs{ out, (L - 1) * N }, sa };
            launch_fwd_cols<LOGN>(io1, tab, mm, ny, nz, st, "rescale");
            launch_fwd_contig<LOGN>(io2, tab, mm, ny, nz, st, "rescale");
        }
        else
        {
            IoPair<RsLoad, RsStore> io{ la, sa };
            launch_fwd_contig<LOGN>(io, tab, mm, ny, nz, st, "rescale");
        }
    }
} // namespace gsb

extern "C" int gsb_rescale_noalloc_dev(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream)
{
    using namespace gsb;
    return guarded([&] {
        check_batch(ctx, batch, L, 2); // evaluator.cpp:845-848: needs a next level
        require(in && out && size >= 1 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        const size_t n_polys = batch * size;
        require(n_polys * ctx->N <= GSB_NOALLOC_MAX_WORDS, GSB_E_RANGE,
                "gsb_rescale_noalloc_dev: batch * size * N exceeds the static scratch (524288 words = 4 MiB); use gsb_rescale_dev");
        require(L - 1 <= 64, GSB_E_RANGE, "gsb_rescale_noalloc_dev: more than 64 remaining limbs");
        uint64_t *base = noalloc_scratch_ptr(ctx->device);
        // rescale constants of this level (util/rns.cpp:719-728, 824), host arithmetic only
        NoallocConsts c{};
        c.n = static_cast<uint32_t>(L - 1);
        const host::ModConst &ml = ctx->tables[L - 1].mod;
        for (size_t i = 0; i + 1 < L; i++)
        {
            const host::ModConst &mi = ctx->tables[i].mod;
            uint64_t inv = 0;
            require(host::inv_mod(ml.q, mi.q, inv), GSB_E_LOGIC, "invalid rns bases");
            c.inv_qlast[2 * i] = inv;
            c.inv_qlast[2 * i + 1] = host::shoup(inv, mi.q);
            c.neg_half[i] = mi.q - host::barrett63(ml.q >> 1, mi);
        }
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        GSB_LAUNCH("noalloc_consts", "", noalloc_consts_kernel, dim3(1), dim3(128), 0, st, c);
        GSB_DISPATCH_LOGN(ctx->logn, rescale_noalloc_core<LOGN>(ctx, base + GSB_NOALLOC_HDR_WORDS, base, base + 128, in, out, n_polys, L, st));
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

#endif // GSB_CUSIM
