// Build-mode shim.  The product is compiled by nvcc for sm_100a.  The same kernel sources can also
// be compiled by g++ against tests/cusim/cusim.h (GSB_CUSIM), a TEST-ONLY fiber emulation of the
// CUDA execution model used to debug kernel indexing on the GPU-less dev container.  The simulator
// is never shipped, never loaded by the package, and is not a CPU fallback: gemini_seal_b200/_lib.py
// only ever loads the nvcc-built library and raises if it is missing.
#pragma once
#include <cstddef>
#include <cstdint>

#ifdef GSB_CUSIM
#include "cusim.h"
#define GSB_LAUNCH(pass, tag, kfn, grid, block, smem, stream, ...) \
    (++::gsb::launch_counter(), ::cusim::launch((grid), (block), (smem), [&]() { kfn(__VA_ARGS__); }))
#else
#include <cuda_runtime.h>
// Every launch goes through here: counted (gsb_launch_count) and, when gsb_prof_enable(1) is on,
// bracketed by CUDA events on the launching stream (gsb_prof_report), labelled "<pass>/<tag>".
#define GSB_LAUNCH(pass, tag, kfn, grid, block, smem, stream, ...)       \
    do                                                                   \
    {                                                                    \
        ++::gsb::launch_counter();                                       \
        const bool prof__ = ::gsb::prof_on();                            \
        if (prof__)                                                      \
            ::gsb::prof_begin((pass), (tag), (stream));                  \
        kfn<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);         \
        if (prof__)                                                      \
            ::gsb::prof_end((stream));                                   \
    } while (0)
#endif

#define GSB_DEV __device__ __forceinline__
// Pins a per-polynomial base pointer in a register pair, so that the compiler adds the per-thread word offsets to it
// with LEA / LEA.HI.X on the ALU pipe instead of folding the base's own multiply, or the scaling of the word index,
// into one IMAD.WIDE per load on the multiplier pipe that bounds the transforms (scripts/sass_pipes.py).  The
// pointer comes out of the asm without its address space; the assumption gives it back (LDG / STG, not generic LD).
#ifdef GSB_CUSIM
#define GSB_PIN_GLOBAL(p) ((void)0)
#else
#define GSB_PIN_GLOBAL(p)                   \
    do                                      \
    {                                       \
        asm("" : "+l"(p));                  \
        __builtin_assume(__isGlobal(p));    \
    } while (0)
#endif

// Range checks that only exist in the test-only simulator build (they compile to nothing for nvcc):
// used to validate the no-wrap bounds of the wide-range butterflies on real inputs.
#ifdef GSB_CUSIM
#define GSB_SIM_ASSERT(cond)                                                              \
    do                                                                                    \
    {                                                                                     \
        if (!(cond))                                                                      \
        {                                                                                 \
            fprintf(stderr, "GSB_SIM_ASSERT failed: %s (%s:%d)\n", #cond, __FILE__, __LINE__); \
            abort();                                                                      \
        }                                                                                 \
    } while (0)
#else
#define GSB_SIM_ASSERT(cond) ((void)0)
#endif

namespace gsb
{
    // Kernel launches issued by this library in this process (gsb_launch_count()).
    unsigned long long &launch_counter();
    bool prof_on();
    void prof_begin(const char *pass, const char *tag, cudaStream_t st);
    void prof_end(cudaStream_t st);

    // Dynamic shared memory base (one extern array per translation unit).
    GSB_DEV uint64_t *dyn_smem()
    {
#ifdef GSB_CUSIM
        return reinterpret_cast<uint64_t *>(::cusim::dyn_smem_ptr());
#else
        extern __shared__ __align__(16) uint64_t gsb_dyn_smem[];
        return gsb_dyn_smem;
#endif
    }
} // namespace gsb

// ---- TMA 1-D bulk copy (cp.async.bulk) + mbarrier helpers -------------------------------------------
// Used to stage the twiddle segments of a tile into shared memory asynchronously.  One thread arms
// the barrier with the byte count and issues the copies; every thread waits on phase parity 0.
// Under the test-only simulator the copy is a plain memcpy and the wait is a __syncthreads().
namespace gsb
{
#ifdef GSB_CUSIM
    GSB_DEV void mbar_init(uint64_t *) {}
    GSB_DEV void mbar_expect(uint64_t *, uint32_t) {}
    GSB_DEV void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *)
    {
        memcpy(dst, src, bytes);
    }
    GSB_DEV void mbar_wait(uint64_t *, uint32_t)
    {
        __syncthreads();
    }
    GSB_DEV void fence_proxy_async() {}
    GSB_DEV void prefetch_l2(const void *) {}
#else
    GSB_DEV uint32_t smem_u32(const void *p)
    {
        return static_cast<uint32_t>(__cvta_generic_to_shared(p));
    }
    GSB_DEV void mbar_init(uint64_t *bar)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    GSB_DEV void mbar_expect(uint64_t *bar, uint32_t bytes)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    }
    GSB_DEV void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
    {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                     "l"(src), "r"(bytes), "r"(smem_u32(bar))
                     : "memory");
    }
    // orders this thread's generic-proxy shared-memory accesses with later async-proxy (bulk copy) ones
    GSB_DEV void fence_proxy_async()
    {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    // pulls the line holding p towards L2 / L1 without a destination register
    GSB_DEV void prefetch_l2(const void *p)
    {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
    GSB_DEV void mbar_wait(uint64_t *bar, uint32_t parity)
    {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "WAIT_LOOP:\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
            "@!p bra WAIT_LOOP;\n"
            "}\n" ::"r"(smem_u32(bar)),
            "r"(parity)
            : "memory");
    }
#endif
} // namespace gsb
