// TEST INFRASTRUCTURE ONLY.  A minimal fiber-based emulation of the CUDA execution model so the
// kernel sources under gemini_seal_b200/csrc can be compiled by g++ (-DGSB_CUSIM) and their
// indexing / synchronisation logic checked against the oracle on a machine without a GPU.
// It is not shipped, not loaded by the package, and is NOT a CPU fallback for the product.
//
// Model: blocks run one after another; the threads of a block are ucontext fibers scheduled
// round-robin; __syncthreads() yields until every live fiber of the block has arrived.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <ucontext.h>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __launch_bounds__(...)
#define __shared__ static
#define __restrict__ __restrict

struct dim3
{
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct ulonglong2
{
    unsigned long long x, y;
};
static inline ulonglong2 make_ulonglong2(unsigned long long a, unsigned long long b)
{
    ulonglong2 r;
    r.x = a;
    r.y = b;
    return r;
}

namespace cusim
{
    struct State
    {
        dim3 grid, block, bidx, tidx;
        char *smem = nullptr;
        size_t smem_cap = 0;
        // scheduler
        ucontext_t sched;
        std::vector<ucontext_t> ctx;
        std::vector<char *> stacks;
        std::vector<int> status; // 0 runnable, 1 at barrier, 2 done
        int cur = 0;
        const std::function<void()> *body = nullptr;
        unsigned long long launches = 0;
    };
    inline State &S()
    {
        static State s;
        return s;
    }
    inline void *dyn_smem_ptr()
    {
        return S().smem;
    }
    inline void trampoline()
    {
        State &s = S();
        (*s.body)();
        s.status[s.cur] = 2;
        swapcontext(&s.ctx[s.cur], &s.sched);
    }
    inline void set_tid(State &s, unsigned t)
    {
        s.tidx.x = t % s.block.x;
        s.tidx.y = (t / s.block.x) % s.block.y;
        s.tidx.z = t / (s.block.x * s.block.y);
    }
    inline void syncthreads()
    {
        State &s = S();
        s.status[s.cur] = 1;
        swapcontext(&s.ctx[s.cur], &s.sched);
    }
    static const size_t kStack = 96 * 1024;
    template <typename F>
    inline void launch(dim3 grid, dim3 block, size_t smem, F &&f)
    {
        State &s = S();
        s.launches++;
        s.grid = grid;
        s.block = block;
        if (smem + 16 > s.smem_cap)
        {
            free(s.smem);
            s.smem_cap = smem + 16;
            s.smem = static_cast<char *>(aligned_alloc(128, (s.smem_cap + 127) / 128 * 128));
        }
        unsigned nt = block.x * block.y * block.z;
        if (s.stacks.size() < nt)
        {
            size_t old = s.stacks.size();
            s.stacks.resize(nt);
            for (size_t i = old; i < nt; i++)
                s.stacks[i] = static_cast<char *>(malloc(kStack));
        }
        s.ctx.resize(nt);
        s.status.assign(nt, 0);
        std::function<void()> body = f;
        s.body = &body;
        for (unsigned bz = 0; bz < grid.z; bz++)
            for (unsigned by = 0; by < grid.y; by++)
                for (unsigned bx = 0; bx < grid.x; bx++)
                {
                    s.bidx = dim3(bx, by, bz);
                    memset(s.smem, 0xA5, smem); // poison: kernels must not rely on initial smem
                    for (unsigned t = 0; t < nt; t++)
                    {
                        getcontext(&s.ctx[t]);
                        s.ctx[t].uc_stack.ss_sp = s.stacks[t];
                        s.ctx[t].uc_stack.ss_size = kStack;
                        s.ctx[t].uc_link = &s.sched;
                        makecontext(&s.ctx[t], (void (*)())trampoline, 0);
                        s.status[t] = 0;
                    }
                    for (;;)
                    {
                        unsigned done = 0, waiting = 0;
                        for (unsigned t = 0; t < nt; t++)
                        {
                            if (s.status[t] == 0)
                            {
                                s.cur = t;
                                set_tid(s, t);
                                swapcontext(&s.sched, &s.ctx[t]);
                            }
                            if (s.status[t] == 2)
                                done++;
                            else if (s.status[t] == 1)
                                waiting++;
                        }
                        if (done == nt)
                            break;
                        if (done + waiting == nt)
                        {
                            if (done != 0)
                            {
                                fprintf(stderr, "cusim: __syncthreads() divergence (some threads exited)\n");
                                abort();
                            }
                            for (unsigned t = 0; t < nt; t++)
                                s.status[t] = 0;
                        }
                    }
                }
        s.body = nullptr;
    }
} // namespace cusim

#define threadIdx (::cusim::S().tidx)
#define blockIdx (::cusim::S().bidx)
#define blockDim (::cusim::S().block)
#define gridDim (::cusim::S().grid)
#define __syncthreads() ::cusim::syncthreads()

static inline int __clz(int x)
{
    return x == 0 ? 32 : __builtin_clz(static_cast<unsigned int>(x));
}
static inline unsigned int __umulhi(unsigned int a, unsigned int b)
{
    return static_cast<unsigned int>((static_cast<unsigned long long>(a) * b) >> 32);
}
static inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b)
{
    return (unsigned long long)(((unsigned __int128)a * b) >> 64);
}
template <typename T>
static inline T __ldg(const T *p)
{
    return *p;
}

// ---- minimal host runtime ---------------------------------------------------------------------
typedef int cudaError_t;
typedef void *cudaStream_t;
typedef void *cudaEvent_t;
enum
{
    cudaSuccess = 0
};
enum cudaMemcpyKind
{
    cudaMemcpyHostToDevice,
    cudaMemcpyDeviceToHost,
    cudaMemcpyDeviceToDevice,
    cudaMemcpyDefault
};
enum cudaFuncAttribute
{
    cudaFuncAttributeMaxDynamicSharedMemorySize
};
static inline cudaError_t cudaMalloc(void **p, size_t n)
{
    *p = aligned_alloc(256, (n + 255) / 256 * 256);
    return *p ? 0 : 2;
}
static inline cudaError_t cudaFree(void *p)
{
    free(p);
    return 0;
}
static inline cudaError_t cudaMallocHost(void **p, size_t n)
{
    return cudaMalloc(p, n);
}
static inline cudaError_t cudaFreeHost(void *p)
{
    free(p);
    return 0;
}
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind)
{
    memmove(d, s, n);
    return 0;
}
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t)
{
    memmove(d, s, n);
    return 0;
}
static inline cudaError_t cudaMemcpy2DAsync(
    void *d, size_t dpitch, const void *s, size_t spitch, size_t width, size_t height, cudaMemcpyKind, cudaStream_t)
{
    for (size_t r = 0; r < height; r++)
        memmove(static_cast<char *>(d) + r * dpitch, static_cast<const char *>(s) + r * spitch, width);
    return 0;
}
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t)
{
    memset(d, v, n);
    return 0;
}
static inline cudaError_t cudaMemset(void *d, int v, size_t n)
{
    memset(d, v, n);
    return 0;
}
static inline cudaError_t cudaStreamSynchronize(cudaStream_t)
{
    return 0;
}
static inline cudaError_t cudaDeviceSynchronize()
{
    return 0;
}
static inline cudaError_t cudaEventCreate(cudaEvent_t *e)
{
    *e = nullptr;
    return 0;
}
enum
{
    cudaStreamNonBlocking = 1,
    cudaEventDisableTiming = 2
};
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned)
{
    *s = nullptr;
    return 0;
}
static inline cudaError_t cudaStreamDestroy(cudaStream_t)
{
    return 0;
}
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned)
{
    return 0;
}
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned)
{
    *e = nullptr;
    return 0;
}
static inline cudaError_t cudaEventDestroy(cudaEvent_t)
{
    return 0;
}
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t)
{
    return 0;
}
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t, cudaEvent_t)
{
    *ms = 0.f;
    return 0;
}
static inline cudaError_t cudaGetLastError()
{
    return 0;
}
static inline cudaError_t cudaSetDevice(int)
{
    return 0;
}
static inline cudaError_t cudaGetDevice(int *d)
{
    *d = 0;
    return 0;
}
static inline const char *cudaGetErrorString(cudaError_t)
{
    return "cusim";
}
template <typename K>
static inline cudaError_t cudaFuncSetAttribute(K, cudaFuncAttribute, int)
{
    return 0;
}
