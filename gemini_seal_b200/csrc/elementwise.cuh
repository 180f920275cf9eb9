// Elementwise RNS kernels: dyadic product, add, sub, negate, scalar multiply
// (reference: util/polyarithsmallmod.cpp:15-117, util/polyarithsmallmod.h:176-470), the CKKS
// tensor product (evaluator.cpp:447-527) and the NTT-form Galois gather (util/galois.cpp:188-214).
// All are pure HBM streams: 128-bit accesses, one pass, grid sized to the data.
#pragma once
#include "ntt_kernels.cuh"
#include "modarith.cuh"

namespace gsb
{
    struct PolyOpArgs
    {
        const uint64_t *a, *b;
        uint64_t *out;
        const LimbC *limb;
        uint32_t N, n_limbs, limb_offset;
        size_t total; // words
        uint64_t scalar[64];
    };

    template <int OP>
    GSB_DEV uint64_t poly_op_one(uint64_t x, uint64_t y, uint64_t s, const LimbC &c)
    {
        switch (OP)
        {
        case GSB_OP_DYADIC: return mul_mod(x, y, c.q, c.cr0, c.cr1);
        case GSB_OP_ADD: return add_mod(x, y, c.q);
        case GSB_OP_SUB: return sub_mod(x, y, c.q);
        case GSB_OP_NEGATE: return neg_mod(x, c.q);
        default: return mul_mod(x, s, c.q, c.cr0, c.cr1);
        }
    }

    template <int OP>
    __global__ void __launch_bounds__(256) poly_op_kernel(PolyOpArgs p)
    {
        const size_t pairs = p.total >> 1;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < pairs;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t w = i << 1;
            const uint32_t y = static_cast<uint32_t>((w / p.N) % p.n_limbs);
            const LimbC c = p.limb[p.limb_offset + y];
            const uint64_t s = OP == GSB_OP_MUL_SCALAR ? p.scalar[y] : 0;
            const ulonglong2 va = *reinterpret_cast<const ulonglong2 *>(p.a + w);
            ulonglong2 vb = make_ulonglong2(0, 0);
            if (OP == GSB_OP_DYADIC || OP == GSB_OP_ADD || OP == GSB_OP_SUB)
                vb = *reinterpret_cast<const ulonglong2 *>(p.b + w);
            *reinterpret_cast<ulonglong2 *>(p.out + w) =
                make_ulonglong2(poly_op_one<OP>(va.x, vb.x, s, c), poly_op_one<OP>(va.y, vb.y, s, c));
        }
    }

    inline unsigned grid_for(size_t items, unsigned block)
    {
        size_t g = (items + block - 1) / block;
        const size_t cap = sm_count() * 16u; // grid-stride kernels: 16 CTAs of 256 threads per SM
        return static_cast<unsigned>(g < 1 ? 1 : g > cap ? cap : g);
    }

    inline void launch_poly_op(int op, const PolyOpArgs &args, cudaStream_t st)
    {
        const unsigned grid = grid_for(args.total >> 1, 256);
        switch (op)
        {
        case GSB_OP_DYADIC: GSB_LAUNCH("poly_op", "", poly_op_kernel<GSB_OP_DYADIC>, dim3(grid), dim3(256), 0, st, args); break;
        case GSB_OP_ADD: GSB_LAUNCH("poly_op", "", poly_op_kernel<GSB_OP_ADD>, dim3(grid), dim3(256), 0, st, args); break;
        case GSB_OP_SUB: GSB_LAUNCH("poly_op", "", poly_op_kernel<GSB_OP_SUB>, dim3(grid), dim3(256), 0, st, args); break;
        case GSB_OP_NEGATE: GSB_LAUNCH("poly_op", "", poly_op_kernel<GSB_OP_NEGATE>, dim3(grid), dim3(256), 0, st, args); break;
        default: GSB_LAUNCH("poly_op", "", poly_op_kernel<GSB_OP_MUL_SCALAR>, dim3(grid), dim3(256), 0, st, args); break;
        }
    }
} // namespace gsb
