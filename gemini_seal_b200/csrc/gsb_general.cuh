// General-size forms of the evaluator hot functions: tensor product of ciphertexts of any sizes
// (dest_size = s1 + s2 - 1; evaluator.cpp:470-520 CKKS, :274-445 BFV) and relinearization of a ciphertext of
// any size down to 2 (relinearize_internal, evaluator.cpp:772-827).  Size-2 x size-2 and size-3 calls keep
// the specialised kernels of gsb_evaluator.cuh; these serve callers that multiply before they relinearize.
// Included at the end of gsb_api.cu.
#pragma once
#include "gsb_evaluator.cuh"

namespace gsb
{
    // out[I] = sum_{j} a[j] (.) b[I - j] mod q_i over the index window of evaluator.cpp:493-519, every dyadic
    // product reduced on its own and accumulated with add_poly_coeffmod, as the reference does.
    struct TensorArgs
    {
        const uint64_t *a, *b;
        uint64_t *out;
        const LimbC *limb;
        uint32_t N, L, s1, s2;
        size_t total; // B * L * N
    };

    template <bool FASTRED>
    __global__ void __launch_bounds__(256) tensor_general_kernel(TensorArgs p)
    {
        const size_t LN = static_cast<size_t>(p.L) * p.N;
        const uint32_t sd = p.s1 + p.s2 - 1;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < p.total;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t ct = i / LN, rem = i - ct * LN;
            const LimbC c = p.limb[rem / p.N];
            const uint64_t *pa = p.a + ct * p.s1 * LN + rem, *pb = p.b + ct * p.s2 * LN + rem;
            uint64_t *po = p.out + ct * sd * LN + rem;
            for (uint32_t I = 0; I < sd; I++)
            {
                const uint32_t last1 = I < p.s1 - 1 ? I : p.s1 - 1, first2 = I < p.s2 - 1 ? I : p.s2 - 1, first1 = I - first2;
                uint64_t acc = 0;
                for (uint32_t j = first1; j <= last1; j++)
                    acc = add_mod(acc, prod_mod<FASTRED>(pa[j * LN], pb[(I - j) * LN], c), c.q);
                po[I * LN] = acc;
            }
        }
    }

    static void tensor_general_core(
        const LimbC *limb, const uint64_t *a, size_t s1, const uint64_t *b, size_t s2, uint64_t *out, size_t B, size_t L, size_t N,
        bool fastred, cudaStream_t st)
    {
        TensorArgs p{ a, b, out, limb, static_cast<uint32_t>(N), static_cast<uint32_t>(L), static_cast<uint32_t>(s1), static_cast<uint32_t>(s2), B * L * N };
        if (fastred)
            GSB_LAUNCH("tensor_general", "", tensor_general_kernel<true>, dim3(grid_for(p.total, 256)), dim3(256), 0, st, p);
        else
            GSB_LAUNCH("tensor_general", "bfv", tensor_general_kernel<false>, dim3(grid_for(p.total, 256)), dim3(256), 0, st, p);
    }

    static void check_sizes(size_t s1, size_t s2)
    {
        // product_fits_in / SEAL_CIPHERTEXT_SIZE_MAX = 16 (util/defines.h:60)
        require(s1 >= 2 && s2 >= 2 && s1 + s2 - 1 <= 16, GSB_E_INVALID_ARG, "invalid ciphertext size");
    }
} // namespace gsb

extern "C" {

int gsb_ckks_multiply_sizes_dev(
    gsb_ctx *ctx, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(a && b && out, GSB_E_INVALID_ARG, "null operand");
        check_sizes(size_a, size_b);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        if (size_a == 2 && size_b == 2)
            ckks_multiply_core(ctx, a, b, out, batch, L, st);
        else
            tensor_general_core(ctx->d_limb, a, size_a, b, size_b, out, batch, L, ctx->N, true, st);
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_bfv_multiply_sizes_dev(
    gsb_ctx *ctx, gsb_ws *ws, uint64_t plain_modulus, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out,
    size_t batch, size_t L, void *stream)
{
    if (size_a == 2 && size_b == 2)
        return gsb_bfv_multiply_dev(ctx, ws, plain_modulus, a, b, out, batch, L, stream);
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(a && b && out, GSB_E_INVALID_ARG, "null operand");
        check_sizes(size_a, size_b);
        const BfvTool &tool = ctx->bfv_tool(L, plain_modulus);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, M = L + tool.nBsk, LN = L * N, MN = M * N, sd = size_a + size_b - 1;
        const size_t per = (size_a + size_b + sd) * MN, chunk = chunk_for(per, batch);
        void *scratch = w->reserve((chunk * per + 2048) * sizeof(uint64_t));
        const NttTables tab{ tool.d_WW, tool.d_VV, tool.d_limb, static_cast<uint32_t>(N), ctx->logn };
        const ModMap mm{ 1 << 30, 0, 0 };
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            Arena ar(scratch);
            uint64_t *ea = ar.words(B * size_a * MN), *eb = ar.words(B * size_b * MN), *dest = ar.words(B * sd * MN);
            // (1)-(3): lift every polynomial of both operands to q u Bsk, lazy NTT (evaluator.cpp:330-367)
            const size_t la = B * size_a * N, lb = B * size_b * N;
            GSB_LAUNCH("behz_lift", "", behz_lift_kernel, dim3(grid_for(la, 256)), dim3(256), 0, st, tool.dev, a + b0 * size_a * LN, ea, la);
            GSB_LAUNCH("behz_lift", "", behz_lift_kernel, dim3(grid_for(lb, 256)), dim3(256), 0, st, tool.dev, b + b0 * size_b * LN, eb, lb);
            GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, ea, MN, ea, MN, static_cast<int>(M), static_cast<int>(B * size_a), true, st, "bfv"));
            GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, eb, MN, eb, MN, static_cast<int>(M), static_cast<int>(B * size_b), true, st, "bfv"));
            // (4): tensor product over all M limbs (:376-420); operands are lazy NTT words -> literal Barrett-128
            tensor_general_core(tool.d_limb, ea, size_a, eb, size_b, dest, B, M, N, false, st);
            // (5): strict inverse NTT; (6)-(8): times t, floor, Shenoy-Kumaresan back to q
            GSB_DISPATCH_LOGN(ctx->logn, ntt_inverse_plain<LOGN>(tab, mm, dest, MN, dest, MN, static_cast<int>(M), static_cast<int>(B * sd), false, st, "bfv"));
            const size_t fl = B * sd * N;
            GSB_LAUNCH("behz_floor", "", behz_floor_kernel, dim3(grid_for(fl, 256)), dim3(256), 0, st, tool.dev, dest, out + b0 * sd * LN, fl);
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// relinearize_internal (evaluator.cpp:772-827) with destination_size = 2: size - 2 key switches.  As in the
// reference, EVERY step switches the LAST component (its iterator is set to component size-1 at :806-807 and is
// not moved inside the loop) with the key RelinKeys::get_index(size - 1 - I) (:811-814).
int gsb_relinearize_sizes_dev(
    gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, size_t size, uint64_t *out, size_t batch, size_t L, int bfv, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(in && out && size >= 2 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, LN = L * N;
        if (size == 2)
        {
            GSB_CUDA_CHECK(cudaMemcpyAsync(out, in, batch * 2 * LN * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
            return;
        }
        std::vector<const SwitchKey *> keys;
        for (size_t I = 0; I + 2 < size; I++)
            keys.push_back(&find_key(ctx, GSB_KEY_RELIN, static_cast<uint32_t>(size - 1 - I - 2))); // get_index(p) = p - 2
        gsb_ws *w = pick_ws(ctx, ws);
        const size_t per = ks_words_per_ct(ctx, L) + size * LN, chunk = chunk_for(per, batch);
        void *scratch = w->reserve((chunk * per + 2048) * sizeof(uint64_t));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            Arena ar(scratch);
            uint64_t *work = ar.words(B * size * LN);
            GSB_CUDA_CHECK(cudaMemcpyAsync(work, in + b0 * size * LN, B * size * LN * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
            for (size_t I = 0; I + 2 < size; I++)
            {
                const KsOutput o{ work, size * LN, LN, work, size * LN, LN, nullptr, false, bfv != 0 };
                GSB_DISPATCH_LOGN(ctx->logn, switch_key_core<LOGN>(ctx, ar.base + ar.off, work + (size - 1) * LN, size * LN, *keys[I], o, B, L, st));
            }
            GSB_CUDA_CHECK(cudaMemcpy2DAsync(
                out + b0 * 2 * LN, 2 * LN * sizeof(uint64_t), work, size * LN * sizeof(uint64_t), 2 * LN * sizeof(uint64_t), B,
                cudaMemcpyDeviceToDevice, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// Evaluator::mod_switch_to_next, CKKS branch = mod_switch_drop_to_next (evaluator.cpp:894-957): a copy of the
// first L-1 limbs of every polynomial.  in = [n_polys][L][N] -> out = [n_polys][L-1][N].
int gsb_drop_last_limb_dev(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t n_polys, size_t L, void *stream)
{
    return guarded([&] {
        check_limbs(ctx, n_polys, L, 0);
        require(in && out && L >= 2, GSB_E_INVALID_ARG, "end of modulus switching chain reached");
        const size_t N = ctx->N;
        GSB_CUDA_CHECK(cudaMemcpy2DAsync(
            out, (L - 1) * N * sizeof(uint64_t), in, L * N * sizeof(uint64_t), (L - 1) * N * sizeof(uint64_t), n_polys, cudaMemcpyDeviceToDevice,
            static_cast<cudaStream_t>(stream)));
    });
}

int gsb_ckks_multiply_sizes_host(
    gsb_ctx *ctx, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, a, size_a * w, b, size_b * w, out, (size_a + size_b - 1) * w, batch,
                     [&](gsb_ws *, uint64_t *da, uint64_t *db, uint64_t *dout, size_t B, void *st) {
                         return gsb_ckks_multiply_sizes_dev(ctx, da, size_a, db, size_b, dout, B, L, st);
                     });
}

int gsb_bfv_multiply_sizes_host(
    gsb_ctx *ctx, uint64_t plain_modulus, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch,
    size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, a, size_a * w, b, size_b * w, out, (size_a + size_b - 1) * w, batch,
                     [&](gsb_ws *ws, uint64_t *da, uint64_t *db, uint64_t *dout, size_t B, void *st) {
                         return gsb_bfv_multiply_sizes_dev(ctx, ws, plain_modulus, da, size_a, db, size_b, dout, B, L, st);
                     });
}

int gsb_relinearize_sizes_host(gsb_ctx *ctx, const uint64_t *in, size_t size, uint64_t *out, size_t batch, size_t L, int bfv)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, size * w, nullptr, 0, out, 2 * w, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_relinearize_sizes_dev(ctx, ws, di, size, dout, B, L, bfv, st);
    });
}

} // extern "C"
