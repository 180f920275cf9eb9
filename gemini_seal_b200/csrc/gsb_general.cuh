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

    // ---- plaintext operands -----------------------------------------------------------
    // Lift of a coefficient-form BFV plaintext into RNS form under using_fast_plain_lift (every prime larger than
    // t; evaluator.cpp:1577-1589, 1722-1736): out[j][i] = m_i + (increment_j if m_i >= threshold), 0 past `count`.
    struct PlainLiftArgs
    {
        const uint64_t *plain; // [count]
        uint64_t *out;         // [L][N]
        uint64_t threshold;
        uint64_t increment[64]; // plain_upper_half_increment in RNS form
        uint32_t N, L, count;
    };
    __global__ void __launch_bounds__(256) plain_lift_kernel(PlainLiftArgs p)
    {
        const size_t total = static_cast<size_t>(p.L) * p.N;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const uint32_t j = static_cast<uint32_t>(i / p.N), d = static_cast<uint32_t>(i - static_cast<size_t>(j) * p.N);
            uint64_t v = 0;
            if (d < p.count)
            {
                v = p.plain[d];
                v += v >= p.threshold ? p.increment[j] : 0;
            }
            p.out[i] = v;
        }
    }

    // multiply_add / multiply_sub_plain_with_scaling_variant (util/scalingvariant.cpp:15-95): for i < count,
    //   fix = floor(((q mod t) m_i + floor((t+1)/2)) / t);  dst[j][i] = dst[j][i] +- (floor(q/t)_j m_i + fix mod q_j).
    struct ScalingVariantArgs
    {
        const uint64_t *plain; // [count]
        uint64_t *dst;         // first polynomial of the ciphertext, [L][N]
        const LimbC *limb;
        uint64_t q_mod_t, threshold, t;
        uint64_t coeff_div_plain_modulus[64];
        uint32_t N, L, count;
        int subtract;
    };
    __global__ void __launch_bounds__(256) scaling_variant_kernel(ScalingVariantArgs p)
    {
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < p.count; i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const uint64_t m = p.plain[i];
            const unsigned __int128 num = static_cast<unsigned __int128>(m) * p.q_mod_t + p.threshold;
            const uint64_t fix = static_cast<uint64_t>(num / p.t);
            for (uint32_t j = 0; j < p.L; j++)
            {
                const LimbC c = p.limb[j];
                // multiply_add_uint_mod(coeff_div_plain_modulus[j], m, fix) = barrett_reduce_128(a*b + c)
                uint64_t lo = p.coeff_div_plain_modulus[j] * m, hi = mulhi(p.coeff_div_plain_modulus[j], m);
                const uint64_t s = lo + fix;
                hi += s < lo ? 1 : 0;
                const uint64_t val = barrett128(s, hi, c.q, c.cr0, c.cr1);
                uint64_t *d = p.dst + static_cast<size_t>(j) * p.N + i;
                *d = p.subtract ? sub_mod(*d, val, c.q) : add_mod(val, *d, c.q);
            }
        }
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

// ---- plaintext operands, BFV ----------------------------------------------------------
// Evaluator::add_plain_inplace / sub_plain_inplace, BFV branch (evaluator.cpp:1290-1437 ->
// util/scalingvariant.cpp:15-95): enc0 = first polynomial of the ciphertext [L][N] (device, updated in place),
// plain = [count] coefficients mod t (device).  coeff_div_plain_modulus[L], q_mod_t, threshold = (t+1)/2 are the
// ContextData constants of the ciphertext's level (context.h: coeff_div_plain_modulus(),
// coeff_modulus_mod_plain_modulus(), plain_upper_half_threshold()).
int gsb_bfv_add_plain_dev(
    gsb_ctx *ctx, uint64_t *enc0, const uint64_t *plain, size_t count, int subtract, const uint64_t *coeff_div_plain_modulus, uint64_t q_mod_t,
    uint64_t threshold, uint64_t plain_modulus, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, 1, L, 1);
        require(enc0 && plain && coeff_div_plain_modulus, GSB_E_INVALID_ARG, "null operand");
        require(count <= ctx->N && plain_modulus >= 2 && L <= 64, GSB_E_INVALID_ARG, "plain is not valid for encryption parameters");
        if (count == 0)
            return;
        ScalingVariantArgs a{};
        a.plain = plain, a.dst = enc0, a.limb = ctx->d_limb;
        a.q_mod_t = q_mod_t, a.threshold = threshold, a.t = plain_modulus;
        for (size_t j = 0; j < L; j++)
            a.coeff_div_plain_modulus[j] = coeff_div_plain_modulus[j];
        a.N = static_cast<uint32_t>(ctx->N), a.L = static_cast<uint32_t>(L), a.count = static_cast<uint32_t>(count), a.subtract = subtract;
        GSB_LAUNCH("scaling_variant", "", scaling_variant_kernel, dim3(grid_for(count, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), a);
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// Evaluator::transform_to_ntt_inplace(Plaintext, parms_id) under using_fast_plain_lift (evaluator.cpp:1648-1744):
// plain = [count] coefficients mod t -> out = [L][N], the lifted plaintext in NTT form.  threshold =
// plain_upper_half_threshold(), increment[L] = plain_upper_half_increment() (RNS form).  ntt = 0 stops after the lift.
int gsb_bfv_plain_to_rns_dev(
    gsb_ctx *ctx, const uint64_t *plain, size_t count, uint64_t threshold, const uint64_t *increment, uint64_t *out, size_t L, int ntt, void *stream)
{
    return guarded([&] {
        check_batch(ctx, 1, L, 1);
        require(plain && increment && out, GSB_E_INVALID_ARG, "null operand");
        require(count <= ctx->N && L <= 64, GSB_E_INVALID_ARG, "plain is not valid for encryption parameters");
        PlainLiftArgs a{};
        a.plain = plain, a.out = out, a.threshold = threshold;
        for (size_t j = 0; j < L; j++)
            a.increment[j] = increment[j];
        a.N = static_cast<uint32_t>(ctx->N), a.L = static_cast<uint32_t>(L), a.count = static_cast<uint32_t>(count);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        GSB_LAUNCH("plain_lift", "", plain_lift_kernel, dim3(grid_for(L * ctx->N, 256)), dim3(256), 0, st, a);
        if (ntt)
        {
            const NttTables tab = ctx->ntt_tables();
            const ModMap mm{ 1 << 30, 0, 0 };
            GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, out, L * ctx->N, out, L * ctx->N, static_cast<int>(L), 1, false, st, "plain"));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// Evaluator::multiply_plain_inplace with a coefficient-form plaintext = multiply_plain_normal (evaluator.cpp:
// 1475-1603), generic branch: enc = [batch][size][L][N] coefficient form, updated in place; plain_ntt = [L][N] the
// lifted plaintext in NTT form (gsb_bfv_plain_to_rns_dev); per polynomial: lazy NTT, dyadic product, strict
// inverse NTT (:1594-1601).  (A monomial plaintext goes through the same route: the caller lifts it as :1512-1546
// does, the negacyclic product is the same canonical word.)
int gsb_multiply_plain_coeff_dev(gsb_ctx *ctx, uint64_t *enc, const uint64_t *plain_ntt, size_t batch, size_t size, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(enc && plain_ntt && size >= 1 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t LN = L * ctx->N, polys = batch * size;
        const NttTables tab = ctx->ntt_tables();
        const ModMap mm{ 1 << 30, 0, 0 };
        GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, enc, LN, enc, LN, static_cast<int>(L), static_cast<int>(polys), true, st, "plain"));
        for (size_t p = 0; p < polys; p++)
        {
            PolyOpArgs a{};
            a.a = enc + p * LN, a.b = plain_ntt, a.out = enc + p * LN, a.limb = ctx->d_limb;
            a.N = static_cast<uint32_t>(ctx->N), a.n_limbs = static_cast<uint32_t>(L), a.limb_offset = 0, a.total = LN;
            launch_poly_op(GSB_OP_DYADIC, a, st);
        }
        GSB_DISPATCH_LOGN(ctx->logn, ntt_inverse_plain<LOGN>(tab, mm, enc, LN, enc, LN, static_cast<int>(L), static_cast<int>(polys), false, st, "plain"));
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// Host-buffer forms of the three (what the C++ shim calls): copy in, run, copy out, synchronise.
int gsb_bfv_add_plain_host(
    gsb_ctx *ctx, uint64_t *enc0, const uint64_t *plain, size_t count, int subtract, const uint64_t *coeff_div_plain_modulus, uint64_t q_mod_t,
    uint64_t threshold, uint64_t plain_modulus, size_t L)
{
    int rc = GSB_OK;
    int g = guarded([&] {
        check_batch(ctx, 1, L, 1);
        require(enc0 && plain, GSB_E_INVALID_ARG, "null operand");
        require(count <= ctx->N, GSB_E_INVALID_ARG, "plain is not valid for encryption parameters");
        std::lock_guard<std::mutex> lock(ctx->host_mu);
        const size_t LN = L * ctx->N;
        uint64_t *d = static_cast<uint64_t *>(ctx->host_io.reserve((LN + ctx->N) * sizeof(uint64_t)));
        GSB_CUDA_CHECK(cudaMemcpyAsync(d, enc0, LN * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
        if (count)
            GSB_CUDA_CHECK(cudaMemcpyAsync(d + LN, plain, count * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
        rc = gsb_bfv_add_plain_dev(ctx, d, d + LN, count, subtract, coeff_div_plain_modulus, q_mod_t, threshold, plain_modulus, L, nullptr);
        if (rc == GSB_OK)
        {
            GSB_CUDA_CHECK(cudaMemcpyAsync(enc0, d, LN * sizeof(uint64_t), cudaMemcpyDeviceToHost, nullptr));
            GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
        }
    });
    return g != GSB_OK ? g : rc;
}

int gsb_bfv_plain_to_rns_host(gsb_ctx *ctx, const uint64_t *plain, size_t count, uint64_t threshold, const uint64_t *increment, uint64_t *out, size_t L, int ntt)
{
    int rc = GSB_OK;
    int g = guarded([&] {
        check_batch(ctx, 1, L, 1);
        require(plain && out, GSB_E_INVALID_ARG, "null operand");
        require(count <= ctx->N, GSB_E_INVALID_ARG, "plain is not valid for encryption parameters");
        std::lock_guard<std::mutex> lock(ctx->host_mu);
        const size_t LN = L * ctx->N;
        uint64_t *d = static_cast<uint64_t *>(ctx->host_io.reserve((LN + ctx->N) * sizeof(uint64_t)));
        if (count)
            GSB_CUDA_CHECK(cudaMemcpyAsync(d + LN, plain, count * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
        rc = gsb_bfv_plain_to_rns_dev(ctx, d + LN, count, threshold, increment, d, L, ntt, nullptr);
        if (rc == GSB_OK)
        {
            GSB_CUDA_CHECK(cudaMemcpyAsync(out, d, LN * sizeof(uint64_t), cudaMemcpyDeviceToHost, nullptr));
            GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
        }
    });
    return g != GSB_OK ? g : rc;
}

// enc = [size][L][N] (host, in place); plain = [count] coefficient-form plaintext as the caller lifts it:
// lift == 1 applies the upper-half increment (generic branch), lift == 0 takes the words as they are (monomial branch)
int gsb_multiply_plain_coeff_host(
    gsb_ctx *ctx, uint64_t *enc, size_t size, size_t L, const uint64_t *plain, size_t count, uint64_t threshold, const uint64_t *increment, int lift)
{
    int rc = GSB_OK;
    int g = guarded([&] {
        check_batch(ctx, 1, L, 1);
        require(enc && plain && increment && size >= 1 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        require(count <= ctx->N, GSB_E_INVALID_ARG, "plain is not valid for encryption parameters");
        std::lock_guard<std::mutex> lock(ctx->host_mu);
        const size_t LN = L * ctx->N;
        uint64_t *d = static_cast<uint64_t *>(ctx->host_io.reserve(((size + 1) * LN + ctx->N) * sizeof(uint64_t)));
        uint64_t *dp = d + size * LN, *draw = dp + LN;
        GSB_CUDA_CHECK(cudaMemcpyAsync(d, enc, size * LN * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
        if (count)
            GSB_CUDA_CHECK(cudaMemcpyAsync(draw, plain, count * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
        rc = gsb_bfv_plain_to_rns_dev(ctx, draw, count, lift ? threshold : ~uint64_t(0), increment, dp, L, 1, nullptr);
        if (rc == GSB_OK)
            rc = gsb_multiply_plain_coeff_dev(ctx, d, dp, 1, size, L, nullptr);
        if (rc == GSB_OK)
        {
            GSB_CUDA_CHECK(cudaMemcpyAsync(enc, d, size * LN * sizeof(uint64_t), cudaMemcpyDeviceToHost, nullptr));
            GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
        }
    });
    return g != GSB_OK ? g : rc;
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
