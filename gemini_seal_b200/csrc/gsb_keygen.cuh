// Switching-key generation on the device: KeyGenerator::generate_one_kswitch_key
// (keygenerator.cpp:325-369) over util::encrypt_zero_symmetric (util/rlwe.cpp:204-304), batched over keys.
//
// Randomness stays with the caller (the reference draws `a` from a fresh BlakePRNG and `e` from its clipped
// normal sampler, rlwe.cpp:232-258): the entry points take the SAMPLED buffers
//     a = [n_keys][beta][K][N]  uniform residues, NTT form (becomes component 1 of every bundle as it is)
//     e = [n_keys][beta][N]     the noise polynomial as signed integers (sample_poly_normal writes the same
//                               small integer into every limb: e or q_i - |e|, rlwe.cpp:60-84)
// and everything arithmetic runs here: the noise's RNS expansion and strict NTT (:263-266), c0 =
// -(a (.) s + NTT(e)) (:259-276), the new key (s^power by dyadic products, keygenerator.cpp:256-323, or the
// Galois-permuted secret key, :99-101) and the (prod_k p_k mod q_i) * new_key added to limb i of c0 for the
// limbs of bundle j (:346-366).  Same sampled buffers in -> the reference's key words out.
// Included at the end of gsb_api.cu.
#pragma once
#include "gsb_evaluator.cuh"

namespace gsb
{
    struct KeygenArgs
    {
        const uint64_t *sk;      // [K][N] NTT form
        const uint64_t *new_key; // [n_keys][K][N] NTT form
        const uint64_t *a;       // [n_keys][beta][K][N]
        const uint64_t *e_ntt;   // [n_keys][beta][K][N] strict NTT of the expanded noise
        uint64_t *out;           // [n_keys][beta][2][K][N]
        const LimbC *limb;
        uint64_t factor[64]; // prod_k p_k mod q_i for i < L0
        uint32_t N, K, L0, k, beta;
        size_t total; // n_keys * beta * K * N
    };

    __global__ void __launch_bounds__(256) keygen_kernel(KeygenArgs p)
    {
        const size_t KN = static_cast<size_t>(p.K) * p.N;
        for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < p.total; t += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t kb = t / KN, rem = t - kb * KN; // kb = key * beta + j
            const uint32_t i = static_cast<uint32_t>(rem / p.N), j = static_cast<uint32_t>(kb % p.beta);
            const size_t key = kb / p.beta;
            const LimbC c = p.limb[i];
            const uint64_t av = p.a[t];
            // dyadic_product_coeffmod(s, c1) ; + NTT(e) ; negate (rlwe.cpp:262-276)
            uint64_t v = mul_mod(p.sk[rem], av, c.q, c.cr0, c.cr1);
            v = add_mod(p.e_ntt[t], v, c.q);
            v = neg_mod(v, c.q);
            // keygenerator.cpp:346-366: limbs of bundle j get factor * new_key
            if (i >= j * p.k && i < j * p.k + p.k && i < p.L0)
                v = add_mod(v, mul_mod(p.new_key[key * KN + rem], p.factor[i], c.q, c.cr0, c.cr1), c.q);
            uint64_t *o = p.out + kb * 2 * KN;
            o[rem] = v;
            o[KN + rem] = av;
        }
    }

    // noise e (signed) -> RNS form [K][N] per (key, bundle): e >= 0 ? e : q_i - |e| (rlwe.cpp:68-84)
    struct NoiseExpandArgs
    {
        const int64_t *e; // [rows][N]
        uint64_t *out;    // [rows][K][N]
        const LimbC *limb;
        uint32_t N, K;
        size_t total; // rows * K * N
    };
    __global__ void __launch_bounds__(256) noise_expand_kernel(NoiseExpandArgs p)
    {
        const size_t KN = static_cast<size_t>(p.K) * p.N;
        for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < p.total; t += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t row = t / KN, rem = t - row * KN;
            const uint32_t i = static_cast<uint32_t>(rem / p.N), d = static_cast<uint32_t>(rem - static_cast<size_t>(i) * p.N);
            const int64_t v = p.e[row * p.N + d];
            p.out[t] = v >= 0 ? static_cast<uint64_t>(v) : p.limb[i].q - static_cast<uint64_t>(-v);
        }
    }
} // namespace gsb

extern "C" {

// kind GSB_KEY_RELIN: ids[n] = key power (>= 2; the key of s^power, RelinKeys::get_index(power) = power - 2);
// kind GSB_KEY_GALOIS: ids[n] = Galois element.  a, e, out are DEVICE buffers (out may be NULL with install = 1);
// install = 1 also places every key in the context's key store (as gsb_key_upload would).
int gsb_kswitch_keygen_dev(
    gsb_ctx *ctx, gsb_ws *ws, int kind, const uint32_t *ids, size_t n_keys, const uint64_t *a, const int64_t *e, uint64_t *out, int install,
    void *stream)
{
    return guarded([&] {
        require(ctx != nullptr && ids != nullptr && a != nullptr && e != nullptr, GSB_E_INVALID_ARG, "null argument");
        require(kind == GSB_KEY_RELIN || kind == GSB_KEY_GALOIS, GSB_E_INVALID_ARG, "key kind");
        require(ctx->K > ctx->k, GSB_E_LOGIC, "keyswitching is not supported by the context");
        require(n_keys >= 1 && n_keys <= 4096, GSB_E_INVALID_ARG, "n_keys out of range");
        require(out != nullptr || install, GSB_E_INVALID_ARG, "null argument");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, K = ctx->K, KN = K * N, beta = (ctx->L0 + ctx->k - 1) / ctx->k, rows = n_keys * beta;
        size_t max_pow = 1;
        for (size_t n = 0; n < n_keys; n++)
        {
            if (kind == GSB_KEY_RELIN)
            {
                require(ids[n] >= 2 && ids[n] <= 16, GSB_E_INVALID_ARG, "key_power cannot be less than 2");
                max_pow = std::max<size_t>(max_pow, ids[n]);
            }
            else
                (void)ctx->galois_table(ids[n]); // validates the element
        }
        const uint64_t *skp = secret_key_powers(ctx, max_pow, st); // needs gsb_secret_key_set
        gsb_ws *w = pick_ws(ctx, ws);
        const size_t words = n_keys * KN + rows * KN + (out ? 0 : rows * 2 * KN);
        Arena ar(w->reserve((words + 4096) * sizeof(uint64_t)));
        uint64_t *new_key = ar.words(n_keys * KN), *e_ntt = ar.words(rows * KN);
        uint64_t *dst = out ? out : ar.words(rows * 2 * KN);
        // the new keys
        for (size_t n = 0; n < n_keys; n++)
        {
            if (kind == GSB_KEY_RELIN)
                GSB_CUDA_CHECK(cudaMemcpyAsync(new_key + n * KN, skp + (ids[n] - 1) * KN, KN * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
            else
                gather_core(ctx, skp, 0, new_key + n * KN, ctx->galois_table(ids[n]), 1, K, st);
        }
        // noise: RNS expansion + strict forward NTT of every limb (rlwe.cpp:263-266)
        {
            NoiseExpandArgs x{ e, e_ntt, ctx->d_limb, static_cast<uint32_t>(N), static_cast<uint32_t>(K), rows * KN };
            GSB_LAUNCH("keygen_noise", "", noise_expand_kernel, dim3(grid_for(x.total, 256)), dim3(256), 0, st, x);
            const NttTables tab = ctx->ntt_tables();
            const ModMap mm{ 1 << 30, 0, 0 };
            GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, e_ntt, KN, e_ntt, KN, static_cast<int>(K), static_cast<int>(rows), false, st, "keygen"));
        }
        KeygenArgs p{};
        p.sk = skp, p.new_key = new_key, p.a = a, p.e_ntt = e_ntt, p.out = dst, p.limb = ctx->d_limb;
        for (size_t i = 0; i < ctx->L0; i++)
        {
            const host::ModConst &qi = ctx->tables[i].mod;
            uint64_t f = 1;
            for (size_t s = 0; s < ctx->k; s++)
                f = host::mul_mod(f, ctx->tables[ctx->L0 + s].mod.q % qi.q, qi);
            p.factor[i] = f;
        }
        p.N = static_cast<uint32_t>(N), p.K = static_cast<uint32_t>(K), p.L0 = static_cast<uint32_t>(ctx->L0), p.k = static_cast<uint32_t>(ctx->k);
        p.beta = static_cast<uint32_t>(beta), p.total = rows * KN;
        GSB_LAUNCH("keygen", "", keygen_kernel, dim3(grid_for(p.total, 256)), dim3(256), 0, st, p);
        if (install)
        {
            const size_t per_key = beta * 2 * KN;
            for (size_t n = 0; n < n_keys; n++)
            {
                uint64_t *d = nullptr;
                GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), per_key * sizeof(uint64_t) + 256));
                int *d_bad = reinterpret_cast<int *>(d + per_key);
                GSB_CUDA_CHECK(cudaMemcpyAsync(d, dst + n * per_key, per_key * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
                GSB_CUDA_CHECK(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
                KeySwizzleArgs sa{ d, ctx->d_limb, static_cast<uint32_t>(N), static_cast<uint32_t>(K), per_key / 128, d_bad };
                GSB_LAUNCH("key_swizzle", "", key_swizzle_kernel, dim3(grid_for(sa.blocks * 64, 256)), dim3(256), 0, st, sa);
                std::lock_guard<std::mutex> lock(ctx->mu);
                const uint32_t slot_id = kind == GSB_KEY_RELIN ? ids[n] - 2 : ids[n];
                SwitchKey &slot = kind == GSB_KEY_RELIN ? ctx->relin[slot_id] : ctx->galois[slot_id];
                if (slot.d)
                {
                    GSB_CUDA_CHECK(cudaStreamSynchronize(st));
                    cudaFree(slot.d);
                }
                slot.d = d;
                slot.bundles = beta;
            }
            GSB_CUDA_CHECK(cudaStreamSynchronize(st)); // keys are visible to every stream from here on
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// Host buffers: a = [n_keys][beta][K][N], e = [n_keys][beta][N] (signed), out = [n_keys][beta][2][K][N] or NULL.
// Keys are processed in chunks that keep the staging below ~2 GiB.
int gsb_kswitch_keygen_host(
    gsb_ctx *ctx, int kind, const uint32_t *ids, size_t n_keys, const uint64_t *a, const int64_t *e, uint64_t *out, int install)
{
    int rc = GSB_OK;
    int g = guarded([&] {
        require(ctx != nullptr && ids != nullptr && a != nullptr && e != nullptr, GSB_E_INVALID_ARG, "null argument");
        require(ctx->K > ctx->k, GSB_E_LOGIC, "keyswitching is not supported by the context");
        require(n_keys >= 1 && n_keys <= 4096, GSB_E_INVALID_ARG, "n_keys out of range");
        std::lock_guard<std::mutex> lock(ctx->host_mu);
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const size_t N = ctx->N, KN = ctx->K * N, beta = (ctx->L0 + ctx->k - 1) / ctx->k;
        const size_t a_words = beta * KN, e_words = beta * N, o_words = beta * 2 * KN;
        size_t chunk = (size_t(2) << 30) / ((a_words + e_words + o_words) * sizeof(uint64_t));
        chunk = chunk < 1 ? 1 : chunk > n_keys ? n_keys : chunk;
        uint64_t *stage = static_cast<uint64_t *>(ctx->host_io.reserve(chunk * (a_words + e_words + o_words) * sizeof(uint64_t)));
        uint64_t *da = stage, *dout = da + chunk * a_words;
        int64_t *de = reinterpret_cast<int64_t *>(dout + chunk * o_words);
        for (size_t n0 = 0; n0 < n_keys && rc == GSB_OK; n0 += chunk)
        {
            const size_t n = std::min(chunk, n_keys - n0);
            GSB_CUDA_CHECK(cudaMemcpyAsync(da, a + n0 * a_words, n * a_words * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
            GSB_CUDA_CHECK(cudaMemcpyAsync(de, e + n0 * e_words, n * e_words * sizeof(int64_t), cudaMemcpyHostToDevice, nullptr));
            rc = gsb_kswitch_keygen_dev(ctx, nullptr, kind, ids + n0, n, da, de, dout, install, nullptr);
            if (rc == GSB_OK && out)
                GSB_CUDA_CHECK(cudaMemcpyAsync(out + n0 * o_words, dout, n * o_words * sizeof(uint64_t), cudaMemcpyDeviceToHost, nullptr));
            GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
        }
    });
    return g != GSB_OK ? g : rc;
}

} // extern "C"
