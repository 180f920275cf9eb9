// TEST INFRASTRUCTURE (oracle).  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
// --impl reference legs may load this library.  It is never on the product path.
//
// A thin extern "C" harness around the UNMODIFIED reference (Gemini-SEAL = SEAL 3.5.3 + fork delta),
// compiled from /root/reference by oracle/Makefile into oracle/_ref/.  It lets Python (ctypes) feed
// raw limb buffers through the reference's own hot-path entry points:
//   util::ntt_negacyclic_harvey[_lazy] / inverse_...         (native/src/seal/util/ntt.h:189-368)
//   Evaluator::multiply_inplace                              (evaluator.cpp:235)
//   Evaluator::relinearize_inplace                           (evaluator.h:344, evaluator.cpp:772)
//   Evaluator::rescale_to_next_inplace                       (evaluator.cpp:1090)
//   Evaluator::apply_galois_inplace / rotate_vector_inplace  (evaluator.cpp:1841,1945)
// All code in this file is ours; it only calls the reference's public API.
#include "seal/seal.h"
#include "seal/util/ntt.h"
#include "seal/util/polyarithsmallmod.h"
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <random>
#include <string>
#include <thread>
#include <vector>

using namespace seal;
using namespace seal::util;

namespace
{
    thread_local std::string g_err;

    struct RefCtx
    {
        std::shared_ptr<SEALContext> context;
        std::unique_ptr<Evaluator> evaluator;
        std::unique_ptr<KeyGenerator> keygen;
        RelinKeys relin;
        GaloisKeys galois;
        bool have_relin = false, have_galois = false;
        bool bfv = false; // BFV context: coefficient-form ciphertexts, scale 1
        size_t n_relin = 1; // relinearization keys (powers s^2 ..) made by ref_keygen / ref_make_synthetic_keys
        size_t N = 0, K = 0, k = 0, L0 = 0;

        parms_id_type parms_at(size_t L) const
        {
            auto cd = context->first_context_data();
            while (cd)
            {
                if (cd->parms().coeff_modulus().size() == L)
                    return cd->parms_id();
                cd = cd->next_context_data();
            }
            throw std::invalid_argument("no level with that many limbs");
        }

        Ciphertext make_ct(const uint64_t *src, size_t size, size_t L, double scale) const
        {
            Ciphertext ct(context, parms_at(L));
            ct.resize(size);
            ct.is_ntt_form() = !bfv;
            ct.scale() = bfv ? 1.0 : scale;
            std::memcpy(ct.data(), src, size * L * N * sizeof(uint64_t));
            return ct;
        }
    };

    template <typename F>
    int guarded(F &&f)
    {
        try
        {
            f();
            return 0;
        }
        catch (const std::invalid_argument &e)
        {
            g_err = std::string("invalid_argument: ") + e.what();
            return -1;
        }
        catch (const std::logic_error &e)
        {
            g_err = std::string("logic_error: ") + e.what();
            return -2;
        }
        catch (const std::exception &e)
        {
            g_err = std::string("exception: ") + e.what();
            return -3;
        }
    }
} // namespace

extern "C"
{
    const char *ref_last_error()
    {
        return g_err.c_str();
    }

    // ---- raw NTT (no context needed) -------------------------------------------------------------
    // kind: 0 = forward lazy, 1 = forward strict, 2 = inverse lazy, 3 = inverse strict.
    // data: n_polys contiguous limbs of N = 2^logn words, all under modulus q.
    int ref_ntt(int logn, uint64_t q, uint64_t *data, size_t n_polys, int kind)
    {
        return guarded([&] {
            Modulus mod(q);
            NTTTables tables(logn, mod);
            size_t n = size_t(1) << logn;
            for (size_t p = 0; p < n_polys; p++)
            {
                uint64_t *x = data + p * n;
                switch (kind)
                {
                case 0: ntt_negacyclic_harvey_lazy(x, tables); break;
                case 1: ntt_negacyclic_harvey(x, tables); break;
                case 2: inverse_ntt_negacyclic_harvey_lazy(x, tables); break;
                case 3: inverse_ntt_negacyclic_harvey(x, tables); break;
                default: throw std::invalid_argument("kind");
                }
            }
        });
    }

    // Dump the reference's NTTTables (ntt.h:46-60,84-92,116) for (logn, q); each array has N words.
    // scalars[0..4] = root, inv_degree_modulo, scaled_inv_degree, reduce_precomp, const_ratio[1]
    int ref_ntt_tables(
        int logn, uint64_t q, uint64_t *root_powers, uint64_t *scaled_root_powers, uint64_t *inv_root_powers,
        uint64_t *scaled_inv_root_powers, uint64_t *scalars)
    {
        return guarded([&] {
            Modulus mod(q);
            NTTTables t(logn, mod);
            size_t n = size_t(1) << logn;
            std::memcpy(root_powers, t.root_powers(), n * 8);
            std::memcpy(scaled_root_powers, t.scaled_root_powers(), n * 8);
            std::memcpy(inv_root_powers, t.inv_root_powers(), n * 8);
            std::memcpy(scaled_inv_root_powers, t.scaled_inv_root_powers(), n * 8);
            scalars[0] = t.get_root();
            scalars[1] = *t.get_inv_degree_modulo();
            scalars[2] = t.get_scaled_inv_degree_modulo();
            scalars[3] = t.get_reduce_precomp();
            scalars[4] = mod.const_ratio()[1];
        });
    }

    // Modulus::const_ratio (modulus.cpp:87-98): out[0..2]
    int ref_const_ratio(uint64_t q, uint64_t *out)
    {
        return guarded([&] {
            Modulus mod(q);
            for (int i = 0; i < 3; i++)
                out[i] = mod.const_ratio()[i];
        });
    }

    // elementwise kernels (polyarithsmallmod.cpp:15-117, .h:176-470). op: 0 dyadic, 1 add, 2 sub, 3 negate(a), 4 a*scalar
    int ref_poly_op(int op, uint64_t q, const uint64_t *a, const uint64_t *b, uint64_t scalar, size_t n, uint64_t *out)
    {
        return guarded([&] {
            Modulus mod(q);
            switch (op)
            {
            case 0: dyadic_product_coeffmod(a, b, n, mod, out); break;
            case 1: add_poly_coeffmod(a, b, n, mod, out); break;
            case 2: sub_poly_coeffmod(a, b, n, mod, out); break;
            case 3: negate_poly_coeffmod(a, n, mod, out); break;
            case 4: multiply_poly_scalar_coeffmod(a, n, scalar, mod, out); break;
            default: throw std::invalid_argument("op");
            }
        });
    }

    // CoeffModulus::BFVDefault(N) (util/globals.cpp:22-160) -- what 7_performance.cpp uses.
    int ref_default_moduli(size_t N, uint64_t *out, size_t cap)
    {
        int count = -1;
        int rc = guarded([&] {
            auto v = CoeffModulus::BFVDefault(N);
            if (v.size() > cap)
                throw std::invalid_argument("cap");
            for (size_t i = 0; i < v.size(); i++)
                out[i] = v[i].value();
            count = static_cast<int>(v.size());
        });
        return rc ? rc : count;
    }

    // CoeffModulus::Create(N, bit_sizes)
    int ref_create_moduli(size_t N, const int *bits, size_t count, uint64_t *out)
    {
        return guarded([&] {
            std::vector<int> b(bits, bits + count);
            auto v = CoeffModulus::Create(N, b);
            for (size_t i = 0; i < v.size(); i++)
                out[i] = v[i].value();
        });
    }

    // ---- CKKS context + Evaluator -------------------------------------------------------------
    // moduli[0..K): key-level coefficient modulus; the last k are special primes (encryptionparams.h:205).
    void *ref_ckks_create(size_t N, const uint64_t *moduli, size_t K, size_t k_special)
    {
        RefCtx *c = nullptr;
        int rc = guarded([&] {
            EncryptionParameters parms(scheme_type::CKKS);
            parms.set_poly_modulus_degree(N);
            std::vector<Modulus> mods;
            for (size_t i = 0; i < K; i++)
                mods.emplace_back(moduli[i]);
            parms.set_coeff_modulus(mods);
            parms.set_n_special_primes(static_cast<int>(k_special));
            auto ctx = SEALContext::Create(parms, true, sec_level_type::none);
            if (!ctx->parameters_set())
                throw std::invalid_argument(std::string("parameters not set: ") + ctx->parameter_error_message());
            c = new RefCtx();
            c->context = ctx;
            c->evaluator.reset(new Evaluator(ctx));
            c->N = N;
            c->K = K;
            c->k = k_special;
            c->L0 = K - k_special;
        });
        if (rc)
        {
            delete c;
            return nullptr;
        }
        return c;
    }

    // BFV context over the same moduli (scheme_type::BFV, evaluator.cpp:274-445 and the BFV branches
    // of switch_key_inplace / apply_galois_inplace); ciphertexts handed to the ops below are then in
    // coefficient form.
    void *ref_bfv_create(size_t N, const uint64_t *moduli, size_t K, size_t k_special, uint64_t plain_modulus)
    {
        RefCtx *c = nullptr;
        int rc = guarded([&] {
            EncryptionParameters parms(scheme_type::BFV);
            parms.set_poly_modulus_degree(N);
            std::vector<Modulus> mods;
            for (size_t i = 0; i < K; i++)
                mods.emplace_back(moduli[i]);
            parms.set_coeff_modulus(mods);
            parms.set_plain_modulus(plain_modulus);
            parms.set_n_special_primes(static_cast<int>(k_special));
            auto ctx = SEALContext::Create(parms, true, sec_level_type::none);
            if (!ctx->parameters_set())
                throw std::invalid_argument(std::string("parameters not set: ") + ctx->parameter_error_message());
            c = new RefCtx();
            c->context = ctx;
            c->evaluator.reset(new Evaluator(ctx));
            c->bfv = true;
            c->N = N;
            c->K = K;
            c->k = k_special;
            c->L0 = K - k_special;
        });
        if (rc)
        {
            delete c;
            return nullptr;
        }
        return c;
    }

    void ref_ckks_destroy(void *h)
    {
        delete static_cast<RefCtx *>(h);
    }

    // Generate a secret key + relin keys + Galois keys for the given elements with the reference's
    // own KeyGenerator (keygenerator.cpp:325-369).  Not reproducible run to run: the
    // limbs are dumped with ref_get_key and fed to both implementations.
    int ref_keygen(void *h, const uint32_t *galois_elts, size_t n_elts)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            c->keygen.reset(new KeyGenerator(c->context));
            c->relin = c->keygen->relin_keys_local();
            // the public KeyGenerator only makes the s^2 key (keygenerator.h:82-85); further powers get the same
            // shape here and their limbs from ref_set_key
            for (size_t p = 1; p < c->n_relin; p++)
                c->relin.data().push_back(c->relin.data()[0]);
            c->have_relin = true;
            if (n_elts)
            {
                std::vector<uint32_t> e(galois_elts, galois_elts + n_elts);
                c->galois = c->keygen->galois_keys_local(e);
                c->have_galois = true;
            }
        });
    }

    // Synthetic switching keys without the (slow, non-reproducible) KeyGenerator: build KSwitchKeys of
    // the shape generate_kswitch_keys produces (keygenerator.cpp:325-369: one PublicKey per bundle, each
    // a size-2 NTT-form ciphertext at key level) and fill every limb with uniform residues from
    // mt19937_64(seed).  The same limbs are read back with ref_get_key and uploaded to the GPU.
    int ref_make_synthetic_keys(void *h, const uint32_t *galois_elts, size_t n_elts, uint64_t seed)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto key_cd = c->context->key_context_data();
            auto &mods = key_cd->parms().coeff_modulus();
            size_t n_bundles = (c->L0 + c->k - 1) / c->k;
            std::mt19937_64 rng(seed);
            auto fill = [&](std::vector<PublicKey> &v) {
                v.resize(n_bundles);
                for (auto &pk : v)
                {
                    Ciphertext &ct = pk.data();
                    ct.resize(c->context, key_cd->parms_id(), 2);
                    ct.is_ntt_form() = true;
                    ct.scale() = 1.0;
                    for (size_t p = 0; p < 2; p++)
                        for (size_t i = 0; i < c->K; i++)
                            for (size_t d = 0; d < c->N; d++)
                                ct.data(p)[i * c->N + d] = rng() % mods[i].value();
                }
            };
            c->relin = RelinKeys();
            c->relin.data().resize(c->n_relin);
            for (auto &v : c->relin.data())
                fill(v);
            c->relin.parms_id() = key_cd->parms_id();
            c->have_relin = true;
            c->galois = GaloisKeys();
            for (size_t e = 0; e < n_elts; e++)
            {
                size_t idx = GaloisKeys::get_index(galois_elts[e]);
                if (c->galois.data().size() <= idx)
                    c->galois.data().resize(idx + 1);
                fill(c->galois.data()[idx]);
            }
            c->galois.parms_id() = key_cd->parms_id();
            c->have_galois = n_elts > 0;
        });
    }

    // Synthetic keys: allocate KSwitchKeys of the right shape by running keygen once, then
    // (optionally) the caller overwrites the limbs with ref_set_key.
    // which: 0 = relin (index 0), 1 = galois for element `elt`.
    // Layout of one switching key: [bundle][component 0/1][K limbs][N]  (kswitchkeys.h:343).
    static std::vector<PublicKey> &key_vec(RefCtx *c, int which, uint32_t elt)
    {
        if (which == 0)
        {
            if (!c->have_relin)
                throw std::logic_error("no relin keys");
            if (elt >= c->relin.data().size())
                throw std::logic_error("no relin key of that power");
            return c->relin.data()[elt]; // elt = RelinKeys::get_index(key_power)
        }
        if (!c->have_galois || !c->galois.has_key(elt))
            throw std::logic_error("no galois key for elt");
        return c->galois.data()[GaloisKeys::get_index(elt)];
    }

    int ref_key_bundles(void *h, int which, uint32_t elt)
    {
        auto *c = static_cast<RefCtx *>(h);
        int n = -1;
        int rc = guarded([&] { n = static_cast<int>(key_vec(c, which, elt).size()); });
        return rc ? rc : n;
    }

    int ref_get_key(void *h, int which, uint32_t elt, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto &v = key_vec(c, which, elt);
            size_t per = 2 * c->K * c->N;
            for (size_t j = 0; j < v.size(); j++)
                std::memcpy(out + j * per, v[j].data().data(), per * 8);
        });
    }

    int ref_set_key(void *h, int which, uint32_t elt, const uint64_t *in)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto &v = key_vec(c, which, elt);
            size_t per = 2 * c->K * c->N;
            for (size_t j = 0; j < v.size(); j++)
                std::memcpy(v[j].data().data(), in + j * per, per * 8);
        });
    }

    // SecretKey::data() of the KeyGenerator made by ref_keygen: NTT form at key level, [K][N]
    int ref_get_secret_key(void *h, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            if (!c->keygen)
                throw std::logic_error("no key generator");
            std::memcpy(out, c->keygen->secret_key().data().data(), c->K * c->N * 8);
        });
    }

    // how many relinearization keys ref_keygen / ref_make_synthetic_keys produce from now on (default 1)
    int ref_set_relin_count(void *h, size_t count)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            if (count < 1 || count > 14)
                throw std::invalid_argument("count");
            c->n_relin = count;
        });
    }

    // multiply_inplace for any sizes: a = [sa][L][N], b = [sb][L][N] -> out [sa + sb - 1][L][N]
    int ref_multiply_sizes(void *h, size_t L, const uint64_t *a, size_t sa, const uint64_t *b, size_t sb, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(a, sa, L, std::pow(2.0, 10));
            auto y = c->make_ct(b, sb, L, std::pow(2.0, 10));
            c->evaluator->multiply_inplace(x, y);
            std::memcpy(out, x.data(), (sa + sb - 1) * L * c->N * 8);
        });
    }

    // relinearize_inplace from any size: in [size][L][N] -> out [2][L][N]
    int ref_relinearize_sizes(void *h, size_t L, size_t size, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(in, size, L, std::pow(2.0, 30));
            c->evaluator->relinearize_inplace(x, c->relin);
            std::memcpy(out, x.data(), 2 * L * c->N * 8);
        });
    }

    // multiply_inplace: a, b = [2][L][N] -> out [3][L][N]
    int ref_multiply(void *h, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(a, 2, L, std::pow(2.0, 20));
            auto y = c->make_ct(b, 2, L, std::pow(2.0, 20));
            c->evaluator->multiply_inplace(x, y);
            std::memcpy(out, x.data(), 3 * L * c->N * 8);
        });
    }

    // relinearize_inplace: in [3][L][N] -> out [2][L][N]
    int ref_relinearize(void *h, size_t L, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(in, 3, L, std::pow(2.0, 40));
            c->evaluator->relinearize_inplace(x, c->relin);
            std::memcpy(out, x.data(), 2 * L * c->N * 8);
        });
    }

    // rescale_to_next_inplace: in [size][L][N] -> out [size][L-1][N]
    int ref_rescale(void *h, size_t L, size_t size, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(in, size, L, std::pow(2.0, 40));
            c->evaluator->rescale_to_next_inplace(x);
            std::memcpy(out, x.data(), size * (L - 1) * c->N * 8);
        });
    }

    // mod_switch_to_next_inplace: in [size][L][N] -> out [size][L-1][N] (BFV: divide and round; CKKS: drop)
    int ref_mod_switch(void *h, size_t L, size_t size, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(in, size, L, std::pow(2.0, 20));
            c->evaluator->mod_switch_to_next_inplace(x);
            std::memcpy(out, x.data(), size * (L - 1) * c->N * 8);
        });
    }

    // apply_galois_inplace: in/out [2][L][N]
    int ref_apply_galois(void *h, size_t L, uint32_t elt, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(in, 2, L, std::pow(2.0, 40));
            c->evaluator->apply_galois_inplace(x, elt, c->galois);
            std::memcpy(out, x.data(), 2 * L * c->N * 8);
        });
    }

    // rotate_vector_inplace(ct, steps): in/out [2][L][N]
    int ref_rotate(void *h, size_t L, int steps, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(in, 2, L, std::pow(2.0, 40));
            c->evaluator->rotate_vector_inplace(x, steps, c->galois);
            std::memcpy(out, x.data(), 2 * L * c->N * 8);
        });
    }

    // mul + relin + rescale in one go (the unit of work of SURVEY 8d(ii)): a,b [2][L][N] -> out [2][L-1][N]
    int ref_mul_relin_rescale(void *h, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            auto x = c->make_ct(a, 2, L, std::pow(2.0, 20));
            auto y = c->make_ct(b, 2, L, std::pow(2.0, 20));
            c->evaluator->multiply_inplace(x, y);
            c->evaluator->relinearize_inplace(x, c->relin);
            c->evaluator->rescale_to_next_inplace(x);
            std::memcpy(out, x.data(), 2 * (L - 1) * c->N * 8);
        });
    }

    uint32_t ref_elt_from_step(void *h, int step)
    {
        auto *c = static_cast<RefCtx *>(h);
        return c->context->key_context_data()->galois_tool()->get_elt_from_step(step);
    }

    // ---- CPU baseline timing (bench.py cpu_baseline / --impl reference) ---------------------------
    // Decryptor::decrypt (decryptor.cpp:60-151) with the secret key given as raw NTT-form limbs at key
    // level, sk_limbs = [K][N] (any canonical residues pass is_valid_for).  in = [size][L][N].
    // CKKS: out = [L][N], the plaintext = dot_product_ct_sk_array's result.  BFV: out = [N] plaintext
    // coefficients mod t (zero-padded past Plaintext::coeff_count()).
    int ref_decrypt(void *h, const uint64_t *sk_limbs, size_t L, size_t size, const uint64_t *in, uint64_t *out)
    {
        auto *c = static_cast<RefCtx *>(h);
        return guarded([&] {
            SecretKey sk;
            sk.data().resize(c->K * c->N);
            std::memcpy(sk.data().data(), sk_limbs, c->K * c->N * sizeof(uint64_t));
            sk.parms_id() = c->context->key_parms_id();
            Decryptor dec(c->context, sk);
            auto x = c->make_ct(in, size, L, std::pow(2.0, 40));
            Plaintext p;
            dec.decrypt(x, p);
            if (c->bfv)
            {
                std::memset(out, 0, c->N * sizeof(uint64_t));
                std::memcpy(out, p.data(), p.coeff_count() * sizeof(uint64_t));
            }
            else
                std::memcpy(out, p.data(), L * c->N * sizeof(uint64_t));
        });
    }

    // Runs `ops` units of work spread over n_threads host threads sharing one Evaluator (const
    // re-entrant) and returns wall seconds, or a negative error code.
    // what: 0 = mul+relin+rescale, 1 = rotate_vector(ct, 1), 2 = strict fwd limb-NTT, 3 = strict inv limb-NTT
    double ref_time_ops(void *h, int what, size_t L, size_t ops, int n_threads, uint64_t seed)
    {
        auto *c = static_cast<RefCtx *>(h);
        double secs = -1.0;
        int rc = guarded([&] {
            auto cd = c->context->get_context_data(c->parms_at(L));
            auto &mods = cd->parms().coeff_modulus();
            size_t N = c->N;
            // one input pair per thread, uniform residues (SURVEY 8d "synthetic inputs")
            std::vector<std::vector<uint64_t>> A(n_threads), B(n_threads);
            for (int t = 0; t < n_threads; t++)
            {
                std::mt19937_64 rng(seed + t);
                A[t].resize(2 * L * N);
                B[t].resize(2 * L * N);
                for (size_t p = 0; p < 2; p++)
                    for (size_t i = 0; i < L; i++)
                        for (size_t d = 0; d < N; d++)
                        {
                            A[t][(p * L + i) * N + d] = rng() % mods[i].value();
                            B[t][(p * L + i) * N + d] = rng() % mods[i].value();
                        }
            }
            std::atomic<size_t> next{ 0 };
            std::atomic<int> failed{ 0 };
            auto worker = [&](int t) {
                try
                {
                    auto tables = cd->small_ntt_tables();
                    std::vector<uint64_t> buf(A[t]);
                    for (;;)
                    {
                        size_t i = next.fetch_add(1);
                        if (i >= ops)
                            break;
                        if (what == 0)
                        {
                            auto x = c->make_ct(A[t].data(), 2, L, std::pow(2.0, 20));
                            auto y = c->make_ct(B[t].data(), 2, L, std::pow(2.0, 20));
                            c->evaluator->multiply_inplace(x, y);
                            c->evaluator->relinearize_inplace(x, c->relin);
                            c->evaluator->rescale_to_next_inplace(x);
                        }
                        else if (what == 1)
                        {
                            auto x = c->make_ct(A[t].data(), 2, L, std::pow(2.0, 40));
                            c->evaluator->rotate_vector_inplace(x, 1, c->galois);
                        }
                        else if (what == 2)
                        {
                            ntt_negacyclic_harvey(buf.data() + (i % (2 * L)) * N, tables[(i % (2 * L)) % L]);
                        }
                        else
                        {
                            inverse_ntt_negacyclic_harvey(buf.data() + (i % (2 * L)) * N, tables[(i % (2 * L)) % L]);
                        }
                    }
                }
                catch (...)
                {
                    failed = 1;
                }
            };
            auto t0 = std::chrono::steady_clock::now();
            std::vector<std::thread> th;
            for (int t = 0; t < n_threads; t++)
                th.emplace_back(worker, t);
            for (auto &x : th)
                x.join();
            auto t1 = std::chrono::steady_clock::now();
            if (failed)
                throw std::logic_error("worker failed");
            secs = std::chrono::duration<double>(t1 - t0).count();
        });
        return rc ? static_cast<double>(rc) : secs;
    }
} // extern "C"
