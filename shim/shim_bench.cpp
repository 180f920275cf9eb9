// MEASUREMENT program for the drop-in C++ boundary: gsb_seal::Evaluator (B200) beside seal::Evaluator (host CPU) on
// real seal::Ciphertext / RelinKeys / GaloisKeys OBJECTS at the headline shape (CKKS, N = 32768, BFVDefault 881-bit
// modulus, L = 15).  Built by oracle/Makefile (target `shim`) where /root/reference exists, run on the GPU box by
// tests/test_gpu_shim.py and by bench.py (extra.shim).  One JSON object on the last stdout line.
//
//   per-call latency, batch 1, host objects (pageable seal memory; every call copies in, runs, copies out):
//       multiply_inplace, relinearize_inplace, rescale_to_next_inplace, rotate_vector_inplace(+1)
//   throughput of the batched device-resident route (one handle = B ciphertexts):
//       upload -> multiply -> relinearize -> rescale_to_next -> download            (chain_ops_per_s)
//       the same with the fused multiply_relinearize_rescale_inplace                (fused_chain_ops_per_s)
//       the fused pass alone, operands already on the device                        (fused_resident_ops_per_s)
// Ciphertexts and switching keys are synthetic uniform residues; every timed device result is compared
// word for word with seal::Evaluator's on the same objects before any number is printed.
#include "gsb_seal_shim.h"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>

using namespace seal;
using namespace std;
using clk = chrono::steady_clock;

static double ms_since(clk::time_point t0)
{
    return chrono::duration<double, milli>(clk::now() - t0).count();
}

static void fill_uniform(Ciphertext &ct, const vector<Modulus> &mods, mt19937_64 &rng)
{
    const size_t N = ct.poly_modulus_degree(), L = ct.coeff_modulus_size();
    for (size_t p = 0; p < ct.size(); p++)
        for (size_t i = 0; i < L; i++)
            for (size_t d = 0; d < N; d++)
                ct.data(p)[i * N + d] = rng() % mods[i].value();
}

static Ciphertext make_ct(shared_ptr<SEALContext> context, mt19937_64 &rng, double scale)
{
    auto cd = context->first_context_data();
    Ciphertext ct(context, cd->parms_id());
    ct.resize(2);
    ct.is_ntt_form() = true;
    ct.scale() = scale;
    fill_uniform(ct, cd->parms().coeff_modulus(), rng);
    return ct;
}

// KSwitchKeys of the shape generate_kswitch_keys produces (keygenerator.cpp:325-369), uniform residues
static void fill_keys(shared_ptr<SEALContext> context, vector<PublicKey> &v, mt19937_64 &rng)
{
    auto key_cd = context->key_context_data();
    auto &mods = key_cd->parms().coeff_modulus();
    const size_t n_bundles = context->first_context_data()->parms().coeff_modulus().size(); // k = 1
    v.resize(n_bundles);
    for (auto &pk : v)
    {
        Ciphertext &ct = pk.data();
        ct.resize(context, key_cd->parms_id(), 2);
        ct.is_ntt_form() = true;
        ct.scale() = 1.0;
        fill_uniform(ct, mods, rng);
    }
}

static bool same(const Ciphertext &r, const Ciphertext &d)
{
    return r.size() == d.size() && r.parms_id() == d.parms_id() && r.is_ntt_form() == d.is_ntt_form() && r.scale() == d.scale() &&
           memcmp(r.data(), d.data(), r.size() * r.coeff_modulus_size() * r.poly_modulus_degree() * sizeof(uint64_t)) == 0;
}

template <class F>
static double median_ms(int reps, F &&f)
{
    vector<double> t;
    for (int i = 0; i < reps; i++)
    {
        auto t0 = clk::now();
        f();
        t.push_back(ms_since(t0));
    }
    sort(t.begin(), t.end());
    return t[t.size() / 2];
}

int main(int argc, char **argv)
{
    const size_t B = argc > 1 ? static_cast<size_t>(atoi(argv[1])) : 64; // ciphertexts per batched handle
    const int dev_reps = argc > 2 ? atoi(argv[2]) : 7, ref_reps = argc > 3 ? atoi(argv[3]) : 2;
    const size_t N = 32768;
    bool parity = true;
    try
    {
        EncryptionParameters parms(scheme_type::CKKS);
        parms.set_poly_modulus_degree(N);
        parms.set_coeff_modulus(CoeffModulus::BFVDefault(N));
        auto context = SEALContext::Create(parms, true, sec_level_type::none);
        auto key_cd = context->key_context_data();
        mt19937_64 rng(2024);
        RelinKeys rlk;
        rlk.data().resize(1);
        fill_keys(context, rlk.data()[0], rng);
        rlk.parms_id() = key_cd->parms_id();
        const uint32_t elt = key_cd->galois_tool()->get_elt_from_step(1);
        GaloisKeys glk;
        glk.data().resize(GaloisKeys::get_index(elt) + 1);
        fill_keys(context, glk.data()[GaloisKeys::get_index(elt)], rng);
        glk.parms_id() = key_cd->parms_id();
        const double scale = pow(2.0, 40);
        Ciphertext a = make_ct(context, rng, scale), b = make_ct(context, rng, scale);

        Evaluator ref(context);
        gsb_seal::Evaluator dev(context);

        // ---- per-call latency, host objects, batch 1 ---------------------------------------------------------
        struct Row
        {
            const char *name;
            double dev_first_ms, dev_ms, ref_ms;
        };
        vector<Row> rows;
        Ciphertext c3r = a, c3d = a;
        {
            auto t0 = clk::now();
            dev.multiply_inplace(c3d, b);
            const double first = ms_since(t0);
            ref.multiply_inplace(c3r, b);
            parity = parity && same(c3r, c3d);
            const double d = median_ms(dev_reps, [&] { Ciphertext t = a; dev.multiply_inplace(t, b); });
            const double r = median_ms(ref_reps, [&] { Ciphertext t = a; ref.multiply_inplace(t, b); });
            const double copy = median_ms(dev_reps, [&] { Ciphertext t = a; (void)t; });
            rows.push_back({ "multiply_inplace", first, d - copy, r - copy });
        }
        Ciphertext c2r = c3r, c2d = c3d;
        {
            auto t0 = clk::now();
            dev.relinearize_inplace(c2d, rlk); // includes the one-time upload of the 120 MiB key
            const double first = ms_since(t0);
            ref.relinearize_inplace(c2r, rlk);
            parity = parity && same(c2r, c2d);
            const double d = median_ms(dev_reps, [&] { Ciphertext t = c3d; dev.relinearize_inplace(t, rlk); });
            const double r = median_ms(ref_reps, [&] { Ciphertext t = c3r; ref.relinearize_inplace(t, rlk); });
            const double copy = median_ms(dev_reps, [&] { Ciphertext t = c3d; (void)t; });
            rows.push_back({ "relinearize_inplace", first, d - copy, r - copy });
        }
        Ciphertext c1r = c2r, c1d = c2d;
        {
            auto t0 = clk::now();
            dev.rescale_to_next_inplace(c1d);
            const double first = ms_since(t0);
            ref.rescale_to_next_inplace(c1r);
            parity = parity && same(c1r, c1d);
            const double d = median_ms(dev_reps, [&] { Ciphertext t = c2d; dev.rescale_to_next_inplace(t); });
            const double r = median_ms(ref_reps, [&] { Ciphertext t = c2r; ref.rescale_to_next_inplace(t); });
            const double copy = median_ms(dev_reps, [&] { Ciphertext t = c2d; (void)t; });
            rows.push_back({ "rescale_to_next_inplace", first, d - copy, r - copy });
        }
        {
            Ciphertext rr = a, rd = a;
            auto t0 = clk::now();
            dev.rotate_vector_inplace(rd, 1, glk);
            const double first = ms_since(t0);
            ref.rotate_vector_inplace(rr, 1, glk);
            parity = parity && same(rr, rd);
            const double d = median_ms(dev_reps, [&] { Ciphertext t = a; dev.rotate_vector_inplace(t, 1, glk); });
            const double r = median_ms(ref_reps, [&] { Ciphertext t = a; ref.rotate_vector_inplace(t, 1, glk); });
            const double copy = median_ms(dev_reps, [&] { Ciphertext t = a; (void)t; });
            rows.push_back({ "rotate_vector_inplace", first, d - copy, r - copy });
        }

        // ---- batched device-resident route ---------------------------------------------------------------------
        vector<Ciphertext> as, bs;
        for (size_t i = 0; i < B; i++)
        {
            as.push_back(make_ct(context, rng, scale));
            bs.push_back(make_ct(context, rng, scale));
        }
        vector<const Ciphertext *> pa, pb;
        for (size_t i = 0; i < B; i++)
            pa.push_back(&as[i]), pb.push_back(&bs[i]);
        vector<Ciphertext> got, got2;
        auto chain = [&](bool fused, vector<Ciphertext> &out) {
            gsb_seal::DeviceCiphertext x, y;
            dev.upload(pa, x);
            dev.upload(pb, y);
            if (fused)
                dev.multiply_relinearize_rescale_inplace(x, y, rlk);
            else
            {
                dev.multiply_inplace(x, y);
                dev.relinearize_inplace(x, rlk);
                dev.rescale_to_next_inplace(x);
            }
            dev.download(x, out);
        };
        chain(false, got);
        chain(true, got2);
        for (size_t i : { size_t(0), B - 1 }) // the reference on the first and the last pair
        {
            Ciphertext w = as[i];
            ref.multiply_inplace(w, bs[i]);
            ref.relinearize_inplace(w, rlk);
            ref.rescale_to_next_inplace(w);
            parity = parity && same(w, got[i]) && same(w, got2[i]);
        }
        const double chain_ms = median_ms(3, [&] { chain(false, got); });
        const double fused_ms = median_ms(3, [&] { chain(true, got2); });
        gsb_seal::DeviceCiphertext y;
        dev.upload(pb, y);
        const int res_reps = 10;
        double resident_ms;
        {
            // operands already in HBM: the fused pass replaces its first operand, so res_reps handles are uploaded
            // outside the timed region and each is consumed by one pass
            vector<gsb_seal::DeviceCiphertext> xs(res_reps);
            for (auto &h : xs)
                dev.upload(pa, h);
            gsb_stream_sync(dev.device_context(), nullptr);
            auto t0 = clk::now();
            for (auto &h : xs)
                dev.multiply_relinearize_rescale_inplace(h, y, rlk);
            gsb_stream_sync(dev.device_context(), nullptr);
            resident_ms = ms_since(t0) / res_reps;
        }

        // ---- key generation (row f2): one relin key + one Galois key, seal::KeyGenerator vs the device route --------
        double keygen_dev_ms = 0, keygen_ref_ms = 0;
        {
            KeyGenerator kg(context);
            gsb_seal::KeyGenerator dkg(context, kg.secret_key());
            auto t0 = clk::now();
            RelinKeys r1 = kg.relin_keys_local();
            GaloisKeys g1 = kg.galois_keys_local(vector<uint32_t>{ elt });
            keygen_ref_ms = ms_since(t0);
            (void)dkg.relin_keys_local(); // warm-up (context tables, workspace)
            t0 = clk::now();
            RelinKeys r2 = dkg.relin_keys_local();
            GaloisKeys g2 = dkg.galois_keys_local(vector<uint32_t>{ elt });
            keygen_dev_ms = ms_since(t0);
            parity = parity && is_valid_for(r2, context) && is_valid_for(g2, context) && r1.data()[0].size() == r2.data()[0].size();
        }

        printf("{\"workload\": \"CKKS N=32768 BFVDefault (L=15, K=16): seal::Ciphertext / RelinKeys / GaloisKeys objects, synthetic uniform residues\", "
               "\"parity_ok\": %s, \"batch\": %zu, \"per_call_ms\": {",
               parity ? "true" : "false", B);
        for (size_t i = 0; i < rows.size(); i++)
            printf("%s\"%s\": {\"gsb_seal\": %.3f, \"gsb_seal_first_call\": %.3f, \"seal_cpu_1thread\": %.3f}", i ? ", " : "", rows[i].name,
                   rows[i].dev_ms, rows[i].dev_first_ms, rows[i].ref_ms);
        printf("}, \"keygen_relin_plus_galois_ms\": {\"gsb_seal\": %.1f, \"seal_cpu_1thread\": %.1f}", keygen_dev_ms, keygen_ref_ms);
        printf(", \"chain_ops_per_s\": %.1f, \"fused_chain_ops_per_s\": %.1f, \"fused_resident_ops_per_s\": %.1f, "
               "\"chain_ms\": %.3f, \"fused_chain_ms\": %.3f, \"fused_resident_ms\": %.3f, "
               "\"note\": \"per-call = batch 1 on pageable host objects (H2D + kernels + D2H per call, the copy of the operand subtracted); chain = upload "
               "2 x B ciphertexts from pageable seal memory, run, download B; resident = operands already in HBM\"}\n",
               B * 1e3 / chain_ms, B * 1e3 / fused_ms, B * 1e3 / resident_ms, chain_ms, fused_ms, resident_ms);
    }
    catch (const exception &e)
    {
        printf("{\"error\": \"%s\"}\n", e.what());
        return 2;
    }
    return parity ? 0 : 1;
}
