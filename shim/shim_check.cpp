// TEST program: the reference's seal::Evaluator (CPU) and gsb_seal::Evaluator (B200) side by side on the
// SAME seal::Ciphertext / RelinKeys / GaloisKeys objects produced by the reference's own KeyGenerator
// and Encryptor; every output ciphertext is compared word for word (plus metadata), then decrypted.
// Built by oracle/Makefile (target `shim`) where /root/reference exists; run on the GPU box by
// tests/test_gpu_shim.py.  Exit code 0 and a final "SHIM_CHECK PASS" line on success.
#include "gsb_seal_shim.h"
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <random>

using namespace seal;
using namespace std;

static int failures = 0;

static void same(const char *what, const Ciphertext &r, const Ciphertext &d)
{
    bool ok = r.size() == d.size() && r.parms_id() == d.parms_id() && r.is_ntt_form() == d.is_ntt_form() && r.scale() == d.scale() &&
              r.coeff_modulus_size() == d.coeff_modulus_size() &&
              memcmp(r.data(), d.data(), r.size() * r.coeff_modulus_size() * r.poly_modulus_degree() * sizeof(uint64_t)) == 0;
    printf("%-34s %s\n", what, ok ? "bit-exact" : "MISMATCH");
    failures += !ok;
}

static void same_plain(const char *what, Decryptor &ref, gsb_seal::Decryptor &dev, const Ciphertext &ct)
{
    Plaintext r, d;
    ref.decrypt(ct, r);
    dev.decrypt(ct, d);
    bool ok = r.coeff_count() == d.coeff_count() && r.parms_id() == d.parms_id() && r.scale() == d.scale() &&
              memcmp(r.data(), d.data(), r.coeff_count() * sizeof(uint64_t)) == 0;
    printf("%-34s %s\n", what, ok ? "bit-exact" : "MISMATCH");
    failures += !ok;
}

template <class F>
static void expect_throw(const char *what, F &&f, int kind) // 0 invalid_argument, 1 logic_error
{
    int got = -1;
    try
    {
        f();
    }
    catch (const invalid_argument &)
    {
        got = 0;
    }
    catch (const logic_error &)
    {
        got = 1;
    }
    catch (...)
    {
        got = 2;
    }
    printf("%-34s %s\n", what, got == kind ? "same exception type" : "WRONG EXCEPTION");
    failures += got != kind;
}

static int run(size_t N, vector<int> bits, int k_special)
{
    printf("---- CKKS N=%zu, %zu primes, n_special_primes=%d\n", N, bits.size(), k_special);
    EncryptionParameters parms(scheme_type::CKKS);
    parms.set_poly_modulus_degree(N);
    parms.set_coeff_modulus(CoeffModulus::Create(N, bits));
    parms.set_n_special_primes(k_special);
    auto context = SEALContext::Create(parms, true, sec_level_type::none);
    KeyGenerator keygen(context);
    auto pk = keygen.public_key();
    auto sk = keygen.secret_key();
    RelinKeys rlk = keygen.relin_keys_local();
    GaloisKeys glk = keygen.galois_keys_local();
    Encryptor enc(context, pk);
    Decryptor dec(context, sk);
    CKKSEncoder encoder(context);
    Evaluator ref(context);
    gsb_seal::Evaluator dev(context);

    const size_t slots = encoder.slot_count();
    mt19937_64 rng(42);
    uniform_real_distribution<double> dist(-1.0, 1.0);
    vector<double> va(slots), vb(slots);
    for (size_t i = 0; i < slots; i++)
        va[i] = dist(rng), vb[i] = dist(rng);
    const double scale = pow(2.0, 40);
    Plaintext pa, pb;
    encoder.encode(va, scale, pa);
    encoder.encode(vb, scale, pb);
    Ciphertext a, b;
    enc.encrypt(pa, a);
    enc.encrypt(pb, b);
    Plaintext pb_ntt = pb; // CKKSEncoder plaintexts are already in NTT form at the first level

    Ciphertext r = a, d = a;
    ref.multiply_inplace(r, b);
    dev.multiply_inplace(d, b);
    same("multiply_inplace", r, d);
    ref.relinearize_inplace(r, rlk);
    dev.relinearize_inplace(d, rlk);
    same("relinearize_inplace", r, d);
    ref.rescale_to_next_inplace(r);
    dev.rescale_to_next_inplace(d);
    same("rescale_to_next_inplace", r, d);
    Ciphertext prod = d;
    {
        // row f4: gsb_seal::Decryptor against seal::Decryptor on the same objects
        gsb_seal::Decryptor ddec(context, sk);
        Ciphertext sq = a;
        ref.multiply_inplace(sq, b); // size 3: needs s^2
        same_plain("Decryptor::decrypt (fresh)", dec, ddec, a);
        same_plain("Decryptor::decrypt (size 3)", dec, ddec, sq);
        same_plain("Decryptor::decrypt (rescaled)", dec, ddec, prod);
        Plaintext bad_sk_plain;
        SecretKey bad = sk;
        bad.data().data()[0] = parms.coeff_modulus()[0].value();
        expect_throw("Decryptor(invalid secret key)", [&] { gsb_seal::Decryptor x(context, bad); }, 0);
    }
    ref.rotate_vector_inplace(r, 1, glk);
    dev.rotate_vector_inplace(d, 1, glk);
    same("rotate_vector_inplace(+1)", r, d);
    ref.rotate_vector_inplace(r, -3, glk); // not a power of two: NAF path, two key switches
    dev.rotate_vector_inplace(d, -3, glk);
    same("rotate_vector_inplace(-3, NAF)", r, d);
    ref.complex_conjugate_inplace(r, glk);
    dev.complex_conjugate_inplace(d, glk);
    same("complex_conjugate_inplace", r, d);
    {
        // row f3: the same chain on device-resident handles -- one upload per operand, one download
        gsb_seal::DeviceCiphertext da, db, dc;
        dev.upload(a, da);
        dev.upload(b, db);
        dev.multiply_inplace(da, db);
        dev.relinearize_inplace(da, rlk);
        dev.rescale_to_next_inplace(da);
        dev.rotate_vector_inplace(da, 1, glk);
        dev.rotate_vector_inplace(da, -3, glk);
        dev.complex_conjugate_inplace(da, glk);
        Ciphertext back;
        dev.download(da, back);
        same("device-resident chain (6 ops)", r, back);
        Ciphertext r5 = a, back5;
        ref.add_inplace(r5, b);
        ref.negate_inplace(r5);
        ref.sub_inplace(r5, a);
        dev.upload(a, dc);
        dev.add_inplace(dc, db);
        dev.negate_inplace(dc);
        gsb_seal::DeviceCiphertext da2;
        dev.upload(a, da2);
        dev.sub_inplace(dc, da2);
        dev.download(dc, back5);
        same("device-resident add / negate / sub", r5, back5);
        {
            // batched handle: three (a_i, b_i) pairs through one batched device pass per operation
            vector<Ciphertext> as(3, a), bs(3, b), want(3), got;
            dev.negate_inplace(as[1]);
            dev.add_inplace(bs[2], a);
            vector<const Ciphertext *> pa2, pb2;
            for (size_t i = 0; i < 3; i++)
            {
                pa2.push_back(&as[i]);
                pb2.push_back(&bs[i]);
                want[i] = as[i];
                ref.multiply_inplace(want[i], bs[i]);
                ref.relinearize_inplace(want[i], rlk);
                ref.rescale_to_next_inplace(want[i]);
                ref.rotate_vector_inplace(want[i], 1, glk);
            }
            gsb_seal::DeviceCiphertext ba, bb;
            dev.upload(pa2, ba);
            dev.upload(pb2, bb);
            dev.multiply_inplace(ba, bb);
            dev.relinearize_inplace(ba, rlk);
            dev.rescale_to_next_inplace(ba);
            dev.rotate_vector_inplace(ba, 1, glk);
            dev.download(ba, got);
            for (size_t i = 0; i < 3; i++)
                same("device-resident batch of 3, element", want[i], got[i]);
        }
        expect_throw("device-resident level mismatch", [&] { dev.add_inplace(da, db); }, 0);
    }
    {
        // general ciphertext sizes (evaluator.cpp:470-520, 772-827): callers that multiply before they relinearize
        Ciphertext a1 = a, b1 = b;
        a1.scale() = b1.scale() = pow(2.0, 8); // scale is metadata; keeps four factors inside the modulus
        Ciphertext r3 = a1, d3 = a1;
        ref.multiply_inplace(r3, b1);
        dev.multiply_inplace(d3, b1);
        Ciphertext r4 = r3, d4 = d3;
        ref.multiply_inplace(r4, b1);
        dev.multiply_inplace(d4, b1);
        same("multiply size 3 x size 2", r4, d4);
        Ciphertext r5 = r3, d5 = d3, r3c = r3;
        r5.scale() = d5.scale() = r3c.scale() = pow(2.0, 4);
        ref.multiply_inplace(r5, r3c);
        dev.multiply_inplace(d5, r3c);
        same("multiply size 3 x size 3", r5, d5);
        Ciphertext r23 = b1, d23 = b1;
        ref.multiply_inplace(r23, r3);
        dev.multiply_inplace(d23, r3);
        same("multiply size 2 x size 3", r23, d23);
        // the public KeyGenerator makes the s^2 key only: size 4 cannot be relinearized, by either evaluator
        expect_throw("relinearize size 4, one key", [&] { Ciphertext t = d4; dev.relinearize_inplace(t, rlk); }, 0);
        // hand-built RelinKeys with keys for s^2, s^3, s^4 (limbs taken from Galois keys: the claim is equal words)
        RelinKeys rl3 = rlk;
        const uint32_t elt1 = context->key_context_data()->galois_tool()->get_elt_from_step(1);
        rl3.data().push_back(glk.data()[GaloisKeys::get_index(elt1)]);
        rl3.data().push_back(glk.data()[GaloisKeys::get_index(static_cast<uint32_t>(2 * N - 1))]);
        Ciphertext r4b = r4, d4b = d4;
        ref.relinearize_inplace(r4b, rl3);
        dev.relinearize_inplace(d4b, rl3);
        same("relinearize size 4 -> 2", r4b, d4b);
        ref.relinearize_inplace(r5, rl3);
        dev.relinearize_inplace(d5, rl3);
        same("relinearize size 5 -> 2", r5, d5);
        gsb_seal::DeviceCiphertext x, y;
        dev.upload(a1, x);
        dev.upload(b1, y);
        dev.multiply_inplace(x, y);
        dev.multiply_inplace(x, y);
        dev.relinearize_inplace(x, rl3);
        Ciphertext back;
        dev.download(x, back);
        same("device-resident size 4 chain", r4b, back);
    }
    {
        // key identity is decided by content, not by address: new key words written into the SAME
        // RelinKeys storage the evaluator has already uploaded from
        RelinKeys rl = rlk;
        Ciphertext r3 = a, d3 = a;
        ref.multiply_inplace(r3, b);
        dev.multiply_inplace(d3, b);
        Ciphertext t1 = d3;
        dev.relinearize_inplace(t1, rl); // uploads rl
        KeyGenerator keygen2(context);
        RelinKeys other = keygen2.relin_keys_local();
        for (size_t j = 0; j < rl.data()[0].size(); j++)
        {
            auto &dst = rl.data()[0][j].data();
            memcpy(dst.data(), other.data()[0][j].data().data(), dst.size() * dst.coeff_modulus_size() * dst.poly_modulus_degree() * sizeof(uint64_t));
        }
        ref.relinearize_inplace(r3, rl);
        dev.relinearize_inplace(d3, rl);
        same("relinearize after in-place key reload", r3, d3);
    }
    {
        // fused device-resident chain and mod_switch_to_next on handles
        gsb_seal::DeviceCiphertext x, y;
        dev.upload(a, x);
        dev.upload(b, y);
        dev.multiply_relinearize_rescale_inplace(x, y, rlk);
        dev.mod_switch_to_next_inplace(x);
        Ciphertext want = a, back;
        ref.multiply_inplace(want, b);
        ref.relinearize_inplace(want, rlk);
        ref.rescale_to_next_inplace(want);
        ref.mod_switch_to_next_inplace(want);
        dev.download(x, back);
        same("handles: fused mul+relin+rescale, mod_switch", want, back);
    }
    {
        Ciphertext r2 = a, d2 = a;
        ref.square_inplace(r2);
        dev.multiply_inplace(d2, d2);
        same("multiply_inplace(x, x) vs square", r2, d2);
    }
    {
        // row f3: thin ops on the same objects (r, d are equal size-2 ciphertexts at the second level)
        Ciphertext r2 = a, d2 = a, r3 = a, d3 = a;
        ref.multiply_inplace(r3, b);
        dev.multiply_inplace(d3, b);
        ref.add_inplace(r2, b);
        dev.add_inplace(d2, b);
        same("add_inplace", r2, d2);
        r3.scale() = d3.scale() = r2.scale(); // scale is metadata only; sub requires equal scales
        ref.sub_inplace(r2, r3);
        dev.sub_inplace(d2, d3);
        same("sub_inplace (size 2 - size 3)", r2, d2);
        ref.negate_inplace(r2);
        dev.negate_inplace(d2);
        same("negate_inplace", r2, d2);
        Ciphertext r4 = a, d4 = a;
        ref.multiply_plain_inplace(r4, pb_ntt);
        dev.multiply_plain_inplace(d4, pb_ntt);
        same("multiply_plain_inplace (NTT)", r4, d4);
        ref.transform_from_ntt_inplace(r4);
        dev.transform_from_ntt_inplace(d4);
        same("transform_from_ntt_inplace", r4, d4);
        ref.transform_to_ntt_inplace(r4);
        dev.transform_to_ntt_inplace(d4);
        same("transform_to_ntt_inplace", r4, d4);
        ref.mod_switch_to_next_inplace(r4);
        dev.mod_switch_to_next_inplace(d4);
        same("mod_switch_to_next_inplace", r4, d4);
        Ciphertext r5 = a, d5 = a;
        ref.square_inplace(r5);
        dev.square_inplace(d5);
        ref.relinearize_inplace(r5, rlk);
        dev.relinearize_inplace(d5, rlk);
        ref.rescale_to_inplace(r5, context->last_parms_id());
        dev.rescale_to_inplace(d5, context->last_parms_id());
        same("square + relin + rescale_to(last)", r5, d5);
    }
    {
        // row f3 remainder: plaintext add / sub, add_many, mod_switch_to (ciphertext and plaintext)
        Ciphertext r2 = a, d2 = a;
        ref.add_plain_inplace(r2, pb_ntt);
        dev.add_plain_inplace(d2, pb_ntt);
        same("add_plain_inplace", r2, d2);
        ref.sub_plain_inplace(r2, pa);
        dev.sub_plain_inplace(d2, pa);
        same("sub_plain_inplace", r2, d2);
        vector<Ciphertext> many{ a, b, r2 };
        Ciphertext rs, ds;
        ref.add_many(many, rs);
        dev.add_many(many, ds);
        same("add_many(3)", rs, ds);
        rs.scale() = ds.scale() = pow(2.0, 20); // metadata only: the last level of the small test modulus has 40 bits
        ref.mod_switch_to_inplace(rs, context->last_parms_id());
        dev.mod_switch_to_inplace(ds, context->last_parms_id());
        same("mod_switch_to_inplace(last)", rs, ds);
        Plaintext p1 = pb, p2 = pb;
        p1.scale() = p2.scale() = pow(2.0, 20);
        ref.mod_switch_to_inplace(p1, context->last_parms_id());
        dev.mod_switch_to_inplace(p2, context->last_parms_id());
        const bool okp = p1.parms_id() == p2.parms_id() && p1.coeff_count() == p2.coeff_count() && p1.scale() == p2.scale() &&
                         memcmp(p1.data(), p2.data(), p1.coeff_count() * sizeof(uint64_t)) == 0;
        printf("%-34s %s\n", "mod_switch_to_inplace(Plaintext)", okp ? "bit-exact" : "MISMATCH");
        failures += !okp;
        expect_throw("add_plain scale mismatch", [&] { Ciphertext t = a; t.scale() = 2.0; dev.add_plain_inplace(t, pb_ntt); }, 0);
    }
    {
        // batched extension vs three reference calls
        vector<Ciphertext> xs(3, a), ys(3, b);
        enc.encrypt(pb, xs[1]);
        enc.encrypt(pa, ys[2]);
        vector<Ciphertext> want = xs;
        for (size_t i = 0; i < 3; i++)
        {
            ref.multiply_inplace(want[i], ys[i]);
            ref.relinearize_inplace(want[i], rlk);
            ref.rescale_to_next_inplace(want[i]);
        }
        vector<Ciphertext *> px{ &xs[0], &xs[1], &xs[2] };
        vector<const Ciphertext *> py{ &ys[0], &ys[1], &ys[2] };
        dev.multiply_relinearize_rescale_inplace(px, py, rlk);
        for (size_t i = 0; i < 3; i++)
            same("batched mul+relin+rescale", want[i], xs[i]);
    }
    {
        // row f2: switching keys generated on the device (reference samplers for the randomness) must work under the
        // REFERENCE's evaluator and decryptor: relinearized product and a rotation decrypt to the right values, and
        // the keys pass the reference's own validity check
        gsb_seal::KeyGenerator dkg(context, sk);
        RelinKeys drlk = dkg.relin_keys_local();
        GaloisKeys dglk = dkg.galois_keys_local(vector<int>{ 1, -2 });
        // same verdicts as for the reference's own keys (this fork's is_metadata_valid_for rejects its own keys when
        // there is more than one special prime), canonical data, same bundle count
        const bool valid = is_valid_for(drlk, context) == is_valid_for(rlk, context) && is_valid_for(dglk, context) == is_valid_for(glk, context) &&
                           is_data_valid_for(drlk, context) && is_data_valid_for(dglk, context) && is_buffer_valid(drlk) && is_buffer_valid(dglk) &&
                           drlk.data()[0].size() == rlk.data()[0].size();
        printf("%-34s %s\n", "device keygen: validity as reference", valid ? "yes" : "NO");
        failures += !valid;
        Ciphertext x = a;
        ref.multiply_inplace(x, b);
        ref.relinearize_inplace(x, drlk);
        ref.rescale_to_next_inplace(x);
        ref.rotate_vector_inplace(x, 1, dglk);
        Plaintext px;
        dec.decrypt(x, px);
        vector<double> ox;
        encoder.decode(px, ox);
        double errk = 0;
        for (size_t i = 0; i < slots; i++)
            errk = max(errk, fabs(ox[i] - va[(i + 1) % slots] * vb[(i + 1) % slots]));
        printf("%-34s max |err| = %.3g %s\n", "device keys: relin + rotate decrypt", errk, errk < 1e-4 ? "ok" : "TOO LARGE");
        failures += !(errk < 1e-4);
    }
    // semantic check: decrypt the device product
    Plaintext pd;
    dec.decrypt(prod, pd);
    vector<double> out;
    encoder.decode(pd, out);
    double err = 0;
    for (size_t i = 0; i < slots; i++)
        err = max(err, fabs(out[i] - va[i] * vb[i]));
    printf("%-34s max |err| = %.3g %s\n", "decrypt(device a*b)", err, err < 1e-4 ? "ok" : "TOO LARGE");
    failures += !(err < 1e-4);

    // util::ntt_* over the context's own NTTTables
    {
        auto tables = context->key_context_data()->small_ntt_tables();
        vector<uint64_t> x(N), y;
        for (auto &v : x)
            v = rng() % tables[0].modulus().value();
        y = x;
        seal::util::ntt_negacyclic_harvey(x.data(), tables[0]);
        gsb_seal::util::ntt_negacyclic_harvey(y.data(), tables[0]);
        bool ok = x == y;
        seal::util::inverse_ntt_negacyclic_harvey_lazy(x.data(), tables[0]);
        gsb_seal::util::inverse_ntt_negacyclic_harvey_lazy(y.data(), tables[0]);
        ok = ok && x == y;
        printf("%-34s %s\n", "util::ntt / inverse lazy", ok ? "bit-exact" : "MISMATCH");
        failures += !ok;
        // the RNSIter and PolyIter overloads of all four families (util/ntt.h:189-368) on a ciphertext's own storage
        auto ct_tables = seal::util::iter(context->first_context_data()->small_ntt_tables());
        const size_t Lc = a.coeff_modulus_size();
        bool ok2 = true;
        for (int kind = 0; kind < 4; kind++)
        {
            Ciphertext p = a, q = a, u = a, v = a; // inputs are canonical residues; lazy outputs are compared as they are
            switch (kind)
            {
            case 0:
                seal::util::inverse_ntt_negacyclic_harvey(seal::util::PolyIter(p), p.size(), ct_tables);
                gsb_seal::util::inverse_ntt_negacyclic_harvey(seal::util::PolyIter(q), q.size(), ct_tables);
                seal::util::inverse_ntt_negacyclic_harvey(*seal::util::PolyIter(u), Lc, ct_tables);
                gsb_seal::util::inverse_ntt_negacyclic_harvey(*seal::util::PolyIter(v), Lc, ct_tables);
                break;
            case 1:
                seal::util::inverse_ntt_negacyclic_harvey_lazy(seal::util::PolyIter(p), p.size(), ct_tables);
                gsb_seal::util::inverse_ntt_negacyclic_harvey_lazy(seal::util::PolyIter(q), q.size(), ct_tables);
                seal::util::inverse_ntt_negacyclic_harvey_lazy(*seal::util::PolyIter(u), Lc, ct_tables);
                gsb_seal::util::inverse_ntt_negacyclic_harvey_lazy(*seal::util::PolyIter(v), Lc, ct_tables);
                break;
            case 2:
                seal::util::ntt_negacyclic_harvey(seal::util::PolyIter(p), p.size(), ct_tables);
                gsb_seal::util::ntt_negacyclic_harvey(seal::util::PolyIter(q), q.size(), ct_tables);
                seal::util::ntt_negacyclic_harvey(*seal::util::PolyIter(u), Lc, ct_tables);
                gsb_seal::util::ntt_negacyclic_harvey(*seal::util::PolyIter(v), Lc, ct_tables);
                break;
            default:
                seal::util::ntt_negacyclic_harvey_lazy(seal::util::PolyIter(p), p.size(), ct_tables);
                gsb_seal::util::ntt_negacyclic_harvey_lazy(seal::util::PolyIter(q), q.size(), ct_tables);
                seal::util::ntt_negacyclic_harvey_lazy(*seal::util::PolyIter(u), Lc, ct_tables);
                gsb_seal::util::ntt_negacyclic_harvey_lazy(*seal::util::PolyIter(v), Lc, ct_tables);
                break;
            }
            const size_t bytes = p.size() * Lc * N * sizeof(uint64_t);
            ok2 = ok2 && memcmp(p.data(), q.data(), bytes) == 0 && memcmp(u.data(), v.data(), bytes) == 0;
        }
        printf("%-34s %s\n", "util::ntt_* RNSIter / PolyIter x 4", ok2 ? "bit-exact" : "MISMATCH");
        failures += !ok2;
    }
    // error convention
    {
        RelinKeys empty;
        Ciphertext c3 = a;
        ref.multiply_inplace(c3, b);
        expect_throw("relinearize with foreign keys", [&] { Ciphertext t = c3; dev.relinearize_inplace(t, empty); }, 0);
        expect_throw("apply_galois even element", [&] { Ciphertext t = a; dev.apply_galois_inplace(t, 4, glk); }, 0);
        expect_throw("apply_galois on size 3", [&] { Ciphertext t = c3; dev.apply_galois_inplace(t, 3, glk); }, 0);
        Plaintext plast;
        encoder.encode(va, context->last_parms_id(), pow(2.0, 15), plast);
        Ciphertext last;
        enc.encrypt(plast, last);
        expect_throw("rescale at last level", [&] { Ciphertext t = last; dev.rescale_to_next_inplace(t); }, 0);
        expect_throw("multiply level mismatch", [&] { Ciphertext t = a; dev.multiply_inplace(t, last); }, 0);
    }
    return failures;
}

// BASELINE config #1: BFV, N = 4096, CoeffModulus::BFVDefault (109 bits), t = 786433 -- what
// native/examples/7_performance.cpp:543-552 runs: multiply, relinearize, rotate rows / columns on real
// encryptions.  (The reference's BFV outputs are not semantically valid at this N, SURVEY F6: the check
// is word-for-word equality with the reference, not decryption.)
static int run_bfv(size_t N)
{
    printf("---- BFV N=%zu, BFVDefault modulus, t=786433\n", N);
    EncryptionParameters parms(scheme_type::BFV);
    parms.set_poly_modulus_degree(N);
    parms.set_coeff_modulus(CoeffModulus::BFVDefault(N));
    parms.set_plain_modulus(786433);
    auto context = SEALContext::Create(parms, true, sec_level_type::none);
    KeyGenerator keygen(context);
    auto pk = keygen.public_key();
    auto sk = keygen.secret_key();
    RelinKeys rlk = keygen.relin_keys_local();
    GaloisKeys glk = keygen.galois_keys_local();
    Encryptor enc(context, pk);
    Decryptor dec(context, sk);
    gsb_seal::Decryptor ddec(context, sk);
    BatchEncoder encoder(context);
    Evaluator ref(context);
    gsb_seal::Evaluator dev(context);
    vector<uint64_t> va(encoder.slot_count()), vb(encoder.slot_count());
    mt19937_64 rng(7);
    for (size_t i = 0; i < va.size(); i++)
        va[i] = rng() % 786433, vb[i] = rng() % 786433;
    Plaintext pa, pb;
    encoder.encode(va, pa);
    encoder.encode(vb, pb);
    Ciphertext a, b;
    enc.encrypt(pa, a);
    enc.encrypt(pb, b);
    Ciphertext r = a, d = a;
    ref.multiply_inplace(r, b);
    dev.multiply_inplace(d, b);
    same("BFV multiply_inplace", r, d);
    same_plain("BFV Decryptor::decrypt (fresh)", dec, ddec, a);
    same_plain("BFV Decryptor::decrypt (size 3)", dec, ddec, d);
    {
        Plaintext p;
        ddec.decrypt(a, p);
        vector<uint64_t> back;
        encoder.decode(p, back);
        const bool ok = back == va;
        printf("%-34s %s\n", "BFV decrypt(encrypt(m)) == m", ok ? "yes" : "NO");
        failures += !ok;
    }
    ref.relinearize_inplace(r, rlk);
    dev.relinearize_inplace(d, rlk);
    same("BFV relinearize_inplace", r, d);
    ref.rotate_rows_inplace(r, 1, glk);
    dev.rotate_rows_inplace(d, 1, glk);
    same("BFV rotate_rows_inplace(+1)", r, d);
    ref.rotate_rows_inplace(r, -5, glk); // NAF path
    dev.rotate_rows_inplace(d, -5, glk);
    same("BFV rotate_rows_inplace(-5, NAF)", r, d);
    ref.rotate_columns_inplace(r, glk);
    dev.rotate_columns_inplace(d, glk);
    same("BFV rotate_columns_inplace", r, d);
    {
        Ciphertext r2 = a, d2 = a;
        ref.square_inplace(r2);
        dev.multiply_inplace(d2, d2);
        same("BFV multiply_inplace(x, x) vs square", r2, d2);
    }
    {
        Ciphertext r2 = a, d2 = a;
        ref.multiply_inplace(r2, b);
        dev.multiply_inplace(d2, b);
        ref.mod_switch_to_next_inplace(r2); // size 3
        dev.mod_switch_to_next_inplace(d2);
        same("BFV mod_switch_to_next_inplace", r2, d2);
        Ciphertext r3 = a, d3 = a;
        ref.add_inplace(r3, b);
        dev.add_inplace(d3, b);
        ref.sub_inplace(r3, a);
        dev.sub_inplace(d3, a);
        ref.negate_inplace(r3);
        dev.negate_inplace(d3);
        same("BFV add / sub / negate", r3, d3);
    }
    {
        // multiply_many (queue of 5 -> waves of 2, 1, 1 products) and exponentiate (x^3, x^4)
        Ciphertext c;
        ref.add(a, b, c);
        vector<Ciphertext> v{ a, b, c, a, b };
        Ciphertext r4, d4;
        ref.multiply_many(v, rlk, r4);
        dev.multiply_many(v, rlk, d4);
        same("BFV multiply_many(5)", r4, d4);
        v.resize(2);
        ref.multiply_many(v, rlk, r4);
        dev.multiply_many(v, rlk, d4);
        same("BFV multiply_many(2)", r4, d4);
        for (uint64_t e : { 3, 4 })
        {
            Ciphertext r5 = a, d5 = a;
            ref.exponentiate_inplace(r5, e, rlk);
            dev.exponentiate_inplace(d5, e, rlk);
            same(e == 3 ? "BFV exponentiate_inplace(3)" : "BFV exponentiate_inplace(4)", r5, d5);
        }
        expect_throw("multiply_many(empty)", [&] { vector<Ciphertext> e; Ciphertext t; dev.multiply_many(e, rlk, t); }, 0);
        expect_throw("exponentiate(0)", [&] { Ciphertext t = a; dev.exponentiate_inplace(t, 0, rlk); }, 0);
    }
    {
        // general sizes and device-resident handles, BFV
        Ciphertext r3 = a, d3 = a;
        ref.multiply_inplace(r3, b);
        dev.multiply_inplace(d3, b);
        Ciphertext r4 = r3, d4 = d3;
        ref.multiply_inplace(r4, b);
        dev.multiply_inplace(d4, b);
        same("BFV multiply size 3 x size 2", r4, d4);
        RelinKeys rl2 = rlk;
        const uint32_t elt1 = context->key_context_data()->galois_tool()->get_elt_from_step(1);
        rl2.data().push_back(glk.data()[GaloisKeys::get_index(elt1)]);
        ref.relinearize_inplace(r4, rl2);
        dev.relinearize_inplace(d4, rl2);
        same("BFV relinearize size 4 -> 2", r4, d4);
        gsb_seal::DeviceCiphertext x, y;
        dev.upload(a, x);
        dev.upload(b, y);
        dev.multiply_inplace(x, y);
        dev.relinearize_inplace(x, rlk);
        dev.mod_switch_to_next_inplace(x);
        dev.apply_galois_inplace(x, elt1, glk);
        dev.add_inplace(x, x);
        Ciphertext want = a, back;
        ref.multiply_inplace(want, b);
        ref.relinearize_inplace(want, rlk);
        ref.mod_switch_to_next_inplace(want);
        ref.apply_galois_inplace(want, elt1, glk);
        ref.add_inplace(want, want);
        dev.download(x, back);
        same("BFV device-resident chain (5 ops)", want, back);
    }
    {
        // row f3 remainder, BFV: scaling-variant plaintext add / sub, coefficient-form multiply_plain (generic and
        // monomial branches), plaintext transform_to_ntt, NTT-form multiply_plain
        Ciphertext r2 = a, d2 = a;
        ref.add_plain_inplace(r2, pb);
        dev.add_plain_inplace(d2, pb);
        same("BFV add_plain_inplace", r2, d2);
        ref.sub_plain_inplace(r2, pa);
        dev.sub_plain_inplace(d2, pa);
        same("BFV sub_plain_inplace", r2, d2);
        ref.multiply_plain_inplace(r2, pb);
        dev.multiply_plain_inplace(d2, pb);
        same("BFV multiply_plain (generic)", r2, d2);
        Plaintext mono_small("5x^3"), mono_big;
        mono_big.resize(8);
        mono_big.set_zero();
        mono_big[7] = 786432; // t - 1: above plain_upper_half_threshold
        ref.multiply_plain_inplace(r2, mono_small);
        dev.multiply_plain_inplace(d2, mono_small);
        same("BFV multiply_plain (monomial)", r2, d2);
        ref.multiply_plain_inplace(r2, mono_big);
        dev.multiply_plain_inplace(d2, mono_big);
        same("BFV multiply_plain (monomial, upper half)", r2, d2);
        Plaintext t1 = pb, t2 = pb;
        ref.transform_to_ntt_inplace(t1, context->first_parms_id());
        dev.transform_to_ntt_inplace(t2, context->first_parms_id());
        const bool okp = t1.parms_id() == t2.parms_id() && t1.coeff_count() == t2.coeff_count() &&
                         memcmp(t1.data(), t2.data(), t1.coeff_count() * sizeof(uint64_t)) == 0;
        printf("%-34s %s\n", "BFV transform_to_ntt(Plaintext)", okp ? "bit-exact" : "MISMATCH");
        failures += !okp;
        Ciphertext r3 = a, d3 = a;
        ref.transform_to_ntt_inplace(r3);
        dev.transform_to_ntt_inplace(d3);
        ref.multiply_plain_inplace(r3, t1);
        dev.multiply_plain_inplace(d3, t2);
        same("BFV multiply_plain (NTT form)", r3, d3);
        vector<Ciphertext> many{ a, b, a, b };
        Ciphertext rs, ds;
        ref.add_many(many, rs);
        dev.add_many(many, ds);
        same("BFV add_many(4)", rs, ds);
        ref.mod_switch_to_inplace(rs, context->last_parms_id());
        dev.mod_switch_to_inplace(ds, context->last_parms_id());
        same("BFV mod_switch_to_inplace(last)", rs, ds);
    }
    expect_throw("BFV rotate_vector (CKKS only)", [&] { Ciphertext t = a; dev.rotate_vector_inplace(t, 1, glk); }, 1);
    return failures;
}

int main()
{
    try
    {
        run(8192, { 50, 40, 40, 40, 50 }, 1);
        run(4096, { 40, 30, 30, 30, 40, 40 }, 2);
        run_bfv(4096);
        run_bfv(8192);
    }
    catch (const exception &e)
    {
        printf("exception: %s\n", e.what());
        return 2;
    }
    if (failures)
    {
        printf("SHIM_CHECK FAIL (%d)\n", failures);
        return 1;
    }
    printf("SHIM_CHECK PASS\n");
    return 0;
}
