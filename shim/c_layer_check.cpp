// TEST program for shim/gsb_seal_c.cpp: everything the evaluator / decryptor do goes through the C entry
// points of the reference's C layer (Evaluator_*, Decryptor_*, SEALContext_Create from the reference's own
// c/sealcontext.cpp), exactly as dotnet/src would P/Invoke them; results are compared word for word with
// seal::Evaluator / seal::Decryptor on the same objects, and the HRESULT conventions are checked.
// Built by oracle/Makefile (target `shim`); run by tests/test_gpu_shim.py.  Ends with "C_LAYER_CHECK PASS".
#include "seal/c/decryptor.h"
#include "seal/c/defines.h"
#include "seal/c/evaluator.h"
#include "seal/c/sealcontext.h"
#include "seal/c/utilities.h"
#include "seal/seal.h"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>

using namespace seal;
using namespace std;

static int failures = 0;

static void expect(const char *what, bool ok)
{
    printf("%-44s %s\n", what, ok ? "ok" : "FAIL");
    failures += !ok;
}

static bool same(const Ciphertext &r, const Ciphertext &d)
{
    return r.size() == d.size() && r.parms_id() == d.parms_id() && r.is_ntt_form() == d.is_ntt_form() && r.scale() == d.scale() &&
           memcmp(r.data(), d.data(), r.size() * r.coeff_modulus_size() * r.poly_modulus_degree() * sizeof(uint64_t)) == 0;
}

int main()
{
    EncryptionParameters parms(scheme_type::CKKS);
    const size_t N = 8192;
    parms.set_poly_modulus_degree(N);
    parms.set_coeff_modulus(CoeffModulus::Create(N, { 50, 40, 40, 50 }));
    void *ctxh = nullptr;
    expect("SEALContext_Create (reference C layer)", SEALContext_Create(&parms, true, 0, &ctxh) == S_OK && ctxh);
    auto context = seal::c::SharedContextFromVoid(ctxh);
    KeyGenerator keygen(context);
    auto pk = keygen.public_key();
    auto sk = keygen.secret_key();
    RelinKeys rlk = keygen.relin_keys_local();
    GaloisKeys glk = keygen.galois_keys_local();
    Encryptor enc(context, pk);
    Decryptor rdec(context, sk);
    Evaluator ref(context);
    CKKSEncoder encoder(context);
    vector<double> va(encoder.slot_count()), vb(encoder.slot_count());
    mt19937_64 rng(5);
    uniform_real_distribution<double> dist(-1.0, 1.0);
    for (size_t i = 0; i < va.size(); i++)
        va[i] = dist(rng), vb[i] = dist(rng);
    Plaintext pa, pb;
    encoder.encode(va, pow(2.0, 40), pa);
    encoder.encode(vb, pow(2.0, 40), pb);
    Ciphertext a, b;
    enc.encrypt(pa, a);
    enc.encrypt(pb, b);

    void *ev = nullptr, *dec = nullptr;
    expect("Evaluator_Create", Evaluator_Create(ctxh, &ev) == S_OK && ev);
    expect("Decryptor_Create", Decryptor_Create(ctxh, &sk, &dec) == S_OK && dec);
    bool ks = false;
    expect("Evaluator_ContextUsingKeyswitching", Evaluator_ContextUsingKeyswitching(ev, &ks) == S_OK && ks == context->using_keyswitching());

    Ciphertext r, d;
    ref.multiply(a, b, r);
    expect("Evaluator_Multiply (out of place)", Evaluator_Multiply(ev, &a, &b, &d, nullptr) == S_OK && same(r, d));
    ref.relinearize_inplace(r, rlk);
    expect("Evaluator_Relinearize (in place)", Evaluator_Relinearize(ev, &d, &rlk, &d, nullptr) == S_OK && same(r, d));
    ref.rescale_to_next_inplace(r);
    expect("Evaluator_RescaleToNext", Evaluator_RescaleToNext(ev, &d, &d, nullptr) == S_OK && same(r, d));
    Ciphertext r2, d2;
    ref.rotate_vector(r, 3, glk, r2);
    expect("Evaluator_RotateVector(3)", Evaluator_RotateVector(ev, &d, 3, &glk, &d2, nullptr) == S_OK && same(r2, d2));
    ref.complex_conjugate_inplace(r2, glk);
    expect("Evaluator_ComplexConjugate", Evaluator_ComplexConjugate(ev, &d2, &glk, &d2, nullptr) == S_OK && same(r2, d2));
    ref.apply_galois_inplace(r2, 5, glk);
    expect("Evaluator_ApplyGalois(5)", Evaluator_ApplyGalois(ev, &d2, 5, &glk, &d2, nullptr) == S_OK && same(r2, d2));
    {
        Ciphertext x, y, s1, s2;
        ref.add(a, b, x);
        expect("Evaluator_Add", Evaluator_Add(ev, &a, &b, &y) == S_OK && same(x, y));
        ref.sub(a, b, x);
        Ciphertext bb = b;
        expect("Evaluator_Sub (destination = operand 2)", Evaluator_Sub(ev, &a, &bb, &bb) == S_OK && same(x, bb));
        ref.negate(a, x);
        expect("Evaluator_Negate", Evaluator_Negate(ev, &a, &y) == S_OK && same(x, y));
        ref.square(a, s1);
        expect("Evaluator_Square", Evaluator_Square(ev, &a, &s2, nullptr) == S_OK && same(s1, s2));
        ref.multiply_plain(a, pb, x);
        expect("Evaluator_MultiplyPlain", Evaluator_MultiplyPlain(ev, &a, &pb, &y, nullptr) == S_OK && same(x, y));
        ref.mod_switch_to_next(a, x);
        expect("Evaluator_ModSwitchToNext1", Evaluator_ModSwitchToNext1(ev, &a, &y, nullptr) == S_OK && same(x, y));
        ref.transform_from_ntt(a, x);
        expect("Evaluator_TransformFromNTT", Evaluator_TransformFromNTT(ev, &a, &y) == S_OK && same(x, y));
        ref.transform_to_ntt_inplace(x);
        expect("Evaluator_TransformToNTT2", Evaluator_TransformToNTT2(ev, &y, &y) == S_OK && same(x, y));
        auto last = context->last_parms_id();
        uint64_t id[4];
        seal::c::CopyParmsId(last, id);
        Ciphertext m = a, n = a;
        ref.multiply_inplace(m, b);
        Evaluator_Multiply(ev, &n, &b, &n, nullptr);
        ref.rescale_to_inplace(m, last);
        expect("Evaluator_RescaleTo(last)", Evaluator_RescaleTo(ev, &n, id, &n, nullptr) == S_OK && same(m, n));
    }
    {
        Plaintext pr, pd;
        rdec.decrypt(d, pr);
        expect("Decryptor_Decrypt", Decryptor_Decrypt(dec, &d, &pd) == S_OK && pr.coeff_count() == pd.coeff_count() &&
                                        pr.parms_id() == pd.parms_id() && pr.scale() == pd.scale() &&
                                        memcmp(pr.data(), pd.data(), pr.coeff_count() * sizeof(uint64_t)) == 0);
        vector<double> out;
        encoder.decode(pd, out);
        double err = 0;
        for (size_t i = 0; i < out.size(); i++)
            err = max(err, fabs(out[i] - va[i] * vb[i]));
        expect("decode(Decryptor_Decrypt(a*b)) == a*b", err < 1e-5);
    }
    {
        // plaintext operands, add_many and level switches to a parms_id through the C entry points
        Ciphertext x, y;
        ref.add_plain(a, pb, x);
        expect("Evaluator_AddPlain", Evaluator_AddPlain(ev, &a, &pb, &y) == S_OK && same(x, y));
        ref.sub_plain(a, pb, x);
        expect("Evaluator_SubPlain", Evaluator_SubPlain(ev, &a, &pb, &y) == S_OK && same(x, y));
        vector<Ciphertext> many{ a, b, a };
        ref.add_many(many, x);
        void *ptrs[3] = { &many[0], &many[1], &many[2] };
        expect("Evaluator_AddMany", Evaluator_AddMany(ev, 3, ptrs, &y) == S_OK && same(x, y));
        auto last = context->last_parms_id();
        uint64_t id[4];
        seal::c::CopyParmsId(last, id);
        ref.mod_switch_to(a, last, x);
        expect("Evaluator_ModSwitchTo1(last)", Evaluator_ModSwitchTo1(ev, &a, id, &y, nullptr) == S_OK && same(x, y));
        Plaintext px, py;
        ref.mod_switch_to_next(pb, px);
        expect("Evaluator_ModSwitchToNext2 (plain)", Evaluator_ModSwitchToNext2(ev, &pb, &py) == S_OK && px.parms_id() == py.parms_id() &&
                                                       px.coeff_count() == py.coeff_count() && px.scale() == py.scale() &&
                                                       memcmp(px.data(), py.data(), px.coeff_count() * sizeof(uint64_t)) == 0);
        ref.mod_switch_to(pb, last, px);
        expect("Evaluator_ModSwitchTo2 (plain, last)", Evaluator_ModSwitchTo2(ev, &pb, id, &py) == S_OK && px.parms_id() == py.parms_id() &&
                                                         px.coeff_count() == py.coeff_count() &&
                                                         memcmp(px.data(), py.data(), px.coeff_count() * sizeof(uint64_t)) == 0);
        Ciphertext lower = a;
        ref.mod_switch_to_next_inplace(lower);
        uint64_t first_id[4];
        seal::c::CopyParmsId(context->first_parms_id(), first_id);
        expect("ModSwitchTo1 to a higher level -> E_INVALIDARG", Evaluator_ModSwitchTo1(ev, &lower, first_id, &y, nullptr) == E_INVALIDARG);
    }
    // HRESULT conventions (c/defines.h:27-56, c/evaluator.cpp:31-58)
    expect("null thisptr -> E_POINTER", Evaluator_Negate(nullptr, &a, &d) == E_POINTER);
    expect("null operand -> E_POINTER", Evaluator_Multiply(ev, &a, nullptr, &d, nullptr) == E_POINTER);
    {
        Ciphertext lower = a, x;
        ref.mod_switch_to_next_inplace(lower);
        expect("parameter mismatch -> E_INVALIDARG", Evaluator_Add(ev, &a, &lower, &x) == E_INVALIDARG);
        Ciphertext three = a;
        ref.multiply_inplace(three, b);
        expect("rotate size-3 ciphertext -> E_INVALIDARG", Evaluator_RotateVector(ev, &three, 1, &glk, &x, nullptr) == E_INVALIDARG);
        expect("BFV-only op on CKKS -> COR_E_INVALIDOPERATION", Evaluator_RotateRows(ev, &a, 1, &glk, &x, nullptr) == COR_E_INVALIDOPERATION);
        expect("CKKS-only op misuse -> E_INVALIDARG", Evaluator_ModSwitchToNext2(ev, &pa, nullptr) == E_POINTER);
        int budget = 0;
        expect("Decryptor_InvariantNoiseBudget -> unserved", Decryptor_InvariantNoiseBudget(dec, &a, &budget) == COR_E_INVALIDOPERATION);
    }
    expect("Decryptor_Destroy", Decryptor_Destroy(dec) == S_OK);
    expect("Evaluator_Destroy", Evaluator_Destroy(ev) == S_OK);
    expect("SEALContext_Destroy", SEALContext_Destroy(ctxh) == S_OK);
    if (failures)
    {
        printf("C_LAYER_CHECK FAIL (%d)\n", failures);
        return 1;
    }
    printf("C_LAYER_CHECK PASS\n");
    return 0;
}
