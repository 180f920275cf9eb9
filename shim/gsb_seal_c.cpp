// C export layer pass-through: the `Evaluator_*` and `Decryptor_*` entry points of
// the reference's C layer (native/src/seal/c/evaluator.h:16-79, c/decryptor.h:16-22 -- what
// dotnet/src/Evaluator.cs and Decryptor.cs P/Invoke) with the SAME names, parameter lists and HRESULT
// conventions (c/defines.h:27-56), served by gsb_seal::Evaluator / gsb_seal::Decryptor, i.e. by the
// B200 library.  A maintainer builds the C layer with this file in place of c/evaluator.cpp and
// c/decryptor.cpp; every other file of the C layer (contexts, keys, ciphertexts, memory pools) stays as
// it is, so the handles a caller already holds keep their meaning:
//   thisptr of Evaluator_* / Decryptor_*  -> gsb_seal::Evaluator / gsb_seal::Decryptor (created here)
//   every other void*                      -> the reference's own object (seal::Ciphertext, RelinKeys, ...)
// Entry points of the reference whose operation is outside the path this library serves report
// COR_E_INVALIDOPERATION (what a std::logic_error maps to) -- never a CPU computation.
#include "gsb_seal_shim.h"
#include "seal/c/defines.h"
#include "seal/c/utilities.h"
#include <new>
#include <stdexcept>

using seal::Ciphertext;
using seal::GaloisKeys;
using seal::Plaintext;
using seal::RelinKeys;
using seal::c::FromVoid;

namespace
{
    // Exception -> HRESULT, the mapping every function of the reference's c/evaluator.cpp spells out
    // (invalid_argument -> E_INVALIDARG, logic_error -> COR_E_INVALIDOPERATION); device failures have no
    // counterpart there and surface as E_UNEXPECTED.
    template <class Fn>
    HRESULT run(Fn &&fn)
    {
        try
        {
            fn();
            return S_OK;
        }
        catch (const std::invalid_argument &)
        {
            return E_INVALIDARG;
        }
        catch (const std::out_of_range &)
        {
            return HRESULT_FROM_WIN32(ERROR_INVALID_INDEX);
        }
        catch (const std::logic_error &)
        {
            return COR_E_INVALIDOPERATION;
        }
        catch (const std::bad_alloc &)
        {
            return E_OUTOFMEMORY;
        }
        catch (const std::exception &)
        {
            return E_UNEXPECTED;
        }
    }

    inline bool any_null()
    {
        return false;
    }
    template <class... Rest>
    inline bool any_null(const void *p, Rest... rest)
    {
        return p == nullptr || any_null(rest...);
    }

    inline gsb_seal::Evaluator *ev(void *p)
    {
        return FromVoid<gsb_seal::Evaluator>(p);
    }
    inline Ciphertext &ct(void *p)
    {
        return *FromVoid<Ciphertext>(p);
    }

    // The reference's out-of-place forms are "destination = encrypted; op_inplace(destination)"
    // (evaluator.h, every *_inplace / out-of-place pair): same here, without the copy when they alias.
    template <class Op>
    HRESULT unary(void *thisptr, void *encrypted, void *destination, Op &&op)
    {
        if (any_null(thisptr, encrypted, destination))
            return E_POINTER;
        return run([&] {
            if (destination != encrypted)
                ct(destination) = ct(encrypted);
            op(*ev(thisptr), ct(destination));
        });
    }

    template <class Op>
    HRESULT binary(void *thisptr, void *encrypted1, void *encrypted2, void *destination, Op &&op)
    {
        if (any_null(thisptr, encrypted1, encrypted2, destination))
            return E_POINTER;
        return run([&] {
            if (destination == encrypted2 && destination != encrypted1)
            {
                Ciphertext tmp = ct(encrypted1); // destination aliases the second operand
                op(*ev(thisptr), tmp, ct(encrypted2));
                ct(destination) = std::move(tmp);
                return;
            }
            if (destination != encrypted1)
                ct(destination) = ct(encrypted1);
            op(*ev(thisptr), ct(destination), ct(encrypted2));
        });
    }

    inline HRESULT not_served(const void *thisptr)
    {
        return thisptr == nullptr ? E_POINTER : COR_E_INVALIDOPERATION;
    }
} // namespace

// ---- Evaluator (c/evaluator.h:16-79) ------------------------------------------------------------------
SEAL_C_FUNC Evaluator_Create(void *sealContext, void **evaluator)
{
    if (any_null(sealContext, evaluator))
        return E_POINTER;
    const auto &shared = seal::c::SharedContextFromVoid(sealContext);
    if (!shared)
        return E_POINTER;
    return run([&] { *evaluator = new gsb_seal::Evaluator(shared); });
}

SEAL_C_FUNC Evaluator_Destroy(void *thisptr)
{
    if (any_null(thisptr))
        return E_POINTER;
    delete ev(thisptr);
    return S_OK;
}

SEAL_C_FUNC Evaluator_Negate(void *thisptr, void *encrypted, void *destination)
{
    return unary(thisptr, encrypted, destination, [](gsb_seal::Evaluator &e, Ciphertext &x) { e.negate_inplace(x); });
}

SEAL_C_FUNC Evaluator_Add(void *thisptr, void *encrypted1, void *encrypted2, void *destination)
{
    return binary(thisptr, encrypted1, encrypted2, destination,
                  [](gsb_seal::Evaluator &e, Ciphertext &x, const Ciphertext &y) { e.add_inplace(x, y); });
}

SEAL_C_FUNC Evaluator_Sub(void *thisptr, void *encrypted1, void *encrypted2, void *destination)
{
    return binary(thisptr, encrypted1, encrypted2, destination,
                  [](gsb_seal::Evaluator &e, Ciphertext &x, const Ciphertext &y) { e.sub_inplace(x, y); });
}

SEAL_C_FUNC Evaluator_Multiply(void *thisptr, void *encrypted1, void *encrypted2, void *destination, void *)
{
    return binary(thisptr, encrypted1, encrypted2, destination,
                  [](gsb_seal::Evaluator &e, Ciphertext &x, const Ciphertext &y) { e.multiply_inplace(x, y); });
}

SEAL_C_FUNC Evaluator_Square(void *thisptr, void *encrypted, void *destination, void *)
{
    return unary(thisptr, encrypted, destination, [](gsb_seal::Evaluator &e, Ciphertext &x) { e.square_inplace(x); });
}

SEAL_C_FUNC Evaluator_Relinearize(void *thisptr, void *encrypted, void *relinKeys, void *destination, void *)
{
    if (any_null(relinKeys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination,
                 [&](gsb_seal::Evaluator &e, Ciphertext &x) { e.relinearize_inplace(x, *FromVoid<RelinKeys>(relinKeys)); });
}

SEAL_C_FUNC Evaluator_MultiplyPlain(void *thisptr, void *encrypted, void *plain, void *destination, void *)
{
    if (any_null(plain))
        return E_POINTER;
    return unary(thisptr, encrypted, destination,
                 [&](gsb_seal::Evaluator &e, Ciphertext &x) { e.multiply_plain_inplace(x, *FromVoid<Plaintext>(plain)); });
}

SEAL_C_FUNC Evaluator_MultiplyMany(void *thisptr, uint64_t count, void **encrypteds, void *relin_keys, void *destination, void *)
{
    if (any_null(thisptr, encrypteds, relin_keys, destination))
        return E_POINTER;
    return run([&] {
        std::vector<Ciphertext> all;
        all.reserve(count);
        for (uint64_t i = 0; i < count; i++)
        {
            if (encrypteds[i] == nullptr)
                throw std::invalid_argument("encrypteds");
            all.push_back(ct(encrypteds[i]));
        }
        ev(thisptr)->multiply_many(all, *FromVoid<RelinKeys>(relin_keys), ct(destination));
    });
}

SEAL_C_FUNC Evaluator_Exponentiate(void *thisptr, void *encrypted, uint64_t exponent, void *relin_keys, void *destination, void *)
{
    if (any_null(relin_keys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) {
        e.exponentiate_inplace(x, exponent, *FromVoid<RelinKeys>(relin_keys));
    });
}

SEAL_C_FUNC Evaluator_ModSwitchToNext1(void *thisptr, void *encrypted, void *destination, void *)
{
    return unary(thisptr, encrypted, destination, [](gsb_seal::Evaluator &e, Ciphertext &x) { e.mod_switch_to_next_inplace(x); });
}

SEAL_C_FUNC Evaluator_RescaleToNext(void *thisptr, void *encrypted, void *destination, void *)
{
    return unary(thisptr, encrypted, destination, [](gsb_seal::Evaluator &e, Ciphertext &x) { e.rescale_to_next_inplace(x); });
}

SEAL_C_FUNC Evaluator_RescaleTo(void *thisptr, void *encrypted, uint64_t *parms_id, void *destination, void *)
{
    if (any_null(parms_id))
        return E_POINTER;
    seal::parms_id_type id;
    seal::c::CopyParmsId(parms_id, id);
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) { e.rescale_to_inplace(x, id); });
}

SEAL_C_FUNC Evaluator_TransformToNTT2(void *thisptr, void *encrypted, void *destination_ntt)
{
    return unary(thisptr, encrypted, destination_ntt, [](gsb_seal::Evaluator &e, Ciphertext &x) { e.transform_to_ntt_inplace(x); });
}

SEAL_C_FUNC Evaluator_TransformFromNTT(void *thisptr, void *encrypted_ntt, void *destination)
{
    return unary(thisptr, encrypted_ntt, destination, [](gsb_seal::Evaluator &e, Ciphertext &x) { e.transform_from_ntt_inplace(x); });
}

SEAL_C_FUNC Evaluator_ApplyGalois(void *thisptr, void *encrypted, uint32_t galois_elt, void *galois_keys, void *destination, void *)
{
    if (any_null(galois_keys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) {
        e.apply_galois_inplace(x, galois_elt, *FromVoid<GaloisKeys>(galois_keys));
    });
}

SEAL_C_FUNC Evaluator_RotateRows(void *thisptr, void *encrypted, int steps, void *galoisKeys, void *destination, void *)
{
    if (any_null(galoisKeys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) {
        e.rotate_rows_inplace(x, steps, *FromVoid<GaloisKeys>(galoisKeys));
    });
}

SEAL_C_FUNC Evaluator_RotateColumns(void *thisptr, void *encrypted, void *galois_keys, void *destination, void *)
{
    if (any_null(galois_keys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) {
        e.rotate_columns_inplace(x, *FromVoid<GaloisKeys>(galois_keys));
    });
}

SEAL_C_FUNC Evaluator_RotateVector(void *thisptr, void *encrypted, int steps, void *galois_keys, void *destination, void *)
{
    if (any_null(galois_keys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) {
        e.rotate_vector_inplace(x, steps, *FromVoid<GaloisKeys>(galois_keys));
    });
}

SEAL_C_FUNC Evaluator_ComplexConjugate(void *thisptr, void *encrypted, void *galois_keys, void *destination, void *)
{
    if (any_null(galois_keys))
        return E_POINTER;
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) {
        e.complex_conjugate_inplace(x, *FromVoid<GaloisKeys>(galois_keys));
    });
}

SEAL_C_FUNC Evaluator_ContextUsingKeyswitching(void *thisptr, bool *using_keyswitching)
{
    if (any_null(thisptr, using_keyswitching))
        return E_POINTER;
    *using_keyswitching = ev(thisptr)->using_keyswitching();
    return S_OK;
}

// Plaintext operands, add_many and level switches to a given parms_id (c/evaluator.cpp:110-212, 387-459, 542-574)
SEAL_C_FUNC Evaluator_AddMany(void *thisptr, uint64_t count, void **encrypteds, void *destination)
{
    if (any_null(thisptr, encrypteds, destination))
        return E_POINTER;
    return run([&] {
        std::vector<Ciphertext> all;
        all.reserve(count);
        for (uint64_t i = 0; i < count; i++)
            all.emplace_back(ct(encrypteds[i]));
        ev(thisptr)->add_many(all, ct(destination));
    });
}

SEAL_C_FUNC Evaluator_AddPlain(void *thisptr, void *encrypted, void *plain, void *destination)
{
    if (any_null(plain))
        return E_POINTER;
    return unary(thisptr, encrypted, destination,
                 [&](gsb_seal::Evaluator &e, Ciphertext &x) { e.add_plain_inplace(x, *FromVoid<Plaintext>(plain)); });
}

SEAL_C_FUNC Evaluator_SubPlain(void *thisptr, void *encrypted, void *plain, void *destination)
{
    if (any_null(plain))
        return E_POINTER;
    return unary(thisptr, encrypted, destination,
                 [&](gsb_seal::Evaluator &e, Ciphertext &x) { e.sub_plain_inplace(x, *FromVoid<Plaintext>(plain)); });
}

SEAL_C_FUNC Evaluator_ModSwitchToNext2(void *thisptr, void *plain, void *destination)
{
    if (any_null(thisptr, plain, destination))
        return E_POINTER;
    return run([&] {
        if (destination != plain)
            *FromVoid<Plaintext>(destination) = *FromVoid<Plaintext>(plain);
        ev(thisptr)->mod_switch_to_next_inplace(*FromVoid<Plaintext>(destination));
    });
}

SEAL_C_FUNC Evaluator_ModSwitchTo1(void *thisptr, void *encrypted, uint64_t *parms_id, void *destination, void *)
{
    if (any_null(parms_id))
        return E_POINTER;
    seal::parms_id_type id;
    seal::c::CopyParmsId(parms_id, id);
    return unary(thisptr, encrypted, destination, [&](gsb_seal::Evaluator &e, Ciphertext &x) { e.mod_switch_to_inplace(x, id); });
}

SEAL_C_FUNC Evaluator_ModSwitchTo2(void *thisptr, void *plain, uint64_t *parms_id, void *destination)
{
    if (any_null(thisptr, plain, parms_id, destination))
        return E_POINTER;
    seal::parms_id_type id;
    seal::c::CopyParmsId(parms_id, id);
    return run([&] {
        if (destination != plain)
            *FromVoid<Plaintext>(destination) = *FromVoid<Plaintext>(plain);
        ev(thisptr)->mod_switch_to_inplace(*FromVoid<Plaintext>(destination), id);
    });
}

SEAL_C_FUNC Evaluator_TransformToNTT1(void *thisptr, void *plain, uint64_t *parms_id, void *destination_ntt, void *)
{
    if (any_null(thisptr, plain, parms_id, destination_ntt))
        return E_POINTER;
    seal::parms_id_type id;
    seal::c::CopyParmsId(parms_id, id);
    return run([&] {
        if (destination_ntt != plain)
            *FromVoid<Plaintext>(destination_ntt) = *FromVoid<Plaintext>(plain);
        ev(thisptr)->transform_to_ntt_inplace(*FromVoid<Plaintext>(destination_ntt), id);
    });
}

// ---- Decryptor (c/decryptor.h:16-22) ---------------------------------------------------------------------
SEAL_C_FUNC Decryptor_Create(void *context, void *secret_key, void **decryptor)
{
    if (any_null(context, secret_key, decryptor))
        return E_POINTER;
    const auto &shared = seal::c::SharedContextFromVoid(context);
    if (!shared)
        return E_POINTER;
    return run([&] { *decryptor = new gsb_seal::Decryptor(shared, *FromVoid<seal::SecretKey>(secret_key)); });
}

SEAL_C_FUNC Decryptor_Destroy(void *thisptr)
{
    if (any_null(thisptr))
        return E_POINTER;
    delete FromVoid<gsb_seal::Decryptor>(thisptr);
    return S_OK;
}

SEAL_C_FUNC Decryptor_Decrypt(void *thisptr, void *encrypted, void *destination)
{
    if (any_null(thisptr, encrypted, destination))
        return E_POINTER;
    return run([&] { FromVoid<gsb_seal::Decryptor>(thisptr)->decrypt(ct(encrypted), *FromVoid<Plaintext>(destination)); });
}

SEAL_C_FUNC Decryptor_InvariantNoiseBudget(void *thisptr, void *, int *)
{
    return not_served(thisptr);
}
