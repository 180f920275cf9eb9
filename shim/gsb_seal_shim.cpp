// See gsb_seal_shim.h.  Host C++ only: validation + metadata exactly as native/src/seal/evaluator.cpp,
// polynomial arithmetic through the C ABI of the B200 library.  No CPU arithmetic fallback.
#include "gsb_seal_shim.h"
#include "seal/util/numth.h"
#include "seal/util/rlwe.h"
#include "seal/randomgen.h"
#include "seal/util/uintcore.h"
#include "seal/valcheck.h"
#include <cmath>
#include <cstring>
#include <atomic>
#include <stdexcept>
#include <thread>
#include <utility>

using namespace seal;
using namespace std;

namespace gsb_seal
{
    namespace
    {
        [[noreturn]] void raise(int rc)
        {
            const string msg = gsb_last_error();
            switch (rc)
            {
            case GSB_E_INVALID_ARG: throw invalid_argument(msg);
            case GSB_E_LOGIC: throw logic_error(msg);
            case GSB_E_RANGE: throw out_of_range(msg);
            default: throw runtime_error("gseal_b200: " + msg);
            }
        }
        inline void check(int rc)
        {
            if (rc != GSB_OK)
                raise(rc);
        }
    } // namespace

    void Evaluator::fail(int rc) const
    {
        raise(rc);
    }

    Evaluator::Evaluator(shared_ptr<SEALContext> context, int device) : context_(move(context))
    {
        // evaluator.cpp:24-36
        if (!context_)
            throw invalid_argument("invalid context");
        if (!context_->parameters_set())
            throw invalid_argument("encryption parameters are not set correctly");
        auto &key_parms = context_->key_context_data()->parms();
        if (key_parms.scheme() != scheme_type::CKKS && key_parms.scheme() != scheme_type::BFV)
            throw logic_error("unsupported scheme");
        bfv_ = key_parms.scheme() == scheme_type::BFV;
        plain_modulus_ = bfv_ ? key_parms.plain_modulus().value() : 0;
        N_ = key_parms.poly_modulus_degree();
        K_ = key_parms.coeff_modulus().size();
        k_ = K_ - context_->first_context_data()->parms().coeff_modulus().size(); // evaluator.cpp:2273-2276
        if (K_ == 1)
            k_ = 1;
        vector<uint64_t> moduli;
        for (auto &m : key_parms.coeff_modulus())
            moduli.push_back(m.value());
        int logn = 0;
        while ((size_t(1) << logn) < N_)
            logn++;
        check(gsb_ctx_create(device, logn, moduli.data(), K_, k_, &ctx_));
    }

    static void drop_pool(gsb_ctx *ctx);
    Evaluator::~Evaluator()
    {
        drop_pool(ctx_);
        gsb_ctx_destroy(ctx_);
    }

    void Evaluator::validate(const Ciphertext &ct, const char *what) const
    {
        if (!is_metadata_valid_for(ct, context_) || !is_buffer_valid(ct))
            throw invalid_argument(string(what) + " is not valid for encryption parameters");
    }

    size_t Evaluator::level_limbs(const Ciphertext &ct) const
    {
        return context_->get_context_data(ct.parms_id())->parms().coeff_modulus().size();
    }

    // ---- multiply (evaluator.cpp:235-272, 447-527) ------------------------------------------------
    void Evaluator::multiply_inplace(Ciphertext &encrypted1, const Ciphertext &encrypted2, MemoryPoolHandle)
    {
        validate(encrypted1, "encrypted1");
        validate(encrypted2, "encrypted2");
        if (encrypted1.parms_id() != encrypted2.parms_id())
            throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
        if (bfv_)
        {
            bfv_multiply(encrypted1, encrypted2);
            return;
        }
        if (!(encrypted1.is_ntt_form() && encrypted2.is_ntt_form()))
            throw invalid_argument("encrypted1 or encrypted2 must be in NTT form");
        auto &context_data = *context_->get_context_data(encrypted1.parms_id());
        const double new_scale = encrypted1.scale() * encrypted2.scale();
        if (new_scale <= 0 || (static_cast<int>(log2(new_scale)) >= context_data.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        const size_t L = level_limbs(encrypted1), s1 = encrypted1.size(), s2 = encrypted2.size();
        const size_t dest_size = seal::util::sub_safe(seal::util::add_safe(s1, s2), size_t(1)); // evaluator.cpp:470
        if (!seal::util::product_fits_in(dest_size, N_, L))
            throw logic_error("invalid parameters");
        vector<uint64_t> b;
        const uint64_t *pb = encrypted2.data();
        if (&encrypted1 == &encrypted2) // square: keep a copy, the destination is resized below
        {
            b.assign(encrypted2.data(), encrypted2.data() + s2 * L * N_);
            pb = b.data();
        }
        // the resize keeps the operand's polynomials at the front of the buffer; the device call reads them before
        // it writes the product there.  A device failure puts the operand back as it was.
        encrypted1.resize(context_, context_data.parms_id(), dest_size);
        const int rc = gsb_ckks_multiply_sizes_host(ctx_, encrypted1.data(), s1, pb, s2, encrypted1.data(), 1, L);
        if (rc != GSB_OK)
        {
            encrypted1.resize(context_, context_data.parms_id(), s1);
            fail(rc);
        }
        encrypted1.scale() = new_scale;
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted1.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    // evaluator.cpp:274-445 (BEHZ); size-2 operands, coefficient form
    void Evaluator::bfv_multiply(Ciphertext &encrypted1, const Ciphertext &encrypted2)
    {
        if (encrypted1.is_ntt_form() || encrypted2.is_ntt_form())
            throw invalid_argument("encrypted1 or encrypted2 cannot be in NTT form");
        auto &context_data = *context_->get_context_data(encrypted1.parms_id());
        const size_t L = level_limbs(encrypted1), s1 = encrypted1.size(), s2 = encrypted2.size();
        const size_t dest_size = seal::util::sub_safe(seal::util::add_safe(s1, s2), size_t(1));
        if (!seal::util::product_fits_in(dest_size, N_, context_data.rns_tool()->base_Bsk_m_tilde()->size())) // evaluator.cpp:297-300
            throw logic_error("invalid parameters");
        vector<uint64_t> b;
        const uint64_t *pb = encrypted2.data();
        if (&encrypted1 == &encrypted2)
        {
            b.assign(encrypted2.data(), encrypted2.data() + s2 * L * N_);
            pb = b.data();
        }
        encrypted1.resize(context_, context_data.parms_id(), dest_size);
        const int rc = gsb_bfv_multiply_sizes_host(ctx_, plain_modulus_, encrypted1.data(), s1, pb, s2, encrypted1.data(), 1, L);
        if (rc != GSB_OK)
        {
            encrypted1.resize(context_, context_data.parms_id(), s1);
            fail(rc);
        }
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted1.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::multiply(const Ciphertext &encrypted1, const Ciphertext &encrypted2, Ciphertext &destination, MemoryPoolHandle pool)
    {
        if (&encrypted2 == &destination) // evaluator.h:262-276
        {
            multiply_inplace(destination, encrypted1, move(pool));
        }
        else
        {
            destination = encrypted1;
            multiply_inplace(destination, encrypted2, move(pool));
        }
    }

    // ---- elementwise ops and form / level changes ---------------------------------
    // evaluator.cpp:65-88
    void Evaluator::negate_inplace(Ciphertext &encrypted)
    {
        validate(encrypted, "encrypted");
        const size_t L = level_limbs(encrypted);
        check(gsb_poly_op_host(ctx_, GSB_OP_NEGATE, encrypted.data(), nullptr, nullptr, encrypted.data(), encrypted.size(), L, 0));
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    // evaluator.cpp:90-152 (add), :174-233 (sub)
    void Evaluator::add_sub(Ciphertext &encrypted1, const Ciphertext &encrypted2, bool subtract)
    {
        validate(encrypted1, "encrypted1");
        validate(encrypted2, "encrypted2");
        if (encrypted1.parms_id() != encrypted2.parms_id())
            throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
        if (encrypted1.is_ntt_form() != encrypted2.is_ntt_form())
            throw invalid_argument("NTT form mismatch");
        if (!seal::util::are_close<double>(encrypted1.scale(), encrypted2.scale())) // are_same_scale, evaluator.cpp:38-41
            throw invalid_argument("scale mismatch");
        auto &context_data = *context_->get_context_data(encrypted1.parms_id());
        const size_t L = level_limbs(encrypted1), LN = L * N_;
        const size_t s1 = encrypted1.size(), s2 = encrypted2.size(), lo = min(s1, s2), hi = max(s1, s2);
        encrypted1.resize(context_, context_data.parms_id(), hi);
        check(gsb_poly_op_host(ctx_, subtract ? GSB_OP_SUB : GSB_OP_ADD, encrypted1.data(), encrypted2.data(), nullptr, encrypted1.data(), lo, L, 0));
        if (s1 < s2)
        {
            if (subtract) // remaining components of encrypted2 are negated (:222-226)
                check(gsb_poly_op_host(ctx_, GSB_OP_NEGATE, encrypted2.data(lo), nullptr, nullptr, encrypted1.data(lo), s2 - lo, L, 0));
            else
                memcpy(encrypted1.data(lo), encrypted2.data(lo), (s2 - lo) * LN * sizeof(uint64_t));
        }
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted1.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::add_inplace(Ciphertext &encrypted1, const Ciphertext &encrypted2)
    {
        add_sub(encrypted1, encrypted2, false);
    }

    void Evaluator::sub_inplace(Ciphertext &encrypted1, const Ciphertext &encrypted2)
    {
        add_sub(encrypted1, encrypted2, true);
    }

    // evaluator.cpp:1438-1475 -> multiply_plain_ntt (:1605-1646); the coefficient-form branch is not served
    void Evaluator::multiply_plain_inplace(Ciphertext &encrypted, const Plaintext &plain, MemoryPoolHandle pool)
    {
        validate(encrypted, "encrypted");
        if (!is_metadata_valid_for(plain, context_) || !is_buffer_valid(plain))
            throw invalid_argument("plain is not valid for encryption parameters");
        if (encrypted.is_ntt_form() != plain.is_ntt_form())
            throw invalid_argument("NTT form mismatch");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        if (!encrypted.is_ntt_form())
        {
            multiply_plain_normal(encrypted, plain);
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
            if (encrypted.is_transparent())
                throw logic_error("result ciphertext is transparent");
#endif
            return;
        }
        if (encrypted.parms_id() != plain.parms_id())
            throw invalid_argument("encrypted_ntt and plain_ntt parameter mismatch");
        auto &context_data = *context_->get_context_data(encrypted.parms_id());
        const double new_scale = encrypted.scale() * plain.scale();
        if (new_scale <= 0 || (static_cast<int>(log2(new_scale)) >= context_data.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        const size_t L = level_limbs(encrypted);
        for (size_t i = 0; i < encrypted.size(); i++)
            check(gsb_poly_op_host(ctx_, GSB_OP_DYADIC, encrypted.data(i), plain.data(), nullptr, encrypted.data(i), 1, L, 0));
        encrypted.scale() = new_scale;
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    // evaluator.cpp:1475-1603 (coefficient-form ciphertext and plaintext: BFV)
    void Evaluator::multiply_plain_normal(Ciphertext &encrypted, const Plaintext &plain)
    {
        auto &context_data = *context_->get_context_data(encrypted.parms_id());
        const size_t L = level_limbs(encrypted), size = encrypted.size();
        if (!seal::util::product_fits_in(size, N_, L))
            throw logic_error("invalid parameters");
        const double new_scale = encrypted.scale() * plain.scale();
        if (new_scale <= 0 || (static_cast<int>(log2(new_scale)) >= context_data.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        const uint64_t threshold = context_data.plain_upper_half_threshold();
        const uint64_t *increment = context_data.plain_upper_half_increment();
        const bool fast_lift = context_data.qualifiers().using_fast_plain_lift;
        if (plain.nonzero_coeff_count() == 1)
        {
            // monomial branch (:1512-1546): the coefficient is used as it is unless it needs the multi-precision lift
            const size_t e = plain.significant_coeff_count() - 1;
            if (plain[e] >= threshold && !fast_lift)
                throw logic_error("device multiply_plain serves contexts with using_fast_plain_lift");
            vector<uint64_t> mono(e + 1, 0);
            mono[e] = plain[e];
            check(gsb_multiply_plain_coeff_host(ctx_, encrypted.data(), size, L, mono.data(), mono.size(), threshold, increment, 0));
        }
        else
        {
            if (!fast_lift) // the multi-precision lift (:1558-1575) is not served
                throw logic_error("device multiply_plain serves contexts with using_fast_plain_lift");
            check(gsb_multiply_plain_coeff_host(ctx_, encrypted.data(), size, L, plain.data(), plain.coeff_count(), threshold, increment, 1));
        }
        encrypted.scale() = new_scale; // set before the arithmetic in the reference (:1506); no effect on the words
    }

    // evaluator.cpp:1290-1437
    void Evaluator::add_sub_plain(Ciphertext &encrypted, const Plaintext &plain, bool subtract)
    {
        validate(encrypted, "encrypted");
        if (!is_metadata_valid_for(plain, context_) || !is_buffer_valid(plain))
            throw invalid_argument("plain is not valid for encryption parameters");
        auto &context_data = *context_->get_context_data(encrypted.parms_id());
        if (bfv_ && encrypted.is_ntt_form())
            throw invalid_argument("BFV encrypted cannot be in NTT form");
        if (!bfv_ && !encrypted.is_ntt_form())
            throw invalid_argument("CKKS encrypted must be in NTT form");
        if (plain.is_ntt_form() != encrypted.is_ntt_form())
            throw invalid_argument("NTT form mismatch");
        if (encrypted.is_ntt_form() && (encrypted.parms_id() != plain.parms_id()))
            throw invalid_argument("encrypted and plain parameter mismatch");
        if (!seal::util::are_close<double>(encrypted.scale(), plain.scale()))
            throw invalid_argument("scale mismatch");
        const size_t L = level_limbs(encrypted);
        if (!seal::util::product_fits_in(N_, L))
            throw logic_error("invalid parameters");
        if (bfv_)
            check(gsb_bfv_add_plain_host(
                ctx_, encrypted.data(), plain.data(), plain.coeff_count(), subtract ? 1 : 0, context_data.coeff_div_plain_modulus(),
                context_data.coeff_modulus_mod_plain_modulus(), context_data.plain_upper_half_threshold(), plain_modulus_, L));
        else
            check(gsb_poly_op_host(ctx_, subtract ? GSB_OP_SUB : GSB_OP_ADD, encrypted.data(), plain.data(), nullptr, encrypted.data(), 1, L, 0));
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::add_plain_inplace(Ciphertext &encrypted, const Plaintext &plain)
    {
        add_sub_plain(encrypted, plain, false);
    }

    void Evaluator::sub_plain_inplace(Ciphertext &encrypted, const Plaintext &plain)
    {
        add_sub_plain(encrypted, plain, true);
    }

    // evaluator.cpp:153-172
    void Evaluator::add_many(const vector<Ciphertext> &encrypteds, Ciphertext &destination)
    {
        if (encrypteds.empty())
            throw invalid_argument("encrypteds cannot be empty");
        for (size_t i = 0; i < encrypteds.size(); i++)
            if (&encrypteds[i] == &destination)
                throw invalid_argument("encrypteds must be different from destination");
        destination = encrypteds[0];
        for (size_t i = 1; i < encrypteds.size(); i++)
            add_inplace(destination, encrypteds[i]);
    }

    // evaluator.cpp:1038-1064
    void Evaluator::mod_switch_to_inplace(Ciphertext &encrypted, parms_id_type parms_id, MemoryPoolHandle pool)
    {
        auto context_data_ptr = context_->get_context_data(encrypted.parms_id());
        auto target_context_data_ptr = context_->get_context_data(parms_id);
        if (!context_data_ptr)
            throw invalid_argument("encrypted is not valid for encryption parameters");
        if (!target_context_data_ptr)
            throw invalid_argument("parms_id is not valid for encryption parameters");
        if (context_data_ptr->chain_index() < target_context_data_ptr->chain_index())
            throw invalid_argument("cannot switch to higher level modulus");
        while (encrypted.parms_id() != parms_id)
            mod_switch_to_next_inplace(encrypted, pool);
    }

    // evaluator.h:430-438 -> mod_switch_drop_to_next(Plaintext &) (evaluator.cpp:958-994): an NTT-form plaintext is
    // limb-major, so dropping the last prime is a resize
    void Evaluator::mod_switch_to_next_inplace(Plaintext &plain)
    {
        if (!is_valid_for(plain, context_))
            throw invalid_argument("plain is not valid for encryption parameters");
        auto context_data_ptr = context_->get_context_data(plain.parms_id());
        if (!plain.is_ntt_form())
            throw invalid_argument("plain is not in NTT form");
        if (!context_data_ptr->next_context_data())
            throw invalid_argument("end of modulus switching chain reached");
        auto &next_context_data = *context_data_ptr->next_context_data();
        if (plain.scale() <= 0 || (static_cast<int>(log2(plain.scale())) >= next_context_data.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        const size_t dest_size = seal::util::mul_safe(next_context_data.parms().coeff_modulus().size(), N_);
        plain.parms_id() = parms_id_zero;
        plain.resize(dest_size);
        plain.parms_id() = next_context_data.parms_id();
    }

    // evaluator.cpp:1066-1088
    void Evaluator::mod_switch_to_inplace(Plaintext &plain, parms_id_type parms_id)
    {
        auto context_data_ptr = context_->get_context_data(plain.parms_id());
        auto target_context_data_ptr = context_->get_context_data(parms_id);
        if (!context_data_ptr)
            throw invalid_argument("plain is not valid for encryption parameters");
        if (!target_context_data_ptr)
            throw invalid_argument("parms_id is not valid for encryption parameters");
        if (!plain.is_ntt_form())
            throw invalid_argument("plain is not in NTT form");
        if (context_data_ptr->chain_index() < target_context_data_ptr->chain_index())
            throw invalid_argument("cannot switch to higher level modulus");
        while (plain.parms_id() != parms_id)
            mod_switch_to_next_inplace(plain);
    }

    // evaluator.cpp:1648-1744 (BFV plaintext -> NTT form at the level parms_id)
    void Evaluator::transform_to_ntt_inplace(Plaintext &plain, parms_id_type parms_id, MemoryPoolHandle pool)
    {
        if (!is_valid_for(plain, context_))
            throw invalid_argument("plain is not valid for encryption parameters");
        auto context_data_ptr = context_->get_context_data(parms_id);
        if (!context_data_ptr)
            throw invalid_argument("parms_id is not valid for the current context");
        if (plain.is_ntt_form())
            throw invalid_argument("plain is already in NTT form");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        auto &context_data = *context_data_ptr;
        const size_t L = context_data.parms().coeff_modulus().size(), count = plain.coeff_count();
        if (!seal::util::product_fits_in(N_, L))
            throw logic_error("invalid parameters");
        if (!context_data.qualifiers().using_fast_plain_lift) // the multi-precision lift (:1694-1717) is not served
            throw logic_error("device transform_to_ntt serves contexts with using_fast_plain_lift");
        vector<uint64_t> raw(plain.data(), plain.data() + count);
        plain.resize(N_ * L);
        const int rc = gsb_bfv_plain_to_rns_host(
            ctx_, raw.data(), count, context_data.plain_upper_half_threshold(), context_data.plain_upper_half_increment(), plain.data(), L, 1);
        if (rc != GSB_OK)
        {
            plain.resize(count);
            fail(rc);
        }
        plain.parms_id() = parms_id;
    }

    // evaluator.cpp:529-558: ckks_square / bfv_square give the words of multiply(x, x)
    void Evaluator::square_inplace(Ciphertext &encrypted, MemoryPoolHandle pool)
    {
        multiply_inplace(encrypted, encrypted, move(pool));
    }

    // evaluator.cpp:1746-1792, :1794-1840
    void Evaluator::transform_to_ntt_inplace(Ciphertext &encrypted)
    {
        validate(encrypted, "encrypted");
        if (encrypted.is_ntt_form())
            throw invalid_argument("encrypted is already in NTT form");
        check(gsb_ntt_forward_host(ctx_, encrypted.data(), encrypted.size(), level_limbs(encrypted), 0, GSB_NTT_STRICT));
        encrypted.is_ntt_form() = true;
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::transform_from_ntt_inplace(Ciphertext &encrypted_ntt)
    {
        validate(encrypted_ntt, "encrypted");
        if (!encrypted_ntt.is_ntt_form())
            throw invalid_argument("encrypted_ntt is not in NTT form");
        check(gsb_ntt_inverse_host(ctx_, encrypted_ntt.data(), encrypted_ntt.size(), level_limbs(encrypted_ntt), 0, GSB_NTT_STRICT));
        encrypted_ntt.is_ntt_form() = false;
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted_ntt.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    // evaluator.cpp:996-1036 -> CKKS: mod_switch_drop_to_next (:894-957), a copy of the first L-1 limbs
    void Evaluator::mod_switch_to_next_inplace(Ciphertext &encrypted, MemoryPoolHandle pool)
    {
        validate(encrypted, "encrypted");
        if (context_->last_parms_id() == encrypted.parms_id())
            throw invalid_argument("end of modulus switching chain reached");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        auto &context_data = *context_->get_context_data(encrypted.parms_id());
        auto &next = *context_data.next_context_data();
        if (bfv_)
        {
            // mod_switch_scale_to_next (evaluator.cpp:829-892) -> divide_and_round_q_last_inplace
            if (encrypted.is_ntt_form())
                throw invalid_argument("BFV encrypted cannot be in NTT form");
            const size_t Lb = level_limbs(encrypted), sz = encrypted.size();
            vector<uint64_t> out(sz * (Lb - 1) * N_);
            check(gsb_bfv_mod_switch_host(ctx_, encrypted.data(), out.data(), 1, sz, Lb));
            encrypted.resize(context_, next.parms_id(), sz);
            encrypted.is_ntt_form() = false;
            memcpy(encrypted.data(), out.data(), out.size() * sizeof(uint64_t));
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
            if (encrypted.is_transparent())
                throw logic_error("result ciphertext is transparent");
#endif
            return;
        }
        if (!encrypted.is_ntt_form())
            throw invalid_argument("CKKS encrypted must be in NTT form");
        if (encrypted.scale() <= 0 || (static_cast<int>(log2(encrypted.scale())) >= next.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        const size_t L = level_limbs(encrypted), size = encrypted.size();
        vector<uint64_t> temp(size * (L - 1) * N_);
        for (size_t i = 0; i < size; i++)
            memcpy(temp.data() + i * (L - 1) * N_, encrypted.data(i), (L - 1) * N_ * sizeof(uint64_t));
        const double scale = encrypted.scale();
        encrypted.resize(context_, next.parms_id(), size);
        encrypted.is_ntt_form() = true;
        encrypted.scale() = scale;
        memcpy(encrypted.data(), temp.data(), temp.size() * sizeof(uint64_t));
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    // evaluator.cpp:1128-1178
    void Evaluator::rescale_to_inplace(Ciphertext &encrypted, parms_id_type parms_id, MemoryPoolHandle pool)
    {
        validate(encrypted, "encrypted");
        auto context_data_ptr = context_->get_context_data(encrypted.parms_id());
        auto target_context_data_ptr = context_->get_context_data(parms_id);
        if (!context_data_ptr)
            throw invalid_argument("encrypted is not valid for encryption parameters");
        if (!target_context_data_ptr)
            throw invalid_argument("parms_id is not valid for encryption parameters");
        if (context_data_ptr->chain_index() < target_context_data_ptr->chain_index())
            throw invalid_argument("cannot switch to higher level modulus");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        if (bfv_)
            throw invalid_argument("unsupported operation for scheme type");
        while (encrypted.parms_id() != parms_id)
            rescale_to_next_inplace(encrypted, pool);
    }

    // ---- key upload ---------------------------------------------------------------------------------
    void Evaluator::upload_key(int kind, uint32_t slot, const vector<PublicKey> &bundles)
    {
        vector<const uint64_t *> ptrs;
        for (size_t j = 0; j < bundles.size(); j++)
        {
            // evaluator.cpp:2283-2290: only the used component is validated
            if (!is_metadata_valid_for(bundles[j], context_) || !is_buffer_valid(bundles[j]) || bundles[j].data().size() != 2)
                throw invalid_argument("kswitch_keys is not valid for encryption parameters");
            ptrs.push_back(bundles[j].data().data());
        }
        check(gsb_key_upload_bundles(ctx_, kind, slot, ptrs.data(), ptrs.size()));
    }

    // Which host key sits in a device slot is decided by CONTENT, not by address: SEAL's MemoryPool hands a freed
    // block to the next allocation of the same size, so a regenerated / reloaded / other client's key can live at
    // the address of the one uploaded earlier.  The fingerprint mixes, for every bundle, its size, its first and
    // last 512 words and one word out of every 1024 (a 120 MiB key costs ~ 20 k words to fingerprint).
    static uint64_t key_fingerprint(const vector<PublicKey> &bundles)
    {
        uint64_t h = 0x9E3779B97F4A7C15ull ^ bundles.size();
        auto mix = [&](uint64_t v) {
            h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
            h *= 0xBF58476D1CE4E5B9ull;
            h ^= h >> 29;
        };
        for (auto &pk : bundles)
        {
            const uint64_t *p = pk.data().data();
            const size_t n = pk.data().size() * pk.data().poly_modulus_degree() * pk.data().coeff_modulus_size();
            mix(n);
            const size_t edge = n < 1024 ? n / 2 : 512;
            for (size_t i = 0; i < edge; i++)
            {
                mix(p[i]);
                mix(p[n - 1 - i]);
            }
            for (size_t i = edge; i + edge < n; i += 1024)
                mix(p[i]);
        }
        return h;
    }

    void Evaluator::ensure_relin(const RelinKeys &keys, size_t index)
    {
        lock_guard<mutex> lock(mu_);
        const auto &bundles = keys.data()[index];
        const uint64_t id = key_fingerprint(bundles);
        auto it = relin_loaded_.find(static_cast<uint32_t>(index));
        if (it != relin_loaded_.end() && it->second == id)
            return;
        upload_key(GSB_KEY_RELIN, static_cast<uint32_t>(index), bundles);
        relin_loaded_[static_cast<uint32_t>(index)] = id;
    }

    void Evaluator::ensure_galois(const GaloisKeys &keys, uint32_t elt)
    {
        lock_guard<mutex> lock(mu_);
        const auto &bundles = keys.data()[GaloisKeys::get_index(elt)];
        const uint64_t id = key_fingerprint(bundles);
        auto it = galois_loaded_.find(elt);
        if (it != galois_loaded_.end() && it->second == id)
            return;
        upload_key(GSB_KEY_GALOIS, elt, bundles);
        galois_loaded_[elt] = id;
    }

    // ---- relinearize (evaluator.cpp:772-827) -----------------------------------------------------------
    void Evaluator::relinearize_inplace(Ciphertext &encrypted, const RelinKeys &relin_keys, MemoryPoolHandle)
    {
        validate(encrypted, "encrypted"); // evaluator.h:344-348 (checked by the public wrapper)
        auto context_data_ptr = context_->get_context_data(encrypted.parms_id());
        if (!context_data_ptr)
            throw invalid_argument("encrypted is not valid for encryption parameters");
        if (relin_keys.parms_id() != context_->key_parms_id())
            throw invalid_argument("relin_keys is not valid for encryption parameters");
        const size_t size = encrypted.size();
        if (relin_keys.size() < (size >= 2 ? size - 2 : 0))
            throw invalid_argument("not enough relinearization keys");
        if (size <= 2)
            return;
        if (encrypted.is_ntt_form() == bfv_) // BFV ciphertexts are in coefficient form, CKKS ones in NTT form
            throw logic_error(bfv_ ? "BFV encrypted cannot be in NTT form" : "CKKS encrypted must be in NTT form");
        const size_t L = level_limbs(encrypted);
        for (size_t p = 2; p < size; p++) // every key the loop at evaluator.cpp:809-815 uses: s^2 .. s^(size-1)
            ensure_relin(relin_keys, RelinKeys::get_index(p));
        // in place on the host buffer: one chunk, so the device has read all `size` polynomials before the two
        // result polynomials are copied back over the front of the buffer; the return code is checked before the
        // ciphertext is resized (no mutation on failure)
        if (size == 3)
            check(bfv_ ? gsb_bfv_relinearize_host(ctx_, encrypted.data(), encrypted.data(), 1, L)
                       : gsb_relinearize_host(ctx_, encrypted.data(), encrypted.data(), 1, L));
        else
            check(gsb_relinearize_sizes_host(ctx_, encrypted.data(), size, encrypted.data(), 1, L, bfv_ ? 1 : 0));
        encrypted.resize(context_, context_data_ptr->parms_id(), 2);
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::relinearize(const Ciphertext &encrypted, const RelinKeys &relin_keys, Ciphertext &destination, MemoryPoolHandle pool)
    {
        destination = encrypted;
        relinearize_inplace(destination, relin_keys, move(pool));
    }

    // ---- multiply_many / exponentiate (evaluator.cpp:1180-1288) ----------------------------------------
    void Evaluator::multiply_many(
        const vector<Ciphertext> &encrypteds, const RelinKeys &relin_keys, Ciphertext &destination, MemoryPoolHandle pool)
    {
        if (encrypteds.size() == 0)
            throw invalid_argument("encrypteds vector must not be empty");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        for (size_t i = 0; i < encrypteds.size(); i++)
            if (&encrypteds[i] == &destination)
                throw invalid_argument("encrypteds must be different from destination");
        auto context_data_ptr = context_->get_context_data(encrypteds[0].parms_id());
        if (!context_data_ptr)
            throw invalid_argument("encrypteds is not valid for encryption parameters");
        if (!bfv_)
            throw logic_error("unsupported scheme");
        if (encrypteds.size() == 1)
        {
            destination = encrypteds[0];
            return;
        }
        // what multiply / relinearize_inplace check per product (:235-272, :772-800)
        for (auto &ct : encrypteds)
        {
            validate(ct, "encrypted1");
            if (ct.parms_id() != encrypteds[0].parms_id())
                throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
            if (ct.is_ntt_form())
                throw invalid_argument("encrypted1 or encrypted2 cannot be in NTT form");
            if (ct.size() != 2)
                throw logic_error("device multiply serves size-2 operands (relinearize first)");
        }
        if (relin_keys.parms_id() != context_->key_parms_id())
            throw invalid_argument("relin_keys is not valid for encryption parameters");
        if (relin_keys.size() < 1)
            throw invalid_argument("not enough relinearization keys");
        ensure_relin(relin_keys, RelinKeys::get_index(2));

        // The reference's queue (:1224-1254): q = [e0*e1, e2*e3, ..., (e_last if odd)], then q.push(q[i]*q[i+1])
        // for i = 0, 2, 4, ... until one product is left.  Entry j of the queue is a pair of earlier entries,
        // so the pairs whose operands already exist form one batch.
        const size_t L = level_limbs(encrypteds[0]), ct_words = 2 * L * N_;
        const size_t n = encrypteds.size();
        vector<vector<uint64_t>> q;                                       // queue entries, [2][L][N] each
        q.reserve(n);
        vector<uint64_t> a, b, c3, c2;
        auto run_wave = [&](const vector<pair<const uint64_t *, const uint64_t *>> &pairs) {
            const size_t w = pairs.size();
            a.resize(w * ct_words), b.resize(w * ct_words), c3.resize(w * 3 * L * N_), c2.resize(w * ct_words);
            for (size_t i = 0; i < w; i++)
            {
                memcpy(a.data() + i * ct_words, pairs[i].first, ct_words * sizeof(uint64_t));
                memcpy(b.data() + i * ct_words, pairs[i].second, ct_words * sizeof(uint64_t));
            }
            check(gsb_bfv_multiply_host(ctx_, plain_modulus_, a.data(), b.data(), c3.data(), w, L));
            check(gsb_bfv_relinearize_host(ctx_, c3.data(), c2.data(), w, L));
            for (size_t i = 0; i < w; i++)
                q.emplace_back(c2.begin() + i * ct_words, c2.begin() + (i + 1) * ct_words);
        };
        vector<pair<const uint64_t *, const uint64_t *>> pairs;
        for (size_t i = 0; i + 1 < n; i += 2)
            pairs.emplace_back(encrypteds[i].data(), encrypteds[i + 1].data());
        run_wave(pairs);
        if (n & 1)
            q.emplace_back(encrypteds.back().data(), encrypteds.back().data() + ct_words);
        for (size_t i = 0; i + 1 < q.size();)
        {
            pairs.clear();
            const size_t avail = q.size();
            for (; i + 1 < avail; i += 2)
                pairs.emplace_back(q[i].data(), q[i + 1].data());
            run_wave(pairs);
        }
        destination.resize(context_, context_data_ptr->parms_id(), 2);
        destination.is_ntt_form() = false;
        memcpy(destination.data(), q.back().data(), ct_words * sizeof(uint64_t));
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (destination.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::exponentiate_inplace(Ciphertext &encrypted, uint64_t exponent, const RelinKeys &relin_keys, MemoryPoolHandle pool)
    {
        if (!context_->get_context_data(encrypted.parms_id()))
            throw invalid_argument("encrypted is not valid for encryption parameters");
        if (!context_->get_context_data(relin_keys.parms_id()))
            throw invalid_argument("relin_keys is not valid for encryption parameters");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        if (exponent == 0)
            throw invalid_argument("exponent cannot be 0");
        if (exponent == 1)
            return;
        vector<Ciphertext> exp_vector(static_cast<size_t>(exponent), encrypted);
        multiply_many(exp_vector, relin_keys, encrypted, move(pool));
    }

    // ---- rescale (evaluator.cpp:1090-1126, 829-892) ---------------------------------------------------
    void Evaluator::rescale_to_next(const Ciphertext &encrypted, Ciphertext &destination, MemoryPoolHandle pool)
    {
        validate(encrypted, "encrypted");
        if (context_->last_parms_id() == encrypted.parms_id())
            throw invalid_argument("end of modulus switching chain reached");
        if (!pool)
            throw invalid_argument("pool is uninitialized");
        if (!encrypted.is_ntt_form())
            throw invalid_argument("CKKS encrypted must be in NTT form");
        auto &context_data = *context_->get_context_data(encrypted.parms_id());
        auto &next_context_data = *context_data.next_context_data();
        const size_t L = context_data.parms().coeff_modulus().size(), size = encrypted.size();
        vector<uint64_t> out(size * (L - 1) * N_);
        check(gsb_rescale_host(ctx_, encrypted.data(), out.data(), 1, size, L));
        const double new_scale = encrypted.scale() / static_cast<double>(context_data.parms().coeff_modulus().back().value());
        const bool ntt = encrypted.is_ntt_form();
        destination.resize(context_, next_context_data.parms_id(), size);
        memcpy(destination.data(), out.data(), out.size() * sizeof(uint64_t));
        destination.is_ntt_form() = ntt;
        destination.scale() = new_scale;
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (destination.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::rescale_to_next_inplace(Ciphertext &encrypted, MemoryPoolHandle pool)
    {
        rescale_to_next(encrypted, encrypted, move(pool));
    }

    // ---- Galois (evaluator.cpp:1841-1943) -----------------------------------------------------------
    void Evaluator::apply_galois_inplace(Ciphertext &encrypted, uint32_t galois_elt, const GaloisKeys &galois_keys, MemoryPoolHandle)
    {
        validate(encrypted, "encrypted");
        if (galois_keys.parms_id() != context_->key_parms_id())
            throw invalid_argument("galois_keys is not valid for encryption parameters");
        if (!galois_keys.has_key(galois_elt))
            throw invalid_argument("Galois key not present");
        if (!(galois_elt & 1) || galois_elt >= 2 * N_)
            throw invalid_argument("Galois element is not valid");
        if (encrypted.size() > 2)
            throw invalid_argument("encrypted size must be 2");
        if (encrypted.is_ntt_form() == bfv_)
            throw logic_error(bfv_ ? "BFV encrypted cannot be in NTT form" : "CKKS encrypted must be in NTT form");
        ensure_galois(galois_keys, galois_elt);
        const size_t L = level_limbs(encrypted);
        vector<uint64_t> out(2 * L * N_);
        check(bfv_ ? gsb_bfv_apply_galois_host(ctx_, galois_elt, encrypted.data(), out.data(), 1, L)
                   : gsb_apply_galois_host(ctx_, galois_elt, encrypted.data(), out.data(), 1, L));
        memcpy(encrypted.data(), out.data(), out.size() * sizeof(uint64_t));
#ifdef SEAL_THROW_ON_TRANSPARENT_CIPHERTEXT
        if (encrypted.is_transparent())
            throw logic_error("result ciphertext is transparent");
#endif
    }

    void Evaluator::apply_galois(
        const Ciphertext &encrypted, uint32_t galois_elt, const GaloisKeys &galois_keys, Ciphertext &destination, MemoryPoolHandle pool)
    {
        destination = encrypted;
        apply_galois_inplace(destination, galois_elt, galois_keys, move(pool));
    }

    // evaluator.cpp:1945-2000
    void Evaluator::rotate_internal(Ciphertext &encrypted, int steps, const GaloisKeys &galois_keys)
    {
        auto context_data_ptr = context_->get_context_data(encrypted.parms_id());
        if (!context_data_ptr)
            throw invalid_argument("encrypted is not valid for encryption parameters");
        if (!context_data_ptr->qualifiers().using_batching)
            throw logic_error("encryption parameters do not support batching");
        if (galois_keys.parms_id() != context_->key_parms_id())
            throw invalid_argument("galois_keys is not valid for encryption parameters");
        if (steps == 0)
            return;
        const size_t coeff_count = context_data_ptr->parms().poly_modulus_degree();
        auto galois_tool = context_data_ptr->galois_tool();
        if (galois_keys.has_key(galois_tool->get_elt_from_step(steps)))
        {
            apply_galois_inplace(encrypted, galois_tool->get_elt_from_step(steps), galois_keys, MemoryManager::GetPool());
            return;
        }
        const vector<int> naf_steps = seal::util::naf(steps);
        if (naf_steps.size() == 1)
            throw invalid_argument("Galois key not present");
        for (int step : naf_steps)
            if (static_cast<size_t>(abs(step)) != (coeff_count >> 1))
                rotate_internal(encrypted, step, galois_keys);
    }

    void Evaluator::rotate_vector_inplace(Ciphertext &encrypted, int steps, const GaloisKeys &galois_keys, MemoryPoolHandle)
    {
        // evaluator.h:1201-1209
        if (context_->key_context_data()->parms().scheme() != scheme_type::CKKS)
            throw logic_error("unsupported scheme");
        rotate_internal(encrypted, steps, galois_keys);
    }

    void Evaluator::rotate_vector(
        const Ciphertext &encrypted, int steps, const GaloisKeys &galois_keys, Ciphertext &destination, MemoryPoolHandle pool)
    {
        destination = encrypted;
        rotate_vector_inplace(destination, steps, galois_keys, move(pool));
    }

    void Evaluator::rotate_rows_inplace(Ciphertext &encrypted, int steps, const GaloisKeys &galois_keys, MemoryPoolHandle)
    {
        // evaluator.h:1049-1057
        if (context_->key_context_data()->parms().scheme() != scheme_type::BFV)
            throw logic_error("unsupported scheme");
        rotate_internal(encrypted, steps, galois_keys);
    }

    void Evaluator::rotate_columns_inplace(Ciphertext &encrypted, const GaloisKeys &galois_keys, MemoryPoolHandle pool)
    {
        // evaluator.h:1125-1133 -> conjugate_internal (:1343-1363): Galois element 2N-1
        if (context_->key_context_data()->parms().scheme() != scheme_type::BFV)
            throw logic_error("unsupported scheme");
        auto context_data_ptr = context_->get_context_data(encrypted.parms_id());
        if (!context_data_ptr)
            throw invalid_argument("encrypted is not valid for encryption parameters");
        if (!context_data_ptr->qualifiers().using_batching)
            throw logic_error("encryption parameters do not support batching");
        apply_galois_inplace(encrypted, static_cast<uint32_t>(2 * N_ - 1), galois_keys, move(pool));
    }

    void Evaluator::complex_conjugate_inplace(Ciphertext &encrypted, const GaloisKeys &galois_keys, MemoryPoolHandle pool)
    {
        // evaluator.h:1343-1352: conjugation is the Galois element 2N-1 (galois.h get_elt_from_step(0))
        apply_galois_inplace(encrypted, static_cast<uint32_t>(2 * N_ - 1), galois_keys, move(pool));
    }

    // ---- batched extension ------------------------------------------------------------------------------
    void Evaluator::multiply_relinearize_rescale_inplace(
        vector<Ciphertext *> &encrypted1, const vector<const Ciphertext *> &encrypted2, const RelinKeys &relin_keys)
    {
        const size_t n = encrypted1.size();
        if (n == 0 || encrypted2.size() != n)
            throw invalid_argument("batch size mismatch");
        if (relin_keys.parms_id() != context_->key_parms_id() || relin_keys.size() < 1)
            throw invalid_argument("relin_keys is not valid for encryption parameters");
        const auto parms_id = encrypted1[0]->parms_id();
        if (context_->last_parms_id() == parms_id)
            throw invalid_argument("end of modulus switching chain reached");
        auto &context_data = *context_->get_context_data(parms_id);
        for (size_t i = 0; i < n; i++)
        {
            validate(*encrypted1[i], "encrypted1");
            validate(*encrypted2[i], "encrypted2");
            if (encrypted1[i]->parms_id() != parms_id || encrypted2[i]->parms_id() != parms_id)
                throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
            if (!encrypted1[i]->is_ntt_form() || !encrypted2[i]->is_ntt_form() || encrypted1[i]->size() != 2 || encrypted2[i]->size() != 2)
                throw invalid_argument("operands must be size-2 NTT-form ciphertexts");
            const double s = encrypted1[i]->scale() * encrypted2[i]->scale();
            if (s <= 0 || (static_cast<int>(log2(s)) >= context_data.total_coeff_modulus_bit_count()))
                throw invalid_argument("scale out of bounds");
        }
        ensure_relin(relin_keys, RelinKeys::get_index(2));
        const size_t L = context_data.parms().coeff_modulus().size(), in_w = 2 * L * N_, out_w = 2 * (L - 1) * N_;
        void *ha = nullptr, *hb = nullptr, *ho = nullptr; // pinned staging so the copies are real DMA
        check(gsb_malloc_host(n * in_w * sizeof(uint64_t), &ha));
        int rc = gsb_malloc_host(n * in_w * sizeof(uint64_t), &hb);
        if (rc == GSB_OK)
            rc = gsb_malloc_host(n * out_w * sizeof(uint64_t), &ho);
        if (rc == GSB_OK)
        {
            for (size_t i = 0; i < n; i++)
            {
                memcpy(static_cast<uint64_t *>(ha) + i * in_w, encrypted1[i]->data(), in_w * sizeof(uint64_t));
                memcpy(static_cast<uint64_t *>(hb) + i * in_w, encrypted2[i]->data(), in_w * sizeof(uint64_t));
            }
            rc = gsb_mul_relin_rescale_host(
                ctx_, static_cast<uint64_t *>(ha), static_cast<uint64_t *>(hb), static_cast<uint64_t *>(ho), n, L);
        }
        if (rc == GSB_OK)
        {
            const double q_last = static_cast<double>(context_data.parms().coeff_modulus().back().value());
            auto next_id = context_data.next_context_data()->parms_id();
            for (size_t i = 0; i < n; i++)
            {
                const double s = encrypted1[i]->scale() * encrypted2[i]->scale() / q_last;
                encrypted1[i]->resize(context_, next_id, 2);
                memcpy(encrypted1[i]->data(), static_cast<uint64_t *>(ho) + i * out_w, out_w * sizeof(uint64_t));
                encrypted1[i]->is_ntt_form() = true;
                encrypted1[i]->scale() = s;
            }
        }
        gsb_free_host(ha);
        gsb_free_host(hb);
        gsb_free_host(ho);
        check(rc);
    }

    // ---- device-resident handles (CKKS) --------------------------------------------------
    // Same checks, same metadata updates and the same device entry points as the host-object overloads
    // above; the limbs never leave HBM between upload() and download().  Everything runs on the default
    // stream; download() synchronises.
    DeviceCiphertext::~DeviceCiphertext()
    {
        release();
    }
    DeviceCiphertext::DeviceCiphertext(DeviceCiphertext &&o) noexcept
    {
        *this = move(o);
    }
    DeviceCiphertext &DeviceCiphertext::operator=(DeviceCiphertext &&o) noexcept
    {
        if (this != &o)
        {
            release();
            ctx_ = o.ctx_, d_ = o.d_, cap_ = o.cap_, size_ = o.size_, L_ = o.L_, batch_ = o.batch_;
            parms_id_ = o.parms_id_, scale_ = o.scale_, ntt_ = o.ntt_;
            o.ctx_ = nullptr, o.d_ = nullptr, o.cap_ = o.size_ = o.L_ = 0, o.batch_ = 1;
        }
        return *this;
    }
    // Buffers of released handles are kept per device context and handed to the next handle that fits, so a chain
    // of handle operations does not pay a cudaMalloc + cudaFree (a device-wide synchronisation) per operation.
    // Everything runs on the default stream, hence reuse is ordered after the previous owner's kernels.
    namespace
    {
        mutex g_pool_mu;
        map<gsb_ctx *, vector<pair<uint64_t *, size_t>>> g_pool; // live contexts only
        constexpr size_t kPoolMax = 8;
    } // namespace

    void DeviceCiphertext::release()
    {
        if (d_)
        {
            lock_guard<mutex> lock(g_pool_mu);
            auto it = g_pool.find(ctx_);
            if (it != g_pool.end() && it->second.size() < kPoolMax)
                it->second.emplace_back(d_, cap_);
            else
                gsb_free(ctx_, d_);
        }
        d_ = nullptr;
        cap_ = 0;
    }
    void DeviceCiphertext::reserve(gsb_ctx *ctx, size_t words)
    {
        if (ctx_ == ctx && cap_ >= words)
            return;
        release();
        {
            lock_guard<mutex> lock(g_pool_mu);
            auto &v = g_pool[ctx];
            size_t best = v.size();
            for (size_t i = 0; i < v.size(); i++)
                if (v[i].second >= words && v[i].second <= 2 * words + 4096 && (best == v.size() || v[i].second < v[best].second))
                    best = i;
            if (best != v.size())
            {
                ctx_ = ctx, d_ = v[best].first, cap_ = v[best].second;
                v.erase(v.begin() + static_cast<ptrdiff_t>(best));
                return;
            }
        }
        void *p = nullptr;
        check(gsb_malloc(ctx, words * sizeof(uint64_t), &p));
        ctx_ = ctx, d_ = static_cast<uint64_t *>(p), cap_ = words;
    }
    static void drop_pool(gsb_ctx *ctx)
    {
        lock_guard<mutex> lock(g_pool_mu);
        auto it = g_pool.find(ctx);
        if (it == g_pool.end())
            return;
        for (auto &b : it->second)
            gsb_free(ctx, b.first);
        g_pool.erase(it); // handles that outlive their evaluator free their buffer directly
    }

    void Evaluator::own(const DeviceCiphertext &ct, const char *what) const
    {
        if (ct.ctx_ != ctx_ || ct.d_ == nullptr || ct.size_ < 2 || ct.batch_ < 1 || !context_->get_context_data(ct.parms_id_))
            throw invalid_argument(string(what) + " is not valid for encryption parameters");
    }

    void Evaluator::upload(const Ciphertext &encrypted, DeviceCiphertext &destination)
    {
        validate(encrypted, "encrypted");
        const size_t L = level_limbs(encrypted), words = encrypted.size() * L * N_;
        destination.reserve(ctx_, words);
        check(gsb_memcpy_h2d(ctx_, destination.d_, encrypted.data(), words * sizeof(uint64_t), nullptr));
        check(gsb_stream_sync(ctx_, nullptr)); // the host object may be changed once this returns
        destination.size_ = encrypted.size(), destination.L_ = L, destination.batch_ = 1;
        destination.parms_id_ = encrypted.parms_id(), destination.scale_ = encrypted.scale(), destination.ntt_ = encrypted.is_ntt_form();
    }

    void Evaluator::upload(const vector<const Ciphertext *> &encrypted, DeviceCiphertext &destination)
    {
        if (encrypted.empty() || encrypted[0] == nullptr)
            throw invalid_argument("encrypted vector must not be empty");
        const Ciphertext &first = *encrypted[0];
        for (const Ciphertext *ct : encrypted)
        {
            if (ct == nullptr)
                throw invalid_argument("encrypted is not valid for encryption parameters");
            validate(*ct, "encrypted");
            if (ct->parms_id() != first.parms_id() || ct->size() != first.size() || ct->is_ntt_form() != first.is_ntt_form() ||
                !seal::util::are_close<double>(ct->scale(), first.scale()))
                throw invalid_argument("batched ciphertexts must share level, size, form and scale");
        }
        const size_t L = level_limbs(first), words = first.size() * L * N_, n = encrypted.size();
        destination.reserve(ctx_, n * words);
        for (size_t i = 0; i < n; i++)
            check(gsb_memcpy_h2d(ctx_, destination.d_ + i * words, encrypted[i]->data(), words * sizeof(uint64_t), nullptr));
        check(gsb_stream_sync(ctx_, nullptr));
        destination.size_ = first.size(), destination.L_ = L, destination.batch_ = n;
        destination.parms_id_ = first.parms_id(), destination.scale_ = first.scale(), destination.ntt_ = first.is_ntt_form();
    }

    void Evaluator::download(const DeviceCiphertext &encrypted, vector<Ciphertext> &destination)
    {
        own(encrypted, "encrypted");
        const size_t words = encrypted.size_ * encrypted.L_ * N_;
        destination.resize(encrypted.batch_);
        for (size_t i = 0; i < encrypted.batch_; i++)
        {
            destination[i].resize(context_, encrypted.parms_id_, encrypted.size_);
            check(gsb_memcpy_d2h(ctx_, destination[i].data(), encrypted.d_ + i * words, words * sizeof(uint64_t), nullptr));
            destination[i].is_ntt_form() = encrypted.ntt_;
            destination[i].scale() = encrypted.scale_;
        }
        check(gsb_stream_sync(ctx_, nullptr));
    }

    void Evaluator::download(const DeviceCiphertext &encrypted, Ciphertext &destination)
    {
        own(encrypted, "encrypted");
        if (encrypted.batch_ != 1)
            throw invalid_argument("handle holds a batch: download into a vector");
        destination.resize(context_, encrypted.parms_id_, encrypted.size_);
        check(gsb_memcpy_d2h(ctx_, destination.data(), encrypted.d_, encrypted.size_ * encrypted.L_ * N_ * sizeof(uint64_t), nullptr));
        check(gsb_stream_sync(ctx_, nullptr));
        destination.is_ntt_form() = encrypted.ntt_;
        destination.scale() = encrypted.scale_;
    }

    void Evaluator::multiply_inplace(DeviceCiphertext &e1, const DeviceCiphertext &e2)
    {
        own(e1, "encrypted1");
        own(e2, "encrypted2");
        if (e1.parms_id_ != e2.parms_id_) // evaluator.cpp:238-249, 449-468
            throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
        if (bfv_) // bfv_multiply (evaluator.cpp:274-445) on coefficient-form handles
        {
            if (e1.ntt_ || e2.ntt_)
                throw invalid_argument("encrypted1 or encrypted2 cannot be in NTT form");
            if (e1.batch_ != e2.batch_)
                throw invalid_argument("encrypted1 and encrypted2 batch mismatch");
            auto &cd = *context_->get_context_data(e1.parms_id_);
            const size_t ds = seal::util::sub_safe(seal::util::add_safe(e1.size_, e2.size_), size_t(1));
            if (!seal::util::product_fits_in(ds, N_, cd.rns_tool()->base_Bsk_m_tilde()->size()))
                throw logic_error("invalid parameters");
            DeviceCiphertext out;
            out.reserve(ctx_, e1.batch_ * ds * e1.L_ * N_);
            check(gsb_bfv_multiply_sizes_dev(ctx_, nullptr, plain_modulus_, e1.d_, e1.size_, e2.d_, e2.size_, out.d_, e1.batch_, e1.L_, nullptr));
            out.size_ = ds, out.L_ = e1.L_, out.batch_ = e1.batch_, out.parms_id_ = e1.parms_id_, out.scale_ = e1.scale_, out.ntt_ = false;
            e1 = move(out);
            return;
        }
        if (!(e1.ntt_ && e2.ntt_))
            throw invalid_argument("encrypted1 or encrypted2 must be in NTT form");
        auto &context_data = *context_->get_context_data(e1.parms_id_);
        const double new_scale = e1.scale_ * e2.scale_;
        if (new_scale <= 0 || (static_cast<int>(log2(new_scale)) >= context_data.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        if (e1.batch_ != e2.batch_)
            throw invalid_argument("encrypted1 and encrypted2 batch mismatch");
        const size_t dest_size = seal::util::sub_safe(seal::util::add_safe(e1.size_, e2.size_), size_t(1));
        if (!seal::util::product_fits_in(dest_size, N_, e1.L_))
            throw logic_error("invalid parameters");
        DeviceCiphertext out;
        out.reserve(ctx_, e1.batch_ * dest_size * e1.L_ * N_);
        check(gsb_ckks_multiply_sizes_dev(ctx_, e1.d_, e1.size_, e2.d_, e2.size_, out.d_, e1.batch_, e1.L_, nullptr));
        out.size_ = dest_size, out.L_ = e1.L_, out.batch_ = e1.batch_, out.parms_id_ = e1.parms_id_, out.scale_ = new_scale, out.ntt_ = true;
        e1 = move(out);
    }

    void Evaluator::relinearize_inplace(DeviceCiphertext &e, const RelinKeys &relin_keys)
    {
        own(e, "encrypted");
        if (relin_keys.parms_id() != context_->key_parms_id()) // evaluator.cpp:776-796
            throw invalid_argument("relin_keys is not valid for encryption parameters");
        if (relin_keys.size() < (e.size_ >= 2 ? e.size_ - 2 : 0))
            throw invalid_argument("not enough relinearization keys");
        if (e.size_ <= 2)
            return;
        if (e.ntt_ == bfv_)
            throw logic_error(bfv_ ? "BFV encrypted cannot be in NTT form" : "CKKS encrypted must be in NTT form");
        for (size_t p = 2; p < e.size_; p++)
            ensure_relin(relin_keys, RelinKeys::get_index(p));
        DeviceCiphertext out;
        out.reserve(ctx_, e.batch_ * 2 * e.L_ * N_);
        if (e.size_ == 3)
            check(bfv_ ? gsb_bfv_relinearize_dev(ctx_, nullptr, e.d_, out.d_, e.batch_, e.L_, nullptr)
                       : gsb_relinearize_dev(ctx_, nullptr, e.d_, out.d_, e.batch_, e.L_, nullptr));
        else
            check(gsb_relinearize_sizes_dev(ctx_, nullptr, e.d_, e.size_, out.d_, e.batch_, e.L_, bfv_ ? 1 : 0, nullptr));
        out.size_ = 2, out.L_ = e.L_, out.batch_ = e.batch_, out.parms_id_ = e.parms_id_, out.scale_ = e.scale_, out.ntt_ = e.ntt_;
        e = move(out);
    }

    void Evaluator::rescale_to_next_inplace(DeviceCiphertext &e)
    {
        own(e, "encrypted");
        if (context_->last_parms_id() == e.parms_id_) // evaluator.cpp:1093-1104
            throw invalid_argument("end of modulus switching chain reached");
        if (bfv_) // evaluator.cpp:1106-1122
            throw invalid_argument("unsupported operation for scheme type");
        if (!e.ntt_)
            throw invalid_argument("CKKS encrypted must be in NTT form");
        auto &context_data = *context_->get_context_data(e.parms_id_);
        DeviceCiphertext out;
        out.reserve(ctx_, e.batch_ * e.size_ * (e.L_ - 1) * N_);
        check(gsb_rescale_dev(ctx_, nullptr, e.d_, out.d_, e.batch_, e.size_, e.L_, nullptr));
        out.size_ = e.size_, out.L_ = e.L_ - 1, out.batch_ = e.batch_, out.parms_id_ = context_data.next_context_data()->parms_id();
        out.scale_ = e.scale_ / static_cast<double>(context_data.parms().coeff_modulus().back().value()), out.ntt_ = true;
        e = move(out);
    }

    void Evaluator::apply_galois_inplace(DeviceCiphertext &e, uint32_t galois_elt, const GaloisKeys &galois_keys)
    {
        own(e, "encrypted");
        if (galois_keys.parms_id() != context_->key_parms_id()) // evaluator.cpp:1845-1887
            throw invalid_argument("galois_keys is not valid for encryption parameters");
        if (!galois_keys.has_key(galois_elt))
            throw invalid_argument("Galois key not present");
        if (!(galois_elt & 1) || galois_elt >= 2 * N_)
            throw invalid_argument("Galois element is not valid");
        if (e.size_ > 2)
            throw invalid_argument("encrypted size must be 2");
        if (e.ntt_ == bfv_)
            throw logic_error(bfv_ ? "BFV encrypted cannot be in NTT form" : "CKKS encrypted must be in NTT form");
        ensure_galois(galois_keys, galois_elt);
        DeviceCiphertext out;
        out.reserve(ctx_, e.batch_ * 2 * e.L_ * N_);
        check(bfv_ ? gsb_bfv_apply_galois_dev(ctx_, nullptr, galois_elt, e.d_, out.d_, e.batch_, e.L_, nullptr)
                   : gsb_apply_galois_dev(ctx_, nullptr, galois_elt, e.d_, out.d_, e.batch_, e.L_, nullptr));
        out.size_ = 2, out.L_ = e.L_, out.batch_ = e.batch_, out.parms_id_ = e.parms_id_, out.scale_ = e.scale_, out.ntt_ = e.ntt_;
        e = move(out);
    }

    // mod_switch_to_next (evaluator.cpp:996-1036): BFV divides and rounds by the last prime
    // (mod_switch_scale_to_next, :829-892), CKKS drops the last limb (mod_switch_drop_to_next, :894-957)
    void Evaluator::mod_switch_to_next_inplace(DeviceCiphertext &e)
    {
        own(e, "encrypted");
        if (context_->last_parms_id() == e.parms_id_)
            throw invalid_argument("end of modulus switching chain reached");
        auto &context_data = *context_->get_context_data(e.parms_id_);
        auto &next = *context_data.next_context_data();
        if (e.ntt_ == bfv_)
            throw invalid_argument(bfv_ ? "BFV encrypted cannot be in NTT form" : "CKKS encrypted must be in NTT form");
        if (!bfv_ && (e.scale_ <= 0 || (static_cast<int>(log2(e.scale_)) >= next.total_coeff_modulus_bit_count())))
            throw invalid_argument("scale out of bounds");
        DeviceCiphertext out;
        out.reserve(ctx_, e.batch_ * e.size_ * (e.L_ - 1) * N_);
        check(bfv_ ? gsb_bfv_mod_switch_dev(ctx_, e.d_, out.d_, e.batch_, e.size_, e.L_, nullptr)
                   : gsb_drop_last_limb_dev(ctx_, e.d_, out.d_, e.batch_ * e.size_, e.L_, nullptr));
        out.size_ = e.size_, out.L_ = e.L_ - 1, out.batch_ = e.batch_, out.parms_id_ = next.parms_id(), out.scale_ = e.scale_, out.ntt_ = e.ntt_;
        e = move(out);
    }

    // multiply_inplace + relinearize_inplace + rescale_to_next_inplace of size-2 CKKS handles as ONE device pass
    // (gsb_mul_relin_rescale_dev: fused tail, nothing but the result leaves the pipeline); the checks of the three
    // reference entry points in their order, the metadata of their composition.
    void Evaluator::multiply_relinearize_rescale_inplace(DeviceCiphertext &e1, const DeviceCiphertext &e2, const RelinKeys &relin_keys)
    {
        own(e1, "encrypted1");
        own(e2, "encrypted2");
        if (bfv_)
            throw invalid_argument("unsupported operation for scheme type");
        if (e1.parms_id_ != e2.parms_id_)
            throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
        if (!(e1.ntt_ && e2.ntt_))
            throw invalid_argument("encrypted1 or encrypted2 must be in NTT form");
        auto &context_data = *context_->get_context_data(e1.parms_id_);
        const double new_scale = e1.scale_ * e2.scale_;
        if (new_scale <= 0 || (static_cast<int>(log2(new_scale)) >= context_data.total_coeff_modulus_bit_count()))
            throw invalid_argument("scale out of bounds");
        if (e1.size_ != 2 || e2.size_ != 2 || e1.batch_ != e2.batch_)
            throw invalid_argument("the fused pass serves size-2 operands of equal batch");
        if (relin_keys.parms_id() != context_->key_parms_id())
            throw invalid_argument("relin_keys is not valid for encryption parameters");
        if (relin_keys.size() < 1)
            throw invalid_argument("not enough relinearization keys");
        if (context_->last_parms_id() == e1.parms_id_)
            throw invalid_argument("end of modulus switching chain reached");
        ensure_relin(relin_keys, RelinKeys::get_index(2));
        DeviceCiphertext out;
        out.reserve(ctx_, e1.batch_ * 2 * (e1.L_ - 1) * N_);
        check(gsb_mul_relin_rescale_dev(ctx_, nullptr, e1.d_, e2.d_, out.d_, e1.batch_, e1.L_, nullptr));
        out.size_ = 2, out.L_ = e1.L_ - 1, out.batch_ = e1.batch_, out.parms_id_ = context_data.next_context_data()->parms_id();
        out.scale_ = new_scale / static_cast<double>(context_data.parms().coeff_modulus().back().value()), out.ntt_ = true;
        e1 = move(out);
    }

    void Evaluator::rotate_internal(DeviceCiphertext &e, int steps, const GaloisKeys &galois_keys)
    {
        auto context_data_ptr = context_->get_context_data(e.parms_id_); // evaluator.cpp:1945-2000
        if (!context_data_ptr->qualifiers().using_batching)
            throw logic_error("encryption parameters do not support batching");
        if (galois_keys.parms_id() != context_->key_parms_id())
            throw invalid_argument("galois_keys is not valid for encryption parameters");
        if (steps == 0)
            return;
        auto galois_tool = context_data_ptr->galois_tool();
        if (galois_keys.has_key(galois_tool->get_elt_from_step(steps)))
        {
            apply_galois_inplace(e, galois_tool->get_elt_from_step(steps), galois_keys);
            return;
        }
        const vector<int> naf_steps = seal::util::naf(steps);
        if (naf_steps.size() == 1)
            throw invalid_argument("Galois key not present");
        for (int step : naf_steps)
            if (static_cast<size_t>(abs(step)) != (N_ >> 1))
                rotate_internal(e, step, galois_keys);
    }

    void Evaluator::rotate_vector_inplace(DeviceCiphertext &e, int steps, const GaloisKeys &galois_keys)
    {
        own(e, "encrypted");
        rotate_internal(e, steps, galois_keys);
    }

    void Evaluator::complex_conjugate_inplace(DeviceCiphertext &e, const GaloisKeys &galois_keys)
    {
        own(e, "encrypted");
        apply_galois_inplace(e, static_cast<uint32_t>(2 * N_ - 1), galois_keys); // evaluator.h:1269-1277
    }

    void Evaluator::add_inplace(DeviceCiphertext &e1, const DeviceCiphertext &e2)
    {
        own(e1, "encrypted1");
        own(e2, "encrypted2");
        if (e1.parms_id_ != e2.parms_id_) // evaluator.cpp:90-152
            throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
        if (e1.ntt_ != e2.ntt_)
            throw invalid_argument("NTT form mismatch");
        if (!seal::util::are_close<double>(e1.scale_, e2.scale_))
            throw invalid_argument("scale mismatch");
        if (e1.size_ != e2.size_ || e1.batch_ != e2.batch_)
            throw logic_error("device-resident add / sub serve operands of equal size and batch");
        check(gsb_poly_op_dev(ctx_, GSB_OP_ADD, e1.d_, e2.d_, nullptr, e1.d_, e1.batch_ * e1.size_, e1.L_, 0, nullptr));
    }

    void Evaluator::sub_inplace(DeviceCiphertext &e1, const DeviceCiphertext &e2)
    {
        own(e1, "encrypted1");
        own(e2, "encrypted2");
        if (e1.parms_id_ != e2.parms_id_) // evaluator.cpp:174-233
            throw invalid_argument("encrypted1 and encrypted2 parameter mismatch");
        if (e1.ntt_ != e2.ntt_)
            throw invalid_argument("NTT form mismatch");
        if (!seal::util::are_close<double>(e1.scale_, e2.scale_))
            throw invalid_argument("scale mismatch");
        if (e1.size_ != e2.size_ || e1.batch_ != e2.batch_)
            throw logic_error("device-resident add / sub serve operands of equal size and batch");
        check(gsb_poly_op_dev(ctx_, GSB_OP_SUB, e1.d_, e2.d_, nullptr, e1.d_, e1.batch_ * e1.size_, e1.L_, 0, nullptr));
    }

    void Evaluator::negate_inplace(DeviceCiphertext &e)
    {
        own(e, "encrypted");
        check(gsb_poly_op_dev(ctx_, GSB_OP_NEGATE, e.d_, nullptr, nullptr, e.d_, e.batch_ * e.size_, e.L_, 0, nullptr));
    }

    // ---- KeyGenerator (keygenerator.cpp:51-180, 325-420) ---------------------------------------------------------
    static gsb_ctx *keygen_ctx(const shared_ptr<SEALContext> &context, const SecretKey &sk, int device, size_t &N, size_t &K, size_t &beta)
    {
        auto &key_parms = context->key_context_data()->parms();
        if (!context->using_keyswitching())
            throw logic_error("keyswitching is not supported by the context");
        N = key_parms.poly_modulus_degree();
        K = key_parms.coeff_modulus().size();
        const size_t L0 = context->first_context_data()->parms().coeff_modulus().size(), k = K - L0;
        beta = (L0 + k - 1) / k;
        vector<uint64_t> moduli;
        for (auto &m : key_parms.coeff_modulus())
            moduli.push_back(m.value());
        int logn = 0;
        while ((size_t(1) << logn) < N)
            logn++;
        gsb_ctx *c = nullptr;
        check(gsb_ctx_create(device, logn, moduli.data(), K, k, &c));
        const int rc = gsb_secret_key_set(c, sk.data().data());
        if (rc != GSB_OK)
        {
            gsb_ctx_destroy(c);
            raise(rc);
        }
        return c;
    }

    KeyGenerator::KeyGenerator(shared_ptr<SEALContext> context, int device) : context_(context), host_(context)
    {
        ctx_ = keygen_ctx(context_, host_.secret_key(), device, N_, K_, beta_);
    }

    KeyGenerator::KeyGenerator(shared_ptr<SEALContext> context, const SecretKey &secret_key, int device) : context_(context), host_(context, secret_key)
    {
        ctx_ = keygen_ctx(context_, host_.secret_key(), device, N_, K_, beta_);
    }

    KeyGenerator::~KeyGenerator()
    {
        gsb_ctx_destroy(ctx_);
    }

    // noise is secret-key-equivalent material: cleared through a volatile pointer so the stores are not elided
    static void wipe(void *p, size_t bytes)
    {
        volatile unsigned char *v = static_cast<volatile unsigned char *>(p);
        for (size_t i = 0; i < bytes; i++)
            v[i] = 0;
    }

    // destinations[n] = the bundle vector of key n (KSwitchKeys::data()[index])
    void KeyGenerator::generate(int kind, const vector<uint32_t> &ids, vector<vector<PublicKey> *> &destinations)
    {
        auto &key_cd = *context_->key_context_data();
        auto &parms = key_cd.parms();
        const size_t n = ids.size(), KN = K_ * N_;
        if (n == 0)
            return;
        // the reference's draws, in its order per bundle: a from a fresh Blake PRNG, then e from the parameters'
        // generator (util/rlwe.cpp:232-258)
        vector<uint64_t> a(n * beta_ * KN);
        vector<int64_t> e(n * beta_ * N_);
        const uint64_t q0 = parms.coeff_modulus()[0].value();
        // every bundle draws from its own freshly seeded generators, so the rows are sampled by a few host threads
        const size_t rows = n * beta_;
        const size_t n_threads = max<size_t>(1, min<size_t>({ rows, size_t(thread::hardware_concurrency()), size_t(16) }));
        atomic<size_t> next{ 0 };
        atomic<bool> failed{ false };
        auto worker = [&] {
            try
            {
                vector<uint64_t> noise(KN);
                for (size_t r = next.fetch_add(1); r < rows; r = next.fetch_add(1))
                {
                    auto rng_error = parms.random_generator()->create();
                    shared_ptr<UniformRandomGenerator> rng_ciphertext = BlakePRNGFactory().create();
                    seal::util::sample_poly_uniform(rng_ciphertext, parms, a.data() + r * KN);
                    seal::util::sample_poly_normal(rng_error, parms, noise.data());
                    for (size_t d = 0; d < N_; d++) // limb 0 holds e or q_0 - |e| (rlwe.cpp:68-84)
                        e[r * N_ + d] = noise[d] > (q0 >> 1) ? -static_cast<int64_t>(q0 - noise[d]) : static_cast<int64_t>(noise[d]);
                }
                wipe(noise.data(), noise.size() * sizeof(uint64_t));
            }
            catch (...)
            {
                failed = true;
            }
        };
        vector<thread> pool;
        for (size_t t = 1; t < n_threads; t++)
            pool.emplace_back(worker);
        worker();
        for (auto &t : pool)
            t.join();
        if (failed)
            throw runtime_error("key generation: sampling failed");
        vector<uint64_t> out(n * beta_ * 2 * KN);
        check(gsb_kswitch_keygen_host(ctx_, kind, ids.data(), n, a.data(), e.data(), out.data(), 0));
        wipe(e.data(), e.size() * sizeof(int64_t));
        for (size_t i = 0; i < n; i++)
        {
            auto &dst = *destinations[i];
            dst.resize(beta_);
            for (size_t j = 0; j < beta_; j++)
            {
                Ciphertext &ct = dst[j].data(); // what encrypt_zero_symmetric leaves (rlwe.cpp:228-230)
                ct.resize(context_, key_cd.parms_id(), 2);
                ct.is_ntt_form() = true;
                ct.scale() = 1.0;
                memcpy(ct.data(), out.data() + (i * beta_ + j) * 2 * KN, 2 * KN * sizeof(uint64_t));
            }
        }
    }

    RelinKeys KeyGenerator::relin_keys_local()
    {
        RelinKeys keys; // keygenerator.cpp:138-180 with count = 1
        keys.data().resize(1);
        vector<vector<PublicKey> *> dst{ &keys.data()[0] };
        generate(GSB_KEY_RELIN, { 2 }, dst);
        keys.parms_id() = context_->key_parms_id();
        return keys;
    }

    GaloisKeys KeyGenerator::galois_keys_local(const vector<uint32_t> &galois_elts)
    {
        auto &context_data = *context_->key_context_data(); // keygenerator.cpp:51-116
        if (!context_data.qualifiers().using_batching)
            throw logic_error("encryption parameters do not support batching");
        GaloisKeys keys;
        keys.data().resize(N_);
        vector<uint32_t> ids;
        vector<vector<PublicKey> *> dst;
        for (auto elt : galois_elts)
        {
            if (!(elt & 1) || (elt >= N_ << 1))
                throw invalid_argument("Galois element is not valid");
            const size_t index = GaloisKeys::get_index(elt);
            bool seen = false;
            for (auto id : ids)
                seen = seen || id == elt;
            if (seen) // "do we already have the key?" (:93-96)
                continue;
            ids.push_back(elt);
            dst.push_back(&keys.data()[index]);
        }
        generate(GSB_KEY_GALOIS, ids, dst);
        keys.parms_id() = context_data.parms_id();
        return keys;
    }

    GaloisKeys KeyGenerator::galois_keys_local(const vector<int> &steps)
    {
        if (!context_->key_context_data()->qualifiers().using_batching) // keygenerator.h:176-183
            throw logic_error("encryption parameters do not support batching");
        return galois_keys_local(context_->key_context_data()->galois_tool()->get_elts_from_steps(steps));
    }

    GaloisKeys KeyGenerator::galois_keys_local()
    {
        return galois_keys_local(context_->key_context_data()->galois_tool()->get_elts_all());
    }

    // ---- Decryptor (decryptor.cpp:21-149) ------------------------------------------------------------------
    Decryptor::Decryptor(shared_ptr<SEALContext> context, const SecretKey &secret_key, int device) : context_(move(context))
    {
        // decryptor.cpp:24-37
        if (!context_)
            throw invalid_argument("invalid context");
        if (!context_->parameters_set())
            throw invalid_argument("encryption parameters are not set correctly");
        if (!is_valid_for(secret_key, context_))
            throw invalid_argument("secret key is not valid for encryption parameters");
        auto &key_parms = context_->key_context_data()->parms();
        if (key_parms.scheme() != scheme_type::CKKS && key_parms.scheme() != scheme_type::BFV)
            throw invalid_argument("unsupported scheme");
        bfv_ = key_parms.scheme() == scheme_type::BFV;
        plain_modulus_ = bfv_ ? key_parms.plain_modulus().value() : 0;
        N_ = key_parms.poly_modulus_degree();
        const size_t K = key_parms.coeff_modulus().size();
        size_t k = K - context_->first_context_data()->parms().coeff_modulus().size();
        if (K == 1)
            k = 1;
        vector<uint64_t> moduli;
        for (auto &m : key_parms.coeff_modulus())
            moduli.push_back(m.value());
        int logn = 0;
        while ((size_t(1) << logn) < N_)
            logn++;
        check(gsb_ctx_create(device, logn, moduli.data(), K, k, &ctx_));
        const int rc = gsb_secret_key_set(ctx_, secret_key.data().data()); // decryptor.cpp:46-57
        if (rc != GSB_OK)
        {
            gsb_ctx_destroy(ctx_);
            ctx_ = nullptr;
            raise(rc);
        }
    }

    Decryptor::~Decryptor()
    {
        gsb_ctx_destroy(ctx_);
    }

    void Decryptor::decrypt(const Ciphertext &encrypted, Plaintext &destination)
    {
        // decryptor.cpp:60-75
        if (!is_valid_for(encrypted, context_))
            throw invalid_argument("encrypted is not valid for encryption parameters");
        auto &context_data = *context_->get_context_data(encrypted.parms_id());
        const size_t L = context_data.parms().coeff_modulus().size();
        if (bfv_)
        {
            if (encrypted.is_ntt_form()) // :78-81
                throw invalid_argument("encrypted cannot be in NTT form");
            vector<uint64_t> tmp(N_);
            check(gsb_bfv_decrypt_host(ctx_, plain_modulus_, encrypted.data(), tmp.data(), 1, encrypted.size(), L));
            destination.parms_id() = parms_id_zero;
            destination.resize(N_); // :101
            memcpy(destination.data(), tmp.data(), N_ * sizeof(uint64_t));
            const size_t plain_coeff_count = seal::util::get_significant_uint64_count_uint(destination.data(), N_); // :107
            destination.resize(max(plain_coeff_count, size_t(1)));                                               // :110
            destination.parms_id() = parms_id_zero;
            return;
        }
        if (!encrypted.is_ntt_form()) // :117-120
            throw invalid_argument("encrypted must be in NTT form");
        destination.parms_id() = parms_id_zero; // :139
        destination.resize(L * N_);             // :142
        check(gsb_dot_product_ct_sk_host(ctx_, encrypted.data(), destination.data(), 1, encrypted.size(), L, 1));
        destination.parms_id() = encrypted.parms_id(); // :148-149
        destination.scale() = encrypted.scale();
    }

    // ---- util::ntt_* over NTTTables (util/ntt.h:189-368) -------------------------------------------------
    namespace util
    {
        namespace
        {
            mutex g_mu;
            map<pair<int, uint64_t>, gsb_ctx *> g_ctx; // one single-modulus device context per (log2 N, q)

            gsb_ctx *ctx_for(const seal::util::NTTTables &t)
            {
                lock_guard<mutex> lock(g_mu);
                const auto key = make_pair(t.coeff_count_power(), t.modulus().value());
                auto it = g_ctx.find(key);
                if (it != g_ctx.end())
                    return it->second;
                gsb_ctx *c = nullptr;
                const uint64_t q = key.second;
                check(gsb_ctx_create(0, key.first, &q, 1, 1, &c));
                g_ctx[key] = c;
                return c;
            }

            // one multi-modulus device context per (log2 N, moduli of the tables array): an RNS or Poly overload is
            // ONE batched device pass over all its limbs (and polynomials), not a call per limb
            map<pair<int, vector<uint64_t>>, gsb_ctx *> g_rns_ctx;

            gsb_ctx *ctx_for(seal::util::ConstNTTTablesIter tables, size_t count)
            {
                if (count == 0)
                    throw invalid_argument("coeff_modulus_size");
                const seal::util::NTTTables *t = *tables;
                if (count == 1)
                    return ctx_for(t[0]);
                vector<uint64_t> q(count);
                for (size_t i = 0; i < count; i++)
                {
                    if (t[i].coeff_count_power() != t[0].coeff_count_power())
                        throw invalid_argument("tables");
                    q[i] = t[i].modulus().value();
                }
                lock_guard<mutex> lock(g_mu);
                const auto key = make_pair(t[0].coeff_count_power(), q);
                auto it = g_rns_ctx.find(key);
                if (it != g_rns_ctx.end())
                    return it->second;
                gsb_ctx *c = nullptr;
                check(gsb_ctx_create(0, key.first, q.data(), count, 1, &c));
                g_rns_ctx[key] = c;
                return c;
            }

            void run(bool inverse, int mode, uint64_t *data, size_t n_polys, size_t n_limbs, gsb_ctx *c)
            {
                check(inverse ? gsb_ntt_inverse_host(c, data, n_polys, n_limbs, 0, mode) : gsb_ntt_forward_host(c, data, n_polys, n_limbs, 0, mode));
            }
            // a PolyIter walks polynomials that lie back to back (step = N * coeff_modulus_size, iterator.h:1099-1102)
            void run_poly(bool inverse, int mode, seal::util::PolyIter operand, size_t size, seal::util::ConstNTTTablesIter tables)
            {
                if (size == 0)
                    return;
                run(inverse, mode, **(*operand), size, operand.coeff_modulus_size(), ctx_for(tables, operand.coeff_modulus_size()));
            }
        } // namespace

        // util/ntt.h:189-368: CoeffIter / RNSIter / PolyIter x lazy / strict x forward / inverse
        void ntt_negacyclic_harvey_lazy(seal::util::CoeffIter operand, const seal::util::NTTTables &tables)
        {
            run(false, GSB_NTT_LAZY, *operand, 1, 1, ctx_for(tables));
        }
        void ntt_negacyclic_harvey_lazy(seal::util::RNSIter operand, size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables)
        {
            run(false, GSB_NTT_LAZY, **operand, 1, coeff_modulus_size, ctx_for(tables, coeff_modulus_size));
        }
        void ntt_negacyclic_harvey_lazy(seal::util::PolyIter operand, size_t size, seal::util::ConstNTTTablesIter tables)
        {
            run_poly(false, GSB_NTT_LAZY, operand, size, tables);
        }
        void ntt_negacyclic_harvey(seal::util::CoeffIter operand, const seal::util::NTTTables &tables)
        {
            run(false, GSB_NTT_STRICT, *operand, 1, 1, ctx_for(tables));
        }
        void ntt_negacyclic_harvey(seal::util::RNSIter operand, size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables)
        {
            run(false, GSB_NTT_STRICT, **operand, 1, coeff_modulus_size, ctx_for(tables, coeff_modulus_size));
        }
        void ntt_negacyclic_harvey(seal::util::PolyIter operand, size_t size, seal::util::ConstNTTTablesIter tables)
        {
            run_poly(false, GSB_NTT_STRICT, operand, size, tables);
        }
        void inverse_ntt_negacyclic_harvey_lazy(seal::util::CoeffIter operand, const seal::util::NTTTables &tables)
        {
            run(true, GSB_NTT_LAZY, *operand, 1, 1, ctx_for(tables));
        }
        void inverse_ntt_negacyclic_harvey_lazy(seal::util::RNSIter operand, size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables)
        {
            run(true, GSB_NTT_LAZY, **operand, 1, coeff_modulus_size, ctx_for(tables, coeff_modulus_size));
        }
        void inverse_ntt_negacyclic_harvey_lazy(seal::util::PolyIter operand, size_t size, seal::util::ConstNTTTablesIter tables)
        {
            run_poly(true, GSB_NTT_LAZY, operand, size, tables);
        }
        void inverse_ntt_negacyclic_harvey(seal::util::CoeffIter operand, const seal::util::NTTTables &tables)
        {
            run(true, GSB_NTT_STRICT, *operand, 1, 1, ctx_for(tables));
        }
        void inverse_ntt_negacyclic_harvey(seal::util::RNSIter operand, size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables)
        {
            run(true, GSB_NTT_STRICT, **operand, 1, coeff_modulus_size, ctx_for(tables, coeff_modulus_size));
        }
        void inverse_ntt_negacyclic_harvey(seal::util::PolyIter operand, size_t size, seal::util::ConstNTTTablesIter tables)
        {
            run_poly(true, GSB_NTT_STRICT, operand, size, tables);
        }
    } // namespace util
} // namespace gsb_seal
