// Drop-in C++ host shim: seal::Evaluator / seal::util::ntt_* entry points of the reference, served by
// the B200 library through the C ABI (include/gseal_b200.h).  Compiles against the reference's own
// headers (native/src); existing callers keep their seal::Ciphertext / RelinKeys / GaloisKeys /
// NTTTables objects and switch the evaluator type:
//
//     gsb_seal::Evaluator evaluator(context);          // was: seal::Evaluator evaluator(context);
//     evaluator.multiply_inplace(x, y);                // evaluator.h:246
//     evaluator.relinearize_inplace(x, relin_keys);    // evaluator.h:344
//     evaluator.rescale_to_next_inplace(x);            // evaluator.h:583
//     evaluator.rotate_vector_inplace(x, 1, gal_keys); // evaluator.h:1201
//
// Same argument meaning, same in-place semantics (resize, scale, parms_id, is_ntt_form), same
// exception types and messages as native/src/seal/evaluator.cpp; CUDA failures surface as
// std::runtime_error.  CKKS (the north-star path) and the BFV branches of multiply / relinearize /
// rotate_rows / rotate_columns (BASELINE config #1); anything else throws std::logic_error, it is never
// silently routed to a CPU implementation.
#pragma once
#include "gseal_b200.h"
#include "seal/seal.h"
#include "seal/util/iterator.h"
#include "seal/util/ntt.h"
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

namespace gsb_seal
{
    class Evaluator;

    // Device-resident ciphertext handle (no counterpart in the reference): the limbs
    // of a CKKS or BFV ciphertext stay in HBM across evaluator calls, only the metadata lives on the host.
    // Evaluator::upload validates exactly as the reference's evaluator validates its operands
    // (is_metadata_valid_for + is_buffer_valid); the overloads below then run multiply / relinearize /
    // rescale / rotate / add / sub / negate without a host copy per operation; Evaluator::download gives
    // back a seal::Ciphertext that is word for word what the same calls on seal::Evaluator produce.
    class DeviceCiphertext
    {
    public:
        DeviceCiphertext() = default;
        ~DeviceCiphertext();
        DeviceCiphertext(DeviceCiphertext &&other) noexcept;
        DeviceCiphertext &operator=(DeviceCiphertext &&other) noexcept;
        DeviceCiphertext(const DeviceCiphertext &) = delete;
        DeviceCiphertext &operator=(const DeviceCiphertext &) = delete;
        std::size_t size() const { return size_; }
        std::size_t batch() const { return batch_; } // ciphertexts held back to back (same level, size, scale)
        std::size_t coeff_modulus_size() const { return L_; }
        const seal::parms_id_type &parms_id() const { return parms_id_; }
        double scale() const { return scale_; }
        bool is_ntt_form() const { return ntt_; }
        const std::uint64_t *device_data() const { return d_; }

    private:
        friend class Evaluator;
        void reserve(gsb_ctx *ctx, std::size_t words);
        void release();
        gsb_ctx *ctx_ = nullptr;
        std::uint64_t *d_ = nullptr;
        std::size_t cap_ = 0, size_ = 0, L_ = 0, batch_ = 1;
        seal::parms_id_type parms_id_ = seal::parms_id_zero;
        double scale_ = 1.0;
        bool ntt_ = false;
    };

    class Evaluator
    {
    public:
        explicit Evaluator(std::shared_ptr<seal::SEALContext> context, int device = 0);
        ~Evaluator();
        Evaluator(const Evaluator &) = delete;
        Evaluator &operator=(const Evaluator &) = delete;

        // evaluator.h:246-258 / evaluator.cpp:235-272,447-527
        void multiply_inplace(
            seal::Ciphertext &encrypted1, const seal::Ciphertext &encrypted2,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void multiply(
            const seal::Ciphertext &encrypted1, const seal::Ciphertext &encrypted2, seal::Ciphertext &destination,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:72-206 / evaluator.cpp:65-233: negate, add, sub
        void negate_inplace(seal::Ciphertext &encrypted);
        void add_inplace(seal::Ciphertext &encrypted1, const seal::Ciphertext &encrypted2);
        void sub_inplace(seal::Ciphertext &encrypted1, const seal::Ciphertext &encrypted2);
        // evaluator.h:276-300 / evaluator.cpp:529-558: square = multiply(x, x)
        void square_inplace(seal::Ciphertext &encrypted, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        // evaluator.h:780-815 / evaluator.cpp:1438-1475,1605-1646: NTT-form plaintext multiplication
        void multiply_plain_inplace(
            seal::Ciphertext &encrypted, const seal::Plaintext &plain, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        // evaluator.h:700-770 / evaluator.cpp:1290-1437: plaintext addition / subtraction (CKKS: NTT-form plaintext;
        // BFV: multiply_add / sub_plain_with_scaling_variant)
        void add_plain_inplace(seal::Ciphertext &encrypted, const seal::Plaintext &plain);
        void sub_plain_inplace(seal::Ciphertext &encrypted, const seal::Plaintext &plain);
        // evaluator.h:208-220 / evaluator.cpp:153-172
        void add_many(const std::vector<seal::Ciphertext> &encrypteds, seal::Ciphertext &destination);
        // evaluator.h:460-540 / evaluator.cpp:1038-1088, 958-994: level switches to a given parms_id, and of plaintexts
        void mod_switch_to_inplace(
            seal::Ciphertext &encrypted, seal::parms_id_type parms_id, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void mod_switch_to_next_inplace(seal::Plaintext &plain);
        void mod_switch_to_inplace(seal::Plaintext &plain, seal::parms_id_type parms_id);
        // evaluator.h:820-856 / evaluator.cpp:1648-1744: BFV plaintext to NTT form at a level
        void transform_to_ntt_inplace(
            seal::Plaintext &plain, seal::parms_id_type parms_id, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        // evaluator.h:858-930 / evaluator.cpp:1746-1840
        void transform_to_ntt_inplace(seal::Ciphertext &encrypted);
        void transform_from_ntt_inplace(seal::Ciphertext &encrypted_ntt);
        // evaluator.h:430-470 / evaluator.cpp:894-957,996-1036 (CKKS: drop the last limb)
        void mod_switch_to_next_inplace(seal::Ciphertext &encrypted, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        // evaluator.h:600-640 / evaluator.cpp:1128-1178
        void rescale_to_inplace(
            seal::Ciphertext &encrypted, seal::parms_id_type parms_id, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:344-378 / evaluator.cpp:772-827
        void relinearize_inplace(
            seal::Ciphertext &encrypted, const seal::RelinKeys &relin_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void relinearize(
            const seal::Ciphertext &encrypted, const seal::RelinKeys &relin_keys, seal::Ciphertext &destination,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:480-520 / evaluator.cpp:1180-1288 (BFV only, as in the reference).  The reference's
        // product queue is evaluated wave by wave: all products whose operands exist go to the device
        // as ONE batched multiply + relinearize, so a product of n ciphertexts costs ceil(log2 n) + 1
        // device passes instead of n - 1.
        void multiply_many(
            const std::vector<seal::Ciphertext> &encrypteds, const seal::RelinKeys &relin_keys, seal::Ciphertext &destination,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void exponentiate_inplace(
            seal::Ciphertext &encrypted, std::uint64_t exponent, const seal::RelinKeys &relin_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:565-600 / evaluator.cpp:1090-1126,829-892
        void rescale_to_next(
            const seal::Ciphertext &encrypted, seal::Ciphertext &destination,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void rescale_to_next_inplace(seal::Ciphertext &encrypted, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:984-1023 / evaluator.cpp:1841-1943
        void apply_galois_inplace(
            seal::Ciphertext &encrypted, std::uint32_t galois_elt, const seal::GaloisKeys &galois_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void apply_galois(
            const seal::Ciphertext &encrypted, std::uint32_t galois_elt, const seal::GaloisKeys &galois_keys,
            seal::Ciphertext &destination, seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:1201-1252 / evaluator.cpp:1945-2000 (NAF decomposition when a key is missing)
        void rotate_vector_inplace(
            seal::Ciphertext &encrypted, int steps, const seal::GaloisKeys &galois_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void rotate_vector(
            const seal::Ciphertext &encrypted, int steps, const seal::GaloisKeys &galois_keys, seal::Ciphertext &destination,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // BFV: rotate_rows / rotate_columns (evaluator.h:1049-1176)
        void rotate_rows_inplace(
            seal::Ciphertext &encrypted, int steps, const seal::GaloisKeys &galois_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());
        void rotate_columns_inplace(
            seal::Ciphertext &encrypted, const seal::GaloisKeys &galois_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // evaluator.h:1269-1311
        void complex_conjugate_inplace(
            seal::Ciphertext &encrypted, const seal::GaloisKeys &galois_keys,
            seal::MemoryPoolHandle pool = seal::MemoryManager::GetPool());

        // Batched extension (not in the reference): multiply + relinearize + rescale_to_next on n
        // independent ciphertext pairs in one device pass; results replace encrypted1[i].
        void multiply_relinearize_rescale_inplace(
            std::vector<seal::Ciphertext *> &encrypted1, const std::vector<const seal::Ciphertext *> &encrypted2,
            const seal::RelinKeys &relin_keys);

        // ---- device-resident handles (CKKS) ------------------------------------------------------------
        void upload(const seal::Ciphertext &encrypted, DeviceCiphertext &destination);
        void download(const DeviceCiphertext &encrypted, seal::Ciphertext &destination);
        // Batched form: n ciphertexts of the same level, size, form and scale in one handle; every
        // overload below then runs ONE batched device pass over all of them (this is how the
        // throughput of bench.py is reached from C++).
        void upload(const std::vector<const seal::Ciphertext *> &encrypted, DeviceCiphertext &destination);
        void download(const DeviceCiphertext &encrypted, std::vector<seal::Ciphertext> &destination);
        void multiply_inplace(DeviceCiphertext &encrypted1, const DeviceCiphertext &encrypted2);
        void relinearize_inplace(DeviceCiphertext &encrypted, const seal::RelinKeys &relin_keys);
        void rescale_to_next_inplace(DeviceCiphertext &encrypted);
        void apply_galois_inplace(DeviceCiphertext &encrypted, std::uint32_t galois_elt, const seal::GaloisKeys &galois_keys);
        void rotate_vector_inplace(DeviceCiphertext &encrypted, int steps, const seal::GaloisKeys &galois_keys);
        void complex_conjugate_inplace(DeviceCiphertext &encrypted, const seal::GaloisKeys &galois_keys);
        void add_inplace(DeviceCiphertext &encrypted1, const DeviceCiphertext &encrypted2);
        void sub_inplace(DeviceCiphertext &encrypted1, const DeviceCiphertext &encrypted2);
        void negate_inplace(DeviceCiphertext &encrypted);
        void mod_switch_to_next_inplace(DeviceCiphertext &encrypted);
        // multiply + relinearize + rescale_to_next of size-2 CKKS handles as one fused device pass
        void multiply_relinearize_rescale_inplace(
            DeviceCiphertext &encrypted1, const DeviceCiphertext &encrypted2, const seal::RelinKeys &relin_keys);

        gsb_ctx *device_context() const { return ctx_; }
        // what the C layer's Evaluator_ContextUsingKeyswitching reports (c/evaluator.cpp:23-29)
        bool using_keyswitching() const { return context_->using_keyswitching(); }

    private:
        void validate(const seal::Ciphertext &ct, const char *what) const;
        void bfv_multiply(seal::Ciphertext &encrypted1, const seal::Ciphertext &encrypted2);
        void multiply_plain_normal(seal::Ciphertext &encrypted, const seal::Plaintext &plain);
        void add_sub_plain(seal::Ciphertext &encrypted, const seal::Plaintext &plain, bool subtract);
        void add_sub(seal::Ciphertext &encrypted1, const seal::Ciphertext &encrypted2, bool subtract);
        std::size_t level_limbs(const seal::Ciphertext &ct) const;
        void ensure_relin(const seal::RelinKeys &keys, std::size_t index);
        void ensure_galois(const seal::GaloisKeys &keys, std::uint32_t elt);
        void upload_key(int kind, std::uint32_t slot, const std::vector<seal::PublicKey> &bundles);
        void rotate_internal(seal::Ciphertext &encrypted, int steps, const seal::GaloisKeys &galois_keys);
        void rotate_internal(DeviceCiphertext &encrypted, int steps, const seal::GaloisKeys &galois_keys);
        void own(const DeviceCiphertext &ct, const char *what) const;
        [[noreturn]] void fail(int rc) const;

        std::shared_ptr<seal::SEALContext> context_;
        gsb_ctx *ctx_ = nullptr;
        std::size_t N_ = 0, K_ = 0, k_ = 0;
        bool bfv_ = false;
        std::uint64_t plain_modulus_ = 0;
        std::mutex mu_;
        // content fingerprint of the host key that currently sits in each device key slot
        std::map<std::uint32_t, std::uint64_t> relin_loaded_, galois_loaded_;
    };

    // seal::KeyGenerator's switching-key generation (keygenerator.h:82-230 / keygenerator.cpp:51-180, 325-420; SURVEY
    // 8(f) row f2).  The secret key (and public key) come from the reference's KeyGenerator; randomness comes from the
    // reference's own samplers (util::sample_poly_uniform / sample_poly_normal, util/rlwe.cpp:60-129) exactly as
    // encrypt_zero_symmetric draws it; the arithmetic -- noise NTT, -(a s + e), secret-key powers / Galois-permuted
    // key, the P * s_new add -- runs on the device, all keys of a call in one batched pass.  The returned objects
    // are ordinary seal::RelinKeys / seal::GaloisKeys.
    class KeyGenerator
    {
    public:
        explicit KeyGenerator(std::shared_ptr<seal::SEALContext> context, int device = 0);
        KeyGenerator(std::shared_ptr<seal::SEALContext> context, const seal::SecretKey &secret_key, int device = 0);
        ~KeyGenerator();
        KeyGenerator(const KeyGenerator &) = delete;
        KeyGenerator &operator=(const KeyGenerator &) = delete;
        const seal::SecretKey &secret_key() const { return host_.secret_key(); }
        seal::PublicKey public_key() const { return host_.public_key(); }
        seal::RelinKeys relin_keys_local();
        seal::GaloisKeys galois_keys_local(const std::vector<std::uint32_t> &galois_elts);
        seal::GaloisKeys galois_keys_local(const std::vector<int> &steps);
        seal::GaloisKeys galois_keys_local();

    private:
        void generate(int kind, const std::vector<std::uint32_t> &ids, std::vector<std::vector<seal::PublicKey> *> &destinations);
        std::shared_ptr<seal::SEALContext> context_;
        seal::KeyGenerator host_; // secret / public key and their checks
        gsb_ctx *ctx_ = nullptr;
        std::size_t N_ = 0, K_ = 0, beta_ = 0;
    };

    // seal::Decryptor (decryptor.h:62-120 / decryptor.cpp:21-149, 218-265): the secret
    // key's NTT-form limbs live on the device, decrypt() runs the dot product c_0 + c_1 s + ... there
    // (and, for BFV, RNSTool::decrypt_scale_and_round) and fills the same seal::Plaintext the reference
    // would.  invariant_noise_budget is not served.
    class Decryptor
    {
    public:
        Decryptor(std::shared_ptr<seal::SEALContext> context, const seal::SecretKey &secret_key, int device = 0);
        ~Decryptor();
        Decryptor(const Decryptor &) = delete;
        Decryptor &operator=(const Decryptor &) = delete;
        void decrypt(const seal::Ciphertext &encrypted, seal::Plaintext &destination);

    private:
        std::shared_ptr<seal::SEALContext> context_;
        gsb_ctx *ctx_ = nullptr;
        std::size_t N_ = 0;
        bool bfv_ = false;
        std::uint64_t plain_modulus_ = 0;
    };

    namespace util
    {
        // util/ntt.h:189-368: the four families (forward / inverse x lazy / strict) over CoeffIter (one limb), RNSIter
        // (one polynomial over coeff_modulus_size limbs) and PolyIter (size polynomials back to back), with the
        // reference's own parameter types, so callers of seal::util::ntt_* recompile unchanged.  Tables are matched
        // to a cached device context by (coeff_count_power, moduli); an RNS / Poly call is one batched device pass.
        void ntt_negacyclic_harvey_lazy(seal::util::CoeffIter operand, const seal::util::NTTTables &tables);
        void ntt_negacyclic_harvey_lazy(seal::util::RNSIter operand, std::size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables);
        void ntt_negacyclic_harvey_lazy(seal::util::PolyIter operand, std::size_t size, seal::util::ConstNTTTablesIter tables);
        void ntt_negacyclic_harvey(seal::util::CoeffIter operand, const seal::util::NTTTables &tables);
        void ntt_negacyclic_harvey(seal::util::RNSIter operand, std::size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables);
        void ntt_negacyclic_harvey(seal::util::PolyIter operand, std::size_t size, seal::util::ConstNTTTablesIter tables);
        void inverse_ntt_negacyclic_harvey_lazy(seal::util::CoeffIter operand, const seal::util::NTTTables &tables);
        void inverse_ntt_negacyclic_harvey_lazy(seal::util::RNSIter operand, std::size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables);
        void inverse_ntt_negacyclic_harvey_lazy(seal::util::PolyIter operand, std::size_t size, seal::util::ConstNTTTablesIter tables);
        void inverse_ntt_negacyclic_harvey(seal::util::CoeffIter operand, const seal::util::NTTTables &tables);
        void inverse_ntt_negacyclic_harvey(seal::util::RNSIter operand, std::size_t coeff_modulus_size, seal::util::ConstNTTTablesIter tables);
        void inverse_ntt_negacyclic_harvey(seal::util::PolyIter operand, std::size_t size, seal::util::ConstNTTTablesIter tables);
    } // namespace util
} // namespace gsb_seal
