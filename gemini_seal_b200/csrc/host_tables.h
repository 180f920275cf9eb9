// Host-side (C++) construction of everything the device context holds: per-modulus constants,
// NTT twiddle tables and the key-switch / rescale constants.  Mirrors what SEALContext builds
// (reference: native/src/seal/modulus.cpp:66-104, util/ntt.cpp:37-118, util/numth.cpp:398-425,
// multi_special_primes.cpp:113-126,186-234,292-299, util/rns.cpp:719-728) so the values uploaded
// to the GPU equal the reference's own tables word for word.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace gsb
{
    namespace host
    {
        typedef unsigned __int128 u128;

        struct ModConst
        {
            uint64_t q = 0, cr0 = 0, cr1 = 0, cr2 = 0;
        };

        ModConst make_mod(uint64_t q);
        uint64_t shoup(uint64_t x, uint64_t q);
        uint64_t mul_mod(uint64_t a, uint64_t b, const ModConst &m);
        uint64_t barrett63(uint64_t x, const ModConst &m);
        uint64_t pow_mod(uint64_t a, uint64_t e, const ModConst &m);
        bool inv_mod(uint64_t a, uint64_t q, uint64_t &out);
        uint32_t reverse_bits(uint32_t x, int bits);

        struct LimbTables
        {
            ModConst mod;
            uint64_t root = 0, inv_n = 0, inv_n_s = 0, rdp = 0;
            std::vector<uint64_t> w, ws, v, vs; // N words each
        };

        // false if q does not support a negacyclic NTT of size 2^logn
        bool build_limb_tables(int logn, uint64_t q, LimbTables &out);

        // RNSTool constants BEHZ multiplication needs for base q = q[0..L) and plain modulus t
        // (util/rns.cpp:540-728).  Extended base = q followed by Bsk = B u {m_sk}; M = L + nBsk.
        struct BehzConsts
        {
            size_t L = 0, nB = 0, nBsk = 0;
            uint64_t t = 0, m_sk = 0;
            std::vector<uint64_t> bsk;            // nBsk moduli: the |B| primes, then m_sk
            std::vector<uint64_t> inv_punct_q;    // [L]    (prod_{l != i} q_l)^-1 mod q_i
            std::vector<uint64_t> mat_q;          // [nBsk + 1][L]: prod_{l != i} q_l mod Bsk_j; last row mod m_tilde
            std::vector<uint64_t> inv_punct_b;    // [nB]
            std::vector<uint64_t> mat_b;          // [L + 1][nB]: prod_{l != i} B_l mod q_j; last row mod m_sk
            std::vector<uint64_t> prod_q_mod_bsk, inv_prod_q_mod_bsk, inv_mtilde_mod_bsk; // [nBsk]
            std::vector<uint64_t> prod_b_mod_q;   // [L]
            uint64_t inv_prod_q_mod_mtilde = 0, inv_prod_b_mod_msk = 0;
            // decrypt_scale_and_round (util/rns.cpp:585-590, 603-608, 690-717, 1070-1126); t > 0 only
            uint64_t gamma = 0, inv_gamma_mod_t = 0;
            std::vector<uint64_t> prod_tg_mod_q;  // [L]    t * gamma mod q_i
            std::vector<uint64_t> mat_tg;         // [2][L] prod_{l != i} q_l mod {t, gamma}
            uint64_t neg_inv_q_mod_tg[2] = { 0, 0 }; // -(prod q)^-1 mod {t, gamma}
        };
        // false when the auxiliary base cannot be built (not enough primes / not invertible)
        bool build_behz(int logn, const uint64_t *q, size_t L, uint64_t t, BehzConsts &out);

        // Galois permutation for NTT-form polynomials (util/galois.cpp:18-47)
        void galois_table_ntt(int logn, uint32_t elt, std::vector<uint32_t> &table);
        // util/galois.cpp:49-91; returns 0 for an out-of-range step
        uint32_t galois_elt_from_step(int logn, int step);
    } // namespace host
} // namespace gsb
