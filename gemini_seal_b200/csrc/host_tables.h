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

        // Galois permutation for NTT-form polynomials (util/galois.cpp:18-47)
        void galois_table_ntt(int logn, uint32_t elt, std::vector<uint32_t> &table);
        // util/galois.cpp:49-91; returns 0 for an out-of-range step
        uint32_t galois_elt_from_step(int logn, int step);
    } // namespace host
} // namespace gsb
