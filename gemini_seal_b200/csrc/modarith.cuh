// 64-bit modular arithmetic on the 32-bit integer pipes, mirroring the reference word for word so
// every intermediate u64 equals the CPU's.  Reference paths are relative to
// /root/reference/native/src/seal/.
#pragma once
#include "compat.cuh"

namespace gsb
{
    // Per-modulus constants (device copy of what Modulus / NTTTables hold).
    struct LimbC
    {
        uint64_t q;       // modulus value
        uint64_t two_q;   // Lp of SlothfulNTT (util/ntt.cpp:214)
        uint64_t rdp;     // floor(2^64/q)  (NTTTables::reduce_precomp_, util/ntt.cpp:75)
        uint64_t cr0;     // const_ratio[0] (modulus.cpp:87-98)
        uint64_t cr1;     // const_ratio[1]
        uint64_t inv_n;   // n^-1 mod q     (util/ntt.cpp:69-73)
        uint64_t inv_n_s; // floor(inv_n * 2^64 / q)
        uint64_t pad;
    };

    GSB_DEV uint64_t mulhi(uint64_t a, uint64_t b)
    {
        return __umul64hi(a, b);
    }

    // SlothfulNTT::mulmodLazy (util/ntt.cpp:230-234): x*y - hi64(x*y')*q  (mod 2^64), in [0,2q).
    GSB_DEV uint64_t mulmod_lazy(uint64_t x, uint64_t y, uint64_t ys, uint64_t q)
    {
        // x*y - h*q == x*y + h*(2^64 - q) (mod 2^64): one multiply-add chain, no negate
        return x * y + mulhi(x, ys) * (0 - q);
    }

    // SlothfulNTT::reduceBarrettLazy (util/ntt.cpp:237-241)
    GSB_DEV uint64_t reduce_barrett_lazy(uint64_t x, uint64_t rdp, uint64_t q)
    {
        return x + mulhi(x, rdp) * (0 - q);
    }

    // MulModShoup::operator() (multi_special_primes.cpp:13-19) / FastMulMod (uintarithsmallmod.h:312-322)
    GSB_DEV uint64_t mulmod_shoup(uint64_t x, uint64_t c, uint64_t cs, uint64_t p)
    {
        uint64_t t = x * c + mulhi(x, cs) * (0 - p);
        return t - (t >= p ? p : 0);
    }

    // barrett_reduce_128 (util/uintarithsmallmod.h:140-178): literal 3-multiply form.
    GSB_DEV uint64_t barrett128(uint64_t z0, uint64_t z1, uint64_t q, uint64_t cr0, uint64_t cr1)
    {
        uint64_t carry = mulhi(z0, cr0);
        uint64_t lo = z0 * cr1, hi = mulhi(z0, cr1);
        uint64_t tmp1 = lo + carry;
        uint64_t tmp3 = hi + (tmp1 < carry ? 1 : 0);
        lo = z1 * cr0;
        hi = mulhi(z1, cr0);
        uint64_t s = tmp1 + lo;
        carry = hi + (s < tmp1 ? 1 : 0);
        uint64_t quot = z1 * cr1 + tmp3 + carry;
        uint64_t r = z0 + quot * (0 - q);
        return r - (r >= q ? q : 0);
    }

    // barrett_reduce_63 (util/uintarithsmallmod.h:181-207)
    GSB_DEV uint64_t barrett63(uint64_t x, uint64_t q, uint64_t cr1)
    {
        uint64_t r = x + mulhi(x, cr1) * (0 - q);
        return r - (r >= q ? q : 0);
    }

    // multiply_uint_mod (util/uintarithsmallmod.h:209-221); also the body of dyadic_product_coeffmod /
    // multiply_poly_scalar_coeffmod (util/polyarithsmallmod.cpp:32-60,89-116).
    GSB_DEV uint64_t mul_mod(uint64_t a, uint64_t b, uint64_t q, uint64_t cr0, uint64_t cr1)
    {
        return barrett128(a * b, mulhi(a, b), q, cr0, cr1);
    }

    GSB_DEV uint64_t add_mod(uint64_t a, uint64_t b, uint64_t q) // uintarithsmallmod.h:93-112
    {
        uint64_t s = a + b;
        return s - (s >= q ? q : 0);
    }

    GSB_DEV uint64_t sub_mod(uint64_t a, uint64_t b, uint64_t q) // uintarithsmallmod.h:114-137
    {
        uint64_t d = a - b;
        return d + (a < b ? q : 0);
    }

    GSB_DEV uint64_t neg_mod(uint64_t a, uint64_t q) // uintarithsmallmod.h:51-65
    {
        return a ? q - a : 0;
    }

    // FMAU128 (multi_special_primes.cpp:22-27): (lo,hi) += a*b as an exact 128-bit accumulate.
    GSB_DEV void mac128(uint64_t &lo, uint64_t &hi, uint64_t a, uint64_t b)
    {
        uint64_t pl = a * b, ph = mulhi(a, b);
        lo += pl;
        hi += ph + (lo < pl ? 1 : 0);
    }
} // namespace gsb
