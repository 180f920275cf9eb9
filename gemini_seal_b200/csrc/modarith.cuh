// 64-bit modular arithmetic on the 32-bit integer pipes, mirroring the reference word for word so
// every intermediate u64 equals the CPU's.  Reference paths are relative to
// /root/reference/native/src/seal/.  Three helpers are NOT literal mirrors: mulmod_wide_add (congruent
// word in a wider range), mac128_mid and reduce_2k (same results as FMAU128 / barrett_reduce_128 under
// stated operand bounds); each says where it may be used.
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
        uint64_t pad;     // run-time zero (see fwd_round)
        // reduce_2k: k = bit length of q; mu = floor(2^(k+63) / q) (< 2^64), sh = k - 1; mu = 0 when k > 60
        uint64_t mu, sh;
    };

    // (mu, sh) of LimbC for modulus q (host side of reduce_2k)
    inline void limb_mu(uint64_t q, uint64_t &mu, uint64_t &sh)
    {
        int k = 0;
        for (uint64_t x = q; x; x >>= 1)
            k++;
        mu = 0;
        sh = static_cast<uint64_t>(k - 1);
        if (k >= 2 && k <= 60)
            mu = static_cast<uint64_t>((static_cast<unsigned __int128>(1) << (k + 63)) / q);
    }

    GSB_DEV uint64_t mulhi(uint64_t a, uint64_t b)
    {
        return __umul64hi(a, b);
    }

    // lo64(x*y + h*nq + add) as one multiply-add chain: 2 IMAD.WIDE + 4 IMAD, no separate carry / add
    // instructions (IMAD.WIDE runs at half the IMAD rate on sm_100 and the FMA-heavy pipe is what
    // bounds the butterflies, so nothing but multiplies may land on it).
    GSB_DEV uint64_t mad_lo64_2(uint64_t x, uint64_t y, uint64_t h, uint64_t nq, uint64_t add)
    {
#ifdef GSB_CUSIM
        return x * y + h * nq + add;
#else
        uint64_t r;
        asm("{\n\t"
            ".reg .u32 x0, x1, y0, y1, h0, h1, n0, n1, lo, hi;\n\t"
            ".reg .u64 t;\n\t"
            "mov.b64 {x0, x1}, %1;\n\t"
            "mov.b64 {y0, y1}, %2;\n\t"
            "mov.b64 {h0, h1}, %3;\n\t"
            "mov.b64 {n0, n1}, %4;\n\t"
            "mad.wide.u32 t, x0, y0, %5;\n\t"
            "mad.wide.u32 t, h0, n0, t;\n\t"
            "mov.b64 {lo, hi}, t;\n\t"
            "mad.lo.u32 hi, x0, y1, hi;\n\t"
            "mad.lo.u32 hi, x1, y0, hi;\n\t"
            "mad.lo.u32 hi, h0, n1, hi;\n\t"
            "mad.lo.u32 hi, h1, n0, hi;\n\t"
            "mov.b64 %0, {lo, hi};\n\t"
            "}"
            : "=l"(r)
            : "l"(x), "l"(y), "l"(h), "l"(nq), "l"(add));
        return r;
#endif
    }

    // lo64(h*nq + add): 1 IMAD.WIDE + 2 IMAD
    GSB_DEV uint64_t mad_lo64_1(uint64_t h, uint64_t nq, uint64_t add)
    {
#ifdef GSB_CUSIM
        return h * nq + add;
#else
        uint64_t r;
        asm("{\n\t"
            ".reg .u32 h0, h1, n0, n1, lo, hi;\n\t"
            ".reg .u64 t;\n\t"
            "mov.b64 {h0, h1}, %1;\n\t"
            "mov.b64 {n0, n1}, %2;\n\t"
            "mad.wide.u32 t, h0, n0, %3;\n\t"
            "mov.b64 {lo, hi}, t;\n\t"
            "mad.lo.u32 hi, h0, n1, hi;\n\t"
            "mad.lo.u32 hi, h1, n0, hi;\n\t"
            "mov.b64 %0, {lo, hi};\n\t"
            "}"
            : "=l"(r)
            : "l"(h), "l"(nq), "l"(add));
        return r;
#endif
    }

    // SlothfulNTT::mulmodLazy (util/ntt.cpp:230-234): x*y - hi64(x*y')*q  (mod 2^64), in [0,2q).
    // x*y - h*q == x*y + h*(2^64 - q) (mod 2^64).
    GSB_DEV uint64_t mulmod_lazy(uint64_t x, uint64_t y, uint64_t ys, uint64_t q)
    {
        return mad_lo64_2(x, y, mulhi(x, ys), 0 - q, 0);
    }

    // add + mulmodLazy(x, y, y') (mod 2^64): the butterfly's "u + v" folded into the multiply chain.
    GSB_DEV uint64_t mulmod_lazy_add(uint64_t x, uint64_t y, uint64_t ys, uint64_t q, uint64_t add)
    {
        return mad_lo64_2(x, y, mulhi(x, ys), 0 - q, add);
    }

    // Wide-range lazy multiply for transforms whose words never leave the library (see `FAST` in
    // ntt_core.cuh): add + x*y - h*(2q) (mod 2^64) with h an UNDER-estimate of floor(x*y/(2q)) built
    // from three 32x32 products instead of the four of an exact hi64 and with no carry instructions:
    //     ysh = y' >> 1 = floor(y 2^63 / q) < 2^63,     x < 2^63   (bound kept by the caller)
    //     t   = x0*ysh1 + x1*ysh0              (no 64-bit wrap: both products are < 2^63)
    //     h   = x1*ysh1 + (t >> 32)            in { H, H-1 },  H = floor(x*ysh / 2^64)
    // With Q = x*y/(2q):  Q - 1.5 < H <= Q  (x < 2^63), so Q - 2.5 < h <= Q and the product term
    // x*y - 2q*h lies in [0, 5q): congruent to the reference's mulmodLazy word, in a wider range.
    // 5 wide multiplies + 4 IMAD (the exact form is 6 + 4 and four more carry adds).  In SASS (sm_100a,
    // CUDA 12.9) ptxas issues the wide products without addends (IMAD.WIDE / IMAD.HI) and joins them with
    // three-input IADD3 pairs on the ALU pipe.  `zero` is a run-time zero (LimbC::pad): the sum
    // x1*ysh1 + (t >> 32) is written as 32-bit carry chains with both halves of `zero` as third inputs, because
    // ptxas turns the carry word of every TWO-input 64-bit add into an IMAD.X on the multiplier pipe (the pipe
    // that bounds these kernels; scripts/sass_pipes.py counts them) and can only use IADD3.X for a three-input one.
    // H3 = true: the three-input form above.  H3 = false: the high word of the third product's addend is `zero`'s low
    // half only (mad.wide with a 64-bit addend in PTX), which leaves ptxas free to put that carry on either pipe --
    // for kernels whose ALU pipe is the fuller one (ks_mac, see DESIGN.md "two-pipe model").
    // EARLY: the two narrow products that do not depend on h (x0*y1, x1*y0) go into the high word of the addend BEFORE
    // the wide chain, so only two narrow multiplies stay on the critical path behind h (same instruction counts; faster
    // in the tile kernels, slower in ks_mac, whose register budget is tighter).
    template <bool H3 = true, bool EARLY = true>
    GSB_DEV uint64_t mulmod_wide_add(uint64_t x, uint64_t y, uint64_t ysh, uint64_t neg_two_q, uint64_t add, uint64_t zero)
    {
#ifdef GSB_CUSIM
        GSB_SIM_ASSERT((x >> 63) == 0 && (ysh >> 63) == 0 && zero == 0);
        const uint64_t x0 = static_cast<uint32_t>(x), x1 = x >> 32, s0 = static_cast<uint32_t>(ysh), s1 = ysh >> 32;
        const uint64_t t = x0 * s1 + x1 * s0;
        const uint64_t h = x1 * s1 + (t >> 32);
        return x * y + h * neg_two_q + add;
#else
        uint64_t r;
        if constexpr (H3 && EARLY)
        {
            asm("{\n\t"
                ".reg .u32 x0, x1, y0, y1, s0, s1, h0, h1, n0, n1, lo, hi, z, z1, a0, a1;\n\t"
                ".reg .u64 t, c;\n\t"
                "mov.b64 {x0, x1}, %1;\n\t"
                "mov.b64 {y0, y1}, %2;\n\t"
                "mov.b64 {s0, s1}, %3;\n\t"
                "mov.b64 {n0, n1}, %4;\n\t"
                "mul.wide.u32 t, x0, s1;\n\t"
                "mad.wide.u32 t, x1, s0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mov.b64 {z, z1}, %6;\n\t"
                "mul.wide.u32 t, x1, s1;\n\t"
                "mov.b64 {h0, h1}, t;\n\t"
                "add.cc.u32 h0, h0, hi;\n\t"
                "addc.u32 h1, h1, z;\n\t"
                "add.cc.u32 h0, h0, z1;\n\t"
                "addc.u32 h1, h1, z1;\n\t"
                "mov.b64 {a0, a1}, %5;\n\t"
                "mad.lo.u32 a1, x0, y1, a1;\n\t"
                "mad.lo.u32 a1, x1, y0, a1;\n\t"
                "mov.b64 c, {a0, a1};\n\t"
                "mad.wide.u32 t, x0, y0, c;\n\t"
                "mad.wide.u32 t, h0, n0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mad.lo.u32 hi, h0, n1, hi;\n\t"
                "mad.lo.u32 hi, h1, n0, hi;\n\t"
                "mov.b64 %0, {lo, hi};\n\t"
                "}"
                : "=l"(r)
                : "l"(x), "l"(y), "l"(ysh), "l"(neg_two_q), "l"(add), "l"(zero));
        }
        else if constexpr (H3)
        {
            asm("{\n\t"
                ".reg .u32 x0, x1, y0, y1, s0, s1, h0, h1, n0, n1, lo, hi, z, z1, a0, a1;\n\t"
                ".reg .u64 t, c;\n\t"
                "mov.b64 {x0, x1}, %1;\n\t"
                "mov.b64 {y0, y1}, %2;\n\t"
                "mov.b64 {s0, s1}, %3;\n\t"
                "mov.b64 {n0, n1}, %4;\n\t"
                "mul.wide.u32 t, x0, s1;\n\t"
                "mad.wide.u32 t, x1, s0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mov.b64 {z, z1}, %6;\n\t"
                "mul.wide.u32 t, x1, s1;\n\t"
                "mov.b64 {h0, h1}, t;\n\t"
                "add.cc.u32 h0, h0, hi;\n\t"
                "addc.u32 h1, h1, z;\n\t"
                "add.cc.u32 h0, h0, z1;\n\t"
                "addc.u32 h1, h1, z1;\n\t"
                "mad.wide.u32 t, x0, y0, %5;\n\t"
                "mad.wide.u32 t, h0, n0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mad.lo.u32 hi, x0, y1, hi;\n\t"
                "mad.lo.u32 hi, x1, y0, hi;\n\t"
                "mad.lo.u32 hi, h0, n1, hi;\n\t"
                "mad.lo.u32 hi, h1, n0, hi;\n\t"
                "mov.b64 %0, {lo, hi};\n\t"
                "}"
                : "=l"(r)
                : "l"(x), "l"(y), "l"(ysh), "l"(neg_two_q), "l"(add), "l"(zero));
        }
        else if constexpr (EARLY)
        {
            asm("{\n\t"
                ".reg .u32 x0, x1, y0, y1, s0, s1, h0, h1, n0, n1, lo, hi, z, z1, a0, a1;\n\t"
                ".reg .u64 t, c;\n\t"
                "mov.b64 {x0, x1}, %1;\n\t"
                "mov.b64 {y0, y1}, %2;\n\t"
                "mov.b64 {s0, s1}, %3;\n\t"
                "mov.b64 {n0, n1}, %4;\n\t"
                "mul.wide.u32 t, x0, s1;\n\t"
                "mad.wide.u32 t, x1, s0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mov.b64 {z, z1}, %6;\n\t"
                "mov.b64 c, {hi, z};\n\t"
                "mad.wide.u32 t, x1, s1, c;\n\t"
                "mov.b64 {h0, h1}, t;\n\t"
                "mov.b64 {a0, a1}, %5;\n\t"
                "mad.lo.u32 a1, x0, y1, a1;\n\t"
                "mad.lo.u32 a1, x1, y0, a1;\n\t"
                "mov.b64 c, {a0, a1};\n\t"
                "mad.wide.u32 t, x0, y0, c;\n\t"
                "mad.wide.u32 t, h0, n0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mad.lo.u32 hi, h0, n1, hi;\n\t"
                "mad.lo.u32 hi, h1, n0, hi;\n\t"
                "mov.b64 %0, {lo, hi};\n\t"
                "}"
                : "=l"(r)
                : "l"(x), "l"(y), "l"(ysh), "l"(neg_two_q), "l"(add), "l"(zero));
        }
        else
        {
            asm("{\n\t"
                ".reg .u32 x0, x1, y0, y1, s0, s1, h0, h1, n0, n1, lo, hi, z, z1, a0, a1;\n\t"
                ".reg .u64 t, c;\n\t"
                "mov.b64 {x0, x1}, %1;\n\t"
                "mov.b64 {y0, y1}, %2;\n\t"
                "mov.b64 {s0, s1}, %3;\n\t"
                "mov.b64 {n0, n1}, %4;\n\t"
                "mul.wide.u32 t, x0, s1;\n\t"
                "mad.wide.u32 t, x1, s0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mov.b64 {z, z1}, %6;\n\t"
                "mov.b64 c, {hi, z};\n\t"
                "mad.wide.u32 t, x1, s1, c;\n\t"
                "mov.b64 {h0, h1}, t;\n\t"
                "mad.wide.u32 t, x0, y0, %5;\n\t"
                "mad.wide.u32 t, h0, n0, t;\n\t"
                "mov.b64 {lo, hi}, t;\n\t"
                "mad.lo.u32 hi, x0, y1, hi;\n\t"
                "mad.lo.u32 hi, x1, y0, hi;\n\t"
                "mad.lo.u32 hi, h0, n1, hi;\n\t"
                "mad.lo.u32 hi, h1, n0, hi;\n\t"
                "mov.b64 %0, {lo, hi};\n\t"
                "}"
                : "=l"(r)
                : "l"(x), "l"(y), "l"(ysh), "l"(neg_two_q), "l"(add), "l"(zero));
        }
        return r;
#endif
    }

    // (u + u + k) - s (mod 2^64) as two three-input 64-bit sums (u, u, k) and (., -s, zero): IADD3 / IADD3.X pairs
    // on the ALU pipe, in this association (NVVM otherwise regroups the C++ expression into two-input sums whose
    // carry words become IMAD.X).  `zero` is a run-time zero.
    GSB_DEV uint64_t twice_plus_minus(uint64_t u, uint64_t k, uint64_t s, uint64_t zero)
    {
#ifdef GSB_CUSIM
        GSB_SIM_ASSERT(zero == 0);
        return u + u + k - s;
#else
        uint64_t r;
        asm("{\n\t"
            ".reg .u64 t;\n\t"
            "add.u64 t, %1, %1;\n\t"
            "add.u64 t, t, %2;\n\t"
            "sub.u64 t, t, %3;\n\t"
            "add.u64 %0, t, %4;\n\t"
            "}"
            : "=l"(r)
            : "l"(u), "l"(k), "l"(s), "l"(zero));
        return r;
#endif
    }

    // SlothfulNTT::reduceBarrettLazy (util/ntt.cpp:237-241)
    GSB_DEV uint64_t reduce_barrett_lazy(uint64_t x, uint64_t rdp, uint64_t q)
    {
        return mad_lo64_1(mulhi(x, rdp), 0 - q, x);
    }

    // MulModShoup::operator() (multi_special_primes.cpp:13-19) / FastMulMod (uintarithsmallmod.h:312-322)
    GSB_DEV uint64_t mulmod_shoup(uint64_t x, uint64_t c, uint64_t cs, uint64_t p)
    {
        uint64_t t = mad_lo64_2(x, c, mulhi(x, cs), 0 - p, 0);
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
        uint64_t r = mad_lo64_1(quot, 0 - q, z0);
        return r - (r >= q ? q : 0);
    }

    // barrett_reduce_63 (util/uintarithsmallmod.h:181-207)
    GSB_DEV uint64_t barrett63(uint64_t x, uint64_t q, uint64_t cr1)
    {
        uint64_t r = mad_lo64_1(mulhi(x, cr1), 0 - q, x);
        return r - (r >= q ? q : 0);
    }

    // Canonical residue of z = (z1:z0) < 2^(2k+1), k = bit length of q <= 60 (c.mu != 0): generalized
    // Barrett with one 64x64 high product.  zt = floor(z / 2^(k-1)) < 2^(k+2); h = floor(zt mu / 2^64)
    // satisfies floor(z/q) - 2 <= h <= floor(z/q)  (zt > z/2^(k-1) - 1, mu > 2^(k+63)/q - 1 and
    // z / 2^(k+63) <= 1/4), so z - h q is in [0, 3q): two conditional subtractions.  The result is the
    // same word as barrett_reduce_128's (both are THE residue in [0, q)), with 5 instead of 14 wide
    // multiplies.  Only for products of canonical operands (a b < q^2, a b + c d < 2 q^2).
    GSB_DEV uint64_t reduce_2k(uint64_t z0, uint64_t z1, const LimbC &c)
    {
        const uint32_t sh = static_cast<uint32_t>(c.sh);
        const uint64_t zt = (z0 >> sh) | (z1 << (64u - sh));
        uint64_t r = mad_lo64_1(mulhi(zt, c.mu), 0 - c.q, z0);
        GSB_SIM_ASSERT(sh >= 1 && (z1 >> sh) == 0 && r < c.q + c.two_q);
        r -= (r >= c.two_q) ? c.two_q : 0;
        return r - (r >= c.q ? c.q : 0);
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
#ifdef GSB_CUSIM
        uint64_t pl = a * b, ph = mulhi(a, b);
        lo += pl;
        hi += ph + (lo < pl ? 1 : 0);
#else
        // four 32x32 products; the two aligned ones ride the 64-bit carry chain, the two middle ones
        // are added word-wise (adds on the ALU pipe; the multiplier pipe bounds this kernel)
        asm("{\n"
            ".reg .u32 a0, a1, b0, b1, l0, l1, h0, h1, m0, m1;\n"
            ".reg .u64 p;\n"
            "mov.b64 {a0, a1}, %2;\n"
            "mov.b64 {b0, b1}, %3;\n"
            "mul.wide.u32 p, a0, b0;\n"
            "add.cc.u64 %0, %0, p;\n"
            "mul.wide.u32 p, a1, b1;\n"
            "addc.u64 %1, %1, p;\n"
            "mov.b64 {l0, l1}, %0;\n"
            "mov.b64 {h0, h1}, %1;\n"
            "mul.wide.u32 p, a0, b1;\n"
            "mov.b64 {m0, m1}, p;\n"
            "add.cc.u32 l1, l1, m0;\n"
            "addc.cc.u32 h0, h0, m1;\n"
            "addc.u32 h1, h1, 0;\n"
            "mul.wide.u32 p, a1, b0;\n"
            "mov.b64 {m0, m1}, p;\n"
            "add.cc.u32 l1, l1, m0;\n"
            "addc.cc.u32 h0, h0, m1;\n"
            "addc.u32 h1, h1, 0;\n"
            "mov.b64 %0, {l0, l1};\n"
            "mov.b64 %1, {h0, h1};\n"
            "}"
            : "+l"(lo), "+l"(hi)
            : "l"(a), "l"(b));
#endif
    }

    // mac128 for a < 2^63 and b < 2^63 (operand of a wide-range transform times a key word, both
    // bounded by gsb_ctx::fast_ntt): the two middle products cannot wrap 64 bits when summed
    // (a0*b1 < 2^63 and a1*b0 < 2^63), so they are added to each other inside the multiplier and
    // enter the accumulator with one carry chain instead of two.  Same 128-bit result as mac128.
    GSB_DEV void mac128_mid(uint64_t &lo, uint64_t &hi, uint64_t a, uint64_t b)
    {
#ifdef GSB_CUSIM
        GSB_SIM_ASSERT((a >> 63) == 0 && (b >> 63) == 0);
        mac128(lo, hi, a, b);
#else
        asm("{\n"
            ".reg .u32 a0, a1, b0, b1, l0, l1, h0, h1, m0, m1;\n"
            ".reg .u64 p;\n"
            "mov.b64 {a0, a1}, %2;\n"
            "mov.b64 {b0, b1}, %3;\n"
            "mul.wide.u32 p, a0, b0;\n"
            "add.cc.u64 %0, %0, p;\n"
            "mul.wide.u32 p, a1, b1;\n"
            "addc.u64 %1, %1, p;\n"
            "mov.b64 {l0, l1}, %0;\n"
            "mov.b64 {h0, h1}, %1;\n"
            "mul.wide.u32 p, a0, b1;\n"
            "mad.wide.u32 p, a1, b0, p;\n"
            "mov.b64 {m0, m1}, p;\n"
            "add.cc.u32 l1, l1, m0;\n"
            "addc.cc.u32 h0, h0, m1;\n"
            "addc.u32 h1, h1, 0;\n"
            "mov.b64 %0, {l0, l1};\n"
            "mov.b64 %1, {h0, h1};\n"
            "}"
            : "+l"(lo), "+l"(hi)
            : "l"(a), "l"(b));
#endif
    }
} // namespace gsb
