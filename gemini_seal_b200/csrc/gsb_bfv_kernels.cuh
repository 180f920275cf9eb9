// BEHZ base conversions of Evaluator::bfv_multiply (evaluator.cpp:274-445; util/rns.cpp:441-496,
// 853-1066): the two elementwise stages either side of the NTT-domain tensor product.  An "extended"
// polynomial is [M][N] limbs, M = L + nBsk: the L limbs of base q followed by the nBsk limbs of
// Bsk = B u {m_sk}; the NTTs and the dyadic products over it reuse the CKKS kernels with the
// extended table.  One thread owns one coefficient across all limbs (the conversions mix limbs);
// these stages are HBM-bound reads of L (or M) words and writes of M (or L) words per coefficient.
#pragma once
#include "modarith.cuh"

namespace gsb
{
    constexpr int BEHZ_MAX = 64; // coeff_modulus size limit of the context (gsb_ctx_create)

    struct BehzDev
    {
        const LimbC *limb;               // [M]: q_0..q_{L-1}, Bsk_0..Bsk_{nBsk-1}
        const uint64_t *inv_punct_q;     // [L]
        const uint64_t *mat_q;           // [nBsk + 1][L] (last row: mod m_tilde = 2^32)
        const uint64_t *inv_punct_b;     // [nB]
        const uint64_t *mat_b;           // [L + 1][nB]  (last row: mod m_sk)
        const uint64_t *prod_q_mod_bsk;  // [nBsk]
        const uint64_t *inv_prod_q_mod_bsk;
        const uint64_t *inv_mtilde_mod_bsk;
        const uint64_t *prod_b_mod_q;    // [L]
        uint64_t inv_prod_q_mod_mtilde, inv_prod_b_mod_msk, t;
        uint32_t L, nB, nBsk, N;
    };

    // multiply_add_uint_mod (util/uintarithsmallmod.h:282-290): barrett_reduce_128(a*b + c)
    GSB_DEV uint64_t mul_add_mod(uint64_t a, uint64_t b, uint64_t c, const LimbC &m)
    {
        uint64_t lo = c, hi = 0;
        mac128(lo, hi, a, b);
        return barrett128(lo, hi, m.q, m.cr0, m.cr1);
    }
    GSB_DEV uint64_t mul_mod(uint64_t a, uint64_t b, const LimbC &m)
    {
        return mul_mod(a, b, m.q, m.cr0, m.cr1);
    }

    // Steps (1)-(2) (evaluator.cpp:340-346): in = n_polys x [L][N] coefficient form over q ->
    // ext = n_polys x [M][N]: limbs [0, L) a copy of the input (its lazy NTT follows), limbs [L, M) =
    // sm_mrq(fastbconv_m_tilde(in)) in base Bsk (rns.cpp:1019-1066, 926-977).
    __global__ void __launch_bounds__(256) behz_lift_kernel(BehzDev z, const uint64_t *in, uint64_t *ext, size_t total)
    {
        const uint32_t L = z.L, nBsk = z.nBsk, N = z.N, M = L + nBsk;
        const uint64_t m_tilde = uint64_t(1) << 32;
        for (size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
             idx += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t p = idx / N, d = idx - p * N;
            const uint64_t *src = in + p * L * N + d;
            uint64_t *dst = ext + p * M * N + d;
            uint64_t tl[BEHZ_MAX];
            for (uint32_t i = 0; i < L; i++)
            {
                const uint64_t x = src[static_cast<size_t>(i) * N];
                dst[static_cast<size_t>(i) * N] = x;
                const LimbC m = z.limb[i];
                // multiply_poly_scalar_coeffmod by m_tilde (:1052-1057), then the conversion's
                // multiply by the inverse punctured product (:478-488)
                tl[i] = mul_mod(mul_mod(x, m_tilde, m), __ldg(z.inv_punct_q + i), m);
            }
            // {m_tilde} component, r_m_tilde = -(c * q^-1) mod m_tilde (:949-955); arithmetic mod 2^32
            uint64_t acc = 0;
            for (uint32_t i = 0; i < L; i++)
                acc += tl[i] * __ldg(z.mat_q + static_cast<size_t>(nBsk) * L + i);
            uint64_t r = ((acc & (m_tilde - 1)) * z.inv_prod_q_mod_mtilde) & (m_tilde - 1);
            r = r ? m_tilde - r : 0;
            for (uint32_t k = 0; k < nBsk; k++)
            {
                const LimbC m = z.limb[L + k];
                uint64_t lo = 0, hi = 0;
                for (uint32_t i = 0; i < L; i++)
                    mac128(lo, hi, tl[i], __ldg(z.mat_q + static_cast<size_t>(k) * L + i));
                const uint64_t v = barrett128(lo, hi, m.q, m.cr0, m.cr1);
                uint64_t temp = r; // centered reduction of r_m_tilde modulo Bsk_k (:964-969)
                if (temp >= (m_tilde >> 1))
                    temp += m.q - m_tilde;
                dst[static_cast<size_t>(L + k) * N] =
                    mul_mod(mul_add_mod(__ldg(z.prod_q_mod_bsk + k), temp, v, m), __ldg(z.inv_mtilde_mod_bsk + k), m);
            }
        }
    }

    // Steps (6)-(8) (evaluator.cpp:428-444): dest = n_polys x [M][N] coefficient form over q u Bsk ->
    // out = n_polys x [L][N]: multiply by t, fast_floor (rns.cpp:979-1017), fastbconv_sk (:853-924).
    __global__ void __launch_bounds__(256) behz_floor_kernel(BehzDev z, const uint64_t *dest, uint64_t *out, size_t total)
    {
        const uint32_t L = z.L, nB = z.nB, nBsk = z.nBsk, N = z.N, M = L + nBsk;
        for (size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
             idx += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t p = idx / N, d = idx - p * N;
            const uint64_t *src = dest + p * M * N + d;
            uint64_t *dst = out + p * L * N + d;
            uint64_t tl[BEHZ_MAX], fb[BEHZ_MAX + 2];
            for (uint32_t i = 0; i < L; i++)
            {
                const LimbC m = z.limb[i];
                tl[i] = mul_mod(mul_mod(src[static_cast<size_t>(i) * N], z.t, m), __ldg(z.inv_punct_q + i), m);
            }
            for (uint32_t k = 0; k < nBsk; k++)
            {
                const LimbC m = z.limb[L + k];
                uint64_t lo = 0, hi = 0;
                for (uint32_t i = 0; i < L; i++)
                    mac128(lo, hi, tl[i], __ldg(z.mat_q + static_cast<size_t>(k) * L + i));
                const uint64_t conv = barrett128(lo, hi, m.q, m.cr0, m.cr1);
                const uint64_t tb = mul_mod(src[static_cast<size_t>(L + k) * N], z.t, m);
                fb[k] = mul_mod(tb + (m.q - conv), __ldg(z.inv_prod_q_mod_bsk + k), m); // :1010-1015
            }
            for (uint32_t i = 0; i < nB; i++)
                tl[i] = mul_mod(fb[i], __ldg(z.inv_punct_b + i), z.limb[L + i]);
            const LimbC msk = z.limb[L + nB];
            uint64_t lo = 0, hi = 0;
            for (uint32_t i = 0; i < nB; i++)
                mac128(lo, hi, tl[i], __ldg(z.mat_b + static_cast<size_t>(L) * nB + i));
            const uint64_t to_msk = barrett128(lo, hi, msk.q, msk.cr0, msk.cr1);
            const uint64_t alpha = mul_mod(to_msk + (msk.q - fb[nB]), z.inv_prod_b_mod_msk, msk); // :889-894
            for (uint32_t l = 0; l < L; l++)
            {
                const LimbC m = z.limb[l];
                lo = hi = 0;
                for (uint32_t i = 0; i < nB; i++)
                    mac128(lo, hi, tl[i], __ldg(z.mat_b + static_cast<size_t>(l) * nB + i));
                const uint64_t conv = barrett128(lo, hi, m.q, m.cr0, m.cr1);
                const uint64_t pb = __ldg(z.prod_b_mod_q + l);
                dst[static_cast<size_t>(l) * N] =
                    alpha > (msk.q >> 1) ? mul_add_mod(pb, msk.q - alpha, conv, m) : mul_add_mod(m.q - pb, alpha, conv, m); // :899-922
            }
        }
    }
    // BFV mod_switch_to_next: RNSTool::divide_and_round_q_last_inplace (util/rns.cpp:731-775), coefficient
    // form.  in = n_polys x [L][N] -> out = n_polys x [L-1][N]; one thread per output word.
    struct BfvModSwitchArgs
    {
        const uint64_t *in;
        uint64_t *out;
        const LimbC *limb;          // [K]
        const uint64_t *inv_qlast;  // [L-1][2] q_last^-1 mod q_i (rns.cpp:719-728) and its Shoup scaling
        uint32_t N, L;
        size_t total; // n_polys * (L-1) * N
    };

    __global__ void __launch_bounds__(256) bfv_mod_switch_kernel(BfvModSwitchArgs p)
    {
        const size_t per = static_cast<size_t>(p.L - 1) * p.N;
        const LimbC last = p.limb[p.L - 1];
        const uint64_t half = last.q >> 1;
        for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < p.total;
             t += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t poly = t / per, rem = t - poly * per;
            const uint32_t i = static_cast<uint32_t>(rem / p.N), d = static_cast<uint32_t>(rem - static_cast<size_t>(i) * p.N);
            const uint64_t *src = p.in + poly * static_cast<size_t>(p.L) * p.N;
            const LimbC m = p.limb[i];
            const uint64_t l = barrett63(src[static_cast<size_t>(p.L - 1) * p.N + d] + half, last.q, last.cr1); // :747-751
            uint64_t tmp = barrett63(l, m.q, m.cr1);                                                              // :758
            tmp = sub_mod(tmp, barrett63(half, m.q, m.cr1), m.q);                                               // :760-764
            const uint64_t v = sub_mod(src[static_cast<size_t>(i) * p.N + d], tmp, m.q);                        // :766-767
            p.out[t] = mul_mod(v, __ldg(p.inv_qlast + 2 * i), m);                                                   // :770-772
        }
    }

    // ---- Decryptor --------------------------------------------------------------
    // dot_product_ct_sk_array (decryptor.cpp:218-265): out[b][i][d] = c0 + sum_{p >= 1} c_p (.) s^p mod q_i.
    // polys: polynomial p >= 1 of ciphertext b at polys + b*ct_stride + (p-1)*L*N (NTT form: the
    // ciphertext's own c_1.. ; coefficient form: their lazy forward NTTs in scratch, any word < 2^64 --
    // dyadic_product_coeffmod reduces the full 128-bit product).  c0 == nullptr: no final add (the BFV
    // branch adds c0 after the inverse NTT).  skp = secret-key powers [size-1][K][N] at key level.
    struct DotSkArgs
    {
        const uint64_t *polys;
        size_t ct_stride;
        const uint64_t *c0;
        size_t c0_stride;
        const uint64_t *skp;
        uint64_t *out; // [B][L][N]
        const LimbC *limb;
        uint32_t N, L, K, n_pow;
        size_t pairs; // B*L*N/2
    };

    __global__ void __launch_bounds__(256) dot_sk_kernel(DotSkArgs p)
    {
        const size_t LN = static_cast<size_t>(p.L) * p.N, KN = static_cast<size_t>(p.K) * p.N;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < p.pairs;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t w = i << 1;
            const size_t b = w / LN, rem = w - b * LN;
            const LimbC c = p.limb[rem / p.N];
            ulonglong2 acc = make_ulonglong2(0, 0); // set_zero_poly (:253), then add_poly_coeffmod per power
            for (uint32_t pw = 0; pw < p.n_pow; pw++)
            {
                const ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(p.polys + b * p.ct_stride + pw * LN + rem);
                const ulonglong2 s = *reinterpret_cast<const ulonglong2 *>(p.skp + pw * KN + rem);
                acc.x = add_mod(acc.x, mul_mod(x.x, s.x, c.q, c.cr0, c.cr1), c.q);
                acc.y = add_mod(acc.y, mul_mod(x.y, s.y, c.q, c.cr0, c.cr1), c.q);
            }
            if (p.c0)
            {
                const ulonglong2 z = *reinterpret_cast<const ulonglong2 *>(p.c0 + b * p.c0_stride + rem);
                acc.x = add_mod(acc.x, z.x, c.q);
                acc.y = add_mod(acc.y, z.y, c.q);
            }
            *reinterpret_cast<ulonglong2 *>(p.out + w) = acc;
        }
    }

    // RNSTool::decrypt_scale_and_round (util/rns.cpp:1070-1126): dot = [B][L][N] coefficient form ->
    // out = [B][N] plaintext coefficients mod t.  One thread per coefficient across the L limbs.
    struct DecryptRoundArgs
    {
        const uint64_t *dot;
        uint64_t *out;
        const LimbC *limb;             // [L] base q
        const uint64_t *prod_tg_mod_q; // [L]
        const uint64_t *inv_punct_q;   // [L]
        const uint64_t *mat_tg;        // [2][L]
        LimbC t, gamma;                // q / cr0 / cr1 of the two output moduli
        uint64_t neg_inv_q_t, neg_inv_q_gamma, inv_gamma_mod_t;
        uint32_t N, L;
        size_t total; // B*N
    };

    __global__ void __launch_bounds__(256) decrypt_round_kernel(DecryptRoundArgs p)
    {
        for (size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; idx < p.total;
             idx += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t b = idx / p.N, d = idx - b * p.N;
            const uint64_t *src = p.dot + b * static_cast<size_t>(p.L) * p.N + d;
            uint64_t lo0 = 0, hi0 = 0, lo1 = 0, hi1 = 0;
            for (uint32_t i = 0; i < p.L; i++)
            {
                const LimbC m = p.limb[i];
                // |gamma t|_qi * ct(s) (:1077-1082), then fast_convert_array's multiply by the inverse
                // punctured product (rns.cpp:478-488) and its dot products with the conversion matrix
                const uint64_t v = mul_mod(mul_mod(src[static_cast<size_t>(i) * p.N], __ldg(p.prod_tg_mod_q + i), m), __ldg(p.inv_punct_q + i), m);
                mac128(lo0, hi0, v, __ldg(p.mat_tg + i));
                mac128(lo1, hi1, v, __ldg(p.mat_tg + p.L + i));
            }
            const uint64_t at = mul_mod(barrett128(lo0, hi0, p.t.q, p.t.cr0, p.t.cr1), p.neg_inv_q_t, p.t);          // :1091-1096
            const uint64_t ag = mul_mod(barrett128(lo1, hi1, p.gamma.q, p.gamma.cr0, p.gamma.cr1), p.neg_inv_q_gamma, p.gamma);
            uint64_t r;
            if (ag > (p.gamma.q >> 1)) // :1107-1117
                r = add_mod(at, (p.gamma.q - ag) % p.t.q, p.t.q);
            else
                r = sub_mod(at, ag % p.t.q, p.t.q);
            if (r != 0)
                r = mul_mod(r, p.inv_gamma_mod_t, p.t); // :1120-1124
            p.out[idx] = r;
        }
    }
} // namespace gsb
