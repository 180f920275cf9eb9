// See host_tables.h.  Setup-time only; nothing here is on the timed path.
#include "host_tables.h"

namespace gsb
{
    namespace host
    {
        static inline uint64_t hi64(uint64_t a, uint64_t b)
        {
            return static_cast<uint64_t>((static_cast<u128>(a) * b) >> 64);
        }

        // Modulus::set_value (modulus.cpp:87-98): const_ratio = floor(2^128 / q), remainder.
        ModConst make_mod(uint64_t q)
        {
            ModConst m;
            m.q = q;
            const u128 top = ~static_cast<u128>(0); // 2^128 - 1
            u128 quo = top / q;
            uint64_t rem = static_cast<uint64_t>(top % q) + 1;
            if (rem == q)
            {
                quo += 1;
                rem = 0;
            }
            m.cr0 = static_cast<uint64_t>(quo);
            m.cr1 = static_cast<uint64_t>(quo >> 64);
            m.cr2 = rem;
            return m;
        }

        // shoupify (util/ntt.cpp:19-24)
        uint64_t shoup(uint64_t x, uint64_t q)
        {
            return static_cast<uint64_t>((static_cast<u128>(x) << 64) / q);
        }

        // multiply_uint_mod via barrett_reduce_128 (util/uintarithsmallmod.h:140-178,209-221)
        uint64_t mul_mod(uint64_t a, uint64_t b, const ModConst &m)
        {
            const u128 z = static_cast<u128>(a) * b;
            const uint64_t z0 = static_cast<uint64_t>(z), z1 = static_cast<uint64_t>(z >> 64);
            uint64_t carry = hi64(z0, m.cr0);
            u128 t = static_cast<u128>(z0) * m.cr1;
            uint64_t tmp1 = static_cast<uint64_t>(t) + carry;
            uint64_t tmp3 = static_cast<uint64_t>(t >> 64) + (tmp1 < carry ? 1 : 0);
            t = static_cast<u128>(z1) * m.cr0;
            uint64_t s = tmp1 + static_cast<uint64_t>(t);
            carry = static_cast<uint64_t>(t >> 64) + (s < tmp1 ? 1 : 0);
            uint64_t quot = z1 * m.cr1 + tmp3 + carry;
            uint64_t r = z0 - quot * m.q;
            return r >= m.q ? r - m.q : r;
        }

        uint64_t barrett63(uint64_t x, const ModConst &m)
        {
            uint64_t r = x - hi64(x, m.cr1) * m.q;
            return r >= m.q ? r - m.q : r;
        }

        uint64_t pow_mod(uint64_t a, uint64_t e, const ModConst &m)
        {
            uint64_t r = 1 % m.q, b = a % m.q;
            for (; e; e >>= 1)
            {
                if (e & 1)
                    r = mul_mod(r, b, m);
                b = mul_mod(b, b, m);
            }
            return r;
        }

        bool inv_mod(uint64_t a, uint64_t q, uint64_t &out)
        {
            a %= q;
            if (!a)
                return false;
            __int128 t0 = 0, t1 = 1;
            uint64_t r0 = q, r1 = a;
            while (r1)
            {
                uint64_t d = r0 / r1, r2 = r0 - d * r1;
                __int128 t2 = t0 - static_cast<__int128>(d) * t1;
                r0 = r1, r1 = r2, t0 = t1, t1 = t2;
            }
            if (r0 != 1)
                return false;
            out = static_cast<uint64_t>(t0 < 0 ? t0 + q : t0);
            return true;
        }

        uint32_t reverse_bits(uint32_t x, int bits)
        {
            uint32_t r = 0;
            for (int i = 0; i < bits; i++)
                r |= ((x >> i) & 1u) << (bits - 1 - i);
            return r;
        }

        // The reference takes the numerically smallest primitive 2N-th root (numth.cpp:398-425).
        // All primitive 2N-th roots are the odd powers of any one of them, so scan those.
        static bool minimal_root(uint64_t two_n, const ModConst &m, uint64_t &root)
        {
            if ((m.q - 1) % two_n)
                return false;
            const uint64_t cofactor = (m.q - 1) / two_n;
            uint64_t g = 0;
            for (uint64_t cand = 2; cand < 4096 && !g; cand++)
            {
                uint64_t r = pow_mod(cand, cofactor, m);
                if (pow_mod(r, two_n >> 1, m) == m.q - 1)
                    g = r;
            }
            if (!g)
                return false;
            const uint64_t g2 = mul_mod(g, g, m);
            uint64_t best = g, cur = g;
            for (uint64_t i = 1; i < (two_n >> 1); i++)
            {
                cur = mul_mod(cur, g2, m);
                if (cur < best)
                    best = cur;
            }
            root = best;
            return true;
        }

        bool build_limb_tables(int logn, uint64_t q, LimbTables &t)
        {
            const size_t n = size_t(1) << logn;
            t.mod = make_mod(q);
            if (!minimal_root(2 * static_cast<uint64_t>(n), t.mod, t.root))
                return false;
            uint64_t inv_root = 0;
            if (!inv_mod(t.root, q, inv_root) || !inv_mod(static_cast<uint64_t>(n), q, t.inv_n))
                return false;
            t.inv_n_s = shoup(t.inv_n, q);
            t.rdp = shoup(1, q);
            t.w.assign(n, 0);
            t.ws.assign(n, 0);
            t.v.assign(n, 0);
            t.vs.assign(n, 0);
            // forward: w[bitrev(i)] = root^i (util/ntt.cpp:101-111)
            std::vector<uint64_t> inv_br(n);
            uint64_t pf = 1, pi = 1;
            t.w[0] = 1;
            inv_br[0] = 1;
            for (size_t i = 1; i < n; i++)
            {
                pf = mul_mod(pf, t.root, t.mod);
                pi = mul_mod(pi, inv_root, t.mod);
                const uint32_t r = reverse_bits(static_cast<uint32_t>(i), logn);
                t.w[r] = pf;
                inv_br[r] = pi;
            }
            // inverse: stage-major re-ordering (util/ntt.cpp:85-96) and n^-1 merged into the last
            // entry (util/ntt.cpp:98)
            t.v[0] = inv_br[0];
            size_t pos = 1;
            for (size_t m = n >> 1; m >= 1; m >>= 1)
                for (size_t i = 0; i < m; i++)
                    t.v[pos++] = inv_br[m + i];
            t.v[n - 1] = mul_mod(t.v[n - 1], t.inv_n, t.mod);
            for (size_t i = 0; i < n; i++)
            {
                t.ws[i] = shoup(t.w[i], q);
                t.vs[i] = shoup(t.v[i], q);
            }
            return true;
        }

        void galois_table_ntt(int logn, uint32_t elt, std::vector<uint32_t> &table)
        {
            const uint32_t n = 1u << logn;
            table.resize(n);
            for (uint32_t i = 0; i < n; i++)
            {
                const uint32_t rev = reverse_bits(n + i, logn + 1);
                const uint64_t raw = (static_cast<uint64_t>(elt) * rev) >> 1;
                table[i] = reverse_bits(static_cast<uint32_t>(raw & (n - 1)), logn);
            }
        }

        uint32_t galois_elt_from_step(int logn, int step)
        {
            const uint32_t n = 1u << logn;
            const uint64_t m = 2ull * n;
            if (step == 0)
                return static_cast<uint32_t>(m - 1);
            const bool neg = step < 0;
            const uint32_t pos = static_cast<uint32_t>(neg ? -static_cast<int64_t>(step) : step);
            if (pos >= (n >> 1))
                return 0;
            uint32_t count = neg ? (n >> 1) - pos : pos;
            uint64_t e = 1;
            while (count--)
                e = (e * 5u) & (m - 1); // GaloisTool::generator_ = 5 (util/galois.h:169)
            return static_cast<uint32_t>(e);
        }
        // ---- BEHZ auxiliary base and constants (util/rns.cpp:540-728, util/numth.cpp:179-322) ---------
        static uint64_t mulmod128(uint64_t a, uint64_t b, uint64_t m)
        {
            return static_cast<uint64_t>((static_cast<u128>(a) * b) % m);
        }

        // Primality of a 64-bit integer; the reference's Miller-Rabin draws random bases
        // (numth.cpp:253-274), the answer does not depend on them.
        static bool is_prime64(uint64_t n)
        {
            static const uint64_t bases[12] = { 2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37 };
            if (n < 2)
                return false;
            for (uint64_t b : bases)
            {
                if (n == b)
                    return true;
                if (n % b == 0)
                    return false;
            }
            uint64_t d = n - 1;
            int r = 0;
            for (; !(d & 1); d >>= 1)
                r++;
            for (uint64_t b : bases)
            {
                uint64_t x = 1, base = b % n, e = d;
                for (; e; e >>= 1, base = mulmod128(base, base, n))
                    if (e & 1)
                        x = mulmod128(x, base, n);
                if (x == 1 || x == n - 1)
                    continue;
                bool witness = true;
                for (int j = 1; j < r && witness; j++)
                {
                    x = mulmod128(x, x, n);
                    witness = x != n - 1;
                }
                if (witness)
                    return false;
            }
            return true;
        }

        // prod_{l in [0, n), l != skip} v[l] mod m  (skip = n: the whole product)
        static uint64_t punctured(const std::vector<uint64_t> &v, size_t skip, uint64_t m)
        {
            uint64_t p = 1 % m;
            for (size_t l = 0; l < v.size(); l++)
                if (l != skip)
                    p = mulmod128(p, v[l] % m, m);
            return p;
        }

        bool build_behz(int logn, const uint64_t *q, size_t L, uint64_t t, BehzConsts &o)
        {
            const uint64_t n = uint64_t(1) << logn, m_tilde = uint64_t(1) << 32;
            o = BehzConsts();
            o.L = L;
            o.t = t;
            // |B| = |q|, one more when K*n*t*q^2 would not fit (rns.cpp:566-573); needs the bit length
            // of prod(q): multiply it out word by word
            std::vector<uint64_t> w(1, 1);
            for (size_t i = 0; i < L; i++)
            {
                uint64_t carry = 0;
                for (auto &x : w)
                {
                    const u128 z = static_cast<u128>(x) * q[i] + carry;
                    x = static_cast<uint64_t>(z);
                    carry = static_cast<uint64_t>(z >> 64);
                }
                if (carry)
                    w.push_back(carry);
            }
            int q_bits = static_cast<int>(w.size() - 1) * 64, t_bits = 0;
            for (uint64_t x = w.back(); x; x >>= 1)
                q_bits++;
            for (uint64_t x = t; x; x >>= 1)
                t_bits++;
            o.nB = L + ((32 + t_bits + q_bits >= 61 * static_cast<int>(L) + 61) ? 1 : 0);
            o.nBsk = o.nB + 1;
            // get_primes(n, SEAL_USER_MOD_BIT_COUNT_MAX + 1 = 60, |B| + 2): m_sk, gamma, B... (rns.cpp:585-590)
            std::vector<uint64_t> primes;
            const uint64_t factor = 2 * n;
            for (uint64_t v = (uint64_t(1) << 60) - factor + 1; primes.size() < o.nB + 2 && v > (uint64_t(1) << 59); v -= factor)
                if (is_prime64(v))
                    primes.push_back(v);
            if (primes.size() < o.nB + 2)
                return false;
            o.m_sk = primes[0];
            const std::vector<uint64_t> B(primes.begin() + 2, primes.end()), Q(q, q + L);
            o.bsk = B;
            o.bsk.push_back(o.m_sk);
            o.inv_punct_q.resize(L);
            o.inv_punct_b.resize(o.nB);
            for (size_t i = 0; i < L; i++)
                if (!inv_mod(punctured(Q, i, q[i]), q[i], o.inv_punct_q[i]))
                    return false;
            for (size_t i = 0; i < o.nB; i++)
                if (!inv_mod(punctured(B, i, B[i]), B[i], o.inv_punct_b[i]))
                    return false;
            o.mat_q.resize((o.nBsk + 1) * L);
            for (size_t j = 0; j <= o.nBsk; j++)
                for (size_t i = 0; i < L; i++)
                    o.mat_q[j * L + i] = punctured(Q, i, j < o.nBsk ? o.bsk[j] : m_tilde);
            o.mat_b.resize((L + 1) * o.nB);
            for (size_t j = 0; j <= L; j++)
                for (size_t i = 0; i < o.nB; i++)
                    o.mat_b[j * o.nB + i] = punctured(B, i, j < L ? q[j] : o.m_sk);
            o.prod_q_mod_bsk.resize(o.nBsk);
            o.inv_prod_q_mod_bsk.resize(o.nBsk);
            o.inv_mtilde_mod_bsk.resize(o.nBsk);
            for (size_t j = 0; j < o.nBsk; j++)
            {
                o.prod_q_mod_bsk[j] = punctured(Q, L, o.bsk[j]);
                if (!inv_mod(o.prod_q_mod_bsk[j], o.bsk[j], o.inv_prod_q_mod_bsk[j]) ||
                    !inv_mod(m_tilde % o.bsk[j], o.bsk[j], o.inv_mtilde_mod_bsk[j]))
                    return false;
            }
            o.prod_b_mod_q.resize(L);
            for (size_t i = 0; i < L; i++)
                o.prod_b_mod_q[i] = punctured(B, o.nB, q[i]);
            if (!inv_mod(punctured(B, o.nB, o.m_sk), o.m_sk, o.inv_prod_b_mod_msk) ||
                !inv_mod(punctured(Q, L, m_tilde), m_tilde, o.inv_prod_q_mod_mtilde))
                return false;
            // {t, gamma} side (rns.cpp:603-608, 690-717); gamma is the second sampled prime (:590)
            o.gamma = primes[1];
            const uint64_t tg[2] = { t, o.gamma };
            o.prod_tg_mod_q.resize(L);
            for (size_t i = 0; i < L; i++)
                o.prod_tg_mod_q[i] = mul_mod(t % q[i], o.gamma % q[i], make_mod(q[i]));
            o.mat_tg.resize(2 * L);
            for (size_t j = 0; j < 2; j++)
            {
                for (size_t i = 0; i < L; i++)
                    o.mat_tg[j * L + i] = punctured(Q, i, tg[j]);
                uint64_t inv = 0;
                if (!inv_mod(punctured(Q, L, tg[j]), tg[j], inv))
                    return false;
                o.neg_inv_q_mod_tg[j] = inv ? tg[j] - inv : 0;
            }
            return inv_mod(o.gamma % t, t, o.inv_gamma_mod_t);
        }
    } // namespace host
} // namespace gsb
