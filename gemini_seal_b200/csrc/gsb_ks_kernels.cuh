// Device kernels and NTT prologue / epilogue functors of the CKKS evaluator path.
// See gsb_evaluator.cuh for the pipeline that launches them.
#pragma once
#include "ntt_kernels.cuh"

namespace gsb
{
    // Geometry + constants of one key switch at level L, passed to kernels by value.
    // ext slot kap in [0, E): kap < L is ciphertext limb kap (modulus index kap), kap >= L is
    // special prime kap-L (modulus index L0 + kap - L)  (evaluator.cpp:2316-2317).
    struct KsDev
    {
        const LimbC *limb;       // [K]
        const uint64_t *modup_a; // [beta][k][2]
        const uint64_t *modup_b; // [beta][k][E]
        const uint64_t *inv_hat; // [k][2]
        const uint64_t *neg_hat; // [L][k]
        const uint64_t *inv_p;   // [L][2]
        uint32_t L, L0, k, E, beta, N;
        uint32_t beta_magic; // ceil(2^32 / beta), 0 for beta = 1: z / beta = hi32(z * beta_magic) while z * beta < 2^32 (make_ksdev)
        // Slack in the lazy ranges, used only where a bound fixed at context creation proves that no
        // 64- or 128-bit intermediate can wrap (gsb_ctx::lazy_modup / lazy_mac): every word then stays
        // congruent to the reference's, and since the inner product ends in a canonical Barrett-128 the
        // outputs are identical.
        //   lazy_modup: a single-limb bundle enters the NTT mod q_kap as it is (< q_max instead of
        //               < q_kap): barrett_reduce_63 of multi_special_primes.cpp:104-107 is skipped
        //               (compile-time LAZY of ModupCoreT, selected by the host from this flag).
        //   lazy_mac:   the operand's last NTT stage skips reduceBarrettLazy (util/ntt.cpp:250-253).
        uint32_t lazy_modup, lazy_mac;
        // z = b * beta + j (one multiply on the uniform path instead of a run-time division per polynomial)
        GSB_DEV void split_z(int z, int &b, int &j) const
        {
            b = z_b(z);
            j = z - b * static_cast<int>(beta);
        }
        GSB_DEV int z_b(int z) const { return beta_magic ? static_cast<int>(__umulhi(static_cast<uint32_t>(z), beta_magic)) : z; }
        GSB_DEV int z_j(int z) const { return z - z_b(z) * static_cast<int>(beta); }
        GSB_DEV int ext_mod(int kap) const { return kap < static_cast<int>(L) ? kap : static_cast<int>(L0) + kap - static_cast<int>(L); }
    };

    // ---- combinator: any prologue with any epilogue -------------------------------------------------
    // A functor type F provides `struct Args` (trivially copyable, with `bool skip(y, z) const` for
    // prologues) and a constructor F(const Args&, y, z, mod, const LimbC&, N).
    template <class ProF, class EpiF>
    struct IoPair
    {
        typename ProF::Args pa;
        typename EpiF::Args ea;
        GSB_DEV bool skip(int y, int z) const { return pa.skip(y, z); }
        struct Pro : ProF
        {
            GSB_DEV Pro(const IoPair &io, int y, int z, int mod, const LimbC &c, uint32_t N) : ProF(io.pa, y, z, mod, c, N) {}
        };
        struct Epi : EpiF
        {
            GSB_DEV Epi(const IoPair &io, int y, int z, int mod, const LimbC &c, uint32_t N) : EpiF(io.ea, y, z, mod, c, N) {}
        };
    };

    struct PlainArgs
    {
        const uint64_t *base;
        size_t zs;
        GSB_DEV bool skip(int, int) const { return false; }
    };
    struct PlainLoad : PlainPro
    {
        typedef PlainArgs Args;
        GSB_DEV PlainLoad(const Args &a, int y, int z, int, const LimbC &, uint32_t N) : PlainPro(a.base, a.zs, y, z, N) {}
    };
    struct PlainStoreArgs
    {
        uint64_t *base;
        size_t zs;
    };
    template <int MODE>
    struct PlainStore : PlainEpi<MODE>
    {
        typedef PlainStoreArgs Args;
        GSB_DEV PlainStore(const Args &a, int y, int z, int, const LimbC &c, uint32_t N) : PlainEpi<MODE>(a.base, a.zs, y, z, c, N) {}
    };

    // ---- mod-up prologue (modup_to_single_rns, multi_special_primes.cpp:80-148) --------------------
    // Source = coefficient-form bundle limbs coef[b][s0 .. s0+cnt); destination modulus = ext slot kap.
    // K1: the context has one special prime, so every bundle is a single limb and only the
    // copy / Barrett-63 path exists (keeps the k > 1 conversion code out of the k = 1 kernels).
    template <bool K1, bool LAZY = false>
    struct ModupCoreT
    {
        const uint64_t *src; // coef + (b*L + s0)*N
        const uint64_t *a;   // modup_a + (j*k)*2
        const uint64_t *bb;  // modup_b + (j*k)*E + kap
        const LimbC *limb;   // limb + s0
        uint64_t q, cr0, cr1;
        uint32_t N, E;
        int cnt;
        bool reduce; // single-limb bundle with q_src > q_dst -> barrett_reduce_63 (:104-107)
        GSB_DEV ModupCoreT(const KsDev &ks, const uint64_t *coef, int b, int j, int kap, const LimbC &c)
        {
            const int s0 = j * static_cast<int>(ks.k);
            const int s1 = s0 + static_cast<int>(ks.k) < static_cast<int>(ks.L) ? s0 + static_cast<int>(ks.k) : static_cast<int>(ks.L);
            N = ks.N;
            E = ks.E;
            cnt = s1 - s0;
            src = coef + (static_cast<size_t>(b) * ks.L + s0) * N;
            GSB_PIN_GLOBAL(src);
            a = ks.modup_a + static_cast<size_t>(j) * ks.k * 2;
            bb = ks.modup_b + static_cast<size_t>(j) * ks.k * E + kap;
            limb = ks.limb + s0;
            q = c.q;
            cr0 = c.cr0;
            cr1 = c.cr1;
            reduce = !LAZY && cnt == 1 && limb[0].q > q;
        }
        GSB_DEV uint64_t convert(uint32_t d) const
        {
            uint64_t lo = 0, hi = 0;
            for (int i = 0; i < cnt; i++)
            {
                const uint64_t t = mulmod_shoup(src[static_cast<size_t>(i) * N + d], __ldg(a + 2 * i), __ldg(a + 2 * i + 1), limb[i].q);
                mac128(lo, hi, t, __ldg(bb + static_cast<size_t>(i) * E));
            }
            return barrett128(lo, hi, q, cr0, cr1);
        }
        // single-limb bundle: raw = the load, xform = the optional Barrett-63; multi-limb bundle: the
        // whole conversion happens in raw (several source limbs per word)
        GSB_DEV uint64_t raw(uint32_t d) const { return (K1 || cnt == 1) ? src[d] : convert(d); }
        GSB_DEV void raw2(uint32_t d, uint64_t &x, uint64_t &y) const
        {
            if (K1 || cnt == 1)
            {
                const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(src + d);
                x = v.x;
                y = v.y;
            }
            else
            {
                x = convert(d);
                y = convert(d + 1);
            }
        }
        GSB_DEV uint64_t xform(uint64_t v) const { return reduce ? barrett63(v, q, cr1) : v; }
    };
    typedef ModupCoreT<false> ModupCore;

    // grid: y = ext slot kap, z = b*beta + j; slots inside the source bundle are skipped.
    struct ModupArgs
    {
        KsDev ks;
        const uint64_t *coef;
        GSB_DEV bool skip(int y, int z) const
        {
            int b, j;
            ks.split_z(z, b, j);
            const int s0 = j * static_cast<int>(ks.k);
            return y >= s0 && y < s0 + static_cast<int>(ks.k) && y < static_cast<int>(ks.L);
        }
    };
    template <bool K1, bool LAZY = false>
    struct ModupLoadT : ModupCoreT<K1, LAZY>
    {
        typedef ModupArgs Args;
        GSB_DEV ModupLoadT(const Args &a, int y, int z, int, const LimbC &c, uint32_t)
            : ModupCoreT<K1, LAZY>(a.ks, a.coef, a.ks.z_b(z), a.ks.z_j(z), y, c)
        {}
    };

    // ---- mod-down prologue (rescale_special_rns_inplace steps 1-2, multi_special_primes.cpp:253-282)
    // grid: y = ciphertext limb i, z = b*2 + l.  red = [B*2][E][N]; its special slots hold the lazy
    // inverse NTT of the reduced accumulators (evaluator.cpp:2353-2357).
    struct MdLoadArgs
    {
        KsDev ks;
        const uint64_t *red;
        GSB_DEV bool skip(int, int) const { return false; }
    };
    // FAST (k = 1 contexts under gsb_ctx::fast_ntt): the wide-range transform that follows accepts any
    // word < 2 q_max, so -[c]_p, an integer in [0, p), enters the NTT mod q_i as it is; and the special
    // slot's lazy inverse NTT word is < 2p, so its barrett_reduce_63 is one conditional subtraction.
    // Both are congruent to the reference's words and cost no multiply.
    template <bool K1, bool FAST = false>
    struct MdLoadT
    {
        typedef MdLoadArgs Args;
        const uint64_t *sp;  // red + (z*E + L)*N
        const uint64_t *ih;  // inv_hat
        const uint64_t *nh;  // neg_hat + i*k
        const LimbC *plimb;  // limb + L0
        uint64_t q, cr0, cr1, p0, p0cr1;
        uint32_t N;
        int k;
        GSB_DEV MdLoadT(const Args &a, int y, int z, int, const LimbC &c, uint32_t N_)
        {
            sp = a.red + (static_cast<size_t>(z) * a.ks.E + a.ks.L) * N_;
            ih = a.ks.inv_hat;
            nh = a.ks.neg_hat + static_cast<size_t>(y) * a.ks.k;
            plimb = a.ks.limb + a.ks.L0;
            q = c.q;
            cr0 = c.cr0;
            cr1 = c.cr1;
            N = N_;
            k = static_cast<int>(a.ks.k);
            p0 = plimb[0].q;
            p0cr1 = plimb[0].cr1;
        }
        GSB_DEV uint64_t convert(uint32_t d) const
        {
            uint64_t lo = 0, hi = 0;
            for (int j = 0; j < k; j++)
            {
                const uint64_t t = mulmod_shoup(sp[static_cast<size_t>(j) * N + d], __ldg(ih + 2 * j), __ldg(ih + 2 * j + 1), plimb[j].q);
                mac128(lo, hi, t, __ldg(nh + j));
            }
            return barrett128(lo, hi, q, cr0, cr1);
        }
        GSB_DEV uint64_t raw(uint32_t d) const { return (K1 || k == 1) ? sp[d] : convert(d); }
        GSB_DEV void raw2(uint32_t d, uint64_t &x, uint64_t &y) const
        {
            x = raw(d);
            y = raw(d + 1);
        }
        GSB_DEV uint64_t xform(uint64_t v) const
        {
            if (!K1 && k != 1)
                return v;
            if (K1 && FAST)
            {
                GSB_SIM_ASSERT(v < 2 * p0);
                return neg_mod(v - (v >= p0 ? p0 : 0), p0);
            }
            // :268-275: negate(barrett_reduce_63(c_p)) then barrett_reduce_128 of (v, 0)
            const uint64_t t = neg_mod(barrett63(v, p0, p0cr1), p0);
            return barrett128(t, 0, q, cr0, cr1);
        }
    };

    // ---- mod-down epilogue (multi_special_primes.cpp:291-302 + evaluator.cpp:2363-2366) -------------
    // val = MulModShoup_{q_i}(red[z][i][d] + corr_ntt[d]; P^-1);  dst = add_mod(val, addend).
    // addend: add_src[b][l][i][d]; with a Galois table, component 0 is gathered through it and
    // component 1 is zero (apply_galois_inplace, evaluator.cpp:1913-1928).
    struct MdStoreArgs
    {
        KsDev ks;
        const uint64_t *red;
        const uint64_t *add_src;
        size_t add_bs, add_ls; // word strides: ciphertext, component
        uint64_t *dst;
        size_t dst_bs, dst_ls;
        const uint32_t *table; // nullptr unless (NTT-form) Galois
        bool zero_l1;          // component 1 has no addend (apply_galois_inplace wipes it, :1928)
    };
    struct MdStore
    {
        typedef MdStoreArgs Args;
        const uint64_t *r;
        const uint64_t *s; // nullptr -> addend 0
        uint64_t *o;
        const uint32_t *table;
        uint64_t q, ip, ips;
        GSB_DEV MdStore(const Args &a, int y, int z, int, const LimbC &c, uint32_t N)
        {
            const int b = z >> 1, l = z & 1;
            r = a.red + (static_cast<size_t>(z) * a.ks.E + y) * N;
            o = a.dst + static_cast<size_t>(b) * a.dst_bs + static_cast<size_t>(l) * a.dst_ls + static_cast<size_t>(y) * N;
            table = a.table;
            s = (a.zero_l1 && l == 1) ? nullptr
                                      : a.add_src + static_cast<size_t>(b) * a.add_bs + static_cast<size_t>(l) * a.add_ls + static_cast<size_t>(y) * N;
            q = c.q;
            ip = a.ks.inv_p[2 * y];
            ips = a.ks.inv_p[2 * y + 1];
        }
        GSB_DEV uint64_t fin(uint32_t d, uint64_t sum) const
        {
            const uint64_t val = mulmod_shoup(sum, ip, ips, q);
            const uint64_t addend = s ? (table ? s[__ldg(table + d)] : s[d]) : 0;
            return add_mod(val, addend, q);
        }
        GSB_DEV void prefetch(uint32_t d) const
        {
            prefetch_l2(r + d);
            if (s && !table)
                prefetch_l2(s + d);
        }
        GSB_DEV uint64_t one(uint32_t d, uint64_t v) const { return fin(d, r[d] + v); }
        GSB_DEV void store(uint32_t d, uint64_t v, int) const { o[d] = one(d, v); }
        GSB_DEV void store2(uint32_t d, uint64_t a, uint64_t b, int) const
        {
            *reinterpret_cast<ulonglong2 *>(o + d) = make_ulonglong2(one(d, a), one(d + 1, b));
        }
    };

    // ---- BFV mod-down epilogue (multi_special_primes.cpp:287-302 with is_ckks == false) -------------
    // The transform is the lazy INVERSE NTT of red[z][i] (:288); the correction term of steps 1-2 stays in
    // coefficient form and is computed here from the special slots: val = MulModShoup(c_i + temp; P^-1).
    struct MdStoreBfvArgs
    {
        MdLoadArgs la;
        MdStoreArgs sa;
    };
    struct MdStoreBfv
    {
        typedef MdStoreBfvArgs Args;
        MdLoadT<false> ml;
        MdStore ms;
        GSB_DEV MdStoreBfv(const Args &a, int y, int z, int mod, const LimbC &c, uint32_t N)
            : ml(a.la, y, z, mod, c, N), ms(a.sa, y, z, mod, c, N)
        {}
        GSB_DEV uint64_t one(uint32_t d, uint64_t v) const { return ms.fin(d, v + ml.xform(ml.raw(d))); }
        GSB_DEV void store(uint32_t d, uint64_t v, int) const { ms.o[d] = one(d, v); }
        GSB_DEV void store2(uint32_t d, uint64_t a, uint64_t b, int) const
        {
            *reinterpret_cast<ulonglong2 *>(ms.o + d) = make_ulonglong2(one(d, a), one(d + 1, b));
        }
    };

    // ---- rescale (RNSTool::divide_and_round_q_last_ntt_inplace, util/rns.cpp:777-851) --------------
    // Stage 1 epilogue: strict inverse NTT of the last limb, then barrett_reduce_63(x + q_last/2)
    // (:798-806).  t = [n_polys][N].
    struct RsLastStore
    {
        typedef PlainStoreArgs Args;
        uint64_t *p;
        uint64_t q, cr1, half;
        GSB_DEV RsLastStore(const Args &a, int y, int z, int, const LimbC &c, uint32_t N)
            : p(a.base + static_cast<size_t>(z) * a.zs + static_cast<size_t>(y) * N), q(c.q), cr1(c.cr1), half(c.q >> 1)
        {}
        GSB_DEV uint64_t one(uint64_t v) const
        {
            v -= (v >= q) ? q : 0; // util/ntt.h:328-333
            return barrett63(v + half, q, cr1);
        }
        GSB_DEV void store(uint32_t d, uint64_t v, int) const { p[d] = one(v); }
        GSB_DEV void store2(uint32_t d, uint64_t a, uint64_t b, int) const
        {
            *reinterpret_cast<ulonglong2 *>(p + d) = make_ulonglong2(one(a), one(b));
        }
    };

    // Stage 2 prologue (:811-827): (t mod q_i if q_i < q_last) + (q_i - half mod q_i).
    // grid: y = limb i in [0, L-1), z = polynomial.
    struct RsLoadArgs
    {
        const uint64_t *t;        // [n_polys][N]
        const uint64_t *neg_half; // [L-1]
        uint64_t q_last;
        GSB_DEV bool skip(int, int) const { return false; }
    };
    // FAST: t < q_last and the wide-range transform accepts any word < 2 q_max, so the reduction
    // modulo q_i is skipped (congruent word).
    template <bool FAST>
    struct RsLoadT
    {
        typedef RsLoadArgs Args;
        const uint64_t *t;
        uint64_t q, cr1, nh;
        bool reduce;
        GSB_DEV RsLoadT(const Args &a, int y, int z, int, const LimbC &c, uint32_t N)
            : t(a.t + static_cast<size_t>(z) * N), q(c.q), cr1(c.cr1), nh(a.neg_half[y]), reduce(!FAST && c.q < a.q_last)
        {}
        GSB_DEV uint64_t xform(uint64_t v) const { return (reduce ? barrett63(v, q, cr1) : v) + nh; }
        GSB_DEV uint64_t raw(uint32_t d) const { return t[d]; }
        GSB_DEV void raw2(uint32_t d, uint64_t &x, uint64_t &y) const
        {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(t + d);
            x = v.x;
            y = v.y;
        }
    };

    typedef RsLoadT<false> RsLoad;

    // Stage 2 epilogue (:836-848): out = multiply_uint_mod(in + 4q - ntt_lazy, q_last^-1 mod q_i).
    // The constant factor comes with its Shoup scaling: x*c - floor(x*c'/2^64)*q lies in [0, 2q) for any
    // 64-bit x, so one conditional subtraction yields the same canonical residue as the reference's
    // Barrett-128 of the full product, with 6 instead of 18 wide multiplies.
    struct RsStoreArgs
    {
        const uint64_t *in; // [n_polys][L][N]
        uint64_t *out;      // [n_polys][L-1][N]
        const uint64_t *inv_qlast; // [L-1][2]
        uint32_t L;
    };
    struct RsStore
    {
        typedef RsStoreArgs Args;
        const uint64_t *in;
        uint64_t *out;
        uint64_t q, four_q, inv, inv_s;
        GSB_DEV RsStore(const Args &a, int y, int z, int, const LimbC &c, uint32_t N)
            : in(a.in + (static_cast<size_t>(z) * a.L + y) * N), out(a.out + (static_cast<size_t>(z) * (a.L - 1) + y) * N),
              q(c.q), four_q(c.q << 2), inv(a.inv_qlast[2 * y]), inv_s(a.inv_qlast[2 * y + 1])
        {}
        GSB_DEV void prefetch(uint32_t d) const { prefetch_l2(in + d); }
        GSB_DEV uint64_t one(uint64_t u, uint64_t v) const { return mulmod_shoup(u + four_q - v, inv, inv_s, q); }
        GSB_DEV void store(uint32_t d, uint64_t v, int) const { out[d] = one(in[d], v); }
        GSB_DEV void store2(uint32_t d, uint64_t a, uint64_t b, int) const
        {
            const ulonglong2 u = *reinterpret_cast<const ulonglong2 *>(in + d);
            *reinterpret_cast<ulonglong2 *>(out + d) = make_ulonglong2(one(u.x, a), one(u.y, b));
        }
    };

    // ---- fused tail of multiply + relinearize + rescale (k = 1 contexts under gsb_ctx::fast_ntt) --------
    // gsb_mul_relin_rescale_* only hands out the rescaled ciphertext, so the mod-down of the key switch
    // (multi_special_primes.cpp:237-304), the final add (evaluator.cpp:2363-2366) and the rescale
    // (util/rns.cpp:777-851) are evaluated as ONE forward transform per remaining limb instead of two.
    // With c_p = [c]_p (special slot, coefficient form), t1 = -c_p in [0,p), P = p mod q_i, u = red_i
    // (reduced inner product, NTT form) and c3 = the tensor product's component (the key switch's addend),
    // the reference computes, all mod q_i,
    //     dst_i = (u_i + NTT_i(t1)) P^-1 + c3_i                                  (mod-down + add)
    //     t     = barrett63(strict_iNTT(dst_{L-1}) + q_{L-1}/2)                  (rescale, last limb)
    //     out_i = (dst_i - NTT_i(t + (q_i - half))) q_{L-1}^-1                   (rescale, limb i < L-1)
    // and canonicalises out_i.  The NTT is linear mod q_i, hence
    //     strict_iNTT(dst_{L-1}) = canon( iNTT(u_{L-1} P^-1 + c3_{L-1}) + P^-1 t1 )
    //     out_i = canon( ((u_i + NTT_i(t1 - P (t + q_i - half))) P^-1 + c3_i) q_{L-1}^-1 ):
    // the same canonical words from 2 inverse + (L-1) forward transforms per polynomial instead of
    // 2 inverse + (2L-1) forward ones.  Every intermediate word is congruent to the reference's and inside
    // the ranges the wide-range transforms accept (inverse input < 2q, forward input < 2 q_max).
    //
    // Inverse transforms, grid y in {0, 1}, z = b*2 + l: y = 0 is limb L-1 with the prologue
    // u P^-1 + c3 (< 2q), y = 1 is the special slot as it is.  ModMap{L, L0, L-1} maps them to their moduli.
    struct TailInvArgs
    {
        KsDev ks;
        const uint64_t *red; // [B*2][E][N]
        const uint64_t *c3;  // tensor product, ciphertext b at c3 + b*c3_bs, component l at + l*c3_ls
        size_t c3_bs, c3_ls;
        GSB_DEV bool skip(int, int) const { return false; }
    };
    struct TailInvLoad
    {
        typedef TailInvArgs Args;
        const uint64_t *r, *c;
        uint64_t q, ip, ips;
        bool last;
        GSB_DEV TailInvLoad(const Args &a, int y, int z, int, const LimbC &lc, uint32_t N)
        {
            const int b = z >> 1, l = z & 1, L = static_cast<int>(a.ks.L);
            last = y == 0;
            r = a.red + (static_cast<size_t>(z) * a.ks.E + (L - 1 + y)) * N;
            c = a.c3 + static_cast<size_t>(b) * a.c3_bs + static_cast<size_t>(l) * a.c3_ls + static_cast<size_t>(L - 1) * N;
            q = lc.q;
            ip = a.ks.inv_p[2 * (L - 1)];
            ips = a.ks.inv_p[2 * (L - 1) + 1];
        }
        GSB_DEV uint64_t raw(uint32_t d) const { return last ? mulmod_shoup(r[d], ip, ips, q) + c[d] : r[d]; }
        GSB_DEV void raw2(uint32_t d, uint64_t &x, uint64_t &y) const
        {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(r + d);
            x = v.x;
            y = v.y;
            if (last)
            {
                const ulonglong2 w = *reinterpret_cast<const ulonglong2 *>(c + d);
                x = mulmod_shoup(x, ip, ips, q) + w.x;
                y = mulmod_shoup(y, ip, ips, q) + w.y;
            }
        }
        GSB_DEV uint64_t xform(uint64_t v) const { return v; }
    };

    // Between the inverse and the forward transforms: buf = [B*2][2][N] holds, per polynomial, the lazy
    // inverse transforms u' = iNTT(u_{L-1} P^-1 + c3_{L-1}) in [0,2q) and s = iNTT(special slot) in [0,2p);
    // they are replaced in place by t (rescale's rounded last limb, canonical mod q_{L-1}) and t1 = -[c]_p.
    struct TailMidArgs
    {
        uint64_t *buf;
        LimbC last, sp;
        uint64_t ip, ips; // P^-1 mod q_{L-1} with its Shoup scaling
        uint32_t N;
        size_t total; // B*2*N
    };
    __global__ void __launch_bounds__(256) tail_mid_kernel(TailMidArgs a)
    {
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < a.total;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t z = i >> (31 - __clz(a.N)), d = i & (a.N - 1); // N is a power of two
            uint64_t *pu = a.buf + (z * 2) * a.N + d, *ps = pu + a.N;
            const uint64_t s = *ps;
            GSB_SIM_ASSERT(s < a.sp.two_q && *pu < a.last.two_q);
            const uint64_t t1 = neg_mod(s - (s >= a.sp.q ? a.sp.q : 0), a.sp.q); // :268-275, -[c]_p in [0,p)
            uint64_t y = *pu + mulmod_shoup(t1, a.ip, a.ips, a.last.q);             // < 3q
            y -= (y >= a.last.two_q) ? a.last.two_q : 0;
            y -= (y >= a.last.q) ? a.last.q : 0;                                    // = strict_iNTT(dst_{L-1})
            *pu = barrett63(y + (a.last.q >> 1), a.last.q, a.last.cr1);             // util/rns.cpp:798-806
            *ps = t1;
        }
    }

    // Forward transforms, grid y = limb i < L-1, z = b*2 + l.  Prologue: t1 + (t + (q_i - half)) (-P) mod q_i
    // (< p + q_i <= 2 q_max, the input range of a wide-range transform).
    struct TailLoadArgs
    {
        const uint64_t *buf;      // [B*2][2][N]: t, t1
        const uint64_t *neg_half; // [L-1]
        const uint64_t *neg_p;    // [L][2]: -p mod q_i with its Shoup scaling
        GSB_DEV bool skip(int, int) const { return false; }
    };
    struct TailLoad
    {
        typedef TailLoadArgs Args;
        const uint64_t *t, *t1;
        uint64_t q, nh, np, nps;
        GSB_DEV TailLoad(const Args &a, int y, int z, int, const LimbC &lc, uint32_t N)
            : t(a.buf + static_cast<size_t>(z) * 2 * N), t1(t + N), q(lc.q), nh(a.neg_half[y]), np(a.neg_p[2 * y]), nps(a.neg_p[2 * y + 1])
        {}
        GSB_DEV uint64_t one(uint64_t tv, uint64_t t1v) const { return t1v + mulmod_shoup(tv + nh, np, nps, q); }
        GSB_DEV uint64_t raw(uint32_t d) const { return one(t[d], t1[d]); }
        GSB_DEV void raw2(uint32_t d, uint64_t &x, uint64_t &y) const
        {
            const ulonglong2 a = *reinterpret_cast<const ulonglong2 *>(t + d), b = *reinterpret_cast<const ulonglong2 *>(t1 + d);
            x = one(a.x, b.x);
            y = one(a.y, b.y);
        }
        GSB_DEV uint64_t xform(uint64_t v) const { return v; }
    };

    // Epilogue: out_i = MulModShoup(MulModShoup(u_i + ntt; P^-1) + c3_i; q_{L-1}^-1), canonical.
    struct TailStoreArgs
    {
        KsDev ks;
        const uint64_t *red;
        const uint64_t *c3;
        size_t c3_bs, c3_ls;
        uint64_t *out;             // [B][2][L-1][N]
        const uint64_t *inv_qlast; // [L-1][2]
    };
    struct TailStore
    {
        typedef TailStoreArgs Args;
        const uint64_t *r, *c;
        uint64_t *o;
        uint64_t q, ip, ips, iq, iqs;
        GSB_DEV TailStore(const Args &a, int y, int z, int, const LimbC &lc, uint32_t N)
        {
            const int b = z >> 1, l = z & 1;
            r = a.red + (static_cast<size_t>(z) * a.ks.E + y) * N;
            c = a.c3 + static_cast<size_t>(b) * a.c3_bs + static_cast<size_t>(l) * a.c3_ls + static_cast<size_t>(y) * N;
            o = a.out + (static_cast<size_t>(z) * (a.ks.L - 1) + y) * N;
            q = lc.q;
            ip = a.ks.inv_p[2 * y];
            ips = a.ks.inv_p[2 * y + 1];
            iq = a.inv_qlast[2 * y];
            iqs = a.inv_qlast[2 * y + 1];
        }
        GSB_DEV void prefetch(uint32_t d) const
        {
            prefetch_l2(r + d);
            prefetch_l2(c + d);
        }
        GSB_DEV uint64_t one(uint64_t u, uint64_t c3v, uint64_t v) const
        {
            return mulmod_shoup(mulmod_shoup(u + v, ip, ips, q) + c3v, iq, iqs, q);
        }
        GSB_DEV void store(uint32_t d, uint64_t v, int) const { o[d] = one(r[d], c[d], v); }
        GSB_DEV void store2(uint32_t d, uint64_t a, uint64_t b, int) const
        {
            const ulonglong2 u = *reinterpret_cast<const ulonglong2 *>(r + d), w = *reinterpret_cast<const ulonglong2 *>(c + d);
            *reinterpret_cast<ulonglong2 *>(o + d) = make_ulonglong2(one(u.x, w.x, a), one(u.y, w.y, b));
        }
    };

    // ---- epilogue that leaves the transform in registers (used by the fused inner product) --------
    struct KeepInRegs
    {
        GSB_DEV void store(uint32_t, uint64_t, int) const {}
        GSB_DEV void store2(uint32_t, uint64_t, uint64_t, int) const {}
    };

    // ---- fused (last NTT pass + key inner product + lazy reduction) --------------------------------
    // evaluator.cpp:2315-2335 (step 3) and :2341-2350 (step 4, Barrett part).  One CTA owns a tile of
    // ext slot kap of ciphertext b and loops over the bundles: the lazy forward NTT of the mod-up'd
    // bundle (its last pass when N >= 2^13, the whole transform otherwise) finishes in registers and
    // is multiplied straight into two 128-bit accumulators per coefficient (FMAU128), so neither the
    // transformed operand nor the accumulators ever touch HBM.  Bundles that contain kap use the
    // original NTT-form target limb (:2319-2320).  grid = (B, tiles per limb, E): ciphertexts of a
    // batch run side by side on the same key tile, so a key word is fetched from HBM once per batch.
    struct KsMacArgs
    {
        KsDev ks;
        const uint64_t *target; // ciphertext b at target + b*target_stride, [L][N], NTT form
        size_t target_stride;
        const uint64_t *coef; // [B][L][N] coefficient form (single-pass prologue)
        const uint64_t *ext1; // [B][beta][E][N] column-pass output (two-pass)
        const uint64_t *key;  // [beta0][2][K][N]
        uint64_t *red;        // [B][2][E][N]
        uint32_t K;
    };

    template <int LOGN>
    struct MacPlan
    {
        static constexpr bool two_pass = Split<LOGN>::two_pass;
#ifndef GSB_MAC_T
#define GSB_MAC_T 7 // log2(threads) of a two-pass ks_mac CTA (tile = 8 << T words): measured 3.10 ms at 7, 3.14 at 6, 3.22 at 8
#endif
        static constexpr int T = two_pass ? GSB_MAC_T : Split<LOGN>::T1;
        static constexpr int min_ctas = T > 8 ? 1 : T == 8 ? 2 : T == 7 ? 4 : 8; // 512 resident threads per SM
        static constexpr int G_HI = two_pass ? LOGN - Split<LOGN>::A - 1 : LOGN - 1;
        typedef TilePlan<T, G_HI, 0> TP;
        static constexpr uint32_t TILE = 8u << T;
        // two-pass plans stage the operand tile (TILE words) and the two key tiles (2*TILE words)
        // through shared memory with bulk copies, behind three mbarriers
        static constexpr size_t stage_words = two_pass ? size_t(3) * TILE : 0;
        static constexpr size_t smem_bytes = TP::smem_bytes + stage_words * sizeof(uint64_t) + 16;
    };

    // reads of a staged (linear) operand tile in the first round's register layout
    struct StagedPro
    {
        const uint64_t *st;
        GSB_DEV uint64_t raw(uint32_t d) const { return st[d]; }
        GSB_DEV void raw2(uint32_t d, uint64_t &a, uint64_t &b) const
        {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(st + d);
            a = v.x;
            b = v.y;
        }
        GSB_DEV uint64_t xform(uint64_t v) const { return v; }
    };

    // LAZY (see KsDev::lazy_mac): the operand's last stage skips reduceBarrettLazy.
#ifndef GSB_MAC_FWD_FORM
#define GSB_MAC_FWD_FORM 3 // fwd_round's FAST form for the operand transform: measured 3.22 ms for D3 (3), 3.34 plain (1), 3.47 H3 + D3 (4), 3.56 H3 (2), 3.32 D3 + EARLY (7)
#endif
    template <int LOGN, bool K1, bool LAZY, int FORM = GSB_MAC_FWD_FORM>
    __global__ void __launch_bounds__(1 << MacPlan<LOGN>::T, MacPlan<LOGN>::min_ctas) ks_mac_kernel(KsMacArgs p, NttTables tab)
    {
        typedef MacPlan<LOGN> MP;
        typedef typename MP::TP TP;
        constexpr int T = MP::T;
        const KsDev &ks = p.ks;
        const int b = blockIdx.x, tile = blockIdx.y, kap = blockIdx.z;
        const uint32_t tid = threadIdx.x, N = tab.N;
        const int mod = ks.ext_mod(kap);
        const LimbC c = tab.limb[mod];
        const Twiddle *WW = tab.WW + static_cast<size_t>(mod) * N;
        const ContigMap<T> map{ N };
        const uint32_t d0 = map.coeff(tile, tid * 8);
        uint64_t *smem = dyn_smem();
        Twiddle *tws = reinterpret_cast<Twiddle *>(smem + TP::exch_words);
        uint64_t *st_op = reinterpret_cast<uint64_t *>(tws + TP::PAIRS);
        uint64_t *st_key = st_op + (MP::two_pass ? MP::TILE : 0);
        uint64_t *bar = st_op + MP::stage_words; // [0] twiddles, [1] operand tile, [2] key tiles
        const int nb = static_cast<int>(ks.beta), k = static_cast<int>(ks.k), L = static_cast<int>(ks.L);
        auto in_bundle = [&](int j) { return kap >= j * k && kap < j * k + k && kap < L; };
        constexpr uint32_t tile_bytes = MP::TILE * static_cast<uint32_t>(sizeof(uint64_t));
        // bulk copy of bundle j's operand tile: the column-pass output, or the original NTT-form
        // target limb for the bundle that contains kap (:2319-2320)
        auto issue_operand = [&](int j) {
            const uint64_t *src = in_bundle(j)
                                      ? p.target + static_cast<size_t>(b) * p.target_stride + static_cast<size_t>(kap) * N
                                      : p.ext1 + (static_cast<size_t>(b * nb + j) * ks.E + kap) * N;
            mbar_expect(bar + 1, tile_bytes);
            bulk_g2s(st_op, src + map.off(tile), tile_bytes, bar + 1);
        };
        auto issue_key = [&](int j) {
            const uint64_t *k0 = p.key + (static_cast<size_t>(j * 2) * p.K + mod) * N + map.off(tile);
            mbar_expect(bar + 2, 2 * tile_bytes);
            bulk_g2s(st_key, k0, tile_bytes, bar + 2);
            bulk_g2s(st_key + MP::TILE, k0 + static_cast<size_t>(p.K) * N, tile_bytes, bar + 2);
        };
        if (tid == 0)
        {
            mbar_init(bar);
            if constexpr (MP::two_pass)
            {
                mbar_init(bar + 1);
                mbar_init(bar + 2);
            }
        }
        __syncthreads();
        if (tid == 0)
        {
            if constexpr (MP::two_pass)
                issue_operand(0);
            TP::TwC::stage(tws, WW, N, map.off(tile), false, bar);
        }
        const typename TP::TwC tw{ tws, TwGlobal{ WW, N, map.off(tile), N } };
        uint64_t lo[2][8], hi[2][8];
#pragma unroll
        for (int r = 0; r < 8; r++)
            lo[0][r] = hi[0][r] = lo[1][r] = hi[1][r] = 0;
        mbar_wait(bar, 0);
        if constexpr (LAZY)
        {
            // every stage of this kernel's transform is a wide-range stage: shift the staged Shoup halves
            // once per CTA instead of once per use (15 bundles x 36 butterflies per thread)
            for (uint32_t i = tid; i < TP::PAIRS; i += (1u << T))
                tws[i].y >>= 1;
            __syncthreads();
        }
        int parity = 0;
        for (int j = 0; j < nb; j++)
        {
            uint64_t x[8];
            const bool direct = in_bundle(j);
            if constexpr (MP::two_pass)
            {
                mbar_wait(bar + 1, j & 1);
                if (direct)
                {
#pragma unroll
                    for (int r = 0; r < 8; r += 2)
                    {
                        const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(st_op + tid * 8 + r);
                        x[r] = v.x;
                        x[r + 1] = v.y;
                    }
                }
                else
                    xf_load_raw<T, MP::G_HI, 0, false>(x, tid, 0, map, StagedPro{ st_op });
                // every thread holds its operand words and is past the previous bundle's key reads:
                // both staging areas may be refilled while the butterflies run.  The refill is an
                // async-proxy write after generic-proxy reads of the same bytes, which only a proxy
                // fence orders (without it a late LDS occasionally saw the next bundle's tile).
                fence_proxy_async();
                __syncthreads();
                if (tid == 0)
                {
                    issue_key(j);
                    if (j + 1 < nb)
                        issue_operand(j + 1);
                }
            }
            else if (direct)
            {
                const uint64_t *t = p.target + static_cast<size_t>(b) * p.target_stride + static_cast<size_t>(kap) * N + d0;
#pragma unroll
                for (int r = 0; r < 8; r += 2)
                {
                    const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(t + r);
                    x[r] = v.x;
                    x[r + 1] = v.y;
                }
            }
            else
            {
                const ModupCoreT<K1, LAZY> pro(ks, p.coef, b, j, kap, c);
                xf_load_raw<T, MP::G_HI, 0, false>(x, tid, tile, map, pro);
#pragma unroll
                for (int r = 0; r < 8; r++)
                    x[r] = pro.xform(x[r]);
            }
            if (!direct)
            {
                if constexpr (LAZY)
                    fwd_compute<T, MP::G_HI, 0, false, 0, FORM>(x, smem, tid, TwPreshifted<typename TP::TwC>{ tw }, c, parity);
                else
                    fwd_compute<T, MP::G_HI, 0, true, 0, false>(x, smem, tid, tw, c, parity);
            }
            // key tiles are stored with their 16-byte units swizzled (key_word_pos): the thread's
            // words 2r, 2r+1 are unit swz8(4 tid + r) of the tile -- conflict-free for a warp
            const ulonglong2 *k0, *k1;
            if constexpr (MP::two_pass)
            {
                mbar_wait(bar + 2, j & 1);
                k0 = reinterpret_cast<const ulonglong2 *>(st_key);
                k1 = k0 + MP::TILE / 2;
            }
            else
            {
                k0 = reinterpret_cast<const ulonglong2 *>(p.key + (static_cast<size_t>(j * 2) * p.K + mod) * N + map.off(tile));
                k1 = k0 + static_cast<size_t>(p.K) * N / 2;
            }
#pragma unroll
            for (int r = 0; r < 4; r++)
            {
                const uint32_t unit = swz8(tid * 4 + r);
                const ulonglong2 u = k0[unit];
                const ulonglong2 v = k1[unit];
                if constexpr (LAZY) // operand and key words are < 2^63 (gsb_ctx::fast_ntt)
                {
                    mac128_mid(lo[0][2 * r], hi[0][2 * r], x[2 * r], u.x);
                    mac128_mid(lo[0][2 * r + 1], hi[0][2 * r + 1], x[2 * r + 1], u.y);
                    mac128_mid(lo[1][2 * r], hi[1][2 * r], x[2 * r], v.x);
                    mac128_mid(lo[1][2 * r + 1], hi[1][2 * r + 1], x[2 * r + 1], v.y);
                }
                else
                {
                    mac128(lo[0][2 * r], hi[0][2 * r], x[2 * r], u.x);
                    mac128(lo[0][2 * r + 1], hi[0][2 * r + 1], x[2 * r + 1], u.y);
                    mac128(lo[1][2 * r], hi[1][2 * r], x[2 * r], v.x);
                    mac128(lo[1][2 * r + 1], hi[1][2 * r + 1], x[2 * r + 1], v.y);
                }
            }
        }
#pragma unroll
        for (int l = 0; l < 2; l++)
        {
            uint64_t *o = p.red + ((static_cast<size_t>(b) * 2 + l) * ks.E + kap) * N + d0;
#pragma unroll
            for (int r = 0; r < 8; r += 2)
                *reinterpret_cast<ulonglong2 *>(o + r) = make_ulonglong2(
                    barrett128(lo[l][r], hi[l][r], c.q, c.cr0, c.cr1), barrett128(lo[l][r + 1], hi[l][r + 1], c.q, c.cr0, c.cr1));
        }
    }

    // ---- GaloisTool::apply_galois, coefficient form (util/galois.cpp:144-186) ------------------------
    // in = [n][L][N] at in + p*in_zs -> out [n][L][N]: out[(i*elt) mod N] = +-in[i], negated when
    // bit logN of i*elt is set.
    struct GaloisCoeffArgs
    {
        const uint64_t *in;
        size_t in_zs;
        uint64_t *out;
        const LimbC *limb;
        uint32_t elt, N, L;
        int logn;
        size_t total; // n*L*N
    };

    __global__ void __launch_bounds__(256) galois_coeff_kernel(GaloisCoeffArgs p)
    {
        const size_t LN = static_cast<size_t>(p.L) * p.N;
        for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < p.total;
             t += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t poly = t / LN, rem = t - poly * LN;
            const uint32_t limb = static_cast<uint32_t>(rem / p.N), i = static_cast<uint32_t>(rem - static_cast<size_t>(limb) * p.N);
            const uint64_t raw = static_cast<uint64_t>(i) * p.elt;
            uint64_t v = p.in[poly * p.in_zs + rem];
            if ((raw >> p.logn) & 1)
                v = neg_mod(v, p.limb[limb].q);
            p.out[poly * LN + static_cast<size_t>(limb) * p.N + (raw & (p.N - 1))] = v;
        }
    }

    // ---- CKKS tensor product of two size-2 ciphertexts (evaluator.cpp:493-520) ---------------------
    // a, b = [B][2][L][N] -> out = [B][3][L][N]: c0 = a0*b0, c1 = a0*b1 + a1*b0, c2 = a1*b1.
    struct CkksMulArgs
    {
        const uint64_t *a, *b;
        uint64_t *out;
        const LimbC *limb;
        uint32_t N, L;
        size_t pairs; // B*L*N/2
        // word w of the batch is limb (w >> logn) % L of ciphertext (w >> logn) / L: a shift and a multiply by
        // l_magic = ceil(2^32 / L) (0 for L = 1; exact while (B L) L < 2^32) instead of two 64-bit divisions per pair
        uint32_t logn, l_magic;
        CkksMulArgs(const uint64_t *a_, const uint64_t *b_, uint64_t *out_, const LimbC *limb_, uint32_t N_, uint32_t L_, size_t pairs_)
            : a(a_), b(b_), out(out_), limb(limb_), N(N_), L(L_), pairs(pairs_), logn(0),
              l_magic(L_ > 1 ? static_cast<uint32_t>(((uint64_t(1) << 32) + L_ - 1) / L_) : 0u)
        {
            while ((1u << logn) < N_)
                logn++;
        }
    };

    // FASTRED: operands are canonical residues and the limb has c.mu (every modulus of <= 60 bits): the
    // products are reduced with reduce_2k; otherwise (BFV tensor product over lazy NTT words, 61-bit
    // moduli) with the literal barrett_reduce_128.  Same canonical words either way.
    template <bool FASTRED>
    GSB_DEV uint64_t prod_mod(uint64_t a, uint64_t b, const LimbC &c)
    {
        const uint64_t lo = a * b, hi = mulhi(a, b);
        return (FASTRED && c.mu) ? reduce_2k(lo, hi, c) : barrett128(lo, hi, c.q, c.cr0, c.cr1);
    }
    // a*b + x*y mod q: the reference reduces both products and adds mod q (evaluator.cpp:503-512); the
    // canonical residue of the exact 128-bit sum (< 2^123 for 61-bit moduli) is the same word
    template <bool FASTRED>
    GSB_DEV uint64_t cross_mod(uint64_t a, uint64_t b, uint64_t x, uint64_t y, const LimbC &c)
    {
        uint64_t lo = a * b, hi = mulhi(a, b);
        mac128(lo, hi, x, y);
        return (FASTRED && c.mu) ? reduce_2k(lo, hi, c) : barrett128(lo, hi, c.q, c.cr0, c.cr1);
    }

    template <bool FASTRED>
    __global__ void __launch_bounds__(256) ckks_multiply_kernel(CkksMulArgs p)
    {
        const size_t LN = static_cast<size_t>(p.L) * p.N;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < p.pairs;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t w = i << 1;
            const uint32_t row = static_cast<uint32_t>(w >> p.logn); // ct * L + limb
            const uint32_t ct32 = p.l_magic ? __umulhi(row, p.l_magic) : row;
            const size_t ct = ct32, rem = w - ct * LN;
            const LimbC c = p.limb[row - ct32 * p.L];
            const uint64_t *pa = p.a + ct * 2 * LN + rem, *pb = p.b + ct * 2 * LN + rem;
            uint64_t *po = p.out + ct * 3 * LN + rem;
            const ulonglong2 a0 = *reinterpret_cast<const ulonglong2 *>(pa), a1 = *reinterpret_cast<const ulonglong2 *>(pa + LN);
            const ulonglong2 b0 = *reinterpret_cast<const ulonglong2 *>(pb), b1 = *reinterpret_cast<const ulonglong2 *>(pb + LN);
            ulonglong2 c0, c1, c2;
            c0.x = prod_mod<FASTRED>(a0.x, b0.x, c);
            c0.y = prod_mod<FASTRED>(a0.y, b0.y, c);
            c1.x = cross_mod<FASTRED>(a1.x, b0.x, a0.x, b1.x, c);
            c1.y = cross_mod<FASTRED>(a1.y, b0.y, a0.y, b1.y, c);
            c2.x = prod_mod<FASTRED>(a1.x, b1.x, c);
            c2.y = prod_mod<FASTRED>(a1.y, b1.y, c);
            *reinterpret_cast<ulonglong2 *>(po) = c0;
            *reinterpret_cast<ulonglong2 *>(po + LN) = c1;
            *reinterpret_cast<ulonglong2 *>(po + 2 * LN) = c2;
        }
    }

    // ---- switching-key upload: validity check + device layout ------------------------------------------
    // key = [rows][N] words as copied from the host (rows = bundles * 2 * K, row r uses modulus r % K).  64
    // threads own one block of 64 16-byte units: each loads its unit, checks both words against the modulus
    // (is_buffer_valid / is_data_valid_for, valcheck.cpp) and, after the CTA barrier, stores it at unit
    // swz8(u) of the same block (key_word_pos, ntt_core.cuh) -- an in-place permutation inside the block.
    struct KeySwizzleArgs
    {
        uint64_t *key;
        const LimbC *limb;
        uint32_t N, K;
        size_t blocks; // rows * N / 128
        int *bad;      // set to 1 if any word is not a canonical residue
    };

    __global__ void __launch_bounds__(256) key_swizzle_kernel(KeySwizzleArgs p)
    {
        const size_t units = p.blocks * 64, stride = static_cast<size_t>(gridDim.x) * blockDim.x;
        // every thread of a CTA runs the same number of iterations (units is a multiple of 64, the stride of 256)
        for (size_t base = static_cast<size_t>(blockIdx.x) * blockDim.x; base < units; base += stride)
        {
            const size_t u = base + threadIdx.x;
            ulonglong2 v = make_ulonglong2(0, 0);
            const bool live = u < units;
            if (live)
            {
                v = *reinterpret_cast<const ulonglong2 *>(p.key + 2 * u);
                const uint64_t q = p.limb[((2 * u) / p.N) % p.K].q;
                if (v.x >= q || v.y >= q)
                    *p.bad = 1;
            }
            __syncthreads();
            if (live)
            {
                const size_t dst = (u & ~static_cast<size_t>(63)) | swz8(static_cast<uint32_t>(u & 63));
                *reinterpret_cast<ulonglong2 *>(p.key + 2 * dst) = v;
            }
            __syncthreads();
        }
    }

    // ---- NTT-form Galois gather (GaloisTool::apply_galois_ntt, util/galois.cpp:188-214) ------------
    // out[z][i][d] = in[z*in_zs + i*N + table[d]] for z < n, i < L.
    struct GatherArgs
    {
        const uint64_t *in;
        size_t in_zs;
        uint64_t *out; // [n][L][N]
        const uint32_t *table;
        uint32_t N, L;
        size_t total; // n*L*N
    };

    __global__ void __launch_bounds__(256) galois_gather_kernel(GatherArgs p)
    {
        const size_t LN = static_cast<size_t>(p.L) * p.N;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < p.total;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            // N is a power of two: row = z * L + limb by a shift, z by one 32-bit division (not two 64-bit ones)
            const uint32_t row = static_cast<uint32_t>(i >> (31 - __clz(p.N))), zz = row / p.L;
            const size_t z = zz, rem = i - z * LN;
            const uint32_t d = static_cast<uint32_t>(rem & (p.N - 1));
            p.out[i] = p.in[z * p.in_zs + (rem - d) + __ldg(p.table + d)];
        }
    }
} // namespace gsb
