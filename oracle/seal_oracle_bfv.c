/* TEST INFRASTRUCTURE -- CPU oracle, BFV BEHZ multiplication (BASELINE config #1).
 *
 * Plain-C restatement of Evaluator::bfv_multiply (evaluator.cpp:274-445) and of the RNSTool /
 * BaseConverter pieces it calls (util/rns.cpp:441-496 fast_convert_array, :540-728 initialize,
 * :853-1068 fastbconv_sk / sm_mrq / fast_floor / fastbconv_m_tilde; util/numth.cpp:179-322 is_prime /
 * get_primes).  Same rules as seal_oracle.c: loaded only by tests/, smoke() and bench.py's CPU legs.
 * Parity status: PINNED against the unmodified reference (tests/test_oracle_pins.py, BFV cases).
 *
 * Every constant the reference derives from multi-precision products (punctured products, prod(q) mod
 * m, ...) is a residue modulo one word-size modulus, so it is computed here by modular multiplication;
 * the one quantity that needs the full product -- the bit length of prod(q) that decides |B| (rns.cpp:
 * 566-573) -- uses a small multiply-by-word routine.
 */
#include "seal_oracle.h"
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

static uint64_t mulmod(uint64_t a, uint64_t b, uint64_t m)
{
    return (uint64_t)(((u128)a * b) % m);
}

static uint64_t powmod(uint64_t a, uint64_t e, uint64_t m)
{
    uint64_t r = 1;
    a %= m;
    while (e)
    {
        if (e & 1)
            r = mulmod(r, a, m);
        a = mulmod(a, a, m);
        e >>= 1;
    }
    return r;
}

/* numth.cpp:179-275 is a Miller-Rabin test with random bases; primality itself is deterministic, so
 * the restatement uses the twelve bases that are exact for 64-bit integers. */
static int is_prime_u64(uint64_t n)
{
    static const uint64_t small[] = { 2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37 };
    if (n < 2)
        return 0;
    for (size_t i = 0; i < 12; i++)
    {
        if (n == small[i])
            return 1;
        if (n % small[i] == 0)
            return 0;
    }
    uint64_t d = n - 1;
    int r = 0;
    while (!(d & 1))
    {
        d >>= 1;
        r++;
    }
    for (size_t i = 0; i < 12; i++)
    {
        uint64_t x = powmod(small[i], d, n);
        if (x == 1 || x == n - 1)
            continue;
        int comp = 1;
        for (int j = 1; j < r; j++)
        {
            x = mulmod(x, x, n);
            if (x == n - 1)
            {
                comp = 0;
                break;
            }
        }
        if (comp)
            return 0;
    }
    return 1;
}

/* numth.cpp:277-322: descending from 2^bits - 2n + 1 in steps of 2n */
static int get_primes(size_t n, int bits, size_t count, uint64_t *out)
{
    const uint64_t factor = 2 * (uint64_t)n;
    uint64_t value = ((uint64_t)1 << bits) - factor + 1;
    const uint64_t lower = (uint64_t)1 << (bits - 1);
    size_t got = 0;
    while (got < count && value > lower)
    {
        if (is_prime_u64(value))
            out[got++] = value;
        value -= factor;
    }
    return got == count ? 0 : -1;
}

/* bit length of prod(q[0..L)) */
static int product_bit_count(const uint64_t *q, size_t L)
{
    uint64_t w[64] = { 1 };
    size_t n = 1;
    for (size_t i = 0; i < L; i++)
    {
        uint64_t carry = 0;
        for (size_t j = 0; j < n; j++)
        {
            u128 t = (u128)w[j] * q[i] + carry;
            w[j] = (uint64_t)t;
            carry = (uint64_t)(t >> 64);
        }
        if (carry)
            w[n++] = carry;
    }
    int bits = 0;
    for (uint64_t top = w[n - 1]; top; top >>= 1)
        bits++;
    return (int)(n - 1) * 64 + bits;
}

#define MAXB 40

typedef struct mod_t
{
    uint64_t v;
    uint64_t cr[3];
} mod_t;

static void mod_set(mod_t *m, uint64_t v)
{
    m->v = v;
    so_const_ratio(v, m->cr);
}
static uint64_t mm(uint64_t a, uint64_t b, const mod_t *m) /* multiply_uint_mod */
{
    return so_multiply_uint_mod(a, b, m->v, m->cr);
}
static uint64_t mam(uint64_t a, uint64_t b, uint64_t c, const mod_t *m) /* multiply_add_uint_mod */
{
    return so_multiply_add_uint_mod(a, b, c, m->v, m->cr);
}

/* BaseConverter (rns.cpp:498-523): inv_punct[i] = (prod_{l != i} ibase_l)^-1 mod ibase_i;
 * matrix[j][i] = prod_{l != i} ibase_l mod obase_j */
typedef struct conv_t
{
    size_t ni, no;
    mod_t ib[MAXB], ob[MAXB];
    uint64_t inv_punct[MAXB];
    uint64_t matrix[MAXB][MAXB];
} conv_t;

static int conv_init(conv_t *c, const uint64_t *ibase, size_t ni, const uint64_t *obase, size_t no)
{
    c->ni = ni;
    c->no = no;
    for (size_t i = 0; i < ni; i++)
        mod_set(&c->ib[i], ibase[i]);
    for (size_t j = 0; j < no; j++)
        mod_set(&c->ob[j], obase[j]);
    for (size_t i = 0; i < ni; i++)
    {
        uint64_t p = 1;
        for (size_t l = 0; l < ni; l++)
            if (l != i)
                p = mulmod(p, ibase[l] % ibase[i], ibase[i]);
        if (!so_try_invert_uint_mod(p, ibase[i], &c->inv_punct[i]))
            return -1;
    }
    for (size_t j = 0; j < no; j++)
        for (size_t i = 0; i < ni; i++)
        {
            uint64_t p = 1 % obase[j];
            for (size_t l = 0; l < ni; l++)
                if (l != i)
                    p = mulmod(p, ibase[l] % obase[j], obase[j]);
            c->matrix[j][i] = p;
        }
    return 0;
}

/* fast_convert_array (rns.cpp:467-496); in = [ni][n], out = [no][n]; dot_product_mod
 * (uintarithsmallmod.cpp:110-173) = exact 128-bit sum, one Barrett-128 at the end */
static void conv_array(const conv_t *c, const uint64_t *in, size_t n, uint64_t *out)
{
    uint64_t *temp = (uint64_t *)malloc(c->ni * n * 8);
    for (size_t i = 0; i < c->ni; i++)
        for (size_t k = 0; k < n; k++)
            temp[k * c->ni + i] = mm(in[i * n + k], c->inv_punct[i], &c->ib[i]);
    for (size_t j = 0; j < c->no; j++)
        for (size_t k = 0; k < n; k++)
        {
            u128 acc = 0;
            for (size_t i = 0; i < c->ni; i++)
                acc += (u128)temp[k * c->ni + i] * c->matrix[j][i];
            out[j * n + k] = so_barrett_reduce_128((uint64_t)acc, (uint64_t)(acc >> 64), c->ob[j].v, c->ob[j].cr);
        }
    free(temp);
}

typedef struct behz_t
{
    size_t n, Lq, nB, nBsk;
    mod_t q[MAXB], Bsk[MAXB], m_tilde, m_sk, t;
    uint64_t qv[MAXB], Bskv[MAXB];
    conv_t q_to_Bsk, q_to_mtilde, B_to_q, B_to_msk;
    uint64_t prod_B_mod_q[MAXB], inv_prod_q_mod_Bsk[MAXB], inv_prod_B_mod_msk, inv_mtilde_mod_Bsk[MAXB];
    uint64_t inv_prod_q_mod_mtilde, prod_q_mod_Bsk[MAXB];
    so_tables qtab[MAXB], Bsktab[MAXB];
} behz_t;

/* RNSTool::initialize (rns.cpp:540-728), the parts bfv_multiply uses */
static int behz_init(behz_t *b, int logn, const uint64_t *q, size_t L, uint64_t t)
{
    memset(b, 0, sizeof(*b));
    const size_t n = (size_t)1 << logn;
    if (L < 1 || L + 3 > MAXB)
        return -1;
    b->n = n;
    b->Lq = L;
    mod_set(&b->t, t);
    int t_bits = 0;
    for (uint64_t x = t; x; x >>= 1)
        t_bits++;
    size_t nB = L; /* :566-573 */
    if (32 + t_bits + product_bit_count(q, L) >= 61 * (int)L + 61)
        nB++;
    b->nB = nB;
    b->nBsk = nB + 1;
    uint64_t primes[MAXB];
    /* :585: SEAL_USER_MOD_BIT_COUNT_MAX + 1 = 60 bits in this fork (util/defines.h:40); m_sk, gamma, then B */
    if (get_primes(n, 60, nB + 2, primes))
        return -1;
    mod_set(&b->m_sk, primes[0]);
    mod_set(&b->m_tilde, (uint64_t)1 << 32);
    uint64_t Bv[MAXB];
    for (size_t i = 0; i < nB; i++)
        Bv[i] = primes[2 + i];
    for (size_t i = 0; i < nB; i++)
        b->Bskv[i] = Bv[i];
    b->Bskv[nB] = primes[0];
    for (size_t i = 0; i < b->nBsk; i++)
        mod_set(&b->Bsk[i], b->Bskv[i]);
    for (size_t i = 0; i < L; i++)
    {
        b->qv[i] = q[i];
        mod_set(&b->q[i], q[i]);
        if (so_tables_init(&b->qtab[i], logn, q[i]))
            return -1;
    }
    for (size_t i = 0; i < b->nBsk; i++)
        if (so_tables_init(&b->Bsktab[i], logn, b->Bskv[i]))
            return -1;
    const uint64_t mt = b->m_tilde.v, msk = b->m_sk.v;
    if (conv_init(&b->q_to_Bsk, q, L, b->Bskv, b->nBsk) || conv_init(&b->q_to_mtilde, q, L, &mt, 1) ||
        conv_init(&b->B_to_q, Bv, nB, q, L) || conv_init(&b->B_to_msk, Bv, nB, &msk, 1))
        return -1;
    for (size_t i = 0; i < L; i++) /* :641-646 */
    {
        uint64_t p = 1;
        for (size_t l = 0; l < nB; l++)
            p = mulmod(p, Bv[l] % q[i], q[i]);
        b->prod_B_mod_q[i] = p;
    }
    for (size_t i = 0; i < b->nBsk; i++) /* :648-657, :679-684, :669-677 */
    {
        uint64_t p = 1;
        for (size_t l = 0; l < L; l++)
            p = mulmod(p, q[l] % b->Bskv[i], b->Bskv[i]);
        b->prod_q_mod_Bsk[i] = p;
        if (!so_try_invert_uint_mod(p, b->Bskv[i], &b->inv_prod_q_mod_Bsk[i]))
            return -1;
        if (!so_try_invert_uint_mod(mt % b->Bskv[i], b->Bskv[i], &b->inv_mtilde_mod_Bsk[i]))
            return -1;
    }
    {
        uint64_t p = 1; /* :659-664 */
        for (size_t l = 0; l < nB; l++)
            p = mulmod(p, Bv[l] % msk, msk);
        if (!so_try_invert_uint_mod(p, msk, &b->inv_prod_B_mod_msk))
            return -1;
        p = 1; /* :679-684 */
        for (size_t l = 0; l < L; l++)
            p = mulmod(p, q[l] % mt, mt);
        if (!so_try_invert_uint_mod(p, mt, &b->inv_prod_q_mod_mtilde))
            return -1;
    }
    return 0;
}

static void behz_free(behz_t *b)
{
    for (size_t i = 0; i < b->Lq; i++)
        so_tables_free(&b->qtab[i]);
    for (size_t i = 0; i < b->nBsk; i++)
        so_tables_free(&b->Bsktab[i]);
}

/* fastbconv_m_tilde (rns.cpp:1019-1066): in [Lq][n] -> out [nBsk + 1][n] */
static void fastbconv_m_tilde(const behz_t *b, const uint64_t *in, uint64_t *out)
{
    const size_t n = b->n;
    uint64_t *temp = (uint64_t *)malloc(b->Lq * n * 8);
    for (size_t i = 0; i < b->Lq; i++)
        for (size_t k = 0; k < n; k++)
            temp[i * n + k] = mm(in[i * n + k], b->m_tilde.v, &b->q[i]);
    conv_array(&b->q_to_Bsk, temp, n, out);
    conv_array(&b->q_to_mtilde, temp, n, out + b->nBsk * n);
    free(temp);
}

/* sm_mrq (rns.cpp:926-977): in [nBsk + 1][n] -> out [nBsk][n] */
static void sm_mrq(const behz_t *b, const uint64_t *in, uint64_t *out)
{
    const size_t n = b->n;
    const uint64_t mt = b->m_tilde.v, half = mt >> 1;
    uint64_t *r = (uint64_t *)malloc(n * 8);
    for (size_t i = 0; i < n; i++)
        r[i] = so_negate_uint_mod(mm(in[b->nBsk * n + i], b->inv_prod_q_mod_mtilde, &b->m_tilde), mt);
    for (size_t k = 0; k < b->nBsk; k++)
        for (size_t i = 0; i < n; i++)
        {
            uint64_t temp = r[i];
            if (temp >= half)
                temp += b->Bsk[k].v - mt;
            out[k * n + i] = mm(mam(b->prod_q_mod_Bsk[k], temp, in[k * n + i], &b->Bsk[k]), b->inv_mtilde_mod_Bsk[k], &b->Bsk[k]);
        }
    free(r);
}

/* fast_floor (rns.cpp:979-1017): in [Lq + nBsk][n] -> out [nBsk][n] */
static void fast_floor(const behz_t *b, const uint64_t *in, uint64_t *out)
{
    const size_t n = b->n;
    conv_array(&b->q_to_Bsk, in, n, out);
    in += b->Lq * n;
    for (size_t i = 0; i < b->nBsk; i++)
        for (size_t k = 0; k < n; k++)
            out[i * n + k] = mm(in[i * n + k] + (b->Bsk[i].v - out[i * n + k]), b->inv_prod_q_mod_Bsk[i], &b->Bsk[i]);
}

/* fastbconv_sk (rns.cpp:853-924): in [nBsk][n] -> out [Lq][n] */
static void fastbconv_sk(const behz_t *b, const uint64_t *in, uint64_t *out)
{
    const size_t n = b->n;
    const uint64_t msk = b->m_sk.v, half = msk >> 1;
    conv_array(&b->B_to_q, in, n, out);
    uint64_t *temp = (uint64_t *)malloc(n * 8), *alpha = (uint64_t *)malloc(n * 8);
    conv_array(&b->B_to_msk, in, n, temp);
    for (size_t i = 0; i < n; i++)
        alpha[i] = mm(temp[i] + (msk - in[b->nB * n + i]), b->inv_prod_B_mod_msk, &b->m_sk);
    for (size_t i = 0; i < b->Lq; i++)
        for (size_t k = 0; k < n; k++)
        {
            uint64_t *d = out + i * n + k;
            if (alpha[k] > half)
                *d = mam(b->prod_B_mod_q[i], msk - alpha[k], *d, &b->q[i]);
            else
                *d = mam(b->q[i].v - b->prod_B_mod_q[i], alpha[k], *d, &b->q[i]);
        }
    free(temp);
    free(alpha);
}

/* Evaluator::bfv_multiply (evaluator.cpp:274-445) for ciphertexts of sizes sa, sb:
 * a = [sa][L][n], b = [sb][L][n] coefficient form over q[0..L); out = [sa + sb - 1][L][n]. */
int so_bfv_multiply_sizes(int logn, const uint64_t *q, size_t L, uint64_t plain_modulus, const uint64_t *a, size_t sa, const uint64_t *b, size_t sb, uint64_t *out)
{
    behz_t *z = (behz_t *)malloc(sizeof(behz_t));
    if (!z)
        return -1;
    if (sa < 1 || sb < 1 || behz_init(z, logn, q, L, plain_modulus))
    {
        free(z);
        return -1;
    }
    const size_t n = z->n, nBsk = z->nBsk, sd = sa + sb - 1;
    uint64_t *eq[2], *eB[2];
    const uint64_t *src[2] = { a, b };
    const size_t sz[2] = { sa, sb };
    uint64_t *temp = (uint64_t *)malloc((nBsk + 1) * n * 8);
    for (int s = 0; s < 2; s++)
    {
        eq[s] = (uint64_t *)malloc(sz[s] * L * n * 8);
        eB[s] = (uint64_t *)malloc(sz[s] * nBsk * n * 8);
        for (size_t p = 0; p < sz[s]; p++) /* behz_extend_base_convert_to_ntt (:331-350) */
        {
            const uint64_t *in = src[s] + p * L * n;
            uint64_t *oq = eq[s] + p * L * n, *oB = eB[s] + p * nBsk * n;
            memcpy(oq, in, L * n * 8);
            for (size_t i = 0; i < L; i++)
                so_ntt_fwd_lazy(oq + i * n, &z->qtab[i]);
            fastbconv_m_tilde(z, in, temp);
            sm_mrq(z, temp, oB);
            for (size_t i = 0; i < nBsk; i++)
                so_ntt_fwd_lazy(oB + i * n, &z->Bsktab[i]);
        }
    }
    free(temp);
    uint64_t *dq = (uint64_t *)calloc(sd * L * n, 8), *dB = (uint64_t *)calloc(sd * nBsk * n, 8);
    for (size_t I = 0; I < sd; I++) /* step (4), :372-421 */
    {
        size_t i_last = I < sa - 1 ? I : sa - 1, j_first = I < sb - 1 ? I : sb - 1, i_first = I - j_first;
        for (size_t i = i_first; i <= i_last; i++)
        {
            const size_t j = I - i;
            for (size_t l = 0; l < L; l++)
            {
                const uint64_t *x = eq[0] + (i * L + l) * n, *y = eq[1] + (j * L + l) * n;
                uint64_t *d = dq + (I * L + l) * n;
                for (size_t k = 0; k < n; k++)
                    d[k] = so_add_uint_mod(mm(x[k], y[k], &z->q[l]), d[k], z->q[l].v);
            }
            for (size_t l = 0; l < nBsk; l++)
            {
                const uint64_t *x = eB[0] + (i * nBsk + l) * n, *y = eB[1] + (j * nBsk + l) * n;
                uint64_t *d = dB + (I * nBsk + l) * n;
                for (size_t k = 0; k < n; k++)
                    d[k] = so_add_uint_mod(mm(x[k], y[k], &z->Bsk[l]), d[k], z->Bsk[l].v);
            }
        }
    }
    uint64_t *tqB = (uint64_t *)malloc((L + nBsk) * n * 8), *tB = (uint64_t *)malloc(nBsk * n * 8);
    for (size_t I = 0; I < sd; I++)
    {
        for (size_t l = 0; l < L; l++) /* step (5), :424-425 */
            so_ntt_inv(dq + (I * L + l) * n, &z->qtab[l]);
        for (size_t l = 0; l < nBsk; l++)
            so_ntt_inv(dB + (I * nBsk + l) * n, &z->Bsktab[l]);
        for (size_t l = 0; l < L; l++) /* step (6), :433-435 */
            for (size_t k = 0; k < n; k++)
                tqB[l * n + k] = mm(dq[(I * L + l) * n + k], plain_modulus, &z->q[l]);
        for (size_t l = 0; l < nBsk; l++)
            for (size_t k = 0; k < n; k++)
                tqB[(L + l) * n + k] = mm(dB[(I * nBsk + l) * n + k], plain_modulus, &z->Bsk[l]);
        fast_floor(z, tqB, tB);                 /* step (7) */
        fastbconv_sk(z, tB, out + I * L * n);   /* step (8) */
    }
    free(tqB);
    free(tB);
    free(dq);
    free(dB);
    for (int s = 0; s < 2; s++)
    {
        free(eq[s]);
        free(eB[s]);
    }
    behz_free(z);
    free(z);
    return 0;
}

int so_bfv_multiply(int logn, const uint64_t *q, size_t L, uint64_t plain_modulus, const uint64_t *a, const uint64_t *b, uint64_t *out)
{
    return so_bfv_multiply_sizes(logn, q, L, plain_modulus, a, 2, b, 2, out);
}

/* The auxiliary base the reference derives for (n, q[0..L), t): out[0] = |B|, then the |B| + 1 Bsk
 * moduli (B primes followed by m_sk).  Lets the tests compare the product's base with the oracle's. */
int so_bfv_aux_base(int logn, const uint64_t *q, size_t L, uint64_t plain_modulus, uint64_t *out)
{
    behz_t *z = (behz_t *)malloc(sizeof(behz_t));
    if (!z || behz_init(z, logn, q, L, plain_modulus))
    {
        free(z);
        return -1;
    }
    out[0] = z->nB;
    for (size_t i = 0; i < z->nBsk; i++)
        out[1 + i] = z->Bskv[i];
    behz_free(z);
    free(z);
    return 0;
}

/* ---- Decryptor ---------------------------------------------------------------
 * dot_product_ct_sk_array (decryptor.cpp:218-265), compute_secret_key_array (:151-216), bfv_decrypt
 * (:76-113) and RNSTool::decrypt_scale_and_round (util/rns.cpp:1070-1126) with the {t, gamma}
 * constants of RNSTool::initialize (:585-590, :603-608, :634-638, :690-717).  Pinned against the
 * unmodified reference's Decryptor::decrypt (tests/test_oracle_pins.py). */

/* ct [size][L][N] at level L; sk = NTT-form secret key at key level [K][N]; out [L][N].
 * ntt_form = Ciphertext::is_ntt_form (1 CKKS, 0 BFV). */
int so_dot_product_ct_sk(const so_ctx *c, size_t L, size_t size, const uint64_t *ct, const uint64_t *sk, int ntt_form, uint64_t *out)
{
    const size_t N = c->N, K = c->K;
    if (size < 2 || L < 1 || L > c->L0)
        return -1;
    /* secret key powers s^1 .. s^(size-1) at key level (:191-199) */
    uint64_t *pw = (uint64_t *)malloc((size - 1) * K * N * 8);
    memcpy(pw, sk, K * N * 8);
    for (size_t p = 1; p + 1 < size; p++)
        for (size_t i = 0; i < K; i++)
            so_poly_op(0, c->q[i], pw + ((p - 1) * K + i) * N, pw + i * N, 0, N, pw + (p * K + i) * N);
    uint64_t *copy = (uint64_t *)malloc((size - 1) * L * N * 8);
    memcpy(copy, ct + L * N, (size - 1) * L * N * 8); /* :237-238 */
    for (size_t p = 0; p + 1 < size; p++)
        for (size_t i = 0; i < L; i++)
        {
            uint64_t *x = copy + (p * L + i) * N;
            if (!ntt_form)
                so_ntt_fwd_lazy(x, &c->tab[i]);                                   /* :241-244 */
            so_poly_op(0, c->q[i], x, pw + (p * K + i) * N, 0, N, x);             /* :247-250 */
        }
    memset(out, 0, L * N * 8); /* :253 */
    for (size_t p = 0; p + 1 < size; p++)
        for (size_t i = 0; i < L; i++)
            so_poly_op(1, c->q[i], out + i * N, copy + (p * L + i) * N, 0, N, out + i * N); /* :254-256 */
    for (size_t i = 0; i < L; i++)
    {
        if (!ntt_form)
            so_ntt_inv(out + i * N, &c->tab[i]);                                  /* :258-262 */
        so_poly_op(1, c->q[i], out + i * N, ct + i * N, 0, N, out + i * N);       /* :265 */
    }
    free(copy);
    free(pw);
    return 0;
}

/* out [N] = plaintext coefficients mod t (the reference then trims leading zeros, :107-111) */
int so_bfv_decrypt(const so_ctx *c, size_t L, size_t size, uint64_t plain_modulus, const uint64_t *ct, const uint64_t *sk, uint64_t *out)
{
    const size_t N = c->N;
    if (L + 3 > MAXB)
        return -1;
    uint64_t *dot = (uint64_t *)malloc(L * N * 8);
    if (so_dot_product_ct_sk(c, L, size, ct, sk, 0, dot))
    {
        free(dot);
        return -1;
    }
    /* constants (rns.cpp:585-590: the second sampled prime is gamma) */
    uint64_t primes[MAXB];
    {
        int t_bits = 0;
        for (uint64_t x = plain_modulus; x; x >>= 1)
            t_bits++;
        size_t nB = L;
        if (32 + t_bits + product_bit_count(c->q, L) >= 61 * (int)L + 61)
            nB++;
        if (get_primes(N, 60, nB + 2, primes))
        {
            free(dot);
            return -1;
        }
    }
    mod_t t, gamma, q[MAXB];
    mod_set(&t, plain_modulus);
    mod_set(&gamma, primes[1]);
    const uint64_t tg[2] = { plain_modulus, primes[1] };
    conv_t *cv = (conv_t *)malloc(sizeof(conv_t));
    if (conv_init(cv, c->q, L, tg, 2))
    {
        free(cv);
        free(dot);
        return -1;
    }
    uint64_t inv_gamma_mod_t, neg_inv_q[2], prod_tg_mod_q[MAXB];
    int ok = so_try_invert_uint_mod(gamma.v % t.v, t.v, &inv_gamma_mod_t);          /* :693 */
    for (size_t i = 0; i < L; i++)
    {
        mod_set(&q[i], c->q[i]);
        prod_tg_mod_q[i] = mm(t.v, gamma.v, &q[i]);                                  /* :699-704 */
    }
    for (size_t j = 0; j < 2; j++)                                                    /* :706-716 */
    {
        uint64_t p = 1 % tg[j];
        for (size_t l = 0; l < L; l++)
            p = mulmod(p, c->q[l] % tg[j], tg[j]);
        ok = ok && so_try_invert_uint_mod(p, tg[j], &neg_inv_q[j]);
        neg_inv_q[j] = so_negate_uint_mod(neg_inv_q[j], tg[j]);
    }
    if (!ok)
    {
        free(cv);
        free(dot);
        return -1;
    }
    /* decrypt_scale_and_round (:1070-1126) */
    uint64_t *temp = (uint64_t *)malloc(L * N * 8), *ttg = (uint64_t *)malloc(2 * N * 8);
    for (size_t i = 0; i < L; i++)
        for (size_t d = 0; d < N; d++)
            temp[i * N + d] = mm(dot[i * N + d], prod_tg_mod_q[i], &q[i]);           /* :1077-1082 */
    conv_array(cv, temp, N, ttg);                                                    /* :1088 */
    for (size_t d = 0; d < N; d++)                                                   /* :1091-1096 */
    {
        ttg[d] = mm(ttg[d], neg_inv_q[0], &t);
        ttg[N + d] = mm(ttg[N + d], neg_inv_q[1], &gamma);
    }
    const uint64_t gamma_div_2 = gamma.v >> 1;
    for (size_t d = 0; d < N; d++)                                                   /* :1104-1125 */
    {
        uint64_t v;
        if (ttg[N + d] > gamma_div_2)
            v = so_add_uint_mod(ttg[d], (gamma.v - ttg[N + d]) % t.v, t.v);
        else
            v = so_sub_uint_mod(ttg[d], ttg[N + d] % t.v, t.v);
        if (v != 0)
            v = mm(v, inv_gamma_mod_t, &t);
        out[d] = v;
    }
    free(temp);
    free(ttg);
    free(cv);
    free(dot);
    return 0;
}
