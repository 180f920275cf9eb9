/* TEST INFRASTRUCTURE -- see seal_oracle.h.  Plain-C restatement of the reference hot path.
 * Reference paths are relative to /root/reference/native/src/seal/.                              */
#include "seal_oracle.h"
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* ======================================================================================== */
/* Word arithmetic                                                                          */
/* ======================================================================================== */

/* modulus.cpp:87-98 -- const_ratio = floor(2^128/q) (two words) and 2^128 mod q.
 * q >= 2 never divides 2^128-1+1 exactly unless q is a power of two; handle both. */
void so_const_ratio(uint64_t q, uint64_t cr[3])
{
    u128 all = ~(u128)0; /* 2^128 - 1 */
    u128 quo = all / q;
    uint64_t rem = (uint64_t)(all % q) + 1; /* (2^128-1) mod q + 1 */
    if (rem == q)
    {
        quo += 1;
        rem = 0;
    }
    cr[0] = (uint64_t)quo;
    cr[1] = (uint64_t)(quo >> 64);
    cr[2] = rem;
}

/* ntt.cpp:19-24, multi_special_primes.cpp:6-11 -- floor(x * 2^64 / q) */
uint64_t so_shoupify(uint64_t x, uint64_t q)
{
    return (uint64_t)((((u128)x) << 64) / q);
}

static inline uint64_t mulhi(uint64_t a, uint64_t b)
{
    return (uint64_t)(((u128)a * b) >> 64);
}

/* util/uintarithsmallmod.h:140-178 -- literal three-multiply Barrett, one conditional subtract. */
uint64_t so_barrett_reduce_128(uint64_t z0, uint64_t z1, uint64_t q, const uint64_t cr[3])
{
    uint64_t carry = mulhi(z0, cr[0]);
    u128 t2 = (u128)z0 * cr[1];
    uint64_t tmp1 = (uint64_t)t2 + carry;
    uint64_t tmp3 = (uint64_t)(t2 >> 64) + (tmp1 < carry);
    t2 = (u128)z1 * cr[0];
    uint64_t s = tmp1 + (uint64_t)t2;
    carry = (uint64_t)(t2 >> 64) + (s < tmp1);
    uint64_t quot = z1 * cr[1] + tmp3 + carry;
    uint64_t r = z0 - quot * q;
    return r - (r >= q ? q : 0);
}

/* util/uintarithsmallmod.h:181-207 */
uint64_t so_barrett_reduce_63(uint64_t x, uint64_t q, const uint64_t cr[3])
{
    uint64_t r = x - mulhi(x, cr[1]) * q;
    return r - (r >= q ? q : 0);
}

/* util/uintarithsmallmod.h:209-221 */
uint64_t so_multiply_uint_mod(uint64_t a, uint64_t b, uint64_t q, const uint64_t cr[3])
{
    u128 z = (u128)a * b;
    return so_barrett_reduce_128((uint64_t)z, (uint64_t)(z >> 64), q, cr);
}

/* util/uintarithsmallmod.h:282-290 */
uint64_t so_multiply_add_uint_mod(uint64_t a, uint64_t b, uint64_t c, uint64_t q, const uint64_t cr[3])
{
    u128 z = (u128)a * b + c;
    return so_barrett_reduce_128((uint64_t)z, (uint64_t)(z >> 64), q, cr);
}

/* util/uintarithsmallmod.h:93-112 */
uint64_t so_add_uint_mod(uint64_t a, uint64_t b, uint64_t q)
{
    uint64_t s = a + b;
    return s - (s >= q ? q : 0);
}

/* util/uintarithsmallmod.h:114-137 */
uint64_t so_sub_uint_mod(uint64_t a, uint64_t b, uint64_t q)
{
    uint64_t d = a - b;
    return d + (a < b ? q : 0);
}

/* util/uintarithsmallmod.h:51-65 */
uint64_t so_negate_uint_mod(uint64_t a, uint64_t q)
{
    return a ? q - a : 0;
}

/* util/uintarithsmallmod.h (try_invert_uint_mod -> numth xgcd): canonical inverse in [0,q). */
int so_try_invert_uint_mod(uint64_t a, uint64_t q, uint64_t *out)
{
    if (a == 0)
        return 0;
    __int128 t0 = 0, t1 = 1;
    uint64_t r0 = q, r1 = a % q;
    if (r1 == 0)
        return 0;
    while (r1)
    {
        uint64_t qt = r0 / r1;
        uint64_t r2 = r0 - qt * r1;
        __int128 t2 = t0 - (__int128)qt * t1;
        r0 = r1;
        r1 = r2;
        t0 = t1;
        t1 = t2;
    }
    if (r0 != 1)
        return 0;
    if (t0 < 0)
        t0 += q;
    *out = (uint64_t)t0;
    return 1;
}

uint64_t so_exponentiate_uint_mod(uint64_t a, uint64_t e, uint64_t q)
{
    uint64_t cr[3];
    so_const_ratio(q, cr);
    uint64_t r = 1 % q, b = a % q;
    while (e)
    {
        if (e & 1)
            r = so_multiply_uint_mod(r, b, q, cr);
        b = so_multiply_uint_mod(b, b, q, cr);
        e >>= 1;
    }
    return r;
}

/* util/common.h:345-398 */
uint32_t so_reverse_bits32(uint32_t x, int bits)
{
    if (bits == 0)
        return 0;
    x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
    x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
    x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
    x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
    x = (x >> 16) | (x << 16);
    return x >> (32 - bits);
}

/* ======================================================================================== */
/* NTT tables  (util/ntt.cpp:37-118, util/numth.cpp:352-425)                                */
/* ======================================================================================== */

/* numth.cpp:398-425 picks the numerically smallest primitive 2N-th root; the random start of
 * try_primitive_root (:352-396) does not influence the minimum, so a deterministic search for
 * any primitive root followed by the same minimisation yields the same value. */
static int minimal_primitive_root(uint64_t degree, uint64_t q, const uint64_t cr[3], uint64_t *out)
{
    if ((q - 1) % degree)
        return 0;
    uint64_t quotient = (q - 1) / degree, root = 0;
    for (uint64_t g = 2; g < 2000; g++)
    {
        uint64_t r = so_exponentiate_uint_mod(g, quotient, q);
        if (r && so_exponentiate_uint_mod(r, degree >> 1, q) == q - 1)
        {
            root = r;
            break;
        }
    }
    if (!root)
        return 0;
    uint64_t gsq = so_multiply_uint_mod(root, root, q, cr), cur = root;
    for (uint64_t i = 0; i < degree; i++)
    {
        if (cur < root)
            root = cur;
        cur = so_multiply_uint_mod(cur, gsq, q, cr);
    }
    *out = root;
    return 1;
}

/* ntt.cpp:101-111 -- destination[bitrev(i)] = root^i */
static void powers_bitrev(uint64_t root, int logn, uint64_t q, const uint64_t cr[3], uint64_t *dst)
{
    size_t n = (size_t)1 << logn;
    uint64_t prev = 1;
    dst[0] = 1;
    for (size_t i = 1; i < n; i++)
    {
        prev = so_multiply_uint_mod(prev, root, q, cr);
        dst[so_reverse_bits32((uint32_t)i, logn)] = prev;
    }
}

int so_tables_init(so_tables *t, int logn, uint64_t q)
{
    memset(t, 0, sizeof(*t));
    size_t n = (size_t)1 << logn;
    t->logn = logn;
    t->q = q;
    so_const_ratio(q, t->cr);
    if (!minimal_primitive_root(2 * (uint64_t)n, q, t->cr, &t->root))
        return -1;
    uint64_t inv_root;
    if (!so_try_invert_uint_mod(t->root, q, &inv_root))
        return -1;
    if (!so_try_invert_uint_mod((uint64_t)n, q, &t->inv_n))
        return -1;
    t->inv_n_s = so_shoupify(t->inv_n, q); /* ntt.cpp:73 */
    t->rdp = so_shoupify(1, q);            /* ntt.cpp:75 */
    t->w = (uint64_t *)malloc(n * 8);
    t->ws = (uint64_t *)malloc(n * 8);
    t->v = (uint64_t *)malloc(n * 8);
    t->vs = (uint64_t *)malloc(n * 8);
    uint64_t *tmp = (uint64_t *)malloc(n * 8);
    powers_bitrev(t->root, logn, q, t->cr, t->w);
    powers_bitrev(inv_root, logn, q, t->cr, tmp);
    /* ntt.cpp:85-96: re-order so the inverse transform walks the table sequentially:
     * after index 0 come the m = n/2 entries of the first inverse stage, then n/4, ... , 1. */
    t->v[0] = tmp[0];
    size_t pos = 1;
    for (size_t m = n >> 1; m > 0; m >>= 1)
        for (size_t i = 0; i < m; i++)
            t->v[pos++] = tmp[m + i];
    /* ntt.cpp:98: fold n^-1 into the last entry */
    t->v[n - 1] = so_multiply_uint_mod(t->v[n - 1], t->inv_n, q, t->cr);
    for (size_t i = 0; i < n; i++)
    {
        t->ws[i] = so_shoupify(t->w[i], q);
        t->vs[i] = so_shoupify(t->v[i], q);
    }
    free(tmp);
    return 0;
}

void so_tables_free(so_tables *t)
{
    free(t->w);
    free(t->ws);
    free(t->v);
    free(t->vs);
    memset(t, 0, sizeof(*t));
}

/* ======================================================================================== */
/* NTT  (util/ntt.cpp:213-404, util/ntt.h:225-246,318-334)                                  */
/* ======================================================================================== */

/* ntt.cpp:230-234 -- x*y - floor(x*y'/2^64)*q, all mod 2^64; result in [0,2q) */
static inline uint64_t mulmod_lazy(uint64_t x, uint64_t y, uint64_t ys, uint64_t q)
{
    return x * y - mulhi(x, ys) * q;
}

/* ntt.cpp:292-342.  Stage with m groups and gap h = n/(2m) uses twiddle index m + r for group r.
 * x0 is never reduced (ntt.cpp:245-252) except in the last stage (ForwardLazyLast, :254-261). */
void so_ntt_fwd_lazy(uint64_t *a, const so_tables *t)
{
    const size_t n = (size_t)1 << t->logn;
    const uint64_t q = t->q, two_q = q << 1;
    for (size_t m = 1, h = n >> 1; h >= 1; m <<= 1, h >>= 1)
    {
        for (size_t r = 0; r < m; r++)
        {
            const uint64_t w = t->w[m + r], ws = t->ws[m + r];
            uint64_t *x0 = a + 2 * h * r, *x1 = x0 + h;
            for (size_t i = 0; i < h; i++)
            {
                uint64_t u = x0[i];
                if (h == 1)
                    u = u - mulhi(u, t->rdp) * q; /* reduceBarrettLazy, ntt.cpp:237-241 */
                uint64_t v = mulmod_lazy(x1[i], w, ws, q);
                x0[i] = u + v;
                x1[i] = u - v + two_q;
            }
        }
    }
}

/* ntt.h:225-246 */
void so_ntt_fwd(uint64_t *a, const so_tables *t)
{
    so_ntt_fwd_lazy(a, t);
    const size_t n = (size_t)1 << t->logn;
    const uint64_t q = t->q, two_q = q * 2;
    for (size_t i = 0; i < n; i++)
    {
        if (a[i] >= two_q)
            a[i] -= two_q;
        if (a[i] >= q)
            a[i] -= q;
    }
}

/* ntt.cpp:345-404.  Stages run with m = n/2, n/4, ..., 2 groups (gap h = n/(2m)); the table was
 * re-ordered so the stage with m groups starts at v[n - 2m + 1].  The final stage (m = 1) folds
 * n^-1 into both outputs (BackwardLazyLast, :274-281).  Defined (like the reference) for n >= 8;
 * for n < 8 the reference's fixed three-loop structure is mirrored literally below. */
void so_ntt_inv_lazy(uint64_t *a, const so_tables *t)
{
    const size_t n = (size_t)1 << t->logn;
    const uint64_t q = t->q, two_q = q << 1;
    size_t tw = 1; /* running table index, ntt.cpp:349-350 */
    /* first loop (:354-361) h = 1, second loop (:363-372) h = 2, main loop (:374-389) h >= 4, m > 1 */
    for (int pass = 0; pass < 3; pass++)
    {
        size_t m_first = pass == 0 ? n >> 1 : pass == 1 ? n >> 2 : n >> 3;
        size_t h_first = pass == 0 ? 1 : pass == 1 ? 2 : 4;
        for (size_t m = m_first, h = h_first; pass == 2 ? m > 1 : m == m_first && m > 0; m >>= 1, h <<= 1)
        {
            for (size_t r = 0; r < m; r++, tw++)
            {
                const uint64_t w = t->v[tw], ws = t->vs[tw];
                uint64_t *x0 = a + 2 * h * r, *x1 = x0 + h;
                for (size_t i = 0; i < h; i++)
                {
                    uint64_t u = x0[i], v = x1[i];
                    uint64_t s = u + v;
                    s -= (s >= two_q) ? two_q : 0;
                    x0[i] = s;
                    x1[i] = mulmod_lazy(u - v + two_q, w, ws, q);
                }
            }
            if (pass != 2)
                break;
        }
    }
    /* last layer (:391-402) */
    const uint64_t w = t->v[tw], ws = t->vs[tw];
    uint64_t *x0 = a, *x1 = a + n / 2;
    for (size_t i = 0; i < n / 2; i++)
    {
        uint64_t u = x0[i], v = x1[i];
        uint64_t s = u + v;
        s -= (s >= two_q) ? two_q : 0;
        x0[i] = mulmod_lazy(s, t->inv_n, t->inv_n_s, q);
        x1[i] = mulmod_lazy(u - v + two_q, w, ws, q);
    }
}

/* ntt.h:318-334 */
void so_ntt_inv(uint64_t *a, const so_tables *t)
{
    so_ntt_inv_lazy(a, t);
    const size_t n = (size_t)1 << t->logn;
    for (size_t i = 0; i < n; i++)
        if (a[i] >= t->q)
            a[i] -= t->q;
}

int so_ntt(int logn, uint64_t q, uint64_t *data, size_t n_polys, int kind)
{
    so_tables t;
    /* the forward loops are generic (the reference's N = 2 KAT, tests/seal/util/ntt.cpp:98-102,
     * runs through them); the inverse restatement mirrors the three fixed loops, n >= 8 */
    if (logn < 1 || (kind >= 2 && logn < 3) || so_tables_init(&t, logn, q))
        return -1;
    size_t n = (size_t)1 << logn;
    for (size_t p = 0; p < n_polys; p++)
    {
        uint64_t *x = data + p * n;
        switch (kind)
        {
        case 0: so_ntt_fwd_lazy(x, &t); break;
        case 1: so_ntt_fwd(x, &t); break;
        case 2: so_ntt_inv_lazy(x, &t); break;
        case 3: so_ntt_inv(x, &t); break;
        default: so_tables_free(&t); return -2;
        }
    }
    so_tables_free(&t);
    return 0;
}

/* ======================================================================================== */
/* Elementwise  (util/polyarithsmallmod.cpp:15-117, util/polyarithsmallmod.h:176-470)       */
/* ======================================================================================== */
int so_poly_op(int op, uint64_t q, const uint64_t *a, const uint64_t *b, uint64_t scalar, size_t n, uint64_t *out)
{
    uint64_t cr[3];
    so_const_ratio(q, cr);
    for (size_t i = 0; i < n; i++)
    {
        switch (op)
        {
        case 0: out[i] = so_multiply_uint_mod(a[i], b[i], q, cr); break;   /* .cpp:63-117 */
        case 1: out[i] = so_add_uint_mod(a[i], b[i], q); break;            /* .h:261-300 */
        case 2: out[i] = so_sub_uint_mod(a[i], b[i], q); break;            /* .h:366-405 */
        case 3: out[i] = so_negate_uint_mod(a[i], q); break;               /* .h:176-206 */
        case 4: out[i] = so_multiply_uint_mod(a[i], scalar, q, cr); break; /* .cpp:15-61 */
        default: return -1;
        }
    }
    return 0;
}

/* ======================================================================================== */
/* Galois  (util/galois.cpp:18-91,188-214)                                                  */
/* ======================================================================================== */
void so_galois_table_ntt(int logn, uint32_t elt, uint32_t *table)
{
    const uint32_t n = 1u << logn;
    for (uint32_t i = 0; i < n; i++)
    {
        uint32_t rev = so_reverse_bits32(n + i, logn + 1);
        uint64_t raw = ((uint64_t)elt * rev) >> 1;
        table[i] = so_reverse_bits32((uint32_t)(raw & (n - 1)), logn);
    }
}

/* galois.cpp:49-91; generator_ = 5 in this fork (util/galois.h:169). The reference unit test
 * tests/seal/util/galois.cpp:28-55 still expects the stock generator 3 and is stale; the
 * compiled reference (oracle/_ref) is the authority and returns 5 for step 1. */
uint32_t so_elt_from_step(int logn, int step)
{
    const uint32_t n = 1u << logn;
    const uint64_t m = 2ull * n;
    if (step == 0)
        return (uint32_t)(m - 1);
    int neg = step < 0;
    uint32_t pos = (uint32_t)(neg ? -step : step);
    if (pos >= (n >> 1))
        return 0;
    uint32_t cnt = neg ? (n >> 1) - pos : pos;
    uint64_t e = 1;
    while (cnt--)
        e = (e * 5) & (m - 1);
    return (uint32_t)e;
}

void so_apply_galois_ntt(int logn, uint32_t elt, const uint64_t *in, uint64_t *out)
{
    const size_t n = (size_t)1 << logn;
    uint32_t *table = (uint32_t *)malloc(n * 4);
    so_galois_table_ntt(logn, elt, table);
    for (size_t i = 0; i < n; i++)
        out[i] = in[table[i]];
    free(table);
}

/* ======================================================================================== */
/* Context                                                                                  */
/* ======================================================================================== */
so_ctx *so_ctx_create(int logn, const uint64_t *moduli, size_t K, size_t k_special)
{
    if (logn < 3 || logn > 16 || K < 2 || k_special < 1 || k_special >= K)
        return NULL;
    so_ctx *c = (so_ctx *)calloc(1, sizeof(so_ctx));
    c->logn = logn;
    c->N = (size_t)1 << logn;
    c->K = K;
    c->k = k_special;
    c->L0 = K - k_special;
    c->q = (uint64_t *)malloc(K * 8);
    c->tab = (so_tables *)calloc(K, sizeof(so_tables));
    for (size_t i = 0; i < K; i++)
    {
        c->q[i] = moduli[i];
        if (so_tables_init(&c->tab[i], logn, moduli[i]))
        {
            so_ctx_destroy(c);
            return NULL;
        }
    }
    return c;
}

void so_ctx_destroy(so_ctx *c)
{
    if (!c)
        return;
    for (size_t i = 0; i < c->K; i++)
        if (c->tab[i].w)
            so_tables_free(&c->tab[i]);
    free(c->tab);
    free(c->q);
    free(c);
}

/* ======================================================================================== */
/* CKKS multiply  (evaluator.cpp:447-527)                                                   */
/* ======================================================================================== */
/* General sizes: a = [sa][L][N], b = [sb][L][N] -> out = [sa + sb - 1][L][N] (dest_size, :470-472). */
int so_ckks_multiply_sizes(const so_ctx *c, size_t L, const uint64_t *a, size_t sa, const uint64_t *b, size_t sb, uint64_t *out)
{
    if (L < 1 || L > c->L0 || sa < 1 || sb < 1)
        return -1;
    const size_t N = c->N, sd = sa + sb - 1;
    /* dest[I] = sum over (i + j = I) of a_i o b_j, accumulated with add_poly_coeffmod into a
     * zeroed buffer (:491-520); i ascending, j descending; index window of :497-503. */
    for (size_t I = 0; I < sd; I++)
    {
        for (size_t l = 0; l < L; l++)
        {
            const uint64_t q = c->q[l];
            const uint64_t *cr = c->tab[l].cr;
            uint64_t *dst = out + (I * L + l) * N;
            memset(dst, 0, N * 8);
            size_t i_last = I < sa - 1 ? I : sa - 1, j_first = I < sb - 1 ? I : sb - 1, i_first = I - j_first;
            for (size_t i = i_first; i <= i_last; i++)
            {
                size_t j = I - i;
                const uint64_t *x = a + (i * L + l) * N, *y = b + (j * L + l) * N;
                for (size_t d = 0; d < N; d++)
                    dst[d] = so_add_uint_mod(so_multiply_uint_mod(x[d], y[d], q, cr), dst[d], q);
            }
        }
    }
    return 0;
}

int so_ckks_multiply(const so_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out)
{
    return so_ckks_multiply_sizes(c, L, a, 2, b, 2, out);
}

/* ======================================================================================== */
/* Hybrid key switching  (evaluator.cpp:2259-2368, multi_special_primes.cpp)                */
/* ======================================================================================== */

/* multi_special_primes.cpp:13-19 (MulModShoup::operator()) */
static inline uint64_t mulmod_shoup(uint64_t x, uint64_t cnst, uint64_t cnst_s, uint64_t p)
{
    uint64_t t = x * cnst - mulhi(x, cnst_s) * p;
    return t - (t >= p ? p : 0);
}

/* key-modulus index of extended limb kappa (evaluator.cpp:2317,2343) */
static inline size_t rns_idx_of(const so_ctx *c, size_t L, size_t kappa)
{
    return kappa < L ? kappa : c->L0 + (kappa - L);
}

/* multi_special_primes.cpp:80-148: bundle limbs src[0..nb) (moduli q[idx0+i]) -> one limb mod q[dst_idx] */
static void modup_to_single(const so_ctx *c, const uint64_t *src, size_t idx0, size_t nb, size_t dst_idx, uint64_t *dst)
{
    const size_t N = c->N;
    const uint64_t qd = c->q[dst_idx];
    const uint64_t *crd = c->tab[dst_idx].cr;
    if (nb == 1)
    {
        if (c->q[idx0] <= qd) /* :100-101 */
            memcpy(dst, src, N * 8);
        else /* :103-107 */
            for (size_t d = 0; d < N; d++)
                dst[d] = so_barrett_reduce_63(src[d], qd, crd);
        return;
    }
    uint64_t inv_punch[64], inv_punch_s[64], punch[64];
    for (size_t i = 0; i < nb; i++) /* :113-126 */
    {
        const uint64_t qi = c->q[idx0 + i];
        const uint64_t *cri = c->tab[idx0 + i].cr;
        uint64_t inv_prod = 1, prod = 1;
        for (size_t j = 0; j < nb; j++)
        {
            if (j == i)
                continue;
            prod = so_multiply_uint_mod(prod, c->q[idx0 + j], qd, crd);
            inv_prod = so_multiply_uint_mod(inv_prod, c->q[idx0 + j], qi, cri);
        }
        punch[i] = prod;
        so_try_invert_uint_mod(inv_prod, qi, &inv_punch[i]);
        inv_punch_s[i] = so_shoupify(inv_punch[i], qi);
    }
    for (size_t d = 0; d < N; d++) /* :128-146 */
    {
        u128 acc = 0;
        for (size_t i = 0; i < nb; i++)
            acc += (u128)mulmod_shoup(src[i * N + d], inv_punch[i], inv_punch_s[i], c->q[idx0 + i]) * punch[i];
        dst[d] = so_barrett_reduce_128((uint64_t)acc, (uint64_t)(acc >> 64), qd, crd);
    }
}

/* multi_special_primes.cpp:237-304: ext [L+k][N] (special part in coefficient form) -> first L limbs */
static void rescale_special(const so_ctx *c, size_t L, uint64_t *ext, int is_ckks)
{
    const size_t N = c->N, k = c->k, sp0 = c->L0;
    uint64_t inv_hat[64], inv_hat_s[64];
    for (size_t i = 0; i < k; i++) /* inv_puncture_products_of_special_rns, :211-234 */
    {
        const uint64_t pi = c->q[sp0 + i];
        const uint64_t *cri = c->tab[sp0 + i].cr;
        uint64_t prod = 1;
        for (size_t j = 0; j < k; j++)
            if (j != i)
                prod = so_multiply_uint_mod(prod, c->q[sp0 + j], pi, cri);
        so_try_invert_uint_mod(prod, pi, &inv_hat[i]);
        inv_hat_s[i] = so_shoupify(inv_hat[i], pi);
    }
    uint64_t *temp = (uint64_t *)malloc(N * 8);
    for (size_t i = 0; i < L; i++)
    {
        const uint64_t qi = c->q[i];
        const uint64_t *cri = c->tab[i].cr;
        uint64_t neg_hat[64]; /* neg_puncture_products_of_special_rns, :186-209 */
        for (size_t a = 0; a < k; a++)
        {
            uint64_t prod = 1;
            for (size_t j = 0; j < k; j++)
                if (j != a)
                    prod = so_multiply_uint_mod(prod, c->q[sp0 + j], qi, cri);
            neg_hat[a] = so_negate_uint_mod(prod, qi);
        }
        for (size_t d = 0; d < N; d++) /* steps 1+2, :253-282 */
        {
            u128 acc = 0;
            for (size_t j = 0; j < k; j++)
            {
                uint64_t x = ext[(L + j) * N + d];
                if (k > 1)
                    acc += (u128)mulmod_shoup(x, inv_hat[j], inv_hat_s[j], c->q[sp0 + j]) * neg_hat[j];
                else /* :268-275 */
                    acc += so_negate_uint_mod(so_barrett_reduce_63(x, c->q[sp0], c->tab[sp0].cr), c->q[sp0]);
            }
            temp[d] = so_barrett_reduce_128((uint64_t)acc, (uint64_t)(acc >> 64), qi, cri);
        }
        if (is_ckks) /* step 3, :285-289 */
            so_ntt_fwd_lazy(temp, &c->tab[i]);
        else
            so_ntt_inv_lazy(ext + i * N, &c->tab[i]);
        uint64_t P = 1, invP;              /* step 4, :291-302 */
        for (size_t j = 0; j < k; j++)
            P = so_multiply_uint_mod(P, c->q[sp0 + j], qi, cri);
        so_try_invert_uint_mod(P, qi, &invP);
        uint64_t invP_s = so_shoupify(invP, qi);
        uint64_t *ci = ext + i * N;
        for (size_t d = 0; d < N; d++)
            ci[d] = mulmod_shoup(ci[d] + temp[d], invP, invP_s, qi);
    }
    free(temp);
}

static int switch_key_scheme(const so_ctx *c, size_t L, uint64_t *enc, const uint64_t *target, const uint64_t *key, int is_ckks)
{
    if (L < 1 || L > c->L0)
        return -1;
    const size_t N = c->N, k = c->k, K = c->K, E = L + k;
    const size_t beta = (L + k - 1) / k;
    u128 *acc[2];
    acc[0] = (u128 *)calloc(E * N, sizeof(u128));
    acc[1] = (u128 *)calloc(E * N, sizeof(u128));
    uint64_t *ext = (uint64_t *)malloc(E * N * 8);
    for (size_t j = 0; j < beta; j++)
    {
        const size_t r0 = j * k, r1 = (r0 + k < L) ? r0 + k : L;
        /* step 1 (:2302-2307): strict inverse NTT of the bundle's limbs (CKKS); BFV targets are
           already in coefficient form */
        for (size_t i = r0; i < r1; i++)
        {
            memcpy(ext + i * N, target + i * N, N * 8);
            if (is_ckks)
                so_ntt_inv(ext + i * N, &c->tab[i]);
        }
        /* step 2 (:2310, modup_rns multi_special_primes.cpp:151-185) */
        for (size_t kap = 0; kap < E; kap++)
            if (kap < r0 || kap >= r1)
                modup_to_single(c, ext + r0 * N, r0, r1 - r0, rns_idx_of(c, L, kap), ext + kap * N);
        /* step 3 (:2315-2335) */
        for (size_t kap = 0; kap < E; kap++)
        {
            const size_t ridx = rns_idx_of(c, L, kap);
            const uint64_t *op;
            if (kap >= r0 && kap < r1)
                op = target + kap * N;
            else
            {
                so_ntt_fwd_lazy(ext + kap * N, &c->tab[ridx]);
                op = ext + kap * N;
            }
            for (size_t l = 0; l < 2; l++)
            {
                const uint64_t *kp = key + ((j * 2 + l) * K + ridx) * N;
                u128 *ap = acc[l] + kap * N;
                for (size_t d = 0; d < N; d++)
                    ap[d] += (u128)op[d] * kp[d];
            }
        }
    }
    for (size_t l = 0; l < 2; l++)
    {
        /* step 4 (:2341-2357) */
        for (size_t kap = 0; kap < E; kap++)
        {
            const size_t ridx = rns_idx_of(c, L, kap);
            for (size_t d = 0; d < N; d++)
            {
                u128 z = acc[l][kap * N + d];
                ext[kap * N + d] = so_barrett_reduce_128((uint64_t)z, (uint64_t)(z >> 64), c->q[ridx], c->tab[ridx].cr);
            }
            if (kap >= L)
                so_ntt_inv_lazy(ext + kap * N, &c->tab[ridx]);
        }
        /* step 5 (:2360-2366) */
        rescale_special(c, L, ext, is_ckks);
        uint64_t *e = enc + l * L * N;
        for (size_t i = 0; i < L; i++)
            for (size_t d = 0; d < N; d++)
                e[i * N + d] = so_add_uint_mod(ext[i * N + d], e[i * N + d], c->q[i]);
    }
    free(acc[0]);
    free(acc[1]);
    free(ext);
    return 0;
}

int so_switch_key(const so_ctx *c, size_t L, uint64_t *enc, const uint64_t *target, const uint64_t *key)
{
    return switch_key_scheme(c, L, enc, target, key, 1);
}

/* BFV branch of the same function: is_ckks == false at :2305, :2320 (the in-bundle slot multiplies
   the coefficient-form target as it is) and multi_special_primes.cpp:287-289. */
int so_switch_key_bfv(const so_ctx *c, size_t L, uint64_t *enc, const uint64_t *target, const uint64_t *key)
{
    return switch_key_scheme(c, L, enc, target, key, 0);
}

int so_relinearize_bfv(const so_ctx *c, size_t L, const uint64_t *in, const uint64_t *key, uint64_t *out)
{
    const size_t N = c->N;
    uint64_t *tmp = (uint64_t *)malloc(3 * L * N * 8);
    memcpy(tmp, in, 3 * L * N * 8);
    int rc = so_switch_key_bfv(c, L, tmp, tmp + 2 * L * N, key);
    if (!rc)
        memcpy(out, tmp, 2 * L * N * 8);
    free(tmp);
    return rc;
}

/* GaloisTool::apply_galois, coefficient form (util/galois.cpp:144-186) */
void so_apply_galois_coeff(int logn, uint32_t elt, uint64_t q, const uint64_t *in, uint64_t *out)
{
    const uint64_t n = (uint64_t)1 << logn;
    uint64_t index_raw = 0;
    for (uint64_t i = 0; i < n; i++, index_raw += elt)
    {
        uint64_t v = in[i];
        if ((index_raw >> logn) & 1)
            v = v ? q - v : 0;
        out[index_raw & (n - 1)] = v;
    }
}

/* evaluator.cpp:1841-1943, BFV branch (:1893-1904) */
int so_apply_galois_bfv(const so_ctx *c, size_t L, uint32_t elt, const uint64_t *in, const uint64_t *key, uint64_t *out)
{
    const size_t N = c->N;
    uint64_t *temp = (uint64_t *)malloc(L * N * 8);
    for (size_t i = 0; i < L; i++)
    {
        so_apply_galois_coeff(c->logn, elt, c->q[i], in + i * N, out + i * N);
        so_apply_galois_coeff(c->logn, elt, c->q[i], in + (L + i) * N, temp + i * N);
    }
    memset(out + L * N, 0, L * N * 8);
    int rc = so_switch_key_bfv(c, L, out, temp, key);
    free(temp);
    return rc;
}

/* BFV mod_switch_to_next: mod_switch_scale_to_next (evaluator.cpp:829-892) ->
 * RNSTool::divide_and_round_q_last_inplace (util/rns.cpp:731-775); coefficient form.
 * in [size][L][N] -> out [size][L-1][N] */
int so_bfv_mod_switch(const so_ctx *c, size_t L, size_t size, const uint64_t *in, uint64_t *out)
{
    if (L < 2 || L > c->L0)
        return -1;
    const size_t N = c->N;
    const uint64_t qL = c->q[L - 1], half = qL >> 1;
    uint64_t *last = (uint64_t *)malloc(N * 8);
    for (size_t p = 0; p < size; p++)
    {
        const uint64_t *src = in + p * L * N;
        uint64_t *dst = out + p * (L - 1) * N;
        for (size_t d = 0; d < N; d++) /* :747-751 */
            last[d] = so_barrett_reduce_63(src[(L - 1) * N + d] + half, qL, c->tab[L - 1].cr);
        for (size_t i = 0; i + 1 < L; i++)
        {
            const uint64_t qi = c->q[i];
            const uint64_t *cri = c->tab[i].cr;
            const uint64_t half_mod = so_barrett_reduce_63(half, qi, cri); /* :760 */
            uint64_t inv;
            so_try_invert_uint_mod(qL, qi, &inv); /* rns.cpp:719-728 */
            for (size_t d = 0; d < N; d++)
            {
                uint64_t t = so_barrett_reduce_63(last[d], qi, cri); /* modulo_poly_coeffs_63, :758 */
                t = so_sub_uint_mod(t, half_mod, qi);                  /* :761-764 */
                dst[i * N + d] = so_multiply_uint_mod(so_sub_uint_mod(src[i * N + d], t, qi), inv, qi, cri); /* :766-772 */
            }
        }
    }
    free(last);
    return 0;
}

/* evaluator.cpp:772-827 (size 3 -> 2) */
int so_relinearize(const so_ctx *c, size_t L, const uint64_t *in, const uint64_t *key, uint64_t *out)
{
    const size_t N = c->N;
    uint64_t *tmp = (uint64_t *)malloc(3 * L * N * 8);
    memcpy(tmp, in, 3 * L * N * 8);
    int rc = so_switch_key(c, L, tmp, tmp + 2 * L * N, key);
    if (!rc)
        memcpy(out, tmp, 2 * L * N * 8);
    free(tmp);
    return rc;
}

/* relinearize_internal (evaluator.cpp:772-827) from any size down to 2.  keys = the relinearization keys
 * back to back, key_words each, key p (RelinKeys::get_index(p + 2)) at keys + p * key_words.  The loop at
 * :809-815 switches `*encrypted_iter` with key get_index(size - 1 - I); encrypted_iter was set to the LAST
 * component at :806-807 and is not moved inside the loop, so every step's target is component size - 1. */
int so_relinearize_sizes(const so_ctx *c, size_t L, size_t size, const uint64_t *in, const uint64_t *keys, size_t key_words, int is_ckks, uint64_t *out)
{
    const size_t N = c->N;
    if (size < 2)
        return -1;
    uint64_t *tmp = (uint64_t *)malloc(size * L * N * 8);
    memcpy(tmp, in, size * L * N * 8);
    int rc = 0;
    for (size_t I = 0; !rc && I + 2 < size; I++)
    {
        const uint64_t *key = keys + (size - 1 - I - 2) * key_words;
        rc = is_ckks ? so_switch_key(c, L, tmp, tmp + (size - 1) * L * N, key) : so_switch_key_bfv(c, L, tmp, tmp + (size - 1) * L * N, key);
    }
    if (!rc)
        memcpy(out, tmp, 2 * L * N * 8);
    free(tmp);
    return rc;
}

/* evaluator.cpp:1090-1126,829-892 -> util/rns.cpp:777-851 */
int so_rescale(const so_ctx *c, size_t L, size_t size, const uint64_t *in, uint64_t *out)
{
    if (L < 2 || L > c->L0)
        return -1;
    const size_t N = c->N;
    const uint64_t qL = c->q[L - 1];
    const uint64_t *crL = c->tab[L - 1].cr;
    const uint64_t half = qL >> 1;
    uint64_t *last = (uint64_t *)malloc(N * 8), *temp = (uint64_t *)malloc(N * 8);
    for (size_t p = 0; p < size; p++)
    {
        const uint64_t *src = in + p * L * N;
        uint64_t *dst = out + p * (L - 1) * N;
        memcpy(last, src + (L - 1) * N, N * 8);
        so_ntt_inv(last, &c->tab[L - 1]); /* :798 */
        for (size_t d = 0; d < N; d++)    /* :801-806 */
            last[d] = so_barrett_reduce_63(last[d] + half, qL, crL);
        for (size_t i = 0; i + 1 < L; i++)
        {
            const uint64_t qi = c->q[i];
            const uint64_t *cri = c->tab[i].cr;
            if (qi < qL) /* :814-821 */
                for (size_t d = 0; d < N; d++)
                    temp[d] = so_barrett_reduce_63(last[d], qi, cri);
            else
                memcpy(temp, last, N * 8);
            const uint64_t neg_half = qi - so_barrett_reduce_63(half, qi, cri); /* :824-827 */
            for (size_t d = 0; d < N; d++)
                temp[d] += neg_half;
            so_ntt_fwd_lazy(temp, &c->tab[i]); /* :832 */
            const uint64_t qi_lazy = qi << 2;
            uint64_t inv;
            so_try_invert_uint_mod(qL, qi, &inv); /* rns.cpp:719-728 */
            for (size_t d = 0; d < N; d++)       /* :843-848 */
                dst[i * N + d] = so_multiply_uint_mod(src[i * N + d] + qi_lazy - temp[d], inv, qi, cri);
        }
    }
    free(last);
    free(temp);
    return 0;
}

/* evaluator.cpp:1841-1943 (CKKS branch) */
int so_apply_galois(const so_ctx *c, size_t L, uint32_t elt, const uint64_t *in, const uint64_t *key, uint64_t *out)
{
    const size_t N = c->N;
    uint64_t *temp = (uint64_t *)malloc(L * N * 8);
    for (size_t i = 0; i < L; i++)
    {
        so_apply_galois_ntt(c->logn, elt, in + i * N, out + i * N);        /* :1913-1917 */
        so_apply_galois_ntt(c->logn, elt, in + (L + i) * N, temp + i * N); /* :1920 */
    }
    memset(out + L * N, 0, L * N * 8); /* :1928 */
    int rc = so_switch_key(c, L, out, temp, key);
    free(temp);
    return rc;
}

int so_mul_relin_rescale(const so_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, const uint64_t *key, uint64_t *out)
{
    const size_t N = c->N;
    uint64_t *t3 = (uint64_t *)malloc(3 * L * N * 8), *t2 = (uint64_t *)malloc(2 * L * N * 8);
    int rc = so_ckks_multiply(c, L, a, b, t3);
    if (!rc)
        rc = so_relinearize(c, L, t3, key, t2);
    if (!rc)
        rc = so_rescale(c, L, 2, t2, out);
    free(t3);
    free(t2);
    return rc;
}

/* ======================================================================================== */
/* Switching-key generation                                            */
/* ======================================================================================== */
/* KeyGenerator::generate_one_kswitch_key (keygenerator.cpp:325-369) over util::encrypt_zero_symmetric
 * (util/rlwe.cpp:204-304) with the SAMPLED buffers given: a = [beta][K][N] uniform residues in NTT form (the
 * reference's c1, :232-241), e = [beta][N] the noise as signed integers (sample_poly_normal, rlwe.cpp:60-84,
 * writes e or q_i - |e| into every limb).  sk, new_key = [K][N] NTT form at key level.  out = [beta][2][K][N]. */
int so_kswitch_keygen(const so_ctx *c, const uint64_t *sk, const uint64_t *new_key, const uint64_t *a, const int64_t *e, uint64_t *out)
{
    const size_t N = c->N, K = c->K, KN = K * N, beta = (c->L0 + c->k - 1) / c->k;
    uint64_t *noise = (uint64_t *)malloc(N * 8);
    if (!noise)
        return -1;
    for (size_t j = 0; j < beta; j++)
    {
        uint64_t *c0 = out + j * 2 * KN, *c1 = c0 + KN;
        memcpy(c1, a + j * KN, KN * 8);
        for (size_t i = 0; i < K; i++)
        {
            const uint64_t q = c->q[i];
            const uint64_t *cr = c->tab[i].cr;
            for (size_t d = 0; d < N; d++)
            {
                const int64_t v = e[j * N + d];
                noise[d] = v >= 0 ? (uint64_t)v : q - (uint64_t)(-v);
            }
            so_ntt_fwd(noise, &c->tab[i]); /* rlwe.cpp:263-266 */
            for (size_t d = 0; d < N; d++)
            {
                uint64_t t = so_multiply_uint_mod(sk[i * N + d], c1[i * N + d], q, cr); /* :259-262 */
                t = so_add_uint_mod(noise[d], t, q);                                    /* :271-273 */
                c0[i * N + d] = so_negate_uint_mod(t, q);                               /* :274 */
            }
        }
        /* keygenerator.cpp:346-366 */
        const size_t rns0 = j * c->k, rns1 = rns0 + c->k < c->L0 ? rns0 + c->k : c->L0;
        for (size_t rns = rns0; rns < rns1; rns++)
        {
            const uint64_t q = c->q[rns];
            const uint64_t *cr = c->tab[rns].cr;
            uint64_t factor = 1;
            for (size_t s = 0; s < c->k; s++)
                factor = so_multiply_uint_mod(factor, c->q[c->L0 + s], q, cr);
            for (size_t d = 0; d < N; d++)
                c0[rns * N + d] = so_add_uint_mod(c0[rns * N + d], so_multiply_uint_mod(new_key[rns * N + d], factor, q, cr), q);
        }
    }
    free(noise);
    return 0;
}
