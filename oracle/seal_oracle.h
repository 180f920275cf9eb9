/* TEST INFRASTRUCTURE -- CPU oracle for the Gemini-SEAL RNS hot path.
 *
 * A plain-C restatement of the reference's algorithm (file:line cited at every function in
 * seal_oracle.c).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load it; the product (gemini_seal_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle_pins.py checks this file against (a) the known-answer
 * vectors transcribed from the reference's own unit tests and (b) the unmodified
 * reference compiled into oracle/_ref/ (Makefile target `ref`) on seeded inputs; the committed
 * fixtures under tests/golden/ were produced by the reference itself (tests/golden/make_golden.py).
 *
 * All polynomial buffers are limb-major uint64_t: limb i of a polynomial at ptr + i*N
 * (ciphertext.h:357-381); a ciphertext is [size][L][N]; one switching key is
 * [bundle][component 0/1][K][N] in NTT form (kswitchkeys.h:343, keygenerator.cpp:348-351).
 */
#ifndef SEAL_ORACLE_H
#define SEAL_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct so_tables
{
    int logn;
    uint64_t q;
    uint64_t cr[3];          /* Modulus::const_ratio: floor(2^128/q) lo,hi ; 2^128 mod q */
    uint64_t root;           /* minimal primitive 2N-th root */
    uint64_t inv_n, inv_n_s; /* n^-1 mod q and its Shoup scaling */
    uint64_t rdp;            /* floor(2^64/q) (reduce_precomp) */
    uint64_t *w, *ws;        /* root_powers / scaled_root_powers, N words each */
    uint64_t *v, *vs;        /* inv_root_powers (re-ordered, last *= n^-1) / scaled */
} so_tables;

typedef struct so_ctx
{
    int logn;
    size_t N, K, k, L0;      /* key-level limbs, special primes, first-level limbs = K-k */
    uint64_t *q;             /* K moduli; the special primes are q[L0..K) */
    so_tables *tab;          /* K tables */
} so_ctx;

/* ---- word arithmetic ------------------------------------------------------------------- */
void so_const_ratio(uint64_t q, uint64_t cr[3]);
uint64_t so_shoupify(uint64_t x, uint64_t q);
uint64_t so_barrett_reduce_128(uint64_t lo, uint64_t hi, uint64_t q, const uint64_t cr[3]);
uint64_t so_barrett_reduce_63(uint64_t x, uint64_t q, const uint64_t cr[3]);
uint64_t so_multiply_uint_mod(uint64_t a, uint64_t b, uint64_t q, const uint64_t cr[3]);
uint64_t so_multiply_add_uint_mod(uint64_t a, uint64_t b, uint64_t c, uint64_t q, const uint64_t cr[3]);
uint64_t so_add_uint_mod(uint64_t a, uint64_t b, uint64_t q);
uint64_t so_sub_uint_mod(uint64_t a, uint64_t b, uint64_t q);
uint64_t so_negate_uint_mod(uint64_t a, uint64_t q);
int so_try_invert_uint_mod(uint64_t a, uint64_t q, uint64_t *out);
uint64_t so_exponentiate_uint_mod(uint64_t a, uint64_t e, uint64_t q);
uint32_t so_reverse_bits32(uint32_t x, int bits);

/* ---- tables + NTT ---------------------------------------------------------------------- */
int so_tables_init(so_tables *t, int logn, uint64_t q); /* 0 ok, -1 invalid modulus */
void so_tables_free(so_tables *t);
void so_ntt_fwd_lazy(uint64_t *a, const so_tables *t);
void so_ntt_fwd(uint64_t *a, const so_tables *t);
void so_ntt_inv_lazy(uint64_t *a, const so_tables *t);
void so_ntt_inv(uint64_t *a, const so_tables *t);
/* convenience for ctypes: kind 0 fwd lazy, 1 fwd strict, 2 inv lazy, 3 inv strict */
int so_ntt(int logn, uint64_t q, uint64_t *data, size_t n_polys, int kind);

/* ---- elementwise ----------------------------------------------------------------------- */
/* op: 0 dyadic, 1 add, 2 sub, 3 negate(a), 4 a*scalar */
int so_poly_op(int op, uint64_t q, const uint64_t *a, const uint64_t *b, uint64_t scalar, size_t n, uint64_t *out);

/* ---- Galois ---------------------------------------------------------------------------- */
void so_galois_table_ntt(int logn, uint32_t elt, uint32_t *table);
uint32_t so_elt_from_step(int logn, int step); /* 0 on invalid step */
void so_apply_galois_ntt(int logn, uint32_t elt, const uint64_t *in, uint64_t *out);

/* ---- context + CKKS evaluator ops ------------------------------------------------------- */
so_ctx *so_ctx_create(int logn, const uint64_t *moduli, size_t K, size_t k_special);
void so_ctx_destroy(so_ctx *c);
/* a,b [2][L][N] -> out [3][L][N] */
int so_ckks_multiply(const so_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, uint64_t *out);
int so_ckks_multiply_sizes(const so_ctx *c, size_t L, const uint64_t *a, size_t sa, const uint64_t *b, size_t sb, uint64_t *out);
/* enc [>=2][L][N] (components 0,1 updated in place); target [L][N]; key [beta][2][K][N] */
int so_switch_key(const so_ctx *c, size_t L, uint64_t *enc, const uint64_t *target, const uint64_t *key);
/* in [3][L][N] -> out [2][L][N] */
int so_relinearize(const so_ctx *c, size_t L, const uint64_t *in, const uint64_t *key, uint64_t *out);
/* row f2: generate_one_kswitch_key given the sampled a (NTT form) and e (signed noise) */
int so_kswitch_keygen(const so_ctx *c, const uint64_t *sk, const uint64_t *new_key, const uint64_t *a, const int64_t *e, uint64_t *out);
int so_relinearize_sizes(const so_ctx *c, size_t L, size_t size, const uint64_t *in, const uint64_t *keys, size_t key_words, int is_ckks, uint64_t *out);
/* in [size][L][N] -> out [size][L-1][N] */
int so_rescale(const so_ctx *c, size_t L, size_t size, const uint64_t *in, uint64_t *out);
/* in/out [2][L][N] */
int so_apply_galois(const so_ctx *c, size_t L, uint32_t elt, const uint64_t *in, const uint64_t *key, uint64_t *out);
/* ---- BFV branches (coefficient-form ciphertexts) ----------------------- */
int so_switch_key_bfv(const so_ctx *c, size_t L, uint64_t *enc, const uint64_t *target, const uint64_t *key);
int so_relinearize_bfv(const so_ctx *c, size_t L, const uint64_t *in, const uint64_t *key, uint64_t *out);
/* in [size][L][N] -> out [size][L-1][N] (util/rns.cpp:731-775) */
int so_bfv_mod_switch(const so_ctx *c, size_t L, size_t size, const uint64_t *in, uint64_t *out);
void so_apply_galois_coeff(int logn, uint32_t elt, uint64_t q, const uint64_t *in, uint64_t *out);
int so_apply_galois_bfv(const so_ctx *c, size_t L, uint32_t elt, const uint64_t *in, const uint64_t *key, uint64_t *out);
/* Evaluator::bfv_multiply (BEHZ, evaluator.cpp:274-445; seal_oracle_bfv.c): a, b [2][L][N] in
 * coefficient form over q[0..L) -> out [3][L][N] */
int so_bfv_multiply(int logn, const uint64_t *q, size_t L, uint64_t plain_modulus, const uint64_t *a, const uint64_t *b, uint64_t *out);
int so_bfv_multiply_sizes(int logn, const uint64_t *q, size_t L, uint64_t plain_modulus, const uint64_t *a, size_t sa, const uint64_t *b, size_t sb, uint64_t *out);
/* out[0] = |B|, out[1..] = the |B|+1 moduli of Bsk (rns.cpp:585-603) */
int so_bfv_aux_base(int logn, const uint64_t *q, size_t L, uint64_t plain_modulus, uint64_t *out);
/* ---- Decryptor (seal_oracle_bfv.c) -------------------------------------- */
/* dot_product_ct_sk_array (decryptor.cpp:218-265): ct [size][L][N], sk [K][N] (NTT form, key level) ->
 * out [L][N] = c0 + c1 s + ... ; this is also Decryptor::ckks_decrypt's plaintext (ntt_form = 1) */
int so_dot_product_ct_sk(const so_ctx *c, size_t L, size_t size, const uint64_t *ct, const uint64_t *sk, int ntt_form, uint64_t *out);
/* Decryptor::bfv_decrypt (decryptor.cpp:76-113, util/rns.cpp:1070-1126) -> out [N] coefficients mod t */
int so_bfv_decrypt(const so_ctx *c, size_t L, size_t size, uint64_t plain_modulus, const uint64_t *ct, const uint64_t *sk, uint64_t *out);
/* a,b [2][L][N] -> out [2][L-1][N] */
int so_mul_relin_rescale(const so_ctx *c, size_t L, const uint64_t *a, const uint64_t *b, const uint64_t *key, uint64_t *out);

#ifdef __cplusplus
}
#endif
#endif
