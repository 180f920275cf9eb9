/* gseal_b200.h -- C ABI of the B200-native RNS polynomial-arithmetic hot path for Gemini-SEAL.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ or torch types.  Each entry
 * point names the reference interface it replaces (paths relative to
 * /root/reference/native/src/seal/).  The C++ shim (shim/gsb_seal_shim.{h,cpp}) and the Python host
 * mirror (gemini_seal_b200/evaluator.py) are thin callers of exactly these functions.
 *
 * Conventions (modelled on the reference's own C layer, c/defines.h:27-56, c/evaluator.cpp:31-58):
 *   - every function returns an int status: GSB_OK or a negative GSB_E_* class; no exception ever
 *     crosses the ABI; gsb_last_error() returns the message of the calling thread's last failure;
 *   - the caller owns every buffer it passes; the context owns device copies of tables and keys;
 *   - polynomial data is limb-major uint64_t exactly as in seal::Ciphertext (ciphertext.h:357-381):
 *     a batch of B ciphertexts of `size` polynomials over L limbs is [B][size][L][N];
 *   - *_dev functions take DEVICE pointers and enqueue on `stream` (a cudaStream_t, NULL = default
 *     stream) without synchronising; *_host functions take HOST pointers, copy in, run, copy out
 *     and synchronise before returning;
 *   - a context is immutable after creation/key upload and may be shared by host threads as long
 *     as each thread uses its own stream AND its own workspace (gsb_ws_*).  Calls that pass
 *     ws = NULL share the context's default workspace: they must all be enqueued on one stream
 *     (or be serialised by the caller); the *_host entry points take a per-context lock.
 *   - every entry point makes the context's device the calling thread's current CUDA device (and leaves it
 *     so); gsb_ctx_create restores the creating thread's previous device before it returns.
 *   - ciphertext and key words must be canonical residues (word < its limb's modulus), which is what
 *     the reference's is_buffer_valid (valcheck.cpp) demands before any Evaluator call; the shim and
 *     gsb_key_upload check it, the device entry points do not scan their inputs.  On such inputs
 *     every output word equals the reference's.
 *   - there is no CPU fallback: every compute entry point fails with GSB_E_CUDA if no CUDA
 *     device is usable.
 */
#ifndef GSEAL_B200_H
#define GSEAL_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GSB_OK 0
#define GSB_E_INVALID_ARG (-1) /* std::invalid_argument in the reference (evaluator.cpp:238-249 ...) */
#define GSB_E_LOGIC (-2)       /* std::logic_error in the reference                                 */
#define GSB_E_RANGE (-3)       /* std::out_of_range                                                 */
#define GSB_E_CUDA (-4)        /* CUDA runtime failure / no device (std::runtime_error in the shim) */
#define GSB_E_NOMEM (-5)

typedef struct gsb_ctx gsb_ctx;
typedef struct gsb_ws gsb_ws;

/* gsb_ntt_* `mode` */
#define GSB_NTT_STRICT 0 /* ntt_negacyclic_harvey / inverse_ntt_negacyclic_harvey  (util/ntt.h:225,318) */
#define GSB_NTT_LAZY 1   /* ..._lazy: forward in [0,4q), inverse in [0,2q)         (util/ntt.cpp:292,345) */

/* gsb_poly_op `op` (util/polyarithsmallmod.{h,cpp}) */
#define GSB_OP_DYADIC 0     /* dyadic_product_coeffmod        .cpp:63-117  */
#define GSB_OP_ADD 1        /* add_poly_coeffmod              .h:261-300   */
#define GSB_OP_SUB 2        /* sub_poly_coeffmod              .h:366-405   */
#define GSB_OP_NEGATE 3     /* negate_poly_coeffmod           .h:176-206   */
#define GSB_OP_MUL_SCALAR 4 /* multiply_poly_scalar_coeffmod  .cpp:15-61 (one scalar per limb) */

/* key kinds for gsb_key_upload */
#define GSB_KEY_RELIN 0  /* RelinKeys::get_index(2) -> key for s^2     (relinkeys.h:61)   */
#define GSB_KEY_GALOIS 1 /* GaloisKeys::get_index(elt)                 (galoiskeys.h:52)  */

const char *gsb_last_error(void);
int gsb_version(void);
/* number of kernel launches this library has issued in this process (bench.py "gpu_launches") */
unsigned long long gsb_launch_count(void);
/* Optional per-launch timing for bench.py's roofline: while enabled, every kernel launch is
 * bracketed by CUDA events on its own stream.  gsb_prof_enable(1) clears earlier records.
 * gsb_prof_report synchronises the device and writes a JSON object
 * {"<kernel label>": {"launches": n, "ms": total}, ...} (launch order) into buf; returns the bytes
 * needed including the terminator (call with buf = NULL to size), or -1 on a CUDA error. */
int gsb_prof_enable(int on);
long gsb_prof_report(char *buf, size_t cap);

/* ---- context: replaces SEALContext::ContextData tables (context.cpp:135-420) --------------------
 * moduli[0..K) is the key-level coeff_modulus; its last k_special entries are the special primes
 * (encryptionparams.h:205-214).  Ciphertext levels have L = K-k_special, K-k_special-1, ... limbs
 * over moduli[0..L).  Tables are built on the host exactly as NTTTables::initialize does
 * (util/ntt.cpp:37-99) and uploaded once.  N = 2^logn, 10 <= logn <= 16. */
int gsb_ctx_create(int device, int logn, const uint64_t *moduli, size_t K, size_t k_special, gsb_ctx **out);
void gsb_ctx_destroy(gsb_ctx *ctx);
/* Introspection (tests compare these with the reference's NTTTables accessors, util/ntt.h:46-60,
 * 84-92,116).  Each array argument may be NULL.  scalars[0..6] = q, const_ratio[0], const_ratio[1],
 * root, inv_degree_modulo, scaled_inv_degree, reduce_precomp. */
int gsb_ctx_get_tables(
    const gsb_ctx *ctx, size_t mod_index, uint64_t *root_powers, uint64_t *scaled_root_powers,
    uint64_t *inv_root_powers, uint64_t *scaled_inv_root_powers, uint64_t *scalars);
int gsb_ctx_info(const gsb_ctx *ctx, int *logn, size_t *K, size_t *k_special);

/* ---- workspaces (the device-side stand-in for the MemoryPoolHandle argument) -------------------- */
int gsb_ws_create(gsb_ctx *ctx, gsb_ws **out);
void gsb_ws_destroy(gsb_ws *ws);

/* ---- device memory helpers for callers without their own CUDA allocator ----------------------- */
int gsb_malloc(gsb_ctx *ctx, size_t bytes, void **out);
int gsb_free(gsb_ctx *ctx, void *ptr);
int gsb_malloc_host(size_t bytes, void **out); /* pinned */
int gsb_free_host(void *ptr);
int gsb_memcpy_h2d(gsb_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream);
int gsb_memcpy_d2h(gsb_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream);
int gsb_stream_sync(gsb_ctx *ctx, void *stream);

/* ---- NTT: util::ntt_negacyclic_harvey[_lazy] / inverse_... over RNSIter / PolyIter --------------
 * (util/ntt.h:189-368).  data = [n_polys][n_limbs][N]; limb y of every polynomial uses modulus
 * index limb_offset + y.  In place.  The LAZY transforms carry the reference's words for any 64-bit input (mod-2^64
 * wraps included).  The STRICT transforms return the reference's canonical words for inputs below 4q (forward) /
 * 2q (inverse), the ranges every transform of the reference and of this library produces. */
int gsb_ntt_forward_dev(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode, void *stream);
int gsb_ntt_inverse_dev(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode, void *stream);
int gsb_ntt_forward_host(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode);
int gsb_ntt_inverse_host(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode);

/* ---- elementwise kernels (util/polyarithsmallmod.*) -------------------------------------------
 * a, b, out = [n_polys][n_limbs][N]; b is ignored for NEGATE / MUL_SCALAR; scalars = n_limbs host
 * words (one per limb) for MUL_SCALAR, else NULL.  out may alias a or b. */
int gsb_poly_op_dev(
    gsb_ctx *ctx, int op, const uint64_t *a, const uint64_t *b, const uint64_t *scalars, uint64_t *out,
    size_t n_polys, size_t n_limbs, size_t limb_offset, void *stream);

int gsb_poly_op_host(
    gsb_ctx *ctx, int op, const uint64_t *a, const uint64_t *b, const uint64_t *scalars, uint64_t *out,
    size_t n_polys, size_t n_limbs, size_t limb_offset);

/* ---- switching keys: KSwitchKeys / RelinKeys / GaloisKeys (kswitchkeys.h:343, relinkeys.h:61,
 * galoiskeys.h:52-66, keygenerator.cpp:325-369) --------------------------------------------------
 * key_host = [n_bundles][2][K][N]: bundle j is `keys.data()[index][j].data()` of the reference (a
 * size-2 ciphertext over all K key-level limbs, NTT form).  n_bundles must be ceil((K-k)/k).
 * For kind GSB_KEY_RELIN, `galois_elt` is RelinKeys::get_index(key_power) (0 = the s^2 key that
 * relinearizes size 3 -> 2; relinkeys.h:61).  The context keeps a device copy until gsb_key_drop /
 * gsb_ctx_destroy.  Uploading also builds the Galois permutation table (util/galois.cpp:18-47). */
int gsb_key_upload(gsb_ctx *ctx, int kind, uint32_t galois_elt, const uint64_t *key_host, size_t n_bundles);
/* The same with one pointer per bundle (what the C++ shim passes: keys.data()[index][j].data().data()), so the
 * caller does not have to flatten a 120 MiB key first.  bundles[j] = [2][K][N]. */
int gsb_key_upload_bundles(gsb_ctx *ctx, int kind, uint32_t galois_elt, const uint64_t *const *bundles, size_t n_bundles);
int gsb_key_drop(gsb_ctx *ctx, int kind, uint32_t galois_elt);
/* GaloisTool::get_elt_from_step (util/galois.cpp:49-91); GSB_E_INVALID_ARG for |step| >= N/2 */
int gsb_galois_elt_from_step(const gsb_ctx *ctx, int step, uint32_t *elt);
/* GaloisTool::generate_table_ntt (util/galois.cpp:18-47): table_out[0..N) (host) */
int gsb_galois_table(const gsb_ctx *ctx, uint32_t galois_elt, uint32_t *table_out);

/* ---- Evaluator hot path, batched: every pointer is a batch of `batch` ciphertexts stored back to
 * back; L = number of limbs of the ciphertexts' level (2 <= L <= K-k for rescale, 1 <= L otherwise).
 * All ciphertexts are CKKS, NTT form.  ws = NULL uses the context's default workspace.
 * Every input word must be a canonical residue (< q_i), i.e. the buffers must pass the reference's
 * is_buffer_valid (valcheck.cpp), which the C++ shim checks as seal::Evaluator does; the *_dev entry
 * points do not re-scan device buffers, and their internal no-wrap bounds assume it. -------------- */

/* Evaluator::multiply_inplace -> ckks_multiply (evaluator.cpp:235-272,447-527) for two size-2
 * operands: a, b = [batch][2][L][N] -> out = [batch][3][L][N]. */
int gsb_ckks_multiply_dev(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L, void *stream);

/* Evaluator::switch_key_inplace (evaluator.cpp:2259-2368): enc = [batch][enc_size][L][N], its
 * components 0 and 1 are updated in place; target = one polynomial [L][N] per ciphertext, the b-th
 * at target + b*target_stride words (may point inside enc, e.g. component 2). */
int gsb_switch_key_dev(
    gsb_ctx *ctx, gsb_ws *ws, uint64_t *enc, size_t enc_size, const uint64_t *target, size_t target_stride,
    int key_kind, uint32_t galois_elt, size_t batch, size_t L, void *stream);

/* Evaluator::relinearize_inplace (evaluator.h:344, evaluator.cpp:772-827) of size-3 ciphertexts:
 * in = [batch][3][L][N] -> out = [batch][2][L][N]  (out must not overlap in). */
int gsb_relinearize_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream);

/* Evaluator::rescale_to_next (evaluator.cpp:1090-1126,829-892; util/rns.cpp:777-851):
 * in = [batch][size][L][N] -> out = [batch][size][L-1][N]  (the caller divides `scale` by q_{L-1}).
 * out must not overlap in: the two-pass plans use `out` as the buffer between their passes. */
int gsb_rescale_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream);

/* Evaluator::apply_galois_inplace (evaluator.cpp:1841-1943; rotate_vector = elt from step,
 * complex_conjugate = elt 2N-1): in, out = [batch][2][L][N]  (out must not overlap in). */
int gsb_apply_galois_dev(gsb_ctx *ctx, gsb_ws *ws, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream);

/* multiply_inplace + relinearize_inplace + rescale_to_next_inplace in one call (the unit of work of
 * the headline metric): a, b = [batch][2][L][N] -> out = [batch][2][L-1][N]. */
int gsb_mul_relin_rescale_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L, void *stream);

/* ---- BFV branches of the same functions (BASELINE config #1) -----------------
 * Ciphertexts are in coefficient form (is_ntt_form() == false).  These mirror what the reference
 * computes for scheme_type::BFV word for word, including the in-bundle slot of the inner product
 * reading the coefficient-form target as it is (evaluator.cpp:2319-2320 with is_ckks == false).
 * Evaluator::relinearize_inplace, BFV (evaluator.cpp:772-827 -> :2259-2368, multi_special_primes.cpp:
 * 287-289): in = [batch][3][L][N] -> out = [batch][2][L][N]. */
int gsb_bfv_relinearize_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream);
/* Evaluator::apply_galois_inplace, BFV branch (evaluator.cpp:1893-1904; GaloisTool::apply_galois,
 * util/galois.cpp:144-186; rotate_rows / rotate_columns, evaluator.h:1079-1176): in, out =
 * [batch][2][L][N]. */
int gsb_bfv_apply_galois_dev(gsb_ctx *ctx, gsb_ws *ws, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream);
/* Evaluator::multiply_inplace, BFV branch = bfv_multiply (BEHZ RNS multiplication, evaluator.cpp:
 * 274-445; RNSTool::{fastbconv_m_tilde, sm_mrq, fast_floor, fastbconv_sk}, util/rns.cpp:853-1066):
 * a, b = [batch][2][L][N] coefficient form -> out = [batch][3][L][N].  The auxiliary base Bsk and
 * its constants for (L, plain_modulus) are derived on first use exactly as RNSTool::initialize does
 * (util/rns.cpp:540-728) and cached in the context. */
int gsb_bfv_multiply_dev(gsb_ctx *ctx, gsb_ws *ws, uint64_t plain_modulus, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L, void *stream);
int gsb_bfv_multiply_host(gsb_ctx *ctx, uint64_t plain_modulus, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L);
/* The moduli of Bsk = B u {m_sk} for (L, plain_modulus) (RNSTool::base_Bsk, util/rns.cpp:585-603). */
int gsb_bfv_aux_base(gsb_ctx *ctx, size_t L, uint64_t plain_modulus, uint64_t *moduli_out, size_t cap, size_t *n_out);
/* Evaluator::mod_switch_to_next, BFV branch: mod_switch_scale_to_next (evaluator.cpp:829-892) ->
 * RNSTool::divide_and_round_q_last_inplace (util/rns.cpp:731-775): in = [batch][size][L][N] coefficient
 * form -> out = [batch][size][L-1][N]. */
int gsb_bfv_mod_switch_dev(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream);
int gsb_bfv_mod_switch_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L);
int gsb_bfv_relinearize_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t L);
int gsb_bfv_apply_galois_host(gsb_ctx *ctx, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L);

/* Host-buffer forms (what the C++ shim calls with seal::Ciphertext::data()): copy in, run, copy out,
 * synchronise.  Shapes as above. */
int gsb_ckks_multiply_host(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L);
int gsb_relinearize_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t L);
int gsb_rescale_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L);
int gsb_apply_galois_host(gsb_ctx *ctx, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L);
int gsb_mul_relin_rescale_host(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L);

/* ---- general ciphertext sizes ------------------------------------------------------------------------
 * Evaluator::multiply_inplace for operands of any sizes (dest_size = size_a + size_b - 1, evaluator.cpp:470-472;
 * index window :493-519): a = [batch][size_a][L][N], b = [batch][size_b][L][N] -> out =
 * [batch][size_a + size_b - 1][L][N]; sizes >= 2, size_a + size_b - 1 <= 16 (SEAL_CIPHERTEXT_SIZE_MAX).
 * CKKS = ckks_multiply (NTT form), BFV = bfv_multiply (coefficient form, BEHZ, evaluator.cpp:274-445). */
int gsb_ckks_multiply_sizes_dev(gsb_ctx *ctx, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch, size_t L, void *stream);
int gsb_ckks_multiply_sizes_host(gsb_ctx *ctx, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch, size_t L);
int gsb_bfv_multiply_sizes_dev(gsb_ctx *ctx, gsb_ws *ws, uint64_t plain_modulus, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch, size_t L, void *stream);
int gsb_bfv_multiply_sizes_host(gsb_ctx *ctx, uint64_t plain_modulus, const uint64_t *a, size_t size_a, const uint64_t *b, size_t size_b, uint64_t *out, size_t batch, size_t L);
/* Evaluator::relinearize_inplace = relinearize_internal(destination_size = 2) (evaluator.cpp:772-827) of
 * ciphertexts of any size: in = [batch][size][L][N] -> out = [batch][2][L][N]; needs the keys
 * RelinKeys::get_index(2 .. size-1) (gsb_key_upload with galois_elt = index).  Mirrors the reference's loop
 * word for word: every one of its size - 2 steps switches the LAST component (evaluator.cpp:806-815 never
 * moves the iterator).  bfv = 1: coefficient-form (BFV) branch.  out must not overlap in. */
int gsb_relinearize_sizes_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, size_t size, uint64_t *out, size_t batch, size_t L, int bfv, void *stream);
int gsb_relinearize_sizes_host(gsb_ctx *ctx, const uint64_t *in, size_t size, uint64_t *out, size_t batch, size_t L, int bfv);
/* Evaluator::mod_switch_to_next, CKKS branch = mod_switch_drop_to_next (evaluator.cpp:894-957): in =
 * [n_polys][L][N] -> out = [n_polys][L-1][N], the first L-1 limbs of every polynomial (device buffers). */
int gsb_drop_last_limb_dev(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t n_polys, size_t L, void *stream);

/* ---- plaintext operands, BFV ------------------------------------------------------
 * Evaluator::add_plain_inplace / sub_plain_inplace, BFV branch (evaluator.cpp:1290-1437 ->
 * multiply_add/sub_plain_with_scaling_variant, util/scalingvariant.cpp:15-95): enc0 = the ciphertext's first
 * polynomial [L][N], updated in place; plain = [count] coefficients mod t; the three constants are the level's
 * ContextData::coeff_div_plain_modulus() [L], coeff_modulus_mod_plain_modulus(), plain_upper_half_threshold(). */
int gsb_bfv_add_plain_dev(gsb_ctx *ctx, uint64_t *enc0, const uint64_t *plain, size_t count, int subtract, const uint64_t *coeff_div_plain_modulus, uint64_t q_mod_t, uint64_t threshold, uint64_t plain_modulus, size_t L, void *stream);
int gsb_bfv_add_plain_host(gsb_ctx *ctx, uint64_t *enc0, const uint64_t *plain, size_t count, int subtract, const uint64_t *coeff_div_plain_modulus, uint64_t q_mod_t, uint64_t threshold, uint64_t plain_modulus, size_t L);
/* Evaluator::transform_to_ntt_inplace(Plaintext &, parms_id) for contexts with using_fast_plain_lift
 * (evaluator.cpp:1648-1744): plain = [count] coefficients mod t -> out = [L][N]; threshold =
 * plain_upper_half_threshold(), increment = plain_upper_half_increment() [L] (host words); ntt = 0 stops after
 * the lift (:1722-1736). */
int gsb_bfv_plain_to_rns_dev(gsb_ctx *ctx, const uint64_t *plain, size_t count, uint64_t threshold, const uint64_t *increment, uint64_t *out, size_t L, int ntt, void *stream);
int gsb_bfv_plain_to_rns_host(gsb_ctx *ctx, const uint64_t *plain, size_t count, uint64_t threshold, const uint64_t *increment, uint64_t *out, size_t L, int ntt);
/* Evaluator::multiply_plain_inplace with a coefficient-form plaintext = multiply_plain_normal (evaluator.cpp:
 * 1475-1603): enc = [batch][size][L][N] coefficient form, in place; plain_ntt = [L][N] from gsb_bfv_plain_to_rns_dev.
 * The host form takes the raw plaintext [count]: lift = 1 generic branch (:1556-1592), lift = 0 the monomial
 * branch's unadjusted coefficient (:1534-1546); both give the reference's words (the negacyclic product is exact). */
int gsb_multiply_plain_coeff_dev(gsb_ctx *ctx, uint64_t *enc, const uint64_t *plain_ntt, size_t batch, size_t size, size_t L, void *stream);
int gsb_multiply_plain_coeff_host(gsb_ctx *ctx, uint64_t *enc, size_t size, size_t L, const uint64_t *plain, size_t count, uint64_t threshold, const uint64_t *increment, int lift);

/* ---- switching-key generation -------------------------------------------------------
 * KeyGenerator::generate_kswitch_keys / generate_one_kswitch_key (keygenerator.cpp:325-420) over
 * util::encrypt_zero_symmetric (util/rlwe.cpp:204-304), batched over n_keys keys.  Randomness stays with the
 * caller: a = [n_keys][beta][K][N] uniform residues in NTT form (the reference's sample_poly_uniform
 * output, which becomes component 1 of every bundle), e = [n_keys][beta][N] the noise polynomials as SIGNED
 * integers (sample_poly_normal writes e or q_i - |e| into every limb, rlwe.cpp:60-84).  The secret key must have
 * been given with gsb_secret_key_set.  kind GSB_KEY_RELIN: ids[n] = key power >= 2 (new key = s^power,
 * keygenerator.cpp:256-323; stored under RelinKeys::get_index(power)); kind GSB_KEY_GALOIS: ids[n] = Galois
 * element (new key = apply_galois_ntt(s), keygenerator.cpp:99-101).  out = [n_keys][beta][2][K][N] exactly as
 * KSwitchKeys::data()[index] lays its bundles out, or NULL with install = 1; install = 1 also places the keys in
 * the context's key store (as gsb_key_upload would), so generated keys never have to leave the device.  Same
 * sampled buffers in -> the reference's key words out. */
int gsb_kswitch_keygen_dev(gsb_ctx *ctx, gsb_ws *ws, int kind, const uint32_t *ids, size_t n_keys, const uint64_t *a, const int64_t *e, uint64_t *out, int install, void *stream);
int gsb_kswitch_keygen_host(gsb_ctx *ctx, int kind, const uint32_t *ids, size_t n_keys, const uint64_t *a, const int64_t *e, uint64_t *out, int install);

/* ---- Decryptor (seal::Decryptor, decryptor.cpp) --------------------------------
 * gsb_secret_key_set: SecretKey::data() in NTT form at key level, sk_host = [K][N] canonical residues
 * (Decryptor::Decryptor, decryptor.cpp:21-58; rejects what is_valid_for(SecretKey) rejects).  Powers
 * s^2.. are computed on the device on demand (compute_secret_key_array, decryptor.cpp:151-216).
 * gsb_dot_product_ct_sk_*: Decryptor::dot_product_ct_sk_array (decryptor.cpp:218-265):
 *   in = [batch][size][L][N] -> out = [batch][L][N] = c_0 + c_1 s + ... + c_{size-1} s^{size-1} mod q_i;
 *   ntt_form = Ciphertext::is_ntt_form(): 1 (CKKS; this IS Decryptor::ckks_decrypt's plaintext,
 *   decryptor.cpp:115-149), 0 (BFV: lazy NTT, dyadic products, inverse NTT, + c_0).
 * gsb_bfv_decrypt_*: Decryptor::bfv_decrypt (decryptor.cpp:76-113) = the dot product followed by
 *   RNSTool::decrypt_scale_and_round (util/rns.cpp:1070-1126); out = [batch][N] plaintext coefficients
 *   mod plain_modulus (the caller trims leading zeros as the reference does at :107-111). */
int gsb_secret_key_set(gsb_ctx *ctx, const uint64_t *sk_host);
int gsb_dot_product_ct_sk_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, int ntt_form, void *stream);
int gsb_dot_product_ct_sk_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, int ntt_form);
int gsb_bfv_decrypt_dev(gsb_ctx *ctx, gsb_ws *ws, uint64_t plain_modulus, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream);
int gsb_bfv_decrypt_host(gsb_ctx *ctx, uint64_t plain_modulus, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L);

/* ---- allocation-free rescale ---------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same operation and same output words as gsb_rescale_dev: in = [batch][size][L][N] -> out =
 * [batch][size][L-1][N], out must not overlap in.  Its scratch (the rounded last limb and the level's
 * constants) lives in a file-scope static device array inside the library -- 4 MiB + 2 KiB that every GPU
 * which loads the library holds until the process exits -- so the call issues no cudaMalloc, cudaFree or
 * memcpy: kernels only.  Limit: batch * size * N <= 524288 words (e.g. 8 size-2 ciphertexts at N = 32768,
 * 64 at N = 4096); larger calls fail with GSB_E_RANGE and a message naming the limit.  Every call on a
 * device shares that array and it keeps the last call's contents; nothing needs to start at zero (each call
 * writes every word it later reads), but calls that overlap on the GPU -- different streams, or host
 * threads without a common stream -- race: serialise them on one stream. */
int gsb_rescale_noalloc_dev(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream);

/* 1 when the context's moduli satisfy q_max (5 log2 N + 2) < 2^63, i.e. the key-switch / rescale
 * internals run the wide-range butterflies (identical outputs); else 0. */
int gsb_ctx_fast_ntt(const gsb_ctx *ctx);
/* Testing aid: a context created while the environment variable GSB_LITERAL is "1" takes none of those
 * shortcuts (gsb_ctx_fast_ntt returns 0); outputs are identical either way. */
/* Which congruent-word shortcuts the context's bounds allow (no reference counterpart, outputs are
 * identical with any of them on or off): bit 0 fast_ntt, 1 lazy_modup, 2 lazy_mac, 3 wide_inv, 4 fast_fwd_public.
 * 0 for a null context. */
int gsb_ctx_shortcuts(const gsb_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* GSEAL_B200_H */
