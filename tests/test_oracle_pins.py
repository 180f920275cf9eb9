"""Pins the CPU oracle (oracle/seal_oracle.c) before anything trusts it:
  (a) known-answer vectors transcribed from the reference's own unit tests;
  (b) the UNMODIFIED reference compiled from /root/reference (oracle/_ref) on seeded inputs --
      skipped where the reference is absent (the GPU box);
  (c) the committed golden digests under tests/golden/ that the reference produced
      (tests/golden/make_golden.py).
CPU only."""
import json
import os

import numpy as np
import pytest

from common import digest, mixed_moduli, ntt_primes, rand_key, rand_poly

Q61 = 0xFFFFFFFFFFC0001  # native/tests/seal/util/ntt.cpp:55-75


# ---- (a) KATs from the reference's tests ---------------------------------------------------------
def test_barrett_kats(oracle):
    # native/tests/seal/util/uintarithsmallmod.cpp:142-185 (BarrettReduce128 / 63)
    assert oracle.barrett_reduce_128(0, 0, 2) == 0
    assert oracle.barrett_reduce_128(1, 0, 2) == 1
    assert oracle.barrett_reduce_128(0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFF, 2) == 1
    assert oracle.barrett_reduce_128(0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFF, 3) == 0
    assert oracle.barrett_reduce_128(123, 456, 13131313131313) == 8722750765283
    assert oracle.barrett_reduce_128(24242424242424, 79797979797979, 13131313131313) == 1010101010101
    assert oracle.barrett_reduce_63(0, 2) == 0
    assert oracle.barrett_reduce_63(1, 2) == 1
    assert oracle.barrett_reduce_63(0x7FFFFFFFFFFFFFFF, 2) == 1
    assert oracle.barrett_reduce_63(0x7FFFFFFFFFFFFFFF, 3) == 1
    assert oracle.barrett_reduce_63(1313131313131313, 13131313131313) == 1313131313131313 % 13131313131313


def test_multiply_mod_kats(oracle):
    # native/tests/seal/util/uintarithsmallmod.cpp:186-230 (MultiplyUIntMod / MultiplyAddMod)
    assert oracle.multiply_uint_mod(0, 0, 2) == 0
    assert oracle.multiply_uint_mod(1, 1, 2) == 1
    assert oracle.multiply_uint_mod(7, 7, 10) == 9
    assert oracle.multiply_uint_mod(1152921504606846974, 1152921504606846974, 1152921504606846975) == 1
    assert oracle.multiply_add_uint_mod(1, 1, 1, 7) == 2
    assert oracle.multiply_add_uint_mod(6, 6, 6, 7) == 0
    for a, b, c, q in [(3, 5, 6, 7), (0xFFFFFFFF12345, 0xABCDEF987, 17, 1152921504606846883)]:
        assert oracle.multiply_add_uint_mod(a, b, c, q) == (a * b + c) % q


def test_ntt_kats(oracle):
    # native/tests/seal/util/ntt.cpp:98-102: NTT([1,1]) under q = 0xffffffffffc0001, N = 2
    out = oracle.ntt(1, Q61, np.array([1, 1], dtype=np.uint64), 1)
    assert out.tolist() == [288794978602139553, 864126526004445282]
    # :105-135: random round trip at N = 8
    rng = np.random.default_rng(7)
    x = rng.integers(0, Q61, size=8, dtype=np.uint64)
    assert np.array_equal(oracle.ntt(3, Q61, oracle.ntt(3, Q61, x, 1), 3), x)
    z = np.zeros(8, dtype=np.uint64)
    assert np.array_equal(oracle.ntt(3, Q61, z, 1), z)


def test_galois_kats(oracle):
    # native/tests/seal/util/galois.cpp:79-113 (N = 8, NTT form) and :28-55 (elt from step)
    assert oracle.galois_table_ntt(3, 3).tolist() == [4, 5, 7, 6, 1, 0, 2, 3]
    # GaloisTool(3): N = 8, generator 5, m = 16
    assert oracle.elt_from_step(3, 0) == 15
    assert oracle.elt_from_step(3, 1) == 5      # 5^1
    assert oracle.elt_from_step(3, -3) == 5     # 5^(N/2-3) = 5^1
    assert oracle.elt_from_step(3, 2) == 9      # 25 mod 16
    assert oracle.elt_from_step(3, -2) == 9
    assert oracle.elt_from_step(3, 3) == 13     # 125 mod 16
    assert oracle.elt_from_step(3, -1) == 13


def test_dyadic_kats(oracle):
    # native/tests/seal/util/polyarithsmallmod.cpp:545-640 (DyadicProductCoeffmod, tiny cases)
    a = np.array([1, 1, 1], dtype=np.uint64)
    b = np.array([2, 3, 4], dtype=np.uint64)
    assert oracle.poly_op(0, 13, a, b).tolist() == [2, 3, 4]
    a = np.array([1, 2, 3], dtype=np.uint64)
    assert oracle.poly_op(0, 13, a, b).tolist() == [2, 6, 12]
    assert oracle.poly_op(0, 7, a, b).tolist() == [2, 6, 5]
    # MultiplyPolyScalarCoeffmod (:340-410): [1,3,4] * 3 mod 5
    assert oracle.poly_op(4, 5, np.array([1, 3, 4], dtype=np.uint64), scalar=3).tolist() == [3, 4, 2]


# ---- (b) restatement vs the unmodified reference ---------------------------------------------------
@pytest.mark.parametrize("logn", [3, 10, 12, 13])
@pytest.mark.parametrize("bits", [30, 50, 59, 60])
def test_ntt_vs_reference(oracle, reflib, logn, bits):
    """All four transforms, lazy outputs included; bits=60 exercises the mod-2^64 wrap of the lazy
    forward transform at N >= 4096 -- the restatement must mirror it word for word."""
    q = ntt_primes(logn, bits, 1)[0]
    rng = np.random.default_rng(logn * 100 + bits)
    x = rng.integers(0, q, size=(3, 1 << logn), dtype=np.uint64)
    for kind in range(4):
        assert np.array_equal(oracle.ntt(logn, q, x, kind), reflib.ntt(logn, q, x, kind)), (logn, bits, kind)


def test_const_ratio_vs_reference(oracle, reflib):
    for q in [2, 3, 13131313131313, Q61] + ntt_primes(12, 59, 2) + ntt_primes(10, 36, 2):
        assert list(oracle.const_ratio(q)) == reflib.const_ratio(q).tolist()


@pytest.mark.parametrize("op", [0, 1, 2, 3, 4])
def test_poly_ops_vs_reference(oracle, reflib, op):
    rng = np.random.default_rng(op)
    for q in ntt_primes(10, 59, 1) + ntt_primes(10, 31, 1) + [Q61]:
        a = rng.integers(0, q, size=4096, dtype=np.uint64)
        b = rng.integers(0, q, size=4096, dtype=np.uint64)
        a[:4] = [0, q - 1, 0, q - 1]
        b[:4] = [0, q - 1, q - 1, 0]
        s = int(rng.integers(0, q))
        assert np.array_equal(oracle.poly_op(op, q, a, b, s), reflib.poly_op(op, q, a, b, s))


def test_galois_vs_reference(oracle, reflib):
    mods = ntt_primes(10, 40, 3)
    rc = reflib.context(1024, mods, 1)
    for step in [0, 1, -1, 2, 5, -7, 100, 511, -511]:
        assert oracle.elt_from_step(10, step) == rc.elt_from_step(step)


EVAL_CASES = [
    # (logn, bit sizes of the key-level modulus, k special, level L)
    (10, [40, 40, 40, 41], 1, 3),
    (10, [36, 36, 37], 1, 2),
    (11, [50, 40, 45, 45, 50, 51], 2, 4),
    (11, [50, 40, 45, 45, 50, 51], 2, 3),     # ragged last bundle (single limb)
    (10, [45, 45, 45, 45, 45, 46, 46, 46], 3, 5),  # k = 3, bundles {0,1,2},{3,4}
    (10, [45, 45, 45, 45, 45, 46, 46, 46], 3, 4),  # bundles {0,1,2},{3}
    (12, [55, 55, 55, 56], 1, 3),
    (10, [59, 59, 59, 59], 1, 3),
]


def make_case(logn, bits, k, L, seed):
    N = 1 << logn
    mods = mixed_moduli(logn, bits)
    K = len(mods)
    rng = np.random.default_rng(seed)
    a = rand_poly(rng, mods[:L], (2,), N)
    b = rand_poly(rng, mods[:L], (2,), N)
    nb = (K - k + k - 1) // k
    relin = rand_key(rng, mods, nb, N)
    gal = rand_key(rng, mods, nb, N)
    return N, mods, a, b, relin, gal


@pytest.mark.parametrize("case", EVAL_CASES, ids=lambda c: "n%d_K%d_k%d_L%d" % (c[0], len(c[1]), c[2], c[3]))
def test_evaluator_vs_reference(oracle, reflib, case):
    """multiply / relinearize / rescale / apply_galois / rotate of the restatement against
    seal::Evaluator on identical limb buffers, synthetic keys written into the reference's KSwitchKeys."""
    logn, bits, k, L = case
    N, mods, a, b, relin, gal = make_case(logn, bits, k, L, seed=logn * 1000 + L * 10 + k)
    oc = oracle.context(logn, mods, k)
    rc = reflib.context(N, mods, k)
    elt = oracle.elt_from_step(logn, 1)
    rc.keygen([elt])
    assert rc.get_key(0).shape == relin.shape
    rc.set_key(0, 0, relin)
    rc.set_key(1, elt, gal)
    c3 = rc.multiply(a, b)
    assert np.array_equal(oc.multiply(a, b), c3)
    c2 = rc.relinearize(c3)
    assert np.array_equal(oc.relinearize(c3, relin), c2)
    assert np.array_equal(oc.rescale(c2), rc.rescale(c2))
    assert np.array_equal(oc.rescale(c3), rc.rescale(c3))
    assert np.array_equal(oc.mul_relin_rescale(a, b, relin), rc.mul_relin_rescale(a, b))
    g = rc.apply_galois(a, elt)
    assert np.array_equal(oc.apply_galois(a, elt, gal), g)
    assert np.array_equal(rc.rotate(a, 1), g)


def test_evaluator_real_keys_vs_reference(oracle, reflib):
    """Same, with keys generated by the reference's own KeyGenerator (keygenerator.cpp:325-369)."""
    logn, k, L = 10, 2, 4
    mods = mixed_moduli(logn, [45, 45, 45, 45, 46, 46])
    N = 1 << logn
    rng = np.random.default_rng(5)
    a = rand_poly(rng, mods[:L], (2,), N)
    b = rand_poly(rng, mods[:L], (2,), N)
    oc = oracle.context(logn, mods, k)
    rc = reflib.context(N, mods, k)
    elt = oracle.elt_from_step(logn, -3)
    rc.keygen([elt])
    relin, gal = rc.get_key(0), rc.get_key(1, elt)
    assert np.array_equal(oc.mul_relin_rescale(a, b, relin), rc.mul_relin_rescale(a, b))
    assert np.array_equal(oc.apply_galois(a, elt, gal), rc.rotate(a, -3))


def test_bfv_branches_vs_reference(oracle, reflib):
    """BASELINE config #1 (BFV N = 4096, default 109-bit modulus, t = 786433 as in 7_performance.cpp:543-552)
    and a k = 2 context: bfv_multiply (BEHZ), relinearize and apply_galois on coefficient-form ciphertexts,
    keys from the reference's own KeyGenerator."""
    import golden_cases
    from common import mixed_moduli, rand_poly
    for logn, mods, k, t in [(12, golden_cases.BFV_DEFAULT[4096], 1, 786433), (11, mixed_moduli(11, [50, 40, 45, 45, 50, 51]), 2, 65537)]:
        N, L = 1 << logn, len(mods) - k
        rc = reflib.context(N, mods, k, bfv_plain_modulus=t)
        oc = oracle.context(logn, mods, k)
        elt = oracle.elt_from_step(logn, 1)
        rc.keygen([elt, 2 * N - 1])
        rng = np.random.default_rng(logn)
        a, b = rand_poly(rng, mods[:L], (2,), N), rand_poly(rng, mods[:L], (2,), N)
        c3 = rc.multiply(a, b)
        assert np.array_equal(oracle.bfv_multiply(logn, mods, t, a, b), c3)
        assert np.array_equal(oc.relinearize_bfv(c3, rc.get_key(0)), rc.relinearize(c3))
        for e in (elt, 2 * N - 1):
            assert np.array_equal(oc.apply_galois_bfv(a, e, rc.get_key(1, e)), rc.apply_galois(a, e))


def test_decryptor_vs_reference(oracle, reflib):
    """Row f4: the restated dot_product_ct_sk_array / bfv_decrypt against seal::Decryptor::decrypt of the
    unmodified reference, every level, ciphertext sizes 2 and 3, k = 1 and 2, synthetic secret key limbs."""
    from common import mixed_moduli, rand_poly
    for logn, bits, k, t in [(10, [40, 40, 41, 42], 1, 65537), (11, [50, 50, 50, 51, 51], 2, 65537), (12, [36, 36, 37], 1, 786433)]:
        N, mods = 1 << logn, mixed_moduli(logn, bits)
        rc, rb = reflib.context(N, mods, k), reflib.context(N, mods, k, bfv_plain_modulus=t)
        oc = oracle.context(logn, mods, k)
        rng = np.random.default_rng(logn)
        sk = rand_poly(rng, mods, (), N)
        for size in (2, 3):
            for L in range(1, len(mods) - k + 1):
                c = rand_poly(rng, mods[:L], (size,), N)
                assert np.array_equal(oc.dot_product_ct_sk(c, sk, True), rc.decrypt(c, sk)), (logn, size, L)
                assert np.array_equal(oc.bfv_decrypt(c, sk, t), rb.decrypt(c, sk, bfv=True)), (logn, size, L)


# ---- (c) golden digests produced by the reference ---------------------------------------------------
def test_golden_digests(oracle):
    path = os.path.join(os.path.dirname(__file__), "golden", "evaluator_digests.json")
    with open(path) as f:
        golden = json.load(f)
    import golden_cases
    for name, want in golden["cases"].items():
        got = golden_cases.run_case(name, golden_cases.OracleBackend(oracle))
        assert got == want, name


def test_batch64_digests_sample(oracle):
    """Two ciphertexts of the batch-64 C3 case (the GPU suite checks all 64 of C3 and C4) through the oracle."""
    import golden_cases
    with open(os.path.join(os.path.dirname(__file__), "golden", "batch64_digests.json")) as f:
        want = json.load(f)["cases"]["c3_n16384_default_b64"]
    got = golden_cases.run_case_per_ct("c3_n16384_default_b64", golden_cases.OracleBackend(oracle), only=(0, 63))
    for op, d in got.items():
        assert d == [want[op][0], want[op][63]], op


def test_bench_digests_and_synth(oracle):
    """bench.py's parity-gate inputs: the splitmix residues are the same words from numpy and from torch's int64
    arithmetic, and the committed reference digest of C3 ciphertext 255 is what the oracle computes."""
    import torch
    import bench
    import synth
    mods = [36028797017456641, 72057594037338113, 65537]
    a = synth.residues_numpy(99, 3, 5, 2, mods, 1024)
    b = synth.residues_torch(torch, 99, 3, 5, 2, mods, 1024, device="cpu", chunk=2).numpy().view(np.uint64)
    assert np.array_equal(a, b)
    assert np.array_equal(synth.residues_numpy(99, 5, 1, 2, mods, 1024)[0], a[2])  # a ciphertext does not depend on its batch
    w = bench.workload("c3")
    keys = bench.synthetic_keys(w)
    oc = oracle.context(w.logn, w.mods, w.k)
    ca = synth.residues_numpy(bench.SEED_A, 255, 1, 2, w.mods[:w.L], w.N)[0]
    cb = synth.residues_numpy(bench.SEED_B, 255, 1, 2, w.mods[:w.L], w.N)[0]
    want = bench.load_digests()["c3"]
    assert bench.digest(oc.mul_relin_rescale(ca, cb, keys[0])) == want["mrr"]["255"]
    assert bench.digest(oc.apply_galois(ca, oracle.elt_from_step(w.logn, 1), keys[1])) == want["rot"]["255"]


@pytest.mark.parametrize("name", ["keygen_n1024_k1", "keygen_n2048_k2_ragged"])
def test_keygen_oracle_reproduces_reference_keys(oracle, name):
    """Row f2: the restated generate_one_kswitch_key / encrypt_zero_symmetric, given the buffers the reference sampled
    (committed fixture, tests/golden/make_keygen_fixtures.py), reproduces the reference KeyGenerator's relin and Galois
    keys word for word."""
    from common import load_keygen_fixture, oracle_keygen
    fx = load_keygen_fixture(name)
    got = oracle_keygen(oracle, None, fx["logn"], fx["mods"], fx["k"], fx["sk"], [2], False, fx["a"][:1], fx["e"][:1])
    assert np.array_equal(got[0], fx["keys"][0])
    got = oracle_keygen(oracle, None, fx["logn"], fx["mods"], fx["k"], fx["sk"], fx["elts"], True, fx["a"][1:], fx["e"][1:])
    for n in range(len(fx["elts"])):
        assert np.array_equal(got[n], fx["keys"][1 + n])


def test_keygen_oracle_vs_live_reference(oracle, reflib):
    """The same against keys the reference generates now (fresh randomness), k = 3 with a ragged last bundle."""
    from common import oracle_keygen, reference_keys_for_keygen
    logn, k = 10, 3
    mods = mixed_moduli(logn, [40, 40, 40, 40, 41, 41, 41])
    fx = reference_keys_for_keygen(reflib, oracle, logn, mods, k, [1])
    a, e = np.stack(fx["a"]), np.stack(fx["e"])
    assert np.array_equal(oracle_keygen(oracle, None, logn, mods, k, fx["sk"], [2], False, a[:1], e[:1])[0], fx["keys"][0])
    assert np.array_equal(oracle_keygen(oracle, None, logn, mods, k, fx["sk"], fx["elts"], True, a[1:], e[1:])[0], fx["keys"][1])
