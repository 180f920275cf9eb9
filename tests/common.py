"""Shared helpers for the parity tests: NTT-friendly primes, seeded synthetic ciphertexts / keys
(uniform residues), and the parity cases run against either the CUDA library (-m gpu) or
the test-only simulator build of the same kernel sources (CPU)."""
import hashlib

import numpy as np
import pytest


def is_prime(n):
    if n < 2:
        return False
    small = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)
    for p in small:
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for a in small:
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def ntt_primes(logn, bits, count, skip=0):
    """The `count` largest primes q = 1 (mod 2N) below 2^bits (after skipping `skip` of them) --
    the same family CoeffModulus::Create draws from (util/numth.cpp get_primes)."""
    out, step = [], 2 << logn
    q = ((1 << bits) - 1) // step * step + 1
    while len(out) < count + skip:
        if is_prime(q):
            out.append(q)
        q -= step
    return out[skip:]


def mixed_moduli(logn, bit_sizes):
    """One distinct prime per entry of bit_sizes."""
    used, out = {}, []
    for b in bit_sizes:
        i = used.get(b, 0)
        out.append(ntt_primes(logn, b, 1, skip=i)[0])
        used[b] = i + 1
    return out


def rand_poly(rng, moduli, lead_shape, N):
    """uint64 array lead_shape + [len(moduli), N], limb i uniform in [0, moduli[i])."""
    cols = [rng.integers(0, q, size=tuple(lead_shape) + (N,), dtype=np.uint64) for q in moduli]
    return np.ascontiguousarray(np.stack(cols, axis=-2))


def rand_key(rng, moduli, n_bundles, N):
    """[n_bundles][2][K][N] uniform residues: a synthetic switching key (same limbs fed to both sides)."""
    return rand_poly(rng, moduli, (n_bundles, 2), N)


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()


NTT_KINDS = {  # (inverse, lazy) -> oracle kind
    (False, True): 0, (False, False): 1, (True, True): 2, (True, False): 3,
}


def check_thin_ops(lib, oracle, logn=11, to_device=None):
    """add / sub / negate / multiply_plain / square / mod_switch_to_next /
    transform_{to,from}_ntt of the Evaluator mirror against the oracle's word arithmetic and NTT.
    `to_device` moves a numpy batch to the buffers under test (identity for host buffers)."""
    import gemini_seal_b200 as g
    to_device = to_device or (lambda x: x)
    back = lambda x: x if isinstance(x, np.ndarray) else x.cpu().numpy().view(np.uint64)
    N = 1 << logn
    mods = mixed_moduli(logn, [50, 40, 45, 51])
    L = 3
    ctx = g.Context(N, mods, 1, lib=lib)
    ev = g.Evaluator(ctx)
    rng = np.random.default_rng(logn)
    a, b = rand_poly(rng, mods[:L], (2, 2), N), rand_poly(rng, mods[:L], (2, 2), N)
    plain = rand_poly(rng, mods[:L], (), N)
    da, db = to_device(a), to_device(b)
    per_limb = lambda op, x, y=None: np.stack(
        [oracle.poly_op(op, mods[i], x[..., i, :], None if y is None else y[..., i, :]) for i in range(L)], axis=-2)
    assert np.array_equal(back(ev.add(da, db)), per_limb(1, a, b))
    assert np.array_equal(back(ev.sub(da, db)), per_limb(2, a, b))
    assert np.array_equal(back(ev.negate(da)), per_limb(3, a))
    assert np.array_equal(back(ev.multiply_plain(da, to_device(plain))), per_limb(0, a, np.broadcast_to(plain, a.shape)))
    oc = oracle.context(logn, mods, 1)
    assert np.array_equal(back(ev.square(da)), np.stack([oc.multiply(a[i], a[i]) for i in range(2)]))
    assert np.array_equal(back(ev.mod_switch_to_next(da)), a[:, :, :L - 1, :])
    coef = oracle.ntt_rns(logn, mods, a, 3)
    assert np.array_equal(back(ev.transform_from_ntt(da)), coef)
    assert np.array_equal(back(ev.transform_to_ntt(to_device(coef))), a)
    ctx.close()


def fuzz_against_oracle(lib, oracle, seed, n_configs, max_logn=13):
    """Random parameter sets (N, #primes, bit sizes incl. 57..59-bit ones that disable the lazy-range
    shortcuts, k special primes, level L with ragged last bundle, batch, rotation step; CKKS and BFV)
    through the C ABI against the oracle, array by array."""
    import gemini_seal_b200 as g
    rng = np.random.default_rng(seed)
    for _ in range(n_configs):
        logn = int(rng.integers(10, max_logn + 1))
        N = 1 << logn
        k = int(rng.integers(1, 4))
        L0 = int(rng.integers(1, 7))
        wide = bool(rng.integers(0, 4) == 0)
        bits = [int(b) for b in rng.integers(57, 60, size=L0 + k)] if wide else [int(b) for b in rng.integers(30, 56, size=L0 + k)]
        mods = mixed_moduli(logn, bits)
        L = int(rng.integers(1, L0 + 1))
        batch = int(rng.integers(1, 4))
        step = int(rng.integers(1, N // 2)) * (1 if rng.integers(0, 2) else -1)
        beta0 = (L0 + k - 1) // k
        what = (logn, bits, k, L, batch, step)
        ctx = g.Context(N, mods, k, lib=lib)
        oc = oracle.context(logn, mods, k)
        relin, gal = rand_key(rng, mods, beta0, N), rand_key(rng, mods, beta0, N)
        ctx.set_relin_keys(relin)
        elt = ctx.galois_elt_from_step(step)
        assert elt == oracle.elt_from_step(logn, step), what
        ctx.set_galois_key(elt, gal)
        a, b = rand_poly(rng, mods[:L], (batch, 2), N), rand_poly(rng, mods[:L], (batch, 2), N)
        ev = g.Evaluator(ctx)
        c3 = ev.multiply(a, b)
        for i in range(batch):
            assert np.array_equal(c3[i], oc.multiply(a[i], b[i])), ("multiply", what)
        c2 = ev.relinearize(c3)
        rot = ev.apply_galois(a, elt)
        for i in range(batch):
            assert np.array_equal(c2[i], oc.relinearize(c3[i], relin)), ("relinearize", what)
            assert np.array_equal(rot[i], oc.apply_galois(a[i], elt, gal)), ("apply_galois", what)
        if L >= 2:
            rs = ev.rescale_to_next(c3)
            for i in range(batch):
                assert np.array_equal(rs[i], oc.rescale(c3[i])), ("rescale", what)
        # the BFV branches on the same context (plain modulus below every prime)
        t = int(rng.choice([65537, 786433, 12289]))
        bev = g.BfvEvaluator(ctx, t)
        b3 = bev.multiply(a, b)
        b2 = bev.relinearize(b3)
        brot = bev.apply_galois(a, elt)
        for i in range(batch):
            assert np.array_equal(b3[i], oracle.bfv_multiply(logn, mods, t, a[i], b[i])), ("bfv multiply", what, t)
            assert np.array_equal(b2[i], oc.relinearize_bfv(b3[i], relin)), ("bfv relinearize", what)
            assert np.array_equal(brot[i], oc.apply_galois_bfv(a[i], elt, gal)), ("bfv apply_galois", what)
        if L >= 2:
            ms = bev.mod_switch_to_next(b3)
            for i in range(batch):
                assert np.array_equal(ms[i], oc.mod_switch_bfv(b3[i])), ("bfv mod_switch", what)
        ctx.close()


def check_decrypt(lib, oracle, logn=11, to_device=None):
    """Decryptor mirror (CKKS dot product, BFV dot product + decrypt_scale_and_round)
    against the oracle on every level, sizes 2..4 (secret-key powers up to s^3), batch 3."""
    import gemini_seal_b200 as g
    to_device = to_device or (lambda x: x)
    back = lambda x: x if isinstance(x, np.ndarray) else x.cpu().numpy().view(np.uint64)
    N = 1 << logn
    for bits, k, t in [([50, 40, 45, 51], 1, 65537), ([36, 36, 37, 58, 59], 2, 786433)]:
        mods = mixed_moduli(logn, bits)
        ctx = g.Context(N, mods, k, lib=lib)
        oc = oracle.context(logn, mods, k)
        rng = np.random.default_rng(logn * 7 + k)
        sk = rand_poly(rng, mods, (), N)
        ckks, bfv = g.Decryptor(ctx, sk), g.Decryptor(ctx, sk, plain_modulus=t)
        for size in (3, 2, 4):  # 3 first: the power array is extended, then reused, then extended again
            for L in range(1, len(mods) - k + 1):
                c = rand_poly(rng, mods[:L], (3, size), N)
                got = back(ckks.decrypt(to_device(c)))
                coef = back(bfv.dot_product_ct_sk_array(to_device(c), False))
                plain = back(bfv.decrypt(to_device(c)))
                for i in range(3):
                    assert np.array_equal(got[i], oc.dot_product_ct_sk(c[i], sk, True)), ("ckks", bits, size, L)
                    assert np.array_equal(coef[i], oc.dot_product_ct_sk(c[i], sk, False)), ("bfv dot", bits, size, L)
                    assert np.array_equal(plain[i], oc.bfv_decrypt(c[i], sk, t)), ("bfv decrypt", bits, size, L)
        with pytest.raises(ValueError):  # is_valid_for(SecretKey): non-canonical residue
            bad = sk.copy()
            bad[0, 0] = mods[0]
            g.Decryptor(ctx, bad)
        ctx.close()
