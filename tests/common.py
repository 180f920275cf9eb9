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


def threshold_primes(logn, k_special=1, L0=3):
    """For each no-wrap flag of gsb_ctx_create (fast_ntt: q (5n+2) < 2^63; lazy_mac: beta0 (5n+2) q^2 < 2^128;
    wide_inv: 10 q < 2^63) the largest NTT prime satisfying the bound and the smallest violating it."""
    n = logn
    step = 2 << logn
    beta0 = (L0 + k_special - 1) // k_special

    def around(bound):  # largest prime q = 1 mod 2N with q < bound, smallest with q >= bound
        lo = (bound - 1) // step * step + 1
        while lo >= bound or not is_prime(lo):
            lo -= step
        hi = (bound + step - 1) // step * step + 1
        while hi < bound or not is_prime(hi):
            hi += step
        return lo, hi
    import math
    fast = -(-(1 << 63) // (5 * n + 2))
    mac = math.isqrt((1 << 128) // (beta0 * (5 * n + 2))) + 1
    inv = -(-(1 << 63) // 10)
    return {"fast_ntt": around(fast), "lazy_mac": around(min(mac, fast)), "wide_inv": around(inv)}


def check_threshold_and_patterns(lib, oracle, logn, literal_too=True, light=False):
    """contexts whose largest modulus sits just below / just above each shortcut threshold, fed with
    structured extreme inputs (all q-1, all 0, alternating 0 / q-1; keys full of q-1) next to random ones, through
    multiply / relinearize / rescale / mul_relin_rescale / apply_galois against the oracle; the default build and a
    GSB_LITERAL=1 context must both equal it."""
    import os
    import gemini_seal_b200 as g
    N, k, L0 = 1 << logn, 1, 3
    th = threshold_primes(logn, k, L0)
    primes = sorted({q for pair in th.values() for q in pair})
    small = ntt_primes(logn, 50, L0)
    for qbig in primes:
        # the big prime once as a ciphertext modulus and once as the special prime
        for vi, mods in enumerate((small[:L0 - 1] + [qbig] + ntt_primes(logn, 51, 1), small + [qbig])):
            L = L0
            rng = np.random.default_rng(qbig % (1 << 31))
            top = np.array([q - 1 for q in mods], dtype=np.uint64)
            pats = []
            full = np.broadcast_to(top[:L, None], (L, N))
            alt = np.where((np.arange(N) & 1)[None, :] == 1, full, 0).astype(np.uint64)
            for x in (full, np.zeros((L, N), dtype=np.uint64), alt):
                pats.append(np.ascontiguousarray(np.stack([x, x[:, ::-1]])))
            a = np.ascontiguousarray(np.stack(pats + [rand_poly(rng, mods[:L], (2,), N)]))
            b = np.ascontiguousarray(np.stack([pats[0], pats[2], pats[0], rand_poly(rng, mods[:L], (2,), N)]))
            key_max = np.ascontiguousarray(np.broadcast_to(top[None, None, :, None], (L0, 2, len(mods), N)))
            for ki, key in enumerate((key_max, rand_key(rng, mods, L0, N))):
                if light and ki != vi:  # the simulator run pairs each placement with one key kind
                    continue
                oc = oracle.context(logn, mods, k)
                elt = oracle.elt_from_step(logn, 1)
                want = None
                for literal in ((False, True) if literal_too else (False,)):
                    if literal:
                        os.environ["GSB_LITERAL"] = "1"
                    try:
                        ctx = g.Context(N, mods, k, lib=lib)
                    finally:
                        os.environ.pop("GSB_LITERAL", None)
                    ctx.set_relin_keys(key)
                    ctx.set_galois_key(elt, key)
                    ev = g.Evaluator(ctx)
                    c3 = ev.multiply(a, b)
                    got = [c3, ev.relinearize(c3), ev.rescale_to_next(c3), ev.mul_relin_rescale(a, b), ev.apply_galois(a, elt),
                           g.ntt_negacyclic_harvey(g.inverse_ntt_negacyclic_harvey(a.copy(), ctx), ctx)]
                    ctx.close()
                    if want is None:
                        want = [np.stack([oc.multiply(a[i], b[i]) for i in range(4)])]
                        want.append(np.stack([oc.relinearize(want[0][i], key) for i in range(4)]))
                        want.append(np.stack([oc.rescale(want[0][i]) for i in range(4)]))
                        want.append(np.stack([oc.mul_relin_rescale(a[i], b[i], key) for i in range(4)]))
                        want.append(np.stack([oc.apply_galois(a[i], elt, key) for i in range(4)]))
                        want.append(oracle.ntt_rns(logn, mods, oracle.ntt_rns(logn, mods, a, 3), 1))
                    for i, (x, y) in enumerate(zip(got, want)):
                        assert np.array_equal(x, y), (logn, qbig, literal, i)


def keygen_inputs_from_key(oracle, logn, mods, k, sk, new_key, key):
    """Recover what the reference SAMPLED for a switching key it generated (its samplers are not
    reproducible, but a key stores a = component 1, and e follows from c0 = -(a s + e) + [factor * new_key]):
    returns a [beta][K][N] and e [beta][N] (int64).  Asserts that e is the same small integer under every limb."""
    N, K = 1 << logn, len(mods)
    L0 = K - k
    beta = key.shape[0]
    a = np.ascontiguousarray(key[:, 1])
    e = np.empty((beta, N), dtype=np.int64)
    for j in range(beta):
        per_limb = []
        for i, q in enumerate(mods):
            c0 = key[j, 0, i]
            if j * k <= i < min(j * k + k, L0):
                factor = 1
                for s in range(k):
                    factor = factor * mods[L0 + s] % q
                c0 = oracle.poly_op(2, q, c0, oracle.poly_op(4, q, new_key[i], scalar=factor))       # c0 - factor * new_key
            t = oracle.poly_op(1, q, c0, oracle.poly_op(0, q, sk[i], a[j, i]))                        # c0 + a s = -e
            en = oracle.ntt(logn, q, oracle.poly_op(3, q, t), 3).astype(np.int64)                     # strict inverse NTT of e
            per_limb.append(np.where(en > q // 2, en - q, en))
        for x in per_limb[1:]:
            assert np.array_equal(x, per_limb[0])
        assert np.abs(per_limb[0]).max() < 64  # clipped normal, sigma = 3.2 (util/rlwe.cpp:60-84)
        e[j] = per_limb[0]
    return a, e


def reference_keys_for_keygen(reflib, oracle, logn, mods, k, steps):
    """Keys from the reference's own KeyGenerator plus everything needed to regenerate them: (sk, relin key,
    {elt: galois key}, new keys, sampled a / e per key)."""
    N = 1 << logn
    rc = reflib.context(N, mods, k)
    elts = [rc.elt_from_step(s) for s in steps]
    rc.keygen(elts)
    sk = rc.secret_key()
    out = {"sk": sk, "elts": elts, "keys": [], "new": [], "a": [], "e": []}
    s2 = np.stack([oracle.poly_op(0, q, sk[i], sk[i]) for i, q in enumerate(mods)])
    cases = [(rc.get_key(0, 0), s2)]
    for elt in elts:
        t = oracle.galois_table_ntt(logn, elt)
        cases.append((rc.get_key(1, elt), np.ascontiguousarray(sk[:, t])))
    for key, new in cases:
        a, e = keygen_inputs_from_key(oracle, logn, mods, k, sk, new, key)
        out["keys"].append(key), out["new"].append(new), out["a"].append(a), out["e"].append(e)
    return out


def check_keygen_fixture(lib, fx, to_device=None):
    """Row f2 on the device: the fixture's sampled buffers in -> the reference's key words out (relin key + Galois
    keys, one batched call each), then an installed key must key-switch like the uploaded one."""
    import gemini_seal_b200 as g
    to_device = to_device or (lambda x: x)
    back = lambda x: x if isinstance(x, np.ndarray) else x.cpu().numpy().view(np.uint64)
    mods, k, logn = fx["mods"], fx["k"], fx["logn"]
    N = 1 << logn
    ctx = g.Context(N, mods, k, lib=lib)
    kg = g.KeyGenerator(ctx, fx["sk"])
    dev_e = lambda e: to_device(e) if to_device(e) is e else to_device(e.view(np.uint64))
    relin = kg.relin_keys(to_device(fx["a"][:1]), dev_e(fx["e"][:1]), count=1, install=True)
    assert np.array_equal(back(relin)[0], fx["keys"][0]), "relin key"
    if len(fx["elts"]):
        gal = kg.galois_keys(fx["elts"], to_device(fx["a"][1:]), dev_e(fx["e"][1:]), install=True)
        for n, elt in enumerate(fx["elts"]):
            assert np.array_equal(back(gal)[n], fx["keys"][1 + n]), ("galois key", elt)
    # the installed keys serve key switching exactly like uploaded ones
    L = len(mods) - k
    rng = np.random.default_rng(7)
    c3 = rand_poly(rng, mods[:L], (2, 3), N)
    c2 = rand_poly(rng, mods[:L], (2, 2), N)
    ev = g.Evaluator(ctx)
    got_r = ev.relinearize(c3)
    got_g = [ev.apply_galois(c2, elt) for elt in fx["elts"]]
    ctx2 = g.Context(N, mods, k, lib=lib)
    ctx2.set_relin_keys(fx["keys"][0])
    ev2 = g.Evaluator(ctx2)
    assert np.array_equal(got_r, ev2.relinearize(c3))
    for n, elt in enumerate(fx["elts"]):
        ctx2.set_galois_key(elt, fx["keys"][1 + n])
        assert np.array_equal(got_g[n], ev2.apply_galois(c2, elt))
    ctx.close()
    ctx2.close()


def load_keygen_fixture(name):
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"))
    n = z["keys"].shape[0]
    return {"logn": int(z["logn"]), "mods": [int(q) for q in z["mods"]], "k": int(z["k"]), "sk": z["sk"], "elts": [int(x) for x in z["elts"]],
            "keys": [z["keys"][i] for i in range(n)], "a": z["a"], "e": z["e"]}


def oracle_keygen(oracle, fx_or_ctx, logn, mods, k, sk, ids, galois, a, e):
    """All keys of a batch through the oracle: ids = key powers (galois False) or Galois elements (True)."""
    oc = oracle.context(logn, mods, k)
    out = []
    for n, ident in enumerate(ids):
        if galois:
            new = np.ascontiguousarray(sk[:, oracle.galois_table_ntt(logn, ident)])
        else:
            new = sk
            for _ in range(ident - 1):
                new = np.stack([oracle.poly_op(0, q, new[i], sk[i]) for i, q in enumerate(mods)])
        out.append(oc.kswitch_keygen(sk, new, a[n], e[n]))
    return np.stack(out)
