"""Parity tests proper (-m gpu): the CUDA library through the C ABI against
  * the golden digests produced by the unmodified reference (tests/golden/evaluator_digests.json),
  * the CPU oracle on the same seeded inputs (bit-exact, integer work),
  * size-independent properties at BASELINE.json's full sizes (round trips, batch independence,
    host-buffer vs device-buffer entry points)."""
import json
import os

import numpy as np
import pytest

import golden_cases
from common import ntt_primes, rand_key, rand_poly

pytestmark = pytest.mark.gpu

with open(os.path.join(os.path.dirname(__file__), "golden", "evaluator_digests.json")) as f:
    GOLDEN = json.load(f)["cases"]


@pytest.mark.parametrize("name", list(golden_cases.NTT) + list(golden_cases.EVAL) + list(golden_cases.BFV) + list(golden_cases.GEN))
def test_matches_reference_digests(cudalib, name):
    got = golden_cases.run_case(name, golden_cases.DeviceBackend(cudalib))
    assert got == GOLDEN[name]


with open(os.path.join(os.path.dirname(__file__), "golden", "batch64_digests.json")) as f:
    GOLDEN_B64 = json.load(f)["cases"]


@pytest.mark.parametrize("name", list(golden_cases.EVAL_B64))
def test_batch64_matches_reference_digests(cudalib, name):
    """SURVEY 8(d) parity gate "for every config and batch element": the C3 and C4 workloads at batch 64, every
    ciphertext of every operation against the digest the unmodified reference produced for it."""
    got = golden_cases.run_case_per_ct(name, golden_cases.DeviceBackend(cudalib))
    want = GOLDEN_B64[name]
    for op in want:
        bad = [i for i, (x, y) in enumerate(zip(got[op], want[op])) if x != y]
        assert not bad and len(got[op]) == len(want[op]) == 64, (name, op, bad)


@pytest.mark.parametrize("logn", [12, 13, 14, 15, 16])
@pytest.mark.parametrize("n_limbs", [1, 16, 48])
def test_ntt_grid_vs_oracle(cudalib, oracle, logn, n_limbs):
    """BASELINE config #2, the whole grid: N = 2^12..2^16 x {1, 16, 48} limbs of 59-bit primes (48 limbs also on the
    two-pass plans), lazy + strict, forward + inverse, device buffers and host buffers, against the oracle."""
    import torch
    import gemini_seal_b200 as g
    mods = ntt_primes(logn, 59, n_limbs)
    ctx = g.Context(1 << logn, mods + ntt_primes(logn, 58, 1), 1, lib=cudalib)
    rng = np.random.default_rng(logn * 100 + n_limbs)
    x = rand_poly(rng, mods, (2,), 1 << logn)
    fns = [g.ntt_negacyclic_harvey_lazy, g.ntt_negacyclic_harvey, g.inverse_ntt_negacyclic_harvey_lazy,
           g.inverse_ntt_negacyclic_harvey]
    for kind, fn in enumerate(fns):
        want = oracle.ntt_rns(logn, mods, x, kind)
        assert np.array_equal(fn(x.copy(), ctx), want), (logn, n_limbs, kind, "host")
        d = torch.from_numpy(x.view(np.int64)).cuda()
        assert np.array_equal(fn(d, ctx).cpu().numpy().view(np.uint64), want), (logn, n_limbs, kind, "device")
    ctx.close()


@pytest.mark.parametrize("logn", [12, 13, 15])
def test_threshold_primes_and_extreme_patterns(cudalib, oracle, logn):
    from common import check_threshold_and_patterns
    check_threshold_and_patterns(cudalib, oracle, logn)


@pytest.mark.parametrize("name", ["n2048_k2_L4", "n8192_k2_L3", "c3_n16384_default"])
def test_matches_oracle_arrays(cudalib, oracle, name):
    """Array-level comparison with the oracle (locates a mismatch, unlike a digest)."""
    c = golden_cases.eval_inputs(name)
    dev, orc = golden_cases.DeviceBackend(cudalib), golden_cases.OracleBackend(oracle)
    for be in (dev, orc):
        be.setup(c["logn"], c["mods"], c["k"])
        be.set_keys(c["relin"], be.elt_from_step(c["logn"], c["step"]), c["galois"])
    elt = orc.elt_from_step(c["logn"], c["step"])
    c3 = orc.multiply(c["a"], c["b"])
    assert np.array_equal(dev.multiply(c["a"], c["b"]), c3)
    assert np.array_equal(dev.relinearize(c3), orc.relinearize(c3))
    assert np.array_equal(dev.rescale(c3), orc.rescale(c3))
    assert np.array_equal(dev.apply_galois(c["a"], elt), orc.apply_galois(c["a"], elt))
    assert np.array_equal(dev.mul_relin_rescale(c["a"], c["b"]), orc.mul_relin_rescale(c["a"], c["b"]))


def test_ntt_sweep_limbs_vs_oracle(cudalib, oracle):
    """BASELINE config #2: N = 2^12..2^16 x up to 48 limbs of 59-bit primes, strict + lazy."""
    import gemini_seal_b200 as g
    for logn, n_limbs in [(12, 48), (13, 7), (14, 1), (15, 16), (16, 3)]:
        mods = ntt_primes(logn, 59, n_limbs)
        ctx = g.Context(1 << logn, mods + ntt_primes(logn, 58, 1), 1, lib=cudalib)
        rng = np.random.default_rng(logn)
        x = rand_poly(rng, mods, (2,), 1 << logn)
        fns = [g.ntt_negacyclic_harvey_lazy, g.ntt_negacyclic_harvey, g.inverse_ntt_negacyclic_harvey_lazy,
               g.inverse_ntt_negacyclic_harvey]
        for kind, fn in enumerate(fns):
            assert np.array_equal(fn(x.copy(), ctx), oracle.ntt_rns(logn, mods, x, kind)), (logn, kind)
        # limb_offset: transform limbs 1.. only
        if n_limbs > 2:
            y = np.ascontiguousarray(x[:, 1:, :])
            assert np.array_equal(g.ntt_negacyclic_harvey(y.copy(), ctx, limb_offset=1), oracle.ntt_rns(logn, mods, y, 1, limb_offset=1))
        ctx.close()


def test_poly_ops_vs_oracle(cudalib, oracle):
    import gemini_seal_b200 as g
    logn = 12
    mods = ntt_primes(logn, 59, 2) + ntt_primes(logn, 36, 1)
    ctx = g.Context(1 << logn, mods, 1, lib=cudalib)
    rng = np.random.default_rng(3)
    a, b = rand_poly(rng, mods, (3,), 1 << logn), rand_poly(rng, mods, (3,), 1 << logn)
    a[0, :, 0], b[0, :, 0] = 0, 0
    scal = [int(rng.integers(0, q)) for q in mods]
    for name, op in g.evaluator.OPS.items():
        got = g.poly_op(ctx, name, a, b if op < 3 else None, scal if op == 4 else None)
        for i, q in enumerate(mods):
            want = oracle.poly_op(op, q, a[:, i, :], b[:, i, :], scal[i])
            assert np.array_equal(got[:, i, :], want), (name, i)


def test_full_size_properties(cudalib):
    """N = 2^15, K = 16 (BASELINE config #4/#5) on device buffers, batch 8: forward/inverse round trip,
    batch independence (each ciphertext of a batch equals the same ciphertext run alone), and
    device-pointer entry points == host-buffer entry points."""
    import torch
    import gemini_seal_b200 as g
    mods = golden_cases.BFV_DEFAULT[32768]
    N, L, B = 32768, 15, 8
    ctx = g.Context(N, mods, 1, lib=cudalib)
    ev = g.Evaluator(ctx)
    rng = np.random.default_rng(11)
    ctx.set_relin_keys(rand_key(rng, mods, 15, N))
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, rand_key(rng, mods, 15, N))
    a, b = rand_poly(rng, mods[:L], (B, 2), N), rand_poly(rng, mods[:L], (B, 2), N)
    da = torch.from_numpy(a.view(np.int64)).cuda()
    db = torch.from_numpy(b.view(np.int64)).cuda()
    # round trip on the device
    w = da.clone()
    g.ntt_negacyclic_harvey(g.inverse_ntt_negacyclic_harvey(w, ctx), ctx)
    assert torch.equal(w, da)
    # batched device call vs single-ciphertext host calls
    out = ev.mul_relin_rescale(da, db)
    rot = ev.rotate_vector(da, 1)
    torch.cuda.synchronize()
    out, rot = out.cpu().numpy().view(np.uint64), rot.cpu().numpy().view(np.uint64)
    for i in (0, B - 1):
        assert np.array_equal(ev.mul_relin_rescale(a[i:i + 1], b[i:i + 1])[0], out[i])
        assert np.array_equal(ev.apply_galois(a[i:i + 1], elt)[0], rot[i])
    # composition: fused pipeline == multiply -> relinearize -> rescale
    c2 = ev.rescale_to_next(ev.relinearize(ev.multiply(da, db)))
    assert np.array_equal(c2.cpu().numpy().view(np.uint64), out)
    # rotate by +1 then -1 is NOT the identity on raw limbs with synthetic keys, but conjugating
    # twice permutes back: check the permutation part (component-wise gather) is an involution
    t = ctx.galois_table(2 * N - 1)
    assert np.array_equal(t[t], np.arange(N, dtype=np.uint32))
    ctx.close()


def test_full_size_chunked_batch(cudalib, oracle):
    """A batch larger than one 16 GiB scratch chunk (70 ciphertext pairs at N = 2^15, K = 16 -> two balanced chunks of
    35): elements on both sides of the chunk boundary equal the same ciphertexts run alone, and the first and last
    equal the oracle.  Also the rotation path (its own chunking) and a batch that does not divide evenly."""
    import torch
    import gemini_seal_b200 as g
    import synth
    mods = golden_cases.BFV_DEFAULT[32768]
    N, L, B = 32768, 15, 70
    ctx = g.Context(N, mods, 1, lib=cudalib)
    ev = g.Evaluator(ctx)
    relin, gal = synth.key_numpy(11, mods, 15, N), synth.key_numpy(12, mods, 15, N)
    ctx.set_relin_keys(relin)
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, gal)
    a = synth.residues_torch(torch, 21, 0, B, 2, mods[:L], N)
    b = synth.residues_torch(torch, 22, 0, B, 2, mods[:L], N)
    out = ev.mul_relin_rescale(a, b)
    rot = ev.apply_galois(a, elt)
    torch.cuda.synchronize()
    for i in (0, 34, 35, 69):
        assert torch.equal(ev.mul_relin_rescale(a[i:i + 1].contiguous(), b[i:i + 1].contiguous())[0], out[i]), i
        assert torch.equal(ev.apply_galois(a[i:i + 1].contiguous(), elt)[0], rot[i]), i
    oc = oracle.context(15, mods, 1)
    for i in (0, 69):
        ai, bi = a[i].cpu().numpy().view(np.uint64), b[i].cpu().numpy().view(np.uint64)
        assert np.array_equal(out[i].cpu().numpy().view(np.uint64), oc.mul_relin_rescale(ai, bi, relin)), i
        assert np.array_equal(rot[i].cpu().numpy().view(np.uint64), oc.apply_galois(ai, elt, gal)), i
    ctx.close()


@pytest.mark.parametrize("logn,n_limbs", [(12, 48), (13, 30)])
def test_many_limbs_literal_path(cudalib, oracle, logn, n_limbs):
    """Maximum-size modulus chains: 48 (single-pass N) and 30 (two-pass N) ciphertext limbs of 59-bit primes plus one
    special prime -- beta = 48 / 30 bundles, the literal butterflies (59-bit moduli fail the wide-range bound), every
    evaluator operation against the oracle at the first level and at a ragged lower level."""
    import gemini_seal_b200 as g
    N = 1 << logn
    mods = ntt_primes(logn, 59, n_limbs + 1)
    ctx = g.Context(N, mods, 1, lib=cudalib)
    assert cudalib.gsb_ctx_fast_ntt(ctx.handle) == 0
    oc = oracle.context(logn, mods, 1)
    rng = np.random.default_rng(logn * 7 + n_limbs)
    relin, gal = rand_key(rng, mods, n_limbs, N), rand_key(rng, mods, n_limbs, N)
    ctx.set_relin_keys(relin)
    elt = ctx.galois_elt_from_step(-1)
    ctx.set_galois_key(elt, gal)
    ev = g.Evaluator(ctx)
    for L in (n_limbs, n_limbs - 3):
        a, b = rand_poly(rng, mods[:L], (2, 2), N), rand_poly(rng, mods[:L], (2, 2), N)
        c3 = ev.multiply(a, b)
        c2, rs, rot, mrr = ev.relinearize(c3), ev.rescale_to_next(c3), ev.apply_galois(a, elt), ev.mul_relin_rescale(a, b)
        for i in range(2):
            w3 = oc.multiply(a[i], b[i])
            assert np.array_equal(c3[i], w3)
            assert np.array_equal(c2[i], oc.relinearize(w3, relin)), ("relinearize", L)
            assert np.array_equal(rs[i], oc.rescale(w3)), ("rescale", L)
            assert np.array_equal(rot[i], oc.apply_galois(a[i], elt, gal)), ("rotate", L)
            assert np.array_equal(mrr[i], oc.mul_relin_rescale(a[i], b[i], relin)), ("mrr", L)
    ctx.close()


def test_full_size_shortcuts_equal_literal(cudalib):
    """N = 2^15, K = 16: the context with every congruent-word shortcut (wide-range butterflies, skipped
    range reductions, single-high-product Barrett, ...) and a GSB_LITERAL=1 context that carries the
    reference's own intermediate words must produce identical output words."""
    import gemini_seal_b200 as g
    mods = golden_cases.BFV_DEFAULT[32768]
    N, L, B = 32768, 15, 3
    rng = np.random.default_rng(23)
    relin, gal = rand_key(rng, mods, 15, N), rand_key(rng, mods, 15, N)
    a, b = rand_poly(rng, mods[:L], (B, 2), N), rand_poly(rng, mods[:L], (B, 2), N)
    outs = []
    for literal in (False, True):
        if literal:
            os.environ["GSB_LITERAL"] = "1"
        try:
            ctx = g.Context(N, mods, 1, lib=cudalib)
        finally:
            os.environ.pop("GSB_LITERAL", None)
        assert cudalib.gsb_ctx_fast_ntt(ctx.handle) == (0 if literal else 1)
        # the headline configuration runs every shortcut (a bound that silently stops holding would cost speed, not parity)
        assert cudalib.gsb_ctx_shortcuts(ctx.handle) == (0 if literal else 31)
        ctx.set_relin_keys(relin)
        elt = ctx.galois_elt_from_step(1)
        ctx.set_galois_key(elt, gal)
        ev = g.Evaluator(ctx)
        c3 = ev.multiply(a, b)
        outs.append((c3, ev.relinearize(c3), ev.mul_relin_rescale(a, b), ev.apply_galois(a, elt)))
        ctx.close()
    for x, y in zip(*outs):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("B", [3, 8])
def test_full_size_repeatability(cudalib, B):
    """Race detector for the staged (bulk-copy) key-switch kernel: the same N = 2^15 key switch run 60
    times must give the same words every time (small batches are where a missing proxy fence once
    showed up, in ~10 % of runs; the values themselves are pinned by the golden digests)."""
    import torch
    import gemini_seal_b200 as g
    mods = golden_cases.BFV_DEFAULT[32768]
    N, L = 32768, 15
    ctx = g.Context(N, mods, 1, lib=cudalib)
    ev = g.Evaluator(ctx)
    rng = np.random.default_rng(5)
    ctx.set_relin_keys(rand_key(rng, mods, 15, N))
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, rand_key(rng, mods, 15, N))
    c3 = torch.from_numpy(rand_poly(rng, mods[:L], (B, 3), N).view(np.int64)).cuda()
    first_relin, first_rot = ev.relinearize(c3).clone(), ev.apply_galois(c3[:, :2].contiguous(), elt).clone()
    torch.cuda.synchronize()
    for _ in range(60):
        assert torch.equal(ev.relinearize(c3), first_relin)
        assert torch.equal(ev.apply_galois(c3[:, :2].contiguous(), elt), first_rot)
    ctx.close()


def test_thin_evaluator_ops(cudalib, oracle):
    """Row f3 on host buffers (N = 2^13, two-pass NTT) and on device buffers (N = 2^12)."""
    import torch
    from common import check_thin_ops
    check_thin_ops(cudalib, oracle, logn=13)
    check_thin_ops(cudalib, oracle, logn=12, to_device=lambda x: torch.from_numpy(x.view(np.int64)).cuda())


def test_decryptor(cudalib, oracle):
    """Row f4 on host buffers (N = 2^13, two-pass NTT) and on device buffers (N = 2^11)."""
    import torch
    from common import check_decrypt
    check_decrypt(cudalib, oracle, logn=13)
    check_decrypt(cudalib, oracle, logn=11, to_device=lambda x: torch.from_numpy(x.view(np.int64)).cuda())


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fuzz_against_oracle(cudalib, oracle, seed):
    """Random parameter sets (k = 1..3, ragged bundles, 30..60-bit primes, CKKS + BFV) vs the oracle."""
    from common import fuzz_against_oracle
    fuzz_against_oracle(cudalib, oracle, seed=seed, n_configs=12, max_logn=14)


@pytest.mark.parametrize("name", ["keygen_n1024_k1", "keygen_n2048_k2_ragged"])
def test_keygen_reproduces_reference_keys(cudalib, name):
    """Row f2: the buffers the reference's KeyGenerator sampled in -> its relin and Galois keys out, word for word
    (host buffers and device buffers); installed keys serve key switching like uploaded ones."""
    import torch
    from common import check_keygen_fixture, load_keygen_fixture
    fx = load_keygen_fixture(name)
    check_keygen_fixture(cudalib, fx)
    check_keygen_fixture(cudalib, fx, to_device=lambda x: torch.from_numpy(x.view(np.int64)).cuda())


@pytest.mark.parametrize("tag", ["c3", "c4"])
def test_keygen_full_size_vs_oracle(cudalib, oracle, tag):
    """Row f2 at the C3 / C4 shapes: a relin key and two Galois keys from synthetic sampled buffers, device (one batched
    call) against the oracle."""
    import gemini_seal_b200 as g
    from common import oracle_keygen
    n = {"c3": 16384, "c4": 32768}[tag]
    mods = golden_cases.BFV_DEFAULT[n]
    logn, K = n.bit_length() - 1, len(mods)
    beta = K - 1
    rng = np.random.default_rng(n + 2)
    sk = rand_poly(rng, mods, (), n)
    ctx = g.Context(n, mods, 1, lib=cudalib)
    kg = g.KeyGenerator(ctx, sk)
    elts = [ctx.galois_elt_from_step(1), 2 * n - 1]
    a = rand_poly(rng, mods, (3, beta), n)
    e = rng.integers(-19, 20, size=(3, beta, n), dtype=np.int64)
    got_r = kg.relin_keys(a[:1], e[:1])
    got_g = kg.galois_keys(elts, a[1:], e[1:])
    assert np.array_equal(got_r, oracle_keygen(oracle, None, logn, mods, 1, sk, [2], False, a[:1], e[:1]))
    assert np.array_equal(got_g, oracle_keygen(oracle, None, logn, mods, 1, sk, elts, True, a[1:], e[1:]))
    ctx.close()


def test_error_convention(cudalib):
    """Argument errors -> ValueError (std::invalid_argument), evaluator.cpp:238-249,781-796,1872-1887."""
    import gemini_seal_b200 as g
    mods = ntt_primes(10, 40, 4)
    ctx = g.Context(1024, mods, 1, lib=cudalib)
    ev = g.Evaluator(ctx)
    rng = np.random.default_rng(0)
    a = rand_poly(rng, mods[:3], (1, 2), 1024)
    c3 = rand_poly(rng, mods[:3], (1, 3), 1024)
    with pytest.raises(ValueError):
        ev.relinearize(c3)  # no relin keys uploaded
    with pytest.raises(ValueError):
        ev.apply_galois(a, 3)  # no Galois key
    with pytest.raises(ValueError):
        ev.apply_galois(a, 4)  # even element
    with pytest.raises(ValueError):
        ctx.galois_elt_from_step(512)
    with pytest.raises(ValueError):
        ctx.set_relin_keys(np.zeros((2, 2, 4, 1024), dtype=np.uint64))  # wrong bundle count
    with pytest.raises(ValueError):
        ev.rescale_to_next(rand_poly(rng, mods[:1], (1, 2), 1024))  # last level
    with pytest.raises(ValueError):
        ev.multiply(a, rand_poly(rng, mods[:2], (1, 2), 1024))
    with pytest.raises(ValueError):
        g.Context(1024, [mods[0], mods[0]], 1, lib=cudalib)
    with pytest.raises(ValueError):
        g.Context(1024, [97, 193], 1, lib=cudalib)  # 97 is not 1 mod 2N
    ctx.close()


def test_concurrent_host_threads(cudalib):
    """Thread-safety contract of include/gseal_b200.h: one context shared by four host threads, each with
    its own stream and workspace, running CKKS and BFV key switches at the same time (ctypes drops the
    GIL inside the calls), must give exactly the words the same calls give one after another."""
    import threading
    import torch
    import gemini_seal_b200 as g
    mods = golden_cases.BFV_DEFAULT[8192]
    N, L, B, t = 8192, len(mods) - 1, 6, 65537
    ctx = g.Context(N, mods, 1, lib=cudalib)
    rng = np.random.default_rng(21)
    ctx.set_relin_keys(rand_key(rng, mods, L, N))
    elt = ctx.galois_elt_from_step(3)
    ctx.set_galois_key(elt, rand_key(rng, mods, L, N))
    dev = lambda x: torch.from_numpy(x.view(np.int64)).cuda()
    ins = [(dev(rand_poly(rng, mods[:L], (B, 2), N)), dev(rand_poly(rng, mods[:L], (B, 2), N))) for _ in range(4)]

    def work(i, ev, bfv):
        a, b = ins[i]
        return [ev.mul_relin_rescale(a, b), ev.apply_galois(a, elt), bfv.relinearize(bfv.multiply(a, b)),
                bfv.apply_galois(b, elt)]

    ev0, bfv0 = g.Evaluator(ctx), g.BfvEvaluator(ctx, t)
    want = [work(i, ev0, bfv0) for i in range(4)]
    torch.cuda.synchronize()
    got, errs = [None] * 4, []

    def thread(i):
        try:
            ws, stream = g.Workspace(ctx), torch.cuda.Stream()
            ev, bfv = g.Evaluator(ctx, ws), g.BfvEvaluator(ctx, t, ws)
            with torch.cuda.stream(stream):
                for _ in range(10):
                    got[i] = work(i, ev, bfv)
            stream.synchronize()
            ws.close()
        except Exception as e:  # surfaced below
            errs.append(e)

    ths = [threading.Thread(target=thread, args=(i,)) for i in range(4)]
    [th.start() for th in ths]
    [th.join() for th in ths]
    assert not errs, errs
    for i in range(4):
        for x, y in zip(got[i], want[i]):
            assert torch.equal(x, y)
    ctx.close()
