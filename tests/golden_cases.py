"""Named, seeded parity cases shared by the golden-fixture generator (tests/golden/make_golden.py,
run against the unmodified reference), the oracle pin test and the GPU / simulator parity tests.

A case is fully determined by its name: parameters + seed -> input limbs and synthetic keys.
`run_case(name, backend)` returns {output name: sha256 of the output limbs}.  Backends expose the
same six operations over single ciphertexts [size][L][N] (device backends batch them internally).
"""
import numpy as np

from common import digest, mixed_moduli, ntt_primes, rand_key, rand_poly

# CoeffModulus::BFVDefault(N) of the reference (util/globals.cpp:22-160), as returned by
# oracle/_ref (ref_default_moduli) -- what native/examples/7_performance.cpp runs on.
BFV_DEFAULT = {
    4096: [68719403009, 68719230977, 137438822401],
    8192: [8796092858369, 8796092792833, 17592186028033, 17592185438209, 17592184717313],
    16384: [281474976546817, 281474976317441, 281474975662081, 562949952798721, 562949952700417,
            562949952274433, 562949951979521, 562949951881217, 562949951619073],
    32768: [36028797017456641, 36028797014704129, 36028797014573057, 36028797014376449,
            36028797013327873, 36028797013000193, 36028797012606977, 36028797010444289,
            36028797009985537, 36028797005856769, 36028797005529089, 36028797005135873,
            36028797003694081, 36028797003563009, 36028797001138177, 72057594037338113],
}

# name -> (logn, moduli spec, k special, level L, batch, rotation step)
EVAL = {
    "c1_n4096_default": (12, BFV_DEFAULT[4096], 1, 2, 2, 1),
    "n1024_k1_L3": (10, [40, 40, 40, 41], 1, 3, 2, 1),
    "n2048_k2_L4": (11, [50, 40, 45, 45, 50, 51], 2, 4, 2, -1),
    "n2048_k2_L3_ragged": (11, [50, 40, 45, 45, 50, 51], 2, 3, 1, 3),
    "n1024_k3_L5": (10, [45, 45, 45, 45, 45, 46, 46, 46], 3, 5, 1, 7),
    "n1024_k3_L4_ragged": (10, [45, 45, 45, 45, 45, 46, 46, 46], 3, 4, 1, 2),
    "n8192_default_k1_L4": (13, BFV_DEFAULT[8192], 1, 4, 1, 1),
    "n8192_k2_L3": (13, BFV_DEFAULT[8192], 2, 3, 1, -2),
    "c3_n16384_default": (14, BFV_DEFAULT[16384], 1, 8, 1, 1),
    "c4_n32768_default": (15, BFV_DEFAULT[32768], 1, 15, 1, 1),
    "n32768_k4_L12": (15, BFV_DEFAULT[32768], 4, 12, 1, 1),
    "n65536_k1_L2": (16, [59, 59, 59], 1, 2, 1, 1),
    "n65536_k2_L3": (16, [50, 50, 50, 51, 51], 2, 3, 1, -1),
    "n16384_k3_L5_ragged": (14, BFV_DEFAULT[16384], 3, 5, 1, 5),
}
# Batch-64 cases ("every config and batch element"): per-ciphertext digests of multiply / relinearize /
# rescale / mul_relin_rescale / rotate live in tests/golden/batch64_digests.json.  Too slow for the CPU oracle as a
# whole (the CPU suite pins two ciphertexts of each), so they are kept out of EVAL.
EVAL_B64 = {
    "c3_n16384_default_b64": (14, BFV_DEFAULT[16384], 1, 8, 64, 1),
    "c4_n32768_default_b64": (15, BFV_DEFAULT[32768], 1, 15, 64, 1),
}
# BFV branches (BASELINE config #1 = the first one): coefficient-form ciphertexts.
# name -> (logn, moduli spec, k special, level L, batch, rotation step, plain modulus)
BFV = {
    "c1_bfv_n4096_default": (12, BFV_DEFAULT[4096], 1, 2, 2, 1, 786433),  # 7_performance.cpp:543-552
    "bfv_n2048_k2_L4": (11, [50, 40, 45, 45, 50, 51], 2, 4, 1, -1, 65537),
    "bfv_n1024_k1_L3": (10, [40, 40, 40, 41], 1, 3, 2, 3, 12289),
    "bfv_n8192_default": (13, BFV_DEFAULT[8192], 1, 4, 1, 1, 786433),
    "bfv_n16384_default": (14, BFV_DEFAULT[16384], 1, 8, 1, 1, 786433),
    "bfv_n32768_default": (15, BFV_DEFAULT[32768], 1, 15, 1, 1, 786433),
    "bfv_n4096_L1": (12, BFV_DEFAULT[4096], 1, 1, 1, 1, 65537),
}

# General ciphertext sizes (evaluator.cpp:470-520, :274-445, :772-827): multiply of a size-sa by a size-sb ciphertext,
# then relinearization from size sa + sb - 1 to 2 with synthetic keys for s^2 .. s^(sa+sb-2).
# name -> (logn, moduli spec, k special, level L, batch, sa, sb, plain modulus or None for CKKS)
GEN = {
    "gen_n2048_k1_L3_3x2": (11, [50, 40, 45, 51], 1, 3, 2, 3, 2, None),
    "gen_n1024_k2_L4_2x4": (10, [45, 45, 45, 45, 46, 46], 2, 4, 1, 2, 4, None),
    "gen_n8192_L4_3x3": (13, BFV_DEFAULT[8192], 1, 4, 1, 3, 3, None),
    "gen_n32768_L15_3x2": (15, BFV_DEFAULT[32768], 1, 15, 1, 3, 2, None),
    "gen_bfv_n2048_3x2": (11, [50, 40, 45, 51], 1, 3, 2, 3, 2, 65537),
    "gen_bfv_n4096_2x3": (12, BFV_DEFAULT[4096], 1, 2, 1, 2, 3, 786433),
    "gen_bfv_n8192_3x3": (13, BFV_DEFAULT[8192], 1, 4, 1, 3, 3, 786433),
}

# (logn, bits): 3 polys x 3 limbs, all four transforms
NTT = {"ntt_n%d_b%d" % (ln, b): (ln, b) for ln in (10, 11, 12, 13, 14, 15, 16) for b in (30, 59)}
NTT["ntt_n12_b60_wrap"] = (12, 60)  # lazy forward transform wraps mod 2^64
NTT["ntt_n15_b60_wrap"] = (15, 60)

SMALL = ([n for n, c in EVAL.items() if c[0] <= 12] + [n for n, c in NTT.items() if c[0] <= 13] +
         [n for n, c in BFV.items() if c[0] <= 12] + [n for n, c in GEN.items() if c[0] <= 12])


def case_seed(name):
    return int.from_bytes(name.encode(), "little") % (1 << 31)


def eval_inputs(name):
    logn, spec, k, L, batch, step = EVAL[name] if name in EVAL else EVAL_B64[name]
    N = 1 << logn
    mods = list(spec) if spec[0] > 64 else mixed_moduli(logn, spec)
    rng = np.random.default_rng(case_seed(name))
    K = len(mods)
    nb = (K - k + k - 1) // k
    return dict(
        logn=logn, N=N, mods=mods, k=k, L=L, batch=batch, step=step,
        a=rand_poly(rng, mods[:L], (batch, 2), N), b=rand_poly(rng, mods[:L], (batch, 2), N),
        relin=rand_key(rng, mods, nb, N), galois=rand_key(rng, mods, nb, N),
        sk=rand_poly(rng, mods, (), N))  # drawn last: a synthetic NTT-form secret key at key level (row f4)


def bfv_inputs(name):
    logn, spec, k, L, batch, step, t = BFV[name]
    N = 1 << logn
    mods = list(spec) if spec[0] > 64 else mixed_moduli(logn, spec)
    rng = np.random.default_rng(case_seed(name))
    nb = (len(mods) - k + k - 1) // k
    return dict(
        logn=logn, N=N, mods=mods, k=k, L=L, batch=batch, step=step, t=t,
        a=rand_poly(rng, mods[:L], (batch, 2), N), b=rand_poly(rng, mods[:L], (batch, 2), N),
        relin=rand_key(rng, mods, nb, N), galois=rand_key(rng, mods, nb, N), conj=rand_key(rng, mods, nb, N),
        sk=rand_poly(rng, mods, (), N))


def gen_inputs(name):
    logn, spec, k, L, batch, sa, sb, t = GEN[name]
    N = 1 << logn
    mods = list(spec) if spec[0] > 64 else mixed_moduli(logn, spec)
    rng = np.random.default_rng(case_seed(name))
    nb = (len(mods) - k + k - 1) // k
    return dict(
        logn=logn, N=N, mods=mods, k=k, L=L, batch=batch, t=t,
        a=rand_poly(rng, mods[:L], (batch, sa), N), b=rand_poly(rng, mods[:L], (batch, sb), N),
        relin=[rand_key(rng, mods, nb, N) for _ in range(sa + sb - 3)])  # relin[p] = key of s^(p + 2)


def ntt_inputs(name):
    logn, bits = NTT[name]
    mods = ntt_primes(logn, bits, 3)
    rng = np.random.default_rng(case_seed(name))
    return dict(logn=logn, mods=mods, x=rand_poly(rng, mods, (3,), 1 << logn))


def run_case(name, be):
    out = {}
    if name in NTT:
        c = ntt_inputs(name)
        for kind in range(4):
            out["kind%d" % kind] = digest(be.ntt(c["logn"], c["mods"], c["x"], kind))
        return out
    if name in GEN:
        c = gen_inputs(name)
        be.setup(c["logn"], c["mods"], c["k"], bfv_t=c["t"])
        be.set_relin_powers(c["relin"])
        prod = be.multiply(c["a"], c["b"])
        out["multiply"] = digest(prod)
        out["relinearize"] = digest(be.relinearize(prod))
        return out
    if name in BFV:
        c = bfv_inputs(name)
        be.setup(c["logn"], c["mods"], c["k"], bfv_t=c["t"])
        elt, conj = be.elt_from_step(c["logn"], c["step"]), 2 * c["N"] - 1
        be.set_keys(c["relin"], elt, c["galois"], conj, c["conj"])
        c3 = be.multiply(c["a"], c["b"])
        out["bfv_multiply"] = digest(c3)
        out["bfv_relinearize"] = digest(be.relinearize(c3))
        if c["L"] >= 2:
            out["bfv_mod_switch"] = digest(be.mod_switch(c3))
        out["bfv_rotate_rows"] = digest(be.apply_galois(c["a"], elt))
        out["bfv_rotate_columns"] = digest(be.apply_galois(c["b"], conj))
        out["bfv_decrypt"] = digest(be.decrypt(c["a"], c["sk"]))    # Decryptor::bfv_decrypt, size 2
        out["bfv_decrypt3"] = digest(be.decrypt(c3, c["sk"]))       # size 3: uses s^2
        return out
    c = eval_inputs(name)
    be.setup(c["logn"], c["mods"], c["k"])
    elt = be.elt_from_step(c["logn"], c["step"])
    be.set_keys(c["relin"], elt, c["galois"])
    c3 = be.multiply(c["a"], c["b"])
    c2 = be.relinearize(c3)
    out["multiply"] = digest(c3)
    out["relinearize"] = digest(c2)
    if c["L"] >= 2:
        out["rescale"] = digest(be.rescale(c2))
        out["mul_relin_rescale"] = digest(be.mul_relin_rescale(c["a"], c["b"]))
    out["rotate"] = digest(be.apply_galois(c["a"], elt))
    out["decrypt"] = digest(be.decrypt(c["a"], c["sk"]))            # Decryptor::ckks_decrypt, size 2
    out["decrypt3"] = digest(be.decrypt(c3, c["sk"]))               # size 3: uses s^2
    return out


def run_case_per_ct(name, be, only=None):
    """Per-ciphertext digests of a batch case: {op: [sha256 of ciphertext i's output limbs]}.  `only` restricts the
    batch to the given ciphertext indices (inputs are still drawn for the whole batch, so indices keep their words)."""
    c = eval_inputs(name)
    be.setup(c["logn"], c["mods"], c["k"])
    elt = be.elt_from_step(c["logn"], c["step"])
    be.set_keys(c["relin"], elt, c["galois"])
    a, b = c["a"], c["b"]
    if only is not None:
        a, b = np.ascontiguousarray(a[list(only)]), np.ascontiguousarray(b[list(only)])
    c3 = be.multiply(a, b)
    c2 = be.relinearize(c3)
    res = {"multiply": c3, "relinearize": c2, "rescale": be.rescale(c2), "mul_relin_rescale": be.mul_relin_rescale(a, b),
           "rotate": be.apply_galois(a, elt)}
    return {op: [digest(x[i]) for i in range(x.shape[0])] for op, x in res.items()}


class _PerCt:
    """Maps batched [B][size][L][N] arrays over a per-ciphertext function."""
    @staticmethod
    def each(fn, *arrs):
        return np.ascontiguousarray(np.stack([fn(*[np.ascontiguousarray(a[i]) for a in arrs]) for i in range(arrs[0].shape[0])]))


class OracleBackend(_PerCt):
    def __init__(self, oracle):
        self.o = oracle

    def ntt(self, logn, mods, x, kind):
        return self.o.ntt_rns(logn, mods, x, kind)

    def setup(self, logn, mods, k, bfv_t=None):
        self.logn, self.mods, self.t = logn, mods, bfv_t
        self.c = self.o.context(logn, mods, k)

    def elt_from_step(self, logn, step):
        return self.o.elt_from_step(logn, step)

    def set_keys(self, relin, elt, galois, elt2=None, galois2=None):
        self.relin, self.gal = relin, {elt: galois, elt2: galois2}
        self.relin_all = np.ascontiguousarray(relin[None])

    def set_relin_powers(self, keys):
        self.relin, self.relin_all = keys[0], np.ascontiguousarray(np.stack(keys))

    def multiply(self, a, b):
        if self.t is not None:
            return self.each(lambda x, y: self.o.bfv_multiply_sizes(self.logn, self.mods, self.t, x, y), a, b)
        return self.each(self.c.multiply_sizes, a, b)

    def relinearize(self, c3):
        return self.each(lambda x: self.c.relinearize_sizes(x, self.relin_all, bfv=self.t is not None), c3)

    def rescale(self, c):
        return self.each(self.c.rescale, c)

    def mod_switch(self, c):
        return self.each(self.c.mod_switch_bfv, c)

    def mul_relin_rescale(self, a, b):
        return self.each(lambda x, y: self.c.mul_relin_rescale(x, y, self.relin), a, b)

    def apply_galois(self, c, elt):
        if self.t is not None:
            return self.each(lambda x: self.c.apply_galois_bfv(x, elt, self.gal[elt]), c)
        return self.each(lambda x: self.c.apply_galois(x, elt, self.gal[elt]), c)


    def decrypt(self, c, sk):
        if self.t is not None:
            return self.each(lambda x: self.c.bfv_decrypt(x, sk, self.t), c)
        return self.each(lambda x: self.c.dot_product_ct_sk(x, sk, True), c)


class RefBackend(_PerCt):
    """The unmodified reference (oracle/_ref); used by make_golden.py only."""

    def __init__(self, reflib):
        self.r = reflib

    def ntt(self, logn, mods, x, kind):
        out = x.copy()
        for y, q in enumerate(mods):
            out[:, y, :] = self.r.ntt(logn, q, np.ascontiguousarray(x[:, y, :]), kind)
        return out

    def setup(self, logn, mods, k, bfv_t=None):
        self.bfv = bfv_t is not None
        self.c = self.r.context(1 << logn, mods, k, bfv_plain_modulus=bfv_t)

    def decrypt(self, c, sk):
        return self.each(lambda x: self.c.decrypt(x, sk, bfv=self.bfv), c)

    def elt_from_step(self, logn, step):
        return self.c.elt_from_step(step)

    def set_keys(self, relin, elt, galois, elt2=None, galois2=None):
        self.c.keygen([elt] + ([elt2] if elt2 is not None else []))
        self.c.set_key(0, 0, relin)
        self.c.set_key(1, elt, galois)
        if elt2 is not None:
            self.c.set_key(1, elt2, galois2)

    def set_relin_powers(self, keys):
        self.c.set_relin_count(len(keys))
        self.c.keygen([])
        for p, key in enumerate(keys):
            self.c.set_key(0, p, key)

    def multiply(self, a, b):
        return self.each(self.c.multiply_sizes, a, b)

    def relinearize(self, c3):
        return self.each(self.c.relinearize_sizes, c3)

    def rescale(self, c):
        return self.each(self.c.rescale, c)

    def mod_switch(self, c):
        return self.each(self.c.mod_switch, c)

    def mul_relin_rescale(self, a, b):
        return self.each(self.c.mul_relin_rescale, a, b)

    def apply_galois(self, c, elt):
        return self.each(lambda x: self.c.apply_galois(x, elt), c)


class DeviceBackend:
    """The product path through the C ABI (host-buffer entry points); `lib` is the CUDA library on
    the GPU box or the test-only simulator build on CPU."""

    def __init__(self, lib):
        self.lib = lib
        self.ctx = None

    def _ctx(self, logn, mods, k):
        from gemini_seal_b200 import Context
        return Context(1 << logn, mods, k, lib=self.lib)

    def ntt(self, logn, mods, x, kind):
        import gemini_seal_b200 as g
        ctx = self._ctx(logn, mods, 1)
        fn = [g.ntt_negacyclic_harvey_lazy, g.ntt_negacyclic_harvey, g.inverse_ntt_negacyclic_harvey_lazy,
              g.inverse_ntt_negacyclic_harvey][kind]
        out = fn(x.copy(), ctx)
        ctx.close()
        return out

    def setup(self, logn, mods, k, bfv_t=None):
        from gemini_seal_b200 import BfvEvaluator, Evaluator
        if self.ctx is not None:
            self.ctx.close()
        self.ctx = self._ctx(logn, mods, k)
        self.ev = Evaluator(self.ctx) if bfv_t is None else BfvEvaluator(self.ctx, bfv_t)
        self.t = bfv_t

    def elt_from_step(self, logn, step):
        return self.ctx.galois_elt_from_step(step)

    def set_keys(self, relin, elt, galois, elt2=None, galois2=None):
        self.ctx.set_relin_keys(relin)
        self.ctx.set_galois_key(elt, galois)
        if elt2 is not None:
            self.ctx.set_galois_key(elt2, galois2)

    def set_relin_powers(self, keys):
        for p, key in enumerate(keys):
            self.ctx.set_relin_keys(key, key_power=p + 2)

    def multiply(self, a, b):
        return self.ev.multiply(a, b)

    def relinearize(self, c3):
        return self.ev.relinearize(c3)

    def rescale(self, c):
        return self.ev.rescale_to_next(c)

    def mod_switch(self, c):
        return self.ev.mod_switch_to_next(c)

    def mul_relin_rescale(self, a, b):
        return self.ev.mul_relin_rescale(a, b)

    def apply_galois(self, c, elt):
        return self.ev.apply_galois(c, elt)

    def decrypt(self, c, sk):
        from gemini_seal_b200 import Decryptor
        return Decryptor(self.ctx, sk, plain_modulus=self.t).decrypt(c)
