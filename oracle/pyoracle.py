"""TEST INFRASTRUCTURE -- ctypes wrappers around the CPU oracles.

  Oracle  : oracle/liboracle.so, the plain-C restatement (seal_oracle.c).
  RefLib  : oracle/_ref/libref_harness.so, the UNMODIFIED reference compiled from /root/reference.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_harness.so")
vp, sz, i32, u32, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64


def build_oracle():
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    return ORACLE_SO


def build_ref(jobs=8):
    """Only possible where /root/reference exists (this container)."""
    subprocess.run(["make", "-s", "-j%d" % jobs, "-C", HERE, "ref"], check=True)
    return REF_SO


def _u64arr(vals):
    return (u64 * len(vals))(*[int(v) for v in vals])


def _p(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


class Oracle:
    def __init__(self, path=None):
        p = path or ORACLE_SO
        if not os.path.exists(p):
            build_oracle()
        lib = self.lib = C.CDLL(p)
        lib.so_ntt.argtypes = [i32, u64, vp, sz, i32]
        lib.so_poly_op.argtypes = [i32, u64, vp, vp, u64, sz, vp]
        lib.so_galois_table_ntt.argtypes = [i32, u32, vp]
        lib.so_elt_from_step.argtypes = [i32, i32]
        lib.so_elt_from_step.restype = u32
        lib.so_apply_galois_ntt.argtypes = [i32, u32, vp, vp]
        lib.so_ctx_create.argtypes = [i32, C.POINTER(u64), sz, sz]
        lib.so_ctx_create.restype = vp
        lib.so_ctx_destroy.argtypes = [vp]
        lib.so_ckks_multiply.argtypes = [vp, sz, vp, vp, vp]
        lib.so_ckks_multiply_sizes.argtypes = [vp, sz, vp, sz, vp, sz, vp]
        lib.so_kswitch_keygen.argtypes = [vp, vp, vp, vp, vp, vp]
        lib.so_relinearize_sizes.argtypes = [vp, sz, sz, vp, vp, sz, i32, vp]
        lib.so_bfv_multiply_sizes.argtypes = [i32, C.POINTER(u64), sz, u64, vp, sz, vp, sz, vp]
        lib.so_switch_key.argtypes = [vp, sz, vp, vp, vp]
        lib.so_relinearize.argtypes = [vp, sz, vp, vp, vp]
        lib.so_rescale.argtypes = [vp, sz, sz, vp, vp]
        lib.so_apply_galois.argtypes = [vp, sz, u32, vp, vp, vp]
        lib.so_mul_relin_rescale.argtypes = [vp, sz, vp, vp, vp, vp]
        lib.so_relinearize_bfv.argtypes = [vp, sz, vp, vp, vp]
        lib.so_bfv_mod_switch.argtypes = [vp, sz, sz, vp, vp]
        lib.so_bfv_multiply.argtypes = [i32, C.POINTER(u64), sz, u64, vp, vp, vp]
        lib.so_bfv_aux_base.argtypes = [i32, C.POINTER(u64), sz, u64, vp]
        lib.so_apply_galois_bfv.argtypes = [vp, sz, u32, vp, vp, vp]
        lib.so_apply_galois_coeff.argtypes = [i32, u32, u64, vp, vp]
        lib.so_dot_product_ct_sk.argtypes = [vp, sz, sz, vp, vp, i32, vp]
        lib.so_bfv_decrypt.argtypes = [vp, sz, sz, u64, vp, vp, vp]
        lib.so_const_ratio.argtypes = [u64, C.POINTER(u64)]
        lib.so_barrett_reduce_128.argtypes = [u64, u64, u64, C.POINTER(u64)]
        lib.so_barrett_reduce_128.restype = u64
        lib.so_barrett_reduce_63.argtypes = [u64, u64, C.POINTER(u64)]
        lib.so_barrett_reduce_63.restype = u64
        lib.so_multiply_uint_mod.argtypes = [u64, u64, u64, C.POINTER(u64)]
        lib.so_multiply_uint_mod.restype = u64
        lib.so_multiply_add_uint_mod.argtypes = [u64, u64, u64, u64, C.POINTER(u64)]
        lib.so_multiply_add_uint_mod.restype = u64
        lib.so_tables_init.argtypes = [vp, i32, u64]

    # -- word arithmetic
    def const_ratio(self, q):
        cr = (u64 * 3)()
        self.lib.so_const_ratio(q, cr)
        return cr

    def barrett_reduce_128(self, lo, hi, q):
        return self.lib.so_barrett_reduce_128(lo, hi, q, self.const_ratio(q))

    def barrett_reduce_63(self, x, q):
        return self.lib.so_barrett_reduce_63(x, q, self.const_ratio(q))

    def multiply_uint_mod(self, a, b, q):
        return self.lib.so_multiply_uint_mod(a, b, q, self.const_ratio(q))

    def multiply_add_uint_mod(self, a, b, c, q):
        return self.lib.so_multiply_add_uint_mod(a, b, c, q, self.const_ratio(q))

    # -- NTT: kind 0 fwd lazy, 1 fwd strict, 2 inv lazy, 3 inv strict; data [..., N] under one modulus
    def ntt(self, logn, q, data, kind):
        out = np.ascontiguousarray(data, dtype=np.uint64).copy()
        rc = self.lib.so_ntt(logn, q, _p(out), out.size >> logn, kind)
        if rc:
            raise ValueError("oracle: invalid modulus for NTT")
        return out

    def ntt_rns(self, logn, moduli, data, kind, limb_offset=0):
        """data [..., n_limbs, N]; limb y uses moduli[limb_offset + y]."""
        out = np.ascontiguousarray(data, dtype=np.uint64).copy()
        flat = out.reshape(-1, out.shape[-2], out.shape[-1])
        for y in range(flat.shape[1]):
            col = np.ascontiguousarray(flat[:, y, :])
            rc = self.lib.so_ntt(logn, moduli[limb_offset + y], _p(col), col.shape[0], kind)
            assert rc == 0
            flat[:, y, :] = col
        return out

    def poly_op(self, op, q, a, b=None, scalar=0):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.empty_like(a)
        bb = np.ascontiguousarray(b, dtype=np.uint64) if b is not None else a
        assert self.lib.so_poly_op(op, q, _p(a), _p(bb), scalar, a.size, _p(out)) == 0
        return out

    def bfv_multiply(self, logn, moduli, plain_modulus, a, b):
        """bfv_multiply of two size-2 coefficient-form ciphertexts [2][L][N] -> [3][L][N]."""
        L = a.shape[1]
        out = np.empty((3, L, 1 << logn), dtype=np.uint64)
        assert self.lib.so_bfv_multiply(logn, _u64arr(moduli[:L]), L, plain_modulus, _p(a), _p(b), _p(out)) == 0
        return out

    def bfv_multiply_sizes(self, logn, moduli, plain_modulus, a, b):
        """bfv_multiply of coefficient-form ciphertexts of any sizes [sa][L][N] x [sb][L][N] -> [sa+sb-1][L][N]."""
        L, sa, sb = a.shape[1], a.shape[0], b.shape[0]
        out = np.empty((sa + sb - 1, L, 1 << logn), dtype=np.uint64)
        assert self.lib.so_bfv_multiply_sizes(logn, _u64arr(moduli[:L]), L, plain_modulus, _p(a), sa, _p(b), sb, _p(out)) == 0
        return out

    def bfv_aux_base(self, logn, moduli, plain_modulus):
        out = np.zeros(64, dtype=np.uint64)
        assert self.lib.so_bfv_aux_base(logn, _u64arr(moduli), len(moduli), plain_modulus, _p(out)) == 0
        nb = int(out[0])
        return nb, [int(x) for x in out[1:nb + 2]]

    def galois_table_ntt(self, logn, elt):
        t = np.empty(1 << logn, dtype=np.uint32)
        self.lib.so_galois_table_ntt(logn, elt, t.ctypes.data)
        return t

    def elt_from_step(self, logn, step):
        return self.lib.so_elt_from_step(logn, step)

    def context(self, logn, moduli, k_special):
        return OracleCtx(self, logn, moduli, k_special)


class OracleCtx:
    """so_ctx + the CKKS evaluator ops of the restatement; buffers are single ciphertexts."""

    def __init__(self, orc, logn, moduli, k_special):
        self.lib = orc.lib
        self.logn, self.N, self.moduli, self.k = logn, 1 << logn, list(moduli), k_special
        self.K = len(moduli)
        self.h = self.lib.so_ctx_create(logn, _u64arr(moduli), self.K, k_special)
        if not self.h:
            raise ValueError("oracle: invalid parameters")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.so_ctx_destroy(self.h)
            self.h = None

    def multiply(self, a, b):
        L = a.shape[1]
        out = np.empty((3, L, self.N), dtype=np.uint64)
        assert self.lib.so_ckks_multiply(self.h, L, _p(a), _p(b), _p(out)) == 0
        return out

    def relinearize(self, c3, key):
        L = c3.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        assert self.lib.so_relinearize(self.h, L, _p(c3), _p(key), _p(out)) == 0
        return out

    def kswitch_keygen(self, sk, new_key, a, e):
        """generate_one_kswitch_key from the sampled a [beta][K][N] (NTT form) and e [beta][N] (int64 noise)."""
        a = np.ascontiguousarray(a, dtype=np.uint64)
        e = np.ascontiguousarray(e, dtype=np.int64)
        out = np.empty((a.shape[0], 2, self.K, self.N), dtype=np.uint64)
        assert self.lib.so_kswitch_keygen(self.h, _p(np.ascontiguousarray(sk)), _p(np.ascontiguousarray(new_key)), _p(a), e.ctypes.data, _p(out)) == 0
        return out

    def multiply_sizes(self, a, b):
        L, sa, sb = a.shape[1], a.shape[0], b.shape[0]
        out = np.empty((sa + sb - 1, L, self.N), dtype=np.uint64)
        assert self.lib.so_ckks_multiply_sizes(self.h, L, _p(a), sa, _p(b), sb, _p(out)) == 0
        return out

    def relinearize_sizes(self, c, keys, bfv=False):
        """c = [size][L][N]; keys = [size - 2][n_bundles][2][K][N], keys[p] = RelinKeys::get_index(p + 2)."""
        size, L = c.shape[0], c.shape[1]
        keys = np.ascontiguousarray(keys)
        out = np.empty((2, L, self.N), dtype=np.uint64)
        assert self.lib.so_relinearize_sizes(self.h, L, size, _p(c), _p(keys), keys[0].size, 0 if bfv else 1, _p(out)) == 0
        return out

    def rescale(self, c):
        size, L = c.shape[0], c.shape[1]
        out = np.empty((size, L - 1, self.N), dtype=np.uint64)
        assert self.lib.so_rescale(self.h, L, size, _p(c), _p(out)) == 0
        return out

    def apply_galois(self, c, elt, key):
        L = c.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        assert self.lib.so_apply_galois(self.h, L, elt, _p(c), _p(key), _p(out)) == 0
        return out

    def mul_relin_rescale(self, a, b, key):
        L = a.shape[1]
        out = np.empty((2, L - 1, self.N), dtype=np.uint64)
        assert self.lib.so_mul_relin_rescale(self.h, L, _p(a), _p(b), _p(key), _p(out)) == 0
        return out

    # -- BFV branches: coefficient-form ciphertexts
    def relinearize_bfv(self, c3, key):
        L = c3.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        assert self.lib.so_relinearize_bfv(self.h, L, _p(c3), _p(key), _p(out)) == 0
        return out

    def mod_switch_bfv(self, c):
        size, L = c.shape[0], c.shape[1]
        out = np.empty((size, L - 1, self.N), dtype=np.uint64)
        assert self.lib.so_bfv_mod_switch(self.h, L, size, _p(c), _p(out)) == 0
        return out

    def apply_galois_bfv(self, c, elt, key):
        L = c.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        assert self.lib.so_apply_galois_bfv(self.h, L, elt, _p(c), _p(key), _p(out)) == 0
        return out


    # -- Decryptor (row f4): ct [size][L][N], sk [K][N] NTT form at key level
    def dot_product_ct_sk(self, ct, sk, ntt_form):
        size, L = ct.shape[0], ct.shape[1]
        out = np.empty((L, self.N), dtype=np.uint64)
        assert self.lib.so_dot_product_ct_sk(self.h, L, size, _p(ct), _p(sk), 1 if ntt_form else 0, _p(out)) == 0
        return out

    def bfv_decrypt(self, ct, sk, plain_modulus):
        size, L = ct.shape[0], ct.shape[1]
        out = np.empty((self.N,), dtype=np.uint64)
        assert self.lib.so_bfv_decrypt(self.h, L, size, plain_modulus, _p(ct), _p(sk), _p(out)) == 0
        return out


class RefLib:
    """The unmodified reference behind oracle/ref_harness.cpp.  Raises FileNotFoundError when the
    harness has not been built (it can only be built where /root/reference exists)."""

    def __init__(self, path=None):
        p = path or REF_SO
        if not os.path.exists(p):
            raise FileNotFoundError(p)
        lib = self.lib = C.CDLL(p)
        lib.ref_last_error.restype = C.c_char_p
        lib.ref_ntt.argtypes = [i32, u64, vp, sz, i32]
        lib.ref_ntt_tables.argtypes = [i32, u64, vp, vp, vp, vp, vp]
        lib.ref_const_ratio.argtypes = [u64, vp]
        lib.ref_poly_op.argtypes = [i32, u64, vp, vp, u64, sz, vp]
        lib.ref_default_moduli.argtypes = [sz, vp, sz]
        lib.ref_create_moduli.argtypes = [sz, C.POINTER(i32), sz, vp]
        lib.ref_ckks_create.argtypes = [sz, C.POINTER(u64), sz, sz]
        lib.ref_ckks_create.restype = vp
        lib.ref_bfv_create.argtypes = [sz, C.POINTER(u64), sz, sz, u64]
        lib.ref_bfv_create.restype = vp
        lib.ref_ckks_destroy.argtypes = [vp]
        lib.ref_keygen.argtypes = [vp, C.POINTER(u32), sz]
        lib.ref_make_synthetic_keys.argtypes = [vp, C.POINTER(u32), sz, u64]
        lib.ref_key_bundles.argtypes = [vp, i32, u32]
        lib.ref_get_key.argtypes = [vp, i32, u32, vp]
        lib.ref_set_key.argtypes = [vp, i32, u32, vp]
        lib.ref_multiply.argtypes = [vp, sz, vp, vp, vp]
        lib.ref_set_relin_count.argtypes = [vp, sz]
        lib.ref_get_secret_key.argtypes = [vp, vp]
        lib.ref_multiply_sizes.argtypes = [vp, sz, vp, sz, vp, sz, vp]
        lib.ref_relinearize_sizes.argtypes = [vp, sz, sz, vp, vp]
        lib.ref_relinearize.argtypes = [vp, sz, vp, vp]
        lib.ref_rescale.argtypes = [vp, sz, sz, vp, vp]
        lib.ref_apply_galois.argtypes = [vp, sz, u32, vp, vp]
        lib.ref_mod_switch.argtypes = [vp, sz, sz, vp, vp]
        lib.ref_rotate.argtypes = [vp, sz, i32, vp, vp]
        lib.ref_mul_relin_rescale.argtypes = [vp, sz, vp, vp, vp]
        lib.ref_elt_from_step.argtypes = [vp, i32]
        lib.ref_elt_from_step.restype = u32
        lib.ref_time_ops.argtypes = [vp, i32, sz, sz, i32, u64]
        lib.ref_time_ops.restype = C.c_double
        lib.ref_decrypt.argtypes = [vp, vp, sz, sz, vp, vp]

    def err(self):
        return (self.lib.ref_last_error() or b"").decode()

    def ntt(self, logn, q, data, kind):
        out = np.ascontiguousarray(data, dtype=np.uint64).copy()
        rc = self.lib.ref_ntt(logn, q, _p(out), out.size >> logn, kind)
        if rc:
            raise ValueError(self.err())
        return out

    def ntt_tables(self, logn, q):
        n = 1 << logn
        arrs = [np.empty(n, dtype=np.uint64) for _ in range(4)]
        sc = np.empty(5, dtype=np.uint64)
        rc = self.lib.ref_ntt_tables(logn, q, *[_p(a) for a in arrs], _p(sc))
        if rc:
            raise ValueError(self.err())
        return (*arrs, sc)

    def const_ratio(self, q):
        out = np.empty(3, dtype=np.uint64)
        assert self.lib.ref_const_ratio(q, _p(out)) == 0
        return out

    def poly_op(self, op, q, a, b=None, scalar=0):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.empty_like(a)
        bb = np.ascontiguousarray(b, dtype=np.uint64) if b is not None else a
        assert self.lib.ref_poly_op(op, q, _p(a), _p(bb), scalar, a.size, _p(out)) == 0, self.err()
        return out

    def default_moduli(self, N):
        out = np.empty(64, dtype=np.uint64)
        n = self.lib.ref_default_moduli(N, _p(out), 64)
        assert n > 0, self.err()
        return [int(x) for x in out[:n]]

    def create_moduli(self, N, bits):
        out = np.empty(len(bits), dtype=np.uint64)
        arr = (i32 * len(bits))(*bits)
        assert self.lib.ref_create_moduli(N, arr, len(bits), _p(out)) == 0, self.err()
        return [int(x) for x in out]

    def context(self, N, moduli, k_special, bfv_plain_modulus=None):
        """CKKS context, or a BFV one (coefficient-form ciphertexts) when a plain modulus is given."""
        return RefCtx(self, N, moduli, k_special, bfv_plain_modulus)


class RefCtx:
    def __init__(self, ref, N, moduli, k_special, bfv_plain_modulus=None):
        self.ref, self.lib = ref, ref.lib
        self.N, self.moduli, self.k, self.K = N, list(moduli), k_special, len(moduli)
        if bfv_plain_modulus is None:
            self.h = self.lib.ref_ckks_create(N, _u64arr(moduli), self.K, k_special)
        else:
            self.h = self.lib.ref_bfv_create(N, _u64arr(moduli), self.K, k_special, bfv_plain_modulus)
        if not self.h:
            raise ValueError(ref.err())

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.ref_ckks_destroy(self.h)
            self.h = None

    def keygen(self, galois_elts=()):
        arr = (u32 * max(1, len(galois_elts)))(*galois_elts)
        rc = self.lib.ref_keygen(self.h, arr, len(galois_elts))
        if rc:
            raise RuntimeError(self.ref.err())

    def make_synthetic_keys(self, galois_elts=(), seed=1):
        arr = (u32 * max(1, len(galois_elts)))(*galois_elts)
        rc = self.lib.ref_make_synthetic_keys(self.h, arr, len(galois_elts), seed)
        if rc:
            raise RuntimeError(self.ref.err())

    def get_key(self, which, elt=0):
        nb = self.lib.ref_key_bundles(self.h, which, elt)
        assert nb > 0, self.ref.err()
        out = np.empty((nb, 2, self.K, self.N), dtype=np.uint64)
        assert self.lib.ref_get_key(self.h, which, elt, _p(out)) == 0, self.ref.err()
        return out

    def set_key(self, which, elt, key):
        assert self.lib.ref_set_key(self.h, which, elt, _p(np.ascontiguousarray(key))) == 0, self.ref.err()

    def _run(self, rc):
        if rc == -1:
            raise ValueError(self.ref.err())
        if rc:
            raise RuntimeError(self.ref.err())

    def multiply(self, a, b):
        L = a.shape[1]
        out = np.empty((3, L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_multiply(self.h, L, _p(a), _p(b), _p(out)))
        return out

    def relinearize(self, c3):
        L = c3.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_relinearize(self.h, L, _p(c3), _p(out)))
        return out

    def secret_key(self):
        out = np.empty((self.K, self.N), dtype=np.uint64)
        self._run(self.lib.ref_get_secret_key(self.h, _p(out)))
        return out

    def set_relin_count(self, count):
        self._run(self.lib.ref_set_relin_count(self.h, count))

    def multiply_sizes(self, a, b):
        L, sa, sb = a.shape[1], a.shape[0], b.shape[0]
        out = np.empty((sa + sb - 1, L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_multiply_sizes(self.h, L, _p(a), sa, _p(b), sb, _p(out)))
        return out

    def relinearize_sizes(self, c):
        size, L = c.shape[0], c.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_relinearize_sizes(self.h, L, size, _p(c), _p(out)))
        return out

    def rescale(self, c):
        size, L = c.shape[0], c.shape[1]
        out = np.empty((size, L - 1, self.N), dtype=np.uint64)
        self._run(self.lib.ref_rescale(self.h, L, size, _p(c), _p(out)))
        return out

    def mod_switch(self, c):
        size, L = c.shape[0], c.shape[1]
        out = np.empty((size, L - 1, self.N), dtype=np.uint64)
        self._run(self.lib.ref_mod_switch(self.h, L, size, _p(c), _p(out)))
        return out

    def apply_galois(self, c, elt):
        L = c.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_apply_galois(self.h, L, elt, _p(c), _p(out)))
        return out

    def rotate(self, c, steps):
        L = c.shape[1]
        out = np.empty((2, L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_rotate(self.h, L, steps, _p(c), _p(out)))
        return out

    def mul_relin_rescale(self, a, b):
        L = a.shape[1]
        out = np.empty((2, L - 1, self.N), dtype=np.uint64)
        self._run(self.lib.ref_mul_relin_rescale(self.h, L, _p(a), _p(b), _p(out)))
        return out

    def elt_from_step(self, step):
        return self.lib.ref_elt_from_step(self.h, step)

    def decrypt(self, ct, sk, bfv=False):
        """Decryptor::decrypt with the secret key given as NTT-form limbs [K][N]: CKKS -> [L][N], BFV -> [N]."""
        size, L = ct.shape[0], ct.shape[1]
        out = np.empty((self.N,) if bfv else (L, self.N), dtype=np.uint64)
        self._run(self.lib.ref_decrypt(self.h, _p(np.ascontiguousarray(sk)), L, size, _p(ct), _p(out)))
        return out

    def time_ops(self, what, L, ops, n_threads, seed=1):
        s = self.lib.ref_time_ops(self.h, what, L, ops, n_threads, seed)
        if s < 0:
            raise RuntimeError(self.ref.err())
        return s
