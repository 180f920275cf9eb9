"""Host-side mirror of the reference interface for the RNS hot path, over the C ABI (include/gseal_b200.h).

Names follow the reference (native/src/seal/evaluator.h, util/ntt.h): `Evaluator.multiply`,
`relinearize`, `rescale_to_next`, `apply_galois`, `rotate_vector`, `complex_conjugate`,
`ntt_negacyclic_harvey[_lazy]`, `inverse_ntt_negacyclic_harvey[_lazy]`.  Ciphertexts are raw limb
buffers in the reference's own layout (ciphertext.h:357-381), batched: uint64 arrays of shape
[batch][size][L][N].  Two kinds of buffers are accepted everywhere:
  * numpy uint64 arrays (HOST)   -> the *_host entry points (copy in, run, copy out, synchronise);
  * torch int64 CUDA tensors     -> the *_dev entry points on torch's current stream, no copies.
Errors mirror the reference's convention (evaluator.cpp:238-249): invalid arguments raise
ValueError (std::invalid_argument), state errors raise RuntimeError (std::logic_error), CUDA failures
raise CudaError.  There is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import _lib

GSB_KEY_RELIN, GSB_KEY_GALOIS = 0, 1
OPS = {"dyadic": 0, "add": 1, "sub": 2, "negate": 3, "mul_scalar": 4}


class CudaError(RuntimeError):
    pass


def _check(lib, rc):
    if rc == 0:
        return
    msg = (lib.gsb_last_error() or b"").decode()
    if rc in (-1, -3):
        raise ValueError(msg)
    if rc == -2:
        raise RuntimeError(msg)
    if rc == -5:
        raise MemoryError(msg)
    raise CudaError(msg)


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _ptr(x):
    if _is_torch(x):
        if not x.is_cuda or not x.is_contiguous() or x.element_size() != 8:
            raise ValueError("device buffers must be contiguous 64-bit CUDA tensors")
        return x.data_ptr()
    if not (isinstance(x, np.ndarray) and x.dtype == np.uint64 and x.flags["C_CONTIGUOUS"]):
        raise ValueError("host buffers must be C-contiguous numpy uint64 arrays")
    return x.ctypes.data


def _stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def _empty_like(x, shape):
    if _is_torch(x):
        import torch
        return torch.empty(shape, dtype=x.dtype, device=x.device)
    return np.empty(shape, dtype=np.uint64)


class Context:
    """Device copy of what SEALContext::ContextData holds for the hot path (context.cpp:135-420):
    per-modulus constants and NTT tables for the key-level coeff_modulus, whose last
    `n_special_primes` entries are the special primes (encryptionparams.h:205-214)."""

    def __init__(self, poly_modulus_degree, coeff_modulus, n_special_primes=1, device=0, lib=None):
        self.lib = lib or _lib.load()
        n = int(poly_modulus_degree)
        if n < 2 or n & (n - 1):
            raise ValueError("poly_modulus_degree must be a power of two")
        self.N = n
        self.logn = n.bit_length() - 1
        self.moduli = [int(q) for q in coeff_modulus]
        self.K = len(self.moduli)
        self.k = int(n_special_primes)
        self.L0 = self.K - self.k
        arr = (C.c_uint64 * self.K)(*self.moduli)
        h = C.c_void_p()
        _check(self.lib, self.lib.gsb_ctx_create(int(device), self.logn, arr, self.K, self.k, C.byref(h)))
        self.handle = h
        self._keys = set()

    def close(self):
        if getattr(self, "handle", None):
            self.lib.gsb_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_bundles(self):
        return (self.L0 + self.k - 1) // self.k

    def tables(self, mod_index):
        """(root_powers, scaled_root_powers, inv_root_powers, scaled_inv_root_powers, scalars) as the
        reference's NTTTables accessors return them (util/ntt.h:46-60,84-92,116)."""
        arrs = [np.empty(self.N, dtype=np.uint64) for _ in range(4)]
        sc = np.empty(7, dtype=np.uint64)
        u64p = C.POINTER(C.c_uint64)
        _check(self.lib, self.lib.gsb_ctx_get_tables(
            self.handle, mod_index, *[a.ctypes.data_as(u64p) for a in arrs], sc.ctypes.data_as(u64p)))
        return (*arrs, sc)

    # RelinKeys / GaloisKeys upload: key = [n_bundles][2][K][N] (kswitchkeys.h:343)
    def set_relin_keys(self, key, key_power=2):
        """keys.data()[RelinKeys::get_index(key_power)] (relinkeys.h:61): the key for s^key_power."""
        if key_power < 2:
            raise ValueError("key_power cannot be less than 2")
        self._upload(GSB_KEY_RELIN, int(key_power) - 2, key)

    def set_galois_key(self, galois_elt, key):
        self._upload(GSB_KEY_GALOIS, int(galois_elt), key)

    def _upload(self, kind, elt, key):
        key = np.ascontiguousarray(key, dtype=np.uint64)
        if key.shape != (self.n_bundles, 2, self.K, self.N):
            raise ValueError("kswitch_keys is not valid for encryption parameters")
        _check(self.lib, self.lib.gsb_key_upload(self.handle, kind, elt, key.ctypes.data, key.shape[0]))

    def galois_elt_from_step(self, step):
        e = C.c_uint32()
        _check(self.lib, self.lib.gsb_galois_elt_from_step(self.handle, int(step), C.byref(e)))
        return e.value

    def galois_table(self, galois_elt):
        t = np.empty(self.N, dtype=np.uint32)
        _check(self.lib, self.lib.gsb_galois_table(self.handle, int(galois_elt), t.ctypes.data_as(C.POINTER(C.c_uint32))))
        return t


class Workspace:
    """Device scratch for the *_dev entry points: the stand-in for the MemoryPoolHandle argument of
    the reference's Evaluator methods (evaluator.h:246).  One per host thread / CUDA stream; an
    Evaluator built without one shares the context's default workspace (single stream only)."""

    def __init__(self, context):
        self.ctx = context
        h = C.c_void_p()
        _check(context.lib, context.lib.gsb_ws_create(context.handle, C.byref(h)))
        self.handle = h

    def close(self):
        if getattr(self, "handle", None) and self.ctx.handle:
            self.ctx.lib.gsb_ws_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _ntt(ctx, data, limb_offset, lazy, inverse):
    if data.ndim < 2 or data.shape[-1] != ctx.N:
        raise ValueError("data must be [..., n_limbs, N]")
    n_limbs = data.shape[-2]
    n_polys = int(np.prod(data.shape[:-2])) if data.ndim > 2 else 1
    lib = ctx.lib
    if _is_torch(data):
        fn = lib.gsb_ntt_inverse_dev if inverse else lib.gsb_ntt_forward_dev
        _check(lib, fn(ctx.handle, _ptr(data), n_polys, n_limbs, limb_offset, int(lazy), _stream()))
    else:
        fn = lib.gsb_ntt_inverse_host if inverse else lib.gsb_ntt_forward_host
        _check(lib, fn(ctx.handle, _ptr(data), n_polys, n_limbs, limb_offset, int(lazy)))
    return data


def ntt_negacyclic_harvey(data, ctx, limb_offset=0):
    """util::ntt_negacyclic_harvey over RNSIter / PolyIter (util/ntt.h:225-279), in place."""
    return _ntt(ctx, data, limb_offset, False, False)


def ntt_negacyclic_harvey_lazy(data, ctx, limb_offset=0):
    """util::ntt_negacyclic_harvey_lazy (util/ntt.cpp:292-342): output in [0, 4q)."""
    return _ntt(ctx, data, limb_offset, True, False)


def inverse_ntt_negacyclic_harvey(data, ctx, limb_offset=0):
    """util::inverse_ntt_negacyclic_harvey (util/ntt.h:318-368), in place."""
    return _ntt(ctx, data, limb_offset, False, True)


def inverse_ntt_negacyclic_harvey_lazy(data, ctx, limb_offset=0):
    """util::inverse_ntt_negacyclic_harvey_lazy (util/ntt.cpp:345-404): output in [0, 2q)."""
    return _ntt(ctx, data, limb_offset, True, True)


def poly_op(ctx, op, a, b=None, scalars=None, limb_offset=0):
    """dyadic_product_coeffmod / add_ / sub_ / negate_poly_coeffmod / multiply_poly_scalar_coeffmod
    (util/polyarithsmallmod.{h,cpp}) on [..., n_limbs, N] buffers."""
    n_limbs = a.shape[-2]
    n_polys = int(np.prod(a.shape[:-2])) if a.ndim > 2 else 1
    out = _empty_like(a, a.shape)
    sc = None
    if scalars is not None:
        sc = (C.c_uint64 * n_limbs)(*[int(s) for s in scalars])
    lib = ctx.lib
    pb = _ptr(b) if b is not None else None
    if _is_torch(a):
        _check(lib, lib.gsb_poly_op_dev(ctx.handle, OPS[op], _ptr(a), pb, sc, _ptr(out), n_polys, n_limbs, limb_offset, _stream()))
    else:
        _check(lib, lib.gsb_poly_op_host(ctx.handle, OPS[op], _ptr(a), pb, sc, _ptr(out), n_polys, n_limbs, limb_offset))
    return out


class Evaluator:
    """Mirror of seal::Evaluator for the CKKS hot path (evaluator.h:246,344,565,984,1201,1269)."""

    def __init__(self, context, workspace=None):
        self.ctx = context
        self.lib = context.lib
        self.workspace = workspace  # keeps it alive; None = the context's default workspace

    @property
    def _ws(self):
        return self.workspace.handle if self.workspace is not None else None

    def _shape(self, x, size=None):
        if x.ndim != 4 or x.shape[-1] != self.ctx.N or (size is not None and x.shape[1] != size):
            raise ValueError("encrypted is not valid for encryption parameters")
        return x.shape[0], x.shape[1], x.shape[2]

    def multiply(self, a, b):
        """multiply_inplace of NTT-form ciphertexts of sizes s1, s2 >= 2 (evaluator.cpp:235,447-527) -> size
        s1 + s2 - 1.  Two size-2 operands take the specialised kernel."""
        B, s1, L = self._shape(a)
        if b.ndim != 4 or (b.shape[0], b.shape[2], b.shape[3]) != (B, L, self.ctx.N):
            raise ValueError("encrypted1 and encrypted2 parameter mismatch")
        s2 = b.shape[1]
        out = _empty_like(a, (B, s1 + s2 - 1, L, self.ctx.N))
        if _is_torch(a):
            _check(self.lib, self.lib.gsb_ckks_multiply_sizes_dev(self.ctx.handle, _ptr(a), s1, _ptr(b), s2, _ptr(out), B, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_ckks_multiply_sizes_host(self.ctx.handle, _ptr(a), s1, _ptr(b), s2, _ptr(out), B, L))
        return out

    _BFV = 0

    def relinearize(self, c):
        """relinearize_inplace (evaluator.cpp:772-827) of ciphertexts of any size >= 2 -> size 2; size 3 takes the
        single key switch, larger sizes need the keys of the higher powers (Context.set_relin_keys(key, power))."""
        B, size, L = self._shape(c)
        out = _empty_like(c, (B, 2, L, self.ctx.N))
        if size == 3:
            dev, host = ((self.lib.gsb_bfv_relinearize_dev, self.lib.gsb_bfv_relinearize_host) if self._BFV
                         else (self.lib.gsb_relinearize_dev, self.lib.gsb_relinearize_host))
            if _is_torch(c):
                _check(self.lib, dev(self.ctx.handle, self._ws, _ptr(c), _ptr(out), B, L, _stream()))
            else:
                _check(self.lib, host(self.ctx.handle, _ptr(c), _ptr(out), B, L))
        elif _is_torch(c):
            _check(self.lib, self.lib.gsb_relinearize_sizes_dev(self.ctx.handle, self._ws, _ptr(c), size, _ptr(out), B, L, self._BFV, _stream()))
        else:
            _check(self.lib, self.lib.gsb_relinearize_sizes_host(self.ctx.handle, _ptr(c), size, _ptr(out), B, L, self._BFV))
        return out

    def rescale_to_next(self, c):
        """rescale_to_next (evaluator.cpp:1090-1126): drops the last limb; caller divides scale by q_last."""
        B, size, L = self._shape(c)
        if L < 2:
            raise ValueError("end of modulus switching chain reached")
        out = _empty_like(c, (B, size, L - 1, self.ctx.N))
        if _is_torch(c):
            _check(self.lib, self.lib.gsb_rescale_dev(self.ctx.handle, self._ws, _ptr(c), _ptr(out), B, size, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_rescale_host(self.ctx.handle, _ptr(c), _ptr(out), B, size, L))
        return out

    def apply_galois(self, c, galois_elt):
        """apply_galois_inplace (evaluator.cpp:1841-1943), size-2 NTT-form ciphertexts."""
        B, _, L = self._shape(c, 2)
        out = _empty_like(c, tuple(c.shape))
        if _is_torch(c):
            _check(self.lib, self.lib.gsb_apply_galois_dev(self.ctx.handle, self._ws, int(galois_elt), _ptr(c), _ptr(out), B, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_apply_galois_host(self.ctx.handle, int(galois_elt), _ptr(c), _ptr(out), B, L))
        return out

    # ---- thin compositions; ciphertext batches [B][size][L][N] ------------------
    def add(self, a, b):
        """add_inplace (evaluator.cpp:90-152) for equal-size operands."""
        self._shape(a)
        if tuple(a.shape) != tuple(b.shape):
            raise ValueError("encrypted1 and encrypted2 parameter mismatch")
        return poly_op(self.ctx, "add", a, b)

    def sub(self, a, b):
        """sub_inplace (evaluator.cpp:174-233) for equal-size operands."""
        self._shape(a)
        if tuple(a.shape) != tuple(b.shape):
            raise ValueError("encrypted1 and encrypted2 parameter mismatch")
        return poly_op(self.ctx, "sub", a, b)

    def negate(self, a):
        """negate_inplace (evaluator.cpp:65-88)."""
        self._shape(a)
        return poly_op(self.ctx, "negate", a)

    def multiply_plain(self, c, plain):
        """multiply_plain_ntt (evaluator.cpp:1605-1646): every component times an NTT-form plaintext [L][N]."""
        B, size, L = self._shape(c)
        if tuple(plain.shape) != (L, self.ctx.N):
            raise ValueError("encrypted_ntt and plain_ntt parameter mismatch")
        if _is_torch(c):
            p = plain.expand(B, size, L, self.ctx.N).contiguous()
        else:
            p = np.ascontiguousarray(np.broadcast_to(plain, c.shape))
        return poly_op(self.ctx, "dyadic", c, p)

    def square(self, c):
        """square_inplace (evaluator.cpp:529-558,704-770): the words of multiply(c, c)."""
        return self.multiply(c, c)

    def mod_switch_to_next(self, c):
        """CKKS mod_switch_to_next = mod_switch_drop_to_next (evaluator.cpp:894-957): drops the last limb."""
        B, size, L = self._shape(c)
        if L < 2:
            raise ValueError("end of modulus switching chain reached")
        out = c[:, :, :L - 1, :]
        return out.contiguous() if _is_torch(c) else np.ascontiguousarray(out)

    def transform_to_ntt(self, c):
        """transform_to_ntt_inplace (evaluator.cpp:1746-1792)."""
        self._shape(c)
        return ntt_negacyclic_harvey(c.clone() if _is_torch(c) else c.copy(), self.ctx)

    def transform_from_ntt(self, c):
        """transform_from_ntt_inplace (evaluator.cpp:1794-1840)."""
        self._shape(c)
        return inverse_ntt_negacyclic_harvey(c.clone() if _is_torch(c) else c.copy(), self.ctx)

    def rotate_vector(self, c, steps):
        """rotate_vector (evaluator.cpp:1945-2000) for a step whose Galois key is present."""
        if steps == 0:
            return c
        return self.apply_galois(c, self.ctx.galois_elt_from_step(steps))

    def complex_conjugate(self, c):
        """complex_conjugate (evaluator.h:1343-1363): Galois element 2N-1."""
        return self.apply_galois(c, 2 * self.ctx.N - 1)

    def mul_relin_rescale(self, a, b):
        """multiply + relinearize + rescale_to_next in one call (the headline unit of work)."""
        B, _, L = self._shape(a, 2)
        if tuple(b.shape) != tuple(a.shape):
            raise ValueError("encrypted1 and encrypted2 parameter mismatch")
        if L < 2:
            raise ValueError("end of modulus switching chain reached")
        out = _empty_like(a, (B, 2, L - 1, self.ctx.N))
        if _is_torch(a):
            _check(self.lib, self.lib.gsb_mul_relin_rescale_dev(self.ctx.handle, self._ws, _ptr(a), _ptr(b), _ptr(out), B, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_mul_relin_rescale_host(self.ctx.handle, _ptr(a), _ptr(b), _ptr(out), B, L))
        return out


class BfvEvaluator(Evaluator):
    """The BFV branches of the same seal::Evaluator entry points (coefficient-form ciphertexts):
    relinearize_inplace (evaluator.cpp:772-827 -> :2259-2368 with is_ckks == false) and
    apply_galois_inplace / rotate_rows / rotate_columns (evaluator.cpp:1893-1904, evaluator.h:1079-1176)."""

    _BFV = 1  # relinearize() of the base class then takes the coefficient-form branches

    def apply_galois(self, c, galois_elt):
        B, _, L = self._shape(c, 2)
        out = _empty_like(c, tuple(c.shape))
        if _is_torch(c):
            _check(self.lib, self.lib.gsb_bfv_apply_galois_dev(self.ctx.handle, self._ws, int(galois_elt), _ptr(c), _ptr(out), B, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_bfv_apply_galois_host(self.ctx.handle, int(galois_elt), _ptr(c), _ptr(out), B, L))
        return out

    def rotate_rows(self, c, steps):
        """rotate_rows (evaluator.h:1079-1130): same element as rotate_vector."""
        return self.rotate_vector(c, steps)

    def rotate_columns(self, c):
        """rotate_columns (evaluator.h:1147-1176): Galois element 2N-1."""
        return self.apply_galois(c, 2 * self.ctx.N - 1)

    def __init__(self, context, plain_modulus, workspace=None):
        super().__init__(context, workspace)
        self.plain_modulus = int(plain_modulus)

    def multiply(self, a, b):
        """multiply_inplace -> bfv_multiply (BEHZ, evaluator.cpp:274-445) of ciphertexts of sizes s1, s2 >= 2."""
        B, s1, L = self._shape(a)
        if b.ndim != 4 or (b.shape[0], b.shape[2], b.shape[3]) != (B, L, self.ctx.N):
            raise ValueError("encrypted1 and encrypted2 parameter mismatch")
        s2 = b.shape[1]
        out = _empty_like(a, (B, s1 + s2 - 1, L, self.ctx.N))
        if _is_torch(a):
            _check(self.lib, self.lib.gsb_bfv_multiply_sizes_dev(self.ctx.handle, self._ws, self.plain_modulus, _ptr(a), s1, _ptr(b), s2, _ptr(out), B, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_bfv_multiply_sizes_host(self.ctx.handle, self.plain_modulus, _ptr(a), s1, _ptr(b), s2, _ptr(out), B, L))
        return out

    def aux_base(self, L):
        """Moduli of Bsk for level L (RNSTool::base_Bsk)."""
        import ctypes as C
        buf = (C.c_uint64 * 80)()
        n = C.c_size_t()
        _check(self.lib, self.lib.gsb_bfv_aux_base(self.ctx.handle, L, self.plain_modulus, buf, 80, C.byref(n)))
        return [int(buf[i]) for i in range(n.value)]

    def mod_switch_to_next(self, c):
        """mod_switch_to_next, BFV: divide_and_round_q_last_inplace (util/rns.cpp:731-775)."""
        B, size, L = self._shape(c)
        if L < 2:
            raise ValueError("end of modulus switching chain reached")
        out = _empty_like(c, (B, size, L - 1, self.ctx.N))
        if _is_torch(c):
            _check(self.lib, self.lib.gsb_bfv_mod_switch_dev(self.ctx.handle, _ptr(c), _ptr(out), B, size, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_bfv_mod_switch_host(self.ctx.handle, _ptr(c), _ptr(out), B, size, L))
        return out

    def _unsupported(self, *a, **k):
        raise NotImplementedError("not part of the BFV surface built so far (no CPU fallback)")

    mul_relin_rescale = rescale_to_next = multiply_plain = _unsupported


class KeyGenerator:
    """Mirror of the arithmetic half of seal::KeyGenerator (keygenerator.cpp:138-180, 51-116, 325-420; SURVEY 8(f)
    row f2).  The caller supplies the sampled buffers -- `a` uniform residues in NTT form [n][beta][K][N], `e` the
    noise polynomials as int64 [n][beta][N] -- and gets the switching keys [n][beta][2][K][N]; install=True also
    places them in the context's key store.  `secret_key` is SecretKey::data(): NTT form at key level, [K][N]."""

    def __init__(self, context, secret_key):
        self.ctx, self.lib = context, context.lib
        sk = np.ascontiguousarray(secret_key, dtype=np.uint64)
        if sk.shape != (context.K, context.N):
            raise ValueError("secret key is not valid for encryption parameters")
        _check(self.lib, self.lib.gsb_secret_key_set(context.handle, _ptr(sk)))

    def _run(self, kind, ids, a, e, install):
        n = len(ids)
        ctx = self.ctx
        if tuple(a.shape) != (n, ctx.n_bundles, ctx.K, ctx.N) or tuple(e.shape) != (n, ctx.n_bundles, ctx.N):
            raise ValueError("sampled buffers do not match the encryption parameters")
        arr = (C.c_uint32 * n)(*[int(i) for i in ids])
        out = _empty_like(a, (n, ctx.n_bundles, 2, ctx.K, ctx.N))
        if _is_torch(a):
            _check(self.lib, self.lib.gsb_kswitch_keygen_dev(ctx.handle, None, kind, arr, n, _ptr(a), _ptr(e), _ptr(out), int(install), _stream()))
        else:
            e = np.ascontiguousarray(e, dtype=np.int64)
            _check(self.lib, self.lib.gsb_kswitch_keygen_host(ctx.handle, kind, arr, n, _ptr(a), e.ctypes.data, _ptr(out), int(install)))
        return out

    def relin_keys(self, a, e, count=1, install=False):
        """relin_keys(count) (keygenerator.cpp:138-180): keys for s^2 .. s^(count+1)."""
        return self._run(GSB_KEY_RELIN, [p + 2 for p in range(count)], a, e, install)

    def galois_keys(self, galois_elts, a, e, install=False):
        """galois_keys(galois_elts) (keygenerator.cpp:51-116), one key per element."""
        return self._run(GSB_KEY_GALOIS, list(galois_elts), a, e, install)


class Decryptor:
    """Mirror of seal::Decryptor (decryptor.h / decryptor.cpp) on raw limb buffers.

    `secret_key` is SecretKey::data() as the reference stores it: NTT form at key level, uint64 [K][N].
    CKKS (plain_modulus None): decrypt(c) = dot_product_ct_sk_array(c) (decryptor.cpp:115-149, 218-265),
    [B][size][L][N] NTT form -> plaintexts [B][L][N].  BFV: decrypt(c) = bfv_decrypt (decryptor.cpp:76-113):
    coefficient-form ciphertexts -> [B][N] plaintext coefficients mod t (leading zeros not trimmed)."""

    def __init__(self, context, secret_key, plain_modulus=None, workspace=None):
        self.ctx, self.lib, self.workspace = context, context.lib, workspace
        self.plain_modulus = None if plain_modulus is None else int(plain_modulus)
        sk = np.ascontiguousarray(secret_key, dtype=np.uint64)
        if sk.shape != (context.K, context.N):
            raise ValueError("secret key is not valid for encryption parameters")
        _check(self.lib, self.lib.gsb_secret_key_set(context.handle, _ptr(sk)))

    @property
    def _ws(self):
        return self.workspace.handle if self.workspace is not None else None

    def dot_product_ct_sk_array(self, c, ntt_form):
        if c.ndim != 4 or c.shape[-1] != self.ctx.N or c.shape[1] < 2:
            raise ValueError("encrypted is not valid for encryption parameters")
        B, size, L = c.shape[0], c.shape[1], c.shape[2]
        out = _empty_like(c, (B, L, self.ctx.N))
        if _is_torch(c):
            _check(self.lib, self.lib.gsb_dot_product_ct_sk_dev(self.ctx.handle, self._ws, _ptr(c), _ptr(out), B, size, L, 1 if ntt_form else 0, _stream()))
        else:
            _check(self.lib, self.lib.gsb_dot_product_ct_sk_host(self.ctx.handle, _ptr(c), _ptr(out), B, size, L, 1 if ntt_form else 0))
        return out

    def decrypt(self, c):
        if self.plain_modulus is None:
            return self.dot_product_ct_sk_array(c, True)
        if c.ndim != 4 or c.shape[-1] != self.ctx.N or c.shape[1] < 2:
            raise ValueError("encrypted is not valid for encryption parameters")
        B, size, L = c.shape[0], c.shape[1], c.shape[2]
        out = _empty_like(c, (B, self.ctx.N))
        if _is_torch(c):
            _check(self.lib, self.lib.gsb_bfv_decrypt_dev(self.ctx.handle, self._ws, self.plain_modulus, _ptr(c), _ptr(out), B, size, L, _stream()))
        else:
            _check(self.lib, self.lib.gsb_bfv_decrypt_host(self.ctx.handle, self.plain_modulus, _ptr(c), _ptr(out), B, size, L))
        return out
