"""Shared helpers for the parity tests: NTT-friendly primes, seeded synthetic ciphertexts / keys
(uniform residues), and the parity cases run against either the CUDA library (-m gpu) or
the test-only simulator build of the same kernel sources (CPU)."""
import hashlib

import numpy as np


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
