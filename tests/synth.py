"""Deterministic synthetic residues that the GPU (torch, int64 two's-complement arithmetic) and the CPU
(numpy, uint64) generate identically, so bench.py can draw its inputs on the device and any single
ciphertext of any rank can be regenerated on the host for the parity gate against the oracle.

word(seed, i) = splitmix64(seed + (i + 1) * GOLDEN) >> 1, reduced modulo the limb's modulus.
A tensor [cts][size][L][N] of ciphertexts first..first+cts-1 uses i = the word's flat index in the
(unbounded) array of all ciphertexts, so ciphertext c is the same whatever batch it is drawn in.
"""
import numpy as np

GOLDEN = 0x9E3779B97F4A7C15
M1 = 0xBF58476D1CE4E5B9
M2 = 0x94D049BB133111EB
MASK = (1 << 64) - 1


def _s64(x):
    x &= MASK
    return x - (1 << 64) if x >> 63 else x


def residues_numpy(seed, first, cts, size, mods, N):
    """uint64 [cts][size][len(mods)][N]."""
    L = len(mods)
    per = size * L * N
    idx = np.arange(first * per, (first + cts) * per, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (idx + np.uint64(1)) * np.uint64(GOLDEN) + np.uint64(seed & MASK)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(M1)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(M2)
        z = (z ^ (z >> np.uint64(31))) >> np.uint64(1)
    z = z.reshape(cts, size, L, N)
    for i, q in enumerate(mods):
        z[:, :, i, :] %= np.uint64(q)
    return z


def residues_torch(torch, seed, first, cts, size, mods, N, device="cuda", chunk=8):
    """int64 tensor [cts][size][len(mods)][N] on `device` holding the same words as residues_numpy."""
    L = len(mods)
    per = size * L * N
    out = torch.empty((cts, size, L, N), dtype=torch.int64, device=device)
    q = torch.tensor([int(m) for m in mods], dtype=torch.int64, device=device).view(1, 1, L, 1)

    def lsr(x, s):  # logical shift right on int64
        return (x >> s) & ((1 << (64 - s)) - 1)
    for c0 in range(0, cts, chunk):
        c1 = min(cts, c0 + chunk)
        idx = torch.arange((first + c0) * per, (first + c1) * per, dtype=torch.int64, device=device)
        z = (idx + 1) * _s64(GOLDEN) + _s64(seed)
        z = (z ^ lsr(z, 30)) * _s64(M1)
        z = (z ^ lsr(z, 27)) * _s64(M2)
        z = lsr(z ^ lsr(z, 31), 1)
        out[c0:c1] = z.view(c1 - c0, size, L, N) % q
    return out


def key_numpy(seed, mods, n_bundles, N):
    """[n_bundles][2][K][N] uniform residues: a synthetic switching key."""
    return residues_numpy(seed, 0, n_bundles, 2, mods, N)
