"""Developer probe (not part of the test-suite): raw batched NTT parity + timing on cuda:0.

Usage on the GPU box:  python scripts/ntt_probe.py [logn ...]
Checks the CUDA NTT against oracle/liboracle.so on a small batch, then times a large batch with CUDA
events.  Prints one line per (logn, direction).
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int

PRIMES59 = {}


def find_primes(logn, bits, count):
    """NTT-friendly primes q = 1 mod 2N just below 2^bits (deterministic Miller-Rabin)."""
    def is_prime(n):
        if n < 2:
            return False
        for p in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
            if n % p == 0:
                return n == p
        d, s = n - 1, 0
        while d % 2 == 0:
            d //= 2
            s += 1
        for a in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
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
    out = []
    step = 2 << logn
    q = ((1 << bits) - 1) // step * step + 1
    while len(out) < count:
        if is_prime(q):
            out.append(q)
        q -= step
    return out


def main():
    logns = [int(a) for a in sys.argv[1:]] or [12, 13, 14, 15, 16]
    lib = C.CDLL(os.path.join(ROOT, "gemini_seal_b200", "_C", "libgseal_b200.so"))
    orc = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
    lib.gsb_last_error.restype = C.c_char_p
    lib.gsb_ctx_create.argtypes = [i32, i32, C.POINTER(C.c_uint64), sz, sz, C.POINTER(vp)]
    for f in (lib.gsb_ntt_forward_dev, lib.gsb_ntt_inverse_dev):
        f.argtypes = [vp, vp, sz, sz, sz, i32, vp]
    orc.so_ntt.argtypes = [i32, C.c_uint64, vp, sz, i32]
    torch.cuda.init()
    for logn in logns:
        N = 1 << logn
        K = 4
        mods = find_primes(logn, 59, K)
        arr = (C.c_uint64 * K)(*mods)
        ctx = vp()
        rc = lib.gsb_ctx_create(0, logn, arr, K, 1, C.byref(ctx))
        assert rc == 0, lib.gsb_last_error()
        rng = np.random.default_rng(logn)
        for inverse in (0, 1):
            for mode in (0, 1):
                host = np.stack([rng.integers(0, q, size=(3, N), dtype=np.uint64) for q in mods], axis=1)  # [3][K][N]
                host = np.ascontiguousarray(host)
                want = host.copy()
                kind = (2 if inverse else 0) + (0 if mode else 1)
                for p in range(3):
                    for i, q in enumerate(mods):
                        orc.so_ntt(logn, q, want[p, i].ctypes.data, 1, kind)
                dev = torch.from_numpy(host.view(np.int64)).cuda()
                fn = lib.gsb_ntt_inverse_dev if inverse else lib.gsb_ntt_forward_dev
                rc = fn(ctx, dev.data_ptr(), 3, K, 0, mode, torch.cuda.current_stream().cuda_stream)
                assert rc == 0, lib.gsb_last_error()
                torch.cuda.synchronize()
                got = dev.cpu().numpy().view(np.uint64)
                ok = np.array_equal(got, want)
                print("logn=%d inverse=%d lazy=%d parity=%s" % (logn, inverse, mode, "OK" if ok else "MISMATCH"), flush=True)
        # timing: ~1 GiB of limbs (larger than L2)
        n_polys = max(1, (1 << 30) // (K * N * 8))
        big = torch.randint(0, mods[-1], (n_polys, K, N), dtype=torch.int64, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        for inverse in (0, 1):
            fn = lib.gsb_ntt_inverse_dev if inverse else lib.gsb_ntt_forward_dev
            for _ in range(3):
                fn(ctx, big.data_ptr(), n_polys, K, 0, 0, st)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            reps = 5
            for _ in range(reps):
                fn(ctx, big.data_ptr(), n_polys, K, 0, 0, st)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            limbs = n_polys * K
            print("logn=%d inverse=%d: %.3f ms for %d limb-NTTs -> %.3f M limb-NTT/s, algorithmic %.0f GB/s" % (
                logn, inverse, ms, limbs, limbs / ms / 1e3, limbs * 16 * N / ms / 1e6), flush=True)
        del big


if __name__ == "__main__":
    main()
