"""Regenerates tests/golden/keygen_*.npz: switching keys made by the UNMODIFIED reference's KeyGenerator (relin key +
Galois keys) together with the secret key and the buffers the reference SAMPLED for them (a = component 1 of every
bundle, e = the noise recovered from c0 = -(a s + e) + factor * new_key with the CPU oracle), so that the device key
generation can be checked word for word where the reference is absent (the GPU box).  Small shapes only (k = 1 and
k = 2 with a ragged last bundle): the fixture holds full keys.  Run in the build container:

    make -C oracle ref && python tests/golden/make_keygen_fixtures.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from common import mixed_moduli, reference_keys_for_keygen  # noqa: E402
from oracle import pyoracle  # noqa: E402

CASES = {"keygen_n1024_k1": (10, [40, 40, 41, 42], 1, [1, -3]), "keygen_n2048_k2_ragged": (11, [45, 45, 45, 46, 46], 2, [2])}


def main():
    ref, orc = pyoracle.RefLib(), pyoracle.Oracle()
    for name, (logn, bits, k, steps) in CASES.items():
        mods = mixed_moduli(logn, bits)
        fx = reference_keys_for_keygen(ref, orc, logn, mods, k, steps)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), logn=logn, mods=np.array(mods, dtype=np.uint64), k=k, sk=fx["sk"],
                            elts=np.array(fx["elts"], dtype=np.uint32), keys=np.stack(fx["keys"]), a=np.stack(fx["a"]), e=np.stack(fx["e"]))
        print(name, "ok", np.abs(np.stack(fx["e"])).max())


if __name__ == "__main__":
    main()
