"""Regenerates tests/golden/bench_digests.json: sha256 digests of what the UNMODIFIED reference
(oracle/_ref/libref_harness.so) computes for the ciphertexts bench.py's parity gate checks -- global
ciphertexts r*64 and r*64 + 63 (r = 0..7: ciphertexts 0 and B-1 of every rank's shard at B = 64, for 1, 2, 4
and 8 ranks) of the C4 workload, and 0 and 255 of the C3 workload.  Inputs and keys are the splitmix streams of
tests/synth.py with bench.py's seeds.  Run in the build container only:

    make -C oracle ref && python tests/golden/make_bench_digests.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

import bench  # noqa: E402
import synth  # noqa: E402


def main():
    out = {"generator": "tests/golden/make_bench_digests.py",
           "source": "unmodified reference via oracle/_ref/libref_harness.so (seal::Evaluator multiply_inplace -> relinearize_inplace -> "
                     "rescale_to_next_inplace, and apply_galois_inplace for rotate_vector(1))"}
    for tag, cts in (("c4", sorted({r * 64 + o for r in range(8) for o in (0, 63)})), ("c3", [0, 255])):
        w = bench.workload(tag)
        keys = bench.synthetic_keys(w)
        rc = bench.reference_context(w, keys)
        elt = rc.elt_from_step(1)
        d = {"mrr": {}, "rot": {}}
        for g in cts:
            a = synth.residues_numpy(bench.SEED_A, g, 1, 2, w.mods[:w.L], w.N)[0]
            b = synth.residues_numpy(bench.SEED_B, g, 1, 2, w.mods[:w.L], w.N)[0]
            d["mrr"][str(g)] = bench.digest(rc.mul_relin_rescale(a, b))
            d["rot"][str(g)] = bench.digest(rc.apply_galois(a, elt))
            print(tag, g, "ok", flush=True)
        out[tag] = d
    with open(os.path.join(HERE, "bench_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
