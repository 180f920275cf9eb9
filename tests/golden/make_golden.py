"""Regenerates tests/golden/evaluator_digests.json and batch64_digests.json (per-ciphertext digests of the
batch-64 C3 / C4 cases) by running the UNMODIFIED reference
(oracle/_ref/libref_harness.so, compiled from /root/reference by oracle/Makefile) on the seeded
cases of tests/golden_cases.py.  Run in the build container only:

    make -C oracle ref && python tests/golden/make_golden.py

The fixture stores sha256 digests of the reference's output limbs (inputs are re-derived from the
case name), so it stays a few KiB while pinning full-size (N = 2^15, K = 16) outputs.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import golden_cases  # noqa: E402
from oracle import pyoracle  # noqa: E402


def main():
    ref = pyoracle.RefLib()
    be = golden_cases.RefBackend(ref)
    cases = {}
    for name in list(golden_cases.NTT) + list(golden_cases.EVAL) + list(golden_cases.BFV) + list(golden_cases.GEN):
        cases[name] = golden_cases.run_case(name, be)
        print(name, "ok", flush=True)
    out = {
        "generator": "tests/golden/make_golden.py",
        "source": "unmodified reference via oracle/_ref/libref_harness.so (SEAL 3.5.3 + Gemini fork delta)",
        "cases": cases,
    }
    with open(os.path.join(HERE, "evaluator_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    b64 = {}
    for name in golden_cases.EVAL_B64:
        b64[name] = golden_cases.run_case_per_ct(name, be)
        print(name, "ok", flush=True)
    out["cases"] = b64
    with open(os.path.join(HERE, "batch64_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
