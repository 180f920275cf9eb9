"""CPU-side parity of the KERNEL SOURCES: gemini_seal_b200/csrc compiled by g++ against the test-only
CUDA execution-model emulation (tests/cusim) and driven through the same C ABI and Python mirror as
the product, compared with the reference-generated golden digests and with the oracle.  This checks
indexing / staging / synchronisation logic on the GPU-less build container; the -m gpu tests repeat
the same cases on the real library.  The simulator is never loaded by the package."""
import json
import os

import pytest

import golden_cases

with open(os.path.join(os.path.dirname(__file__), "golden", "evaluator_digests.json")) as f:
    GOLDEN = json.load(f)["cases"]

SIM_CASES = golden_cases.SMALL + ["n8192_k2_L3", "ntt_n14_b59", "bfv_n8192_default"]


@pytest.mark.parametrize("name", SIM_CASES)
def test_sim_matches_reference_digests(simlib, oracle, name):
    got = golden_cases.run_case(name, golden_cases.DeviceBackend(simlib))
    assert got == GOLDEN[name]


def test_sim_exports_every_declared_symbol(simlib):
    from gemini_seal_b200 import _lib
    for sym in _lib.PROTOTYPES:
        assert hasattr(simlib, sym) or sym in _lib.CUDA_ONLY


def test_sim_thin_evaluator_ops(simlib, oracle):
    from common import check_thin_ops
    check_thin_ops(simlib, oracle, logn=11)


def test_sim_fuzz_against_oracle(simlib, oracle):
    from common import fuzz_against_oracle
    fuzz_against_oracle(simlib, oracle, seed=2024, n_configs=24, max_logn=13)


def test_sim_decryptor(simlib, oracle):
    from common import check_decrypt
    check_decrypt(simlib, oracle, logn=10)
