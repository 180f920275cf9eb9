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

SIM_CASES = golden_cases.SMALL + ["n8192_k2_L3", "ntt_n14_b59", "bfv_n8192_default", "gen_n8192_L4_3x3"]


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


def test_sim_fast_flag_and_literal_mode(simlib, oracle):
    """gsb_ctx_fast_ntt follows q_max (5 log2 N + 2) < 2^63, and GSB_LITERAL=1 turns the shortcuts off
    without changing any output word (small case; the GPU suite repeats it at N = 2^15)."""
    import numpy as np
    import gemini_seal_b200 as g
    from common import mixed_moduli, rand_key, rand_poly
    logn, N = 11, 2048
    # gsb_ctx_shortcuts: bit 0 fast_ntt, 1 lazy_modup, 2 lazy_mac, 3 wide_inv, 4 fast_fwd_public
    for bits, want, mask in (([56, 55, 56, 56], 1, 31), ([57, 55, 56, 56], 1, 31), ([58, 55, 56, 56], 0, 8), ([59, 58, 59, 59], 0, 8),
                             ([60, 58, 59, 59], 0, 0)):
        ctx = g.Context(N, mixed_moduli(logn, bits), 1, lib=simlib)
        assert simlib.gsb_ctx_fast_ntt(ctx.handle) == want, bits
        assert simlib.gsb_ctx_shortcuts(ctx.handle) == mask, (bits, simlib.gsb_ctx_shortcuts(ctx.handle))
        ctx.close()
    mods = mixed_moduli(logn, [50, 40, 45, 51])
    rng = np.random.default_rng(3)
    relin = rand_key(rng, mods, 3, N)
    a, b = rand_poly(rng, mods[:3], (2, 2), N), rand_poly(rng, mods[:3], (2, 2), N)
    outs = []
    for literal in (False, True):
        if literal:
            os.environ["GSB_LITERAL"] = "1"
        try:
            ctx = g.Context(N, mods, 1, lib=simlib)
        finally:
            os.environ.pop("GSB_LITERAL", None)
        assert simlib.gsb_ctx_fast_ntt(ctx.handle) == (0 if literal else 1)
        assert simlib.gsb_ctx_shortcuts(ctx.handle) == (0 if literal else 31)
        ctx.set_relin_keys(relin)
        outs.append(g.Evaluator(ctx).mul_relin_rescale(a, b))
        ctx.close()
    assert np.array_equal(outs[0], outs[1])


def test_sim_threshold_primes_and_extreme_patterns(simlib, oracle):
    """Shortcut thresholds and structured extreme inputs under the simulator, whose GSB_SIM_ASSERTs also check every
    no-wrap precondition on these inputs (the GPU suite repeats it at N = 2^12, 2^13, 2^15)."""
    from common import check_threshold_and_patterns
    check_threshold_and_patterns(simlib, oracle, 10, light=True)


@pytest.mark.parametrize("name", ["keygen_n1024_k1", "keygen_n2048_k2_ragged"])
def test_sim_keygen_reproduces_reference_keys(simlib, name):
    """Row f2 under the simulator: sampled buffers of the reference in -> the reference's key words out."""
    from common import check_keygen_fixture, load_keygen_fixture
    check_keygen_fixture(simlib, load_keygen_fixture(name))
