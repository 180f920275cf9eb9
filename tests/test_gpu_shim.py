"""Drop-in check of the C++ shim (shim/gsb_seal_shim.{h,cpp}) on the GPU box: oracle/_ref/shim_check
links the UNMODIFIED reference and the B200 library into one program and compares
seal::Evaluator (CPU) with gsb_seal::Evaluator (device) on the same seal::Ciphertext / RelinKeys /
GaloisKeys objects -- word for word, metadata, decryption, exception types.  The binary is built in
the container that has /root/reference (oracle/Makefile target `shim`) and travels with the repo."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "shim_check")


@pytest.mark.gpu
def test_shim_matches_reference_evaluator():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/shim_check not built (needs /root/reference at build time)")
    p = subprocess.run([BIN], capture_output=True, text=True, timeout=600)
    print(p.stdout[-4000:])
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-1000:]
    assert "SHIM_CHECK PASS" in p.stdout
    assert "MISMATCH" not in p.stdout


def test_shim_compiles_against_reference_headers():
    """CPU-side: where the reference is present, the shim must compile against its headers."""
    if not os.path.isdir("/root/reference/native/src/seal"):
        pytest.skip("/root/reference absent")
    from gemini_seal_b200 import build
    build.build()
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "shim"], check=True)
    assert os.path.exists(BIN)


def test_shim_against_simulator(simlib):
    """CPU-side: the same shim_check binary with the TEST-ONLY simulator build of the kernel sources
    preloaded (its C-ABI symbols take precedence over libgseal_b200.so's), so the shim's host logic,
    the C ABI and the kernels are compared with the unmodified reference without a GPU."""
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/shim_check not built (needs /root/reference at build time)")
    env = dict(os.environ, LD_PRELOAD=simlib._name)
    p = subprocess.run([BIN], capture_output=True, text=True, timeout=900, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-1000:]
    assert "SHIM_CHECK PASS" in p.stdout and "MISMATCH" not in p.stdout


CBIN = os.path.join(ROOT, "oracle", "_ref", "c_layer_check")


@pytest.mark.gpu
def test_c_export_layer_on_device():
    """the reference's C export entry points (Evaluator_*, Decryptor_*; what the .NET
    wrapper P/Invokes) served by shim/gsb_seal_c.cpp, compared with seal::Evaluator / seal::Decryptor."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if not os.path.exists(CBIN):
        pytest.skip("oracle/_ref/c_layer_check not built (needs /root/reference at build time)")
    p = subprocess.run([CBIN], capture_output=True, text=True, timeout=600)
    print(p.stdout[-4000:])
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-1000:]
    assert "C_LAYER_CHECK PASS" in p.stdout and "FAIL " not in p.stdout


def test_c_export_layer_against_simulator(simlib):
    """CPU-side twin of the test above (simulator build preloaded, see test_shim_against_simulator)."""
    if not os.path.exists(CBIN):
        pytest.skip("oracle/_ref/c_layer_check not built (needs /root/reference at build time)")
    env = dict(os.environ, LD_PRELOAD=simlib._name)
    p = subprocess.run([CBIN], capture_output=True, text=True, timeout=900, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-1000:]
    assert "C_LAYER_CHECK PASS" in p.stdout


SBIN = os.path.join(ROOT, "oracle", "_ref", "shim_bench")


@pytest.mark.gpu
def test_shim_bench_on_device():
    """The measurement program of the C++ boundary (shim/shim_bench.cpp) at a reduced batch: every timed result is
    compared with seal::Evaluator inside the program (parity_ok), and the JSON line it prints is what bench.py embeds
    as extra.shim."""
    import json
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if not os.path.exists(SBIN):
        pytest.skip("oracle/_ref/shim_bench not built (needs /root/reference at build time)")
    p = subprocess.run([SBIN, "8", "3", "1"], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-1000:]
    line = json.loads(p.stdout.strip().splitlines()[-1])
    print(json.dumps(line))
    assert line["parity_ok"] is True and line["chain_ops_per_s"] > 0
