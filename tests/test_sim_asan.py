"""Bounds check of the kernel sources without a GPU: the simulator build (tests/cusim) compiled with
-fsanitize=address runs a few representative cases in a subprocess; any out-of-bounds global / shared
access of a kernel (device buffers are plain heap blocks there) aborts it.  compute-sanitizer is not
available on the GPU pool, so this is the memory checker these kernels get."""
import os
import shutil
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "gemini_seal_b200", "csrc")
CASES = ["n1024_k3_L4_ragged", "n2048_k2_L4", "c1_bfv_n4096_default", "bfv_n2048_k2_L4", "n8192_k2_L3", "ntt_n13_b59"]

DRIVER = r"""
import ctypes, sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
from gemini_seal_b200 import _lib
import golden_cases
lib = ctypes.CDLL(%r); _lib.bind(lib)
for name in %r:
    golden_cases.run_case(name, golden_cases.DeviceBackend(lib))
print("ASAN_CASES_OK")
"""


def test_simulator_under_address_sanitizer(tmp_path):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    asan = subprocess.run(["g++", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan not installed")
    lib = str(tmp_path / "libsim_asan.so")
    subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-shared", "-fPIC", "-fsanitize=address", "-fno-omit-frame-pointer",
                    "-DGSB_CUSIM", "-I", os.path.join(HERE, "cusim"), "-I", CSRC, "-x", "c++",
                    os.path.join(CSRC, "gsb_api.cu"), os.path.join(CSRC, "host_tables.cpp"), "-o", lib], check=True)
    env = dict(os.environ, LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0:detect_stack_use_after_return=0")
    r = subprocess.run([sys.executable, "-c", DRIVER % (ROOT, HERE, lib, CASES)], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "ASAN_CASES_OK" in r.stdout, (r.stdout[-2000:], r.stderr[-4000:])
