"""TEST INFRASTRUCTURE: compile the kernel sources with g++ against cusim.h (no GPU needed).

Produces tests/cusim/_build/libgseal_b200_sim.so exporting the same C ABI, used only by
tests/test_sim_*.py to debug kernel indexing on the dev container.  Never loaded by the package.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
CSRC = os.path.join(ROOT, "gemini_seal_b200", "csrc")
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "libgseal_b200_sim.so")


def build(force=False):
    os.makedirs(OUT, exist_ok=True)
    newest = max(os.path.getmtime(os.path.join(d, f)) for d in (CSRC, HERE, os.path.join(ROOT, "include"))
                 for f in os.listdir(d) if not f.startswith("_"))
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= newest:
        return LIB
    cmd = ["g++", "-O2", "-g", "-std=c++17", "-shared", "-fPIC", "-DGSB_CUSIM", "-I", HERE, "-I", CSRC,
           "-x", "c++", os.path.join(CSRC, "gsb_api.cu"), os.path.join(CSRC, "host_tables.cpp"), "-o", LIB]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
