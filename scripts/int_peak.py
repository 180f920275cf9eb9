"""Measurement aid (bench side): the integer-pipe peaks the INT roofline of bench.py / DESIGN.md is quoted
against, from scripts/int_peak.cu (register-resident chains; not part of the product library).

    python scripts/int_peak.py            # full sweep, text report (committed as profiles/r2_int_peak.txt)

measure(device) is what bench.py calls: for each instruction kind it sweeps chains per thread x CTAs per SM
and keeps the maximum.
"""
import ctypes as C
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "int_peak.cu")
LIB = os.path.join(HERE, "_C", "libgsb_intpeak.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
KINDS = {"imad": 0, "imad_wide_acc": 1, "imad_hi": 2, "imad_wide": 3}
Q55 = 36028797017456641  # first BFVDefault(32768) prime


def build(force=False):
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    deps = [SRC] + [os.path.join(HERE, "..", "gemini_seal_b200", "csrc", f) for f in ("ntt_core.cuh", "modarith.cuh", "compat.cuh")]
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(d) for d in deps):
        return LIB
    subprocess.run([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-shared",
                    "-Xcompiler", "-fPIC", "-o", LIB, SRC], check=True)
    build_pipes()
    return LIB


def build_pipes():
    """scripts/exp/pipes.cu -> scripts/_C/pipes.bin (the instruction-mix measurement behind DESIGN.md's two-pipe model):
    `scripts/_C/pipes.bin > profiles/r2_pipes.txt` on the GPU box."""
    src, out = os.path.join(HERE, "exp", "pipes.cu"), os.path.join(HERE, "_C", "pipes.bin")
    if os.path.exists(src) and not (os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(src)):
        subprocess.run([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", out, src], check=True)
    return out


def load():
    lib = C.CDLL(build())
    lib.gsb_bench_chain_rate.restype = C.c_double
    lib.gsb_bench_chain_rate.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
    lib.gsb_bench_butterfly_rate.restype = C.c_double
    lib.gsb_bench_butterfly_rate.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64]
    return lib


def sweep(lib, device, chains=(4, 8, 12, 16), ctas=(1, 2, 3, 4)):
    """{kind: {(chains, ctas): rate}} for the four instruction kinds, {form: {ctas: rate}} for the butterflies."""
    out = {}
    for name, kind in KINDS.items():
        out[name] = {}
        for ch in chains:
            for c in ctas:
                r = lib.gsb_bench_chain_rate(device, kind, ch, c)
                if r <= 0:
                    raise RuntimeError("int_peak: CUDA failure")
                out[name][(ch, c)] = r
    for form, wide in (("butterfly_literal", 0), ("butterfly_wide", 1)):
        out[form] = {c: lib.gsb_bench_butterfly_rate(device, wide, c, Q55) for c in ctas}
    return out


def measure(device=0, quick=True):
    """Maxima over the sweep (quick: chains in {8, 16}, CTAs per SM in {2, 4})."""
    lib = load()
    s = sweep(lib, device, (8, 16), (2, 4)) if quick else sweep(lib, device)
    best = {k: max(v.values()) for k, v in s.items()}
    best["where"] = {k: "%s" % (max(v, key=v.get),) for k, v in s.items()}
    return best, s


def clocks_line(device):
    try:
        out = subprocess.run(["nvidia-smi", "-i", str(device), "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active",
                              "--format=csv,noheader"], capture_output=True, text=True, timeout=10).stdout.strip()
        return out
    except (OSError, subprocess.TimeoutExpired):
        return "nvidia-smi unavailable"


def main():
    device = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    lib = load()
    print("integer-pipe peaks, device %d; clocks before: %s" % (device, clocks_line(device)))
    s = sweep(lib, device)
    for name in KINDS:
        print("\n%s (T/s): rows = chains per thread, columns = CTAs (256 threads) per SM 1..4" % name)
        for ch in (4, 8, 12, 16):
            print("  %2d  " % ch + "  ".join("%6.2f" % (s[name][(ch, c)] / 1e12) for c in (1, 2, 3, 4)))
        print("  max %.3f T/s" % (max(s[name].values()) / 1e12))
    for form in ("butterfly_literal", "butterfly_wide"):
        print("\n%s (G butterflies/s) at 1..4 CTAs per SM: " % form + "  ".join("%.1f" % (s[form][c] / 1e9) for c in (1, 2, 3, 4)))
    print("\nclocks after: %s" % clocks_line(device))


if __name__ == "__main__":
    main()
