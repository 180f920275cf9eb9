"""CPU-only checks of the host side: the C-ABI library builds for sm_100a and exports every symbol
include/gseal_b200.h declares (no compute calls without a GPU), the device tables built by the host
code equal the reference's NTTTables, and the package refuses to run without its CUDA library."""
import ctypes
import os
import re

import numpy as np
import pytest

from common import ntt_primes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cuda_library_builds_and_exports_header_symbols():
    from gemini_seal_b200 import _lib, build
    path = build.build()
    lib = ctypes.CDLL(path)
    hdr = open(os.path.join(ROOT, "include", "gseal_b200.h")).read()
    declared = set(re.findall(r"\b(gsb_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    assert lib.gsb_version() >= 100


def test_no_cpu_fallback_without_device():
    """Without a CUDA device every compute entry point fails loudly (GSB_E_CUDA -> CudaError)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a CUDA device")
    import gemini_seal_b200 as g
    with pytest.raises(g.CudaError):
        g.Context(1024, ntt_primes(10, 40, 3), 1)


def test_missing_library_raises(tmp_path):
    from gemini_seal_b200 import _lib
    with pytest.raises(RuntimeError):
        _lib.load(str(tmp_path / "nope.so"))


def test_tables_equal_reference(simlib, reflib):
    """gsb_ctx_get_tables (host_tables.cpp) == NTTTables accessors of the reference (util/ntt.h:46-60)."""
    import gemini_seal_b200 as g
    for logn, bits in [(10, 36), (12, 59), (13, 55), (12, 60)]:
        mods = ntt_primes(logn, bits, 2)
        ctx = g.Context(1 << logn, mods, 1, lib=simlib)
        for i, q in enumerate(mods):
            w, ws, v, vs, sc = ctx.tables(i)
            rw, rws, rv, rvs, rsc = reflib.ntt_tables(logn, q)
            assert np.array_equal(w, rw) and np.array_equal(ws, rws)
            assert np.array_equal(v, rv) and np.array_equal(vs, rvs)
            cr = reflib.const_ratio(q)
            assert sc.tolist() == [q, int(cr[0]), int(cr[1]), int(rsc[0]), int(rsc[1]), int(rsc[2]), int(rsc[3])]
        ctx.close()


def test_galois_tables_equal_reference_kats(simlib, oracle):
    import gemini_seal_b200 as g
    ctx = g.Context(1024, ntt_primes(10, 40, 2), 1, lib=simlib)
    for elt in (3, 5, 25, 2047):
        assert np.array_equal(ctx.galois_table(elt), oracle.galois_table_ntt(10, elt))
    for step in (0, 1, -1, 7, -100, 511):
        assert ctx.galois_elt_from_step(step) == oracle.elt_from_step(10, step)
    ctx.close()
