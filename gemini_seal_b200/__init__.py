"""B200-native RNS polynomial-arithmetic hot path for Gemini-SEAL callers.

The compute lives in hand-written sm_100a CUDA behind the C ABI of include/gseal_b200.h
(gemini_seal_b200/_C/libgseal_b200.so, built by gemini_seal_b200.build).  This package is the thin
host-side mirror of the reference interface; it never falls back to a CPU implementation.
"""
from .evaluator import (BfvEvaluator, Context, CudaError, Decryptor, Evaluator, KeyGenerator, Workspace, inverse_ntt_negacyclic_harvey,  # noqa: F401
                        inverse_ntt_negacyclic_harvey_lazy, ntt_negacyclic_harvey,
                        ntt_negacyclic_harvey_lazy, poly_op)

__all__ = [
    "Context", "Workspace", "Evaluator", "BfvEvaluator", "Decryptor", "KeyGenerator", "CudaError", "ntt_negacyclic_harvey", "ntt_negacyclic_harvey_lazy",
    "inverse_ntt_negacyclic_harvey", "inverse_ntt_negacyclic_harvey_lazy", "poly_op",
]
