"""Determinism stress: run each device entry point repeatedly on the same full-size inputs and report where
(ciphertext, component, limb, coefficient range) any run differs from the first.  Usage (GPU box):
    python scripts/stress_race.py [reps] [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import gemini_seal_b200 as g
from gemini_seal_b200 import _lib

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
N, L = bench.N, bench.L
lib = _lib.load()
mods = bench.moduli()
ctx = g.Context(N, mods, bench.K_SPECIAL, device=0, lib=lib)
ev = g.Evaluator(ctx)
gen = torch.Generator(device="cuda")
gen.manual_seed(99)
ctx.set_relin_keys(bench.uniform_residues(torch, mods, (L, 2), gen).cpu().numpy().view(np.uint64))
elt = ctx.galois_elt_from_step(1)
ctx.set_galois_key(elt, bench.uniform_residues(torch, mods, (L, 2), gen).cpu().numpy().view(np.uint64))
a = bench.uniform_residues(torch, mods[:L], (B, 2), gen)
b = bench.uniform_residues(torch, mods[:L], (B, 2), gen)
c3 = bench.uniform_residues(torch, mods[:L], (B, 3), gen)


def where(ref, got):
    d = (ref != got)
    idx = d.nonzero()
    print("    %d words differ" % idx.shape[0])
    seen = {}
    for row in idx.tolist():
        key = tuple(row[:-1])
        lo, hi, n = seen.get(key, (1 << 30, -1, 0))
        seen[key] = (min(lo, row[-1]), max(hi, row[-1]), n + 1)
    for key, (lo, hi, n) in list(seen.items())[:12]:
        print("    [ct, comp, limb] = %s coefficients %d..%d (%d words)" % (list(key), lo, hi, n))


cases = {
    "multiply": lambda: ev.multiply(a, b),
    "relinearize": lambda: ev.relinearize(c3),
    "rescale": lambda: ev.rescale_to_next(a),
    "apply_galois": lambda: ev.apply_galois(a, elt),
    "mul_relin_rescale": lambda: ev.mul_relin_rescale(a, b),
    "ntt_fwd": lambda: g.ntt_negacyclic_harvey(a.clone(), ctx),
    "ntt_inv": lambda: g.inverse_ntt_negacyclic_harvey(a.clone(), ctx),
}
only = os.environ.get("STRESS_CASES")
if only:
    cases = {k: v for k, v in cases.items() if k in only.split(",")}
bad = 0
for name, fn in cases.items():
    ref = fn().clone()
    torch.cuda.synchronize()
    fails = 0
    for r in range(reps):
        got = fn()
        torch.cuda.synchronize()
        if not torch.equal(ref, got):
            fails += 1
            if fails <= 2:
                print("  %s: run %d differs" % (name, r))
                where(ref, got)
    print("%s: %d/%d runs differ" % (name, fails, reps))
    bad += fails
sys.exit(1 if bad else 0)
