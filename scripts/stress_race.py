"""Determinism stress: run the fused device entry points repeatedly on the same inputs and report any run whose output
differs from the first (a shared-memory or staging race shows up as run-to-run differences long before it shows up as a
wrong golden digest).  Usage (GPU box):
    python scripts/stress_race.py [reps] [batch] [c4|c3]
(compute-sanitizer is closed on this GPU pool, so this and the simulator build's assertions are the race checks.)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import bench  # noqa: E402
import gemini_seal_b200 as g  # noqa: E402
import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
w = bench.workload(sys.argv[3] if len(sys.argv) > 3 else "c4")
ctx = g.Context(w.N, w.mods, w.k)
lib = ctx.lib
keys = bench.synthetic_keys(w)
ctx.set_relin_keys(keys[0])
elt = ctx.galois_elt_from_step(1)
ctx.set_galois_key(elt, keys[1])
a = synth.residues_torch(torch, bench.SEED_A, 0, B, 2, w.mods[:w.L], w.N)
b = synth.residues_torch(torch, bench.SEED_B, 0, B, 2, w.mods[:w.L], w.N)
stream = torch.cuda.current_stream().cuda_stream


def mrr():
    out = torch.empty((B, 2, w.L - 1, w.N), dtype=torch.int64, device="cuda")
    rc = lib.gsb_mul_relin_rescale_dev(ctx.handle, None, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, w.L, stream)
    assert rc == 0, lib.gsb_last_error().decode()
    return out


def rot():
    out = torch.empty_like(a)
    rc = lib.gsb_apply_galois_dev(ctx.handle, None, elt, a.data_ptr(), out.data_ptr(), B, w.L, stream)
    assert rc == 0, lib.gsb_last_error().decode()
    return out


def ntt():
    x = a.clone()
    rc = lib.gsb_ntt_inverse_dev(ctx.handle, x.data_ptr(), B * 2, w.L, 0, 0, stream)
    rc |= lib.gsb_ntt_forward_dev(ctx.handle, x.data_ptr(), B * 2, w.L, 0, 0, stream)
    assert rc == 0, lib.gsb_last_error().decode()
    return x


bad = 0
for name, fn in (("mul_relin_rescale", mrr), ("rotate_vector(1)", rot), ("strict inverse + forward NTT", ntt)):
    first = fn()
    torch.cuda.synchronize()
    for r in range(reps):
        again = fn()
        torch.cuda.synchronize()
        if not torch.equal(first, again):
            d = (first != again).nonzero()
            print("%s: run %d differs from run 0 at %d words, first at %s" % (name, r + 1, d.shape[0], d[0].tolist()))
            bad += 1
    if name.startswith("strict"):
        assert torch.equal(first, a), "inverse + forward strict NTT is not the identity on canonical input"
    print("%s: %d repeats identical to the first run" % (name, reps) if bad == 0 else "%s: FAILED" % name)
ctx.close()
sys.exit(1 if bad else 0)
