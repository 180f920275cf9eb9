"""One headline step (C4, 64 ciphertext pairs, mul+relin+rescale) bracketed by cudaProfilerStart/Stop after three
warm-up steps, for `ncu --profile-from-start off ...` captures of exactly the kernels of one timed step.
    ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/step python scripts/prof_step.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import bench  # noqa: E402
import gemini_seal_b200 as g  # noqa: E402
import synth  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "mrr"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
w = bench.workload("c4")
ctx = g.Context(w.N, w.mods, w.k)
lib = ctx.lib
keys = bench.synthetic_keys(w)
ctx.set_relin_keys(keys[0])
elt = ctx.galois_elt_from_step(1)
ctx.set_galois_key(elt, keys[1])
a = synth.residues_torch(torch, bench.SEED_A, 0, B, 2, w.mods[:w.L], w.N)
b = synth.residues_torch(torch, bench.SEED_B, 0, B, 2, w.mods[:w.L], w.N)
out = torch.empty((B, 2, w.L - 1, w.N), dtype=torch.int64, device="cuda")
rot = torch.empty_like(a)
stream = torch.cuda.current_stream().cuda_stream


def step():
    if what == "rot":
        rc = lib.gsb_apply_galois_dev(ctx.handle, None, elt, a.data_ptr(), rot.data_ptr(), B, w.L, stream)
    elif what == "ntt":
        rc = lib.gsb_ntt_forward_dev(ctx.handle, a.data_ptr(), B * 2, w.L, 0, 0, stream)
        rc |= lib.gsb_ntt_inverse_dev(ctx.handle, a.data_ptr(), B * 2, w.L, 0, 0, stream)
    else:
        rc = lib.gsb_mul_relin_rescale_dev(ctx.handle, None, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, w.L, stream)
    assert rc == 0, lib.gsb_last_error().decode()


for _ in range(3):
    step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one %s step, batch %d" % (what, B))
