"""Time the fused step without any parity check (for deliberately wrong experimental builds): ms per 64-pair step."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, bench, synth
import gemini_seal_b200 as g
B = 64
w = bench.workload("c4")
ctx = g.Context(w.N, w.mods, w.k)
lib = ctx.lib
keys = bench.synthetic_keys(w)
ctx.set_relin_keys(keys[0])
a = synth.residues_torch(torch, bench.SEED_A, 0, B, 2, w.mods[:w.L], w.N)
b = synth.residues_torch(torch, bench.SEED_B, 0, B, 2, w.mods[:w.L], w.N)
out = torch.empty((B, 2, w.L - 1, w.N), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def step():
    assert lib.gsb_mul_relin_rescale_dev(ctx.handle, None, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, w.L, st) == 0
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): step()
e1.record(); torch.cuda.synchronize()
print(os.environ.get("GSB_B200_LIB", "default"), "%.3f ms/step" % (e0.elapsed_time(e1) / 20))
