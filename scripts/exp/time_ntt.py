"""Strict forward + inverse NTT rate at a given log2 N (experiment): M limb-NTT/s."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import gemini_seal_b200 as g
from common import ntt_primes
logn = int(sys.argv[1]); N = 1 << logn; limbs = 16; polys = (1 << 26) // (N * limbs)
mods = ntt_primes(logn, 55, limbs)
ctx = g.Context(N, mods + ntt_primes(logn, 56, 1), 1)
lib = ctx.lib
x = torch.randint(0, 1 << 40, (polys, limbs, N), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def step():
    assert lib.gsb_ntt_forward_dev(ctx.handle, x.data_ptr(), polys, limbs, 0, 0, st) == 0
    assert lib.gsb_ntt_inverse_dev(ctx.handle, x.data_ptr(), polys, limbs, 0, 0, st) == 0
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): step()
e1.record(); torch.cuda.synchronize()
print(os.path.basename(os.environ.get("GSB_B200_LIB", "default")), "N=%d" % N, "%.2f M limb-NTT/s" % (2 * 20 * polys * limbs / (e0.elapsed_time(e1) * 1e-3) / 1e6))
