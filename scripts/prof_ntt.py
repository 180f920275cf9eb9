"""Developer probe for ncu: a few strict forward + inverse NTT launches at N = 2^15 on [32][2][15][N]."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gemini_seal_b200 as g
import golden_cases

N = 32768
mods = golden_cases.BFV_DEFAULT[N]
ctx = g.Context(N, mods, 1)
x = torch.stack([torch.randint(0, q, (int(sys.argv[1]) if len(sys.argv) > 1 else 32, 2, N), dtype=torch.int64, device="cuda") for q in mods[:15]], dim=-2).contiguous()
ref = x.clone()
for _ in range(3):
    g.ntt_negacyclic_harvey(x, ctx)
    g.inverse_ntt_negacyclic_harvey(x, ctx)
torch.cuda.synchronize()
assert torch.equal(x, ref)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    g.ntt_negacyclic_harvey(x, ctx)
e1.record()
torch.cuda.synchronize()
n = x.shape[0] * 2 * 15 * 5
print("fwd strict: %.3f M limb-NTT/s" % (n / e0.elapsed_time(e1) / 1e3))
