"""gsb_rescale_noalloc_dev (include/gseal_b200.h): the allocation-free variant of gsb_rescale_dev whose scratch is
a static device array inside the library.  Collected last on purpose.

Checked through the C ABI: bit-exact against gsb_rescale_dev (integer work: tolerance 0) and against the oracle, on
two calls in a row that share the static array (different inputs and a different level, so stale contents of the
first call would show), on a single-pass and a two-pass NTT size; the free device memory does not move across a
call; sizes beyond the static array are rejected with GSB_E_RANGE and a message that names the limit."""
import numpy as np
import pytest

import golden_cases
from common import rand_poly

pytestmark = pytest.mark.gpu

GSB_E_RANGE = -3
MAX_WORDS = 1 << 19


@pytest.mark.parametrize("n", [4096, 8192, 32768])
def test_noalloc_rescale_matches_rescale(cudalib, oracle, n):
    import torch
    import gemini_seal_b200 as g
    lib = cudalib
    mods = golden_cases.BFV_DEFAULT[n]
    ctx = g.Context(n, mods, 1, lib=lib)
    logn = n.bit_length() - 1
    oc = oracle.context(logn, mods, 1)
    L0 = len(mods) - 1
    rng = np.random.default_rng(n)
    size = 2
    batch = min(4, MAX_WORDS // (size * n))
    stream = torch.cuda.current_stream().cuda_stream
    # two calls in a row: level L0, then (another input, and where there is one, the next level down)
    plan = [(L0, rand_poly(rng, mods[:L0], (batch, size), n))]
    L2 = L0 - 1 if L0 - 1 >= 2 else L0
    plan.append((L2, rand_poly(rng, mods[:L2], (batch, size), n)))
    devs = [torch.from_numpy(x.view(np.int64)).cuda() for _, x in plan]
    outs = [torch.empty((batch, size, L - 1, n), dtype=torch.int64, device="cuda") for L, _ in plan]
    refs = [torch.empty_like(o) for o in outs]
    torch.cuda.synchronize()
    for i, (L, x) in enumerate(plan):
        if i == 1:
            free0 = torch.cuda.mem_get_info()[0]
        rc = lib.gsb_rescale_noalloc_dev(ctx.handle, devs[i].data_ptr(), outs[i].data_ptr(), batch, size, L, stream)
        assert rc == 0, lib.gsb_last_error().decode()
        if i == 1:
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info()[0] == free0, "the allocation-free entry point changed the free device memory"
    torch.cuda.synchronize()
    for i, (L, x) in enumerate(plan):  # the existing entry point afterwards, so it cannot have prepared anything
        rc = lib.gsb_rescale_dev(ctx.handle, None, devs[i].data_ptr(), refs[i].data_ptr(), batch, size, L, stream)
        assert rc == 0, lib.gsb_last_error().decode()
    torch.cuda.synchronize()
    for i, (L, x) in enumerate(plan):
        assert torch.equal(outs[i], refs[i]), (n, i, "noalloc != gsb_rescale_dev")
        o = outs[i].cpu().numpy().view(np.uint64)
        for b in (0, batch - 1):
            assert np.array_equal(o[b], oc.rescale(x[b])), (n, i, b, "noalloc != oracle")
    # a third call repeating the first input must reproduce the first output (the array keeps call 2's contents)
    again = torch.empty_like(outs[0])
    rc = lib.gsb_rescale_noalloc_dev(ctx.handle, devs[0].data_ptr(), again.data_ptr(), batch, size, plan[0][0], stream)
    assert rc == 0, lib.gsb_last_error().decode()
    torch.cuda.synchronize()
    assert torch.equal(again, outs[0])
    ctx.close()


def test_noalloc_rescale_rejects_oversize(cudalib):
    import torch
    import gemini_seal_b200 as g
    n = 4096
    mods = golden_cases.BFV_DEFAULT[n]
    ctx = g.Context(n, mods, 1, lib=cudalib)
    L = len(mods) - 1
    batch = MAX_WORDS // (2 * n) + 1  # one ciphertext too many
    x = torch.zeros((batch, 2, L, n), dtype=torch.int64, device="cuda")
    out = torch.empty((batch, 2, L - 1, n), dtype=torch.int64, device="cuda")
    rc = cudalib.gsb_rescale_noalloc_dev(ctx.handle, x.data_ptr(), out.data_ptr(), batch, 2, L, None)
    assert rc == GSB_E_RANGE
    assert "524288" in cudalib.gsb_last_error().decode()
    rc = cudalib.gsb_rescale_noalloc_dev(ctx.handle, x.data_ptr(), out.data_ptr(), batch - 1, 2, L, None)  # exactly the limit
    assert rc == 0, cudalib.gsb_last_error().decode()
    torch.cuda.synchronize()
    rc = cudalib.gsb_rescale_noalloc_dev(ctx.handle, x.data_ptr(), out.data_ptr(), 1, 2, 1, None)  # no next level
    assert rc == -1
    ctx.close()
