"""Developer probe: print the measured integer-pipe peaks for several occupancies."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gemini_seal_b200 as g
import golden_cases

ctx = g.Context(32768, golden_cases.BFV_DEFAULT[32768], 1)
for ctas in (1, 2, 3, 4, 6, 8):
    a, b, c = C.c_double(), C.c_double(), C.c_double()
    rc = ctx.lib.gsb_int_peak(ctx.handle, ctas, C.byref(a), C.byref(b), C.byref(c))
    assert rc == 0, ctx.lib.gsb_last_error()
    w = C.c_double()
    assert ctx.lib.gsb_int_peak_wide(ctx.handle, ctas, C.byref(w)) == 0, ctx.lib.gsb_last_error()
    print("ctas/SM=%d  IMAD %.2f T/s  IMAD.WIDE %.2f T/s  butterflies %.1f G/s literal, %.1f G/s wide-range"
          % (ctas, a.value / 1e12, b.value / 1e12, c.value / 1e9, w.value / 1e9))
