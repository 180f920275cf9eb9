import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import int_peak
lib = int_peak.load()
for wide in (0, 1):
    print(wide, lib.gsb_bench_butterfly_rate(0, wide, 4, int_peak.Q55) / 1e9)
