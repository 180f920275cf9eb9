#!/bin/bash
# bench alternative builds of the library (experiment): scripts/exp/variants.sh name1 name2 ...
for v in "$@"; do
  GSB_B200_LIB=$PWD/gemini_seal_b200/_C/$v.so python bench.py --no-cpu-baseline --sustain-s 0 --steps 10 > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/var_$v.json").read().strip().splitlines()[-1])
k=d["kernels"]
print("$v", round(d["ms_per_step"],3), {n: round(x["ms"],3) for n,x in k.items()}, d["parity_checked"]["ok"], round(d["extra"]["limb_ntt_per_s"]/1e6,3))
PY
done
