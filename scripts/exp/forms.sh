#!/bin/bash
# bench the ks_mac add forms (experiment)
for f in 1 2 3 4; do
  GSB_MAC_FORM=$f python bench.py --no-cpu-baseline --sustain-s 0 --steps 10 > gpurun_out/form_$f.json 2> gpurun_out/form_$f.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/form_$f.json").read().strip().splitlines()[-1])
k=d["kernels"]
print("form $f", round(d["ms_per_step"],3), {n: round(v["ms"],3) for n,v in k.items() if n in ("ks_mac","fwd_cols/modup","fwd_contig/tail","fwd_cols/tail")}, d["parity_checked"]["ok"])
PY
done
