#!/bin/bash
# everything the committed round-2 evidence comes from, on one box (experiment helper)
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_latest.log 2>&1; tail -2 gpurun_out/r2_pytest_gpu_latest.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_full.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["two_pipe"], d["extra"]["limb_ntt_per_s"], d["extra"]["c3_n16384"]["ops_per_s"], d["extra"]["rotate_ops_per_s"], {n: round(v["ms"],3) for n,v in d["kernels"].items()})
for k,v in d["extra"]["ntt_sweep"].items(): print(k, round(v["limb_ntt_per_s"]/1e6,3), round(v["int_frac"],3))
print(d["extra"]["bfv_n4096"])
PY
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2_launches_step.csv python scripts/prof_step.py > /dev/null 2>&1
if [ "$1" = "full" ]; then
ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/r2_step_full python scripts/prof_step.py > gpurun_out/r2_step_full.log 2>&1; tail -1 gpurun_out/r2_step_full.log; ls -la gpurun_out/
fi
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustain-s 0 > /dev/null 2>&1
scripts/_C/pipes.bin > gpurun_out/r2_pipes.txt
python scripts/int_peak.py > gpurun_out/r2_int_peak.txt
