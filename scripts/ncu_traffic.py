"""Extract per-launch DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) and duration of the hot-path
kernels from an `ncu --set full` report of one step (scripts/prof_step.py) and write profiles/<round>_ncu_traffic.json,
which bench.py reads for `roofline.traffic`.
Usage: python scripts/ncu_traffic.py gpurun_out/step.ncu-rep <batch_per_launch> [round tag, default r2]"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, batch = sys.argv[1], int(sys.argv[2])
tag = sys.argv[3] if len(sys.argv) > 3 else "r2"
# a .csv argument is the `ncu -i <rep> --page raw --csv` export itself (a report over 64 MiB does not travel back from the GPU box)
raw = open(rep).read() if rep.endswith('.csv') else subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]


def col(row, name):
    v, u = float(row[hdr.index(name)].replace(",", "")), units[hdr.index(name)]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
    return v * scale.get(u, 1)


def label(name):
    """bench.py's kernel labels (direction_pass/tag) for the launches of one mul+relin+rescale step, in order."""
    if "ks_mac" in name:
        return "ks_mac"
    if "ckks_multiply" in name:
        return "ckks_multiply"
    if "ModupLoad" in name:
        return "fwd_cols/modup"
    if "tail_mid" in name:
        return "tail_mid"
    if "TailInvLoad" in name:
        return "inv_contig/tail_last"
    if "TailLoad" in name:
        return "fwd_cols/tail"
    if "TailStore" in name:
        return "fwd_contig/tail"
    if "MdLoad" in name:
        return "fwd_cols/moddown"
    if "MdStore" in name:
        return "fwd_contig/moddown"
    if "RsLoad" in name:
        return "fwd_cols/rescale"
    if "RsStore" in name:
        return "fwd_contig/rescale"
    if "RsLastStore" in name:
        return "inv_cols/rescale_last"
    return None


out = {}
for row in rows[2:]:
    name = row[hdr.index("Kernel Name")]
    lab = label(name)
    if lab is None or lab in out:
        continue
    out[lab] = {
        "kernel": name[:100],
        "batch_per_launch": batch,
        "grid": row[hdr.index("Grid Size")],
        "dram_bytes_per_launch": col(row, "dram__bytes_read.sum") + col(row, "dram__bytes_write.sum"),
        "duration_ms_under_ncu": col(row, "gpu__time_duration.sum"),
        "fmaheavy_pct": float(row[hdr.index("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed")]),
        "lsu_wavefront_pct": float(row[hdr.index("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed")]),
        "issue_active_pct": float(row[hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")]),
        "registers": int(float(row[hdr.index("launch__registers_per_thread")])),
    }
total = sum(col(row, "dram__bytes_read.sum") + col(row, "dram__bytes_write.sum") for row in rows[2:])
doc = {"source": os.path.basename(rep), "command": "ncu --set full --clock-control none --import-source on --profile-from-start off "
       "python scripts/prof_step.py  (the kernels of one 64-ciphertext mul+relin+rescale step after 3 warm-up steps)",
       "launches_in_step": len(rows) - 2, "dram_bytes_per_step": total, "kernels": out}
path = os.path.join(ROOT, "profiles", tag + "_ncu_traffic.json")
with open(path, "w") as f:
    json.dump(doc, f, indent=1)
print(json.dumps(doc, indent=1))
