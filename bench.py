#!/usr/bin/env python
"""Headline benchmark: batched CKKS multiply + relinearize + rescale_to_next at N = 2^15
(BASELINE.json configs #4/#5: 881-bit default modulus, L = 15, K = 16, k = 1), ciphertexts sharded
across GPUs; plus config #3 (N = 2^14), rotations/s, limb-NTT/s (config #2) and BFV config #1 as `extra`.

    python bench.py --gpus 1 --steps 20 --warmup 3             # our arm
    python bench.py --impl reference --steps 2 --warmup 1      # the reference's CPU path on host cores
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field's definition.

Parity gate: before anything is timed, every rank compares ciphertexts 0 and B-1 of its own
shard -- mul+relin+rescale and rotation outputs of the very buffers that are then timed -- with the unmodified
reference (oracle/_ref, where it travelled with the repo) and with digests the reference produced in the build
container (tests/golden/bench_digests.json).  Inputs and switching keys are splitmix64 streams that numpy
regenerates word for word on the host (tests/synth.py).
"""
import argparse
import ctypes as C
import hashlib
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

METRIC = "ckks_mul_relin_rescale_ops_per_s_n32768"
SEED_A, SEED_B, SEED_RELIN, SEED_GALOIS = 0xA11CE, 0xB0B, 0x5EA1, 0x6A101
DIGESTS = os.path.join(ROOT, "tests", "golden", "bench_digests.json")


class Workload:
    def __init__(self, tag, logn, k_special, title):
        import golden_cases
        self.tag, self.logn, self.N, self.k = tag, logn, 1 << logn, k_special
        self.mods = golden_cases.BFV_DEFAULT[self.N]
        self.K = len(self.mods)
        self.L = self.K - k_special
        self.beta = (self.L + k_special - 1) // k_special
        self.E = self.L + k_special
        self.limb_bytes = 8 * self.N
        self.title = title

    def name(self, batch):
        return ("%s: batched multiply_inplace + relinearize_inplace + rescale_to_next_inplace, %d ciphertext pairs per GPU "
                "per step, synthetic uniform-residue ciphertexts and switching key" % (self.title, batch))

    # limb-NTTs per unit of work: key switch beta*L + beta*k + 2k + 2L, rescale 2L
    def ntts_mul_relin_rescale(self):
        return self.beta * self.L + self.beta * self.k + 2 * self.k + 2 * self.L + 2 * self.L

    # SURVEY 8(d) algorithmic bytes of one mul+relin+rescale: 4L in + 2(L-1) out (+ key 2 beta K / B)
    def op_bytes(self, batch):
        return (4 * self.L + 2 * (self.L - 1) + 2.0 * self.beta * self.K / batch) * self.limb_bytes


def workload(tag):
    if tag == "c4":
        return Workload("c4", 15, 1, "C4/C5: CKKS N=32768, 881-bit BFVDefault modulus (L=15, K=16, k=1, beta=15)")
    return Workload("c3", 14, 1, "C3: CKKS N=16384, 438-bit BFVDefault modulus (L=8, K=9, k=1, beta=8)")


def config_dict(w, batch, world):
    """Printed identically by both arms (the driver compares the dicts)."""
    return {"workload": w.name(batch), "batch_per_gpu": batch,
            "parallelism": "ciphertext-sharded x%d, no data-path collective" % world,
            "l2": "inputs %d MiB per step per GPU, larger than the 126 MiB L2" % (2 * batch * 2 * w.L * w.limb_bytes >> 20)}


# ---- clocks sampler (B200_PROFILING.md "clocks DURING the timed region") -----------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, pw, mx, reasons = [], [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or not f[0].isdigit() or int(f[0]) != self.gpu:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "sm_mhz_min": min(sm) if sm else None, "power_w_max": max(pw) if pw else None}


# ---- per-kernel algorithmic HBM bytes -----------------------------------------------
def kernel_bytes(w, label, B):
    """Bytes a launch of `label` must read + write for B ciphertexts given what the ALGORITHM hands it (the
    coefficient- or NTT-form limbs it consumes and the limbs it must produce; intermediates the implementation
    chooses to park in HBM, e.g. the column-pass output `ext1`, are NOT counted)."""
    L, k, E, beta = w.L, w.k, w.E, w.beta
    limbs = {
        "ckks_multiply": B * (4 * L + 3 * L),
        "inv_contig/ks_target": B * 2 * L, "inv_cols/ks_target": B * 2 * L,
        "fwd_cols/modup": B * (L + L),                     # coefficient-form target in; its mod-up'd NTTs are intermediates
        "ks_mac": B * (2 * L + 2 * E) + 2 * beta * E,      # target (both forms) in, reduced accumulators out, key once
        "inv_contig/ks_special": B * 4 * k, "inv_cols/ks_special": B * 4 * k,
        "fwd_cols/moddown": B * 2 * (k + L), "fwd_contig/moddown": B * 2 * 4 * L,
        "inv_contig/rescale_last": B * 4, "inv_cols/rescale_last": B * 4,
        "fwd_cols/rescale": B * 2 * L, "fwd_contig/rescale": B * 2 * 3 * (L - 1),
        "fwd_cols/tail": B * 2 * (2 + L - 1), "fwd_contig/tail": B * 2 * 4 * (L - 1),
        "inv_contig/tail_last": B * 2 * 3, "inv_cols/tail_last": B * 2 * 2,
    }
    return limbs.get(label, 0) * w.limb_bytes


# forward butterfly stages each (bundle, ext slot) pair goes through in the two key-switch kernels
def butterfly_stages(w, label):
    a = {13: 6, 14: 6, 15: 6, 16: 8}.get(w.logn, 0)
    return {"ks_mac": w.logn - a, "fwd_cols/modup": a}.get(label)


def ncu_traffic(label, batch_per_launch):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `label` from the committed ncu --set full
    capture (profiles/r2_ncu_traffic.json, else r1; written by scripts/ncu_traffic.py), or None when that capture
    was taken at another batch per launch."""
    for name in ("r2_ncu_traffic.json", "r1_ncu_traffic.json"):
        path = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(path):
            continue
        with open(path) as f:
            t = json.load(f)
        k = t.get("kernels", {}).get(label)
        if k and k.get("batch_per_launch") == batch_per_launch:
            return k["dram_bytes_per_launch"]
    return None


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def digest(arr):
    import numpy as np
    return hashlib.sha256(np.ascontiguousarray(arr).tobytes()).hexdigest()


# ---- synthetic keys / reference context -------------------------------------------------------------------
def synthetic_keys(w):
    import synth
    return synth.key_numpy(SEED_RELIN, w.mods, w.beta, w.N), synth.key_numpy(SEED_GALOIS, w.mods, w.beta, w.N)


def reference_context(w, keys=None, elt=None):
    """The unmodified reference (oracle/_ref) with the bench's synthetic keys, or None where it did not travel."""
    from oracle import pyoracle
    try:
        ref = pyoracle.RefLib()
    except (FileNotFoundError, OSError):
        return None
    rc = ref.context(w.N, w.mods, w.k)
    elt = elt if elt is not None else rc.elt_from_step(1)
    rc.make_synthetic_keys([elt], seed=7)
    if keys is not None:
        rc.set_key(0, 0, keys[0])
        rc.set_key(1, elt, keys[1])
    return rc


def cpu_port_runner(w):
    """Single-threaded oracle port, used only where oracle/_ref is absent."""
    import numpy as np
    from common import rand_key, rand_poly
    from oracle import pyoracle
    orc = pyoracle.Oracle().context(w.logn, w.mods, w.k)
    rng = np.random.default_rng(1)
    a, b = rand_poly(rng, w.mods[:w.L], (2,), w.N), rand_poly(rng, w.mods[:w.L], (2,), w.N)
    key = rand_key(rng, w.mods, w.beta, w.N)

    def run(ops, threads):
        t0 = time.perf_counter()
        for _ in range(ops):
            orc.mul_relin_rescale(a, b, key)
        return time.perf_counter() - t0
    return run, orc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    w = workload("c4")
    rc = reference_context(w, synthetic_keys(w))
    if rc is not None:
        kind, threads = "reference", host_threads()
        run = lambda ops, th: rc.time_ops(0, w.L, ops, th, 1)
    else:
        kind, threads = "port", 1
        run, _keep = cpu_port_runner(w)
    ops_per_step = max(1, threads)  # one op per thread per step: a bounded sample of the workload
    for _ in range(args.warmup):
        run(ops_per_step, threads)
    secs = 0.0
    for _ in range(args.steps):
        secs += run(ops_per_step, threads)
    value = ops_per_step * args.steps / secs
    sample = "%d steps x %d mul+relin+rescale ops (one per thread) on synthetic inputs of the workload's shape" % (args.steps, ops_per_step)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "ops/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": config_dict(w, args.batch, max(world, args.gpus)),
        "note": "reference CPU path (seal::Evaluator, unmodified, oracle/_ref) on the host cores" if kind == "reference"
                else "oracle port (oracle/seal_oracle.c): oracle/_ref did not travel",
        "cpu_baseline": {"value": value, "unit": "ops/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "ops/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- parity gate --------------------------------------------------------------------------------------------
def load_digests():
    if os.path.exists(DIGESTS):
        with open(DIGESTS) as f:
            return json.load(f)
    return {}


def parity_gate(w, rc, first_ct, B, out_np, rot_np, elt):
    """Ciphertexts 0 and B-1 of this rank's shard (global indices first_ct, first_ct + B - 1): the device outputs
    `out_np[i]` (mul+relin+rescale) and `rot_np[i]` (rotation), i in {0, 1}, against the unmodified reference run
    here on the regenerated inputs and against the committed reference digests.  Returns (ok, how)."""
    import numpy as np
    import synth
    dg = load_digests().get(w.tag, {})
    ok, ref_used, dig_used = True, 0, 0
    for i, local in enumerate((0, B - 1)):
        g = first_ct + local
        got = {"mrr": out_np[i], "rot": rot_np[i]}
        if rc is not None:
            a = synth.residues_numpy(SEED_A, g, 1, 2, w.mods[:w.L], w.N)[0]
            b = synth.residues_numpy(SEED_B, g, 1, 2, w.mods[:w.L], w.N)[0]
            ok &= bool(np.array_equal(rc.mul_relin_rescale(a, b), got["mrr"]))
            ok &= bool(np.array_equal(rc.apply_galois(a, elt), got["rot"]))
            ref_used += 1
        for what in ("mrr", "rot"):
            want = dg.get(what, {}).get(str(g))
            if want is not None:
                ok &= digest(got[what]) == want
                dig_used += 1
    if ref_used == 0 and dig_used < 4:
        return False, "unchecked: neither oracle/_ref nor committed digests for these ciphertexts"
    how = []
    if ref_used:
        how.append("oracle/_ref (unmodified seal::Evaluator, memcmp)")
    if dig_used:
        how.append("committed reference digests (%d)" % dig_used)
    return ok, " + ".join(how)


# ---- our arm ----------------------------------------------------------------------------------------------
def bfv_config1(torch, g, lib, np, timed, batch=256):
    """BASELINE config #1 shape (BFV, N = 4096, BFVDefault 109-bit modulus, t = 786433; parity is pinned by the
    golden digests): device-resident ops/s of bfv_multiply, relinearize and rotate_rows on one GPU."""
    n, mods, t = 4096, [68719403009, 68719230977, 137438822401], 786433
    ctx = g.Context(n, mods, 1, device=torch.cuda.current_device(), lib=lib)
    ev = g.BfvEvaluator(ctx, t)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(4096)

    def residues(ms, lead):
        cols = [torch.randint(0, q, tuple(lead) + (n,), dtype=torch.int64, device="cuda", generator=gen) for q in ms]
        return torch.stack(cols, dim=-2).contiguous()
    ctx.set_relin_keys(residues(mods, (2, 2)).cpu().numpy().view(np.uint64))
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, residues(mods, (2, 2)).cpu().numpy().view(np.uint64))
    a, b = residues(mods[:2], (batch, 2)), residues(mods[:2], (batch, 2))
    c3 = ev.multiply(a, b)
    out = {}
    for name, fn in (("multiply", lambda: ev.multiply(a, b)), ("relinearize", lambda: ev.relinearize(c3)),
                     ("rotate_rows", lambda: ev.apply_galois(a, elt))):
        fn()
        out[name + "_ops_per_s"] = batch * 5 / (timed(fn, 5) / 1e3)
    out["batch"] = batch
    ctx.close()
    return out


def ntt_sweep(torch, g, lib, timed, limbs=16, target_bytes=1 << 29):
    """BASELINE config #2: strict forward + inverse limb-NTTs per second for N = 2^12..2^16 over `limbs` 59-bit
    primes, batch sized to ~512 MiB (> L2).  Reported with the two rooflines of SURVEY 8(d): HBM (16 N bytes per
    limb-NTT against the measured copy peak) and integer (butterflies/s)."""
    from common import ntt_primes
    peak, _ = measured_peaks()
    out = {}
    for logn in (12, 13, 14, 15, 16):
        n = 1 << logn
        mods = ntt_primes(logn, 59, limbs)
        ctx = g.Context(n, mods + ntt_primes(logn, 58, 1), 1, device=torch.cuda.current_device(), lib=lib)
        polys = max(1, target_bytes // (limbs * n * 8))
        cols = [torch.randint(0, q, (polys, n), dtype=torch.int64, device="cuda") for q in mods]
        x = torch.stack(cols, dim=1).contiguous()
        ref = x.clone()
        stream = torch.cuda.current_stream().cuda_stream

        def fwd():
            lib.gsb_ntt_forward_dev(ctx.handle, x.data_ptr(), polys, limbs, 0, 0, stream)

        def inv():
            lib.gsb_ntt_inverse_dev(ctx.handle, x.data_ptr(), polys, limbs, 0, 0, stream)
        fwd()
        inv()
        ms_f, ms_i = timed(fwd, 3), timed(inv, 3)
        # 3 forward transforms were applied in a row and then 3 inverse ones: still a round trip
        assert torch.equal(x, ref), "NTT round trip failed at N=2^%d" % logn
        per_s = 2 * polys * limbs * 3 / ((ms_f + ms_i) / 1e3)
        out["n%d" % n] = {"limb_ntt_per_s": per_s, "fwd_per_s": polys * limbs * 3 / (ms_f / 1e3),
                          "inv_per_s": polys * limbs * 3 / (ms_i / 1e3), "hbm_frac": per_s * 16 * n / 1e9 / peak,
                          "butterflies_per_s": per_s * (n // 2) * logn, "polys": polys, "limbs": limbs}
        ctx.close()
        del x, ref, cols
    return out


def kernel_table(lib, step):
    """Per-kernel CUDA-event times of one more step (gsb_prof_*: events on the launching stream around every launch)."""
    lib.gsb_prof_enable(1)
    step()
    need = lib.gsb_prof_report(None, 0)
    buf = C.create_string_buffer(max(need, 16))
    lib.gsb_prof_report(buf, need)
    lib.gsb_prof_enable(0)
    kernels = json.loads(buf.value.decode())
    total = sum(k["ms"] for k in kernels.values()) or 1.0
    for k in kernels.values():
        k["share"] = k["ms"] / total
    return kernels


def int_roofline(w, lib, ctx, kernels, B, peaks):
    """The binding roofline of the dominant kernel: the multiplies its ALGORITHM needs (butterflies and 64x64->128
    multiply-accumulates only; no address, carry or move instructions), each weighted by its measured
    register-resident issue rate (scripts/int_peak.cu), over the kernel's measured time.  Unit: IMAD.WIDE-equivalents/s;
    it is what ncu reports as sm__pipe_fmaheavy_cycles_active minus the non-multiply instructions on that pipe."""
    top = max(kernels, key=lambda x: kernels[x]["ms"])
    k = kernels[top]
    secs = k["ms"] * 1e-3
    hbm_peak, how = measured_peaks()
    kb = kernel_bytes(w, top, B)
    roof = {"bound": "int", "kernel": top, "kernel_share_of_step": k["share"], "launch_ms": k["ms"] / k["launches"],
            "traffic": ncu_traffic(top, B // k["launches"]),
            "hbm": {"bound": "hbm", "achieved": kb / secs / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": kb / secs / 1e9 / hbm_peak,
                    "bytes_per_launch": kb / k["launches"], "peak_source": how,
                    "what": "algorithmic bytes of this kernel (inputs it is handed + outputs it must produce + the key once per launch) "
                            "over its measured time: far below 1 because the kernel is multiplier-bound"}}
    stages = butterfly_stages(w, top)
    if stages is None or peaks is None:
        roof.update({"achieved": None, "peak": None, "unit": "T IMAD.WIDE-equivalents/s", "frac": None})
        return roof
    wide, imad = peaks["imad_wide"], peaks["imad"]
    fast = bool(lib.gsb_ctx_fast_ntt(ctx.handle))
    pairs = w.beta * w.E - w.L                      # (bundle, ext slot) pairs that go through the NTT (k = 1)
    n_bf = B * pairs * stages * (w.N // 2)
    w_per_bf, i_per_bf = (5, 4) if fast else (6, 4)
    n_wide, n_imad = n_bf * w_per_bf, n_bf * i_per_bf
    if top == "ks_mac":
        n_wide += B * w.beta * w.E * w.N * 2 * 4    # two FMAU128 per (bundle, slot, coefficient)
        if not fast:
            n_wide += B * pairs * w.N // 2 * 5      # reduceBarrettLazy of the last stage's x0
            n_imad += B * pairs * w.N // 2 * 2
    t_peak = n_wide / wide + n_imad / imad
    roof.update({
        "achieved": t_peak * wide / secs / 1e12, "peak": wide / 1e12, "unit": "T IMAD.WIDE-equivalents/s", "frac": t_peak / secs,
        "multiplies_per_launch": {"imad_wide": n_wide / k["launches"], "imad": n_imad / k["launches"]},
        "butterfly": "wide-range (5 IMAD.WIDE + 4 IMAD)" if fast else "literal (6 IMAD.WIDE + 4 IMAD)",
        "butterflies_per_s": n_bf / secs,
        "peak_source": "scripts/int_peak.cu, register-resident chains, max over chains/thread x CTAs/SM: IMAD %.2f T/s, IMAD.WIDE (no addend) "
                       "%.2f T/s, IMAD.WIDE + 64-bit addend %.2f T/s, IMAD.HI %.2f T/s; register-resident butterflies %.0f G/s literal, "
                       "%.0f G/s wide-range" % (imad / 1e12, wide / 1e12, peaks["imad_wide_acc"] / 1e12, peaks["imad_hi"] / 1e12,
                                                peaks["butterfly_literal"] / 1e9, peaks["butterfly_wide"] / 1e9)})
    return roof


def two_pipe_bound(w, roof, B, clocks):
    """The dominant kernel against the two-pipe model of DESIGN.md section 6: scripts/sass_pipes.py counts, from the SASS
    of the library that just ran, the cycles one iteration of the kernel's hot loop keeps the multiplier (FMA-heavy) pipe
    and the ALU pipe busy per SM sub-partition (IMAD.WIDE / IMAD.HI 4 + 2, other IMAD 2 + 0, ALU instructions 0 + 2, measured
    by scripts/exp/pipes.cu); the larger of the two is the floor of an iteration.  frac = that floor over the measured
    cycles per iteration (launch time x SM clock under load / iterations per sub-partition)."""
    try:
        if roof.get("kernel") != "ks_mac" or w.k != 1 or not clocks or not clocks.get("sm_mhz"):
            return None
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import sass_pipes
        from gemini_seal_b200 import _lib
        hot = sass_pipes.hot_loop(_lib.LIB_PATH, ["ks_mac_kernelILi%dELb1ELb1" % w.logn])
        if hot is None:
            return None
        import torch
        sms = torch.cuda.get_device_properties(0).multi_processor_count
        ctas = B * (w.N // 2048) * w.E
        # iterations of the bundle loop per SM sub-partition; for L of the E slots one of the beta iterations has no
        # transform (the bundle that contains the slot): counted as 0.3 of an iteration (its 16 multiply-accumulates)
        warp_iters = ctas * 8 * (w.beta - 0.7 * w.L / w.E) / (sms * 4.0)
        measured = roof["launch_ms"] * 1e-3 * clocks["sm_mhz"] * 1e6 / warp_iters
        floor = max(hot["fma_cycles"], hot["alu_cycles"])
        return {"kernel_loop": "ks_mac bundle loop (one operand tile transform + 16 multiply-accumulates per thread)",
                "instructions_per_iteration": hot["instr"], "fma_pipe_cycles": hot["fma_cycles"], "alu_pipe_cycles": hot["alu_cycles"],
                "floor_cycles": floor, "measured_cycles": measured, "frac": floor / measured, "sm_mhz": clocks["sm_mhz"],
                "what": "static pipe occupancy of the loop body (scripts/sass_pipes.py on the loaded library) over measured cycles per iteration per SM sub-partition"}
    except Exception as e:  # the number is an explanation, never a reason to lose the bench line
        return {"error": repr(e)}


def run_c3(torch, g, lib, np, timed, world_rank_local, peaks, batch=256):
    """BASELINE config #3 as a measured line (1 GPU): device-resident mul+relin+rescale ops/s at N = 16384, with its own
    parity gate, kernel table and rooflines."""
    import synth
    w = workload("c3")
    ctx = g.Context(w.N, w.mods, w.k, device=world_rank_local, lib=lib)
    keys = synthetic_keys(w)
    ctx.set_relin_keys(keys[0])
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, keys[1])
    a = synth.residues_torch(torch, SEED_A, 0, batch, 2, w.mods[:w.L], w.N)
    b = synth.residues_torch(torch, SEED_B, 0, batch, 2, w.mods[:w.L], w.N)
    out = torch.empty((batch, 2, w.L - 1, w.N), dtype=torch.int64, device="cuda")
    rot = torch.empty_like(a)
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        rc = lib.gsb_mul_relin_rescale_dev(ctx.handle, None, a.data_ptr(), b.data_ptr(), out.data_ptr(), batch, w.L, stream)
        if rc != 0:
            raise RuntimeError(lib.gsb_last_error().decode())

    def step_rot():
        rc = lib.gsb_apply_galois_dev(ctx.handle, None, elt, a.data_ptr(), rot.data_ptr(), batch, w.L, stream)
        if rc != 0:
            raise RuntimeError(lib.gsb_last_error().decode())
    step()
    step_rot()
    torch.cuda.synchronize()
    sel = [0, batch - 1]
    rc = reference_context(w, keys, elt)
    ok, how = parity_gate(w, rc, 0, batch, out[sel].cpu().numpy().view(np.uint64), rot[sel].cpu().numpy().view(np.uint64), elt)
    del rc
    for _ in range(3):
        step()
    ms = timed(step, 10)
    ms_rot = timed(step_rot, 5)
    kernels = kernel_table(lib, step)
    res = {"workload": w.name(batch), "ops_per_s": batch * 10 / (ms / 1e3), "ms_per_step": ms / 10,
           "rotate_ops_per_s": batch * 5 / (ms_rot / 1e3), "parity_checked": {"cts": 2, "ok": ok, "against": how},
           "limb_ntts_per_op": w.ntts_mul_relin_rescale(), "kernels": kernels,
           "roofline": int_roofline(w, lib, ctx, kernels, batch, peaks)}
    hbm_peak, _ = measured_peaks()
    res["step_hbm"] = {"achieved": w.op_bytes(batch) * batch / (ms / 10 / 1e3) / 1e9, "peak": hbm_peak, "unit": "GB/s"}
    res["step_hbm"]["frac"] = res["step_hbm"]["achieved"] / hbm_peak
    ctx.close()
    if not ok:
        raise SystemExit("bench.py: C3 parity gate failed (%s)" % how)
    return res


def cpu_legs(w, rc, threads):
    """The reference's CPU path for the other two thirds of the metric, on this box's host cores (bounded samples)."""
    out = {"cores": threads}
    t = rc.time_ops(0, w.L, 2, 1, 1)
    out["mul_relin_rescale_ops_per_s_1thread"] = 2 / t
    n = max(threads, 2)
    rc.time_ops(1, w.L, min(n, threads), threads, 1)
    out["rotate_ops_per_s"] = 2 * n / rc.time_ops(1, w.L, 2 * n, threads, 1)
    out["rotate_ops_per_s_1thread"] = 2 / rc.time_ops(1, w.L, 2, 1, 1)
    for what, name in ((2, "limb_ntt_fwd_per_s"), (3, "limb_ntt_inv_per_s")):
        m = 400 * threads
        rc.time_ops(what, w.L, threads * 8, threads, 1)
        out[name] = m / rc.time_ops(what, w.L, m, threads, 1)
        out[name + "_1thread"] = 400 / rc.time_ops(what, w.L, 400, 1, 1)
    out["sample"] = ("seal::Evaluator / util::ntt_* of oracle/_ref: 2 ops on 1 thread; %d rotations on %d threads; %d strict limb-NTTs "
                     "(N = %d) per direction on %d threads, 400 on 1 thread" % (2 * n, threads, 400 * threads, w.N, threads))
    return out


def shim_numbers():
    """The drop-in C++ boundary measured by shim/shim_bench.cpp (gsb_seal::Evaluator beside seal::Evaluator on real
    seal objects): per-call latencies and the throughput of the batched device-resident route.  None where the
    binary (built against /root/reference) did not travel."""
    exe = os.path.join(ROOT, "oracle", "_ref", "shim_bench")
    if not os.path.exists(exe):
        return None
    try:
        p = subprocess.run([exe, "64", "7", "2"], capture_output=True, text=True, timeout=600)
        return json.loads(p.stdout.strip().splitlines()[-1])
    except (OSError, subprocess.TimeoutExpired, ValueError, IndexError) as e:
        return {"error": str(e)}


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import gemini_seal_b200 as g
    import synth
    from gemini_seal_b200 import _lib, shard

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL prints its version banner to stdout while the
        # communicator is created, so fd 1 points at stderr until that has happened
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    lib = _lib.load()
    w = workload("c4")
    N, L, B = w.N, w.L, args.batch
    ctx = g.Context(N, w.mods, w.k, device=local, lib=lib)
    # switching keys: the same synthetic limbs on every rank and in the reference (splitmix streams)
    keys = synthetic_keys(w)
    ctx.set_relin_keys(keys[0])
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, keys[1])
    # inputs: rank 0 draws the whole job on its GPU and scatters the shards over NCCL (plumbing; timed separately
    # as value_with_scatter_gather)
    ga = gb = None
    if world > 1:
        if rank == 0:
            ga = synth.residues_torch(torch, SEED_A, 0, world * B, 2, w.mods[:L], N)
            gb = synth.residues_torch(torch, SEED_B, 0, world * B, 2, w.mods[:L], N)
        a = shard.scatter_batch(ga, (B, 2, L, N), torch.int64, "cuda")
        b = shard.scatter_batch(gb, (B, 2, L, N), torch.int64, "cuda")
    else:
        a = synth.residues_torch(torch, SEED_A, 0, B, 2, w.mods[:L], N)
        b = synth.residues_torch(torch, SEED_B, 0, B, 2, w.mods[:L], N)
    out = torch.empty((B, 2, L - 1, N), dtype=torch.int64, device="cuda")
    rot_out = torch.empty_like(a)
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        rc = lib.gsb_mul_relin_rescale_dev(ctx.handle, None, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, L, stream)
        if rc != 0:
            raise RuntimeError(lib.gsb_last_error().decode())

    def step_rot():
        rc = lib.gsb_apply_galois_dev(ctx.handle, None, elt, a.data_ptr(), rot_out.data_ptr(), B, L, stream)
        if rc != 0:
            raise RuntimeError(lib.gsb_last_error().decode())

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- parity gate on the buffers that are timed, on every rank
    step()
    step_rot()
    torch.cuda.synchronize()
    sel = [0, B - 1]
    rc = reference_context(w, keys, elt)
    ok, how = parity_gate(w, rc, rank * B, B, out[sel].cpu().numpy().view(np.uint64), rot_out[sel].cpu().numpy().view(np.uint64), elt)
    flag = torch.tensor([1 if ok else 0, 1 if rc is not None else 0], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.SUM)
    parity = {"ranks": world, "cts_per_rank": 2, "ok": int(flag[0].item()) == world, "ranks_ok": int(flag[0].item()),
              "ranks_with_reference": int(flag[1].item()), "against": how, "ops": ["mul_relin_rescale", "rotate_vector(1)"]}
    if not parity["ok"]:
        if rank == 0:
            print(json.dumps({"metric": METRIC, "error": "parity gate failed", "parity_checked": parity}), flush=True)
        raise SystemExit("bench.py: parity gate failed on rank %d (%s)" % (rank, how))

    # ---- headline: device-resident steps
    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = lib.gsb_launch_count()
    ms = timed(step, args.steps)
    launches = lib.gsb_launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    value = world * B * args.steps / (ms / 1e3)

    # ---- end to end through the host-buffer C ABI: pinned host inputs, H2D + compute + D2H every step
    ha, hb = a.cpu().pin_memory(), b.cpu().pin_memory()
    hout = torch.empty((B, 2, L - 1, N), dtype=torch.int64).pin_memory()

    def step_e2e():
        rc_ = lib.gsb_mul_relin_rescale_host(ctx.handle, ha.data_ptr(), hb.data_ptr(), hout.data_ptr(), B, L)
        if rc_ != 0:
            raise RuntimeError(lib.gsb_last_error().decode())
    step_e2e()
    assert torch.equal(hout, out.cpu()), "host-buffer path and device-buffer path disagree"
    e2e_steps = max(1, min(args.steps, 5))
    ms_e2e = timed(step_e2e, e2e_steps)
    e2e_value = world * B * e2e_steps / (ms_e2e / 1e3)
    # the box's PCIe ceiling for exactly this traffic: every rank copies a step's input bytes host->device and a step's
    # output bytes device->host concurrently (two streams, pinned memory, no kernels), all ranks at once
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    da_, db_, dout_ = torch.empty_like(a), torch.empty_like(b), torch.empty_like(out)

    def step_copy():
        s_in.wait_stream(torch.cuda.current_stream())
        s_out.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s_in):
            da_.copy_(ha, non_blocking=True)
            db_.copy_(hb, non_blocking=True)
        with torch.cuda.stream(s_out):
            hout.copy_(dout_, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s_in)
        torch.cuda.current_stream().wait_stream(s_out)
    step_copy()
    ms_copy = timed(step_copy, e2e_steps)
    pcie = {"ms_per_step": ms_copy / e2e_steps, "h2d_gbs_aggregate": world * 2 * B * 2 * L * w.limb_bytes * e2e_steps / (ms_copy / 1e3) / 1e9,
            "d2h_gbs_aggregate": world * B * 2 * (L - 1) * w.limb_bytes * e2e_steps / (ms_copy / 1e3) / 1e9,
            "ops_per_s_ceiling": world * B * e2e_steps / (ms_copy / 1e3),
            "what": "the same pinned H2D + D2H bytes per step on every rank at once, no kernels: the host / PCIe ceiling of the e2e number"}
    del ha, hb, hout, da_, db_, dout_

    # ---- secondary numbers (short, outside the headline region)
    extra = {}
    extra["rotate_ops_per_s"] = world * B * 3 / (timed(step_rot, 3) / 1e3)
    work = a.clone()

    def step_ntt():
        lib.gsb_ntt_forward_dev(ctx.handle, work.data_ptr(), B * 2, L, 0, 0, stream)
        lib.gsb_ntt_inverse_dev(ctx.handle, work.data_ptr(), B * 2, L, 0, 0, stream)
    step_ntt()
    extra["limb_ntt_per_s"] = world * 2 * B * 2 * L * 3 / (timed(step_ntt, 3) / 1e3)
    assert torch.equal(work, a), "forward/inverse NTT round trip failed"
    del work

    # ---- sustained: the headline step for >= 3 s with the clocks sampled
    n_sus = max(args.steps, int(math.ceil(args.sustain_s / (ms / args.steps / 1e3)))) if args.sustain_s > 0 else 0
    if n_sus:
        s2 = ClockSampler(local)
        if rank == 0:
            s2.start()
        ms_sus = timed(step, n_sus)
        extra["sustained"] = {"ops_per_s": world * B * n_sus / (ms_sus / 1e3), "steps": n_sus, "seconds": ms_sus / 1e3,
                              "ms_per_step": ms_sus / n_sus, "clocks": s2.stop() if rank == 0 else None}

    # ---- multi-GPU: the same step with the NCCL scatter of the inputs and the gather of the results inside the
    # timed region (SURVEY 8(e) "with and without transfer time")
    if world > 1:
        def step_sg():
            a.copy_(shard.scatter_batch(ga, (B, 2, L, N), torch.int64, "cuda"))
            b.copy_(shard.scatter_batch(gb, (B, 2, L, N), torch.int64, "cuda"))
            step()
            shard.gather_batch(out)
        step_sg()
        sg_steps = max(1, min(args.steps, 5))
        ms_sg = timed(step_sg, sg_steps)
        extra["value_with_scatter_gather"] = {
            "ops_per_s": world * B * sg_steps / (ms_sg / 1e3), "ms_per_step": ms_sg / sg_steps, "steps": sg_steps,
            "what": "rank 0 scatters 2 x %d MiB of inputs per rank over NCCL, every rank runs its shard, rank 0 gathers %d MiB per rank"
                    % (B * 2 * L * w.limb_bytes >> 20, B * 2 * (L - 1) * w.limb_bytes >> 20)}
    del ga, gb

    peaks = None
    if rank == 0:
        import int_peak
        peaks, _sweep = int_peak.measure(local, quick=True)
    if world == 1:  # `timed` synchronises all ranks, so single-rank extras only run in the 1-GPU bench
        extra["c3_n16384"] = run_c3(torch, g, lib, np, timed, local, peaks)
        extra["bfv_n4096"] = bfv_config1(torch, g, lib, np, timed)
        extra["ntt_sweep"] = ntt_sweep(torch, g, lib, timed)
        # the raw strict NTT sweep against the multiplier-pipe peak: forward and inverse both run the wide-range
        # butterfly (5 IMAD.WIDE + 4 IMAD) for moduli with 10 q < 2^63 except in their global last stage
        for e in extra["ntt_sweep"].values():
            e["int_frac"] = e["butterflies_per_s"] * (5.0 / peaks["imad_wide"] + 4 / peaks["imad"])

    # ---- per-kernel CUDA-event timing of one more step -> roofline of the dominant kernel
    roofline, kernels, step_hbm = None, {}, None
    if rank == 0:
        kernels = kernel_table(lib, step)
        roofline = int_roofline(w, lib, ctx, kernels, B, peaks)
        hbm_peak, how_peak = measured_peaks()
        gbs = w.op_bytes(B) * B / (ms / args.steps / 1e3) / 1e9
        step_hbm = {"achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                    "bytes_per_op": w.op_bytes(B), "what": "SURVEY 8(d): (4L + 2(L-1) + 2 beta K / B) limbs per op over the whole step"}
        roofline["step_hbm"] = step_hbm
        roofline["step_int"] = {
            "limb_ntts_per_op": w.ntts_mul_relin_rescale(),
            "butterflies_per_s": w.ntts_mul_relin_rescale() * (N // 2) * w.logn * B / (ms / args.steps / 1e3),
            "frac_of_register_resident_wide_butterflies": w.ntts_mul_relin_rescale() * (N // 2) * w.logn * B / (ms / args.steps / 1e3) / peaks["butterfly_wide"],
            "what": "SURVEY 8(d) butterflies of the reference algorithm (302 limb-NTTs per op at C4) per second over the whole step, "
                    "against the register-resident wide-range butterfly loop"}
        extra["int_peaks"] = {k: v for k, v in peaks.items()}
        roofline["two_pipe"] = two_pipe_bound(w, roofline, B, clocks)

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = host_threads()
        if rc is not None:
            ops = max(threads, 2)
            rc.time_ops(0, L, min(ops, threads), threads, 1)  # warm-up
            n_ops = ops * 6
            secs = rc.time_ops(0, L, n_ops, threads, 1)
            cpu = {"value": n_ops / secs, "unit": "ops/s", "cores": threads, "kind": "reference",
                   "sample": "%d mul+relin+rescale ops on %d host threads (shared seal::Evaluator, disjoint ciphertexts), %.1f s wall = %.0f s of CPU work"
                             % (n_ops, threads, secs, secs * threads)}
            extra["cpu"] = cpu_legs(w, rc, threads)
        else:
            run, _keep = cpu_port_runner(w)
            secs = run(2, 1)
            cpu = {"value": 2 / secs, "unit": "ops/s", "cores": 1, "kind": "port", "sample": "2 mul+relin+rescale ops, oracle port, 1 thread"}
    del rc

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ctx.close()  # the shim program creates its own context and a 16 GiB workspace
        extra["shim"] = shim_numbers()
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "ops/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic", "config": config_dict(w, B, world),
            "e2e": {"value": e2e_value, "unit": "ops/s", "h2d_bytes_per_step": 2 * B * 2 * L * w.limb_bytes,
                    "d2h_bytes_per_step": B * 2 * (L - 1) * w.limb_bytes, "steps": e2e_steps, "ms_per_step": ms_e2e / e2e_steps,
                    "pcie_ceiling": pcie, "frac_of_pcie_ceiling": e2e_value / pcie["ops_per_s_ceiling"]},
            "gpu_launches": int(launches), "clocks": clocks, "parity_checked": parity, "roofline": roofline, "cpu_baseline": cpu,
            "extra": extra, "kernels": kernels,
        }
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="ciphertext pairs per GPU per step")
    ap.add_argument("--sustain-s", type=float, default=3.0, help="seconds of back-to-back headline steps for extra.sustained (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
