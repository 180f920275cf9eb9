#!/usr/bin/env python
"""Headline benchmark: batched CKKS multiply + relinearize + rescale_to_next at N = 2^15
(BASELINE.json configs #4/#5: 881-bit default modulus, L = 15, K = 16, k = 1), ciphertexts sharded
across GPUs, plus limb-NTT/s and rotations/s as secondary numbers.

    python bench.py --gpus 1 --steps 20 --warmup 3             # our arm
    python bench.py --impl reference --steps 2 --warmup 1      # the reference's CPU path on host cores
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field's definition.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "ckks_mul_relin_rescale_ops_per_s_n32768"
LOGN, N, L, K_KEY, K_SPECIAL = 15, 32768, 15, 16, 1
LIMB_BYTES = 8 * N


def workload_name(batch):
    return ("C4/C5: CKKS N=32768, 881-bit BFVDefault modulus (L=15, K=16, k=1, beta=15): batched multiply_inplace + "
            "relinearize_inplace + rescale_to_next_inplace, %d ciphertext pairs per GPU per step, synthetic "
            "uniform-residue ciphertexts and switching key" % batch)


def moduli():
    import golden_cases
    return golden_cases.BFV_DEFAULT[N]


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
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or not f[0].isdigit() or int(f[0]) != self.gpu:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- algorithmic bytes per kernel launch (DESIGN.md "Kernels and rooflines") ----------------------------
def algorithmic_bytes(label, B):
    """Unique bytes a launch must read + write for a batch of B ciphertexts at (L=15, k=1, beta=15, E=16)."""
    E, beta = L + K_SPECIAL, L
    pairs = beta * E - L  # (bundle, ext slot) pairs that need a mod-up + NTT
    limbs = {
        "ckks_multiply": B * (4 * L + 3 * L),
        "inv_contig/ks_target": B * 2 * L, "inv_cols/ks_target": B * 2 * L,
        "fwd_cols/modup": B * (L + pairs),
        "ks_mac": B * (pairs + L + 2 * E) + 2 * beta * E,  # + the switching key, read once per launch
        "inv_contig/ks_special": B * 2 * 2 * K_SPECIAL, "inv_cols/ks_special": B * 2 * 2 * K_SPECIAL,
        "fwd_cols/moddown": B * 2 * (K_SPECIAL + L),
        "fwd_contig/moddown": B * 2 * (L + L + L + L),  # tmp + red + addend in, result out
        "inv_contig/rescale_last": B * 2 * 2, "inv_cols/rescale_last": B * 2 * 2,
        "fwd_cols/rescale": B * 2 * (1 + (L - 1)),
        "fwd_contig/rescale": B * 2 * 3 * (L - 1),
        "fwd_cols/plain": B * 2, "fwd_contig/plain": B * 2, "inv_cols/plain": B * 2, "inv_contig/plain": B * 2,
    }
    return limbs.get(label, 0) * LIMB_BYTES


# forward butterfly stages each (bundle, ext slot) pair goes through in the two key-switch kernels at N = 2^15
BUTTERFLY_STAGES = {"ks_mac": 9, "fwd_cols/modup": 6}


def ncu_traffic(label, batch_per_launch):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `label` from the committed ncu --set full
    capture (profiles/r1_ncu_traffic.json, written by scripts/ncu_traffic.py), or None when that capture was
    taken at another batch per launch."""
    path = os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        t = json.load(f)
    k = t.get("kernels", {}).get(label)
    if not k or k.get("batch_per_launch") != batch_per_launch:
        return None
    return k["dram_bytes_per_launch"]


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---- reference arm / cpu baseline -------------------------------------------------------------------------
def cpu_runner():
    """(kind, fn(ops, threads) -> seconds) timing mul+relin+rescale on host cores: the unmodified
    reference when oracle/_ref travelled here, else the oracle port (single-threaded)."""
    from oracle import pyoracle
    mods = moduli()
    try:
        ref = pyoracle.RefLib()
        rc = ref.context(N, mods, K_SPECIAL)
        rc.make_synthetic_keys([], seed=7)
        return "reference", lambda ops, threads: rc.time_ops(0, L, ops, threads, 1), rc
    except (FileNotFoundError, OSError):
        import numpy as np
        import golden_cases
        from common import rand_key, rand_poly
        orc = pyoracle.Oracle().context(LOGN, mods, K_SPECIAL)
        rng = np.random.default_rng(1)
        a, b = rand_poly(rng, mods[:L], (2,), N), rand_poly(rng, mods[:L], (2,), N)
        key = rand_key(rng, mods, L, N)

        def run(ops, threads):
            t0 = time.perf_counter()
            for _ in range(ops):
                orc.mul_relin_rescale(a, b, key)
            return time.perf_counter() - t0
        return "port", run, orc


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind, run, _keep = cpu_runner()
    threads = host_threads() if kind == "reference" else 1
    ops_per_step = max(1, threads)  # one op per thread per step: a bounded sample of the workload
    for _ in range(args.warmup):
        run(ops_per_step, threads)
    secs = 0.0
    for _ in range(args.steps):
        secs += run(ops_per_step, threads)
    value = ops_per_step * args.steps / secs
    sample = "%d steps x %d mul+relin+rescale ops (one per thread) on identical synthetic inputs" % (args.steps, ops_per_step)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "ops/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_name(args.batch), "note": "reference CPU path (seal::Evaluator, unmodified) on host cores"},
        "cpu_baseline": {"value": value, "unit": "ops/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "ops/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- our arm ----------------------------------------------------------------------------------------------
def uniform_residues(torch, mods, lead, gen):
    """int64 CUDA tensor lead + [len(mods), N], limb i uniform in [0, mods[i])."""
    cols = [torch.randint(0, q, tuple(lead) + (N,), dtype=torch.int64, device="cuda", generator=gen) for q in mods]
    return torch.stack(cols, dim=-2).contiguous()


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
    sys.path.insert(0, os.path.join(ROOT, "tests"))
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

        def step():
            lib.gsb_ntt_forward_dev(ctx.handle, x.data_ptr(), polys, limbs, 0, 0, stream)
            lib.gsb_ntt_inverse_dev(ctx.handle, x.data_ptr(), polys, limbs, 0, 0, stream)
        step()
        per_s = 2 * polys * limbs * 3 / (timed(step, 3) / 1e3)
        assert torch.equal(x, ref), "NTT round trip failed at N=2^%d" % logn
        out["n%d" % n] = {"limb_ntt_per_s": per_s, "hbm_frac": per_s * 16 * n / 1e9 / peak,
                          "butterflies_per_s": per_s * (n // 2) * logn, "polys": polys, "limbs": limbs}
        ctx.close()
        del x, ref, cols
    return out


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import gemini_seal_b200 as g
    from gemini_seal_b200 import _lib

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
    mods = moduli()
    B = args.batch
    ctx = g.Context(N, mods, K_SPECIAL, device=local, lib=lib)
    ev = g.Evaluator(ctx)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1234)
    # switching keys: the same synthetic limbs on every rank (seeded), uploaded through the C ABI
    relin = uniform_residues(torch, mods, (L, 2), gen)
    ctx.set_relin_keys(relin.cpu().numpy().view(np.uint64))
    elt = ctx.galois_elt_from_step(1)
    ctx.set_galois_key(elt, uniform_residues(torch, mods, (L, 2), gen).cpu().numpy().view(np.uint64))
    del relin
    # inputs: rank 0 draws the whole job and scatters the shards over NCCL (untimed plumbing)
    from gemini_seal_b200 import shard
    if world > 1:
        ga = uniform_residues(torch, mods[:L], (world * B, 2), gen) if rank == 0 else None
        a = shard.scatter_batch(ga, (B, 2, L, N), torch.int64, "cuda")
        del ga
        gb = uniform_residues(torch, mods[:L], (world * B, 2), gen) if rank == 0 else None
        b = shard.scatter_batch(gb, (B, 2, L, N), torch.int64, "cuda")
        del gb
    else:
        a = uniform_residues(torch, mods[:L], (B, 2), gen)
        b = uniform_residues(torch, mods[:L], (B, 2), gen)
    out = torch.empty((B, 2, L - 1, N), dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        rc = lib.gsb_mul_relin_rescale_dev(ctx.handle, None, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, L, stream)
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
        rc = lib.gsb_mul_relin_rescale_host(ctx.handle, ha.data_ptr(), hb.data_ptr(), hout.data_ptr(), B, L)
        if rc != 0:
            raise RuntimeError(lib.gsb_last_error().decode())
    step_e2e()
    assert torch.equal(hout, out.cpu()), "host-buffer path and device-buffer path disagree"
    e2e_steps = max(1, min(args.steps, 5))
    ms_e2e = timed(step_e2e, e2e_steps)
    e2e_value = world * B * e2e_steps / (ms_e2e / 1e3)

    # ---- secondary numbers (short, outside the headline region)
    extra = {}
    rot_out = torch.empty_like(a)

    def step_rot():
        rc = lib.gsb_apply_galois_dev(ctx.handle, None, elt, a.data_ptr(), rot_out.data_ptr(), B, L, stream)
        if rc != 0:
            raise RuntimeError(lib.gsb_last_error().decode())
    step_rot()
    extra["rotate_ops_per_s"] = world * B * 3 / (timed(step_rot, 3) / 1e3)
    work = a.clone()

    def step_ntt():
        lib.gsb_ntt_forward_dev(ctx.handle, work.data_ptr(), B * 2, L, 0, 0, stream)
        lib.gsb_ntt_inverse_dev(ctx.handle, work.data_ptr(), B * 2, L, 0, 0, stream)
    step_ntt()
    extra["limb_ntt_per_s"] = world * 2 * B * 2 * L * 3 / (timed(step_ntt, 3) / 1e3)
    assert torch.equal(work, a), "forward/inverse NTT round trip failed"
    del work, rot_out
    if world == 1:  # `timed` synchronises all ranks, so single-rank extras only run in the 1-GPU bench
        extra["bfv_n4096"] = bfv_config1(torch, g, lib, np, timed)
        extra["ntt_sweep"] = ntt_sweep(torch, g, lib, timed)

    # ---- per-kernel CUDA-event timing of one more step -> roofline of the dominant kernel
    roofline, kernels = None, {}
    if rank == 0:
        lib.gsb_prof_enable(1)
        step()
        need = lib.gsb_prof_report(None, 0)
        buf = C.create_string_buffer(max(need, 16))
        lib.gsb_prof_report(buf, need)
        lib.gsb_prof_enable(0)
        kernels = json.loads(buf.value.decode())
        total = sum(k["ms"] for k in kernels.values()) or 1.0
        peak, how = measured_peaks()
        for lab, k in kernels.items():
            k["share"] = k["ms"] / total
            per = k["ms"] / k["launches"]
            chunk_b = B * 1.0  # launches per label == chunks; bytes scale with B / chunks
            k["algorithmic_gbs"] = algorithmic_bytes(lab, chunk_b) / k["launches"] / (per * 1e-3) / 1e9
        top = max(kernels, key=lambda x: kernels[x]["ms"])
        achieved = kernels[top]["algorithmic_gbs"]
        roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": ncu_traffic(top, B // kernels[top]["launches"]), "peak_source": how,
                    "kernel_share_of_step": kernels[top]["share"]}
        # The binding limit of these kernels is the integer multiplier ("FMA-heavy") pipe, not HBM (DESIGN.md
        # section 3).  int_pipe = the multiplies the kernel's ALGORITHM needs (butterflies and 64x64->128
        # multiply-accumulates only; no address, carry or move instructions), each weighted by its measured
        # register-resident issue rate (gsb_int_peak: IMAD.WIDE.U32/s and IMAD/s on this GPU, now), over the
        # kernel's measured time.  Unit: IMAD.WIDE-equivalents/s; it is the same quantity ncu reports as
        # sm__pipe_fmaheavy_cycles_active minus the non-multiply instructions ptxas places on that pipe.
        imad, wide, bf, bfw = C.c_double(), C.c_double(), C.c_double(), C.c_double()
        if (lib.gsb_int_peak(ctx.handle, 2, C.byref(imad), C.byref(wide), C.byref(bf)) == 0
                and lib.gsb_int_peak_wide(ctx.handle, 2, C.byref(bfw)) == 0 and top in BUTTERFLY_STAGES):
            fast = bool(lib.gsb_ctx_fast_ntt(ctx.handle))
            pairs = L * (L + K_SPECIAL) - L          # (bundle, ext slot) pairs that go through the NTT
            n_bf = B * pairs * BUTTERFLY_STAGES[top] * (N // 2)
            w_per_bf, i_per_bf = (5, 4) if fast else (6, 4)
            n_wide, n_imad = n_bf * w_per_bf, n_bf * i_per_bf
            if top == "ks_mac":
                n_wide += B * L * (L + K_SPECIAL) * N * 2 * 4      # two FMAU128 per (bundle, slot, coefficient)
                if not fast:
                    n_wide += B * pairs * N // 2 * 5               # reduceBarrettLazy of the last stage's x0
                    n_imad += B * pairs * N // 2 * 2
            secs = kernels[top]["ms"] * 1e-3
            t_peak = n_wide / wide.value + n_imad / imad.value
            roofline["int_pipe"] = {
                "achieved": t_peak * wide.value / secs / 1e12, "peak": wide.value / 1e12,
                "unit": "T IMAD.WIDE-equivalents/s", "frac": t_peak / secs,
                "multiplies_per_launch": {"imad_wide": n_wide / kernels[top]["launches"], "imad": n_imad / kernels[top]["launches"]},
                "butterfly": "wide-range (5 IMAD.WIDE + 4 IMAD)" if fast else "literal (6 IMAD.WIDE + 4 IMAD)",
                "butterflies_per_s": n_bf / secs,
                "peak_source": "gsb_int_peak: register-resident chains, 2 CTAs x 256 threads per SM: IMAD %.2f T/s, IMAD.WIDE %.2f T/s; "
                               "register-resident butterflies %.0f G/s literal, %.0f G/s wide-range"
                               % (imad.value / 1e12, wide.value / 1e12, bf.value / 1e9, bfw.value / 1e9)}
            # the raw strict NTT sweep against the same multiplier-pipe peak: literal forward butterflies
            # (6 IMAD.WIDE + 4 IMAD), wide-range inverse ones (5 + 4; 59-bit moduli satisfy 10 q < 2^63)
            for e in extra.get("ntt_sweep", {}).values():
                e["int_frac"] = e["butterflies_per_s"] * (5.5 / wide.value + 4 / imad.value)

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        kind, run, _keep = cpu_runner()
        threads = host_threads() if kind == "reference" else 1
        ops = max(threads, 2) if kind == "reference" else 2
        run(min(ops, threads), threads)  # warm-up
        n_ops = ops * 6 if kind == "reference" else ops  # ~15-20 s of CPU work on the reference, a few s on the port
        secs = run(n_ops, threads)
        cpu = {"value": n_ops / secs, "unit": "ops/s", "cores": threads, "kind": kind,
               "sample": "%d mul+relin+rescale ops on %d host threads (shared Evaluator, disjoint ciphertexts), %.1f s wall = %.0f s of CPU work"
                         % (n_ops, threads, secs, secs * threads)}

    if world > 1:
        # gather every shard's result on rank 0 (NCCL plumbing, untimed) and spot-check one of them
        step()
        allout = shard.gather_batch(out)
        if rank == 0:
            assert allout.shape[0] == world * B and torch.equal(allout[:B], out)
        del allout
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "ops/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": workload_name(B), "batch_per_gpu": B, "parallelism": "ciphertext-sharded x%d, no data-path collective" % world,
                       "l2": "inputs %d MiB per step per GPU, larger than the 126 MiB L2" % (2 * B * 2 * L * LIMB_BYTES >> 20)},
            "e2e": {"value": e2e_value, "unit": "ops/s", "h2d_bytes_per_step": 2 * B * 2 * L * LIMB_BYTES,
                    "d2h_bytes_per_step": B * 2 * (L - 1) * LIMB_BYTES, "steps": e2e_steps, "ms_per_step": ms_e2e / e2e_steps},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
