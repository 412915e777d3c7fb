#!/usr/bin/env python
"""bench.py -- KS+PBS/s at PARAM_MESSAGE_2_CARRY_2 (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" = one pass of the hot path (shortint apply_lookup_table core: keyswitch + programmable bootstrap) over one
batch of `--batch` synthetic ciphertexts PER GPU (weak scaling: independent ciphertexts shard with no collective;
keys are generated on rank 0 and broadcast once over NCCL).  One JSON line on stdout from rank 0:
  value          device-resident throughput (inputs already in HBM), CUDA events on the library's stream, max over ranks
  e2e            the same metric through the host-pointer C-ABI call (pinned host buffers, H2D + D2H inside the timed region)
  roofline       blind-rotation kernel alone: algorithmic FP64 flop per launch / event time, against the FP64 FMA peak
                 measured live by the library's probe kernel (MEASURED_PEAKS.json carries no FP64 figure); also the HBM view
  cpu_baseline   the CPU oracle (a C port of the reference path; the Rust reference cannot be built here) on all host cores
`--impl reference` times that CPU port alone (see DESIGN.md "reference arm").
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PARAM_NAME = "PARAM_MESSAGE_2_CARRY_2_KS_PBS"
ORACLE_NAME = "MESSAGE_2_CARRY_2"
METRIC = "KS+PBS/sec at PARAM_MESSAGE_2_CARRY_2"
UNIT = "KS+PBS/s"


def flop_per_pbs(n, k, N, l):
    """per CMUX (k+1)(l+1)*5*M*log2(M) + (k+1)^2*l*M*8 + (k+1)(l+1)*M*6, M = N/2; times n."""
    M = N // 2
    lg = M.bit_length() - 1
    return n * ((k + 1) * (l + 1) * 5 * M * lg + (k + 1) ** 2 * l * M * 8 + (k + 1) * (l + 1) * M * 6)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_throughput(seconds_target=12.0, cores=None):
    """The CPU oracle's batched KS+PBS (one ciphertext per task over all host cores, the structure of the reference's
    PBS_throughput bench, tfhe/benches/core_crypto/pbs_bench.rs:517-532) on a bounded sample of the workload."""
    import numpy as np
    from oracle.oracle import Oracle
    cores = cores or os.cpu_count() or 1
    o = Oracle(ORACLE_NAME, nthreads=cores).keygen()
    lut = o.lut(lambda x: x)
    rng = np.random.default_rng(0xB2000002)
    probe = max(cores, 4)
    cts = o.encrypt(o.encode(rng.integers(0, o.pmod, probe)), "big")
    t0 = time.perf_counter(); o.ks_pbs(cts, lut); t1 = time.perf_counter()
    per_round = t1 - t0
    count = int(max(probe, min(4096, probe * max(1.0, seconds_target / max(per_round, 1e-3)))))
    count = (count // cores) * cores or cores
    msgs = rng.integers(0, o.pmod, count)
    cts = o.encrypt(o.encode(msgs), "big")
    t0 = time.perf_counter(); out = o.ks_pbs(cts, lut); t1 = time.perf_counter()
    ok = bool((o.decode(o.phase(out, "big")) % o.pmod == msgs).all())
    return {"value": count / (t1 - t0), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{count} ciphertexts of the same workload (KS+PBS, identity LUT), {t1 - t0:.1f} s, decrypt-checked={ok}; "
                      f"the Rust reference is unbuildable here (no rustc); its published single-core figure is 16.6 ms/KS+PBS"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot be built in this image
    (no rustc/cargo, dependencies not vendored), so the timed code is oracle/ (C restatement), all host threads."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = 12.0 / max(1, args.steps + args.warmup)
    vals, last = [], None
    for i in range(args.warmup + args.steps):
        last = cpu_port_throughput(seconds_target=max(2.0, per_step), cores=cores)
        if i >= args.warmup:
            vals.append(last["value"])
    v = sum(vals) / len(vals)
    last["value"] = v
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic",
                      "config": {"workload": f"shortint {PARAM_NAME} apply_lookup_table (KS+PBS), bounded sample per step", "inputs": "host"},
                      "cpu_baseline": last,
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


class DevView:
    """Wrap a raw device pointer so torch can address it (NCCL broadcast / random fill of library-owned key buffers)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 8,), "typestr": "<i8", "data": (ptr, False), "version": 3}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=16384, help="ciphertexts per GPU per step (BASELINE.json configs[1])")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    args.warmup = max(args.warmup, 3)
    # libraries (NCCL's version banner, torchrun notices) write to fd 1: keep stdout for the ONE JSON line
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import numpy as np
    import torch
    import torch.distributed as dist
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import shortint

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    ctx = t.Context(local)
    lib = ctx._lib
    p = t.parameters.get(PARAM_NAME)
    n, k, N, l = p.lwe_dimension, p.glwe_dimension, p.polynomial_size, p.pbs_level
    big, small, glwe = k * N + 1, n + 1, (k + 1) * N
    B = args.batch

    # ---- keys: synthetic (uniform torus values: timing is data independent); rank 0 creates, NCCL broadcasts once ----
    bsk = t.FourierLweBootstrapKey.alloc(ctx, p)
    ksk = t.LweKeyswitchKey.alloc(ctx, p)
    g = torch.Generator(device=dev); g.manual_seed(0xB2000001)
    bptr, bbytes = bsk.device_buffer()
    kptr, kbytes = ksk.device_buffer()
    bview = torch.as_tensor(DevView(bptr, bbytes), device=dev)
    kview = torch.as_tensor(DevView(kptr, kbytes), device=dev)
    if rank == 0:
        std = torch.randint(-2 ** 63, 2 ** 63 - 1, (n * l * (k + 1) ** 2 * N,), dtype=torch.int64, device=dev, generator=g)
        tmp = t.convert_standard_lwe_bootstrap_key_to_fourier_dev(ctx, p, std.data_ptr())
        tptr, _ = tmp.device_buffer()
        bview.copy_(torch.as_tensor(DevView(tptr, bbytes), device=dev))
        torch.cuda.synchronize(); tmp.close(); del std
        kview.copy_(torch.randint(-2 ** 63, 2 ** 63 - 1, (kbytes // 8,), dtype=torch.int64, device=dev, generator=g))
    if world > 1:
        dist.broadcast(bview, 0); dist.broadcast(kview, 0)
    torch.cuda.synchronize()
    ksk.finalize()

    sk = shortint.ServerKey(ksk, bsk, 4, 4, "ks_pbs")
    lut = sk.generate_lookup_table(lambda x: x)
    d_lut = torch.from_numpy(lut.acc.view(np.int64)).to(dev)
    g.manual_seed(0xB2000002 + rank)
    d_in = torch.randint(-2 ** 63, 2 ** 63 - 1, (B, big), dtype=torch.int64, device=dev, generator=g)
    d_out = torch.empty((B, big), dtype=torch.int64, device=dev)
    d_small = torch.randint(-2 ** 63, 2 ** 63 - 1, (B, small), dtype=torch.int64, device=dev, generator=g)
    h_in = torch.empty((B, big), dtype=torch.int64).pin_memory(); h_in.copy_(d_in)
    h_out = torch.empty((B, big), dtype=torch.int64).pin_memory()
    torch.cuda.synchronize()

    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    VP = C.c_void_p

    def step_dev():
        ctx.check(lib.b200tfhe_ks_pbs_batch_dev(ctx._h, ksk._h, bsk._h, VP(d_in.data_ptr()), VP(d_out.data_ptr()),
                                                VP(d_lut.data_ptr()), None, B))

    def step_br():
        ctx.check(lib.b200tfhe_pbs_batch_dev(ctx._h, bsk._h, VP(d_small.data_ptr()), VP(d_out.data_ptr()), VP(d_lut.data_ptr()), None, B))

    def step_ks():
        ctx.check(lib.b200tfhe_keyswitch_batch_dev(ctx._h, ksk._h, VP(d_in.data_ptr()), VP(d_small.data_ptr()), B))

    def step_e2e():
        ctx.check(lib.b200tfhe_ks_pbs_batch_u64(ctx._h, ksk._h, bsk._h, VP(h_in.data_ptr()), VP(h_out.data_ptr()),
                                                VP(lut.acc.ctypes.data), 1, None, B))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(local)
    l0 = ctx.launch_count
    sampler.start()
    ms_total = timed(step_dev, args.steps, args.warmup)
    clocks = sampler.stop()
    launches = (ctx.launch_count - l0) // (args.steps + args.warmup) * args.steps   # kernels of the timed steps only
    ms_e2e = timed(step_e2e, args.steps, args.warmup)
    br_steps = max(2, min(args.steps, 3))
    ms_br = timed(step_br, br_steps, 1) / br_steps
    ms_ks = timed(step_ks, br_steps, 1) / br_steps
    # single-PBS latency (the metric's second half): batch of ONE ciphertext through the same KS+PBS entry point
    lat = []
    lat_sampler = ClockSampler(local)
    lat_sampler.start()
    for it in range(210):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.check(lib.b200tfhe_ks_pbs_batch_dev(ctx._h, ksk._h, bsk._h, VP(d_in.data_ptr()), VP(d_out.data_ptr()),
                                                VP(d_lut.data_ptr()), None, 1))
        e1.record(stream)
        torch.cuda.synchronize()
        if it >= 10:
            lat.append(e0.elapsed_time(e1))
    lat_clocks = lat_sampler.stop()
    lat_order = list(lat)
    lat.sort()
    tf = C.c_double()
    ctx.check(lib.b200tfhe_probe_fp64_tflops(ctx._h, C.c_double(300.0), C.byref(tf)))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_step = ms_total / args.steps
    value = world * B * args.steps / (ms_total * 1e-3)
    e2e = world * B * args.steps / (ms_e2e * 1e-3)
    fl = flop_per_pbs(n, k, N, l)
    achieved = B * fl / (ms_br * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    bsk_bytes, ksk_bytes = bbytes, kbytes
    hbm_alg = (B * (small + big) * 8 + bsk_bytes) / (ms_br * 1e-3) / 1e9   # minimum HBM traffic of one blind-rotation launch
    lat_max = int(os.environ.get("B200TFHE_LAT_MAX", 2 * 148))       # dispatch of csrc/blind_rotate.cu::br_launch
    br_kernel = ("blind_rotate_lat_kernel" if args.batch <= lat_max
                 else "blind_rotate_kernel" if os.environ.get("B200TFHE_NO_TMEM") else "blind_rotate_tmem_kernel")
    traffic = None   # DRAM bytes of one launch of that kernel, from the committed ncu capture (same kernel, same batch)
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json"))).get(br_kernel)
        if tr and tr["batch"] == B:
            traffic = tr["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"shortint {PARAM_NAME} apply_lookup_table (keyswitch + PBS), batch {B} per GPU",
                   "batch_per_gpu": B, "l2_policy": f"inputs+outputs {2 * B * big * 8 / 2 ** 20:.0f} MiB per step > 126 MB L2; keys stay L2-resident by design",
                   "keys": "uniform synthetic BSK/KSK generated on rank 0, NCCL-broadcast once"},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": B * big * 8 + glwe * 8, "d2h_bytes_per_step": B * big * 8,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "latency": {"what": "single KS+PBS (batch 1), device-resident, CUDA events", "p50_ms": lat[len(lat) // 2],
                    "p99_ms": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "min_ms": lat[0], "samples": len(lat),
                    "first20_median_ms": sorted(lat_order[:20])[10], "last20_median_ms": sorted(lat_order[-20:])[10],
                    "sm_mhz_during": lat_clocks.get("sm_mhz"),
                    "reference_published_ms": 16.6},
        "roofline": {"bound": "fp64",
                     "kernel": br_kernel, "achieved": achieved, "peak": tf.value, "unit": "TFLOP/s",
                     "frac": achieved / tf.value if tf.value else None, "traffic": traffic,
                     "peak_source": "live b200tfhe_probe_fp64_tflops (MEASURED_PEAKS.json has no FP64 figure)",
                     "flop_per_pbs": fl, "ms_per_launch": ms_br, "pbs_per_s_kernel_only": B / (ms_br * 1e-3),
                     "hbm": {"achieved": hbm_alg, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_alg / hbm_peak,
                             "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                     "keyswitch_ms_per_launch": ms_ks, "keyswitch_share": ms_ks / (ms_ks + ms_br)},
    }
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_port_throughput(seconds_target=12.0)
    real_stdout.write(json.dumps(out) + "\n")
    real_stdout.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
