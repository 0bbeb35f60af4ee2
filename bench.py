#!/usr/bin/env python
"""Benchmark of the KMeans Lloyd fit hot path (BASELINE.json metric: fit rows*iterations/s).

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port on the host cores

A "step" is one Lloyd iteration (assignment + per-cluster accumulation + all-reduce + centroid update)
over the configuration BASELINE.json's metric is quoted on: C2 = KMeans k=20 on 50M x 100 fp32 Gaussian
blobs, the rows split evenly across the N GPUs (total work fixed -> "strong" scaling).  `value` is the
whole-job rate with X already resident in HBM; `e2e` is the same metric through the plugin's fit-function
slot with HOST (pinned) partitions: host->device load, K iterations, centroids back to the host.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (rows_total, d, k, description)
    "c1": (100_000, 20, 8, "C1: KMeans k=8, 100k x 20 fp32 blobs"),
    "c2": (50_000_000, 100, 20, "C2: KMeans k=20, 50M x 100 fp32 blobs, row-sharded"),
    "c3": (5_000_000, 256, 1000, "C3: KMeans k=1000, 5M x 256 fp32 blobs"),
}
METRIC = "kmeans_fit_rows_iters_per_s"
UNIT = "rows*iters/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--rows", type=int, default=0, help="override total rows (debug only; marks the line)")
    ap.add_argument("--path", default="auto", choices=["auto", "ffma", "tc", "tc_split"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows per rank for the e2e leg (default: the shard)")
    return ap.parse_args()


def load_peaks():
    """(HBM GB/s, dense bf16 TFLOP/s sustained, source): the driver-measured peaks of this pool's B200s."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return (float(j["hbm_gbs"]), float(j.get("bf16_tflops_sustained", j.get("bf16_tflops", 1392.4))),
                "measured (MEASURED_PEAKS.json)")
    return 6650.0, 1392.4, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
def gen_blobs_device(torch, n, d, k, seed, rank, device):
    """Gaussian blobs generated on the device, shard by shard (reference recipe: shared centres,
    per-partition seeds; python/benchmark/gen_data_distributed.py:123-137,152-157)."""
    g0 = torch.Generator(device=device).manual_seed(seed)
    centers = torch.rand((k, d), generator=g0, device=device) * 20.0 - 10.0  # same on every rank
    g = torch.Generator(device=device).manual_seed(seed * 1000003 + rank + 1)
    X = torch.empty((n, d), dtype=torch.float32, device=device)
    chunk = 1 << 20
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        which = torch.randint(0, k, (e - s,), generator=g, device=device)
        X[s:e] = torch.randn((e - s, d), generator=g, device=device)
        X[s:e] += centers[which]
    return X


def run_reference(args):
    """CPU arm: the oracle port of the Lloyd iteration on all host cores, bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline as cb

    rows_total, d, k, desc = CONFIGS[args.config]
    sample_rows = min(rows_total, 2_000_000 if d * k <= 4000 else 200_000)
    iters = max(1, args.steps)
    if args.warmup > 0:
        cb.time_fit(min(sample_rows, 100_000), d, k, min(args.warmup, 2))
    value, secs, threads = cb.time_fit(sample_rows, d, k, iters)
    sample = (f"{sample_rows} of {rows_total} rows x {iters} Lloyd iterations, NumPy float64 oracle port over "
              f"{threads} row blocks; pyspark.ml baseline not runnable on this box: no JVM/pyspark in image")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": iters, "warmup": args.warmup, "ms_per_step": 1e3 * secs / iters * (rows_total / sample_rows),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "rows_total": rows_total, "d": d, "k": k, "sample_rows": sample_rows},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    from spark_rapids_ml_nai_b200 import _native
    from spark_rapids_ml_nai_b200.clustering import kmeans_fit_partitions

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun for --gpus > 1 (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    rows_total, d, k, desc = CONFIGS[args.config]
    if args.rows:
        rows_total = args.rows
        desc += f" (DEBUG rows={rows_total})"
    n_local = rows_total // world + (1 if rank < rows_total % world else 0)

    # ---- communicator bootstrap: rank 0's NCCL id reaches the others through the host-side store ---
    uid = None
    if world > 1:
        obj = [_native.get_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        uid = obj[0]
    h = _native.Handle(device=local_rank, rank=rank, nranks=world, uid=uid)
    h.set_path({"auto": _native.PATH_AUTO, "ffma": _native.PATH_FFMA, "tc": _native.PATH_TC,
                "tc_split": _native.PATH_TC_SPLIT}[args.path])

    X = gen_blobs_device(torch, n_local, d, k, seed=1, rank=rank, device=device)
    init = h.init_centroids(X, k, mode=_native.INIT_RANDOM, seed=1)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up: W untimed iterations -----------------------------------------------------------
    if args.warmup > 0:
        h.fit(X, init, max_iter=args.warmup, tol=0.0, want_inertia=False)
    barrier()

    # ---- timed: exactly K iterations, device-resident -------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    h.profile_enable(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    C, _, n_iter, _ = h.fit(X, init, max_iter=args.steps, tol=0.0, want_inertia=False)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    kern_ms, kern_n, launches = h.profile_read()
    h.profile_enable(False)
    refined_rows, last_kernel = h.screen_stats()
    clocks = sampler.stop() if rank == 0 else None
    assert n_iter == args.steps, (n_iter, args.steps)
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = rows_total * args.steps / (ms_max * 1e-3)
    path = h.last_path()

    # ---- roofline of the dominant kernel (this rank's launches; bytes = 4*d per row) -------------------
    peak, peak_tf, peak_src = load_peaks()
    roofline = None
    if kern_n > 0:
        kms = kern_ms / kern_n
        achieved = (4.0 * d * n_local) / (kms * 1e-3) / 1e9
        traffic = None
        kernel = {0: "assign_ffma_kernel", 1: "kmeans_fused_tc_kernel", 2: "kmeans_screen_kernel"}[last_kernel]
        tpath = os.path.join(ROOT, "profiles", "r01_traffic.json")
        if os.path.exists(tpath):
            t = json.load(open(tpath)).get(kernel)
            if t and (t["d"], t["k"]) == (d, k):  # bytes per row from the committed ncu capture, scaled to this launch
                traffic = (t["dram_bytes_read"] + t["dram_bytes_write"]) / t["rows"] * n_local
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic, "kernel": kernel,
                    "rows_decided_exactly_per_pass": refined_rows / max(1, args.steps),
                    "kernel_ms": kms, "launches_timed": kern_n, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": 4.0 * d * n_local}
        # 2*k*d FLOP against 4*d bytes per row: past the machine balance the contraction, not HBM, bounds the
        # pass (C3 / C5); report it against the measured dense tensor peak instead
        if 2.0 * k * d / (4.0 * d) > peak_tf * 1e3 / peak:
            tf = 2.0 * k * d * n_local / (kms * 1e-3) / 1e12
            roofline.update({"bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
                             "algorithmic_flops_per_launch": 2.0 * k * d * n_local,
                             "note": "peak = measured dense bf16 (sustained); the FP32 FFMA path cannot approach it"})

    # ---- e2e: plugin fit-function slot with host partitions --------------------------------------------
    e2e = None
    if not args.no_e2e:
        import psutil

        e_rows = args.e2e_rows or n_local
        need = e_rows * d * 4
        avail = psutil.virtual_memory().available
        if need * 1.5 > avail:
            e_rows = int(avail / 3 / (d * 4))
        while True:
            try:
                host = torch.empty((e_rows, d), dtype=torch.float32, pin_memory=True)
                break
            except RuntimeError:
                e_rows //= 2  # not enough pinnable host memory on this box: shrink the e2e sample
                if e_rows < 1024:
                    raise
        host.copy_(X[:e_rows])
        torch.cuda.synchronize()
        n_parts = 8
        bounds = [e_rows * i // n_parts for i in range(n_parts + 1)]
        init_host = init.cpu().numpy()
        cuml_params = {"n_clusters": k, "max_iter": args.steps, "tol": 0.0, "init": init_host, "random_state": 1}
        del X
        torch.cuda.empty_cache()

        def one_e2e():
            parts = [host[bounds[i]:bounds[i + 1]].numpy() for i in range(n_parts)]
            return kmeans_fit_partitions(parts, h, cuml_params, num_cols=d)

        one_e2e()  # warm (allocator, staging)
        barrier()
        t0 = time.perf_counter()
        res = one_e2e()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=device)
        rows_e2e = torch.tensor([e_rows], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(rows_e2e, op=dist.ReduceOp.SUM)
        e2e = {"value": float(rows_e2e.item()) * res["n_iter"] / float(t.item()), "unit": UNIT,
               "h2d_bytes_per_step": e_rows * d * 4 / res["n_iter"],
               "d2h_bytes_per_step": (k * d * 4 + 16) / res["n_iter"],
               "seconds": float(t.item()), "rows": int(rows_e2e.item()), "iterations": res["n_iter"],
               "note": "one plugin fit call: pinned host partitions -> device, K Lloyd iterations, centroids -> host"}

    # ---- CPU baseline beside it (rank 0, N=1 only) -------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb

        sample_rows = min(rows_total, 2_000_000 if d * k <= 4000 else 200_000)
        v, secs, threads = cb.time_fit(sample_rows, d, k, 4)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{sample_rows} of {rows_total} rows x 4 Lloyd iterations ({secs:.1f} s), NumPy float64 "
                         f"oracle port; pyspark.ml baseline not runnable on this box: no JVM/pyspark in image"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "rows_total": rows_total, "rows_per_gpu": n_local, "d": d, "k": k,
                       "path": path, "init": "random rows (seed 1), identical on all ranks", "tol": 0.0,
                       "l2": f"per-GPU input {n_local * d * 4 / 1e9:.2f} GB " +
                             ("is larger than the 126 MB L2, no flush needed" if n_local * d * 4 > 4e8
                              else "FITS in L2: latency-bound config, flagged"),
                       "arithmetic": "fp32 data; tc: TF32 tensor-core screen with a rigorous error bound, rows inside "
                                     "it decided exactly (FP32 differences / float64); ffma: FP32 products; "
                                     "float64 sums"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    h.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
