#!/usr/bin/env python
"""Benchmark of the KMeans hot path (BASELINE.json metric: fit rows*iterations/s; predict rows/s for C4).

    python bench.py --gpus N --steps K --warmup W [--config c1|c2|c3|c4|c5]   # our CUDA path (one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...                    # CPU arm on the host cores

A "step" of a fit configuration is one Lloyd iteration (assignment + per-cluster accumulation + all-reduce +
centroid update) over the whole data set, the rows split evenly across the N GPUs (total work fixed -> "strong"
scaling); a step of the predict configuration (C4) is one labelling pass over all rows.  The default is the
configuration BASELINE.json's metric is quoted on: C2 = KMeans k=20 on 50M x 100 fp32 Gaussian blobs.

The synthetic rows are a function of their GLOBAL row index (one generator seed per 2^20-row chunk), so every
world size fits the very same data: the centroid checksum and the inertia of the line agree across N.
`value` is the whole-job rate with X already resident in HBM; `e2e` is the same metric through the plugin's
executor-side path with HOST data: pageable Arrow-sized batches -> pinned staging -> device, K iterations,
centroids back to the host.  After the timed region the line's `parity_check` compares the path's labels (and,
for a fit, one accumulation step) on a row subsample against the float64 oracle.  With the default
configuration the other BASELINE.json configurations that fit this world size are measured too (`also`).
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
    "c1": dict(kind="fit", rows=100_000, d=20, k=8, desc="C1: KMeans k=8, 100k x 20 fp32 blobs"),
    "c2": dict(kind="fit", rows=50_000_000, d=100, k=20, desc="C2: KMeans k=20, 50M x 100 fp32 blobs, row-sharded"),
    "c3": dict(kind="fit", rows=5_000_000, d=256, k=1000, desc="C3: KMeans k=1000, 5M x 256 fp32 blobs"),
    "c4": dict(kind="predict", rows=200_000_000, d=64, k=100,
               desc="C4: KMeansModel.transform predict k=100, 200M x 64 fp32 blobs, row-sharded"),
    "c5": dict(kind="fit", rows=20_000_000, d=1024, k=10_000,
               desc="C5: KMeans k=10000, 20M x 1024 fp32 blobs, row-sharded"),
}
FIT_METRIC, FIT_UNIT = "kmeans_fit_rows_iters_per_s", "rows*iters/s"
PRED_METRIC, PRED_UNIT = "kmeans_predict_rows_per_s", "rows/s"
CHUNK_ROWS = 1 << 20
KERNELS = {0: "assign_ffma_kernel", 1: "kmeans_fused_tc_kernel", 2: "kmeans_screen_kernel", 3: "kmeans_gemm_kernel",
           4: "kmeans_small_fit_kernel"}
NO_SPARK = "pyspark.ml baseline not runnable on this box: no JVM/pyspark in image"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--rows", type=int, default=0, help="override total rows (debug only; marks the line)")
    ap.add_argument("--path", default="auto", choices=["auto", "ffma", "tc", "tc_split", "gemm", "small"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the other BASELINE.json configurations")
    ap.add_argument("--no-parity", action="store_true")
    return ap.parse_args()


def load_peaks():
    """(HBM GB/s, dense bf16 TFLOP/s sustained, source): the driver-measured peaks of this pool's B200s."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return (float(j["hbm_gbs"]), float(j.get("bf16_tflops_sustained", j.get("bf16_tflops", 1392.4))),
                "measured (MEASURED_PEAKS.json)")
    return 6650.0, 1400.0, "fallback (B200_PROFILING.md)"


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
def gen_blobs_device(torch, lo, hi, d, k, seed, device):
    """Rows [lo, hi) of the global Gaussian-blob data set (reference recipe: shared centres, cluster_std 1,
    python/benchmark/gen_data_distributed.py:123-137,152-157).  Every 2^20-row chunk has its own generator seed,
    so a row's value depends on its global index only, not on how the rows are split across ranks."""
    g0 = torch.Generator(device=device).manual_seed(seed)
    centers = torch.rand((k, d), generator=g0, device=device) * 20.0 - 10.0
    X = torch.empty((hi - lo, d), dtype=torch.float32, device=device)
    sub = max(1 << 14, min(CHUNK_ROWS, (1 << 26) // d))  # rows generated at a time (bounds the temporaries)
    for c in range(lo // CHUNK_ROWS, (hi - 1) // CHUNK_ROWS + 1):
        g = torch.Generator(device=device).manual_seed(seed * 1000003 + 7919 * c + 1)
        c_lo = c * CHUNK_ROWS
        for s in range(c_lo, c_lo + CHUNK_ROWS, sub):  # the chunk's stream is consumed in a fixed order
            e = s + sub
            which = torch.randint(0, k, (sub,), generator=g, device=device)
            if e <= lo or s >= hi:
                torch.randn((sub, d), generator=g, device=device)  # keeps the stream position; result dropped
                continue
            block = torch.randn((sub, d), generator=g, device=device)
            block += centers[which]
            a, b = max(s, lo), min(e, hi)
            X[a - lo:b - lo] = block[a - s:b - s]
    return X


def sample_for_cpu(cfg, budget_flops=4e10):
    """(rows, iterations) of the bounded CPU sample of a configuration."""
    rows_total, d, k = cfg["rows"], cfg["d"], cfg["k"]
    iters = 1 if cfg["kind"] == "predict" else (10 if k * d <= 4000 else 2)
    rows = int(min(rows_total, max(2000, budget_flops / (2.0 * k * d * iters))))
    if k * d <= 4000:
        rows = min(rows_total, 2_000_000)
    return rows, iters


def cpu_baseline(cfg, iters=None, rows=None):
    """The path on the host cores, on a bounded sample: scikit-learn's Lloyd (OpenMP, all cores, same initial
    centres, tol=0) and the NumPy float64 oracle port over row blocks; the faster one is reported."""
    import numpy as np
    from oracle import cpu_baseline as cb
    from oracle import kmeans_oracle as ko

    d, k = cfg["d"], cfg["k"]
    s_rows, s_iters = sample_for_cpu(cfg)
    rows = rows or s_rows
    iters = iters or s_iters
    threads = cb.host_threads()
    X, _ = ko.make_blobs(rows, d, k, seed=1)
    init = ko.random_init(X, k, seed=1)
    out = {}
    if cfg["kind"] == "fit":
        cb.lloyd_iterations(X[: max(threads * 1024, 4096)], init, 1, threads)
        t0 = time.perf_counter()
        cb.lloyd_iterations(X, init, iters, threads)
        out["oracle port (NumPy float64)"] = time.perf_counter() - t0
        try:
            from sklearn.cluster import KMeans as SkKMeans

            SkKMeans(n_clusters=k, init=init, n_init=1, max_iter=1, tol=0.0, algorithm="lloyd").fit(X[:8192])
            t0 = time.perf_counter()
            SkKMeans(n_clusters=k, init=init, n_init=1, max_iter=iters, tol=0.0, algorithm="lloyd").fit(X)
            out["scikit-learn Lloyd (float32, OpenMP)"] = time.perf_counter() - t0
        except Exception as e:  # pragma: no cover
            out["scikit-learn failed: " + repr(e)[:60]] = float("inf")
        work = rows * iters
    else:
        t0 = time.perf_counter()
        ko.predict(X, init)
        out["oracle port (NumPy float64)"] = time.perf_counter() - t0
        try:
            from sklearn.cluster import KMeans as SkKMeans

            km = SkKMeans(n_clusters=k, init=init, n_init=1, max_iter=1, tol=0.0, algorithm="lloyd").fit(X[:8192])
            km.cluster_centers_ = init.astype(km.cluster_centers_.dtype)
            km.predict(X[:8192])
            t0 = time.perf_counter()
            km.predict(X)
            out["scikit-learn predict (float32, OpenMP)"] = time.perf_counter() - t0
        except Exception as e:  # pragma: no cover
            out["scikit-learn failed: " + repr(e)[:60]] = float("inf")
        work = rows
    best = min(out, key=out.get)
    unit = FIT_UNIT if cfg["kind"] == "fit" else PRED_UNIT
    detail = "; ".join(f"{nm}: {work / sec:.3g} {unit}" for nm, sec in out.items() if sec != float("inf"))
    what = f"{rows} of {cfg['rows']} rows" + (f" x {iters} Lloyd iterations" if cfg["kind"] == "fit" else ", one pass")
    return {"value": work / out[best], "unit": unit, "cores": threads, "kind": "port",
            "sample": f"{what} ({out[best]:.1f} s), fastest of [{detail}] = {best}; {NO_SPARK}",
            "seconds": out[best], "rows": rows, "iters": iters}


def run_reference(args):
    """CPU arm: the path on all host cores (scikit-learn Lloyd / the oracle port, whichever is faster), on a
    bounded sample of the configuration."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    s_rows, _ = sample_for_cpu(cfg)
    iters = max(1, args.steps) if cfg["kind"] == "fit" else 1
    if cfg["kind"] == "fit" and cfg["k"] * cfg["d"] > 4000:
        iters = min(iters, 3)
    if args.warmup > 0:
        cpu_baseline(cfg, iters=1, rows=min(s_rows, 50_000))
    cpu = cpu_baseline(cfg, iters=iters, rows=s_rows)
    metric, unit = (FIT_METRIC, FIT_UNIT) if cfg["kind"] == "fit" else (PRED_METRIC, PRED_UNIT)
    line = {
        "impl": "reference", "metric": metric, "value": cpu["value"], "unit": unit, "n_gpus": args.gpus,
        "steps": iters, "warmup": args.warmup, "ms_per_step": 1e3 * cpu["seconds"] / iters,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 (scikit-learn) / f64 (oracle port)", "data": "synthetic",
        "config": {"workload": cfg["desc"], "rows_total": cfg["rows"], "d": cfg["d"], "k": cfg["k"],
                   "sample_rows": cpu["rows"],
                   "note": "ms_per_step is the time of one step over the SAMPLE; value = sample rows * steps / time"},
        "cpu_baseline": {kk: cpu[kk] for kk in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": cpu["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def parity_check(torch, h, X, C, cfg):
    """Labels of the path on a row subsample against the float64 oracle (rows outside BASELINE.json's 1e-5 band
    must agree exactly) and, for a fit, one accumulation step (new centroids within 1e-4)."""
    import numpy as np
    from oracle import kmeans_oracle as ko

    d, k = cfg["d"], cfg["k"]
    n = X.shape[0]
    m = int(min(n, max(2000, min(100_000, 3e10 / (k * d)))))
    idx = torch.arange(m, device=X.device, dtype=torch.int64) * (n // m)  # evenly spaced rows
    rows = X[idx].contiguous()
    lab = h.predict(rows, C)[0].cpu().numpy()
    Xh, Ch = rows.cpu().numpy(), C.cpu().numpy()
    l_ref, best, second = ko.assign(Xh, Ch)
    amb = ko.ambiguous_rows(best, second)
    out = {"rows": m, "mismatches_outside_band": int(((lab != l_ref) & ~amb).sum()), "rows_in_band": int(amb.sum()),
           "oracle": "NumPy float64 restatement (oracle/kmeans_oracle.py), subsample = evenly spaced rows of rank 0"}
    if cfg["kind"] == "fit":
        sums, counts, _, _ = h.lloyd_step(rows, C)
        s_ref, c_ref = ko.accumulate(Xh, lab, k)  # sums of the path's own labels
        newc, _ = ko.finalize(s_ref, c_ref, Ch)
        got, _ = ko.finalize(sums.cpu().numpy(), counts.cpu().numpy(), Ch)
        scale = np.maximum(np.abs(newc), np.linalg.norm(newc, axis=1, keepdims=True) / np.sqrt(d) + 1e-30)
        out["step_centroid_max_rel_err"] = float((np.abs(got - newc) / scale).max())
        out["step_counts_equal"] = bool(np.array_equal(counts.cpu().numpy(), c_ref))
    return out


def roofline_of(cfg, n_local, kms, kern_n, last_kernel, refined_rows, steps, peaks):
    peak, peak_tf, peak_src = peaks
    d, k = cfg["d"], cfg["k"]
    if kern_n <= 0:
        return None
    bytes_row = 4.0 * d + (4.0 if cfg["kind"] == "predict" else 0.0)
    achieved = bytes_row * n_local / (kms * 1e-3) / 1e9
    kernel = KERNELS[last_kernel]
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(tpath):
        t = json.load(open(tpath)).get(kernel)
        if t and (t["d"], t["k"]) == (d, k):  # bytes per row of the committed ncu capture, scaled to this launch
            traffic = (t["dram_bytes_read"] + t["dram_bytes_write"]) / t["rows"] * n_local
    r = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
         "traffic": traffic, "kernel": kernel, "rows_decided_exactly_per_pass": refined_rows / max(1, steps),
         "kernel_ms": kms, "launches_timed": kern_n, "peak_source": peak_src,
         "algorithmic_bytes_per_launch": bytes_row * n_local}
    # 2*k*d FLOP against 4*d bytes per row: past the machine balance the contraction, not HBM, bounds the pass
    # (C3 / C5).  kind::tf32 runs at half the bf16 rate, so the dense TF32 peak is taken as half the measured
    # bf16 figure; the tile GEMM issues two MMAs (c_hi, c_lo) per algorithmic one.
    tf32_peak = peak_tf / 2.0
    if 2.0 * k * d / (4.0 * d) > tf32_peak * 1e3 / peak:
        tf = 2.0 * k * d * n_local / (kms * 1e-3) / 1e12
        r.update({"bound": "tensor", "achieved": tf, "peak": tf32_peak, "unit": "TFLOP/s", "frac": tf / tf32_peak,
                  "algorithmic_flops_per_launch": 2.0 * k * d * n_local,
                  "tensor_pipe_frac_estimate": (2.0 if last_kernel == 3 else 1.0) * tf / tf32_peak,
                  "note": "peak = half the measured dense bf16 rate (sustained); achieved counts 2*n*k*d once, the "
                          "kernel issues the c_hi and c_lo products (2x) - tensor_pipe_frac_estimate; the measured "
                          "sm__pipe_tensor_cycles_active is in profiles/"})
    return r


def run_config(torch, dist, _native, h, name, args, steps, warmup, world, rank, device, peaks, with_e2e, with_cpu,
               with_parity, sampler=None):
    """Generates the configuration's shard, times `steps` steps device-resident, and returns the result dict."""
    cfg = dict(CONFIGS[name])
    if args.rows and name == args.config:
        cfg["rows"] = args.rows
        cfg["desc"] += f" (DEBUG rows={args.rows})"
    rows_total, d, k = cfg["rows"], cfg["d"], cfg["k"]
    lo = rows_total * rank // world
    hi = rows_total * (rank + 1) // world
    n_local = hi - lo
    X = gen_blobs_device(torch, lo, hi, d, k, seed=1, device=device)
    init = h.init_centroids(X, k, mode=_native.INIT_RANDOM, seed=1)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        t = torch.tensor([v], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    is_fit = cfg["kind"] == "fit"
    labels = None if is_fit else torch.empty(n_local, dtype=torch.int32, device=device)
    if warmup > 0:
        if is_fit:
            h.fit(X, init, max_iter=warmup, tol=0.0, want_inertia=False)
        else:
            for _ in range(warmup):
                h.predict(X, init, out=labels)
    barrier()
    if sampler is not None:
        sampler.start()
    h.profile_enable(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    if is_fit:
        C, _, n_iter, _ = h.fit(X, init, max_iter=steps, tol=0.0, want_inertia=False)
        assert n_iter == steps, (n_iter, steps)
    else:
        C = init
        for _ in range(steps):
            h.predict(X, C, out=labels)
    ev1.record()
    barrier()
    ms_max = allmax(ev0.elapsed_time(ev1))
    kern_ms, kern_n, launches = h.profile_read()
    h.profile_enable(False)
    refined_rows, last_kernel = h.screen_stats()
    clocks = sampler.stop() if sampler is not None else None
    work = rows_total * steps
    value = work / (ms_max * 1e-3)
    path = h.last_path()
    kms = kern_ms / max(1, kern_n)
    if last_kernel == 4:
        kms /= steps  # one launch runs every iteration: report the kernel time per iteration
    roofline = roofline_of(cfg, n_local, kms, kern_n, last_kernel, refined_rows, steps, peaks)
    if roofline and last_kernel == 4:
        roofline["note"] = ("one cooperative launch runs all iterations with the shard resident in shared memory: no HBM "
                            "traffic per iteration, kernel_ms is the launch time / iterations (latency-bound config)")

    # ---- result fingerprint: identical for every world size (the data is a function of the global row index)
    if is_fit:
        _, inertia, _, _ = h.fit(X, C, max_iter=0, tol=0.0, want_inertia=True)
        result = {"centroid_checksum": float(C.double().sum().item()), "inertia": float(inertia)}
    else:
        hist = torch.bincount(labels.long(), minlength=k).double()
        if world > 1:
            dist.all_reduce(hist)
        w = torch.arange(1, k + 1, dtype=torch.float64, device=device)
        result = {"label_checksum": float((hist * w).sum().item()), "rows_labelled": float(hist.sum().item())}

    parity = None
    if with_parity and rank == 0:
        parity = parity_check(torch, h, X, C, cfg)

    # ---- e2e: the plugin's executor-side path with HOST data ----------------------------------------------
    e2e = None
    if with_e2e:
        e2e = run_e2e(torch, dist, h, X, C if not is_fit else init, cfg, steps, world, device, barrier, allmax)
    del X
    torch.cuda.empty_cache()

    cpu = None
    if with_cpu and rank == 0 and world == 1:
        cpu = cpu_baseline(cfg)
        cpu = {kk: cpu[kk] for kk in ("value", "unit", "cores", "kind", "sample")}

    metric, unit = (FIT_METRIC, FIT_UNIT) if is_fit else (PRED_METRIC, PRED_UNIT)
    return {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_max / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["desc"], "rows_total": rows_total, "rows_per_gpu": n_local, "d": d, "k": k,
                   "path": path, "allreduce": h.last_reduce() if is_fit else "none (predict shards need no exchange)", "init": "random rows (seed 1), identical on all ranks", "tol": 0.0,
                   "data_keying": "row values are a function of the global row index (one seed per 2^20-row chunk)",
                   "l2": f"per-GPU input {n_local * d * 4 / 1e9:.2f} GB " +
                         ("is larger than the 126 MB L2, no flush needed" if n_local * d * 4 > 4e8
                          else "FITS in L2: latency-bound config, flagged"),
                   "arithmetic": "fp32 data; tc / gemm: TF32 tensor-core screen with a rigorous error bound, rows "
                                 "inside it decided exactly (FP32 differences / float64); ffma: FP32 products; "
                                 "float64 sums",
                   "result": result},
        "roofline": roofline, "parity_check": parity, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clocks,
    }


def run_e2e(torch, dist, h, X, C0, cfg, steps, world, device, barrier, allmax):
    """One call of the executor-side plugin path per rank, from pageable host memory: the shard arrives as
    Arrow-sized batches (spark.sql.execution.arrow.maxRecordsPerBatch = 10000 rows, the reference's default:
    python/src/spark_rapids_ml/clustering.py:776) of a multi-column numeric pandas frame, goes through the same
    batch -> ndarray helper as _train_udf / the transform loop, the packing loader and the C ABI."""
    import numpy as np
    import pandas as pd
    import psutil

    from spark_rapids_ml_nai_b200.clustering import _DevicePredictor, kmeans_fit_partitions
    from spark_rapids_ml_nai_b200.core import _batch_features

    d, k = cfg["d"], cfg["k"]
    n_local = X.shape[0]
    e_rows = n_local
    avail = psutil.virtual_memory().available / max(1, world)
    if e_rows * d * 4 * 1.3 > avail:
        e_rows = int(avail / 1.5 / (d * 4))
    host = np.empty((e_rows, d), dtype=np.float32)  # pageable, as a Spark Python worker holds its Arrow batches
    step_rows = 1 << 22
    for s in range(0, e_rows, step_rows):
        host[s:s + step_rows] = X[s:min(e_rows, s + step_rows)].cpu().numpy()
    cols = [f"c{i}" for i in range(d)]
    pdf = pd.DataFrame(host, columns=cols, copy=False)
    batch = 10_000
    C_host = C0.cpu().numpy()
    is_fit = cfg["kind"] == "fit"

    def one():
        feats = [_batch_features(pdf.iloc[s:s + batch], None, cols, "C", np.dtype(np.float32))
                 for s in range(0, e_rows, batch)]
        if is_fit:
            params = {"n_clusters": k, "max_iter": steps, "tol": 0.0, "init": C_host, "random_state": 1}
            res = kmeans_fit_partitions(feats, h, params, num_cols=d)
            return res["n_iter"], None
        pred = _DevicePredictor.__new__(_DevicePredictor)  # the task's backend object, on this rank's handle
        pred.handle, pred.device, pred.centers = h, device, C0
        out = []
        for s in range(0, len(feats), 128):  # the transform loop's coalescing: >= 2^20 rows per device round trip
            out.append(pred.predict_many(feats[s:s + 128]))
        return 1, out

    one()  # warm (staging buffers, allocator)
    barrier()
    t0 = time.perf_counter()
    n_it, _ = one()
    torch.cuda.synchronize()
    dt = allmax(time.perf_counter() - t0)
    rows_e2e = torch.tensor([e_rows], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(rows_e2e, op=dist.ReduceOp.SUM)
    tot = float(rows_e2e.item())
    if is_fit:
        return {"value": tot * n_it / dt, "unit": FIT_UNIT, "h2d_bytes_per_step": e_rows * d * 4 / n_it,
                "d2h_bytes_per_step": (k * d * 4 + 16) / n_it, "seconds": dt, "rows": int(tot), "iterations": n_it,
                "note": "one executor-side fit per rank: pageable Arrow-sized batches (10000 rows) via the "
                        "_train_udf batch helper -> packed pinned staging -> device, K Lloyd iterations, centroids -> host"}
    return {"value": tot / dt, "unit": PRED_UNIT, "h2d_bytes_per_step": e_rows * d * 4, "d2h_bytes_per_step": e_rows * 4,
            "seconds": dt, "rows": int(tot),
            "note": "one executor-side transform per rank: pageable Arrow-sized batches (10000 rows) coalesced to "
                    ">= 2^20 rows per device round trip, labels -> host"}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    from spark_rapids_ml_nai_b200 import _native

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch with torchrun for --gpus > 1 (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    # ---- communicator bootstrap: rank 0's NCCL id reaches the others through the host-side store ---
    uid = None
    if world > 1:
        obj = [_native.get_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        uid = obj[0]
    h = _native.Handle(device=local_rank, rank=rank, nranks=world, uid=uid)
    h.set_path({"auto": _native.PATH_AUTO, "ffma": _native.PATH_FFMA, "tc": _native.PATH_TC,
                "tc_split": _native.PATH_TC_SPLIT, "gemm": _native.PATH_GEMM, "small": _native.PATH_SMALL}[args.path])
    peaks = load_peaks()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    line = run_config(torch, dist, _native, h, args.config, args, args.steps, args.warmup, world, rank, device, peaks,
                      with_e2e=not args.no_e2e, with_cpu=not args.no_cpu, with_parity=not args.no_parity,
                      sampler=sampler)

    # ---- the other BASELINE.json configurations that this world size can hold (short runs, same contract) ----
    if args.config == "c2" and not args.no_also and not args.rows:
        also = {}
        plan = [("c4", 3, 1)]
        if world == 1:
            plan += [("c3", 3, 1), ("c1", 20, 3)]
        if world == 8:
            plan += [("c5", 2, 1)]
        for name, st, wu in plan:
            try:
                r = run_config(torch, dist, _native, h, name, args, st, wu, world, rank, device, peaks, with_e2e=False,
                               with_cpu=False, with_parity=not args.no_parity)
                also[name] = {kk: r[kk] for kk in ("metric", "value", "unit", "steps", "warmup", "ms_per_step", "config",
                                                   "roofline", "parity_check", "gpu_launches")}
            except Exception as e:  # a failure here must not lose the headline line
                also[name] = {"error": repr(e)[:300]}
        line["also"] = also

    if rank == 0:
        print(json.dumps(line), flush=True)
    h.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
