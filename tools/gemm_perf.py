"""Times the tile-GEMM assignment pass (labels only) at a given shape: python tools/gemm_perf.py n d k [path]."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from spark_rapids_ml_nai_b200 import _native

n, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
path = sys.argv[4] if len(sys.argv) > 4 else "gemm"
blobs = int(sys.argv[5]) if len(sys.argv) > 5 else k
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
centers = torch.rand((blobs, d), generator=g, device=dev) * 20 - 10
X = torch.empty((n, d), dtype=torch.float32, device=dev)
for s in range(0, n, 1 << 18):
    e = min(n, s + (1 << 18))
    which = torch.randint(0, blobs, (e - s,), generator=g, device=dev)
    X[s:e] = torch.randn((e - s, d), generator=g, device=dev) + centers[which]
idx = torch.randperm(n, generator=g, device=dev)[:k]
C = X[idx].contiguous()
h = _native.Handle(0)
h.set_path({"gemm": _native.PATH_GEMM, "ffma": _native.PATH_FFMA, "auto": _native.PATH_AUTO}[path])
out = torch.empty(n, dtype=torch.int32, device=dev)
for _ in range(2):
    h.predict(X, C, out=out)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
reps = 3
ev[0].record()
for _ in range(reps):
    h.predict(X, C, out=out)
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / reps
refined, kern = h.screen_stats()
fl = 2.0 * n * k * d
print(f"n={n} d={d} k={k} path={h.last_path()} predict {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s algorithmic  "
      f"{n * d * 4 / ms / 1e6:.0f} GB/s  open rows {refined} ({100.0 * refined / n:.2f} %)")
# fit step timing (assignment + sums)
C2, _, it, _ = h.fit(X, C, max_iter=2, tol=0.0, want_inertia=False)
torch.cuda.synchronize()
ev[0].record()
C2, _, it, _ = h.fit(X, C, max_iter=4, tol=0.0, want_inertia=False)
ev[1].record()
torch.cuda.synchronize()
print(f"fit: {ev[0].elapsed_time(ev[1]) / 4:.3f} ms per iteration")
