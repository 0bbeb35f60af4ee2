"""Times KMeansModel.transform's device work (b200km_predict, labels only) on C4-like data: k=100, d=64
(development aid).  usage: predict_perf.py [rows] [d] [k]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from spark_rapids_ml_nai_b200 import _native

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
k = int(sys.argv[3]) if len(sys.argv) > 3 else 100
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
centers = torch.rand((k, d), generator=g, device=dev) * 20.0 - 10.0
X = torch.empty((n, d), dtype=torch.float32, device=dev)
step = 1 << 22
for s in range(0, n, step):
    e = min(n, s + step)
    lab = torch.randint(0, k, (e - s,), generator=g, device=dev)
    X[s:e] = torch.randn((e - s, d), generator=g, device=dev) + centers[lab]
h = _native.Handle(0)
out = torch.empty(n, dtype=torch.int32, device=dev)
for _ in range(2):
    h.predict(X, centers, out=out)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
reps = 5
ev[0].record()
for _ in range(reps):
    h.predict(X, centers, out=out)
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / reps
print(f"predict n={n} d={d} k={k}: {ms:.3f} ms  {n / ms / 1e6:.2f} G rows/s  {n * (d * 4 + 4) / ms / 1e6:.0f} GB/s  "
      f"path {h.last_path()} kernel {h.screen_stats()[1]}", flush=True)
h.destroy()
