"""Times single Lloyd passes (b200km_lloyd_step) on C2-like blobs (development aid).
usage: pass_perf.py [rows] [d] [k] [path]  - prints ms per pass and the float64 re-decision count for random-row
initial centroids and for the true blob centres."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from spark_rapids_ml_nai_b200 import _native

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 100
k = int(sys.argv[3]) if len(sys.argv) > 3 else 20
path = int(sys.argv[4]) if len(sys.argv) > 4 else _native.PATH_TC
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
centers = torch.rand((k, d), generator=g, device=dev) * 20.0 - 10.0
X = torch.empty((n, d), dtype=torch.float32, device=dev)
step = 1 << 20
for s in range(0, n, step):
    e = min(n, s + step)
    lab = torch.randint(0, k, (e - s,), generator=g, device=dev)
    X[s:e] = torch.randn((e - s, d), generator=g, device=dev) + centers[lab]
idx = torch.randperm(n, generator=g, device=dev)[:k]
C_rand = X[idx].clone()
h = _native.Handle(0)
h.set_path(path)
for name, C in (("random rows", C_rand), ("true centres", centers.contiguous()), ("uniform noise", None)):
    if C is None:
        X.uniform_(-1, 1, generator=g)
        C = X[idx].clone()
    for _ in range(2):
        h.lloyd_step(X, C)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    reps = 5
    ev[0].record()
    for _ in range(reps):
        h.lloyd_step(X, C)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    refined, kern = h.screen_stats()
    print(f"{name:14s}: {ms:7.3f} ms/pass  {n * d * 4 / ms / 1e6:7.1f} GB/s  refined rows/pass {refined} "
          f"({100.0 * refined / n:.3f}%)  kernel {kern}", flush=True)
h.destroy()
