"""A traced 3-iteration fit on C2-like blobs from the true centres (or, with a 2nd argument, random rows) (development aid; needs a B200KM_ENABLE_TRACE build)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from spark_rapids_ml_nai_b200 import _native
n, d, k = int(os.environ.get('TRACE_ROWS', 4_000_000)), 100, 20
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
centers = torch.rand((k, d), generator=g, device=dev) * 20.0 - 10.0
X = torch.empty((n, d), dtype=torch.float32, device=dev)
for s0 in range(0, n, 1 << 20):
    e0 = min(n, s0 + (1 << 20))
    lab = torch.randint(0, k, (e0 - s0,), generator=g, device=dev)
    X[s0:e0] = torch.randn((e0 - s0, d), generator=g, device=dev) + centers[lab]
h = _native.Handle(0)
h.set_path(int(sys.argv[1]) if len(sys.argv) > 1 else _native.PATH_TC)
C = centers.contiguous() if len(sys.argv) < 3 else X[torch.randperm(n, generator=g, device=dev)[:k]].clone()
h.fit(X, C, max_iter=3, tol=0.0, want_inertia=False)  # the trace file holds the last pass (row norms read back)
torch.cuda.synchronize()
