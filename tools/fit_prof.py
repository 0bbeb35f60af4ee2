"""A 3-iteration fit on C2-like blobs from random-row centroids (development aid: `ncu -k regex:kmeans_screen -s 1 -c 1`
then profiles a later pass, which reads the row norms back instead of computing them)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from spark_rapids_ml_nai_b200 import _native
n, d, k = 4_000_000, 100, 20
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
centers = torch.rand((k, d), generator=g, device=dev) * 20.0 - 10.0
lab = torch.randint(0, k, (n,), generator=g, device=dev)
X = torch.randn((n, d), generator=g, device=dev) + centers[lab]
h = _native.Handle(0)
h.set_path(int(sys.argv[1]) if len(sys.argv) > 1 else _native.PATH_TC)
C = X[torch.randperm(n, generator=g, device=dev)[:k]].clone()
h.fit(X, C, max_iter=3, tol=0.0)
torch.cuda.synchronize()
