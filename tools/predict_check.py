import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import kmeans_oracle as ko
from spark_rapids_ml_nai_b200 import _native
h = _native.Handle(0)
X, tc = ko.make_blobs(20000, 16, 10, seed=4)
Xd = torch.from_numpy(X).cuda()
for k in [1, 2, 21, 45, 100, 161, 200]:
    C = X[np.random.default_rng(k).choice(20000, k, replace=False)]
    for path in [_native.PATH_FFMA, _native.PATH_AUTO]:
        h.set_path(path)
        lab, dist = h.predict(Xd, torch.from_numpy(C).cuda(), want_dist=True)
        l_ref, best, second = ko.assign(X, C)
        amb = ko.ambiguous_rows(best, second)
        bad = ((lab.cpu().numpy() != l_ref) & ~amb).sum()
        print(f"k={k} path={h.last_path()} mismatches={bad} max|dist err|={np.abs(dist.cpu().numpy()-best).max():.3e} sum dist={dist.double().sum().item():.6e} ref={best.sum():.6e}")
