import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import kmeans_oracle as ko
from spark_rapids_ml_nai_b200 import _native
from spark_rapids_ml_nai_b200.init import kmeans_parallel_init
h = _native.Handle(0)
X, tc = ko.make_blobs(20000, 16, 10, seed=4)
Xd = torch.from_numpy(X).cuda()
C = kmeans_parallel_init(h, Xd, 10, seed=7)
print("init covers blobs:", sorted(ko.predict(tc, C.cpu().numpy()).tolist()))
ref = ko.lloyd_fit(X, C.cpu().numpy(), max_iter=20, tol=1e-4)
print("oracle from this init: inertia", ref.inertia, "iters", ref.n_iter)
for path in (_native.PATH_FFMA, _native.PATH_TC):
    h.set_path(path)
    Cf, inertia, it, _ = h.fit(Xd, C, max_iter=20, tol=1e-4)
    print(h.last_path(), "fit: inertia", inertia, "iters", it, "max|C-ref|", np.abs(Cf.cpu().numpy() - ref.centers).max())
    for mi in (1, 2, 3):
        Cf, inertia, it, _ = h.fit(Xd, C, max_iter=mi, tol=1e-20)
        r = ko.lloyd_fit(X, C.cpu().numpy(), max_iter=mi, tol=1e-20)
        print("   max_iter", mi, "iters", it, "inertia", inertia, "ref", r.inertia, "max|dC|", np.abs(Cf.cpu().numpy() - r.centers).max())
