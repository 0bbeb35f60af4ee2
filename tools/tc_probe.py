"""GPU probe of the tcgen05 path on small shapes (development aid; prints error statistics)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import kmeans_oracle as ko
from spark_rapids_ml_nai_b200 import _native

shapes = [(1000, 64, 16), (128, 32, 16), (4096, 100, 20), (5000, 64, 100), (3000, 128, 128), (2000, 256, 200), (777, 36, 5)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
h = _native.Handle(0)
h.set_path(_native.PATH_TC)
for (n, d, k) in shapes:
    rng = np.random.default_rng(n + d + k)
    X, _ = ko.make_blobs(n, d, k, seed=n)
    C = ko.random_init(X, k, seed=3)
    Xd, Cd = torch.from_numpy(X).cuda(), torch.from_numpy(C).cuda()
    t0 = time.time()
    labels, dist = h.predict(Xd, Cd, want_dist=True)
    torch.cuda.synchronize()
    l_ref, best, second = ko.assign(X, C)
    lab = labels.cpu().numpy(); dd = dist.cpu().numpy()
    amb = ko.ambiguous_rows(best, second)
    bad = (lab != l_ref) & ~amb
    print(f"predict n={n} d={d} k={k}: path={h.last_path()} mismatches={bad.sum()} (ambiguous {amb.sum()}) "
          f"max|dist-best|={np.abs(dd - best).max():.3e} max best={best.max():.3e} t={time.time()-t0:.3f}s", flush=True)
    lab_s, _ = h.predict(Xd, Cd, want_dist=False)   # labels only: the screen kernel
    torch.cuda.synchronize()
    bad_s = (lab_s.cpu().numpy() != l_ref) & ~amb
    print(f"  screen predict: mismatches={bad_s.sum()}", flush=True)
    sums, counts, inertia, lab2 = h.lloyd_step(Xd, Cd, want_labels=True)
    torch.cuda.synchronize()
    lab2 = lab2.cpu().numpy()
    s_own, c_own = ko.accumulate(X, lab2, k)
    print(f"  step: labels==predict {np.array_equal(lab2, lab)} counts ok {np.array_equal(counts.cpu().numpy(), c_own)} "
          f"sums maxerr {np.abs(sums.cpu().numpy() - s_own).max():.3e} inertia {inertia.item():.6e} ref {best.sum():.6e}", flush=True)
h.destroy()
print("probe done")
