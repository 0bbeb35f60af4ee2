"""Prints the per-tile event timeline written by B200KM_TRACE (development aid)."""
import sys
import numpy as np
a = np.loadtxt(sys.argv[1], dtype=np.uint64).astype(np.int64)
a = a[a[:, 7] > 0]
t0 = a[0, 0]
names = ["tma_issue", "x_arrived", "xform_done", "mma_commit", "epi_start", "labels_done", "sort_done", "released", "mma_dfree", "mma_a0", "mma_alast", "upd_start", "xf_w4", "xf_w5", "xf_w6", "xf_w7"]
print("tile " + " ".join(f"{n:>11}" for n in names))
for i in list(range(0, min(12, len(a)))) + list(range(100, min(112, len(a)))):
    print(f"{i:4d} " + " ".join(f"{int(v - t0):11d}" for v in a[i][:16]))
d = a[100:200] if len(a) > 200 else a
print("rows handled by update warp 0 per tile: mean", d[:,14].mean() if d.shape[1]>14 else None, "max", d[:,14].max())
print("steady-state mean cycles: tile period", np.diff(d[:, 7]).mean())
for j, k, nm in [(0, 1, "tma latency"), (1, 2, "transform"), (2, 3, "xform_done->mma_commit"), (3, 4, "commit->epi_start"),
                 (4, 5, "argmin+refine"), (5, 6, "sort"), (11, 7, "update (service)"), (6, 11, "sort_done->upd_start"), (0, 7, "stage held"), (1, 9, "x_arrived->mma_a0"), (9, 10, "mma a0->alast"), (10, 3, "alast->commit"), (8, 9, "dfree->a0")]:
    print(f"  {nm:26s} {(d[:, k] - d[:, j]).mean():9.0f}")
