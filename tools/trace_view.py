"""Prints the per-tile event timeline written by B200KM_TRACE (development aid)."""
import sys
import numpy as np
a = np.loadtxt(sys.argv[1], dtype=np.uint64).astype(np.int64)
a = a[a[:, 7] > 0]
t0 = a[0, 0]
names = ["tma_issue", "x_arrived", "xform_done", "mma_commit", "epi_start", "labels_done", "sort_done", "released"]
print("tile " + " ".join(f"{n:>11}" for n in names))
for i in list(range(0, min(12, len(a)))) + list(range(100, min(112, len(a)))):
    print(f"{i:4d} " + " ".join(f"{int(v - t0):11d}" for v in a[i]))
d = a[100:200] if len(a) > 200 else a
print("steady-state mean cycles: tile period", np.diff(d[:, 7]).mean())
for j, k, nm in [(0, 1, "tma latency"), (1, 2, "transform"), (2, 3, "xform_done->mma_commit"), (3, 4, "commit->epi_start"),
                 (4, 5, "argmin+refine"), (5, 6, "sort"), (6, 7, "update"), (0, 7, "stage held")]:
    print(f"  {nm:26s} {(d[:, k] - d[:, j]).mean():9.0f}")
