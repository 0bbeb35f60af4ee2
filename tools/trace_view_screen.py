"""Prints the per-tile event timeline of the screen kernel written by B200KM_TRACE (development aid).
events: 0 TMA issued, 1 MMA sees the tile, 2 MMA committed, 4 assign sees the scores, 5 labels done,
6 masks written, 7 update starts, 3 row list built, 8 update done (stage released)."""
import sys
import numpy as np
a = np.loadtxt(sys.argv[1], dtype=np.uint64).astype(np.int64)
a = a[a[:, 8] > 0]
t0 = a[0, 0]
ev = [0, 1, 2, 4, 5, 6, 7, 3, 8]
names = ["tma_issue", "mma_start", "mma_commit", "scores_seen", "labels_done", "sort_done", "upd_start", "list_built", "upd_done"]
print("tile " + " ".join(f"{n:>11}" for n in names))
for i in list(range(0, min(10, len(a)))) + list(range(100, min(110, len(a)))):
    print(f"{i:4d} " + " ".join(f"{int(a[i][e] - t0):11d}" for e in ev))
d = a[100:200] if len(a) > 200 else a
if a.shape[1] >= 16:
    for i in range(100, min(108, len(a))):
        print(f"{i:4d} update warps done (rel. to warp 0): " + " ".join(f"{int(a[i][8 + u] - a[i][8]):6d}" for u in range(8)))
print("steady-state mean cycles: tile period", np.diff(d[:, 8]).mean())
for j, k, nm in [(0, 1, "tma issue->mma start"), (1, 2, "mma issue"), (2, 4, "commit->scores seen"), (4, 5, "argmin+refine"),
                 (5, 6, "sort"), (6, 7, "sort_done->upd_start"), (7, 3, "list build"), (3, 8, "update"), (0, 8, "stage held")]:
    print(f"  {nm:26s} {(d[:, k] - d[:, j]).mean():9.0f}")
