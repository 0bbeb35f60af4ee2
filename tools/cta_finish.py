"""Spread of the per-CTA finish times written by a B200KM_ENABLE_TRACE build (development aid)."""
import sys
import numpy as np
a = np.loadtxt(sys.argv[1], dtype=np.uint64).astype(np.int64).reshape(-1)
t = a[3900:3900 + 148]
t = t[t > 0]
print(f"{len(t)} CTAs: finish spread {(t.max() - t.min()) / 1e3:.1f} us; mean lag behind the last {(t.max() - t).mean() / 1e3:.1f} us")
print("slowest SM-order positions:", np.argsort(-a[3900:3900 + 148])[:10])
