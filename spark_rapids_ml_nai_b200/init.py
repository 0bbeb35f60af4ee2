"""k-means|| ("scalable k-means++") initialisation over the row-sharded data.

The reference passes ``initMode="k-means||"`` straight through as cuML's ``init`` with
``oversampling_factor`` (python/src/spark_rapids_ml/clustering.py:72,117-119; default init mode, :290), and
cuML's multi-GPU fit runs the algorithm of Bahmani et al.: start from one random row; for a few rounds
sample every row with probability l*d^2(x)/phi (l = oversampling_factor * k), adding the sampled rows to
the candidate set; weight each candidate by the number of rows it is nearest to; reduce the weighted
candidates to k centres with k-means++ seeding and a short weighted Lloyd run.

The passes over X (nearest-candidate distances) run through the C ABI's predict kernel; the candidate
bookkeeping works on a few thousand rows with torch tensors on the device.  Across ranks the scalars and
the sampled rows travel through the barrier context's allGather (the control plane), as the reference
does for its own small exchanges (utils.py:206, classification.py:1009).
"""
from __future__ import annotations

import base64
import json
from typing import Any, List, Optional

import numpy as np


def _ctx(nranks: int) -> Optional[Any]:
    if nranks <= 1:
        return None
    from .runtime import BarrierTaskContext

    return BarrierTaskContext.get()


def _allgather_tensor(ctx: Any, t) -> List[Any]:
    """All ranks' 2-D float32 tensors (different row counts allowed), via base64 strings."""
    import torch

    arr = t.detach().cpu().numpy().astype(np.float32)
    msg = json.dumps({"shape": list(arr.shape), "data": base64.b64encode(arr.tobytes()).decode("ascii")})
    outs = []
    for m in ctx.allGather(msg):
        o = json.loads(m)
        a = np.frombuffer(base64.b64decode(o["data"]), dtype=np.float32).reshape(o["shape"])
        outs.append(torch.from_numpy(a.copy()).to(t.device))
    return outs


def _allreduce_sum(ctx: Any, values: np.ndarray) -> np.ndarray:
    msgs = ctx.allGather(json.dumps([float(v) for v in np.asarray(values, dtype=np.float64).ravel()]))
    return np.sum([np.array(json.loads(m), dtype=np.float64) for m in msgs], axis=0).reshape(np.shape(values))


def kmeans_parallel_init(handle: Any, X, k: int, seed: int = 1, oversampling: float = 2.0, max_rounds: int = 8):
    """Returns k initial centroids ([k,d] float32 CUDA tensor), identical on every rank."""
    import torch

    nranks, rank = handle.nranks, handle.rank
    ctx = _ctx(nranks)
    n, d = X.shape
    dev = X.device
    g = torch.Generator(device=dev).manual_seed(int(seed) * 1000003 + rank)
    g_shared = np.random.default_rng(int(seed))  # same stream on every rank

    # first centre: a uniformly random global row
    sizes = np.array([n], dtype=np.float64)
    if ctx is not None:
        sizes = np.array([float(json.loads(m)) for m in ctx.allGather(json.dumps(float(n)))])
    total = int(sizes.sum())
    if total < k:
        raise ValueError(f"k={k} exceeds the number of rows {total}")
    first = int(g_shared.integers(0, total))
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    mine = X[first - offs[rank]: first - offs[rank] + 1] if offs[rank] <= first < offs[rank + 1] else X[:0]
    cands = torch.cat(_allgather_tensor(ctx, mine), dim=0) if ctx is not None else mine.clone()

    ell = max(1.0, float(oversampling) * k)
    _, dist = handle.predict(X, cands.contiguous(), want_dist=True)
    phi = float(dist.double().sum().item())
    if ctx is not None:
        phi = float(_allreduce_sum(ctx, np.array([phi]))[0])
    rounds = int(min(max_rounds, max(1, np.ceil(np.log(max(phi, 2.0))))))
    for _ in range(rounds):
        if phi <= 0.0:
            break
        prob = torch.clamp(dist * (ell / phi), max=1.0)
        pick = torch.rand(n, generator=g, device=dev) < prob
        new = X[pick]
        if ctx is not None:
            new = torch.cat(_allgather_tensor(ctx, new), dim=0)
        if new.shape[0] == 0:
            continue
        cands = torch.cat([cands, new], dim=0)
        _, dist = handle.predict(X, cands.contiguous(), want_dist=True)
        phi = float(dist.double().sum().item())
        if ctx is not None:
            phi = float(_allreduce_sum(ctx, np.array([phi]))[0])

    m = cands.shape[0]
    if m < k:  # too few candidates (tiny data): top up with distinct random rows, same choice on all ranks
        extra_idx = g_shared.choice(total, size=k - m, replace=False)
        parts = [X[i - offs[rank]: i - offs[rank] + 1] for i in sorted(extra_idx) if offs[rank] <= i < offs[rank + 1]]
        extra = torch.cat(parts, dim=0) if parts else X[:0]
        if ctx is not None:
            extra = torch.cat(_allgather_tensor(ctx, extra), dim=0)
        return torch.cat([cands, extra], dim=0)[:k].contiguous()
    if m == k:
        return cands.contiguous()

    # weights = number of rows nearest to each candidate
    labels, _ = handle.predict(X, cands.contiguous())
    w = torch.bincount(labels.long(), minlength=m).double()
    if ctx is not None:
        w = torch.from_numpy(_allreduce_sum(ctx, w.cpu().numpy())).to(dev)

    # weighted greedy k-means++ seeding on the candidates (identical random stream on every rank): each step
    # draws a few candidates with probability ~ w*d^2 and keeps the one that lowers the weighted potential most
    C64 = cands.double()
    trials = 2 + int(np.log(k))
    chosen = [int(g_shared.choice(m, p=(w / w.sum()).cpu().numpy()))]
    d2 = ((C64 - C64[chosen[0]]) ** 2).sum(dim=1)
    for _ in range(1, k):
        p = (w * d2).cpu().numpy()
        s = p.sum()
        if s <= 0:
            chosen.append(int(g_shared.integers(0, m)))
            continue
        tries = g_shared.choice(m, size=trials, p=p / s)
        dist_try = torch.cdist(C64[torch.as_tensor(tries, device=dev)], C64) ** 2          # [trials, m]
        pot = (torch.minimum(dist_try, d2[None, :]) * w[None, :]).sum(dim=1)
        best = int(torch.argmin(pot).item())
        chosen.append(int(tries[best]))
        d2 = torch.minimum(d2, dist_try[best])
    centers = C64[chosen].clone()
    # short weighted Lloyd run over the candidates
    for _ in range(10):
        dd = torch.cdist(C64, centers) ** 2
        lab = dd.argmin(dim=1)
        sums = torch.zeros_like(centers).index_add_(0, lab, C64 * w[:, None])
        cnt = torch.zeros(k, dtype=torch.float64, device=dev).index_add_(0, lab, w)
        newc = torch.where(cnt[:, None] > 0, sums / cnt.clamp(min=1e-300)[:, None], centers)
        shift = float(((newc - centers) ** 2).sum().item())
        centers = newc
        if shift == 0.0:
            break
    return centers.float().contiguous()
