"""Initial centres for KMeans: `random` and `k-means||` (SURVEY.md A.2 'init', 8(f) row 3).

Statistically -- not bit -- equivalent to Spark: the JVM's XORShiftRandom streams and RDD
partitioning cannot be reproduced off-JVM.  `initialModel` (exact, index-stable) is the
path used for parity.  Heavy work (distance passes) runs through b2f_kmeans_predict;
sampling uses torch device ops as plumbing; the final weighted k-means++ over the <= few
thousand candidates runs on the host.
"""
import numpy as np


def _torch():
    import torch

    return torch


def gather_rows(X, local_idx, comm):
    """Rows X[local_idx] from every rank, concatenated in rank order -> host fp64 [m, d]."""
    torch = _torch()
    dev = X.data.device
    cnt = torch.zeros(comm.world, dtype=torch.int64, device=dev)
    cnt[comm.rank] = int(local_idx.numel())
    comm.all_reduce_sum(cnt)
    cnt_h = cnt.cpu().numpy()
    total = int(cnt_h.sum())
    out = torch.zeros((total, X.d), dtype=torch.float64, device=dev)
    off = int(cnt_h[: comm.rank].sum())
    if local_idx.numel():
        out[off: off + local_idx.numel()] = X.rows(local_idx.long()).double()
    comm.all_reduce_sum(out)
    return out.cpu().numpy()


def init_random(X, n_total, k, seed, comm):
    """takeSample(withReplacement=False, k, seed): k distinct global rows."""
    torch = _torch()
    rng = np.random.RandomState(seed & 0x7FFFFFFF)
    if n_total < k:
        raise ValueError("KMeans: number of rows (%d) is smaller than k (%d)" % (n_total, k))
    chosen = []
    seen = set()
    while len(chosen) < k:
        for v in rng.randint(0, n_total, size=k - len(chosen)):
            if int(v) not in seen:
                seen.add(int(v))
                chosen.append(int(v))
    g = np.array(chosen, dtype=np.int64)
    mine = g[(g >= X.row0) & (g < X.row0 + X.n)] - X.row0
    # rows arrive in rank order; the order of random initial centres carries no meaning
    return gather_rows(X, torch.from_numpy(np.sort(mine)).to(X.data.device), comm)


def _local_kmeans_pp(points, weights, k, seed, max_iter=30):
    """LocalKMeans.kMeansPlusPlus: weighted k-means++ seeding then <= 30 Lloyd iterations (host)."""
    rng = np.random.RandomState(seed & 0x7FFFFFFF)
    m, d = points.shape
    w = np.asarray(weights, dtype=np.float64)
    centers = np.empty((k, d))
    centers[0] = points[rng.choice(m, p=w / w.sum())]
    cost = ((points - centers[0]) ** 2).sum(1)
    for i in range(1, k):
        p = cost * w
        s = p.sum()
        j = rng.choice(m, p=p / s) if s > 0 else rng.randint(m)
        centers[i] = points[j]
        cost = np.minimum(cost, ((points - centers[i]) ** 2).sum(1))
    for _ in range(max_iter):
        D = (points * points).sum(1)[:, None] - 2.0 * points @ centers.T + (centers * centers).sum(1)[None, :]
        lab = D.argmin(1)
        moved = False
        for j in range(k):
            sel = lab == j
            if sel.any():
                nc = (points[sel] * w[sel, None]).sum(0) / w[sel].sum()
                if not np.array_equal(nc, centers[j]):
                    moved = True
                centers[j] = nc
        if not moved:
            break
    return centers


def init_kmeans_parallel(X, n_total, k, seed, init_steps, comm, predict_fn):
    """k-means|| (Bahmani et al.): one random centre, `init_steps` rounds of oversampling with
    probability 2k * d^2 / cost, candidates weighted by closest-row counts, then local k-means++."""
    torch = _torch()
    dev = X.data.device
    first = init_random(X, n_total, 1, seed, comm)
    cands = first
    for step in range(init_steps):
        _, dist = predict_fn(X, cands, True)
        cost = dist.double().sum().reshape(1)
        comm.all_reduce_sum(cost)
        c = float(cost.item())
        if not c > 0:
            break
        gen = torch.Generator(device=dev)
        gen.manual_seed((seed ^ ((step + 1) << 16) ^ comm.rank) & 0x7FFFFFFFFFFF)
        u = torch.rand(X.n, device=dev, generator=gen)
        pick = torch.nonzero(u < (2.0 * k / c) * dist, as_tuple=False).flatten()
        new = gather_rows(X, pick, comm)
        if len(new):
            cands = np.concatenate([cands, new], axis=0)
    cands = np.unique(cands, axis=0)
    if len(cands) <= k:
        # not enough distinct candidates: top up with random rows (duplicates allowed, as Spark does)
        extra = init_random(X, n_total, k, seed + 1, comm)
        return np.concatenate([cands, extra], axis=0)[:k]
    lab = predict_fn(X, cands, False)
    wts = torch.bincount(lab.long(), minlength=len(cands)).double()
    comm.all_reduce_sum(wts)
    return _local_kmeans_pp(cands, np.maximum(wts.cpu().numpy(), 1e-12), k, seed)
