"""Oracle (test infrastructure): MLlib KMeans Lloyd iterations, fp64 numpy.

Restates SURVEY.md Appendix A.2 (mllib/clustering/KMeans.scala,
DistanceMeasure.scala, util/MLUtils.scala -- upstream Spark, un-vendored and
un-pinned).  Reference call site: `method_instance.fit(self.in_df, params=...)`
orangecontrib/spark/base/spark_ml_estimator.py:22 with
`module = clustering` orangecontrib/spark/widgets/ml/spark_ml_clustering.py:14.
"""
import numpy as np

EPS = np.finfo(np.float64).eps  # 2.22e-16, MLUtils.EPSILON


def fast_squared_distance(v1, n1, v2, n2, precision=1e-6):
    """MLUtils.fastSquaredDistance, dense case: use n1^2 + n2^2 - 2 v1.v2 when its
    relative error bound is below `precision`, else the exact sum (v1-v2)^2."""
    s = n1 * n1 + n2 * n2
    nd = n1 - n2
    bound = 2.0 * EPS * s / (nd * nd + EPS)
    if bound < precision:
        d = s - 2.0 * float(np.dot(v1, v2))
    else:
        diff = np.asarray(v1, dtype=np.float64) - np.asarray(v2, dtype=np.float64)
        d = float(np.dot(diff, diff))
    return d


def find_closest_scalar(x, xn, C, Cn):
    """DistanceMeasure.findClosest, literal loop (small cases only): centres in
    index order, lower-bound skip, strict '<' so ties keep the lowest index."""
    best = np.inf
    bi = 0
    for j in range(C.shape[0]):
        lb = Cn[j] - xn
        lb = lb * lb
        if lb < best:
            d = fast_squared_distance(C[j], Cn[j], x, xn)
            if d < best:
                best = d
                bi = j
    return bi, best


def kmeans_assign(X, C, Xn=None, chunk=16384, return_second=False):
    """Vectorised findClosest over all rows (result-equivalent to the scalar loop:
    the lower-bound skip can never change an argmin).  Returns labels (int32) and
    squared distances (fp64); optionally the second-best distance per row (used by
    tests to decide which rows are allowed to disagree: top-2 gap < 1e-6 rel)."""
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    n = X.shape[0]
    Cn = np.sqrt((C * C).sum(axis=1))
    labels = np.empty(n, dtype=np.int32)
    dist = np.empty(n, dtype=np.float64)
    second = np.empty(n, dtype=np.float64) if return_second else None
    Cn2 = Cn * Cn
    order = np.argsort(Cn, kind="stable")
    Cn_sorted = Cn[order]
    cn_max = Cn.max() if len(Cn) else 0.0
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        Xc = X[a:b]
        xn = np.sqrt((Xc * Xc).sum(axis=1)) if Xn is None else Xn[a:b]
        D = Xc @ C.T
        D *= -2.0
        D += Cn2[None, :]
        D += (xn * xn)[:, None]
        # fastSquaredDistance switches to the exact form when bound = 2 EPS s / (nd^2 + EPS) >= 1e-6, i.e. when
        # the two norms agree to ~2e-5 relative.  Find those (rare) pairs through the sorted centre norms
        # instead of forming the full bound matrix, then apply the literal test to them.
        thr = 2.2e-5 * (xn + cn_max) + 1e-300
        lo = np.searchsorted(Cn_sorted, xn - thr, side="left")
        hi = np.searchsorted(Cn_sorted, xn + thr, side="right")
        for i in np.nonzero(hi > lo)[0]:
            for j in order[lo[i]:hi[i]]:
                s_ = xn[i] * xn[i] + Cn2[j]
                nd = xn[i] - Cn[j]
                if 2.0 * EPS * s_ / (nd * nd + EPS) >= 1e-6:
                    diff = Xc[i] - C[j]
                    D[i, j] = diff @ diff
        lab = np.argmin(D, axis=1)  # first minimum == lowest index on ties
        labels[a:b] = lab
        dist[a:b] = D[np.arange(b - a), lab]
        if return_second:
            if C.shape[0] > 1:
                D[np.arange(b - a), lab] = np.inf
                second[a:b] = D.min(axis=1)
            else:
                second[a:b] = np.inf
    if return_second:
        return labels, dist, second
    return labels, dist


def kmeans_step(X, C, w=None, chunk=16384):
    """One Lloyd assignment+accumulate pass: labels, cost, sums[k,d], counts[k]."""
    X = np.asarray(X, dtype=np.float64)
    k, d = C.shape
    labels, dist = kmeans_assign(X, C, chunk=chunk)
    ww = np.ones(X.shape[0]) if w is None else np.asarray(w, dtype=np.float64)
    cost = float((np.maximum(dist, 0.0) * ww).sum())
    # sums[j] = sum of w_i x_i over rows with label j: a one-hot (k x n) sparse product, the same
    # block formulation Spark >= 3.4 hands to BLAS (far faster than np.add.at, identical result)
    n = X.shape[0]
    try:
        from scipy import sparse

        onehot = sparse.csr_matrix((ww, (labels, np.arange(n))), shape=(k, n))
        sums = np.asarray(onehot @ X)
    except ImportError:  # pragma: no cover
        sums = np.zeros((k, d))
        np.add.at(sums, labels, X * ww[:, None])
    counts = np.bincount(labels, weights=ww, minlength=k).astype(np.float64)
    return labels, cost, sums, counts


def kmeans_update(C, sums, counts, tol):
    """Centre update after the global reduce: clusters with count > 0 move to
    sums/count, EMPTY clusters keep their previous centre; converged iff every
    moved centre moved by <= tol (squared distance <= tol^2)."""
    newC = C.copy()
    converged = True
    for j in range(C.shape[0]):
        if counts[j] > 0:
            nc = sums[j] / counts[j]
            diff = nc - C[j]
            if float(diff @ diff) > tol * tol:
                converged = False
            newC[j] = nc
    return newC, converged


def kmeans_lloyd(X, C0, max_iter=20, tol=1e-4, w=None, chunk=16384):
    """A.2 iteration loop with a fixed initial model.  Returns a dict with
    per-iteration centres (after each update), costs (cost of each assignment
    step, i.e. measured before that iteration's update), counts, numIter and
    trainingCost (= cost of the last executed iteration's assignment)."""
    C = np.asarray(C0, dtype=np.float64).copy()
    centers_hist, cost_hist, counts_hist = [], [], []
    labels = None
    it = 0
    converged = False
    while it < max_iter and not converged:
        labels, cost, sums, counts = kmeans_step(X, C, w=w, chunk=chunk)
        C, converged = kmeans_update(C, sums, counts, tol)
        centers_hist.append(C.copy())
        cost_hist.append(cost)
        counts_hist.append(counts)
        it += 1
    return {
        "centers": C,
        "centers_hist": centers_hist,
        "cost_hist": cost_hist,
        "counts_hist": counts_hist,
        "labels_last_assign": labels,
        "numIter": it,
        "trainingCost": cost_hist[-1] if cost_hist else float("nan"),
        "converged": converged,
    }
