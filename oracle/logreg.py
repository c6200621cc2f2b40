"""Oracle (test infrastructure): MLlib LogisticRegression, fp64 numpy.

Restates SURVEY.md Appendix A.3 (ml/classification/LogisticRegression.scala,
ml/optim/aggregator/*LogisticAggregator*, ml/optim/loss/{RDDLossFunction,
L2Regularization} -- upstream Spark, un-vendored, un-pinned).  Reference call
site: `method_instance.fit(self.in_df, params=paramMap)`
orangecontrib/spark/base/spark_ml_estimator.py:22 with
`module = classification` orangecontrib/spark/widgets/ml/spark_ml_classification.py:15.

Optimisation runs in standardized space x~_j = x_j / sigma_j WITHOUT centering
(Spark <= 3.1 behaviour; matches the doctest KATs).  `coef` below always means
the standardized-space coefficients the optimizer sees.
"""
import numpy as np

from .assemble import summarize
from .optim import LBFGS, OWLQN


def _log1pexp(m):
    # stable log(1 + e^m)
    return np.where(m > 0, m + np.log1p(np.exp(-np.abs(m))), np.log1p(np.exp(-np.abs(m))))


def lr_binary_lossgrad(X, y, w, coef, intercept, inv_std, chunk=1 << 16):
    """Data part of the binary objective at standardized coefficients `coef`:
      loss = (1/W) sum w_i [log(1+e^{m_i}) - y_i m_i],  m_i = x_i . (coef*inv_std) + b
      grad_j = (1/W) sum w_i (sigma(m_i) - y_i) x_ij * inv_std_j ; grad_b = (1/W) sum w_i (sigma - y)
    Returns (loss, grad_coef[d], grad_intercept)."""
    X = np.asarray(X)
    n, d = X.shape
    ww = np.ones(n) if w is None else np.asarray(w, dtype=np.float64)
    W = ww.sum()
    beta = np.asarray(coef, dtype=np.float64) * inv_std
    loss = 0.0
    g = np.zeros(d)
    gb = 0.0
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        Xc = X[a:b].astype(np.float64, copy=False)
        yc = np.asarray(y[a:b], dtype=np.float64)
        m = Xc @ beta + intercept
        loss += float((ww[a:b] * (_log1pexp(m) - yc * m)).sum())
        mult = ww[a:b] * (1.0 / (1.0 + np.exp(-m)) - yc)
        g += mult @ Xc
        gb += float(mult.sum())
    return loss / W, g * inv_std / W, gb / W


def lr_multinomial_lossgrad(X, y, w, coef, intercepts, inv_std, chunk=1 << 15):
    """Softmax objective, coef[K,d] in standardized space, intercepts[K].
      loss = (1/W) sum w_i (log sum_k e^{m_ik} - m_{i,y_i})
      grad_k = (1/W) sum_i w_i (softmax_ik - 1[y_i = k]) x_i * inv_std"""
    X = np.asarray(X)
    n, d = X.shape
    K = coef.shape[0]
    ww = np.ones(n) if w is None else np.asarray(w, dtype=np.float64)
    W = ww.sum()
    B = (np.asarray(coef, dtype=np.float64) * inv_std[None, :]).T  # [d,K]
    loss = 0.0
    G = np.zeros((K, d))
    gb = np.zeros(K)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        Xc = X[a:b].astype(np.float64, copy=False)
        yc = np.asarray(y[a:b]).astype(np.int64)
        M = Xc @ B + intercepts[None, :]
        mx = np.maximum(M.max(axis=1), 0.0)  # "subtract max if > 0"
        E = np.exp(M - mx[:, None])
        s = E.sum(axis=1)
        rows = np.arange(b - a)
        loss += float((ww[a:b] * (np.log(s) + mx - M[rows, yc])).sum())
        P = E / s[:, None]
        P[rows, yc] -= 1.0
        P *= ww[a:b, None]
        G += P.T @ Xc
        gb += P.sum(axis=0)
    return loss / W, G * inv_std[None, :] / W, gb / W


def lr_fit(X, y, w=None, max_iter=100, reg_param=0.0, elastic_net=0.0, tol=1e-6, fit_intercept=True,
           standardization=True, family="auto", m=10, return_states=False):
    """LogisticRegression.train: summarizer pass, L-BFGS (or OWL-QN when the L1
    part is non-zero) in standardized space, then back-transform."""
    X = np.asarray(X)
    n, d = X.shape
    y = np.asarray(y)
    ww = np.ones(n) if w is None else np.asarray(w, dtype=np.float64)
    mean, std, W = summarize(X.astype(np.float64, copy=False), ww)
    num_classes = int(y.max()) + 1 if n else 0
    hist = np.bincount(y.astype(np.int64), weights=ww, minlength=max(num_classes, 2))
    if family == "auto":
        binomial = num_classes <= 2
    else:
        binomial = family == "binomial"
    if binomial and num_classes > 2:
        raise ValueError("Binomial family only supports 1 or 2 outcome classes but found %d" % num_classes)
    K = 1 if binomial else max(num_classes, 2)
    if fit_intercept and np.count_nonzero(hist) == 1:
        # A.3 / LogisticRegression.scala: "All labels are the same value and fitIntercept=true, so the coefficients
        # will be zeros. Training is not needed."  Intercept +inf on the constant class (binomial: +inf when the
        # label is always 1 -- numClasses == 2 -- and -inf when it is always 0); empty objective history.
        j = int(np.argmax(hist))
        if binomial:
            icpt = np.inf if num_classes == 2 else -np.inf
            coef = np.zeros(d)
        else:
            icpt = np.zeros(K)
            icpt[j] = np.inf
            coef = np.zeros((K, d))
        return {"coefficients": coef, "intercept": icpt, "objectiveHistory": [], "totalIterations": 0,
                "numClasses": 2 if binomial else K, "binomial": binomial, "mean": mean, "std": std, "evals": 0}
    inv_std = np.where(std != 0.0, 1.0 / np.where(std != 0.0, std, 1.0), 0.0)
    lam2 = reg_param * (1.0 - elastic_net)
    lam1 = reg_param * elastic_net
    # per-coefficient regularisation weights (intercepts: none)
    if standardization:
        l2w = np.full(d, lam2)
        l1w = np.full(d, lam1)
    else:
        l2w = lam2 * inv_std ** 2
        l1w = lam1 * inv_std
    ncoef = K * d
    nint = K if fit_intercept else 0

    def unpack(x):
        C = x[:ncoef].reshape(K, d)
        b = x[ncoef:] if fit_intercept else np.zeros(K)
        return C, b

    def f(x):
        C, b = unpack(x)
        C = C * (std != 0.0)  # constant columns: coefficient pinned to 0
        if binomial:
            loss, g, gb = lr_binary_lossgrad(X, y, ww, C[0], float(b[0]), inv_std)
            G = g[None, :]
            GB = np.array([gb])
        else:
            loss, G, GB = lr_multinomial_lossgrad(X, y, ww, C, b, inv_std)
        loss = loss + 0.5 * float((l2w[None, :] * C * C).sum())
        G = G + l2w[None, :] * C
        grad = np.concatenate([G.ravel(), GB]) if fit_intercept else G.ravel()
        return loss, grad

    x0 = np.zeros(ncoef + nint)
    if fit_intercept:
        if binomial:
            x0[ncoef] = np.log(hist[1] / hist[0])
        else:
            raw = np.log1p(hist[:K])
            x0[ncoef:] = raw - raw.mean()
    if lam1 == 0.0:
        opt = LBFGS(max_iter, m, tol)
    else:
        l1 = np.concatenate([np.tile(l1w, K), np.zeros(nint)])
        opt = OWLQN(max_iter, m, l1, tol)
    last, states = opt.minimize(f, x0)
    C, b = unpack(last["x"])
    C = C * (std != 0.0)
    coef = C * inv_std[None, :]
    b = np.array(b, dtype=np.float64)
    if not binomial:
        if fit_intercept:
            b = b - b.mean()
        if reg_param == 0.0:
            coef = coef - coef.mean(axis=0, keepdims=True)
    out = {
        "coefficients": coef[0] if binomial else coef,
        "intercept": float(b[0]) if binomial else b,
        "objectiveHistory": [s["adj_value"] for s in states],
        "totalIterations": last["iter"],
        "numClasses": 2 if binomial else K,
        "binomial": binomial,
        "mean": mean,
        "std": std,
        "evals": last["evals"],
    }
    if return_states:
        out["states"] = states
    return out


def lr_predict(X, coef, intercept, threshold=0.5):
    """A.3 prediction: binary raw=[-m,m], prob=[1-s,s], pred = 1 if s > threshold;
    multinomial raw=margins, prob=softmax, pred=argmax."""
    X = np.asarray(X, dtype=np.float64)
    coef = np.asarray(coef, dtype=np.float64)
    if coef.ndim == 1:
        m = X @ coef + intercept
        s = 1.0 / (1.0 + np.exp(-m))
        raw = np.stack([-m, m], axis=1)
        prob = np.stack([1.0 - s, s], axis=1)
        return raw, prob, (s > threshold).astype(np.float64)
    M = X @ coef.T + np.asarray(intercept)[None, :]
    E = np.exp(M - M.max(axis=1, keepdims=True))
    P = E / E.sum(axis=1, keepdims=True)
    return M, P, np.argmax(M, axis=1).astype(np.float64)
