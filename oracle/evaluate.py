"""CPU restatement of the pyspark.ml.evaluation metrics the Evaluation widget lists -- TEST INFRASTRUCTURE.

Reference call site: orangecontrib/spark/widgets/ml/spark_ml_evaluation.py:42-55 (module = evaluation, :7,19).
The arithmetic lives in Apache Spark MLlib (absent, un-pinned; see oracle/__init__.py):
  ml/evaluation/ClusteringEvaluator.scala (SquaredEuclideanSilhouette), mllib/evaluation/BinaryClassificationMetrics.scala,
  mllib/evaluation/MulticlassMetrics.scala.  PARITY UNPINNED against Spark itself; the silhouette and the ROC area are
pinned against scikit-learn in tests/test_oracle_kat.py.
"""
import numpy as np


def silhouette_sqeuclid(X, labels, w=None):
    """Spark's squared-Euclidean silhouette from per-cluster statistics:
    m_c(x) = ||x||^2 + Psi_c/N_c - 2 x.S_c/N_c;  a = m_own N_own/(N_own-1);  b = min_{c != own} m_c;
    s = 0 (singleton) | 1 - a/b (a < b) | b/a - 1 (a > b) | 0."""
    X = np.asarray(X, dtype=np.float64)
    labels = np.asarray(labels).astype(np.int64)
    n = X.shape[0]
    ww = np.ones(n) if w is None else np.asarray(w, dtype=np.float64)
    k = int(labels.max()) + 1
    xn = (X * X).sum(axis=1)
    N = np.bincount(labels, weights=ww, minlength=k)
    S = np.zeros((k, X.shape[1]))
    np.add.at(S, labels, X * ww[:, None])
    Psi = np.bincount(labels, weights=ww * xn, minlength=k)
    live = N > 0
    M = np.full((n, k), np.inf)
    M[:, live] = xn[:, None] + (Psi[live] / N[live])[None, :] - 2.0 * X @ (S[live] / N[live][:, None]).T
    M = np.maximum(M, 0.0)
    own = M[np.arange(n), labels]
    No = N[labels]
    a = np.where(No > 1, own * No / np.where(No > 1, No - 1, 1.0), -1.0)
    Mo = M.copy()
    Mo[np.arange(n), labels] = np.inf
    b = Mo.min(axis=1)
    s = np.zeros(n)
    ok = (a >= 0) & np.isfinite(b)
    lt = ok & (a < b)
    gt = ok & (a > b)
    s[lt] = 1.0 - a[lt] / b[lt]
    s[gt] = b[gt] / a[gt] - 1.0
    return float((ww * s).sum() / ww.sum())


def _binary_curve_counts(score, y, w=None):
    """Cumulative weighted true / false positive counts at each DISTINCT threshold, descending score."""
    score = np.asarray(score, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    ww = np.ones_like(score) if w is None else np.asarray(w, dtype=np.float64)
    order = np.argsort(-score, kind="stable")
    s, pos, neg = score[order], (ww * (y > 0.5))[order], (ww * (y <= 0.5))[order]
    last = np.r_[s[1:] != s[:-1], True]  # last element of each run of equal scores
    tp = np.cumsum(pos)[last]
    fp = np.cumsum(neg)[last]
    return tp, fp


def area_under_roc(score, y, w=None):
    """BinaryClassificationMetrics.areaUnderROC with numBins = 0: trapezoid over (0,0), (FPR, TPR)..., (1,1)."""
    tp, fp = _binary_curve_counts(score, y, w)
    P, Nn = tp[-1], fp[-1]
    tpr = np.r_[0.0, tp / P, 1.0]
    fpr = np.r_[0.0, fp / Nn, 1.0]
    return float(np.sum((fpr[1:] - fpr[:-1]) * (tpr[1:] + tpr[:-1]) * 0.5))


def area_under_pr(score, y, w=None):
    """BinaryClassificationMetrics.areaUnderPR with numBins = 0: trapezoid over (recall, precision) points, the
    curve starting at (0, precision of the first point)."""
    tp, fp = _binary_curve_counts(score, y, w)
    P = tp[-1]
    rec = tp / P
    prec = tp / (tp + fp)
    rec = np.r_[0.0, rec]
    prec = np.r_[prec[0], prec]
    return float(np.sum((rec[1:] - rec[:-1]) * (prec[1:] + prec[:-1]) * 0.5))


def multiclass_metrics(pred, y, w=None):
    """MulticlassMetrics: accuracy, weightedPrecision, weightedRecall, f1 (weighted by true-class frequency)."""
    pred = np.asarray(pred).astype(np.int64)
    y = np.asarray(y).astype(np.int64)
    ww = np.ones(len(y)) if w is None else np.asarray(w, dtype=np.float64)
    K = int(max(pred.max(), y.max())) + 1
    cm = np.zeros((K, K))
    np.add.at(cm, (y, pred), ww)
    tot = cm.sum()
    tp = np.diag(cm)
    by_label = cm.sum(axis=1)
    by_pred = cm.sum(axis=0)
    prec = np.where(by_pred > 0, tp / np.where(by_pred > 0, by_pred, 1.0), 0.0)
    rec = np.where(by_label > 0, tp / np.where(by_label > 0, by_label, 1.0), 0.0)
    f1 = np.where(prec + rec > 0, 2 * prec * rec / np.where(prec + rec > 0, prec + rec, 1.0), 0.0)
    wt = by_label / tot
    return {"accuracy": float(tp.sum() / tot), "weightedPrecision": float((wt * prec).sum()),
            "weightedRecall": float((wt * rec).sum()), "f1": float((wt * f1).sum())}
