"""Oracle (test infrastructure): VectorAssembler, Summarizer, StandardScaler.

Restates SURVEY.md Appendix A.1 / A.3 (pass 0) / A.6.  Reference call site:
orangecontrib/spark/widgets/ml/spark_ml_dataset.py:569-582 (commit):
    VA = VectorAssembler(inputCols=attributes, outputCol='features')   :575
    out_df = VA.transform(in_df)                                       :576
    out_df.withColumn('label', col.cast('double'))                     :577-578
"""
import numpy as np


def assemble(columns, handle_invalid="error"):
    """A.1: concatenate the ordered scalar numeric/boolean columns, each cast to
    double, into one dense row vector per row.  handleInvalid (VectorAssembler.scala,
    `assemble(..., keepInvalid)`): 'error' (the default) rejects null AND NaN
    ("Encountered NaN while assembling a row with handleInvalid = "error""), 'keep' lets
    NaN through; infinities are ordinary values in both modes.  ('skip', which drops
    rows, is not restated: the reference never sets it.)"""
    if handle_invalid not in ("error", "keep"):
        raise ValueError("handle_invalid must be 'error' or 'keep'")
    cols = []
    n = None
    for c in columns:
        a = np.asarray(c)
        if a.dtype == object:
            if any(v is None for v in a.ravel()):
                raise ValueError("VectorAssembler: null value in input column")
            a = a.astype(np.float64)
        if a.ndim == 1:
            a = a[:, None]  # scalar column: one slot
        elif a.ndim != 2:
            raise ValueError("VectorAssembler: column must be scalar or vector valued")
        if n is None:
            n = a.shape[0]
        elif a.shape[0] != n:
            raise ValueError("VectorAssembler: ragged columns")
        a = a.astype(np.float64)
        if handle_invalid == "error" and np.isnan(a).any():
            raise ValueError("VectorAssembler: encountered NaN while assembling a row with handleInvalid='error'")
        cols.append(a)
    if not cols:
        return np.zeros((0, 0))
    return np.concatenate(cols, axis=1)


def summarize(X, w=None):
    """A.3 pass 0 / A.6: weighted mean and UNBIASED weighted variance
    sigma_j^2 = sum w (x - mu)^2 / (W - sum w^2 / W)   (n-1 for unit weights).
    Returns (mean, std, W)."""
    X = np.asarray(X, dtype=np.float64)
    n = X.shape[0]
    w = np.ones(n) if w is None else np.asarray(w, dtype=np.float64)
    W = w.sum()
    mean = (w[:, None] * X).sum(axis=0) / W
    denom = W - (w * w).sum() / W
    if denom > 0:
        var = (w[:, None] * (X - mean) ** 2).sum(axis=0) / denom
    else:
        var = np.zeros(X.shape[1])
    return mean, np.sqrt(np.maximum(var, 0.0)), W


def standard_scaler_fit(X, w=None):
    mean, std, _ = summarize(X, w)
    return mean, std


def standard_scaler_transform(X, mean, std, with_mean=False, with_std=True):
    """A.6: withStd multiplies by 1/sigma_j (0 when sigma_j == 0); withMean
    subtracts mu_j first.  Defaults match pyspark (withMean=False, withStd=True)."""
    X = np.asarray(X, dtype=np.float64)
    out = X - mean if with_mean else X.copy()
    if with_std:
        inv = np.where(std != 0.0, 1.0 / np.where(std != 0.0, std, 1.0), 0.0)
        out = out * inv
    return out
