"""How far are the GPU loss / gradient from the fp64 oracle's AT THE ORACLE'S OWN ITERATES?  (dev tool)

  python tools/lr_noise_check.py

Near the optimum the gradient is a small residue of large cancelling terms, so its absolute error decides which
iteration first satisfies Breeze's gradient-convergence test; this prints the error per state next to that threshold."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orange3-spark_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import gpu_util as G
import oracle
from oracle.logreg import summarize
from test_gpu_logreg import _binary_data

Xh, yh = _binary_data(5000, 12, 5)
Xh[:, 3] = 1.5
X, _, ld = G.upload_matrix(Xh)
y = torch.from_numpy(yh).to(G.dev())
n, d = Xh.shape
reg = 0.05
ref = oracle.lr_fit(Xh, yh.astype(np.int64), None, return_states=True, max_iter=10, reg_param=reg, tol=1e-6,
                    fit_intercept=False)
mean, std, W = summarize(Xh.astype(np.float64), np.ones(n))
inv_std = np.where(std != 0, 1.0 / np.where(std != 0, std, 1.0), 0.0)
for st in ref["states"]:
    C = st["x"] * (std != 0)
    loss, g, gb, _ = G.lr_lossgrad_raw(X, y, None, n, d, ld, 2, C * inv_std, 0.0)
    grad = g[0] * inv_std + reg * C
    val = loss + 0.5 * reg * float((C * C).sum())
    print("iter %d  f_oracle %.16g  f_gpu-f %.3e  |g_oracle| %.3e  |g_gpu| %.3e  |g_gpu-g_oracle| %.3e  threshold %.3e" % (
        st["iter"], st["value"], val - st["value"], np.linalg.norm(st["grad"]), np.linalg.norm(grad),
        np.linalg.norm(grad - st["grad"]), max(1e-6 * st["value"], 1e-8)))

from b200fit.classification import LogisticRegression
from b200fit.frame import ColumnarFrame

m = LogisticRegression(fitIntercept=False, regParam=reg, maxIter=10).fit(
    ColumnarFrame({"features": Xh, "label": yh.astype(np.float64)}))
s = m.summary
print("product: iterations", s.totalIterations, "passes at state", s.passesAtState)
print("product objective history", s.objectiveHistory)
