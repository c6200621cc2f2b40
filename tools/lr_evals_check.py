"""Dev check: L-BFGS evaluation counts / objective history of the GPU fit vs the CPU oracle on the same data."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "orange3-spark_b200"))
import numpy as np

import oracle
from b200fit.classification import LogisticRegression
from b200fit.frame import ColumnarFrame

rng = np.random.default_rng(20240612)
n, d = int(sys.argv[1]) if len(sys.argv) > 1 else 200000, 256
scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d))
w = rng.standard_normal(d) / np.sqrt(d)
X = (rng.standard_normal((n, d)) * scale).astype(np.float32)
y = (rng.uniform(size=n) < 1 / (1 + np.exp(-(X @ w + 0.3)))).astype(np.int64)
r = oracle.lr_fit(X, y, None, reg_param=0.01, max_iter=100, tol=1e-6, return_states=True)
print("oracle", r["totalIterations"], [s["evals"] for s in r["states"]])
print("oracle", ["%.9f" % v for v in r["objectiveHistory"]])
m = LogisticRegression(regParam=0.01, maxIter=100, tol=1e-6).fit(ColumnarFrame({"features": X, "label": y.astype(np.float64)}))
print("gpu   ", m.summary.totalIterations, m.summary.lossGradPasses)
print("gpu   ", ["%.9f" % v for v in m.summary.objectiveHistory])
